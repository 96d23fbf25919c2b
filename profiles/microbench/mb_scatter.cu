// mb_scatter.cu -- design-input microbenchmark for the partition scatter: cost of writing 2^28
// 16-byte tuples to pseudo-random places of a 4 GiB buffer, as a function of the write granularity.
//   a  16 B per lane, every lane its own random 16-byte slot          (partial sector writes)
//   b  32 B per lane pair: lanes 2i, 2i+1 write the halves of one random sector
//   c  64 B per 4 lanes, d  128 B per 8 lanes (one full line)
//   e  one st.global.v8.b32 (256-bit) per lane to a random sector      (half the lanes, 2 tuples each)
//   s  sequential 16 B per lane (the streaming ceiling)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_scatter mb_scatter.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL; x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL; x ^= x >> 32; return x;
}

// GROUP lanes share one random block of GROUP*16 bytes; GROUP = 0 means sequential.
template <int GROUP>
__global__ void __launch_bounds__(256) k_scatter(uint4* out, uint64_t n) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t dst;
        if (GROUP == 0) dst = i;
        else {
            const uint64_t blk = i / GROUP;                       // a bijection on block indices: odd multiplier mod 2^k
            const uint64_t nblk = n / GROUP;                      // n is a power of two
            dst = ((blk * 0x9E3779B97F4A7C15ULL + 12345) & (nblk - 1)) * GROUP + (i % GROUP);
        }
        out[dst] = make_uint4((uint32_t)i, 1u, 2u, 3u);
    }
}
__global__ void __launch_bounds__(256) k_scatter_v8(uint4* out, uint64_t n) {
    const uint64_t npair = n / 2;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < npair; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t dst = ((i * 0x9E3779B97F4A7C15ULL + 12345) & (npair - 1)) * 2;
        uint32_t a = (uint32_t)i;
        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                     :: "l"(out + dst), "r"(a), "r"(a + 1), "r"(a + 2), "r"(a + 3), "r"(a + 4), "r"(a + 5), "r"(a + 6), "r"(a + 7) : "memory");
    }
}
// random 16-byte reads (gather), and 32-byte reads by lane pairs, for the phase-2 side
template <int GROUP>
__global__ void __launch_bounds__(256) k_gather(const uint4* in, uint64_t n, unsigned long long* out) {
    uint32_t acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t blk = i / GROUP, nblk = n / GROUP;
        const uint64_t src = ((blk * 0x9E3779B97F4A7C15ULL + 12345) & (nblk - 1)) * GROUP + (i % GROUP);
        const uint4 v = in[src];
        acc += v.x ^ v.w;
    }
    if (acc == 0x12345u) *out = acc;
}

template <class F> float timeit(F f, int reps = 3) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    const uint64_t n = 1ull << 28;          // 4 GiB of tuples
    uint4* buf; unsigned long long* out;
    CK(cudaMalloc(&buf, n * 16)); CK(cudaMalloc(&out, 8));
    CK(cudaMemset(buf, 0, n * 16));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * 16;
    auto rep = [&](const char* name, float ms) { printf("%-44s %8.3f ms  %8.2f Gtuples/s  %7.1f GB/s\n", name, ms, n / ms * 1e-6, n * 16 / ms * 1e-6); fflush(stdout); };
    rep("sequential 16 B/lane", timeit([&] { k_scatter<0><<<grid, 256>>>(buf, n); }));
    rep("random 16 B (1 lane)", timeit([&] { k_scatter<1><<<grid, 256>>>(buf, n); }));
    rep("random 32 B (2 lanes, full sector)", timeit([&] { k_scatter<2><<<grid, 256>>>(buf, n); }));
    rep("random 64 B (4 lanes)", timeit([&] { k_scatter<4><<<grid, 256>>>(buf, n); }));
    rep("random 128 B (8 lanes, full line)", timeit([&] { k_scatter<8><<<grid, 256>>>(buf, n); }));
    rep("random 512 B (32 lanes)", timeit([&] { k_scatter<32><<<grid, 256>>>(buf, n); }));
    rep("random 32 B, one 256-bit store per lane", timeit([&] { k_scatter_v8<<<grid, 256>>>(buf, n); }));
    rep("gather random 16 B", timeit([&] { k_gather<1><<<grid, 256>>>(buf, n, out); }));
    rep("gather random 32 B (2 lanes)", timeit([&] { k_gather<2><<<grid, 256>>>(buf, n, out); }));
    rep("gather random 128 B (8 lanes)", timeit([&] { k_gather<8><<<grid, 256>>>(buf, n, out); }));
    CK(cudaDeviceSynchronize());
    printf("done\n");
    return 0;
}
