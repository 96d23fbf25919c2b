// mb_frontier.cu -- design-input microbenchmark for k_hc_scatter: n 32-byte tuples appended to P = n/3600
// hash partitions (one L2 atomic per tuple for the position), two layouts of the partition storage:
//   a  fixed   partition p owns tuples [p * 4096, (p + 1) * 4096): P frontiers spread over the whole buffer
//   b  chunked partitions take 16-tuple (512 B) chunks from one bump allocator: all frontiers live in a
//              compact window that moves through the buffer (TLB / DRAM-row locality)
//   s  sequential 32 B per thread (streaming ceiling)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_frontier mb_frontier.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
__device__ __forceinline__ uint64_t mix64(uint64_t x) { x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL; x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL; x ^= x >> 32; return x; }
__device__ __forceinline__ void st32(uint32_t* d, uint32_t a) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %1, %1, %1};" :: "l"(d), "r"(a) : "memory");
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %1, %1, %1};" :: "l"(d + 4), "r"(a) : "memory");
}
constexpr uint32_t CAP = 4096, CH = 16, NCH = CAP / CH;
template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t* tuples, uint32_t* cursor, uint32_t* index, uint32_t* next_chunk, uint32_t n, uint32_t P) {
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n / 4; q += gridDim.x * blockDim.x) {
        uint32_t p[4], pos[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) { p[r] = __umulhi((uint32_t)(mix64(q * 4ull + r) >> 32), P); pos[r] = MODE == 2 ? 0 : (MODE == 4 ? (uint32_t)(mix64(q * 4ull + r + 77) & 4095u) : atomicAdd(cursor + p[r], 1u)); }
        if (MODE == 3) {                                          // atomics, but sequential stores
#pragma unroll
            for (int r = 0; r < 4; ++r) st32(tuples + (size_t(q) * 4 + r) * 8, pos[r]);
        } else if (MODE == 2) {
#pragma unroll
            for (int r = 0; r < 4; ++r) st32(tuples + (size_t(q) * 4 + r) * 8, q);
        } else if (MODE == 0 || MODE == 4) {
#pragma unroll
            for (int r = 0; r < 4; ++r) if (pos[r] < CAP) st32(tuples + (size_t(p[r]) * CAP + pos[r]) * 8, q);
        } else {
            uint32_t ch[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {                         // allocators first: nobody waits before allocating
                ch[r] = 0xffffffffu;
                if (pos[r] < CAP && pos[r] % CH == 0) { ch[r] = atomicAdd(next_chunk, 1u); *reinterpret_cast<volatile uint32_t*>(index + size_t(p[r]) * NCH + pos[r] / CH) = ch[r]; }
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (pos[r] >= CAP) continue;
                if (ch[r] == 0xffffffffu) { volatile uint32_t* ip = index + size_t(p[r]) * NCH + pos[r] / CH; do { ch[r] = *ip; } while (ch[r] == 0xffffffffu); }
                st32(tuples + (size_t(ch[r]) * CH + pos[r] % CH) * 8, q);
            }
        }
    }
}
int main(int argc, char** argv) {
    const uint32_t n = argc > 1 ? (uint32_t)atof(argv[1]) : (1u << 29);
    const uint32_t P = n / 3600;
    uint32_t *tuples, *cursor, *index, *next;
    const size_t tbytes = size_t(P) * CAP * 32;
    CK(cudaMalloc(&tuples, tbytes)); CK(cudaMalloc(&cursor, size_t(P) * 4)); CK(cudaMalloc(&index, size_t(P) * NCH * 4)); CK(cudaMalloc(&next, 4));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const char* names[5] = {"fixed partitions", "chunked (16-tuple chunks, bump allocator)", "sequential", "atomics + sequential stores", "no atomics + random sector stores"};
    for (int mode = 0; mode < 5; ++mode) for (int rep = 0; rep < 2; ++rep) {
        CK(cudaMemset(cursor, 0, size_t(P) * 4)); CK(cudaMemset(index, 0xff, size_t(P) * NCH * 4)); CK(cudaMemset(next, 0, 4));
        cudaEventRecord(a);
        if (mode == 0) k<0><<<148 * 16, 256>>>(tuples, cursor, index, next, n, P);
        if (mode == 1) k<1><<<148 * 16, 256>>>(tuples, cursor, index, next, n, P);
        if (mode == 2) k<2><<<148 * 16, 256>>>(tuples, cursor, index, next, n, P);
        if (mode == 3) k<3><<<148 * 16, 256>>>(tuples, cursor, index, next, n, P);
        if (mode == 4) k<4><<<148 * 16, 256>>>(tuples, cursor, index, next, n, P);
        cudaEventRecord(b); CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep) printf("%-45s n=%u P=%u  %8.3f ms  %7.2f Gtuples/s  %7.1f GB/s\n", names[mode], n, P, ms, n / ms * 1e-6, n * 32.0 / ms * 1e-6);
    }
    return 0;
}
