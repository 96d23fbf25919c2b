// mb_random.cu -- design-input microbenchmark: random read-modify-write into a table of 32-byte
// slots as a function of the table size (L2-resident .. far beyond the TLB reach), with the access
// patterns the global group table uses.  Not part of the library.
//   A  red:        NACC x RED.ADD.u64 into the row's slot (fire and forget)
//   B  probe-seq:  ld slot word 0 -> (CAS if empty) -> NACC x RED, one row after the other (4 rows/thread)
//   C  probe-ilp:  the 4 loads of a thread issued together, then the 4 CAS, then the REDs
//   D  ld-only:    one 8-byte load per row from a random slot (no write-back traffic)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mb_random mb_random.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL; x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL; x ^= x >> 32; return x;
}
// slot of row i: uniform in [0, nslots)
__device__ __forceinline__ size_t slot_of(uint64_t i, uint64_t nslots) {
    return (size_t)(((unsigned __int128)mix64(i + 0x9e3779b97f4a7c15ULL) * nslots) >> 64);
}

template <int NACC>
__global__ void __launch_bounds__(256) k_red(unsigned long long* tab, uint64_t nslots, uint64_t n) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        unsigned long long* s = tab + slot_of(i, nslots) * 4;
#pragma unroll
        for (int a = 0; a < NACC; ++a) atomicAdd(s + 1 + a, 1ull);
    }
}

template <int NACC, bool ILP>
__global__ void __launch_bounds__(256) k_probe(unsigned long long* tab, uint64_t nslots, uint64_t n) {
    const uint64_t nq = n / 4;
    for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < nq; q += (uint64_t)gridDim.x * blockDim.x) {
        unsigned long long* s[4];
        unsigned long long cur[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) s[r] = tab + slot_of(q * 4 + r, nslots) * 4;
        if (ILP) {
#pragma unroll
            for (int r = 0; r < 4; ++r) cur[r] = *reinterpret_cast<volatile unsigned long long*>(s[r]);
#pragma unroll
            for (int r = 0; r < 4; ++r) if (cur[r] == 0ull) atomicCAS(s[r], 0ull, q * 4 + r + 1);
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int a = 0; a < NACC; ++a) atomicAdd(s[r] + 1 + a, 1ull);
        } else {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                cur[r] = *reinterpret_cast<volatile unsigned long long*>(s[r]);
                if (cur[r] == 0ull) { if (atomicCAS(s[r], 0ull, q * 4 + r + 1) == 12345ull) return; }
#pragma unroll
                for (int a = 0; a < NACC; ++a) atomicAdd(s[r] + 1 + a, 1ull);
            }
        }
    }
}

__global__ void __launch_bounds__(256) k_ld(const unsigned long long* tab, uint64_t nslots, uint64_t n, unsigned long long* out) {
    unsigned long long acc = 0;
    const uint64_t nq = n / 4;
    for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < nq; q += (uint64_t)gridDim.x * blockDim.x) {
        unsigned long long v[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) v[r] = __ldg(tab + slot_of(q * 4 + r, nslots) * 4);
        acc += v[0] ^ v[1] ^ v[2] ^ v[3];
    }
    if (acc == 0x123456789ull) *out = acc;
}

template <class F> float timeit(F f, int reps = 2) {
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
    const uint64_t n = 1ull << 28;
    const size_t max_bytes = (size_t)64 << 30;
    unsigned long long* tab; unsigned long long* out;
    CK(cudaMalloc(&tab, max_bytes)); CK(cudaMalloc(&out, 8));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms * 8;
    printf("%-12s %10s %10s %10s %10s %10s %10s   (Grows/s, n = 2^28 rows, 32-byte slots)\n", "table", "red x1", "red x3", "probe-seq3", "probe-ilp3", "probe-ilp1", "ld-only");
    for (size_t mb : {32ull, 64ull, 128ull, 256ull, 512ull, 1024ull, 4096ull, 16384ull, 65536ull}) {
        const size_t bytes = mb << 20;
        const uint64_t nslots = bytes / 32;
        CK(cudaMemset(tab, 0, bytes));
        float t1 = timeit([&] { k_red<1><<<grid, 256>>>(tab, nslots, n); });
        float t3 = timeit([&] { k_red<3><<<grid, 256>>>(tab, nslots, n); });
        CK(cudaMemset(tab, 0, bytes));
        float ps = timeit([&] { k_probe<3, false><<<grid, 256>>>(tab, nslots, n); });
        CK(cudaMemset(tab, 0, bytes));
        float pi = timeit([&] { k_probe<3, true><<<grid, 256>>>(tab, nslots, n); });
        CK(cudaMemset(tab, 0, bytes));
        float p1 = timeit([&] { k_probe<1, true><<<grid, 256>>>(tab, nslots, n); });
        float ld = timeit([&] { k_ld<<<grid, 256>>>(tab, nslots, n, out); });
        CK(cudaGetLastError());
        printf("%8zu MiB %10.2f %10.2f %10.2f %10.2f %10.2f %10.2f\n", mb, n / t1 * 1e-6, n / t3 * 1e-6, n / ps * 1e-6, n / pi * 1e-6, n / p1 * 1e-6, n / ld * 1e-6);
        fflush(stdout);
    }
    CK(cudaDeviceSynchronize());
    printf("done\n");
    return 0;
}
