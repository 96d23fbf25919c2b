// mb_atomics.cu -- design-input microbenchmarks for the group-by kernels (not part of the library).
// Measures, full chip, rows/s of: shared-memory atomics, lane-private shared RMW, global RED to an
// L2-resident / DRAM-resident table, and __match_any_sync.  Build: nvcc -gencode
// arch=compute_100a,code=sm_100a -O3 -o mb_atomics mb_atomics.cu ; run: ./mb_atomics
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}

// keys: n uint32 in [0, G)
__global__ void k_gen(uint32_t* k, size_t n, uint32_t G) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        k[i] = (uint32_t)(((uint64_t)mix32((uint32_t)i * 2654435761u + 12345u) * G) >> 32);
}

// (1) shared atomics, 32-bit add, table of G cells per CTA
template <int NACC>
__global__ void __launch_bounds__(512) k_smem_atomic32(const uint32_t* __restrict__ k, size_t n, uint32_t G, unsigned long long* out) {
    extern __shared__ uint32_t tab[];
    for (uint32_t i = threadIdx.x; i < G * NACC; i += blockDim.x) tab[i] = 0;
    __syncthreads();
    const size_t nq = n / 4;
    for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < nq; q += (size_t)gridDim.x * blockDim.x) {
        const uint4 kk = reinterpret_cast<const uint4*>(k)[q];
#pragma unroll
        for (int a = 0; a < NACC; ++a) {
            atomicAdd(&tab[kk.x * NACC + a], kk.x + a); atomicAdd(&tab[kk.y * NACC + a], kk.y + a);
            atomicAdd(&tab[kk.z * NACC + a], kk.z + a); atomicAdd(&tab[kk.w * NACC + a], kk.w + a);
        }
    }
    __syncthreads();
    unsigned long long s = 0;
    for (uint32_t i = threadIdx.x; i < G * NACC; i += blockDim.x) s += tab[i];
    if (s) atomicAdd(out, s);
}

// (2) shared atomics, 64-bit add
__global__ void __launch_bounds__(512) k_smem_atomic64(const uint32_t* __restrict__ k, size_t n, uint32_t G, unsigned long long* out) {
    extern __shared__ unsigned long long tab64[];
    for (uint32_t i = threadIdx.x; i < G; i += blockDim.x) tab64[i] = 0;
    __syncthreads();
    const size_t nq = n / 4;
    for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < nq; q += (size_t)gridDim.x * blockDim.x) {
        const uint4 kk = reinterpret_cast<const uint4*>(k)[q];
        atomicAdd(&tab64[kk.x], (unsigned long long)kk.x); atomicAdd(&tab64[kk.y], (unsigned long long)kk.y);
        atomicAdd(&tab64[kk.z], (unsigned long long)kk.z); atomicAdd(&tab64[kk.w], (unsigned long long)kk.w);
    }
    __syncthreads();
    unsigned long long s = 0;
    for (uint32_t i = threadIdx.x; i < G; i += blockDim.x) s += tab64[i];
    if (s) atomicAdd(out, s);
}

// (3) lane-private columns: tab[g * T + t], plain LDS/ADD/STS, 4 rows per thread per iteration with a
// duplicate check (equal keys inside the quad are chained through registers)
template <int T>
__global__ void __launch_bounds__(T) k_smem_private(const uint32_t* __restrict__ k, size_t n, uint32_t G, unsigned long long* out) {
    extern __shared__ uint32_t tab[];
    for (uint32_t i = threadIdx.x; i < G * T; i += T) tab[i] = 0;
    __syncthreads();
    const size_t nq = n / 4;
    const uint32_t t = threadIdx.x;
    for (size_t q = blockIdx.x * (size_t)T + t; q < nq; q += (size_t)gridDim.x * T) {
        const uint4 kk = reinterpret_cast<const uint4*>(k)[q];
        uint32_t* p0 = &tab[kk.x * T + t]; uint32_t* p1 = &tab[kk.y * T + t];
        uint32_t* p2 = &tab[kk.z * T + t]; uint32_t* p3 = &tab[kk.w * T + t];
        if (kk.x != kk.y && kk.x != kk.z && kk.x != kk.w && kk.y != kk.z && kk.y != kk.w && kk.z != kk.w) {
            const uint32_t a0 = *p0, a1 = *p1, a2 = *p2, a3 = *p3;
            *p0 = a0 + kk.x; *p1 = a1 + kk.y; *p2 = a2 + kk.z; *p3 = a3 + kk.w;
        } else {
            *(volatile uint32_t*)p0 += kk.x; *(volatile uint32_t*)p1 += kk.y;
            *(volatile uint32_t*)p2 += kk.z; *(volatile uint32_t*)p3 += kk.w;
        }
    }
    __syncthreads();
    unsigned long long s = 0;
    for (uint32_t i = threadIdx.x; i < G * T; i += T) s += tab[i];
    if (s) atomicAdd(out, s);
}

// (4) global RED 64-bit add into a table of G slots of `stride` u64 words, NACC reds per row
template <int NACC>
__global__ void __launch_bounds__(256) k_global_red(const uint32_t* __restrict__ k, size_t n, unsigned long long* tab, uint32_t stride) {
    const size_t nq = n / 4;
    for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < nq; q += (size_t)gridDim.x * blockDim.x) {
        const uint4 kk = reinterpret_cast<const uint4*>(k)[q];
#pragma unroll
        for (int a = 0; a < NACC; ++a) {
            atomicAdd(tab + (size_t)kk.x * stride + a, 1ull); atomicAdd(tab + (size_t)kk.y * stride + a, 1ull);
            atomicAdd(tab + (size_t)kk.z * stride + a, 1ull); atomicAdd(tab + (size_t)kk.w * stride + a, 1ull);
        }
    }
}

// (5) match_any throughput
__global__ void __launch_bounds__(256) k_match(const uint32_t* __restrict__ k, size_t n, unsigned long long* out) {
    unsigned acc = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        acc += __popc(__match_any_sync(0xffffffffu, k[i]));
    if (acc == 0xffffffffu) atomicAdd(out, 1ull);
}

// (6) plain streaming read (the ceiling)
__global__ void __launch_bounds__(256) k_read(const uint32_t* __restrict__ k, size_t n, unsigned long long* out) {
    unsigned acc = 0;
    const size_t nq = n / 4;
    for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < nq; q += (size_t)gridDim.x * blockDim.x) {
        const uint4 kk = reinterpret_cast<const uint4*>(k)[q];
        acc += kk.x ^ kk.y ^ kk.z ^ kk.w;
    }
    if (acc == 0x12345678u) atomicAdd(out, 1ull);
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
    const size_t n = 1ull << 28;   // 268M keys = 1 GiB
    uint32_t* k; unsigned long long* out; unsigned long long* tab;
    CK(cudaMalloc(&k, n * 4)); CK(cudaMalloc(&out, 8)); CK(cudaMemset(out, 0, 8));
    const size_t tab_bytes = (size_t)1 << 31;
    CK(cudaMalloc(&tab, tab_bytes));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    auto report = [&](const char* name, uint32_t G, float ms) {
        printf("%-34s G=%-9u %8.3f ms  %8.2f Grows/s  %6.3f rows/clk/SM@1.9GHz\n", name, G, ms, n / ms * 1e-6, n / (ms * 1e-3) / sms / 1.9e9);
    };
    CK(cudaFuncSetAttribute(k_smem_atomic32<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_smem_atomic32<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_smem_atomic64, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_smem_private<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_smem_private<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (uint32_t G : {100u, 1000u, 10000u, 40000u}) {
        k_gen<<<sms * 8, 256>>>(k, n, G); CK(cudaDeviceSynchronize());
        report("read-only", G, timeit([&] { k_read<<<sms * 8, 256>>>(k, n, out); }));
        report("match_any", G, timeit([&] { k_match<<<sms * 8, 256>>>(k, n, out); }));
        report("smem atomicAdd u32 x1 (2 CTA/SM)", G, timeit([&] { k_smem_atomic32<1><<<sms * 2, 512, G * 4>>>(k, n, G, out); }));
        if (G * 12 <= 200 * 1024) report("smem atomicAdd u32 x3", G, timeit([&] { k_smem_atomic32<3><<<sms * (G * 12 > 100 * 1024 ? 1 : 2), 512, G * 12>>>(k, n, G, out); }));
        if (G * 8 <= 200 * 1024) report("smem atomicAdd u64 x1", G, timeit([&] { k_smem_atomic64<<<sms * (G * 8 > 100 * 1024 ? 1 : 2), 512, G * 8>>>(k, n, G, out); }));
        if (G * 256 * 4 <= 200 * 1024) report("smem lane-private RMW T=256", G, timeit([&] { k_smem_private<256><<<sms, 256, G * 256 * 4>>>(k, n, G, out); }));
        if (G * 128 * 4 <= 200 * 1024) report("smem lane-private RMW T=128", G, timeit([&] { k_smem_private<128><<<sms, 128, G * 128 * 4>>>(k, n, G, out); }));
    }
    for (uint32_t G : {100u, 10000u, 1000000u, 10000000u, 50000000u}) {
        k_gen<<<sms * 8, 256>>>(k, n, G); CK(cudaDeviceSynchronize());
        CK(cudaMemset(tab, 0, tab_bytes));
        report("global RED u64 x1 stride 4 (32B slot)", G, timeit([&] { k_global_red<1><<<sms * 16, 256>>>(k, n, tab, 4); }));
        report("global RED u64 x3 stride 4 (32B slot)", G, timeit([&] { k_global_red<3><<<sms * 16, 256>>>(k, n, tab, 4); }));
    }
    CK(cudaDeviceSynchronize());
    printf("done\n");
    return 0;
}
