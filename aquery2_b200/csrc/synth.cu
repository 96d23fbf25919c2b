// synth.cu -- counter-based synthetic columns, bit-identical to aquery2_b200/synth.py.
// Shapes: h2o G1 numeric-id table (reference benchmark/h2o/groupby_prep.sql:1) and the trade table
// of reference datagen.cpp:90-107.  Bench / test support, not part of the query path.
#include "common.cuh"

namespace aq {

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t uniform_k(uint64_t x, uint32_t k) { return ((x >> 32) * uint64_t(k)) >> 32; }

__global__ void __launch_bounds__(256)
k_synth(int kind, uint64_t key0, uint64_t key1, uint64_t key2, uint32_t k, uint32_t aux_k,
        uint64_t row_lo, uint32_t n, uint32_t* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint64_t row = row_lo + i;
        const uint64_t x = splitmix64(row + key0);
        uint32_t r;
        switch (kind) {
        case 0: r = 1u + (uint32_t)uniform_k(x, k); break;
        case 1: r = (uint32_t)uniform_k(x, k); break;
        case 2: r = __float_as_uint((float)((double)uniform_k(x, k) / 1e6)); break;
        case 3: r = (uint32_t)(row + 1); break;
        case 4: {   // trade price in [50, 500]: per-symbol base + small noise
            const uint64_t sym = uniform_k(splitmix64(row + key1), k);
            const uint64_t base = uniform_k(splitmix64(sym + key2), 441);
            r = 50u + (uint32_t)base + (uint32_t)uniform_k(x, 11);
            break; }
        case 5: r = aux_k + (uint32_t)uniform_k(x, k); break;     // aux_k + U[0,k)
        case 6: r = 1u + (uint32_t)((row * 40353607ull + 12345ull) % (uint64_t)k); break;   // a permutation of 1..k when gcd(7^9, k) = 1
        default: r = 0;
        }
        out[i] = r;
    }
}

} // namespace aq

extern "C" int aqgpu_synth_column(int kind, uint64_t key0, uint64_t key1, uint64_t key2, uint32_t k, uint32_t aux_k,
                                  uint64_t row_lo, uint32_t n, void* d_out) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    if (!aq::is_device_ptr(d_out)) return aq::fail("synth: output must be a device pointer");
    AQ_LAUNCH(aq::k_synth, aq::grid_for(n, 256, 4, 16), 256, 0, kind, key0, key1, key2, k, aux_k, row_lo, n,
              static_cast<uint32_t*>(d_out));
    return 0;
}
