// cells.cuh -- compact accumulator cells for shared-memory group tables (lc.cu, groupby_p2.cu).
//
// Shared memory on B200 has native 32-bit atomics only (64-bit adds are CAS loops: 195-380 Grows/s against
// 650-1500 for u32, profiles/r01/microbench.txt), and the tables that decide whether a query runs in ONE
// streaming pass are sized by bytes per group.  So every accumulator is made of 32-bit words:
//
//   CNT     1 word.
//   ISUM1   1 word (biased by 2^31) + a global spill word counted in units of 2^32: the sum of a group inside
//           one CTA is word - 2^31 + 2^32 * spill.  A row adds the low 32 bits of |x| to (or subtracts them from)
//           the word; the carry / borrow and the high bits of |x| go to the spill (an L2 atomic that small
//           values almost never execute).  Exact for any int64 x.
//   ISUM2   2 words (lo, hi): for columns whose sampled magnitudes make the spill frequent.
//   FX      float32 values summed EXACTLY in 128-bit fixed point: unit 2^(B-150), word 0 (bits 0..31) in a
//           global spill, words 1..3 in shared memory (word 3 biased).  The 24-bit mantissa of a value lands
//           in at most two adjacent words; carries ripple upwards through the returned old values.  B is
//           chosen from the sampled maximum exponent so that typical mantissas sit in word 2; a value that
//           cannot be represented (below the unit with non-zero low bits, above bit 94, inf / nan) raises
//           FLAG_ABORT = 4 and the query re-runs on the generic path.  The sum is rounded ONCE, at the end
//           (the reference rounds after every addition, aggregations.h:19-27 in double: both are within
//           1e-12 of each other, the exact one is the more accurate).
//   MIN32 / MAX32   order-preserving 32-bit image, atomicMin / atomicMax after a plain pre-check.
#pragma once
#include "groupby.cuh"

namespace aq {

constexpr uint32_t CELL_BIAS = 0x80000000u;
constexpr int FX_S_TOP = 66;        // bit position of the mantissa LSB of a value with the sampled max exponent
constexpr int FX_S_MAX = 70;        // mantissa must end below bit 94: 2^32 rows of such values still fit 127 bits

#ifdef __CUDACC__
// ---- ISUM1 ---------------------------------------------------------------------------------------
__device__ __forceinline__ void isum1_add(uint32_t* cell, unsigned long long* spill, long long x) {
    if (x >= 0) {
        const uint32_t al = (uint32_t)x;
        const unsigned long long ah = (unsigned long long)x >> 32;
        const uint32_t old = atomicAdd(cell, al);
        const unsigned long long d = ah + (((uint32_t)(old + al) < old) ? 1ull : 0ull);
        if (d) atomicAdd(spill, d);
    } else {
        const unsigned long long m = 0ull - (unsigned long long)x;
        const uint32_t al = (uint32_t)m;
        const unsigned long long ah = m >> 32;
        const uint32_t old = atomicSub(cell, al);
        const unsigned long long d = ah + ((old < al) ? 1ull : 0ull);
        if (d) atomicAdd(spill, 0ull - d);
    }
}
// ---- ISUM2 ---------------------------------------------------------------------------------------
__device__ __forceinline__ void isum2_add(uint32_t* lo, uint32_t* hi, long long x) {
    const uint32_t xl = (uint32_t)x;
    const uint32_t old = atomicAdd(lo, xl);
    const uint32_t d = (uint32_t)((unsigned long long)x >> 32) + (((uint32_t)(old + xl) < old) ? 1u : 0u);
    if (d) atomicAdd(hi, d);
}
// ---- FX ------------------------------------------------------------------------------------------
// false: the value has no exact image (the caller raises the abort flag)
__device__ __forceinline__ bool fx_add(uint32_t* w1, uint32_t* w2, uint32_t* w3, unsigned long long* glo,
                                       unsigned long long* ghi, float x, int B) {
    const uint32_t bits = __float_as_uint(x);
    uint32_t m = bits & 0x7fffffu;
    int e = (int)((bits >> 23) & 0xffu);
    if (e == 0) { if (m == 0) return true; e = 1; }
    else { if (e == 255) return false; m |= 0x800000u; }
    int s = e - B;
    if (s < 0) {
        const int r = -s;
        if (r >= 24 || (m & ((1u << r) - 1u)) != 0u) return false;
        m >>= r; s = 0;
    }
    if (s > FX_S_MAX) return false;
    const int q = s >> 5;
    const unsigned long long lo64 = (unsigned long long)m << (s & 31);
    const uint32_t pa = (uint32_t)lo64, pb = (uint32_t)(lo64 >> 32);
    uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    if (q == 0) { a0 = pa; a1 = pb; } else if (q == 1) { a1 = pa; a2 = pb; } else { a2 = pa; a3 = pb; }
    if (!(bits >> 31)) {
        if (a0) { const unsigned long long old = atomicAdd(glo, (unsigned long long)a0); if (old + a0 < old) atomicAdd(ghi, 1ull); }
        uint32_t c = 0;
        if (a1) { const uint32_t old = atomicAdd(w1, a1); c = ((uint32_t)(old + a1) < old) ? 1u : 0u; }
        a2 += c; c = 0;
        if (a2) { const uint32_t old = atomicAdd(w2, a2); c = ((uint32_t)(old + a2) < old) ? 1u : 0u; }
        a3 += c;
        if (a3) atomicAdd(w3, a3);
    } else {
        if (a0) { const unsigned long long old = atomicAdd(glo, 0ull - (unsigned long long)a0); if (old < a0) atomicAdd(ghi, ~0ull); }
        uint32_t b = 0;
        if (a1) { const uint32_t old = atomicSub(w1, a1); b = (old < a1) ? 1u : 0u; }
        a2 += b; b = 0;
        if (a2) { const uint32_t old = atomicSub(w2, a2); b = (old < a2) ? 1u : 0u; }
        a3 += b;
        if (a3) atomicSub(w3, a3);
    }
    return true;
}
// a CTA's shared words of one FX cell -> the global 128-bit sum (glo, ghi)
__device__ __forceinline__ void fx_merge(uint32_t w1, uint32_t w2, uint32_t w3, unsigned long long* glo, unsigned long long* ghi) {
    if (w1 == 0u && w2 == 0u && w3 == CELL_BIAS) return;
    const unsigned long long lo = (unsigned long long)w1 << 32;
    unsigned long long hi = (unsigned long long)w2 | ((unsigned long long)(uint32_t)(w3 - CELL_BIAS) << 32);
    if (lo) { const unsigned long long old = atomicAdd(glo, lo); if (old + lo < old) hi += 1ull; }
    if (hi) atomicAdd(ghi, hi);
}
// the global 128-bit fixed-point sum -> double (one rounding)
__device__ __forceinline__ double fx_to_double(unsigned long long glo, unsigned long long ghi, int B) {
    const __int128 v = (__int128)(((unsigned __int128)ghi << 64) | (unsigned __int128)glo);
    return scalbn((double)v, B - 150);
}
// B for a column whose largest sampled biased exponent is emax (1..254; 0 = no non-zero sample)
__host__ __device__ __forceinline__ int fx_base(int emax) { return (emax > 0 ? emax : 127) - FX_S_TOP; }

// ---- MIN32 / MAX32 --------------------------------------------------------------------------------
__device__ __forceinline__ int ordered32(uint32_t raw, int vtype) {
    const int b = (int)raw;
    return vtype == AQ_INT32 ? b : (b ^ ((b >> 31) & 0x7fffffff));
}
__device__ __forceinline__ void min32_upd(int* p, int img) { if (img < *reinterpret_cast<volatile int*>(p)) atomicMin(p, img); }
__device__ __forceinline__ void max32_upd(int* p, int img) { if (img > *reinterpret_cast<volatile int*>(p)) atomicMax(p, img); }
// 64-bit image the global tables / emit_one use for MIN / MAX cells (ACC_MIN_I64 / ACC_MAX_I64)
__device__ __forceinline__ long long image64_of32(int img, int vtype) {
    if (vtype == AQ_INT32) return (long long)img;
    const int b = img ^ ((img >> 31) & 0x7fffffff);
    return f64_to_ordered((double)__int_as_float(b));
}
#endif

} // namespace aq
