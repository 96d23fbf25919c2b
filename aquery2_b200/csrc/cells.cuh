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
//   FX      float32 values summed EXACTLY in fixed point.  The global record holds a 128-bit sum in units of
//           2^(B-150) (B chosen from the sampled maximum exponent so that a value with that exponent has its
//           mantissa LSB at bit FX_S_TOP).  The shared cell is 2 words: a 64-bit sum in units 2^Q times
//           coarser (Q = FX_S_TOP - R), which holds every value whose mantissa LSB lies within R bits below
//           the sampled maximum -- one returning add for the low word (carry) and one plain add for the high
//           word, no carry chain.  R = 39 - log2(rows one CTA can see), so the 64-bit cell cannot overflow.
//           Values outside that window (tiny ones with low mantissa bits set, ones above the sampled maximum)
//           are added straight to the 128-bit global sum; a value that has no exact image at all (below the
//           unit with non-zero low bits, above bit 94, inf / nan) raises FLAG_ABORT = 4 and the query re-runs
//           on the generic path.  The sum is rounded ONCE, at the end (the reference rounds after every
//           addition, aggregations.h:19-27 in double: both are within 1e-12 of each other, the exact one is
//           the more accurate).
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
// m * 2^s (sign neg) added to the 128-bit global sum; s in [0, FX_S_MAX]
__device__ __forceinline__ void fx_global_add(unsigned long long* glo, unsigned long long* ghi, uint32_t m, int s, bool neg) {
    unsigned long long lo, hi;
    if (s < 64) { lo = (unsigned long long)m << s; hi = s ? ((unsigned long long)m >> (64 - s)) : 0ull; }
    else { lo = 0ull; hi = (unsigned long long)m << (s - 64); }
    if (neg) { lo = 0ull - lo; hi = ~hi + (lo == 0ull ? 1ull : 0ull); }
    if (lo) { const unsigned long long old = atomicAdd(glo, lo); if (old + lo < old) hi += 1ull; }
    if (hi) atomicAdd(ghi, hi);
}
// false: the value has no exact image (the caller raises the abort flag).  Q, R: see the header.
__device__ __forceinline__ bool fx_add(uint32_t* lo, uint32_t* hi, unsigned long long* glo, unsigned long long* ghi,
                                       float x, int B, int Q, int R) {
    // fast path: x scaled to the cell's unit 2^(B+Q-150) is an integer of at most 24 + R bits.  Power-of-two scaling
    // in double is exact, and so is the round trip through int64 exactly when the scaled value is an integer (this
    // also admits values below the window whose low mantissa bits are zero); inf / nan fail the comparison.
    {
        const double sc = __longlong_as_double((long long)(1023 + 150 - B - Q) << 52);
        const double lim = __longlong_as_double((long long)(1023 + 24 + R) << 52);
        const double xd = (double)x * sc;
        const long long X = __double2ll_rn(xd);
        if ((double)X == xd && fabs(xd) <= lim) { isum2_add(lo, hi, X); return true; }
    }
    const uint32_t bits = __float_as_uint(x);
    uint32_t m = bits & 0x7fffffu;
    int e = (int)((bits >> 23) & 0xffu);
    if (e == 0) { if (m == 0) return true; e = 1; }
    else { if (e == 255) return false; m |= 0x800000u; }
    const int t = e - B - Q;
    if ((unsigned)t <= (unsigned)R) {                         // the common case: inside the cell's window
        long long X = (long long)((unsigned long long)m << t);
        if (bits >> 31) X = -X;
        isum2_add(lo, hi, X);
        return true;
    }
    int s = e - B;
    if (s < 0) {
        const int r = -s;
        if (r >= 24 || (m & ((1u << r) - 1u)) != 0u) return false;
        m >>= r; s = 0;
    }
    if (s > FX_S_MAX) return false;
    fx_global_add(glo, ghi, m, s, (bits >> 31) != 0u);
    return true;
}
// a CTA's shared 64-bit cell (units 2^Q) -> the global 128-bit sum
__device__ __forceinline__ void fx_merge(uint32_t lo, uint32_t hi, int Q, unsigned long long* glo, unsigned long long* ghi) {
    const long long v = (long long)((unsigned long long)lo | ((unsigned long long)hi << 32));
    if (v == 0) return;
    const unsigned long long l = (unsigned long long)v << Q;
    unsigned long long h = (unsigned long long)(v >> (64 - Q));          // arithmetic shift: sign-extends; Q in [1, 63]
    if (l) { const unsigned long long old = atomicAdd(glo, l); if (old + l < old) h += 1ull; }
    if (h) atomicAdd(ghi, h);
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
