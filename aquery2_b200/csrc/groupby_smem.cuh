// groupby_smem.cuh -- shared-memory table cells used by groupby_smem.cu (per-CTA pre-aggregation) and
// groupby_part.cu (per-partition aggregation): cell kinds, layout, per-row update and merge.
#pragma once
#include "groupby.cuh"

namespace aq {

enum SmemMode : int { SM_DIRECT = 0, SM_HASH32 = 1, SM_HASH64 = 2 };
enum CellKind : int { CK_COUNT = 0, CK_ADD_I64, CK_ADD_F64, CK_MIN32, CK_MAX32, CK_MIN64, CK_MAX64, CK_MAXROW, CK_TOP2 };

constexpr int SMEM_THREADS = 512;
constexpr uint32_t EMPTY32 = 0xffffffffu;
constexpr unsigned long long EMPTY64 = 0xffffffffffffffffULL;

struct SmemLayout {
    int mode;
    uint32_t cap, mask, limit;     // per replica
    uint32_t reps;                 // power of two: warp w works on replica w & (reps - 1)
    uint32_t off_key, off_minrow;
    uint32_t off_cell[MAX_ACC];
    int kind[MAX_ACC];
    uint32_t total_bytes;
};

__device__ __forceinline__ int f32_ordered(float f) {
    const int b = __float_as_int(f);
    return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ordered_f32(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

__device__ __forceinline__ void smem_acc_row(const AccOp& o, int kind, unsigned char* cell0, uint32_t cap, uint32_t s,
                                             uint32_t row, long long xi, double xd) {
    const bool isint = (o.vtype == AQ_INT32 || o.vtype == AQ_INT64);
    switch (kind) {
    case CK_COUNT: atomicAdd(reinterpret_cast<uint32_t*>(cell0) + s, 1u); break;
    case CK_ADD_I64: {
        uint32_t* lo = reinterpret_cast<uint32_t*>(cell0) + s;
        const uint32_t xl = (uint32_t)xi;
        const uint32_t old = atomicAdd(lo, xl);
        const uint32_t d = (uint32_t)((unsigned long long)xi >> 32) + (((uint32_t)(old + xl) < old) ? 1u : 0u);
        if (d) atomicAdd(lo + cap, d);
        break; }
    case CK_ADD_F64: {
        unsigned long long* p = reinterpret_cast<unsigned long long*>(cell0) + s;
        const double x = isint ? (double)xi : xd;
        unsigned long long old = *reinterpret_cast<volatile unsigned long long*>(p), assumed;
        do {
            assumed = old;
            old = atomicCAS(p, assumed, (unsigned long long)__double_as_longlong(__longlong_as_double((long long)assumed) + x));
        } while (old != assumed);
        break; }
    case CK_MIN32: case CK_MAX32: {
        int* p = reinterpret_cast<int*>(cell0) + s;
        const int img = o.vtype == AQ_INT32 ? (int)xi : f32_ordered((float)xd);
        const int cur = *reinterpret_cast<volatile int*>(p);
        if (kind == CK_MIN32) { if (img < cur) atomicMin(p, img); }
        else { if (img > cur) atomicMax(p, img); }
        break; }
    case CK_MIN64: case CK_MAX64: {
        unsigned long long* p = reinterpret_cast<unsigned long long*>(cell0) + s;
        const long long img = isint ? xi : f64_to_ordered(xd);
        unsigned long long old = *reinterpret_cast<volatile unsigned long long*>(p);
        while (kind == CK_MIN64 ? img < (long long)old : img > (long long)old) {
            const unsigned long long seen = atomicCAS(p, old, (unsigned long long)img);
            if (seen == old) break;
            old = seen;
        }
        break; }
    case CK_MAXROW: {
        uint32_t* p = reinterpret_cast<uint32_t*>(cell0) + s;
        if (row > *reinterpret_cast<volatile uint32_t*>(p) || row == 0u) atomicMax(p, row);
        break; }
    case CK_TOP2: top2_update(reinterpret_cast<unsigned long long*>(cell0) + s, top2_key(o.xform, row)); break;
    default: break;
    }
}

// merge one shared-memory cell into cell c of the global slot
__device__ __forceinline__ void smem_merge(const AccOp& o, int kind, const unsigned char* cell0, uint32_t cap, uint32_t s,
                                           unsigned long long* c) {
    switch (kind) {
    case CK_COUNT: atomicAdd(c, (unsigned long long)(reinterpret_cast<const uint32_t*>(cell0)[s])); break;
    case CK_ADD_I64: {
        const uint32_t* lo = reinterpret_cast<const uint32_t*>(cell0) + s;
        atomicAdd(c, ((unsigned long long)lo[cap] << 32) | lo[0]);
        break; }
    case CK_ADD_F64: atomicAdd(reinterpret_cast<double*>(c), __longlong_as_double((long long)reinterpret_cast<const unsigned long long*>(cell0)[s])); break;
    case CK_MIN32: case CK_MAX32: {
        const int k = reinterpret_cast<const int*>(cell0)[s];
        const long long img = o.vtype == AQ_INT32 ? (long long)k : f64_to_ordered((double)ordered_f32(k));
        if (kind == CK_MIN32) atomicMin(reinterpret_cast<long long*>(c), img);
        else atomicMax(reinterpret_cast<long long*>(c), img);
        break; }
    case CK_MIN64: atomicMin(reinterpret_cast<long long*>(c), (long long)reinterpret_cast<const unsigned long long*>(cell0)[s]); break;
    case CK_MAX64: atomicMax(reinterpret_cast<long long*>(c), (long long)reinterpret_cast<const unsigned long long*>(cell0)[s]); break;
    case CK_MAXROW: atomicMax(c, (unsigned long long)reinterpret_cast<const uint32_t*>(cell0)[s]); break;
    case CK_TOP2: {
        const unsigned long long v = reinterpret_cast<const unsigned long long*>(cell0)[s];
        if (v >> 32) top2_update(c, v >> 32);
        if (v & 0xffffffffull) top2_update(c, v & 0xffffffffull);
        break; }
    default: break;
    }
}

__device__ __forceinline__ uint4 pick_col(const uint4 (&a)[MAX_VCOL], int ci) {
    uint4 r = a[0];
    if (ci == 1) r = a[1];
    if (ci == 2) r = a[2];
    if (ci == 3) r = a[3];
    return r;
}
__device__ __forceinline__ uint32_t lane4(const uint4& v, int r) { return r == 0 ? v.x : r == 1 ? v.y : r == 2 ? v.z : v.w; }

__device__ __forceinline__ uint4 load_quad(const void* base, uint32_t row0, int cnt, bool vec) {
    uint32_t t[4];
    load4_b32(base, row0, cnt, vec, t);
    return make_uint4(t[0], t[1], t[2], t[3]);
}

} // namespace aq
