// groupby_smem.cu -- shared-memory pre-aggregation in front of the global group table.
//
// For group-bys whose groups fit one SM's shared memory (h2o q1, q2, q4: 1e2..1e4 groups) every CTA
// keeps a private open-addressing (or directly indexed) table of 32-bit cells in shared memory and
// aggregates its rows there with native 32-bit shared atomics (measured on B200: they keep up with
// the HBM stream, profiles/microbench/mb_atomics.cu), then merges its <= cap groups into the global
// table of groupby.cu once, at the end.  Rows that find no room (table beyond its fill limit, or a key
// equal to the EMPTY marker) go straight to the global table, so the result never depends on the
// capacity guess.  Replaces the same reference loops as groupby.cu (server/hasher.h:176-198,
// server/table.h:184-189, server/aggregations.h scalar aggregates).
//
// Cells (SoA arrays of `cap` entries, so neighbouring lanes hit different banks):
//   key      u32 (exact 32-bit tag) | u64 (two 32-bit words / one int64 key) | none (direct index)
//   minrow   u32, first-appearance row of the group inside this CTA (atomicMin after a plain pre-check)
//   COUNT    u32
//   ADD_I64  u32 lo + u32 hi: lo += (u32)x with the carry (and x's high word) added to hi only when
//            non-zero, i.e. one shared atomic per row for small values
//   ADD_F64  u64, compare-and-swap loop (there is no native 64-bit shared add)
//   MIN/MAX  s32 native for int32 / float (order-preserving image), CAS loop on u64 otherwise
//   MAXROW   u32 atomicMax;  TOP2ROW u64 CAS loop
#include "groupby.cuh"
#include <cstring>

namespace aq {

enum SmemMode : int { SM_DIRECT = 0, SM_HASH32 = 1, SM_HASH64 = 2 };
enum CellKind : int { CK_COUNT = 0, CK_ADD_I64, CK_ADD_F64, CK_MIN32, CK_MAX32, CK_MIN64, CK_MAX64, CK_MAXROW, CK_TOP2 };

constexpr int SMEM_THREADS = 512;
constexpr uint32_t EMPTY32 = 0xffffffffu;
constexpr unsigned long long EMPTY64 = 0xffffffffffffffffULL;

struct SmemLayout {
    int mode;
    uint32_t cap, mask, limit;
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

// 64-bit exact key image of 4 consecutive rows: one int64 column (stride-2 words) or two 32-bit columns
__device__ __forceinline__ void key64_4(const KeySpec& ks, uint32_t row0, int cnt, unsigned long long k[4]) {
    if (ks.stride[0] == 2) {
        load4_b64(ks.base[0], row0, cnt, ks.vec_ok, k);
    } else {
        uint32_t a[4], b[4];
        load4_b32(ks.base[0], row0, cnt, ks.vec_ok, a);
        load4_b32(ks.base[1], row0, cnt, ks.vec_ok, b);
#pragma unroll
        for (int r = 0; r < 4; ++r) k[r] = (unsigned long long)a[r] | ((unsigned long long)b[r] << 32);
    }
}

template <int MODE>
__global__ void __launch_bounds__(SMEM_THREADS, 1)
k_smem_agg(const KeySpec ks, const uint32_t nrows, const uint32_t gmask, const uint32_t max_probe,
           const AccProg prog, const SmemLayout lay, uint32_t* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ uint32_t n_used;
    constexpr bool EXACT = MODE != SM_HASH64;
    const uint32_t cap = lay.cap;
    uint32_t* const minrow = reinterpret_cast<uint32_t*>(sm + lay.off_minrow);
    uint32_t* const key32 = reinterpret_cast<uint32_t*>(sm + lay.off_key);
    unsigned long long* const key64 = reinterpret_cast<unsigned long long*>(sm + lay.off_key);

    // ---- init: keys EMPTY, minrow = none, accumulators = identities
    for (uint32_t s = threadIdx.x; s < cap; s += SMEM_THREADS) {
        minrow[s] = EMPTY32;
        if (MODE == SM_HASH32) key32[s] = EMPTY32;
        if (MODE == SM_HASH64) key64[s] = EMPTY64;
        for (int a = 0; a < prog.n; ++a) {
            unsigned char* c0 = sm + lay.off_cell[a];
            switch (lay.kind[a]) {
            case CK_ADD_I64: reinterpret_cast<uint32_t*>(c0)[s] = 0u; reinterpret_cast<uint32_t*>(c0)[s + cap] = 0u; break;
            case CK_ADD_F64: case CK_TOP2: reinterpret_cast<unsigned long long*>(c0)[s] = 0ull; break;
            case CK_MIN32: reinterpret_cast<int*>(c0)[s] = 0x7fffffff; break;
            case CK_MAX32: reinterpret_cast<int*>(c0)[s] = (int)0x80000000u; break;
            case CK_MIN64: reinterpret_cast<unsigned long long*>(c0)[s] = 0x7fffffffffffffffULL; break;
            case CK_MAX64: reinterpret_cast<unsigned long long*>(c0)[s] = 0x8000000000000000ULL; break;
            default: reinterpret_cast<uint32_t*>(c0)[s] = 0u; break;     // COUNT, MAXROW
            }
        }
    }
    if (threadIdx.x == 0) n_used = 0;
    __syncthreads();

    const uint32_t nquad = (nrows + 3u) >> 2;
    const uint32_t stride = gridDim.x * SMEM_THREADS;
    for (uint32_t q = blockIdx.x * SMEM_THREADS + threadIdx.x; q < nquad; q += stride) {
        if (*reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT)) break;
        const uint32_t row0 = q << 2;
        const int cnt = min(4u, nrows - row0);
        // next iteration's columns -> L2 while this one is being aggregated
        if (q + stride < nquad) {
            const size_t nxt = size_t(q + stride) << 2;
            for (int w = 0; w < ks.nwords; ++w)
                asm volatile("prefetch.global.L2 [%0];" :: "l"(ks.base[w] + nxt * ks.stride[w]));
            for (int a = 0; a < prog.n; ++a) {
                const AccOp& o = prog.op[a];
                if (!acc_needs_value(o)) continue;
                const size_t esz = (o.vtype == AQ_INT32 || o.vtype == AQ_FLOAT) ? 4 : 8;
                asm volatile("prefetch.global.L2 [%0];" :: "l"(static_cast<const char*>(o.v) + nxt * esz));
                if (o.v2) asm volatile("prefetch.global.L2 [%0];" :: "l"(static_cast<const char*>(o.v2) + nxt * esz));
            }
        }
        uint32_t tag[4], hsh[4], sslot[4], gslot[4];
        unsigned long long k64[4];
        if (MODE == SM_HASH64) {
            key64_4(ks, row0, cnt, k64);
        } else if (!key_tags4(ks, row0, cnt, tag, hsh)) {
            atomicExch(flags + FLAG_ABORT, 3u);
            break;
        }
        bool bad = false;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            sslot[r] = NOSLOT; gslot[r] = NOSLOT;
            if (r >= cnt) continue;
            const uint32_t row = row0 + r;
            uint32_t s = NOSLOT;
            if (MODE == SM_DIRECT) {
                s = tag[r];                                     // < cap by construction (bounds checked above)
            } else if (MODE == SM_HASH32) {
                const uint32_t mine = tag[r];
                if (mine != EMPTY32) {
                    uint32_t p = hsh[r] & lay.mask;
                    for (int step = 0; step < 32; ++step) {
                        uint32_t cur = *reinterpret_cast<volatile uint32_t*>(key32 + p);
                        if (cur == EMPTY32) {
                            if (*reinterpret_cast<volatile uint32_t*>(&n_used) >= lay.limit) break;
                            cur = atomicCAS(key32 + p, EMPTY32, mine);
                            if (cur == EMPTY32) { atomicAdd(&n_used, 1u); s = p; break; }
                        }
                        if (cur == mine) { s = p; break; }
                        p = (p + 1u) & lay.mask;
                    }
                }
            } else {
                const unsigned long long mine = k64[r];
                if (mine != EMPTY64) {
                    uint32_t p = (uint32_t)mix64(mine) & lay.mask;
                    for (int step = 0; step < 32; ++step) {
                        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(key64 + p);
                        if (cur == EMPTY64) {
                            if (*reinterpret_cast<volatile uint32_t*>(&n_used) >= lay.limit) break;
                            cur = atomicCAS(key64 + p, EMPTY64, mine);
                            if (cur == EMPTY64) { atomicAdd(&n_used, 1u); s = p; break; }
                        }
                        if (cur == mine) { s = p; break; }
                        p = (p + 1u) & lay.mask;
                    }
                }
            }
            if (s != NOSLOT) {
                sslot[r] = s;
                if (row < *reinterpret_cast<volatile uint32_t*>(minrow + s)) atomicMin(minrow + s, row);
            } else {                                            // no room in shared memory: global table
                uint32_t t, h;
                if (MODE == SM_HASH64) key_tag1(ks, row, t, h); else { t = tag[r]; h = hsh[r]; }
                gslot[r] = probe_insert<EXACT>(ks, prog.tab, prog.sw, gmask, max_probe, h, t, row);
                bad |= gslot[r] == NOSLOT;
            }
        }
        if (bad) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); break; }
        for (int a = 0; a < prog.n; ++a) {
            const AccOp& o = prog.op[a];
            long long xi[4] = {0, 0, 0, 0}; double xd[4] = {0, 0, 0, 0};
            if (acc_needs_value(o)) acc_inputs(o, row0, cnt, xi, xd);
            unsigned char* c0 = sm + lay.off_cell[a];
            const int kind = lay.kind[a];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (r >= cnt) continue;
                if (sslot[r] != NOSLOT) smem_acc_row(o, kind, c0, cap, sslot[r], row0 + r, xi[r], xd[r]);
                else acc_global_row(o, slot_word(prog, gslot[r]) + 1 + a, row0 + r, xi[r], xd[r]);
            }
        }
    }
    __syncthreads();
    if (*reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT)) return;

    // ---- merge this CTA's groups into the global table
    for (uint32_t s = threadIdx.x; s < cap; s += SMEM_THREADS) {
        const uint32_t row = minrow[s];
        if (row == EMPTY32) continue;
        uint32_t t, h;
        if (MODE == SM_DIRECT) { t = s; h = mix32(t); }
        else if (MODE == SM_HASH32) { t = key32[s]; h = mix32(t); }
        else key_tag1(ks, row, t, h);
        const uint32_t g = probe_insert<EXACT>(ks, prog.tab, prog.sw, gmask, max_probe, h, t, row);
        if (g == NOSLOT) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); return; }
        for (int a = 0; a < prog.n; ++a)
            smem_merge(prog.op[a], lay.kind[a], sm + lay.off_cell[a], cap, s, slot_word(prog, g) + 1 + a);
    }
}

static uint32_t pow2_floor(uint64_t x) { uint32_t p = 1; while ((uint64_t(p) << 1) <= x) p <<= 1; return p; }

int launch_smem_probe(const KeySpec& ks, uint32_t nrows, uint32_t gcap, uint32_t max_probe,
                      const AccProg& prog, uint32_t groups_hint, bool* used) {
    *used = false;
    Context& c = ctx();
    SmemLayout lay;
    memset(&lay, 0, sizeof(lay));
    // ---- mode
    const bool direct_ok = ks.exact == 2 && ks.pack_bits <= 16;
    if (ks.exact == 1 || ks.exact == 2) lay.mode = SM_HASH32;
    else if (ks.nwords == 2 && (ks.stride[0] == 1) == (ks.stride[1] == 1)) lay.mode = SM_HASH64;
    else return 0;
    // ---- bytes per slot
    uint32_t cell_bytes[MAX_ACC];
    uint32_t bps = 4;                                             // minrow
    for (int a = 0; a < prog.n; ++a) {
        const AccOp& o = prog.op[a];
        const bool v32 = o.vtype == AQ_INT32 || o.vtype == AQ_FLOAT;
        switch (o.kind) {
        case ACC_COUNT: lay.kind[a] = CK_COUNT; cell_bytes[a] = 4; break;
        case ACC_ADD_I64: lay.kind[a] = CK_ADD_I64; cell_bytes[a] = 8; break;
        case ACC_ADD_F64: lay.kind[a] = CK_ADD_F64; cell_bytes[a] = 8; break;
        case ACC_MIN_I64: lay.kind[a] = v32 ? CK_MIN32 : CK_MIN64; cell_bytes[a] = v32 ? 4 : 8; break;
        case ACC_MAX_I64: lay.kind[a] = v32 ? CK_MAX32 : CK_MAX64; cell_bytes[a] = v32 ? 4 : 8; break;
        case ACC_MAXROW: lay.kind[a] = CK_MAXROW; cell_bytes[a] = 4; break;
        case ACC_TOP2ROW: lay.kind[a] = CK_TOP2; cell_bytes[a] = 8; break;
        default: return 0;
        }
        bps += cell_bytes[a];
    }
    const size_t budget = (c.smem_optin > 2048 ? c.smem_optin : 48 * 1024) - 1024;
    // ---- capacity
    if (direct_ok && (size_t(bps) << ks.pack_bits) <= budget) {
        lay.mode = SM_DIRECT;
        lay.cap = 1u << ks.pack_bits;
        if (lay.cap < 32) lay.cap = 32;
        lay.limit = lay.cap;
    } else {
        bps += lay.mode == SM_HASH64 ? 8 : 4;
        const uint32_t cap_max = pow2_floor(budget / bps);
        if (cap_max < 4096) return 0;
        uint32_t cap = 4096;
        if (groups_hint) {
            while (cap < cap_max && (uint64_t(cap) * 3) / 4 < groups_hint) cap <<= 1;
            if ((uint64_t(cap) * 3) / 4 < groups_hint) return 0;      // more groups than one SM can hold
        } else {
            cap = cap_max;
        }
        lay.cap = cap;
        lay.limit = cap / 4 * 3;
    }
    lay.mask = lay.cap - 1;
    // ---- offsets: 64-bit arrays first
    uint32_t off = 0;
    if (lay.mode == SM_HASH64) { lay.off_key = off; off += 8 * lay.cap; }
    for (int a = 0; a < prog.n; ++a)
        if (cell_bytes[a] == 8 && lay.kind[a] != CK_ADD_I64) { lay.off_cell[a] = off; off += 8 * lay.cap; }
    for (int a = 0; a < prog.n; ++a)
        if (lay.kind[a] == CK_ADD_I64) { lay.off_cell[a] = off; off += 8 * lay.cap; }
    if (lay.mode == SM_HASH32) { lay.off_key = off; off += 4 * lay.cap; }
    lay.off_minrow = off; off += 4 * lay.cap;
    for (int a = 0; a < prog.n; ++a)
        if (cell_bytes[a] == 4) { lay.off_cell[a] = off; off += 4 * lay.cap; }
    lay.total_bytes = off;
    if (off > budget) return 0;

    const uint32_t nquad = (nrows + 3) / 4;
    int grid = (int)((nquad + SMEM_THREADS - 1) / SMEM_THREADS);
    if (grid > c.sm_count) grid = c.sm_count;
    static bool attr_done = false;
    if (!attr_done) {
        AQ_CUDA(cudaFuncSetAttribute(k_smem_agg<SM_DIRECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        AQ_CUDA(cudaFuncSetAttribute(k_smem_agg<SM_HASH32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        AQ_CUDA(cudaFuncSetAttribute(k_smem_agg<SM_HASH64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        attr_done = true;
    }
    switch (lay.mode) {
    case SM_DIRECT: AQ_LAUNCH(k_smem_agg<SM_DIRECT>, grid, SMEM_THREADS, lay.total_bytes, ks, nrows, gcap - 1, max_probe, prog, lay, c.d_flags); break;
    case SM_HASH32: AQ_LAUNCH(k_smem_agg<SM_HASH32>, grid, SMEM_THREADS, lay.total_bytes, ks, nrows, gcap - 1, max_probe, prog, lay, c.d_flags); break;
    default: AQ_LAUNCH(k_smem_agg<SM_HASH64>, grid, SMEM_THREADS, lay.total_bytes, ks, nrows, gcap - 1, max_probe, prog, lay, c.d_flags); break;
    }
    *used = true;
    return 0;
}

} // namespace aq
