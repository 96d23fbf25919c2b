// lc.cu -- low-cardinality group-by: ONE streaming pass with a directly indexed shared-memory table.
//
// Replaces, for key tuples whose value ranges multiply to at most LC_MAX_G slots (h2o q1, q2, q4, q6, q9),
// the reference's per-row hashtable_push + ht_postproc + gather + aggregate loops (server/hasher.h:176-198,
// server/table.h:184-189, server/aggregations.h:10-86, 332-415) with
//
//   k_lc_agg    every needed column is read exactly once through a bulk-copy ring (stream.cuh: cp.async.bulk
//               + mbarrier, one producer lane, STAGES tiles in flight per CTA); the consumer warps index the
//               table with the mixed-radix image of the key tuple (the reference's PerfectHashTable idea,
//               server/hasher.h:226-243, without rounding the ranges up to powers of two) and update compact
//               32-bit cells (cells.cuh) with native shared atomics.  Tables too fat for one SM are split BY
//               ACCUMULATOR over a thread-block cluster of 2 or 4 CTAs: the CTAs of a cluster read the same
//               tiles and each keeps its share of the accumulators, so no atomic ever leaves the SM.
//   k_lc_count  non-empty slots -> compact list, group count.
//   k_lc_first  first-appearance row of every group: rows are scanned from row 0 in ticket order until all
//               groups have been seen -- a few hundred thousand rows for 1e4 groups, not 1e9.
//   k_lc_rank   rank of each group's first row among all first rows = the insertion index m_values has in the
//               reference (server/unordered_dense.h:1142-1146).
//   k_lc_emit   keys decoded from the slot index, aggregates finalised (emit_one), written at the rank.
//
// Everything speculated from the sample (key ranges, fixed-point exponent base) is checked per row; a
// violation raises FLAG_ABORT and the caller re-runs the query on the generic path.
#include "lc.cuh"
#include "stream.cuh"
#include "cells.cuh"
#include "emit.cuh"
#include "comm.cuh"
#include <algorithm>
#include <cstring>

namespace aq {

constexpr int LC_RING_OFF = 1024;       // barriers + accumulator descriptors live in the first KB of dynamic shared memory
constexpr int LC_DESC_OFF = 256;

// One accumulator of THIS CTA, as the consumer loop wants it (shared memory, read with uniform loads).
struct LcAccS {
    int kind, xform, vtype, fxB;
    uint32_t tab;           // byte offset (from the start of dynamic shared memory) of word array 0
    uint32_t c0, c1;        // byte offset of the column of v / v2 inside a ring stage
    uint32_t gword;
};
static_assert(sizeof(LcAccS) * MAX_ACC + LC_DESC_OFF <= LC_RING_OFF, "descriptor area");

// Per-thread state that outlives a tile: what the speculation checks have seen.
struct LcChecks {
    uint32_t kd[LC_MAX_KEYS];       // largest key - klo per key column (> krange: key outside the sampled bounds)
    uint32_t vabs32;                // largest |x| added to an ISUM1 cell (32-bit values / 64-bit products)
    unsigned long long vabs;
    bool fx_bad;
};

// RPT consecutive rows of a column tile
template <int RPT>
__device__ __forceinline__ void lc_ldrows(const unsigned char* p, uint32_t (&v)[RPT]) {
    if (RPT == 4) { const uint4 q = lds128(p); v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w; }
    else { const uint2 q = *reinterpret_cast<const uint2*>(p); v[0] = q.x; v[1] = q.y; }
}

// RPT rows of one accumulator.  s[j]: slot of row j; u / u2: the rows' raw 32-bit values.
template <int RPT>
__device__ __forceinline__ void lc_accn(const LcParams& P, const LcAccS& A, unsigned char* sm, const uint32_t (&s)[RPT], const uint32_t (&u)[RPT],
                                        const uint32_t (&u2)[RPT], LcChecks& ck) {
    const uint32_t G = P.G;
    switch (A.kind) {
    case LCK_SEEN: {
        unsigned char* b = sm + A.tab;
#pragma unroll
        for (int j = 0; j < RPT; ++j) b[s[j]] = 1;
        break; }
    case LCK_CNT: {
        uint32_t* c = reinterpret_cast<uint32_t*>(sm + A.tab);
#pragma unroll
        for (int j = 0; j < RPT; ++j) atomicAdd(c + s[j], 1u);
        break; }
    case LCK_ISUM1: {
        uint32_t* c = reinterpret_cast<uint32_t*>(sm + A.tab);
        if (A.xform == X_IDENT) {
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                const int x = (int)u[j];
                const uint32_t ax = (uint32_t)(x < 0 ? -x : x);
                ck.vabs32 = max(ck.vabs32, ax);
                atomicAdd(c + s[j], (uint32_t)x);
            }
        } else {
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                const long long x = (long long)(int)u[j] * (long long)(int)u2[j];
                const unsigned long long ax = (unsigned long long)(x < 0 ? -x : x);
                if (ax > ck.vabs) ck.vabs = ax;
                atomicAdd(c + s[j], (uint32_t)x);
            }
        }
        break; }
    case LCK_ISUM2: {
        uint32_t* lo = reinterpret_cast<uint32_t*>(sm + A.tab);
        uint32_t* hi = lo + G;
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
            const long long x = A.xform == X_IDENT ? (long long)(int)u[j] : (long long)(int)u[j] * (long long)(int)u2[j];
            isum2_add(lo + s[j], hi + s[j], x);
        }
        break; }
    case LCK_FX: {
        uint32_t* lo = reinterpret_cast<uint32_t*>(sm + A.tab);
        uint32_t* hi = lo + G;
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
            const float f = __uint_as_float(u[j]);
            const float x = A.xform == X_IDENT ? f : __fmul_rn(f, __uint_as_float(u2[j]));   // product in T (aggregations.h:339-345)
            unsigned long long* g = P.gtab + size_t(s[j]) * P.gw + A.gword;
            if (!fx_add(lo + s[j], hi + s[j], g, g + 1, x, A.fxB, P.fxQ, P.fxR)) ck.fx_bad = true;
        }
        break; }
    case LCK_MIN32: {
        int* c = reinterpret_cast<int*>(sm + A.tab);
#pragma unroll
        for (int j = 0; j < RPT; ++j) min32_upd(c + s[j], ordered32(u[j], A.vtype));
        break; }
    case LCK_MAX32: {
        int* c = reinterpret_cast<int*>(sm + A.tab);
#pragma unroll
        for (int j = 0; j < RPT; ++j) max32_upd(c + s[j], ordered32(u[j], A.vtype));
        break; }
    default: break;
    }
}

// One row (the ragged tail of the table only): values straight from global memory.
__device__ __forceinline__ void lc_acc1(const LcParams& P, const LcAccS& A, unsigned char* sm, uint32_t slot, uint32_t row, LcChecks& ck) {
    const uint32_t tile_bytes = P.tile_rows * 4u;
    uint32_t u = 0, u2 = 0;
    if (A.kind != LCK_CNT && A.kind != LCK_SEEN) {
        u = __ldg(P.cols[A.c0 / tile_bytes] + row);
        u2 = A.xform == X_PRODUCT ? __ldg(P.cols[A.c1 / tile_bytes] + row) : u;
    }
    const uint32_t G = P.G;
    switch (A.kind) {
    case LCK_SEEN: (sm + A.tab)[slot] = 1; break;
    case LCK_CNT: atomicAdd(reinterpret_cast<uint32_t*>(sm + A.tab) + slot, 1u); break;
    case LCK_ISUM1: {
        const long long x = A.xform == X_IDENT ? (long long)(int)u : (long long)(int)u * (long long)(int)u2;
        const unsigned long long ax = (unsigned long long)(x < 0 ? -x : x);
        if (ax > ck.vabs) ck.vabs = ax;
        atomicAdd(reinterpret_cast<uint32_t*>(sm + A.tab) + slot, (uint32_t)x);
        break; }
    case LCK_ISUM2: {
        const long long x = A.xform == X_IDENT ? (long long)(int)u : (long long)(int)u * (long long)(int)u2;
        uint32_t* lo = reinterpret_cast<uint32_t*>(sm + A.tab);
        isum2_add(lo + slot, lo + G + slot, x);
        break; }
    case LCK_FX: {
        const float f = __uint_as_float(u);
        const float x = A.xform == X_IDENT ? f : __fmul_rn(f, __uint_as_float(u2));
        uint32_t* lo = reinterpret_cast<uint32_t*>(sm + A.tab);
        unsigned long long* g = P.gtab + size_t(slot) * P.gw + A.gword;
        if (!fx_add(lo + slot, lo + G + slot, g, g + 1, x, A.fxB, P.fxQ, P.fxR)) ck.fx_bad = true;
        break; }
    case LCK_MIN32: min32_upd(reinterpret_cast<int*>(sm + A.tab) + slot, ordered32(u, A.vtype)); break;
    case LCK_MAX32: max32_upd(reinterpret_cast<int*>(sm + A.tab) + slot, ordered32(u, A.vtype)); break;
    default: break;
    }
}

// NK: number of key columns (compile time: the key loop is straight-line code); CL: CTAs per cluster.
template <int NK, int CL, int RPT>
__global__ void __launch_bounds__(1024, 1)
k_lc_agg(const LcParams P) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t* const full = reinterpret_cast<uint64_t*>(sm);
    uint64_t* const empty = full + LC_MAX_STAGES;
    LcAccS* const sacc = reinterpret_cast<LcAccS*>(sm + LC_DESC_OFF);
    __shared__ int s_nown;
    const uint32_t rank = CL > 1 ? cluster_ctarank() : 0u;
    const uint32_t cid = blockIdx.x / CL, ncl = gridDim.x / CL;
    const int ncols = P.ncols;
    const uint32_t tile_bytes = P.tile_rows * 4u;
    const uint32_t stage_bytes = tile_bytes * (uint32_t)ncols;
    unsigned char* const ring = sm + LC_RING_OFF;
    const uint32_t table_off = LC_RING_OFF + (uint32_t)P.stages * stage_bytes;
    const uint32_t G = P.G;
    const int ncw = (int)(blockDim.x >> 5) - 1;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t nfull = P.nrows / P.tile_rows;               // tiles the ring carries; the ragged rest is read directly

    if (threadIdx.x == 0) {
        for (int s = 0; s < P.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, (uint32_t)(ncw * CL)); }
        mbar_fence_init();
        int n = 0;
        for (int a = 0; a < P.nacc; ++a) {
            const LcAcc& A = P.acc[a];
            if ((uint32_t)A.owner != rank) continue;
            LcAccS d;
            d.kind = A.kind; d.xform = A.xform; d.vtype = A.vtype; d.fxB = A.fxB;
            d.tab = table_off + A.off; d.c0 = (uint32_t)A.c0 * tile_bytes; d.c1 = (uint32_t)A.c1 * tile_bytes; d.gword = A.gword;
            sacc[n++] = d;
        }
        s_nown = n;
    }
    __syncthreads();
    const int nown = s_nown;
    for (int a = 0; a < nown; ++a) {
        const LcAccS A = sacc[a];
        uint32_t* w = reinterpret_cast<uint32_t*>(sm + A.tab);
        if (A.kind == LCK_SEEN) { for (uint32_t g = threadIdx.x; g < (G + 3u) / 4u; g += blockDim.x) w[g] = 0u; continue; }
        for (uint32_t g = threadIdx.x; g < G; g += blockDim.x) {
            switch (A.kind) {
            case LCK_ISUM2: case LCK_FX: w[g] = 0u; w[G + g] = 0u; break;
            case LCK_MIN32: w[g] = 0x7fffffffu; break;
            case LCK_MAX32: w[g] = 0x80000000u; break;
            default: w[g] = 0u; break;
            }
        }
    }
    __syncthreads();
    if (CL > 1) cluster_sync_all();                              // barriers of every CTA are initialised before anyone copies into it

    LcChecks ck;
#pragma unroll
    for (int k = 0; k < LC_MAX_KEYS; ++k) ck.kd[k] = 0u;
    ck.vabs = 0ull; ck.vabs32 = 0u; ck.fx_bad = false;

    if (warp == 0) {
        // ---- producer: one lane arms the barrier and issues the bulk copies of a tile (its share of every column,
        //      delivered to all CTAs of the cluster)
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            const uint32_t part_rows = P.tile_rows / CL, part_bytes = tile_bytes / CL;
            for (uint32_t t = cid; t < nfull; t += ncl) {
                if (CL > 1) mbar_wait_cluster(empty + s, ph ^ 1u); else mbar_wait(empty + s, ph ^ 1u);
                mbar_arrive_expect_tx(full + s, stage_bytes);
                const uint32_t base = t * P.tile_rows + rank * part_rows;
                unsigned char* dst = ring + size_t(s) * stage_bytes + rank * part_bytes;
                for (int c = 0; c < ncols; ++c) {
                    if (CL > 1) bulk_g2s_multicast(dst + size_t(c) * tile_bytes, P.cols[c] + base, part_bytes, full + s, (uint16_t)((1u << CL) - 1u));
                    else bulk_g2s(dst + size_t(c) * tile_bytes, P.cols[c] + base, part_bytes, full + s);
                }
                if (++s == P.stages) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        // ---- consumers: RPT consecutive rows per thread and tile
        const uint32_t r0b = (threadIdx.x - 32u) * (4u * RPT);
        int s = 0; uint32_t ph = 0;
        for (uint32_t t = cid; t < nfull; t += ncl) {
            mbar_wait(full + s, ph);
            const unsigned char* const stage = ring + size_t(s) * stage_bytes + r0b;
            uint32_t slot[RPT];
#pragma unroll
            for (int j = 0; j < RPT; ++j) slot[j] = 0u;
#pragma unroll
            for (int k = 0; k < NK; ++k) {
                uint32_t kv[RPT];
                lc_ldrows<RPT>(stage + size_t(k) * tile_bytes, kv);
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    uint32_t d = kv[j] - P.klo[k];
                    ck.kd[k] = max(ck.kd[k], d);
                    d = min(d, P.krange[k]);                     // an out-of-range key aborts the query; keep the slot inside the table
                    slot[j] += d * P.kmul[k];
                }
            }
            for (int a = 0; a < nown; ++a) {
                const LcAccS A = sacc[a];
                uint32_t u[RPT], u2[RPT];
#pragma unroll
                for (int j = 0; j < RPT; ++j) u[j] = u2[j] = 0u;
                if (A.kind != LCK_CNT && A.kind != LCK_SEEN) {
                    lc_ldrows<RPT>(stage + A.c0, u);
                    if (A.xform == X_PRODUCT) lc_ldrows<RPT>(stage + A.c1, u2);
                    else {
#pragma unroll
                        for (int j = 0; j < RPT; ++j) u2[j] = u[j];
                    }
                }
                lc_accn<RPT>(P, A, sm, slot, u, u2, ck);
            }
            __syncwarp();
            if (lane == 0) {
                if (CL > 1) { for (uint32_t r = 0; r < (uint32_t)CL; ++r) mbar_arrive_cluster(empty + s, r); }
                else mbar_arrive(empty + s);
            }
            if (++s == P.stages) { s = 0; ph ^= 1u; }
        }
        // ---- ragged tail (fewer rows than a tile): the cluster whose turn it would be reads it directly
        if (nfull % ncl == cid) {
            for (uint32_t row = nfull * P.tile_rows + (threadIdx.x - 32u); row < P.nrows; row += blockDim.x - 32u) {
                uint32_t slot = 0;
#pragma unroll
                for (int k = 0; k < NK; ++k) {
                    uint32_t d = __ldg(P.cols[k] + row) - P.klo[k];
                    ck.kd[k] = max(ck.kd[k], d);
                    d = min(d, P.krange[k]);
                    slot += d * P.kmul[k];
                }
                for (int a = 0; a < nown; ++a) lc_acc1(P, sacc[a], sm, slot, row, ck);
            }
        }
        bool bad_key = false;
#pragma unroll
        for (int k = 0; k < NK; ++k) bad_key |= ck.kd[k] > P.krange[k];
        if (bad_key) atomicExch(P.flags + FLAG_ABORT, 3u);
        else if (ck.fx_bad) atomicCAS(P.flags + FLAG_ABORT, 0u, 4u);
        else if (ck.vabs > (unsigned long long)P.isum_lim || ck.vabs32 > P.isum_lim) atomicCAS(P.flags + FLAG_ABORT, 0u, 2u);
    }
    __syncthreads();

    // ---- merge this CTA's cells into the global direct table
    for (int a = 0; a < nown; ++a) {
        const LcAccS A = sacc[a];
        const uint32_t* w = reinterpret_cast<const uint32_t*>(sm + A.tab);
        for (uint32_t g = threadIdx.x; g < G; g += blockDim.x) {
            unsigned long long* gw = P.gtab + size_t(g) * P.gw + A.gword;
            switch (A.kind) {
            case LCK_SEEN: if ((sm + A.tab)[g]) *reinterpret_cast<volatile unsigned long long*>(gw) = 1ull; break;
            case LCK_CNT: { const uint32_t c = w[g]; if (c) atomicAdd(gw, (unsigned long long)c); break; }
            case LCK_ISUM1: { const int c = (int)w[g]; if (c) atomicAdd(gw, (unsigned long long)(long long)c); break; }
            case LCK_ISUM2: {
                const unsigned long long v = (unsigned long long)w[g] | ((unsigned long long)w[G + g] << 32);
                if (v) atomicAdd(gw, v);
                break; }
            case LCK_FX: fx_merge(w[g], w[G + g], P.fxQ, gw, gw + 1); break;
            case LCK_MIN32: {
                const int c = (int)w[g];
                if (c != 0x7fffffff) atomicMin(reinterpret_cast<long long*>(gw), image64_of32(c, A.vtype));
                break; }
            case LCK_MAX32: {
                const int c = (int)w[g];
                if (c != (int)0x80000000u) atomicMax(reinterpret_cast<long long*>(gw), image64_of32(c, A.vtype));
                break; }
            default: break;
            }
        }
    }
    if (CL > 1) cluster_sync_all();                              // nobody leaves while a peer may still signal its barriers
}

// global table: word pattern of one group record, first rows = none
__global__ void k_lc_init(unsigned long long* __restrict__ gtab, uint32_t G, uint32_t gw, const LcInit pat,
                          uint32_t* __restrict__ gfirst, uint32_t* __restrict__ ctl) {
    const size_t n = size_t(G) * gw;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) gtab[i] = pat.w[i % gw];
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < G; g += gridDim.x * blockDim.x) gfirst[g] = 0xffffffffu;
    if (blockIdx.x == 0 && threadIdx.x < 8) ctl[threadIdx.x] = 0u;
}

// ctl[0] = number of non-empty slots, slots[] = their indices (any order)
__global__ void k_lc_count(const unsigned long long* __restrict__ gtab, uint32_t G, uint32_t gw, uint32_t cnt_word,
                           uint32_t* __restrict__ slots, uint32_t* __restrict__ ctl) {
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < G; g += gridDim.x * blockDim.x)
        if (gtab[size_t(g) * gw + cnt_word] != 0ull) slots[atomicAdd(ctl + 0, 1u)] = g;
}

// ctl[1] = ticket, ctl[2] = groups whose first row is known
constexpr int LCF_THREADS = 256;
constexpr int LCF_ROWS = LCF_THREADS * 4;
__global__ void __launch_bounds__(LCF_THREADS)
k_lc_first(const LcKeys K, const uint32_t nrows, uint32_t* __restrict__ gfirst, uint32_t* __restrict__ ctl) {
    __shared__ uint32_t s_tile;
    const uint32_t ntiles = (nrows + LCF_ROWS - 1) / LCF_ROWS;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint32_t found = *reinterpret_cast<volatile uint32_t*>(ctl + 2), need = *reinterpret_cast<volatile uint32_t*>(ctl + 0);
            s_tile = found >= need ? 0xffffffffu : atomicAdd(ctl + 1, 1u);
        }
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= ntiles) return;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t row = tile * LCF_ROWS + j * LCF_THREADS + threadIdx.x;
            if (row >= nrows) continue;
            uint32_t slot = 0; bool ok = true;
            for (int k = 0; k < K.nkeys; ++k) {
                const uint32_t d = __ldg(K.key[k] + row) - K.klo[k];
                ok &= d <= K.krange[k];
                slot += d * K.kmul[k];
            }
            if (!ok) continue;                                   // k_lc_agg has raised the flag
            const uint32_t cur = *reinterpret_cast<volatile uint32_t*>(gfirst + slot);
            if (row < cur) {
                const uint32_t old = atomicMin(gfirst + slot, row);
                if (old == 0xffffffffu) atomicAdd(ctl + 2, 1u);
            }
        }
    }
}

// rank[i] = number of groups whose first row precedes that of group slots[i]
constexpr int LCR_THREADS = 256;
constexpr int LCR_CHUNK = 4096;
__global__ void __launch_bounds__(LCR_THREADS)
k_lc_rank(const uint32_t* __restrict__ slots, const uint32_t* __restrict__ gfirst, const uint32_t* __restrict__ ctl,
          uint32_t* __restrict__ rank) {
    __shared__ __align__(16) uint32_t sf[LCR_CHUNK];
    const uint32_t ne = ctl[0];
    const uint32_t i = blockIdx.x * LCR_THREADS + threadIdx.x;
    if (blockIdx.x * LCR_THREADS >= ne) return;
    const uint32_t mine = i < ne ? gfirst[slots[i]] : 0u;
    uint32_t r = 0;
    for (uint32_t c0 = 0; c0 < ne; c0 += LCR_CHUNK) {
        __syncthreads();
        for (uint32_t j = threadIdx.x; j < LCR_CHUNK; j += LCR_THREADS) sf[j] = c0 + j < ne ? gfirst[slots[c0 + j]] : 0xffffffffu;
        __syncthreads();
        const uint32_t lim = min((uint32_t)LCR_CHUNK, ne - c0);
        for (uint32_t j = 0; j < lim; j += 4) {
            const uint4 f = *reinterpret_cast<const uint4*>(sf + j);
            r += (f.x < mine) + (f.y < mine) + (f.z < mine) + (f.w < mine);
        }
    }
    if (i < ne) rank[i] = r;
}

__global__ void __launch_bounds__(256)
k_lc_emit(const LcParams P, const LcKeys K, const uint32_t* __restrict__ slots, const uint32_t* __restrict__ rank,
          const uint32_t* __restrict__ gfirst, const KeyOut ko, uint32_t* __restrict__ first_row_out, const EmitProg ep,
          const uint32_t max_groups, uint32_t* __restrict__ flags) {
    const uint32_t ne = flags[FLAG_NGROUPS];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ne; i += gridDim.x * blockDim.x) {
        const uint32_t s = slots[i], gid = rank[i], fr = gfirst[s];
        if (gid >= max_groups) continue;
        if (first_row_out) first_row_out[gid] = fr;
        for (int k = 0; k < K.nkeys; ++k) {
            if (!ko.dst[k]) continue;
            const uint32_t digit = (s / K.kmul[k]) % (K.krange[k] + 1u);
            static_cast<uint32_t*>(ko.dst[k])[gid] = digit + K.klo[k];
        }
        unsigned long long cell[1 + MAX_ACC];
        const unsigned long long* g = P.gtab + size_t(s) * P.gw;
        for (int a = 0; a < P.nacc; ++a) {
            const LcAcc& A = P.acc[a];
            unsigned long long v = g[A.gword];
            if (A.kind == LCK_FX) v = (unsigned long long)__double_as_longlong(fx_to_double(v, g[A.gword + 1], A.fxB));
            cell[1 + a] = v;
        }
        for (int a = 0; a < ep.n; ++a) emit_one(ep.agg[a], cell, gid, fr, ep.row_order);
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
template <int NK, int CL, int RPT>
static int launch_lc_agg_r(const LcParams& P, int grid, int threads, size_t smem) {
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        AQ_CUDA(cudaFuncSetAttribute(k_lc_agg<NK, CL, RPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem; cfg.stream = ctx().stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (ctx().profiling) profile_pre("k_lc_agg");
    const cudaError_t e = cudaLaunchKernelEx(&cfg, k_lc_agg<NK, CL, RPT>, P);
    if (ctx().profiling) profile_post();
    ctx().launches.fetch_add(1, std::memory_order_relaxed);
    if (e != cudaSuccess) return fail("%s:%d: launch k_lc_agg (cluster %d, %d threads, %zu B smem) -> %s", __FILE__, __LINE__, CL, threads,
                                      smem, cudaGetErrorString(e));
    return 0;
}
template <int NK, int CL>
static int launch_lc_agg_t(const LcParams& P, int grid, int threads, size_t smem) {
    const uint32_t rpt = P.tile_rows / (uint32_t)(threads - 32);
    if (rpt == 4) return launch_lc_agg_r<NK, CL, 4>(P, grid, threads, smem);
    if (rpt == 2) return launch_lc_agg_r<NK, CL, 2>(P, grid, threads, smem);
    return fail("internal: %u rows per thread on the direct-table path", rpt);
}
template <int NK>
static int launch_lc_agg_k(const LcParams& P, int grid, int threads, size_t smem) {
    switch (P.cluster) {
    case 1: return launch_lc_agg_t<NK, 1>(P, grid, threads, smem);
    case 2: return launch_lc_agg_t<NK, 2>(P, grid, threads, smem);
    case 4: return launch_lc_agg_t<NK, 4>(P, grid, threads, smem);
    default: return fail("internal: cluster size %d", P.cluster);
    }
}
static int launch_lc_agg(const LcParams& P, int grid, int threads, size_t smem) {
    switch (P.nkeys) {
    case 1: return launch_lc_agg_k<1>(P, grid, threads, smem);
    case 2: return launch_lc_agg_k<2>(P, grid, threads, smem);
    case 3: return launch_lc_agg_k<3>(P, grid, threads, smem);
    case 4: return launch_lc_agg_k<4>(P, grid, threads, smem);
    default: return fail("internal: %d key columns on the direct-table path", P.nkeys);
    }
}

static uint32_t lc_bytes_of(int kind, uint32_t G) {
    if (kind == LCK_SEEN) return (G + 15u) & ~15u;
    const uint32_t words = (kind == LCK_FX || kind == LCK_ISUM2) ? 2u : 1u;
    return words * 4u * G;
}

// Decides whether the low-cardinality path takes the query and lays the table out.  st: key bounds and
// per-accumulator magnitudes from the caller / the sample (plan.cu).
bool lc_layout(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, const PlanStats& st, int marker_acc, LcParams& P, LcKeys& K,
               int* threads_out, size_t* smem_out) {
    Context& c = ctx();
    memset(&P, 0, sizeof(P)); memset(&K, 0, sizeof(K));
    const uint32_t nrows_local = nrows;
    if (dist().active) nrows = dist().rows_max;                  // row shards: the same layout on every rank
    if (!st.have_bounds || kc.n > LC_MAX_KEYS || nrows < 4096) return false;
    unsigned long long G = 1;
    for (int k = 0; k < kc.n; ++k) {
        if (kc.type[k] != AQ_INT32 && kc.type[k] != AQ_UINT32) return false;
        if (reinterpret_cast<uintptr_t>(kc.col[k].d) & 15) return false;
        const long long range = st.kmax[k] - st.kmin[k];
        if (range < 0 || range > 0xffffffffLL) return false;
        P.klo[k] = (uint32_t)st.kmin[k]; P.krange[k] = (uint32_t)range; P.kmul[k] = (uint32_t)G;
        G *= (unsigned long long)range + 1ull;
        if (G > LC_MAX_G) return false;
    }
    P.nkeys = kc.n; P.G = (uint32_t)G; P.nrows = nrows_local; P.nacc = plan.n_acc; P.flags = c.d_flags;
    K.nkeys = kc.n;
    for (int k = 0; k < kc.n; ++k) { K.key[k] = static_cast<const uint32_t*>(kc.col[k].d); K.klo[k] = P.klo[k]; K.krange[k] = P.krange[k]; K.kmul[k] = P.kmul[k]; }
    // ---- ring columns: keys, then every value column once
    P.ncols = 0;
    for (int k = 0; k < kc.n; ++k) P.cols[P.ncols++] = static_cast<const uint32_t*>(kc.col[k].d);
    bool ok = true;
    auto col_of = [&](const void* p) -> int {
        if (!p) return 0;
        for (int i = 0; i < P.ncols; ++i) if (P.cols[i] == p) return i;
        if (P.ncols >= LC_MAX_COLS) { ok = false; return 0; }
        P.cols[P.ncols] = static_cast<const uint32_t*>(p);
        return P.ncols++;
    };
    bool have_cnt = false;
    for (int a = 0; a < plan.n_acc; ++a) {
        const AccOp& o = plan.prog.op[a];
        LcAcc& A = P.acc[a];
        A.xform = o.xform; A.vtype = o.vtype; A.fxB = 0;
        if (o.kind != ACC_COUNT) {
            const bool v32 = o.vtype == AQ_INT32 || o.vtype == AQ_FLOAT;
            if (!v32 || (reinterpret_cast<uintptr_t>(o.v) & 15) || (o.v2 && (reinterpret_cast<uintptr_t>(o.v2) & 15))) return false;
            A.c0 = col_of(o.v);
            A.c1 = col_of(o.xform == X_PRODUCT ? o.v2 : nullptr);
        } else have_cnt = true;
    }
    if (!have_cnt || !ok) return false;                          // the caller adds a COUNT accumulator first
    // ---- cluster size, thread count, cell kinds, accumulator owners
    const size_t budget = c.smem_optin - 1024;                  // static shared memory of the kernel + slack
    for (int cl = 1; cl <= 4; cl *= 2) {
        for (int variant = 0; variant < 3; ++variant) {               // (threads, rows per thread and tile): fewer rows = thinner ring stages
            const int threads = variant < 2 ? 1024 : 512;
            const uint32_t rpt = variant == 1 ? 2u : 4u;
            const uint32_t tile_rows = (uint32_t)(threads - 32) * rpt;
            const uint32_t nfull = nrows / tile_rows;
            const uint32_t nclusters = std::max<uint32_t>(1u, std::min<uint32_t>((uint32_t)c.sm_count / cl, std::max<uint32_t>(nfull, 1u)));
            // rows one CTA can see: its share of the full tiles and the ragged tail
            const unsigned long long rows_cta = (unsigned long long)((nfull + nclusters - 1) / nclusters) * tile_rows + tile_rows;
            const unsigned long long lim = std::min<unsigned long long>(0x7fffffffULL / rows_cta, 0x7fffffffULL);
            uint32_t gw = 0, bytes[MAX_ACC];
            for (int a = 0; a < plan.n_acc; ++a) {
                const AccOp& o = plan.prog.op[a];
                LcAcc& A = P.acc[a];
                A.gword = gw;
                switch (o.kind) {
                case ACC_COUNT: A.kind = a == marker_acc ? LCK_SEEN : LCK_CNT; gw += 1; break;
                case ACC_ADD_I64:
                    if (o.vtype != AQ_INT32) return false;
                    A.kind = (lim > 0 && st.isum_maxabs[a] <= lim) ? LCK_ISUM1 : LCK_ISUM2; gw += 1; break;
                case ACC_ADD_F64:
                    if (o.vtype != AQ_FLOAT || st.fx_maxexp[a] < 0) return false;
                    A.kind = LCK_FX; A.fxB = fx_base(st.fx_maxexp[a]); gw += 2; break;
                case ACC_MIN_I64: A.kind = LCK_MIN32; gw += 1; break;
                case ACC_MAX_I64: A.kind = LCK_MAX32; gw += 1; break;
                default: return false;                           // MAXROW / TOP2ROW: generic path
                }
                bytes[a] = lc_bytes_of(A.kind, P.G);
            }
            uint32_t load[4] = {0, 0, 0, 0};
            int owner[MAX_ACC];
            bool used[MAX_ACC];
            for (int a = 0; a < plan.n_acc; ++a) used[a] = false;
            for (int n = 0; n < plan.n_acc; ++n) {               // largest first into the least loaded CTA
                int best = -1;
                for (int a = 0; a < plan.n_acc; ++a) if (!used[a] && (best < 0 || bytes[a] > bytes[best])) best = a;
                int r = 0;
                for (int x = 1; x < cl; ++x) if (load[x] < load[r]) r = x;
                owner[best] = r; load[r] += bytes[best]; used[best] = true;
            }
            uint32_t tab = 0;
            for (int r = 0; r < cl; ++r) tab = std::max(tab, load[r]);
            tab = (tab + 15u) & ~15u;
            const size_t stage = size_t(P.ncols) * tile_rows * 4;
            if (LC_RING_OFF + tab + 2 * stage > budget) continue;
            int stages = (int)std::min<size_t>((budget - LC_RING_OFF - tab) / stage, (size_t)LC_MAX_STAGES);
            // ---- fits: fill in
            P.cluster = cl; P.stages = stages; P.tile_rows = tile_rows; P.ntiles = nfull; P.gw = gw;
            P.isum_lim = (uint32_t)lim;
            int lg = 0;
            while ((1ull << lg) < rows_cta) ++lg;
            P.fxR = std::max(0, std::min(39 - lg, FX_S_TOP - 1));       // |sum of a CTA's cell| < rows_cta * 2^(24+R) <= 2^63
            P.fxQ = FX_S_TOP - P.fxR;
            uint32_t off[4] = {0, 0, 0, 0};
            for (int a = 0; a < plan.n_acc; ++a) { P.acc[a].owner = owner[a]; P.acc[a].off = off[owner[a]]; off[owner[a]] += bytes[a]; }
            *threads_out = threads;
            *smem_out = LC_RING_OFF + size_t(stages) * stage + tab;
            return true;
        }
    }
    return false;
}

// Runs the path.  *used = false (and nothing emitted) when a per-row check failed: the caller falls back.
int run_lc(const KeyCols& kc, uint32_t nrows, AggPlan& plan, LcParams& P, const LcKeys& K, int threads, size_t smem,
           OutSink& sink, uint32_t* ngroups, bool* used) {
    Context& c = ctx();
    *used = false;
    const int cnt_acc = plan.find(ACC_COUNT);
    DevBuf gtab, gfirst, slots, rank;
    AQ_TRY(gtab.alloc(size_t(P.G) * P.gw * 8));
    AQ_TRY(gfirst.alloc(size_t(P.G) * 4));
    AQ_TRY(slots.alloc(size_t(P.G) * 4));
    AQ_TRY(rank.alloc(size_t(P.G) * 4));
    P.gtab = gtab.as<unsigned long long>();
    uint32_t* const ctl = c.d_flags + 16;
    LcInit pat;
    memset(&pat, 0, sizeof(pat));
    for (int a = 0; a < P.nacc; ++a) {
        if (P.acc[a].kind == LCK_MIN32) pat.w[P.acc[a].gword] = 0x7fffffffffffffffULL;
        if (P.acc[a].kind == LCK_MAX32) pat.w[P.acc[a].gword] = 0x8000000000000000ULL;
    }
    AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 8 * sizeof(uint32_t), c.stream));
    AQ_LAUNCH(k_lc_init, grid_for(size_t(P.G) * P.gw, 256, 1, 4), 256, 0, P.gtab, P.G, P.gw, pat, gfirst.as<uint32_t>(), ctl);
    int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c.sm_count / P.cluster, (uint64_t)P.ntiles)) * P.cluster;
    AQ_TRY(launch_lc_agg(P, grid, threads, smem));
    AQ_LAUNCH(k_lc_count, grid_for(P.G, 256, 1, 4), 256, 0, P.gtab, P.G, P.gw, P.acc[cnt_acc].gword, slots.as<uint32_t>(), ctl);
    AQ_LAUNCH(k_lc_first, c.sm_count * 4, LCF_THREADS, 0, K, nrows, gfirst.as<uint32_t>(), ctl);
    const bool sharded = dist().active;
    if (sharded) {
        // row shards: all-gather the record tables, combine element-wise (exact integer / fixed-point cells), recount
        RecLayout L;
        memset(&L, 0, sizeof(L));
        L.gw = P.gw; L.n = P.nacc; L.cnt_word = P.acc[cnt_acc].gword;
        for (int a = 0; a < P.nacc; ++a) {
            L.word[a] = P.acc[a].gword;
            L.kind[a] = P.acc[a].kind == LCK_FX ? RK_ADD128 : (P.acc[a].kind == LCK_MIN32 ? RK_MIN64 : (P.acc[a].kind == LCK_MAX32 ? RK_MAX64 : RK_ADD64));
        }
        AQ_TRY(dist_combine_replicated(P.gtab, gfirst.as<uint32_t>(), P.G, L, c.d_flags, nullptr));
        AQ_CUDA(cudaMemsetAsync(ctl, 0, 4, c.stream));
        AQ_LAUNCH(k_lc_count, grid_for(P.G, 256, 1, 4), 256, 0, P.gtab, P.G, P.gw, P.acc[cnt_acc].gword, slots.as<uint32_t>(), ctl);
    }
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 24 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) return 0;                     // speculation failed (on some rank): generic path
    const uint32_t ng = c.h_flags[16];
    if (!sharded && c.h_flags[18] != ng) return fail("internal: low-cardinality path found first rows of %u of %u groups", c.h_flags[18], ng);
    if (!sink.res && ng > sink.max_groups) return fail("capacity: %u groups exceed max_groups=%u", ng, sink.max_groups);
    AQ_TRY(sink.prepare(ng, plan));
    KeyOut ko;
    memset(&ko, 0, sizeof(ko));
    ko.n = kc.n;
    for (int k = 0; k < kc.n; ++k) { ko.src[k] = kc.col[k].d; ko.size[k] = 4; ko.dst[k] = sink.key_dst[k]; }
    for (int a = 0; a < plan.emit.n; ++a) plan.emit.agg[a].out = sink.agg_dst[a];
    AQ_CUDA(cudaMemcpyAsync(c.d_flags + FLAG_NGROUPS, c.d_flags + 16, 4, cudaMemcpyDeviceToDevice, c.stream));
    if (ng) {
        AQ_LAUNCH(k_lc_rank, (ng + LCR_THREADS - 1) / LCR_THREADS, LCR_THREADS, 0, slots.as<uint32_t>(), gfirst.as<uint32_t>(), ctl, rank.as<uint32_t>());
        AQ_LAUNCH(k_lc_emit, grid_for(ng, 256, 1, 4), 256, 0, P, K, slots.as<uint32_t>(), rank.as<uint32_t>(), gfirst.as<uint32_t>(), ko,
                  sink.fr_dst, plan.emit, sink.cap, c.d_flags);
    }
    AQ_TRY(sink.finish(ng, plan));
    *ngroups = ng;
    *used = true;
    return 0;
}

} // namespace aq
