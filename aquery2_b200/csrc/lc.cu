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
#include <algorithm>
#include <cstring>

namespace aq {

constexpr int LC_RING_OFF = 1024;       // barriers live in the first KB of dynamic shared memory

__device__ __forceinline__ uint32_t* lc_words(unsigned char* table, const LcAcc& a, uint32_t G, int w) {
    return reinterpret_cast<uint32_t*>(table + a.off) + size_t(w) * G;
}

__global__ void __launch_bounds__(1024, 1)
k_lc_agg(const LcParams P) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t* const full = reinterpret_cast<uint64_t*>(sm);
    uint64_t* const empty = full + LC_MAX_STAGES;
    const uint32_t rank = P.cluster > 1 ? cluster_ctarank() : 0u;
    const uint32_t cid = blockIdx.x / P.cluster, ncl = gridDim.x / P.cluster;
    const int ncols = P.ncols[rank];
    const uint32_t tile_bytes = P.tile_rows * 4u;
    unsigned char* const ring = sm + LC_RING_OFF;
    unsigned char* const table = ring + size_t(P.stages) * ncols * tile_bytes;
    const uint32_t G = P.G;
    const int ncw = (int)(blockDim.x >> 5) - 1;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;

    if (threadIdx.x == 0) {
        for (int s = 0; s < P.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, (uint32_t)ncw); }
        mbar_fence_init();
    }
    for (int a = 0; a < P.nacc; ++a) {
        const LcAcc& A = P.acc[a];
        if ((uint32_t)A.owner != rank) continue;
        for (uint32_t g = threadIdx.x; g < G; g += blockDim.x) {
            switch (A.kind) {
            case LCK_ISUM1: lc_words(table, A, G, 0)[g] = CELL_BIAS; break;
            case LCK_ISUM2: lc_words(table, A, G, 0)[g] = 0u; lc_words(table, A, G, 1)[g] = 0u; break;
            case LCK_FX: lc_words(table, A, G, 0)[g] = 0u; lc_words(table, A, G, 1)[g] = 0u; lc_words(table, A, G, 2)[g] = CELL_BIAS; break;
            case LCK_MIN32: lc_words(table, A, G, 0)[g] = 0x7fffffffu; break;
            case LCK_MAX32: lc_words(table, A, G, 0)[g] = 0x80000000u; break;
            default: lc_words(table, A, G, 0)[g] = 0u; break;
            }
        }
    }
    __syncthreads();

    if (warp == 0) {
        // ---- producer: one lane arms the barrier and issues one bulk copy per column and tile
        uint32_t it = 0;
        for (uint32_t t = cid; t < P.ntiles; t += ncl, ++it) {
            if (lane == 0) {
                const int s = (int)(it % (uint32_t)P.stages);
                const uint32_t ph = (it / (uint32_t)P.stages) & 1u;
                mbar_wait(empty + s, ph ^ 1u);
                const uint32_t base = t * P.tile_rows;
                const uint32_t valid = min(P.tile_rows, P.nrows - base);
                const uint32_t bytes = (valid * 4u) & ~15u;
                mbar_arrive_expect_tx(full + s, bytes * (uint32_t)ncols);
                if (bytes)
                    for (int c = 0; c < ncols; ++c)
                        bulk_g2s(ring + (size_t(s) * ncols + c) * tile_bytes, P.cols[rank][c] + base, bytes, full + s);
            }
            __syncwarp();
        }
    } else {
        // ---- consumers: 4 consecutive rows per thread and tile
        const uint32_t r0 = (threadIdx.x - 32u) * 4u;
        uint32_t it = 0;
        bool bad_key = false, bad_fx = false;
        for (uint32_t t = cid; t < P.ntiles; t += ncl, ++it) {
            const int s = (int)(it % (uint32_t)P.stages);
            const uint32_t ph = (it / (uint32_t)P.stages) & 1u;
            mbar_wait(full + s, ph);
            const uint32_t base = t * P.tile_rows;
            const uint32_t valid = min(P.tile_rows, P.nrows - base);
            const int nv = r0 < valid ? (int)min(4u, valid - r0) : 0;
            if (nv) {
                const unsigned char* const stage = ring + size_t(s) * ncols * tile_bytes;
                auto quad = [&](int c) -> uint4 {
                    if (nv == 4) return lds128(stage + size_t(c) * tile_bytes + r0 * 4u);
                    const uint32_t* g = P.cols[rank][c] + base + r0;      // ragged tail: not covered by the bulk copy
                    uint4 q = make_uint4(0, 0, 0, 0);
                    q.x = __ldg(g);
                    if (nv > 1) q.y = __ldg(g + 1);
                    if (nv > 2) q.z = __ldg(g + 2);
                    return q;
                };
                uint32_t slot[4] = {0, 0, 0, 0};
                uint32_t okm = (1u << nv) - 1u;
#pragma unroll
                for (int k = 0; k < LC_MAX_KEYS; ++k) {
                    if (k >= P.nkeys) break;
                    const uint4 kq = quad(k);
                    const uint32_t kv[4] = {kq.x, kq.y, kq.z, kq.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t d = kv[j] - P.klo[k];
                        if (d > P.krange[k]) { if (okm & (1u << j)) bad_key = true; okm &= ~(1u << j); }
                        slot[j] += d * P.kmul[k];
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) if (!(okm & (1u << j))) slot[j] = 0;
                for (int a = 0; a < P.nacc; ++a) {
                    const LcAcc& A = P.acc[a];
                    if ((uint32_t)A.owner != rank) continue;
                    uint32_t u[4] = {0, 0, 0, 0}, u2[4] = {0, 0, 0, 0};
                    if (A.kind != LCK_CNT) {
                        const uint4 q = quad(A.c0);
                        u[0] = q.x; u[1] = q.y; u[2] = q.z; u[3] = q.w;
                        if (A.xform == X_PRODUCT) { const uint4 q2 = quad(A.c1); u2[0] = q2.x; u2[1] = q2.y; u2[2] = q2.z; u2[3] = q2.w; }
                        else { u2[0] = u[0]; u2[1] = u[1]; u2[2] = u[2]; u2[3] = u[3]; }
                    }
                    switch (A.kind) {
                    case LCK_CNT: {
                        uint32_t* c = lc_words(table, A, G, 0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) if (okm & (1u << j)) atomicAdd(c + slot[j], 1u);
                        break; }
                    case LCK_ISUM1: {
                        uint32_t* c = lc_words(table, A, G, 0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (!(okm & (1u << j))) continue;
                            const long long x = A.xform == X_IDENT ? (long long)(int)u[j] : (long long)(int)u[j] * (long long)(int)u2[j];
                            isum1_add(c + slot[j], P.gtab + size_t(slot[j]) * P.gw + A.gword + 1, x);
                        }
                        break; }
                    case LCK_ISUM2: {
                        uint32_t* lo = lc_words(table, A, G, 0);
                        uint32_t* hi = lc_words(table, A, G, 1);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (!(okm & (1u << j))) continue;
                            const long long x = A.xform == X_IDENT ? (long long)(int)u[j] : (long long)(int)u[j] * (long long)(int)u2[j];
                            isum2_add(lo + slot[j], hi + slot[j], x);
                        }
                        break; }
                    case LCK_FX: {
                        uint32_t* w1 = lc_words(table, A, G, 0);
                        uint32_t* w2 = lc_words(table, A, G, 1);
                        uint32_t* w3 = lc_words(table, A, G, 2);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (!(okm & (1u << j))) continue;
                            const float f = __uint_as_float(u[j]);
                            const float x = A.xform == X_IDENT ? f : __fmul_rn(f, __uint_as_float(u2[j]));   // product in T (aggregations.h:339-345)
                            unsigned long long* g = P.gtab + size_t(slot[j]) * P.gw + A.gword;
                            if (!fx_add(w1 + slot[j], w2 + slot[j], w3 + slot[j], g, g + 1, x, A.fxB)) bad_fx = true;
                        }
                        break; }
                    case LCK_MIN32: {
                        int* c = reinterpret_cast<int*>(lc_words(table, A, G, 0));
#pragma unroll
                        for (int j = 0; j < 4; ++j) if (okm & (1u << j)) min32_upd(c + slot[j], ordered32(u[j], A.vtype));
                        break; }
                    case LCK_MAX32: {
                        int* c = reinterpret_cast<int*>(lc_words(table, A, G, 0));
#pragma unroll
                        for (int j = 0; j < 4; ++j) if (okm & (1u << j)) max32_upd(c + slot[j], ordered32(u[j], A.vtype));
                        break; }
                    default: break;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty + s);
        }
        if (bad_key) atomicExch(P.flags + FLAG_ABORT, 3u);
        else if (bad_fx) atomicCAS(P.flags + FLAG_ABORT, 0u, 4u);
    }
    __syncthreads();

    // ---- merge this CTA's cells into the global direct table
    for (int a = 0; a < P.nacc; ++a) {
        const LcAcc& A = P.acc[a];
        if ((uint32_t)A.owner != rank) continue;
        for (uint32_t g = threadIdx.x; g < G; g += blockDim.x) {
            unsigned long long* gw = P.gtab + size_t(g) * P.gw + A.gword;
            switch (A.kind) {
            case LCK_CNT: { const uint32_t c = lc_words(table, A, G, 0)[g]; if (c) atomicAdd(gw, (unsigned long long)c); break; }
            case LCK_ISUM1: {
                const uint32_t c = lc_words(table, A, G, 0)[g];
                if (c != CELL_BIAS) atomicAdd(gw, (unsigned long long)((long long)c - (long long)CELL_BIAS));
                break; }
            case LCK_ISUM2: {
                const unsigned long long v = (unsigned long long)lc_words(table, A, G, 0)[g] | ((unsigned long long)lc_words(table, A, G, 1)[g] << 32);
                if (v) atomicAdd(gw, v);
                break; }
            case LCK_FX: fx_merge(lc_words(table, A, G, 0)[g], lc_words(table, A, G, 1)[g], lc_words(table, A, G, 2)[g], gw, gw + 1); break;
            case LCK_MIN32: {
                const int c = (int)lc_words(table, A, G, 0)[g];
                if (c != 0x7fffffff) atomicMin(reinterpret_cast<long long*>(gw), image64_of32(c, A.vtype));
                break; }
            case LCK_MAX32: {
                const int c = (int)lc_words(table, A, G, 0)[g];
                if (c != (int)0x80000000u) atomicMax(reinterpret_cast<long long*>(gw), image64_of32(c, A.vtype));
                break; }
            default: break;
            }
        }
    }
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
            if (A.kind == LCK_ISUM1 || A.kind == LCK_ISUM2) {
                if (A.kind == LCK_ISUM1) {
                    const __int128 t = (__int128)(long long)v + ((__int128)(long long)g[A.gword + 1] << 32);
                    if (t != (__int128)(long long)t) atomicCAS(flags + FLAG_ABORT, 0u, 5u);   // beyond int64: generic path decides
                    v = (unsigned long long)(long long)t;
                }
            } else if (A.kind == LCK_FX) {
                v = (unsigned long long)__double_as_longlong(fx_to_double(v, g[A.gword + 1], A.fxB));
            }
            cell[1 + a] = v;
        }
        for (int a = 0; a < ep.n; ++a) emit_one(ep.agg[a], cell, gid, fr, ep.row_order);
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
static int launch_lc_agg(const LcParams& P, int grid, int threads, size_t smem) {
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        AQ_CUDA(cudaFuncSetAttribute(k_lc_agg, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem; cfg.stream = ctx().stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)P.cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (ctx().profiling) profile_pre("k_lc_agg");
    const cudaError_t e = cudaLaunchKernelEx(&cfg, k_lc_agg, P);
    if (ctx().profiling) profile_post();
    ctx().launches.fetch_add(1, std::memory_order_relaxed);
    if (e != cudaSuccess) return fail("%s:%d: launch k_lc_agg (cluster %d, %d threads, %zu B smem) -> %s", __FILE__, __LINE__, P.cluster, threads,
                                      smem, cudaGetErrorString(e));
    return 0;
}

static uint32_t lc_words_of(int kind) { return kind == LCK_FX ? 3u : (kind == LCK_ISUM2 ? 2u : 1u); }

// Decides whether the low-cardinality path takes the query and lays the table out.  st: key bounds and
// per-accumulator magnitudes from the caller / the sample (plan.cu).
bool lc_layout(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, const PlanStats& st, LcParams& P, LcKeys& K,
               int* threads_out, size_t* smem_out) {
    Context& c = ctx();
    memset(&P, 0, sizeof(P)); memset(&K, 0, sizeof(K));
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
    P.nkeys = kc.n; P.G = (uint32_t)G; P.nrows = nrows; P.nacc = plan.n_acc; P.flags = c.d_flags;
    K.nkeys = kc.n;
    for (int k = 0; k < kc.n; ++k) { K.key[k] = static_cast<const uint32_t*>(kc.col[k].d); K.klo[k] = P.klo[k]; K.krange[k] = P.krange[k]; K.kmul[k] = P.kmul[k]; }
    // ---- accumulators -> cell kinds
    uint32_t gw = 0, bytes[MAX_ACC];
    bool have_cnt = false;
    for (int a = 0; a < plan.n_acc; ++a) {
        const AccOp& o = plan.prog.op[a];
        LcAcc& A = P.acc[a];
        A.xform = o.xform; A.vtype = o.vtype; A.gword = gw; A.fxB = 0;
        const bool v32 = o.vtype == AQ_INT32 || o.vtype == AQ_FLOAT;
        if (o.kind != ACC_COUNT) {
            if (!v32 || (reinterpret_cast<uintptr_t>(o.v) & 15) || (o.v2 && (reinterpret_cast<uintptr_t>(o.v2) & 15))) return false;
        }
        switch (o.kind) {
        case ACC_COUNT: A.kind = LCK_CNT; gw += 1; have_cnt = true; break;
        case ACC_ADD_I64:
            if (o.vtype != AQ_INT32) return false;
            A.kind = st.isum_maxabs[a] < (1ull << 24) ? LCK_ISUM1 : LCK_ISUM2; gw += 2; break;
        case ACC_ADD_F64:
            if (o.vtype != AQ_FLOAT || st.fx_maxexp[a] < 0) return false;
            A.kind = LCK_FX; A.fxB = fx_base(st.fx_maxexp[a]); gw += 2; break;
        case ACC_MIN_I64: A.kind = LCK_MIN32; gw += 1; break;
        case ACC_MAX_I64: A.kind = LCK_MAX32; gw += 1; break;
        default: return false;                                   // MAXROW / TOP2ROW: generic path
        }
        bytes[a] = lc_words_of(A.kind) * 4u * P.G;
    }
    if (!have_cnt) return false;                                 // the caller adds a COUNT accumulator first
    P.gw = gw;
    // ---- cluster size, accumulator owners, ring
    const size_t budget = c.smem_optin;
    for (int cl = 1; cl <= 4; cl *= 2) {
        for (int threads = 1024; threads >= 512; threads /= 2) {
            uint32_t load[4] = {0, 0, 0, 0};
            int owner[MAX_ACC];
            bool used[MAX_ACC];
            for (int a = 0; a < plan.n_acc; ++a) used[a] = false;
            for (int n = 0; n < plan.n_acc; ++n) {                 // largest first into the least loaded CTA
                int best = -1;
                for (int a = 0; a < plan.n_acc; ++a) if (!used[a] && (best < 0 || bytes[a] > bytes[best])) best = a;
                int r = 0;
                for (int x = 1; x < cl; ++x) if (load[x] < load[r]) r = x;
                owner[best] = r; load[r] += bytes[best]; used[best] = true;
            }
            const uint32_t tile_rows = (uint32_t)(threads / 32 - 1) * 128u;
            bool ok = true;
            int stages = LC_MAX_STAGES;
            int ncols[4]; const uint32_t* cols[4][LC_MAX_COLS]; int c0[MAX_ACC], c1[MAX_ACC];
            for (int r = 0; r < cl && ok; ++r) {
                ncols[r] = 0;
                for (int k = 0; k < kc.n; ++k) cols[r][ncols[r]++] = static_cast<const uint32_t*>(kc.col[k].d);
                for (int a = 0; a < plan.n_acc && ok; ++a) {
                    if (owner[a] != r) continue;
                    const AccOp& o = plan.prog.op[a];
                    auto slot_of = [&](const void* p) -> int {
                        if (!p) return -1;
                        for (int i = 0; i < ncols[r]; ++i) if (cols[r][i] == p) return i;
                        if (ncols[r] >= LC_MAX_COLS) { ok = false; return -1; }
                        cols[r][ncols[r]] = static_cast<const uint32_t*>(p);
                        return ncols[r]++;
                    };
                    c0[a] = slot_of(o.kind == ACC_COUNT ? nullptr : o.v);
                    c1[a] = slot_of(o.xform == X_PRODUCT ? o.v2 : nullptr);
                }
                if (!ok) break;
                const size_t tab = (size_t(load[r]) + 15) & ~size_t(15);
                if (LC_RING_OFF + tab + 2 * size_t(ncols[r]) * tile_rows * 4 > budget) { ok = false; break; }
                const int fit = (int)((budget - LC_RING_OFF - tab) / (size_t(ncols[r]) * tile_rows * 4));
                if (fit < stages) stages = fit;
            }
            if (!ok || stages < 2) continue;
            // ---- fits: fill in
            P.cluster = cl; P.stages = stages; P.tile_rows = tile_rows; P.ntiles = (nrows + tile_rows - 1) / tile_rows;
            size_t smem = 0;
            for (int r = 0; r < cl; ++r) {
                P.ncols[r] = ncols[r];
                for (int i = 0; i < ncols[r]; ++i) P.cols[r][i] = cols[r][i];
                uint32_t off = 0;
                for (int a = 0; a < plan.n_acc; ++a) if (owner[a] == r) { P.acc[a].off = off; off += bytes[a]; }
                const size_t need = LC_RING_OFF + size_t(stages) * ncols[r] * tile_rows * 4 + ((size_t(load[r]) + 15) & ~size_t(15));
                if (need > smem) smem = need;
            }
            for (int a = 0; a < plan.n_acc; ++a) { P.acc[a].owner = owner[a]; P.acc[a].c0 = c0[a] < 0 ? 0 : c0[a]; P.acc[a].c1 = c1[a] < 0 ? 0 : c1[a]; }
            *threads_out = threads; *smem_out = smem;
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
    int grid = (int)std::min<uint64_t>((uint64_t)c.sm_count / P.cluster, (uint64_t)P.ntiles) * P.cluster;
    if (grid < P.cluster) grid = P.cluster;
    AQ_TRY(launch_lc_agg(P, grid, threads, smem));
    AQ_LAUNCH(k_lc_count, grid_for(P.G, 256, 1, 4), 256, 0, P.gtab, P.G, P.gw, P.acc[cnt_acc].gword, slots.as<uint32_t>(), ctl);
    AQ_LAUNCH(k_lc_first, c.sm_count * 4, LCF_THREADS, 0, K, nrows, gfirst.as<uint32_t>(), ctl);
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 24 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) return 0;                     // speculation failed: generic path
    const uint32_t ng = c.h_flags[16];
    if (c.h_flags[18] != ng) return fail("internal: low-cardinality path found first rows of %u of %u groups", c.h_flags[18], ng);
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
        bool big = false;
        for (int a = 0; a < P.nacc; ++a) big |= P.acc[a].kind == LCK_ISUM1;
        if (big) {                                                // a sum beyond int64 is reported, never wrapped
            AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
            AQ_CUDA(cudaStreamSynchronize(c.stream));
            if (c.h_flags[FLAG_ABORT] == 5) return fail("integer sum exceeds the int64 range of the accumulator cells");
        }
    }
    AQ_TRY(sink.finish(ng, plan));
    *ngroups = ng;
    *used = true;
    return 0;
}

} // namespace aq
