// mc.cu -- mid-cardinality group-by (h2o q3, q5, q7, q8: 1e7 groups of ~100 rows): range-partitioned
// direct tables.  ONE partitioning pass whose writes are row-free compact tuples, then ONE aggregation pass
// in which every partition is owned by exactly one CTA (no merges, no global atomics on the common path).
//
// Replaces, for bounded 32-bit key tuples whose mixed-radix image has up to ~2.6e7 slots, the reference's
// per-row hashtable_push + ht_postproc + gather + aggregate loops (server/hasher.h:176-198,
// server/table.h:184-189, server/aggregations.h:10-86, 332-415).  Why partition at all: a DRAM-resident
// group table costs one random sector read-modify-write per row (~2e10/s on B200, profiles/r01/
// microbench.txt), shared-memory atomics run 50x faster, so the rows are first brought to the SM that owns
// their key range.
//
//   k_mc_hist     rows per (partition, scatter CTA): one pass over the key columns (scatter CTA c owns the
//                 contiguous row range [c * rpc, (c + 1) * rpc)); exclusive scan -> exact, deterministic
//                 positions: no cursors, no global atomics, no speculative capacities.
//   k_mc_scatter  persistent, one CTA per SM: a producer lane streams the needed columns through a bulk-copy
//                 ring (stream.cuh: cp.async.bulk + mbarrier); the consumer warps turn every row into a tuple
//                 (16-bit slot inside the partition + 32-bit values; NO row id; 8 or 16 bytes) and append it
//                 to the partition's 128-byte staging line in shared memory.  ONLY WHOLE LINES are written
//                 to HBM (8 lanes x 16 bytes = one request): B200's L2 takes ~4e10 write requests per second
//                 whatever their size (profiles/r02/ncu_mc_v1_scatter.txt: 6.2e8 sector-sized pieces of
//                 sorted runs cost 15 ms where the bytes alone would take 4), so the line is the unit.
//   k_mc_agg      one CTA per partition: directly indexed shared-memory table of compact 32-bit cells
//                 (cells.cuh), tuples streamed with 128-bit loads one batch ahead; the finished groups are
//                 written to a dense global record table with plain stores.
//   k_mc_first    first-appearance row of every group: rows are scanned from row 0 in ticket order until every
//                 group has been seen (uniform keys: ~17 % of the rows at 100 rows per group).
//   k_mc_mark, bitmap_rank, k_mc_emit
//                 first rows -> bitmap -> prefix popcount = the insertion index m_values has in the reference
//                 (server/unordered_dense.h:1142-1146); the emit walks the bitmap in row order, so every
//                 output column is written coalesced at consecutive gids.
//
// Everything speculated from the planner's sample (key ranges, fixed-point exponent base, 32-bit integer
// cells) is checked per row; a violation raises FLAG_ABORT and the caller re-runs the query on the generic path.
#include "lc.cuh"
#include "stream.cuh"
#include "cells.cuh"
#include "emit.cuh"
#include "sort.cuh"
#include "comm.cuh"
#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace aq {

constexpr int MC_THREADS = 1024;
constexpr int MC_CONS = MC_THREADS - 32;        // consumer threads of k_mc_scatter (warp 0 is the producer)
constexpr int MC_SUB = MC_CONS * 2;             // rows per ring stage = per tile: 2 per consumer
constexpr int MC_MAX_KEYS = 4;
constexpr int MC_MAX_V = 4;                     // value fields of a tuple (a generated row id counts)
constexpr uint32_t MC_PMAX = 1024;              // partitions (one 128-byte staging line each)
constexpr int MC_HSPLIT = 4;                    // k_mc_hist CTAs per scatter CTA
constexpr int MC_MAX_STAGES = 8;
constexpr uint32_t MC_INVALID = 0xffffffffu;
constexpr int MC_AGG_TPT = 8;                   // consecutive tuples per 128-bit slot load of k_mc_agg
constexpr int MC_AGG_THREADS = 512;
constexpr uint32_t MC_GMAX = 1u << 28;

enum McKind : int { MCK_CNT = 0, MCK_ISUM1, MCK_ISUM2, MCK_FX, MCK_MIN32, MCK_MAX32, MCK_MAXROW, MCK_TOP2 };

struct McAcc {
    int kind, xform, vtype, fxB;
    int c0, c1;             // tuple value field of v / v2
    uint32_t off;           // byte offset of the first word array in the shared-memory table
    uint32_t gword;         // first u64 word of this accumulator in a group's global record
};

struct McParams {
    uint32_t nrows, G, R, P, C, rpc;            // rpc: rows per scatter CTA (multiple of MC_SUB)
    unsigned long long magic;                   // ceil(2^48 / R): partition = (slot * magic) >> 48
    int nkeys, nacc, nv, row_field, cnt_acc;    // nv: value fields of a tuple; row_field: the generated row id or -1
    uint32_t klo[MC_MAX_KEYS], krange[MC_MAX_KEYS], kmul[MC_MAX_KEYS];
    const uint32_t* keys[MC_MAX_KEYS];
    const uint32_t* vals[MC_MAX_V];             // nullptr for the row field
    McAcc acc[MAX_ACC];
    uint32_t* tuples;                           // tw 32-bit words per tuple: word 0 = slot (0xffff: padding), then the value fields
    int tw, lb;                                 // words per tuple (2 or 4), tuples per 128-byte line (16 or 8)
    uint32_t* starts;                           // [P * C + 1]: line-padded counts per (partition, scatter CTA), then their exclusive scan
    unsigned long long* gtab;
    uint32_t gw;
    uint32_t* gfirst;
    uint32_t* flags;
    uint32_t* ctl;                              // [0] groups, [1] ticket, [2] groups whose first row is known
    int stages;
    // row shards only: row-valued aggregates (FIRST / LAST / SECOND) carry their 32-bit value with the row key in
    // extra record words, so that the owner of a group never has to read another rank's rows
    int nx;
    struct Extra { int agg, mode, acc; uint32_t word; const uint32_t* v; } x[MAX_AGGS];   // mode 0: value at the largest row (MAXROW cell),
                                                                                          // 1: at the first row, 2: top-2 rows (two words)
    int xof[MAX_AGGS];                          // extra of emit aggregate a, or -1
    unsigned long long row_base;
};

__device__ __forceinline__ uint2 lds64(const void* p) {
    uint2 r;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(smem_u32(p)));
    return r;
}
__device__ __forceinline__ void cons_bar() { asm volatile("bar.sync 1, %0;" :: "n"(MC_CONS) : "memory"); }

// ---- phase 0: rows per (partition, scatter CTA) -----------------------------------------------------------
template <int NK>
__global__ void __launch_bounds__(512)
k_mc_hist(const McParams M) {
    extern __shared__ uint32_t mh_hist[];
    for (uint32_t b = threadIdx.x; b < M.P; b += blockDim.x) mh_hist[b] = 0u;
    __syncthreads();
    const uint32_t c = blockIdx.x / MC_HSPLIT, h = blockIdx.x % MC_HSPLIT;
    const unsigned long long lo64 = (unsigned long long)c * M.rpc;
    bool bad = false;
    if (lo64 < M.nrows) {
        const uint32_t lo = (uint32_t)lo64;
        const uint32_t hi = (uint32_t)min((unsigned long long)M.nrows, lo64 + M.rpc);
        const uint32_t nquad = (hi - lo + 3u) >> 2;
        for (uint32_t q = h * blockDim.x + threadIdx.x; q < nquad; q += blockDim.x * MC_HSPLIT) {
            const uint32_t row0 = lo + (q << 2);
            const int cnt = (int)min(4u, hi - row0);
            uint32_t slot[4] = {0u, 0u, 0u, 0u};
            bool ok[4] = {true, true, true, true};
#pragma unroll
            for (int k = 0; k < NK; ++k) {
                uint32_t kv[4];
                load4_b32(M.keys[k], row0, cnt, true, kv);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t d = kv[r] - M.klo[k];
                    ok[r] &= d <= M.krange[k];
                    slot[r] += d * M.kmul[k];
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (r >= cnt) continue;
                if (!ok[r]) { bad = true; continue; }
                const uint32_t p = (uint32_t)(((unsigned long long)slot[r] * M.magic) >> 48);
                atomicAdd(&mh_hist[p], 1u);
            }
        }
    }
    if (bad) atomicExch(M.flags + FLAG_ABORT, 3u);          // a key outside the sampled / supplied bounds
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < M.P; b += blockDim.x) {
        const uint32_t v = mh_hist[b];
        if (v) atomicAdd(M.starts + size_t(b) * M.C + c, v);
    }
}

// every (partition, scatter CTA) range holds whole 128-byte lines
__global__ void k_mc_round(uint32_t* __restrict__ cnt, uint32_t n, uint32_t lb) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        cnt[i] = (cnt[i] + lb - 1u) / lb * lb;
}

// ---- phase 1: scatter ----------------------------------------------------------------------------------------
struct McScatterLayout { uint32_t tab_off, ring_off, stage_bytes, st_off, total; };
__host__ __device__ inline McScatterLayout mc_scatter_layout(uint32_t P, int ncol, int stages) {
    McScatterLayout L;
    const uint32_t Pp = (P + 31u) & ~31u;
    L.tab_off = 128;
    L.ring_off = (L.tab_off + (3u * Pp + 32u) * 4u + 127u) & ~127u;
    L.stage_bytes = (uint32_t)ncol * MC_SUB * 4u;
    L.st_off = L.ring_off + (uint32_t)stages * L.stage_bytes;
    L.total = L.st_off + Pp * 128u;
    return L;
}
// barrier of the consumer warps with an OR-reduction of `pred`
__device__ __forceinline__ bool cons_bar_or(bool pred) {
    uint32_t r;
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %1, 0;\n\tbar.red.or.pred p, 1, %2, q;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(r) : "r"((uint32_t)pred), "n"(MC_CONS) : "memory");
    return r != 0u;
}

// NK key columns, NV value columns read from memory, HASROW: one more tuple field holding the row id
template <int NK, int NV, bool HASROW>
__global__ void __launch_bounds__(MC_THREADS, 1)
k_mc_scatter(const McParams M) {
    constexpr int NCOL = NK + NV;
    constexpr int NT = NV + (HASROW ? 1 : 0);
    constexpr int TW = NT <= 1 ? 2 : 4;
    constexpr uint32_t LB = 32 / TW;
    constexpr uint32_t DONE = 0xffffffffu;
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t* const full = reinterpret_cast<uint64_t*>(sm);
    uint64_t* const empty = full + MC_MAX_STAGES;
    const McScatterLayout L = mc_scatter_layout(M.P, NCOL, M.stages);
    const uint32_t Pp = (M.P + 31u) & ~31u;
    uint32_t* const fill = reinterpret_cast<uint32_t*>(sm + L.tab_off);     // tuples appended to the partition's current line (may exceed LB inside a tile)
    uint32_t* const wpos = fill + Pp;                                      // next line of this CTA's range of the partition (tuple index)
    uint32_t* const wl = wpos + Pp;                                        // partitions whose line is complete
    uint32_t* const wl_n = wl + Pp;                                        // [2]
    unsigned char* const ring = sm + L.ring_off;
    uint32_t* const stage = reinterpret_cast<uint32_t*>(sm + L.st_off);     // [P][32 words]

    const uint32_t c = blockIdx.x;
    if (*reinterpret_cast<volatile uint32_t*>(M.flags + FLAG_ABORT)) return;
    for (uint32_t b = threadIdx.x; b < Pp; b += MC_THREADS) {
        fill[b] = 0u;
        wpos[b] = b < M.P ? M.starts[size_t(b) * M.C + c] : 0u;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < M.stages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, MC_CONS / 32); }
        mbar_fence_init();
        wl_n[0] = wl_n[1] = 0u;
    }
    __syncthreads();
    const unsigned long long lo64 = (unsigned long long)c * M.rpc;
    if (lo64 >= M.nrows) return;
    const uint32_t lo = (uint32_t)lo64;
    const uint32_t hi = (uint32_t)min((unsigned long long)M.nrows, lo64 + M.rpc);
    const uint32_t nsub = (hi - lo + MC_SUB - 1) / MC_SUB;
    const uint32_t lane = threadIdx.x & 31u;

    if (threadIdx.x < 32) {
        // ---- producer warp: full stages by bulk copy, the ragged last stage by plain loads
        int s = 0; uint32_t ph = 0;
        for (uint32_t k = 0; k < nsub; ++k) {
            if (lane == 0) mbar_wait(empty + s, ph ^ 1u);
            __syncwarp();
            const uint32_t row0 = lo + k * MC_SUB;
            const uint32_t rows = min((uint32_t)MC_SUB, hi - row0);
            unsigned char* dst = ring + size_t(s) * L.stage_bytes;
            if (rows == MC_SUB) {
                if (lane == 0) {
                    mbar_arrive_expect_tx(full + s, L.stage_bytes);
#pragma unroll
                    for (int col = 0; col < NCOL; ++col) {
                        const uint32_t* src = (col < NK ? M.keys[col] : M.vals[col - NK]) + row0;
                        bulk_g2s(dst + size_t(col) * MC_SUB * 4, src, MC_SUB * 4, full + s);
                    }
                }
            } else {
#pragma unroll
                for (int col = 0; col < NCOL; ++col) {
                    const uint32_t* src = (col < NK ? M.keys[col] : M.vals[col - NK]) + row0;
                    uint32_t* d = reinterpret_cast<uint32_t*>(dst + size_t(col) * MC_SUB * 4);
                    for (uint32_t i = lane; i < rows; i += 32) d[i] = __ldg(src + i);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(full + s);
            }
            if (++s == M.stages) { s = 0; ph ^= 1u; }
        }
        return;
    }

    // ---- consumers: one ring stage = one tile, 2 rows per thread
    const uint32_t t = threadIdx.x - 32u;
    const uint32_t cw = t >> 5;                                 // consumer warp 0..30
    constexpr uint32_t NCW = MC_CONS / 32;
    const uint32_t sub = lane >> 3, l8 = lane & 7u;             // a quarter warp copies one line
    // copies the complete line of partition p to its place (quarter warp `sub`); act: this quarter has work
    auto flush_line = [&](uint32_t p, bool act) {
        uint32_t dst = 0;
        if (act) {
            dst = wpos[p];
            const uint4 v = *reinterpret_cast<const uint4*>(stage + size_t(p) * 32u + l8 * 4u);
            *reinterpret_cast<uint4*>(M.tuples + size_t(dst) * TW + l8 * 4u) = v;
        }
        __syncwarp();
        if (act && l8 == 0) { wpos[p] = dst + LB; fill[p] -= LB; }
    };
    int s = 0; uint32_t ph = 0, par = 0;
    for (uint32_t k = 0; k < nsub; ++k) {
        mbar_wait(full + s, ph);
        const unsigned char* stg = ring + size_t(s) * L.stage_bytes + t * 8u;
        uint2 kk[NK], vv[NV > 0 ? NV : 1];
#pragma unroll
        for (int q = 0; q < NK; ++q) kk[q] = lds64(stg + size_t(q) * MC_SUB * 4);
#pragma unroll
        for (int q = 0; q < NV; ++q) vv[q] = lds64(stg + size_t(NK + q) * MC_SUB * 4);
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + s);
        if (++s == M.stages) { s = 0; ph ^= 1u; }
        const uint32_t row0 = lo + k * MC_SUB + t * 2u;
        uint32_t part[2], pos[2], w[2][TW];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            uint32_t slot = 0; bool ok = row0 + r < hi;
#pragma unroll
            for (int q = 0; q < NK; ++q) {
                const uint32_t d = (r ? kk[q].y : kk[q].x) - M.klo[q];
                ok &= d <= M.krange[q];
                slot += d * M.kmul[q];
            }
#pragma unroll
            for (int x = 0; x < TW; ++x) w[r][x] = 0u;
#pragma unroll
            for (int q = 0; q < NV; ++q) w[r][1 + q] = r ? vv[q].y : vv[q].x;
            if (HASROW) w[r][NT] = row0 + r;
            part[r] = 0u; pos[r] = DONE;
            if (ok) {
                const uint32_t p = (uint32_t)(((unsigned long long)slot * M.magic) >> 48);
                w[r][0] = slot - p * M.R;
                part[r] = p;
                pos[r] = atomicAdd(&fill[p], 1u);
            }
        }
        // rounds: a tuple whose line is full waits for the flush and tries again (its position is already reserved)
        for (;;) {
            bool deferred = false;
            uint32_t* const my_n = wl_n + par;
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                if (pos[r] == DONE) continue;
                if (pos[r] >= LB) { pos[r] -= LB; deferred = true; continue; }
                uint32_t* d = stage + size_t(part[r]) * 32u + pos[r] * TW;
                if (TW == 4) *reinterpret_cast<uint4*>(d) = make_uint4(w[r][0], w[r][1], w[r][2 % TW], w[r][3 % TW]);
                else *reinterpret_cast<uint2*>(d) = make_uint2(w[r][0], w[r][1]);
                if (pos[r] == LB - 1u) wl[atomicAdd(my_n, 1u)] = part[r];          // this tuple completes the line
                pos[r] = DONE;
            }
            const bool more = cons_bar_or(deferred);                          // staged tuples + work list visible
            const uint32_t n = *reinterpret_cast<volatile uint32_t*>(my_n);
            if (t == 0) wl_n[par ^ 1u] = 0u;
            for (uint32_t e0 = cw * 4u; e0 < n; e0 += NCW * 4u) {
                const uint32_t e = e0 + sub;
                const bool act = e < n;
                flush_line(act ? wl[e] : 0u, act);
            }
            cons_bar();
            par ^= 1u;
            if (!more) break;
        }
    }
    // ---- the partly filled lines: complete with padding tuples (slot 0xffff), write them out
    for (uint32_t p0 = cw * 4u; p0 < M.P; p0 += NCW * 4u) {
        const uint32_t p = p0 + sub;
        const uint32_t f = p < M.P ? fill[p] : 0u;
        const bool act = f > 0u;
        if (act) {
            uint32_t* line = stage + size_t(p) * 32u;
            for (uint32_t x = f * TW + l8; x < 32u; x += 8u) line[x] = 0xffffu;
        }
        __syncwarp();
        flush_line(p < M.P ? p : 0u, act);
    }
}

// ---- phase 2: one CTA per partition ------------------------------------------------------------------------------
struct McChecks { uint32_t vabs32; unsigned long long vabs; bool fx_bad; };

// 8 tuples of one accumulator.  s[j]: slot; m: valid mask; u / u2: the raw 32-bit value fields.
__device__ __forceinline__ void mc_accn(const McParams& M, const McAcc& A, unsigned char* sm, const uint32_t Rp, const uint32_t gbase,
                                        const uint32_t (&s)[MC_AGG_TPT], const uint32_t m, const uint32_t (&u)[MC_AGG_TPT],
                                        const uint32_t (&u2)[MC_AGG_TPT], const int fxQ, const int fxR, McChecks& ck) {
    uint32_t* const w0 = reinterpret_cast<uint32_t*>(sm + A.off);
    switch (A.kind) {
    case MCK_CNT:
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) if (m & (1u << j)) atomicAdd(w0 + s[j], 1u);
        break;
    case MCK_ISUM1:
        if (A.xform == X_IDENT) {
#pragma unroll
            for (int j = 0; j < MC_AGG_TPT; ++j) {
                if (!(m & (1u << j))) continue;
                const int x = (int)u[j];
                ck.vabs32 = max(ck.vabs32, (uint32_t)(x < 0 ? -x : x));
                atomicAdd(w0 + s[j], (uint32_t)x);
            }
        } else {
#pragma unroll
            for (int j = 0; j < MC_AGG_TPT; ++j) {
                if (!(m & (1u << j))) continue;
                const long long x = (long long)(int)u[j] * (long long)(int)u2[j];
                const unsigned long long ax = (unsigned long long)(x < 0 ? -x : x);
                if (ax > ck.vabs) ck.vabs = ax;
                atomicAdd(w0 + s[j], (uint32_t)x);
            }
        }
        break;
    case MCK_ISUM2:
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) {
            if (!(m & (1u << j))) continue;
            const long long x = A.xform == X_IDENT ? (long long)(int)u[j] : (long long)(int)u[j] * (long long)(int)u2[j];
            isum2_add(w0 + s[j], w0 + Rp + s[j], x);
        }
        break;
    case MCK_FX:
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) {
            if (!(m & (1u << j))) continue;
            const float f = __uint_as_float(u[j]);
            const float x = A.xform == X_IDENT ? f : __fmul_rn(f, __uint_as_float(u2[j]));     // product in T (aggregations.h:339-345)
            unsigned long long* g = M.gtab + size_t(gbase + s[j]) * M.gw + A.gword;
            if (!fx_add(w0 + s[j], w0 + Rp + s[j], g, g + 1, x, A.fxB, fxQ, fxR)) ck.fx_bad = true;
        }
        break;
    case MCK_MIN32:
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) if (m & (1u << j)) min32_upd(reinterpret_cast<int*>(w0) + s[j], ordered32(u[j], A.vtype));
        break;
    case MCK_MAX32:
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) if (m & (1u << j)) max32_upd(reinterpret_cast<int*>(w0) + s[j], ordered32(u[j], A.vtype));
        break;
    case MCK_MAXROW:
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) {
            if (!(m & (1u << j))) continue;
            if (u[j] > *reinterpret_cast<volatile uint32_t*>(w0 + s[j])) atomicMax(w0 + s[j], u[j]);
        }
        break;
    case MCK_TOP2:
        // best / second-best row key in two 32-bit cells: every key loses exactly one comparison against the
        // running best (atomicMax returns the old best), and the loser is offered to the second cell
#pragma unroll
        for (int j = 0; j < MC_AGG_TPT; ++j) {
            if (!(m & (1u << j))) continue;
            const uint32_t key = A.xform == AQ_ROWS_DESC ? u[j] + 1u : 0xffffffffu - u[j];
            const uint32_t old = atomicMax(w0 + s[j], key);
            const uint32_t loser = min(old, key);
            if (loser > *reinterpret_cast<volatile uint32_t*>(w0 + Rp + s[j])) atomicMax(w0 + Rp + s[j], loser);
        }
        break;
    default: break;
    }
}

template <int TW>
__global__ void __launch_bounds__(MC_AGG_THREADS, 1)
k_mc_agg(const McParams M) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ uint32_t s_groups;
    if (*reinterpret_cast<volatile uint32_t*>(M.flags + FLAG_ABORT)) return;
    const uint32_t p = blockIdx.x;
    const uint32_t base = M.starts[size_t(p) * M.C];
    const uint32_t n = M.starts[size_t(p + 1) * M.C] - base;       // tuples of the partition, padding included (whole lines)
    const uint32_t R = M.R, Rp = R + 8u;
    const uint32_t gbase = p * R;
    if (threadIdx.x == 0) s_groups = 0u;
    for (int a = 0; a < M.nacc; ++a) {
        const McAcc A = M.acc[a];
        uint32_t* w = reinterpret_cast<uint32_t*>(sm + A.off);
        const bool two = A.kind == MCK_ISUM2 || A.kind == MCK_FX || A.kind == MCK_TOP2;
        const uint32_t init = A.kind == MCK_MIN32 ? 0x7fffffffu : (A.kind == MCK_MAX32 ? 0x80000000u : 0u);
        for (uint32_t g = threadIdx.x; g < Rp; g += MC_AGG_THREADS) { w[g] = init; if (two) w[Rp + g] = 0u; }
    }
    __syncthreads();
    int lg = 0;
    while ((1ull << lg) < (unsigned long long)n + 1ull) ++lg;
    const int fxR = max(0, min(39 - lg, FX_S_TOP - 1));            // |a cell's sum| < n * 2^(24 + R) <= 2^63
    const int fxQ = FX_S_TOP - fxR;
    McChecks ck; ck.vabs32 = 0u; ck.vabs = 0ull; ck.fx_bad = false;

    // 16-byte vectors: one tuple (TW = 4) or two (TW = 2); consecutive threads read consecutive vectors
    constexpr int U = MC_AGG_TPT;
    constexpr uint32_t BATCH = MC_AGG_THREADS * U;
    const uint4* const src = reinterpret_cast<const uint4*>(M.tuples + size_t(base) * TW);
    const uint32_t nvec = n / (4 / TW);
    uint4 nx[U];
    auto load = [&](uint32_t b0) {
#pragma unroll
        for (int j = 0; j < U; ++j) {
            const uint32_t i = b0 + j * MC_AGG_THREADS + threadIdx.x;
            if (i < nvec) { const int4 q = ld_stream_v4(src + i); nx[j] = make_uint4((uint32_t)q.x, (uint32_t)q.y, (uint32_t)q.z, (uint32_t)q.w); }
            else nx[j] = make_uint4(0xffffu, 0u, 0xffffu, 0u);
        }
    };
    load(0);
    for (uint32_t b0 = 0; b0 < nvec; b0 += BATCH) {
        uint4 cv[U];
#pragma unroll
        for (int j = 0; j < U; ++j) cv[j] = nx[j];
        if (b0 + BATCH < nvec) load(b0 + BATCH);
#pragma unroll
        for (int half = 0; half < 4 / TW; ++half) {
            uint32_t s[U], m = 0u;
#pragma unroll
            for (int j = 0; j < U; ++j) {
                s[j] = (half ? cv[j].z : cv[j].x) & 0xffffu;
                if (s[j] < R) m |= 1u << j; else s[j] = R;         // padding (0xffff): masked, trash entry
            }
            if (TW == 4 && b0 + (U - 1) * MC_AGG_THREADS + threadIdx.x >= nvec) {
#pragma unroll
                for (int j = 0; j < U; ++j) if (b0 + j * MC_AGG_THREADS + threadIdx.x >= nvec) m &= ~(1u << j);
            }
            for (int a = 0; a < M.nacc; ++a) {
                const McAcc A = M.acc[a];
                uint32_t u[U], u2[U];
                // value fields are picked out of the tuple only where the accumulator reads them (counts read none)
                const bool need1 = A.kind != MCK_CNT;
                const bool need2 = need1 && A.xform != X_IDENT && A.kind != MCK_TOP2 && A.kind != MCK_MAXROW;
#pragma unroll
                for (int j = 0; j < U; ++j) {
                    u[j] = u2[j] = 0u;
                    if (TW == 4) {
                        if (need1) u[j] = A.c0 == 0 ? cv[j].y : (A.c0 == 1 ? cv[j].z : cv[j].w);
                        if (need2) u2[j] = A.c1 == 0 ? cv[j].y : (A.c1 == 1 ? cv[j].z : cv[j].w);
                    } else if (need1) {
                        u[j] = u2[j] = half ? cv[j].w : cv[j].y;
                    }
                }
                mc_accn(M, A, sm, Rp, gbase, s, m, u, u2, fxQ, fxR, ck);
            }
        }
    }
    {
        const unsigned long long lim = 0x7fffffffULL / ((unsigned long long)n + 1ull);
        if (ck.fx_bad) atomicCAS(M.flags + FLAG_ABORT, 0u, 4u);
        else if (ck.vabs > lim || (unsigned long long)ck.vabs32 > lim) atomicCAS(M.flags + FLAG_ABORT, 0u, 2u);
    }
    __syncthreads();
    // ---- the partition's groups -> global records (this CTA is their only writer)
    const uint32_t* cntw = reinterpret_cast<const uint32_t*>(sm + M.acc[M.cnt_acc].off);
    uint32_t mine = 0;
    for (uint32_t sl = threadIdx.x; sl < R; sl += MC_AGG_THREADS) {
        const uint32_t g = gbase + sl;
        if (g >= M.G) break;
        if (cntw[sl] == 0u) continue;
        ++mine;
        unsigned long long* rec = M.gtab + size_t(g) * M.gw;
        for (int a = 0; a < M.nacc; ++a) {
            const McAcc A = M.acc[a];
            const uint32_t* w = reinterpret_cast<const uint32_t*>(sm + A.off);
            switch (A.kind) {
            case MCK_CNT: rec[A.gword] = (unsigned long long)w[sl]; break;
            case MCK_ISUM1: rec[A.gword] = (unsigned long long)(long long)(int)w[sl]; break;
            case MCK_ISUM2: rec[A.gword] = (unsigned long long)w[sl] | ((unsigned long long)w[Rp + sl] << 32); break;
            case MCK_FX: fx_merge(w[sl], w[Rp + sl], fxQ, rec + A.gword, rec + A.gword + 1); break;
            case MCK_MIN32: case MCK_MAX32: rec[A.gword] = (unsigned long long)image64_of32((int)w[sl], A.vtype); break;
            case MCK_MAXROW: rec[A.gword] = (unsigned long long)w[sl]; break;
            case MCK_TOP2: rec[A.gword] = ((unsigned long long)w[sl] << 32) | (unsigned long long)w[Rp + sl]; break;
            default: break;
            }
        }
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_groups, mine);
    __syncthreads();
    if (threadIdx.x == 0 && s_groups) atomicAdd(M.ctl + 0, s_groups);
}

// ---- first rows -------------------------------------------------------------------------------------------------
constexpr int MCF_THREADS = 256;
constexpr int MCF_Q = 4;                                  // quads of rows per thread and tile: the ticket / counter round trips are per tile
constexpr int MCF_ROWS = MCF_THREADS * 4 * MCF_Q;
template <int NK>
__global__ void __launch_bounds__(MCF_THREADS)
k_mc_first(const McParams M) {
    __shared__ uint32_t s_tile;
    if (*reinterpret_cast<volatile uint32_t*>(M.flags + FLAG_ABORT)) return;
    const uint32_t nquad = (M.nrows + 3u) >> 2;
    const uint32_t tile_quads = MCF_THREADS * MCF_Q;
    const uint32_t ntiles = (nquad + tile_quads - 1) / tile_quads;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint32_t found = *reinterpret_cast<volatile uint32_t*>(M.ctl + 2), need = *reinterpret_cast<volatile uint32_t*>(M.ctl + 0);
            s_tile = found >= need ? 0xffffffffu : atomicAdd(M.ctl + 1, 1u);
        }
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= ntiles) return;
        uint32_t slot[MCF_Q][4];
        uint32_t okm = 0;                                  // bit 4 * i + r: row r of quad i is a row with in-range keys
#pragma unroll
        for (int i = 0; i < MCF_Q; ++i) {
            const uint32_t q = tile * tile_quads + i * MCF_THREADS + threadIdx.x;
            const uint32_t row0 = q << 2;
            const int cnt = q < nquad ? (int)min(4u, M.nrows - row0) : 0;      // (no early exit: the warp reduction below is full-mask)
            bool ok[4] = {cnt > 0, cnt > 1, cnt > 2, cnt > 3};
#pragma unroll
            for (int r = 0; r < 4; ++r) slot[i][r] = 0u;
#pragma unroll
            for (int k = 0; k < NK; ++k) {
                uint32_t kv[4] = {0u, 0u, 0u, 0u};
                if (cnt) load4_b32(M.keys[k], row0, cnt, true, kv);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t d = kv[r] - M.klo[k];
                    ok[r] &= d <= M.krange[k];
                    slot[i][r] += d * M.kmul[k];
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) if (ok[r]) okm |= 1u << (4 * i + r);
        }
        uint32_t cur[MCF_Q][4];
#pragma unroll
        for (int i = 0; i < MCF_Q; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r)
                cur[i][r] = (okm >> (4 * i + r)) & 1u ? *reinterpret_cast<volatile uint32_t*>(M.gfirst + slot[i][r]) : 0u;
        uint32_t fresh = 0;
#pragma unroll
        for (int i = 0; i < MCF_Q; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const uint32_t row = ((tile * tile_quads + i * MCF_THREADS + threadIdx.x) << 2) + r;
                if (((okm >> (4 * i + r)) & 1u) && row < cur[i][r]) {
                    const uint32_t old = atomicMin(M.gfirst + slot[i][r], row);
                    if (old == 0xffffffffu) ++fresh;
                }
            }
        fresh = __reduce_add_sync(0xffffffffu, fresh);
        if ((threadIdx.x & 31) == 0 && fresh) atomicAdd(M.ctl + 2, fresh);
    }
}

__global__ void k_mc_mark(const uint32_t* __restrict__ gfirst, uint32_t G, uint32_t* __restrict__ bitmap) {
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < G; g += gridDim.x * blockDim.x) {
        const uint32_t fr = gfirst[g];
        if (fr != 0xffffffffu) atomicOr(bitmap + (fr >> 5), 1u << (fr & 31));
    }
}

// Emit in ROW order (like k_emit_rows of groupby.cu): a warp walks 32 bitmap words; a set bit = the first row of
// a group, gid = rank of the bit.  The group's record is found from the row's own key (direct index).
__global__ void __launch_bounds__(256)
k_mc_emit(const McParams M, const uint32_t* __restrict__ bitmap, const uint32_t* __restrict__ wordprefix, const uint32_t nwords,
          const KeyOut ko, uint32_t* __restrict__ first_row_out, const EmitProg ep, const uint32_t max_groups) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nchunks = (nwords + 31u) >> 5;
    for (uint32_t chunk = gwarp; chunk < nchunks; chunk += nwarps) {
        const uint32_t wi = chunk * 32u + lane;
        const uint32_t word = wi < nwords ? bitmap[wi] : 0u;
        const uint32_t pfx = wi < nwords ? wordprefix[wi] : 0u;
        uint32_t mm = __ballot_sync(0xffffffffu, word != 0u);
        while (mm) {
            const int j = __ffs(mm) - 1;
            mm &= mm - 1u;
            const uint32_t w = __shfl_sync(0xffffffffu, word, j);
            const uint32_t pf = __shfl_sync(0xffffffffu, pfx, j);
            if (!((w >> lane) & 1u)) continue;
            const uint32_t fr = (chunk * 32u + j) * 32u + lane;
            const uint32_t gid = pf + __popc(w & ((1u << lane) - 1u));
            if (gid >= max_groups) continue;
            uint32_t g = 0;
            for (int k = 0; k < M.nkeys; ++k) {
                const uint32_t kv = __ldg(M.keys[k] + fr);
                g += (kv - M.klo[k]) * M.kmul[k];
                if (ko.dst[k]) static_cast<uint32_t*>(ko.dst[k])[gid] = kv;
            }
            if (first_row_out) first_row_out[gid] = fr;
            const unsigned long long* rec = M.gtab + size_t(g) * M.gw;
            unsigned long long cell[1 + MAX_ACC];
            for (int a = 0; a < M.nacc; ++a) {
                const McAcc& A = M.acc[a];
                unsigned long long v = rec[A.gword];
                if (A.kind == MCK_FX) v = (unsigned long long)__double_as_longlong(fx_to_double(v, rec[A.gword + 1], A.fxB));
                cell[1 + a] = v;
            }
            for (int a = 0; a < ep.n; ++a) emit_one(ep.agg[a], cell, gid, fr, ep.row_order);
        }
    }
}

// ---- row shards (dist.cu): the groups of the slot range this rank owns, in ascending GLOBAL first-row order ------
__global__ void k_mc_count_range(const unsigned long long* __restrict__ gtab, uint32_t gw, uint32_t cnt_word, uint32_t lo, uint32_t n,
                                 uint32_t* __restrict__ ctl) {
    uint32_t mine = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (gtab[size_t(lo + i) * gw + cnt_word] != 0ull) ++mine;
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(ctl + 0, mine);
}
// order[gid] = slot, gid = rank of the group's first row among the first rows marked in the bitmap
__global__ void k_mc_order(const uint32_t* __restrict__ gfirst, uint32_t lo, uint32_t n, const uint32_t* __restrict__ bitmap,
                           const uint32_t* __restrict__ wordprefix, uint32_t* __restrict__ order) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t fr = gfirst[lo + i];
        if (fr == 0xffffffffu) continue;
        const uint32_t w = fr >> 5, b = fr & 31u;
        order[wordprefix[w] + __popc(bitmap[w] & ((1u << b) - 1u))] = lo + i;
    }
}
// local row-valued cells -> (global row key << 32 | value) words, before the record tables are exchanged
__global__ void __launch_bounds__(256)
k_mc_localize(const McParams M) {
    const uint32_t base = (uint32_t)M.row_base;
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < M.G; g += gridDim.x * blockDim.x) {
        unsigned long long* rec = M.gtab + size_t(g) * M.gw;
        if (rec[M.acc[M.cnt_acc].gword] == 0ull) continue;
        for (int i = 0; i < M.nx; ++i) {
            const McParams::Extra& X = M.x[i];
            if (X.mode == 0) {
                const uint32_t r = (uint32_t)rec[M.acc[X.acc].gword];
                rec[X.word] = ((unsigned long long)(r + base + 1u) << 32) | X.v[r];
            } else if (X.mode == 1) {
                const uint32_t r = M.gfirst[g];                       // still the LOCAL first row here
                rec[X.word] = ((unsigned long long)(r + base) << 32) | X.v[r];
            } else {
                const unsigned long long c = rec[M.acc[X.acc].gword];
                const uint32_t k0 = (uint32_t)(c >> 32), k1 = (uint32_t)c;
                const bool desc = M.acc[X.acc].xform == AQ_ROWS_DESC;
                auto one = [&](uint32_t k) -> unsigned long long {
                    if (k == 0u) return 0ull;
                    const uint32_t r = desc ? k - 1u : 0xffffffffu - k;
                    const uint32_t gk = desc ? k + base : k - base;
                    return ((unsigned long long)gk << 32) | X.v[r];
                };
                rec[X.word] = one(k0);
                rec[X.word + 1] = one(k1);
            }
        }
    }
}
__global__ void __launch_bounds__(256)
k_mc_emit_ord(const McParams M, const uint32_t* __restrict__ order, const uint32_t ng, const KeyOut ko, uint32_t* __restrict__ first_row_out,
              const EmitProg ep, const uint32_t max_groups) {
    for (uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x; gid < ng && gid < max_groups; gid += gridDim.x * blockDim.x) {
        const uint32_t g = order[gid];
        const uint32_t fr = M.gfirst[g];
        for (int k = 0; k < M.nkeys; ++k) {
            if (!ko.dst[k]) continue;
            const uint32_t digit = (g / M.kmul[k]) % (M.krange[k] + 1u);
            static_cast<uint32_t*>(ko.dst[k])[gid] = digit + M.klo[k];
        }
        if (first_row_out) first_row_out[gid] = fr;
        const unsigned long long* rec = M.gtab + size_t(g) * M.gw;
        unsigned long long cell[1 + MAX_ACC];
        for (int a = 0; a < M.nacc; ++a) {
            const McAcc& A = M.acc[a];
            unsigned long long v = rec[A.gword];
            if (A.kind == MCK_FX) v = (unsigned long long)__double_as_longlong(fx_to_double(v, rec[A.gword + 1], A.fxB));
            cell[1 + a] = v;
        }
        for (int a = 0; a < ep.n; ++a) {
            const int xi = M.xof[a];
            if (xi < 0) { emit_one(ep.agg[a], cell, gid, fr, ep.row_order); continue; }
            const McParams::Extra& X = M.x[xi];                      // the value came with the row key
            const unsigned long long w = rec[X.mode == 2 ? X.word + 1 : X.word];
            static_cast<uint32_t*>(ep.agg[a].out)[gid] = (X.mode == 2 && (w >> 32) == 0ull) ? 0u : (uint32_t)w;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
static uint32_t mc_cell_bytes(int kind) { return (kind == MCK_ISUM2 || kind == MCK_FX || kind == MCK_TOP2) ? 8u : 4u; }

template <int NK, int NV, bool HASROW>
static int launch_scatter_t(const McParams& M, size_t smem) {
    static size_t attr = 0;
    if (smem > attr) { AQ_CUDA(cudaFuncSetAttribute(k_mc_scatter<NK, NV, HASROW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = smem; }
    AQ_LAUNCH((k_mc_scatter<NK, NV, HASROW>), M.C, MC_THREADS, smem, M);
    return 0;
}
template <int NK>
static int launch_scatter_k(const McParams& M, size_t smem) {
    const bool hr = M.row_field >= 0;
    const int nvm = M.nv - (hr ? 1 : 0);
    switch (nvm * 2 + (hr ? 1 : 0)) {
    case 0: return launch_scatter_t<NK, 0, false>(M, smem);
    case 1: return launch_scatter_t<NK, 0, true>(M, smem);
    case 2: return launch_scatter_t<NK, 1, false>(M, smem);
    case 3: return launch_scatter_t<NK, 1, true>(M, smem);
    case 4: return launch_scatter_t<NK, 2, false>(M, smem);
    case 5: return launch_scatter_t<NK, 2, true>(M, smem);
    case 6: return launch_scatter_t<NK, 3, false>(M, smem);
    default: return fail("internal: %d tuple fields in k_mc_scatter", M.nv);
    }
}
static int launch_scatter(const McParams& M, size_t smem) {
    switch (M.nkeys) {
    case 1: return launch_scatter_k<1>(M, smem);
    case 2: return launch_scatter_k<2>(M, smem);
    default: return fail("internal: %d key columns in k_mc_scatter", M.nkeys);
    }
}
template <int TW>
static int launch_agg_t(const McParams& M, size_t smem) {
    static size_t attr = 0;
    if (smem > attr) { AQ_CUDA(cudaFuncSetAttribute(k_mc_agg<TW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = smem; }
    AQ_LAUNCH((k_mc_agg<TW>), M.P, MC_AGG_THREADS, smem, M);
    return 0;
}
static int launch_agg(const McParams& M, size_t smem) {
    return M.tw == 2 ? launch_agg_t<2>(M, smem) : launch_agg_t<4>(M, smem);
}

static uint32_t mc_min_rows() {                // AQGPU_MC_MIN_ROWS: tests drive the path at small sizes (read per call)
    const char* e = getenv("AQGPU_MC_MIN_ROWS");
    long v = e ? atol(e) : (1L << 22);
    return (uint32_t)(v < 4096 ? 4096 : v);
}

// Decides whether the range-partitioned path takes the query; fills everything but the buffers.
bool mc_layout(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, const PlanStats& st, McParams& M, size_t* smem_scatter, size_t* smem_agg) {
    Context& c = ctx();
    memset(&M, 0, sizeof(M));
    const uint32_t nrows_local = nrows;
    const bool sharded = dist().active;
    if (sharded) nrows = dist().rows_max;                                  // row shards: the same layout on every rank
    if (!st.have_bounds || kc.n > 2 || nrows < mc_min_rows() || st.est_groups <= 0.0) return false;
    if (st.est_groups * 4.0 > (double)nrows) return false;                  // fewer than ~4 rows per group: not this path
    unsigned long long G = 1;
    for (int k = 0; k < kc.n; ++k) {
        if (kc.type[k] != AQ_INT32 && kc.type[k] != AQ_UINT32) return false;
        if (reinterpret_cast<uintptr_t>(kc.col[k].d) & 15) return false;
        long long lo = st.kmin[k], hi = st.kmax[k];
        if (!st.bounds_given) {                                            // a sample rarely sees the extreme keys
            const long long margin = (hi - lo) / 64 + 16;
            const long long tlo = kc.type[k] == AQ_INT32 ? -2147483648LL : 0LL, thi = kc.type[k] == AQ_INT32 ? 2147483647LL : 4294967295LL;
            lo = std::max(tlo, lo - margin); hi = std::min(thi, hi + margin);
        }
        const long long range = hi - lo;
        if (range < 0 || range > 0xffffffffLL) return false;
        M.klo[k] = (uint32_t)lo; M.krange[k] = (uint32_t)range; M.kmul[k] = (uint32_t)G;
        G *= (unsigned long long)range + 1ull;
        if (G > MC_GMAX) return false;
        M.keys[k] = static_cast<const uint32_t*>(kc.col[k].d);
    }
    if ((double)G > 16.0 * st.est_groups + 65536.0) return false;           // sparse key domain: hash paths
    M.nkeys = kc.n; M.G = (uint32_t)G; M.nrows = nrows_local; M.nacc = plan.n_acc; M.flags = c.d_flags; M.ctl = c.d_flags + 16;
    M.C = (uint32_t)c.sm_count;
    // ---- tuple fields and accumulator cells
    M.nv = 0; M.row_field = -1; M.cnt_acc = -1;
    bool ok = true;
    auto field_of = [&](const void* p) -> int {
        for (int i = 0; i < M.nv; ++i) if (M.vals[i] == p && i != M.row_field) return i;
        if (M.nv >= MC_MAX_V) { ok = false; return 0; }
        M.vals[M.nv] = static_cast<const uint32_t*>(p);
        return M.nv++;
    };
    uint32_t bpg = 0, gw = 0;
    for (int a = 0; a < plan.n_acc; ++a) {
        const AccOp& o = plan.prog.op[a];
        McAcc& A = M.acc[a];
        A.xform = o.xform; A.vtype = o.vtype; A.fxB = 0; A.c0 = A.c1 = -1; A.gword = gw;
        const bool v32 = o.vtype == AQ_INT32 || o.vtype == AQ_FLOAT;
        switch (o.kind) {
        case ACC_COUNT: A.kind = MCK_CNT; if (M.cnt_acc < 0) M.cnt_acc = a; break;
        case ACC_ADD_I64: if (o.vtype != AQ_INT32) return false; A.kind = MCK_ISUM1; break;     // ISUM2 decided below
        case ACC_ADD_F64: if (o.vtype != AQ_FLOAT || st.fx_maxexp[a] < 0) return false; A.kind = MCK_FX; A.fxB = fx_base(st.fx_maxexp[a]); break;
        case ACC_MIN_I64: if (!v32) return false; A.kind = MCK_MIN32; break;
        case ACC_MAX_I64: if (!v32) return false; A.kind = MCK_MAX32; break;
        case ACC_MAXROW: A.kind = MCK_MAXROW; break;
        case ACC_TOP2ROW: A.kind = MCK_TOP2; break;
        default: return false;
        }
        if (A.kind == MCK_MAXROW || A.kind == MCK_TOP2) {
            if (M.row_field < 0) { if (M.nv >= MC_MAX_V) return false; M.row_field = M.nv; M.vals[M.nv++] = nullptr; }
            A.c0 = M.row_field;
        } else if (A.kind != MCK_CNT) {
            if ((reinterpret_cast<uintptr_t>(o.v) & 15) || (o.v2 && (reinterpret_cast<uintptr_t>(o.v2) & 15))) return false;
            A.c0 = field_of(o.v);
            if (o.xform == X_PRODUCT) A.c1 = field_of(o.v2);
            else if (o.xform == X_SQUARE) A.c1 = A.c0;
        }
        gw += A.kind == MCK_FX ? 2u : 1u;
    }
    if (!ok || M.cnt_acc < 0) return false;
    // the row field must be the last one (k_mc_scatter generates it after the loaded fields)
    if (M.row_field >= 0 && M.row_field != M.nv - 1) {
        const int rf = M.row_field, last = M.nv - 1;
        std::swap(M.vals[rf], M.vals[last]);
        for (int a = 0; a < plan.n_acc; ++a) {
            McAcc& A = M.acc[a];
            auto fix = [&](int& f) { if (f == rf) f = last; else if (f == last) f = rf; };
            fix(A.c0); fix(A.c1);
        }
        M.row_field = last;
    }
    const int nvm = M.nv - (M.row_field >= 0 ? 1 : 0);
    if (M.nv > 3) return false;                                           // a tuple is a slot word + at most 3 value words
    M.tw = M.nv <= 1 ? 2 : 4; M.lb = 32 / M.tw;
    M.nx = 0; M.row_base = sharded ? dist().row_base : 0ull;
    for (int a = 0; a < MAX_AGGS; ++a) M.xof[a] = -1;
    if (sharded) {
        for (int a = 0; a < plan.emit.n; ++a) {
            const EmitAgg& e = plan.emit.agg[a];
            if (e.op != AQ_FIRST && e.op != AQ_LAST && e.op != AQ_SECOND) continue;
            if (e.vtype != AQ_INT32 && e.vtype != AQ_FLOAT) return false;
            McParams::Extra& X = M.x[M.nx];
            X.agg = a; X.v = static_cast<const uint32_t*>(e.v); X.word = gw;
            if (e.op == AQ_SECOND) { X.mode = 2; X.acc = e.top2; gw += 2; }
            else if ((e.op == AQ_FIRST) == (plan.emit.row_order == AQ_ROWS_DESC)) { X.mode = 0; X.acc = e.maxrow; gw += 1; }
            else { X.mode = 1; X.acc = -1; gw += 1; }
            if (X.mode != 1 && X.acc < 0) return false;
            M.xof[a] = M.nx++;
        }
    }
    M.gw = gw;
    // ---- table of k_mc_agg: R groups per partition
    const size_t budget = c.smem_optin - 1024;
    for (int pass = 0; pass < 2; ++pass) {
        bpg = 0;
        for (int a = 0; a < plan.n_acc; ++a) bpg += mc_cell_bytes(M.acc[a].kind);
        uint32_t R = (uint32_t)(budget / bpg);
        R = R > 8 ? ((R - 8u) & ~7u) : 0u;
        if (R > 65528u) R = 65528u;
        if (R < 1024u) return false;
        // one CTA aggregates one partition: few slots still want at least ~2 partitions per SM
        const uint32_t Rpar = (uint32_t)(((G + 2ull * M.C - 1) / (2ull * M.C) + 7ull) & ~7ull);
        if (Rpar < R) R = Rpar < 64u ? 64u : Rpar;
        const uint32_t P = (uint32_t)((G + R - 1) / R);
        if (P > MC_PMAX) return false;
        if (P < 32u) return false;                                        // a handful of slots: the shared-memory paths
        // spread the slots evenly over the partitions (smaller tables, same partition count)
        uint32_t Re = (uint32_t)((G + P - 1) / P);
        Re = (Re + 7u) & ~7u;
        if (Re > R) Re = R;
        M.R = Re; M.P = (uint32_t)((G + Re - 1) / Re);
        M.magic = ((1ull << 48) + Re - 1) / Re;
        if (pass == 1) break;
        // 32-bit integer cells only where a partition's rows cannot overflow them (4x the mean as the estimate; checked exactly per row)
        const unsigned long long rows_part = std::min<unsigned long long>(nrows, (unsigned long long)nrows / M.P * 4ull + 65536ull);
        const unsigned long long lim = 0x7fffffffULL / (rows_part + 1ull);
        bool changed = false;
        for (int a = 0; a < plan.n_acc; ++a)
            if (M.acc[a].kind == MCK_ISUM1 && st.isum_maxabs[a] > lim) { M.acc[a].kind = MCK_ISUM2; changed = true; }
        if (!changed) break;
    }
    {
        const uint32_t Rp = M.R + 8u;
        uint32_t off = 0;
        for (int a = 0; a < plan.n_acc; ++a) { M.acc[a].off = off; off += mc_cell_bytes(M.acc[a].kind) * Rp; }
        *smem_agg = off;
        if (off > budget) return false;
    }
    // ---- k_mc_scatter: ring depth
    const int ncol = M.nkeys + nvm;
    M.stages = 0;
    for (int st = MC_MAX_STAGES; st >= 2; --st)
        if (mc_scatter_layout(M.P, ncol, st).total <= budget) { M.stages = st; break; }
    if (!M.stages) return false;
    *smem_scatter = mc_scatter_layout(M.P, ncol, M.stages).total;
    const uint64_t per = (uint64_t(nrows_local) + M.C - 1) / M.C;
    M.rpc = (uint32_t)((per + MC_SUB - 1) / MC_SUB * MC_SUB);
    return true;
}

// Runs the path.  *used = false (and nothing emitted) when a per-row check failed: the caller falls back.
int run_mc(const KeyCols& kc, uint32_t nrows, AggPlan& plan, McParams& M, size_t smem_scatter, size_t smem_agg,
           OutSink& sink, uint32_t* ngroups, bool* used, uint32_t* abort_code) {
    Context& c = ctx();
    *used = false; *abort_code = 0;
    const size_t nstart = size_t(M.P) * M.C + 1;
    const size_t ntup = size_t(nrows) + size_t(M.P) * M.C * (M.lb - 1) + M.lb;      // every (partition, CTA) range is padded to whole lines
    if (ntup >= 0xffffffffull) return 0;
    const uint32_t nwords_bm = (nrows + 31) / 32 + 1;
    DevBuf starts, tuples, gtab, gfirst, bitmap, wordprefix;
    AQ_TRY(starts.alloc_zero(nstart * 4));
    AQ_TRY(tuples.alloc(ntup * M.tw * 4));
    AQ_TRY(gtab.alloc_zero(size_t(M.G) * M.gw * 8));
    AQ_TRY(gfirst.alloc(size_t(M.G) * 4));
    M.starts = starts.as<uint32_t>(); M.tuples = tuples.as<uint32_t>();
    M.gtab = gtab.as<unsigned long long>(); M.gfirst = gfirst.as<uint32_t>();
    AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 24 * sizeof(uint32_t), c.stream));
    AQ_CUDA(cudaMemsetAsync(gfirst.p, 0xff, size_t(M.G) * 4, c.stream));
    if (M.nkeys == 1) AQ_LAUNCH(k_mc_hist<1>, M.C * MC_HSPLIT, 512, size_t(M.P) * 4, M);
    else AQ_LAUNCH(k_mc_hist<2>, M.C * MC_HSPLIT, 512, size_t(M.P) * 4, M);
    AQ_LAUNCH(k_mc_round, grid_for(nstart - 1, 256), 256, 0, M.starts, (uint32_t)(nstart - 1), (uint32_t)M.lb);
    AQ_TRY(exclusive_scan_u32(M.starts, M.starts, nstart, nullptr));
    AQ_TRY(launch_scatter(M, smem_scatter));
    AQ_TRY(launch_agg(M, smem_agg));
    if (M.nkeys == 1) AQ_LAUNCH(k_mc_first<1>, c.sm_count * 8, MCF_THREADS, 0, M);
    else AQ_LAUNCH(k_mc_first<2>, c.sm_count * 8, MCF_THREADS, 0, M);
    const bool sharded = dist().active;
    uint32_t own_lo = 0, own_n = M.G;
    if (sharded) {
        // row shards: rank d owns the slots [d * Gd, (d + 1) * Gd); every rank's slice of that range goes to d in one
        // grouped send / receive, d combines the records element-wise (exact cells) and keeps the smallest global first row
        const uint32_t nr = (uint32_t)dist().nranks;
        const uint32_t Gd = (uint32_t)(((unsigned long long)M.G + nr - 1) / nr + 7ull) & ~7u;
        RecLayout L;
        memset(&L, 0, sizeof(L));
        L.gw = M.gw; L.n = M.nacc; L.cnt_word = M.acc[M.cnt_acc].gword;
        for (int a = 0; a < M.nacc; ++a) {
            L.word[a] = M.acc[a].gword;
            const int kd = M.acc[a].kind;
            L.kind[a] = kd == MCK_FX ? RK_ADD128 : (kd == MCK_MIN32 ? RK_MIN64 : (kd == MCK_MAX32 ? RK_MAX64 : RK_ADD64));
            if (kd == MCK_MAXROW || kd == MCK_TOP2) L.kind[a] = RK_UMAX64;          // local row ids: meaningless after the merge (the extras carry the values)
        }
        for (int i = 0; i < M.nx; ++i) {
            L.word[L.n] = M.x[i].word;
            L.kind[L.n++] = M.x[i].mode == 0 ? RK_UMAX64 : (M.x[i].mode == 1 ? RK_UMIN64 : RK_TOP2);
        }
        if (M.nx) AQ_LAUNCH(k_mc_localize, grid_for(M.G, 256, 2), 256, 0, M);
        AQ_TRY(dist_combine_owned(M.gtab, M.gfirst, M.G, Gd, L, c.d_flags, nullptr));
        own_lo = (uint32_t)std::min<unsigned long long>(M.G, (unsigned long long)dist().rank * Gd);
        own_n = (uint32_t)std::min<unsigned long long>(M.G, (unsigned long long)(dist().rank + 1) * Gd) - own_lo;
        AQ_CUDA(cudaMemsetAsync(M.ctl, 0, 4, c.stream));
        if (own_n) AQ_LAUNCH(k_mc_count_range, grid_for(own_n, 256, 4), 256, 0, M.gtab, M.gw, M.acc[M.cnt_acc].gword, own_lo, own_n, M.ctl);
    }
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 24 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) { *abort_code = c.h_flags[FLAG_ABORT]; return 0; }    // speculation failed (on some rank): generic path
    const uint32_t ng = c.h_flags[16];
    if (!sharded && c.h_flags[18] != ng) return fail("internal: range-partitioned path found first rows of %u of %u groups", c.h_flags[18], ng);
    if (!sink.res && ng > sink.max_groups) return fail("capacity: %u groups exceed max_groups=%u", ng, sink.max_groups);
    AQ_TRY(sink.prepare(ng, plan));
    KeyOut ko;
    memset(&ko, 0, sizeof(ko));
    ko.n = kc.n;
    for (int k = 0; k < kc.n; ++k) { ko.src[k] = kc.col[k].d; ko.size[k] = 4; ko.dst[k] = sink.key_dst[k]; }
    for (int a = 0; a < plan.emit.n; ++a) plan.emit.agg[a].out = sink.agg_dst[a];
    if (ng && sharded) {
        const uint32_t nw = (uint32_t)((dist().rows_total + 31) / 32 + 1);            // first rows are GLOBAL row ids
        DevBuf order;
        AQ_TRY(bitmap.alloc_zero(size_t(nw) * 4));
        AQ_TRY(wordprefix.alloc(size_t(nw) * 4));
        AQ_TRY(order.alloc(size_t(ng) * 4));
        AQ_LAUNCH(k_mc_mark, grid_for(own_n, 256, 4), 256, 0, M.gfirst + own_lo, own_n, bitmap.as<uint32_t>());
        AQ_TRY(bitmap_rank(bitmap.as<uint32_t>(), nw, wordprefix.as<uint32_t>()));
        AQ_LAUNCH(k_mc_order, grid_for(own_n, 256, 4), 256, 0, M.gfirst, own_lo, own_n, bitmap.as<uint32_t>(), wordprefix.as<uint32_t>(), order.as<uint32_t>());
        AQ_LAUNCH(k_mc_emit_ord, grid_for(ng, 256, 1, 8), 256, 0, M, order.as<uint32_t>(), ng, ko, sink.fr_dst, plan.emit, sink.cap);
    } else if (ng) {
        AQ_TRY(bitmap.alloc_zero(size_t(nwords_bm) * 4));
        AQ_TRY(wordprefix.alloc(size_t(nwords_bm) * 4));
        AQ_LAUNCH(k_mc_mark, grid_for(M.G, 256, 4), 256, 0, M.gfirst, M.G, bitmap.as<uint32_t>());
        AQ_TRY(bitmap_rank(bitmap.as<uint32_t>(), nwords_bm, wordprefix.as<uint32_t>()));
        const uint32_t nwords = (nrows + 31) / 32;
        AQ_LAUNCH(k_mc_emit, grid_for((nwords + 31) / 32, 8, 1, 8), 256, 0, M, bitmap.as<uint32_t>(), wordprefix.as<uint32_t>(), nwords, ko,
                  sink.fr_dst, plan.emit, sink.cap);
    }
    AQ_TRY(sink.finish(ng, plan));
    *ngroups = ng;
    *used = true;
    return 0;
}

// plan.cu entry: declines (used = false) or completes the query
int try_mc(const KeyCols& kc, uint32_t nrows, AggPlan& plan, const PlanStats& st, OutSink& sink, uint32_t* ngroups, bool* used,
           uint32_t* abort_code) {
    *used = false; *abort_code = 0;
    McParams M; size_t ss = 0, sa = 0;
    if (!mc_layout(kc, nrows, plan, st, M, &ss, &sa)) return 0;
    return run_mc(kc, nrows, plan, M, ss, sa, sink, ngroups, used, abort_code);
}

} // namespace aq
