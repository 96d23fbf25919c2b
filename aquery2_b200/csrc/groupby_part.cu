// groupby_part.cu -- partitioned aggregation for group-bys whose groups do not fit one SM's shared
// memory (h2o q3, q5, q7, q8: 1e7 groups; q6, q9: 1e4 groups with wide accumulators).
//
// Why: a DRAM-resident group table costs one random 32-byte sector read-modify-write per row, and
// B200 sustains only ~2e10 of those per second (profiles/r01/microbench.txt) -- 50x below the
// streaming rate.  So the rows are first hash-partitioned into P buckets whose groups DO fit shared
// memory (one streaming read of the needed columns, one write and one read of 16..32-byte tuples), and
// every bucket is then aggregated exactly like groupby_smem.cu does, in shared memory.
//
//   k_part_hist     row counts per (partition, scatter CTA): one pass over the key columns
//   k_part_pad, exclusive_scan_u32, k_part_units
//                   fixed tuple range of every (partition, CTA), partition offsets, work units
//   k_part_scatter  tuple [row, tag | k0 k1, v0..] of every row -> staged in shared memory per
//                   partition, written out in whole sector-aligned blocks only
//   k_part_agg      one CTA per work unit: shared-memory table, merged into the global table of
//                   groupby.cu (which then ranks the groups by first row and emits, unchanged)
//
// Replaces the same reference loops as groupby.cu (server/hasher.h:176-198, server/table.h:184-189,
// server/aggregations.h scalar aggregates); the partitioning itself has no reference counterpart.
#include "groupby_smem.cuh"
#include "sort.cuh"
#include <cstring>

namespace aq {

enum PartMode : int { PM_DIRECT = 0, PM_HASH32 = 1, PM_HASH64 = 2 };
constexpr int PART_THREADS = 512;
constexpr int PART_MAX = 8192;            // partitions
constexpr int PART_TILE_QUADS = 2;        // quads of rows per thread per tile
constexpr int PART_TILE_ROWS = PART_THREADS * 4 * PART_TILE_QUADS;
constexpr int PART_HIST_SPLIT = 4;        // CTAs of k_part_hist that share the tiles of one scatter CTA
constexpr int PART_UNR = 4;               // tuples per thread per iteration of k_part_agg
constexpr uint32_t PAD_ROW = 0xffffffffu; // row id of a padding tuple

struct PartSpec {
    int mode;
    uint32_t nparts;
    uint32_t range;                 // PM_DIRECT: tags per partition; slot = tag - partition * range
    unsigned long long magic;       // ceil(2^48 / range): partition = (tag * magic) >> 48, exact for tag < 2^31
    uint32_t chunk;                 // tuples per work unit of k_part_agg
    int nkw;                        // key words read from the columns (1 or 2)
    uint32_t bsz;                   // tuples per staging block (one flush = bsz tuples = a whole number of sectors)
    uint32_t ncta;                  // CTAs of k_part_scatter; (partition, CTA) owns a fixed range of tuples
};

__device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) { return __umulhi(a, b); }

// tag / slot hash of one row from its raw key words -- the same values key_tags4 / key_tag1 produce
__device__ __forceinline__ void part_tag(const KeySpec& ks, uint32_t k0, uint32_t k1, uint32_t& tag, uint32_t& hsh, bool& oob) {
    if (ks.exact == 1) { tag = k0; hsh = mix32(k0); }
    else if (ks.exact == 2) {
        const uint32_t d0 = k0 - ks.pack_lo[0];
        oob |= d0 > ks.pack_range[0];
        uint32_t t = d0 << ks.pack_shift[0];
        if (ks.nwords == 2) {
            const uint32_t d1 = k1 - ks.pack_lo[1];
            oob |= d1 > ks.pack_range[1];
            t |= d1 << ks.pack_shift[1];
        }
        tag = t; hsh = mix32(t);
    } else {
        uint64_t h = mix64(0x9e3779b97f4a7c15ULL ^ k0);
        if (ks.nwords == 2) h = mix64(h ^ k1);
        tag = (uint32_t)(h >> 32); hsh = (uint32_t)h;
    }
}
template <int MODE>
__device__ __forceinline__ uint32_t part_of(const PartSpec& ps, uint32_t tag, uint32_t hsh) {
    if (MODE == PM_DIRECT) return (uint32_t)(((unsigned long long)tag * ps.magic) >> 48);
    return mulhi32(hsh, ps.nparts);
}
__device__ __forceinline__ uint32_t local_start(uint32_t hsh, uint32_t cap) { return mulhi32(hsh * 0x9E3779B1u, cap); }

// ---- phase 0: per-(partition, scatter CTA) row counts ------------------------------------------------
// Tile t (PART_TILE_ROWS consecutive rows) belongs to scatter CTA t % ncta; the PART_HIST_SPLIT hist
// CTAs of one scatter CTA take its tiles round robin and add their counts into cnt[p * ncta + c].
template <int MODE>
__global__ void __launch_bounds__(PART_THREADS)
k_part_hist(const KeySpec ks, const uint32_t nrows, const PartSpec ps, uint32_t* __restrict__ cnt, uint32_t* __restrict__ flags) {
    extern __shared__ uint32_t hist[];
    for (uint32_t b = threadIdx.x; b < ps.nparts; b += PART_THREADS) hist[b] = 0;
    __syncthreads();
    const uint32_t c = blockIdx.x % ps.ncta, h = blockIdx.x / ps.ncta;
    const uint32_t nquad = (nrows + 3u) >> 2;
    const uint32_t tile_quads = PART_THREADS * PART_TILE_QUADS;
    const uint32_t ntiles = (nquad + tile_quads - 1) / tile_quads;
    bool bad = false;
    for (uint32_t tile = c + h * ps.ncta; tile < ntiles; tile += ps.ncta * PART_HIST_SPLIT) {
        uint4 kr[PART_TILE_QUADS][2]; int cnts[PART_TILE_QUADS];
#pragma unroll
        for (int i = 0; i < PART_TILE_QUADS; ++i) {
            const uint32_t q = tile * tile_quads + i * PART_THREADS + threadIdx.x;
            cnts[i] = q < nquad ? (int)min(4u, nrows - (q << 2)) : 0;
#pragma unroll
            for (int w = 0; w < 2; ++w)
                kr[i][w] = (w < ps.nkw && cnts[i]) ? load_quad(ks.base[w], q << 2, cnts[i], ks.vec_ok) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int i = 0; i < PART_TILE_QUADS; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (r >= cnts[i]) continue;
                uint32_t tag, hsh; bool oob = false;
                part_tag(ks, lane4(kr[i][0], r), lane4(kr[i][1], r), tag, hsh, oob);
                const uint32_t p = part_of<MODE>(ps, tag, hsh);
                if (oob || p >= ps.nparts) { bad = true; continue; }
                atomicAdd(&hist[p], 1u);
            }
    }
    if (bad) atomicExch(flags + FLAG_ABORT, 3u);      // a key outside the caller's bounds
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < ps.nparts; b += PART_THREADS) {
        const uint32_t v = hist[b];
        if (v) atomicAdd(cnt + size_t(b) * ps.ncta + c, v);
    }
}

// counts -> sizes of the (partition, CTA) ranges: rounded up to whole staging blocks
__global__ void k_part_pad(uint32_t* __restrict__ cnt, uint32_t n, uint32_t bsz) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        cnt[i] = (cnt[i] + bsz - 1u) / bsz * bsz;
}

// ---- phase 0b: partition offsets and work units (single CTA) ------------------------------------------
constexpr int PSCAN_THREADS = 1024;
__device__ __forceinline__ uint32_t pscan_block_excl(uint32_t v, uint32_t* sm /* 33 words */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
    if (lane == 31) sm[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = sm[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += y; }
        sm[lane] = s;
    }
    __syncthreads();
    const uint32_t r = (wid ? sm[wid - 1] : 0u) + x - v;
    __syncthreads();
    return r;
}
// pos = exclusive scan of the padded counts (nparts * ncta + 1 entries, partition-major)
__global__ void __launch_bounds__(PSCAN_THREADS)
k_part_units(const uint32_t* __restrict__ pos, uint32_t nparts, uint32_t ncta, uint32_t chunk,
             uint32_t* __restrict__ offsets, uint32_t* __restrict__ unit_off) {
    __shared__ uint32_t sm[33];
    constexpr int PER = PART_MAX / PSCAN_THREADS;
    uint32_t u[PER], su = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const uint32_t b = threadIdx.x * PER + i;
        u[i] = 0;
        if (b < nparts) {
            const uint32_t lo = pos[size_t(b) * ncta], hi = pos[size_t(b + 1) * ncta];
            offsets[b] = lo;
            if (b + 1 == nparts) offsets[nparts] = hi;
            u[i] = (hi - lo + chunk - 1u) / chunk;
        }
        su += u[i];
    }
    uint32_t eu = pscan_block_excl(su, sm);
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const uint32_t b = threadIdx.x * PER + i;
        if (b < nparts) unit_off[b] = eu;
        eu += u[i];
        if (b + 1 == nparts) unit_off[nparts] = eu;
    }
}

// ---- phase 1: scatter -------------------------------------------------------------------------------
template <int TW>
__device__ __forceinline__ void load_tuple(const uint32_t* __restrict__ src, uint32_t (&w)[TW]) {
    if (TW == 4) { const int4 t = ld_stream_v4(src); w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w; }
    else if (TW == 8) {
        const int4 a = ld_stream_v4(src), b = ld_stream_v4(src + 4);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4 % TW] = b.x; w[5 % TW] = b.y; w[6 % TW] = b.z; w[7 % TW] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < TW; i += 2) { const uint2 t = __ldg(reinterpret_cast<const uint2*>(src) + (i >> 1)); w[i] = t.x; w[(i + 1) % TW] = t.y; }
    }
}

// Random 16-byte stores reach 23 Gtuples/s on B200 (partial sectors are read back from DRAM), whole
// 128-byte lines 286 Gtuples/s (profiles/r01/microbench.txt: mb_scatter).  So every CTA stages the
// tuples per partition in shared memory, bsz tuples per partition, and only ever writes whole blocks
// (bsz * TW * 4 bytes, a multiple of 32 and sector-aligned) to the range it owns inside the partition
// (fixed by k_part_hist + scan: no cursors, no global atomics, deterministic layout).  The row that
// takes the last position of a block puts the partition on a work list; after a barrier the warps
// copy the listed blocks out.  Rows that found their block full wait for the next round of the same
// tile.  The last, partly filled block of every (partition, CTA) is completed with padding tuples
// (row = PAD_ROW).  The columns of the next tile are loaded while the current one is staged.
template <int MODE, int TW>
__global__ void __launch_bounds__(PART_THREADS, 1)
k_part_scatter(const KeySpec ks, const AccProg prog, const uint32_t nrows, const PartSpec ps,
               const uint32_t* __restrict__ pos, uint32_t* __restrict__ tuples, const uint32_t* __restrict__ flags) {
    constexpr int NKT = MODE == PM_HASH64 ? 2 : 1;
    constexpr int VB = 1 + NKT;
    constexpr int NV = TW - VB < MAX_VCOL ? TW - VB : MAX_VCOL;     // value columns a tuple can carry
    constexpr int NQ = PART_TILE_QUADS;
    constexpr int NW = PART_THREADS / 32;
    extern __shared__ __align__(16) uint32_t psm[];
    __shared__ uint32_t wl_n[2];
    if (flags[FLAG_ABORT]) return;
    const uint32_t P = ps.nparts, B = ps.bsz;
    uint32_t* const fill = psm;
    uint32_t* const wpos = psm + P;
    uint32_t* const wl = psm + 2 * P;                           // work list: partitions with a complete block
    uint32_t* const stage = psm + ((3 * P + 3) & ~3u);          // [P][B][TW] words, 16-byte aligned
    for (uint32_t b = threadIdx.x; b < P; b += PART_THREADS) { fill[b] = 0; wpos[b] = pos[size_t(b) * ps.ncta + blockIdx.x]; }
    if (threadIdx.x < 2) wl_n[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t U = B * TW / 4;                              // 16-byte units per block
    const uint32_t G = U <= 32 ? 32 / U : 1;                    // blocks a warp copies per step
    const uint32_t lane_b = U <= 32 ? lane / U : 0, lane_u = U <= 32 ? lane % U : lane;
    // copies block `b` (whole) to its place; lanes [0, U) of the group (or the whole warp, strided)
    auto copy_block = [&](uint32_t b, bool act) {
        uint32_t dst = 0;
        if (act) {
            dst = wpos[b];
            const uint4* s4 = reinterpret_cast<const uint4*>(stage) + size_t(b) * U;
            uint4* d4 = reinterpret_cast<uint4*>(tuples + size_t(dst) * TW);
            if (U <= 32) d4[lane_u] = s4[lane_u];
            else for (uint32_t u = lane; u < U; u += 32) d4[u] = s4[u];
        }
        __syncwarp();
        if (act && lane_u == 0) { const uint32_t f = fill[b]; fill[b] = f >= B ? f - B : 0u; wpos[b] = dst + B; }
    };
    const uint32_t nquad = (nrows + 3u) >> 2;
    const uint32_t tile_quads = PART_THREADS * NQ;
    const uint32_t ntiles = (nquad + tile_quads - 1) / tile_quads;
    uint4 kr[NQ][2], vr[NQ][NV > 0 ? NV : 1];
    int cnts[NQ];
    auto load_tile = [&](uint32_t tile, uint4 (&k)[NQ][2], uint4 (&v)[NQ][NV > 0 ? NV : 1], int (&c)[NQ]) {
#pragma unroll
        for (int i = 0; i < NQ; ++i) {
            const uint32_t q = tile * tile_quads + i * PART_THREADS + threadIdx.x;
            c[i] = (tile < ntiles && q < nquad) ? (int)min(4u, nrows - (q << 2)) : 0;
#pragma unroll
            for (int w = 0; w < 2; ++w)
                k[i][w] = (w < ps.nkw && c[i]) ? load_quad(ks.base[w], q << 2, c[i], ks.vec_ok) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int j = 0; j < (NV > 0 ? NV : 1); ++j)
                v[i][j] = (j < prog.nv && j < NV && c[i])
                    ? load_quad(prog.vcol[j], q << 2, c[i], (reinterpret_cast<uintptr_t>(prog.vcol[j]) & 15) == 0) : make_uint4(0, 0, 0, 0);
        }
    };
    load_tile(blockIdx.x, kr, vr, cnts);
    uint32_t round = 0;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        // partition and position inside the partition's staging sequence (carries over between tiles)
        uint32_t part[NQ][4], spos[NQ][4], ktag[NQ][4];
#pragma unroll
        for (int i = 0; i < NQ; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                part[i][r] = 0; spos[i][r] = PAD_ROW; ktag[i][r] = 0;       // PAD_ROW = nothing (left) to place
                if (r >= cnts[i]) continue;
                uint32_t tag, hsh; bool oob = false;
                part_tag(ks, lane4(kr[i][0], r), lane4(kr[i][1], r), tag, hsh, oob);
                const uint32_t p = part_of<MODE>(ps, tag, hsh);
                if (oob || p >= P) continue;                              // k_part_hist already raised the flag
                part[i][r] = p; ktag[i][r] = tag;
                spos[i][r] = atomicAdd(&fill[p], 1u);
            }
        uint4 nkr[NQ][2], nvr[NQ][NV > 0 ? NV : 1];
        int ncnts[NQ];
        load_tile(tile + gridDim.x, nkr, nvr, ncnts);                     // in flight while this tile is staged
        for (;;) {
            bool deferred = false;
            uint32_t* const my_n = &wl_n[round & 1];
#pragma unroll
            for (int i = 0; i < NQ; ++i) {
                const uint32_t row0 = (tile * tile_quads + i * PART_THREADS + threadIdx.x) << 2;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t sp = spos[i][r];
                    if (sp == PAD_ROW) continue;
                    if (sp >= B) { spos[i][r] = sp - B; deferred = true; continue; }
                    uint32_t w[TW];
#pragma unroll
                    for (int x = 0; x < TW; ++x) w[x] = 0;
                    w[0] = row0 + r;
                    if (MODE == PM_HASH64) { w[1] = lane4(kr[i][0], r); w[2 % TW] = lane4(kr[i][1], r); }
                    else w[1] = ktag[i][r];
#pragma unroll
                    for (int j = 0; j < NV; ++j) w[(VB + j) % TW] = lane4(vr[i][j], r);
                    uint32_t* d = stage + (size_t(part[i][r]) * B + sp) * TW;
                    if (TW % 4 == 0) {
#pragma unroll
                        for (int x = 0; x < TW; x += 4) *reinterpret_cast<uint4*>(d + x) = make_uint4(w[x], w[(x + 1) % TW], w[(x + 2) % TW], w[(x + 3) % TW]);
                    } else {
#pragma unroll
                        for (int x = 0; x < TW; x += 2) *reinterpret_cast<uint2*>(d + x) = make_uint2(w[x], w[(x + 1) % TW]);
                    }
                    if (sp == B - 1u) wl[atomicAdd(my_n, 1u)] = part[i][r];   // this row completes the block
                    spos[i][r] = PAD_ROW;
                }
            }
            const int more = __syncthreads_or(deferred ? 1 : 0);          // staged tuples + work list visible
            const uint32_t n = *reinterpret_cast<volatile uint32_t*>(my_n);
            if (threadIdx.x == 0) wl_n[(round + 1) & 1] = 0;
            for (uint32_t e0 = wid * G; e0 < n; e0 += NW * G) {
                const uint32_t e = e0 + lane_b;
                const bool act = lane_b < G && e < n;
                copy_block(act ? wl[e] : 0u, act);
            }
            __syncthreads();
            ++round;
            if (!more) break;
        }
#pragma unroll
        for (int i = 0; i < NQ; ++i) {
            cnts[i] = ncnts[i];
#pragma unroll
            for (int w = 0; w < 2; ++w) kr[i][w] = nkr[i][w];
#pragma unroll
            for (int j = 0; j < (NV > 0 ? NV : 1); ++j) vr[i][j] = nvr[i][j];
        }
    }
    // ---- the partly filled blocks: complete with padding tuples, then write them out
    for (uint32_t i = threadIdx.x; i < P * B; i += PART_THREADS) {
        const uint32_t b = i / B, t = i % B;
        if (t >= fill[b]) stage[size_t(i) * TW] = PAD_ROW;
    }
    __syncthreads();
    for (uint32_t b0 = wid * G; b0 < P; b0 += NW * G) {
        const uint32_t b = b0 + lane_b;
        const bool act = lane_b < G && b < P && fill[b < P ? b : 0] > 0;
        copy_block(b, act);
    }
}

// ---- phase 2: per-unit shared-memory aggregation --------------------------------------------------------
template <int TW, int VB>
__device__ __forceinline__ uint32_t tup_val(const uint32_t (&w)[TW], int ci) {
    uint32_t r = w[VB < TW ? VB : 0];
    if (VB + 1 < TW && ci == 1) r = w[VB + 1 < TW ? VB + 1 : 0];
    if (VB + 2 < TW && ci == 2) r = w[VB + 2 < TW ? VB + 2 : 0];
    if (VB + 3 < TW && ci == 3) r = w[VB + 3 < TW ? VB + 3 : 0];
    return r;
}
// x for accumulator o from a tuple (same meaning as decode_x of groupby_smem.cu)
template <int TW, int VB>
__device__ __forceinline__ void tup_x(const AccOp& o, const uint32_t (&w)[TW], long long& xi, double& xd) {
    xi = 0; xd = 0.0;
    if (!acc_needs_value(o)) return;
    const uint32_t u = tup_val<TW, VB>(w, o.ci);
    const uint32_t u2 = o.xform == X_PRODUCT ? tup_val<TW, VB>(w, o.ci2) : u;
    if (o.vtype == AQ_INT32) {
        xi = (long long)(int)u;
        if (o.xform != X_IDENT) xi *= (long long)(int)u2;
    } else {
        const float f = __uint_as_float(u);
        xd = o.xform == X_IDENT ? (double)f : (double)__fmul_rn(f, __uint_as_float(u2));
    }
}

// lay.cap = slots of the table; every array has lay.cap + 8 entries, entry lay.cap is a trash slot that
// absorbs the (zero-valued) updates of tail lanes and of rows that were sent to the global table.
template <int MODE, int TW>
__global__ void __launch_bounds__(PART_THREADS, 1)
k_part_agg(const KeySpec ks, const uint32_t* __restrict__ tuples, const uint32_t* __restrict__ offsets,
           const uint32_t* __restrict__ unit_off, const PartSpec ps, const uint32_t gmask, const uint32_t max_probe,
           const AccProg prog, const SmemLayout lay, uint32_t* __restrict__ flags) {
    constexpr int NKT = MODE == PM_HASH64 ? 2 : 1;
    constexpr int VB = 1 + NKT;
    constexpr bool EXACT = MODE != PM_HASH64;
    constexpr int UNR = PART_UNR;
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ uint32_t n_used;
    __shared__ uint32_t s_part;
    if (flags[FLAG_ABORT]) return;
    const uint32_t unit = blockIdx.x;
    if (unit >= unit_off[ps.nparts]) return;
    if (threadIdx.x == 0) {                      // partition of this unit: last p with unit_off[p] <= unit
        uint32_t lo = 0, hi = ps.nparts;         // invariant: unit_off[lo] <= unit < unit_off[hi]
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (unit_off[mid] <= unit) lo = mid; else hi = mid; }
        s_part = lo; n_used = 0;
    }
    const uint32_t cap = lay.cap, capa = lay.cap + 8u, trash = lay.cap;
    uint32_t* const minrow = reinterpret_cast<uint32_t*>(sm + lay.off_minrow);
    uint32_t* const key32 = reinterpret_cast<uint32_t*>(sm + lay.off_key);
    unsigned long long* const key64 = reinterpret_cast<unsigned long long*>(sm + lay.off_key);
    for (uint32_t s = threadIdx.x; s < capa; s += PART_THREADS) {
        minrow[s] = EMPTY32;
        if (MODE == PM_HASH32) key32[s] = EMPTY32;
        if (MODE == PM_HASH64) key64[s] = EMPTY64;
        for (int a = 0; a < prog.n; ++a) {
            unsigned char* c0 = sm + lay.off_cell[a];
            switch (lay.kind[a]) {
            case CK_ADD_I64: reinterpret_cast<uint32_t*>(c0)[s] = 0u; reinterpret_cast<uint32_t*>(c0)[s + capa] = 0u; break;
            case CK_ADD_F64: case CK_TOP2: reinterpret_cast<unsigned long long*>(c0)[s] = 0ull; break;
            case CK_MIN32: reinterpret_cast<int*>(c0)[s] = 0x7fffffff; break;
            case CK_MAX32: reinterpret_cast<int*>(c0)[s] = (int)0x80000000u; break;
            case CK_MIN64: reinterpret_cast<unsigned long long*>(c0)[s] = 0x7fffffffffffffffULL; break;
            case CK_MAX64: reinterpret_cast<unsigned long long*>(c0)[s] = 0x8000000000000000ULL; break;
            default: reinterpret_cast<uint32_t*>(c0)[s] = 0u; break;
            }
        }
    }
    __syncthreads();
    const uint32_t part = s_part;
    const uint32_t t_lo = offsets[part] + (unit - unit_off[part]) * ps.chunk;
    const uint32_t t_hi = min(offsets[part + 1], t_lo + ps.chunk);
    const uint32_t part_base = part * ps.range;      // PM_DIRECT: first tag of this partition
    bool dead = false;

    for (uint32_t base = t_lo; base < t_hi && !dead; base += PART_THREADS * UNR) {
        uint32_t w[UNR][TW];
        bool valid[UNR];
#pragma unroll
        for (int j = 0; j < UNR; ++j) {
            const uint32_t t = base + j * PART_THREADS + threadIdx.x;
            valid[j] = t < t_hi;
            if (valid[j]) load_tuple<TW>(tuples + size_t(t) * TW, w[j]);
            valid[j] = valid[j] && w[j][0] != PAD_ROW;
            if (!valid[j]) {
#pragma unroll
                for (int x = 0; x < TW; ++x) w[j][x] = 0;
            }
        }
        uint32_t sl[UNR];
#pragma unroll
        for (int j = 0; j < UNR; ++j) {
            sl[j] = trash;
            if (!valid[j]) continue;
            const uint32_t row = w[j][0];
            uint32_t s = NOSLOT, tag = w[j][1], hsh = 0;
            if (MODE == PM_DIRECT) {
                s = tag - part_base;
            } else if (MODE == PM_HASH32) {
                hsh = mix32(tag);
                if (tag != EMPTY32) {
                    uint32_t p = local_start(hsh, cap);
                    for (int step = 0; step < 64; ++step) {
                        uint32_t cur = *reinterpret_cast<volatile uint32_t*>(key32 + p);
                        if (cur == EMPTY32) {
                            if (*reinterpret_cast<volatile uint32_t*>(&n_used) >= lay.limit) break;
                            cur = atomicCAS(key32 + p, EMPTY32, tag);
                            if (cur == EMPTY32) { atomicAdd(&n_used, 1u); s = p; break; }
                        }
                        if (cur == tag) { s = p; break; }
                        p = (p + 1u == cap) ? 0u : p + 1u;
                    }
                }
            } else {
                const unsigned long long mine = (unsigned long long)w[j][1] | ((unsigned long long)w[j][2 % TW] << 32);
                bool oob = false;
                part_tag(ks, w[j][1], w[j][2 % TW], tag, hsh, oob);
                if (mine != EMPTY64) {
                    uint32_t p = local_start(hsh, cap);
                    for (int step = 0; step < 64; ++step) {
                        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(key64 + p);
                        if (cur == EMPTY64) {
                            if (*reinterpret_cast<volatile uint32_t*>(&n_used) >= lay.limit) break;
                            cur = atomicCAS(key64 + p, EMPTY64, mine);
                            if (cur == EMPTY64) { atomicAdd(&n_used, 1u); s = p; break; }
                        }
                        if (cur == mine) { s = p; break; }
                        p = (p + 1u == cap) ? 0u : p + 1u;
                    }
                }
            }
            if (s != NOSLOT) {
                sl[j] = s;
                if (row < *reinterpret_cast<volatile uint32_t*>(minrow + s)) atomicMin(minrow + s, row);
            } else {
                // no room in shared memory (or a key equal to the EMPTY marker): straight to the global table
                const uint32_t g = probe_insert<EXACT>(ks, prog.tab, prog.sw, gmask, max_probe, hsh, tag, row);
                if (g == NOSLOT) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); dead = true; break; }
                for (int a = 0; a < prog.n; ++a) {
                    long long xi; double xd;
                    tup_x<TW, VB>(prog.op[a], w[j], xi, xd);
                    acc_global_row(prog.op[a], slot_word(prog, g) + 1 + a, row, xi, xd);
                }
#pragma unroll
                for (int x = 1; x < TW; ++x) w[j][x] = 0;      // its update of the trash slot adds nothing
            }
        }
        if (dead) break;
        // ---- accumulate: dispatch on the cell kind once per accumulator, straight-line row code
        for (int a = 0; a < prog.n; ++a) {
            const int kind = lay.kind[a];
            unsigned char* const c0 = sm + lay.off_cell[a];
            const int vtype = prog.op[a].vtype, xform = prog.op[a].xform;
            uint32_t u[UNR], u2[UNR];
            if (kind != CK_COUNT && kind != CK_MAXROW && kind != CK_TOP2) {
#pragma unroll
                for (int j = 0; j < UNR; ++j) {
                    u[j] = tup_val<TW, VB>(w[j], prog.op[a].ci);
                    u2[j] = xform == X_PRODUCT ? tup_val<TW, VB>(w[j], prog.op[a].ci2) : u[j];
                }
            }
            switch (kind) {
            case CK_COUNT:
#pragma unroll
                for (int j = 0; j < UNR; ++j) atomicAdd(reinterpret_cast<uint32_t*>(c0) + sl[j], 1u);
                break;
            case CK_ADD_I64:
#pragma unroll
                for (int j = 0; j < UNR; ++j) {
                    const long long x = xform == X_IDENT ? (long long)(int)u[j] : (long long)(int)u[j] * (long long)(int)u2[j];
                    uint32_t* lo = reinterpret_cast<uint32_t*>(c0) + sl[j];
                    const uint32_t xl = (uint32_t)x;
                    const uint32_t old = atomicAdd(lo, xl);
                    const uint32_t d = (uint32_t)((unsigned long long)x >> 32) + (((uint32_t)(old + xl) < old) ? 1u : 0u);
                    if (d) atomicAdd(lo + capa, d);
                }
                break;
            case CK_ADD_F64: {
                double x[UNR];
                unsigned long long old[UNR];
                unsigned pending = 0;
#pragma unroll
                for (int j = 0; j < UNR; ++j) {
                    if (vtype == AQ_INT32) x[j] = xform == X_IDENT ? (double)(int)u[j] : (double)((long long)(int)u[j] * (long long)(int)u2[j]);
                    else x[j] = xform == X_IDENT ? (double)__uint_as_float(u[j]) : (double)__fmul_rn(__uint_as_float(u[j]), __uint_as_float(u2[j]));
                    old[j] = *(reinterpret_cast<volatile unsigned long long*>(c0) + sl[j]);
                    pending |= 1u << j;
                }
                while (pending) {
                    unsigned long long seen[UNR];
#pragma unroll
                    for (int j = 0; j < UNR; ++j)
                        if (pending & (1u << j))
                            seen[j] = atomicCAS(reinterpret_cast<unsigned long long*>(c0) + sl[j], old[j],
                                                (unsigned long long)__double_as_longlong(__longlong_as_double((long long)old[j]) + x[j]));
#pragma unroll
                    for (int j = 0; j < UNR; ++j)
                        if (pending & (1u << j)) {
                            if (seen[j] == old[j]) pending &= ~(1u << j);
                            else old[j] = seen[j];
                        }
                }
                break; }
            case CK_MIN32: case CK_MAX32:
#pragma unroll
                for (int j = 0; j < UNR; ++j) {
                    if (sl[j] == trash) continue;
                    int* p = reinterpret_cast<int*>(c0) + sl[j];
                    const int b = (int)u[j];
                    const int img = vtype == AQ_INT32 ? b : (b ^ ((b >> 31) & 0x7fffffff));
                    const int cur = *reinterpret_cast<volatile int*>(p);
                    if (kind == CK_MIN32) { if (img < cur) atomicMin(p, img); }
                    else { if (img > cur) atomicMax(p, img); }
                }
                break;
            default:
#pragma unroll
                for (int j = 0; j < UNR; ++j) {
                    if (sl[j] == trash) continue;
                    long long xi; double xd;
                    tup_x<TW, VB>(prog.op[a], w[j], xi, xd);
                    smem_acc_row(prog.op[a], kind, c0, capa, sl[j], w[j][0], xi, xd);
                }
                break;
            }
        }
    }
    __syncthreads();
    if (*reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT)) return;

    // ---- merge this unit's groups into the global table
    for (uint32_t s = threadIdx.x; s < cap; s += PART_THREADS) {
        const uint32_t row = minrow[s];
        if (row == EMPTY32) continue;
        uint32_t t, h;
        if (MODE == PM_DIRECT) { t = part_base + s; h = mix32(t); }
        else if (MODE == PM_HASH32) { t = key32[s]; h = mix32(t); }
        else { const unsigned long long k = key64[s]; bool oob = false; part_tag(ks, (uint32_t)k, (uint32_t)(k >> 32), t, h, oob); }
        const uint32_t g = probe_insert<EXACT>(ks, prog.tab, prog.sw, gmask, max_probe, h, t, row);
        if (g == NOSLOT) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); return; }
        for (int a = 0; a < prog.n; ++a)
            smem_merge(prog.op[a], lay.kind[a], sm + lay.off_cell[a], capa, s, slot_word(prog, g) + 1 + a);
    }
}

// ---- host ---------------------------------------------------------------------------------------------
template <int MODE, int TW>
static int run_part(const KeySpec& ks, uint32_t nrows, const PartSpec& ps, const AccProg& prog, const SmemLayout& lay,
                    uint32_t gmask, uint32_t max_probe, size_t stage_bytes) {
    Context& c = ctx();
    const size_t P = ps.nparts, C = ps.ncta;
    const size_t cap_tuples = size_t(nrows) + P * C * (ps.bsz - 1);           // every (partition, CTA) range is padded to whole blocks
    if (cap_tuples >= 0xffffffffull) return fail("partitioned group-by: %zu tuples exceed the 32-bit index space", cap_tuples);
    DevBuf tuples, cnt, small;
    AQ_TRY(tuples.alloc(cap_tuples * TW * 4));
    AQ_TRY(cnt.alloc_zero((P * C + 1) * 4));
    AQ_TRY(small.alloc((2 * P + 2) * 4));
    uint32_t* offsets = small.as<uint32_t>();
    uint32_t* unit_off = offsets + P + 1;
    AQ_LAUNCH((k_part_hist<MODE>), (unsigned)(C * PART_HIST_SPLIT), PART_THREADS, P * 4, ks, nrows, ps, cnt.as<uint32_t>(), c.d_flags);
    AQ_LAUNCH(k_part_pad, grid_for(P * C, 256), 256, 0, cnt.as<uint32_t>(), (uint32_t)(P * C), ps.bsz);
    AQ_TRY(exclusive_scan_u32(cnt.as<uint32_t>(), cnt.as<uint32_t>(), P * C + 1, nullptr));
    AQ_LAUNCH(k_part_units, 1, PSCAN_THREADS, 0, cnt.as<uint32_t>(), ps.nparts, ps.ncta, ps.chunk, offsets, unit_off);
    static bool attr_done = false;
    if (!attr_done) {
        AQ_CUDA(cudaFuncSetAttribute(k_part_scatter<MODE, TW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(c.smem_optin - 1024)));
        AQ_CUDA(cudaFuncSetAttribute(k_part_agg<MODE, TW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(c.smem_optin - 1024)));
        attr_done = true;
    }
    AQ_LAUNCH((k_part_scatter<MODE, TW>), (unsigned)C, PART_THREADS, stage_bytes, ks, prog, nrows, ps, cnt.as<uint32_t>(),
              tuples.as<uint32_t>(), c.d_flags);
    const uint32_t max_units = (uint32_t)(cap_tuples / ps.chunk + P + 1);
    AQ_LAUNCH((k_part_agg<MODE, TW>), max_units, PART_THREADS, lay.total_bytes, ks, tuples.as<uint32_t>(), offsets, unit_off, ps,
              gmask, max_probe, prog, lay, c.d_flags);
    return 0;
}

int launch_part_probe(const KeySpec& ks, uint32_t nrows, uint32_t gcap, uint32_t max_probe,
                      const AccProg& prog, uint32_t groups_hint, bool* used) {
    *used = false;
    Context& c = ctx();
    if (groups_hint == 0 || nrows < 4096) return 0;
    if (ks.nwords < 1 || ks.nwords > 2 || ks.stride[0] != 1 || (ks.nwords == 2 && ks.stride[1] != 1)) return 0;
    if (prog.nv < 0) return 0;
    for (int j = 0; j < prog.nv; ++j) if (prog.vctype[j] != AQ_INT32 && prog.vctype[j] != AQ_FLOAT) return 0;
    if (c.smem_optin < 96 * 1024) return 0;
    PartSpec ps;
    memset(&ps, 0, sizeof(ps));
    ps.nkw = ks.nwords;
    ps.ncta = (uint32_t)c.sm_count;
    ps.mode = ks.exact ? PM_HASH32 : PM_HASH64;
    if (!ks.exact && ks.nwords != 2) return 0;
    SmemLayout lay;
    memset(&lay, 0, sizeof(lay));
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
    const size_t budget = c.smem_optin - 1024 - 64;
    // ---- tuple width
    const int nkt = ps.mode == PM_HASH64 ? 2 : 1;
    if (1 + nkt + prog.nv > 8) return 0;
    int tw = 1 + nkt + prog.nv;
    tw = tw <= 4 ? 4 : (tw <= 6 ? 6 : 8);
    // ---- partitions and table capacity
    bool direct = false;
    if (ks.exact == 2 && ks.pack_bits >= 1 && ks.pack_bits <= 31) {
        uint64_t max_tag = 0;                                     // largest tag the bounds allow
        for (int w = 0; w < ks.nwords; ++w) max_tag |= uint64_t(ks.pack_range[w]) << ks.pack_shift[w];
        uint32_t range = (uint32_t)(budget / bps);
        range = (range - 8) & ~7u;
        if (range > 65536) range = 65536;
        const uint64_t np = max_tag / range + 1;
        if (range >= 256 && np <= PART_MAX) {
            direct = true;
            ps.mode = PM_DIRECT; ps.range = range; ps.nparts = (uint32_t)np;
            ps.magic = ((1ull << 48) + range - 1) / range;
            lay.cap = range; lay.limit = range;
        }
    }
    if (!direct) {
        const uint32_t kbytes = ps.mode == PM_HASH64 ? 8 : 4;
        bps += kbytes;
        uint32_t cap = (uint32_t)(budget / bps);
        cap = (cap - 8) & ~7u;
        if (cap < 1024) return 0;
        lay.cap = cap; lay.limit = (uint32_t)(uint64_t(cap) * 85 / 100);
        const uint64_t per = uint64_t(cap) * 60 / 100;
        const uint64_t np = (uint64_t(groups_hint) + per - 1) / per;
        if (np > PART_MAX) return 0;
        ps.nparts = (uint32_t)(np ? np : 1);
    }
    // ---- staging blocks of the scatter kernel: bsz tuples = a whole number of 32-byte sectors; as large as
    // shared memory allows (few partitions: long blocks, so that a tile rarely overflows a block twice)
    uint32_t bsz = 0;
    size_t stage_bytes = 0;
    for (uint32_t b = 2048; b >= 1; b >>= 1) {
        const size_t blk = size_t(b) * tw * 4;
        if (blk % 32 != 0) continue;
        const size_t need = ((3 * size_t(ps.nparts) + 3) & ~size_t(3)) * 4 + size_t(ps.nparts) * blk;
        if (need <= budget) { bsz = b; stage_bytes = need; break; }
    }
    if (bsz == 0) return 0;                                       // too many partitions for one SM's staging area
    ps.bsz = bsz;
    {   // work-unit size: a few waves of units, but few enough that merging stays cheap; whole blocks
        uint64_t chunk = uint64_t(nrows) / (uint64_t(c.sm_count) * 6);
        if (chunk < 16384) chunk = 16384;
        if (chunk > (1u << 20)) chunk = 1u << 20;
        ps.chunk = (uint32_t)(chunk / 8 * 8);
    }
    // ---- shared-memory layout of k_part_agg: every array has cap + 8 entries (trash slot at index cap)
    const uint32_t capa = lay.cap + 8;
    lay.mode = ps.mode; lay.mask = 0; lay.reps = 1;
    uint32_t off = 0;
    if (ps.mode == PM_HASH64) { lay.off_key = off; off += 8 * capa; }
    for (int a = 0; a < prog.n; ++a) if (cell_bytes[a] == 8) { lay.off_cell[a] = off; off += 8 * capa; }
    if (ps.mode == PM_HASH32) { lay.off_key = off; off += 4 * capa; }
    lay.off_minrow = off; off += 4 * capa;
    for (int a = 0; a < prog.n; ++a) if (cell_bytes[a] == 4) { lay.off_cell[a] = off; off += 4 * capa; }
    lay.total_bytes = off;
    if (off > c.smem_optin - 1024) return 0;
    const uint32_t gmask = gcap - 1;
#define AQ_PART_CASE(M, T) if (ps.mode == M && tw == T) { AQ_TRY((run_part<M, T>(ks, nrows, ps, prog, lay, gmask, max_probe, stage_bytes))); *used = true; return 0; }
    AQ_PART_CASE(PM_DIRECT, 4) AQ_PART_CASE(PM_DIRECT, 6) AQ_PART_CASE(PM_DIRECT, 8)
    AQ_PART_CASE(PM_HASH32, 4) AQ_PART_CASE(PM_HASH32, 6) AQ_PART_CASE(PM_HASH32, 8)
    AQ_PART_CASE(PM_HASH64, 4) AQ_PART_CASE(PM_HASH64, 6) AQ_PART_CASE(PM_HASH64, 8)
#undef AQ_PART_CASE
    return 0;
}

} // namespace aq
