// groupby_part.cu -- partitioned aggregation for group-bys whose groups do not fit one SM's shared
// memory (h2o q3, q5, q7, q8: 1e7 groups; q6, q9: 1e4 groups with wide accumulators).
//
// Why: a DRAM-resident group table costs one random 32-byte sector read-modify-write per row, and
// B200 sustains only ~2e10 of those per second (profiles/r01/microbench.txt) -- 50x below the
// streaming rate.  So the rows are first hash-partitioned into P buckets whose groups DO fit shared
// memory (one streaming read of the needed columns, one write and one read of 16..32-byte tuples), and
// every bucket is then aggregated exactly like groupby_smem.cu does, in shared memory.
//
//   k_part_hist     per-partition row counts (one pass over the key columns)
//   k_part_scan     exclusive scan -> partition offsets, write cursors, work units of <= chunk tuples
//   k_part_scatter  tuple [row, tag | k0 k1, v0..] of every row -> its partition's range
//   k_part_agg      one CTA per work unit: shared-memory table, merged into the global table of
//                   groupby.cu (which then ranks the groups by first row and emits, unchanged)
//
// Replaces the same reference loops as groupby.cu (server/hasher.h:176-198, server/table.h:184-189,
// server/aggregations.h scalar aggregates); the partitioning itself has no reference counterpart.
#include "groupby_smem.cuh"
#include <cstring>

namespace aq {

enum PartMode : int { PM_DIRECT = 0, PM_HASH32 = 1, PM_HASH64 = 2 };
constexpr int PART_THREADS = 512;
constexpr int PART_MAX = 8192;            // partitions (shared histogram of the scatter kernel)
constexpr int PART_TILE_QUADS = 2;        // quads of rows per thread per scatter tile
constexpr int PART_UNR = 4;               // tuples per thread per iteration of k_part_agg

struct PartSpec {
    int mode;
    uint32_t nparts;
    uint32_t dbits;        // PM_DIRECT: slot = tag & ((1 << dbits) - 1), partition = tag >> dbits
    uint32_t chunk;        // tuples per work unit
    int nkw;               // key words read from the columns (1 or 2)
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
    if (MODE == PM_DIRECT) return tag >> ps.dbits;
    return mulhi32(hsh, ps.nparts);
}
__device__ __forceinline__ uint32_t local_start(uint32_t hsh, uint32_t cap) { return mulhi32(hsh * 0x9E3779B1u, cap); }

// ---- phase 0: histogram ---------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(PART_THREADS)
k_part_hist(const KeySpec ks, const uint32_t nrows, const PartSpec ps, uint32_t* __restrict__ ghist, uint32_t* __restrict__ flags) {
    extern __shared__ uint32_t hist[];
    for (uint32_t b = threadIdx.x; b < ps.nparts; b += PART_THREADS) hist[b] = 0;
    __syncthreads();
    const uint32_t nquad = (nrows + 3u) >> 2;
    const uint32_t stride = gridDim.x * PART_THREADS;
    bool bad = false;
    for (uint32_t q0 = blockIdx.x * PART_THREADS + threadIdx.x; q0 < nquad; q0 += 2 * stride) {
        uint4 kr[2][2]; int cnts[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const uint32_t q = q0 + i * stride;
            cnts[i] = q < nquad ? (int)min(4u, nrows - (q << 2)) : 0;
#pragma unroll
            for (int w = 0; w < 2; ++w)
                kr[i][w] = (w < ps.nkw && cnts[i]) ? load_quad(ks.base[w], q << 2, cnts[i], ks.vec_ok) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
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
        const uint32_t c = hist[b];
        if (c) atomicAdd(ghist + b, c);
    }
}

// ---- phase 0b: offsets, cursors, work units (single CTA) --------------------------------------------
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
__global__ void __launch_bounds__(PSCAN_THREADS)
k_part_scan(const uint32_t* __restrict__ ghist, uint32_t nparts, uint32_t chunk, uint32_t* __restrict__ offsets,
            uint32_t* __restrict__ cursor, uint32_t* __restrict__ unit_off) {
    __shared__ uint32_t sm[33];
    constexpr int PER = PART_MAX / PSCAN_THREADS;
    uint32_t c[PER], u[PER], sc = 0, su = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const uint32_t b = threadIdx.x * PER + i;
        c[i] = b < nparts ? ghist[b] : 0u;
        u[i] = (c[i] + chunk - 1u) / chunk;
        sc += c[i]; su += u[i];
    }
    uint32_t ec = pscan_block_excl(sc, sm);
    uint32_t eu = pscan_block_excl(su, sm);
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const uint32_t b = threadIdx.x * PER + i;
        if (b < nparts) { offsets[b] = ec; cursor[b] = ec; unit_off[b] = eu; }
        ec += c[i]; eu += u[i];
        if (b + 1 == nparts) { offsets[nparts] = ec; unit_off[nparts] = eu; }
    }
}

// ---- phase 1: scatter -------------------------------------------------------------------------------
template <int TW>
__device__ __forceinline__ void store_tuple(uint32_t* __restrict__ dst, const uint32_t (&w)[TW]) {
    if (TW == 4) *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    else if (TW == 8) {
        reinterpret_cast<uint4*>(dst)[0] = make_uint4(w[0], w[1], w[2], w[3]);
        reinterpret_cast<uint4*>(dst)[1] = make_uint4(w[4 % TW], w[5 % TW], w[6 % TW], w[7 % TW]);
    } else {
#pragma unroll
        for (int i = 0; i < TW; i += 2) reinterpret_cast<uint2*>(dst)[i >> 1] = make_uint2(w[i], w[(i + 1) % TW]);
    }
}
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

// One tile = PART_THREADS * 4 * PART_TILE_QUADS consecutive rows.  rank inside (tile, partition) = value
// returned by the shared atomicAdd; one global atomicAdd per non-empty (tile, partition) reserves
// the run in the partition's range.  NKT = key words carried in the tuple.
template <int MODE, int TW>
__global__ void __launch_bounds__(PART_THREADS, 2)
k_part_scatter(const KeySpec ks, const AccProg prog, const uint32_t nrows, const PartSpec ps,
               uint32_t* __restrict__ cursor, uint32_t* __restrict__ tuples, const uint32_t* __restrict__ flags) {
    constexpr int NKT = MODE == PM_HASH64 ? 2 : 1;
    constexpr int VB = 1 + NKT;
    constexpr int NQ = PART_TILE_QUADS;
    extern __shared__ uint32_t hist[];
    if (flags[FLAG_ABORT]) return;
    for (uint32_t b = threadIdx.x; b < ps.nparts; b += PART_THREADS) hist[b] = 0;
    __syncthreads();
    const uint32_t nquad = (nrows + 3u) >> 2;
    const uint32_t tile_quads = PART_THREADS * NQ;
    const uint32_t ntiles = (nquad + tile_quads - 1) / tile_quads;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        uint4 kr[NQ][2], vr[NQ][MAX_VCOL];
        int cnts[NQ];
#pragma unroll
        for (int i = 0; i < NQ; ++i) {
            const uint32_t q = tile * tile_quads + i * PART_THREADS + threadIdx.x;
            cnts[i] = q < nquad ? (int)min(4u, nrows - (q << 2)) : 0;
#pragma unroll
            for (int w = 0; w < 2; ++w)
                kr[i][w] = (w < ps.nkw && cnts[i]) ? load_quad(ks.base[w], q << 2, cnts[i], ks.vec_ok) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int j = 0; j < MAX_VCOL; ++j)
                vr[i][j] = (j < prog.nv && j < TW - VB && cnts[i])
                    ? load_quad(prog.vcol[j], q << 2, cnts[i], (reinterpret_cast<uintptr_t>(prog.vcol[j]) & 15) == 0) : make_uint4(0, 0, 0, 0);
        }
        uint32_t part[NQ][4], rank[NQ][4], ktag[NQ][4];
#pragma unroll
        for (int i = 0; i < NQ; ++i)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                part[i][r] = 0xffffffffu; rank[i][r] = 0; ktag[i][r] = 0;
                if (r >= cnts[i]) continue;
                uint32_t tag, hsh; bool oob = false;
                part_tag(ks, lane4(kr[i][0], r), lane4(kr[i][1], r), tag, hsh, oob);
                const uint32_t p = part_of<MODE>(ps, tag, hsh);
                if (oob || p >= ps.nparts) continue;              // k_part_hist already raised the flag
                part[i][r] = p; ktag[i][r] = tag;
                rank[i][r] = atomicAdd(&hist[p], 1u);
            }
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < ps.nparts; b += PART_THREADS) {
            const uint32_t c = hist[b];
            if (c) hist[b] = atomicAdd(cursor + b, c);
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < NQ; ++i) {
            const uint32_t row0 = (tile * tile_quads + i * PART_THREADS + threadIdx.x) << 2;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                if (part[i][r] == 0xffffffffu) continue;
                uint32_t w[TW];
#pragma unroll
                for (int x = 0; x < TW; ++x) w[x] = 0;
                w[0] = row0 + r;
                if (MODE == PM_HASH64) { w[1] = lane4(kr[i][0], r); w[2] = lane4(kr[i][1], r); }
                else w[1] = ktag[i][r];
#pragma unroll
                for (int j = 0; j < MAX_VCOL; ++j) if (j < TW - VB) w[VB + (j < TW - VB ? j : 0)] = lane4(vr[i][j], r);
                const uint32_t pos = hist[part[i][r]] + rank[i][r];
                store_tuple<TW>(tuples + size_t(pos) * TW, w);
            }
        }
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < ps.nparts; b += PART_THREADS) hist[b] = 0;
        __syncthreads();
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
    const uint32_t dmask = (1u << ps.dbits) - 1u;
    bool dead = false;

    for (uint32_t base = t_lo; base < t_hi && !dead; base += PART_THREADS * UNR) {
        uint32_t w[UNR][TW];
        bool valid[UNR];
#pragma unroll
        for (int j = 0; j < UNR; ++j) {
            const uint32_t t = base + j * PART_THREADS + threadIdx.x;
            valid[j] = t < t_hi;
            if (valid[j]) load_tuple<TW>(tuples + size_t(t) * TW, w[j]);
            else {
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
                s = tag & dmask;
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
        if (MODE == PM_DIRECT) { t = (part << ps.dbits) | s; h = mix32(t); }
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
                    uint32_t gmask, uint32_t max_probe) {
    Context& c = ctx();
    DevBuf tuples, small;
    AQ_TRY(tuples.alloc(size_t(nrows) * TW * 4));
    const size_t P = ps.nparts;
    AQ_TRY(small.alloc_zero((P + (P + 1) + P + (P + 1)) * 4));
    uint32_t* ghist = small.as<uint32_t>();
    uint32_t* offsets = ghist + P;
    uint32_t* cursor = offsets + P + 1;
    uint32_t* unit_off = cursor + P;
    const size_t hist_bytes = P * 4;
    const uint32_t nquad = (nrows + 3) / 4;
    int grid_h = (int)((nquad + PART_THREADS - 1) / PART_THREADS);
    if (grid_h > c.sm_count * 4) grid_h = c.sm_count * 4;
    AQ_LAUNCH((k_part_hist<MODE>), grid_h, PART_THREADS, hist_bytes, ks, nrows, ps, ghist, c.d_flags);
    AQ_LAUNCH(k_part_scan, 1, PSCAN_THREADS, 0, ghist, ps.nparts, ps.chunk, offsets, cursor, unit_off);
    const uint32_t tile_quads = PART_THREADS * PART_TILE_QUADS;
    int grid_s = (int)((nquad + tile_quads - 1) / tile_quads);
    if (grid_s > c.sm_count * 3) grid_s = c.sm_count * 3;
    AQ_LAUNCH((k_part_scatter<MODE, TW>), grid_s, PART_THREADS, hist_bytes, ks, prog, nrows, ps, cursor, tuples.as<uint32_t>(), c.d_flags);
    static bool attr_done = false;
    if (!attr_done) {
        AQ_CUDA(cudaFuncSetAttribute(k_part_agg<MODE, TW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(c.smem_optin - 1024)));
        attr_done = true;
    }
    const uint32_t max_units = nrows / ps.chunk + ps.nparts + 1;
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
    if (c.smem_optin < 64 * 1024) return 0;
    PartSpec ps;
    memset(&ps, 0, sizeof(ps));
    ps.nkw = ks.nwords;
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
    // ---- partitions and table capacity
    bool direct = false;
    if (ks.exact == 2 && ks.pack_bits >= 1 && ks.pack_bits <= 31) {
        uint32_t dbits = 0;
        while (dbits < 20 && (size_t(bps) * ((size_t(2) << dbits) + 8)) <= budget) ++dbits;   // largest 2^dbits that fits
        if (dbits > (uint32_t)ks.pack_bits) dbits = ks.pack_bits;
        const uint64_t np = (uint64_t(1) << ks.pack_bits) >> dbits;
        if (dbits >= 8 && np <= PART_MAX) {
            direct = true;
            ps.mode = PM_DIRECT; ps.dbits = dbits; ps.nparts = (uint32_t)(np ? np : 1);
            lay.cap = 1u << dbits; lay.limit = lay.cap;
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
    {   // work-unit size: a few waves of units, but few enough that merging stays cheap
        uint64_t chunk = uint64_t(nrows) / (uint64_t(c.sm_count) * 6);
        if (chunk < 16384) chunk = 16384;
        if (chunk > (1u << 20)) chunk = 1u << 20;
        ps.chunk = (uint32_t)chunk;
    }
    // ---- shared-memory layout: every array has cap + 8 entries (trash slot at index cap)
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
    // ---- tuple width
    const int nkt = ps.mode == PM_HASH64 ? 2 : 1;
    int tw = 1 + nkt + prog.nv;
    tw = tw <= 4 ? 4 : (tw <= 6 ? 6 : 8);
    if (1 + nkt + prog.nv > 8) return 0;
    const uint32_t gmask = gcap - 1;
#define AQ_PART_CASE(M, T) if (ps.mode == M && tw == T) { AQ_TRY((run_part<M, T>(ks, nrows, ps, prog, lay, gmask, max_probe))); *used = true; return 0; }
    AQ_PART_CASE(PM_DIRECT, 4) AQ_PART_CASE(PM_DIRECT, 6) AQ_PART_CASE(PM_DIRECT, 8)
    AQ_PART_CASE(PM_HASH32, 4) AQ_PART_CASE(PM_HASH32, 6) AQ_PART_CASE(PM_HASH32, 8)
    AQ_PART_CASE(PM_HASH64, 4) AQ_PART_CASE(PM_HASH64, 6) AQ_PART_CASE(PM_HASH64, 8)
#undef AQ_PART_CASE
    return 0;
}

} // namespace aq
