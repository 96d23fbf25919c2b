// window.cu -- per-group running / windowed / neighbour aggregates of ALL groups and SEVERAL expressions in one
// streaming pass (the "stock / moving average" query shape: avgs(5,price), sums(5,quantity), deltas(price),
// mins(5,price), max(price - mins(price)) by sym; benchmark/quries/Aquery/q7.a, q10.a, tests/moving_avg.a).
//
// The reference evaluates every expression group by group on a gathered vector col[vecs[g]]
// (engine/ast.py:696-718; loops of server/aggregations.h:89-281, 439-462).  Here:
//
//   1. the rows are brought into CSR order ONCE, carrying their values: 16-byte tuples (gid, v0, v1, v2) go
//      through a stable LSD radix sort on gid (k_ws_hist / k_ws_scan / k_ws_scatter: one or two passes of
//      <= 11 bits).  Stability keeps the row order inside a group (ascending, or descending for the
//      reference's ht_postproc order, server/hasher.h:181-198); the second pass needs no histogram because
//      the final position of a (low digit, high digit) run IS the CSR offset of that group.
//      A handle with its own order (order_by) skips the sort: the tuples are gathered through its row list.
//   2. k_win streams the sorted tuples tile by tile (one bulk copy per tile, double buffered) and evaluates
//      every requested expression from shared memory: windows look back at most WIN_MAX_W elements into the
//      halo, running aggregates are segmented scans inside the tile plus a carry that open segments fetch from
//      the tiles before (each tile publishes its tail aggregate; no tile waits for another one's carry),
//      `v - scan(v)` and a per-group max / min of the vector are folded in (one atomic per segment and tile).
//
// Algorithmic bytes per row: 4 (gid) + 4 per value column read, outputs written once; the sort adds
// 2 x (16 + 16) bytes of tuple traffic.
#include "window.cuh"
#include "groupby.cuh"
#include "stream.cuh"
#include "sort.cuh"
#include <algorithm>
#include <cstring>

namespace aq {

// ====================================================================================================
// 1. stable radix scatter of (gid, v0, v1, v2) tuples
// ====================================================================================================
constexpr int WS_THREADS = 512;
constexpr int WS_WARPS = WS_THREADS / 32;
constexpr int WS_RPT = 8;
constexpr int WS_T = WS_THREADS * WS_RPT;         // 4096 tuples per tile
constexpr uint32_t WS_MAXBITS = 11;

struct WsArgs {
    uint32_t n, bins, shift;          // digit = (gid >> shift) & (bins - 1); bins is a power of two >= 32
    int pass;                         // 1: source = columns in table order; 2: source = tuples of pass 1
    int reverse;                      // pass 1: element i of the processing order is row n-1-i
    const uint32_t* gid;
    const uint32_t* col[WIN_MAX_COLS];
    int ncols;
    const uint4* in;
    uint4* out;
    const uint32_t* base;             // pass 1: [nchunks][bins] first output position of (chunk, digit); pass 2: CSR offsets
    const uint32_t* bstart;           // pass 2: [nchunks + 1] bucket boundaries in `in`
    uint32_t chunk_rows, nchunks;
    uint32_t b1, ngroups;             // pass 2: gid = digit << b1 | chunk
    uint32_t* ticket;
};

__device__ __forceinline__ uint32_t ws_row(const WsArgs& A, uint32_t i) { return A.reverse ? A.n - 1u - i : i; }

// per-chunk digit histogram of pass 1
__global__ void __launch_bounds__(WS_THREADS)
k_ws_hist(const WsArgs A, uint32_t* __restrict__ H) {
    extern __shared__ uint32_t ws_h[];
    for (uint32_t c = blockIdx.x; c < A.nchunks; c += gridDim.x) {
        for (uint32_t d = threadIdx.x; d < A.bins; d += WS_THREADS) ws_h[d] = 0u;
        __syncthreads();
        const uint32_t lo = c * A.chunk_rows, hi = min(A.n, lo + A.chunk_rows);
        for (uint32_t i = lo + threadIdx.x; i < hi; i += WS_THREADS)
            atomicAdd(&ws_h[(ld_stream_u32(A.gid + ws_row(A, i)) >> A.shift) & (A.bins - 1u)], 1u);
        __syncthreads();
        for (uint32_t d = threadIdx.x; d < A.bins; d += WS_THREADS) H[size_t(c) * A.bins + d] = ws_h[d];
        __syncthreads();
    }
}

// H[c][d] := first output position of (chunk c, digit d); bstart[d] := start of digit d's bucket
__global__ void __launch_bounds__(1024)
k_ws_scan(uint32_t* __restrict__ H, const uint32_t nchunks, const uint32_t bins, uint32_t* __restrict__ bstart, const uint32_t n) {
    __shared__ uint32_t tot[1u << WS_MAXBITS];
    __shared__ uint32_t wsum[32];
    for (uint32_t d = threadIdx.x; d < bins; d += 1024) {
        uint32_t run = 0;
        for (uint32_t c = 0; c < nchunks; ++c) { const uint32_t t = H[size_t(c) * bins + d]; H[size_t(c) * bins + d] = run; run += t; }
        tot[d] = run;
    }
    __syncthreads();
    // exclusive scan of tot[0..bins): PER consecutive digits per thread
    const uint32_t PER = (bins + 1023u) / 1024u;
    uint32_t loc[2] = {0, 0}, sum = 0;
    for (uint32_t k = 0; k < PER; ++k) { const uint32_t d = threadIdx.x * PER + k; loc[k] = d < bins ? tot[d] : 0u; sum += loc[k]; }
    uint32_t x = sum;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, s); if (lane >= (uint32_t)s) x += y; }
    if (lane == 31) wsum[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t v = wsum[lane], sc = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, sc, s); if (lane >= (uint32_t)s) sc += y; }
        wsum[lane] = sc - v;
    }
    __syncthreads();
    uint32_t ex = wsum[wid] + x - sum;
    for (uint32_t k = 0; k < PER; ++k) { const uint32_t d = threadIdx.x * PER + k; if (d < bins) { tot[d] = ex; bstart[d] = ex; } ex += loc[k]; }
    if (threadIdx.x == 0) bstart[bins] = n;
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < bins; d += 1024) {
        const uint32_t b = tot[d];
        for (uint32_t c = 0; c < nchunks; ++c) H[size_t(c) * bins + d] += b;
    }
}

// dynamic shared memory: stage[WS_T] uint4 | soff[bins] loff[bins] gpos[bins] u32 | whist[WS_WARPS][bins] u16
static size_t ws_smem(uint32_t bins) { return size_t(WS_T) * 16 + size_t(bins) * 12 + size_t(WS_WARPS) * bins * 2 + 256; }

__global__ void __launch_bounds__(WS_THREADS, 2)
k_ws_scatter(const WsArgs A) {
    extern __shared__ __align__(16) unsigned char ws_sm[];
    uint4* const stage = reinterpret_cast<uint4*>(ws_sm);
    uint32_t* const soff = reinterpret_cast<uint32_t*>(stage + WS_T);
    uint32_t* const loff = soff + A.bins;
    uint32_t* const gpos = loff + A.bins;
    unsigned short* const whist = reinterpret_cast<unsigned short*>(gpos + A.bins);
    uint32_t* const wsum = reinterpret_cast<uint32_t*>(whist + size_t(WS_WARPS) * A.bins);   // 32 words + chunk id
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const uint32_t bins = A.bins, dmask = bins - 1u;
    const uint32_t PER = (bins + WS_THREADS - 1u) / WS_THREADS;          // consecutive digits per thread (<= 4)
    for (;;) {
        __syncthreads();
        if (tid == 0) wsum[40] = atomicAdd(A.ticket, 1u);
        __syncthreads();
        const uint32_t c = wsum[40];
        if (c >= A.nchunks) return;
        uint32_t lo, hi;
        if (A.pass == 1) { lo = c * A.chunk_rows; hi = min(A.n, lo + A.chunk_rows); }
        else { lo = A.bstart[c]; hi = A.bstart[c + 1]; }
        for (uint32_t d = tid; d < bins; d += WS_THREADS) {
            if (A.pass == 1) soff[d] = A.base[size_t(c) * bins + d];
            else { const uint32_t g = (d << A.b1) | c; soff[d] = g < A.ngroups ? A.base[g] : 0u; }
        }
        for (uint32_t i = tid; i < WS_WARPS * bins / 2u; i += WS_THREADS) reinterpret_cast<uint32_t*>(whist)[i] = 0u;
        __syncthreads();
        for (uint32_t tlo = lo; tlo < hi; tlo += WS_T) {
            const uint32_t nvalid = min((uint32_t)WS_T, hi - tlo);
            // ---- digits of this warp's 256 consecutive elements, stable rank inside the warp's share
            uint32_t dg[WS_RPT], rk[WS_RPT];
#pragma unroll
            for (int r = 0; r < WS_RPT; ++r) {
                const uint32_t e = wid * (32u * WS_RPT) + r * 32u + lane;
                dg[r] = 0xffffffffu;
                if (e < nvalid) {
                    const uint32_t g = A.pass == 1 ? ld_stream_u32(A.gid + ws_row(A, tlo + e)) : __ldg(reinterpret_cast<const uint32_t*>(A.in + tlo + e));
                    dg[r] = (g >> A.shift) & dmask;
                }
            }
#pragma unroll
            for (int r = 0; r < WS_RPT; ++r) {
                const uint32_t peers = __match_any_sync(0xffffffffu, dg[r]);
                const uint32_t before = __popc(peers & ((1u << lane) - 1u));
                uint32_t old = 0;
                if (before == 0 && dg[r] != 0xffffffffu) {
                    unsigned short* hcell = whist + size_t(wid) * bins + dg[r];
                    old = *hcell;
                    *hcell = (unsigned short)(old + __popc(peers));
                }
                old = __shfl_sync(0xffffffffu, old, __ffs(peers) - 1);
                rk[r] = old + before;
                __syncwarp();
            }
            __syncthreads();
            // ---- per digit: exclusive prefix over the warps, tile count; exclusive scan of the counts over the digits
            uint32_t cnt[4] = {0, 0, 0, 0}, sum = 0;
            for (uint32_t k = 0; k < PER; ++k) {
                const uint32_t d = tid * PER + k;
                if (d < bins) {
                    uint32_t run = 0;
#pragma unroll
                    for (int w = 0; w < WS_WARPS; ++w) { unsigned short* hcell = whist + size_t(w) * bins + d; const uint32_t t = *hcell; *hcell = (unsigned short)run; run += t; }
                    cnt[k] = run;
                }
                sum += cnt[k];
            }
            uint32_t x = sum;
#pragma unroll
            for (int s = 1; s < 32; s <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, s); if (lane >= (uint32_t)s) x += y; }
            if (lane == 31) wsum[wid] = x;
            __syncthreads();
            if (wid == 0) {
                uint32_t v = lane < WS_WARPS ? wsum[lane] : 0u, sc = v;
#pragma unroll
                for (int s = 1; s < 32; s <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, sc, s); if (lane >= (uint32_t)s) sc += y; }
                if (lane < WS_WARPS) wsum[lane] = sc - v;
            }
            __syncthreads();
            uint32_t ex = wsum[wid] + x - sum;
            for (uint32_t k = 0; k < PER; ++k) {
                const uint32_t d = tid * PER + k;
                if (d < bins) { loff[d] = ex; gpos[d] = soff[d]; soff[d] += cnt[k]; }
                ex += cnt[k];
            }
            __syncthreads();
            // ---- tuples into the stage at their sorted place
#pragma unroll
            for (int r = 0; r < WS_RPT; ++r) {
                if (dg[r] == 0xffffffffu) continue;
                const uint32_t e = wid * (32u * WS_RPT) + r * 32u + lane;
                uint4 t;
                if (A.pass == 1) {
                    const uint32_t row = ws_row(A, tlo + e);
                    t.x = ld_stream_u32(A.gid + row);
                    t.y = A.col[0] ? ld_stream_u32(A.col[0] + row) : 0u;
                    t.z = A.col[1] ? ld_stream_u32(A.col[1] + row) : 0u;
                    t.w = A.col[2] ? ld_stream_u32(A.col[2] + row) : 0u;
                } else {
                    t = __ldg(A.in + tlo + e);
                }
                stage[loff[dg[r]] + whist[size_t(wid) * bins + dg[r]] + rk[r]] = t;
            }
            __syncthreads();
            // ---- runs of equal digits out, consecutive threads = consecutive tuples
            for (uint32_t j = tid; j < nvalid; j += WS_THREADS) {
                const uint4 t = stage[j];
                const uint32_t d = (t.x >> A.shift) & dmask;
                A.out[size_t(gpos[d]) + (j - loff[d])] = t;
            }
            for (uint32_t i = tid; i < WS_WARPS * bins / 2u; i += WS_THREADS) reinterpret_cast<uint32_t*>(whist)[i] = 0u;
            __syncthreads();
        }
    }
}

// a handle with its own row order: tuples straight through its row list
__global__ void __launch_bounds__(256)
k_ws_gather(const uint32_t n, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ gid, const uint32_t* __restrict__ c0,
            const uint32_t* __restrict__ c1, const uint32_t* __restrict__ c2, uint4* __restrict__ out) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t row = perm[p];
        uint4 t;
        t.x = gid[row]; t.y = c0 ? c0[row] : 0u; t.z = c1 ? c1[row] : 0u; t.w = c2 ? c2[row] : 0u;
        out[p] = t;
    }
}

// ====================================================================================================
// 2. the streaming window kernel
// ====================================================================================================
constexpr int WK_THREADS = 1024;
constexpr int WK_EPT = 4;
constexpr int WK_T = WK_THREADS * WK_EPT;          // 4096 elements per tile
constexpr int WK_H = (int)WIN_MAX_W;               // halo before the tile
constexpr int WK_STAGE = WK_H + WK_T;              // tuples per stage
constexpr int WK_MAXRUN = 2;                       // running (unbounded) expressions per call

struct WkOp {
    int op, vtype, col, minus, reduce, out_type, run;     // run: index among the running expressions or -1
    uint32_t w;
    void* out;
};
struct WkArgs {
    uint32_t n, ntiles, ngroups;
    const uint4* tup;
    const uint32_t* offsets;
    int nops, nrun;
    WkOp ops[WIN_MAX_OPS];
    uint32_t* tflag;                  // [ntiles] 0 = not published, 1 = no head in the tile, 2 = the tile has a head
    unsigned long long* tval;         // [ntiles][WK_MAXRUN] tail aggregate (image) of every running expression
};

enum { WOP_ADD = 0, WOP_MIN = 1, WOP_MAX = 2 };
__device__ __forceinline__ int wk_kind(int op) { return (op == AQ_SUMS || op == AQ_AVGS) ? WOP_ADD : (op == AQ_MINS ? WOP_MIN : WOP_MAX); }

// accumulator images: ints as long long, floats as double (bit pattern in the 64-bit word)
template <bool FP> struct WkAcc;
template <> struct WkAcc<false> {
    typedef long long T;
    __device__ static T load(uint32_t raw) { return (long long)(int)raw; }
    __device__ static T id(int k) { return k == WOP_ADD ? 0ll : (k == WOP_MIN ? 0x7fffffffffffffffLL : (long long)0x8000000000000000ULL); }
    __device__ static T f(int k, T a, T b) { return k == WOP_ADD ? a + b : (k == WOP_MIN ? (b < a ? b : a) : (b > a ? b : a)); }
    __device__ static unsigned long long bits(T v) { return (unsigned long long)v; }
    __device__ static T from(unsigned long long b) { return (long long)b; }
};
template <> struct WkAcc<true> {
    typedef double T;
    __device__ static T load(uint32_t raw) { return (double)__uint_as_float(raw); }
    __device__ static T id(int k) {
        return k == WOP_ADD ? 0.0 : __longlong_as_double(k == WOP_MIN ? 0x7ff0000000000000LL : (long long)0xfff0000000000000ULL);
    }
    __device__ static T f(int k, T a, T b) { return k == WOP_ADD ? a + b : (k == WOP_MIN ? (b < a ? b : a) : (b > a ? b : a)); }
    __device__ static unsigned long long bits(T v) { return (unsigned long long)__double_as_longlong(v); }
    __device__ static T from(unsigned long long b) { return __longlong_as_double((long long)b); }
};

struct alignas(16) WkI128 { unsigned long long lo; long long hi; };

struct WkShared {
    uint64_t full[2];
    uint32_t wflag[WK_THREADS / 32];
    unsigned long long wval[WK_THREADS / 32];
    uint32_t open_len;                            // elements of the tile's first group that lie before the tile
    uint32_t tile_has_head;
    unsigned long long carry[WK_MAXRUN];
    unsigned long long tail[WK_MAXRUN];
    uint32_t whead[WK_THREADS / 32];
};

// Block-wide segmented exclusive scan of one (flag, value) pair per thread, threads in element order.
// flag: the thread's elements contain a head; val: aggregate of its elements after its last head (all of them
// if it has none).  Returns for this thread: ex_flag = a head lies in an earlier thread, ex_val = aggregate of the
// elements since that head (or since the tile start).  *tot_flag / *tot_val: the same over the whole block.
template <bool FP>
__device__ __forceinline__ void wk_block_segscan(int kind, bool flag, typename WkAcc<FP>::T val, bool& ex_flag, typename WkAcc<FP>::T& ex_val,
                                                 bool& tot_flag, typename WkAcc<FP>::T& tot_val, WkShared& S) {
    typedef WkAcc<FP> C;
    typedef typename C::T T;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    bool f = flag; T v = val;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const bool f2 = __shfl_up_sync(0xffffffffu, (int)f, d) != 0;
        const unsigned long long b2 = __shfl_up_sync(0xffffffffu, C::bits(v), d);
        if (lane >= (uint32_t)d) { if (!f) v = C::f(kind, C::from(b2), v); f = f || f2; }
    }
    __syncthreads();                              // S.wflag / S.wval free again
    if (lane == 31) { S.wflag[wid] = f ? 1u : 0u; S.wval[wid] = C::bits(v); }
    __syncthreads();
    if (wid == 0) {
        bool wf = S.wflag[lane] != 0u; T wv = C::from(S.wval[lane]);
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const bool f2 = __shfl_up_sync(0xffffffffu, (int)wf, d) != 0;
            const unsigned long long b2 = __shfl_up_sync(0xffffffffu, C::bits(wv), d);
            if (lane >= (uint32_t)d) { if (!wf) wv = C::f(kind, C::from(b2), wv); wf = wf || f2; }
        }
        S.wflag[lane] = wf ? 1u : 0u; S.wval[lane] = C::bits(wv);
    }
    __syncthreads();
    bool ef = __shfl_up_sync(0xffffffffu, (int)f, 1) != 0;
    unsigned long long eb = __shfl_up_sync(0xffffffffu, C::bits(v), 1);
    T ev = C::from(eb);
    if (lane == 0) { ef = false; ev = C::id(kind); }
    if (wid > 0) {
        const bool pf = S.wflag[wid - 1] != 0u; const T pv = C::from(S.wval[wid - 1]);
        if (!ef) ev = C::f(kind, pv, ev);
        ef = ef || pf;
    }
    ex_flag = ef; ex_val = ev;
    tot_flag = S.wflag[WK_THREADS / 32 - 1] != 0u; tot_val = C::from(S.wval[WK_THREADS / 32 - 1]);
}

// ---- per-thread view of a tile: WK_EPT consecutive elements, everything in registers ---------------------
struct WkThread {
    uint32_t base;                 // index of the thread's first element in the SoA arrays (halo included)
    uint32_t p0;                   // its CSR position
    uint32_t nval;                 // how many of its elements exist (ragged last tile)
    bool head[WK_EPT];
    uint32_t idx[WK_EPT];          // index inside the group's row list
    bool last[WK_EPT];             // last element of its segment inside this tile
    uint32_t gid[WK_EPT];
};

template <class T> __device__ __forceinline__ T wk_val(uint32_t raw);
template <> __device__ __forceinline__ int wk_val<int>(uint32_t raw) { return (int)raw; }
template <> __device__ __forceinline__ float wk_val<float>(uint32_t raw) { return __uint_as_float(raw); }
template <class T> struct WkLong { typedef long long type; };
template <> struct WkLong<float> { typedef double type; };

// WK_EPT consecutive result cells of one thread: vector stores when all four exist
__device__ __forceinline__ unsigned long long wk_bits64(double v) { return (unsigned long long)__double_as_longlong(v); }
__device__ __forceinline__ unsigned long long wk_bits64(long long v) { return (unsigned long long)v; }
__device__ __forceinline__ uint32_t wk_bits32(int v) { return (uint32_t)v; }
__device__ __forceinline__ uint32_t wk_bits32(float v) { return __float_as_uint(v); }
template <class V> __device__ __forceinline__ void wk_store4_32(V* dst, const V (&x)[WK_EPT], uint32_t nval) {
    if (nval == WK_EPT) *reinterpret_cast<uint4*>(dst) = make_uint4(wk_bits32(x[0]), wk_bits32(x[1]), wk_bits32(x[2]), wk_bits32(x[3]));
    else {
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) if ((uint32_t)j < nval) dst[j] = x[j];
    }
}
template <class V> __device__ __forceinline__ void wk_store4_64(V* dst, const V (&x)[WK_EPT], uint32_t nval) {
    if (nval == WK_EPT) {
        reinterpret_cast<ulonglong2*>(dst)[0] = make_ulonglong2(wk_bits64(x[0]), wk_bits64(x[1]));
        reinterpret_cast<ulonglong2*>(dst)[1] = make_ulonglong2(wk_bits64(x[2]), wk_bits64(x[3]));
    } else {
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) if ((uint32_t)j < nval) dst[j] = x[j];
    }
}
__device__ __forceinline__ void wk_store4(int* dst, const int (&x)[WK_EPT], uint32_t nval) { wk_store4_32(dst, x, nval); }
__device__ __forceinline__ void wk_store4(float* dst, const float (&x)[WK_EPT], uint32_t nval) { wk_store4_32(dst, x, nval); }
__device__ __forceinline__ void wk_store4(double* dst, const double (&x)[WK_EPT], uint32_t nval) { wk_store4_64(dst, x, nval); }
__device__ __forceinline__ void wk_store4(long long* dst, const long long (&x)[WK_EPT], uint32_t nval) { wk_store4_64(dst, x, nval); }
// sums of 32-bit integers as __int128 cells (sign-extended)
__device__ __forceinline__ void wk_store4_i128(WkI128* dst, const long long (&x)[WK_EPT], uint32_t nval) {
#pragma unroll
    for (int j = 0; j < WK_EPT; ++j)
        if ((uint32_t)j < nval) reinterpret_cast<ulonglong2*>(dst)[j] = make_ulonglong2((unsigned long long)x[j], (unsigned long long)(x[j] >> 63));
}

// per-group max / min of the thread values r[] (int32): one atomic per segment and tile
__device__ __forceinline__ void wk_reduce_tile(const WkOp& o, const WkThread& t, const int (&r)[WK_EPT], WkShared& S) {
    const int kind = o.reduce == AQ_MAX ? WOP_MAX : WOP_MIN;
    typedef WkAcc<false> C;
    long long x[WK_EPT];
    bool any = false;
    long long run = C::id(kind);
#pragma unroll
    for (int j = 0; j < WK_EPT; ++j) {
        if ((uint32_t)j < t.nval) {
            const long long e = (long long)r[j];
            if (t.head[j]) { run = e; any = true; } else run = C::f(kind, run, e);
        }
        x[j] = run;
    }
    bool ef, tf; long long ev, tv;
    wk_block_segscan<false>(kind, any, run, ef, ev, tf, tv, S);
    bool seen = false;
#pragma unroll
    for (int j = 0; j < WK_EPT; ++j) {
        if ((uint32_t)j >= t.nval) break;
        if (t.head[j]) seen = true;
        if (!t.last[j]) continue;
        const long long v = seen ? x[j] : C::f(kind, ev, x[j]);
        int* dst = static_cast<int*>(o.out) + t.gid[j];
        if (kind == WOP_MAX) atomicMax(dst, (int)v); else atomicMin(dst, (int)v);
    }
}

// one running expression over the tile: scan, carry, output (or a per-group reduce)
template <bool FP>
__device__ __forceinline__ void wk_running(const WkOp& o, const WkThread& t, const uint32_t (&cur)[WK_EPT], WkShared& S, bool first_is_head) {
    typedef WkAcc<FP> C;
    typedef typename C::T T;
    const int kind = wk_kind(o.op);
    T x[WK_EPT];
    bool any = false;
    T run = C::id(kind);
#pragma unroll
    for (int j = 0; j < WK_EPT; ++j) {
        if ((uint32_t)j < t.nval) {
            const T e = C::load(cur[j]);
            if (t.head[j]) { run = e; any = true; } else run = C::f(kind, run, e);
        }
        x[j] = run;
    }
    bool ef, tf; T ev, tv;
    wk_block_segscan<FP>(kind, any, run, ef, ev, tf, tv, S);
    const T carry = C::from(S.carry[o.run]);
    const T prefix = ef ? ev : (first_is_head ? ev : C::f(kind, carry, ev));   // elements before the tile's first head continue the open segment
    bool seen = false;
#pragma unroll
    for (int j = 0; j < WK_EPT; ++j) { if (t.head[j]) seen = true; if (!seen) x[j] = C::f(kind, prefix, x[j]); }
    if (o.op == AQ_SUMS) {
        if (FP) {
            double r[WK_EPT];
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) r[j] = (double)x[j];
            wk_store4(static_cast<double*>(o.out) + t.p0, r, t.nval);
        }
        else {
            long long r[WK_EPT];
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) r[j] = (long long)x[j];
            if (o.out_type == AQ_INT128) wk_store4_i128(static_cast<WkI128*>(o.out) + t.p0, r, t.nval);
            else wk_store4(static_cast<long long*>(o.out) + t.p0, r, t.nval);
        }
    } else if (o.op == AQ_AVGS) {
        double r[WK_EPT];
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) r[j] = (double)x[j] / (double)(t.idx[j] + 1u);     // aggregations.h:226
        wk_store4(static_cast<double*>(o.out) + t.p0, r, t.nval);
    } else if (FP) {
        float r[WK_EPT];
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) {
            float f = (float)x[j];
            if (o.op == AQ_MAXS && !(f > 1.17549435e-38f)) f = 1.17549435e-38f;             // aggregations.h:111
            r[j] = o.minus ? __uint_as_float(cur[j]) - f : f;
        }
        wk_store4(static_cast<float*>(o.out) + t.p0, r, t.nval);
    } else {
        int r[WK_EPT];
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) { const int m = (int)(long long)x[j]; r[j] = o.minus ? (int)cur[j] - m : m; }
        if (o.reduce >= 0) wk_reduce_tile(o, t, r, S);
        else wk_store4(static_cast<int*>(o.out) + t.p0, r, t.nval);
    }
}

// windowed / neighbour expressions over the thread's elements.  v[]: the column's values at base-8 .. base+3
// (own elements at 8..11); longer windows read shared memory (colp = the column's SoA array).
constexpr int WK_REGW = 9;            // windows up to this length are served from registers
template <class T>
__device__ __forceinline__ void wk_windowed(const WkOp& o, const WkThread& t, const uint32_t (&v)[12], const uint32_t* colp, WkShared& S) {
    typedef typename WkLong<T>::type L;
    constexpr bool FP = sizeof(T) == 4 && ((T)0.5 != (T)0);
    T res[WK_EPT];
    switch (o.op) {
    case AQ_SUMW: case AQ_AVGW: {          // aggregations.h:238-249, 258-272: the last min(w, idx+1) elements
        L ws[WK_EPT]; uint32_t cnt[WK_EPT];
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) {
            cnt[j] = min(o.w, t.idx[j] + 1u);
            ws[j] = 0;
            if (o.w <= (uint32_t)WK_REGW) {
#pragma unroll
                for (int k = WK_REGW - 1; k >= 0; --k) if ((uint32_t)k < cnt[j]) ws[j] += (L)wk_val<T>(v[8 + j - k]);     // oldest first
            } else {
                for (uint32_t k = cnt[j]; k > 0; --k) ws[j] += (L)wk_val<T>(colp[t.base + j + 1 - k]);
            }
        }
        if (o.op == AQ_AVGW) {
            double r[WK_EPT];
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) r[j] = (double)ws[j] / (double)cnt[j];
            wk_store4(static_cast<double*>(o.out) + t.p0, r, t.nval);
        } else if (FP) {
            double r[WK_EPT];
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) r[j] = (double)ws[j];
            wk_store4(static_cast<double*>(o.out) + t.p0, r, t.nval);
        } else {
            long long r[WK_EPT];
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) r[j] = (long long)ws[j];
            if (o.out_type == AQ_INT128) wk_store4_i128(static_cast<WkI128*>(o.out) + t.p0, r, t.nval);
            else wk_store4(static_cast<long long*>(o.out) + t.p0, r, t.nval);
        }
        return; }
    case AQ_MINW: case AQ_MAXW: {          // aggregations.h:127-167: extreme of the last w elements
        const bool is_min = o.op == AQ_MINW;
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) {
            const uint32_t cnt = min(o.w, t.idx[j] + 1u);
            T m;
            if (o.w <= (uint32_t)WK_REGW) {
                m = wk_val<T>(v[8 + j]);
                // oldest to newest like the reference's deque: ties keep the newer element, nan never wins a `<`
                bool have = false;
#pragma unroll
                for (int k = WK_REGW - 1; k >= 0; --k) {
                    if ((uint32_t)k >= cnt) continue;
                    const T e = wk_val<T>(v[8 + j - k]);
                    if (!have) { m = e; have = true; }
                    else if (is_min ? (e < m) : !(e < m)) m = e;
                }
            } else {
                m = wk_val<T>(colp[t.base + j + 1 - cnt]);
                for (uint32_t k = cnt - 1u; k > 0; --k) { const T e = wk_val<T>(colp[t.base + j + 1 - k]); if (is_min ? (e < m) : !(e < m)) m = e; }
            }
            res[j] = m;
        }
        break; }
    case AQ_DELTAS:
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) res[j] = t.idx[j] ? (T)(wk_val<T>(v[8 + j]) - wk_val<T>(v[7 + j])) : (T)0;       // aggregations.h:442-446
        break;
    case AQ_PREV:
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) res[j] = t.idx[j] ? wk_val<T>(v[7 + j]) : wk_val<T>(v[8 + j]);                    // aggregations.h:458-462
        break;
    default: return;
    }
    if (o.minus) {
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) res[j] = (T)(wk_val<T>(v[8 + j]) - res[j]);
    }
    if (o.reduce >= 0) {
        int r[WK_EPT];
#pragma unroll
        for (int j = 0; j < WK_EPT; ++j) r[j] = (int)res[j];
        wk_reduce_tile(o, t, r, S);
    } else wk_store4(static_cast<T*>(o.out) + t.p0, res, t.nval);
}

// dynamic shared memory: the tile as it arrives (16-byte tuples) | the same as four word arrays (gid, v0, v1, v2)
constexpr int WK_SOA = WK_STAGE + 4;               // words per array (16-byte aligned rows, one element of lookahead)
constexpr size_t WK_SMEM = size_t(WK_STAGE + 1) * 16 + size_t(4) * WK_SOA * 4;

__global__ void __launch_bounds__(WK_THREADS, 1)
k_win(const WkArgs A) {
    extern __shared__ __align__(128) unsigned char wk_sm[];
    __shared__ WkShared S;
    uint4* const aos = reinterpret_cast<uint4*>(wk_sm);
    uint32_t* const soa = reinterpret_cast<uint32_t*>(aos + (WK_STAGE + 1));
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    if (tid == 0) { mbar_init(&S.full[0], 1); mbar_fence_init(); }
    __syncthreads();
    auto issue = [&](uint32_t tile) {
        // elements [tile*WK_T - WK_H, tile*WK_T + WK_T] (halo, tile, one element of lookahead), clipped to [0, n)
        const long long first = (long long)tile * WK_T - WK_H;
        long long last = (long long)tile * WK_T + WK_T + 1;
        if (last > (long long)A.n) last = (long long)A.n;
        const long long lo = first < 0 ? 0 : first;
        const uint32_t bytes = (uint32_t)(last - lo) * 16u;
        mbar_arrive_expect_tx(&S.full[0], bytes);
        bulk_g2s(aos + (lo - first), A.tup + lo, bytes, &S.full[0]);
    };
    if (tid == 0 && blockIdx.x < A.ntiles) issue(blockIdx.x);
    uint32_t k = 0;
    for (uint32_t tile = blockIdx.x; tile < A.ntiles; tile += gridDim.x, ++k) {
        mbar_wait(&S.full[0], k & 1u);
        const uint32_t tile_lo = tile * WK_T;
        const uint32_t nvalid = min((uint32_t)WK_T, A.n - tile_lo);
        // ---- tuples -> word arrays (conflict-free on both sides); then the arrival buffer is free for the next tile
        {
            const uint32_t e_lo = tile_lo == 0u ? (uint32_t)WK_H : 0u;
            const uint32_t e_hi = WK_H + nvalid + ((tile_lo + nvalid < A.n) ? 1u : 0u);
            for (uint32_t e = e_lo + tid; e < e_hi; e += WK_THREADS) {
                const uint4 t4 = aos[e];
                soa[e] = t4.x; soa[WK_SOA + e] = t4.y; soa[2 * WK_SOA + e] = t4.z; soa[3 * WK_SOA + e] = t4.w;
            }
        }
        __syncthreads();
        if (tid == 0 && tile + gridDim.x < A.ntiles) issue(tile + gridDim.x);
        // ---- heads and the index of every element inside its group
        WkThread t;
        t.base = WK_H + tid * WK_EPT;
        t.p0 = tile_lo + tid * WK_EPT;
        t.nval = tid * WK_EPT < nvalid ? min((uint32_t)WK_EPT, nvalid - tid * WK_EPT) : 0u;
        {
            const uint4 g4 = *reinterpret_cast<const uint4*>(soa + t.base);
            t.gid[0] = g4.x; t.gid[1] = g4.y; t.gid[2] = g4.z; t.gid[3] = g4.w;
            const uint32_t gprev = (t.p0 == 0u) ? ~g4.x : soa[t.base - 1];
            const uint32_t gnext = soa[t.base + WK_EPT];
            int lasth = -1;
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) {
                const uint32_t pj = j ? t.gid[j - 1] : gprev;
                t.head[j] = (uint32_t)j < t.nval && t.gid[j] != pj;
                if (t.head[j]) lasth = (int)(tid * WK_EPT) + j;
                const uint32_t nj = j + 1 < WK_EPT ? t.gid[j + 1] : gnext;
                t.last[j] = (uint32_t)j < t.nval && ((tid * WK_EPT + j + 1 == nvalid) || nj != t.gid[j]);
            }
            // exclusive max-scan of the last head index over the threads
            int m = lasth;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, m, d); if (lane >= (uint32_t)d) m = max(m, y); }
            if (lane == 31) S.whead[wid] = (uint32_t)m;
            if (tid == 0) S.open_len = tile_lo - A.offsets[soa[WK_H]];
            __syncthreads();
            if (wid == 0) {
                int v = (int)S.whead[lane];
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, v, d); if (lane >= (uint32_t)d) v = max(v, y); }
                S.whead[lane] = (uint32_t)v;             // inclusive over the warps
            }
            __syncthreads();
            int cur = __shfl_up_sync(0xffffffffu, m, 1);
            if (lane == 0) cur = -1;
            if (wid > 0) cur = max(cur, (int)S.whead[wid - 1]);
            const uint32_t open_len = S.open_len;
#pragma unroll
            for (int j = 0; j < WK_EPT; ++j) {
                const uint32_t i = tid * WK_EPT + j;
                if (t.head[j]) cur = (int)i;
                t.idx[j] = cur >= 0 ? i - (uint32_t)cur : open_len + i;
            }
        }
        const bool tile_has_head = (int)S.whead[WK_THREADS / 32 - 1] >= 0;
        const bool first_is_head = (tile_lo == 0u) || (soa[WK_H] != soa[WK_H - 1]);
        // ---- running expressions: publish this tile's tail aggregates, fetch the carry of an open first segment
        if (A.nrun > 0) {
            for (int q = 0; q < A.nops; ++q) {
                const WkOp& o = A.ops[q];
                if (o.run < 0) continue;
                const int kind = wk_kind(o.op);
                const uint4 c4 = *reinterpret_cast<const uint4*>(soa + (1 + o.col) * WK_SOA + t.base);
                const uint32_t cur[WK_EPT] = {c4.x, c4.y, c4.z, c4.w};
                unsigned long long tailbits;
                if (o.vtype == AQ_FLOAT) {
                    typedef WkAcc<true> C;
                    bool any = false; double run = C::id(kind);
#pragma unroll
                    for (int j = 0; j < WK_EPT; ++j) {
                        if ((uint32_t)j >= t.nval) break;
                        const double e = C::load(cur[j]);
                        if (t.head[j]) { run = e; any = true; } else run = C::f(kind, run, e);
                    }
                    bool ef, tf; double ev, tv;
                    wk_block_segscan<true>(kind, any, run, ef, ev, tf, tv, S);
                    tailbits = C::bits(tv);
                } else {
                    typedef WkAcc<false> C;
                    bool any = false; long long run = C::id(kind);
#pragma unroll
                    for (int j = 0; j < WK_EPT; ++j) {
                        if ((uint32_t)j >= t.nval) break;
                        const long long e = C::load(cur[j]);
                        if (t.head[j]) { run = e; any = true; } else run = C::f(kind, run, e);
                    }
                    bool ef, tf; long long ev, tv;
                    wk_block_segscan<false>(kind, any, run, ef, ev, tf, tv, S);
                    tailbits = C::bits(tv);
                }
                if (tid == 0) A.tval[size_t(tile) * WK_MAXRUN + o.run] = tailbits;
            }
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(A.tflag + tile) = tile_has_head ? 2u : 1u;
                // carry = aggregate of the open segment's elements in the tiles before this one
                for (int r = 0; r < A.nrun; ++r) S.carry[r] = 0ull;
                if (!first_is_head && tile > 0) {
                    bool init[WK_MAXRUN] = {false, false};
                    for (long long pt = (long long)tile - 1; pt >= 0; --pt) {
                        uint32_t fl;
                        uint32_t spins = 0;
                        do { fl = *reinterpret_cast<volatile uint32_t*>(A.tflag + pt); if (++spins > (1u << 28)) __trap(); } while (fl == 0u);
                        __threadfence();
                        for (int q = 0; q < A.nops; ++q) {
                            const WkOp& o = A.ops[q];
                            if (o.run < 0) continue;
                            const unsigned long long b = *reinterpret_cast<volatile unsigned long long*>(A.tval + size_t(pt) * WK_MAXRUN + o.run);
                            const int kind = wk_kind(o.op);
                            if (!init[o.run]) { S.carry[o.run] = b; init[o.run] = true; }
                            else if (o.vtype == AQ_FLOAT) S.carry[o.run] = WkAcc<true>::bits(WkAcc<true>::f(kind, WkAcc<true>::from(b), WkAcc<true>::from(S.carry[o.run])));
                            else S.carry[o.run] = WkAcc<false>::bits(WkAcc<false>::f(kind, WkAcc<false>::from(b), WkAcc<false>::from(S.carry[o.run])));
                        }
                        if (fl == 2u) break;               // that tile's tail starts at a head: the segment begins there
                    }
                }
            }
            __syncthreads();
        }
        // ---- the expressions
        for (int q = 0; q < A.nops; ++q) {
            const WkOp& o = A.ops[q];
            const uint32_t* colp = soa + (1 + o.col) * WK_SOA;
            const uint4 c4 = *reinterpret_cast<const uint4*>(colp + t.base);
            if (o.run >= 0) {
                const uint32_t cur[WK_EPT] = {c4.x, c4.y, c4.z, c4.w};
                if (o.vtype == AQ_FLOAT) wk_running<true>(o, t, cur, S, first_is_head);
                else wk_running<false>(o, t, cur, S, first_is_head);
            } else {
                const uint4 p1 = *reinterpret_cast<const uint4*>(colp + t.base - 4), p2 = *reinterpret_cast<const uint4*>(colp + t.base - 8);
                const uint32_t v[12] = {p2.x, p2.y, p2.z, p2.w, p1.x, p1.y, p1.z, p1.w, c4.x, c4.y, c4.z, c4.w};
                if (o.vtype == AQ_FLOAT) wk_windowed<float>(o, t, v, colp, S);
                else wk_windowed<int>(o, t, v, colp, S);
            }
        }
        __syncthreads();                                   // the word arrays are rewritten by the next iteration
    }
}

template <class T>
__global__ void k_fill(T* __restrict__ p, T v, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = v;
}

// ====================================================================================================
// 1b. ASSUMING order inside the call: every group's tuples sorted by the key they carry (field w)
// ====================================================================================================
// One CTA per group: the tuples and a composite (key image << 32 | position) go to shared memory, a bitonic network
// sorts the composites (the position makes the order total, i.e. the sort stable), the tuples are written back in that
// order.  Replaces, for groups of up to SEG_MAX rows, the 7 global radix passes of aqgpu_groups_order_by AND the random
// gathers through the resulting row list (k_ws_gather: 96 bytes of DRAM traffic per 4-byte cell).
constexpr uint32_t SEG_MAX = 8192;
constexpr int SEG_THREADS = 256;
__global__ void k_seg_maxlen(const uint32_t* __restrict__ offsets, uint32_t ngroups, uint32_t* __restrict__ out) {
    uint32_t m = 0;
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) m = max(m, offsets[g + 1] - offsets[g]);
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}
__global__ void __launch_bounds__(SEG_THREADS)
k_seg_sort(uint4* __restrict__ tup, const uint32_t* __restrict__ offsets, const uint32_t ngroups, const uint32_t cap,
           const uint32_t xor_mask, const int float_flip) {
    extern __shared__ __align__(16) unsigned char seg_sm[];
    uint4* const st = reinterpret_cast<uint4*>(seg_sm);                                   // cap tuples
    unsigned long long* const comp = reinterpret_cast<unsigned long long*>(st + cap);   // cap composites
    for (uint32_t g = blockIdx.x; g < ngroups; g += gridDim.x) {
        const uint32_t lo = offsets[g], n = offsets[g + 1] - lo;
        if (n < 2u) continue;                                   // (uniform per CTA)
        uint32_t m = 64;                                        // >= 64: a warp's 32 pairs span 64 elements
        while (m < n) m <<= 1;
        for (uint32_t i = threadIdx.x; i < m; i += SEG_THREADS) {
            unsigned long long c = ~0ull;
            if (i < n) {
                const uint4 t = tup[size_t(lo) + i];
                st[i] = t;
                uint32_t k = t.w;
                if (float_flip) k ^= (uint32_t)((int)k >> 31) | 0x80000000u;
                c = ((unsigned long long)(k ^ xor_mask) << 32) | i;
            }
            comp[i] = c;
        }
        __syncthreads();
        for (uint32_t k = 2; k <= m; k <<= 1) {
            uint32_t j = k >> 1;
            for (; j > 32u; j >>= 1) {                           // partners in other warps' elements: CTA-wide steps
                for (uint32_t p = threadIdx.x; p < (m >> 1); p += SEG_THREADS) {
                    const uint32_t i = ((p & ~(j - 1u)) << 1) | (p & (j - 1u));
                    const uint32_t l = i | j;
                    const unsigned long long a = comp[i], b = comp[l];
                    const bool up = (i & k) == 0u;
                    if ((a > b) == up) { comp[i] = b; comp[l] = a; }
                }
                __syncthreads();
            }
            // j <= 32: the 32 consecutive pairs of a warp stay inside one aligned block of 64 elements -> warp-level steps
            for (uint32_t p = threadIdx.x; p < (m >> 1); p += SEG_THREADS) {
                for (uint32_t jj = j; jj > 0; jj >>= 1) {
                    const uint32_t i = ((p & ~(jj - 1u)) << 1) | (p & (jj - 1u));
                    const uint32_t l = i | jj;
                    const unsigned long long a = comp[i], b = comp[l];
                    const bool up = (i & k) == 0u;
                    if ((a > b) == up) { comp[i] = b; comp[l] = a; }
                    __syncwarp();
                }
            }
            __syncthreads();
        }
        for (uint32_t i = threadIdx.x; i < n; i += SEG_THREADS) tup[size_t(lo) + i] = st[(uint32_t)comp[i]];
        __syncthreads();
    }
}

// ====================================================================================================
// host
// ====================================================================================================
static bool win_supported(const WinOp& o) {
    if (o.vtype != AQ_INT32 && o.vtype != AQ_FLOAT) return false;
    switch (o.op) {
    case AQ_SUMS: case AQ_AVGS: if (o.minus || o.reduce >= 0) return false; break;
    case AQ_MINS: case AQ_MAXS: break;
    case AQ_SUMW: case AQ_AVGW: if (o.minus || o.reduce >= 0 || o.w == 0 || o.w > WIN_MAX_W) return false; break;
    case AQ_MINW: case AQ_MAXW: if (o.w == 0 || o.w > WIN_MAX_W) return false; break;
    case AQ_DELTAS: case AQ_PREV: break;
    default: return false;
    }
    if (o.reduce >= 0 && (o.vtype != AQ_INT32 || (o.reduce != AQ_MAX && o.reduce != AQ_MIN))) return false;
    return true;
}

int window_fused(const WinGroups& G, const WinOp* ops, int n_ops, bool* handled) {
    Context& c = ctx();
    *handled = false;
    if (n_ops < 1 || n_ops > WIN_MAX_OPS || G.nrows < 2 || G.ngroups == 0) return 0;
    if (path_disabled("win")) return 0;
    // ---- value columns
    const uint32_t* cols[WIN_MAX_COLS] = {nullptr, nullptr, nullptr};
    int ncols = 0, nrun = 0;
    WkArgs K;
    memset(&K, 0, sizeof(K));
    for (int q = 0; q < n_ops; ++q) {
        if (!win_supported(ops[q])) return 0;
        int ci = -1;
        for (int i = 0; i < ncols; ++i) if (cols[i] == ops[q].v) ci = i;
        if (ci < 0) { if (ncols == WIN_MAX_COLS) return 0; cols[ncols] = static_cast<const uint32_t*>(ops[q].v); ci = ncols++; }
        WkOp& o = K.ops[q];
        o.op = ops[q].op; o.vtype = ops[q].vtype; o.col = ci; o.minus = ops[q].minus; o.reduce = ops[q].reduce;
        o.out_type = ops[q].out_type; o.w = ops[q].w; o.out = ops[q].out; o.run = -1;
        if (o.op == AQ_SUMS || o.op == AQ_AVGS || o.op == AQ_MINS || o.op == AQ_MAXS) { if (nrun == WK_MAXRUN) return 0; o.run = nrun++; }
    }
    int bits = 0;
    while ((uint64_t(1) << bits) < uint64_t(G.ngroups)) ++bits;
    if (bits > 2 * (int)WS_MAXBITS) return 0;
    const uint32_t n = G.nrows;
    uint32_t seg_cap = 0;
    if (G.order_key) {                                          // ASSUMING order inside the call
        if (G.perm || ncols >= WIN_MAX_COLS) return 0;          // the key needs the third tuple field
        DevBuf mx;
        AQ_TRY(mx.alloc_zero(4));
        AQ_LAUNCH(k_seg_maxlen, grid_for(G.ngroups, 256, 4), 256, 0, G.offsets, G.ngroups, mx.as<uint32_t>());
        uint32_t longest = 0;
        AQ_CUDA(cudaMemcpyAsync(&longest, mx.p, 4, cudaMemcpyDeviceToHost, c.stream));
        AQ_CUDA(cudaStreamSynchronize(c.stream));
        if (longest > SEG_MAX) return 0;                        // a group too long for one SM's shared memory: the caller sorts globally
        for (seg_cap = 256; seg_cap < longest; seg_cap <<= 1) {}
        cols[WIN_MAX_COLS - 1] = G.order_key;
    }
    // ---- 1. tuples in CSR order
    DevBuf ta, tb, hist, bstart, tick;
    AQ_TRY(ta.alloc(size_t(n) * 16 + 16));
    AQ_TRY(tick.alloc_zero(16));
    const uint4* sorted = nullptr;
    if (G.perm) {
        AQ_LAUNCH(k_ws_gather, grid_for(n, 256, 1, 32), 256, 0, n, G.perm, G.gid, cols[0], cols[1], cols[2], ta.as<uint4>());
        sorted = ta.as<uint4>();
    } else {
        WsArgs W;
        memset(&W, 0, sizeof(W));
        W.n = n; W.gid = G.gid; W.ncols = G.order_key ? WIN_MAX_COLS : ncols; W.ngroups = G.ngroups;
        for (int i = 0; i < WIN_MAX_COLS; ++i) W.col[i] = cols[i];
        W.reverse = G.row_order == AQ_ROWS_DESC;
        const int npass = bits > (int)WS_MAXBITS ? 2 : 1;
        const uint32_t b1 = npass == 2 ? (uint32_t)(bits + 1) / 2u : (uint32_t)std::max(bits, 5);
        const uint32_t b2 = npass == 2 ? (uint32_t)std::max(bits - (int)b1, 5) : 0u;
        // pass 1: digit = low b1 bits, chunks = contiguous shares of the processing order
        W.pass = 1; W.bins = 1u << b1; W.shift = 0;
        const uint32_t want_chunks = (uint32_t)c.sm_count * 8u;
        uint32_t chunk_rows = (n + want_chunks - 1) / want_chunks;
        chunk_rows = std::max<uint32_t>((chunk_rows + WS_T - 1) / WS_T * WS_T, WS_T);
        W.chunk_rows = chunk_rows; W.nchunks = (n + chunk_rows - 1) / chunk_rows;
        AQ_TRY(hist.alloc(size_t(W.nchunks) * W.bins * 4));
        AQ_TRY(bstart.alloc((size_t(W.bins) + 1) * 4));
        if (npass == 2) AQ_TRY(tb.alloc(size_t(n) * 16 + 16));
        static size_t attr = 0;
        const size_t smem1 = ws_smem(W.bins), smem2 = npass == 2 ? ws_smem(1u << b2) : 0;
        if (std::max(smem1, smem2) > attr) {
            AQ_CUDA(cudaFuncSetAttribute(k_ws_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(smem1, smem2)));
            attr = std::max(smem1, smem2);
        }
        AQ_LAUNCH(k_ws_hist, (int)std::min<uint32_t>(W.nchunks, (uint32_t)c.sm_count * 4), WS_THREADS, size_t(W.bins) * 4, W, hist.as<uint32_t>());
        AQ_LAUNCH(k_ws_scan, 1, 1024, 0, hist.as<uint32_t>(), W.nchunks, W.bins, bstart.as<uint32_t>(), n);
        W.base = hist.as<uint32_t>(); W.ticket = tick.as<uint32_t>();
        W.out = npass == 2 ? tb.as<uint4>() : ta.as<uint4>();
        AQ_LAUNCH(k_ws_scatter, (int)std::min<uint32_t>(W.nchunks, (uint32_t)c.sm_count * 2), WS_THREADS, smem1, W);
        if (npass == 2) {
            // pass 2: chunks = the buckets of pass 1, digit = the high bits; a run's place is its group's CSR offset
            WsArgs V = W;
            V.pass = 2; V.in = tb.as<uint4>(); V.out = ta.as<uint4>(); V.bins = 1u << b2; V.shift = b1; V.b1 = b1;
            V.nchunks = 1u << b1; V.bstart = bstart.as<uint32_t>(); V.base = G.offsets; V.ticket = tick.as<uint32_t>() + 1;
            AQ_LAUNCH(k_ws_scatter, (int)std::min<uint32_t>(V.nchunks, (uint32_t)c.sm_count * 2), WS_THREADS, smem2, V);
        }
        sorted = ta.as<uint4>();
        if (G.order_key) {
            static size_t seg_attr = 0;
            const size_t seg_smem = size_t(seg_cap) * 24;
            if (seg_smem > seg_attr) { AQ_CUDA(cudaFuncSetAttribute(k_seg_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)seg_smem)); seg_attr = seg_smem; }
            const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (c.smem_optin - 1024) / (seg_smem + 1024)));
            AQ_LAUNCH(k_seg_sort, (int)std::min<uint32_t>(G.ngroups, (uint32_t)(c.sm_count * per_sm * 4)), SEG_THREADS, seg_smem, ta.as<uint4>(), G.offsets,
                      G.ngroups, seg_cap, G.order_xor, G.order_float);
        }
    }
    // ---- 2. the expressions
    K.n = n; K.ntiles = (n + WK_T - 1) / WK_T; K.ngroups = G.ngroups; K.tup = sorted; K.offsets = G.offsets;
    K.nops = n_ops; K.nrun = nrun;
    DevBuf tflag, tval;
    AQ_TRY(tflag.alloc_zero(size_t(K.ntiles) * 4));
    AQ_TRY(tval.alloc(size_t(K.ntiles) * WK_MAXRUN * 8));
    K.tflag = tflag.as<uint32_t>(); K.tval = tval.as<unsigned long long>();
    for (int q = 0; q < n_ops; ++q) {
        if (ops[q].reduce < 0) continue;
        const int idv = ops[q].reduce == AQ_MAX ? (int)0x80000000u : 0x7fffffff;
        AQ_LAUNCH(k_fill<int>, grid_for(G.ngroups, 256, 1, 8), 256, 0, static_cast<int*>(ops[q].out), idv, G.ngroups);
    }
    static bool attr_done = false;
    if (!attr_done) { AQ_CUDA(cudaFuncSetAttribute(k_win, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WK_SMEM)); attr_done = true; }
    AQ_LAUNCH(k_win, (int)std::min<uint32_t>(K.ntiles, (uint32_t)c.sm_count), WK_THREADS, WK_SMEM, K);
    *handled = true;
    return 0;
}

} // namespace aq
