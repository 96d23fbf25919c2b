// scan.cu -- per-group running / windowed / neighbour aggregates over CSR row lists.
//
// Replaces the loops of reference server/aggregations.h that the generated code runs once per group
// on the gathered vector col[vecs[g]] (engine/ast.py:696-718, 720-790):
//   sums :203-210  avgs :219-228  mins :89-98  maxs :108-117      -> segmented inclusive scans
//   sumw :238-249  avgw :258-272  minw :127-138  maxw :148-159    -> windows over the last w rows
//   deltas :439-446  prev :455-462  aggnext :471-478  ratiow :169-184 -> neighbours inside a group
// for ALL groups at once.  Position p of the flat CSR output belongs to the group whose row list
// contains perm[p]; start[p] = first position of that group (computed once per handle).
//
// Segmented scan = 3 phases per level (tile aggregates -> recursive scan of the aggregates -> tile
// scans with carry); no spin-waiting between CTAs, deterministic summation order.
#include "scan.cuh"
#include <type_traits>

namespace aq {

constexpr int SS_THREADS = 256;
constexpr int SS_ITEMS = 8;
constexpr int SS_TILE = SS_THREADS * SS_ITEMS;     // 2048

enum { OP_ADD = 0, OP_MIN = 1, OP_MAX = 2 };

template <class A, int OP> struct Comb;
template <class A> struct Comb<A, OP_ADD> {
    __device__ static A id() { return A(0); }
    __device__ static A f(A a, A b) { return a + b; }
};
template <> struct Comb<long long, OP_MIN> {
    __device__ static long long id() { return 0x7fffffffffffffffLL; }
    __device__ static long long f(long long a, long long b) { return a < b ? a : b; }
};
template <> struct Comb<long long, OP_MAX> {
    __device__ static long long id() { return (long long)0x8000000000000000ULL; }
    __device__ static long long f(long long a, long long b) { return a > b ? a : b; }
};
// fp min/max follow the reference's compare-and-keep loops: `if (arr[i] < min) min = arr[i]`
template <> struct Comb<double, OP_MIN> {
    __device__ static double id() { return __longlong_as_double(0x7ff0000000000000LL); }
    __device__ static double f(double a, double b) { return b < a ? b : a; }
};
template <> struct Comb<double, OP_MAX> {
    __device__ static double id() { return __longlong_as_double((long long)0xfff0000000000000ULL); }
    __device__ static double f(double a, double b) { return b > a ? b : a; }
};

// ---- loaders / storers ---------------------------------------------------------------------------
template <class A> struct ColLoader {           // level 1: the gathered column, any value dtype
    const void* gv; int vtype;
    __device__ A operator()(uint32_t p) const {
        switch (vtype) {
        case AQ_INT32: return (A) static_cast<const int*>(gv)[p];
        case AQ_INT64: return (A) static_cast<const long long*>(gv)[p];
        case AQ_FLOAT: return (A) static_cast<const float*>(gv)[p];
        default:       return (A) static_cast<const double*>(gv)[p];
        }
    }
};
template <class A> struct SqLoader {            // x * x in the accumulator type (moments)
    const void* gv; int vtype;
    __device__ A operator()(uint32_t p) const {
        A x;
        switch (vtype) {
        case AQ_INT32: x = (A) static_cast<const int*>(gv)[p]; break;
        case AQ_INT64: x = (A) static_cast<const long long*>(gv)[p]; break;
        case AQ_FLOAT: x = (A) static_cast<const float*>(gv)[p]; break;
        default:       x = (A) static_cast<const double*>(gv)[p]; break;
        }
        return x * x;
    }
};
template <class A> struct ArrLoader {            // level >= 2: tile aggregates
    const A* a;
    __device__ A operator()(uint32_t p) const { return a[p]; }
};

enum StoreMode { ST_RAW = 0, ST_I128 = 1, ST_AVGS = 2, ST_T = 3 };
struct alignas(16) I128s { unsigned long long lo; long long hi; };

template <class A> struct OutStorer {
    void* out; int mode; int vtype; int is_max;
    __device__ void operator()(uint32_t p, A v, uint32_t start) const {
        switch (mode) {
        case ST_RAW: static_cast<A*>(out)[p] = v; break;
        case ST_I128: { I128s x; x.lo = (unsigned long long)(long long)v; x.hi = ((long long)v) >> 63; static_cast<I128s*>(out)[p] = x; break; }
        case ST_AVGS: static_cast<double*>(out)[p] = (double)v / (double)(p - start + 1u); break;   // aggregations.h:226
        default:   // ST_T: running min / max back in the column's type
            switch (vtype) {
            case AQ_INT32: static_cast<int*>(out)[p] = (int)v; break;
            case AQ_INT64: static_cast<long long*>(out)[p] = (long long)v; break;
            case AQ_FLOAT: { float f = (float)v;       // aggregations.h:111: maxs starts at numeric_limits<T>::min()
                if (is_max && !(f > 1.17549435e-38f)) f = 1.17549435e-38f;
                static_cast<float*>(out)[p] = f; break; }
            default: { double d = (double)v;
                if (is_max && !(d > 2.2250738585072014e-308)) d = 2.2250738585072014e-308;
                static_cast<double*>(out)[p] = d; break; }
            }
        }
    }
};

// ---- block-level segmented scan of (flag, value) pairs -------------------------------------------
// returns the EXCLUSIVE prefix of this thread: (any head strictly before this thread's chunk in the
// block?, combined value since that head / since the block start)
template <class A, int OP>
__device__ __forceinline__ void block_seg_exclusive(bool flag, A val, bool& ex_flag, A& ex_val,
                                                    bool* sm_flag, A* sm_val) {
    typedef Comb<A, OP> C;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    bool f = flag; A v = val;                      // inclusive within the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const bool f2 = __shfl_up_sync(0xffffffffu, (int)f, d) != 0;
        const A v2 = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) { if (!f) v = C::f(v2, v); f = f || f2; }
    }
    if (lane == 31) { sm_flag[wid] = f; sm_val[wid] = v; }
    __syncthreads();
    if (wid == 0) {
        bool wf = lane < SS_THREADS / 32 ? sm_flag[lane] : false;
        A wv = lane < SS_THREADS / 32 ? sm_val[lane] : C::id();
#pragma unroll
        for (int d = 1; d < SS_THREADS / 32; d <<= 1) {
            const bool f2 = __shfl_up_sync(0xffffffffu, (int)wf, d) != 0;
            const A v2 = __shfl_up_sync(0xffffffffu, wv, d);
            if (lane >= d) { if (!wf) wv = C::f(v2, wv); wf = wf || f2; }
        }
        if (lane < SS_THREADS / 32) { sm_flag[lane] = wf; sm_val[lane] = wv; }
    }
    __syncthreads();
    // exclusive of this thread inside its warp
    bool ef = __shfl_up_sync(0xffffffffu, (int)f, 1) != 0;
    A ev = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) { ef = false; ev = C::id(); }
    if (wid > 0) {
        const bool pf = sm_flag[wid - 1]; const A pv = sm_val[wid - 1];
        if (!ef) ev = C::f(pv, ev);
        ef = ef || pf;
    }
    ex_flag = ef; ex_val = ev;
    __syncthreads();
}

// Phase A: aggregate of the tile's tail segment + the tile index where that segment starts.
template <class A, int OP, class Loader>
__global__ void __launch_bounds__(SS_THREADS)
k_seg_tile_agg(const Loader load, const uint32_t* __restrict__ start, uint32_t n,
               A* __restrict__ agg, uint32_t* __restrict__ tstart) {
    typedef Comb<A, OP> C;
    __shared__ A sm[SS_THREADS / 32];
    const uint32_t t0 = blockIdx.x * SS_TILE;
    const uint32_t last = min(t0 + SS_TILE, n) - 1u;
    const uint32_t s0 = start[last];
    const uint32_t lo = max(s0, t0);
    A v = C::id();
    for (uint32_t p = lo + threadIdx.x; p <= last; p += SS_THREADS) v = C::f(v, load(p));
    // plain block reduction (order inside the tail differs from the scan order only for fp ADD;
    // phase C recomputes nothing from it except carries, which are themselves tolerance-level)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = C::f(v, __shfl_down_sync(0xffffffffu, v, d));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        A r = sm[0];
        for (int w = 1; w < SS_THREADS / 32; ++w) r = C::f(r, sm[w]);
        agg[blockIdx.x] = r;
        tstart[blockIdx.x] = s0 / SS_TILE;
    }
}

// Phase C: in-tile segmented inclusive scan with the carry of earlier tiles.
template <class A, int OP, class Loader, class Storer>
__global__ void __launch_bounds__(SS_THREADS)
k_seg_tile_scan(const Loader load, const uint32_t* __restrict__ start, uint32_t n,
                const A* __restrict__ chain, const Storer store) {
    typedef Comb<A, OP> C;
    __shared__ bool sm_flag[SS_THREADS / 32];
    __shared__ A sm_val[SS_THREADS / 32];
    const uint32_t t0 = blockIdx.x * SS_TILE;
    const uint32_t p0 = t0 + threadIdx.x * SS_ITEMS;
    A x[SS_ITEMS]; uint32_t st[SS_ITEMS];
    bool any_head = false;
    A run = C::id();
#pragma unroll
    for (int i = 0; i < SS_ITEMS; ++i) {
        const uint32_t p = p0 + i;
        if (p < n) {
            st[i] = start[p];
            const A e = load(p);
            if (st[i] == p) { run = e; any_head = true; } else run = C::f(run, e);
            x[i] = run;                    // inclusive inside the thread since its last head
        } else { st[i] = p; x[i] = C::id(); }
    }
    bool ef; A ev;
    block_seg_exclusive<A, OP>(any_head, run, ef, ev, sm_flag, sm_val);
    // carry from earlier tiles applies while no head has been seen in this tile
    A carry = C::id();
    const bool tile_open = blockIdx.x > 0 && start[t0] < t0;
    if (tile_open) carry = chain[blockIdx.x - 1];
    A prefix = ef ? ev : (tile_open ? C::f(carry, ev) : ev);
    bool seen = false;
#pragma unroll
    for (int i = 0; i < SS_ITEMS; ++i) {
        const uint32_t p = p0 + i;
        if (p >= n) break;
        if (st[i] == p) seen = true;
        const A v = seen ? x[i] : C::f(prefix, x[i]);
        store(p, v, st[i]);
    }
}

template <class A, int OP, class Loader, class Storer>
int seg_scan(uint32_t n, const uint32_t* d_start, const Loader& load, const Storer& store) {
    if (n == 0) return 0;
    const uint32_t ntiles = (n + SS_TILE - 1) / SS_TILE;
    if (ntiles == 1) {
        AQ_LAUNCH((k_seg_tile_scan<A, OP, Loader, Storer>), 1, SS_THREADS, 0, load, d_start, n, (const A*)nullptr, store);
        return 0;
    }
    DevBuf agg, tstart, chain;
    AQ_TRY(agg.alloc(size_t(ntiles) * sizeof(A)));
    AQ_TRY(tstart.alloc(size_t(ntiles) * 4));
    AQ_TRY(chain.alloc(size_t(ntiles) * sizeof(A)));
    AQ_LAUNCH((k_seg_tile_agg<A, OP, Loader>), ntiles, SS_THREADS, 0, load, d_start, n, agg.as<A>(), tstart.as<uint32_t>());
    ArrLoader<A> l2{agg.as<A>()};
    OutStorer<A> s2{chain.p, ST_RAW, 0, 0};
    AQ_TRY((seg_scan<A, OP, ArrLoader<A>, OutStorer<A>>(ntiles, tstart.as<uint32_t>(), l2, s2)));
    AQ_LAUNCH((k_seg_tile_scan<A, OP, Loader, Storer>), ntiles, SS_THREADS, 0, load, d_start, n, chain.as<A>(), store);
    return 0;
}

// ---- segment starts ------------------------------------------------------------------------------
__global__ void k_mark_heads(const uint32_t* __restrict__ offsets, uint32_t ngroups, uint32_t* __restrict__ start) {
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
        const uint32_t o = offsets[g];
        start[o] = o;
    }
}
// every position is its own "segment" for this helper scan: a plain running max over the marks
__global__ void k_iota_zero(uint32_t* __restrict__ z, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) z[i] = 0;
}
struct MarkLoader { const uint32_t* m; __device__ long long operator()(uint32_t p) const { return (long long)m[p]; } };
struct StartStorer { uint32_t* out; __device__ void operator()(uint32_t p, long long v, uint32_t) const { out[p] = (uint32_t)v; } };

int compute_segment_starts(const uint32_t* d_offsets, uint32_t ngroups, uint32_t n, uint32_t* d_start) {
    if (n == 0) return 0;
    DevBuf marks, zeros;
    AQ_TRY(marks.alloc_zero(size_t(n) * 4));
    AQ_TRY(zeros.alloc_zero(size_t(n) * 4));      // "start" array of one segment covering everything
    AQ_LAUNCH(k_mark_heads, grid_for(ngroups, 256), 256, 0, d_offsets, ngroups, marks.as<uint32_t>());
    MarkLoader ld{marks.as<uint32_t>()};
    StartStorer stv{d_start};
    return seg_scan<long long, OP_MAX, MarkLoader, StartStorer>(n, zeros.as<uint32_t>(), ld, stv);
}

// ---- gather ---------------------------------------------------------------------------------------
template <class W>
__global__ void k_gather(const W* __restrict__ v, const uint32_t* __restrict__ perm, uint32_t n, W* __restrict__ out) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) out[p] = v[perm[p]];
}

int gather_rows(const void* d_v, int elem_size, const uint32_t* d_perm, uint32_t n, void* d_out) {
    if (n == 0) return 0;
    if (elem_size == 4) AQ_LAUNCH(k_gather<uint32_t>, grid_for(n, 256, 1, 32), 256, 0, static_cast<const uint32_t*>(d_v), d_perm, n, static_cast<uint32_t*>(d_out));
    else AQ_LAUNCH(k_gather<unsigned long long>, grid_for(n, 256, 1, 32), 256, 0, static_cast<const unsigned long long*>(d_v), d_perm, n, static_cast<unsigned long long*>(d_out));
    return 0;
}

// ---- windows / neighbours -------------------------------------------------------------------------
template <class T> struct LongOf { typedef long long type; };
template <> struct LongOf<float> { typedef double type; };
template <> struct LongOf<double> { typedef double type; };
template <class T> struct FPOf { typedef float type; };              // types.h:199-204 GetFPType
template <> struct FPOf<long long> { typedef double type; };
template <> struct FPOf<double> { typedef double type; };

template <class L> __device__ __forceinline__ void store_long(void* out, int out_type, uint32_t p, L v);
template <> __device__ __forceinline__ void store_long<long long>(void* out, int out_type, uint32_t p, long long v) {
    if (out_type == AQ_INT128) { I128s x; x.lo = (unsigned long long)v; x.hi = v >> 63; static_cast<I128s*>(out)[p] = x; }
    else static_cast<long long*>(out)[p] = v;
}
template <> __device__ __forceinline__ void store_long<double>(void* out, int, uint32_t p, double v) {
    static_cast<double*>(out)[p] = v;
}

// gv: gathered column; S: segmented inclusive prefix sums in LongType (SUMW / AVGW only)
template <class T>
__global__ void __launch_bounds__(256)
k_window(int op, uint32_t w, const T* __restrict__ gv, const typename LongOf<T>::type* __restrict__ S,
         const uint32_t* __restrict__ start, const uint32_t* __restrict__ seg_len, uint32_t n,
         void* __restrict__ out, int out_type) {
    typedef typename LongOf<T>::type L;
    typedef typename FPOf<T>::type FP;
    constexpr bool kIsInt = sizeof(L) == 8 && ((T)0.5 == (T)0);
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t s = start[p];
        const uint32_t i = p - s;                       // index inside the group's row list
        switch (op) {
        case AQ_SUMW: case AQ_AVGW: {
            // window = the last min(w, i+1) elements (aggregations.h:238-249, 258-272)
            L ws;
            uint32_t cnt;
            if (i < w) { ws = S[p]; cnt = i + 1; }
            else {
                cnt = w;
                if (kIsInt || w > 64) ws = S[p] - S[p - w];      // ints: exact prefix difference
                else { ws = 0; for (uint32_t j = 0; j < w; ++j) ws += (L)gv[p - w + 1 + j]; }  // fp: sum the window itself
            }
            if (op == AQ_SUMW) store_long<L>(out, out_type, p, ws);
            else static_cast<double*>(out)[p] = (double)ws / (double)cnt;
            break; }
        case AQ_MINW: case AQ_MAXW: {
            // aggregations.h:127-167: front of the monotonic deque = extreme of the last w elements
            const uint32_t lo = (i >= w) ? p - w + 1 : s;
            T m = gv[lo];
            if (op == AQ_MINW) { for (uint32_t j = lo + 1; j <= p; ++j) { const T e = gv[j]; if (e < m) m = e; } }
            else               { for (uint32_t j = lo + 1; j <= p; ++j) { const T e = gv[j]; if (!(e < m)) m = e; } }
            static_cast<T*>(out)[p] = m;
            break; }
        case AQ_DELTAS: static_cast<T*>(out)[p] = i ? (T)(gv[p] - gv[p - 1]) : (T)0; break;   // aggregations.h:442-446
        case AQ_PREV:   static_cast<T*>(out)[p] = i ? gv[p - 1] : gv[p]; break;               // aggregations.h:458-462
        case AQ_NEXT: {                                                                       // aggregations.h:474-477
            const bool is_last = (p + 1 == n) || (start[p + 1] == p + 1);
            static_cast<T*>(out)[p] = is_last ? gv[p] : gv[p + 1];
            break; }
        case AQ_RATIOW: {                                                                     // aggregations.h:169-184
            const uint32_t len = seg_len[p];
            const uint32_t we = (len <= w) ? 1u : w;
            const T den = (i < we) ? gv[s] : gv[p - we];
            static_cast<FP*>(out)[p] = gv[p] / (FP)den;
            break; }
        default: break;
        }
    }
}

__global__ void k_seg_len(const uint32_t* __restrict__ perm, const uint32_t* __restrict__ gid,
                          const uint32_t* __restrict__ offsets, uint32_t n, uint32_t* __restrict__ len) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t g = gid[perm[p]];
        len[p] = offsets[g + 1] - offsets[g];
    }
}

// running (w == 0) / sliding population variance from the segmented prefix sums of x (S) and x * x (S2):
// var = (c * sum(x^2) - sum(x)^2) / c^2 over the last c = min(w, i + 1) elements; integers in 128-bit arithmetic
// floating-point columns with short windows (w <= VAR_DIRECT_W) are evaluated from the window's own elements (two
// passes: the difference of two long prefix sums would carry the rounding of the whole prefix into a short window)
constexpr uint32_t VAR_DIRECT_W = 64;
template <class L>
__global__ void __launch_bounds__(256)
k_var_window(const uint32_t w, const int sd, const L* __restrict__ S, const L* __restrict__ S2, const uint32_t* __restrict__ start,
             const uint32_t n, double* __restrict__ out, const void* __restrict__ gv, const int vtype) {
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        const uint32_t i = p - start[p];
        const uint32_t c = (w == 0u || w > i + 1u) ? i + 1u : w;
        if (!std::is_integral<L>::value && w != 0u && w <= VAR_DIRECT_W) {
            auto at = [&](uint32_t q) -> double { return vtype == AQ_FLOAT ? (double)static_cast<const float*>(gv)[q] : static_cast<const double*>(gv)[q]; };
            double m = 0.0, q2 = 0.0;
            for (uint32_t j = p + 1u - c; j <= p; ++j) m += at(j);
            m /= (double)c;
            for (uint32_t j = p + 1u - c; j <= p; ++j) { const double d = at(j) - m; q2 += d * d; }
            q2 /= (double)c;
            out[p] = sd ? sqrt(q2) : q2;
            continue;
        }
        L s = S[p], s2 = S2[p];
        if (i >= c) { s -= S[p - c]; s2 -= S2[p - c]; }
        double v;
        if (std::is_integral<L>::value) {                                // integer sums: exact numerator
            const __int128 num = (__int128)c * (__int128)(long long)s2 - (__int128)(long long)s * (__int128)(long long)s;
            v = (double)num / ((double)c * (double)c);
        } else {
            const double ds = (double)s, ds2 = (double)s2;
            v = (ds2 - ds * ds / (double)c) / (double)c;
        }
        if (v < 0.0) v = 0.0;
        out[p] = sd ? sqrt(v) : v;
    }
}

template <class T>
static int window_t(int op, uint32_t w, const void* gv, const void* S, const uint32_t* start,
                    const uint32_t* seg_len, uint32_t n, void* out, int out_type) {
    AQ_LAUNCH(k_window<T>, grid_for(n, 256, 1, 32), 256, 0, op, w, static_cast<const T*>(gv),
              static_cast<const typename LongOf<T>::type*>(S), start, seg_len, n, out, out_type);
    return 0;
}

// Entry used by groups.cu.  d_v is the (device) value column in table order; outputs are device.
int group_scan_device(int op, uint32_t w, int vtype, const void* d_v, uint32_t n, uint32_t ngroups,
                      const uint32_t* d_gid, const uint32_t* d_offsets, const uint32_t* d_perm,
                      const uint32_t* d_start, int out_type, void* d_out) {
    if (n == 0) return 0;
    const bool isint = is_int_type(vtype);
    const int esz = (int)dtype_size(vtype);
    DevBuf gv;
    AQ_TRY(gv.alloc(size_t(n) * esz));
    AQ_TRY(gather_rows(d_v, esz, d_perm, n, gv.p));

    switch (op) {
    case AQ_SUMS: case AQ_AVGS: {
        const int mode = op == AQ_AVGS ? ST_AVGS : (out_type == AQ_INT128 ? ST_I128 : ST_RAW);
        if (isint) {
            ColLoader<long long> ld{gv.p, vtype};
            OutStorer<long long> st{d_out, mode, vtype, 0};
            return seg_scan<long long, OP_ADD, ColLoader<long long>, OutStorer<long long>>(n, d_start, ld, st);
        }
        ColLoader<double> ld{gv.p, vtype};
        OutStorer<double> st{d_out, mode, vtype, 0};
        return seg_scan<double, OP_ADD, ColLoader<double>, OutStorer<double>>(n, d_start, ld, st);
    }
    case AQ_MINS: case AQ_MAXS: {
        const int is_max = op == AQ_MAXS;
        if (isint) {
            ColLoader<long long> ld{gv.p, vtype};
            OutStorer<long long> st{d_out, ST_T, vtype, is_max};
            if (is_max) return seg_scan<long long, OP_MAX, ColLoader<long long>, OutStorer<long long>>(n, d_start, ld, st);
            return seg_scan<long long, OP_MIN, ColLoader<long long>, OutStorer<long long>>(n, d_start, ld, st);
        }
        ColLoader<double> ld{gv.p, vtype};
        OutStorer<double> st{d_out, ST_T, vtype, is_max};
        if (is_max) return seg_scan<double, OP_MAX, ColLoader<double>, OutStorer<double>>(n, d_start, ld, st);
        return seg_scan<double, OP_MIN, ColLoader<double>, OutStorer<double>>(n, d_start, ld, st);
    }
    default: break;
    }

    if (op == AQ_VARS || op == AQ_STDDEVS || op == AQ_VARW || op == AQ_STDDEVW) {
        if ((op == AQ_VARW || op == AQ_STDDEVW) && w == 0) return fail("window length 0");
        const uint32_t ww = (op == AQ_VARS || op == AQ_STDDEVS) ? 0u : w;
        const int sd = (op == AQ_STDDEVS || op == AQ_STDDEVW) ? 1 : 0;
        DevBuf S1, S2;
        AQ_TRY(S1.alloc(size_t(n) * 8));
        AQ_TRY(S2.alloc(size_t(n) * 8));
        if (isint) {
            ColLoader<long long> ld{gv.p, vtype}; SqLoader<long long> lq{gv.p, vtype};
            OutStorer<long long> st1{S1.p, ST_RAW, vtype, 0}, st2{S2.p, ST_RAW, vtype, 0};
            AQ_TRY((seg_scan<long long, OP_ADD, ColLoader<long long>, OutStorer<long long>>(n, d_start, ld, st1)));
            AQ_TRY((seg_scan<long long, OP_ADD, SqLoader<long long>, OutStorer<long long>>(n, d_start, lq, st2)));
            AQ_LAUNCH(k_var_window<long long>, grid_for(n, 256, 1, 32), 256, 0, ww, sd, S1.as<long long>(), S2.as<long long>(), d_start, n, static_cast<double*>(d_out), gv.p, vtype);
        } else {
            ColLoader<double> ld{gv.p, vtype}; SqLoader<double> lq{gv.p, vtype};
            OutStorer<double> st1{S1.p, ST_RAW, vtype, 0}, st2{S2.p, ST_RAW, vtype, 0};
            AQ_TRY((seg_scan<double, OP_ADD, ColLoader<double>, OutStorer<double>>(n, d_start, ld, st1)));
            AQ_TRY((seg_scan<double, OP_ADD, SqLoader<double>, OutStorer<double>>(n, d_start, lq, st2)));
            AQ_LAUNCH(k_var_window<double>, grid_for(n, 256, 1, 32), 256, 0, ww, sd, S1.as<double>(), S2.as<double>(), d_start, n, static_cast<double*>(d_out), gv.p, vtype);
        }
        return 0;
    }

    if (w == 0 && (op == AQ_SUMW || op == AQ_AVGW || op == AQ_MINW || op == AQ_MAXW || op == AQ_RATIOW))
        return fail("window length 0 (the reference reads ret[-1] for w == 0)");

    DevBuf S, seg_len;
    if (op == AQ_SUMW || op == AQ_AVGW) {          // segmented prefix sums in LongType
        AQ_TRY(S.alloc(size_t(n) * 8));
        if (isint) {
            ColLoader<long long> ld{gv.p, vtype};
            OutStorer<long long> st{S.p, ST_RAW, vtype, 0};
            AQ_TRY((seg_scan<long long, OP_ADD, ColLoader<long long>, OutStorer<long long>>(n, d_start, ld, st)));
        } else {
            ColLoader<double> ld{gv.p, vtype};
            OutStorer<double> st{S.p, ST_RAW, vtype, 0};
            AQ_TRY((seg_scan<double, OP_ADD, ColLoader<double>, OutStorer<double>>(n, d_start, ld, st)));
        }
    }
    if (op == AQ_RATIOW) {
        AQ_TRY(seg_len.alloc(size_t(n) * 4));
        AQ_LAUNCH(k_seg_len, grid_for(n, 256, 1, 32), 256, 0, d_perm, d_gid, d_offsets, n, seg_len.as<uint32_t>());
    }
    (void)ngroups;
    switch (vtype) {
    case AQ_INT32: return window_t<int>(op, w, gv.p, S.p, d_start, seg_len.as<uint32_t>(), n, d_out, out_type);
    case AQ_INT64: return window_t<long long>(op, w, gv.p, S.p, d_start, seg_len.as<uint32_t>(), n, d_out, out_type);
    case AQ_FLOAT: return window_t<float>(op, w, gv.p, S.p, d_start, seg_len.as<uint32_t>(), n, d_out, out_type);
    case AQ_DOUBLE: return window_t<double>(op, w, gv.p, S.p, d_start, seg_len.as<uint32_t>(), n, d_out, out_type);
    default: return fail("unsupported value dtype %d", vtype);
    }
}

} // namespace aq
