// vecops.cu -- element-wise vector arithmetic and comparison: the operators the generated code applies
// to columns and to vector-valued aggregate results, e.g. `price - mins(price)` (tests/stock.a:22-35) or
// `sums(deltas(_time) > 120)` used as a group key (tests/network.a:8-11).
//
// Reference (server/vector_type.hpp:372-396, server/table.h:826-973):
//   * every cell is computed as the C++ expression `lhs[i] op rhs[i]` -- i.e. in the USUAL ARITHMETIC CONVERSION type of
//     the two element types (int32 op int32 wraps in 32 bits, int / int is INTEGER division, int64 op float is float
//     arithmetic) -- and then converted to the element type of the result container;
//   * the result container's element type depends on the operand FORM (probed with the reference build, the table is
//     pinned by tests/test_gpu_vecops.py against oracle/_ref):
//       vector op vector  (member templates, vector_type.hpp:386-396)   + - * /  -> Coercion<T1, T2>
//       vector op scalar, scalar op vector (table.h:826-925)             + -      -> Coercion<T1, T2>
//                                                                        *        -> GetLongType<Coercion>  (__int128 | double)
//                                                                        /        -> GetFPType<Coercion>    (float for int32)
//       comparisons (vector_type.hpp:372-385, table.h:927-973 aqop_*)             -> bool (1-byte cells)
//     Coercion (server/types.h:264-275) for {int32, int64, float, double}: the wider type; equal widths: the floating one;
//     any floating operand makes it GetFPType of that (so int64 with float is DOUBLE, although the cell was computed in
//     float);
//   * aqop_x(lhs, rhs, ret) (table.h:949-973, what the emitter produces at the root of an expression,
//     engine/expr.py:373-380) writes into a container whose element type the CALLER chose: out_type is explicit.
// Waived quirk: the reference's member comparison `vector < scalar` (and <=, >=, ==, !=) all evaluate `>`
// (vector_type.hpp:372-385 expands `container[i] > v` for every operator) -- an evident slip; here every comparison means
// what it says, and `>` agrees with the reference.  Integer division by zero traps in the reference; here the cell is 0.
//
// One kernel, templated on the compute type only; operand / result cell types are uniform run-time switches (the kernel
// moves 8..32 bytes per row and is bound by HBM).
#include "common.cuh"
#include <cfloat>
#include <cmath>
#include <type_traits>

namespace aq {

struct VecArgs {
    int op;
    int ta, tb, tout;
    const void* a; const void* b;            // null = scalar operand
    long long sa_i, sb_i; double sa_f, sb_f; // scalar operands, already converted to the compute type
    void* out;
    uint32_t n;
};

struct I128v { unsigned long long lo; long long hi; };

template <class C> struct VecIsInt { static constexpr bool v = false; };
template <> struct VecIsInt<int> { static constexpr bool v = true; };
template <> struct VecIsInt<long long> { static constexpr bool v = true; };

template <class C> __device__ __forceinline__ C vec_load(const void* p, int t, uint32_t i) {
    switch (t) {
    case AQ_INT32: return (C)static_cast<const int*>(p)[i];
    case AQ_INT64: return (C)static_cast<const long long*>(p)[i];
    case AQ_FLOAT: return (C)static_cast<const float*>(p)[i];
    case AQ_DOUBLE: return (C)static_cast<const double*>(p)[i];
    default: return (C)static_cast<const unsigned char*>(p)[i];          // AQ_BOOL
    }
}
template <class C> __device__ __forceinline__ void vec_store(void* out, int t, uint32_t i, C r) {
    switch (t) {
    case AQ_INT32: static_cast<int*>(out)[i] = (int)r; break;
    case AQ_INT64: static_cast<long long*>(out)[i] = (long long)r; break;
    case AQ_FLOAT: static_cast<float*>(out)[i] = (float)r; break;
    case AQ_DOUBLE: static_cast<double*>(out)[i] = (double)r; break;
    case AQ_INT128: { const long long v = (long long)r; I128v x; x.lo = (unsigned long long)v; x.hi = v >> 63; static_cast<I128v*>(out)[i] = x; break; }
    default: static_cast<unsigned char*>(out)[i] = r != (C)0 ? 1 : 0; break;   // AQ_BOOL
    }
}
// wrapping integer arithmetic (signed overflow is undefined in the reference; two's complement here)
__device__ __forceinline__ int wrap_add(int x, int y) { return (int)((unsigned)x + (unsigned)y); }
__device__ __forceinline__ int wrap_sub(int x, int y) { return (int)((unsigned)x - (unsigned)y); }
__device__ __forceinline__ int wrap_mul(int x, int y) { return (int)((unsigned)x * (unsigned)y); }
__device__ __forceinline__ long long wrap_add(long long x, long long y) { return (long long)((unsigned long long)x + (unsigned long long)y); }
__device__ __forceinline__ long long wrap_sub(long long x, long long y) { return (long long)((unsigned long long)x - (unsigned long long)y); }
__device__ __forceinline__ long long wrap_mul(long long x, long long y) { return (long long)((unsigned long long)x * (unsigned long long)y); }
__device__ __forceinline__ float wrap_add(float x, float y) { return __fadd_rn(x, y); }
__device__ __forceinline__ float wrap_sub(float x, float y) { return __fsub_rn(x, y); }
__device__ __forceinline__ float wrap_mul(float x, float y) { return __fmul_rn(x, y); }
__device__ __forceinline__ double wrap_add(double x, double y) { return __dadd_rn(x, y); }
__device__ __forceinline__ double wrap_sub(double x, double y) { return __dsub_rn(x, y); }
__device__ __forceinline__ double wrap_mul(double x, double y) { return __dmul_rn(x, y); }
__device__ __forceinline__ int vdiv(int x, int y) { return y == 0 ? 0 : (y == -1 ? wrap_sub(0, x) : x / y); }
__device__ __forceinline__ long long vdiv(long long x, long long y) { return y == 0 ? 0 : (y == -1 ? wrap_sub(0ll, x) : x / y); }
__device__ __forceinline__ float vdiv(float x, float y) { return __fdiv_rn(x, y); }
__device__ __forceinline__ double vdiv(double x, double y) { return __ddiv_rn(x, y); }

template <class C>
__global__ void __launch_bounds__(256) k_vec_binary(const VecArgs A) {
    const C sa = VecIsInt<C>::v ? (C)A.sa_i : (C)A.sa_f;
    const C sb = VecIsInt<C>::v ? (C)A.sb_i : (C)A.sb_f;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < A.n; i += gridDim.x * blockDim.x) {
        const C x = A.a ? vec_load<C>(A.a, A.ta, i) : sa;
        const C y = A.b ? vec_load<C>(A.b, A.tb, i) : sb;
        if (A.op >= AQ_V_GT) {
            bool r;
            switch (A.op) {
            case AQ_V_GT: r = x > y; break;
            case AQ_V_LT: r = x < y; break;
            case AQ_V_GTE: r = x >= y; break;
            case AQ_V_LTE: r = x <= y; break;
            case AQ_V_EQ: r = x == y; break;
            default: r = x != y; break;
            }
            vec_store<int>(A.out, A.tout, i, r ? 1 : 0);
            continue;
        }
        C r;
        switch (A.op) {
        case AQ_V_ADD: r = wrap_add(x, y); break;
        case AQ_V_SUB: r = wrap_sub(x, y); break;
        case AQ_V_MUL: r = wrap_mul(x, y); break;
        case AQ_V_DIV: r = vdiv(x, y); break;
        default:
            if constexpr (VecIsInt<C>::v) r = A.op == AQ_V_AND ? (x & y) : (A.op == AQ_V_OR ? (x | y) : (x ^ y));
            else r = (C)0;
            break;
        }
        vec_store<C>(A.out, A.tout, i, r);
    }
}

// out[j] = col[idx[j]] for cells of 1, 2, 4, 8, 12 or 16 bytes; an index beyond the column raises the flag (cell left zero)
template <class T>
__global__ void __launch_bounds__(256)
k_take(const T* __restrict__ col, const uint32_t rows, const uint32_t* __restrict__ idx, const uint32_t n, T* __restrict__ out, uint32_t* __restrict__ bad) {
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const uint32_t i = idx[j];
        if (i < rows) out[j] = col[i];
        else { out[j] = T{}; atomicExch(bad, 1u); }
    }
}
struct Cell12 { uint32_t w[3]; };

// sqrt(v) -> double cells; truncate(v, precision) -> cells of v's type (server/aggregations.h:34-69).
//   sqrt: the reference's `ret[i] = sqrt(v[i])` resolves to the float overload for float cells (the root is taken in float, then
//         widened) and to sqrt(double) for integer cells -- probed with the reference build, pinned by tests/test_gpu_vecops.py;
//   truncate: v[i] < max(T) / 10^p ? round(v[i] * 10^p) / 10^p : v[i], evaluated in double (the multiplier is a double), rounded
//         half away from zero; a copy when aq_fp_precision<T> (float 7, double 16, integers 0; server/types.h:473-475) <= p.
template <class T>
__global__ void __launch_bounds__(256)
k_vec_unary(const int op, const T* __restrict__ x, const double mult, const double max_trunc, void* __restrict__ out, const uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const T v = x[i];
        if (op == AQ_U_SQRT) {
            double r;
            if constexpr (std::is_same<T, float>::value) r = (double)__fsqrt_rn(v);
            else r = __dsqrt_rn((double)v);
            static_cast<double*>(out)[i] = r;
        } else {
            T r = v;
            if constexpr (std::is_floating_point<T>::value)
                if (mult > 0.0 && (double)v < max_trunc) r = (T)__ddiv_rn(round(__dmul_rn((double)v, mult)), mult);
            static_cast<T*>(out)[i] = r;
        }
    }
}

// 0 int32, 1 int64, 2 float, 3 double
inline int vec_rank(int t) { return t == AQ_INT32 || t == AQ_BOOL ? 0 : t == AQ_INT64 ? 1 : t == AQ_FLOAT ? 2 : t == AQ_DOUBLE ? 3 : -1; }
inline int vec_compute_rank(int ra, int rb) {            // usual arithmetic conversions
    if (ra == 3 || rb == 3) return 3;
    if (ra == 2 || rb == 2) return 2;
    return ra == 1 || rb == 1 ? 1 : 0;
}
inline int vec_coercion(int ra, int rb) {                // types::Coercion<T1, T2> (server/types.h:264-275) as a dtype code
    const bool fp = ra >= 2 || rb >= 2;
    const bool wide = ra == 1 || ra == 3 || rb == 1 || rb == 3;
    return fp ? (wide ? AQ_DOUBLE : AQ_FLOAT) : (wide ? AQ_INT64 : AQ_INT32);
}

} // namespace aq

using namespace aq;

extern "C" int aqgpu_vec_out_type(int op, int a_type, int a_scalar, int b_type, int b_scalar) {
    const int ra = vec_rank(a_type), rb = vec_rank(b_type);
    if (ra < 0 || rb < 0 || op < AQ_V_ADD || op > AQ_V_NEQ) return -1;
    if (op >= AQ_V_GT) return AQ_BOOL;
    const int k = vec_coercion(ra, rb);
    if (op >= AQ_V_AND) return (k == AQ_INT32 || k == AQ_INT64) ? k : -1;
    if (!a_scalar && !b_scalar) return k;
    if (op == AQ_V_MUL) return (k == AQ_INT32 || k == AQ_INT64) ? AQ_INT128 : AQ_DOUBLE;
    if (op == AQ_V_DIV) return k == AQ_INT32 ? AQ_FLOAT : (k == AQ_INT64 ? AQ_DOUBLE : k);
    return k;
}

extern "C" int aqgpu_vec_binary(int op, int a_type, const void* a, int a_scalar, int b_type, const void* b, int b_scalar,
                                int out_type, void* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    const int ra = vec_rank(a_type), rb = vec_rank(b_type);
    if (ra < 0 || rb < 0) return fail("vec_binary: operand dtypes %d, %d (int32, int64, float, double, bool)", a_type, b_type);
    if (op < AQ_V_ADD || op > AQ_V_NEQ) return fail("vec_binary: unknown op %d", op);
    if (a_scalar && b_scalar) return fail("vec_binary: at least one operand must be a vector");
    if (out_type < 0) out_type = aqgpu_vec_out_type(op, a_type, a_scalar, b_type, b_scalar);
    const size_t osz = out_type == AQ_BOOL ? 1 : dtype_size(out_type);
    if (out_type < 0 || !osz || out_type == AQ_UINT32) return fail("vec_binary: no result type for op %d on dtypes %d, %d", op, a_type, b_type);
    const int rc = vec_compute_rank(ra, rb);
    if (op >= AQ_V_AND && op <= AQ_V_XOR && rc >= 2) return fail("vec_binary: bitwise op %d on floating-point operands", op);
    if (n == 0) return 0;
    if (!a || !b || !out) return fail("vec_binary: null argument");
    auto cell = [](int t) { return t == AQ_BOOL ? size_t(1) : dtype_size(t); };
    VecArgs A{};
    A.op = op; A.ta = a_type; A.tb = b_type; A.tout = out_type; A.n = n;
    InCol ia, ib; OutCol oc;
    auto scalar = [&](const void* p, int t, long long& si, double& sf) {      // C++ conversion of the scalar to the compute type
        long long vi = 0; double vf = 0; bool is_fp = false;
        switch (t) {
        case AQ_INT32: vi = *static_cast<const int*>(p); break;
        case AQ_INT64: vi = *static_cast<const long long*>(p); break;
        case AQ_FLOAT: vf = *static_cast<const float*>(p); is_fp = true; break;
        case AQ_DOUBLE: vf = *static_cast<const double*>(p); is_fp = true; break;
        default: vi = *static_cast<const unsigned char*>(p) ? 1 : 0; break;
        }
        if (rc == 2) sf = is_fp ? (double)(float)vf : (double)(float)vi;
        else if (rc == 3) sf = is_fp ? vf : (double)vi;
        else si = is_fp ? (long long)vf : (rc == 0 ? (long long)(int)vi : vi);
    };
    if (a_scalar) { if (is_device_ptr(a)) return fail("vec_binary: a scalar operand is a host cell"); scalar(a, a_type, A.sa_i, A.sa_f); }
    else { AQ_TRY(ia.bind(a, cell(a_type) * n)); A.a = ia.d; }
    if (b_scalar) { if (is_device_ptr(b)) return fail("vec_binary: a scalar operand is a host cell"); scalar(b, b_type, A.sb_i, A.sb_f); }
    else { AQ_TRY(ib.bind(b, cell(b_type) * n)); A.b = ib.d; }
    AQ_TRY(oc.bind(out, osz * n));
    A.out = oc.d;
    const int grid = grid_for(n, 256, 1, 16);
    switch (rc) {
    case 0: AQ_LAUNCH(k_vec_binary<int>, grid, 256, 0, A); break;
    case 1: AQ_LAUNCH(k_vec_binary<long long>, grid, 256, 0, A); break;
    case 2: AQ_LAUNCH(k_vec_binary<float>, grid, 256, 0, A); break;
    default: AQ_LAUNCH(k_vec_binary<double>, grid, 256, 0, A); break;
    }
    AQ_TRY(oc.finish(osz * n));
    if (oc.host) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

extern "C" int aqgpu_take(int elem_bytes, const void* col, uint32_t col_rows, const uint32_t* idx, uint32_t nidx, void* out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4 && elem_bytes != 8 && elem_bytes != 12 && elem_bytes != 16)
        return fail("take: %d-byte cells (1, 2, 4, 8, 12 or 16)", elem_bytes);
    if (nidx == 0) return 0;
    if (!idx || !out || (col_rows && !col)) return fail("take: null argument");
    if (col_rows == 0) return fail("take: %u indices into an empty column", nidx);
    InCol ic, ii; OutCol oc;
    AQ_TRY(ic.bind(col, size_t(elem_bytes) * col_rows));
    AQ_TRY(ii.bind(idx, size_t(nidx) * 4));
    AQ_TRY(oc.bind(out, size_t(elem_bytes) * nidx));
    DevBuf bad;
    AQ_TRY(bad.alloc_zero(16));
    const int grid = grid_for(nidx, 256, 1, 16);
    const uint32_t* di = static_cast<const uint32_t*>(ii.d);
    switch (elem_bytes) {
    case 1: AQ_LAUNCH(k_take<unsigned char>, grid, 256, 0, static_cast<const unsigned char*>(ic.d), col_rows, di, nidx, static_cast<unsigned char*>(oc.d), bad.as<uint32_t>()); break;
    case 2: AQ_LAUNCH(k_take<unsigned short>, grid, 256, 0, static_cast<const unsigned short*>(ic.d), col_rows, di, nidx, static_cast<unsigned short*>(oc.d), bad.as<uint32_t>()); break;
    case 4: AQ_LAUNCH(k_take<uint32_t>, grid, 256, 0, static_cast<const uint32_t*>(ic.d), col_rows, di, nidx, static_cast<uint32_t*>(oc.d), bad.as<uint32_t>()); break;
    case 8: AQ_LAUNCH(k_take<unsigned long long>, grid, 256, 0, static_cast<const unsigned long long*>(ic.d), col_rows, di, nidx, static_cast<unsigned long long*>(oc.d), bad.as<uint32_t>()); break;
    case 12: AQ_LAUNCH(k_take<Cell12>, grid, 256, 0, static_cast<const Cell12*>(ic.d), col_rows, di, nidx, static_cast<Cell12*>(oc.d), bad.as<uint32_t>()); break;
    default: AQ_LAUNCH(k_take<uint4>, grid, 256, 0, static_cast<const uint4*>(ic.d), col_rows, di, nidx, static_cast<uint4*>(oc.d), bad.as<uint32_t>()); break;
    }
    uint32_t flag = 0;
    AQ_CUDA(cudaMemcpyAsync(&flag, bad.p, 4, cudaMemcpyDeviceToHost, ctx().stream));
    AQ_TRY(oc.finish(size_t(elem_bytes) * nidx));
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    if (flag) return fail("take: an index is beyond the column's %u rows", col_rows);
    return 0;
}

extern "C" int aqgpu_vec_unary(int op, int type, const void* x, uint32_t param, void* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (op != AQ_U_SQRT && op != AQ_U_TRUNCATE) return fail("vec_unary: unknown op %d", op);
    if (!is_val_type(type)) return fail("vec_unary: dtype %d (int32, int64, float, double)", type);
    if (n == 0) return 0;
    if (!x || !out) return fail("vec_unary: null argument");
    const size_t isz = dtype_size(type), osz = op == AQ_U_SQRT ? 8 : isz;
    InCol ic; OutCol oc;
    AQ_TRY(ic.bind(x, isz * n));
    AQ_TRY(oc.bind(out, osz * n));
    // truncate: a copy when the type's precision (float 7, double 16, integers 0) does not exceed the requested one
    const uint32_t prec = type == AQ_FLOAT ? 7u : (type == AQ_DOUBLE ? 16u : 0u);
    double mult = 0.0, max_trunc = 0.0;
    if (op == AQ_U_TRUNCATE && prec > param) {
        mult = std::pow(10.0, (double)param);
        max_trunc = (type == AQ_FLOAT ? (double)FLT_MAX : DBL_MAX) / mult;
    }
    const int grid = grid_for(n, 256, 1, 16);
    switch (type) {
    case AQ_INT32: AQ_LAUNCH(k_vec_unary<int>, grid, 256, 0, op, static_cast<const int*>(ic.d), mult, max_trunc, oc.d, n); break;
    case AQ_INT64: AQ_LAUNCH(k_vec_unary<long long>, grid, 256, 0, op, static_cast<const long long*>(ic.d), mult, max_trunc, oc.d, n); break;
    case AQ_FLOAT: AQ_LAUNCH(k_vec_unary<float>, grid, 256, 0, op, static_cast<const float*>(ic.d), mult, max_trunc, oc.d, n); break;
    default: AQ_LAUNCH(k_vec_unary<double>, grid, 256, 0, op, static_cast<const double*>(ic.d), mult, max_trunc, oc.d, n); break;
    }
    AQ_TRY(oc.finish(osz * n));
    if (oc.host) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}
