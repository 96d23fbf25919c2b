// emit.cuh -- finalisation of one requested aggregate from a group's accumulator cells (device code shared by
// the emit kernels of groupby.cu, lc.cu, hcfp.cu): result types and formulas of server/aggregations.h.
#pragma once
#include "groupby.cuh"

namespace aq {

#ifdef __CUDACC__
struct alignas(16) I128 { unsigned long long lo; long long hi; };

static __device__ __forceinline__ double i128_to_f64(__int128 x) { return (double)x; }

static __device__ void emit_one(const EmitAgg& e, const unsigned long long* __restrict__ sl, uint32_t gid,
                         uint32_t first_row, int row_order) {
    auto cell = [&](int a) -> unsigned long long { return sl[1 + a]; };
    const bool isint = (e.vtype == AQ_INT32 || e.vtype == AQ_INT64);
    const uint32_t n = e.cnt >= 0 ? (uint32_t)cell(e.cnt) : 0u;
    switch (e.op) {
    case AQ_SUM: case AQ_SUMSQ: case AQ_SUMXY:
        if (isint) {
            const long long s = (long long)cell(e.a[0]);
            if (e.out_type == AQ_INT128) { I128 v; v.lo = (unsigned long long)s; v.hi = s >> 63; static_cast<I128*>(e.out)[gid] = v; }
            else static_cast<long long*>(e.out)[gid] = s;
        } else {
            static_cast<double*>(e.out)[gid] = __longlong_as_double((long long)cell(e.a[0]));
        }
        break;
    case AQ_AVG: {     // aggregations.h:30: sum / static_cast<double>(size)
        const double s = isint ? (double)(long long)cell(e.a[0]) : __longlong_as_double((long long)cell(e.a[0]));
        static_cast<double*>(e.out)[gid] = s / (double)n;
        break; }
    case AQ_MIN: case AQ_MAX: {
        const long long k = (long long)cell(e.a[0]);
        if (e.vtype == AQ_INT32) static_cast<int*>(e.out)[gid] = (int)k;
        else if (e.vtype == AQ_INT64) static_cast<long long*>(e.out)[gid] = k;
        else {
            double d = ordered_to_f64(k);
            if (e.vtype == AQ_FLOAT) {
                float f = (float)d;
                // aggregations.h:73: max starts at numeric_limits<T>::min() = smallest positive normal
                if (e.op == AQ_MAX && !(f > 1.17549435e-38f)) f = 1.17549435e-38f;
                static_cast<float*>(e.out)[gid] = f;
            } else {
                if (e.op == AQ_MAX && !(d > 2.2250738585072014e-308)) d = 2.2250738585072014e-308;
                static_cast<double*>(e.out)[gid] = d;
            }
        }
        break; }
    case AQ_COUNT:
        if (e.out_type == AQ_INT64) static_cast<long long*>(e.out)[gid] = (long long)n;
        else static_cast<uint32_t*>(e.out)[gid] = n;
        break;
    case AQ_FIRST: case AQ_LAST: {
        const bool want_max = (e.op == AQ_FIRST) == (row_order == AQ_ROWS_DESC);
        const uint32_t r = want_max ? (uint32_t)cell(e.maxrow) : first_row;
        if (e.vtype == AQ_INT32 || e.vtype == AQ_FLOAT) static_cast<uint32_t*>(e.out)[gid] = static_cast<const uint32_t*>(e.v)[r];
        else static_cast<unsigned long long*>(e.out)[gid] = static_cast<const unsigned long long*>(e.v)[r];
        break; }
    case AQ_SECOND: {                // element [1] of the row list; 0 for one-row groups
        const uint32_t k2 = (uint32_t)(cell(e.top2) & 0xffffffffull);
        const bool have = k2 != 0;
        const uint32_t r = row_order == AQ_ROWS_DESC ? k2 - 1u : 0xffffffffu - k2;
        if (e.vtype == AQ_INT32 || e.vtype == AQ_FLOAT) static_cast<uint32_t*>(e.out)[gid] = have ? static_cast<const uint32_t*>(e.v)[r] : 0u;
        else static_cast<unsigned long long*>(e.out)[gid] = have ? static_cast<const unsigned long long*>(e.v)[r] : 0ull;
        break; }
    case AQ_VAR: case AQ_STDDEV: {   // aggregations.h:332-348: (ssq - s*s/(len+1)) / (len+1)
        const double np1 = (double)(n + 1u);
        double r;
        if (isint) {
            const long long s = (long long)cell(e.a[0]), ssq = (long long)cell(e.a[1]);
            r = ((double)ssq - i128_to_f64((__int128)s * (__int128)s) / np1) / np1;
        } else {
            const double s = __longlong_as_double((long long)cell(e.a[0])), ssq = __longlong_as_double((long long)cell(e.a[1]));
            r = (ssq - s * s / np1) / np1;
        }
        static_cast<double*>(e.out)[gid] = e.op == AQ_STDDEV ? sqrt(r) : r;
        break; }
    case AQ_CORR: {                  // aggregations.h:400-406
        const long long sx = (long long)cell(e.a[0]), sy = (long long)cell(e.a[1]), sxy = (long long)cell(e.a[2]),
                        sx2 = (long long)cell(e.a[3]), sy2 = (long long)cell(e.a[4]);
        const __int128 len = (__int128)n;
        const double num = i128_to_f64(len * sxy) - i128_to_f64((__int128)sx * sy);
        const double dx = i128_to_f64(len * sx2) - i128_to_f64((__int128)sx * sx);
        const double dy = i128_to_f64(len * sy2) - i128_to_f64((__int128)sy * sy);
        static_cast<double*>(e.out)[gid] = num / sqrt(dx * dy);
        break; }
    default: break;
    }
}
#endif

} // namespace aq
