#include <fcntl.h>
#include <unistd.h>
/*
 * ref_harness.cpp -- the REFERENCE'S OWN post-processor library compiled as a CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/aqo.h).  This file contains no algorithm of its own: it is a
 * unity build that #includes the reference's sources where they lie under /root/reference
 * (never copied into this repository) and drives them exactly the way the reference's generated
 * post-processor does (engine/ast.py:664-790, saved sample mem_opt.cpp:31-63):
 *
 *   per row   g.hashtable_push(forward_as_tuple(k0[i], k1[i], ...), i)      server/hasher.h:176-179
 *   then      vecs = g.ht_postproc(n)                                       server/hasher.h:181-198
 *   per group out[i] = agg(col[vecs[i]]); GC::scratch_space->release()      engine/ast.py:720-790
 *   vector-valued outputs: cell i is a view at buf + offsets[i]             engine/ast.py:696-718
 *
 * Why a unity build: ScratchSpace::alloc is `inline` inside server/libaquery.cpp:464 but called from
 * server/vector_type.hpp:73, so separate translation units do not link; Context needs Trigger
 * complete, hence -DTHREADING.  Built by oracle/Makefile into
 * oracle/_ref/libaqref.so (git-ignored, travels to the GPU box with the snapshot).
 */
#define THREADING 1
#include "server/libaquery.cpp"
#include "server/threading.cpp"
/* server/monetdb_conn.cpp:426-443 defines these; MonetDB is not on the hot path and is absent here. */
int execTriggerPayload(void*) { return 0; }
int execTriggerPayloadCond(void*) { return 0; }

#include "server/hasher.h"
#include "server/aggregations.h"

#include <chrono>
#include <tuple>
#include <utility>
#include <cstring>

#include "aqo.h"

namespace {

GC* ensure_gc() {
    if (!GC::gc_handle) new GC();          /* the engine creates one per Context (server/libaquery.cpp) */
    return GC::gc_handle;
}

template <class T, size_t> using same_t = T;

/* hashtable_push + ht_postproc for a K-column key of element type KT. */
template <class KT, size_t... I>
int groupby_impl(std::index_sequence<I...>, const void* const* keys, uint32_t n,
                 uint32_t* ngroups, uint32_t* gid, uint32_t* offsets, uint32_t* perm) {
    typedef record<same_t<KT, I>...> rec_t;
    if (n == 0) { *ngroups = 0; if (offsets) offsets[0] = 0; return 0; }
    const KT* k[] = { static_cast<const KT*>(keys[I])... };
    AQHashTable<rec_t, transTypes<rec_t, hasher>> g{n};
    for (uint32_t i = 0; i < n; ++i)
        g.hashtable_push(std::forward_as_tuple(k[I][i]...), i);
    auto vecs = g.ht_postproc(n);
    const uint32_t ng = static_cast<uint32_t>(g.size());
    *ngroups = ng;
    if (gid) memcpy(gid, g.reversemap, sizeof(uint32_t) * n);
    if (perm) memcpy(perm, g.mapbase, sizeof(uint32_t) * n);
    if (offsets) {
        for (uint32_t j = 0; j < ng; ++j) offsets[j] = g.ht_base[j];
        offsets[ng] = n;
    }
    /* sanity: the CSR views the emitter uses are exactly (mapbase + start, count) */
    for (uint32_t j = 0; j < ng; ++j)
        if (vecs[j].container != g.mapbase + g.ht_base[j]) return -3;
    free(vecs); free(g.reversemap); free(g.ht_base);
    return 0;
}

template <class KT>
int groupby_k(int nkeys, const void* const* keys, uint32_t n,
              uint32_t* ngroups, uint32_t* gid, uint32_t* offsets, uint32_t* perm) {
    switch (nkeys) {
    case 1: return groupby_impl<KT>(std::make_index_sequence<1>{}, keys, n, ngroups, gid, offsets, perm);
    case 2: return groupby_impl<KT>(std::make_index_sequence<2>{}, keys, n, ngroups, gid, offsets, perm);
    case 3: return groupby_impl<KT>(std::make_index_sequence<3>{}, keys, n, ngroups, gid, offsets, perm);
    case 4: return groupby_impl<KT>(std::make_index_sequence<4>{}, keys, n, ngroups, gid, offsets, perm);
    case 5: return groupby_impl<KT>(std::make_index_sequence<5>{}, keys, n, ngroups, gid, offsets, perm);
    case 6: return groupby_impl<KT>(std::make_index_sequence<6>{}, keys, n, ngroups, gid, offsets, perm);
    default: return -2;
    }
}

/* Key tuples of mixed / non-integer element types (astring_view, types::date_t, types::time_t, types::timestamp_t):
 * the same hashtable_push + ht_postproc, the reference's own hashes (server/hasher.h:97-132) and operator==. */
template <class... KT, size_t... I>
int groupby_mixed(std::index_sequence<I...>, const void* const* keys, uint32_t n,
                  uint32_t* ngroups, uint32_t* gid, uint32_t* offsets, uint32_t* perm) {
    typedef record<KT...> rec_t;
    if (n == 0) { *ngroups = 0; if (offsets) offsets[0] = 0; return 0; }
    std::tuple<const KT*...> k{ static_cast<const KT*>(keys[I])... };
    AQHashTable<rec_t, transTypes<rec_t, hasher>> g{n};
    for (uint32_t i = 0; i < n; ++i)
        g.hashtable_push(std::forward_as_tuple(std::get<I>(k)[i]...), i);
    auto vecs = g.ht_postproc(n);
    const uint32_t ng = static_cast<uint32_t>(g.size());
    *ngroups = ng;
    if (gid) memcpy(gid, g.reversemap, sizeof(uint32_t) * n);
    if (perm) memcpy(perm, g.mapbase, sizeof(uint32_t) * n);
    if (offsets) {
        for (uint32_t j = 0; j < ng; ++j) offsets[j] = g.ht_base[j];
        offsets[ng] = n;
    }
    free(vecs); free(g.reversemap); free(g.ht_base);
    return 0;
}
static_assert(sizeof(astring_view) == sizeof(const char*), "a string key column is an array of char pointers");
static_assert(sizeof(types::date_t) == 4 && sizeof(types::time_t) == 8 && sizeof(types::timestamp_t) == 12, "key cell sizes");

/* The emitted per-group loop: gather through the row list, hand the gathered vector to f. */
template <class T, class F>
void per_group(const void* v, uint32_t ng, const uint32_t* off, const uint32_t* perm, F f) {
    ensure_gc();
    ColRef<T> col(ng ? off[ng] : 0, const_cast<void*>(v));
    GC::scratch_space = GC::gc_handle ? &(GC::gc_handle->scratch) : nullptr;
    for (uint32_t g = 0; g < ng; ++g) {
        vector_type<uint32_t> val;
        val.init_from(off[g + 1] - off[g], const_cast<uint32_t*>(perm + off[g]));
        f(g, col, val);
        GC::scratch_space->release();
    }
    GC::scratch_space = nullptr;
}

template <class T>
int reduce_t(int op, const void* v, const void* v2, uint32_t ng, const uint32_t* off,
             const uint32_t* perm, void* out) {
    typedef types::GetLongType<T> LongT;
    switch (op) {
    case AQO_SUM:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<LongT*>(out)[g] = sum(c[val]); });
        return 0;
    case AQO_AVG:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<double*>(out)[g] = avg(c[val]); });
        return 0;
    case AQO_MIN:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<T*>(out)[g] = min(c[val]); });
        return 0;
    case AQO_MAX:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<T*>(out)[g] = max(c[val]); });
        return 0;
    case AQO_FIRST:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<T*>(out)[g] = first(c[val]); });
        return 0;
    case AQO_LAST:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<T*>(out)[g] = last(c[val]); });
        return 0;
    case AQO_SECOND:                    /* subvec(v, 0, 2) of the gathered vector, element [1] */
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            if (val.size < 2) { static_cast<T*>(out)[g] = 0; return; }
            auto gathered = c[val];
            auto sv = gathered.subvec_memcpy(0, 2);
            static_cast<T*>(out)[g] = sv[1]; });
        return 0;
    case AQO_VAR:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<double*>(out)[g] = var(c[val]); });
        return 0;
    case AQO_STDDEV:
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<double*>(out)[g] = stddev(c[val]); });
        return 0;
    case AQO_CORR: {
        ColRef<T> c2(ng ? off[ng] : 0, const_cast<void*>(v2));
        per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
            static_cast<double*>(out)[g] = corr(c[val], c2[val]); });
        return 0; }
    default: return -2;
    }
}

template <class T, class OutT, class F>
void per_group_vec(const void* v, uint32_t ng, const uint32_t* off, const uint32_t* perm, void* out, F f) {
    per_group<T>(v, ng, off, perm, [&](uint32_t g, ColRef<T>& c, vector_type<uint32_t>& val) {
        vector_type<OutT> cell;                                   /* engine/ast.py:713-716 */
        cell.init_from(val.size, static_cast<OutT*>(out) + off[g]);
        auto gathered = c[val];
        f(gathered, cell);
    });
}

template <class T>
int scan_t(int op, uint32_t w, const void* v, uint32_t ng, const uint32_t* off,
           const uint32_t* perm, void* out) {
    typedef types::GetLongType<T> LongT;
    typedef types::GetFPType<LongT> LFP;
    typedef std::decay_t<types::GetFPType<T>> FP;
    typedef vector_type<T> V;
    switch (op) {
    case AQO_SUMS:   per_group_vec<T, LongT>(v, ng, off, perm, out, [&](V& a, vector_type<LongT>& r) { sums(a, r); }); return 0;
    case AQO_AVGS:   per_group_vec<T, LFP>(v, ng, off, perm, out, [&](V& a, vector_type<LFP>& r) { avgs(a, r); }); return 0;
    case AQO_MINS:   per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { mins(a, r); }); return 0;
    case AQO_MAXS:   per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { maxs(a, r); }); return 0;
    case AQO_SUMW:   per_group_vec<T, LongT>(v, ng, off, perm, out, [&](V& a, vector_type<LongT>& r) { sumw(w, a, r); }); return 0;
    case AQO_AVGW:   per_group_vec<T, LFP>(v, ng, off, perm, out, [&](V& a, vector_type<LFP>& r) { avgw(w, a, r); }); return 0;
    case AQO_MINW:   per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { minw(w, a, r); }); return 0;
    case AQO_MAXW:   per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { maxw(w, a, r); }); return 0;
    case AQO_DELTAS: per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { deltas(a, r); }); return 0;
    case AQO_PREV:   per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { prev(a, r); }); return 0;
    case AQO_NEXT:   per_group_vec<T, T>(v, ng, off, perm, out, [&](V& a, vector_type<T>& r) { aggnext(a, r); }); return 0;
    case AQO_RATIOW: per_group_vec<T, FP>(v, ng, off, perm, out, [&](V& a, vector_type<FP>& r) { ratiow(w, a, r); }); return 0;
    case AQO_VARS: case AQO_STDDEVS: {
        // the reference's loop printf()s every element (aggregations.h:368): stdout is parked on /dev/null meanwhile
        fflush(stdout);
        const int keep = ::dup(1), nul = ::open("/dev/null", O_WRONLY);
        dup2(nul, 1);
        if (op == AQO_VARS) per_group_vec<T, LFP>(v, ng, off, perm, out, [&](V& a, vector_type<LFP>& r) { vars<T, vector_type, vector_type<LFP>, false>(a, r); });
        else per_group_vec<T, LFP>(v, ng, off, perm, out, [&](V& a, vector_type<LFP>& r) { vars<T, vector_type, vector_type<LFP>, true>(a, r); });
        fflush(stdout);
        dup2(keep, 1); close(keep); close(nul);
        return 0; }
    default: return -2;                            // AQO_VARW / AQO_STDDEVW: the reference reads arr[-1] (aggregations.h:311): no defined result
    }
}

template <class T> struct code_of;
template <> struct code_of<int> { enum { v = AQO_INT32 }; };
template <> struct code_of<long> { enum { v = AQO_INT64 }; };
template <> struct code_of<float> { enum { v = AQO_FLOAT }; };
template <> struct code_of<double> { enum { v = AQO_DOUBLE }; };
template <> struct code_of<__int128_t> { enum { v = AQO_INT128 }; };
template <> struct code_of<unsigned> { enum { v = AQO_UINT32 }; };

template <class T> int reduce_out_t(int op) {
    switch (op) {
    case AQO_SUM: return code_of<types::GetLongType<T>>::v;
    case AQO_AVG: case AQO_VAR: case AQO_STDDEV: case AQO_CORR: return AQO_DOUBLE;
    case AQO_MIN: case AQO_MAX: case AQO_FIRST: case AQO_LAST: case AQO_SECOND: return code_of<T>::v;
    case AQO_COUNT: return AQO_UINT32;
    default: return -2;
    }
}
template <class T> int scan_out_t(int op) {
    switch (op) {
    case AQO_SUMS: case AQO_SUMW: return code_of<types::GetLongType<T>>::v;
    case AQO_AVGS: case AQO_AVGW: case AQO_VARS: case AQO_STDDEVS: return code_of<types::GetFPType<types::GetLongType<T>>>::v;
    case AQO_RATIOW: return code_of<std::decay_t<types::GetFPType<T>>>::v;
    case AQO_MINS: case AQO_MAXS: case AQO_MINW: case AQO_MAXW:
    case AQO_DELTAS: case AQO_PREV: case AQO_NEXT: return code_of<T>::v;
    default: return -2;
    }
}

/* Element-wise vector arithmetic / compare: the reference's own operators (server/table.h:820-973).  The element type of
 * the result is whatever the reference's templates deduce (Coercion / GetLongType / GetFPType, server/types.h:194-275);
 * its Type_t code is reported with the cells. */
template <class R> int vec_emit(const vector_type<R>& r, void* out, int* out_type) {
    *out_type = std::is_same<R, bool>::value ? AQO_BOOL : (int)types::Types<R>::getType();
    if (out) memcpy(out, r.container, sizeof(R) * r.size);
    return (int)sizeof(R);
}
template <class A, class B>
int vec_binary_t(int op, const void* a, int a_scalar, const void* b, int b_scalar, uint32_t n, void* out, int* out_type) {
    ensure_gc();
    GC::scratch_space = nullptr;                                   /* results on the heap */
    vector_type<A> va(a_scalar ? 0 : n, const_cast<void*>(a));
    vector_type<B> vb(b_scalar ? 0 : n, const_cast<void*>(b));
    const A sa = *static_cast<const A*>(a);
    const B sb = *static_cast<const B*>(b);
    if (a_scalar && b_scalar) return -2;
    switch (op) {
#define AQ_INFIX(CODE, OP)                                                                              \
    case CODE:                                                                                          \
        if (b_scalar) return vec_emit(va OP sb, out, out_type);                                         \
        if (a_scalar) return vec_emit(sa OP vb, out, out_type);                                         \
        return vec_emit(va OP vb, out, out_type);
    AQ_INFIX(AQO_V_ADD, +) AQ_INFIX(AQO_V_SUB, -) AQ_INFIX(AQO_V_MUL, *) AQ_INFIX(AQO_V_DIV, /)
#undef AQ_INFIX
    case AQO_V_GT:
        /* vector > scalar resolves to the member template (server/vector_type.hpp:372-381); vector > vector does not
         * compile as an infix expression at HEAD (the member wins and assigns a vector to a bool), so two vectors go
         * through aqop_gt (server/table.h:949-958) like the other comparisons */
        if (b_scalar) return vec_emit(va > sb, out, out_type);
        if (a_scalar) return -2;
        { vector_type<bool> r(n); aqop_gt(va, vb, r); return vec_emit(r, out, out_type); }
#define AQ_FN(CODE, NAME)                                                                               \
    case CODE: {                                                                                        \
        if (a_scalar || b_scalar) return -2;                      /* aqop_x takes two vectors */        \
        vector_type<bool> r(n);                                                                         \
        aqop_##NAME(va, vb, r);                                                                         \
        return vec_emit(r, out, out_type); }
    AQ_FN(AQO_V_LT, lt) AQ_FN(AQO_V_GTE, gte) AQ_FN(AQO_V_LTE, lte) AQ_FN(AQO_V_EQ, eq) AQ_FN(AQO_V_NEQ, neq)
#undef AQ_FN
    default: return -2;
    }
}
template <class A>
int vec_binary_a(int op, const void* a, int a_scalar, int tb, const void* b, int b_scalar, uint32_t n, void* out, int* out_type) {
    switch (tb) {
    case AQO_INT32:  return vec_binary_t<A, int>(op, a, a_scalar, b, b_scalar, n, out, out_type);
    case AQO_INT64:  return vec_binary_t<A, long>(op, a, a_scalar, b, b_scalar, n, out, out_type);
    case AQO_FLOAT:  return vec_binary_t<A, float>(op, a, a_scalar, b, b_scalar, n, out, out_type);
    case AQO_DOUBLE: return vec_binary_t<A, double>(op, a, a_scalar, b, b_scalar, n, out, out_type);
    default: return -2;
    }
}

/* sqrt / truncate over a vector (server/aggregations.h:34-69) */
template <class T> int vec_unary_t(int op, const void* x, uint32_t param, uint32_t n, void* out) {
    ensure_gc();
    GC::scratch_space = nullptr;
    vector_type<T> v(n, const_cast<void*>(x));
    if (op == AQO_U_SQRT) { auto r = sqrt(v); memcpy(out, r.container, sizeof(double) * n); return AQO_DOUBLE; }
    if (op == AQO_U_TRUNCATE) { auto r = truncate(v, param); memcpy(out, r.container, sizeof(T) * n); return (int)types::Types<T>::getType(); }
    return -2;
}

#define DISPATCH_VTYPE(vtype, CALL)                 \
    switch (vtype) {                                \
    case AQO_INT32:  return CALL(int);              \
    case AQO_INT64:  return CALL(long);             \
    case AQO_FLOAT:  return CALL(float);            \
    case AQO_DOUBLE: return CALL(double);           \
    default: return -2;                             \
    }

} // namespace

extern "C" {

const char* aqo_kind(void) { return "reference"; }

int aqo_groupby(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                uint32_t* ngroups, uint32_t* gid, uint32_t* offsets, uint32_t* perm) {
    if (nkeys < 1) return -2;
    if (nkeys == 1 && key_types[0] == AQO_STR)
        return groupby_mixed<astring_view>(std::make_index_sequence<1>{}, keys, nrows, ngroups, gid, offsets, perm);
    if (nkeys == 1 && key_types[0] == AQO_DATE)
        return groupby_mixed<types::date_t>(std::make_index_sequence<1>{}, keys, nrows, ngroups, gid, offsets, perm);
    if (nkeys == 1 && key_types[0] == AQO_TIME)
        return groupby_mixed<types::time_t>(std::make_index_sequence<1>{}, keys, nrows, ngroups, gid, offsets, perm);
    if (nkeys == 1 && key_types[0] == AQO_TIMESTAMP)
        return groupby_mixed<types::timestamp_t>(std::make_index_sequence<1>{}, keys, nrows, ngroups, gid, offsets, perm);
    if (nkeys == 2 && key_types[0] == AQO_STR && key_types[1] == AQO_INT32)
        return groupby_mixed<astring_view, int>(std::make_index_sequence<2>{}, keys, nrows, ngroups, gid, offsets, perm);
    for (int k = 1; k < nkeys; ++k) if (key_types[k] != key_types[0]) return -2;
    if (key_types[0] == AQO_INT32) return groupby_k<int>(nkeys, keys, nrows, ngroups, gid, offsets, perm);
    if (key_types[0] == AQO_INT64) return groupby_k<long>(nkeys, keys, nrows, ngroups, gid, offsets, perm);
    return -2;
}

int aqo_vec_binary(int op, int ta, const void* a, int a_scalar, int tb, const void* b, int b_scalar, uint32_t n,
                   void* out, int* out_type) {
#define CALL(T) vec_binary_a<T>(op, a, a_scalar, tb, b, b_scalar, n, out, out_type)
    DISPATCH_VTYPE(ta, CALL)
#undef CALL
}

int aqo_vec_unary(int op, int t, const void* x, uint32_t param, uint32_t n, void* out) {
#define CALL(T) vec_unary_t<T>(op, x, param, n, out)
    DISPATCH_VTYPE(t, CALL)
#undef CALL
}
int aqo_fp_precision(int t) {
#define CALL(T) (int)aq_fp_precision<T>
    DISPATCH_VTYPE(t, CALL)
#undef CALL
}

int aqo_reduce_out_type(int op, int vtype) {
#define CALL(T) reduce_out_t<T>(op)
    DISPATCH_VTYPE(vtype, CALL)
#undef CALL
}
int aqo_scan_out_type(int op, int vtype) {
#define CALL(T) scan_out_t<T>(op)
    DISPATCH_VTYPE(vtype, CALL)
#undef CALL
}

int aqo_reduce(int op, int vtype, const void* v, const void* v2,
               uint32_t ngroups, const uint32_t* offsets, const uint32_t* perm, void* out) {
    if (op == AQO_COUNT) {                       /* engine/ast.py:763: count(*) -> val.size */
        for (uint32_t g = 0; g < ngroups; ++g) static_cast<uint32_t*>(out)[g] = offsets[g + 1] - offsets[g];
        return 0;
    }
#define CALL(T) reduce_t<T>(op, v, v2, ngroups, offsets, perm, out)
    DISPATCH_VTYPE(vtype, CALL)
#undef CALL
}

int aqo_scan(int op, uint32_t w, int vtype, const void* v,
             uint32_t ngroups, const uint32_t* offsets, const uint32_t* perm, void* out) {
#define CALL(T) scan_t<T>(op, w, v, ngroups, offsets, perm, out)
    DISPATCH_VTYPE(vtype, CALL)
#undef CALL
}

int64_t aqo_time_groupby_reduce(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                                int n_aggs, const int* ops, const int* vtypes,
                                const void* const* vals, const void* const* vals2,
                                void* const* outs, uint32_t* ngroups) {
    ensure_gc();
    uint32_t* gid = static_cast<uint32_t*>(malloc(sizeof(uint32_t) * (size_t(nrows) * 3 + 2)));
    uint32_t* perm = gid + nrows;
    uint32_t* offsets = perm + nrows;
    void** tmp = static_cast<void**>(calloc(n_aggs ? n_aggs : 1, sizeof(void*)));
    for (int a = 0; a < n_aggs; ++a)
        if (!outs || !outs[a]) tmp[a] = malloc(size_t(nrows ? nrows : 1) * 16);
    auto t0 = std::chrono::high_resolution_clock::now();
    uint32_t ng = 0;
    int rc = aqo_groupby(keys, key_types, nkeys, nrows, &ng, gid, offsets, perm);
    for (int a = 0; rc == 0 && a < n_aggs; ++a)
        rc = aqo_reduce(ops[a], vtypes[a], vals[a], vals2 ? vals2[a] : nullptr, ng, offsets, perm,
                        tmp[a] ? tmp[a] : outs[a]);
    auto t1 = std::chrono::high_resolution_clock::now();
    for (int a = 0; a < n_aggs; ++a) free(tmp[a]);
    free(tmp); free(gid);
    if (ngroups) *ngroups = ng;
    if (rc) return rc;
    return std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
}

} // extern "C"
