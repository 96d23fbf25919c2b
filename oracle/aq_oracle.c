/*
 * aq_oracle.c -- plain-C restatement of AQuery++'s post-processor hot path (CPU oracle, "port").
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/aqo.h): loaded by tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py, never by the product library.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_ref.py checks every function below against
 * oracle/_ref/libaqref.so (the reference's own headers compiled from /root/reference, see
 * ref_harness.cpp) on seeded inputs, and tests/test_oracle_golden.py checks it against the fixtures
 * under tests/golden/ that oracle/gen_golden.py generated with that reference build.  The reference's
 * own tests hold no expected outputs for this path, so the reference build is
 * the only possible anchor.
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference).
 */
#define _POSIX_C_SOURCE 200809L
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <limits.h>
#include <time.h>
#include "aqo.h"

typedef __int128 i128;

const char* aqo_kind(void) { return "port"; }

/* ------------------------------------------------------------------------------------------------
 * Group build.  server/hasher.h:176-179 (AQHashTable::hashtable_push) on top of
 * server/unordered_dense.h:1117-1147: a key tuple not seen before is appended to m_values and its
 * index (= number of distinct tuples seen so far) is the group id; reversemap[i] = gid;
 * ++ht_base[gid].  The probing scheme (robin hood, wyhash) does not influence any output, so a plain
 * linear-probing table keyed on a representative row is used here.
 * ---------------------------------------------------------------------------------------------- */
static inline uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}

static inline uint64_t key_at(const void* col, int type, uint32_t i) {
    return type == AQO_INT64 ? (uint64_t)((const int64_t*)col)[i]
                             : (uint64_t)(int64_t)((const int32_t*)col)[i];
}

/* Keys that are not plain integers.  The reference's table decides with the types' operator==:
 *   AQO_STR        column of `const char*` (astring_view, server/types.h:280-284): same characters up to the NUL
 *                  (server/types.h:299-309)
 *   AQO_DATE       4-byte cells {day u8, month u8, year i16}; year, month, day equal (server/libaquery.cpp:112-114)
 *   AQO_TIME       8-byte cells {ms u32, seconds u8, minutes u8, hours u8, padding}; the four fields equal
 *                  (server/libaquery.cpp:162-164) -- the padding byte is not compared
 *   AQO_TIMESTAMP  12-byte cells {date, time}; both equal (server/libaquery.cpp:201-203)
 * The hash values (server/hasher.h:97-132) are not observable. */
static inline int key_is_plain(int type) { return type == AQO_INT32 || type == AQO_INT64; }
static inline int key_known(int type) {
    return key_is_plain(type) || type == AQO_STR || type == AQO_DATE || type == AQO_TIME || type == AQO_TIMESTAMP;
}
static inline uint64_t time_image(const unsigned char* c) {       /* ms, seconds, minutes, hours; padding dropped */
    uint32_t ms; memcpy(&ms, c, 4);
    return (uint64_t)ms | ((uint64_t)c[4] << 32) | ((uint64_t)c[5] << 40) | ((uint64_t)c[6] << 48);
}
static inline uint64_t key_hash1(const void* col, int type, uint32_t i) {
    switch (type) {
    case AQO_STR: {
        const unsigned char* p = (const unsigned char*)((const char* const*)col)[i];
        uint64_t h = 0xcbf29ce484222325ULL;
        for (; *p; ++p) h = (h ^ *p) * 0x100000001b3ULL;
        return h; }
    case AQO_DATE: { uint32_t w; memcpy(&w, (const unsigned char*)col + (size_t)i * 4, 4); return w; }
    case AQO_TIME: return time_image((const unsigned char*)col + (size_t)i * 8);
    case AQO_TIMESTAMP: {
        const unsigned char* c = (const unsigned char*)col + (size_t)i * 12;
        uint32_t w; memcpy(&w, c, 4);
        return mix64(w) ^ time_image(c + 4); }
    default: return key_at(col, type, i);
    }
}
static inline int key_eq1(const void* col, int type, uint32_t i, uint32_t r) {
    switch (type) {
    case AQO_STR: return strcmp(((const char* const*)col)[i], ((const char* const*)col)[r]) == 0;
    case AQO_DATE: return memcmp((const unsigned char*)col + (size_t)i * 4, (const unsigned char*)col + (size_t)r * 4, 4) == 0;
    case AQO_TIME: return time_image((const unsigned char*)col + (size_t)i * 8) == time_image((const unsigned char*)col + (size_t)r * 8);
    case AQO_TIMESTAMP: {
        const unsigned char* a = (const unsigned char*)col + (size_t)i * 12;
        const unsigned char* b = (const unsigned char*)col + (size_t)r * 12;
        return memcmp(a, b, 4) == 0 && time_image(a + 4) == time_image(b + 4); }
    default: return key_at(col, type, i) == key_at(col, type, r);
    }
}

int aqo_groupby(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                uint32_t* ngroups, uint32_t* gid_out, uint32_t* offsets, uint32_t* perm) {
    if (nkeys < 1 || nkeys > 16) return -2;
    for (int k = 0; k < nkeys; ++k)
        if (!key_known(key_types[k])) return -2;
    if (nrows == 0) { *ngroups = 0; if (offsets) offsets[0] = 0; return 0; }

    uint64_t cap = 16;
    while (cap < (uint64_t)nrows * 2) cap <<= 1;
    const uint64_t mask = cap - 1;
    uint32_t* slot_gid = (uint32_t*)malloc(cap * sizeof(uint32_t));     /* 0xffffffff = empty */
    uint32_t* rep_row = (uint32_t*)malloc((size_t)nrows * sizeof(uint32_t)); /* gid -> first row */
    uint32_t* gid = gid_out ? gid_out : (uint32_t*)malloc((size_t)nrows * sizeof(uint32_t));
    uint32_t* count = (uint32_t*)calloc((size_t)nrows + 1, sizeof(uint32_t));  /* hasher.h:157 ht_base */
    if (!slot_gid || !rep_row || !gid || !count) return -1;
    memset(slot_gid, 0xff, cap * sizeof(uint32_t));

    uint32_t ng = 0;
    for (uint32_t i = 0; i < nrows; ++i) {
        uint64_t h = 0x9e3779b97f4a7c15ULL;
        for (int k = 0; k < nkeys; ++k) h = mix64(h ^ key_hash1(keys[k], key_types[k], i));
        uint64_t s = h & mask;
        uint32_t g;
        for (;;) {
            g = slot_gid[s];
            if (g == 0xffffffffu) {                 /* unseen tuple: next dense id (unordered_dense.h:1142-1146) */
                g = ng++;
                slot_gid[s] = g;
                rep_row[g] = i;
                break;
            }
            const uint32_t r = rep_row[g];
            int eq = 1;
            for (int k = 0; k < nkeys && eq; ++k)
                eq = key_eq1(keys[k], key_types[k], i, r);
            if (eq) break;
            s = (s + 1) & mask;
        }
        gid[i] = g;                                 /* hasher.h:177 */
        ++count[g];                                 /* hasher.h:178 */
    }
    *ngroups = ng;

    /* ht_postproc, server/hasher.h:181-198: inclusive prefix sum of the counts, then
     * mapbase[--ht_base[gid]] = i for ascending i, i.e. every group's row list ends up in
     * DESCENDING row order and ht_base[g] ends as the start offset of group g. */
    if (offsets || perm) {
        for (uint32_t g = 1; g < ng; ++g) count[g] += count[g - 1];            /* hasher.h:188-191 */
        if (perm) {
            for (uint32_t i = 0; i < nrows; ++i) perm[--count[gid[i]]] = i;    /* hasher.h:193-196 */
        } else {                                   /* caller wants offsets only: inclusive -> exclusive */
            for (uint32_t g = ng; g-- > 1;) count[g] = count[g - 1];
            count[0] = 0;
        }
        if (offsets) {
            for (uint32_t g = 0; g < ng; ++g) offsets[g] = count[g];
            offsets[ng] = nrows;
        }
    }
    free(slot_gid); free(rep_row); free(count);
    if (!gid_out) free(gid);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Result types.  server/types.h:199-210: GetLongType (signed int -> __int128, fp -> double),
 * GetFPType (4-byte -> float, else double); avg is always double (aggregations.h:30).
 * ---------------------------------------------------------------------------------------------- */
static int is_int(int t) { return t == AQO_INT32 || t == AQO_INT64; }
static int is_val(int t) { return is_int(t) || t == AQO_FLOAT || t == AQO_DOUBLE; }

int aqo_reduce_out_type(int op, int vtype) {
    if (!is_val(vtype)) return -2;
    switch (op) {
    case AQO_SUM: return is_int(vtype) ? AQO_INT128 : AQO_DOUBLE;
    case AQO_AVG: case AQO_VAR: case AQO_STDDEV: case AQO_CORR: return AQO_DOUBLE;
    case AQO_MIN: case AQO_MAX: case AQO_FIRST: case AQO_LAST: case AQO_SECOND: return vtype;
    case AQO_COUNT: return AQO_UINT32;
    default: return -2;
    }
}
int aqo_scan_out_type(int op, int vtype) {
    if (!is_val(vtype)) return -2;
    switch (op) {
    case AQO_SUMS: case AQO_SUMW: return is_int(vtype) ? AQO_INT128 : AQO_DOUBLE;
    case AQO_AVGS: case AQO_AVGW: return AQO_DOUBLE;
    case AQO_VARS: case AQO_STDDEVS: case AQO_VARW: case AQO_STDDEVW: return AQO_DOUBLE;
    case AQO_RATIOW: return (vtype == AQO_INT32 || vtype == AQO_FLOAT) ? AQO_FLOAT : AQO_DOUBLE;
    case AQO_MINS: case AQO_MAXS: case AQO_MINW: case AQO_MAXW:
    case AQO_DELTAS: case AQO_PREV: case AQO_NEXT: return vtype;
    default: return -2;
    }
}

/* ------------------------------------------------------------------------------------------------
 * Per-type kernels.  GEN(T, L, FP, TMIN, TMAX, SFX):
 *   T = element type, L = GetLongType<T>, FP = GetFPType<T> (ratios), TMIN = numeric_limits<T>::min()
 *   (for floating point that is the smallest POSITIVE normal -- the reference's max/maxs start there,
 *   aggregations.h:73,111, so all-negative fp input yields 2.2e-308/1.2e-38; kept on purpose).
 * `a` below is the gathered vector col[vecs[g]] (server/table.h:184-189): a[j] = v[rows[j]].
 * ---------------------------------------------------------------------------------------------- */
#define GEN(T, L, FP, TMIN, TMAX, SFX)                                                              \
/* aggregations.h:19-27 */                                                                          \
static L sum_##SFX(const T* v, const uint32_t* r, uint32_t n) {                                     \
    L s = 0; for (uint32_t j = 0; j < n; ++j) s += v[r[j]]; return s; }                             \
/* aggregations.h:28-32 */                                                                          \
static double avg_##SFX(const T* v, const uint32_t* r, uint32_t n) {                                \
    return sum_##SFX(v, r, n) / (double)n; }                                                        \
/* aggregations.h:71-78 */                                                                          \
static T max_##SFX(const T* v, const uint32_t* r, uint32_t n) {                                     \
    T m = TMIN; for (uint32_t j = 0; j < n; ++j) m = m > v[r[j]] ? m : v[r[j]]; return m; }         \
/* aggregations.h:79-86 */                                                                          \
static T min_##SFX(const T* v, const uint32_t* r, uint32_t n) {                                     \
    T m = TMAX; for (uint32_t j = 0; j < n; ++j) m = m < v[r[j]] ? m : v[r[j]]; return m; }         \
/* aggregations.h:332-348 (divides by len+1; kept) */                                               \
static double var_##SFX(const T* v, const uint32_t* r, uint32_t n) {                                \
    L s = 0, ssq = 0;                                                                               \
    if (n) { s = v[r[0]]; ssq = v[r[0]] * v[r[0]]; }                                                \
    for (uint32_t j = 1; j < n; ++j) { s += v[r[j]]; ssq += v[r[j]] * v[r[j]]; }                    \
    return (ssq - s * s / (double)(n + 1)) / (double)(n + 1); }                                     \
/* aggregations.h:383-407.  `types::Coercion<...>` is used there without ::type, so InnerType is a  \
 * struct, GetLongType of it is __int128 and ALL five sums are __int128 for every input type: fp     \
 * inputs are truncated at each `+=` (usual arithmetic conversions, then back to __int128). */      \
static double corr_##SFX(const T* x, const T* y, const uint32_t* r, uint32_t n) {                   \
    i128 sx = 0, sy = 0, sxy = 0, sx2 = 0, sy2 = 0;                                                 \
    for (uint32_t j = 0; j < n; ++j) {                                                              \
        const T a = x[r[j]], b = y[r[j]];                                                           \
        sx += a; sx2 += a * a; sy += b; sxy += a * b; sy2 += b * b; }                               \
    return (n * sxy - (double)(sx * sy)) /                                                          \
           sqrt((n * sx2 - (double)(sx * sx)) * (n * sy2 - (double)(sy * sy))); }                   \
static int reduce_##SFX(int op, const T* v, const T* v2, uint32_t ng, const uint32_t* off,          \
                        const uint32_t* perm, void* out) {                                          \
    for (uint32_t g = 0; g < ng; ++g) {                                                             \
        const uint32_t* r = perm + off[g]; const uint32_t n = off[g + 1] - off[g];                  \
        switch (op) {                                                                               \
        case AQO_SUM: ((L*)out)[g] = sum_##SFX(v, r, n); break;                                     \
        case AQO_AVG: ((double*)out)[g] = avg_##SFX(v, r, n); break;                                \
        case AQO_MIN: ((T*)out)[g] = min_##SFX(v, r, n); break;                                     \
        case AQO_MAX: ((T*)out)[g] = max_##SFX(v, r, n); break;                                     \
        case AQO_FIRST: ((T*)out)[g] = n ? v[r[0]] : 0; break;        /* aggregations.h:493-497 */  \
        case AQO_LAST: ((T*)out)[g] = n ? v[r[n - 1]] : 0; break;     /* aggregations.h:487-491 */  \
        case AQO_SECOND: ((T*)out)[g] = n > 1 ? v[r[1]] : 0; break;   /* vector_type.hpp:348-355 */ \
        case AQO_VAR: ((double*)out)[g] = var_##SFX(v, r, n); break;                                \
        case AQO_STDDEV: ((double*)out)[g] = sqrt(var_##SFX(v, r, n)); break; /* :413-415 */        \
        case AQO_CORR: ((double*)out)[g] = corr_##SFX(v, v2, r, n); break;                          \
        default: return -2; } }                                                                     \
    return 0; }                                                                                     \
/* sliding min/max keep a monotonic deque of (value, index), aggregations.h:127-167; the window is   \
 * the last w elements because the front is dropped when its index equals i - w. */                 \
static void minmaxw_##SFX(int is_max, uint32_t w, const T* v, const uint32_t* r, uint32_t n, T* ret, \
                          T* dq_v, uint32_t* dq_i) {                                                \
    uint32_t head = 0, tail = 0;                                                                    \
    for (uint32_t i = 0; i < n; ++i) {                                                              \
        const T x = v[r[i]];                                                                        \
        if (tail > head && dq_i[head] == i - w) ++head;                                             \
        if (is_max) while (tail > head && dq_v[tail - 1] < x) --tail;                               \
        else        while (tail > head && dq_v[tail - 1] > x) --tail;                               \
        dq_v[tail] = x; dq_i[tail] = i; ++tail;                                                     \
        ret[i] = dq_v[head]; } }                                                                    \
static int scan_##SFX(int op, uint32_t w0, const T* v, uint32_t ng, const uint32_t* off,            \
                      const uint32_t* perm, void* out) {                                            \
    uint32_t maxn = 0;                                                                              \
    for (uint32_t g = 0; g < ng; ++g) if (off[g + 1] - off[g] > maxn) maxn = off[g + 1] - off[g];   \
    T* dq_v = NULL; uint32_t* dq_i = NULL;                                                          \
    if (op == AQO_MINW || op == AQO_MAXW) {                                                         \
        dq_v = (T*)malloc(sizeof(T) * (maxn + 1)); dq_i = (uint32_t*)malloc(4 * ((size_t)maxn + 1)); } \
    for (uint32_t g = 0; g < ng; ++g) {                                                             \
        const uint32_t* r = perm + off[g]; const uint32_t n = off[g + 1] - off[g];                  \
        const size_t o = off[g]; uint32_t w = w0;                                                   \
        switch (op) {                                                                               \
        case AQO_SUMS: { L* ret = (L*)out + o;                       /* aggregations.h:203-210 */   \
            if (n) ret[0] = v[r[0]];                                                                \
            for (uint32_t i = 1; i < n; ++i) ret[i] = ret[i - 1] + v[r[i]]; } break;                \
        case AQO_AVGS: { double* ret = (double*)out + o; L s = 0;    /* aggregations.h:219-228 */   \
            if (n) { s = v[r[0]]; ret[0] = v[r[0]]; }                                               \
            for (uint32_t i = 1; i < n; ++i) ret[i] = (s += v[r[i]]) / (double)(i + 1); } break;    \
        case AQO_MINS: { T* ret = (T*)out + o; T m = TMAX;           /* aggregations.h:89-98 */     \
            for (uint32_t i = 0; i < n; ++i) { if (v[r[i]] < m) m = v[r[i]]; ret[i] = m; } } break; \
        case AQO_MAXS: { T* ret = (T*)out + o; T m = TMIN;           /* aggregations.h:108-117 */   \
            for (uint32_t i = 0; i < n; ++i) { if (v[r[i]] > m) m = v[r[i]]; ret[i] = m; } } break; \
        case AQO_SUMW: { L* ret = (L*)out + o;                       /* aggregations.h:238-249 */   \
            w = w > n ? n : w;                                                                      \
            if (n) ret[0] = v[r[0]];                                                                \
            for (uint32_t i = 1; i < w; ++i) ret[i] = ret[i - 1] + v[r[i]];                         \
            for (uint32_t i = w; i < n; ++i) ret[i] = ret[i - 1] + v[r[i]] - v[r[i - w]]; } break;  \
        case AQO_AVGW: { double* ret = (double*)out + o; L s = 0;    /* aggregations.h:258-272 */   \
            w = w > n ? n : w;                                                                      \
            if (n) { s = v[r[0]]; ret[0] = v[r[0]]; }                                               \
            for (uint32_t i = 1; i < w; ++i) ret[i] = (s += v[r[i]]) / (double)(i + 1);             \
            for (uint32_t i = w; i < n; ++i)                                                        \
                ret[i] = ret[i - 1] + (v[r[i]] - v[r[i - w]]) / (double)w; } break;                 \
        case AQO_MINW: minmaxw_##SFX(0, w, v, r, n, (T*)out + o, dq_v, dq_i); break;                \
        case AQO_MAXW: minmaxw_##SFX(1, w, v, r, n, (T*)out + o, dq_v, dq_i); break;                \
        case AQO_DELTAS: { T* ret = (T*)out + o;                     /* aggregations.h:439-446 */   \
            if (n) ret[0] = 0;                                                                      \
            for (uint32_t i = 1; i < n; ++i) ret[i] = v[r[i]] - v[r[i - 1]]; } break;               \
        case AQO_PREV: { T* ret = (T*)out + o;                       /* aggregations.h:455-462 */   \
            if (n) ret[0] = v[r[0]];                                                                \
            for (uint32_t i = 1; i < n; ++i) ret[i] = v[r[i - 1]]; } break;                         \
        case AQO_NEXT: { T* ret = (T*)out + o;                       /* aggregations.h:471-478 */   \
            for (uint32_t i = 1; i < n; ++i) ret[i - 1] = v[r[i]];                                  \
            if (n) ret[n - 1] = v[r[n - 1]]; } break;                                               \
        case AQO_RATIOW: { FP* ret = (FP*)out + o;                   /* aggregations.h:169-184 */   \
            uint32_t len = n; if (n <= w) len = 1;                                                  \
            w = w > len ? len : w;                                                                  \
            if (n) ret[0] = 0;                                                                      \
            for (uint32_t i = 0; i < w && i < n; ++i) ret[i] = v[r[i]] / (FP)v[r[0]];               \
            for (uint32_t i = w; i < n; ++i) ret[i] = v[r[i]] / (FP)v[r[i - w]]; } break;           \
        case AQO_VARS: case AQO_STDDEVS: { double* ret = (double*)out + o; /* aggregations.h:350-381, 419-425: Welford;   \
            the reference's loop also printf()s every element (:368, a debug leftover) -- not restated */                  \
            L s = 0; double MnX = 0, EnX = 0;                                                       \
            if (n) { s = v[r[0]]; EnX = v[r[0]]; ret[0] = 0; }                                      \
            for (uint32_t i = 1; i < n; ++i) {                                                      \
                s += v[r[i]];                                                                       \
                const double e2 = s / (double)(i + 1);                                              \
                MnX += (v[r[i]] - EnX) * (v[r[i]] - e2); EnX = e2;                                  \
                ret[i] = MnX / (double)(i + 1);                                                     \
                if (op == AQO_STDDEVS) ret[i] = sqrt(ret[i]); } } break;                            \
        case AQO_VARW: case AQO_STDDEVW: { double* ret = (double*)out + o;                          \
            /* NOT a reference result: aggregations.h:283-330 reads arr[i-w-1] (arr[-1] at i == w) and its stddev   \
               form feeds sqrt'ed cells back into the recurrence.  This states what its warm-up loop defines:      \
               the population variance of the last min(w, i+1) elements, two-pass per window. */                  \
            if (w == 0) return -2;                                                                  \
            for (uint32_t i = 0; i < n; ++i) {                                                      \
                const uint32_t c = w > i + 1 ? i + 1 : w; double m = 0, q = 0;                      \
                for (uint32_t j = i + 1 - c; j <= i; ++j) m += (double)v[r[j]];                     \
                m /= (double)c;                                                                     \
                for (uint32_t j = i + 1 - c; j <= i; ++j) q += ((double)v[r[j]] - m) * ((double)v[r[j]] - m); \
                q /= (double)c;                                                                     \
                ret[i] = op == AQO_STDDEVW ? sqrt(q) : q; } } break;                                \
        default: free(dq_v); free(dq_i); return -2; } }                                             \
    free(dq_v); free(dq_i);                                                                         \
    return 0; }

GEN(int32_t, i128, float, INT32_MIN, INT32_MAX, i32)
GEN(int64_t, i128, double, INT64_MIN, INT64_MAX, i64)
GEN(float, double, float, FLT_MIN, FLT_MAX, f32)
GEN(double, double, double, DBL_MIN, DBL_MAX, f64)

int aqo_reduce(int op, int vtype, const void* v, const void* v2,
               uint32_t ngroups, const uint32_t* offsets, const uint32_t* perm, void* out) {
    if (op == AQO_COUNT) {                     /* engine/ast.py:763: count(*) is emitted as val.size */
        for (uint32_t g = 0; g < ngroups; ++g) ((uint32_t*)out)[g] = offsets[g + 1] - offsets[g];
        return 0;
    }
    switch (vtype) {
    case AQO_INT32: return reduce_i32(op, (const int32_t*)v, (const int32_t*)v2, ngroups, offsets, perm, out);
    case AQO_INT64: return reduce_i64(op, (const int64_t*)v, (const int64_t*)v2, ngroups, offsets, perm, out);
    case AQO_FLOAT: return reduce_f32(op, (const float*)v, (const float*)v2, ngroups, offsets, perm, out);
    case AQO_DOUBLE: return reduce_f64(op, (const double*)v, (const double*)v2, ngroups, offsets, perm, out);
    default: return -2;
    }
}

int aqo_scan(int op, uint32_t w, int vtype, const void* v,
             uint32_t ngroups, const uint32_t* offsets, const uint32_t* perm, void* out) {
    if ((op == AQO_SUMW || op == AQO_AVGW || op == AQO_MINW || op == AQO_MAXW || op == AQO_RATIOW) && w == 0)
        return -2;                             /* the reference indexes ret[-1] for w == 0 */
    switch (vtype) {
    case AQO_INT32: return scan_i32(op, w, (const int32_t*)v, ngroups, offsets, perm, out);
    case AQO_INT64: return scan_i64(op, w, (const int64_t*)v, ngroups, offsets, perm, out);
    case AQO_FLOAT: return scan_f32(op, w, (const float*)v, ngroups, offsets, perm, out);
    case AQO_DOUBLE: return scan_f64(op, w, (const double*)v, ngroups, offsets, perm, out);
    default: return -2;
    }
}

/* The reference's `postproc_time` bracket (server/server.cpp:303-305): hash build + row lists +
 * per-group aggregation, single thread, as the generated code runs them. */
int64_t aqo_time_groupby_reduce(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                                int n_aggs, const int* ops, const int* vtypes,
                                const void* const* vals, const void* const* vals2,
                                void* const* outs, uint32_t* ngroups) {
    uint32_t* gid = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)nrows * 3 + 2));
    if (!gid) return -1;
    uint32_t* perm = gid + nrows;
    uint32_t* offsets = perm + nrows;
    void** tmp = (void**)calloc(n_aggs ? n_aggs : 1, sizeof(void*));
    for (int a = 0; a < n_aggs; ++a)
        if (!outs || !outs[a]) tmp[a] = malloc((size_t)(nrows ? nrows : 1) * 16);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    uint32_t ng = 0;
    int rc = aqo_groupby(keys, key_types, nkeys, nrows, &ng, gid, offsets, perm);
    for (int a = 0; rc == 0 && a < n_aggs; ++a)
        rc = aqo_reduce(ops[a], vtypes[a], vals[a], vals2 ? vals2[a] : NULL, ng, offsets, perm,
                        tmp[a] ? tmp[a] : outs[a]);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    for (int a = 0; a < n_aggs; ++a) free(tmp[a]);
    free(tmp); free(gid);
    if (ngroups) *ngroups = ng;
    if (rc) return rc;
    return (int64_t)(t1.tv_sec - t0.tv_sec) * 1000000000LL + (t1.tv_nsec - t0.tv_nsec);
}

/* ------------------------------------------------------------------------------------------------
 * Element-wise vector arithmetic / compare and sqrt / truncate.
 * server/vector_type.hpp:372-396 (member operators: vector op vector -> Coercion<T1,T2>; vector > scalar),
 * server/table.h:826-925 (vector op scalar, scalar op vector: + - -> Coercion, * -> GetLongType, / -> GetFPType),
 * server/table.h:949-973 (aqop_x(lhs, rhs, ret): bool cells here), server/aggregations.h:34-69 (sqrt, truncate).
 * Every cell is the C expression `lhs[i] op rhs[i]` in the usual arithmetic conversion type of the two cell types, then
 * converted to the result cell type.  Pinned against the reference build by tests/test_oracle_vs_ref.py.
 * ---------------------------------------------------------------------------------------------- */
static int vrank(int t) { return t == AQO_INT32 ? 0 : t == AQO_INT64 ? 1 : t == AQO_FLOAT ? 2 : t == AQO_DOUBLE ? 3 : -1; }
static int vcoercion(int ra, int rb) {                   /* types::Coercion (server/types.h:264-275) as a Type_t code */
    const int fp = ra >= 2 || rb >= 2, wide = ra == 1 || ra == 3 || rb == 1 || rb == 3;
    return fp ? (wide ? AQO_DOUBLE : AQO_FLOAT) : (wide ? AQO_INT64 : AQO_INT32);
}
static int vec_out_type(int op, int ta, int a_scalar, int tb, int b_scalar) {
    const int ra = vrank(ta), rb = vrank(tb);
    if (ra < 0 || rb < 0) return -1;
    if (op >= AQO_V_GT) return AQO_BOOL;
    const int k = vcoercion(ra, rb);
    if (!a_scalar && !b_scalar) return k;
    if (op == AQO_V_MUL) return (k == AQO_INT32 || k == AQO_INT64) ? AQO_INT128 : AQO_DOUBLE;
    if (op == AQO_V_DIV) return k == AQO_INT32 ? AQO_FLOAT : (k == AQO_INT64 ? AQO_DOUBLE : k);
    return k;
}
#define VLOAD(C, p, t, i) ((t) == AQO_INT32 ? (C)((const int32_t*)(p))[i] : (t) == AQO_INT64 ? (C)((const int64_t*)(p))[i] : \
                           (t) == AQO_FLOAT ? (C)((const float*)(p))[i] : (C)((const double*)(p))[i])
#define VSTORE(out, t, i, r)                                                                              \
    do {                                                                                                  \
        switch (t) {                                                                                      \
        case AQO_INT32: ((int32_t*)(out))[i] = (int32_t)(r); break;                                       \
        case AQO_INT64: ((int64_t*)(out))[i] = (int64_t)(r); break;                                       \
        case AQO_FLOAT: ((float*)(out))[i] = (float)(r); break;                                           \
        case AQO_DOUBLE: ((double*)(out))[i] = (double)(r); break;                                        \
        case AQO_INT128: ((i128*)(out))[i] = (i128)(r); break;                                            \
        default: ((unsigned char*)(out))[i] = (r) != 0; break;                                            \
        }                                                                                                 \
    } while (0)
#define VLOOP(C, WRAP)                                                                                    \
    for (uint32_t i = 0; i < n; ++i) {                                                                    \
        const C x = VLOAD(C, a, ta, a_scalar ? 0 : i), y = VLOAD(C, b, tb, b_scalar ? 0 : i);             \
        switch (op) {                                                                                     \
        case AQO_V_ADD: { const C r = WRAP(x, +, y); VSTORE(out, ot, i, r); break; }                      \
        case AQO_V_SUB: { const C r = WRAP(x, -, y); VSTORE(out, ot, i, r); break; }                      \
        case AQO_V_MUL: { const C r = WRAP(x, *, y); VSTORE(out, ot, i, r); break; }                      \
        case AQO_V_DIV: { const C r = x / y; VSTORE(out, ot, i, r); break; }                              \
        case AQO_V_GT: ((unsigned char*)out)[i] = x > y; break;                                           \
        case AQO_V_LT: ((unsigned char*)out)[i] = x < y; break;                                           \
        case AQO_V_GTE: ((unsigned char*)out)[i] = x >= y; break;                                         \
        case AQO_V_LTE: ((unsigned char*)out)[i] = x <= y; break;                                         \
        case AQO_V_EQ: ((unsigned char*)out)[i] = x == y; break;                                          \
        case AQO_V_NEQ: ((unsigned char*)out)[i] = x != y; break;                                         \
        default: return -2;                                                                               \
        }                                                                                                 \
    }
#define WRAP32(x, o, y) ((int32_t)((uint32_t)(x) o (uint32_t)(y)))       /* two's complement wrap, as the compiled reference does */
#define WRAP64(x, o, y) ((int64_t)((uint64_t)(x) o (uint64_t)(y)))
#define PLAIN(x, o, y) ((x) o (y))

int aqo_vec_binary(int op, int ta, const void* a, int a_scalar, int tb, const void* b, int b_scalar, uint32_t n,
                   void* out, int* out_type) {
    const int ra = vrank(ta), rb = vrank(tb);
    if (ra < 0 || rb < 0 || (a_scalar && b_scalar)) return -2;
    if (op > AQO_V_DIV && op < AQO_V_GT) return -2;                      /* and / or / xor: not driven through the oracle */
    if (op == AQO_V_GT ? a_scalar : (op > AQO_V_GT && (a_scalar || b_scalar))) return -2;   /* forms the reference cannot evaluate */
    const int ot = vec_out_type(op, ta, a_scalar, tb, b_scalar);
    if (out_type) *out_type = ot;
    const int osz = ot == AQO_BOOL ? 1 : ot == AQO_INT128 ? 16 : (ot == AQO_INT32 || ot == AQO_FLOAT) ? 4 : 8;
    if (!out) return osz;
    const int rc = (ra == 3 || rb == 3) ? 3 : (ra == 2 || rb == 2) ? 2 : (ra == 1 || rb == 1) ? 1 : 0;
    switch (rc) {
    case 0: VLOOP(int32_t, WRAP32) break;
    case 1: VLOOP(int64_t, WRAP64) break;
    case 2: VLOOP(float, PLAIN) break;
    default: VLOOP(double, PLAIN) break;
    }
    return osz;
}

/* sqrt: float cells take the root in float (the overload the reference's call resolves to), integers in double; cells of double.
 * truncate: v < max(T) / 10^p ? round(v * 10^p) / 10^p : v in double; a copy when aq_fp_precision<T> <= p (float 7, double 16,
 * integers 0: server/types.h:473-475). */
int aqo_fp_precision(int t) { return t == AQO_FLOAT ? 7 : t == AQO_DOUBLE ? 16 : vrank(t) >= 0 ? 0 : -2; }
int aqo_vec_unary(int op, int t, const void* x, uint32_t param, uint32_t n, void* out) {
    if (vrank(t) < 0) return -2;
    if (op == AQO_U_SQRT) {
        for (uint32_t i = 0; i < n; ++i)
            ((double*)out)[i] = t == AQO_FLOAT ? (double)sqrtf(((const float*)x)[i]) : sqrt(VLOAD(double, x, t, i));
        return AQO_DOUBLE;
    }
    if (op != AQO_U_TRUNCATE) return -2;
    const int esz = (t == AQO_INT32 || t == AQO_FLOAT) ? 4 : 8;
    if ((uint32_t)aqo_fp_precision(t) <= param) { memcpy(out, x, (size_t)esz * n); return t; }
    const double mult = pow(10, param);
    if (t == AQO_FLOAT) {
        const double mx = FLT_MAX / mult;
        for (uint32_t i = 0; i < n; ++i) { const float v = ((const float*)x)[i]; ((float*)out)[i] = v < mx ? (float)(round(v * mult) / mult) : v; }
    } else {
        const double mx = DBL_MAX / mult;
        for (uint32_t i = 0; i < n; ++i) { const double v = ((const double*)x)[i]; ((double*)out)[i] = v < mx ? round(v * mult) / mult : v; }
    }
    return t;
}
