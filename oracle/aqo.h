/*
 * aqo.h -- C interface shared by the two CPU oracles of the group-by / aggregation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load these
 * libraries, and only as the checker or the timed CPU baseline.  The product path
 * (aquery2_b200/libaqgpu.so) never links, loads or calls them and has no CPU fallback.
 *
 * Two implementations export exactly these symbols:
 *   oracle/_ref/libaqref.so   -- ref_harness.cpp: the REFERENCE'S OWN headers (server/hasher.h,
 *                                server/table.h, server/aggregations.h ...) compiled in a unity
 *                                build from /root/reference where they lie (oracle/Makefile).
 *   oracle/libaqoracle.so     -- aq_oracle.c: a plain-C restatement of the same algorithms that
 *                                travels to the GPU box (the reference tree does not).
 * tests/test_oracle_vs_ref.py pins the restatement against the reference build, and
 * tests/golden/ holds outputs of the reference build as committed fixtures.
 *
 * dtype codes are the reference's types::Type_t values (server/aquery_types.h:1-5).
 */
#ifndef AQO_H
#define AQO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum aqo_dtype {
    AQO_INT32 = 0, AQO_FLOAT = 1, AQO_DOUBLE = 3, AQO_INT64 = 5, AQO_INT128 = 6, AQO_UINT32 = 11,
    /* key columns only (aqo_groupby): ASTR = column of const char*, ADATE 4-byte, ATIME 8-byte, ATIMESTAMP 12-byte cells */
    AQO_STR = 2, AQO_DATE = 8, AQO_TIME = 9, AQO_TIMESTAMP = 18
};

/* per-group scalar aggregates (server/aggregations.h:10-32,71-86,332-348,383-407,413-415,487-497) */
enum aqo_reduce_op {
    AQO_SUM = 0, AQO_AVG = 1, AQO_MIN = 2, AQO_MAX = 3, AQO_COUNT = 4, AQO_FIRST = 5, AQO_LAST = 6,
    AQO_VAR = 7, AQO_STDDEV = 8, AQO_CORR = 9,
    AQO_SECOND = 10   /* subvec(v, 0, 2)[1] (server/vector_type.hpp:348-355), 0 for one-row groups */
};

/* per-group vector-valued aggregates (server/aggregations.h:89-125 running min/max, :127-167 windowed
 * min/max, :169-201 ratios, :203-281 running/windowed sum/avg, :439-485 deltas/prev/next) */
enum aqo_scan_op {
    AQO_SUMS = 0, AQO_AVGS = 1, AQO_MINS = 2, AQO_MAXS = 3,
    AQO_SUMW = 4, AQO_AVGW = 5, AQO_MINW = 6, AQO_MAXW = 7,
    AQO_DELTAS = 8, AQO_PREV = 9, AQO_NEXT = 10, AQO_RATIOW = 11,
    AQO_VARS = 12, AQO_STDDEVS = 13, AQO_VARW = 14, AQO_STDDEVW = 15
};

/* element-wise vector arithmetic / compare (server/table.h:820-973) */
enum aqo_vec_op {
    AQO_V_ADD = 0, AQO_V_SUB = 1, AQO_V_MUL = 2, AQO_V_DIV = 3, AQO_V_AND = 4, AQO_V_OR = 5, AQO_V_XOR = 6,
    AQO_V_GT = 7, AQO_V_LT = 8, AQO_V_GTE = 9, AQO_V_LTE = 10, AQO_V_EQ = 11, AQO_V_NEQ = 12
};
#define AQO_BOOL 16     /* types::ABOOL: 1-byte cells */

/* out[i] = a[i] op b[i] with the reference's operators: infix + - * / > for vector-vector, vector-scalar and
 * scalar-vector operands (a scalar operand is a pointer to one cell), aqop_lt/gte/lte/eq/neq for two vectors with a
 * bool result.  *out_type = Type_t code of the result cells as the reference's templates deduce it; returns the
 * cell size in bytes, or < 0.  out may be NULL (type query).  Both oracles export it; forms the reference cannot evaluate
 * (scalar > vector, other comparisons with a scalar: its member operators all compute `>`) return < 0. */
int aqo_vec_binary(int op, int ta, const void* a, int a_scalar, int tb, const void* b, int b_scalar, uint32_t n,
                   void* out, int* out_type);

/* sqrt(v) -> double cells, truncate(v, precision) -> cells of v's type (server/aggregations.h:34-69); returns the Type_t
 * code of the result cells.  aqo_fp_precision: aq_fp_precision<T> (server/types.h:473-475). */
enum aqo_vec_unary_op { AQO_U_SQRT = 0, AQO_U_TRUNCATE = 1 };
int aqo_vec_unary(int op, int t, const void* x, uint32_t param, uint32_t n, void* out);
int aqo_fp_precision(int t);

/* Which oracle is this?  "reference" (built from /root/reference) or "port" (C restatement). */
const char* aqo_kind(void);

/* Hash group-by: AQHashTable::hashtable_push per row + ht_postproc (server/hasher.h:146-199).
 *   keys[k]      : key column k, nrows elements of key_types[k] (AQO_INT32 or AQO_INT64; the port also takes any mix
 *                  with AQO_STR / AQO_DATE / AQO_TIME / AQO_TIMESTAMP, the reference build a single such key or
 *                  the pair (AQO_STR, AQO_INT32))
 *   ngroups      : out, number of distinct key tuples
 *   gid[nrows]   : out, "reversemap": dense group id of every row, ids in FIRST-APPEARANCE order
 *   offsets[ngroups+1] : out, CSR starts ("ht_base" after ht_postproc) plus the terminal nrows
 *   perm[nrows]  : out, "mapbase": row ids grouped by gid, DESCENDING row order inside a group
 * offsets must have room for nrows+1 entries.  Returns 0, or <0 for an unsupported key shape. */
int aqo_groupby(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                uint32_t* ngroups, uint32_t* gid, uint32_t* offsets, uint32_t* perm);

/* dtype of one output cell of aqo_reduce / aqo_scan for a value column of type vtype. */
int aqo_reduce_out_type(int op, int vtype);
int aqo_scan_out_type(int op, int vtype);

/* out[g] = op(v[perm[offsets[g] .. offsets[g+1])])   -- gather (server/table.h:184-189) then the
 * aggregate, one cell per group.  v2 is the second column of AQO_CORR (else NULL).
 * AQO_COUNT ignores v and writes uint32 group sizes (engine/ast.py:763 emits val.size). */
int aqo_reduce(int op, int vtype, const void* v, const void* v2,
               uint32_t ngroups, const uint32_t* offsets, const uint32_t* perm, void* out);

/* out[offsets[g] + j] = op(w, gathered group g)[j]   -- the CSR layout the emitter builds for
 * vector-valued group outputs (engine/ast.py:696-718).  w is ignored by the un-windowed ops. */
int aqo_scan(int op, uint32_t w, int vtype, const void* v,
             uint32_t ngroups, const uint32_t* offsets, const uint32_t* perm, void* out);

/* Whole post-processor bracket (server/server.cpp:303-305) for the CPU baseline: group-by on the key
 * columns, then n_aggs scalar aggregates; returns elapsed nanoseconds (or <0).  Outputs are written to
 * outs[a] when non-NULL (sized for nrows groups by the caller), ngroups to *ngroups. */
int64_t aqo_time_groupby_reduce(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                                int n_aggs, const int* ops, const int* vtypes,
                                const void* const* vals, const void* const* vals2,
                                void* const* outs, uint32_t* ngroups);
#ifdef __cplusplus
}
#endif
#endif
