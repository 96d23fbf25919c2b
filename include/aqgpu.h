/*
 * aqgpu.h -- C ABI of libaqgpu: AQuery++'s post-processor hot path (hash group-by over column
 * batches + per-group scalar / running / windowed aggregation) on one B200 (sm_100a).
 *
 * This is the drop-in boundary.  The reference runs this path inside its generated dll.so by
 * instantiating header templates (server/hasher.h, server/table.h, server/aggregations.h); there is
 * no FFI in the reference for it, so every entry point below cites the reference template / loop it
 * replaces (paths relative to the reference tree).  The C++ drop-in headers under
 * aquery2_b200/host/server/ keep the reference's names (ColRef, vector_type, TableInfo,
 * HashTableFactory, sum/avg/...) and forward here; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ / torch types cross this boundary;
 *   - every function returns 0 on success, non-zero on error (server/monetdb_conn.cpp:393-396 uses
 *     the int result as a boolean); aqgpu_last_error() gives the message; nothing throws;
 *   - column pointers may be DEVICE pointers (HBM-resident column store, the normal case) or HOST
 *     pointers (pageable or pinned): host inputs are staged to HBM and host outputs copied back by
 *     the call itself, on the library stream;
 *   - row counts and row ids are uint32_t as in the reference (server/vector_type.hpp:66,
 *     server/hasher.h:149-198); byte counts are size_t;
 *   - dtype codes are the reference's types::Type_t values (server/aquery_types.h:1-5);
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef AQGPU_H
#define AQGPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AQGPU_ABI_VERSION 1

/* ---- dtypes: types::Type_t (server/aquery_types.h:1-5) ------------------------------------- */
enum aqgpu_dtype {
    AQ_INT32 = 0, AQ_FLOAT = 1, AQ_DOUBLE = 3, AQ_INT64 = 5, AQ_INT128 = 6, AQ_UINT32 = 11,
    AQ_BOOL = 16        /* 1-byte cells: results of comparisons (aqgpu_vec_binary only) */
};

/* ---- per-group scalar aggregates (server/aggregations.h) ------------------------------------
 * Result cell types follow the reference (server/types.h:199-210):
 *   SUM(int32|int64) -> AQ_INT128 (or AQ_INT64 on request), SUM(float|double) -> AQ_DOUBLE,
 *   (integer sums, sums of squares and of products accumulate in 64 bits and are widened on write: |sum| < 2^63 is
 *   the documented limit -- the reference's __int128 accumulator, server/types.h:155, has none in practice either)
 *   AVG/VAR/STDDEV/CORR -> AQ_DOUBLE, MIN/MAX/FIRST/LAST -> the input type, COUNT -> AQ_UINT32.  */
enum aqgpu_reduce_op {
    AQ_SUM = 0,     /* aggregations.h:19-27  */
    AQ_AVG = 1,     /* aggregations.h:28-32  sum / double(size) */
    AQ_MIN = 2,     /* aggregations.h:79-86  */
    AQ_MAX = 3,     /* aggregations.h:71-78  identity numeric_limits<T>::min(): fp quirk kept */
    AQ_COUNT = 4,   /* engine/ast.py:763     count(*) == vecs[g].size */
    AQ_FIRST = 5,   /* aggregations.h:493-497 first element of the group's row list */
    AQ_LAST = 6,    /* aggregations.h:487-491 */
    AQ_VAR = 7,     /* aggregations.h:332-348 (divides by len+1, kept) */
    AQ_STDDEV = 8,  /* aggregations.h:413-415 */
    AQ_CORR = 9,    /* aggregations.h:383-407 integer inputs only */
    AQ_SECOND = 10, /* element [1] of the group's row list (0 when the group has one row): with
                       AQ_FIRST this is subvec(v, 0, 2) of h2o q8 (benchmark/h2o/groupby.sql:17,
                       server/vector_type.hpp:348-355) */
    /* partial moments, the decomposable halves of VAR / STDDEV / CORR for multi-GPU merges
     * (result cells like AQ_SUM): sum of v*v and of v*v2, products formed in the column's own type as
     * aggregations.h:339-345, 392-398 do */
    AQ_SUMSQ = 11,
    AQ_SUMXY = 12
};

/* ---- per-group vector-valued aggregates, CSR output (engine/ast.py:696-718) ------------------ */
enum aqgpu_scan_op {
    AQ_SUMS = 0,    /* aggregations.h:203-210 */
    AQ_AVGS = 1,    /* aggregations.h:219-228 */
    AQ_MINS = 2,    /* aggregations.h:89-98   */
    AQ_MAXS = 3,    /* aggregations.h:108-117 */
    AQ_SUMW = 4,    /* aggregations.h:238-249 SQL sums(w, x) */
    AQ_AVGW = 5,    /* aggregations.h:258-272 SQL avgs(w, x) */
    AQ_MINW = 6,    /* aggregations.h:127-138 SQL mins(w, x) */
    AQ_MAXW = 7,    /* aggregations.h:148-159 SQL maxs(w, x) */
    AQ_DELTAS = 8,  /* aggregations.h:439-446 */
    AQ_PREV = 9,    /* aggregations.h:455-462 */
    AQ_NEXT = 10,   /* aggregations.h:471-478 (aggnext) */
    AQ_RATIOW = 11, /* aggregations.h:169-184 ratios(x) == ratiow(1, x) */
    /* moments, second wave (benchmark/quries/Aquery/q10.a: MAX(stddevs(3, price))); results AQ_DOUBLE.
     * AQ_VARS / AQ_STDDEVS: running population variance of the group's first i+1 elements (aggregations.h:350-381,
     * 419-425: Welford; evaluated here from exact prefix sums of x and x*x, <= 1e-12 apart).
     * AQ_VARW / AQ_STDDEVW: population variance of the last min(w, i+1) elements.  The reference's recurrence
     * (aggregations.h:283-330) reads arr[i-w-1] -- arr[-1] at i == w -- and mixes sqrt'ed and plain cells in its
     * stddev form, so it has no defined result to match: this is the definition its warm-up loop states. */
    AQ_VARS = 12, AQ_STDDEVS = 13, AQ_VARW = 14, AQ_STDDEVW = 15
};

/* Order of the rows inside a group's row list. */
enum aqgpu_row_order {
    AQ_ROWS_DESC = 0,  /* reference AQHashTable::ht_postproc (server/hasher.h:193-196): descending */
    AQ_ROWS_ASC = 1    /* PerfectHashTable / legacy emplace_back(i) (hasher.h:292-295,
                          common/groupby.py:42-46): ascending */
};

/* ---- context -------------------------------------------------------------------------------- */
int aqgpu_abi_version(void);
/* Binds the calling process to CUDA device `device`, creates the library stream and memory pool. */
int aqgpu_init(int device);
int aqgpu_shutdown(void);
const char* aqgpu_last_error(void);
/* Use `cuda_stream` (a cudaStream_t) for all subsequent work; NULL restores the library stream. */
int aqgpu_set_stream(void* cuda_stream);
int aqgpu_sync(void);
/* Number of kernels this library has launched since aqgpu_init (bench.py's gpu_launches). */
uint64_t aqgpu_launch_count(void);
/* CUDA-event timing on the library stream: begin/end return elapsed milliseconds in *ms. */
int aqgpu_timer_begin(void);
int aqgpu_timer_end(float* ms);

/* Per-kernel timing with CUDA events recorded on the library stream around every launch (off by
 * default).  aqgpu_profile_tag prefixes subsequent records ("q3:k_probe_agg<true>");
 * aqgpu_profile_report synchronises and writes "name<TAB>calls<TAB>total_ms" lines, then clears. */
int aqgpu_profile_enable(int on);
int aqgpu_profile_tag(const char* tag);
int aqgpu_profile_report(char* buf, size_t cap);

/* ---- HBM column store (replaces the malloc'ed / MonetDB-owned buffers ColRef points into:
 *      server/table.h:96-103 ColRef(size, void*), server/monetdb_conn.cpp:203-224 getCol) -------- */
int aqgpu_col_alloc(size_t bytes, void** dptr);
int aqgpu_col_free(void* dptr);
int aqgpu_col_upload(void* dptr, const void* host, size_t bytes);
int aqgpu_col_download(void* host, const void* dptr, size_t bytes);
/* stream-ordered copies: return at once, complete in stream order (aqgpu_sync / the next synchronous call) */
int aqgpu_col_upload_async(void* dptr, const void* pinned_host, size_t bytes);
int aqgpu_col_download_async(void* pinned_host, const void* dptr, size_t bytes);
/* Ingest and write-back on dedicated copy streams (server/monetdb_conn.cpp:203-224 getCol, the 'C' cache
 * message server/server.cpp:552-583, result write-back server/table_ext_monetdb.hpp:35-86): both directions of the host
 * link run while the library stream computes.  All ordering is by events on the device:
 *   aqgpu_ingest_async     H2D copy; starts after the work already enqueued on the library stream, which may still read the
 *                          column's previous contents
 *   aqgpu_ingest_fence     library work enqueued from now on waits for every ingest enqueued so far
 *   aqgpu_writeback_async  D2H copy; starts after the work already enqueued on the library stream (the producing query)
 *   aqgpu_writeback_fence  library work enqueued from now on waits for every write-back enqueued so far (call before a
 *                          query overwrites result buffers that may still be on their way out)
 *   aqgpu_writeback_wait   blocks the host until every write-back has landed
 * Host buffers must be pinned (aqgpu_host_alloc) for the copies to overlap.  Measured on B200 / PCIe 5: the two directions
 * overlap (45 + 45 GB/s), but copies of ONE direction execute in submission order across streams, the library's own few-byte
 * read-backs included -- enqueue the large write-backs last in a step. */
int aqgpu_ingest_async(void* dptr, const void* pinned_host, size_t bytes);
int aqgpu_ingest_fence(void);
int aqgpu_writeback_async(void* pinned_host, const void* dptr, size_t bytes);
int aqgpu_writeback_fence(void);
int aqgpu_writeback_wait(void);
/* Returns the unused memory of the library's stream-ordered pool to the device (synchronises). */
int aqgpu_trim_pool(void);
int aqgpu_host_alloc(size_t bytes, void** hptr);      /* pinned host memory for staging */
int aqgpu_host_free(void* hptr);
/* 1 if p points into device memory (a column of the HBM store), 0 for host memory. */
int aqgpu_ptr_is_device(const void* p);

/* ---- fused group-by + scalar aggregation -----------------------------------------------------
 * Replaces, in one pass over the columns and without materialising row lists:
 *   AQHashTable::hashtable_push per row     server/hasher.h:176-179
 *   AQHashTable::ht_postproc                server/hasher.h:181-198   (only its observable order)
 *   gather col[vecs[g]]                     server/table.h:184-189
 *   agg(...) per group                      server/aggregations.h, loop at engine/ast.py:720-790
 * Groups come out in FIRST-APPEARANCE order of their key tuple, as m_values does
 * (server/unordered_dense.h:1142-1146).                                                        */
typedef struct aqgpu_agg {
    int op;            /* aqgpu_reduce_op */
    int vtype;         /* dtype of v (and v2); ignored for AQ_COUNT */
    const void* v;     /* value column, nrows elements (device or host) */
    const void* v2;    /* second column of AQ_CORR, else NULL */
    int out_type;      /* requested cell type; 0-initialised structs get the reference default via
                          aqgpu_reduce_out_type(); AQ_INT64 is accepted for integer SUM */
    void* out;         /* result column, room for max_groups cells (device or host) */
} aqgpu_agg_t;

int aqgpu_reduce_out_type(int op, int vtype);
int aqgpu_scan_out_type(int op, int vtype);
size_t aqgpu_dtype_size(int dtype);

/*   keys[k]        key column k: nrows elements of key_types[k] (AQ_INT32, AQ_UINT32, AQ_INT64)
 *   key_min/max    optional per-column inclusive bounds of the key values (both NULL = unknown): the
 *                  column statistics the reference keeps in ColRef::stats (server/table.h:60-91) and
 *                  HashTableFactory::get consults (server/hasher.h:335-343).  With bounds whose
 *                  bit-widths sum to <= 32 the tuple is bit-packed like PerfectHashTable does
 *                  (hasher.h:226-243); a key outside its bounds is detected and the call silently
 *                  re-runs without them.  Output order is first-appearance either way.
 *   row_order      aqgpu_row_order: decides what FIRST/LAST mean (reference: AQ_ROWS_DESC)
 *   groups_hint    expected number of groups (table sizing); 0 = unknown (adaptive growth)
 *   max_groups     capacity, in cells, of out_keys[k], out_first_row and aggs[a].out
 *   out_keys[k]    key column k of the result, first-appearance order (may be NULL)
 *   out_first_row  row id of each group's first appearance (may be NULL)
 *   ngroups        host out
 * Fails with a "capacity" error if the number of groups exceeds max_groups.                      */
int aqgpu_groupby_agg(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                      const int64_t* key_min, const int64_t* key_max,
                      int row_order, uint32_t groups_hint,
                      const aqgpu_agg_t* aggs, int n_aggs, uint32_t max_groups,
                      void* const* out_keys, uint32_t* out_first_row, uint32_t* ngroups);

/* Same computation with the result columns ALLOCATED BY THE LIBRARY in HBM once the number of groups is
 * known -- the order of events of the generated code itself: group build, then col.resize(g.size), then the
 * per-group loop (engine/ast.py:681-701, 720-790).  The caller passes no capacities; key_min / key_max /
 * groups_hint may all be 0 / NULL, in which case the library samples the key columns (strided, ~1M rows)
 * inside this call to choose between its low-cardinality, partitioned and groups ~ rows paths; everything it
 * speculates from the sample is verified by the scan itself.  aggs[a].out is ignored.                      */
typedef struct aqgpu_result aqgpu_result_t;
int aqgpu_groupby_run(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                      const int64_t* key_min, const int64_t* key_max,
                      int row_order, uint32_t groups_hint,
                      const aqgpu_agg_t* aggs, int n_aggs, int want_first_row, aqgpu_result_t** out);
int aqgpu_result_size(const aqgpu_result_t* r, uint32_t* ngroups);
/* device pointers to ngroups cells each (owned by the result; valid until aqgpu_result_free) */
int aqgpu_result_key(const aqgpu_result_t* r, int k, void** dptr, int* dtype);
int aqgpu_result_agg(const aqgpu_result_t* r, int a, void** dptr, int* dtype);
int aqgpu_result_first_row(const aqgpu_result_t* r, uint32_t** dptr);
int aqgpu_result_free(aqgpu_result_t* r);

/* ---- multi-GPU: one process per GPU, rows sharded contiguously by global row index.
 * The reference is a single process (server/server.cpp:236-238); these entry points have no counterpart there.
 * aqgpu_comm_unique_id: rank 0 obtains the 128-byte NCCL id and the host program hands it to every rank (any
 * channel); aqgpu_comm_init: every rank, after aqgpu_init on its own device.  libnccl.so.2 is bound at run time. */
int aqgpu_comm_unique_id(void* id128);
int aqgpu_comm_init(int nranks, int rank, const void* id128);
int aqgpu_comm_destroy(void);
int aqgpu_comm_size(int* nranks, int* rank);
/* aqgpu_groupby_run over a row SHARD, called collectively by every rank: keys / value columns hold the rows
 * [row_base, row_base + nrows_local) of a table of rows_total rows (< 2^32).  Every rank aggregates its shard into
 * a dense record table, the tables are combined over NVLink (all-gather for few groups; an all-to-all of slot
 * ranges otherwise) and the result is what aqgpu_groupby_run gives on the whole table: first rows are GLOBAL row
 * ids and groups come in global first-appearance order.  aqgpu_result_replicated tells how the result is spread:
 * 1 = every rank holds all groups; 0 = a rank holds the groups of the key-slot range it owns (the ranks' results
 * merged by first row give the reference's order); 2 = groups ~ rows: only (row, fingerprint) tuples were exchanged
 * and a rank holds the groups whose FIRST ROW lies in its shard, in row order (the results concatenated in rank order
 * are in the reference's order).  *used = 0 (and no result): the shape has no sharded path (64-bit value columns,
 * sparse wide keys with many rows per group, ...) -- the caller exchanges rows or partial results itself
 * (aquery2_b200/dist.py).  A speculation failure on ANY rank makes every rank decline. */
int aqgpu_groupby_run_dist(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows_local,
                           uint64_t row_base, uint64_t rows_total, int row_order,
                           const aqgpu_agg_t* aggs, int n_aggs, int want_first_row, aqgpu_result_t** out, int* used);
int aqgpu_result_replicated(const aqgpu_result_t* r, int* replicated);
/* Order-sensitive (windowed) aggregates over row shards: the ROWS are repartitioned by hash(key tuple) mod ranks so
 * that every group lies whole on one rank; called collectively.  cols[k]: device columns of
 * nrows cells of elem_bytes[k] (4, 8 or 16) -- pass the key columns among them if they are needed afterwards.
 * out_cols[0..ncols) receive library-allocated device columns (aqgpu_col_free) of *out_nrows cells; out_cols[ncols] is
 * an extra uint32 column with the rows' GLOBAL ids.  The received rows are in ascending global row order, so a table
 * that was in time order still is inside every group; aqgpu_groupby_build / aqgpu_groups_order_by / aqgpu_group_window
 * then run locally and the results stay partitioned by key owner. */
int aqgpu_repartition_rows(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows, uint64_t row_base,
                           const void* const* cols, const int* elem_bytes, int ncols, void** out_cols, uint32_t* out_nrows);

/* ---- group handle: ids, row lists, per-group vectors ----------------------------------------
 * aqgpu_groupby_build replaces HashTableFactory<...>::get(cols...) / AQHashTable (server/hasher.h:
 * 146-199, 324-357): the handle is the HashTableComponents {size, keys, values, offsets}
 * (server/hasher.h:201-207) kept in HBM.                                                         */
typedef struct aqgpu_groups aqgpu_groups_t;

int aqgpu_groupby_build(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                        const int64_t* key_min, const int64_t* key_max,
                        int row_order, uint32_t groups_hint, aqgpu_groups_t** out);
/* Inclusive [min, max] of an integer column (AQ_INT32, AQ_UINT32, AQ_INT64): ColRef::populate_stats
 * (server/table.h:76-91).  One streaming pass; callers cache the result with the column. */
int aqgpu_col_minmax(int dtype, const void* col, uint32_t nrows, int64_t* out_min, int64_t* out_max);
int aqgpu_groups_free(aqgpu_groups_t* g);
int aqgpu_groups_size(aqgpu_groups_t* g, uint32_t* ngroups);                 /* .size           */
int aqgpu_groups_keys(aqgpu_groups_t* g, int k, void* out_keys);             /* .keys, column k */
int aqgpu_groups_first_row(aqgpu_groups_t* g, uint32_t* out);
int aqgpu_groups_gid(aqgpu_groups_t* g, uint32_t* out_gid);                  /* reversemap      */
/* offsets[ngroups+1] (.offsets / ht_base + terminal) and perm[nrows] (mapbase, the storage all
 * .values[g] views alias); either may be NULL. */
int aqgpu_groups_csr(aqgpu_groups_t* g, uint32_t* out_offsets, uint32_t* out_perm);
/* Device-resident views for callers that keep working on the GPU (NULL until first needed). */
int aqgpu_groups_device_views(aqgpu_groups_t* g, const uint32_t** d_gid, const uint32_t** d_offsets,
                              const uint32_t** d_perm);

/* out[g] = op(v[rows of g])  -- server/aggregations.h scalar aggregates over col[vecs[g]]. */
int aqgpu_group_reduce(aqgpu_groups_t* g, int op, int vtype, const void* v, const void* v2,
                       int out_type, void* out);
/* out[g] = median of v over the rows of g, fp64 (mean of the two middle values for an even count).
 * benchmark/h2o/groupby.sql:11-12 asks for median(v3) (the line is commented out there); the reference's library has
 * NO median (server/aggregations.h has none, though common/types.py:343 names it), so this is not a reference result:
 * it is checked against numpy.  A stable sort of the row lists by v (the segmented radix sort of
 * aqgpu_groups_order_by, on a temporary copy: the handle's order is untouched), then the middle element(s).
 * v: AQ_INT32, AQ_UINT32 or AQ_FLOAT. */
int aqgpu_group_median(aqgpu_groups_t* g, int vtype, const void* v, double* out);
/* out[offsets[g] + j] = op(w, v[rows of g])[j]  -- running / windowed / neighbour aggregates in
 * the CSR layout of engine/ast.py:696-718.  out has nrows cells of out_type. */
int aqgpu_group_scan(aqgpu_groups_t* g, int op, uint32_t w, int vtype, const void* v,
                     int out_type, void* out);

/* Several vector-valued expressions of one group-by in ONE call -- the body of a generated query such as
 *   SELECT sym, avgs(5, price), sums(5, quantity), deltas(price), mins(5, price), max(price - mins(price))
 *   FROM trade GROUP BY sym              (benchmark/quries/Aquery/q7.a, q10.a; tests/moving_avg.a)
 * which the reference evaluates group by group on gathered vectors (engine/ast.py:696-718).  Each entry is
 * one aqgpu_group_scan; `minus_from_v` turns it into the element-wise v - scan(v) of server/table.h:820-973
 * (only for ops whose result has v's type: MINS MAXS MINW MAXW DELTAS PREV) and `reduce` folds the vector into
 * one cell per group with AQ_MAX / AQ_MIN (server/aggregations.h:71-86) instead of writing it.
 * Vector results: nrows cells in CSR order (like aqgpu_group_scan); reduced results: ngroups cells of v's type.
 * All entries share one pass over the rows: they are brought into group order once, with their values. */
typedef struct aqgpu_window {
    int op;            /* aqgpu_scan_op */
    uint32_t w;        /* window length for AQ_SUMW / AVGW / MINW / MAXW / RATIOW, else 0 */
    int vtype;
    const void* v;     /* nrows cells, table row order; device or host */
    int minus_from_v;  /* 0 / 1 */
    int reduce;        /* -1: vector result; AQ_MAX or AQ_MIN: one cell per group */
    int out_type;      /* 0 = default of the op (aqgpu_scan_out_type) */
    void* out;         /* device or host */
} aqgpu_window_t;
int aqgpu_group_window(aqgpu_groups_t* g, const aqgpu_window_t* ops, int n_ops);
/* The same with ASSUMING order evaluated inside the call: every group's rows are first ordered by order_col (AQ_INT32,
 * AQ_UINT32 or AQ_FLOAT; desc != 0: descending; ties keep the row lists' order), exactly as aqgpu_groups_order_by +
 * aqgpu_group_window would (common/groupby.py:54-59 -> TableInfo::order_by, server/table.h:447-461), but for groups of
 * up to 8192 rows the key rides through the group scatter as a tuple field and every group is sorted in one SM's shared
 * memory -- no global sort passes, no gathers through a row list.  Other shapes take the two-call route internally
 * (the handle's row lists are then left re-ordered; on the fused route they are untouched). */
int aqgpu_group_window_ordered(aqgpu_groups_t* g, const void* order_col, int order_type, int desc,
                               const aqgpu_window_t* ops, int n_ops);
/* out[offsets[g] + j] = v[row j of group g]  -- the gather col[vecs[g]] of server/table.h:184-189 for
 * ALL groups at once, in the CSR order of the row lists.  out has nrows cells of vtype. */
int aqgpu_group_gather(aqgpu_groups_t* g, int vtype, const void* v, void* out);
/* sqrt(v) -> AQ_DOUBLE cells (float cells: the root is taken in float, then widened, as the reference's overload resolution
 * does); truncate(v, precision) -> cells of v's type: round(v * 10^p) / 10^p in double, half away from zero, a copy for
 * integer cells and when precision >= 7 (float) / 16 (double) (server/aggregations.h:34-69, server/types.h:473-475).
 * param = precision for AQ_U_TRUNCATE. */
enum aqgpu_vec_unary_op { AQ_U_SQRT = 0, AQ_U_TRUNCATE = 1 };
int aqgpu_vec_unary(int op, int type, const void* x, uint32_t param, void* out, uint32_t n);
/* out[j] = col[idx[j]], j < nidx: the gather ColRef<T>::operator[](vector_type<uint32_t>&) (server/table.h:184-189) and
 * TableInfo::materialize (server/table.h:429-440) for an explicit index vector.  Cells of 1, 2, 4, 8, 12 or 16 bytes;
 * an index >= col_rows fails the call (the reference reads out of bounds). */
int aqgpu_take(int elem_bytes, const void* col, uint32_t col_rows, const uint32_t* idx, uint32_t nidx, void* out);
/* ASSUMING-ordered evaluation for unsorted input (common/groupby.py:54-59 ->
 * TableInfo::order_by, server/table.h:447-461): re-orders every group's row list by the given
 * columns (desc[c] != 0 = descending), ties by the current order.  Integer / float keys. */
int aqgpu_groups_order_by(aqgpu_groups_t* g, const void* const* cols, const int* col_types,
                          const int* desc, int ncols);

/* ---- element-wise helpers used by generated code around the aggregates ----------------------- */
/* out[i] = a[i] - b[i] (same dtype); q7 max(v1) - min(v2) (server/table.h:820-973 vector arith). */
int aqgpu_vec_sub(int dtype, const void* a, const void* b, void* out, uint32_t n);
/* out[i] = a[i] op b[i]: the element-wise operators of server/vector_type.hpp:372-396 and server/table.h:826-973
 * (`price - mins(price)`, `deltas(_time) > 120`, ...).  Operand cells: AQ_INT32, AQ_INT64, AQ_FLOAT, AQ_DOUBLE (and
 * AQ_BOOL, promoted to int like C++ does); with x_scalar != 0 the operand is a pointer to ONE host cell.  Every
 * cell is computed like the C++ expression `lhs[i] op rhs[i]` (usual arithmetic conversions: int32 wraps, int / int
 * is integer division, int64 with float is float arithmetic) and converted to out_type.  out_type < 0 selects the
 * element type the reference's templates give the result, aqgpu_vec_out_type(): Coercion<T1,T2> for vector-vector
 * forms; with a scalar operand `*` gives GetLongType (AQ_INT128 | AQ_DOUBLE) and `/` GetFPType (AQ_FLOAT for int32);
 * comparisons give AQ_BOOL.  An explicit out_type is the `Ret&` of aqop_x(lhs, rhs, ret) (table.h:949-973).
 * Deviations: the reference's `vector < scalar` (<=, >=, ==, !=) all evaluate `>` (vector_type.hpp:372-385);
 * here each comparison means what it says.  Integer division by zero gives 0 (the reference traps). */
enum aqgpu_vec_op {
    AQ_V_ADD = 0, AQ_V_SUB = 1, AQ_V_MUL = 2, AQ_V_DIV = 3, AQ_V_AND = 4, AQ_V_OR = 5, AQ_V_XOR = 6,
    AQ_V_GT = 7, AQ_V_LT = 8, AQ_V_GTE = 9, AQ_V_LTE = 10, AQ_V_EQ = 11, AQ_V_NEQ = 12
};
int aqgpu_vec_out_type(int op, int a_type, int a_scalar, int b_type, int b_scalar);
int aqgpu_vec_binary(int op, int a_type, const void* a, int a_scalar, int b_type, const void* b, int b_scalar,
                     int out_type, void* out, uint32_t n);
/* out[i] = sum[i] / double(count[i]): AVG finalisation after a multi-GPU merge. sum is AQ_INT64,
 * AQ_INT128 or AQ_DOUBLE. */
int aqgpu_avg_finalize(int sum_type, const void* sum, const uint32_t* count, double* out, uint32_t n);
/* Finalisation of merged partial states (multi-GPU); n cells each.
 *   AQ_AVG     out f64 = s[0] / count                                  aggregations.h:28-32
 *   AQ_VAR / AQ_STDDEV   from s[0] = sum, s[1] = sum of squares        aggregations.h:332-348, 413-415
 *   AQ_CORR    from s = {sum x, sum y, sum xy, sum xx, sum yy}          aggregations.h:383-407
 *   AQ_COUNT   out u32 = count
 * count is AQ_INT64; every s[k] has dtype s_type (AQ_INT64, AQ_INT128 or AQ_DOUBLE). */
int aqgpu_finalize(int op, uint32_t n, const int64_t* count, int s_type, const void* const* s, void* out);
/* out[i] = (int64) in[i] + add: local first rows -> global row ids of a shard. */
int aqgpu_rows_to_global(const uint32_t* in, int64_t add, int64_t* out, uint32_t n);
/* owner[i] = hash(key tuple of row i) % nparts: the rank that merges the group (hash-partitioned
 * exchange of partial aggregates). */
int aqgpu_key_owner(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                    uint32_t nparts, uint32_t* owner);
/* Stable split of device-resident columns by owner[i] in [0, nparts), nparts <= 64: out_cols[k] receives
 * column k with the rows of owner 0 first, then owner 1, ..., each in their original order; counts[p]
 * (host) = rows of owner p.  elem_bytes[k] in {4, 8, 16}.  The step before the all-to-all exchange. */
int aqgpu_partition_rows(const uint32_t* owner, uint32_t nrows, uint32_t nparts, const void* const* cols,
                         const int* elem_bytes, int ncols, void* const* out_cols, uint32_t* counts);
/* Allocation-free variant of the counting step of aqgpu_partition_rows: out_counts[p] (DEVICE memory, nparts x
 * uint32) = rows of owner p.  No counterpart in the reference; added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The
 * call performs no cudaMalloc / cudaFree / pool allocation: its scratch is a static device array of 2 MiB
 * (64 owners x 8192 chunks of 2048 rows) that every GPU loading the library holds until exit.
 *   limits: nrows <= 16,777,216 and nparts <= 64, beyond that the call fails with a message (use
 *           aqgpu_partition_rows); owner and out_counts must be device pointers; aqgpu_init must have run.
 *   Asynchronous on the library stream (aqgpu_sync / stream order before reading out_counts).
 *   Every call shares the one scratch array: calls that overlap on the GPU race -- serialise them.  The first
 *   call on a device resolves the array's address (cudaGetSymbolAddress); make it outside a graph capture. */
int aqgpu_partition_counts_noalloc(const uint32_t* owner, uint32_t nrows, uint32_t nparts, uint32_t* out_counts);
/* AQ_SECOND across shards: out = cnt_top >= 2 ? second_top : (nparts >= 2 ? first_prev : 0), cells of
 * vtype; cnt_top and nparts are AQ_INT64 columns. */
int aqgpu_second_merge(int vtype, const int64_t* cnt_top, const int64_t* nparts, const void* second_top,
                       const void* first_prev, void* out, uint32_t n);
/* ---- key columns that are not plain integers -----------------------------------------
 * The reference groups date / time / timestamp / string columns through AQHashTable with the hashes of
 * server/hasher.h:97-132 and the types' operator== (server/libaquery.cpp:112-114,162-164,201-203,
 * server/types.h:299-309).  Only equality and first-appearance order are observable, so such a column is
 * lowered to integer key columns that are equal exactly when operator== holds, and passed to the group-by
 * entry points above:
 *   AQ_KEY_DATE       types::date_t, 4-byte cells {day, month, year}     -> out_a: AQ_UINT32 (the image; the
 *                     column may equally be passed as AQ_UINT32 without calling anything)
 *   AQ_KEY_TIME       types::time_t, 8-byte cells {ms, s, m, h, padding} -> out_a: AQ_INT64, padding cleared
 *   AQ_KEY_TIMESTAMP  types::timestamp_t, 12-byte cells {date, time}     -> out_a: AQ_UINT32, out_b: AQ_INT64
 * aqgpu_key_join is the inverse, for the key columns of a result (padding bytes come back as zero). */
enum aqgpu_key_kind { AQ_KEY_DATE = 1, AQ_KEY_TIME = 2, AQ_KEY_TIMESTAMP = 3 };
size_t aqgpu_key_cell_size(int key_kind);
int aqgpu_key_split(int key_kind, const void* col, uint32_t nrows, void* out_a, void* out_b);
int aqgpu_key_join(int key_kind, const void* a, const void* b, uint32_t nrows, void* out);
/* String key column -> dictionary codes: out_codes[i] = index of row i's string among the distinct strings in
 * FIRST-APPEARANCE order, i.e. the group id AQHashTable<astring_view> gives it (server/hasher.h:97-106,176-179);
 * out_first_row[c] = first row holding string c (may be NULL; room for max_distinct cells, else a "capacity"
 * error).  Strings are Arrow-style: row i = bytes[offsets[i] .. offsets[i+1]) (no terminator; the reference's
 * astring_view is a char pointer into the data source's heap, which packing replaces at ingest).  Two 64-bit
 * hashes per string are grouped by the integer path and every row is then compared byte for byte with its
 * group's first row: a hash collision fails the call, it never merges two strings.  Group by the codes
 * (AQ_UINT32 keys) alone or with other key columns; map result codes back through out_first_row. */
int aqgpu_string_dict(const void* bytes, uint64_t nbytes, const uint64_t* offsets, uint32_t nrows,
                      uint32_t* out_codes, uint32_t max_distinct, uint32_t* out_first_row, uint32_t* ndistinct);
/* out[i] = pow(x[i], p) in fp64 (q9 pow(corr, 2); aggregations.h:409-411). */
int aqgpu_vec_pow(const double* x, double p, double* out, uint32_t n);

/* ---- synthetic column generators (bench / tests; same streams as aquery2_b200/synth.py) ------ */
/* kind: 0 = 1+U[0,k) int32, 1 = U[0,k) int32, 2 = U[0,k)/1e6 as float32 (k = 1e8 gives
 * round(U(0,100),6)), 3 = row+1 int32, 4 = trade price in [50,500] (key1 = symbol stream, key2 =
 * per-symbol base key, k = number of symbols), 5 = aux_k + U[0,k) int32.
 * Rows are GLOBAL ids [row_lo, row_lo+n); key0 = stream key of the column (synth.stream_key). */
int aqgpu_synth_column(int kind, uint64_t key0, uint64_t key1, uint64_t key2, uint32_t k, uint32_t aux_k,
                       uint64_t row_lo, uint32_t n, void* d_out);

#ifdef __cplusplus
}
#endif
#endif /* AQGPU_H */
