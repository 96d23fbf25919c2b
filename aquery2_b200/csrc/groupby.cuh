// groupby.cuh -- internal structures of the group-by engine (groupby.cu, groupby_smem.cu, groups.cu).
#pragma once
#include "common.cuh"

struct aqgpu_result;

namespace aq {

constexpr int MAX_KEY_COLS = 12;
constexpr int MAX_KEY_WORDS = 12;     // 32-bit words per key tuple (an int64 column is 2 words)
constexpr int MAX_ACC = 15;           // accumulator cells per group (slot = 1 + cells, <= 16 words)
constexpr int MAX_AGGS = 16;
constexpr unsigned long long SLOT_EMPTY = 0xffffffffffffffffULL;
constexpr uint32_t NOSLOT = 0xffffffffu;

// flags[] protocol between kernels and host (ctx().d_flags)
enum : int {
    FLAG_ABORT = 0,     // 0 ok, 1 table too full (grow), 2 CTA-local table overflow (use global path),
                        // 3 key outside the supplied bounds (drop the bounds), 4 fp fixed-point inexact
    FLAG_NGROUPS = 1
};

// Key tuple as a list of 32-bit words: word w of row i is base[w][i * stride[w]].
struct KeySpec {
    const uint32_t* base[MAX_KEY_WORDS];
    uint32_t stride[MAX_KEY_WORDS];
    int nwords;
    int vec_ok;          // every stride-1 column is 16-byte aligned: 128-bit loads allowed
    // exact 32-bit image of the key tuple ("tag"): 0 = none (tag = hash bits, verified against the
    // key columns), 1 = the single 32-bit key itself, 2 = bit-packed (k - lo) fields, the scheme of
    // the reference's PerfectHashTable (server/hasher.h:226-243) from per-column bounds.
    int exact;
    uint32_t pack_lo[MAX_KEY_WORDS];
    uint32_t pack_range[MAX_KEY_WORDS];   // hi - lo
    uint32_t pack_shift[MAX_KEY_WORDS];
    int pack_bits;
    // "wide" exact image for tuples whose packed fields exceed 32 bits but fit 96 (h2o q10: 76 bits):
    // every column goes whole into the 32-bit tag (wide_part = 0) or into a 64-bit second word kept in
    // slot word 1 (wide_part = 1), so membership is decided by the slot's first 16 bytes alone
    // (k_probe_wide: one 128-bit CAS inserts both) -- no gathers from the key columns.
    int wide;
    uint32_t wide_part[MAX_KEY_WORDS];
    uint32_t wide_shift[MAX_KEY_WORDS];
};

// One 64-bit accumulator cell per group and AccOp.
enum AccKind : int {
    ACC_COUNT = 0,     // += 1
    ACC_ADD_I64 = 1,   // += (int64) x
    ACC_ADD_F64 = 2,   // += (double) x
    ACC_MIN_I64 = 3,   // min of the int64 image (ints: value, fp: order-preserving bits of the double)
    ACC_MAX_I64 = 4,
    ACC_MAXROW = 5,    // max row id of the group
    ACC_TOP2ROW = 6    // (best << 32 | second best) row keys; key = row + 1 (DESC lists: two largest
                       // rows) or 0xffffffff - row (ASC lists: two smallest rows); 0 = none
};
enum AccXform : int {
    X_IDENT = 0,       // x = v[i]
    X_SQUARE = 1,      // x = v[i] * v[i]   (fp: product rounded in the column's own type, as
                       //                    aggregations.h:339-345 computes arr[i]*arr[i] in T)
    X_PRODUCT = 2      // x = v[i] * v2[i]
};
struct AccOp {
    int kind;
    int xform;
    int vtype;
    int vec_ok;        // v (and v2) 16-byte aligned
    const void* v;
    const void* v2;
    int ci, ci2;       // index of v / v2 in AccProg::vcol (-1: none or table full)
};
constexpr int MAX_VCOL = 4;
// Global table: cap slots of `sw` 64-bit words; word 0 = (tag << 32 | min row) or SLOT_EMPTY,
// word 1 + a = accumulator a.  sw is a power of two so that a slot never straddles a 32 B sector
// boundary more than necessary (sw = 4 -> exactly one sector).
struct AccProg {
    AccOp op[MAX_ACC];
    int n;
    unsigned long long* tab;
    uint32_t sw;
    // distinct value columns (kernels that load every column once per row batch, up front)
    const void* vcol[MAX_VCOL];
    int vctype[MAX_VCOL];
    int nv;            // -1: more than MAX_VCOL distinct columns
};
__host__ __device__ __forceinline__ unsigned long long* slot_word(const AccProg& p, uint32_t slot) {
    return p.tab + size_t(slot) * p.sw;
}

// How one requested aggregate is produced from accumulator cells at emit time.
struct EmitAgg {
    int op;            // aqgpu_reduce_op
    int vtype;
    int out_type;
    int a[5];          // accumulator indices (meaning depends on op)
    int cnt;           // index of the COUNT accumulator or -1
    int maxrow;        // index of the MAXROW accumulator or -1
    int top2;          // index of the TOP2ROW accumulator or -1
    const void* v;     // value column for FIRST / LAST gathers
    void* out;
};
struct EmitProg {
    EmitAgg agg[MAX_AGGS];
    int n;
    int row_order;
};

struct KeyOut {
    const void* src[MAX_KEY_COLS];
    void* dst[MAX_KEY_COLS];
    int size[MAX_KEY_COLS];    // 4 or 8
    int n;
};

// Host-side description of the key columns after pointer binding.
struct KeyCols {
    int n = 0;
    int type[MAX_KEY_COLS];
    InCol col[MAX_KEY_COLS];
    int bind(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows);
    int make_spec(const int64_t* key_min, const int64_t* key_max, KeySpec& ks) const;
};

struct AggPlan {
    AccProg prog;
    EmitProg emit;
    int n_acc = 0;
    AggPlan();
    int add_acc(int kind, int xform, int vtype, const void* v, const void* v2);
    int add_count() { return add_acc(ACC_COUNT, X_IDENT, 0, nullptr, nullptr); }
    int add_maxrow() { return add_acc(ACC_MAXROW, X_IDENT, 0, nullptr, nullptr); }
    int add_top2(int row_order) { return add_acc(ACC_TOP2ROW, row_order, 0, nullptr, nullptr); }
    int find(int kind) const;
    int plan(const aqgpu_agg_t& a, const void* dv, const void* dv2, void* dout);
};

// Where the result columns of one fused group-by go: caller-provided columns (aqgpu_groupby_agg: device or
// host pointers with room for max_groups cells) or columns the library allocates once it knows the group
// count (aqgpu_groupby_run -> struct aqgpu_result).  prepare() is called after the group count (or an upper
// bound of it) is known and before anything is emitted.
struct OutSink {
    // caller-provided
    const aqgpu_agg_t* aggs = nullptr;
    void* const* out_keys = nullptr;
    uint32_t* out_first_row = nullptr;
    uint32_t max_groups = 0;
    std::vector<OutCol> vout, kout;
    OutCol fr;
    // library-allocated
    ::aqgpu_result* res = nullptr;
    bool want_first_row = false;
    // common
    int n_aggs = 0, nkeys = 0;
    const int* key_types = nullptr;
    uint32_t cap = 0;                     // cells every destination below can hold
    void* agg_dst[MAX_AGGS] = {};
    void* key_dst[MAX_KEY_COLS] = {};
    uint32_t* fr_dst = nullptr;
    int prepare(uint32_t cap_cells, const AggPlan& plan);
    int finish(uint32_t ngroups, const AggPlan& plan);
};

// Sampling planner + the fast paths it can choose (plan.cu): low-cardinality direct tables (lc.cu) and the
// fingerprint-partitioned path for groups ~ rows (hcfp.cu).  *used = false: fall through to hash_aggregate.
// AQGPU_DISABLE=lc,hcfp,plan,smem,part,hc switches paths off at run time (A/B measurements, fault isolation)
bool path_disabled(const char* name);
struct PlanHints {                        // what the generic paths get when the planner declines
    uint32_t groups_hint = 0;
    bool have_bounds = false;
    int64_t kmin[MAX_KEY_COLS], kmax[MAX_KEY_COLS];
};
int run_planned_paths(KeyCols& kc, uint32_t nrows, const int64_t* key_min, const int64_t* key_max,
                      uint32_t groups_hint, AggPlan& plan, OutSink& sink, uint32_t* ngroups, bool* used,
                      PlanHints* hints_out);
int plan_hints_only(KeyCols& kc, uint32_t nrows, const int64_t* key_min, const int64_t* key_max, uint32_t groups_hint,
                    PlanHints* out);

// Result of the hash phase, consumed by emit / handle construction.
struct HashResult {
    uint32_t cap = 0;
    uint32_t sw = 0;
    KeySpec ks;                   // the key spec the table was finally built with (bounds may have been dropped)
    DevBuf tab;                   // cap x sw u64
    DevBuf bitmap;                // nrows bits: first-appearance rows
    DevBuf wordprefix;            // per bitmap word: number of groups before it
    // fused path (global table only): every row's slot, ranked and emitted by ONE row-order pass
    // (k_emit_fused) instead of k_mark / k_rank_* / k_emit_rows; ngroups is known after the emit.
    bool fused = false;
    int acc_off = 1;              // accumulator a lives in slot word acc_off + a (2 for wide keys)
    DevBuf row_slot;
    DevBuf rflags;                // groupby_hc.cu: per row 0 = first row of a one-row group (no record written),
                                  // 1 = first row with a record in tab, 2 = not a first row; empty = records only
};

// Hash build + fused accumulation + ranking of the groups by first row.  Synchronises once to
// return *ngroups.  d_row_slot (optional, nrows) receives every row's table slot.
int hash_aggregate(KeySpec ks, uint32_t nrows, uint32_t groups_hint, AggPlan& plan,
                   uint32_t* d_row_slot, HashResult& hr, uint32_t* ngroups, bool allow_fused = false);
// fused rank + emit of a HashResult with hr.fused set; synchronises to return *ngroups
int emit_fused(const KeySpec& ks, uint32_t nrows, const HashResult& hr, const KeyOut& ko, uint32_t* d_first_row,
               const AggPlan& plan, uint32_t max_groups, uint32_t* ngroups);

// wordprefix[w] = set bits of bitmap before word w; total -> flags[FLAG_NGROUPS] (groupby.cu)
int bitmap_rank(const uint32_t* bitmap, uint32_t nwords, uint32_t* wordprefix);

// Writes keys / first rows / aggregates in first-appearance order (gid = rank of the first row).
int emit_groups(const KeySpec& ks, uint32_t nrows, const HashResult& hr, const KeyOut& ko, uint32_t* d_first_row,
                uint32_t* d_counts, uint32_t* d_last_row, uint32_t* d_slot_gid, const AggPlan& plan, uint32_t max_groups);

// smem-privatised scan (groupby_smem.cu).  *used = false if the request does not qualify.
int launch_smem_probe(const KeySpec& ks, uint32_t nrows, uint32_t cap, uint32_t max_probe,
                      const AccProg& prog, uint32_t groups_hint, bool* used);
// partitioned aggregation (groupby_part.cu) for groups beyond one SM's shared memory; same contract.
int launch_part_probe(const KeySpec& ks, uint32_t nrows, uint32_t cap, uint32_t max_probe,
                      const AccProg& prog, uint32_t groups_hint, bool* used);
// high-cardinality path for wide exact keys (groupby_hc.cu): rows -> hash partitions -> per-CTA
// L2-resident tables -> one record per ROW (rec[i] = the slot of the group whose first row is i,
// else all ones), which k_emit_fused then ranks and emits.  *used = false if it does not qualify.
bool hc_qualifies(const KeySpec& ks, uint32_t nrows, const AccProg& prog, uint32_t sw, uint32_t groups_hint);
int launch_hc(const KeySpec& ks, uint32_t nrows, const AccProg& prog, uint32_t sw, const unsigned long long* pattern,
              uint32_t groups_hint, DevBuf& rec, DevBuf& rflags, bool* used);

#ifdef __CUDACC__
// ---- device helpers shared by the probe kernels ------------------------------------------------
__device__ __forceinline__ void load4_b32(const void* base, uint32_t row0, int cnt, bool vec, uint32_t out[4]) {
    const uint32_t* p = static_cast<const uint32_t*>(base) + row0;
    if (vec && cnt == 4) {
        int4 t = ld_stream_v4(p);
        out[0] = t.x; out[1] = t.y; out[2] = t.z; out[3] = t.w;
    } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) out[r] = r < cnt ? ld_stream_u32(p + r) : 0u;
    }
}
__device__ __forceinline__ void load4_b64(const void* base, uint32_t row0, int cnt, bool vec, unsigned long long out[4]) {
    const unsigned long long* p = static_cast<const unsigned long long*>(base) + row0;
    if (vec && cnt == 4) {
        int4 t0 = ld_stream_v4(p), t1 = ld_stream_v4(p + 2);
        out[0] = (unsigned long long)(uint32_t)t0.x | ((unsigned long long)(uint32_t)t0.y << 32);
        out[1] = (unsigned long long)(uint32_t)t0.z | ((unsigned long long)(uint32_t)t0.w << 32);
        out[2] = (unsigned long long)(uint32_t)t1.x | ((unsigned long long)(uint32_t)t1.y << 32);
        out[3] = (unsigned long long)(uint32_t)t1.z | ((unsigned long long)(uint32_t)t1.w << 32);
    } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) out[r] = r < cnt ? __ldg(p + r) : 0ull;
    }
}

__device__ __forceinline__ bool keys_equal(const KeySpec& ks, uint32_t a, uint32_t b) {
    for (int w = 0; w < ks.nwords; ++w) {
        const uint32_t* p = ks.base[w];
        if (__ldg(p + size_t(a) * ks.stride[w]) != __ldg(p + size_t(b) * ks.stride[w])) return false;
    }
    return true;
}

// Loads the key words of 4 consecutive rows; produces the 32-bit tag and the slot hash of each.
// Returns false if a key lies outside the supplied bounds (packed mode).
__device__ __forceinline__ bool key_tags4(const KeySpec& ks, uint32_t row0, int cnt, uint32_t tag[4], uint32_t hsh[4]) {
    bool ok = true;
    if (ks.exact) {
        uint32_t t[4] = {0, 0, 0, 0};
        for (int w = 0; w < ks.nwords; ++w) {
            uint32_t kw[4];
            load4_b32(ks.base[w], row0, cnt, ks.vec_ok, kw);
            if (ks.exact == 1) {
#pragma unroll
                for (int r = 0; r < 4; ++r) t[r] = kw[r];
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t d = kw[r] - ks.pack_lo[w];
                    ok &= (r >= cnt) || (d <= ks.pack_range[w]);
                    t[r] |= d << ks.pack_shift[w];
                }
            }
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) { tag[r] = t[r]; hsh[r] = mix32(t[r]); }
    } else {
        uint64_t h[4] = {0x9e3779b97f4a7c15ULL, 0x9e3779b97f4a7c15ULL, 0x9e3779b97f4a7c15ULL, 0x9e3779b97f4a7c15ULL};
        for (int w = 0; w < ks.nwords; ++w) {
            uint32_t kw[4];
            if (ks.stride[w] == 1) {
                load4_b32(ks.base[w], row0, cnt, ks.vec_ok, kw);
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r) kw[r] = r < cnt ? __ldg(ks.base[w] + size_t(row0 + r) * ks.stride[w]) : 0u;
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) h[r] = mix64(h[r] ^ kw[r]);
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) { tag[r] = (uint32_t)(h[r] >> 32); hsh[r] = (uint32_t)h[r]; }
    }
    return ok;
}

// tag / slot hash of ONE row (same values as key_tags4 produces for it); bounds are not re-checked.
__device__ __forceinline__ void key_tag1(const KeySpec& ks, uint32_t row, uint32_t& tag, uint32_t& hsh) {
    if (ks.exact) {
        uint32_t t = 0;
        for (int w = 0; w < ks.nwords; ++w) {
            const uint32_t kw = __ldg(ks.base[w] + row);
            if (ks.exact == 1) t = kw;
            else t |= (kw - ks.pack_lo[w]) << ks.pack_shift[w];
        }
        tag = t; hsh = mix32(t);
    } else {
        uint64_t h = 0x9e3779b97f4a7c15ULL;
        for (int w = 0; w < ks.nwords; ++w) h = mix64(h ^ __ldg(ks.base[w] + size_t(row) * ks.stride[w]));
        tag = (uint32_t)(h >> 32); hsh = (uint32_t)h;
    }
}

__device__ __forceinline__ void cas128(unsigned long long* p, unsigned long long c0, unsigned long long c1,
                                       unsigned long long n0, unsigned long long n1,
                                       unsigned long long& o0, unsigned long long& o1) {
    asm volatile("{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 n, {%4, %5};\n\t"
                 "atom.relaxed.gpu.global.cas.b128 o, [%6], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
                 : "=l"(o0), "=l"(o1) : "l"(c0), "l"(c1), "l"(n0), "l"(n1), "l"(p) : "memory");
}
__device__ __forceinline__ void ld128_volatile(const unsigned long long* p, unsigned long long& a, unsigned long long& b) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ uint32_t wide_hash(uint32_t tag, unsigned long long hi) {
    return (uint32_t)(mix64(hi * 0x9e3779b97f4a7c15ULL + tag) >> 16);
}
// wide exact image of 4 consecutive rows (KeySpec::wide, at most WIDE_MAX_COLS columns): every column's
// 128-bit load is issued before the first one is used.  false if a key lies outside the bounds.
constexpr int WIDE_MAX_COLS = 8;
__device__ __forceinline__ bool wide_pack4(const KeySpec& ks, uint32_t row0, int cnt, uint32_t tag[4], unsigned long long hi[4]) {
    uint32_t kw[WIDE_MAX_COLS][4];
#pragma unroll
    for (int w = 0; w < WIDE_MAX_COLS; ++w) {
        if (w < ks.nwords) load4_b32(ks.base[w], row0, cnt, ks.vec_ok, kw[w]);
        else { kw[w][0] = kw[w][1] = kw[w][2] = kw[w][3] = 0u; }
    }
    bool ok = true;
#pragma unroll
    for (int r = 0; r < 4; ++r) { tag[r] = 0; hi[r] = 0; }
#pragma unroll
    for (int w = 0; w < WIDE_MAX_COLS; ++w) {
        if (w >= ks.nwords) break;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const uint32_t d = kw[w][r] - ks.pack_lo[w];
            ok &= (r >= cnt) || (d <= ks.pack_range[w]);
            if (ks.wide_part[w]) hi[r] |= (unsigned long long)d << ks.wide_shift[w];
            else tag[r] |= d << ks.wide_shift[w];
        }
    }
    return ok;
}

// find-or-insert in the global table; keeps the smallest row id in the low half of word 0.
template <bool EXACT>
__device__ __forceinline__ uint32_t probe_insert(const KeySpec& ks, unsigned long long* tab, uint32_t sw,
                                                 uint32_t mask, uint32_t max_probe, uint32_t h, uint32_t tag,
                                                 uint32_t row) {
    const unsigned long long mine = ((unsigned long long)tag << 32) | row;
    uint32_t slot = h & mask;
    for (uint32_t step = 0; step <= max_probe; ++step) {
        unsigned long long* w0 = tab + size_t(slot) * sw;
        unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(w0);
        if (cur == SLOT_EMPTY) {
            const unsigned long long old = atomicCAS(w0, SLOT_EMPTY, mine);
            if (old == SLOT_EMPTY) return slot;
            cur = old;
        }
        if ((uint32_t)(cur >> 32) == tag) {
            if (EXACT || keys_equal(ks, row, (uint32_t)cur)) {
                if (row < (uint32_t)cur) atomicMin(w0, mine);
                return slot;
            }
        }
        slot = (slot + 1) & mask;
    }
    return NOSLOT;
}
// value column -> 4 (int64 image | double image) values
__device__ __forceinline__ void load4_val(const void* v, int vtype, uint32_t row0, int cnt, bool vec,
                                          long long iv[4], double dv[4], float fv[4]) {
    if (vtype == AQ_INT32 || vtype == AQ_FLOAT) {
        uint32_t u[4];
        load4_b32(v, row0, cnt, vec, u);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (vtype == AQ_INT32) { iv[r] = (long long)(int)u[r]; }
            else { fv[r] = __uint_as_float(u[r]); dv[r] = (double)fv[r]; }
        }
    } else {
        unsigned long long u[4];
        load4_b64(v, row0, cnt, vec, u);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (vtype == AQ_INT64) iv[r] = (long long)u[r];
            else dv[r] = __longlong_as_double((long long)u[r]);
        }
    }
}

// x[r] for one accumulator: integer image in xi (int columns), double image in xd (fp columns)
__device__ __forceinline__ void acc_inputs(const AccOp& o, uint32_t row0, int cnt, long long xi[4], double xd[4]) {
    long long iv[4] = {0, 0, 0, 0}; double dv[4] = {0, 0, 0, 0}; float fv[4] = {0, 0, 0, 0};
    load4_val(o.v, o.vtype, row0, cnt, o.vec_ok, iv, dv, fv);
    if (o.xform == X_IDENT) {
#pragma unroll
        for (int r = 0; r < 4; ++r) { xi[r] = iv[r]; xd[r] = dv[r]; }
        return;
    }
    long long iw[4] = {0, 0, 0, 0}; double dw[4] = {0, 0, 0, 0}; float fw[4] = {0, 0, 0, 0};
    if (o.xform == X_PRODUCT) {
        load4_val(o.v2, o.vtype, row0, cnt, o.vec_ok, iw, dw, fw);
    } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) { iw[r] = iv[r]; dw[r] = dv[r]; fw[r] = fv[r]; }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        xi[r] = iv[r] * iw[r];
        // the reference multiplies in the column's own type before widening (aggregations.h:339-345)
        xd[r] = o.vtype == AQ_FLOAT ? (double)__fmul_rn(fv[r], fw[r]) : __dmul_rn(dv[r], dw[r]);
    }
}

// packed top-2 cell update (see ACC_TOP2ROW): k = row key, larger is better
__device__ __forceinline__ void top2_update(unsigned long long* c, unsigned long long k) {
    unsigned long long old = *reinterpret_cast<volatile unsigned long long*>(c);
    for (;;) {
        const unsigned long long best = old >> 32, second = old & 0xffffffffull;
        unsigned long long nw;
        if (k > best) nw = (k << 32) | best;
        else if (k > second && k != best) nw = (best << 32) | k;
        else break;
        const unsigned long long seen = atomicCAS(c, old, nw);
        if (seen == old) break;
        old = seen;
    }
}
__device__ __forceinline__ unsigned long long top2_key(int row_order, uint32_t row) {
    return row_order == AQ_ROWS_DESC ? (unsigned long long)row + 1ull : (unsigned long long)(0xffffffffu - row);
}

// one row's contribution to accumulator cell c of the GLOBAL table
__device__ __forceinline__ void acc_global_row(const AccOp& o, unsigned long long* c, uint32_t row, long long xi, double xd) {
    const bool isint = (o.vtype == AQ_INT32 || o.vtype == AQ_INT64);
    switch (o.kind) {
    case ACC_COUNT: atomicAdd(c, 1ull); break;
    case ACC_MAXROW: atomicMax(c, (unsigned long long)row); break;
    case ACC_TOP2ROW: top2_update(c, top2_key(o.xform, row)); break;
    case ACC_ADD_I64: atomicAdd(c, (unsigned long long)xi); break;
    case ACC_ADD_F64: atomicAdd(reinterpret_cast<double*>(c), isint ? (double)xi : xd); break;
    case ACC_MIN_I64: atomicMin(reinterpret_cast<long long*>(c), isint ? xi : f64_to_ordered(xd)); break;
    case ACC_MAX_I64: atomicMax(reinterpret_cast<long long*>(c), isint ? xi : f64_to_ordered(xd)); break;
    default: break;
    }
}
// x of accumulator o for ONE row (same values acc_inputs produces for it): integer image xi, double image xd
__device__ __forceinline__ void acc_row_x(const AccOp& o, uint32_t row, long long& xi, double& xd) {
    xi = 0; xd = 0.0;
    auto ld = [&](const void* v, long long& i, double& d, float& f) {
        i = 0; d = 0.0; f = 0.f;
        switch (o.vtype) {
        case AQ_INT32: i = (long long)static_cast<const int*>(v)[row]; break;
        case AQ_INT64: i = static_cast<const long long*>(v)[row]; break;
        case AQ_FLOAT: f = static_cast<const float*>(v)[row]; d = (double)f; break;
        default: d = static_cast<const double*>(v)[row]; break;
        }
    };
    long long i1, i2; double d1, d2; float f1, f2;
    ld(o.v, i1, d1, f1);
    if (o.xform == X_IDENT) { xi = i1; xd = d1; }
    else {
        if (o.xform == X_PRODUCT) ld(o.v2, i2, d2, f2); else { i2 = i1; d2 = d1; f2 = f1; }
        xi = i1 * i2;
        xd = o.vtype == AQ_FLOAT ? (double)__fmul_rn(f1, f2) : __dmul_rn(d1, d2);
    }
}
// the cell of accumulator o after exactly ONE row (a one-row group): identity (+) that row, no atomics
__device__ __forceinline__ unsigned long long acc_single_row(const AccOp& o, uint32_t row) {
    if (o.kind == ACC_COUNT) return 1ull;
    if (o.kind == ACC_MAXROW) return (unsigned long long)row;
    if (o.kind == ACC_TOP2ROW) return top2_key(o.xform, row) << 32;
    long long xi = 0; double xd = 0.0;
    const bool isint = (o.vtype == AQ_INT32 || o.vtype == AQ_INT64);
    acc_row_x(o, row, xi, xd);
    switch (o.kind) {
    case ACC_ADD_I64: return (unsigned long long)xi;
    case ACC_ADD_F64: return (unsigned long long)__double_as_longlong(0.0 + (isint ? (double)xi : xd));
    case ACC_MIN_I64: case ACC_MAX_I64: return (unsigned long long)(isint ? xi : f64_to_ordered(xd));
    default: return 0ull;
    }
}
__device__ __forceinline__ bool acc_needs_value(const AccOp& o) {
    return o.kind != ACC_COUNT && o.kind != ACC_MAXROW && o.kind != ACC_TOP2ROW;
}

__device__ __forceinline__ void acc_apply(const AccProg& p, int a, const uint32_t slot[4], uint32_t row0, int cnt) {
    const AccOp& o = p.op[a];
    long long xi[4] = {0, 0, 0, 0}; double xd[4] = {0, 0, 0, 0};
    if (acc_needs_value(o)) acc_inputs(o, row0, cnt, xi, xd);
#pragma unroll
    for (int r = 0; r < 4; ++r)
        if (r < cnt) acc_global_row(o, slot_word(p, slot[r]) + 1 + a, row0 + r, xi[r], xd[r]);
}

#endif

} // namespace aq
