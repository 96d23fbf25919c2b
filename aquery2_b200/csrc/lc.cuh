// lc.cuh -- parameters of the low-cardinality path (lc.cu) and the statistics the planner hands to it (plan.cu).
#pragma once
#include "groupby.cuh"

namespace aq {

constexpr int LC_MAX_KEYS = 4;
constexpr int LC_MAX_COLS = 8;          // ring columns per CTA (keys + the values of its accumulators)
constexpr int LC_MAX_STAGES = 8;
constexpr uint32_t LC_MAX_G = 65536;    // slots of the direct table

enum LcKind : int { LCK_CNT = 0, LCK_ISUM1, LCK_ISUM2, LCK_FX, LCK_MIN32, LCK_MAX32, LCK_SEEN };

struct LcAcc {
    int kind, xform, vtype;
    int owner;              // CTA rank inside the cluster that keeps this accumulator
    int c0, c1;             // ring column of v / v2
    uint32_t off;           // byte offset of the first word array in the owner's table (arrays of G words, SoA)
    int fxB;                // LCK_FX: biased exponent of the fixed-point unit (cells.cuh)
    uint32_t gword;         // first u64 word of this accumulator in a group's global record
};

struct LcParams {
    uint32_t nrows, G, tile_rows, ntiles;
    int nkeys, nacc, cluster, stages;
    uint32_t klo[LC_MAX_KEYS], krange[LC_MAX_KEYS], kmul[LC_MAX_KEYS];
    const uint32_t* cols[LC_MAX_COLS];  // ring columns: the keys first, then every value column once (same ring in every CTA)
    int ncols;
    int fxQ, fxR;                       // LCK_FX: the shared cell counts units of 2^fxQ and takes mantissas shifted by <= fxR (cells.cuh)
    uint32_t isum_lim;                  // LCK_ISUM1: |x| <= isum_lim keeps a CTA's 32-bit cell exact (checked per row)
    LcAcc acc[MAX_ACC];
    unsigned long long* gtab;           // G records of gw u64 words
    uint32_t gw;
    uint32_t* flags;
};
struct LcKeys {
    int nkeys;
    const uint32_t* key[LC_MAX_KEYS];
    uint32_t klo[LC_MAX_KEYS], krange[LC_MAX_KEYS], kmul[LC_MAX_KEYS];
};
struct LcInit { unsigned long long w[2 * MAX_ACC]; };
struct HcIdentity { unsigned long long w[MAX_ACC]; };    // identity cell of every accumulator (hcfp.cu records)

// What the planner knows about the inputs: given by the caller (ColRef::stats, server/table.h:60-91) or taken
// from a strided sample inside the call.  Everything sampled is a speculation the kernels verify per row.
struct PlanStats {
    bool have_bounds = false;           // kmin / kmax valid for every key column
    bool bounds_given = false;          // ... and supplied by the caller (exact by contract)
    long long kmin[MAX_KEY_COLS], kmax[MAX_KEY_COLS];
    uint32_t sample_rows = 0, sample_distinct = 0;
    double est_groups = 0.0;            // 0 = unknown
    int fx_maxexp[MAX_ACC];             // float accumulators: largest biased exponent of x (after its xform); -1 = unusable
    unsigned long long isum_maxabs[MAX_ACC];   // integer accumulators: largest |x| (after its xform)
};

// marker_acc: index of a COUNT accumulator no output reads (it only tells empty slots apart), or -1
bool lc_layout(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, const PlanStats& st, int marker_acc, LcParams& P, LcKeys& K,
               int* threads_out, size_t* smem_out);
int run_lc(const KeyCols& kc, uint32_t nrows, AggPlan& plan, LcParams& P, const LcKeys& K, int threads, size_t smem,
           OutSink& sink, uint32_t* ngroups, bool* used);

// mid cardinality, bounded keys: range-partitioned direct tables (mc.cu).  *used = false: declined, or a per-row
// check failed (*abort_code = the FLAG_ABORT value; 3 = a key outside the sampled bounds)
int try_mc(const KeyCols& kc, uint32_t nrows, AggPlan& plan, const PlanStats& st, OutSink& sink, uint32_t* ngroups, bool* used,
           uint32_t* abort_code);

// groups ~ rows: fingerprint-partitioned path (hcfp.cu)
bool hcfp_qualifies(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, const PlanStats& st);
int run_hcfp(const KeyCols& kc, uint32_t nrows, AggPlan& plan, OutSink& sink, uint32_t* ngroups, bool* used);
// row shards: only (global row, fingerprint) tuples travel; every rank emits the groups whose first row it holds
bool hcfp_sharded_qualifies(const KeyCols& kc, const AggPlan& plan, const PlanStats& st);
int run_hcfp_sharded(const KeyCols& kc, uint32_t nrows, AggPlan& plan, OutSink& sink, uint32_t* ngroups, bool* used);

} // namespace aq
