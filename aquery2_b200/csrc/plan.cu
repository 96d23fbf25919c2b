// plan.cu -- sampling planner in front of the group-by paths.
//
// The reference's emitter never tells the library how many groups to expect nor what the key ranges are
// (ColRef::populate_stats exists, server/table.h:76-91, but nothing calls it), so the
// library finds out by itself, inside the call: ONE small kernel reads a strided sample of the rows (256
// chunks of 4096 consecutive rows, ~1M rows = a few tens of microseconds at HBM speed) and returns
//   * min / max of every 32-bit key column                          -> direct-table extents, bit packing
//   * the number of distinct key tuples in the sample (64-bit fingerprints in an L2-resident table)
//                                                                   -> estimated group count (occupancy inversion)
//   * per accumulator: largest |x| (integers) / largest exponent (floats) after the accumulator's transform
//                                                                   -> cell kinds, fixed-point base (cells.cuh)
// and the planner picks: lc.cu (few groups), hcfp.cu (groups ~ rows) or the hash / partitioned paths of
// groupby.cu with the estimate as the table-sizing hint.  Every number taken from the sample is a SPECULATION:
// the scan kernels check each row against it and raise FLAG_ABORT, upon which the query re-runs on a path that
// assumes nothing.  Bounds or hints passed by the caller replace the sampled ones.
#include "lc.cuh"
#include "comm.cuh"
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace aq {

constexpr uint32_t SAMPLE_CHUNKS = 256, SAMPLE_CHUNK_ROWS = 4096;
constexpr uint32_t SAMPLE_TABLE_BITS = 21;
constexpr int PLAN_WORDS = 2 * MAX_KEY_WORDS + 2 * MAX_ACC + 4;     // u64 result words

struct SampleAcc { int kind; int xform; const uint32_t* v; const uint32_t* v2; };   // kind: 0 none, 1 int32, 2 float
struct SampleSpec {
    int nwords;
    const uint32_t* kbase[MAX_KEY_WORDS];
    uint32_t kstride[MAX_KEY_WORDS];
    uint32_t kflip[MAX_KEY_WORDS];        // 0x80000000 for signed columns: order-preserving u32 image
    int nacc;
    SampleAcc acc[MAX_ACC];
    uint32_t nrows, nsample, chunk_rows;
    unsigned long long chunk_stride;
    unsigned long long* table;
    unsigned long long* out;              // [2w] min image, [2w+1] max image, [2*MAXW + 2a] maxabs / maxexp, [.. + 1] bad flag, last: distinct
};

__device__ __forceinline__ uint64_t hash_words(const uint32_t* kw, int n) {
    uint64_t h = 0x9e3779b97f4a7c15ULL;
    for (int w = 0; w < n; ++w) { h = (h ^ kw[w]) * 0xff51afd7ed558ccdULL; h ^= h >> 32; }
    h *= 0xc4ceb9fe1a85ec53ULL;
    return h ^ (h >> 29);
}

__global__ void __launch_bounds__(256)
k_sample(const SampleSpec S) {
    uint32_t kmin[MAX_KEY_WORDS], kmax[MAX_KEY_WORDS];
    unsigned long long amax[MAX_ACC];
    uint32_t abad = 0, distinct = 0;
#pragma unroll
    for (int w = 0; w < MAX_KEY_WORDS; ++w) { kmin[w] = 0xffffffffu; kmax[w] = 0u; }
#pragma unroll
    for (int a = 0; a < MAX_ACC; ++a) amax[a] = 0ull;
    const uint32_t mask = (1u << SAMPLE_TABLE_BITS) - 1u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < S.nsample; i += gridDim.x * blockDim.x) {
        const unsigned long long row = (unsigned long long)(i / S.chunk_rows) * S.chunk_stride + (i % S.chunk_rows);
        uint32_t kw[MAX_KEY_WORDS];
#pragma unroll
        for (int w = 0; w < MAX_KEY_WORDS; ++w) {
            kw[w] = 0;
            if (w < S.nwords) {
                kw[w] = __ldg(S.kbase[w] + row * S.kstride[w]);
                const uint32_t img = kw[w] ^ S.kflip[w];
                kmin[w] = min(kmin[w], img); kmax[w] = max(kmax[w], img);
            }
        }
        uint64_t h = hash_words(kw, S.nwords);
        if (h == 0) h = 1;
        uint32_t s = (uint32_t)(h >> 24) & mask;
        for (int step = 0; step < 128; ++step) {
            const unsigned long long old = atomicCAS(S.table + s, 0ull, (unsigned long long)h);
            if (old == 0ull) { ++distinct; break; }
            if (old == h) break;
            s = (s + 1u) & mask;
        }
#pragma unroll
        for (int a = 0; a < MAX_ACC; ++a) {
            if (a >= S.nacc || S.acc[a].kind == 0) continue;
            const uint32_t u = __ldg(S.acc[a].v + row);
            const uint32_t u2 = S.acc[a].xform == X_PRODUCT ? __ldg(S.acc[a].v2 + row) : u;
            if (S.acc[a].kind == 1) {
                long long x = (long long)(int)u;
                if (S.acc[a].xform != X_IDENT) x *= (long long)(int)u2;
                const unsigned long long m = x < 0 ? 0ull - (unsigned long long)x : (unsigned long long)x;
                amax[a] = max(amax[a], m);
            } else {
                const float f = __uint_as_float(u);
                const float x = S.acc[a].xform == X_IDENT ? f : __fmul_rn(f, __uint_as_float(u2));
                const uint32_t e = (__float_as_uint(x) >> 23) & 0xffu;
                if (e == 255u) abad |= 1u << a;
                else if ((__float_as_uint(x) & 0x7fffffffu) != 0u) amax[a] = max(amax[a], (unsigned long long)(e ? e : 1u));
            }
        }
    }
    // warp reductions, one atomic per warp and quantity
#pragma unroll
    for (int w = 0; w < MAX_KEY_WORDS; ++w) {
        if (w >= S.nwords) continue;
        const uint32_t lo = __reduce_min_sync(0xffffffffu, kmin[w]), hi = __reduce_max_sync(0xffffffffu, kmax[w]);
        if ((threadIdx.x & 31) == 0) { atomicMin(S.out + 2 * w, (unsigned long long)lo); atomicMax(S.out + 2 * w + 1, (unsigned long long)hi); }
    }
#pragma unroll
    for (int a = 0; a < MAX_ACC; ++a) {
        if (a >= S.nacc || S.acc[a].kind == 0) continue;
        unsigned long long m = amax[a];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_down_sync(0xffffffffu, m, d); m = max(m, o); }
        if ((threadIdx.x & 31) == 0) atomicMax(S.out + 2 * MAX_KEY_WORDS + 2 * a, m);
    }
    const uint32_t bad = __reduce_or_sync(0xffffffffu, abad);
    const uint32_t dsum = __reduce_add_sync(0xffffffffu, distinct);
    if ((threadIdx.x & 31) == 0) {
        for (int a = 0; a < S.nacc; ++a) if (bad & (1u << a)) S.out[2 * MAX_KEY_WORDS + 2 * a + 1] = 1ull;
        if (dsum) atomicAdd(S.out + PLAN_WORDS - 1, (unsigned long long)dsum);
    }
}

// number of groups G whose expected distinct count in a sample of S rows is d:  d = G (1 - exp(-S / G))
static double invert_occupancy(double S, double d, double cap) {
    if (d >= S) return cap;
    double lo = d, hi = cap > d ? cap : d;
    if (hi * (1.0 - std::exp(-S / hi)) <= d) return hi;
    for (int it = 0; it < 200; ++it) {
        const double mid = std::sqrt(lo * hi);
        if (mid * (1.0 - std::exp(-S / mid)) < d) lo = mid; else hi = mid;
        if (hi / lo < 1.0005) break;
    }
    return std::sqrt(lo * hi);
}

bool path_disabled(const char* name) {
    const char* e = getenv("AQGPU_DISABLE");
    if (!e) return false;
    const size_t n = strlen(name);
    for (const char* p = e; *p;) {
        const char* q = strchr(p, ',');
        const size_t len = q ? size_t(q - p) : strlen(p);
        if (len == n && strncmp(p, name, n) == 0) return true;
        if (!q) break;
        p = q + 1;
    }
    return false;
}

static int sample_inputs(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, PlanStats& st) {
    Context& c = ctx();
    SampleSpec S;
    memset(&S, 0, sizeof(S));
    int w = 0;
    bool all32 = true;
    for (int k = 0; k < kc.n; ++k) {
        const uint32_t* p = static_cast<const uint32_t*>(kc.col[k].d);
        if (kc.type[k] == AQ_INT64) {
            if (w + 2 > MAX_KEY_WORDS) return fail("key tuple wider than %d words", MAX_KEY_WORDS);
            S.kbase[w] = p; S.kstride[w] = 2; S.kflip[w] = 0; ++w;
            S.kbase[w] = p + 1; S.kstride[w] = 2; S.kflip[w] = 0; ++w;
            all32 = false;
        } else {
            if (w + 1 > MAX_KEY_WORDS) return fail("key tuple wider than %d words", MAX_KEY_WORDS);
            S.kbase[w] = p; S.kstride[w] = 1; S.kflip[w] = kc.type[k] == AQ_INT32 ? 0x80000000u : 0u; ++w;
        }
    }
    S.nwords = w;
    S.nacc = plan.n_acc;
    for (int a = 0; a < plan.n_acc; ++a) {
        const AccOp& o = plan.prog.op[a];
        S.acc[a].kind = 0;
        if ((o.kind == ACC_ADD_I64 && o.vtype == AQ_INT32) || (o.kind == ACC_ADD_F64 && o.vtype == AQ_FLOAT)) {
            S.acc[a].kind = o.vtype == AQ_INT32 ? 1 : 2;
            S.acc[a].xform = o.xform;
            S.acc[a].v = static_cast<const uint32_t*>(o.v);
            S.acc[a].v2 = static_cast<const uint32_t*>(o.v2);
        }
    }
    S.nrows = nrows;
    if (nrows <= SAMPLE_CHUNKS * SAMPLE_CHUNK_ROWS) { S.nsample = nrows; S.chunk_rows = nrows; S.chunk_stride = 0; }
    else { S.nsample = SAMPLE_CHUNKS * SAMPLE_CHUNK_ROWS; S.chunk_rows = SAMPLE_CHUNK_ROWS; S.chunk_stride = nrows / SAMPLE_CHUNKS; }
    DevBuf table, out;
    AQ_TRY(table.alloc_zero(size_t(8) << SAMPLE_TABLE_BITS));
    AQ_TRY(out.alloc(PLAN_WORDS * 8));
    unsigned long long init[PLAN_WORDS];
    memset(init, 0, sizeof(init));
    for (int i = 0; i < MAX_KEY_WORDS; ++i) init[2 * i] = ~0ull;
    AQ_CUDA(cudaMemcpyAsync(out.p, init, sizeof(init), cudaMemcpyHostToDevice, c.stream));
    S.table = table.as<unsigned long long>();
    S.out = out.as<unsigned long long>();
    AQ_LAUNCH(k_sample, grid_for(S.nsample, 256, 4, 4), 256, 0, S);
    unsigned long long res[PLAN_WORDS];
    AQ_CUDA(cudaMemcpyAsync(res, out.p, sizeof(res), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    st.sample_rows = S.nsample;
    st.sample_distinct = (uint32_t)res[PLAN_WORDS - 1];
    if (!st.bounds_given) {
        st.have_bounds = all32;
        for (int k = 0; k < kc.n && all32; ++k) {
            const uint32_t lo = (uint32_t)res[2 * k] ^ S.kflip[k], hi = (uint32_t)res[2 * k + 1] ^ S.kflip[k];
            st.kmin[k] = kc.type[k] == AQ_INT32 ? (long long)(int)lo : (long long)lo;
            st.kmax[k] = kc.type[k] == AQ_INT32 ? (long long)(int)hi : (long long)hi;
        }
    }
    for (int a = 0; a < plan.n_acc; ++a) {
        st.fx_maxexp[a] = -1; st.isum_maxabs[a] = ~0ull;
        if (S.acc[a].kind == 1) st.isum_maxabs[a] = res[2 * MAX_KEY_WORDS + 2 * a];
        if (S.acc[a].kind == 2 && res[2 * MAX_KEY_WORDS + 2 * a + 1] == 0) st.fx_maxexp[a] = (int)res[2 * MAX_KEY_WORDS + 2 * a];
    }
    if (st.est_groups <= 0.0) {
        const double Sn = (double)st.sample_rows, d = (double)st.sample_distinct;
        st.est_groups = S.nsample == nrows ? d : invert_occupancy(Sn, d, (double)nrows);
    }
    return 0;
}

// Row shards (dist.cu): every rank has sampled its own rows; one small all-reduce (max) makes the statistics -- and with
// them every path / layout decision -- identical on all ranks.
static int merge_stats_across_ranks(const KeyCols& kc, const AggPlan& plan, PlanStats& st, uint32_t nrows) {
    constexpr long long BIG = 0x7fffffffffffffffLL;
    long long v[4 + 2 * MAX_KEY_COLS + 3 * MAX_ACC];
    int n = 0;
    v[n++] = st.have_bounds ? 0 : 1;
    v[n++] = -(long long)nrows;
    v[n++] = (long long)nrows;
    v[n++] = (long long)std::ceil(st.est_groups);
    for (int k = 0; k < kc.n; ++k) { v[n++] = st.have_bounds ? -st.kmin[k] : 0; v[n++] = st.have_bounds ? st.kmax[k] : 0; }
    for (int a = 0; a < plan.n_acc; ++a) {
        v[n++] = st.fx_maxexp[a] < 0 ? 1 : 0;
        v[n++] = st.fx_maxexp[a];
        v[n++] = st.isum_maxabs[a] > (unsigned long long)BIG ? BIG : (long long)st.isum_maxabs[a];
    }
    AQ_TRY(comm_allreduce_max_i64(v, n));
    n = 0;
    st.have_bounds = v[n++] == 0;
    dist().rows_min = (uint32_t)(-v[n++]);
    dist().rows_max = (uint32_t)v[n++];
    st.est_groups = (double)v[n++];
    for (int k = 0; k < kc.n; ++k) { st.kmin[k] = -v[n++]; st.kmax[k] = v[n++]; }
    for (int a = 0; a < plan.n_acc; ++a) {
        const bool unusable = v[n++] != 0;
        st.fx_maxexp[a] = unusable ? -1 : (int)v[n];
        ++n;
        st.isum_maxabs[a] = v[n] == BIG ? ~0ull : (unsigned long long)v[n];
        ++n;
    }
    return 0;
}

// what the generic paths get: the sampled group estimate and key bounds (expanded: a sample rarely sees the extreme keys)
static void fill_hints(const PlanStats& st, const KeyCols& kc, uint32_t nrows, uint32_t groups_hint, PlanHints& out) {
    PlanHints* hints_out = &out;
    // ---- generic paths: hand over what the sample says (expanded: a sample rarely sees the extreme keys)
    {
        hints_out->groups_hint = groups_hint;
        if (!groups_hint && st.est_groups > 0.0) {
            const double g = st.est_groups * 1.125 + 64.0;
            hints_out->groups_hint = (uint32_t)(g < (double)nrows ? g : (double)nrows);
        }
        hints_out->have_bounds = false;
        if (st.have_bounds) {
            hints_out->have_bounds = true;
            for (int k = 0; k < kc.n; ++k) {
                long long lo = st.kmin[k], hi = st.kmax[k];
                if (!st.bounds_given) {
                    const long long margin = (hi - lo) / 64 + 16;
                    const long long tlo = kc.type[k] == AQ_INT32 ? -2147483648LL : 0LL, thi = kc.type[k] == AQ_INT32 ? 2147483647LL : 4294967295LL;
                    lo = lo - margin < tlo ? tlo : lo - margin;
                    hi = hi + margin > thi ? thi : hi + margin;
                }
                hints_out->kmin[k] = lo; hints_out->kmax[k] = hi;
            }
        }
    }
}

// Sampling only (aqgpu_groupby_build): group estimate + key bounds for hash_aggregate.
int plan_hints_only(KeyCols& kc, uint32_t nrows, const int64_t* key_min, const int64_t* key_max, uint32_t groups_hint,
                    PlanHints* out) {
    out->groups_hint = groups_hint; out->have_bounds = false;
    if (nrows < 4096 || path_disabled("plan")) return 0;
    PlanStats st;
    if (key_min && key_max) {
        st.have_bounds = st.bounds_given = true;
        for (int k = 0; k < kc.n; ++k) { st.kmin[k] = key_min[k]; st.kmax[k] = key_max[k]; }
    }
    if (groups_hint) st.est_groups = (double)groups_hint;
    AggPlan none;
    AQ_TRY(sample_inputs(kc, nrows, none, st));
    fill_hints(st, kc, nrows, groups_hint, *out);
    return 0;
}

int run_planned_paths(KeyCols& kc, uint32_t nrows, const int64_t* key_min, const int64_t* key_max,
                      uint32_t groups_hint, AggPlan& plan, OutSink& sink, uint32_t* ngroups, bool* used,
                      PlanHints* hints_out) {
    *used = false;
    const bool sharded = dist().active;
    if (!sharded && (nrows < 4096 || path_disabled("plan"))) return 0;
    // value columns the sampler / the fast paths read as 4-byte cells must be 4-byte types; others keep their kind 0
    PlanStats st;
    if (key_min && key_max) {
        st.have_bounds = st.bounds_given = true;
        for (int k = 0; k < kc.n; ++k) { st.kmin[k] = key_min[k]; st.kmax[k] = key_max[k]; }
    }
    if (groups_hint) st.est_groups = (double)groups_hint;
    const bool had_count = plan.find(ACC_COUNT) >= 0;
    if (!had_count && plan.n_acc < MAX_ACC) plan.add_count();      // the direct-table path tells empty slots by their count
    AQ_TRY(sample_inputs(kc, nrows, plan, st));
    if (sharded) AQ_TRY(merge_stats_across_ranks(kc, plan, st, nrows));

    if (!path_disabled("lc")) {
        LcParams P; LcKeys K; int threads = 0; size_t smem = 0;
        if (lc_layout(kc, nrows, plan, st, had_count ? -1 : plan.find(ACC_COUNT), P, K, &threads, &smem)) {
            AQ_TRY(run_lc(kc, nrows, plan, P, K, threads, smem, sink, ngroups, used));
            if (*used) return 0;
            st.have_bounds = st.bounds_given;                        // a sampled bound was wrong: drop them
        }
    }
    if (!path_disabled("mc")) {
        uint32_t code = 0;
        AQ_TRY(try_mc(kc, nrows, plan, st, sink, ngroups, used, &code));
        if (*used) return 0;
        if (code == 3) st.have_bounds = st.bounds_given;            // a sampled bound was wrong: drop them
    }
    if (sharded) {                                              // row shards: record tables above, fingerprints here, nothing else
        if (!path_disabled("hcfp") && hcfp_sharded_qualifies(kc, plan, st)) AQ_TRY(run_hcfp_sharded(kc, nrows, plan, sink, ngroups, used));
        return 0;
    }
    if (!path_disabled("hcfp") && hcfp_qualifies(kc, nrows, plan, st)) {
        AQ_TRY(run_hcfp(kc, nrows, plan, sink, ngroups, used));
        if (*used) return 0;
    }
    if (hints_out) fill_hints(st, kc, nrows, groups_hint, *hints_out);
    return 0;
}

} // namespace aq
