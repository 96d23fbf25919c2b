// groups.cu -- the group handle: HashTableComponents {size, keys, values, offsets} of the reference
// (server/hasher.h:201-207) kept in HBM, plus the per-group reduce / scan / order_by entry points.
#include "groupby.cuh"
#include "sort.cuh"
#include "scan.cuh"
#include "window.cuh"
#include <cstring>
#include <mutex>

using namespace aq;

struct aqgpu_groups {
    uint32_t nrows = 0, ngroups = 0;
    int row_order = AQ_ROWS_DESC;
    int nkeys = 0;
    int key_type[MAX_KEY_COLS];
    DevBuf keys[MAX_KEY_COLS];      // ngroups cells each, first-appearance order
    DevBuf first_row, last_row;     // ngroups
    DevBuf counts;                  // ngroups + 1 (last = 0)
    DevBuf gid;                     // nrows: "reversemap"
    DevBuf offsets;                 // ngroups + 1: "ht_base" after ht_postproc + terminal
    DevBuf perm;                    // nrows: "mapbase"
    DevBuf segstart;                // nrows: first CSR position of the owning group
    bool has_offsets = false, has_perm = false, has_segstart = false;
    bool custom_order = false;      // order_by has replaced the row lists
    aqgpu_groups() {                // the handle outlives the call that builds it: pool memory, not the scratch arena
        for (auto& b : keys) b.persistent = true;
        for (DevBuf* b : {&first_row, &last_row, &counts, &gid, &offsets, &perm, &segstart}) b->persistent = true;
    }
};

namespace aq {

__global__ void k_map_gid(const uint32_t* __restrict__ row_slot, const uint32_t* __restrict__ slot_gid,
                          uint32_t n, uint32_t* __restrict__ gid) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        gid[i] = slot_gid[row_slot[i]];
}

// ---- direct build: ONE 32-bit key column whose (sampled, verified) value range is small enough for a table ----
// The reference's PerfectHashTable idea (server/hasher.h:226-243) with the contract's order: slot = key - lo,
// one pass counts the rows of every slot and keeps its smallest row (a plain read after the first few hundred
// thousand rows: the table is L2-resident), the non-empty slots are ranked by that first row (= insertion index
// of m_values, server/unordered_dense.h:1142-1146), and a second pass maps every row to its group id.
// A key outside the sampled range raises FLAG_ABORT and the build falls back to the hash path.
constexpr uint32_t GD_MAX_RANGE = 1u << 24;
__global__ void __launch_bounds__(256)
k_gd_count(const uint32_t* __restrict__ key, const uint32_t n, const uint32_t lo, const uint32_t range, uint32_t* __restrict__ cnt,
           uint32_t* __restrict__ first, uint32_t* __restrict__ flags) {
    const uint32_t nq = n >> 2;
    bool bad = false;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += gridDim.x * blockDim.x) {
        const uint4 k4 = __ldg(reinterpret_cast<const uint4*>(key) + q);
        const uint32_t kv[4] = {k4.x, k4.y, k4.z, k4.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t d = kv[j] - lo, row = q * 4u + j;
            if (d >= range) { bad = true; continue; }
            atomicAdd(cnt + d, 1u);
            if (row < *reinterpret_cast<volatile uint32_t*>(first + d)) atomicMin(first + d, row);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3u)) {
        const uint32_t row = (n & ~3u) + threadIdx.x, d = key[row] - lo;
        if (d >= range) bad = true;
        else { atomicAdd(cnt + d, 1u); atomicMin(first + d, row); }
    }
    if (bad) atomicExch(flags + FLAG_ABORT, 3u);
}
__global__ void k_gd_ngroups(const uint32_t* __restrict__ cnt, const uint32_t range, uint32_t* __restrict__ flags) {
    uint32_t c = 0;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < range; s += gridDim.x * blockDim.x) c += cnt[s] != 0u;
#pragma unroll
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(flags + FLAG_NGROUPS, c);
}
__global__ void k_gd_emit(const uint32_t* __restrict__ order, const uint32_t ng, const uint32_t lo, const uint32_t* __restrict__ first,
                          const uint32_t* __restrict__ cnt, uint32_t* __restrict__ keys, uint32_t* __restrict__ first_row,
                          uint32_t* __restrict__ counts, uint32_t* __restrict__ gid_of_slot) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ng; i += gridDim.x * blockDim.x) {
        const uint32_t s = order[i];
        keys[i] = lo + s; first_row[i] = first[s]; counts[i] = cnt[s]; gid_of_slot[s] = i;
    }
}
__global__ void __launch_bounds__(256)
k_gd_gid(const uint32_t* __restrict__ key, const uint32_t n, const uint32_t lo, const uint32_t* __restrict__ gid_of_slot, uint32_t* __restrict__ gid) {
    const uint32_t nq = n >> 2;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += gridDim.x * blockDim.x) {
        const uint4 k4 = __ldg(reinterpret_cast<const uint4*>(key) + q);
        uint4 o;
        o.x = __ldg(gid_of_slot + (k4.x - lo)); o.y = __ldg(gid_of_slot + (k4.y - lo));
        o.z = __ldg(gid_of_slot + (k4.z - lo)); o.w = __ldg(gid_of_slot + (k4.w - lo));
        reinterpret_cast<uint4*>(gid)[q] = o;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3u)) { const uint32_t row = (n & ~3u) + threadIdx.x; gid[row] = gid_of_slot[key[row] - lo]; }
}

static int ensure_offsets(aqgpu_groups* g) {
    if (g->has_offsets) return 0;
    AQ_TRY(g->offsets.alloc(size_t(g->ngroups + 1) * 4));
    AQ_TRY(exclusive_scan_u32(g->counts.as<uint32_t>(), g->offsets.as<uint32_t>(), size_t(g->ngroups) + 1, nullptr));
    g->has_offsets = true;
    return 0;
}

// ht_postproc (hasher.h:181-198): rows grouped by gid; descending row ids inside a group in the
// reference order, ascending in AQ_ROWS_ASC.  Stable LSD radix sort of the (reversed) row list.
static int ensure_perm(aqgpu_groups* g) {
    if (g->has_perm) return 0;
    AQ_TRY(ensure_offsets(g));
    AQ_TRY(g->perm.alloc(size_t(g->nrows) * 4));
    SortJob job;
    job.n = g->nrows;
    job.rows_in = nullptr;
    job.reverse = g->row_order == AQ_ROWS_DESC;
    job.col = g->gid.as<uint32_t>();
    int bits = 0;
    while (bits < 32 && (uint64_t(1) << bits) < uint64_t(g->ngroups)) ++bits;
    job.key_bits = bits;
    job.rows_out = g->perm.as<uint32_t>();
    AQ_TRY(sort_rows_by_key(job));
    g->has_perm = true;
    return 0;
}

static int ensure_segstart(aqgpu_groups* g) {
    if (g->has_segstart) return 0;
    AQ_TRY(ensure_offsets(g));
    AQ_TRY(g->segstart.alloc(size_t(g->nrows) * 4));
    AQ_TRY(compute_segment_starts(g->offsets.as<uint32_t>(), g->ngroups, g->nrows, g->segstart.as<uint32_t>()));
    g->has_segstart = true;
    return 0;
}

static int copy_out(void* dst, const void* d_src, size_t bytes) {
    if (!dst) return fail("null output pointer");
    if (bytes == 0) return 0;
    const bool dev = is_device_ptr(dst);
    AQ_CUDA(cudaMemcpyAsync(dst, d_src, bytes, dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx().stream));
    if (!dev) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

// ---- min / max of an integer column --------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(256)
k_minmax(const T* __restrict__ v, uint32_t n, long long* __restrict__ out /* [min, max] */) {
    long long lo = 0x7fffffffffffffffLL, hi = (long long)0x8000000000000000ULL;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const long long x = (long long)v[i];
        lo = x < lo ? x : lo; hi = x > hi ? x : hi;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const long long l2 = __shfl_down_sync(0xffffffffu, lo, d), h2 = __shfl_down_sync(0xffffffffu, hi, d);
        lo = l2 < lo ? l2 : lo; hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(out, lo); atomicMax(out + 1, hi); }
}

// ---- element-wise helpers -------------------------------------------------------------------------
template <class T>
__global__ void k_vec_sub(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ o, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) o[i] = a[i] - b[i];
}
__global__ void k_avg_finalize(int sum_type, const void* __restrict__ sum, const uint32_t* __restrict__ cnt,
                               double* __restrict__ o, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double s;
        if (sum_type == AQ_DOUBLE) s = static_cast<const double*>(sum)[i];
        else if (sum_type == AQ_INT64) s = (double)static_cast<const long long*>(sum)[i];
        else s = (double)(static_cast<const __int128*>(sum)[i]);
        o[i] = s / (double)cnt[i];
    }
}
__global__ void k_vec_pow(const double* __restrict__ x, double p, double* __restrict__ o, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) o[i] = pow(x[i], p);
}

} // namespace aq

// ---- several expressions at once -------------------------------------------------------------------------------
namespace aq {
template <class T>
__global__ void k_sub_from(const T* __restrict__ gv, T* __restrict__ io, uint32_t n) {      // io = gv - io
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) io[i] = (T)(gv[i] - io[i]);
}
// one thread per group over its CSR range; identities of server/aggregations.h:71-86
template <class T>
__global__ void k_seg_extreme(const T* __restrict__ vec, const uint32_t* __restrict__ offsets, uint32_t ngroups, int is_max, T ident,
                              T* __restrict__ out) {
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
        T m = ident;
        for (uint32_t p = offsets[g]; p < offsets[g + 1]; ++p) { const T e = vec[p]; if (is_max ? (e > m) : (e < m)) m = e; }
        out[g] = m;
    }
}
template <class T>
static int window_post(int minus, int reduce, const void* gv, void* vec, const uint32_t* offsets, uint32_t n, uint32_t ngroups,
                       T max_ident, T min_ident, void* out) {
    if (minus) AQ_LAUNCH(k_sub_from<T>, grid_for(n, 256, 1, 32), 256, 0, static_cast<const T*>(gv), static_cast<T*>(vec), n);
    if (reduce >= 0)
        AQ_LAUNCH(k_seg_extreme<T>, grid_for(ngroups, 128, 1, 32), 128, 0, static_cast<const T*>(vec), offsets, ngroups, reduce == AQ_MAX ? 1 : 0,
                  reduce == AQ_MAX ? max_ident : min_ident, static_cast<T*>(out));
    return 0;
}
} // namespace aq


static int build_direct(aqgpu_groups* g, const KeyCols& kc, const PlanHints& ph, bool* used) {
    *used = false;
    Context& c = ctx();
    const uint32_t n = g->nrows;
    if (kc.n != 1 || (kc.type[0] != AQ_INT32 && kc.type[0] != AQ_UINT32) || !ph.have_bounds || n < (1u << 20)) return 0;
    if (path_disabled("gd") || (reinterpret_cast<uintptr_t>(kc.col[0].d) & 15)) return 0;
    const long long span = ph.kmax[0] - ph.kmin[0] + 1;
    if (span < 1 || span > (long long)GD_MAX_RANGE || span > 4ll * n) return 0;
    const uint32_t range = (uint32_t)span, lo = (uint32_t)ph.kmin[0];
    const uint32_t* key = static_cast<const uint32_t*>(kc.col[0].d);
    DevBuf cnt, first, order, gos;
    AQ_TRY(cnt.alloc_zero(size_t(range) * 4));
    AQ_TRY(first.alloc(size_t(range) * 4));
    AQ_TRY(order.alloc(size_t(range) * 4));
    AQ_TRY(gos.alloc(size_t(range) * 4));
    AQ_CUDA(cudaMemsetAsync(first.p, 0xff, size_t(range) * 4, c.stream));
    AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 8 * sizeof(uint32_t), c.stream));
    AQ_LAUNCH(k_gd_count, c.sm_count * 8, 256, 0, key, n, lo, range, cnt.as<uint32_t>(), first.as<uint32_t>(), c.d_flags);
    AQ_LAUNCH(k_gd_ngroups, grid_for(range, 256, 1, 4), 256, 0, cnt.as<uint32_t>(), range, c.d_flags);
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) return 0;                      // a key outside the sampled range: hash path
    const uint32_t ng = c.h_flags[FLAG_NGROUPS];
    SortJob job;                                                  // slots by first row; empty slots (0xffffffff) go last
    job.n = range; job.rows_in = nullptr; job.col = first.as<uint32_t>(); job.key_bits = 32; job.rows_out = order.as<uint32_t>();
    AQ_TRY(sort_rows_by_key(job));
    g->ngroups = ng;
    AQ_TRY(g->keys[0].alloc(size_t(ng) * 4));
    AQ_TRY(g->first_row.alloc(size_t(ng) * 4));
    AQ_TRY(g->counts.alloc_zero(size_t(ng + 1) * 4));
    AQ_TRY(g->gid.alloc(size_t(n) * 4));
    if (ng) AQ_LAUNCH(k_gd_emit, grid_for(ng, 256, 1, 8), 256, 0, order.as<uint32_t>(), ng, lo, first.as<uint32_t>(), cnt.as<uint32_t>(),
                      g->keys[0].as<uint32_t>(), g->first_row.as<uint32_t>(), g->counts.as<uint32_t>(), gos.as<uint32_t>());
    AQ_LAUNCH(k_gd_gid, c.sm_count * 8, 256, 0, key, n, lo, gos.as<uint32_t>(), g->gid.as<uint32_t>());
    *used = true;
    return 0;
}

// perm: row lists ordered by (group, v); the middle element(s) of every list
template <class T>
__global__ void k_median_pick(const T* __restrict__ v, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ offsets,
                              uint32_t ngroups, double* __restrict__ out) {
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += gridDim.x * blockDim.x) {
        const uint32_t lo = offsets[g], n = offsets[g + 1] - lo;
        if (n == 0) { out[g] = 0.0; continue; }
        const double a = (double)v[perm[lo + (n - 1) / 2]], b = (double)v[perm[lo + n / 2]];
        out[g] = (n & 1u) ? a : 0.5 * (a + b);
    }
}

extern "C" {

int aqgpu_groupby_build(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                        const int64_t* key_min, const int64_t* key_max,
                        int row_order, uint32_t groups_hint, aqgpu_groups_t** out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!out) return fail("null handle pointer");
    *out = nullptr;
    if (row_order != AQ_ROWS_DESC && row_order != AQ_ROWS_ASC) return fail("bad row_order %d", row_order);
    if (!keys || !key_types || nkeys < 1 || nkeys > MAX_KEY_COLS) return fail("need 1..%d key columns", MAX_KEY_COLS);
    aqgpu_groups* g = new aqgpu_groups();
    g->nrows = nrows; g->row_order = row_order; g->nkeys = nkeys;
    for (int k = 0; k < nkeys; ++k) g->key_type[k] = key_types[k];
    struct Guard { aqgpu_groups* g; ~Guard() { delete g; } } guard{g};
    if (nrows == 0) {
        AQ_TRY(g->counts.alloc_zero(4));
        guard.g = nullptr; *out = g;
        return 0;
    }
    KeyCols kc;
    AQ_TRY(kc.bind(keys, key_types, nkeys, nrows));
    PlanHints ph;                         // sampled group estimate / key bounds when the caller gave none (plan.cu)
    AQ_TRY(plan_hints_only(kc, nrows, key_min, key_max, groups_hint, &ph));
    KeySpec ks;
    if (!(key_min && key_max) && ph.have_bounds) AQ_TRY(kc.make_spec(ph.kmin, ph.kmax, ks));
    else AQ_TRY(kc.make_spec(key_min, key_max, ks));
    groups_hint = ph.groups_hint;
    {
        bool direct = false;
        AQ_TRY(build_direct(g, kc, ph, &direct));
        if (direct) {
            AQ_CUDA(cudaStreamSynchronize(ctx().stream));
            guard.g = nullptr; *out = g;
            return 0;
        }
    }
    AggPlan plan;
    plan.emit.row_order = row_order;
    if (plan.add_count() < 0 || plan.add_maxrow() < 0) return fail("internal: accumulator allocation");
    DevBuf row_slot;
    AQ_TRY(row_slot.alloc(size_t(nrows) * 4));
    HashResult hr;
    uint32_t ng = 0;
    AQ_TRY(hash_aggregate(ks, nrows, groups_hint, plan, row_slot.as<uint32_t>(), hr, &ng));
    g->ngroups = ng;

    KeyOut ko;
    memset(&ko, 0, sizeof(ko));
    ko.n = nkeys;
    for (int k = 0; k < nkeys; ++k) {
        ko.size[k] = (int)dtype_size(key_types[k]);
        AQ_TRY(g->keys[k].alloc(size_t(ng) * ko.size[k]));
        ko.src[k] = kc.col[k].d;
        ko.dst[k] = g->keys[k].p;
    }
    AQ_TRY(g->first_row.alloc(size_t(ng) * 4));
    AQ_TRY(g->last_row.alloc(size_t(ng) * 4));
    AQ_TRY(g->counts.alloc_zero(size_t(ng + 1) * 4));
    DevBuf slot_gid;
    AQ_TRY(slot_gid.alloc(size_t(hr.cap) * 4));
    AQ_TRY(emit_groups(hr.ks, nrows, hr, ko, g->first_row.as<uint32_t>(), g->counts.as<uint32_t>(), g->last_row.as<uint32_t>(),
                       slot_gid.as<uint32_t>(), plan, ng));
    AQ_TRY(g->gid.alloc(size_t(nrows) * 4));
    AQ_LAUNCH(k_map_gid, grid_for(nrows, 256, 1, 32), 256, 0, row_slot.as<uint32_t>(), slot_gid.as<uint32_t>(), nrows,
              g->gid.as<uint32_t>());
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));    // staged key columns (host inputs) are released on return
    guard.g = nullptr;
    *out = g;
    return 0;
}

int aqgpu_groups_free(aqgpu_groups_t* g) {
    delete g;
    return 0;
}

int aqgpu_groups_size(aqgpu_groups_t* g, uint32_t* ngroups) {
    if (!g || !ngroups) return fail("null argument");
    *ngroups = g->ngroups;
    return 0;
}

int aqgpu_groups_keys(aqgpu_groups_t* g, int k, void* out_keys) {
    if (!g || k < 0 || k >= g->nkeys) return fail("bad key column index %d", k);
    return copy_out(out_keys, g->keys[k].p, size_t(g->ngroups) * dtype_size(g->key_type[k]));
}

int aqgpu_groups_first_row(aqgpu_groups_t* g, uint32_t* out) {
    if (!g) return fail("null handle");
    return copy_out(out, g->first_row.p, size_t(g->ngroups) * 4);
}

int aqgpu_groups_gid(aqgpu_groups_t* g, uint32_t* out_gid) {
    if (!g) return fail("null handle");
    return copy_out(out_gid, g->gid.p, size_t(g->nrows) * 4);
}

int aqgpu_groups_csr(aqgpu_groups_t* g, uint32_t* out_offsets, uint32_t* out_perm) {
    if (!g) return fail("null handle");
    if (g->nrows == 0) {
        if (out_offsets) { uint32_t z = 0; if (is_device_ptr(out_offsets)) AQ_CUDA(cudaMemsetAsync(out_offsets, 0, 4, ctx().stream)); else *out_offsets = z; }
        return 0;
    }
    if (out_offsets) { AQ_TRY(ensure_offsets(g)); AQ_TRY(copy_out(out_offsets, g->offsets.p, size_t(g->ngroups + 1) * 4)); }
    if (out_perm) { AQ_TRY(ensure_perm(g)); AQ_TRY(copy_out(out_perm, g->perm.p, size_t(g->nrows) * 4)); }
    return 0;
}

int aqgpu_groups_device_views(aqgpu_groups_t* g, const uint32_t** d_gid, const uint32_t** d_offsets,
                              const uint32_t** d_perm) {
    if (!g) return fail("null handle");
    if (g->nrows) {
        if (d_offsets) AQ_TRY(ensure_offsets(g));
        if (d_perm) AQ_TRY(ensure_perm(g));
    }
    if (d_gid) *d_gid = g->gid.as<uint32_t>();
    if (d_offsets) *d_offsets = g->offsets.as<uint32_t>();
    if (d_perm) *d_perm = g->perm.as<uint32_t>();
    return 0;
}

// Scalar aggregate per group == group-by on the dense gid column: ids were handed out in
// first-appearance order, so the fused engine returns the cells already in gid order.
int aqgpu_group_reduce(aqgpu_groups_t* g, int op, int vtype, const void* v, const void* v2,
                       int out_type, void* out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!g) return fail("null handle");
    if (g->nrows == 0) return 0;
    const void* keys[1] = { g->gid.p };
    const int kt[1] = { AQ_UINT32 };
    aqgpu_agg_t a;
    a.op = op; a.vtype = vtype; a.v = v; a.v2 = v2; a.out_type = out_type; a.out = out;
    uint32_t ng = 0;
    AQ_TRY(aqgpu_groupby_agg(keys, kt, 1, g->nrows, nullptr, nullptr, g->row_order, g->ngroups, &a, 1, g->ngroups,
                             nullptr, nullptr, &ng));
    if (ng != g->ngroups) return fail("internal: reduce saw %u groups, handle has %u", ng, g->ngroups);
    return 0;
}

int aqgpu_group_median(aqgpu_groups_t* g, int vtype, const void* v, double* out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!g || !v || !out) return fail("null argument");
    if (vtype != AQ_INT32 && vtype != AQ_UINT32 && vtype != AQ_FLOAT) return fail("median: unsupported value dtype %d (int32, uint32, float)", vtype);
    if (g->ngroups == 0) return 0;
    const uint32_t n = g->nrows;
    AQ_TRY(ensure_offsets(g));
    InCol in;
    AQ_TRY(in.bind(v, size_t(4) * n));
    OutCol oc;
    AQ_TRY(oc.bind(out, size_t(8) * g->ngroups));
    DevBuf byval, bygroup;
    AQ_TRY(byval.alloc(size_t(n) * 4));
    AQ_TRY(bygroup.alloc(size_t(n) * 4));
    SortJob job;                                              // rows by value ...
    job.n = n; job.rows_in = nullptr; job.col = static_cast<const uint32_t*>(in.d); job.col_stride = 1; job.key_bits = 32;
    job.float_flip = vtype == AQ_FLOAT;
    job.xor_mask = vtype == AQ_INT32 ? 0x80000000u : 0u;
    job.rows_out = byval.as<uint32_t>();
    AQ_TRY(sort_rows_by_key(job));
    SortJob rg;                                               // ... then (stable) by group
    rg.n = n; rg.rows_in = byval.as<uint32_t>(); rg.col = g->gid.as<uint32_t>();
    int bits = 0;
    while (bits < 32 && (uint64_t(1) << bits) < uint64_t(g->ngroups)) ++bits;
    rg.key_bits = bits;
    rg.rows_out = bygroup.as<uint32_t>();
    const uint32_t* perm = bygroup.as<uint32_t>();
    if (bits == 0) perm = byval.as<uint32_t>(); else AQ_TRY(sort_rows_by_key(rg));
    if (vtype == AQ_FLOAT) AQ_LAUNCH(k_median_pick<float>, grid_for(g->ngroups, 256), 256, 0, static_cast<const float*>(in.d), perm, g->offsets.as<uint32_t>(), g->ngroups, static_cast<double*>(oc.d));
    else if (vtype == AQ_INT32) AQ_LAUNCH(k_median_pick<int>, grid_for(g->ngroups, 256), 256, 0, static_cast<const int*>(in.d), perm, g->offsets.as<uint32_t>(), g->ngroups, static_cast<double*>(oc.d));
    else AQ_LAUNCH(k_median_pick<uint32_t>, grid_for(g->ngroups, 256), 256, 0, static_cast<const uint32_t*>(in.d), perm, g->offsets.as<uint32_t>(), g->ngroups, static_cast<double*>(oc.d));
    AQ_TRY(oc.finish(size_t(8) * g->ngroups));
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_group_scan(aqgpu_groups_t* g, int op, uint32_t w, int vtype, const void* v, int out_type, void* out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!g) return fail("null handle");
    if (g->nrows == 0) return 0;
    if (!is_val_type(vtype)) return fail("unsupported value dtype %d", vtype);
    const int def = aqgpu_scan_out_type(op, vtype);
    if (def < 0) return fail("unknown scan op %d", op);
    if (out_type == 0 && def != 0) out_type = def;
    if (out_type != def && !(def == AQ_INT128 && out_type == AQ_INT64))
        return fail("scan op %d on dtype %d cannot produce dtype %d", op, vtype, out_type);
    AQ_TRY(ensure_perm(g));
    AQ_TRY(ensure_segstart(g));
    InCol in;
    AQ_TRY(in.bind(v, dtype_size(vtype) * g->nrows));
    OutCol oc;
    AQ_TRY(oc.bind(out, dtype_size(out_type) * g->nrows));
    AQ_TRY(group_scan_device(op, w, vtype, in.d, g->nrows, g->ngroups, g->gid.as<uint32_t>(), g->offsets.as<uint32_t>(),
                             g->perm.as<uint32_t>(), g->segstart.as<uint32_t>(), out_type, oc.d));
    AQ_TRY(oc.finish(dtype_size(out_type) * g->nrows));
    if (oc.host || in.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_group_gather(aqgpu_groups_t* g, int vtype, const void* v, void* out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!g) return fail("null handle");
    if (g->nrows == 0) return 0;
    const size_t esz = dtype_size(vtype);
    if (esz != 4 && esz != 8) return fail("gather: 4- or 8-byte cells only");
    AQ_TRY(ensure_perm(g));
    InCol in; OutCol oc;
    AQ_TRY(in.bind(v, esz * g->nrows));
    AQ_TRY(oc.bind(out, esz * g->nrows));
    AQ_TRY(gather_rows(in.d, (int)esz, g->perm.as<uint32_t>(), g->nrows, oc.d));
    AQ_TRY(oc.finish(esz * g->nrows));
    if (oc.host || in.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

// TableInfo::order_by per group (table.h:447-461, common/groupby.py:54-59) as one segmented sort:
// stable LSD passes over the order columns (last column first), then over gid.
int aqgpu_groups_order_by(aqgpu_groups_t* g, const void* const* cols, const int* col_types,
                          const int* desc, int ncols) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!g || !cols || !col_types || ncols < 1) return fail("bad order_by arguments");
    if (g->nrows == 0) return 0;
    const uint32_t n = g->nrows;
    // ties keep the current order of the row lists: the lists built so far, or -- when none has been materialised yet --
    // the implicit ascending / descending row order (no need to sort by group first: the regroup pass below is stable)
    const bool implicit = !g->has_perm;
    if (implicit) { AQ_TRY(ensure_offsets(g)); AQ_TRY(g->perm.alloc(size_t(n) * 4)); }
    DevBuf a, b;
    AQ_TRY(a.alloc(size_t(n) * 4));
    AQ_TRY(b.alloc(size_t(n) * 4));
    const uint32_t* cur = implicit ? nullptr : g->perm.as<uint32_t>();
    bool first_pass = true;
    uint32_t* bufs[2] = { a.as<uint32_t>(), b.as<uint32_t>() };
    int which = 0;
    std::vector<InCol> in(ncols);
    for (int c = ncols - 1; c >= 0; --c) {
        const int t = col_types[c];
        const bool d = desc && desc[c];
        AQ_TRY(in[c].bind(cols[c], dtype_size(t) * n));
        SortJob job;
        job.n = n; job.key_bits = 32;
        if (t == AQ_INT32 || t == AQ_UINT32 || t == AQ_FLOAT) {
            job.rows_in = cur; job.col = static_cast<const uint32_t*>(in[c].d); job.col_stride = 1;
            job.reverse = (first_pass && implicit && g->row_order == AQ_ROWS_DESC) ? 1 : 0; first_pass = false;
            job.float_flip = t == AQ_FLOAT;
            const uint32_t asc_mask = t == AQ_INT32 ? 0x80000000u : 0u;
            job.xor_mask = d ? ~asc_mask : asc_mask;
            job.rows_out = bufs[which];
            AQ_TRY(sort_rows_by_key(job));
            cur = bufs[which]; which ^= 1;
        } else if (t == AQ_INT64) {
            for (int half = 0; half < 2; ++half) {           // low word (unsigned) then high word (signed)
                job.rows_in = cur; job.col = static_cast<const uint32_t*>(in[c].d) + half; job.col_stride = 2;
                job.reverse = (first_pass && implicit && g->row_order == AQ_ROWS_DESC) ? 1 : 0; first_pass = false;
                const uint32_t asc_mask = half ? 0x80000000u : 0u;
                job.xor_mask = d ? ~asc_mask : asc_mask;
                job.rows_out = bufs[which];
                AQ_TRY(sort_rows_by_key(job));
                cur = bufs[which]; which ^= 1;
            }
        } else {
            return fail("order_by: unsupported column dtype %d (int32, uint32, float, int64)", t);
        }
    }
    SortJob job;                                              // regroup
    job.n = n; job.rows_in = cur; job.col = g->gid.as<uint32_t>();
    int bits = 0;
    while (bits < 32 && (uint64_t(1) << bits) < uint64_t(g->ngroups)) ++bits;
    job.key_bits = bits;
    job.rows_out = g->perm.as<uint32_t>();
    if (bits == 0) AQ_CUDA(cudaMemcpyAsync(g->perm.p, cur, size_t(n) * 4, cudaMemcpyDeviceToDevice, ctx().stream));
    else AQ_TRY(sort_rows_by_key(job));
    g->custom_order = true;
    g->has_perm = true;
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

static int group_window_impl(aqgpu_groups_t* g, const aqgpu_window_t* ops, int n_ops, const void* ocol, int ocol_type, int odesc);
int aqgpu_group_window(aqgpu_groups_t* g, const aqgpu_window_t* ops, int n_ops) { return group_window_impl(g, ops, n_ops, nullptr, 0, 0); }
int aqgpu_group_window_ordered(aqgpu_groups_t* g, const void* order_col, int order_type, int desc, const aqgpu_window_t* ops, int n_ops) {
    if (!order_col) return fail("window: null order column");
    if (order_type != AQ_INT32 && order_type != AQ_UINT32 && order_type != AQ_FLOAT) return fail("window: order column dtype %d (int32, uint32, float)", order_type);
    return group_window_impl(g, ops, n_ops, order_col, order_type, desc);
}
static int group_window_impl(aqgpu_groups_t* g, const aqgpu_window_t* ops, int n_ops, const void* ocol, int ocol_type, int odesc) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!g || !ops || n_ops < 1 || n_ops > 64) return fail("bad window arguments");
    if (g->nrows == 0) return 0;
    const uint32_t n = g->nrows, ng = g->ngroups;
    std::vector<InCol> in(n_ops);
    std::vector<OutCol> oc(n_ops);
    std::vector<int> otype(n_ops);
    std::vector<size_t> obytes(n_ops);
    bool any_host = false;
    for (int q = 0; q < n_ops; ++q) {
        const aqgpu_window_t& o = ops[q];
        if (!is_val_type(o.vtype)) return fail("window %d: unsupported value dtype %d", q, o.vtype);
        const int def = aqgpu_scan_out_type(o.op, o.vtype);
        if (def < 0) return fail("window %d: unknown scan op %d", q, o.op);
        int ot = o.out_type ? o.out_type : def;
        if (ot != def && !(def == AQ_INT128 && ot == AQ_INT64)) return fail("window %d: op %d on dtype %d cannot produce dtype %d", q, o.op, o.vtype, ot);
        if ((o.minus_from_v || o.reduce >= 0) && def != o.vtype)
            return fail("window %d: v - scan(v) and per-group max / min need a scan whose result has v's type", q);
        if (o.reduce >= 0 && o.reduce != AQ_MAX && o.reduce != AQ_MIN) return fail("window %d: reduce must be AQ_MAX or AQ_MIN", q);
        if ((o.op == AQ_SUMW || o.op == AQ_AVGW || o.op == AQ_MINW || o.op == AQ_MAXW || o.op == AQ_RATIOW) && o.w == 0)
            return fail("window length 0 (the reference reads ret[-1] for w == 0)");
        otype[q] = ot;
        obytes[q] = dtype_size(ot) * (o.reduce >= 0 ? size_t(ng) : size_t(n));
        AQ_TRY(in[q].bind(o.v, dtype_size(o.vtype) * n));
        AQ_TRY(oc[q].bind(o.out, obytes[q]));
        any_host |= oc[q].host != nullptr || in[q].tmp.p != nullptr;
    }
    AQ_TRY(ensure_offsets(g));
    InCol oin;
    if (ocol) AQ_TRY(oin.bind(ocol, size_t(4) * n));
    bool handled = false;
    for (int attempt = 0; attempt < 2 && !handled; ++attempt) {
        if (attempt == 1) {                // the fused path declined the in-call ordering (a very long group, three value columns,
            if (!ocol) break;              // a handle with its own order ...): sort the row lists globally, then carry on as usual
            const void* oc1[1] = { oin.d }; const int ot1[1] = { ocol_type }; const int od1[1] = { odesc };
            AQ_TRY(aqgpu_groups_order_by(g, oc1, ot1, od1, 1));
        }
        if (n_ops > WIN_MAX_OPS) continue;
        WinGroups G;
        G.nrows = n; G.ngroups = ng; G.row_order = g->row_order; G.gid = g->gid.as<uint32_t>(); G.offsets = g->offsets.as<uint32_t>();
        G.perm = g->custom_order ? g->perm.as<uint32_t>() : nullptr;
        if (ocol && attempt == 0) {
            if (g->custom_order) continue;
            G.order_key = static_cast<const uint32_t*>(oin.d);
            const uint32_t asc = ocol_type == AQ_INT32 ? 0x80000000u : 0u;
            G.order_xor = odesc ? ~asc : asc;
            G.order_float = ocol_type == AQ_FLOAT;
        }
        WinOp w[WIN_MAX_OPS];
        for (int q = 0; q < n_ops; ++q) {
            w[q].op = ops[q].op; w[q].w = ops[q].w; w[q].vtype = ops[q].vtype; w[q].v = in[q].d; w[q].minus = ops[q].minus_from_v;
            w[q].reduce = ops[q].reduce; w[q].out_type = otype[q]; w[q].out = oc[q].d;
        }
        AQ_TRY(window_fused(G, w, n_ops, &handled));
    }
    if (!handled) {                       // one expression at a time over the handle's row lists
        AQ_TRY(ensure_perm(g));
        AQ_TRY(ensure_segstart(g));
        for (int q = 0; q < n_ops; ++q) {
            const aqgpu_window_t& o = ops[q];
            DevBuf vec, gv;
            void* dst = oc[q].d;
            if (o.reduce >= 0) { AQ_TRY(vec.alloc(dtype_size(otype[q]) * n)); dst = vec.p; }
            AQ_TRY(group_scan_device(o.op, o.w, o.vtype, in[q].d, n, ng, g->gid.as<uint32_t>(), g->offsets.as<uint32_t>(),
                                     g->perm.as<uint32_t>(), g->segstart.as<uint32_t>(), otype[q], dst));
            if (!o.minus_from_v && o.reduce < 0) continue;
            if (o.minus_from_v) {
                AQ_TRY(gv.alloc(dtype_size(o.vtype) * n));
                AQ_TRY(gather_rows(in[q].d, (int)dtype_size(o.vtype), g->perm.as<uint32_t>(), n, gv.p));
            }
            const uint32_t* off = g->offsets.as<uint32_t>();
            switch (o.vtype) {
            case AQ_INT32: AQ_TRY(window_post<int>(o.minus_from_v, o.reduce, gv.p, dst, off, n, ng, (int)0x80000000u, 0x7fffffff, oc[q].d)); break;
            case AQ_INT64: AQ_TRY(window_post<long long>(o.minus_from_v, o.reduce, gv.p, dst, off, n, ng, (long long)0x8000000000000000ULL, 0x7fffffffffffffffLL, oc[q].d)); break;
            case AQ_FLOAT: AQ_TRY(window_post<float>(o.minus_from_v, o.reduce, gv.p, dst, off, n, ng, 1.17549435e-38f, 3.402823466e+38f, oc[q].d)); break;
            default: AQ_TRY(window_post<double>(o.minus_from_v, o.reduce, gv.p, dst, off, n, ng, 2.2250738585072014e-308, 1.7976931348623157e+308, oc[q].d)); break;
            }
        }
    }
    for (int q = 0; q < n_ops; ++q) AQ_TRY(oc[q].finish(obytes[q]));
    if (any_host) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_col_minmax(int dtype, const void* col, uint32_t nrows, int64_t* out_min, int64_t* out_max) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!out_min || !out_max) return fail("null output");
    if (nrows == 0) { *out_min = 0; *out_max = -1; return 0; }
    const size_t sz = dtype_size(dtype);
    if (dtype != AQ_INT32 && dtype != AQ_UINT32 && dtype != AQ_INT64) return fail("minmax: integer columns only");
    InCol in;
    AQ_TRY(in.bind(col, sz * nrows));
    DevBuf mm;
    AQ_TRY(mm.alloc(16));
    const long long init[2] = { 0x7fffffffffffffffLL, (long long)0x8000000000000000ULL };
    AQ_CUDA(cudaMemcpyAsync(mm.p, init, 16, cudaMemcpyHostToDevice, ctx().stream));
    const int grid = grid_for(nrows, 256, 8, 8);
    if (dtype == AQ_INT32) AQ_LAUNCH(k_minmax<int>, grid, 256, 0, static_cast<const int*>(in.d), nrows, mm.as<long long>());
    else if (dtype == AQ_UINT32) AQ_LAUNCH(k_minmax<uint32_t>, grid, 256, 0, static_cast<const uint32_t*>(in.d), nrows, mm.as<long long>());
    else AQ_LAUNCH(k_minmax<long long>, grid, 256, 0, static_cast<const long long*>(in.d), nrows, mm.as<long long>());
    long long res[2];
    AQ_CUDA(cudaMemcpyAsync(res, mm.p, 16, cudaMemcpyDeviceToHost, ctx().stream));
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    *out_min = res[0]; *out_max = res[1];
    return 0;
}

int aqgpu_vec_sub(int dtype, const void* a, const void* b, void* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    const size_t sz = dtype_size(dtype);
    if (!is_val_type(dtype)) return fail("vec_sub: unsupported dtype %d", dtype);
    InCol ia, ib; OutCol oc;
    AQ_TRY(ia.bind(a, sz * n)); AQ_TRY(ib.bind(b, sz * n)); AQ_TRY(oc.bind(out, sz * n));
    const int grid = grid_for(n, 256, 1, 16);
    switch (dtype) {
    case AQ_INT32: AQ_LAUNCH(k_vec_sub<int>, grid, 256, 0, (const int*)ia.d, (const int*)ib.d, (int*)oc.d, n); break;
    case AQ_INT64: AQ_LAUNCH(k_vec_sub<long long>, grid, 256, 0, (const long long*)ia.d, (const long long*)ib.d, (long long*)oc.d, n); break;
    case AQ_FLOAT: AQ_LAUNCH(k_vec_sub<float>, grid, 256, 0, (const float*)ia.d, (const float*)ib.d, (float*)oc.d, n); break;
    default: AQ_LAUNCH(k_vec_sub<double>, grid, 256, 0, (const double*)ia.d, (const double*)ib.d, (double*)oc.d, n); break;
    }
    AQ_TRY(oc.finish(sz * n));
    if (oc.host || ia.tmp.p || ib.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_avg_finalize(int sum_type, const void* sum, const uint32_t* count, double* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    if (sum_type != AQ_DOUBLE && sum_type != AQ_INT64 && sum_type != AQ_INT128) return fail("avg_finalize: bad sum dtype %d", sum_type);
    InCol is, ic; OutCol oc;
    AQ_TRY(is.bind(sum, dtype_size(sum_type) * n)); AQ_TRY(ic.bind(count, size_t(4) * n)); AQ_TRY(oc.bind(out, size_t(8) * n));
    AQ_LAUNCH(k_avg_finalize, grid_for(n, 256, 1, 16), 256, 0, sum_type, is.d, (const uint32_t*)ic.d, (double*)oc.d, n);
    AQ_TRY(oc.finish(size_t(8) * n));
    if (oc.host || is.tmp.p || ic.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_vec_pow(const double* x, double p, double* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    InCol ix; OutCol oc;
    AQ_TRY(ix.bind(x, size_t(8) * n)); AQ_TRY(oc.bind(out, size_t(8) * n));
    AQ_LAUNCH(k_vec_pow, grid_for(n, 256, 1, 16), 256, 0, (const double*)ix.d, p, (double*)oc.d, n);
    AQ_TRY(oc.finish(size_t(8) * n));
    if (oc.host || ix.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

} // extern "C"
