// merge.cu -- the G-sized steps of a multi-GPU group-by: finalisation of merged
// partial states, local -> global row ids, owner rank of a key tuple.  Rows are sharded by global row
// index; every rank aggregates its shard into decomposable partial states, the partial tables are
// exchanged (all-gather, or hash-partitioned all-to-all) and merged by a second aqgpu_groupby_agg whose
// input is the concatenation of the partial tables in rank order -- which keeps the reference's group
// order: first-appearance order over contiguous shards = first appearance in that concatenation.
// Finalisation formulas follow server/aggregations.h (cited per case).
#include "groupby.cuh"

namespace aq {

__device__ __forceinline__ double cell_f64(int t, const void* p, uint32_t i) {
    if (t == AQ_DOUBLE) return static_cast<const double*>(p)[i];
    if (t == AQ_INT64) return (double)static_cast<const long long*>(p)[i];
    return (double)static_cast<const __int128*>(p)[i];
}
__device__ __forceinline__ __int128 cell_i128(int t, const void* p, uint32_t i) {
    if (t == AQ_INT64) return (__int128)static_cast<const long long*>(p)[i];
    return static_cast<const __int128*>(p)[i];
}

struct FinArgs { const void* s[5]; };

__global__ void k_finalize(int op, uint32_t n, const long long* __restrict__ count, int st, const FinArgs a, void* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const long long cnt = count[i];
        switch (op) {
        case AQ_COUNT: static_cast<uint32_t*>(out)[i] = (uint32_t)cnt; break;
        case AQ_AVG: static_cast<double*>(out)[i] = cell_f64(st, a.s[0], i) / (double)cnt; break;              // aggregations.h:30
        case AQ_VAR: case AQ_STDDEV: {                                                                         // :332-348 (len + 1 kept)
            const double np1 = (double)(cnt + 1);
            double r;
            if (st == AQ_DOUBLE) { const double s = cell_f64(st, a.s[0], i), q = cell_f64(st, a.s[1], i); r = (q - s * s / np1) / np1; }
            else { const __int128 s = cell_i128(st, a.s[0], i); r = (cell_f64(st, a.s[1], i) - (double)(s * s) / np1) / np1; }
            static_cast<double*>(out)[i] = op == AQ_STDDEV ? sqrt(r) : r;
            break; }
        case AQ_CORR: {                                                                                        // :400-406
            const __int128 len = (__int128)cnt;
            const __int128 sx = cell_i128(st, a.s[0], i), sy = cell_i128(st, a.s[1], i), sxy = cell_i128(st, a.s[2], i),
                           sx2 = cell_i128(st, a.s[3], i), sy2 = cell_i128(st, a.s[4], i);
            const double num = (double)(len * sxy) - (double)(sx * sy);
            const double dx = (double)(len * sx2) - (double)(sx * sx);
            const double dy = (double)(len * sy2) - (double)(sy * sy);
            static_cast<double*>(out)[i] = num / sqrt(dx * dy);
            break; }
        default: break;
        }
    }
}

__global__ void k_rows_to_global(const uint32_t* __restrict__ in, long long add, long long* __restrict__ out, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = (long long)in[i] + add;
}

__global__ void k_key_owner(const KeySpec ks, uint32_t n, uint32_t nparts, uint32_t* __restrict__ owner) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint64_t h = 0x9e3779b97f4a7c15ULL;
        for (int w = 0; w < ks.nwords; ++w) h = mix64(h ^ __ldg(ks.base[w] + size_t(i) * ks.stride[w]));
        owner[i] = (uint32_t)(((h >> 32) * nparts) >> 32);
    }
}

template <class W>
__global__ void k_second_merge(const long long* __restrict__ cnt_top, const long long* __restrict__ nparts,
                               const W* __restrict__ second_top, const W* __restrict__ first_prev, W* __restrict__ out, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = cnt_top[i] >= 2 ? second_top[i] : (nparts[i] >= 2 ? first_prev[i] : W(0));
}

} // namespace aq

using namespace aq;

extern "C" {

int aqgpu_finalize(int op, uint32_t n, const int64_t* count, int s_type, const void* const* s, void* out) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    int ns = 0, osz = 8;
    switch (op) {
    case AQ_COUNT: ns = 0; osz = 4; break;
    case AQ_AVG: ns = 1; break;
    case AQ_VAR: case AQ_STDDEV: ns = 2; break;
    case AQ_CORR: ns = 5; if (s_type == AQ_DOUBLE) return fail("finalize: corr takes integer sums (aggregations.h:387-389)"); break;
    default: return fail("finalize: unsupported op %d", op);
    }
    if (ns && s_type != AQ_INT64 && s_type != AQ_INT128 && s_type != AQ_DOUBLE) return fail("finalize: bad sum dtype %d", s_type);
    if (!count || !out || (ns && !s)) return fail("finalize: null argument");
    InCol ic, is[5]; OutCol oc;
    AQ_TRY(ic.bind(count, size_t(8) * n));
    FinArgs fa{};
    bool staged = ic.tmp.p != nullptr;
    for (int k = 0; k < ns; ++k) {
        if (!s[k]) return fail("finalize: null sum column %d", k);
        AQ_TRY(is[k].bind(s[k], dtype_size(s_type) * n));
        fa.s[k] = is[k].d;
        staged |= is[k].tmp.p != nullptr;
    }
    AQ_TRY(oc.bind(out, size_t(osz) * n));
    AQ_LAUNCH(k_finalize, grid_for(n, 256, 1, 16), 256, 0, op, n, (const long long*)ic.d, s_type, fa, oc.d);
    AQ_TRY(oc.finish(size_t(osz) * n));
    if (oc.host || staged) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_rows_to_global(const uint32_t* in, int64_t add, int64_t* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    InCol ii; OutCol oc;
    AQ_TRY(ii.bind(in, size_t(4) * n)); AQ_TRY(oc.bind(out, size_t(8) * n));
    AQ_LAUNCH(k_rows_to_global, grid_for(n, 256, 1, 16), 256, 0, (const uint32_t*)ii.d, (long long)add, (long long*)oc.d, n);
    AQ_TRY(oc.finish(size_t(8) * n));
    if (oc.host || ii.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

int aqgpu_key_owner(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows, uint32_t nparts, uint32_t* owner) {
    AQ_REQUIRE_INIT();
    if (nparts == 0 || !owner) return fail("key_owner: bad arguments");
    if (nrows == 0) return 0;
    KeyCols kc;
    AQ_TRY(kc.bind(keys, key_types, nkeys, nrows));
    KeySpec ks;
    AQ_TRY(kc.make_spec(nullptr, nullptr, ks));
    OutCol oc;
    AQ_TRY(oc.bind(owner, size_t(4) * nrows));
    AQ_LAUNCH(k_key_owner, grid_for(nrows, 256, 1, 16), 256, 0, ks, nrows, nparts, (uint32_t*)oc.d);
    AQ_TRY(oc.finish(size_t(4) * nrows));
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));         // staged key columns are released on return
    return 0;
}

int aqgpu_second_merge(int vtype, const int64_t* cnt_top, const int64_t* nparts, const void* second_top,
                       const void* first_prev, void* out, uint32_t n) {
    AQ_REQUIRE_INIT();
    if (n == 0) return 0;
    const size_t sz = dtype_size(vtype);
    if (sz != 4 && sz != 8) return fail("second_merge: 4- or 8-byte cells only");
    InCol a, b, c, d; OutCol oc;
    AQ_TRY(a.bind(cnt_top, size_t(8) * n)); AQ_TRY(b.bind(nparts, size_t(8) * n));
    AQ_TRY(c.bind(second_top, sz * n)); AQ_TRY(d.bind(first_prev, sz * n)); AQ_TRY(oc.bind(out, sz * n));
    const int grid = grid_for(n, 256, 1, 16);
    if (sz == 4) AQ_LAUNCH(k_second_merge<uint32_t>, grid, 256, 0, (const long long*)a.d, (const long long*)b.d, (const uint32_t*)c.d, (const uint32_t*)d.d, (uint32_t*)oc.d, n);
    else AQ_LAUNCH(k_second_merge<unsigned long long>, grid, 256, 0, (const long long*)a.d, (const long long*)b.d, (const unsigned long long*)c.d, (const unsigned long long*)d.d, (unsigned long long*)oc.d, n);
    AQ_TRY(oc.finish(sz * n));
    if (oc.host || a.tmp.p || b.tmp.p || c.tmp.p || d.tmp.p) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

} // extern "C"
