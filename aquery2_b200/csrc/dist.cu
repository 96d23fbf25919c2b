// dist.cu -- multi-GPU group-by: row shards, one process per GPU, dense record tables exchanged over NVLink.
//
// rows are sharded contiguously by global row index; every rank aggregates its shard,
// partial states are exchanged, the owner merges.  The reference has no multi-process path (single engine thread,
// server/server.cpp:236-238), so there is no reference code to cite for the exchange itself; results are those of
// the reference's single-process loops (server/hasher.h:176-198, server/aggregations.h) on the whole table.
//
// The design rests on the single-GPU paths keeping their partial state in DENSE, DIRECTLY INDEXED record tables
// (lc.cu: one record per mixed-radix key slot; mc.cu: the same, 1e7 slots) whose cells are integers / 128-bit fixed
// point (cells.cuh): the sum of two ranks' records is exact and order-independent, so "merge" is an element-wise
// combination of equally laid out arrays -- no second hash build, no partial-row repartitioning:
//   * few groups (lc.cu):  all-gather of the G-record tables (<= 4 MB per rank), every rank combines all of them
//     and emits the whole result (replicated);
//   * 1e5..2.6e7 slots (mc.cu): rank d owns the slot range [d * Gd, (d + 1) * Gd); one grouped ncclSend/ncclRecv
//     all-to-all moves every rank's slice of that range to d (contiguous, nothing to pack), d combines and emits
//     the groups it owns in ascending GLOBAL first-row order.
// First-appearance order: every rank finds the first LOCAL row of its groups, adds its shard's row offset, and the
// combination keeps the minimum: the global first-appearance row, hence the reference's group order
// (server/unordered_dense.h:1142-1146).  What the sample-driven planner speculates (key bounds, fixed-point base)
// is agreed on by one small all-reduce before anything runs; a rank whose rows violate it makes EVERY rank decline.
#include "comm.cuh"
#include "groupby.cuh"
#include <cstring>

namespace aq {

DistCtx& dist() { static DistCtx d; return d; }
constexpr int REC_WORDS_MAX = 24;            // u64 words of a record (accumulator cells + row-valued extras)

__global__ void k_dist_globalize(uint32_t* __restrict__ gfirst, uint32_t G, uint32_t base) {
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < G; g += gridDim.x * blockDim.x) {
        const uint32_t f = gfirst[g];
        if (f != 0xffffffffu) gfirst[g] = f + base;
    }
}
__global__ void k_dist_or_flags(const uint32_t* __restrict__ all, int n, uint32_t* __restrict__ flags) {
    uint32_t v = 0;
    for (int i = 0; i < n; ++i) if (all[i]) v = all[i];
    if (threadIdx.x == 0 && blockIdx.x == 0 && v) flags[FLAG_ABORT] = v;
}

// out record i = combination of record i of each of the nr input tables (nr x n records, table r at in_tab + r * stride_rec
// records); empty records (cnt word 0) are skipped.  first rows: minimum.
__global__ void __launch_bounds__(256)
k_dist_combine(unsigned long long* __restrict__ out_tab, uint32_t* __restrict__ out_first, const unsigned long long* __restrict__ in_tab,
               const uint32_t* __restrict__ in_first, const uint32_t n, const int nr, const size_t stride_rec, const RecLayout L) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned long long acc[REC_WORDS_MAX];
#pragma unroll
        for (int w = 0; w < REC_WORDS_MAX; ++w) acc[w] = 0ull;
        for (int a = 0; a < L.n; ++a) {
            if (L.kind[a] == RK_MIN64) acc[L.word[a]] = 0x7fffffffffffffffULL;
            if (L.kind[a] == RK_MAX64) acc[L.word[a]] = 0x8000000000000000ULL;
            if (L.kind[a] == RK_UMIN64) acc[L.word[a]] = ~0ull;
        }
        uint32_t first = 0xffffffffu;
        for (int r = 0; r < nr; ++r) {
            const unsigned long long* rec = in_tab + (size_t(r) * stride_rec + i) * L.gw;
            first = min(first, in_first[size_t(r) * stride_rec + i]);
            if (rec[L.cnt_word] == 0ull) continue;
            for (int a = 0; a < L.n; ++a) {
                const uint32_t w = L.word[a];
                switch (L.kind[a]) {
                case RK_ADD64: acc[w] += rec[w]; break;
                case RK_ADD128: {
                    const unsigned long long lo = acc[w] + rec[w];
                    acc[w + 1] += rec[w + 1] + (lo < acc[w] ? 1ull : 0ull);
                    acc[w] = lo;
                    break; }
                case RK_MIN64: acc[w] = (unsigned long long)min((long long)acc[w], (long long)rec[w]); break;
                case RK_MAX64: acc[w] = (unsigned long long)max((long long)acc[w], (long long)rec[w]); break;
                case RK_UMAX64: acc[w] = max(acc[w], rec[w]); break;
                case RK_UMIN64: acc[w] = min(acc[w], rec[w]); break;
                case RK_TOP2: {
                    const unsigned long long a0 = acc[w], a1 = acc[w + 1], b0 = rec[w], b1 = rec[w + 1];
                    acc[w] = max(a0, b0);
                    acc[w + 1] = max(min(a0, b0), max(a1, b1));
                    break; }
                default: break;
                }
            }
        }
        unsigned long long* o = out_tab + size_t(i) * L.gw;
        for (uint32_t w = 0; w < L.gw; ++w) o[w] = acc[w];
        out_first[i] = first;
    }
}

int dist_exchange_flags(uint32_t* flags) {
    DistCtx& d = dist();
    DevBuf all;
    AQ_TRY(all.alloc(size_t(d.nranks) * 4));
    AQ_TRY(comm_allgather(flags + FLAG_ABORT, all.p, 4));
    AQ_LAUNCH(k_dist_or_flags, 1, 32, 0, all.as<uint32_t>(), d.nranks, flags);
    return 0;
}

int dist_combine_replicated(unsigned long long* gtab, uint32_t* gfirst, uint32_t G, const RecLayout& L, uint32_t* flags, uint32_t* any_abort) {
    (void)any_abort;
    if (L.gw > REC_WORDS_MAX) return fail("internal: %u-word records", L.gw);
    DistCtx& d = dist();
    const int nr = d.nranks;
    AQ_LAUNCH(k_dist_globalize, grid_for(G, 256), 256, 0, gfirst, G, (uint32_t)d.row_base);
    AQ_TRY(dist_exchange_flags(flags));
    DevBuf tabs, firsts;
    const size_t tb = size_t(G) * L.gw * 8, fb = size_t(G) * 4;
    AQ_TRY(tabs.alloc(tb * nr));
    AQ_TRY(firsts.alloc(fb * nr));
    AQ_TRY(comm_allgather(gtab, tabs.p, tb));
    AQ_TRY(comm_allgather(gfirst, firsts.p, fb));
    AQ_LAUNCH(k_dist_combine, grid_for(G, 256), 256, 0, gtab, gfirst, tabs.as<unsigned long long>(), firsts.as<uint32_t>(), G, nr, (size_t)G, L);
    d.replicated = true;
    d.exchange_bytes += (tb + fb) * (nr - 1);
    return 0;
}

int dist_combine_owned(unsigned long long* gtab, uint32_t* gfirst, uint32_t G, uint32_t Gd, const RecLayout& L, uint32_t* flags, uint32_t* any_abort) {
    (void)any_abort;
    if (L.gw > REC_WORDS_MAX) return fail("internal: %u-word records", L.gw);
    DistCtx& d = dist();
    const int nr = d.nranks, me = d.rank;
    AQ_LAUNCH(k_dist_globalize, grid_for(G, 256), 256, 0, gfirst, G, (uint32_t)d.row_base);
    AQ_TRY(dist_exchange_flags(flags));
    auto lo_of = [&](int r) { return (uint32_t)std::min<unsigned long long>(G, (unsigned long long)r * Gd); };
    const uint32_t my_lo = lo_of(me), my_n = lo_of(me + 1) - my_lo;
    DevBuf tabs, firsts;
    AQ_TRY(tabs.alloc(size_t(Gd) * L.gw * 8 * nr + 16));
    AQ_TRY(firsts.alloc(size_t(Gd) * 4 * nr + 16));
    size_t so[64], sb[64], ro[64], rb[64];
    if (nr > 64) return fail("multi-GPU: more than 64 ranks");
    for (int r = 0; r < nr; ++r) {
        const uint32_t lo = lo_of(r), n = lo_of(r + 1) - lo;
        so[r] = size_t(lo) * L.gw * 8; sb[r] = size_t(n) * L.gw * 8;
        ro[r] = size_t(r) * Gd * L.gw * 8; rb[r] = size_t(my_n) * L.gw * 8;
    }
    AQ_TRY(comm_alltoallv(gtab, so, sb, tabs.p, ro, rb));
    size_t moved = 0;
    for (int r = 0; r < nr; ++r) if (r != me) moved += rb[r];
    for (int r = 0; r < nr; ++r) {
        const uint32_t lo = lo_of(r), n = lo_of(r + 1) - lo;
        so[r] = size_t(lo) * 4; sb[r] = size_t(n) * 4;
        ro[r] = size_t(r) * Gd * 4; rb[r] = size_t(my_n) * 4;
    }
    AQ_TRY(comm_alltoallv(gfirst, so, sb, firsts.p, ro, rb));
    for (int r = 0; r < nr; ++r) if (r != me) moved += rb[r];
    if (my_n)
        AQ_LAUNCH(k_dist_combine, grid_for(my_n, 256), 256, 0, gtab + size_t(my_lo) * L.gw, gfirst + my_lo, tabs.as<unsigned long long>(),
                  firsts.as<uint32_t>(), my_n, nr, (size_t)Gd, L);
    d.replicated = false;
    d.exchange_bytes += moved;
    return 0;
}

__global__ void k_dist_iota(uint32_t* __restrict__ out, uint32_t n, uint32_t base) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = base + i;
}

} // namespace aq

using namespace aq;

// Rows of a sharded table repartitioned by hash(key tuple) mod ranks, so that every group lies whole on one rank:
// SURVEY.md section 8(e), "windowed / order-sensitive aggregates".  A stable split by owner (aqgpu_key_owner +
// aqgpu_partition_rows), the per-destination row counts all-gathered, then ONE grouped ncclSend / ncclRecv per column.
// Blocks arrive in source-rank order and the split is stable, so the received rows are in ascending GLOBAL row order:
// a table that was in time order still is, inside every group.  The extra output column holds the global row ids.
extern "C" int aqgpu_repartition_rows(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows, uint64_t row_base,
                                      const void* const* cols, const int* elem_bytes, int ncols, void** out_cols, uint32_t* out_nrows) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!comm_ready()) return fail("aqgpu_repartition_rows: call aqgpu_comm_init first");
    if (!keys || !key_types || nkeys < 1 || !cols || !elem_bytes || ncols < 1 || ncols > 30 || !out_cols || !out_nrows)
        return fail("aqgpu_repartition_rows: bad arguments");
    if (row_base + nrows >= 0xffffffffull) return fail("aqgpu_repartition_rows: global row ids must fit 32 bits");
    Context& c = ctx();
    const int N = comm_nranks(), me = comm_rank();
    if (N > 64) return fail("multi-GPU: more than 64 ranks");
    for (int k = 0; k < ncols + 1; ++k) out_cols[k] = nullptr;
    // ---- owner of every row, stable split of the columns (+ the global row ids)
    DevBuf owner, grow, counts_d, counts_all;
    AQ_TRY(owner.alloc(size_t(nrows ? nrows : 1) * 4));
    AQ_TRY(grow.alloc(size_t(nrows ? nrows : 1) * 4));
    if (nrows) {
        AQ_TRY(aqgpu_key_owner(keys, key_types, nkeys, nrows, (uint32_t)N, owner.as<uint32_t>()));
        AQ_LAUNCH(k_dist_iota, grid_for(nrows, 256, 4), 256, 0, grow.as<uint32_t>(), nrows, (uint32_t)row_base);
    }
    std::vector<DevBuf> parted(ncols + 1);
    std::vector<const void*> src(ncols + 1);
    std::vector<void*> dst(ncols + 1);
    std::vector<int> eb(ncols + 1);
    for (int k = 0; k < ncols; ++k) { src[k] = cols[k]; eb[k] = elem_bytes[k]; }
    src[ncols] = grow.p; eb[ncols] = 4;
    for (int k = 0; k <= ncols; ++k) { AQ_TRY(parted[k].alloc(size_t(nrows ? nrows : 1) * eb[k])); dst[k] = parted[k].p; }
    uint32_t send[64];
    AQ_TRY(aqgpu_partition_rows(owner.as<uint32_t>(), nrows, (uint32_t)N, src.data(), eb.data(), ncols + 1, dst.data(), send));
    // ---- who sends how much to whom
    AQ_TRY(counts_d.alloc(size_t(N) * 4));
    AQ_TRY(counts_all.alloc(size_t(N) * N * 4));
    AQ_CUDA(cudaMemcpyAsync(counts_d.p, send, size_t(N) * 4, cudaMemcpyHostToDevice, c.stream));
    AQ_TRY(comm_allgather(counts_d.p, counts_all.p, size_t(N) * 4));
    std::vector<uint32_t> m(size_t(N) * N);
    AQ_CUDA(cudaMemcpyAsync(m.data(), counts_all.p, m.size() * 4, cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    size_t soff[64], roff[64], nrecv = 0, acc = 0;
    for (int r = 0; r < N; ++r) { soff[r] = acc; acc += send[r]; }
    for (int s2 = 0; s2 < N; ++s2) { roff[s2] = nrecv; nrecv += m[size_t(s2) * N + me]; }
    if (nrecv >= 0xffffffffull) return fail("aqgpu_repartition_rows: %zu rows on one rank", nrecv);
    // ---- one grouped exchange per column, into columns that outlive the call (aqgpu_col_free)
    for (int k = 0; k <= ncols; ++k) {
        void* p = nullptr;
        AQ_TRY(aqgpu_col_alloc((nrecv ? nrecv : 1) * eb[k], &p));
        out_cols[k] = p;
        size_t so[64], sb[64], ro[64], rb[64];
        for (int r = 0; r < N; ++r) {
            so[r] = soff[r] * eb[k]; sb[r] = size_t(send[r]) * eb[k];
            ro[r] = roff[r] * eb[k]; rb[r] = size_t(m[size_t(r) * N + me]) * eb[k];
        }
        AQ_TRY(comm_alltoallv(parted[k].p, so, sb, p, ro, rb));
    }
    AQ_CUDA(cudaStreamSynchronize(c.stream));                    // the split copies live in the call's scratch
    *out_nrows = (uint32_t)nrecv;
    return 0;
}
