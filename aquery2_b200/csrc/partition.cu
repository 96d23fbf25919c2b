// partition.cu -- stable split of rows (or partial-aggregate rows) by owner rank, the step before the
// hash-partitioned all-to-all of the multi-GPU group-by (no reference counterpart: the
// reference is single-process).  Rows keep their order inside an owner, so an owner that receives the
// blocks in source-rank order sees every group's rows in ascending global row order -- first-appearance
// order survives the exchange without carrying a sort key.
//
//   k_split_count   rows of every (owner, chunk of SPLIT_ROWS rows)      -> cnt[owner * nchunks + chunk]
//   exclusive scan  (owner-major)                                        -> start of every (owner, chunk)
//   k_split_dest    destination of every row: start + stable rank inside the chunk (match.any per warp,
//                   warp totals in shared memory)
//   k_split_move    out[dest[i]] = in[i], one launch per column (4-, 8- or 16-byte cells)
#include "common.cuh"
#include "sort.cuh"

namespace aq {

constexpr int SPLIT_THREADS = 256;
constexpr int SPLIT_ITERS = 8;
constexpr int SPLIT_ROWS = SPLIT_THREADS * SPLIT_ITERS;
constexpr int SPLIT_MAX_PARTS = 64;

__global__ void __launch_bounds__(SPLIT_THREADS)
k_split_count(const uint32_t* __restrict__ owner, const uint32_t n, const uint32_t nparts, const uint32_t nchunks,
              uint32_t* __restrict__ cnt) {
    __shared__ uint32_t h[SPLIT_MAX_PARTS];
    for (uint32_t chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
        if (threadIdx.x < SPLIT_MAX_PARTS) h[threadIdx.x] = 0;
        __syncthreads();
#pragma unroll
        for (int it = 0; it < SPLIT_ITERS; ++it) {
            const uint32_t i = chunk * SPLIT_ROWS + it * SPLIT_THREADS + threadIdx.x;
            const bool ok = i < n;
            const uint32_t p = ok ? owner[i] : 0u;
            const uint32_t m = __match_any_sync(0xffffffffu, ok ? p : 0xffffffffu);
            if (ok && (__ffs(m) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&h[p], __popc(m));
        }
        __syncthreads();
        if (threadIdx.x < nparts) cnt[size_t(threadIdx.x) * nchunks + chunk] = h[threadIdx.x];
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SPLIT_THREADS)
k_split_dest(const uint32_t* __restrict__ owner, const uint32_t n, const uint32_t nparts, const uint32_t nchunks,
             const uint32_t* __restrict__ pos, uint32_t* __restrict__ dest) {
    constexpr int NW = SPLIT_THREADS / 32;
    __shared__ uint32_t base[SPLIT_MAX_PARTS];
    __shared__ uint32_t wcnt[NW][SPLIT_MAX_PARTS];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (uint32_t chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
        if (threadIdx.x < nparts) base[threadIdx.x] = pos[size_t(threadIdx.x) * nchunks + chunk];
        for (int it = 0; it < SPLIT_ITERS; ++it) {
            for (uint32_t x = threadIdx.x; x < NW * SPLIT_MAX_PARTS; x += SPLIT_THREADS) (&wcnt[0][0])[x] = 0;
            __syncthreads();
            const uint32_t i = chunk * SPLIT_ROWS + it * SPLIT_THREADS + threadIdx.x;
            const bool ok = i < n;
            const uint32_t p = ok ? owner[i] : 0u;
            const uint32_t m = __match_any_sync(0xffffffffu, ok ? p : 0xffffffffu);
            const uint32_t rank = __popc(m & ((1u << lane) - 1u));
            if (ok && rank == 0) wcnt[wid][p] = __popc(m);
            __syncthreads();
            if (ok) {
                uint32_t before = 0;
                for (uint32_t w = 0; w < wid; ++w) before += wcnt[w][p];
                dest[i] = base[p] + before + rank;
            }
            __syncthreads();
            if (threadIdx.x < nparts) {
                uint32_t tot = 0;
#pragma unroll
                for (int w = 0; w < NW; ++w) tot += wcnt[w][threadIdx.x];
                base[threadIdx.x] += tot;
            }
            __syncthreads();                      // wcnt is cleared by the next iteration
        }
    }
}

template <class T>
__global__ void __launch_bounds__(256)
k_split_move(const T* __restrict__ in, const uint32_t* __restrict__ dest, T* __restrict__ out, const uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[dest[i]] = in[i];
}

} // namespace aq

using namespace aq;

extern "C" int aqgpu_partition_rows(const uint32_t* owner, uint32_t nrows, uint32_t nparts, const void* const* cols,
                                    const int* elem_bytes, int ncols, void* const* out_cols, uint32_t* counts) {
    AQ_REQUIRE_INIT();
    if (nparts == 0 || nparts > SPLIT_MAX_PARTS) return fail("partition_rows: 1..%d owners", SPLIT_MAX_PARTS);
    if (!counts || ncols < 0 || (ncols && (!cols || !out_cols || !elem_bytes))) return fail("partition_rows: null argument");
    for (uint32_t p = 0; p < nparts; ++p) counts[p] = 0;
    if (nrows == 0) return 0;
    if (!owner || !is_device_ptr(owner)) return fail("partition_rows: owner must be a device column");
    Context& c = ctx();
    const uint32_t nchunks = (nrows + SPLIT_ROWS - 1) / SPLIT_ROWS;
    const size_t ncnt = size_t(nparts) * nchunks;
    DevBuf cnt, dest;
    AQ_TRY(cnt.alloc((ncnt + 1) * 4));
    AQ_TRY(dest.alloc(size_t(nrows) * 4));
    const int grid = (int)std::min<size_t>(nchunks, size_t(c.sm_count) * 8);
    AQ_LAUNCH(k_split_count, grid, SPLIT_THREADS, 0, owner, nrows, nparts, nchunks, cnt.as<uint32_t>());
    AQ_CUDA(cudaMemsetAsync(cnt.as<uint32_t>() + ncnt, 0, 4, c.stream));
    AQ_TRY(exclusive_scan_u32(cnt.as<uint32_t>(), cnt.as<uint32_t>(), ncnt + 1, nullptr));
    AQ_LAUNCH(k_split_dest, grid, SPLIT_THREADS, 0, owner, nrows, nparts, nchunks, cnt.as<uint32_t>(), dest.as<uint32_t>());
    for (int k = 0; k < ncols; ++k) {
        if (!cols[k] || !out_cols[k] || !is_device_ptr(cols[k]) || !is_device_ptr(out_cols[k]))
            return fail("partition_rows: column %d must be device-resident", k);
        const int g = grid_for(nrows, 256, 1, 16);
        switch (elem_bytes[k]) {
        case 4: AQ_LAUNCH(k_split_move<uint32_t>, g, 256, 0, (const uint32_t*)cols[k], dest.as<uint32_t>(), (uint32_t*)out_cols[k], nrows); break;
        case 8: AQ_LAUNCH(k_split_move<unsigned long long>, g, 256, 0, (const unsigned long long*)cols[k], dest.as<uint32_t>(), (unsigned long long*)out_cols[k], nrows); break;
        case 16: AQ_LAUNCH(k_split_move<uint4>, g, 256, 0, (const uint4*)cols[k], dest.as<uint32_t>(), (uint4*)out_cols[k], nrows); break;
        default: return fail("partition_rows: column %d has %d-byte cells (4, 8 or 16 only)", k, elem_bytes[k]);
        }
    }
    // rows per owner = differences of the owner-major starts
    std::vector<uint32_t> starts(nparts + 1);
    for (uint32_t p = 0; p <= nparts; ++p)
        AQ_CUDA(cudaMemcpyAsync(&starts[p], cnt.as<uint32_t>() + size_t(p) * nchunks, 4, cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    for (uint32_t p = 0; p < nparts; ++p) counts[p] = starts[p + 1] - starts[p];
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the counting step (no counterpart in the reference):
// some callers cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops,
// harnesses that forbid it.  k_split_count's scratch (cnt[owner * nchunks + chunk], what aqgpu_partition_rows
// takes from the arena) lives in a file-scope device array instead, so the entry point below performs no
// cudaMalloc / cudaFree / pool or arena allocation at all.
//   limit: NOALLOC_MAX_ROWS rows (= NOALLOC_MAX_CHUNKS chunks of SPLIT_ROWS rows) x SPLIT_MAX_PARTS owners
//          -> 64 x 8192 x 4 B = 2 MiB of device memory, held by every GPU that loads the library until exit.
//   Every call shares the array and it keeps the previous call's contents.  Nothing in it has to start at zero:
//   k_split_count overwrites every cell [0, nparts * nchunks) that k_split_total then reads.  Calls that overlap
//   on the GPU (two streams, two host threads) race on it: callers serialise them.
// In host code the array's NAME is a host-side stand-in, not a device address; the address a kernel may be given
// comes from cudaGetSymbolAddress, resolved once per device with that device current.
// ------------------------------------------------------------------------------------------------------------
namespace aq {

constexpr uint32_t NOALLOC_MAX_CHUNKS = 8192;
constexpr uint32_t NOALLOC_MAX_ROWS = NOALLOC_MAX_CHUNKS * SPLIT_ROWS;             // 16,777,216 rows
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) uint32_t g_split_scratch[size_t(SPLIT_MAX_PARTS) * NOALLOC_MAX_CHUNKS];
static uint32_t* g_split_scratch_ptr[NOALLOC_MAX_DEVICES];                         // device address, per device

// rows of owner p = sum over chunks of cnt[p * nchunks + chunk]; one CTA per owner
__global__ void __launch_bounds__(SPLIT_THREADS)
k_split_total(const uint32_t* __restrict__ cnt, const uint32_t nchunks, uint32_t* __restrict__ out_counts) {
    __shared__ uint32_t ws[SPLIT_THREADS / 32];
    const uint32_t* row = cnt + size_t(blockIdx.x) * nchunks;
    uint32_t s = 0;
    for (uint32_t c = threadIdx.x; c < nchunks; c += SPLIT_THREADS) s += row[c];
    s = __reduce_add_sync(0xffffffffu, s);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < SPLIT_THREADS / 32; ++w) t += ws[w];
        out_counts[blockIdx.x] = t;
    }
}

} // namespace aq

extern "C" int aqgpu_partition_counts_noalloc(const uint32_t* owner, uint32_t nrows, uint32_t nparts, uint32_t* out_counts) {
    Context& c = ctx();
    if (!c.ready) return fail("partition_counts_noalloc: call aqgpu_init first (set-up allocates, this call does not)");
    if (nparts == 0 || nparts > SPLIT_MAX_PARTS) return fail("partition_counts_noalloc: 1..%d owners", SPLIT_MAX_PARTS);
    if (nrows > NOALLOC_MAX_ROWS)
        return fail("partition_counts_noalloc: %u rows exceed the maximum of %u rows of the static scratch (use aqgpu_partition_rows)",
                    nrows, NOALLOC_MAX_ROWS);
    if (!out_counts || !is_device_ptr(out_counts)) return fail("partition_counts_noalloc: out_counts must be device memory (%u x uint32)", nparts);
    if (nrows && (!owner || !is_device_ptr(owner))) return fail("partition_counts_noalloc: owner must be a device column");
    int cur = -1;
    AQ_CUDA(cudaGetDevice(&cur));
    if (cur != c.device) return fail("partition_counts_noalloc: device %d is current, the library runs on device %d", cur, c.device);
    if (cur < 0 || cur >= NOALLOC_MAX_DEVICES) return fail("partition_counts_noalloc: device ordinal %d", cur);
    if (!g_split_scratch_ptr[cur]) {                                 // set-up, once per device, that device current
        uint32_t* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_split_scratch);
        if (e != cudaSuccess || !p) return fail("partition_counts_noalloc: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
        g_split_scratch_ptr[cur] = p;
    }
    uint32_t* cnt = g_split_scratch_ptr[cur];
    if (nrows == 0) {
        AQ_CUDA(cudaMemsetAsync(out_counts, 0, size_t(nparts) * 4, c.stream));
        return 0;
    }
    const uint32_t nchunks = (nrows + SPLIT_ROWS - 1) / SPLIT_ROWS;
    const int grid = (int)std::min<size_t>(nchunks, size_t(c.sm_count) * 8);
    // plain launches: no profiling hooks (they create events), same launch counter as AQ_LAUNCH
    k_split_count<<<grid, SPLIT_THREADS, 0, c.stream>>>(owner, nrows, nparts, nchunks, cnt);
    k_split_total<<<(int)nparts, SPLIT_THREADS, 0, c.stream>>>(cnt, nchunks, out_counts);
    c.launches.fetch_add(2, std::memory_order_relaxed);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("partition_counts_noalloc: launch -> %s", cudaGetErrorString(e));
    return 0;
}
