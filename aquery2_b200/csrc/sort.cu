// sort.cu -- stable LSD radix sort of (32-bit key, row id) pairs and generic scans.
//
// Used for (a) the CSR row lists of a group-by -- the reference's ht_postproc counting sort
// (server/hasher.h:181-198): rows grouped by gid, DESCENDING row id inside a group -- and (b)
// ASSUMING-ordered evaluation, the per-group std::sort of TableInfo::order_by
// (server/table.h:447-461), as a segmented sort: LSD passes over the order columns, then over gid.
//
// One pass = 3 kernels: per-tile digit histogram (digit-major), exclusive scan of the histogram,
// stable scatter.  Stability inside a tile: every warp owns a contiguous run of the tile, ranks its
// 32 keys per round with match.any, keeps per-(warp,digit) running counts in shared memory; the
// tile is then reordered through shared memory so that global writes are runs of equal digits.
#include "sort.cuh"

namespace aq {

// ------------------------------------------------------------------------------------------------
// generic exclusive scan over uint32 (3 kernels)
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t block_scan_excl(uint32_t v, uint32_t* total, uint32_t* smem /*>= 8 words*/) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) smem[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = lane < SCAN_THREADS / 32 ? smem[lane] : 0;
#pragma unroll
        for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += y;
        }
        if (lane < SCAN_THREADS / 32) smem[lane] = s;
    }
    __syncthreads();
    const uint32_t base = wid ? smem[wid - 1] : 0;
    if (total) *total = smem[SCAN_THREADS / 32 - 1];
    __syncthreads();
    return base + x - v;
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_tilesum(const uint32_t* __restrict__ in, size_t n, uint32_t* __restrict__ tilesum) {
    __shared__ uint32_t sm[8];
    const size_t i0 = size_t(blockIdx.x) * SCAN_TILE + size_t(threadIdx.x) * SCAN_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) if (i0 + i < n) s += in[i0 + i];
    uint32_t total;
    block_scan_excl(s, &total, sm);
    if (threadIdx.x == 0) tilesum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_tiles_serial(uint32_t* __restrict__ tilesum, uint32_t ntiles, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t sm[8];
    uint32_t carry = 0;
    for (uint32_t b0 = 0; b0 < ntiles; b0 += SCAN_THREADS) {
        const uint32_t i = b0 + threadIdx.x;
        const uint32_t v = i < ntiles ? tilesum[i] : 0;
        uint32_t total;
        const uint32_t ex = block_scan_excl(v, &total, sm);
        if (i < ntiles) tilesum[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_apply(const uint32_t* __restrict__ in, size_t n, const uint32_t* __restrict__ tilesum, uint32_t* __restrict__ out) {
    __shared__ uint32_t sm[8];
    const size_t i0 = size_t(blockIdx.x) * SCAN_TILE + size_t(threadIdx.x) * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) { v[i] = (i0 + i < n) ? in[i0 + i] : 0; s += v[i]; }
    uint32_t ex = block_scan_excl(s, nullptr, sm) + tilesum[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        if (i0 + i < n) out[i0 + i] = ex;
        ex += v[i];
    }
}

// out[i] = sum(in[0..i)), i < n; *d_total (device, optional) = sum of all.  in == out is allowed.
// For more than 2^19 tiles the tile sums are scanned recursively.
int exclusive_scan_u32(const uint32_t* d_in, uint32_t* d_out, size_t n, uint32_t* d_total) {
    if (n == 0) {
        if (d_total) AQ_CUDA(cudaMemsetAsync(d_total, 0, 4, ctx().stream));
        return 0;
    }
    const size_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    DevBuf tilesum;
    AQ_TRY(tilesum.alloc(ntiles * 4));
    AQ_LAUNCH(k_scan_tilesum, (unsigned)ntiles, SCAN_THREADS, 0, d_in, n, tilesum.as<uint32_t>());
    if (ntiles <= (1u << 16)) {
        AQ_LAUNCH(k_scan_tiles_serial, 1, SCAN_THREADS, 0, tilesum.as<uint32_t>(), (uint32_t)ntiles, d_total);
    } else {
        AQ_TRY(exclusive_scan_u32(tilesum.as<uint32_t>(), tilesum.as<uint32_t>(), ntiles, d_total));
    }
    AQ_LAUNCH(k_scan_apply, (unsigned)ntiles, SCAN_THREADS, 0, d_in, n, tilesum.as<uint32_t>(), d_out);
    return 0;
}

// ------------------------------------------------------------------------------------------------
// radix sort pass
// ------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 512;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ROUNDS = 16;
constexpr int RS_WARP_ITEMS = 32 * RS_ROUNDS;
constexpr int RS_TILE = RS_THREADS * RS_ROUNDS;       // 8192
constexpr int RS_RADIX = 256;

// Where pass input j comes from.
//   rows_in == nullptr: the implicit row list (reverse ? n-1-j : j)
//   keys_in != nullptr: key of element j is keys_in[j] (carried along), else key = xform(col[row])
struct SortSrc {
    const uint32_t* rows_in;
    const uint32_t* keys_in;
    const uint32_t* col;       // gathered through the row id when keys_in is null
    uint32_t col_stride;
    uint32_t n;
    int reverse;
    uint32_t xor_mask;         // key transform: k ^ xor_mask, float keys via float_flip
    int float_flip;            // 1: IEEE float -> order-preserving uint first
    int shift;
    uint32_t digit_mask;
};

__device__ __forceinline__ uint32_t sort_row(const SortSrc& s, uint32_t j) {
    return s.rows_in ? s.rows_in[j] : (s.reverse ? s.n - 1u - j : j);
}
__device__ __forceinline__ uint32_t sort_key(const SortSrc& s, uint32_t j, uint32_t row) {
    if (s.keys_in) return s.keys_in[j];
    uint32_t k = s.col[size_t(row) * s.col_stride];
    if (s.float_flip) k ^= (uint32_t)((int)k >> 31) | 0x80000000u;   // negative: flip all, else flip sign
    return k ^ s.xor_mask;
}

__global__ void __launch_bounds__(RS_THREADS)
k_radix_hist(const SortSrc s, uint32_t ntiles, uint32_t* __restrict__ hist /*[RADIX][ntiles]*/) {
    __shared__ uint32_t h[RS_RADIX];
    for (int i = threadIdx.x; i < RS_RADIX; i += RS_THREADS) h[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE;
    for (int r = 0; r < RS_ROUNDS; ++r) {
        const uint32_t j = base + r * RS_THREADS + threadIdx.x;
        if (j < s.n) {
            const uint32_t k = sort_key(s, j, sort_row(s, j));
            atomicAdd(&h[(k >> s.shift) & s.digit_mask], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RS_RADIX; i += RS_THREADS) hist[size_t(i) * ntiles + blockIdx.x] = h[i];
}

__global__ void __launch_bounds__(RS_THREADS)
k_radix_scatter(const SortSrc s, uint32_t ntiles, const uint32_t* __restrict__ hist_scanned,
                uint32_t* __restrict__ rows_out, uint32_t* __restrict__ keys_out) {
    __shared__ uint32_t warp_cnt[RS_WARPS][RS_RADIX];     // counts, then exclusive bases
    __shared__ uint32_t digit_start[RS_RADIX];            // start of each digit inside the tile
    __shared__ uint32_t gbase[RS_RADIX];                  // global destination of each digit's run
    extern __shared__ uint32_t rs_dyn[];                  // 2 x RS_TILE words (dynamic: above 48 KB)
    uint32_t* st_key = rs_dyn;
    uint32_t* st_row = rs_dyn + RS_TILE;
    __shared__ uint32_t scan_sm[RS_WARPS];

    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();

    const uint32_t tile0 = blockIdx.x * RS_TILE;
    const uint32_t wbase = tile0 + wid * RS_WARP_ITEMS;
    uint32_t key[RS_ROUNDS], row[RS_ROUNDS], lrank[RS_ROUNDS];
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; ++r) {
        const uint32_t j = wbase + r * 32 + lane;
        const bool valid = j < s.n;
        uint32_t d = RS_RADIX;                            // invalid lanes form their own class
        if (valid) {
            row[r] = sort_row(s, j);
            key[r] = sort_key(s, j, row[r]);
            d = (key[r] >> s.shift) & s.digit_mask;
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        const uint32_t before = __popc(peers & ((1u << lane) - 1u));
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = warp_cnt[wid][d];
            warp_cnt[wid][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        lrank[r] = old + before;
        __syncwarp();
    }
    __syncthreads();

    // per digit: exclusive prefix over warps, digit totals, tile-level digit starts
    uint32_t tot = 0;
    if (threadIdx.x < RS_RADIX) {
        const int d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) { const uint32_t c = warp_cnt[w][d]; warp_cnt[w][d] = run; run += c; }
        tot = run;
        gbase[d] = hist_scanned[size_t(d) * ntiles + blockIdx.x];
    }
    {   // block exclusive scan of tot over the first 256 threads (all 512 take part)
        uint32_t x = tot;
#pragma unroll
        for (int dd = 1; dd < 32; dd <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, dd);
            if (lane >= dd) x += y;
        }
        if (lane == 31) scan_sm[wid] = x;
        __syncthreads();
        if (wid == 0) {
            uint32_t v = lane < RS_WARPS ? scan_sm[lane] : 0;
#pragma unroll
            for (int dd = 1; dd < RS_WARPS; dd <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, v, dd);
                if (lane >= dd) v += y;
            }
            if (lane < RS_WARPS) scan_sm[lane] = v;
        }
        __syncthreads();
        const uint32_t ex = (wid ? scan_sm[wid - 1] : 0) + x - tot;
        if (threadIdx.x < RS_RADIX) digit_start[threadIdx.x] = ex;
    }
    __syncthreads();

    // reorder the tile through shared memory
#pragma unroll
    for (int r = 0; r < RS_ROUNDS; ++r) {
        const uint32_t j = wbase + r * 32 + lane;
        if (j < s.n) {
            const uint32_t d = (key[r] >> s.shift) & s.digit_mask;
            const uint32_t pos = digit_start[d] + warp_cnt[wid][d] + lrank[r];
            st_key[pos] = key[r];
            st_row[pos] = row[r];
        }
    }
    __syncthreads();
    const uint32_t tile_n = min((uint32_t)RS_TILE, s.n - tile0);
    for (uint32_t p = threadIdx.x; p < tile_n; p += RS_THREADS) {
        const uint32_t k = st_key[p];
        const uint32_t d = (k >> s.shift) & s.digit_mask;
        const uint32_t dst = gbase[d] + (p - digit_start[d]);
        rows_out[dst] = st_row[p];
        if (keys_out) keys_out[dst] = k;
    }
}

static int radix_pass(const SortSrc& s, uint32_t* rows_out, uint32_t* keys_out) {
    const uint32_t ntiles = (s.n + RS_TILE - 1) / RS_TILE;
    DevBuf hist;
    AQ_TRY(hist.alloc(size_t(RS_RADIX) * ntiles * 4));
    AQ_LAUNCH(k_radix_hist, ntiles, RS_THREADS, 0, s, ntiles, hist.as<uint32_t>());
    AQ_TRY(exclusive_scan_u32(hist.as<uint32_t>(), hist.as<uint32_t>(), size_t(RS_RADIX) * ntiles, nullptr));
    static bool attr_set = false;
    const size_t dyn = size_t(2) * RS_TILE * sizeof(uint32_t);
    if (!attr_set) {
        AQ_CUDA(cudaFuncSetAttribute(k_radix_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        attr_set = true;
    }
    AQ_LAUNCH(k_radix_scatter, ntiles, RS_THREADS, dyn, s, ntiles, hist.as<uint32_t>(), rows_out, keys_out);
    return 0;
}

// Stable sort of the row list by a 32-bit key held in a column; see sort.cuh.
int sort_rows_by_key(SortJob& job) {
    const uint32_t n = job.n;
    if (n == 0) return 0;
    DevBuf alt_rows, keys_a, keys_b;
    AQ_TRY(alt_rows.alloc(size_t(n) * 4));
    const int passes = (job.key_bits + 7) / 8;
    if (passes == 0) {                 // a single key value: the order is the input order
        if (job.rows_in) { AQ_CUDA(cudaMemcpyAsync(job.rows_out, job.rows_in, size_t(n) * 4, cudaMemcpyDeviceToDevice, ctx().stream)); return 0; }
        SortSrc s{};                   // materialise the implicit list with a 0-bit pass
        s.rows_in = nullptr; s.keys_in = nullptr; s.col = job.col; s.col_stride = job.col_stride; s.n = n; s.reverse = job.reverse;
        s.xor_mask = job.xor_mask; s.float_flip = job.float_flip; s.shift = 0; s.digit_mask = 0;
        return radix_pass(s, job.rows_out, nullptr);
    }
    if (passes > 1) { AQ_TRY(keys_a.alloc(size_t(n) * 4)); AQ_TRY(keys_b.alloc(size_t(n) * 4)); }
    // ping-pong so that the last pass lands in rows_out
    uint32_t* bufs[2] = { job.rows_out, alt_rows.as<uint32_t>() };
    uint32_t* kbufs[2] = { keys_a.as<uint32_t>(), keys_b.as<uint32_t>() };
    int cur = (passes & 1) ? 0 : 1;    // destination of pass 0
    const uint32_t* rows_in = job.rows_in;
    const uint32_t* keys_in = nullptr;
    for (int p = 0; p < passes; ++p) {
        SortSrc s{};
        s.rows_in = rows_in; s.keys_in = keys_in; s.col = job.col; s.col_stride = job.col_stride; s.n = n;
        s.reverse = (p == 0 && !job.rows_in) ? job.reverse : 0;
        s.xor_mask = job.xor_mask; s.float_flip = job.float_flip;
        s.shift = p * 8;
        const int bits = job.key_bits - p * 8 >= 8 ? 8 : job.key_bits - p * 8;
        s.digit_mask = (1u << bits) - 1u;
        uint32_t* kout = (p + 1 < passes) ? kbufs[p & 1] : nullptr;
        AQ_TRY(radix_pass(s, bufs[cur], kout));
        rows_in = bufs[cur];
        keys_in = kout;
        cur ^= 1;
    }
    return 0;
}

} // namespace aq
