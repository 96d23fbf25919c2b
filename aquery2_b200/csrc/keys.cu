// keys.cu -- key columns that are not plain integers: strings and date / time / timestamp cells.
//
// The reference groups such columns through the same AQHashTable with type-specific hashes and equality:
//   astring_view   server/hasher.h:97-106 (hash of the characters), server/types.h:299-309 (equality: same characters)
//   date_t         server/hasher.h:108-113, {day u8, month u8, year i16} = 4 bytes (server/types.h:82-86);
//                  equality field by field (server/libaquery.cpp:112-114) = equality of the 4-byte image
//   time_t         server/hasher.h:115-123, {ms u32, seconds u8, minutes u8, hours u8} + 1 padding byte = 8 bytes
//                  (server/types.h:105-109); equality field by field (server/libaquery.cpp:162-164): the padding byte
//                  does NOT take part
//   timestamp_t    server/hasher.h:125-131, {date_t, time_t} = 12 bytes (server/types.h:128-130)
// Only equality and first-appearance order are observable (the hash values are not), so here every such column is
// LOWERED to one or two integer key columns that are equal exactly when the reference's operator== says so, and the
// integer group-by paths (plan.cu: lc / mc / hcfp / generic) do the rest:
//   date_t       -> the 4-byte image as AQ_UINT32 (no kernel: pass the column as AQ_UINT32)
//   time_t       -> int64 = the 8-byte image with the padding byte cleared (also order-preserving: h, m, s, ms)
//   timestamp_t  -> AQ_UINT32 date image + int64 time image
//   strings      -> dictionary codes in FIRST-APPEARANCE order (code = insertion index of the string, what m_values
//                   holds for a one-column string key): two independent 64-bit hashes per string are grouped by the
//                   integer path, then EVERY row's characters are compared with its group's first row -- a 128-bit
//                   collision cannot go unnoticed (the call fails loudly instead of merging two strings).
// Strings arrive Arrow-style (one byte buffer + nrows+1 offsets): a pointer array into a database heap
// (ColRef<astring_view>, server/monetdb_conn.cpp:203-224) means nothing on the device; packing is the ingest step.
#include "common.cuh"

namespace aq {

// ---- time_t / timestamp_t <-> integer images ------------------------------------------------------------------
constexpr unsigned long long TIME_MASK = 0x00ffffffffffffffULL;      // byte 7 = padding

__global__ void __launch_bounds__(256)
k_time_split(const unsigned long long* __restrict__ in, const uint32_t n, long long* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = (long long)(in[i] & TIME_MASK);
}
// 12-byte cells {date, ms, s|m<<8|h<<16|pad<<24}: three 32-bit loads per row, consecutive rows are contiguous
__global__ void __launch_bounds__(256)
k_ts_split(const uint32_t* __restrict__ in, const uint32_t n, uint32_t* __restrict__ date, long long* __restrict__ time) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t* c = in + size_t(i) * 3;
        const uint32_t d = c[0], ms = c[1], hms = c[2] & 0x00ffffffu;
        date[i] = d;
        time[i] = (long long)(((unsigned long long)hms << 32) | ms);
    }
}
__global__ void __launch_bounds__(256)
k_ts_join(const uint32_t* __restrict__ date, const long long* __restrict__ time, const uint32_t n, uint32_t* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned long long t = (unsigned long long)time[i];
        uint32_t* c = out + size_t(i) * 3;
        c[0] = date[i]; c[1] = (uint32_t)t; c[2] = (uint32_t)(t >> 32) & 0x00ffffffu;
    }
}

// ---- strings ------------------------------------------------------------------------------------------------------
// Two independent 64-bit hashes of the characters [off[i], off[i + 1]) (the length is mixed in: "a" + "" != "" + "a"
// across rows is not an issue, but a trailing zero byte would otherwise hash like a shorter string under h1's start).
__global__ void __launch_bounds__(256)
k_str_hash(const unsigned char* __restrict__ bytes, const unsigned long long* __restrict__ off, const uint32_t n,
           long long* __restrict__ h1, long long* __restrict__ h2) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned long long b = off[i], e = off[i + 1];
        unsigned long long a = 0xcbf29ce484222325ULL, c = 0x9e3779b97f4a7c15ULL;
        for (unsigned long long p = b; p < e; ++p) {
            const unsigned long long ch = bytes[p];
            a = (a ^ ch) * 0x100000001b3ULL;                                   // FNV-1a
            c = (c + ch + 1ull) * 0xd6e8feb86659fd93ULL; c ^= c >> 29;         // multiply-xorshift, different constants
        }
        const unsigned long long len = e - b;
        h1[i] = (long long)mix64(a ^ (len * 0xff51afd7ed558ccdULL));
        h2[i] = (long long)mix64(c + len);
    }
}
// every row's characters against those of the first row of its group: equal hashes must mean equal strings
__global__ void __launch_bounds__(256)
k_str_verify(const unsigned char* __restrict__ bytes, const unsigned long long* __restrict__ off, const uint32_t* __restrict__ code,
             const uint32_t* __restrict__ first_row, const uint32_t n, uint32_t* __restrict__ bad) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t f = first_row[code[i]];
        if (f == i) continue;
        const unsigned long long b = off[i], e = off[i + 1], fb = off[f], fe = off[f + 1];
        bool same = (e - b) == (fe - fb);
        for (unsigned long long k = 0; same && k < e - b; ++k) same = bytes[b + k] == bytes[fb + k];
        if (!same) atomicExch(bad, 1u);
    }
}
// offsets must ascend and stay inside the byte buffer
__global__ void __launch_bounds__(256)
k_str_check_offsets(const unsigned long long* __restrict__ off, const uint32_t n, const unsigned long long nbytes, uint32_t* __restrict__ bad) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (off[i] > off[i + 1] || off[i + 1] > nbytes) atomicExch(bad, 2u);
}

struct GroupsGuard {
    aqgpu_groups_t* g = nullptr;
    ~GroupsGuard() { if (g) aqgpu_groups_free(g); }
};

} // namespace aq

using namespace aq;

extern "C" size_t aqgpu_key_cell_size(int key_kind) {
    switch (key_kind) {
    case AQ_KEY_DATE: return 4;
    case AQ_KEY_TIME: return 8;
    case AQ_KEY_TIMESTAMP: return 12;
    default: return 0;
    }
}

extern "C" int aqgpu_key_split(int key_kind, const void* col, uint32_t nrows, void* out_a, void* out_b) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    const size_t cell = aqgpu_key_cell_size(key_kind);
    if (!cell) return fail("key_split: unknown key kind %d", key_kind);
    if (nrows == 0) return 0;
    if (!col || !out_a || (key_kind == AQ_KEY_TIMESTAMP && !out_b)) return fail("key_split: null argument");
    InCol in; OutCol oa, ob;
    AQ_TRY(in.bind(col, cell * nrows));
    const int grid = grid_for(nrows, 256, 1, 16);
    if (key_kind == AQ_KEY_DATE) {                          // the image IS the key
        AQ_TRY(oa.bind(out_a, size_t(nrows) * 4));
        AQ_CUDA(cudaMemcpyAsync(oa.d, in.d, size_t(nrows) * 4, cudaMemcpyDeviceToDevice, ctx().stream));
        AQ_TRY(oa.finish(size_t(nrows) * 4));
    } else if (key_kind == AQ_KEY_TIME) {
        AQ_TRY(oa.bind(out_a, size_t(nrows) * 8));
        AQ_LAUNCH(k_time_split, grid, 256, 0, static_cast<const unsigned long long*>(in.d), nrows, static_cast<long long*>(oa.d));
        AQ_TRY(oa.finish(size_t(nrows) * 8));
    } else {
        AQ_TRY(oa.bind(out_a, size_t(nrows) * 4));
        AQ_TRY(ob.bind(out_b, size_t(nrows) * 8));
        AQ_LAUNCH(k_ts_split, grid, 256, 0, static_cast<const uint32_t*>(in.d), nrows, static_cast<uint32_t*>(oa.d), static_cast<long long*>(ob.d));
        AQ_TRY(oa.finish(size_t(nrows) * 4));
        AQ_TRY(ob.finish(size_t(nrows) * 8));
    }
    if (oa.host || ob.host) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

extern "C" int aqgpu_key_join(int key_kind, const void* a, const void* b, uint32_t nrows, void* out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    const size_t cell = aqgpu_key_cell_size(key_kind);
    if (!cell) return fail("key_join: unknown key kind %d", key_kind);
    if (nrows == 0) return 0;
    if (!a || !out || (key_kind == AQ_KEY_TIMESTAMP && !b)) return fail("key_join: null argument");
    InCol ia, ib; OutCol oc;
    AQ_TRY(oc.bind(out, cell * nrows));
    if (key_kind == AQ_KEY_TIMESTAMP) {
        AQ_TRY(ia.bind(a, size_t(nrows) * 4));
        AQ_TRY(ib.bind(b, size_t(nrows) * 8));
        AQ_LAUNCH(k_ts_join, grid_for(nrows, 256, 1, 16), 256, 0, static_cast<const uint32_t*>(ia.d), static_cast<const long long*>(ib.d), nrows, static_cast<uint32_t*>(oc.d));
    } else {                                                // date: the image; time: the image with a zero padding byte
        AQ_TRY(ia.bind(a, cell * nrows));
        AQ_CUDA(cudaMemcpyAsync(oc.d, ia.d, cell * nrows, cudaMemcpyDeviceToDevice, ctx().stream));
    }
    AQ_TRY(oc.finish(cell * nrows));
    if (oc.host) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

extern "C" int aqgpu_string_dict(const void* bytes, uint64_t nbytes, const uint64_t* offsets, uint32_t nrows,
                                 uint32_t* out_codes, uint32_t max_distinct, uint32_t* out_first_row, uint32_t* ndistinct) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!ndistinct) return fail("string_dict: null ndistinct");
    *ndistinct = 0;
    if (nrows == 0) return 0;
    if (!offsets || !out_codes || (nbytes && !bytes)) return fail("string_dict: null argument");
    Context& c = ctx();
    InCol ib, io; OutCol oc, of;
    AQ_TRY(ib.bind(nbytes ? bytes : (const void*)offsets, nbytes ? (size_t)nbytes : 8));
    AQ_TRY(io.bind(offsets, (size_t(nrows) + 1) * 8));
    AQ_TRY(oc.bind(out_codes, size_t(nrows) * 4));
    const unsigned char* d_bytes = static_cast<const unsigned char*>(ib.d);
    const unsigned long long* d_off = static_cast<const unsigned long long*>(io.d);
    DevBuf h1, h2, bad;
    AQ_TRY(h1.alloc(size_t(nrows) * 8));
    AQ_TRY(h2.alloc(size_t(nrows) * 8));
    AQ_TRY(bad.alloc_zero(16));
    const int grid = grid_for(nrows, 256, 1, 16);
    AQ_LAUNCH(k_str_check_offsets, grid, 256, 0, d_off, nrows, (unsigned long long)nbytes, bad.as<uint32_t>());
    uint32_t flag = 0;
    AQ_CUDA(cudaMemcpyAsync(&flag, bad.p, 4, cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (flag) return fail("string_dict: offsets must ascend and end inside the %llu-byte buffer", (unsigned long long)nbytes);
    AQ_LAUNCH(k_str_hash, grid, 256, 0, d_bytes, d_off, nrows, h1.as<long long>(), h2.as<long long>());
    const void* keys[2] = { h1.p, h2.p };
    const int types[2] = { AQ_INT64, AQ_INT64 };
    GroupsGuard gg;
    AQ_TRY(aqgpu_groupby_build(keys, types, 2, nrows, nullptr, nullptr, AQ_ROWS_DESC, 0, &gg.g));
    uint32_t ng = 0;
    AQ_TRY(aqgpu_groups_size(gg.g, &ng));
    DevBuf first;
    AQ_TRY(first.alloc(size_t(ng) * 4 + 16));
    AQ_TRY(aqgpu_groups_gid(gg.g, static_cast<uint32_t*>(oc.d)));
    AQ_TRY(aqgpu_groups_first_row(gg.g, first.as<uint32_t>()));
    AQ_LAUNCH(k_str_verify, grid, 256, 0, d_bytes, d_off, static_cast<const uint32_t*>(oc.d), first.as<uint32_t>(), nrows, bad.as<uint32_t>());
    AQ_CUDA(cudaMemcpyAsync(&flag, bad.p, 4, cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (flag) return fail("string_dict: two different strings share both 64-bit hashes (128-bit collision); refusing to merge them");
    *ndistinct = ng;
    if (out_first_row) {
        if (ng > max_distinct) return fail("string_dict: capacity: %u distinct strings, room for %u", ng, max_distinct);
        AQ_TRY(of.bind(out_first_row, size_t(ng) * 4));
        AQ_CUDA(cudaMemcpyAsync(of.d, first.p, size_t(ng) * 4, cudaMemcpyDeviceToDevice, c.stream));
        AQ_TRY(of.finish(size_t(ng) * 4));
    }
    AQ_TRY(oc.finish(size_t(nrows) * 4));
    if (oc.host || of.host) AQ_CUDA(cudaStreamSynchronize(c.stream));
    return 0;
}
