// common.cuh -- shared declarations of libaqgpu (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include <string>
#include <vector>
#include "../../include/aqgpu.h"

namespace aq {

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
struct Context {
    bool ready = false;
    int device = -1;
    int sm_count = 148;
    size_t smem_optin = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;      // the stream every call works on
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::atomic<uint64_t> launches{0};
    bool profiling = false;             // aqgpu_profile_enable
    // small device scratch reused by every call (flags, counters), and a pinned mirror
    uint32_t* d_flags = nullptr;        // 64 words
    uint32_t* h_flags = nullptr;        // pinned, 64 words
};
Context& ctx();

void set_error(const char* fmt, ...);
int fail(const char* fmt, ...);         // sets the message, returns -1

#define AQ_CUDA(expr)                                                                       \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess)                                                              \
            return ::aq::fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
    } while (0)

#define AQ_TRY(expr)                                                                        \
    do {                                                                                    \
        int _rc = (expr);                                                                   \
        if (_rc != 0) return _rc;                                                           \
    } while (0)

#define AQ_REQUIRE_INIT()                                                                   \
    do {                                                                                    \
        if (!::aq::ctx().ready) {                                                           \
            int _rc = aqgpu_init(-1);                                                       \
            if (_rc != 0) return _rc;                                                       \
        }                                                                                   \
    } while (0)

// Per-kernel CUDA-event timing on the launching stream (aqgpu_profile_*): off by default.
void profile_pre(const char* kernel_name);
void profile_post();

// kernel launch with bookkeeping (gpu_launches in bench.py is this counter)
#define AQ_LAUNCH(kernel, grid, block, smem, ...)                                           \
    do {                                                                                    \
        if (::aq::ctx().profiling) ::aq::profile_pre(#kernel);                              \
        kernel<<<(grid), (block), (smem), ::aq::ctx().stream>>>(__VA_ARGS__);               \
        if (::aq::ctx().profiling) ::aq::profile_post();                                    \
        ::aq::ctx().launches.fetch_add(1, std::memory_order_relaxed);                       \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess)                                                              \
            return ::aq::fail("%s:%d: launch %s -> %s", __FILE__, __LINE__, #kernel,        \
                              cudaGetErrorString(_e));                                      \
    } while (0)

// ------------------------------------------------------------------------------------------------
// stream-ordered device temporaries (cudaMallocAsync pool, release threshold = never)
// ------------------------------------------------------------------------------------------------
// Scratch of one API call comes from a device ARENA (the device counterpart of the reference's per-query
// ScratchSpace bump allocator, server/gc.h:7-35 / server/libaquery.cpp:464-494): one block that grows to the
// largest call seen and is handed out by bumping an offset, reset when the outermost ArenaScope closes.
// Work is stream-ordered, so the next call may reuse the bytes at once.  Buffers that outlive a call (result
// sets, group handles) are marked persistent and come from the stream-ordered pool (cudaMallocAsync).
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    bool persistent = false;
    bool from_arena = false;
    DevBuf() = default;
    explicit DevBuf(bool persist) : persistent(persist) {}
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes), persistent(o.persistent), from_arena(o.from_arena) { o.p = nullptr; o.bytes = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        release(); p = o.p; bytes = o.bytes; persistent = o.persistent; from_arena = o.from_arena; o.p = nullptr; o.bytes = 0; return *this;
    }
    ~DevBuf() { release(); }
    int alloc(size_t n);                 // uninitialised
    int alloc_zero(size_t n);
    void release();
    template <class T> T* as() const { return static_cast<T*>(p); }
};
struct ArenaScope {                      // opened by the entry points of the C ABI that allocate scratch
    ArenaScope();
    ~ArenaScope();
    ArenaScope(const ArenaScope&) = delete;              // holds the API lock: one lock, one unlock
    ArenaScope& operator=(const ArenaScope&) = delete;
};

bool is_device_ptr(const void* p);

// Input column: device pointers pass through, host pointers are staged to HBM on the stream.
struct InCol {
    const void* d = nullptr;
    DevBuf tmp;
    int bind(const void* p, size_t bytes);
};
// Output column: device pointers pass through, host pointers get a device temporary that
// finish() copies back (the caller synchronises the stream afterwards).
struct OutCol {
    void* d = nullptr;
    void* host = nullptr;
    size_t bytes = 0;
    DevBuf tmp;
    int bind(void* p, size_t bytes);
    int finish(size_t used_bytes);
};

inline size_t dtype_size(int t) {
    switch (t) {
    case AQ_INT32: case AQ_UINT32: case AQ_FLOAT: return 4;
    case AQ_INT64: case AQ_DOUBLE: return 8;
    case AQ_INT128: return 16;
    default: return 0;
    }
}
inline bool is_int_type(int t) { return t == AQ_INT32 || t == AQ_INT64; }
inline bool is_val_type(int t) { return is_int_type(t) || t == AQ_FLOAT || t == AQ_DOUBLE; }

inline int grid_for(size_t work_items, int block, int per_thread = 1, int waves = 8) {
    size_t per_block = size_t(block) * per_thread;
    size_t need = (work_items + per_block - 1) / per_block;
    size_t cap = size_t(ctx().sm_count) * waves;
    if (need < 1) need = 1;
    return int(need < cap ? need : cap);
}

// ------------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL;
    x ^= x >> 32; x *= 0xd6e8feb86659fd93ULL;
    x ^= x >> 32;
    return x;
}
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
// streaming 128-bit load that does not allocate in L1 (columns are read exactly once)
__device__ __forceinline__ int4 ld_stream_v4(const void* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const void* p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
// order-preserving map double <-> int64 (signed compare of the keys == numeric compare)
__device__ __forceinline__ long long f64_to_ordered(double d) {
    long long b = __double_as_longlong(d);
    return b ^ ((b >> 63) & 0x7fffffffffffffffLL);
}
__device__ __forceinline__ double ordered_to_f64(long long k) {
    return __longlong_as_double(k ^ ((k >> 63) & 0x7fffffffffffffffLL));
}
#endif

} // namespace aq
