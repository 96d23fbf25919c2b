// context.cu -- process context, error reporting, HBM column store primitives of libaqgpu.
#include "common.cuh"
#include <cuda.h>
#include <algorithm>
#include <cstdarg>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace aq {

static Context g_ctx;
static thread_local char g_err[1024] = "";
static std::mutex g_init_mu;

Context& ctx() { return g_ctx; }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return -1;
}

// ---- scratch arena ---------------------------------------------------------------------------------
struct Arena {
    char* base = nullptr;
    size_t cap = 0, off = 0;
    size_t overflow = 0, need = 0;       // bytes served by the pool in this scope / high-water mark of the scope
    int depth = 0;
    std::vector<void*> spilled;
};
static Arena g_arena;
static constexpr size_t ARENA_ALIGN = 256;

// The arena's block is a virtual-address reservation that is backed on demand (driver virtual memory management:
// cuMemAddressReserve + cuMemCreate / cuMemMap, bound at run time through cudaGetDriverEntryPoint) and only ever GROWS,
// in place.  Freeing and re-allocating a larger block through the stream-ordered pool (the first version) fragmented
// the pool: the first handful of 1e9-row steps each stalled for 0.2-0.8 s while the pool re-shuffled tens of GB
// (profiles/r02/arena_warmup.txt).  Memory is returned by aqgpu_trim_pool.
struct VmmArena {
    bool tried = false, ok = false;
    CUresult (*AddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*GetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    CUdeviceptr va = 0;
    size_t reserved = 0, mapped = 0, gran = 0;
    CUmemAllocationProp prop;
    std::vector<std::pair<CUmemGenericAllocationHandle, size_t>> chunks;
};
static VmmArena g_vmm;

static bool vmm_init() {
    VmmArena& v = g_vmm;
    if (v.tried) return v.ok;
    v.tried = true;
    if (getenv("AQGPU_NO_VMM")) return false;
    auto sym = [](const char* name, void** fn) {
        cudaDriverEntryPointQueryResult st;
        return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &st) == cudaSuccess && st == cudaDriverEntryPointSuccess && *fn;
    };
    if (!sym("cuMemAddressReserve", (void**)&v.AddressReserve) || !sym("cuMemCreate", (void**)&v.Create) || !sym("cuMemMap", (void**)&v.Map) ||
        !sym("cuMemSetAccess", (void**)&v.SetAccess) || !sym("cuMemUnmap", (void**)&v.Unmap) || !sym("cuMemRelease", (void**)&v.Release) ||
        !sym("cuMemGetAllocationGranularity", (void**)&v.GetGranularity)) { cudaGetLastError(); return false; }
    memset(&v.prop, 0, sizeof(v.prop));
    v.prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    v.prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    v.prop.location.id = ctx().device;
    if (v.GetGranularity(&v.gran, &v.prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || v.gran == 0) return false;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return false; }
    v.reserved = (total_b + v.gran - 1) / v.gran * v.gran;
    if (v.AddressReserve(&v.va, v.reserved, 0, 0, 0) != CUDA_SUCCESS) return false;
    v.ok = true;
    return true;
}
// backs the reservation up to at least `want` bytes; false: out of memory (the caller falls back to the pool)
static bool vmm_grow(size_t want) {
    VmmArena& v = g_vmm;
    if (!vmm_init() || want > v.reserved) return false;
    while (v.mapped < want) {
        size_t chunk = want - v.mapped;
        const size_t step = size_t(1) << 30;
        if (chunk < step) chunk = step;
        chunk = (chunk + v.gran - 1) / v.gran * v.gran;
        if (v.mapped + chunk > v.reserved) chunk = v.reserved - v.mapped;
        CUmemGenericAllocationHandle h;
        if (v.Create(&h, chunk, &v.prop, 0) != CUDA_SUCCESS) return false;
        if (v.Map(v.va + v.mapped, chunk, 0, h, 0) != CUDA_SUCCESS) { v.Release(h); return false; }
        CUmemAccessDesc acc;
        memset(&acc, 0, sizeof(acc));
        acc.location = v.prop.location;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (v.SetAccess(v.va + v.mapped, chunk, &acc, 1) != CUDA_SUCCESS) { v.Unmap(v.va + v.mapped, chunk); v.Release(h); return false; }
        v.chunks.push_back({h, chunk});
        v.mapped += chunk;
    }
    return true;
}
static void vmm_release_all() {
    VmmArena& v = g_vmm;
    size_t off = 0;
    for (auto& c : v.chunks) { v.Unmap(v.va + off, c.second); v.Release(c.first); off += c.second; }
    v.chunks.clear();
    v.mapped = 0;
}

// Entry points that use scratch are serialised per process (= per device: one process per GPU): the reference runs one query at a
// time on its engine thread, but triggers may call c(cxt) from pool threads (server/monetdb_conn.cpp:379-443).  The lock is
// recursive because entry points nest (aqgpu_string_dict -> aqgpu_groupby_build, ...) and is held for the scope's lifetime.
static std::recursive_mutex g_api_mu;
ArenaScope::ArenaScope() { g_api_mu.lock(); ++g_arena.depth; }
ArenaScope::~ArenaScope() {
    struct Unlock { ~Unlock() { g_api_mu.unlock(); } } unlock_at_exit;      // runs on every path out of this destructor
    Arena& a = g_arena;
    if (--a.depth > 0) return;
    for (void* q : a.spilled) cudaFreeAsync(q, ctx().stream);
    a.spilled.clear();
    if (a.need > a.cap && !g_vmm.ok) {   // no virtual memory management: grow to what this call needed through the pool
        if (a.base) cudaFreeAsync(a.base, ctx().stream);
        a.base = nullptr; a.cap = 0;
        const size_t want = a.need + a.need / 16 + (size_t(1) << 20);
        void* q = nullptr;
        if (cudaMallocAsync(&q, want, ctx().stream) == cudaSuccess) { a.base = static_cast<char*>(q); a.cap = want; }
        else cudaGetLastError();
    }
    a.off = 0; a.overflow = 0; a.need = 0;
}

int DevBuf::alloc(size_t n) {
    release();
    if (n == 0) n = 16;
    Arena& a = g_arena;
    if (!persistent && a.depth > 0) {
        const size_t m = (n + ARENA_ALIGN - 1) & ~(ARENA_ALIGN - 1);
        if (a.off + m <= a.cap) {
            p = a.base + a.off; a.off += m; bytes = n; from_arena = true;
            if (a.off + a.overflow > a.need) a.need = a.off + a.overflow;
            return 0;
        }
        // does not fit: back more of the reservation (in place, at least a quarter more than before)
        if ((a.base == nullptr || g_vmm.ok) && vmm_grow(std::max(a.off + m, a.cap + a.cap / 4))) {
            a.base = reinterpret_cast<char*>(g_vmm.va); a.cap = g_vmm.mapped;
            p = a.base + a.off; a.off += m; bytes = n; from_arena = true;
            if (a.off + a.overflow > a.need) a.need = a.off + a.overflow;
            return 0;
        }
        // no more memory to map (or no VMM): serve from the pool now
        cudaError_t e = cudaMallocAsync(&p, n, ctx().stream);
        if (e != cudaSuccess) { p = nullptr; return fail("cudaMallocAsync(%zu bytes): %s", n, cudaGetErrorString(e)); }
        a.spilled.push_back(p);
        a.overflow += m;
        if (a.off + a.overflow > a.need) a.need = a.off + a.overflow;
        bytes = n; from_arena = true;    // owned by the scope
        return 0;
    }
    cudaError_t e = cudaMallocAsync(&p, n, ctx().stream);
    if (e != cudaSuccess) {
        p = nullptr;
        return fail("cudaMallocAsync(%zu bytes): %s", n, cudaGetErrorString(e));
    }
    bytes = n; from_arena = false;
    return 0;
}

int DevBuf::alloc_zero(size_t n) {
    AQ_TRY(alloc(n));
    AQ_CUDA(cudaMemsetAsync(p, 0, bytes, ctx().stream));
    return 0;
}

void DevBuf::release() {
    if (p) {
        if (from_arena) {                // bump allocator: only the top block can be handed back
            Arena& a = g_arena;
            const size_t m = (bytes + ARENA_ALIGN - 1) & ~(ARENA_ALIGN - 1);
            if (a.base && static_cast<char*>(p) + m == a.base + a.off) a.off -= m;
        } else {
            cudaFreeAsync(p, ctx().stream);
        }
    }
    p = nullptr;
    bytes = 0;
    from_arena = false;
}

// ---- per-kernel event timing ---------------------------------------------------------------------
struct ProfRec { std::string name; cudaEvent_t e0, e1; };
static std::vector<ProfRec> g_prof;
static std::vector<cudaEvent_t> g_ev_pool;
static std::string g_prof_tag;

static cudaEvent_t prof_event() {
    if (!g_ev_pool.empty()) { cudaEvent_t e = g_ev_pool.back(); g_ev_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
void profile_pre(const char* kernel_name) {
    ProfRec r;
    r.name = g_prof_tag.empty() ? std::string(kernel_name) : g_prof_tag + ":" + kernel_name;
    r.e0 = prof_event(); r.e1 = prof_event();
    cudaEventRecord(r.e0, ctx().stream);
    g_prof.push_back(r);
}
void profile_post() {
    if (!g_prof.empty()) cudaEventRecord(g_prof.back().e1, ctx().stream);
}

bool is_device_ptr(const void* p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

int InCol::bind(const void* p, size_t bytes) {
    if (!p) return fail("null input column");
    if (is_device_ptr(p)) { d = p; return 0; }
    AQ_TRY(tmp.alloc(bytes));
    AQ_CUDA(cudaMemcpyAsync(tmp.p, p, bytes, cudaMemcpyHostToDevice, ctx().stream));
    d = tmp.p;
    return 0;
}

int OutCol::bind(void* p, size_t n) {
    if (!p) return fail("null output column");
    bytes = n;
    if (is_device_ptr(p)) { d = p; host = nullptr; return 0; }
    AQ_TRY(tmp.alloc(n));
    d = tmp.p;
    host = p;
    return 0;
}

int OutCol::finish(size_t used) {
    if (!host || used == 0) return 0;
    if (used > bytes) used = bytes;
    AQ_CUDA(cudaMemcpyAsync(host, d, used, cudaMemcpyDeviceToHost, ctx().stream));
    return 0;
}

} // namespace aq

using namespace aq;

extern "C" {

int aqgpu_abi_version(void) { return AQGPU_ABI_VERSION; }

const char* aqgpu_last_error(void) { return g_err; }

int aqgpu_init(int device) {
    std::lock_guard<std::mutex> lk(g_init_mu);
    Context& c = g_ctx;
    if (c.ready && (device < 0 || device == c.device)) return 0;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail("libaqgpu needs a CUDA device (sm_100a); none is visible (%s). There is no CPU fallback.",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0) {
        if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    }
    if (device >= n) return fail("device %d out of range (%d visible)", device, n);
    AQ_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    AQ_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail("libaqgpu is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
    c.device = device;
    c.sm_count = prop.multiProcessorCount;
    c.smem_optin = prop.sharedMemPerBlockOptin;
    if (!c.own_stream) AQ_CUDA(cudaStreamCreateWithFlags(&c.own_stream, cudaStreamNonBlocking));
    c.stream = c.own_stream;
    if (!c.ev0) AQ_CUDA(cudaEventCreate(&c.ev0));
    if (!c.ev1) AQ_CUDA(cudaEventCreate(&c.ev1));
    cudaMemPool_t pool;
    AQ_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t thr = UINT64_MAX;
    AQ_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    if (!c.d_flags) AQ_CUDA(cudaMalloc(&c.d_flags, 64 * sizeof(uint32_t)));
    if (!c.h_flags) AQ_CUDA(cudaHostAlloc(&c.h_flags, 64 * sizeof(uint32_t), cudaHostAllocDefault));
    c.ready = true;
    return 0;
}

int aqgpu_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_init_mu);
    Context& c = g_ctx;
    if (!c.ready) return 0;
    cudaStreamSynchronize(c.stream);
    if (c.d_flags) cudaFree(c.d_flags);
    if (c.h_flags) cudaFreeHost(c.h_flags);
    if (c.ev0) cudaEventDestroy(c.ev0);
    if (c.ev1) cudaEventDestroy(c.ev1);
    if (c.own_stream) cudaStreamDestroy(c.own_stream);
    c.d_flags = nullptr; c.h_flags = nullptr; c.ev0 = c.ev1 = nullptr;
    c.own_stream = c.stream = nullptr;
    c.ready = false;
    return 0;
}

int aqgpu_set_stream(void* s) {
    AQ_REQUIRE_INIT();
    g_ctx.stream = s ? static_cast<cudaStream_t>(s) : g_ctx.own_stream;
    return 0;
}

int aqgpu_sync(void) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaStreamSynchronize(g_ctx.stream));
    return 0;
}

uint64_t aqgpu_launch_count(void) { return g_ctx.launches.load(); }

int aqgpu_timer_begin(void) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaEventRecord(g_ctx.ev0, g_ctx.stream));
    return 0;
}

int aqgpu_timer_end(float* ms) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaEventRecord(g_ctx.ev1, g_ctx.stream));
    AQ_CUDA(cudaEventSynchronize(g_ctx.ev1));
    AQ_CUDA(cudaEventElapsedTime(ms, g_ctx.ev0, g_ctx.ev1));
    return 0;
}

int aqgpu_profile_enable(int on) {
    AQ_REQUIRE_INIT();
    g_ctx.profiling = on != 0;
    return 0;
}

int aqgpu_profile_tag(const char* tag) {
    g_prof_tag = tag ? tag : "";
    return 0;
}

/* Synchronises the stream and writes one line per (tag:kernel): "name calls total_ms\n". */
int aqgpu_profile_report(char* buf, size_t cap) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaStreamSynchronize(g_ctx.stream));
    std::vector<std::string> names; std::vector<double> ms; std::vector<unsigned> calls;
    for (auto& r : g_prof) {
        float t = 0.f;
        cudaEventElapsedTime(&t, r.e0, r.e1);
        size_t i = 0;
        for (; i < names.size(); ++i) if (names[i] == r.name) break;
        if (i == names.size()) { names.push_back(r.name); ms.push_back(0.0); calls.push_back(0); }
        ms[i] += t; calls[i] += 1;
        g_ev_pool.push_back(r.e0); g_ev_pool.push_back(r.e1);
    }
    g_prof.clear();
    std::string out;
    char line[512];
    for (size_t i = 0; i < names.size(); ++i) {
        snprintf(line, sizeof(line), "%s\t%u\t%.6f\n", names[i].c_str(), calls[i], ms[i]);
        out += line;
    }
    if (buf && cap) {
        const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
        memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    return 0;
}

int aqgpu_col_alloc(size_t bytes, void** dptr) {
    AQ_REQUIRE_INIT();
    if (!dptr) return fail("null dptr");
    // stream-ordered pool (the one DevBuf uses): no device-wide synchronisation, no ms-scale cudaMalloc /
    // cudaFree per column -- the multi-GPU flows allocate and drop dozens of GB-sized columns per query
    AQ_CUDA(cudaMallocAsync(dptr, bytes ? bytes : 16, g_ctx.stream));
    return 0;
}

int aqgpu_col_free(void* dptr) {
    if (dptr) AQ_CUDA(cudaFreeAsync(dptr, g_ctx.stream));
    return 0;
}

int aqgpu_col_upload(void* dptr, const void* host, size_t bytes) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaMemcpyAsync(dptr, host, bytes, cudaMemcpyHostToDevice, g_ctx.stream));
    AQ_CUDA(cudaStreamSynchronize(g_ctx.stream));
    return 0;
}

int aqgpu_col_download(void* host, const void* dptr, size_t bytes) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaMemcpyAsync(host, dptr, bytes, cudaMemcpyDeviceToHost, g_ctx.stream));
    AQ_CUDA(cudaStreamSynchronize(g_ctx.stream));
    return 0;
}

int aqgpu_col_upload_async(void* dptr, const void* host, size_t bytes) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaMemcpyAsync(dptr, host, bytes, cudaMemcpyHostToDevice, g_ctx.stream));
    return 0;
}

int aqgpu_col_download_async(void* host, const void* dptr, size_t bytes) {
    AQ_REQUIRE_INIT();
    AQ_CUDA(cudaMemcpyAsync(host, dptr, bytes, cudaMemcpyDeviceToHost, g_ctx.stream));
    return 0;
}

// ---- ingest / write-back on their own copy streams ------------------------------------------------
// The reference hands the post-processor pointers into the data source's buffers (server/monetdb_conn.cpp:203-224, the 'C'
// cache message server/server.cpp:552-583) and appends result tables back (server/table_ext_monetdb.hpp:35-86); here the two
// directions of the host link run on dedicated streams, so that PCIe carries columns in and results out at the same time
// while the library stream computes.  Ordering is by events, never by the host:
//   ingest_async     the copy starts after everything ALREADY enqueued on the library stream (queries still reading the
//                    column's previous contents) and runs on the ingest stream
//   ingest_fence     work enqueued on the library stream from now on waits for every ingest enqueued so far
//   writeback_async  the copy starts after everything already enqueued on the library stream (the query that produced the
//                    column) and runs on the write-back stream
//   writeback_fence  work enqueued on the library stream from now on waits for every write-back enqueued so far (call it
//                    before a query overwrites result buffers whose write-back may still be in flight)
//   writeback_wait   blocks the host until every write-back has landed in host memory
static cudaStream_t g_ingest_stream = nullptr, g_writeback_stream = nullptr;
static int copy_streams() {
    if (!g_ingest_stream) AQ_CUDA(cudaStreamCreateWithFlags(&g_ingest_stream, cudaStreamNonBlocking));
    if (!g_writeback_stream) AQ_CUDA(cudaStreamCreateWithFlags(&g_writeback_stream, cudaStreamNonBlocking));
    return 0;
}
static int order_after(cudaStream_t waiter, cudaStream_t producer) {      // waiter's next work starts after producer's current tail
    cudaEvent_t e;
    AQ_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    cudaError_t rc = cudaEventRecord(e, producer);
    if (rc == cudaSuccess) rc = cudaStreamWaitEvent(waiter, e, 0);
    cudaEventDestroy(e);                                                  // released once the wait has been satisfied
    AQ_CUDA(rc);
    return 0;
}
// Measured on the B200 box (profiles/r02/e2e_pipeline_trace.txt): copies of ONE direction execute in submission order ACROSS
// streams -- a query's 100-byte flag read-back enqueued on the library stream after a 36 GB write-back completes only when the
// write-back does, whether the write-back is one copy or 16 MB pieces -- while the two directions do overlap (8 GiB each way:
// 55 / 52 GB/s alone, 45 + 45 GB/s together, profiles/microbench/pcie_duplex.py).  So callers enqueue the big write-backs LAST
// in a step (they then overlap the NEXT step's ingest), not first.
int aqgpu_ingest_async(void* dptr, const void* pinned_host, size_t bytes) {
    AQ_REQUIRE_INIT();
    if (!bytes) return 0;
    if (!dptr || !pinned_host) return fail("ingest_async: null argument");
    AQ_TRY(copy_streams());
    AQ_TRY(order_after(g_ingest_stream, g_ctx.stream));
    AQ_CUDA(cudaMemcpyAsync(dptr, pinned_host, bytes, cudaMemcpyHostToDevice, g_ingest_stream));
    return 0;
}
int aqgpu_ingest_fence(void) {
    AQ_REQUIRE_INIT();
    AQ_TRY(copy_streams());
    return order_after(g_ctx.stream, g_ingest_stream);
}
int aqgpu_writeback_async(void* pinned_host, const void* dptr, size_t bytes) {
    AQ_REQUIRE_INIT();
    if (!bytes) return 0;
    if (!dptr || !pinned_host) return fail("writeback_async: null argument");
    AQ_TRY(copy_streams());
    AQ_TRY(order_after(g_writeback_stream, g_ctx.stream));
    AQ_CUDA(cudaMemcpyAsync(pinned_host, dptr, bytes, cudaMemcpyDeviceToHost, g_writeback_stream));
    return 0;
}
int aqgpu_writeback_fence(void) {
    AQ_REQUIRE_INIT();
    AQ_TRY(copy_streams());
    return order_after(g_ctx.stream, g_writeback_stream);
}
int aqgpu_writeback_wait(void) {
    AQ_REQUIRE_INIT();
    if (g_writeback_stream) AQ_CUDA(cudaStreamSynchronize(g_writeback_stream));
    return 0;
}

int aqgpu_trim_pool(void) {
    AQ_REQUIRE_INIT();
    if (g_arena.depth == 0 && g_arena.base && !g_vmm.ok) { cudaFreeAsync(g_arena.base, g_ctx.stream); g_arena.base = nullptr; g_arena.cap = 0; g_arena.off = 0; }
    AQ_CUDA(cudaStreamSynchronize(g_ctx.stream));
    if (g_arena.depth == 0 && g_vmm.ok) { vmm_release_all(); g_arena.base = nullptr; g_arena.cap = 0; g_arena.off = 0; }
    cudaMemPool_t pool;
    AQ_CUDA(cudaDeviceGetDefaultMemPool(&pool, g_ctx.device));
    AQ_CUDA(cudaMemPoolTrimTo(pool, 0));
    return 0;
}

int aqgpu_host_alloc(size_t bytes, void** hptr) {
    AQ_REQUIRE_INIT();
    if (!hptr) return fail("null hptr");
    AQ_CUDA(cudaHostAlloc(hptr, bytes ? bytes : 16, cudaHostAllocDefault));
    return 0;
}

int aqgpu_ptr_is_device(const void* p) { return aq::is_device_ptr(p) ? 1 : 0; }

int aqgpu_host_free(void* hptr) {
    if (hptr) AQ_CUDA(cudaFreeHost(hptr));
    return 0;
}

size_t aqgpu_dtype_size(int dtype) { return dtype_size(dtype); }

int aqgpu_reduce_out_type(int op, int vtype) {
    if (op == AQ_COUNT) return AQ_UINT32;
    if (!is_val_type(vtype)) return -2;
    switch (op) {
    case AQ_SUM: case AQ_SUMSQ: case AQ_SUMXY: return is_int_type(vtype) ? AQ_INT128 : AQ_DOUBLE;     // types.h:205-210
    case AQ_AVG: case AQ_VAR: case AQ_STDDEV: case AQ_CORR: return AQ_DOUBLE;
    case AQ_MIN: case AQ_MAX: case AQ_FIRST: case AQ_LAST: case AQ_SECOND: return vtype;
    default: return -2;
    }
}

int aqgpu_scan_out_type(int op, int vtype) {
    if (!is_val_type(vtype)) return -2;
    switch (op) {
    case AQ_SUMS: case AQ_SUMW: return is_int_type(vtype) ? AQ_INT128 : AQ_DOUBLE;
    case AQ_AVGS: case AQ_AVGW: return AQ_DOUBLE;                       // GetFPType<GetLongType<T>>
    case AQ_VARS: case AQ_STDDEVS: case AQ_VARW: case AQ_STDDEVW: return AQ_DOUBLE;
    case AQ_RATIOW: return (vtype == AQ_INT32 || vtype == AQ_FLOAT) ? AQ_FLOAT : AQ_DOUBLE;  // types.h:199-204
    case AQ_MINS: case AQ_MAXS: case AQ_MINW: case AQ_MAXW:
    case AQ_DELTAS: case AQ_PREV: case AQ_NEXT: return vtype;
    default: return -2;
    }
}

} // extern "C"
