// comm.cu -- one communicator per process for the multi-GPU group-by (dist.cu): NCCL over NVLink 5 / NVSwitch,
// bound at run time (dlopen of libnccl.so.2, the library torch itself has loaded or the system one), so
// libaqgpu.so has no link-time dependency and single-GPU users never touch it.
//
// The reference has no multi-process path at all; the contract is SURVEY.md section 8(e):
// row shards, local partial aggregation, exchange over NVLink, merge.  Entry points (include/aqgpu.h):
//   aqgpu_comm_unique_id   rank 0 creates the 128-byte NCCL id, the host program broadcasts it (any channel)
//   aqgpu_comm_init        every rank: ncclCommInitRank on the library's device
//   aqgpu_comm_destroy
// All collectives run on the library stream, byte-typed.
#include "comm.cuh"
#include <dlfcn.h>
#include <cstring>

namespace aq {

namespace {
typedef struct { char internal[128]; } NcclId;          // ncclUniqueId (nccl.h:37-38)
typedef void* NcclComm;
enum { NCCL_UINT8 = 1, NCCL_INT64 = 4 };                 // ncclDataType_t (nccl.h:276-290)
enum { NCCL_SUM = 0, NCCL_MAX = 2, NCCL_MIN = 3 };       // ncclRedOp_t (nccl.h:260-263)

struct Nccl {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
Nccl g_nccl;
NcclComm g_comm = nullptr;
int g_nranks = 1, g_rank = 0;

int load_nccl() {
    if (g_nccl.lib) return 0;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);     // the copy the process already has (torch's)
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail("multi-GPU: cannot load libnccl.so.2 (%s)", dlerror());
#define AQ_SYM(field, name)                                                                         \
    *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);                                      \
    if (!g_nccl.field) return fail("multi-GPU: libnccl has no symbol %s", name);
    AQ_SYM(GetUniqueId, "ncclGetUniqueId") AQ_SYM(CommInitRank, "ncclCommInitRank") AQ_SYM(CommDestroy, "ncclCommDestroy")
    AQ_SYM(AllGather, "ncclAllGather") AQ_SYM(AllReduce, "ncclAllReduce") AQ_SYM(Send, "ncclSend") AQ_SYM(Recv, "ncclRecv")
    AQ_SYM(GroupStart, "ncclGroupStart") AQ_SYM(GroupEnd, "ncclGroupEnd") AQ_SYM(GetErrorString, "ncclGetErrorString")
#undef AQ_SYM
    g_nccl.lib = h;
    return 0;
}
} // namespace

#define AQ_NCCL(expr)                                                                               \
    do {                                                                                            \
        const int _r = (expr);                                                                      \
        if (_r != 0) return ::aq::fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, g_nccl.GetErrorString(_r)); \
    } while (0)

int comm_nranks() { return g_comm ? g_nranks : 1; }
int comm_rank() { return g_comm ? g_rank : 0; }
bool comm_ready() { return g_comm != nullptr; }

int comm_allgather(const void* d_send, void* d_recv, size_t bytes_per_rank) {
    if (!g_comm) return fail("multi-GPU: aqgpu_comm_init has not been called");
    AQ_NCCL(g_nccl.AllGather(d_send, d_recv, bytes_per_rank, NCCL_UINT8, g_comm, ctx().stream));
    return 0;
}

// rank r sends send_bytes[d] bytes at d_send + send_off[d] to every rank d and receives recv_bytes[s] bytes from rank s
// at d_recv + recv_off[s]; the own block goes through a device copy.  One grouped call: all transfers run concurrently.
int comm_alltoallv(const void* d_send, const size_t* send_off, const size_t* send_bytes,
                   void* d_recv, const size_t* recv_off, const size_t* recv_bytes) {
    if (!g_comm) return fail("multi-GPU: aqgpu_comm_init has not been called");
    const char* s = static_cast<const char*>(d_send);
    char* r = static_cast<char*>(d_recv);
    if (send_bytes[g_rank] != recv_bytes[g_rank]) return fail("multi-GPU: own block sizes differ");
    if (send_bytes[g_rank])
        AQ_CUDA(cudaMemcpyAsync(r + recv_off[g_rank], s + send_off[g_rank], send_bytes[g_rank], cudaMemcpyDeviceToDevice, ctx().stream));
    AQ_NCCL(g_nccl.GroupStart());
    for (int k = 1; k < g_nranks; ++k) {
        const int to = (g_rank + k) % g_nranks, from = (g_rank - k + g_nranks) % g_nranks;
        if (send_bytes[to]) AQ_NCCL(g_nccl.Send(s + send_off[to], send_bytes[to], NCCL_UINT8, to, g_comm, ctx().stream));
        if (recv_bytes[from]) AQ_NCCL(g_nccl.Recv(r + recv_off[from], recv_bytes[from], NCCL_UINT8, from, g_comm, ctx().stream));
    }
    AQ_NCCL(g_nccl.GroupEnd());
    return 0;
}

// element-wise sum of n uint32 values over all ranks, in place on the device
int comm_allreduce_sum_u32(uint32_t* d_vals, size_t n) {
    if (!g_comm) return fail("multi-GPU: aqgpu_comm_init has not been called");
    AQ_NCCL(g_nccl.AllReduce(d_vals, d_vals, n, 3 /* ncclUint32 */, NCCL_SUM, g_comm, ctx().stream));
    return 0;
}

// element-wise max of n int64 values over all ranks, result on the host (synchronises the stream)
int comm_allreduce_max_i64(long long* h_vals, int n) {
    if (!g_comm) return 0;
    DevBuf d;
    AQ_TRY(d.alloc(size_t(n) * 8));
    AQ_CUDA(cudaMemcpyAsync(d.p, h_vals, size_t(n) * 8, cudaMemcpyHostToDevice, ctx().stream));
    AQ_NCCL(g_nccl.AllReduce(d.p, d.p, (size_t)n, NCCL_INT64, NCCL_MAX, g_comm, ctx().stream));
    AQ_CUDA(cudaMemcpyAsync(h_vals, d.p, size_t(n) * 8, cudaMemcpyDeviceToHost, ctx().stream));
    AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

} // namespace aq

using namespace aq;

extern "C" int aqgpu_comm_unique_id(void* id128) {
    if (!id128) return fail("null id buffer");
    AQ_TRY(load_nccl());
    NcclId id;
    AQ_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return 0;
}

extern "C" int aqgpu_comm_init(int nranks, int rank, const void* id128) {
    AQ_REQUIRE_INIT();
    if (nranks < 1 || rank < 0 || rank >= nranks || !id128) return fail("aqgpu_comm_init: bad arguments (nranks %d, rank %d)", nranks, rank);
    if (g_comm) return fail("aqgpu_comm_init: a communicator already exists");
    AQ_TRY(load_nccl());
    NcclId id;
    memcpy(&id, id128, sizeof(id));
    AQ_CUDA(cudaSetDevice(ctx().device));
    NcclComm c = nullptr;
    AQ_NCCL(g_nccl.CommInitRank(&c, nranks, id, rank));
    g_comm = c; g_nranks = nranks; g_rank = rank;
    return 0;
}

extern "C" int aqgpu_comm_destroy(void) {
    if (g_comm) {
        cudaStreamSynchronize(ctx().stream);
        g_nccl.CommDestroy(g_comm);
        g_comm = nullptr; g_nranks = 1; g_rank = 0;
    }
    return 0;
}

extern "C" int aqgpu_comm_size(int* nranks, int* rank) {
    if (nranks) *nranks = comm_nranks();
    if (rank) *rank = comm_rank();
    return 0;
}
