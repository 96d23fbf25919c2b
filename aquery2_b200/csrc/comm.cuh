// comm.cuh -- the process-wide NCCL communicator of the multi-GPU paths (comm.cu) and the hook through which the
// single-GPU group-by paths (lc.cu, mc.cu) exchange their dense record tables (dist.cu).
#pragma once
#include "common.cuh"

namespace aq {

bool comm_ready();
int comm_nranks();
int comm_rank();
int comm_allgather(const void* d_send, void* d_recv, size_t bytes_per_rank);
int comm_alltoallv(const void* d_send, const size_t* send_off, const size_t* send_bytes,
                   void* d_recv, const size_t* recv_off, const size_t* recv_bytes);
int comm_allreduce_max_i64(long long* h_vals, int n);
int comm_allreduce_sum_u32(uint32_t* d_vals, size_t n);

// How one accumulator's u64 words in a dense group record combine across ranks.
enum RecKind : int {
    RK_ADD64 = 0, RK_ADD128 = 1, RK_MIN64 = 2, RK_MAX64 = 3,
    RK_UMAX64 = 4, RK_UMIN64 = 5,    // (row key << 32 | 32-bit value): the value travels with the largest / smallest row key
    RK_TOP2 = 6                       // two such words, best and second best: the pair of the two largest over all ranks
};
constexpr int REC_MAX_FIELDS = 40;
struct RecLayout {
    uint32_t gw;                 // u64 words per record
    int n;                       // fields
    int kind[REC_MAX_FIELDS];
    uint32_t word[REC_MAX_FIELDS];
    uint32_t cnt_word;           // a record with cnt_word == 0 is empty
};

// Set by aqgpu_groupby_run_dist while the query runs: the group-by paths then work on a row SHARD.
struct DistCtx {
    bool active = false;
    int nranks = 1, rank = 0;
    unsigned long long row_base = 0;      // global row id of local row 0
    unsigned long long rows_total = 0;
    uint32_t rows_min = 0, rows_max = 0;  // smallest / largest shard: path and layout decisions must agree on every rank
    bool replicated = false;              // out: every rank holds all groups (else: the groups whose slot range it owns)
    int by_first_row = 0;                 // out: 1 = a rank holds the groups whose first row lies in its shard (hcfp.cu)
    double exchange_ms = 0.0;
    size_t exchange_bytes = 0;
};
DistCtx& dist();

// flags[FLAG_ABORT] of every rank -> the same non-zero value on all ranks if any rank raised it (stream-ordered)
int dist_exchange_flags(uint32_t* flags);

// ---- dist.cu: record-table exchanges used by lc.cu / mc.cu ------------------------------------------------
// every rank ends up with the element-wise combination of all ranks' tables (G records) and first rows (global ids,
// 0xffffffff = none); *any_abort = some rank raised FLAG_ABORT (then nothing else is valid)
int dist_combine_replicated(unsigned long long* gtab, uint32_t* gfirst, uint32_t G, const RecLayout& L, uint32_t* flags, uint32_t* any_abort);
// rank d ends up with the combination of all ranks' records of the slots [d * Gd, min(G, (d + 1) * Gd)), in place at
// the start of that range of its own gtab / gfirst
int dist_combine_owned(unsigned long long* gtab, uint32_t* gfirst, uint32_t G, uint32_t Gd, const RecLayout& L, uint32_t* flags, uint32_t* any_abort);

} // namespace aq
