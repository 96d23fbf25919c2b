// window.cuh -- fused per-group window / running aggregates (window.cu).
#pragma once
#include "common.cuh"

namespace aq {

constexpr int WIN_MAX_OPS = 8;
constexpr int WIN_MAX_COLS = 3;         // distinct 4-byte value columns one call can carry through the sort
constexpr uint32_t WIN_MAX_W = 64;      // longest window the fused kernel serves (its halo)

struct WinOp {
    int op;             // aqgpu_scan_op
    uint32_t w;
    int vtype;          // AQ_INT32 / AQ_FLOAT on the fused path
    const void* v;      // device, nrows cells, table row order
    int minus;          // the vector is v - scan(v)
    int reduce;         // -1: vector output (nrows cells, CSR order); AQ_MAX / AQ_MIN / AQ_SUM: one cell per group
    int out_type;
    void* out;          // device
};

struct WinGroups {
    uint32_t nrows, ngroups;
    int row_order;
    const uint32_t* gid;        // nrows
    const uint32_t* offsets;    // ngroups + 1
    const uint32_t* perm;       // nullptr: row lists in table order (row_order); else the handle's own order (order_by)
    // ASSUMING order evaluated inside the call (only with perm == nullptr): every group's rows are ordered by this 4-byte
    // column (ties: table order) -- it rides through the scatter as a tuple field and the groups are sorted in shared memory
    const uint32_t* order_key = nullptr;
    uint32_t order_xor = 0;     // key image = (float_flip(k)) ^ order_xor, ascending unsigned (sort.cuh's convention)
    int order_float = 0;
};

// *handled = false: the combination is not served here (the caller runs the per-op path).
int window_fused(const WinGroups& G, const WinOp* ops, int n_ops, bool* handled);

} // namespace aq
