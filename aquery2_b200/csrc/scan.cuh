// scan.cuh -- per-group vector-valued aggregates (scan.cu).
#pragma once
#include "common.cuh"

namespace aq {

// start[p] = first CSR position of the group that owns position p.
int compute_segment_starts(const uint32_t* d_offsets, uint32_t ngroups, uint32_t n, uint32_t* d_start);
// out[p] = v[perm[p]] for 4- or 8-byte elements.
int gather_rows(const void* d_v, int elem_size, const uint32_t* d_perm, uint32_t n, void* d_out);
// All of aqgpu_scan_op on device-resident inputs; d_out has n cells of out_type.
int group_scan_device(int op, uint32_t w, int vtype, const void* d_v, uint32_t n, uint32_t ngroups,
                      const uint32_t* d_gid, const uint32_t* d_offsets, const uint32_t* d_perm,
                      const uint32_t* d_start, int out_type, void* d_out);

} // namespace aq
