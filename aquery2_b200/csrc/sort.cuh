// sort.cuh -- device sort / scan utilities (sort.cu).
#pragma once
#include "common.cuh"

namespace aq {

// out[i] = sum(in[0..i)); *d_total (device pointer, optional) = total.  in == out allowed.
int exclusive_scan_u32(const uint32_t* d_in, uint32_t* d_out, size_t n, uint32_t* d_total);

// Stable LSD radix sort of a row list by a 32-bit key fetched from `col` through the row id.
struct SortJob {
    uint32_t n = 0;
    const uint32_t* rows_in = nullptr;   // nullptr: implicit list 0..n-1 (reverse: n-1..0)
    int reverse = 0;
    const uint32_t* col = nullptr;       // key of a row = xform(col[row * col_stride])
    uint32_t col_stride = 1;             // 2 = one 32-bit half of a 64-bit column
    int key_bits = 32;                   // only the low key_bits of the transformed key are sorted
    uint32_t xor_mask = 0;               // 0x80000000 = signed ascending, 0x7fffffff = signed descending,
                                         // 0 / 0xffffffff = unsigned ascending / descending
    int float_flip = 0;                  // IEEE float column
    uint32_t* rows_out = nullptr;        // n row ids
};
int sort_rows_by_key(SortJob& job);

} // namespace aq
