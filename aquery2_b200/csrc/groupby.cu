// groupby.cu -- hash group-by with fused scalar aggregation: planning, global-table path, ranking, emit.
//
// Replaces AQHashTable::hashtable_push (reference server/hasher.h:176-179), the observable order of
// ht_postproc (hasher.h:181-198), the gather col[vecs[g]] (server/table.h:184-189) and the scalar
// aggregates of server/aggregations.h in one pass over the key and value columns.
//
// Table: open addressing, linear probing, array-of-structs slots of `sw` 64-bit words (one 32-byte
// sector for up to 3 accumulators): word 0 = (tag << 32 | min row id), words 1.. = accumulators.
//   * tag = the single 32-bit key, or the bit-packed (k - lo) fields when per-column bounds are
//     known (the reference's PerfectHashTable packing, hasher.h:226-243): membership is decided by
//     word 0 alone;
//   * otherwise tag = high hash bits and a tag match is verified against the key columns at the
//     slot's representative row (any row of the group works: word 0 only ever moves to smaller
//     rows of the same group, via atomicMin).
// The low half therefore ends as the group's FIRST-APPEARANCE row.  Group ids are the rank of that
// row among all first rows -- a prefix-popcount over an nrows-bit bitmap, no sort of the groups --
// which is exactly the insertion index m_values gives (server/unordered_dense.h:1142-1146).
#include "groupby.cuh"
#include "comm.cuh"
#include "emit.cuh"
#include <algorithm>
#include <cstring>

namespace aq {

// ------------------------------------------------------------------------------------------------
// host-side planning
// ------------------------------------------------------------------------------------------------
int KeyCols::bind(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows) {
    if (!keys || !key_types || nkeys < 1 || nkeys > MAX_KEY_COLS) return fail("need 1..%d key columns", MAX_KEY_COLS);
    n = nkeys;
    for (int k = 0; k < nkeys; ++k) {
        type[k] = key_types[k];
        if (type[k] != AQ_INT32 && type[k] != AQ_UINT32 && type[k] != AQ_INT64)
            return fail("unsupported key dtype %d (int32, uint32, int64 keys only)", type[k]);
        AQ_TRY(col[k].bind(keys[k], dtype_size(type[k]) * nrows));
    }
    return 0;
}

int KeyCols::make_spec(const int64_t* key_min, const int64_t* key_max, KeySpec& ks) const {
    memset(&ks, 0, sizeof(ks));
    int w = 0;
    int vec_ok = 1;
    bool all32 = true;
    for (int k = 0; k < n; ++k) {
        const uint32_t* p = static_cast<const uint32_t*>(col[k].d);
        if (type[k] == AQ_INT32 || type[k] == AQ_UINT32) {
            if (w + 1 > MAX_KEY_WORDS) return fail("key tuple wider than %d words", MAX_KEY_WORDS);
            ks.base[w] = p; ks.stride[w] = 1; ++w;
            if (reinterpret_cast<uintptr_t>(p) & 15) vec_ok = 0;
        } else {
            if (w + 2 > MAX_KEY_WORDS) return fail("key tuple wider than %d words", MAX_KEY_WORDS);
            ks.base[w] = p; ks.stride[w] = 2; ++w;
            ks.base[w] = p + 1; ks.stride[w] = 2; ++w;
            all32 = false;
        }
    }
    ks.nwords = w;
    ks.vec_ok = vec_ok;
    ks.exact = (w == 1) ? 1 : 0;
    if (key_min && key_max && all32) {
        // PerfectHashTable-style packing (hasher.h:226-243): sum of field widths must fit 32 bits
        int bits = 0;
        bool ok = true;
        for (int k = 0; k < n && ok; ++k) {
            const int64_t lo = key_min[k], hi = key_max[k];
            if (hi < lo || hi - lo > 0xffffffffLL) { ok = false; break; }
            const uint64_t range = uint64_t(hi - lo);
            int b = 0;
            while (b < 32 && (range >> b) != 0) ++b;
            ks.pack_lo[k] = (uint32_t)lo;            // modular: (uint32)k - (uint32)lo == k - lo
            ks.pack_range[k] = (uint32_t)range;
            ks.pack_shift[k] = (uint32_t)bits;
            bits += b;
            if (bits > 32) ok = false;
        }
        if (ok) { ks.exact = 2; ks.pack_bits = bits; }
        else if (n <= MAX_KEY_WORDS) {
            // wide image: first-fit of whole columns into the 32-bit tag or the 64-bit second word
            uint32_t used[2] = {0, 0};
            const uint32_t room[2] = {32, 64};
            bool wide = true;
            for (int k = 0; k < n && wide; ++k) {
                const int64_t lo = key_min[k], hi = key_max[k];
                if (hi < lo || hi - lo > 0xffffffffLL) { wide = false; break; }
                const uint64_t range = uint64_t(hi - lo);
                uint32_t b = 0;
                while (b < 32 && (range >> b) != 0) ++b;
                const int part = used[0] + b <= room[0] ? 0 : (used[1] + b <= room[1] ? 1 : -1);
                if (part < 0) { wide = false; break; }
                ks.pack_lo[k] = (uint32_t)lo;
                ks.pack_range[k] = (uint32_t)range;
                ks.wide_part[k] = (uint32_t)part;
                ks.wide_shift[k] = used[part];
                used[part] += b;
            }
            // all-ones in the second word marks an empty slot: keep one bit of it clear
            ks.wide = (wide && used[1] < 64 && n <= WIDE_MAX_COLS) ? 1 : 0;
        }
    }
    return 0;
}

AggPlan::AggPlan() {
    memset(&prog, 0, sizeof(prog));
    memset(&emit, 0, sizeof(emit));
}

int AggPlan::add_acc(int kind, int xform, int vtype, const void* v, const void* v2) {
    for (int i = 0; i < n_acc; ++i) {           // share identical accumulators (e.g. COUNT)
        const AccOp& o = prog.op[i];
        if (o.kind == kind && o.xform == xform && o.vtype == vtype && o.v == v && o.v2 == v2) return i;
    }
    if (n_acc >= MAX_ACC) return -1;
    AccOp& o = prog.op[n_acc];
    o.kind = kind; o.xform = xform; o.vtype = vtype; o.v = v; o.v2 = v2;
    o.vec_ok = ((reinterpret_cast<uintptr_t>(v) & 15) == 0) && ((reinterpret_cast<uintptr_t>(v2) & 15) == 0);
    auto col_index = [&](const void* p) -> int {
        if (!p || prog.nv < 0) return -1;
        for (int j = 0; j < prog.nv; ++j) if (prog.vcol[j] == p) return j;
        if (prog.nv >= MAX_VCOL) { prog.nv = -1; return -1; }
        prog.vcol[prog.nv] = p; prog.vctype[prog.nv] = vtype;
        return prog.nv++;
    };
    o.ci = col_index(v); o.ci2 = col_index(v2);
    prog.n = ++n_acc;
    return n_acc - 1;
}

int AggPlan::find(int kind) const {
    for (int i = 0; i < n_acc; ++i) if (prog.op[i].kind == kind) return i;
    return -1;
}

int AggPlan::plan(const aqgpu_agg_t& a, const void* dv, const void* dv2, void* dout) {
    if (emit.n >= MAX_AGGS) return fail("more than %d aggregates in one call", MAX_AGGS);
    EmitAgg& e = emit.agg[emit.n];
    memset(&e, 0, sizeof(e));
    e.op = a.op; e.vtype = a.vtype; e.v = dv; e.out = dout;
    e.cnt = e.maxrow = e.top2 = -1;
    for (int i = 0; i < 5; ++i) e.a[i] = -1;
    const int def = aqgpu_reduce_out_type(a.op, a.vtype);
    if (def < 0) return fail("unsupported aggregate op %d on dtype %d", a.op, a.vtype);
    e.out_type = a.out_type ? a.out_type : def;
    if (e.out_type != def && !(def == AQ_INT128 && e.out_type == AQ_INT64) &&
        !(a.op == AQ_COUNT && e.out_type == AQ_INT64))
        return fail("aggregate op %d on dtype %d cannot produce dtype %d", a.op, a.vtype, e.out_type);
    const bool isint = is_int_type(a.vtype);
    const int add = isint ? ACC_ADD_I64 : ACC_ADD_F64;
    bool ok = true;
    switch (a.op) {
    case AQ_SUM: ok = (e.a[0] = add_acc(add, X_IDENT, a.vtype, dv, nullptr)) >= 0; break;
    case AQ_SUMSQ: ok = (e.a[0] = add_acc(add, X_SQUARE, a.vtype, dv, nullptr)) >= 0; break;
    case AQ_SUMXY:
        if (!dv2) return fail("sumxy needs a second column");
        ok = (e.a[0] = add_acc(add, X_PRODUCT, a.vtype, dv, dv2)) >= 0; break;
    case AQ_AVG: ok = (e.a[0] = add_acc(add, X_IDENT, a.vtype, dv, nullptr)) >= 0 && (e.cnt = add_count()) >= 0; break;
    case AQ_MIN: ok = (e.a[0] = add_acc(ACC_MIN_I64, X_IDENT, a.vtype, dv, nullptr)) >= 0; break;
    case AQ_MAX: ok = (e.a[0] = add_acc(ACC_MAX_I64, X_IDENT, a.vtype, dv, nullptr)) >= 0; break;
    case AQ_COUNT: ok = (e.cnt = add_count()) >= 0; break;
    case AQ_FIRST: case AQ_LAST: {
        // DESC row lists (reference): first = largest row id, last = first-appearance row.
        const bool need_max = (a.op == AQ_FIRST) == (emit.row_order == AQ_ROWS_DESC);
        if (need_max) ok = (e.maxrow = add_maxrow()) >= 0;
        break; }
    case AQ_SECOND: ok = (e.top2 = add_top2(emit.row_order)) >= 0; break;
    case AQ_VAR: case AQ_STDDEV:
        ok = (e.a[0] = add_acc(add, X_IDENT, a.vtype, dv, nullptr)) >= 0 &&
             (e.a[1] = add_acc(add, X_SQUARE, a.vtype, dv, nullptr)) >= 0 && (e.cnt = add_count()) >= 0;
        break;
    case AQ_CORR:
        if (!isint) return fail("corr: integer inputs only (the reference truncates fp inputs to __int128, aggregations.h:387-389)");
        if (!dv2) return fail("corr needs a second column");
        ok = (e.a[0] = add_acc(add, X_IDENT, a.vtype, dv, nullptr)) >= 0 &&
             (e.a[1] = add_acc(add, X_IDENT, a.vtype, dv2, nullptr)) >= 0 &&
             (e.a[2] = add_acc(add, X_PRODUCT, a.vtype, dv, dv2)) >= 0 &&
             (e.a[3] = add_acc(add, X_SQUARE, a.vtype, dv, nullptr)) >= 0 &&
             (e.a[4] = add_acc(add, X_SQUARE, a.vtype, dv2, nullptr)) >= 0 && (e.cnt = add_count()) >= 0;
        break;
    default: return fail("unknown aggregate op %d", a.op);
    }
    if (!ok) return fail("too many accumulators (max %d per call)", MAX_ACC);
    ++emit.n;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// device code
// ------------------------------------------------------------------------------------------------
struct InitPattern { unsigned long long w[16]; };

// slot word 0 = SLOT_EMPTY, accumulator words = their identities
__global__ void k_init_table(unsigned long long* __restrict__ tab, size_t nwords, uint32_t sw, const InitPattern pat) {
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nwords; i += size_t(gridDim.x) * blockDim.x)
        tab[i] = pat.w[i & (sw - 1)];
}

// One thread = 4 consecutive rows per iteration: 128-bit coalesced loads of every column.
template <bool EXACT>
__global__ void __launch_bounds__(256)
k_probe_agg(const KeySpec ks, const uint32_t nrows, const uint32_t mask, const uint32_t max_probe,
            const AccProg prog, uint32_t* __restrict__ row_slot, uint32_t* __restrict__ flags) {
    const uint32_t nquad = (nrows + 3u) >> 2;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nquad; q += gridDim.x * blockDim.x) {
        if (*reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT)) return;   // another CTA gave up
        const uint32_t row0 = q << 2;
        const int cnt = min(4u, nrows - row0);
        uint32_t tag[4], hsh[4], slot[4];
        if (!key_tags4(ks, row0, cnt, tag, hsh)) { atomicExch(flags + FLAG_ABORT, 3u); return; }
        bool bad = false;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            slot[r] = 0;
            if (r < cnt) {
                slot[r] = probe_insert<EXACT>(ks, prog.tab, prog.sw, mask, max_probe, hsh[r], tag[r], row0 + r);
                bad |= slot[r] == NOSLOT;
            }
        }
        if (bad) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); return; }
        if (row_slot) {
            if (cnt == 4) *reinterpret_cast<uint4*>(row_slot + row0) = make_uint4(slot[0], slot[1], slot[2], slot[3]);
            else for (int r = 0; r < cnt; ++r) row_slot[row0 + r] = slot[r];
        }
        for (int a = 0; a < prog.n; ++a) acc_apply(prog, a, slot, row0, cnt);
    }
}

// ---- wide exact keys: slot words (0, 1) = (tag << 32 | min row, second key word), accumulators from word 2
// one thread = 4 consecutive rows, every key column read once with a 128-bit load
__global__ void __launch_bounds__(256)
k_probe_wide(const KeySpec ks, const uint32_t nrows, const uint32_t mask, const uint32_t max_probe,
             const AccProg prog, uint32_t* __restrict__ row_slot, uint32_t* __restrict__ flags) {
    const uint32_t nquad = (nrows + 3u) >> 2;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nquad; q += gridDim.x * blockDim.x) {
        if (*reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT)) return;
        const uint32_t row0 = q << 2;
        const int cnt = min(4u, nrows - row0);
        uint32_t tag[4], slot[4];
        unsigned long long hi[4];
        const bool ok = wide_pack4(ks, row0, cnt, tag, hi);
        if (!ok) { atomicExch(flags + FLAG_ABORT, 3u); return; }
        // first probe of all four rows in flight together
        uint32_t s0[4];
        unsigned long long c0[4], c1[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            s0[r] = wide_hash(tag[r], hi[r]) & mask;
            if (r < cnt) ld128_volatile(prog.tab + size_t(s0[r]) * prog.sw, c0[r], c1[r]);
        }
        bool bad = false;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            slot[r] = 0;
            if (r >= cnt) continue;
            const uint32_t row = row0 + r;
            const unsigned long long mine = ((unsigned long long)tag[r] << 32) | row;
            uint32_t s = s0[r], found = NOSLOT;
            unsigned long long a = c0[r], b = c1[r];
            for (uint32_t step = 0; step <= max_probe; ++step) {
                unsigned long long* w0 = prog.tab + size_t(s) * prog.sw;
                if (step) ld128_volatile(w0, a, b);
                if (a == SLOT_EMPTY) {
                    unsigned long long o0, o1;
                    cas128(w0, SLOT_EMPTY, SLOT_EMPTY, mine, hi[r], o0, o1);
                    if (o0 == SLOT_EMPTY) { found = s; break; }
                    a = o0; b = o1;
                }
                if ((uint32_t)(a >> 32) == tag[r] && b == hi[r]) {
                    if (row < (uint32_t)a) atomicMin(w0, mine);
                    found = s; break;
                }
                s = (s + 1) & mask;
            }
            slot[r] = found;
            bad |= found == NOSLOT;
        }
        if (bad) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); return; }
        if (row_slot) {
            if (cnt == 4) *reinterpret_cast<uint4*>(row_slot + row0) = make_uint4(slot[0], slot[1], slot[2], slot[3]);
            else for (int r = 0; r < cnt; ++r) row_slot[row0 + r] = slot[r];
        }
        for (int a = 0; a < prog.n; ++a) {
            const AccOp& o = prog.op[a];
            long long xi[4] = {0, 0, 0, 0}; double xd[4] = {0, 0, 0, 0};
            if (acc_needs_value(o)) acc_inputs(o, row0, cnt, xi, xd);
#pragma unroll
            for (int r = 0; r < 4; ++r)
                if (r < cnt) acc_global_row(o, slot_word(prog, slot[r]) + 2 + a, row0 + r, xi[r], xd[r]);
        }
    }
}

// first-appearance rows -> bitmap
__global__ void k_mark(const unsigned long long* __restrict__ tab, uint32_t cap, uint32_t sw, uint32_t* __restrict__ bitmap) {
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < cap; s += gridDim.x * blockDim.x) {
        const unsigned long long w = tab[size_t(s) * sw];
        if (w != SLOT_EMPTY) {
            const uint32_t fr = (uint32_t)w;
            atomicOr(bitmap + (fr >> 5), 1u << (fr & 31));
        }
    }
}

// prefix-popcount over the bitmap, 3 kernels: block sums, scan of block sums, per-word prefixes.
constexpr int RANK_THREADS = 256;
constexpr int RANK_WORDS = RANK_THREADS * 4;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[RANK_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = lane < RANK_THREADS / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int d = 1; d < RANK_THREADS / 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += y;
        }
        if (lane < RANK_THREADS / 32) warp_sums[lane] = s;
    }
    __syncthreads();
    const uint32_t base = wid ? warp_sums[wid - 1] : 0;
    if (total) *total = warp_sums[RANK_THREADS / 32 - 1];
    __syncthreads();
    return base + x - v;
}

__global__ void __launch_bounds__(RANK_THREADS)
k_rank_blocksum(const uint32_t* __restrict__ bitmap, uint32_t nwords, uint32_t* __restrict__ blocksum) {
    const uint32_t w0 = blockIdx.x * RANK_WORDS + threadIdx.x * 4;
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) if (w0 + i < nwords) c += __popc(bitmap[w0 + i]);
    uint32_t total;
    block_exclusive_scan(c, &total);
    if (threadIdx.x == 0) blocksum[blockIdx.x] = total;
}

// single block; exclusive scan in place, total number of groups -> flags[FLAG_NGROUPS]
__global__ void __launch_bounds__(RANK_THREADS)
k_rank_scan_blocks(uint32_t* __restrict__ blocksum, uint32_t nblocks, uint32_t* __restrict__ flags) {
    uint32_t carry = 0;
    for (uint32_t b0 = 0; b0 < nblocks; b0 += RANK_THREADS) {
        const uint32_t i = b0 + threadIdx.x;
        const uint32_t v = i < nblocks ? blocksum[i] : 0;
        uint32_t total;
        const uint32_t ex = block_exclusive_scan(v, &total);
        if (i < nblocks) blocksum[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) flags[FLAG_NGROUPS] = carry;
}

__global__ void __launch_bounds__(RANK_THREADS)
k_rank_wordprefix(const uint32_t* __restrict__ bitmap, uint32_t nwords, const uint32_t* __restrict__ blocksum,
                  uint32_t* __restrict__ wordprefix) {
    const uint32_t w0 = blockIdx.x * RANK_WORDS + threadIdx.x * 4;
    uint32_t c[4];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { c[i] = (w0 + i < nwords) ? __popc(bitmap[w0 + i]) : 0; s += c[i]; }
    uint32_t ex = block_exclusive_scan(s, nullptr) + blocksum[blockIdx.x];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (w0 + i < nwords) wordprefix[w0 + i] = ex;
        ex += c[i];
    }
}

// Prefix popcount of a row bitmap (nwords words): wordprefix[w] = set bits before word w; the total goes to
// flags[FLAG_NGROUPS].  Shared by hash_aggregate and the range-partitioned path (mc.cu).
int bitmap_rank(const uint32_t* bitmap, uint32_t nwords, uint32_t* wordprefix) {
    const uint32_t rank_blocks = (nwords + RANK_WORDS - 1) / RANK_WORDS;
    DevBuf blocksum;
    AQ_TRY(blocksum.alloc(size_t(rank_blocks) * 4));
    AQ_LAUNCH(k_rank_blocksum, rank_blocks, RANK_THREADS, 0, bitmap, nwords, blocksum.as<uint32_t>());
    AQ_LAUNCH(k_rank_scan_blocks, 1, RANK_THREADS, 0, blocksum.as<uint32_t>(), rank_blocks, ctx().d_flags);
    AQ_LAUNCH(k_rank_wordprefix, rank_blocks, RANK_THREADS, 0, bitmap, nwords, blocksum.as<uint32_t>(), wordprefix);
    return 0;
}

// Emit in ROW order: a warp walks 32 bitmap words (1024 rows) at a time; lane b of the warp owns row
// 32*word + b.  A set bit = the first row of a group, whose gid is the rank of that bit.  The key
// columns are read where they lie (coalesced for dense bitmaps), the group's slot is found again by
// hashing the row's key (word 0 of the slot == tag << 32 | row exactly), and every output column is
// written at consecutive gids by consecutive lanes.  One random 32-byte sector read per group.
__global__ void __launch_bounds__(256)
k_emit_rows(const KeySpec ks, const unsigned long long* __restrict__ tab, const uint32_t mask, const uint32_t sw,
            const uint32_t* __restrict__ bitmap, const uint32_t* __restrict__ wordprefix, const uint32_t nwords,
            const KeyOut ko, uint32_t* __restrict__ first_row_out, uint32_t* __restrict__ counts_out,
            uint32_t* __restrict__ last_row_out, uint32_t* __restrict__ slot_gid, const EmitProg ep,
            const int cnt_acc, const int maxrow_acc, const uint32_t max_groups, const uint32_t* __restrict__ flags) {
    if (flags[FLAG_ABORT] != 0 || flags[FLAG_NGROUPS] > max_groups) return;   // host reports the error
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nchunks = (nwords + 31u) >> 5;
    for (uint32_t chunk = gwarp; chunk < nchunks; chunk += nwarps) {
        const uint32_t wi = chunk * 32u + lane;
        const uint32_t word = wi < nwords ? bitmap[wi] : 0u;
        const uint32_t pfx = wi < nwords ? wordprefix[wi] : 0u;
        uint32_t m = __ballot_sync(0xffffffffu, word != 0u);
        while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1u;
            const uint32_t w = __shfl_sync(0xffffffffu, word, j);
            const uint32_t p = __shfl_sync(0xffffffffu, pfx, j);
            if (!((w >> lane) & 1u)) continue;
            const uint32_t fr = (chunk * 32u + j) * 32u + lane;
            const uint32_t gid = p + __popc(w & ((1u << lane) - 1u));
            uint32_t tag, hsh;
            key_tag1(ks, fr, tag, hsh);
            const unsigned long long want = ((unsigned long long)tag << 32) | fr;
            uint32_t s = hsh & mask;
            while (tab[size_t(s) * sw] != want) s = (s + 1u) & mask;
            const unsigned long long* sl = tab + size_t(s) * sw;
            if (slot_gid) slot_gid[s] = gid;
            if (first_row_out) first_row_out[gid] = fr;
            if (counts_out) counts_out[gid] = (uint32_t)sl[1 + cnt_acc];
            if (last_row_out) last_row_out[gid] = (uint32_t)sl[1 + maxrow_acc];
            for (int k = 0; k < ko.n; ++k) {
                if (!ko.dst[k]) continue;
                if (ko.size[k] == 4) static_cast<uint32_t*>(ko.dst[k])[gid] = static_cast<const uint32_t*>(ko.src[k])[fr];
                else static_cast<unsigned long long*>(ko.dst[k])[gid] = static_cast<const unsigned long long*>(ko.src[k])[fr];
            }
            for (int a = 0; a < ep.n; ++a) emit_one(ep.agg[a], sl, gid, fr, ep.row_order);
        }
    }
}

// ---- fused rank + emit (global-table path): ONE pass in row order ---------------------------------------
// row_slot[i] is row i's slot; row i is its group's first appearance iff the slot's min row == i.  A
// tile of EF_ROWS rows counts its first rows, gets the number of first rows before it by decoupled
// look-back over per-tile states (tiles are handed out by a ticket counter, so a tile's predecessors
// have always started), and writes keys / first rows / aggregates at gid = that rank: the insertion
// index of m_values in the reference (server/unordered_dense.h:1142-1146).  One random 32-byte sector
// read per ROW replaces k_mark + k_rank_* + the re-probe of k_emit_rows when groups ~ rows (h2o q10).
constexpr int EF_THREADS = 256;
constexpr int EF_ROWS = EF_THREADS * 4;
constexpr unsigned long long EF_AGG = 1ull << 62, EF_PREFIX = 2ull << 62, EF_VAL = (1ull << 62) - 1;

__global__ void __launch_bounds__(EF_THREADS)
k_emit_fused(const unsigned long long* __restrict__ tab, const uint32_t sw, const int acc_off,
             const uint32_t* __restrict__ row_slot, const uint32_t nrows, const uint32_t ntiles,
             unsigned long long* __restrict__ tile_state, uint32_t* __restrict__ ticket,
             const KeyOut ko, uint32_t* __restrict__ first_row_out, const EmitProg ep,
             const unsigned char* __restrict__ rflags, const AccProg prog,
             const uint32_t max_groups, uint32_t* __restrict__ flags) {
    __shared__ uint32_t s_tile, s_excl, warp_sums[EF_THREADS / 32];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (;;) {
        if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= ntiles) return;
        const uint32_t row0 = tile * EF_ROWS + threadIdx.x * 4;
        const int cnt = row0 < nrows ? (int)min(4u, nrows - row0) : 0;
        uint32_t sl[4] = {0, 0, 0, 0};
        if (!row_slot) { for (int r = 0; r < 4; ++r) sl[r] = row0 + r; }          // one record per row (groupby_hc.cu)
        else if (cnt == 4) { const int4 t = ld_stream_v4(row_slot + row0); sl[0] = t.x; sl[1] = t.y; sl[2] = t.z; sl[3] = t.w; }
        else for (int r = 0; r < cnt; ++r) sl[r] = row_slot[row0 + r];
        uint32_t fl = 0, single = 0;
        if (rflags) {
            // groupby_hc.cu: one byte per row -- 0 first row of a one-row group (cells rebuilt below from the
            // row's own values, nothing was written for it), 1 first row with a record, 2 not a first row
            uint32_t rf = 0x02020202u;
            if (cnt == 4) rf = *reinterpret_cast<const uint32_t*>(rflags + row0);
            else for (int r = 0; r < cnt; ++r) rf = (rf & ~(0xffu << (8 * r))) | ((uint32_t)rflags[row0 + r] << (8 * r));
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const uint32_t b = (rf >> (8 * r)) & 0xffu;
                if (r < cnt && b != 2u) { fl |= 1u << r; if (b == 0u) single |= 1u << r; }
            }
        } else {
            unsigned long long w0[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) w0[r] = r < cnt ? tab[size_t(sl[r]) * sw] : SLOT_EMPTY;
#pragma unroll
            for (int r = 0; r < 4; ++r) if (r < cnt && (uint32_t)w0[r] == row0 + r) fl |= 1u << r;
        }
        const uint32_t mine = __popc(fl);
        // block exclusive scan of `mine`
        uint32_t x = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) warp_sums[wid] = x;
        __syncthreads();
        if (wid == 0) {
            uint32_t v = lane < EF_THREADS / 32 ? warp_sums[lane] : 0u, sc = v;
#pragma unroll
            for (int d = 1; d < EF_THREADS / 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, sc, d); if (lane >= d) sc += y; }
            if (lane < EF_THREADS / 32) warp_sums[lane] = sc - v;            // exclusive per warp
            const uint32_t total = __shfl_sync(0xffffffffu, sc, EF_THREADS / 32 - 1);
            // ---- decoupled look-back
            volatile unsigned long long* st = tile_state;
            if (lane == 0) st[tile] = (tile == 0 ? EF_PREFIX : EF_AGG) | total;
            uint32_t excl = 0;
            if (tile > 0) {
                long long look = (long long)tile - 1;
                for (;;) {
                    const long long idx = look - lane;
                    unsigned long long v2 = EF_PREFIX;
                    if (idx >= 0) { do { v2 = st[idx]; } while ((v2 >> 62) == 0); }
                    const uint32_t pm = __ballot_sync(0xffffffffu, (v2 >> 62) == 2);
                    const int first = pm ? __ffs(pm) - 1 : 32;
                    uint32_t part = ((int)lane <= first) ? (uint32_t)(v2 & EF_VAL) : 0u;
#pragma unroll
                    for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
                    excl += part;
                    if (pm) break;
                    look -= 32;
                }
                if (lane == 0) st[tile] = EF_PREFIX | (unsigned long long)(excl + total);
            }
            if (lane == 0) {
                s_excl = excl;
                if (tile + 1 == ntiles) flags[FLAG_NGROUPS] = excl + total;
            }
        }
        __syncthreads();
        uint32_t gid = s_excl + warp_sums[wid] + x - mine;
        const uint32_t gid0 = gid;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (!((fl >> r) & 1u)) continue;
            const uint32_t g = gid++;
            if (g >= max_groups) continue;                                    // the host reports the overflow
            const uint32_t fr = row0 + r;
            const unsigned long long* slp = tab + size_t(sl[r]) * sw + (acc_off - 1);
            unsigned long long one[1 + MAX_ACC];
            if ((single >> r) & 1u) {
                for (int a = 0; a < prog.n; ++a) one[1 + a] = acc_single_row(prog.op[a], fr);
                slp = one;
            }
            if (first_row_out) first_row_out[g] = fr;
            for (int a = 0; a < ep.n; ++a) emit_one(ep.agg[a], slp, g, fr, ep.row_order);
        }
        // key columns.  4-byte keys (the usual case): ONE 128-bit load per column for the thread's four
        // rows, up to 8 columns in flight before the first store
        bool keys32 = cnt == 4 && fl != 0;
        for (int k = 0; k < ko.n; ++k) keys32 &= ko.size[k] == 4 && (reinterpret_cast<uintptr_t>(ko.src[k]) & 15) == 0;
        for (int k0 = 0; keys32 && k0 < ko.n; k0 += 8) {
            uint32_t kv[8][4];
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const int k = k0 + kk;
                if (k >= ko.n || !ko.dst[k]) continue;
                const int4 t = ld_stream_v4(static_cast<const uint32_t*>(ko.src[k]) + row0);
                kv[kk][0] = t.x; kv[kk][1] = t.y; kv[kk][2] = t.z; kv[kk][3] = t.w;
            }
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const int k = k0 + kk;
                if (k >= ko.n || !ko.dst[k]) continue;
                uint32_t g = gid0;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if (!((fl >> r) & 1u)) continue;
                    if (g < max_groups) static_cast<uint32_t*>(ko.dst[k])[g] = kv[kk][r];
                    ++g;
                }
            }
        }
        // general case: all loads of a batch of 4 columns in flight before the first store
        for (int k0 = 0; !keys32 && k0 < ko.n; k0 += 4) {
            unsigned long long kv[4][4];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int k = k0 + kk;
                if (k >= ko.n || !ko.dst[k]) continue;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if (!((fl >> r) & 1u)) continue;
                    kv[kk][r] = ko.size[k] == 4 ? (unsigned long long)static_cast<const uint32_t*>(ko.src[k])[row0 + r]
                                                : static_cast<const unsigned long long*>(ko.src[k])[row0 + r];
                }
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int k = k0 + kk;
                if (k >= ko.n || !ko.dst[k]) continue;
                uint32_t g = gid0;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if (!((fl >> r) & 1u)) continue;
                    if (g < max_groups) {
                        if (ko.size[k] == 4) static_cast<uint32_t*>(ko.dst[k])[g] = (uint32_t)kv[kk][r];
                        else static_cast<unsigned long long*>(ko.dst[k])[g] = kv[kk][r];
                    }
                    ++g;
                }
            }
        }
        __syncthreads();                                                      // s_tile / warp_sums reused
    }
}

// ------------------------------------------------------------------------------------------------
// host pipeline
// ------------------------------------------------------------------------------------------------
static uint32_t pow2_ceil(uint64_t x) {
    uint64_t p = 1024;
    while (p < x && p < (1ull << 31)) p <<= 1;
    return (uint32_t)p;
}

static unsigned long long acc_identity(int kind) {
    switch (kind) {
    case ACC_MIN_I64: return 0x7fffffffffffffffULL;
    case ACC_MAX_I64: return 0x8000000000000000ULL;
    default: return 0ull;
    }
}

int hash_aggregate(KeySpec ks, uint32_t nrows, uint32_t groups_hint, AggPlan& plan,
                   uint32_t* d_row_slot, HashResult& hr, uint32_t* ngroups, bool allow_fused) {
    Context& c = ctx();
    const uint32_t max_cap = pow2_ceil(2ull * (nrows ? nrows : 1));
    uint32_t cap = groups_hint ? pow2_ceil(2ull * groups_hint) : (1u << 16);
    if (cap > max_cap) cap = max_cap;
    if (d_row_slot) allow_fused = false;
    uint32_t sw = 1;
    const uint32_t nwords_bm = (nrows + 31) / 32 + 1;
    const uint32_t rank_blocks = (nwords_bm + RANK_WORDS - 1) / RANK_WORDS;
    DevBuf blocksum;
    AQ_TRY(hr.bitmap.alloc(size_t(nwords_bm) * 4));
    AQ_TRY(hr.wordprefix.alloc(size_t(nwords_bm) * 4));
    AQ_TRY(blocksum.alloc(size_t(rank_blocks) * 4));
    bool allow_smem = d_row_slot == nullptr && !path_disabled("smem"), allow_part = d_row_slot == nullptr && !path_disabled("part"),
         allow_hc = !path_disabled("hc");

    for (int attempt = 0; attempt < 16; ++attempt) {
        // wide exact keys (only the fused path knows the shifted slot layout; the shared-memory and
        // partitioned kernels never take more than 2 key words, so they cannot meet such a table)
        // ... and only when the second key word is free (the slot does not grow) or the partitioned path
        // takes the query, which needs no table at all: a 2^30-slot table must not double to 64 GB
        bool wide = allow_fused && ks.exact == 0 && ks.wide && ks.nwords > 2 && plan.n_acc + 2 <= 16;
        if (wide) {
            uint32_t sw1 = 1, sw2 = 1;
            while (sw1 < uint32_t(1 + plan.n_acc)) sw1 <<= 1;
            while (sw2 < uint32_t(2 + plan.n_acc)) sw2 <<= 1;
            if (sw2 != sw1 && !(allow_hc && hc_qualifies(ks, nrows, plan.prog, sw2, groups_hint))) wide = false;
        }
        const int acc_off = wide ? 2 : 1;
        for (sw = 1; sw < uint32_t(acc_off + plan.n_acc); sw <<= 1) {}
        hr.sw = sw; hr.acc_off = acc_off;
        InitPattern pat;
        for (int i = 0; i < 16; ++i) pat.w[i] = 0;
        pat.w[0] = SLOT_EMPTY;
        if (wide) pat.w[1] = SLOT_EMPTY;
        for (int a = 0; a < plan.n_acc; ++a) pat.w[acc_off + a] = acc_identity(plan.prog.op[a].kind);
        if (wide && allow_hc && nrows) {                       // partitioned, L2-resident tables, one record per row
            bool used_hc = false;
            AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 8 * sizeof(uint32_t), c.stream));
            AQ_TRY(launch_hc(ks, nrows, plan.prog, sw, pat.w, groups_hint, hr.tab, hr.rflags, &used_hc));
            if (used_hc) {
                AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
                AQ_CUDA(cudaStreamSynchronize(c.stream));
                const uint32_t st = c.h_flags[FLAG_ABORT];
                if (st == 0) { *ngroups = 0; hr.ks = ks; hr.fused = true; hr.cap = nrows; hr.row_slot.release(); return 0; }
                hr.rflags.release(); hr.tab.release();
                if (st == 3) ks.wide = 0; else allow_hc = false;
                continue;
            }
        }
        hr.cap = cap;
        const uint32_t max_probe = cap == max_cap ? cap : 256u;
        const size_t tab_words = size_t(cap) * sw;
        AQ_TRY(hr.tab.alloc(tab_words * 8));
        plan.prog.tab = hr.tab.as<unsigned long long>();
        plan.prog.sw = sw;
        AQ_LAUNCH(k_init_table, grid_for(tab_words, 256, 4), 256, 0, plan.prog.tab, tab_words, sw, pat);
        AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 8 * sizeof(uint32_t), c.stream));
        AQ_CUDA(cudaMemsetAsync(hr.bitmap.p, 0, size_t(nwords_bm) * 4, c.stream));

        bool used_smem = false, used_part = false;
        if (allow_smem && nrows) AQ_TRY(launch_smem_probe(ks, nrows, cap, max_probe, plan.prog, groups_hint, &used_smem));
        if (!used_smem && allow_part && nrows) AQ_TRY(launch_part_probe(ks, nrows, cap, max_probe, plan.prog, groups_hint, &used_part));
        const bool fused = allow_fused && !used_smem && !used_part && nrows;
        uint32_t* rs = d_row_slot;
        if (fused) { AQ_TRY(hr.row_slot.alloc(size_t(nrows) * 4)); rs = hr.row_slot.as<uint32_t>(); }
        if (!used_smem && !used_part && nrows) {
            const int grid = grid_for((nrows + 3) / 4, 256, 1, 16);
            if (wide) AQ_LAUNCH(k_probe_wide, grid, 256, 0, ks, nrows, cap - 1, max_probe, plan.prog, rs, c.d_flags);
            else if (ks.exact) AQ_LAUNCH(k_probe_agg<true>, grid, 256, 0, ks, nrows, cap - 1, max_probe, plan.prog, rs, c.d_flags);
            else AQ_LAUNCH(k_probe_agg<false>, grid, 256, 0, ks, nrows, cap - 1, max_probe, plan.prog, rs, c.d_flags);
        }
        if (!fused) {
        AQ_LAUNCH(k_mark, grid_for(cap, 256, 4), 256, 0, plan.prog.tab, cap, sw, hr.bitmap.as<uint32_t>());
        AQ_LAUNCH(k_rank_blocksum, rank_blocks, RANK_THREADS, 0, hr.bitmap.as<uint32_t>(), nwords_bm, blocksum.as<uint32_t>());
        AQ_LAUNCH(k_rank_scan_blocks, 1, RANK_THREADS, 0, blocksum.as<uint32_t>(), rank_blocks, c.d_flags);
        AQ_LAUNCH(k_rank_wordprefix, rank_blocks, RANK_THREADS, 0, hr.bitmap.as<uint32_t>(), nwords_bm,
                  blocksum.as<uint32_t>(), hr.wordprefix.as<uint32_t>());
        }
        AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        AQ_CUDA(cudaStreamSynchronize(c.stream));
        const uint32_t st = c.h_flags[FLAG_ABORT];
        if (st == 0) { *ngroups = fused ? 0u : c.h_flags[FLAG_NGROUPS]; hr.ks = ks; hr.fused = fused; return 0; }
        if (st == 2 || st == 4) { if (used_part) allow_part = false; else allow_smem = false; continue; }   // CTA-local overflow / inexact fp: global path
        if (st == 3) {                                              // a key outside the caller's bounds
            if (ks.exact != 2 && !wide) return fail("internal: bounds flag without packed keys");
            if (wide) ks.wide = 0; else { ks.exact = 0; ks.pack_bits = 0; }
            continue;
        }
        if (cap == max_cap) return fail("hash table overflow at maximum capacity (%u slots)", cap);
        const uint64_t next = uint64_t(cap) * 16;                   // grow and retry
        cap = next > max_cap ? max_cap : (uint32_t)next;
    }
    return fail("hash build did not converge");
}

int emit_groups(const KeySpec& ks, uint32_t nrows, const HashResult& hr, const KeyOut& ko, uint32_t* d_first_row,
                uint32_t* d_counts, uint32_t* d_last_row, uint32_t* d_slot_gid, const AggPlan& plan, uint32_t max_groups) {
    const int cnt_acc = plan.find(ACC_COUNT), maxrow_acc = plan.find(ACC_MAXROW);
    if (d_counts && cnt_acc < 0) return fail("internal: counts requested without a COUNT accumulator");
    if (d_last_row && maxrow_acc < 0) return fail("internal: last rows requested without a MAXROW accumulator");
    const uint32_t nwords = (nrows + 31) / 32;
    AQ_LAUNCH(k_emit_rows, grid_for((nwords + 31) / 32, 8, 1, 8), 256, 0, ks, hr.tab.as<unsigned long long>(), hr.cap - 1, hr.sw,
              hr.bitmap.as<uint32_t>(), hr.wordprefix.as<uint32_t>(), nwords, ko, d_first_row, d_counts, d_last_row,
              d_slot_gid, plan.emit, cnt_acc, maxrow_acc, max_groups, ctx().d_flags);
    return 0;
}

int emit_fused(const KeySpec& ks, uint32_t nrows, const HashResult& hr, const KeyOut& ko, uint32_t* d_first_row,
               const AggPlan& plan, uint32_t max_groups, uint32_t* ngroups) {
    (void)ks;
    Context& c = ctx();
    const uint32_t ntiles = (nrows + EF_ROWS - 1) / EF_ROWS;
    DevBuf state;
    AQ_TRY(state.alloc_zero(size_t(ntiles) * 8 + 8));
    unsigned long long* st = state.as<unsigned long long>();
    uint32_t* ticket = reinterpret_cast<uint32_t*>(st + ntiles);
    const int grid = (int)std::min<uint64_t>(ntiles, uint64_t(c.sm_count) * 8);
    AQ_LAUNCH(k_emit_fused, grid, EF_THREADS, 0, hr.tab.as<unsigned long long>(), hr.sw, hr.acc_off, hr.row_slot.as<uint32_t>(),
              nrows, ntiles, st, ticket, ko, d_first_row, plan.emit, hr.rflags.as<unsigned char>(), plan.prog, max_groups, c.d_flags);
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    *ngroups = c.h_flags[FLAG_NGROUPS];
    return 0;
}

} // namespace aq

using namespace aq;

// Result columns allocated by the library (aqgpu_groupby_run): the reference's order of events --
// group build first, then col.resize(g.size) (engine/ast.py:681-701).
struct aqgpu_result {
    uint32_t ngroups = 0;
    int replicated = 1;                   // multi-GPU: 1 = every rank holds all groups, 0 = the groups this rank owns
    int nkeys = 0, naggs = 0;
    int key_type[MAX_KEY_COLS];
    int agg_type[MAX_AGGS];
    DevBuf keys[MAX_KEY_COLS];
    DevBuf aggs[MAX_AGGS];
    DevBuf first_row;
    aqgpu_result() {                      // result columns outlive the call: never from the scratch arena
        for (auto& b : keys) b.persistent = true;
        for (auto& b : aggs) b.persistent = true;
        first_row.persistent = true;
    }
};

namespace aq {

int OutSink::prepare(uint32_t cap_cells, const AggPlan& plan) {
    cap = cap_cells ? cap_cells : 1;
    if (res) {
        for (int a = 0; a < n_aggs; ++a) {
            res->agg_type[a] = plan.emit.agg[a].out_type;
            AQ_TRY(res->aggs[a].alloc(dtype_size(res->agg_type[a]) * size_t(cap)));
            agg_dst[a] = res->aggs[a].p;
        }
        for (int k = 0; k < nkeys; ++k) {
            AQ_TRY(res->keys[k].alloc(dtype_size(key_types[k]) * size_t(cap)));
            key_dst[k] = res->keys[k].p;
        }
        fr_dst = nullptr;
        if (want_first_row) { AQ_TRY(res->first_row.alloc(size_t(4) * cap)); fr_dst = res->first_row.as<uint32_t>(); }
        return 0;
    }
    cap = max_groups;
    vout.resize(n_aggs); kout.resize(nkeys);
    for (int a = 0; a < n_aggs; ++a) {
        AQ_TRY(vout[a].bind(aggs[a].out, dtype_size(plan.emit.agg[a].out_type) * size_t(cap)));
        agg_dst[a] = vout[a].d;
    }
    for (int k = 0; k < nkeys; ++k) {
        key_dst[k] = nullptr;
        if (out_keys && out_keys[k]) {
            AQ_TRY(kout[k].bind(out_keys[k], dtype_size(key_types[k]) * size_t(cap)));
            key_dst[k] = kout[k].d;
        }
    }
    fr_dst = nullptr;
    if (out_first_row) { AQ_TRY(fr.bind(out_first_row, size_t(4) * cap)); fr_dst = static_cast<uint32_t*>(fr.d); }
    return 0;
}

int OutSink::finish(uint32_t ng, const AggPlan& plan) {
    if (res) { res->ngroups = ng; return 0; }
    bool any_host = false;
    for (int a = 0; a < n_aggs; ++a) { AQ_TRY(vout[a].finish(dtype_size(plan.emit.agg[a].out_type) * size_t(ng))); any_host |= vout[a].host != nullptr; }
    for (int k = 0; k < nkeys; ++k) if (key_dst[k]) { AQ_TRY(kout[k].finish(dtype_size(key_types[k]) * size_t(ng))); any_host |= kout[k].host != nullptr; }
    if (fr_dst) { AQ_TRY(fr.finish(size_t(4) * ng)); any_host |= fr.host != nullptr; }
    // host outputs are complete on return; device outputs are stream-ordered
    if (any_host) AQ_CUDA(cudaStreamSynchronize(ctx().stream));
    return 0;
}

// The whole fused group-by: bind inputs, plan, pick a path, emit into `sink`.
static int groupby_core(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                        const int64_t* key_min, const int64_t* key_max, int row_order, uint32_t groups_hint,
                        const aqgpu_agg_t* aggs, int n_aggs, OutSink& sink, uint32_t* ngroups, bool* declined = nullptr) {
    if (n_aggs < 0 || n_aggs > MAX_AGGS || (n_aggs && !aggs)) return fail("need 0..%d aggregates", MAX_AGGS);
    if (!ngroups) return fail("ngroups must not be null");
    if (row_order != AQ_ROWS_DESC && row_order != AQ_ROWS_ASC) return fail("bad row_order %d", row_order);
    *ngroups = 0;
    if (!keys || !key_types || nkeys < 1 || nkeys > MAX_KEY_COLS) return fail("need 1..%d key columns", MAX_KEY_COLS);
    sink.n_aggs = n_aggs; sink.nkeys = nkeys; sink.key_types = key_types;
    KeyCols kc;
    if (nrows == 0) {
        for (int k = 0; k < nkeys; ++k)
            if (key_types[k] != AQ_INT32 && key_types[k] != AQ_UINT32 && key_types[k] != AQ_INT64)
                return fail("unsupported key dtype %d (int32, uint32, int64 keys only)", key_types[k]);
        if (sink.res) {
            AggPlan plan0; plan0.emit.row_order = row_order;
            for (int a = 0; a < n_aggs; ++a) AQ_TRY(plan0.plan(aggs[a], aggs[a].v, aggs[a].v2, nullptr));
            AQ_TRY(sink.prepare(1, plan0));
            AQ_TRY(sink.finish(0, plan0));
        }
        return 0;
    }
    AQ_TRY(kc.bind(keys, key_types, nkeys, nrows));
    if (!sink.res && sink.max_groups == 0) sink.max_groups = nrows;

    AggPlan plan;
    plan.emit.row_order = row_order;
    std::vector<InCol> vin(n_aggs), vin2(n_aggs);
    for (int a = 0; a < n_aggs; ++a) {
        const aqgpu_agg_t& g = aggs[a];
        const void* dv = nullptr; const void* dv2 = nullptr;
        if (g.op != AQ_COUNT) {
            if (!is_val_type(g.vtype)) return fail("aggregate %d: unsupported value dtype %d", a, g.vtype);
            AQ_TRY(vin[a].bind(g.v, dtype_size(g.vtype) * nrows));
            dv = vin[a].d;
            if (g.v2) { AQ_TRY(vin2[a].bind(g.v2, dtype_size(g.vtype) * nrows)); dv2 = vin2[a].d; }
        }
        const int ot = g.out_type ? g.out_type : aqgpu_reduce_out_type(g.op, g.vtype);
        if (ot < 0 || dtype_size(ot) == 0) return fail("aggregate %d: bad output dtype", a);
        AQ_TRY(plan.plan(g, dv, dv2, nullptr));
    }
    auto set_outputs = [&](KeyOut& ko) {
        for (int a = 0; a < n_aggs; ++a) plan.emit.agg[a].out = sink.agg_dst[a];
        memset(&ko, 0, sizeof(ko));
        ko.n = nkeys;
        for (int k = 0; k < nkeys; ++k) {
            ko.src[k] = kc.col[k].d;
            ko.size[k] = (int)dtype_size(key_types[k]);
            ko.dst[k] = sink.key_dst[k];
        }
    };

    // ---- planned fast paths (plan.cu / lc.cu / hcfp.cu); each declines (used = false) or completes the query
    PlanHints ph;
    ph.groups_hint = groups_hint;
    {
        bool used = false;
        AQ_TRY(run_planned_paths(kc, nrows, key_min, key_max, groups_hint, plan, sink, ngroups, &used, &ph));
        if (used) {
            bool staged = false;
            for (int k = 0; k < kc.n; ++k) staged |= kc.col[k].tmp.p != nullptr;
            if (staged || std::any_of(vin.begin(), vin.end(), [](const InCol& c) { return c.tmp.p != nullptr; }))
                AQ_CUDA(cudaStreamSynchronize(ctx().stream));
            return 0;
        }
        if (dist().active) {                                     // row shards: only the record-table paths know how to merge
            if (declined) *declined = true;
            return 0;
        }
    }

    KeySpec ks;
    if (!(key_min && key_max) && ph.have_bounds) AQ_TRY(kc.make_spec(ph.kmin, ph.kmax, ks));
    else AQ_TRY(kc.make_spec(key_min, key_max, ks));
    HashResult hr;
    uint32_t ng = 0;
    AQ_TRY(hash_aggregate(ks, nrows, ph.groups_hint, plan, nullptr, hr, &ng, true));
    if (!sink.res && !hr.fused && ng > sink.max_groups) return fail("capacity: %u groups exceed max_groups=%u", ng, sink.max_groups);
    AQ_TRY(sink.prepare(hr.fused ? nrows : ng, plan));
    KeyOut ko;
    set_outputs(ko);
    if (hr.fused) {
        AQ_TRY(emit_fused(hr.ks, nrows, hr, ko, sink.fr_dst, plan, sink.cap, &ng));
        if (ng > sink.cap) return fail("capacity: %u groups exceed max_groups=%u", ng, sink.cap);
    } else {
        AQ_TRY(emit_groups(hr.ks, nrows, hr, ko, sink.fr_dst, nullptr, nullptr, nullptr, plan, sink.cap));
    }
    AQ_TRY(sink.finish(ng, plan));
    *ngroups = ng;
    bool staged = false;
    for (int k = 0; k < kc.n; ++k) staged |= kc.col[k].tmp.p != nullptr;
    if (staged || std::any_of(vin.begin(), vin.end(), [](const InCol& c) { return c.tmp.p != nullptr; }))
        AQ_CUDA(cudaStreamSynchronize(ctx().stream));      // staged host inputs are released on return
    return 0;
}

} // namespace aq

extern "C" int aqgpu_groupby_agg(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                                 const int64_t* key_min, const int64_t* key_max,
                                 int row_order, uint32_t groups_hint,
                                 const aqgpu_agg_t* aggs, int n_aggs, uint32_t max_groups,
                                 void* const* out_keys, uint32_t* out_first_row, uint32_t* ngroups) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    OutSink sink;
    sink.aggs = aggs; sink.out_keys = out_keys; sink.out_first_row = out_first_row; sink.max_groups = max_groups;
    return groupby_core(keys, key_types, nkeys, nrows, key_min, key_max, row_order, groups_hint, aggs, n_aggs, sink, ngroups);
}

extern "C" int aqgpu_groupby_run(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows,
                                 const int64_t* key_min, const int64_t* key_max,
                                 int row_order, uint32_t groups_hint,
                                 const aqgpu_agg_t* aggs, int n_aggs, int want_first_row, aqgpu_result_t** out) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!out) return fail("null result pointer");
    *out = nullptr;
    aqgpu_result* r = new aqgpu_result();
    struct Guard { aqgpu_result* r; ~Guard() { delete r; } } guard{r};
    r->nkeys = nkeys > 0 && nkeys <= MAX_KEY_COLS ? nkeys : 0;
    r->naggs = n_aggs > 0 && n_aggs <= MAX_AGGS ? n_aggs : 0;
    for (int k = 0; k < r->nkeys; ++k) r->key_type[k] = key_types ? key_types[k] : 0;
    OutSink sink;
    sink.res = r; sink.want_first_row = want_first_row != 0;
    uint32_t ng = 0;
    AQ_TRY(groupby_core(keys, key_types, nkeys, nrows, key_min, key_max, row_order, groups_hint, aggs, n_aggs, sink, &ng));
    r->ngroups = ng;
    guard.r = nullptr;
    *out = r;
    return 0;
}

extern "C" int aqgpu_groupby_run_dist(const void* const* keys, const int* key_types, int nkeys, uint32_t nrows_local,
                                      uint64_t row_base, uint64_t rows_total, int row_order,
                                      const aqgpu_agg_t* aggs, int n_aggs, int want_first_row, aqgpu_result_t** out, int* used) {
    AQ_REQUIRE_INIT();
    ArenaScope scope;
    if (!out || !used) return fail("null result pointer");
    *out = nullptr; *used = 0;
    if (!comm_ready()) return fail("aqgpu_groupby_run_dist: call aqgpu_comm_init first");
    if (rows_total >= 0xffffffffull || row_base + nrows_local > rows_total) return fail("aqgpu_groupby_run_dist: global row ids must fit 32 bits");
    DistCtx& d = dist();
    d = DistCtx();
    d.active = true; d.nranks = comm_nranks(); d.rank = comm_rank(); d.row_base = row_base; d.rows_total = rows_total;
    struct Off { ~Off() { dist().active = false; } } off;
    aqgpu_result* r = new aqgpu_result();
    struct Guard { aqgpu_result* r; ~Guard() { delete r; } } guard{r};
    r->nkeys = nkeys > 0 && nkeys <= MAX_KEY_COLS ? nkeys : 0;
    r->naggs = n_aggs > 0 && n_aggs <= MAX_AGGS ? n_aggs : 0;
    for (int k = 0; k < r->nkeys; ++k) r->key_type[k] = key_types ? key_types[k] : 0;
    OutSink sink;
    sink.res = r; sink.want_first_row = want_first_row != 0;
    uint32_t ng = 0;
    bool declined = false;
    AQ_TRY(groupby_core(keys, key_types, nkeys, nrows_local, nullptr, nullptr, row_order, 0, aggs, n_aggs, sink, &ng, &declined));
    if (declined) return 0;                                      // *used = 0: the caller takes another strategy
    r->ngroups = ng;
    r->replicated = d.replicated ? 1 : (d.by_first_row ? 2 : 0);
    guard.r = nullptr;
    *out = r; *used = 1;
    return 0;
}
extern "C" int aqgpu_result_replicated(const aqgpu_result_t* r, int* replicated) {
    if (!r || !replicated) return fail("null argument");
    *replicated = r->replicated;
    return 0;
}

extern "C" int aqgpu_result_size(const aqgpu_result_t* r, uint32_t* ngroups) {
    if (!r || !ngroups) return fail("null argument");
    *ngroups = r->ngroups;
    return 0;
}
extern "C" int aqgpu_result_key(const aqgpu_result_t* r, int k, void** dptr, int* dtype) {
    if (!r || !dptr || k < 0 || k >= r->nkeys) return fail("bad key column index %d", k);
    *dptr = r->keys[k].p;
    if (dtype) *dtype = r->key_type[k];
    return 0;
}
extern "C" int aqgpu_result_agg(const aqgpu_result_t* r, int a, void** dptr, int* dtype) {
    if (!r || !dptr || a < 0 || a >= r->naggs) return fail("bad aggregate index %d", a);
    *dptr = r->aggs[a].p;
    if (dtype) *dtype = r->agg_type[a];
    return 0;
}
extern "C" int aqgpu_result_first_row(const aqgpu_result_t* r, uint32_t** dptr) {
    if (!r || !dptr) return fail("null argument");
    *dptr = r->first_row.as<uint32_t>();
    return 0;
}
extern "C" int aqgpu_result_free(aqgpu_result_t* r) {
    delete r;
    return 0;
}
