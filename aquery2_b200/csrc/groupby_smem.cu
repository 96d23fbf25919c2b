// groupby_smem.cu -- shared-memory pre-aggregation in front of the global group table.
//
// For group-bys whose groups fit one SM's shared memory (h2o q1, q2, q4: 1e2..1e4 groups) every CTA
// keeps a private open-addressing (or directly indexed) table of 32-bit cells in shared memory and
// aggregates its rows there with native 32-bit shared atomics (measured on B200: they keep up with
// the HBM stream, profiles/microbench/mb_atomics.cu), then merges its <= cap groups into the global
// table of groupby.cu once, at the end.  Rows that find no room (table beyond its fill limit, or a key
// equal to the EMPTY marker) go straight to the global table, so the result never depends on the
// capacity guess.  Replaces the same reference loops as groupby.cu (server/hasher.h:176-198,
// server/table.h:184-189, server/aggregations.h scalar aggregates).
//
// Cells (SoA arrays of `cap` entries, so neighbouring lanes hit different banks):
//   key      u32 (exact 32-bit tag) | u64 (two 32-bit words / one int64 key) | none (direct index)
//   minrow   u32, first-appearance row of the group inside this CTA (atomicMin after a plain pre-check)
//   COUNT    u32
//   ADD_I64  u32 lo + u32 hi: lo += (u32)x with the carry (and x's high word) added to hi only when
//            non-zero, i.e. one shared atomic per row for small values
//   ADD_F64  u64, compare-and-swap loop (there is no native 64-bit shared add)
//   MIN/MAX  s32 native for int32 / float (order-preserving image), CAS loop on u64 otherwise
//   MAXROW   u32 atomicMax;  TOP2ROW u64 CAS loop
#include "groupby_smem.cuh"
#include <cstring>

namespace aq {


// x for accumulator o and row r of a quad whose value columns are in raw[]
__device__ __forceinline__ void decode_x(const AccOp& o, const uint4 (&raw)[MAX_VCOL], int r, long long& xi, double& xd) {
    xi = 0; xd = 0.0;
    if (!acc_needs_value(o)) return;
    const uint32_t u = lane4(pick_col(raw, o.ci), r);
    const uint32_t u2 = o.xform == X_PRODUCT ? lane4(pick_col(raw, o.ci2), r) : u;
    if (o.vtype == AQ_INT32) {
        xi = (long long)(int)u;
        if (o.xform != X_IDENT) xi *= (long long)(int)u2;
    } else {
        const float f = __uint_as_float(u);
        xd = o.xform == X_IDENT ? (double)f : (double)__fmul_rn(f, __uint_as_float(u2));   // product in T (aggregations.h:339-345)
    }
}

constexpr int QPI = 2;      // quads per thread per iteration: all their loads are issued before any use

// Key words: 1 or 2 stride-1 32-bit columns; value columns: <= MAX_VCOL 4-byte columns (prog.vcol).
template <int MODE>
__global__ void __launch_bounds__(SMEM_THREADS, 1)
k_smem_agg(const KeySpec ks, const uint32_t nrows, const uint32_t gmask, const uint32_t max_probe,
           const AccProg prog, const SmemLayout lay, uint32_t* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ uint32_t n_used[16];
    constexpr bool EXACT = MODE != SM_HASH64;
    const uint32_t cap = lay.cap, cap_all = lay.cap * lay.reps;
    uint32_t* const minrow = reinterpret_cast<uint32_t*>(sm + lay.off_minrow);
    uint32_t* const key32 = reinterpret_cast<uint32_t*>(sm + lay.off_key);
    unsigned long long* const key64 = reinterpret_cast<unsigned long long*>(sm + lay.off_key);

    // ---- init: keys EMPTY, minrow = none, accumulators = identities
    for (uint32_t s = threadIdx.x; s < cap_all; s += SMEM_THREADS) {
        minrow[s] = EMPTY32;
        if (MODE == SM_HASH32) key32[s] = EMPTY32;
        if (MODE == SM_HASH64) key64[s] = EMPTY64;
        for (int a = 0; a < prog.n; ++a) {
            unsigned char* c0 = sm + lay.off_cell[a];
            switch (lay.kind[a]) {
            case CK_ADD_I64: reinterpret_cast<uint32_t*>(c0)[s] = 0u; reinterpret_cast<uint32_t*>(c0)[s + cap_all] = 0u; break;
            case CK_ADD_F64: case CK_TOP2: reinterpret_cast<unsigned long long*>(c0)[s] = 0ull; break;
            case CK_MIN32: reinterpret_cast<int*>(c0)[s] = 0x7fffffff; break;
            case CK_MAX32: reinterpret_cast<int*>(c0)[s] = (int)0x80000000u; break;
            case CK_MIN64: reinterpret_cast<unsigned long long*>(c0)[s] = 0x7fffffffffffffffULL; break;
            case CK_MAX64: reinterpret_cast<unsigned long long*>(c0)[s] = 0x8000000000000000ULL; break;
            default: reinterpret_cast<uint32_t*>(c0)[s] = 0u; break;     // COUNT, MAXROW
            }
        }
    }
    if (threadIdx.x < 16) n_used[threadIdx.x] = 0;
    __syncthreads();

    const uint32_t rep = (threadIdx.x >> 5) & (lay.reps - 1u);
    const uint32_t rep_off = rep * cap;
    const uint32_t nquad = (nrows + 3u) >> 2;
    const uint32_t stride = gridDim.x * SMEM_THREADS;
    bool dead = false;
    for (uint32_t q0 = blockIdx.x * SMEM_THREADS + threadIdx.x; q0 < nquad && !dead; q0 += QPI * stride) {
        // ---- all loads of this iteration first
        uint4 kraw[QPI][2];
        uint4 raw[QPI][MAX_VCOL];
        int cnts[QPI];
#pragma unroll
        for (int i = 0; i < QPI; ++i) {
            const uint32_t q = q0 + i * stride;
            cnts[i] = q < nquad ? (int)min(4u, nrows - (q << 2)) : 0;
            const uint32_t row0 = q << 2;
#pragma unroll
            for (int w = 0; w < 2; ++w)
                kraw[i][w] = (w < ks.nwords && cnts[i]) ? load_quad(ks.base[w], row0, cnts[i], ks.vec_ok) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int j = 0; j < MAX_VCOL; ++j)
                raw[i][j] = (j < prog.nv && cnts[i]) ? load_quad(prog.vcol[j], row0, cnts[i], (reinterpret_cast<uintptr_t>(prog.vcol[j]) & 15) == 0)
                                                     : make_uint4(0, 0, 0, 0);
        }
        uint32_t sslot[QPI][4], gslot[QPI][4];
        bool full[QPI];
#pragma unroll
        for (int i = 0; i < QPI; ++i) {
            const int cnt = cnts[i];
            const uint32_t row0 = (q0 + i * stride) << 2;
            bool bad = false, oob = false;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                sslot[i][r] = NOSLOT; gslot[i][r] = NOSLOT;
                if (r >= cnt) continue;
                const uint32_t row = row0 + r;
                const uint32_t k0 = lane4(kraw[i][0], r), k1 = lane4(kraw[i][1], r);
                uint32_t tag = k0, s = NOSLOT;
                if (EXACT && ks.exact == 2) {
                    const uint32_t d0 = k0 - ks.pack_lo[0];
                    oob |= d0 > ks.pack_range[0];
                    tag = d0 << ks.pack_shift[0];
                    if (ks.nwords == 2) {
                        const uint32_t d1 = k1 - ks.pack_lo[1];
                        oob |= d1 > ks.pack_range[1];
                        tag |= d1 << ks.pack_shift[1];
                    }
                }
                if (MODE == SM_DIRECT) {
                    s = oob ? NOSLOT : rep_off + tag;
                } else if (MODE == SM_HASH32) {
                    if (tag != EMPTY32) {
                        uint32_t p = mix32(tag) & lay.mask;
                        for (int step = 0; step < 32; ++step) {
                            uint32_t cur = *reinterpret_cast<volatile uint32_t*>(key32 + rep_off + p);
                            if (cur == EMPTY32) {
                                if (*reinterpret_cast<volatile uint32_t*>(&n_used[rep]) >= lay.limit) break;
                                cur = atomicCAS(key32 + rep_off + p, EMPTY32, tag);
                                if (cur == EMPTY32) { atomicAdd(&n_used[rep], 1u); s = rep_off + p; break; }
                            }
                            if (cur == tag) { s = rep_off + p; break; }
                            p = (p + 1u) & lay.mask;
                        }
                    }
                } else {
                    const unsigned long long mine = (unsigned long long)k0 | ((unsigned long long)k1 << 32);
                    if (mine != EMPTY64) {
                        uint32_t p = (uint32_t)mix64(mine) & lay.mask;
                        for (int step = 0; step < 32; ++step) {
                            unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(key64 + rep_off + p);
                            if (cur == EMPTY64) {
                                if (*reinterpret_cast<volatile uint32_t*>(&n_used[rep]) >= lay.limit) break;
                                cur = atomicCAS(key64 + rep_off + p, EMPTY64, mine);
                                if (cur == EMPTY64) { atomicAdd(&n_used[rep], 1u); s = rep_off + p; break; }
                            }
                            if (cur == mine) { s = rep_off + p; break; }
                            p = (p + 1u) & lay.mask;
                        }
                    }
                }
                if (s != NOSLOT) {
                    sslot[i][r] = s;
                    if (row < *reinterpret_cast<volatile uint32_t*>(minrow + s)) atomicMin(minrow + s, row);
                } else if (!oob) {                                  // no room in shared memory: global table
                    uint32_t t = tag, h = mix32(tag);
                    if (MODE == SM_HASH64) key_tag1(ks, row, t, h);
                    gslot[i][r] = probe_insert<EXACT>(ks, prog.tab, prog.sw, gmask, max_probe, h, t, row);
                    bad |= gslot[i][r] == NOSLOT;
                }
            }
            if (oob) { atomicExch(flags + FLAG_ABORT, 3u); dead = true; break; }     // key outside the caller's bounds
            if (bad) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); dead = true; break; }
            full[i] = cnt == 4 && sslot[i][0] != NOSLOT && sslot[i][1] != NOSLOT && sslot[i][2] != NOSLOT && sslot[i][3] != NOSLOT;
        }
        if (dead) break;
        bool all_full = true;
#pragma unroll
        for (int i = 0; i < QPI; ++i) all_full &= full[i];
        if (all_full) {
            // ---- fast path: QPI full quads, every row has a shared-memory slot.  The dispatch on the
            // accumulator kind happens once per accumulator, the row code is straight-line.
            for (int a = 0; a < prog.n; ++a) {
                const int kind = lay.kind[a];
                unsigned char* const c0 = sm + lay.off_cell[a];
                const int vtype = prog.op[a].vtype, xform = prog.op[a].xform;
                uint32_t u[QPI][4], u2[QPI][4];
                if (kind != CK_COUNT && kind != CK_MAXROW && kind != CK_TOP2) {
                    const int ci = prog.op[a].ci, ci2 = prog.op[a].ci2;
#pragma unroll
                    for (int i = 0; i < QPI; ++i) {
                        const uint4 c = pick_col(raw[i], ci);
                        u[i][0] = c.x; u[i][1] = c.y; u[i][2] = c.z; u[i][3] = c.w;
                        const uint4 d = xform == X_PRODUCT ? pick_col(raw[i], ci2) : c;
                        u2[i][0] = d.x; u2[i][1] = d.y; u2[i][2] = d.z; u2[i][3] = d.w;
                    }
                }
                switch (kind) {
                case CK_COUNT:
#pragma unroll
                    for (int i = 0; i < QPI; ++i)
#pragma unroll
                        for (int r = 0; r < 4; ++r) atomicAdd(reinterpret_cast<uint32_t*>(c0) + sslot[i][r], 1u);
                    break;
                case CK_ADD_I64:
                    if (xform == X_IDENT) {          // int32 value: high word is the sign
#pragma unroll
                        for (int i = 0; i < QPI; ++i)
#pragma unroll
                            for (int r = 0; r < 4; ++r) {
                                uint32_t* lo = reinterpret_cast<uint32_t*>(c0) + sslot[i][r];
                                const uint32_t xl = u[i][r];
                                const uint32_t old = atomicAdd(lo, xl);
                                const uint32_t d = (uint32_t)((int)xl >> 31) + (((uint32_t)(old + xl) < old) ? 1u : 0u);
                                if (d) atomicAdd(lo + cap_all, d);
                            }
                    } else {
#pragma unroll
                        for (int i = 0; i < QPI; ++i)
#pragma unroll
                            for (int r = 0; r < 4; ++r) {
                                const long long x = (long long)(int)u[i][r] * (long long)(int)u2[i][r];
                                uint32_t* lo = reinterpret_cast<uint32_t*>(c0) + sslot[i][r];
                                const uint32_t xl = (uint32_t)x;
                                const uint32_t old = atomicAdd(lo, xl);
                                const uint32_t d = (uint32_t)((unsigned long long)x >> 32) + (((uint32_t)(old + xl) < old) ? 1u : 0u);
                                if (d) atomicAdd(lo + cap_all, d);
                            }
                    }
                    break;
                case CK_ADD_F64: {
                    // compare-and-swap rounds over all QPI*4 rows at once: the loads, adds and CAS of a
                    // round are independent, so a round costs one shared-memory round trip, and the number
                    // of rounds is the largest number of rows of this warp that share a slot
                    double x[QPI][4];
                    unsigned long long old[QPI][4];
                    unsigned pending = 0;
#pragma unroll
                    for (int i = 0; i < QPI; ++i)
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            if (vtype == AQ_INT32) x[i][r] = xform == X_IDENT ? (double)(int)u[i][r] : (double)((long long)(int)u[i][r] * (long long)(int)u2[i][r]);
                            else x[i][r] = xform == X_IDENT ? (double)__uint_as_float(u[i][r]) : (double)__fmul_rn(__uint_as_float(u[i][r]), __uint_as_float(u2[i][r]));
                            old[i][r] = *(reinterpret_cast<volatile unsigned long long*>(c0) + sslot[i][r]);
                            pending |= 1u << (i * 4 + r);
                        }
                    while (pending) {
                        unsigned long long seen[QPI][4];
#pragma unroll
                        for (int i = 0; i < QPI; ++i)
#pragma unroll
                            for (int r = 0; r < 4; ++r)
                                if (pending & (1u << (i * 4 + r)))
                                    seen[i][r] = atomicCAS(reinterpret_cast<unsigned long long*>(c0) + sslot[i][r], old[i][r],
                                                           (unsigned long long)__double_as_longlong(__longlong_as_double((long long)old[i][r]) + x[i][r]));
#pragma unroll
                        for (int i = 0; i < QPI; ++i)
#pragma unroll
                            for (int r = 0; r < 4; ++r)
                                if (pending & (1u << (i * 4 + r))) {
                                    if (seen[i][r] == old[i][r]) pending &= ~(1u << (i * 4 + r));
                                    else old[i][r] = seen[i][r];
                                }
                    }
                    break; }
                case CK_MIN32: case CK_MAX32:
#pragma unroll
                    for (int i = 0; i < QPI; ++i)
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            int* p = reinterpret_cast<int*>(c0) + sslot[i][r];
                            const int b = (int)u[i][r];
                            const int img = vtype == AQ_INT32 ? b : (b ^ ((b >> 31) & 0x7fffffff));
                            const int cur = *reinterpret_cast<volatile int*>(p);
                            if (kind == CK_MIN32) { if (img < cur) atomicMin(p, img); }
                            else { if (img > cur) atomicMax(p, img); }
                        }
                    break;
                default:
#pragma unroll
                    for (int i = 0; i < QPI; ++i) {
                        const uint32_t row0 = (q0 + i * stride) << 2;
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            long long xi; double xd;
                            decode_x(prog.op[a], raw[i], r, xi, xd);
                            smem_acc_row(prog.op[a], kind, c0, cap_all, sslot[i][r], row0 + r, xi, xd);
                        }
                    }
                    break;
                }
            }
        } else {
            // ---- slow path: tails, rows that went to the global table
#pragma unroll
            for (int i = 0; i < QPI; ++i) {
                const int cnt = cnts[i];
                const uint32_t row0 = (q0 + i * stride) << 2;
                for (int a = 0; a < prog.n; ++a) {
                    const AccOp& o = prog.op[a];
                    unsigned char* c0 = sm + lay.off_cell[a];
                    const int kind = lay.kind[a];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        if (r >= cnt) continue;
                        long long xi; double xd;
                        decode_x(o, raw[i], r, xi, xd);
                        if (sslot[i][r] != NOSLOT) smem_acc_row(o, kind, c0, cap_all, sslot[i][r], row0 + r, xi, xd);
                        else acc_global_row(o, slot_word(prog, gslot[i][r]) + 1 + a, row0 + r, xi, xd);
                    }
                }
            }
        }
    }
    __syncthreads();
    if (*reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT)) return;

    // ---- merge this CTA's groups into the global table
    for (uint32_t s = threadIdx.x; s < cap_all; s += SMEM_THREADS) {
        const uint32_t row = minrow[s];
        if (row == EMPTY32) continue;
        uint32_t t, h;
        if (MODE == SM_DIRECT) { t = s & lay.mask; h = mix32(t); }
        else if (MODE == SM_HASH32) { t = key32[s]; h = mix32(t); }
        else key_tag1(ks, row, t, h);
        const uint32_t g = probe_insert<EXACT>(ks, prog.tab, prog.sw, gmask, max_probe, h, t, row);
        if (g == NOSLOT) { atomicCAS(flags + FLAG_ABORT, 0u, 1u); return; }
        for (int a = 0; a < prog.n; ++a)
            smem_merge(prog.op[a], lay.kind[a], sm + lay.off_cell[a], cap_all, s, slot_word(prog, g) + 1 + a);
    }
}

static uint32_t pow2_floor(uint64_t x) { uint32_t p = 1; while ((uint64_t(p) << 1) <= x) p <<= 1; return p; }

int launch_smem_probe(const KeySpec& ks, uint32_t nrows, uint32_t gcap, uint32_t max_probe,
                      const AccProg& prog, uint32_t groups_hint, bool* used) {
    *used = false;
    Context& c = ctx();
    SmemLayout lay;
    memset(&lay, 0, sizeof(lay));
    // ---- what this kernel handles: 1-2 stride-1 32-bit key words, <= MAX_VCOL 4-byte value columns
    if (ks.nwords < 1 || ks.nwords > 2 || ks.stride[0] != 1 || (ks.nwords == 2 && ks.stride[1] != 1)) return 0;
    if (prog.nv < 0) return 0;
    for (int j = 0; j < prog.nv; ++j) if (prog.vctype[j] != AQ_INT32 && prog.vctype[j] != AQ_FLOAT) return 0;
    if (ks.exact == 1 || ks.exact == 2) lay.mode = SM_HASH32;
    else if (ks.nwords == 2) lay.mode = SM_HASH64;
    else return 0;
    const bool direct_ok = ks.exact == 2 && ks.pack_bits <= 16;
    // ---- bytes per slot
    uint32_t cell_bytes[MAX_ACC];
    uint32_t bps = 4;                                             // minrow
    for (int a = 0; a < prog.n; ++a) {
        const AccOp& o = prog.op[a];
        const bool v32 = o.vtype == AQ_INT32 || o.vtype == AQ_FLOAT;
        switch (o.kind) {
        case ACC_COUNT: lay.kind[a] = CK_COUNT; cell_bytes[a] = 4; break;
        case ACC_ADD_I64: lay.kind[a] = CK_ADD_I64; cell_bytes[a] = 8; break;
        case ACC_ADD_F64: lay.kind[a] = CK_ADD_F64; cell_bytes[a] = 8; break;
        case ACC_MIN_I64: lay.kind[a] = v32 ? CK_MIN32 : CK_MIN64; cell_bytes[a] = v32 ? 4 : 8; break;
        case ACC_MAX_I64: lay.kind[a] = v32 ? CK_MAX32 : CK_MAX64; cell_bytes[a] = v32 ? 4 : 8; break;
        case ACC_MAXROW: lay.kind[a] = CK_MAXROW; cell_bytes[a] = 4; break;
        case ACC_TOP2ROW: lay.kind[a] = CK_TOP2; cell_bytes[a] = 8; break;
        default: return 0;
        }
        bps += cell_bytes[a];
    }
    const size_t budget = (c.smem_optin > 2048 ? c.smem_optin : 48 * 1024) - 1024;
    // ---- capacity and replication
    if (direct_ok && (size_t(bps) << ks.pack_bits) <= budget) {
        lay.mode = SM_DIRECT;
        lay.cap = 1u << ks.pack_bits;
        if (lay.cap < 32) lay.cap = 32;
        lay.limit = lay.cap;
    } else {
        bps += lay.mode == SM_HASH64 ? 8 : 4;
        const uint32_t cap_max = pow2_floor(budget / bps);
        if (cap_max < 4096) return 0;
        uint32_t cap = (groups_hint && groups_hint <= 512) ? 1024 : 4096;
        if (groups_hint) {
            while (cap < cap_max && (uint64_t(cap) * 3) / 4 < groups_hint) cap <<= 1;
            if ((uint64_t(cap) * 3) / 4 < groups_hint) return 0;      // more groups than one SM can hold
        } else {
            cap = cap_max;
        }
        lay.cap = cap;
        lay.limit = cap / 4 * 3;
    }
    lay.mask = lay.cap - 1;
    lay.reps = pow2_floor(budget / (size_t(bps) * lay.cap));
    if (lay.reps > 16) lay.reps = 16;
    if (lay.reps < 1) return 0;
    if (lay.mode != SM_DIRECT && lay.cap / 4 < SMEM_THREADS / lay.reps) {   // insertion slack >= threads per replica
        if (size_t(bps) * 4096 > budget) return 0;
        lay.cap = 4096; lay.mask = 4095; lay.limit = 3072;
        lay.reps = pow2_floor(budget / (size_t(bps) * lay.cap));
        if (lay.reps > 16) lay.reps = 16;
    }
    const uint32_t cap_all = lay.cap * lay.reps;
    // ---- offsets: 64-bit arrays first
    uint32_t off = 0;
    if (lay.mode == SM_HASH64) { lay.off_key = off; off += 8 * cap_all; }
    for (int a = 0; a < prog.n; ++a)
        if (cell_bytes[a] == 8) { lay.off_cell[a] = off; off += 8 * cap_all; }    // u64 arrays and (lo[], hi[]) pairs
    if (lay.mode == SM_HASH32) { lay.off_key = off; off += 4 * cap_all; }
    lay.off_minrow = off; off += 4 * cap_all;
    for (int a = 0; a < prog.n; ++a)
        if (cell_bytes[a] == 4) { lay.off_cell[a] = off; off += 4 * cap_all; }
    lay.total_bytes = off;
    if (off > budget) return 0;

    const uint32_t nquad = (nrows + 3) / 4;
    int grid = (int)((nquad + SMEM_THREADS - 1) / SMEM_THREADS);
    if (grid > c.sm_count) grid = c.sm_count;
    static bool attr_done = false;
    if (!attr_done) {
        AQ_CUDA(cudaFuncSetAttribute(k_smem_agg<SM_DIRECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        AQ_CUDA(cudaFuncSetAttribute(k_smem_agg<SM_HASH32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        AQ_CUDA(cudaFuncSetAttribute(k_smem_agg<SM_HASH64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)budget));
        attr_done = true;
    }
    switch (lay.mode) {
    case SM_DIRECT: AQ_LAUNCH(k_smem_agg<SM_DIRECT>, grid, SMEM_THREADS, lay.total_bytes, ks, nrows, gcap - 1, max_probe, prog, lay, c.d_flags); break;
    case SM_HASH32: AQ_LAUNCH(k_smem_agg<SM_HASH32>, grid, SMEM_THREADS, lay.total_bytes, ks, nrows, gcap - 1, max_probe, prog, lay, c.d_flags); break;
    default: AQ_LAUNCH(k_smem_agg<SM_HASH64>, grid, SMEM_THREADS, lay.total_bytes, ks, nrows, gcap - 1, max_probe, prog, lay, c.d_flags); break;
    }
    *used = true;
    return 0;
}

} // namespace aq
