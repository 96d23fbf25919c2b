// groupby_hc.cu -- high-cardinality group-by for wide exact keys (h2o q10: six key columns, groups ~ rows).
//
// Why: with ~1e9 groups no table fits anywhere but DRAM, and B200 sustains only 7-20 G random 32-byte
// read-modify-writes per second there (profiles/r01/microbench.txt: mb_random) -- k_probe_wide spends
// 145 ms on 1e9 rows, 20x the streaming time of the same bytes.  Random accesses that stay inside L2 run
// at 40-190 G/s.  So:
//
//   k_hc_scatter  every row -> one 32-byte tuple [row, tag, key word 2 (64 bit), up to 4 value words],
//                 appended to one of P hash partitions (P ~ rows / 3600; fixed capacity HC_CAP, the
//                 cursor of a partition is one L2 atomic).  32 bytes = one DRAM sector, so a tuple is a
//                 whole-sector write and a partition's frontier line completes in L2 before it leaves.
//   k_hc_agg      one CTA per partition at a time, with a private 8192-slot table in a global scratch
//                 that never leaves L2 (148 x 256 KB): insert / accumulate exactly like k_probe_wide, then
//                 every tuple writes ONE 32-byte record at rec[row]: the whole slot if `row` is the
//                 group's first appearance (slot min row == row), all ones otherwise; the first row
//                 clears its slot for the next partition.
//   k_emit_fused  (groupby.cu) walks rec[] in row order: rank by look-back, emit at gid.
//
// Every byte moves in streams or whole sectors; nothing random touches DRAM.  Replaces the same
// reference loops as groupby.cu (server/hasher.h:176-198, server/table.h:184-189, scalar aggregates of
// server/aggregations.h); the partitioning has no reference counterpart.
#include "groupby.cuh"
#include <cstring>

namespace aq {

constexpr int HC_THREADS = 1024;
constexpr int HC_TPT = 4;                              // tuples per thread of k_hc_agg
constexpr uint32_t HC_CAP = HC_THREADS * HC_TPT;       // tuples a partition can hold
constexpr uint32_t HC_MEAN = 3600;                     // target tuples per partition: HC_CAP is +8 sigma
constexpr uint32_t HC_SLOTS = 8192;                    // slots of a CTA's table (load <= 0.5)
constexpr int HC_TW = 8;                               // 32-bit words per tuple
constexpr int HC_SCATTER_THREADS = 256;

__device__ __forceinline__ void st_stream_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void st128_volatile(unsigned long long* p, unsigned long long a, unsigned long long b) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" :: "l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ uint32_t hc_part(uint32_t tag, unsigned long long hi, uint32_t nparts, uint32_t& slot_hash) {
    const uint64_t h = mix64(hi * 0x9e3779b97f4a7c15ULL + tag);
    slot_hash = (uint32_t)h;
    return __umulhi((uint32_t)(h >> 32), nparts);
}

// ---- phase 1: rows -> tuples, grouped by hash partition ------------------------------------------------
__global__ void __launch_bounds__(HC_SCATTER_THREADS)
k_hc_scatter(const KeySpec ks, const AccProg prog, const uint32_t nrows, const uint32_t nparts,
             uint32_t* __restrict__ cursor, uint32_t* __restrict__ tuples, uint32_t* __restrict__ flags) {
    const uint32_t nquad = (nrows + 3u) >> 2;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nquad; q += gridDim.x * blockDim.x) {
        const uint32_t row0 = q << 2;
        const int cnt = min(4u, nrows - row0);
        uint32_t tag[4];
        unsigned long long hi[4];
        if (!wide_pack4(ks, row0, cnt, tag, hi)) { atomicExch(flags + FLAG_ABORT, 3u); return; }
        uint32_t val[MAX_VCOL][4];
#pragma unroll
        for (int j = 0; j < MAX_VCOL; ++j) {
            if (j < prog.nv) load4_b32(prog.vcol[j], row0, cnt, (reinterpret_cast<uintptr_t>(prog.vcol[j]) & 15) == 0, val[j]);
            else { val[j][0] = val[j][1] = val[j][2] = val[j][3] = 0u; }
        }
        uint32_t part[4], pos[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            uint32_t sh;
            part[r] = hc_part(tag[r], hi[r], nparts, sh);
            pos[r] = r < cnt ? atomicAdd(cursor + part[r], 1u) : 0u;
        }
        bool over = false;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (r >= cnt) continue;
            if (pos[r] >= HC_CAP) { over = true; continue; }
            uint32_t* d = tuples + (size_t(part[r]) * HC_CAP + pos[r]) * HC_TW;
            st_stream_v4(d, row0 + r, tag[r], (uint32_t)hi[r], (uint32_t)(hi[r] >> 32));
            st_stream_v4(d + 4, val[0][r], val[1][r], val[2][r], val[3][r]);
        }
        if (over) atomicCAS(flags + FLAG_ABORT, 0u, 2u);           // a partition is full: this path does not fit the data
    }
}

// x for accumulator o from the value words of a tuple (same meaning as acc_inputs)
__device__ __forceinline__ void hc_x(const AccOp& o, const uint32_t (&v)[MAX_VCOL], long long& xi, double& xd) {
    xi = 0; xd = 0.0;
    if (!acc_needs_value(o)) return;
    auto pick = [&](int i) { return i == 0 ? v[0] : (i == 1 ? v[1] : (i == 2 ? v[2] : v[3])); };
    const uint32_t u = pick(o.ci);
    const uint32_t u2 = o.xform == X_PRODUCT ? pick(o.ci2) : u;
    if (o.vtype == AQ_INT32) {
        xi = (long long)(int)u;
        if (o.xform != X_IDENT) xi *= (long long)(int)u2;
    } else {
        const float f = __uint_as_float(u);
        xd = o.xform == X_IDENT ? (double)f : (double)__fmul_rn(f, __uint_as_float(u2));
    }
}

// ---- phase 2: per-partition aggregation in an L2-resident table, one record per row --------------------
// scratch: gridDim.x tables of HC_SLOTS slots x sw words, initialised to `pat` once by the CTA itself and
// left clean after every partition (the first row of a group clears the slot it reads).
struct HcPattern { unsigned long long w[16]; };

__global__ void __launch_bounds__(HC_THREADS, 1)
k_hc_agg(const AccProg prog, const uint32_t sw, const HcPattern pat, const uint32_t nparts,
         const uint32_t* __restrict__ cursor, const uint32_t* __restrict__ tuples,
         unsigned long long* __restrict__ scratch, unsigned long long* __restrict__ rec, uint32_t* __restrict__ flags) {
    __shared__ unsigned long long spat[16];
    if (flags[FLAG_ABORT]) return;
    unsigned long long* const tab = scratch + size_t(blockIdx.x) * HC_SLOTS * sw;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) spat[i] = pat.w[i];
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < HC_SLOTS * sw; i += HC_THREADS) tab[i] = spat[i & (sw - 1)];
    __syncthreads();
    constexpr uint32_t mask = HC_SLOTS - 1;
    for (uint32_t p = blockIdx.x; p < nparts; p += gridDim.x) {
        const uint32_t n = min(cursor[p], HC_CAP);
        const uint32_t* const base = tuples + size_t(p) * HC_CAP * HC_TW;
        uint32_t row[HC_TPT], slot[HC_TPT];
        // ---- probe + accumulate
        int4 ta[HC_TPT], tb[HC_TPT];
#pragma unroll
        for (int j = 0; j < HC_TPT; ++j) {
            const uint32_t t = j * HC_THREADS + threadIdx.x;
            if (t < n) { ta[j] = ld_stream_v4(base + size_t(t) * HC_TW); tb[j] = ld_stream_v4(base + size_t(t) * HC_TW + 4); }
        }
#pragma unroll
        for (int j = 0; j < HC_TPT; ++j) {
            const uint32_t t = j * HC_THREADS + threadIdx.x;
            slot[j] = NOSLOT; row[j] = 0;
            if (t >= n) continue;
            row[j] = (uint32_t)ta[j].x;
            const uint32_t tag = (uint32_t)ta[j].y;
            const unsigned long long hi = (unsigned long long)(uint32_t)ta[j].z | ((unsigned long long)(uint32_t)ta[j].w << 32);
            const unsigned long long mine = ((unsigned long long)tag << 32) | row[j];
            uint32_t sh;
            hc_part(tag, hi, nparts, sh);
            uint32_t s = sh & mask;
            for (uint32_t step = 0; step < HC_SLOTS; ++step) {
                unsigned long long* w0 = tab + size_t(s) * sw;
                unsigned long long a, b;
                ld128_volatile(w0, a, b);
                if (a == SLOT_EMPTY) {
                    unsigned long long o0, o1;
                    cas128(w0, SLOT_EMPTY, SLOT_EMPTY, mine, hi, o0, o1);
                    if (o0 == SLOT_EMPTY) { slot[j] = s; break; }
                    a = o0; b = o1;
                }
                if ((uint32_t)(a >> 32) == tag && b == hi) {
                    if (row[j] < (uint32_t)a) atomicMin(w0, mine);
                    slot[j] = s; break;
                }
                s = (s + 1) & mask;
            }
            const uint32_t v[MAX_VCOL] = {(uint32_t)tb[j].x, (uint32_t)tb[j].y, (uint32_t)tb[j].z, (uint32_t)tb[j].w};
            for (int a2 = 0; a2 < prog.n; ++a2) {
                long long xi; double xd;
                hc_x(prog.op[a2], v, xi, xd);
                acc_global_row(prog.op[a2], tab + size_t(slot[j]) * sw + 2 + a2, row[j], xi, xd);
            }
        }
        __syncthreads();                                   // every insert / accumulate of the partition is done (L2)
        // ---- one record per row; the first row takes the slot and clears it
#pragma unroll
        for (int j = 0; j < HC_TPT; ++j) {
            if (slot[j] == NOSLOT) continue;
            unsigned long long* const sl = tab + size_t(slot[j]) * sw;
            unsigned long long* const out = rec + size_t(row[j]) * sw;
            unsigned long long a, b;
            ld128_volatile(sl, a, b);
            if ((uint32_t)a == row[j]) {
                for (uint32_t w = 0; w < sw; w += 2) {
                    if (w) ld128_volatile(sl + w, a, b);
                    st_stream_v4(out + w, (uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
                }
                for (uint32_t w = sw; w > 0; w -= 2)        // word 0 last: the slot reads EMPTY only when all of it is clean
                    st128_volatile(sl + w - 2, spat[w - 2], spat[w - 1]);
            } else {
                for (uint32_t w = 0; w < sw; w += 2) st_stream_v4(out + w, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
            }
        }
        __syncthreads();                                   // the table is clean again
    }
}

// ---- phase 2, shared-memory tables (slots of 4 words: h2o q10) ------------------------------------------
// Same protocol as k_hc_agg with the table in the CTA's shared memory (HC_SM_SLOTS x 32 B = 192 KB): the
// L2 then only carries the two streams (tuples in, records out).
constexpr uint32_t HC_SM_SLOTS = 6144;
constexpr uint32_t HC_SM_SW = 4;

__device__ __forceinline__ void cas128_shared(uint32_t saddr, unsigned long long c0, unsigned long long c1,
                                              unsigned long long n0, unsigned long long n1,
                                              unsigned long long& o0, unsigned long long& o1) {
    asm volatile("{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 n, {%4, %5};\n\t"
                 "atom.relaxed.cta.shared.cas.b128 o, [%6], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
                 : "=l"(o0), "=l"(o1) : "l"(c0), "l"(c1), "l"(n0), "l"(n1), "r"(saddr) : "memory");
}
__device__ __forceinline__ void ld128_shared_volatile(uint32_t saddr, unsigned long long& a, unsigned long long& b) {
    asm volatile("ld.volatile.shared.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(saddr) : "memory");
}

__global__ void __launch_bounds__(HC_THREADS, 1)
k_hc_agg_smem(const AccProg prog, const HcPattern pat, const uint32_t nparts,
              const uint32_t* __restrict__ cursor, const uint32_t* __restrict__ tuples,
              unsigned long long* __restrict__ rec, unsigned char* __restrict__ rflags, const int cnt_acc,
              uint32_t* __restrict__ flags) {
    extern __shared__ __align__(16) unsigned long long tab[];          // HC_SM_SLOTS x 4 words
    if (flags[FLAG_ABORT]) return;
    const unsigned long long p0 = pat.w[0], p1 = pat.w[1], p2 = pat.w[2], p3 = pat.w[3];
    for (uint32_t i = threadIdx.x; i < HC_SM_SLOTS; i += HC_THREADS) {
        tab[i * 4 + 0] = p0; tab[i * 4 + 1] = p1; tab[i * 4 + 2] = p2; tab[i * 4 + 3] = p3;
    }
    __syncthreads();
    const uint32_t tab_s = (uint32_t)__cvta_generic_to_shared(tab);
    uint32_t p = blockIdx.x;
    // tuples of the first partition
    int4 ta[HC_TPT], tb[HC_TPT];
    uint32_t n = p < nparts ? min(cursor[p], HC_CAP) : 0u;
#pragma unroll
    for (int j = 0; j < HC_TPT; ++j) {
        const uint32_t t = j * HC_THREADS + threadIdx.x;
        if (t < n) {
            const uint32_t* src = tuples + (size_t(p) * HC_CAP + t) * HC_TW;
            ta[j] = ld_stream_v4(src); tb[j] = ld_stream_v4(src + 4);
        }
    }
    for (; p < nparts; p += gridDim.x) {
        uint32_t row[HC_TPT], slot[HC_TPT];
#pragma unroll
        for (int j = 0; j < HC_TPT; ++j) {
            const uint32_t t = j * HC_THREADS + threadIdx.x;
            slot[j] = NOSLOT; row[j] = 0;
            if (t >= n) continue;
            row[j] = (uint32_t)ta[j].x;
            const uint32_t tag = (uint32_t)ta[j].y;
            const unsigned long long hi = (unsigned long long)(uint32_t)ta[j].z | ((unsigned long long)(uint32_t)ta[j].w << 32);
            const unsigned long long mine = ((unsigned long long)tag << 32) | row[j];
            uint32_t sh;
            hc_part(tag, hi, nparts, sh);
            uint32_t s = __umulhi(sh, HC_SM_SLOTS);
            for (uint32_t step = 0; step < HC_SM_SLOTS; ++step) {
                const uint32_t sa = tab_s + s * 32u;
                unsigned long long a, b;
                ld128_shared_volatile(sa, a, b);
                if (a == SLOT_EMPTY) {
                    unsigned long long o0, o1;
                    cas128_shared(sa, SLOT_EMPTY, SLOT_EMPTY, mine, hi, o0, o1);
                    if (o0 == SLOT_EMPTY) { slot[j] = s; break; }
                    a = o0; b = o1;
                }
                if ((uint32_t)(a >> 32) == tag && b == hi) {
                    if (row[j] < (uint32_t)a) atomicMin(tab + s * 4, mine);
                    slot[j] = s; break;
                }
                s = s + 1 == HC_SM_SLOTS ? 0u : s + 1;
            }
            const uint32_t v[MAX_VCOL] = {(uint32_t)tb[j].x, (uint32_t)tb[j].y, (uint32_t)tb[j].z, (uint32_t)tb[j].w};
            for (int a2 = 0; a2 < prog.n; ++a2) {
                long long xi; double xd;
                hc_x(prog.op[a2], v, xi, xd);
                acc_global_row(prog.op[a2], tab + slot[j] * 4 + 2 + a2, row[j], xi, xd);
            }
        }
        // the next partition's tuples travel while this one is finished
        const uint32_t pn = p + gridDim.x;
        const uint32_t nn = pn < nparts ? min(cursor[pn], HC_CAP) : 0u;
#pragma unroll
        for (int j = 0; j < HC_TPT; ++j) {
            const uint32_t t = j * HC_THREADS + threadIdx.x;
            if (t < nn) {
                const uint32_t* src = tuples + (size_t(pn) * HC_CAP + t) * HC_TW;
                ta[j] = ld_stream_v4(src); tb[j] = ld_stream_v4(src + 4);
            }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < HC_TPT; ++j) {
            if (slot[j] == NOSLOT) continue;
            unsigned long long* const sl = tab + slot[j] * 4;
            unsigned long long* const out = rec + size_t(row[j]) * 4;
            const unsigned long long a = *reinterpret_cast<volatile unsigned long long*>(sl);
            if ((uint32_t)a == row[j]) {
                const unsigned long long b = sl[1], c2 = sl[2], d = sl[3];
                // rflags (needs a COUNT accumulator): a one-row group writes NOTHING -- k_emit_fused rebuilds
                // its cells from the row's own values; rflags[row] stays 0 = "first row, singleton"
                const bool single = rflags && (cnt_acc == 0 ? c2 : d) == 1ull;
                if (!single) {
                    st_stream_v4(out, (uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
                    st_stream_v4(out + 2, (uint32_t)c2, (uint32_t)(c2 >> 32), (uint32_t)d, (uint32_t)(d >> 32));
                    if (rflags) rflags[row[j]] = 1;
                }
                sl[3] = p3; sl[2] = p2; sl[1] = p1;
                *reinterpret_cast<volatile unsigned long long*>(sl) = p0;
            } else if (rflags) {
                rflags[row[j]] = 2;
            } else {
                st_stream_v4(out, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
                st_stream_v4(out + 2, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
            }
        }
        __syncthreads();
        n = nn;
    }
}

// ---- host ---------------------------------------------------------------------------------------------
bool hc_qualifies(const KeySpec& ks, uint32_t nrows, const AccProg& prog, uint32_t sw, uint32_t groups_hint) {
    if (!ks.wide || ks.exact != 0 || nrows < (1u << 20)) return false;
    if (groups_hint < nrows / 8) return false;             // few groups: tables elsewhere are small and cheap
    if (sw < 2 || sw > 16 || prog.nv < 0 || prog.nv > MAX_VCOL) return false;
    for (int j = 0; j < prog.nv; ++j) if (prog.vctype[j] != AQ_INT32 && prog.vctype[j] != AQ_FLOAT) return false;
    for (int w = 0; w < ks.nwords; ++w) if (ks.stride[w] != 1) return false;
    return true;
}

int launch_hc(const KeySpec& ks, uint32_t nrows, const AccProg& prog, uint32_t sw, const unsigned long long* pattern,
              uint32_t groups_hint, DevBuf& rec, DevBuf& rflags, bool* used) {
    *used = false;
    Context& c = ctx();
    if (!hc_qualifies(ks, nrows, prog, sw, groups_hint)) return 0;
    const uint32_t nparts = (nrows + HC_MEAN - 1) / HC_MEAN;
    const int grid = c.sm_count;
    DevBuf tuples, cursor, scratch;
    AQ_TRY(tuples.alloc(size_t(nparts) * HC_CAP * HC_TW * 4));
    AQ_TRY(cursor.alloc_zero(size_t(nparts) * 4));
    AQ_TRY(scratch.alloc(size_t(grid) * HC_SLOTS * sw * 8));
    AQ_TRY(rec.alloc(size_t(nrows) * sw * 8));
    HcPattern pat;
    for (int i = 0; i < 16; ++i) pat.w[i] = pattern[i];
    AQ_LAUNCH(k_hc_scatter, grid_for((nrows + 3) / 4, HC_SCATTER_THREADS, 1, 16), HC_SCATTER_THREADS, 0, ks, prog, nrows, nparts,
              cursor.as<uint32_t>(), tuples.as<uint32_t>(), c.d_flags);
    if (sw == HC_SM_SW && c.smem_optin >= HC_SM_SLOTS * 32 + 1024) {
        const size_t smem = size_t(HC_SM_SLOTS) * 32;
        static bool attr_done = false;
        if (!attr_done) { AQ_CUDA(cudaFuncSetAttribute(k_hc_agg_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr_done = true; }
        int cnt_acc = -1;
        for (int a = 0; a < prog.n && a < 2; ++a) if (prog.op[a].kind == ACC_COUNT) cnt_acc = a;
        if (cnt_acc >= 0) AQ_TRY(rflags.alloc_zero(size_t(nrows)));
        AQ_LAUNCH(k_hc_agg_smem, grid, HC_THREADS, smem, prog, pat, nparts, cursor.as<uint32_t>(), tuples.as<uint32_t>(),
                  rec.as<unsigned long long>(), cnt_acc >= 0 ? rflags.as<unsigned char>() : nullptr, cnt_acc, c.d_flags);
        *used = true;
        return 0;
    }
    AQ_LAUNCH(k_hc_agg, grid, HC_THREADS, 0, prog, sw, pat, nparts, cursor.as<uint32_t>(), tuples.as<uint32_t>(),
              scratch.as<unsigned long long>(), rec.as<unsigned long long>(), c.d_flags);
    *used = true;
    return 0;
}

} // namespace aq
