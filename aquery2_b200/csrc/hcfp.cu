// hcfp.cu -- group-by for groups ~ rows (h2o q10: six key columns, every row its own group): duplicate
// detection by 64-bit fingerprint, partitioned so that every random access stays inside one SM.
//
// Replaces the reference's per-row hashtable_push / ht_postproc / gather / aggregate loops
// (server/hasher.h:176-198, server/table.h:184-189, server/aggregations.h) for ANY key tuple (no bounds
// needed: the columns are hashed, never packed).  A DRAM-resident table costs one random sector
// read-modify-write per row (~2e10/s on B200, profiles/r01/microbench.txt); here the only per-row traffic
// beyond the columns themselves is an 8-byte tuple (row id, 32-bit fingerprint) written and read twice:
//
//   k_fp_part1   rows -> tuples, split into P1 partitions by fingerprint bits.  A CTA sorts a tile of 8192
//                tuples by partition in shared memory (histogram with returned ranks, scan, scatter) and
//                writes every partition's run as whole 32-byte sectors (runs padded with invalid tuples, one
//                L2 atomic per run reserves the space): DRAM sees full-sector, mostly full-line writes.
//   k_fp_part2   the same split of every level-1 partition into P2 sub-partitions (~3800 tuples each).
//   k_fp_agg     one CTA per final partition: a shared-memory table keyed by the fingerprint.  A fingerprint
//                that is already present is verified against the key columns (gather of both rows' keys), so
//                the result is exact whatever the hash does.  Rows that turn out to share their key tuple
//                (rare by construction: the planner sends a query here only when its sample says groups ~
//                rows) accumulate into a compact record list through global atomics; every other row writes
//                NOTHING -- rflags[row] stays 0 = "first row of a one-row group".
//   k_fp_emit    one pass in row order: a tile counts its first rows, gets the number of first rows before it
//                by decoupled look-back, and every first row writes its keys (read where they lie, coalesced)
//                and its aggregates (from its own values, or from its record) at gid = that rank: the
//                insertion index m_values has in the reference (server/unordered_dense.h:1142-1146).
#include "lc.cuh"
#include "emit.cuh"
#include "sort.cuh"
#include "comm.cuh"
#include <cmath>
#include <cstring>

namespace aq {

constexpr int FP_THREADS = 1024;
constexpr int FP_RPT = 8;                               // tuples per thread and tile
constexpr int FP_T = FP_THREADS * FP_RPT;               // 8192
constexpr int FP_PMAX = 1024;
constexpr uint32_t FP_STAGE = FP_T + 3 * FP_PMAX;       // staged tuples incl. padding
constexpr unsigned long long FP_INVALID = 0xffffffffffffffffULL;   // row = 0xffffffff
constexpr uint32_t FP_NOPOS = 0xffffffffu;
constexpr uint32_t FP_MEAN = 3800;                      // target tuples per final partition
constexpr uint32_t FP_SLOTS = 8192;                     // table of k_fp_agg
constexpr int FPA_THREADS = 512;
constexpr int FPA_TPT = 10;                             // a final partition holds at most 5120 tuples

struct FpSmem {
    uint32_t hist[FP_PMAX];
    uint32_t loff[FP_PMAX + 1];
    uint32_t gpos[FP_PMAX];
    uint32_t wsum[32];
    uint32_t total;
    unsigned short secp[FP_STAGE / 4];
    alignas(16) unsigned long long stage[FP_STAGE];
};

__device__ __forceinline__ uint64_t fp_hash(const uint32_t* kw, int n) {
    uint64_t h = 0x9e3779b97f4a7c15ULL;
    for (int w = 0; w < n; ++w) { h = (h ^ kw[w]) * 0xff51afd7ed558ccdULL; h ^= h >> 32; }
    h *= 0xc4ceb9fe1a85ec53ULL;
    return h ^ (h >> 29);
}

// Sorts the CTA's tile (RPT tuples per thread, part[i] < P, valid bit i of vmask) by partition in shared
// memory and appends every partition's run, padded to whole sectors, to its region out + p * cap.
__device__ __forceinline__ void fp_tile_scatter(FpSmem& S, const unsigned long long (&tup)[FP_RPT], const uint32_t (&part)[FP_RPT],
                                                uint32_t vmask, uint32_t P, uint32_t* __restrict__ cursor,
                                                unsigned long long* __restrict__ out, uint32_t cap, uint32_t* __restrict__ flags) {
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    uint32_t pr[FP_RPT];
#pragma unroll
    for (int i = 0; i < FP_RPT; ++i) {
        pr[i] = 0;
        if (vmask & (1u << i)) pr[i] = (part[i] << 16) | atomicAdd(&S.hist[part[i]], 1u);
    }
    __syncthreads();
    // ---- padded exclusive scan over the partitions (one per thread), space reservation, sector map
    {
        const uint32_t c = tid < P ? S.hist[tid] : 0u;
        const uint32_t pc = (c + 3u) & ~3u;
        uint32_t x = pc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= (uint32_t)d) x += y; }
        if (lane == 31) S.wsum[wid] = x;
        __syncthreads();
        if (wid == 0) {
            uint32_t v = S.wsum[lane], sc = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, sc, d); if (lane >= (uint32_t)d) sc += y; }
            S.wsum[lane] = sc - v;
            if (lane == 31) S.total = sc;
        }
        __syncthreads();
        const uint32_t ex = S.wsum[wid] + x - pc;
        if (tid < P) {
            S.loff[tid] = ex;
            uint32_t gp = FP_NOPOS;
            if (pc) {
                gp = atomicAdd(cursor + tid, pc);
                if (gp + pc > cap) { gp = FP_NOPOS; atomicCAS(flags + FLAG_ABORT, 0u, 2u); }     // region full: not this path
            }
            S.gpos[tid] = gp;
            for (uint32_t s = ex >> 2; s < (ex + pc) >> 2; ++s) S.secp[s] = (unsigned short)tid;
            for (uint32_t i = c; i < pc; ++i) S.stage[ex + i] = FP_INVALID;
            S.hist[tid] = 0u;                                   // ready for the next tile
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < FP_RPT; ++i)
        if (vmask & (1u << i)) S.stage[S.loff[pr[i] >> 16] + (pr[i] & 0xffffu)] = tup[i];
    __syncthreads();
    // ---- copy out: one 32-byte sector per thread and step, consecutive threads = consecutive sectors
    const uint32_t nsec = S.total >> 2;
    for (uint32_t s = tid; s < nsec; s += FP_THREADS) {
        const uint32_t p = S.secp[s];
        const uint32_t gp = S.gpos[p];
        if (gp == FP_NOPOS) continue;
        const uint4* src = reinterpret_cast<const uint4*>(S.stage + size_t(s) * 4);
        uint4* dst = reinterpret_cast<uint4*>(out + size_t(p) * cap + gp + (s * 4u - S.loff[p]));
        const uint4 a = src[0], b = src[1];
        dst[0] = a; dst[1] = b;
    }
    __syncthreads();
}

// ---- level 1: key columns -> (row, fingerprint) tuples in P1 partitions -------------------------------------
__global__ void __launch_bounds__(FP_THREADS, 1)
k_fp_part1(const KeySpec ks, const uint32_t nrows, const uint32_t P1, uint32_t* __restrict__ cursor,
           unsigned long long* __restrict__ out, const uint32_t cap, uint32_t* __restrict__ flags, const uint32_t row_base) {
    extern __shared__ __align__(16) unsigned char fp_sm[];
    FpSmem& S = *reinterpret_cast<FpSmem*>(fp_sm);
    for (uint32_t i = threadIdx.x; i < FP_PMAX; i += FP_THREADS) S.hist[i] = 0u;
    __syncthreads();
    const uint32_t ntiles = (nrows + FP_T - 1) / FP_T;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        unsigned long long tup[FP_RPT];
        uint32_t part[FP_RPT];
        uint32_t vmask = 0;
        // two quads of consecutive rows per thread; the hash state of the eight rows lives in registers and the
        // key words stream through it, two columns (four 128-bit loads) in flight at a time
        uint64_t hs[FP_RPT];
        uint32_t row0[2]; int cnt[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t q = tile * (FP_T / 4) + h * FP_THREADS + threadIdx.x;
            row0[h] = q << 2;
            cnt[h] = row0[h] < nrows ? (int)min(4u, nrows - row0[h]) : 0;
        }
#pragma unroll
        for (int i = 0; i < FP_RPT; ++i) hs[i] = 0x9e3779b97f4a7c15ULL;
#pragma unroll 2
        for (int w = 0; w < ks.nwords; ++w) {
            uint32_t kw[2][4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                kw[h][0] = kw[h][1] = kw[h][2] = kw[h][3] = 0u;
                if (cnt[h] == 0) continue;
                if (ks.stride[w] == 1) load4_b32(ks.base[w], row0[h], cnt[h], ks.vec_ok, kw[h]);
                else {
#pragma unroll
                    for (int r = 0; r < 4; ++r) if (r < cnt[h]) kw[h][r] = __ldg(ks.base[w] + size_t(row0[h] + r) * ks.stride[w]);
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    uint64_t x = (hs[h * 4 + r] ^ kw[h][r]) * 0xff51afd7ed558ccdULL;
                    hs[h * 4 + r] = x ^ (x >> 32);
                }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int i = h * 4 + r;
                tup[i] = 0; part[i] = 0;
                if (r >= cnt[h]) continue;
                uint64_t hsh = hs[i] * 0xc4ceb9fe1a85ec53ULL;
                hsh ^= hsh >> 29;
                part[i] = __umulhi((uint32_t)(hsh >> 32), P1);
                tup[i] = ((unsigned long long)(uint32_t)hsh << 32) | (unsigned long long)(row_base + row0[h] + r);      // row shards: GLOBAL row id
                vmask |= 1u << i;
            }
        fp_tile_scatter(S, tup, part, vmask, P1, cursor, out, cap, flags);
    }
}

// ---- level 2: every level-1 partition -> P2 sub-partitions ---------------------------------------------------
__global__ void __launch_bounds__(FP_THREADS, 1)
k_fp_part2(const unsigned long long* __restrict__ in, const uint32_t* __restrict__ cursor1, const uint32_t cap1,
           const uint32_t P1, const uint32_t P2, uint32_t* __restrict__ cursor2, unsigned long long* __restrict__ out,
           const uint32_t cap2, uint32_t* __restrict__ flags, const uint32_t PG) {
    // PG: input region p1 feeds the sub-partitions of group p1 % PG (one GPU: PG = P1; row shards: the regions of
    // one level-1 partition received from every rank share a group)
    extern __shared__ __align__(16) unsigned char fp_sm[];
    FpSmem& S = *reinterpret_cast<FpSmem*>(fp_sm);
    for (uint32_t i = threadIdx.x; i < FP_PMAX; i += FP_THREADS) S.hist[i] = 0u;
    if (threadIdx.x == 0) S.total = *reinterpret_cast<volatile uint32_t*>(flags + FLAG_ABORT);
    __syncthreads();
    if (S.total) return;                                        // level 1 gave up (uniform: read once per CTA)
    __syncthreads();
    uint32_t tix = 0;                                           // running tile index over all level-1 partitions
    for (uint32_t p1 = 0; p1 < P1; ++p1) {
        const uint32_t n = min(__ldg(cursor1 + p1), cap1);
        const uint32_t nt = (n + FP_T - 1) / FP_T;
        uint32_t j = (blockIdx.x + gridDim.x - tix % gridDim.x) % gridDim.x;     // first tile of p1 that is mine
        for (; j < nt; j += gridDim.x) {
            const unsigned long long* src = in + size_t(p1) * cap1 + size_t(j) * FP_T;
            const uint32_t left = n - j * FP_T;
            unsigned long long tup[FP_RPT];
            uint32_t part[FP_RPT];
            uint32_t vmask = 0;
#pragma unroll
            for (int i = 0; i < FP_RPT / 2; ++i) {
                const uint32_t idx = (i * FP_THREADS + threadIdx.x) * 2u;       // regions and caps are multiples of 4 tuples
                uint4 v = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
                if (idx < left) v = *reinterpret_cast<const uint4*>(src + idx);
                tup[2 * i] = (unsigned long long)v.x | ((unsigned long long)v.y << 32);
                tup[2 * i + 1] = (unsigned long long)v.z | ((unsigned long long)v.w << 32);
            }
#pragma unroll
            for (int i = 0; i < FP_RPT; ++i) {
                part[i] = 0;
                if ((uint32_t)tup[i] != 0xffffffffu) { part[i] = __umulhi((uint32_t)(tup[i] >> 32), P2); vmask |= 1u << i; }
            }
            fp_tile_scatter(S, tup, part, vmask, P2, cursor2 + size_t(p1 % PG) * P2, out + size_t(p1 % PG) * P2 * cap2, cap2, flags);
        }
        tix += nt;
    }
}

// ---- level 3: per-partition duplicate detection ----------------------------------------------------------------
struct FpAggArgs {
    const unsigned long long* tuples;     // npart regions of cap tuples
    const uint32_t* cursor;
    uint32_t npart, cap;
    unsigned char* rflags;                // nrows, zero-filled: 0 first row of a one-row group, 1 first row with a record, 2 not a first row
    uint32_t* rec_of_row;                 // nrows, sparse: record index of a first row with a record
    unsigned long long* rec;              // rec_cap records of nacc cells
    uint32_t rec_cap;
    uint32_t* nrec;
    uint32_t* flags;
    uint32_t* plist;                      // partitions with duplicated key tuples (k_fp_screen -> k_fp_agg)
    uint32_t* nlist;
    // row shards: rows are GLOBAL ids and may live on another rank, so equal fingerprints cannot be settled here:
    // the pair goes on a candidate list that every rank resolves after a small exchange (k_fp_resolve)
    uint32_t* cand;                       // nullptr on one GPU
    uint32_t* ncand;
    uint32_t cand_cap;
};

// Screen: almost every final partition holds distinct key tuples only.  One 32-bit compare-and-swap per
// tuple into a table of (19 fingerprint bits, 13-bit tuple index) entries finds the candidates; a candidate
// pair is settled by the full fingerprint and then by the key columns themselves.  Partitions with a real
// duplicate go on plist for k_fp_agg; for all the others rflags stays 0 and nothing is written.
constexpr int FPS_THREADS = 512;
constexpr uint32_t FPS_EMPTY = 0xffffffffu;
constexpr size_t FPS_SMEM = size_t(FP_SLOTS) * 4 + size_t(FPS_THREADS) * FPA_TPT * 8;
__global__ void __launch_bounds__(FPS_THREADS, 3)
k_fp_screen(const KeySpec ks, const FpAggArgs A) {
    extern __shared__ __align__(16) unsigned char fps_sm[];
    unsigned long long* const stage = reinterpret_cast<unsigned long long*>(fps_sm);   // tuple i of the partition at stage[i]
    uint32_t* const tab = reinterpret_cast<uint32_t*>(stage + size_t(FPS_THREADS) * FPA_TPT);
    __shared__ uint32_t s_abort, s_dup;
    if (threadIdx.x == 0) s_abort = *reinterpret_cast<volatile uint32_t*>(A.flags + FLAG_ABORT);
    __syncthreads();
    if (s_abort) return;
    constexpr uint32_t mask = FP_SLOTS - 1;
    for (uint32_t p = blockIdx.x; p < A.npart; p += gridDim.x) {
        const uint32_t n = min(__ldg(A.cursor + p), A.cap);
        const unsigned long long* src = A.tuples + size_t(p) * A.cap;
#pragma unroll
        for (int j = 0; j < FPA_TPT; ++j) {
            const uint32_t idx = j * FPS_THREADS + threadIdx.x;
            stage[idx] = idx < n ? __ldg(src + idx) : FP_INVALID;
        }
        for (uint32_t i = threadIdx.x; i < FP_SLOTS; i += FPS_THREADS) tab[i] = FPS_EMPTY;
        if (threadIdx.x == 0) s_dup = 0u;
        __syncthreads();
        bool dup = false;
        // every lane walks its own tuples at its own pace: one probe per iteration, the next tuple as soon as one is placed
        uint32_t idx = threadIdx.x, s = 0, mine = 0;
        unsigned long long t = 0;
        bool fresh = true;
        while (idx < n) {
            if (fresh) {
                t = stage[idx];
                if ((uint32_t)t == 0xffffffffu) { idx += FPS_THREADS; continue; }          // sector padding
                const uint32_t fp = (uint32_t)(t >> 32);
                s = (fp * 0x9E3779B1u) >> 19;
                mine = (fp << 13) | idx;
                if (mine == FPS_EMPTY) mine ^= 1u << 13;         // never the empty pattern (the full compare below settles it)
                fresh = false;
            }
            const uint32_t old = atomicCAS(tab + s, FPS_EMPTY, mine);
            if (old == FPS_EMPTY) { idx += FPS_THREADS; fresh = true; continue; }
            if (((old ^ mine) >> 14) == 0u) {                     // 18 tag bits agree: look at the other tuple
                const unsigned long long other = stage[old & 0x1fffu];
                if ((uint32_t)(other >> 32) == (uint32_t)(t >> 32)) {
                    if (A.cand) {
                        const uint32_t ci = atomicAdd(A.ncand, 2u);
                        if (ci + 1u < A.cand_cap) { A.cand[ci] = (uint32_t)t; A.cand[ci + 1u] = (uint32_t)other; }
                        else atomicCAS(A.flags + FLAG_ABORT, 0u, 2u);            // many equal keys: not this path
                    } else if (keys_equal(ks, (uint32_t)t, (uint32_t)other)) dup = true;
                }
            }
            s = (s + 1u) & mask;
        }
        if (dup) s_dup = 1u;
        __syncthreads();
        if (threadIdx.x == 0 && s_dup) A.plist[atomicAdd(A.nlist, 1u)] = p;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(FPA_THREADS, 2)
k_fp_agg(const KeySpec ks, const AccProg prog, const FpAggArgs A, const HcIdentity ident) {
    extern __shared__ __align__(16) unsigned char fpa_sm[];
    unsigned long long* const tab = reinterpret_cast<unsigned long long*>(fpa_sm);
    uint32_t* const multi = reinterpret_cast<uint32_t*>(tab + FP_SLOTS);
    __shared__ uint32_t s_abort;
    if (threadIdx.x == 0) s_abort = *reinterpret_cast<volatile uint32_t*>(A.flags + FLAG_ABORT);
    __syncthreads();
    if (s_abort) return;
    constexpr uint32_t mask = FP_SLOTS - 1;
    const uint32_t nlist = min(*A.nlist, A.npart);
    for (uint32_t li = blockIdx.x; li < nlist; li += gridDim.x) {
        const uint32_t p = A.plist[li];
        for (uint32_t i = threadIdx.x; i < FP_SLOTS; i += FPA_THREADS) tab[i] = FP_INVALID;
        for (uint32_t i = threadIdx.x; i < FP_SLOTS / 32; i += FPA_THREADS) multi[i] = 0u;
        const uint32_t n = min(__ldg(A.cursor + p), A.cap);
        const unsigned long long* src = A.tuples + size_t(p) * A.cap;
        unsigned long long tup[FPA_TPT];
#pragma unroll
        for (int j = 0; j < FPA_TPT; ++j) {
            const uint32_t idx = j * FPA_THREADS + threadIdx.x;
            tup[j] = idx < n ? __ldg(src + idx) : FP_INVALID;
        }
        __syncthreads();
        uint32_t myslot[FPA_TPT];
        bool any = false;
#pragma unroll
        for (int j = 0; j < FPA_TPT; ++j) {
            myslot[j] = FP_NOPOS;
            const uint32_t row = (uint32_t)tup[j];
            if (row == 0xffffffffu) continue;
            const uint32_t fp = (uint32_t)(tup[j] >> 32);
            const unsigned long long mine = tup[j];
            uint32_t s = (fp * 0x9E3779B1u) >> 19;                 // 13 bits
            for (uint32_t step = 0; step < FP_SLOTS; ++step) {
                unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(tab + s);
                if (cur == FP_INVALID) {
                    cur = atomicCAS(tab + s, FP_INVALID, mine);
                    if (cur == FP_INVALID) { myslot[j] = s; break; }
                }
                if ((uint32_t)(cur >> 32) == fp && keys_equal(ks, row, (uint32_t)cur)) {
                    atomicMin(tab + s, mine);                         // same fingerprint: orders by row
                    atomicOr(multi + (s >> 5), 1u << (s & 31u));
                    myslot[j] = s; any = true; break;
                }
                s = (s + 1u) & mask;
            }
        }
        // ---- groups with more than one row (rare): records through global atomics
        if (__syncthreads_or(any ? 1 : 0)) {
            uint32_t role[FPA_TPT];                                // 0 none, 1 first row of a multi-row group, 2 other row of one
#pragma unroll
            for (int j = 0; j < FPA_TPT; ++j) {
                role[j] = 0;
                const uint32_t s = myslot[j];
                if (s == FP_NOPOS || !((multi[s >> 5] >> (s & 31u)) & 1u)) continue;
                role[j] = (uint32_t)tab[s] == (uint32_t)tup[j] ? 1u : 2u;
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < FPA_TPT; ++j) {
                if (role[j] != 1u) continue;
                const uint32_t ridx = atomicAdd(A.nrec, 1u);
                const uint32_t row = (uint32_t)tup[j];
                if (ridx < A.rec_cap) {
                    for (int a = 0; a < prog.n; ++a) A.rec[size_t(ridx) * prog.n + a] = ident.w[a];
                    A.rec_of_row[row] = ridx;
                    A.rflags[row] = 1;
                } else {
                    atomicCAS(A.flags + FLAG_ABORT, 0u, 2u);           // more duplicated groups than this path provides for
                }
                tab[myslot[j]] = (unsigned long long)ridx;
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < FPA_TPT; ++j) {
                if (role[j] == 0u) continue;
                const uint32_t row = (uint32_t)tup[j];
                const uint32_t ridx = (uint32_t)tab[myslot[j]];
                if (role[j] == 2u) A.rflags[row] = 2;
                if (ridx >= A.rec_cap) continue;
                for (int a = 0; a < prog.n; ++a) {
                    long long xi = 0; double xd = 0.0;
                    if (acc_needs_value(prog.op[a])) acc_row_x(prog.op[a], row, xi, xd);
                    acc_global_row(prog.op[a], A.rec + size_t(ridx) * prog.n + a, row, xi, xd);
                }
            }
        }
        __syncthreads();
    }
}

// ---- emit: one pass in row order ------------------------------------------------------------------------------
// k_fp_tilecount: first rows per tile (a first row is any row whose flag is not 2); an exclusive scan of the
// tile counts (sort.cu) gives every tile the rank of its first first-row; k_fp_emit then streams: flags ->
// ranks inside the tile by ballots -> for every output column, the tile's rows are read where they lie and
// written at their ranks (both coalesced: consecutive lanes = consecutive rows = consecutive ranks).
constexpr int FPE_THREADS = 256;
constexpr int FPE_RPT = 4;
constexpr int FPE_ROWS = FPE_THREADS * FPE_RPT;          // 1024

__global__ void __launch_bounds__(FPE_THREADS)
k_fp_tilecount(const uint32_t nrows, const unsigned char* __restrict__ rflags, uint32_t* __restrict__ tile_count,
               const uint32_t* __restrict__ flags) {
    __shared__ uint32_t wsum[FPE_THREADS / 32];
    if (*reinterpret_cast<const volatile uint32_t*>(flags + FLAG_ABORT)) return;
    static_assert(FPE_RPT == 4, "one 32-bit word of flags per thread");
    const uint32_t base = blockIdx.x * FPE_ROWS + threadIdx.x * FPE_RPT;       // 4 consecutive flags per thread
    uint32_t c = 0;
    if (base + FPE_RPT <= nrows) {
        const uint32_t v = *reinterpret_cast<const uint32_t*>(rflags + base);
        c = FPE_RPT - (uint32_t)__popc(v & 0x02020202u);         // bytes are 0, 1 or 2: bit 1 marks "not a first row"
    } else {
        for (uint32_t r = base; r < nrows; ++r) c += rflags[r] != 2 ? 1u : 0u;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < FPE_THREADS / 32; ++w) t += wsum[w];
        tile_count[blockIdx.x] = t;
    }
}

// One output column of a one-row group: dst[rank] = conv(src[row]).
enum FpConv : int { FC_COPY4 = 0, FC_COPY8, FC_I32_I128, FC_I32_I64, FC_I32_F64, FC_F32_F64, FC_MAXF32, FC_ONE_I64, FC_ONE_U32,
                    FC_ZERO4, FC_ZERO8, FC_ROW };
constexpr int FP_MAX_OUTCOLS = MAX_KEY_COLS + MAX_AGGS + 1;
struct FpOutCols {
    int n;
    int simple;                           // every requested aggregate has a column op (else: emit_one per row)
    int conv[FP_MAX_OUTCOLS];
    const void* src[FP_MAX_OUTCOLS];
    void* dst[FP_MAX_OUTCOLS];
};

template <int CONV>
__device__ __forceinline__ void fp_emit_col(const void* __restrict__ src, void* __restrict__ dst, const uint32_t (&row)[FPE_RPT],
                                            const uint32_t (&g)[FPE_RPT], const uint32_t okm) {
    if (CONV == FC_COPY8) {
        unsigned long long v[FPE_RPT];
#pragma unroll
        for (int j = 0; j < FPE_RPT; ++j) if (okm & (1u << j)) v[j] = __ldg(static_cast<const unsigned long long*>(src) + row[j]);
#pragma unroll
        for (int j = 0; j < FPE_RPT; ++j) if (okm & (1u << j)) static_cast<unsigned long long*>(dst)[g[j]] = v[j];
        return;
    }
    uint32_t v[FPE_RPT];
    if (CONV != FC_ONE_I64 && CONV != FC_ONE_U32 && CONV != FC_ZERO4 && CONV != FC_ZERO8 && CONV != FC_ROW) {
#pragma unroll
        for (int j = 0; j < FPE_RPT; ++j) if (okm & (1u << j)) v[j] = ld_stream_u32(static_cast<const uint32_t*>(src) + row[j]);
    }
#pragma unroll
    for (int j = 0; j < FPE_RPT; ++j) {
        if (!(okm & (1u << j))) continue;
        const uint32_t gi = g[j];
        switch (CONV) {
        case FC_COPY4: static_cast<uint32_t*>(dst)[gi] = v[j]; break;
        case FC_I32_I128: { const long long x = (long long)(int)v[j]; I128 o; o.lo = (unsigned long long)x; o.hi = x >> 63; static_cast<I128*>(dst)[gi] = o; break; }
        case FC_I32_I64: static_cast<long long*>(dst)[gi] = (long long)(int)v[j]; break;
        case FC_I32_F64: static_cast<double*>(dst)[gi] = (double)(int)v[j]; break;
        case FC_F32_F64: static_cast<double*>(dst)[gi] = (double)__uint_as_float(v[j]); break;
        case FC_MAXF32: { float f = __uint_as_float(v[j]); if (!(f > 1.17549435e-38f)) f = 1.17549435e-38f; static_cast<float*>(dst)[gi] = f; break; }   // aggregations.h:73
        case FC_ONE_I64: static_cast<long long*>(dst)[gi] = 1ll; break;
        case FC_ONE_U32: static_cast<uint32_t*>(dst)[gi] = 1u; break;
        case FC_ZERO4: static_cast<uint32_t*>(dst)[gi] = 0u; break;
        case FC_ZERO8: static_cast<unsigned long long*>(dst)[gi] = 0ull; break;
        case FC_ROW: static_cast<uint32_t*>(dst)[gi] = row[j]; break;
        default: break;
        }
    }
}

// rows of this thread in the tile, their flags and ranks (the rank of a row that is not a first row is unused)
__device__ __forceinline__ void fp_tile_ranks(const uint32_t nrows, const unsigned char* __restrict__ rflags, const uint32_t* __restrict__ tile_excl,
                                              uint32_t (&wcnt)[FPE_RPT][FPE_THREADS / 32], uint32_t (&row)[FPE_RPT], uint32_t (&g)[FPE_RPT],
                                              uint32_t (&fl)[FPE_RPT]) {
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    constexpr int NW = FPE_THREADS / 32;
    const uint32_t base = blockIdx.x * FPE_ROWS;
#pragma unroll
    for (int j = 0; j < FPE_RPT; ++j) {
        row[j] = base + j * FPE_THREADS + threadIdx.x;
        fl[j] = row[j] < nrows ? (uint32_t)rflags[row[j]] : 2u;
        const uint32_t m = __ballot_sync(0xffffffffu, fl[j] != 2u);
        g[j] = __popc(m & ((1u << lane) - 1u));
        if (lane == 0) wcnt[j][wid] = __popc(m);
    }
    __syncthreads();
    static_assert(FPE_RPT * NW == 32, "one warp count per lane");
    if (wid == 0) {                       // exclusive scan of the 4 x NW warp counts (row order: j major, warp minor)
        const uint32_t v = wcnt[lane / NW][lane % NW];
        uint32_t sc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, sc, d); if (lane >= (uint32_t)d) sc += y; }
        wcnt[lane / NW][lane % NW] = sc - v;
    }
    __syncthreads();
    const uint32_t excl = tile_excl[blockIdx.x];
#pragma unroll
    for (int j = 0; j < FPE_RPT; ++j) g[j] += excl + wcnt[j][wid];
}

__global__ void __launch_bounds__(FPE_THREADS, 4)
k_fp_emit(const uint32_t nrows, const unsigned char* __restrict__ rflags, const uint32_t* __restrict__ tile_excl,
          const FpOutCols oc, const uint32_t max_groups, const uint32_t* __restrict__ flags) {
    __shared__ uint32_t wcnt[FPE_RPT][FPE_THREADS / 32];
    if (*reinterpret_cast<const volatile uint32_t*>(flags + FLAG_ABORT)) return;
    uint32_t row[FPE_RPT], g[FPE_RPT], fl[FPE_RPT];
    fp_tile_ranks(nrows, rflags, tile_excl, wcnt, row, g, fl);
    uint32_t okm = 0, allm = 0;           // one-row groups the column ops finish / every first row
#pragma unroll
    for (int j = 0; j < FPE_RPT; ++j) {
        if (fl[j] != 2u && g[j] < max_groups) {           // the host reports an overflow of max_groups
            allm |= 1u << j;
            if (fl[j] == 0u) okm |= 1u << j;
        }
    }
    for (int c = 0; c < oc.n; ++c) {
        const void* src = oc.src[c]; void* dst = oc.dst[c];
        const uint32_t m = oc.conv[c] < 0 ? allm : okm;  // keys and first rows: every first row alike
        switch (oc.conv[c] < 0 ? -1 - oc.conv[c] : oc.conv[c]) {
        case FC_COPY4: fp_emit_col<FC_COPY4>(src, dst, row, g, m); break;
        case FC_COPY8: fp_emit_col<FC_COPY8>(src, dst, row, g, m); break;
        case FC_I32_I128: fp_emit_col<FC_I32_I128>(src, dst, row, g, m); break;
        case FC_I32_I64: fp_emit_col<FC_I32_I64>(src, dst, row, g, m); break;
        case FC_I32_F64: fp_emit_col<FC_I32_F64>(src, dst, row, g, m); break;
        case FC_F32_F64: fp_emit_col<FC_F32_F64>(src, dst, row, g, m); break;
        case FC_MAXF32: fp_emit_col<FC_MAXF32>(src, dst, row, g, m); break;
        case FC_ONE_I64: fp_emit_col<FC_ONE_I64>(src, dst, row, g, m); break;
        case FC_ONE_U32: fp_emit_col<FC_ONE_U32>(src, dst, row, g, m); break;
        case FC_ZERO4: fp_emit_col<FC_ZERO4>(src, dst, row, g, m); break;
        case FC_ZERO8: fp_emit_col<FC_ZERO8>(src, dst, row, g, m); break;
        case FC_ROW: fp_emit_col<FC_ROW>(src, dst, row, g, m); break;
        default: break;
        }
    }
}

// The aggregates of groups with a record -- and of every group when some requested aggregate has no column
// op (all = 1) -- through the generic finaliser.  Does nothing when there is no record and all = 0.
__global__ void __launch_bounds__(FPE_THREADS)
k_fp_emit_rest(const uint32_t nrows, const unsigned char* __restrict__ rflags, const uint32_t* __restrict__ tile_excl,
               const uint32_t* __restrict__ rec_of_row, const unsigned long long* __restrict__ rec, const uint32_t* __restrict__ nrec,
               const int all, const __grid_constant__ EmitProg ep, const __grid_constant__ AccProg prog, const uint32_t max_groups,
               const uint32_t* __restrict__ flags) {
    __shared__ uint32_t wcnt[FPE_RPT][FPE_THREADS / 32];
    if (*reinterpret_cast<const volatile uint32_t*>(flags + FLAG_ABORT)) return;
    if (!all && *nrec == 0u) return;
    uint32_t row[FPE_RPT], g[FPE_RPT], fl[FPE_RPT];
    fp_tile_ranks(nrows, rflags, tile_excl, wcnt, row, g, fl);
#pragma unroll 1
    for (int j = 0; j < FPE_RPT; ++j) {
        uint32_t r = 0, gi = 0, f = 2;
#pragma unroll
        for (int i = 0; i < FPE_RPT; ++i) if (i == j) { r = row[i]; gi = g[i]; f = fl[i]; }
        if (f == 2u || gi >= max_groups || (f == 0u && !all)) continue;
        unsigned long long one[1 + MAX_ACC];
        const unsigned long long* cells = one;
        if (f == 0u) {
            for (int a = 0; a < prog.n; ++a) one[1 + a] = acc_single_row(prog.op[a], r);
        } else {
            cells = rec + size_t(rec_of_row[r]) * prog.n - 1;
        }
        for (int a = 0; a < ep.n; ++a) emit_one(ep.agg[a], cells, gi, r, ep.row_order);
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
bool hcfp_qualifies(const KeyCols& kc, uint32_t nrows, const AggPlan& plan, const PlanStats& st) {
    (void)kc; (void)plan;
    if (nrows < (1u << 21)) return false;
    if (st.est_groups < 0.9 * (double)nrows) return false;
    // the sample must itself be (nearly) duplicate free: S^2 / (2 G) collisions expected at G = 0.9 rows
    return true;
}

static uint32_t region_cap(double mean, double runs) {
    const double c = mean + 8.0 * std::sqrt(mean) + 1.5 * 1.25 * runs + 64.0;
    return ((uint32_t)c + 3u) & ~3u;
}

int run_hcfp(const KeyCols& kc, uint32_t nrows, AggPlan& plan, OutSink& sink, uint32_t* ngroups, bool* used) {
    Context& c = ctx();
    *used = false;
    KeySpec ks;
    AQ_TRY(kc.make_spec(nullptr, nullptr, ks));
    // ---- partitions
    const double parts = std::ceil((double)nrows / FP_MEAN);
    uint32_t P1, P2;
    if (parts <= FP_PMAX) { P1 = (uint32_t)parts; P2 = 1; }
    else {
        P1 = (uint32_t)std::ceil(std::sqrt(parts));
        if (P1 > FP_PMAX) return 0;                              // > 4e9 rows: not reachable with 32-bit row ids
        P2 = (uint32_t)std::ceil(parts / P1);
    }
    if (P1 < 1) P1 = 1;
    const double ntiles1 = std::ceil((double)nrows / FP_T);
    const uint32_t cap1 = region_cap((double)nrows / P1, ntiles1);
    const double tiles2 = std::ceil((double)cap1 / FP_T);       // tiles a level-1 partition is read in
    const uint32_t cap2 = P2 > 1 ? region_cap((double)nrows / P1 / P2, tiles2) : cap1;
    if (cap2 > (uint32_t)FPA_THREADS * FPA_TPT) return 0;
    const uint32_t nfinal = P1 * P2;
    DevBuf part1, part2, cur, rflags, rec_of_row, rec, state;
    AQ_TRY(part1.alloc(size_t(P1) * cap1 * 8));
    if (P2 > 1) AQ_TRY(part2.alloc(size_t(nfinal) * cap2 * 8));
    AQ_TRY(cur.alloc_zero((size_t(P1) + nfinal + 8) * 4));
    AQ_TRY(rflags.alloc_zero(size_t(nrows)));
    AQ_TRY(rec_of_row.alloc(size_t(nrows) * 4));
    const uint32_t rec_cap = std::max<uint32_t>(1u << 16, nrows / 16);
    AQ_TRY(rec.alloc(size_t(rec_cap) * std::max(plan.n_acc, 1) * 8));
    const uint32_t ntiles_e = (nrows + FPE_ROWS - 1) / FPE_ROWS;
    AQ_TRY(state.alloc(size_t(ntiles_e) * 8 + 8));
    DevBuf plist;
    AQ_TRY(plist.alloc(size_t(nfinal) * 4));
    uint32_t* cursor1 = cur.as<uint32_t>();
    uint32_t* cursor2 = cursor1 + P1;
    uint32_t* nrec = cursor2 + nfinal;
    uint32_t* nlist = nrec + 1;
    AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 8 * sizeof(uint32_t), c.stream));
    static bool attr_done = false;
    if (!attr_done) {
        AQ_CUDA(cudaFuncSetAttribute(k_fp_part1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FpSmem)));
        AQ_CUDA(cudaFuncSetAttribute(k_fp_part2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FpSmem)));
        AQ_CUDA(cudaFuncSetAttribute(k_fp_agg, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(size_t(FP_SLOTS) * 8 + FP_SLOTS / 8)));
        AQ_CUDA(cudaFuncSetAttribute(k_fp_screen, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FPS_SMEM));
        attr_done = true;
    }
    AQ_LAUNCH(k_fp_part1, c.sm_count, FP_THREADS, sizeof(FpSmem), ks, nrows, P1, cursor1, part1.as<unsigned long long>(), cap1, c.d_flags, 0u);
    const unsigned long long* fin = part1.as<unsigned long long>();
    const uint32_t* fcur = cursor1;
    if (P2 > 1) {
        AQ_LAUNCH(k_fp_part2, c.sm_count, FP_THREADS, sizeof(FpSmem), part1.as<unsigned long long>(), cursor1, cap1, P1, P2, cursor2,
                  part2.as<unsigned long long>(), cap2, c.d_flags, P1);
        fin = part2.as<unsigned long long>(); fcur = cursor2;
    }
    FpAggArgs A;
    A.tuples = fin; A.cursor = fcur; A.npart = nfinal; A.cap = cap2;
    A.rflags = rflags.as<unsigned char>(); A.rec_of_row = rec_of_row.as<uint32_t>();
    A.rec = rec.as<unsigned long long>(); A.rec_cap = rec_cap; A.nrec = nrec; A.flags = c.d_flags;
    A.plist = plist.as<uint32_t>(); A.nlist = nlist;
    A.cand = nullptr; A.ncand = nullptr; A.cand_cap = 0;
    HcIdentity ident;
    for (int a = 0; a < MAX_ACC; ++a) ident.w[a] = 0ull;
    for (int a = 0; a < plan.n_acc; ++a) {
        if (plan.prog.op[a].kind == ACC_MIN_I64) ident.w[a] = 0x7fffffffffffffffULL;
        if (plan.prog.op[a].kind == ACC_MAX_I64) ident.w[a] = 0x8000000000000000ULL;
    }
    const size_t agg_smem = size_t(FP_SLOTS) * 8 + FP_SLOTS / 8;
    AQ_LAUNCH(k_fp_screen, (int)std::min<uint32_t>(nfinal, (uint32_t)c.sm_count * 3), FPS_THREADS, FPS_SMEM, ks, A);
    AQ_LAUNCH(k_fp_agg, (int)std::min<uint32_t>(nfinal, (uint32_t)c.sm_count * 2), FPA_THREADS, agg_smem, ks, plan.prog, A, ident);
    // ---- emit (the kernel does nothing when an earlier stage gave up)
    AQ_TRY(sink.prepare(sink.res ? nrows : sink.max_groups, plan));
    KeyOut ko;
    memset(&ko, 0, sizeof(ko));
    ko.n = kc.n;
    for (int k = 0; k < kc.n; ++k) { ko.src[k] = kc.col[k].d; ko.size[k] = (int)dtype_size(kc.type[k]); ko.dst[k] = sink.key_dst[k]; }
    for (int a = 0; a < plan.emit.n; ++a) plan.emit.agg[a].out = sink.agg_dst[a];
    FpOutCols oc;
    memset(&oc, 0, sizeof(oc));
    oc.simple = 1;
    auto add_col = [&](int conv, const void* src, void* dst) { oc.conv[oc.n] = conv; oc.src[oc.n] = src; oc.dst[oc.n] = dst; ++oc.n; };
    for (int k = 0; k < kc.n; ++k)
        if (ko.dst[k]) add_col(-1 - (ko.size[k] == 4 ? FC_COPY4 : FC_COPY8), ko.src[k], ko.dst[k]);
    if (sink.fr_dst) add_col(-1 - FC_ROW, nullptr, sink.fr_dst);
    const int nfixed = oc.n;
    for (int a = 0; a < plan.emit.n && oc.simple; ++a) {
        const EmitAgg& e = plan.emit.agg[a];
        const AccOp* o = (e.a[0] >= 0 && e.a[0] < plan.n_acc) ? &plan.prog.op[e.a[0]] : nullptr;
        const bool v4 = e.vtype == AQ_INT32 || e.vtype == AQ_FLOAT;
        int conv = -1; const void* src = nullptr;
        switch (e.op) {
        case AQ_SUM:
            if (o && o->xform == X_IDENT && v4) { src = o->v; conv = e.vtype == AQ_FLOAT ? FC_F32_F64 : (e.out_type == AQ_INT128 ? FC_I32_I128 : FC_I32_I64); }
            break;
        case AQ_AVG:
            if (o && o->xform == X_IDENT && v4) { src = o->v; conv = e.vtype == AQ_FLOAT ? FC_F32_F64 : FC_I32_F64; }
            break;
        case AQ_MIN: case AQ_MAX:
            if (o && o->xform == X_IDENT && v4) { src = o->v; conv = (e.op == AQ_MAX && e.vtype == AQ_FLOAT) ? FC_MAXF32 : FC_COPY4; }
            break;
        case AQ_COUNT: conv = e.out_type == AQ_INT64 ? FC_ONE_I64 : FC_ONE_U32; break;
        case AQ_FIRST: case AQ_LAST: src = e.v; conv = v4 ? FC_COPY4 : FC_COPY8; break;
        case AQ_SECOND: conv = v4 ? FC_ZERO4 : FC_ZERO8; break;
        default: break;
        }
        if (conv < 0) { oc.simple = 0; break; }
        add_col(conv, src, e.out);
    }
    if (!oc.simple) oc.n = nfixed;
    uint32_t* tile_count = state.as<uint32_t>();
    uint32_t* tile_excl = tile_count + ntiles_e + 1;
    AQ_LAUNCH(k_fp_tilecount, (int)ntiles_e, FPE_THREADS, 0, nrows, rflags.as<unsigned char>(), tile_count, c.d_flags);
    AQ_TRY(exclusive_scan_u32(tile_count, tile_excl, ntiles_e, c.d_flags + FLAG_NGROUPS));
    AQ_LAUNCH(k_fp_emit, (int)ntiles_e, FPE_THREADS, 0, nrows, rflags.as<unsigned char>(), tile_excl, oc, sink.cap, c.d_flags);
    AQ_LAUNCH(k_fp_emit_rest, (int)ntiles_e, FPE_THREADS, 0, nrows, rflags.as<unsigned char>(), tile_excl, rec_of_row.as<uint32_t>(),
              rec.as<unsigned long long>(), nrec, oc.simple ? 0 : 1, plan.emit, plan.prog, sink.cap, c.d_flags);
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) return 0;                     // a region or the record list overflowed: generic path
    const uint32_t ng = c.h_flags[FLAG_NGROUPS];
    if (ng > sink.cap) return fail("capacity: %u groups exceed max_groups=%u", ng, sink.cap);
    AQ_TRY(sink.finish(ng, plan));
    *ngroups = ng;
    *used = true;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// row shards (dist.cu): groups ~ rows over several GPUs WITHOUT moving the rows
// ------------------------------------------------------------------------------------------------
// Only the 8-byte (global row, fingerprint) tuples travel: rank d owns the level-1 partitions [d * PG, (d + 1) * PG),
// every rank sends its regions of those partitions to d (one grouped send / receive), d splits them further and
// screens every final partition for equal fingerprints.  Equal fingerprints are CANDIDATES (~n^2 / 2^51 false ones
// for all-distinct keys, a few hundred at 1e9 rows): the candidate rows are all-gathered, their home ranks publish
// key tuples and values (an all-reduce over a table that is zero elsewhere), and every rank settles the candidates
// identically (k_fp_resolve): a row that is not its key's first appearance gets flag 2 on its home rank, a first row
// of a multi-row group gets a record.  Every other row stays what the zero-filled flags say -- the first row of a
// one-row group -- and the emit is the single-GPU one over the LOCAL rows: a rank emits the groups whose first row it
// holds, in row order, so the ranks' results concatenated in rank order are in global first-appearance order.
// Key tuples that really repeat more than FPC_TOTAL / 2 times overall make every rank decline.
constexpr uint32_t FPC_TOTAL = 65536;           // candidate list entries over all ranks (the exact resolve is O(entries^2): ~5 ms at the cap)

// entry i of the gathered candidate list: key words and value words of row cand[i] if it lives here, zeros otherwise
__global__ void k_fp_cand_fill(const KeySpec ks, const AccProg prog, const uint32_t* __restrict__ cand, const uint32_t m,
                               const uint32_t row_base, const uint32_t nrows, uint32_t* __restrict__ tab, const uint32_t stride) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const uint32_t r = cand[i] - row_base;
        uint32_t* t = tab + size_t(i) * stride;
        for (uint32_t w = 0; w < stride; ++w) t[w] = 0u;
        if (r >= nrows) continue;
        for (int w = 0; w < ks.nwords; ++w) t[w] = ks.base[w][size_t(r) * ks.stride[w]];
        for (int c = 0; c < prog.nv; ++c) t[ks.nwords + c] = static_cast<const uint32_t*>(prog.vcol[c])[r];
    }
}

// first filter of the candidates: a second, independent 64-bit hash of the key tuple, computed by the row's home rank
// (zeros elsewhere, so an all-reduce sum publishes it); list slot e holds a row id or 0xffffffff
__global__ void k_fp_cand_hash(const KeySpec ks, const uint32_t* __restrict__ slots, const uint32_t nslots, const uint32_t row_base,
                               const uint32_t nrows, uint32_t* __restrict__ h2) {
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < nslots; e += gridDim.x * blockDim.x) {
        const uint32_t r = slots[e] - row_base;
        uint32_t lo = 0u, hi = 0u;
        if (slots[e] != 0xffffffffu && r < nrows) {
            uint64_t h = 0x2545F4914F6CDD1DULL;
            for (int w = 0; w < ks.nwords; ++w) { h = (h ^ ks.base[w][size_t(r) * ks.stride[w]]) * 0x9E3779B97F4A7C15ULL; h ^= h >> 29; }
            h *= 0xD6E8FEB86659FD93ULL; h ^= h >> 32;
            lo = (uint32_t)h; hi = (uint32_t)(h >> 32);
        }
        h2[2 * size_t(e)] = lo; h2[2 * size_t(e) + 1] = hi;
    }
}
// my pairs whose second hashes agree too -> the short list that is settled exactly
__global__ void k_fp_cand_filter(const uint32_t* __restrict__ mine, const uint32_t cap_in, const uint32_t* __restrict__ h2_mine,
                                 uint32_t* __restrict__ out, const uint32_t cap_out, uint32_t* __restrict__ flags) {
    const uint32_t n = min(mine[0], cap_in) / 2u;                    // pairs; mine[1 + 2i], mine[2 + 2i]
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const size_t a = 1 + 2 * size_t(i), b = a + 1;
        if (h2_mine[2 * a] != h2_mine[2 * b] || h2_mine[2 * a + 1] != h2_mine[2 * b + 1]) continue;
        const uint32_t j = atomicAdd(out, 2u);
        if (j + 1u < cap_out) { out[1 + j] = mine[a]; out[2 + j] = mine[b]; }
        else atomicCAS(flags + FLAG_ABORT, 0u, 2u);                  // key tuples really repeat a lot: not this path
    }
}

// canon[i] = 1 iff entry i is the first occurrence of its row in the list
__global__ void k_fp_canon(const uint32_t* __restrict__ cand, const uint32_t m, unsigned char* __restrict__ canon) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const uint32_t row = cand[i];
        unsigned char c = 1;
        for (uint32_t j = 0; j < i; ++j) if (cand[j] == row) { c = 0; break; }
        canon[i] = c;
    }
}

// one thread per list entry; identical on every rank
__global__ void __launch_bounds__(256)
k_fp_resolve(const AccProg prog, const HcIdentity ident, const uint32_t* __restrict__ cand, const unsigned char* __restrict__ canon, const uint32_t m, const uint32_t* __restrict__ tab,
             const uint32_t stride, const int nkw, const uint32_t row_base, const uint32_t nrows, unsigned char* __restrict__ rflags,
             uint32_t* __restrict__ rec_of_row, unsigned long long* __restrict__ rec, const uint32_t rec_cap, uint32_t* __restrict__ nrec,
             uint32_t* __restrict__ flags) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const uint32_t row = cand[i];
        const uint32_t local = row - row_base;
        if (local >= nrows) continue;                               // another rank's row: its home settles it
        if (!canon[i]) continue;
        const uint32_t* ki = tab + size_t(i) * stride;
        uint32_t first = row, members = 0;
        for (uint32_t j = 0; j < m; ++j) {
            const uint32_t* kj = tab + size_t(j) * stride;
            bool eq = true;
            for (int w = 0; w < nkw; ++w) eq &= ki[w] == kj[w];
            if (!eq) continue;
            if (!canon[j]) continue;                                  // count every row once
            ++members;
            first = min(first, cand[j]);
        }
        if (members < 2u) continue;                                  // a false candidate: equal fingerprints, different keys
        if (row != first) { rflags[local] = 2; continue; }
        const uint32_t ridx = atomicAdd(nrec, 1u);
        if (ridx >= rec_cap) { atomicCAS(flags + FLAG_ABORT, 0u, 2u); continue; }
        unsigned long long cell[MAX_ACC];
        for (int a = 0; a < prog.n; ++a) cell[a] = ident.w[a];
        for (uint32_t j = 0; j < m; ++j) {
            const uint32_t* kj = tab + size_t(j) * stride;
            bool eq = true;
            for (int w = 0; w < nkw; ++w) eq &= ki[w] == kj[w];
            if (!eq) continue;
            if (!canon[j]) continue;
            for (int a = 0; a < prog.n; ++a) {
                const AccOp& o = prog.op[a];
                long long xi = 0; double xd = 0.0;
                if (acc_needs_value(o)) {
                    const uint32_t u = kj[nkw + o.ci];
                    const uint32_t u2 = o.xform == X_PRODUCT ? kj[nkw + o.ci2] : u;
                    if (o.vtype == AQ_INT32) { xi = (long long)(int)u; if (o.xform != X_IDENT) xi *= (long long)(int)u2; }
                    else { const float f = __uint_as_float(u); xd = o.xform == X_IDENT ? (double)f : (double)__fmul_rn(f, __uint_as_float(u2)); }
                }
                const bool isint = o.vtype == AQ_INT32;
                switch (o.kind) {
                case ACC_COUNT: cell[a] += 1ull; break;
                case ACC_ADD_I64: cell[a] += (unsigned long long)xi; break;
                case ACC_ADD_F64: cell[a] = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)cell[a]) + (isint ? (double)xi : xd)); break;
                case ACC_MIN_I64: cell[a] = (unsigned long long)min((long long)cell[a], isint ? xi : f64_to_ordered(xd)); break;
                case ACC_MAX_I64: cell[a] = (unsigned long long)max((long long)cell[a], isint ? xi : f64_to_ordered(xd)); break;
                default: break;
                }
            }
        }
        for (int a = 0; a < prog.n; ++a) rec[size_t(ridx) * prog.n + a] = cell[a];
        rec_of_row[local] = ridx;
        rflags[local] = 1;
    }
}

__global__ void k_fp_add_base(uint32_t* __restrict__ rows, const uint32_t* __restrict__ n_dev, uint32_t base) {
    const uint32_t n = *n_dev;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) rows[i] += base;
}

bool hcfp_sharded_qualifies(const KeyCols& kc, const AggPlan& plan, const PlanStats& st) {
    (void)kc;
    const DistCtx& d = dist();
    if (d.rows_min < (1u << 18)) return false;
    if (st.est_groups < 0.9 * (double)d.rows_max) return false;
    if (plan.prog.nv < 0) return false;
    for (int c = 0; c < plan.prog.nv; ++c) if (plan.prog.vctype[c] != AQ_INT32 && plan.prog.vctype[c] != AQ_FLOAT) return false;
    for (int a = 0; a < plan.n_acc; ++a) {
        const int k = plan.prog.op[a].kind;
        if (k == ACC_MAXROW || k == ACC_TOP2ROW) return false;       // row-valued cells would need remote rows
        if (k != ACC_COUNT && plan.prog.op[a].vtype != AQ_INT32 && plan.prog.op[a].vtype != AQ_FLOAT) return false;
    }
    return true;
}

int run_hcfp_sharded(const KeyCols& kc, uint32_t nrows, AggPlan& plan, OutSink& sink, uint32_t* ngroups, bool* used) {
    Context& c = ctx();
    DistCtx& D = dist();
    *used = false;
    const uint32_t N = (uint32_t)D.nranks;
    KeySpec ks;
    AQ_TRY(kc.make_spec(nullptr, nullptr, ks));
    // ---- partitions: sized from the WHOLE table, level 1 split evenly over the ranks
    const double ntot = (double)D.rows_total;
    const double parts = std::ceil(ntot / FP_MEAN);
    uint32_t P1 = parts <= FP_PMAX ? (uint32_t)parts : (uint32_t)std::ceil(std::sqrt(parts));
    if (P1 < N) P1 = N;
    P1 = (P1 + N - 1) / N * N;
    if (P1 > FP_PMAX) return 0;
    const uint32_t P2 = std::max<uint32_t>(1u, (uint32_t)std::ceil(parts / P1));
    if (P2 > FP_PMAX) return 0;
    const uint32_t PG = P1 / N;                                   // level-1 partitions this rank owns
    const double nmax = (double)D.rows_max;
    const uint32_t cap1 = region_cap(nmax / P1, std::ceil(nmax / FP_T));            // a region holds ONE rank's tuples of a partition
    const double tiles2 = std::ceil((double)cap1 / FP_T) * N;
    const uint32_t cap2 = region_cap(ntot / P1 / P2, tiles2);
    if (cap2 > (uint32_t)FPA_THREADS * FPA_TPT) return 0;
    const uint32_t nfinal = PG * P2;
    const uint32_t cand_cap = FPC_TOTAL / N;
    // equal 32-bit fingerprints inside a final partition: its sub-partition index has used log2(P2) of the bits, so
    // ~ rows x FP_MEAN / 2 / 2^(32 - log2 P2) pairs are expected for all-distinct keys (2e5 at 1e9 rows)
    const double exp_pairs = ntot * FP_MEAN / 2.0 / (4294967296.0 / P2);
    const uint32_t cand1_cap = (uint32_t)std::min<double>(64.0 * 1024 * 1024, 8.0 * exp_pairs / N + 16384.0) & ~1u;
    DevBuf part1, recv1, part2, cur, rflags, rec_of_row, rec, state, plist, cand, candall, cand1, cand1all, h2;
    AQ_TRY(part1.alloc(size_t(P1) * cap1 * 8));
    AQ_TRY(recv1.alloc(size_t(P1) * cap1 * 8));
    AQ_TRY(part2.alloc(size_t(nfinal) * cap2 * 8));
    AQ_TRY(cur.alloc_zero((size_t(2) * P1 + nfinal + 8) * 4));
    AQ_TRY(rflags.alloc_zero(size_t(nrows) + 4));
    AQ_TRY(rec_of_row.alloc(size_t(nrows) * 4 + 4));
    const uint32_t rec_cap = FPC_TOTAL;
    AQ_TRY(rec.alloc(size_t(rec_cap) * std::max(plan.n_acc, 1) * 8));
    const uint32_t ntiles_e = (nrows + FPE_ROWS - 1) / FPE_ROWS;
    AQ_TRY(state.alloc(size_t(ntiles_e) * 8 + 16));
    AQ_TRY(plist.alloc(size_t(nfinal) * 4 + 4));
    AQ_TRY(cand.alloc_zero(size_t(cand_cap + 1) * 4));           // [0] = count, then the rows
    AQ_TRY(candall.alloc(size_t(cand_cap + 1) * 4 * N));
    AQ_TRY(cand1.alloc(size_t(cand1_cap + 1) * 4));
    AQ_TRY(cand1all.alloc(size_t(cand1_cap + 1) * 4 * N));
    AQ_TRY(h2.alloc(size_t(cand1_cap + 1) * 8 * N));
    AQ_CUDA(cudaMemsetAsync(cand1.p, 0xff, size_t(cand1_cap + 1) * 4, c.stream));
    AQ_CUDA(cudaMemsetAsync(cand1.p, 0, 4, c.stream));
    uint32_t* cursor1 = cur.as<uint32_t>();
    uint32_t* rcursor1 = cursor1 + P1;
    uint32_t* cursor2 = rcursor1 + P1;
    uint32_t* nrec = cursor2 + nfinal;
    uint32_t* nlist = nrec + 1;
    AQ_CUDA(cudaMemsetAsync(c.d_flags, 0, 8 * sizeof(uint32_t), c.stream));
    static bool attr_done = false;
    if (!attr_done) {
        AQ_CUDA(cudaFuncSetAttribute(k_fp_part1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FpSmem)));
        AQ_CUDA(cudaFuncSetAttribute(k_fp_part2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FpSmem)));
        AQ_CUDA(cudaFuncSetAttribute(k_fp_screen, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FPS_SMEM));
        attr_done = true;
    }
    AQ_LAUNCH(k_fp_part1, c.sm_count, FP_THREADS, sizeof(FpSmem), ks, nrows, P1, cursor1, part1.as<unsigned long long>(), cap1, c.d_flags,
              (uint32_t)D.row_base);
    // ---- regions of the partitions rank d owns -> d (tuples and fill counts)
    {
        size_t so[64], sb[64], ro[64], rb[64];
        if (N > 64) return fail("multi-GPU: more than 64 ranks");
        for (uint32_t r = 0; r < N; ++r) { so[r] = size_t(r) * PG * cap1 * 8; sb[r] = size_t(PG) * cap1 * 8; ro[r] = so[r]; rb[r] = sb[r]; }
        AQ_TRY(comm_alltoallv(part1.p, so, sb, recv1.p, ro, rb));
        for (uint32_t r = 0; r < N; ++r) { so[r] = size_t(r) * PG * 4; sb[r] = size_t(PG) * 4; ro[r] = so[r]; rb[r] = sb[r]; }
        AQ_TRY(comm_alltoallv(cursor1, so, sb, rcursor1, ro, rb));
        D.exchange_bytes += size_t(N - 1) * PG * (size_t(cap1) * 8 + 4);
    }
    // ---- level 2 over the N x PG received regions (region s * PG + j feeds group j)
    AQ_LAUNCH(k_fp_part2, c.sm_count, FP_THREADS, sizeof(FpSmem), recv1.as<unsigned long long>(), rcursor1, cap1, P1, P2, cursor2,
              part2.as<unsigned long long>(), cap2, c.d_flags, PG);
    FpAggArgs A;
    A.tuples = part2.as<unsigned long long>(); A.cursor = cursor2; A.npart = nfinal; A.cap = cap2;
    A.rflags = rflags.as<unsigned char>(); A.rec_of_row = rec_of_row.as<uint32_t>();
    A.rec = rec.as<unsigned long long>(); A.rec_cap = rec_cap; A.nrec = nrec; A.flags = c.d_flags;
    A.plist = plist.as<uint32_t>(); A.nlist = nlist;
    A.cand = cand1.as<uint32_t>() + 1; A.ncand = cand1.as<uint32_t>(); A.cand_cap = cand1_cap;
    AQ_LAUNCH(k_fp_screen, (int)std::min<uint32_t>(nfinal, (uint32_t)c.sm_count * 3), FPS_THREADS, FPS_SMEM, ks, A);
    // ---- candidates, first filter: the rows' home ranks publish a second 64-bit hash; pairs that still agree go on
    {
        const uint32_t nslots = (cand1_cap + 1) * N;
        AQ_TRY(comm_allgather(cand1.p, cand1all.p, size_t(cand1_cap + 1) * 4));
        AQ_LAUNCH(k_fp_cand_hash, grid_for(nslots, 256, 1, 16), 256, 0, ks, cand1all.as<uint32_t>(), nslots, (uint32_t)D.row_base, nrows, h2.as<uint32_t>());
        AQ_TRY(comm_allreduce_sum_u32(h2.as<uint32_t>(), size_t(nslots) * 2));
        AQ_LAUNCH(k_fp_cand_filter, grid_for(cand1_cap / 2 + 1, 256, 1, 16), 256, 0, cand1.as<uint32_t>(), cand1_cap,
                  h2.as<uint32_t>() + size_t(D.rank) * (cand1_cap + 1) * 2, cand.as<uint32_t>(), cand_cap, c.d_flags);
    }
    // ---- second step: all-gather the short lists, publish keys / values of the local rows, settle exactly
    AQ_TRY(comm_allgather(cand.p, candall.p, size_t(cand_cap + 1) * 4));
    AQ_TRY(dist_exchange_flags(c.d_flags));
    std::vector<uint32_t> hc(size_t(cand_cap + 1) * N);
    AQ_CUDA(cudaMemcpyAsync(hc.data(), candall.p, hc.size() * 4, cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) return 0;                     // on every rank alike: a region or a list overflowed
    std::vector<uint32_t> list;
    for (uint32_t r = 0; r < N; ++r) {
        const uint32_t* p = hc.data() + size_t(r) * (cand_cap + 1);
        const uint32_t cnt = std::min(p[0], cand_cap);
        list.insert(list.end(), p + 1, p + 1 + cnt);
    }
    const uint32_t m = (uint32_t)list.size();
    HcIdentity ident;
    for (int a = 0; a < MAX_ACC; ++a) ident.w[a] = 0ull;
    for (int a = 0; a < plan.n_acc; ++a) {
        if (plan.prog.op[a].kind == ACC_MIN_I64) ident.w[a] = 0x7fffffffffffffffULL;
        if (plan.prog.op[a].kind == ACC_MAX_I64) ident.w[a] = 0x8000000000000000ULL;
    }
    if (m) {
        const uint32_t stride = (uint32_t)(ks.nwords + std::max(plan.prog.nv, 0));
        DevBuf dl, tab, canon;
        AQ_TRY(dl.alloc(size_t(m) * 4));
        AQ_TRY(canon.alloc(size_t(m) + 4));
        AQ_TRY(tab.alloc(size_t(m) * stride * 4));
        AQ_CUDA(cudaMemcpyAsync(dl.p, list.data(), size_t(m) * 4, cudaMemcpyHostToDevice, c.stream));
        AQ_LAUNCH(k_fp_cand_fill, grid_for(m, 256), 256, 0, ks, plan.prog, dl.as<uint32_t>(), m, (uint32_t)D.row_base, nrows, tab.as<uint32_t>(), stride);
        AQ_TRY(comm_allreduce_sum_u32(tab.as<uint32_t>(), size_t(m) * stride));
        AQ_LAUNCH(k_fp_canon, grid_for(m, 256), 256, 0, dl.as<uint32_t>(), m, canon.as<unsigned char>());
        AQ_LAUNCH(k_fp_resolve, grid_for(m, 256), 256, 0, plan.prog, ident, dl.as<uint32_t>(), canon.as<unsigned char>(), m, tab.as<uint32_t>(), stride, ks.nwords,
                  (uint32_t)D.row_base, nrows, rflags.as<unsigned char>(), rec_of_row.as<uint32_t>(), rec.as<unsigned long long>(), rec_cap, nrec, c.d_flags);
        AQ_CUDA(cudaStreamSynchronize(c.stream));                 // `list` is read by the copy above
    }
    // ---- emit: the single-GPU one over the local rows
    AQ_TRY(sink.prepare(sink.res ? nrows : sink.max_groups, plan));
    KeyOut ko;
    memset(&ko, 0, sizeof(ko));
    ko.n = kc.n;
    for (int k = 0; k < kc.n; ++k) { ko.src[k] = kc.col[k].d; ko.size[k] = (int)dtype_size(kc.type[k]); ko.dst[k] = sink.key_dst[k]; }
    for (int a = 0; a < plan.emit.n; ++a) plan.emit.agg[a].out = sink.agg_dst[a];
    FpOutCols oc;
    memset(&oc, 0, sizeof(oc));
    oc.simple = 1;
    auto add_col = [&](int conv, const void* src, void* dst) { oc.conv[oc.n] = conv; oc.src[oc.n] = src; oc.dst[oc.n] = dst; ++oc.n; };
    for (int k = 0; k < kc.n; ++k)
        if (ko.dst[k]) add_col(-1 - (ko.size[k] == 4 ? FC_COPY4 : FC_COPY8), ko.src[k], ko.dst[k]);
    if (sink.fr_dst) add_col(-1 - FC_ROW, nullptr, sink.fr_dst);
    const int nfixed = oc.n;
    for (int a = 0; a < plan.emit.n && oc.simple; ++a) {
        const EmitAgg& e = plan.emit.agg[a];
        const AccOp* o = (e.a[0] >= 0 && e.a[0] < plan.n_acc) ? &plan.prog.op[e.a[0]] : nullptr;
        const bool v4 = e.vtype == AQ_INT32 || e.vtype == AQ_FLOAT;
        int conv = -1; const void* src = nullptr;
        switch (e.op) {
        case AQ_SUM:
            if (o && o->xform == X_IDENT && v4) { src = o->v; conv = e.vtype == AQ_FLOAT ? FC_F32_F64 : (e.out_type == AQ_INT128 ? FC_I32_I128 : FC_I32_I64); }
            break;
        case AQ_AVG:
            if (o && o->xform == X_IDENT && v4) { src = o->v; conv = e.vtype == AQ_FLOAT ? FC_F32_F64 : FC_I32_F64; }
            break;
        case AQ_MIN: case AQ_MAX:
            if (o && o->xform == X_IDENT && v4) { src = o->v; conv = (e.op == AQ_MAX && e.vtype == AQ_FLOAT) ? FC_MAXF32 : FC_COPY4; }
            break;
        case AQ_COUNT: conv = e.out_type == AQ_INT64 ? FC_ONE_I64 : FC_ONE_U32; break;
        default: break;
        }
        if (conv < 0) { oc.simple = 0; break; }
        add_col(conv, src, e.out);
    }
    if (!oc.simple) oc.n = nfixed;
    uint32_t* tile_count = state.as<uint32_t>();
    uint32_t* tile_excl = tile_count + ntiles_e + 1;
    AQ_LAUNCH(k_fp_tilecount, (int)ntiles_e, FPE_THREADS, 0, nrows, rflags.as<unsigned char>(), tile_count, c.d_flags);
    AQ_TRY(exclusive_scan_u32(tile_count, tile_excl, ntiles_e, c.d_flags + FLAG_NGROUPS));
    AQ_LAUNCH(k_fp_emit, (int)ntiles_e, FPE_THREADS, 0, nrows, rflags.as<unsigned char>(), tile_excl, oc, sink.cap, c.d_flags);
    AQ_LAUNCH(k_fp_emit_rest, (int)ntiles_e, FPE_THREADS, 0, nrows, rflags.as<unsigned char>(), tile_excl, rec_of_row.as<uint32_t>(),
              rec.as<unsigned long long>(), nrec, oc.simple ? 0 : 1, plan.emit, plan.prog, sink.cap, c.d_flags);
    if (sink.fr_dst && D.row_base)
        AQ_LAUNCH(k_fp_add_base, c.sm_count * 4, 256, 0, sink.fr_dst, c.d_flags + FLAG_NGROUPS, (uint32_t)D.row_base);
    AQ_TRY(dist_exchange_flags(c.d_flags));
    AQ_CUDA(cudaMemcpyAsync(c.h_flags, c.d_flags, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    AQ_CUDA(cudaStreamSynchronize(c.stream));
    if (c.h_flags[FLAG_ABORT] != 0) return 0;
    const uint32_t ng = c.h_flags[FLAG_NGROUPS];
    if (ng > sink.cap) return fail("capacity: %u groups exceed max_groups=%u", ng, sink.cap);
    AQ_TRY(sink.finish(ng, plan));
    D.replicated = false; D.by_first_row = 1;
    *ngroups = ng;
    *used = true;
    return 0;
}

} // namespace aq
