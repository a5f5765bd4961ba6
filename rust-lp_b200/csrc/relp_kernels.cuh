// CUDA kernels (sm_100a) of the exact revised-simplex iteration: SURVEY.md section 8(a) rows a3-a10.
//
//   stage A  price        K3  pivot_rule.rs:38-127, tableau/mod.rs:102-108, carry/mod.rs:572-577,
//                             dense.rs:81-92             (fused with the block-level argmin)
//   stage C  gen_column   K2  basis_inverse_rows.rs:144-173, sparse.rs:82-106
//            + ratios     K4  tableau/mod.rs:221-247      (fused, block-level argmin)
//   stage D  pivot_row    K1a basis_inverse_rows.rs:42-51, sparse.rs:291-301, carry/mod.rs:283-295
//   stage E  update       K1b basis_inverse_rows.rs:63-83 -> sparse.rs:213-248, carry/mod.rs:296-333
//   stage F  update_big   K1c the same elements, multi-limb tier
//   K5       recompute_cost_row   carry/mod.rs:214-271 (phase hand-over)
//
// Two drivers share these stage functions:
//   * k_simplex_loop -- ONE persistent cooperative kernel (one CTA per SM) that runs whole
//     iterations back to back with grid-wide barriers between the stages; (j, r, status) never
//     leave the device.  Single-GPU solves use this.
//   * one __global__ wrapper per stage -- the step-wise trait API (relp_price, ...) and the
//     multi-GPU path, where NCCL collectives sit between the stages.
//
// HBM layout.  The Carry block of this GPU is two dense COLUMN-major int64 planes xn, xd
// (numerators / denominators): element (local row li, column k) at k * ld + li, k = 0 is b,
// k = 1 + c is column c of B^-1.  Column generation reads whole columns (coalesced); the rank-1
// update touches the T x S block of rows with d_i != 0 and columns with p_k != 0.  The cost row
// [-obj | -pi] is a separate contiguous vector zn, zd (replicated on every GPU): pricing gathers
// from it.  A denominator of 0 tags an element escalated to the multi-limb heap (relp_arith.cuh);
// heap offsets are never 0, so "numerator == 0" alone means the reference's absent sparse entry
// (sparse.rs:235-237).  A is stored twice: CSC (column generation) and SELL-32 (pricing: slice s
// holds columns 32s..32s+31 interleaved, so a warp reads 128 contiguous bytes of row indices per
// step and each lane owns one column).
#pragma once
#ifdef RELP_EMUL
#include "cuda_emul.h"   // tests/native: sequential stand-in, CPU-only CI of the host logic
#else
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#endif
#include "relp_arith.cuh"

namespace RELP_NS {

enum { RULE_FIRST = 0, RULE_FIRST_MEMORY = 1, RULE_DANTZIG = 2 };
enum { ARG_DANTZIG = 0, ARG_FIRST = 1, ARG_RATIO = 2 };
enum { ST_RUNNING = 0, ST_NO_COLUMN = 1, ST_NO_ROW = 2 };
enum { NSTAGE = 8 };

struct LogEntry { int tag, j, r, leaving, T, S; };

struct DState {
    int m, mloc, row_lo, ld;       // global rows, local rows, first local row, leading dimension
    int n, nr_art, phase, ncols;   // provider columns, artificials, 1|2, tableau columns
    int sl_lo, sl_hi;              // SELL slices priced by this rank
    int art_lo, art_hi;            // artificial columns priced by this rank
    int status, err;
    int j, r, leaving;
    int T, S, wl_count;
    int last_selected;             // FirstProfitableWithMemory (pivot_rule.rs:62-94)
    int tag;                       // log tag of the running loop
    int compacting;
    int a_int;                     // every A value has denominator 1
    int go;                        // persistent loop: continue? (written by one thread between barriers)
    int rc_valid;                  // the stored reduced-cost row (Dev::rn, rd) is current
    int W;                         // sum of the CSR row lengths over the pivot row's nonzero columns
    int wl_cap;                    // capacity of the escalation worklist
    int narrow_verdict;            // persistent loop: outcome of CTA 0's attempt at the current iteration
    long long k1_min;              // k_update: T x S from which the Carry update is tiled (dense d: staged through shared memory)
    long long k1_exit;             // persistent loop: T x S from which a dense update is handed to k_update
    int want_wide;                 // persistent loop left with the pivot row ready and the update pending
    int xflags;                    // peer mode: flags of the last exchange (written by one thread between barriers)
    unsigned long long xseq;       // peer mode: exchanges done so far (replicated)
    int narrow_skip, narrow_backoff;   // after a pivot that was too big for one CTA, do not try again for a while
    unsigned long long narrow_done;   // iterations completed by CTA 0 alone
    unsigned long long iter, iter_limit;
    unsigned long long pool_top, pool_limit, pool_base;   // active persistent region of the limb heap
    unsigned long long alt_base, alt_limit;               // the other persistent region (compaction target)
    unsigned long long tr_top, tr_base, tr_limit;         // per-iteration transient region
    unsigned long long escalated, sum_T, sum_S, sum_TS, pool_peak, compactions;
    unsigned long long stage_ns[NSTAGE];                  // persistent kernel: time per stage (globaltimer)
    Val rc;
    // (appended, so that the words the loop reads every iteration stay where they were)
    int long_seen;                 // a value longer than RELP_LONG_TRIGGER limbs has been stored (Arena::flag points here)
    int want_long;                 // persistent loop (copy without the arbitrary-length tier) left because long_seen is set
};

struct Cand { Val v; int key; int idx; };

// ---- peer exchange (multi-GPU persistent loop) ----------------------------------------------------
// Every rank owns one exchange buffer that all its peers map (CUDA IPC over NVLink).  A message is
// written INTO the receiver's buffer with plain stores, followed by a system-scope fence and a
// release store of its sequence number; the receiver polls its own memory.  Sequence numbers are
// (epoch << 40) | xseq with xseq a replicated counter that every rank advances at every exchange, so
// they never repeat and nothing has to be reset; messages go through a ring of 4 (a sender cannot be
// more than one exchange ahead of a receiver that has not read yet).
#define RELP_MAX_WORLD 8
#define RELP_XNARROW_S 64
enum { XF_TOOBIG = 1, XF_ERR = 2 };
struct XSlot { Cand c; int T; int flags; int ok, S, W, b_nz; unsigned long long seq; unsigned long long pad; };      // 64 bytes
struct XRowHdr { int ok, S, W, b_nz; unsigned long long seq; unsigned long long pad; };              // 32 bytes
struct XNarrowRow { int k[RELP_XNARROW_S]; Val p[RELP_XNARROW_S]; };
struct XHead { XSlot slot[4][RELP_MAX_WORLD]; XRowHdr rowhdr[4]; XNarrowRow nrow[4][RELP_MAX_WORLD]; unsigned long long hello[2][RELP_MAX_WORLD]; };
struct XPeer {                       // one rank's exchange buffer as mapped in THIS process
    XHead* head;
    int64_t* pn; int64_t* pd; int64_t* psn; int64_t* psd; int* slist; uint32_t* pbits;
};

struct Dev {
    DState* st;
    uint32_t* heap;
    int64_t* xn; int64_t* xd;                 // Carry planes, column-major
    int64_t* zn; int64_t* zd;                 // cost row [-obj | -pi], m + 1
    const long long* col_ptr; const int* row_idx;
    const int64_t* an; const int64_t* ad;     // CSC values
    const long long* sell_ptr; const int* sell_row;
    const int64_t* sell_an; const int64_t* sell_ad;
    const int64_t* cn; const int64_t* cd;     // cost
    int* basis;                               // basis_indices[m], tableau numbering, replicated
    unsigned char* in_basis;                  // per tableau column
    const int* art_row;                       // artificial column -> row (Partially::column_to_row)
    int64_t* dn; int64_t* dd;                 // generated column, local rows
    int64_t* pn; int64_t* pd;                 // normalised pivot row (m + 1), replicated
    int64_t* psn; int64_t* psd;               // the same values compacted in slist order (wide update: contiguous reads)
    int* tlist; int* slist;
    int2* worklist;
    Cand* cand;                               // per-block partial argmin: pricing
    Cand* cand2;                              // per-block partial argmin: ratio test
    Cand* gather;                             // world candidates (multi-GPU all-gather buffer)
    LogEntry* log; long long log_cap;
    uint32_t* zbits;                          // bit i set <=> -pi_i != 0 (maintained with the cost row)
    // Reduced-cost row, kept by the persistent loop (see stage_update_rc): one rational per tableau column
    int64_t* rn; int64_t* rd;
    const long long* row_ptr; const int4* csr_ent;  // CSR copy of A's pattern: per entry (column, its CSC start, CSC end, CSC position of this entry)
    unsigned* claim;                          // per provider column: ticket of the last reduced-cost update round that took it
    // Nonzero pattern of the Carry block, kept with it (stage_update): colbits word (k, li / 32) has bit
    // li % 32 set <=> X[li][k] != 0 (column generation tests 32 rows of a B^-1 column with one word);
    // rowbits word (li, k / 32) is the transpose (the pivot row's nonzero columns come from one
    // contiguous read, and that row of rowbits IS the bitmap "p_k != 0" of the reduced-cost update).
    uint32_t* colbits; uint32_t* rowbits;
    uint32_t* colsum;                         // per column: which blocks of 32 rows MAY be nonzero (bits are only ever set)
    const int* row_art;                       // row -> its artificial column, or -1
    unsigned long long* trace;                // debug: per-CTA arrival time at each barrier (16 slots per CTA)
    // multi-GPU persistent loop: world > 1 and peers mapped (0 / 1 otherwise)
    int xworld, xrank; unsigned long long xepoch;
    uint32_t* xbits;                          // bitmap p_k != 0 of the current pivot row (replicated copy)
    const ScratchExt* ext;                    // scratch of the cooperative multi-limb operations (slabs + locks), in device memory
    int allow_long;                           // this launch's code contains the arbitrary-length tier (else the loop leaves when long_seen)
    XPeer xp[RELP_MAX_WORLD];                 // xp[xrank] is this rank's own buffer
};

#define RELP_THREADS 256          // stage wrappers
#define RELP_UTHREADS 512         // k_update: one CTA per SM (its shared-memory ring takes most of the SM's)
#ifndef RELP_PTHREADS
#define RELP_PTHREADS 512         // persistent kernel: one CTA per SM; 512 threads leave 128 registers each
#endif
#ifndef RELP_CHP
#define RELP_CHP 8                // pricing: row indices / -pi gathers in flight per lane
#endif
#ifndef RELP_CHG
#define RELP_CHG 8                // column generation: B^-1 column reads in flight per row
#endif
#define RELP_MAXWARPS 32
#ifndef RELP_COOP_FEW
#define RELP_COOP_FEW 32              // stage_update_big with few queued elements (<= 2 per warp): operations longer than this many limbs are shared by the warp (below ~1000 bits one thread is as fast, profiles/arith_time_r02.jsonl)
#endif

struct Grid { int tid, nth, bid, nb; };
#if !defined(RELP_EMUL) && defined(RELP_WTRACE)
__device__ __forceinline__ unsigned long long gtime0() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define WTRACE(k) do { if (D.trace) atomicMax(&D.trace[(size_t)g.bid * 16 + (k)], gtime0()); } while (0)
#else
#define WTRACE(k) do {} while (0)
#endif
// loop invariants of DState, read once per kernel (every barrier invalidates L1, so re-reading
// them from global memory in every stage costs an L2 round trip each time)
struct Inv { int m, mloc, row_lo, ld, n, nr_art, phase, ncols, sl_lo, sl_hi, art_lo, art_hi, a_int, wpc, wpr, wsc; };
__device__ __forceinline__ Grid this_grid_dims() {
    Grid g; g.tid = blockIdx.x * blockDim.x + threadIdx.x; g.nth = gridDim.x * blockDim.x;
    g.bid = blockIdx.x; g.nb = gridDim.x; return g;
}

__device__ __forceinline__ Inv load_inv(const DState* st) {
    Inv I; I.m = st->m; I.mloc = st->mloc; I.row_lo = st->row_lo; I.ld = st->ld; I.n = st->n; I.nr_art = st->nr_art;
    I.phase = st->phase; I.ncols = st->ncols; I.sl_lo = st->sl_lo; I.sl_hi = st->sl_hi; I.art_lo = st->art_lo;
    I.art_hi = st->art_hi; I.a_int = st->a_int;
    I.wpc = st->ld >> 5; I.wpr = (st->m + 32) >> 5;   // words per column of colbits / per row of rowbits
    I.wsc = (I.wpc + 31) >> 5;                        // words per column of colsum
    return I;
}
__device__ __forceinline__ Arena transient(const Dev& D) {
    Arena a; a.top = &D.st->tr_top; a.limit = D.st->tr_limit; a.err = &D.st->err; a.flag = &D.st->long_seen; return a;
}
__device__ __forceinline__ Arena persistent(const Dev& D) {
    Arena a; a.top = &D.st->pool_top; a.limit = D.st->pool_limit; a.err = &D.st->err; a.flag = &D.st->long_seen; return a;
}
__device__ __forceinline__ bool halted(const DState* st) {
    return st->status != ST_RUNNING || st->err != 0 || st->iter >= st->iter_limit;
}
__device__ __forceinline__ Val ld_val(const int64_t* n, const int64_t* d, size_t i) {
    Val v; v.n = n[i]; v.d = d[i]; return v;
}
__device__ __forceinline__ void st_val(int64_t* n, int64_t* d, size_t i, const Val& v) {
    n[i] = v.n; d[i] = v.d;
}

// position in a list shared by the whole grid; one atomic per warp
__device__ __forceinline__ int list_append(int* counter) {
#ifdef RELP_EMUL
    return atomicAdd(counter, 1);
#else
    cooperative_groups::coalesced_group g = cooperative_groups::coalesced_threads();
    int base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(counter, (int)g.size());
    return g.shfl(base, 0) + (int)g.thread_rank();
#endif
}

// ---------------------------------------------------------------------------------------------
// Exact lexicographic argmin used by pricing and the ratio test.
//   ARG_DANTZIG : candidates rc_j < 0, order (rc_j, j)            pivot_rule.rs:111-127
//   ARG_FIRST   : candidates rc_j < 0, order cyclic distance from `start`   :53-57, :78-93
//   ARG_RATIO   : candidates present,  order (ratio_i, basis_indices[i])    tableau/mod.rs:221-247
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ Cand cand_none() { Cand c; c.idx = -1; c.key = 0; c.v = val_absent(); return c; }

__device__ __forceinline__ bool cand_better(const Cand& a, const Cand& b, int mode, const uint32_t* heap) {
    if (a.idx < 0) return false;
    if (b.idx < 0) return true;
    if (mode != ARG_FIRST) {
        int c = val_cmp(a.v, b.v, heap);
        if (c) return c < 0;
    }
    return a.key < b.key;
}

__device__ __forceinline__ Cand cand_shfl(const Cand& c, int delta) {
    Cand o;
    o.v.n = __shfl_down_sync(0xffffffffu, c.v.n, delta);
    o.v.d = __shfl_down_sync(0xffffffffu, c.v.d, delta);
    o.key = __shfl_down_sync(0xffffffffu, c.key, delta);
    o.idx = __shfl_down_sync(0xffffffffu, c.idx, delta);
    return o;
}

// Block-wide argmin; the result is returned to EVERY thread of the block.
// Warp argmin, result in every lane.  Most warps hold no candidate at all (a few thousand reduced
// costs of 200 000 are negative) or a single one: both cases are settled by one ballot.
__device__ __forceinline__ Cand warp_argmin(Cand best, int mode, const uint32_t* heap) {
#ifndef RELP_EMUL
    const unsigned have = __ballot_sync(0xffffffffu, best.idx >= 0);
    if (have == 0) return best;
    if ((have & (have - 1)) != 0) {
        for (int d = 16; d > 0; d >>= 1) {
            Cand o = cand_shfl(best, d);
            if (cand_better(o, best, mode, heap)) best = o;
        }
    }
    const int src = (have & (have - 1)) != 0 ? 0 : ctz32(have);
    best.v.n = __shfl_sync(0xffffffffu, best.v.n, src); best.v.d = __shfl_sync(0xffffffffu, best.v.d, src);
    best.key = __shfl_sync(0xffffffffu, best.key, src); best.idx = __shfl_sync(0xffffffffu, best.idx, src);
#endif
    return best;
}

// Block-wide argmin; the result is returned to EVERY thread of the block.  Two block barriers: the
// warp winners go through shared memory and every warp reduces them again for itself.
__device__ Cand block_argmin(Cand best, int mode, const uint32_t* heap) {
    __shared__ Cand sh[RELP_MAXWARPS];
    best = warp_argmin(best, mode, heap);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();                      // the previous call's readers are done with sh
    if (lane == 0) sh[w] = best;
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5;
    if (lane < nw) best = sh[lane]; else best = cand_none();
    return warp_argmin(best, mode, heap);
}

// Reduce `count` candidates of `src` (every block can do this redundantly: same input, same
// deterministic order, same winner).
__device__ Cand reduce_candidates(const Cand* src, int count, int mode, const uint32_t* heap) {
    Cand best = cand_none();
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        Cand c = src[i];
        if (cand_better(c, best, mode, heap)) best = c;
    }
    return block_argmin(best, mode, heap);
}

// element (li, k) of the Carry block changed between zero and nonzero
__device__ __forceinline__ void pattern_set(const Dev& D, const Inv& I, int li, int k, bool nonzero) {
    uint32_t* c = D.colbits + (size_t)k * I.wpc + (li >> 5);
    uint32_t* r = D.rowbits + (size_t)li * I.wpr + (k >> 5);
    if (nonzero) {
        atomicOr(c, 1u << (li & 31)); atomicOr(r, 1u << (k & 31));
        atomicOr(D.colsum + (size_t)k * I.wsc + (li >> 10), 1u << ((li >> 5) & 31));
    }
    else { atomicAnd(c, ~(1u << (li & 31))); atomicAnd(r, ~(1u << (k & 31))); }
}

// ---------------------------------------------------------------------------------------------
// Stage A -- K3 pricing: rc_j = c_j + sum_i (-pi_i) A_ij for every non-basic tableau column of
// this rank, fused with the argmin (no reduced-cost vector is ever stored).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ Val price_column_csc(const Dev& D, const DState* st, int j, const Arena& ar) {
    const int64_t* zn = D.zn + 1; const int64_t* zd = D.zd + 1;
    if (st->phase == 1 && j < st->nr_art) {
        // artificial column e_row with Cost::One: -pi_row + 1   (big/special.rs:36-48)
        int row = D.art_row[j];
        Val one = make_small(1, 1);
        return val_fma(ld_val(zn, zd, row), one, one, 1, D.heap, ar);
    }
    int jj = st->phase == 1 ? j - st->nr_art : j;
    Val acc = st->phase == 1 ? val_zero() : make_small(D.cn[jj], D.cd[jj]);
    long long e = D.col_ptr[jj + 1];
    for (long long t = D.col_ptr[jj]; t < e; t++) {
        int row = D.row_idx[t];
        int64_t z = zn[row];
        if (z == 0) continue;
        Val zv; zv.n = z; zv.d = zd[row];
        acc = val_fma(acc, zv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
    }
    return acc;
}

__device__ __forceinline__ void price_consider(const Dev& D, Val rc, int j, int mode, int start, int ncols, Cand& best) {
    if (val_sign(rc, D.heap) >= 0) return;
    Cand c; c.v = rc; c.idx = j;
    c.key = mode == ARG_FIRST ? (j >= start ? j - start : j - start + ncols) : j;
    if (cand_better(c, best, mode, D.heap)) best = c;
}

#ifndef RELP_EMUL
extern __shared__ uint32_t relp_dyn_smem[];
// staging ring of the dense Carry update: `off` bytes into the dynamic shared memory, < 0 = none
__device__ __forceinline__ unsigned char* k1_region(int off) {
    if (off < 0) return nullptr;
    unsigned char* p = (unsigned char*)relp_dyn_smem + off;
    return p + ((128u - ((uint32_t)__cvta_generic_to_shared(p) & 127u)) & 127u);     // bulk copies want 16-byte aligned shared addresses
}
#else
static inline unsigned char* k1_region(int) { return nullptr; }
#endif

// One pass over this rank's SELL slices, one lane per column jj.  `init(jj, acc)` says whether the
// column takes part and seeds the sum; the sum then collects v[row + voff] * a over the entries
// whose bit (row + voff) is set in `bits`; `done(jj, acc)` receives it.
// Pass 1 (streaming): every entry's row is tested against the bitmap and the hits of a lane are
// remembered as a bit mask; the row indices of the next chunk are requested before this chunk is
// consumed.  Pass 2: the lanes walk their own hits in lockstep, two at a time, so a warp pays
// (max hits per lane)/2 dependent round trips instead of one per hit position.  (A 32-lane
// gather from global memory costs 32 L1 wavefronts; measured, testing -pi itself instead of a
// shared-memory bitmap was the whole pricing time.)
template <class Init, class Done>
__device__ __forceinline__ void sell_scan(const Dev& D, const Grid& g, const Inv& I, const uint32_t* bits, int voff,
                                          const int64_t* vn, const int64_t* vd, const Arena& ar, Init init, Done done) {
    const int n = I.n;
#ifdef RELP_EMUL
    const int gw = 0, nw = 1;        // sequential stand-in: one thread plays every lane in turn
    for (int lane = 0; lane < 32; lane++)
#else
    // slices are dealt round-robin over the CTAs so that every SM's load/store unit shares the work
    const int lane = threadIdx.x & 31, gw = (threadIdx.x >> 5) * g.nb + g.bid, nw = g.nth >> 5;
#endif
    for (int s = I.sl_lo + gw; s < I.sl_hi; s += nw) {
        const int a_int = I.a_int;
        int jj = s * 32 + lane;
        long long base = D.sell_ptr[s];
        int width = (int)((D.sell_ptr[s + 1] - base) >> 5);
        Val acc = val_zero();
        bool live = jj < n && init(jj, acc);
        const int* __restrict__ rows = D.sell_row + base + lane;
        for (int seg = 0; seg < width; seg += 64) {
            const int wseg = width - seg < 64 ? width - seg : 64;
            unsigned long long hits = 0;
            int rr[RELP_CHP], nx[RELP_CHP];
#pragma unroll
            for (int u = 0; u < RELP_CHP; u++) rr[u] = (live && u < wseg) ? rows[(size_t)(seg + u) * 32] : -1;
            for (int t = 0; t < wseg; t += RELP_CHP) {
#pragma unroll
                for (int u = 0; u < RELP_CHP; u++) nx[u] = (live && t + RELP_CHP + u < wseg) ? rows[(size_t)(seg + t + RELP_CHP + u) * 32] : -1;
#pragma unroll
                for (int u = 0; u < RELP_CHP; u++) {
                    int row = rr[u] + voff;
                    if (rr[u] >= 0 && ((bits[row >> 5] >> (row & 31)) & 1u)) hits |= 1ull << (t + u);
                }
#pragma unroll
                for (int u = 0; u < RELP_CHP; u++) rr[u] = nx[u];
            }
            while (hits) {
                int ta = ctz64(hits); hits &= hits - 1;
                int tb = hits ? ctz64(hits) : -1; if (tb >= 0) hits &= hits - 1;
                size_t ea = (size_t)base + (size_t)(seg + ta) * 32 + lane;
                size_t eb = tb >= 0 ? (size_t)base + (size_t)(seg + tb) * 32 + lane : ea;
                int ra = D.sell_row[ea] + voff, rb = D.sell_row[eb] + voff;
                Val za = ld_val(vn, vd, ra), zb2 = ld_val(vn, vd, rb);
                Val aa = make_small(D.sell_an[ea], a_int ? 1 : D.sell_ad[ea]);
                Val ab = make_small(D.sell_an[eb], a_int ? 1 : D.sell_ad[eb]);
                acc = val_fma(acc, za, aa, 1, D.heap, ar);
                if (tb >= 0) acc = val_fma(acc, zb2, ab, 1, D.heap, ar);
            }
        }
        if (jj < n) done(jj, live, acc);
    }
}

// copy a bitmap into the CTA's dynamic shared memory (smem_words == 0: too large, or the CPU
// stand-in -- the kernels then test the global bitmap)
__device__ __forceinline__ const uint32_t* stage_bits(const uint32_t* global_bits, int smem_words) {
#ifndef RELP_EMUL
    if (smem_words > 0) {
        __syncthreads();
        for (int w = threadIdx.x; w < smem_words; w += blockDim.x) relp_dyn_smem[w] = global_bits[w];
        __syncthreads();
        return relp_dyn_smem;
    }
#endif
    return global_bits;
}

// Full pricing pass.  `fill`: also store every reduced cost (0 for basic columns) in the
// reduced-cost row, from where the persistent loop keeps it current (stage_update_rc).
__device__ void stage_price(const Dev& D, const Grid& g, const Inv& I, int mode, int smem_words, int fill) {
    const DState* st = D.st;
    Arena ar = fill ? persistent(D) : transient(D);
    const int phase = I.phase, nr_art = I.nr_art, ncols = I.ncols;
    const int start = (mode == ARG_FIRST && st->last_selected >= 0) ? st->last_selected : 0;
    const int off = phase == 1 ? nr_art : 0;
    const int64_t* zn = D.zn + 1; const int64_t* zd = D.zd + 1;   // mutable between stages: no __restrict__
    const uint32_t* zb = stage_bits(D.zbits, smem_words);
    Cand best = cand_none();
    if (phase == 1) {
        Val one = make_small(1, 1);
        for (int a = I.art_lo + g.tid; a < I.art_hi; a += g.nth) {
            bool basic = D.in_basis[a];
            Val rc = basic ? val_zero() : val_fma(ld_val(zn, zd, D.art_row[a]), one, one, 1, D.heap, ar);
            if (fill) st_val(D.rn, D.rd, a, rc);
            if (!basic) price_consider(D, rc, a, mode, start, ncols, best);
        }
    }
    sell_scan(D, g, I, zb, 0, zn, zd, ar,
              [&](int jj, Val& acc) {
                  if (D.in_basis[off + jj]) return false;
                  if (phase == 2) acc = make_small(D.cn[jj], D.cd[jj]);
                  return true;
              },
              [&](int jj, bool live, const Val& acc) {
                  if (fill) st_val(D.rn, D.rd, off + jj, live ? acc : val_zero());
                  if (live) price_consider(D, acc, off + jj, mode, start, ncols, best);
              });
    best = block_argmin(best, mode, D.heap);
    if (threadIdx.x == 0) D.cand[g.bid] = best;
}

// Stage A of the persistent loop: the entering column from the stored reduced-cost row.  Basic
// columns hold exactly 0 there, so "rc_j < 0" alone is the reference's filter (pivot_rule.rs:38-127).
__device__ void stage_select_rc(const Dev& D, const Grid& g, const Inv& I, int mode) {
    const DState* st = D.st;
    const int start = (mode == ARG_FIRST && st->last_selected >= 0) ? st->last_selected : 0;
    Cand best = cand_none();
    for (int j = g.tid; j < I.ncols; j += g.nth) {
        Val rc = ld_val(D.rn, D.rd, j);
        if (rc.n == 0) continue;
        price_consider(D, rc, j, mode, start, I.ncols, best);
    }
    best = block_argmin(best, mode, D.heap);
    if (threadIdx.x == 0) D.cand[g.bid] = best;
}

// Reduced-cost row after a pivot.  -pi' = -pi - rc_e p (carry/mod.rs:326-333) gives, for EVERY
// column, rc_j' = rc_j - rc_e (p . A_j) with p the normalised pivot row of B^-1: exact arithmetic
// makes this the same reduced rational the reference recomputes from scratch (tableau/mod.rs:
// 102-108), basic columns included (p . A_j is 0 for them, 1 for the entering column, so it drops
// to exactly 0).  Only columns with an entry in a row where p != 0 change.  Sparse pivot row: those
// columns are enumerated from the CSR pattern of A; otherwise one
// SELL pass with the pivot row's bitmap.  Runs in the same barrier interval as the Carry update,
// from the other end of the grid.
// rc_jj -= rc * (p . A_jj), done by the caller that came from the SMALLEST row k with p_k != 0 in
// column jj (`from_k`), so a column touched from several rows is updated exactly once, no atomics.
// `ent` = (jj, CSC start, CSC end) from the CSR copy, so the chain is csr_ent -> row indices (all of a
// segment of RELP_RCH requested together) -> values of the hits.
#ifndef RELP_RCH
#define RELP_RCH 24
#endif
__device__ __forceinline__ void rc_column_update(const Dev& D, const int4& ent, int from_k, const Val& rc, int off, const uint32_t* pbits, const Arena& ar) {
    const int jj = ent.x;
    Val old = ld_val(D.rn, D.rd, off + jj);            // requested early, used last
    Val alpha = val_zero();
    int first = 0x7fffffff;
    for (int seg = ent.y; seg < ent.z; seg += RELP_RCH) {
        const int wseg = ent.z - seg < RELP_RCH ? ent.z - seg : RELP_RCH;
        int kk[RELP_RCH];
#pragma unroll
        for (int u = 0; u < RELP_RCH; u++) kk[u] = u < wseg ? D.row_idx[seg + u] + 1 : -1;
        unsigned hits = 0;
#pragma unroll
        for (int u = 0; u < RELP_RCH; u++) if (kk[u] >= 0 && ((pbits[kk[u] >> 5] >> (kk[u] & 31)) & 1u)) {
            hits |= 1u << u;
            if (kk[u] < first) first = kk[u];
        }
        if (first < from_k) return;                    // another caller owns this column
        while (hits) {
            int u = ctz32(hits); hits &= hits - 1;
            int k = D.row_idx[seg + u] + 1;
            alpha = val_fma(alpha, ld_val(D.pn, D.pd, k), make_small(D.an[seg + u], D.ad[seg + u]), 1, D.heap, ar);
        }
    }
    if (first == from_k && alpha.n != 0)
        st_val(D.rn, D.rd, off + jj, val_fma(old, rc, alpha, -1, D.heap, ar));
}

// one column k = slist[s] of the pivot row, i.e. row k - 1 of A: one thread of the CTA per entry
__device__ __forceinline__ void rc_sparse_row(const Dev& D, const Inv& I, int s, const Val& rc, int off, const uint32_t* pbits, const Arena& ar) {
    const int k = D.slist[s];
    if (k == 0) return;
    if (threadIdx.x == 0 && I.phase == 1) {      // artificial column e_i (cost one): p . e_i = p_k
        int a = D.row_art[k - 1];
        if (a >= 0) st_val(D.rn, D.rd, a, val_fma(ld_val(D.rn, D.rd, a), rc, ld_val(D.pn, D.pd, k), -1, D.heap, ar));
    }
    const long long lo = D.row_ptr[k - 1], hi = D.row_ptr[k];
    for (long long e = lo + threadIdx.x; e < hi; e += blockDim.x) rc_column_update(D, D.csr_ent[e], k, rc, off, pbits, ar);
}

#ifndef RELP_RC_SPARSE_S
#define RELP_RC_SPARSE_S 32
#endif
__device__ void stage_update_rc(const Dev& D, const Grid& g, const Inv& I, int r, Val rc, int smem_words) {
    const DState* st = D.st;
    if (rc.n == 0) return;
    const int S = st->S, W = st->W;
    const int off = I.phase == 1 ? I.nr_art : 0;
    Arena ar = persistent(D);
    // p_k != 0 <=> X[r][k] != 0; with peers the owner of row r has sent every rank a copy of that row of rowbits
    const uint32_t* pbits = D.xworld > 1 ? D.xbits : D.rowbits + (size_t)(r - I.row_lo) * I.wpr;
    long long nnz = D.col_ptr[I.n];
    if (S <= RELP_RC_SPARSE_S && (long long)W * 32 <= nnz + 1024
#ifndef RELP_EMUL
        && S <= g.nb
#endif
        ) {
        // CTA (counted from the far end of the grid) s takes column k = slist[s] of the pivot row,
        // i.e. row k - 1 of A, one thread per entry of that row.  A column with entries in several
        // such rows is updated by the thread that came from its smallest one: no atomics, and every
        // load chain is slist -> row_ptr -> csr_col -> col_ptr -> row_idx -> pbits -> p.
#ifdef RELP_EMUL
        for (int s = 0; s < S; s++) rc_sparse_row(D, I, s, rc, off, pbits, ar);
#else
        const int s = g.nb - 1 - g.bid;
        if (s < S) rc_sparse_row(D, I, s, rc, off, pbits, ar);
#endif
        return;
    }
    if (I.phase == 1) {
        for (int s = g.tid; s < S; s += g.nth) {
            int k = D.slist[s];
            int a = k > 0 ? D.row_art[k - 1] : -1;
            if (a >= 0) st_val(D.rn, D.rd, a, val_fma(ld_val(D.rn, D.rd, a), rc, ld_val(D.pn, D.pd, k), -1, D.heap, ar));
        }
    }
    const uint32_t* pb = stage_bits(pbits, smem_words);
    sell_scan(D, g, I, pb, 1, D.pn, D.pd, ar,
              [&](int, Val&) { return true; },
              [&](int jj, bool, const Val& alpha) {
                  if (alpha.n != 0)
                      st_val(D.rn, D.rd, off + jj, val_fma(ld_val(D.rn, D.rd, off + jj), rc, alpha, -1, D.heap, ar));
              });
}

// ---------------------------------------------------------------------------------------------
// Stage C -- K2 + K4: d_i = sum_k Binv[i][k] a_kj for every local row, the list of rows with
// d_i != 0, and the exact argmin of b_i / d_i over rows with d_i > 0 (block level).
// ---------------------------------------------------------------------------------------------
struct ColumnRef { int unit_row; long long t0, t1; };
__device__ __forceinline__ ColumnRef column_ref(const Dev& D, const Inv& I, int j) {
    ColumnRef c; c.unit_row = -1; c.t0 = 0; c.t1 = 0;
    if (I.phase == 1 && j < I.nr_art) c.unit_row = D.art_row[j];
    else { int jj = I.phase == 1 ? j - I.nr_art : j; c.t0 = D.col_ptr[jj]; c.t1 = D.col_ptr[jj + 1]; }
    return c;
}

#ifndef RELP_EMUL
// d_i for the 32 rows of row block `rb` (one warp, lane = row).  Lane u fetches entry u of a_j and the
// ONE word of colbits that says which of the 32 rows are nonzero in that column of B^-1; the words
// are passed round by shuffle, so a block of rows that the column does not touch costs col_ptr ->
// row_idx -> colbits and no read of B^-1 at all.  A row's own hits are then fetched four at a time.
__device__ __forceinline__ Val gen_rows_block(const Dev& D, const Inv& I, int unit_row, const int* ents, int nent,
                                              const int64_t* an, const int64_t* ad, int rb, int lane, const Arena& ar) {
    const size_t ld = I.ld;
    const int64_t* xn = D.xn + ld;     // column 1 + k at xn + k * ld
    const int64_t* xd = D.xd + ld;
    const uint32_t* cb = D.colbits + I.wpc;          // column 1 + k at cb + k * wpc
    const int li = rb * 32 + lane;
    Val acc = val_zero();
    if (unit_row >= 0) {
        uint32_t word = cb[(size_t)unit_row * I.wpc + rb];
        if ((word >> lane) & 1u) acc = copy_val(ld_val(xn, xd, (size_t)unit_row * ld + li), D.heap, ar);
        return acc;
    }
    for (int seg = 0; seg < nent; seg += 32) {
        const int cnt = nent - seg < 32 ? nent - seg : 32;
        const int kk = lane < cnt ? ents[seg + lane] : -1;
        const uint32_t word = kk >= 0 ? cb[(size_t)kk * I.wpc + rb] : 0u;
        unsigned act = __ballot_sync(0xffffffffu, word != 0u);
        unsigned mine = 0;                      // entries of this segment that hit MY row
        while (act) {
            const int u = ctz32(act); act &= act - 1;
            const uint32_t w = __shfl_sync(0xffffffffu, word, u);
            if ((w >> lane) & 1u) mine |= 1u << u;
        }
        while (mine) {
            int uu[4]; Val xx[4], aa[4];
#pragma unroll
            for (int q = 0; q < 4; q++) { uu[q] = mine ? ctz32(mine) : -1; if (mine) mine &= mine - 1; }
#pragma unroll
            for (int q = 0; q < 4; q++) if (uu[q] >= 0) {
                xx[q] = ld_val(xn, xd, (size_t)ents[seg + uu[q]] * ld + li);
                aa[q] = make_small(an[seg + uu[q]], ad[seg + uu[q]]);
            }
#pragma unroll
            for (int q = 0; q < 4; q++) if (uu[q] >= 0) acc = val_fma(acc, xx[q], aa[q], 1, D.heap, ar);
        }
    }
    return acc;
}
#endif

// row li has d_i = acc: record it, and offer it to the ratio test when d_i > 0
__device__ __forceinline__ void gen_row_done(const Dev& D, const Inv& I, int li, const Val& acc, int* t_counter, Cand& best, const Arena& ar) {
    st_val(D.dn, D.dd, li, acc);
    if (acc.n == 0) return;
    int p = list_append(t_counter); D.tlist[p] = li;
    if (val_sign(acc, D.heap) > 0) {
        Cand c; c.v = val_ratio(ld_val(D.xn, D.xd, li), acc, D.heap, ar);
        c.idx = I.row_lo + li; c.key = D.basis[I.row_lo + li];
        if (cand_better(c, best, ARG_RATIO, D.heap)) best = c;
    }
}

__device__ void stage_gen_column(const Dev& D, const Grid& g, const Inv& I, int j) {
    DState* st = D.st;
    Arena ar = transient(D);
    const ColumnRef c = column_ref(D, I, j);
    const int mloc = I.mloc;
    Cand best = cand_none();
#ifdef RELP_EMUL
    const size_t ld = I.ld;
    const int64_t* xn = D.xn + ld; const int64_t* xd = D.xd + ld;
    const uint32_t* cb = D.colbits + I.wpc;
    for (int li = 0; li < mloc; li++) {
        Val acc = val_zero();
        if (c.unit_row >= 0) {
            if ((cb[(size_t)c.unit_row * I.wpc + (li >> 5)] >> (li & 31)) & 1u)
                acc = copy_val(ld_val(xn, xd, (size_t)c.unit_row * ld + li), D.heap, ar);
        } else for (long long t = c.t0; t < c.t1; t++) {
            int kk = D.row_idx[t];
            if ((cb[(size_t)kk * I.wpc + (li >> 5)] >> (li & 31)) & 1u)
                acc = val_fma(acc, ld_val(xn, xd, (size_t)kk * ld + li), make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
        }
        gen_row_done(D, I, li, acc, &st->T, best, ar);
    }
#else
    const int lane = threadIdx.x & 31, gw = (threadIdx.x >> 5) * g.nb + g.bid, nw = g.nth >> 5;
    for (int rb = gw; rb * 32 < mloc; rb += nw) {
        Val acc = gen_rows_block(D, I, c.unit_row, D.row_idx + c.t0, (int)(c.t1 - c.t0), D.an + c.t0, D.ad + c.t0, rb, lane, ar);
        if (rb * 32 + lane < mloc) gen_row_done(D, I, rb * 32 + lane, acc, &st->T, best, ar);
    }
#endif
    best = block_argmin(best, ARG_RATIO, D.heap);
    if (threadIdx.x == 0) D.cand2[g.bid] = best;
}

// ---------------------------------------------------------------------------------------------
// Stage D -- K1a (owner of row r): p_k = X[r][k] / d_r for k = 0..m, written back to row r and
// staged in the pivot-row buffer; list of columns with p_k != 0.
// ---------------------------------------------------------------------------------------------
// multi: 0 = one GPU; 1 = NCCL path (dense row, summed by an all-reduce); 2 = peer path (the owner
// writes the row's entries into every rank's exchange buffer, the header follows in k_simplex_loop)
__device__ void stage_pivot_row(const Dev& D, const Grid& g, const Inv& I, int r, int multi) {
    DState* st = D.st;
    int lr = r - I.row_lo;
    bool owner = lr >= 0 && lr < I.mloc;
    if (!owner) {   // NCCL path: contribute zeros to the sum all-reduce that replicates the row
        if (multi == 1) for (int k = g.tid; k <= I.m; k += g.nth) { D.pn[k] = 0; D.pd[k] = 0; }
        return;
    }
    Val pivot = ld_val(D.dn, D.dd, lr);
    if (pivot.n == 0) { if (g.tid == 0) atomicOr(&st->err, ERR_ZERO_PIVOT); return; }
    Arena ar = persistent(D);
    const size_t ld = I.ld;
    if (multi == 1) {    // the all-reduce needs p densely
        for (int k = g.tid; k <= I.m; k += g.nth) {
            size_t e = (size_t)k * ld + lr;
            int64_t x = D.xn[e];
            if (x == 0) { D.pn[k] = 0; D.pd[k] = 1; continue; }
            Val xv; xv.n = x; xv.d = D.xd[e];
            Val p = val_div(xv, pivot, D.heap, ar);
            st_val(D.xn, D.xd, e, p);
            st_val(D.pn, D.pd, k, p);
            if (p.d == 0) atomicOr(&st->err, ERR_LIMB_CAPACITY);   // escalated pivot rows do not cross GPUs yet
        }
        return;
    }
    // the row's nonzero columns come from its row of rowbits ((m + 1) / 32 contiguous words; a warp shares
    // one word); one thread per nonzero column, so a dense row of multi-limb entries is divided in parallel
    const uint32_t* rb = D.rowbits + (size_t)lr * I.wpr;
    int len = 0;
    for (int k = g.tid; k <= I.m; k += g.nth) {
        if (!((rb[k >> 5] >> (k & 31)) & 1u)) {
#ifndef RELP_EMUL
            if (multi == 2 && k == 0) for (int w = 0; w < D.xworld; w++) { D.xp[w].pn[0] = 0; D.xp[w].pd[0] = 1; }   // b_r = 0: bookkeeping reads p_0 everywhere
#endif
            continue;
        }
        size_t e = (size_t)k * ld + lr;
        Val p = val_div(ld_val(D.xn, D.xd, e), pivot, D.heap, ar);
        st_val(D.xn, D.xd, e, p);
        int s = list_append(&st->S);
#ifndef RELP_EMUL
        if (multi == 2) {
            if (p.d == 0) atomicOr(&st->err, ERR_PEER_BIG);           // a multi-limb p_k cannot cross GPUs
            for (int w = 0; w < D.xworld; w++) {                      // own buffer included: D.pn, D.slist ... point into it
                const XPeer& P = D.xp[w];
                P.pn[k] = p.n; P.pd[k] = p.d; P.slist[s] = k; P.psn[s] = p.n; P.psd[s] = p.d;
            }
        } else
#endif
        { st_val(D.pn, D.pd, k, p); D.slist[s] = k; D.psn[s] = p.n; D.psd[s] = p.d; }
        if (k > 0) len += (int)(D.row_ptr[k] - D.row_ptr[k - 1]);
    }
#ifndef RELP_EMUL
    if (multi == 2) {   // the row's bitmap (reduced-cost update), then make everything visible before the header goes out
        for (int w0 = g.tid; w0 < I.wpr; w0 += g.nth) { const uint32_t bits = rb[w0]; for (int w = 0; w < D.xworld; w++) D.xp[w].pbits[w0] = bits; }
        __threadfence_system();
    }
#endif
    if (len) atomicAdd(&st->W, len);
}

// multi-GPU: every rank derives the column list from the replicated pivot row
__device__ void stage_build_slist(const Dev& D, const Grid& g, const Inv& I) {
    DState* st = D.st;
    for (int k = g.tid; k <= I.m; k += g.nth)
        if (D.pn[k] != 0) { int s = list_append(&st->S); D.slist[s] = k; D.psn[s] = D.pn[k]; D.psd[s] = D.pd[k]; }
}

// ---------------------------------------------------------------------------------------------
// Stage E -- K1b: x_ik -= d_i * p_k over the T x S block (i != r), and z_k -= rc * p_k on the cost
// row.  Fixed-width tier in place; elements whose operands or result leave int64/int64 are queued
// for the multi-limb stage.  One thread does the basis bookkeeping (tableau/mod.rs:72-84).
// ---------------------------------------------------------------------------------------------
// x_ik -= d_i p_k for one element whose old value `x` is already in registers (fixed-width tier in
// place, else queued for the multi-limb stage).  The denominator plane is only written when the
// denominator changes: on integer data (den = 1 throughout) the update moves 24 bytes per element, not 32.
__device__ __forceinline__ void carry_elem_apply(const Dev& D, const Inv& I, int li, int k, const Val& x, const Val& dv, const Val& p) {
    DState* st = D.st;
    size_t e = (size_t)k * I.ld + li;
    Val out;
    if (x.d > 0 && dv.d > 0 && p.d > 0 && small_fma(x, dv, p, -1, &out)) {
        D.xn[e] = out.n;
        if (out.d != x.d) D.xd[e] = out.d;
        if ((x.n != 0) != (out.n != 0)) pattern_set(D, I, li, k, out.n != 0);
    } else {
        int q = atomicAdd(&st->wl_count, 1);
        if (q < st->wl_cap) D.worklist[q] = make_int2(li, k); else atomicOr(&st->err, ERR_HEAP_EXHAUSTED);
    }
}
__device__ __forceinline__ void carry_elem_update(const Dev& D, const Inv& I, int li, int k, const Val& dv, const Val& p) {
    carry_elem_apply(D, I, li, k, ld_val(D.xn, D.xd, (size_t)k * I.ld + li), dv, p);
}
// z_k -= rc p_k on the cost row [-obj | -pi] and its bitmap      (carry/mod.rs:326-333)
__device__ __forceinline__ void cost_row_update(const Dev& D, int k, const Val& rc, const Val& p, const Arena& ar) {
    Val z = val_fma(ld_val(D.zn, D.zd, k), rc, p, -1, D.heap, ar);
    st_val(D.zn, D.zd, k, z);
    if (k > 0) {
        uint32_t bit = 1u << ((k - 1) & 31);
        if (z.n != 0) atomicOr(&D.zbits[(k - 1) >> 5], bit); else atomicAnd(&D.zbits[(k - 1) >> 5], ~bit);
    }
}
// basis bookkeeping and the log entry of this pivot (tableau/mod.rs:72-84); one thread
__device__ __forceinline__ void pivot_bookkeeping(const Dev& D, int j, int r, int T, int S, int b_nz, const Val& rc, int keeps_rc) {
    DState* st = D.st;
    int leaving = D.basis[r];
    D.basis[r] = j;
    if (leaving >= 0) D.in_basis[leaving] = 0;
    D.in_basis[j] = 1;
    st->leaving = leaving; st->j = j; st->r = r;
    if (!keeps_rc) st->rc_valid = 0;     // step-wise / multi-GPU drivers price from scratch
    int Tl = T - 1;                       // reference's col.len() - 1 (rows eliminated) on this rank
    int Sl = S - b_nz + 1;                // nnz(B^-1 row r) + 1
    if ((long long)st->iter < D.log_cap) {
        LogEntry e; e.tag = st->tag; e.j = j; e.r = r; e.leaving = leaving; e.T = Tl; e.S = Sl;
        D.log[st->iter] = e;
    }
    unsigned long long rows = (unsigned long long)((T > 0 ? T - 1 : 0) + (rc.n != 0));     // T = 0: a peer rank without a touched row
    st->sum_T += rows; st->sum_S += (unsigned long long)S; st->sum_TS += rows * (unsigned long long)S;
}

#ifndef RELP_EMUL
// ---- bulk asynchronous copies (TMA, 1-D) global -> shared, completion on an mbarrier ----------
#ifndef RELP_TMA_INLINE
#define RELP_TMA_INLINE __forceinline__
#endif
#ifndef RELP_K1_STAGES
#define RELP_K1_STAGES 6           // tiles in the shared-memory ring (STAGES - 1 loads in flight, 32 KB each)
#endif
#define RELP_K1_NC 2               // columns per tile
__host__ __device__ __forceinline__ size_t k1_smem_bytes(int threads) {
    return 128 + (size_t)RELP_K1_STAGES * 2 * RELP_K1_NC * (2 * (size_t)threads) * 8;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_inval(uint64_t* b) {
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" :: "r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    const uint32_t a = smem_u32(b);
    uint32_t ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// one element of the pair a thread owns in the dense-row update: returns the new value, or the old one
// when the element went to the multi-limb queue (stage_update_big reads it from memory)
__device__ __noinline__ Val dense_elem(const Dev& D, const Inv& I, int li, int k, const Val& x, const Val& dv, const Val& p) {
    Val out;
    if (x.d > 0 && dv.d > 0 && p.d > 0 && small_fma(x, dv, p, -1, &out)) {
        if ((x.n != 0) != (out.n != 0)) pattern_set(D, I, li, k, out.n != 0);
        return out;
    }
    DState* st = D.st;
    int w = atomicAdd(&st->wl_count, 1);
    if (w < st->wl_cap) D.worklist[w] = make_int2(li, k); else atomicOr(&st->err, ERR_HEAP_EXHAUSTED);
    return x;
}

// K1 when d is dense, staged through shared memory: a tile is 2 * blockDim.x consecutive rows of
// RELP_K1_NC pivot-row columns, both planes; one thread issues the bulk copies of the tile STAGES - 1
// ahead (8 KB contiguous per column and plane at 512 threads), every thread owns two consecutive rows
// (16-byte shared loads, 16-byte global stores; a denominator pair is stored only if it changed).  The
// loads in flight are set by the ring (96 KB per SM), not by registers or occupancy.  Every CTA takes
// a contiguous run of tiles with the column chunks fastest, so d_i stays in registers.
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(b)) : "memory");
}
__device__ RELP_TMA_INLINE void update_dense_tma(const Dev& D, const Grid& g, const Inv& I, int S, int lr, unsigned char* smem) {
    const int NS = RELP_K1_STAGES, NC = RELP_K1_NC;
    const int R = blockDim.x * 2, nwarps = blockDim.x >> 5;
    uint64_t* full = (uint64_t*)smem;                      // tile landed (bulk copies complete)
    uint64_t* empty = full + NS;                           // tile consumed (one arrival per warp)
    int64_t* buf = (int64_t*)(smem + 128);                 // [NS][plane][NC][R]
    const int stage_elems = 2 * NC * R;
    const size_t ld = I.ld;
    const int ntc = ((int)ld + R - 1) / R, nsc = (S + NC - 1) / NC;
    const long long ntiles = (long long)ntc * nsc;
    const long long lo = ntiles * g.bid / g.nb, hi = ntiles * (g.bid + 1) / g.nb;
    const int count = (int)(hi - lo);                      // tiles of this CTA: (tc, sc) runs on from (lo / nsc, lo % nsc)
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], nwarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // producer state (thread 0): next tile to request
    int p_it = 0, p_tc = (int)(lo / nsc), p_sc = (int)(lo - (long long)p_tc * nsc), p_st = 0, p_use = 0;
    auto issue = [&]() {
        if (p_use > 0) mbar_wait(&empty[p_st], (uint32_t)((p_use - 1) & 1));    // every warp is done with the tile that was there
        const int row0 = p_tc * R;
        const int rows = (int)ld - row0 < R ? (int)ld - row0 : R;
        const int ncols = S - p_sc * NC < NC ? S - p_sc * NC : NC;
        const uint32_t bytes = (uint32_t)rows * 8u;
        mbar_expect_tx(&full[p_st], 2u * (uint32_t)ncols * bytes);
        int64_t* dst = buf + p_st * stage_elems;
        for (int q = 0; q < ncols; q++) {
            const size_t src = (size_t)D.slist[p_sc * NC + q] * ld + row0;
            bulk_g2s(dst + q * R, D.xn + src, bytes, &full[p_st]);
            bulk_g2s(dst + (NC + q) * R, D.xd + src, bytes, &full[p_st]);
        }
        p_it++;
        if (++p_sc == nsc) { p_sc = 0; p_tc++; }
        if (++p_st == NS) { p_st = 0; p_use++; }
    };
    if (threadIdx.x == 0) for (int i = 0; i < NS - 1 && i < count; i++) issue();
    int tc = (int)(lo / nsc), sc = (int)(lo - (long long)tc * nsc), st = 0;
    uint32_t parity = 0;
    int cur_tc = -1, li0 = 0;
    bool a0 = false, a1 = false, f0 = false, f1 = false;
    Val d0 = val_zero(), d1 = val_zero();
    for (int it = 0; it < count; it++) {
        if (threadIdx.x == 0 && p_it < count) issue();
        if (tc != cur_tc) {
            cur_tc = tc; li0 = tc * R + threadIdx.x * 2;
            longlong2 dn2 = make_longlong2(0, 0), dd2 = make_longlong2(1, 1);
            if (li0 < I.mloc) { dn2 = *(const longlong2*)(D.dn + li0); dd2 = *(const longlong2*)(D.dd + li0); }
            if (li0 == lr || li0 >= I.mloc) dn2.x = 0;
            if (li0 + 1 == lr || li0 + 1 >= I.mloc) dn2.y = 0;
            a0 = dn2.x != 0; a1 = dn2.y != 0;
            d0.n = dn2.x; d0.d = dd2.x; d1.n = dn2.y; d1.d = dd2.y;
            f0 = d0.d == 1 && d0.n == (int64_t)(int32_t)d0.n; f1 = d1.d == 1 && d1.n == (int64_t)(int32_t)d1.n;
        }
        // the tile's columns and their p_k (compacted in slist order by the pivot-row stage): independent
        // loads, requested before the wait
        const int ncols = S - sc * NC < NC ? S - sc * NC : NC;
        int kk[NC]; Val pp[NC];
#pragma unroll
        for (int q = 0; q < NC; q++) if (q < ncols) { kk[q] = D.slist[sc * NC + q]; pp[q] = ld_val(D.psn, D.psd, sc * NC + q); }
        mbar_wait(&full[st], parity);
        const int64_t* sb = buf + st * stage_elems + threadIdx.x * 2;
        longlong2 xn2[NC], xd2[NC];
#pragma unroll
        for (int q = 0; q < NC; q++) if (q < ncols) { xn2[q] = *(const longlong2*)(sb + q * R); xd2[q] = *(const longlong2*)(sb + (NC + q) * R); }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[st]);
        if (a0 || a1) {
#pragma unroll
            for (int q = 0; q < NC; q++) if (q < ncols) {
                const int k = kk[q];
                longlong2 on = xn2[q], od = xd2[q];
                // common case in line: integer element, d_i and p_k fit 32 bits -> one IMAD.WIDE, a subtraction and
                // a sign test; everything else (rationals, overflow, zero <-> nonzero) goes through dense_elem
                const bool pf = pp[q].d == 1 && pp[q].n == (int64_t)(int32_t)pp[q].n;
                if (a0) {
                    const int64_t t = (int64_t)(int32_t)d0.n * (int64_t)(int32_t)pp[q].n;
                    const int64_t r = (int64_t)((uint64_t)xn2[q].x - (uint64_t)t);
                    if (f0 && pf && xd2[q].x == 1 && xn2[q].x != 0 && r != 0 && r != INT64_MIN && ((xn2[q].x ^ t) & (xn2[q].x ^ r)) >= 0) on.x = r;
                    else { Val x; x.n = xn2[q].x; x.d = xd2[q].x; Val o = dense_elem(D, I, li0, k, x, d0, pp[q]); on.x = o.n; od.x = o.d; }
                }
                if (a1) {
                    const int64_t t = (int64_t)(int32_t)d1.n * (int64_t)(int32_t)pp[q].n;
                    const int64_t r = (int64_t)((uint64_t)xn2[q].y - (uint64_t)t);
                    if (f1 && pf && xd2[q].y == 1 && xn2[q].y != 0 && r != 0 && r != INT64_MIN && ((xn2[q].y ^ t) & (xn2[q].y ^ r)) >= 0) on.y = r;
                    else { Val x; x.n = xn2[q].y; x.d = xd2[q].y; Val o = dense_elem(D, I, li0 + 1, k, x, d1, pp[q]); on.y = o.n; od.y = o.d; }
                }
                if (on.x != xn2[q].x || on.y != xn2[q].y) *(longlong2*)(D.xn + (size_t)k * ld + li0) = on;
                if (od.x != xd2[q].x || od.y != xd2[q].y) *(longlong2*)(D.xd + (size_t)k * ld + li0) = od;
            }
        }
        if (++sc == nsc) { sc = 0; tc++; }
        if (++st == NS) { st = 0; parity ^= 1u; }
    }
    __syncthreads();
    if (threadIdx.x == 0) for (int s = 0; s < NS; s++) { mbar_inval(&full[s]); mbar_inval(&empty[s]); }
    __syncthreads();
}
#endif

// Large T x S blocks are cut into tiles of blockDim.x consecutive entries of tlist x RELP_US entries of
// slist: a thread keeps its row's d_i in registers, has RELP_US independent element loads in flight, and a
// warp reads 32 rows of one column (list_append hands a warp consecutive slots, so when d is dense they are
// 32 consecutive rows = 256 contiguous bytes per plane).  Small blocks keep one element per thread.
#ifndef RELP_US
#define RELP_US 8
#endif

// TILED = false: the persistent loop's instance (one element per thread; keeping the tiled code out of
// that kernel keeps its register allocation -- and the narrow iterations -- as they were; a pivot that is
// large enough to matter is handed to k_update, see k_simplex_loop).
template <bool TILED>
__device__ void stage_update(const Dev& D, const Grid& g, const Inv& I, int j, int r, Val rc, int keeps_rc, unsigned char* k1_smem) {
    DState* st = D.st;
    const int T = st->T, S = st->S;
    const int lr = r - I.row_lo;
    const long long total = (long long)T * S;
#ifndef RELP_EMUL
    if (TILED && k1_smem && total >= st->k1_min && 2 * T >= I.mloc) {
        // d is dense: walk the rows themselves (rows with d_i = 0 are carried along unchanged;
        // stage_gen_column wrote d_i for every local row)
        update_dense_tma(D, g, I, S, lr, k1_smem);
    } else if (TILED && total >= st->k1_min) {
        const int TT = blockDim.x;
        const int ntc = (T + TT - 1) / TT, nsc = (S + RELP_US - 1) / RELP_US;
        const size_t ld = I.ld;
        for (long long tile = g.bid; tile < (long long)ntc * nsc; tile += g.nb) {
            const int sc = (int)(tile / ntc), tc = (int)(tile - (long long)sc * ntc);
            const int t = tc * TT + threadIdx.x, s0 = sc * RELP_US;
            const int li = t < T ? D.tlist[t] : -1;
            const bool act = li >= 0 && li != lr;
            Val dv = val_zero();
            if (act) dv = ld_val(D.dn, D.dd, li);
            int kk[RELP_US]; Val xx[RELP_US], pp[RELP_US];
#pragma unroll
            for (int q = 0; q < RELP_US; q++) kk[q] = s0 + q < S ? D.slist[s0 + q] : -1;
#pragma unroll
            for (int q = 0; q < RELP_US; q++) if (act && kk[q] >= 0) xx[q] = ld_val(D.xn, D.xd, (size_t)kk[q] * ld + li);
#pragma unroll
            for (int q = 0; q < RELP_US; q++) if (kk[q] >= 0) pp[q] = ld_val(D.pn, D.pd, kk[q]);
#pragma unroll
            for (int q = 0; q < RELP_US; q++) if (act && kk[q] >= 0) carry_elem_apply(D, I, li, kk[q], xx[q], dv, pp[q]);
        }
    } else
#endif
    for (long long w = g.tid; w < total; w += g.nth) {
        int s = (int)(w / T), t = (int)(w - (long long)s * T);
        int li = D.tlist[t];
        if (li == lr) continue;
        int k = D.slist[s];
        carry_elem_update(D, I, li, k, ld_val(D.dn, D.dd, li), ld_val(D.pn, D.pd, k));
    }
    if (rc.n != 0) {   // rc == 0 only for the zero-cost clean-up pivots
        Arena ar = persistent(D);
        for (int s = g.tid; s < S; s += g.nth) { int k = D.slist[s]; cost_row_update(D, k, rc, ld_val(D.pn, D.pd, k), ar); }
    }
    if (g.tid == 0) {
        int b_nz = (lr >= 0 && lr < I.mloc) ? (D.xn[lr] != 0) : (D.pn[0] != 0);   // p_0 = b_r / pivot
        pivot_bookkeeping(D, j, r, T, S, b_nz, rc, keeps_rc);
    }
}

// Stage F -- K1c: the queued elements in the multi-limb tier.  Every warp takes 32 consecutive entries of the queue:
// elements whose operands are short (<= `coop_from` limbs, at most RELP_OPL) are done by their own lane with the
// per-thread routines; the long ones are then done one after the other by the WHOLE warp (relp_wcb.cuh: limbs across
// lanes, carries by ballot, Lehmer steps on the leading limbs) -- the warp is converged here by construction, and a
// long operation's latency is what the stage waits for when few elements are queued.
#ifdef RELP_NO_WCB
__device__ void stage_update_big(const Dev& D, const Grid& g, const Inv& I) {
#else
__device__ __noinline__ void stage_update_big(const Dev& D, const Grid& g, const Inv& I) {
#endif
    DState* st = D.st;
    const int n = st->wl_count < st->wl_cap ? st->wl_count : st->wl_cap; const size_t ld = I.ld;
    Arena ar = persistent(D);
#if defined(RELP_EMUL) || defined(RELP_NO_WCB)
    for (int w = g.tid; w < n; w += g.nth) {
        int2 it = D.worklist[w];
        size_t e = (size_t)it.y * ld + it.x;
        Val x = ld_val(D.xn, D.xd, e);
        Val out = big_fma(x, ld_val(D.dn, D.dd, it.x), ld_val(D.pn, D.pd, it.y), -1, D.heap, ar);
        st_val(D.xn, D.xd, e, out);
        if ((x.n != 0) != (out.n != 0)) pattern_set(D, I, it.x, it.y, out.n != 0);
    }
#else
    if (n == 0) return;
    const int lane = threadIdx.x & 31, nwarps = g.nth >> 5;
    // Measured (profiles/arith_time_r02.jsonl): a warp finishes ONE long operation sooner than a thread (1.2x at 1 600 bits,
    // 4-6x beyond 2 000), but a GPU full of single threads finishes 12-60x more operations per second.
    if (n > 2 * nwarps) {
        // throughput regime: every element keeps its own thread, whatever its length (big_fma: per-thread buffers up to
        // RELP_OPL limbs, the template routines on a group of one beyond)
        for (int w = g.tid; w < n; w += g.nth) {
            const int2 it = D.worklist[w];
            const size_t e = (size_t)it.y * ld + it.x;
            const Val x = ld_val(D.xn, D.xd, e);
            const Val out = big_fma(x, ld_val(D.dn, D.dd, it.x), ld_val(D.pn, D.pd, it.y), -1, D.heap, ar);
            st_val(D.xn, D.xd, e, out);
            if ((x.n != 0) != (out.n != 0)) pattern_set(D, I, it.x, it.y, out.n != 0);
        }
        return;
    }
    // latency regime (few elements for the machine: everybody would wait for the longest chain): ONE ELEMENT PER WARP, the
    // whole warp reads the same element; a long one is computed by its 32 lanes together, a short one by lane 0
    const int thr = D.ext->coop_from < RELP_COOP_FEW ? D.ext->coop_from : RELP_COOP_FEW;
    const WarpGroup wg;
    for (int w = g.tid >> 5; w < n; w += nwarps) {
        const int2 it = D.worklist[w];
        const size_t e = (size_t)it.y * ld + it.x;
        const Val x = ld_val(D.xn, D.xd, e), dv = ld_val(D.dn, D.dd, it.x), pv = ld_val(D.pn, D.pd, it.y);
        const bool lng = imax3(ref_max_limbs(x, D.heap), ref_max_limbs(dv, D.heap), ref_max_limbs(pv, D.heap)) > thr;      // warp-uniform
        Val out = val_zero();
        if (lng) out = wcb_fma(wg, x, dv, pv, -1, D.heap, ar);
        else if (lane == 0) out = big_fma(x, dv, pv, -1, D.heap, ar);
        __syncwarp();
        if (lane == 0) {
            st_val(D.xn, D.xd, e, out);
            if ((x.n != 0) != (out.n != 0)) pattern_set(D, I, it.x, it.y, out.n != 0);
        }
    }
#endif
}

// one thread, after stage F has completed everywhere
__device__ void end_of_iteration(const Dev& D) {
    DState* st = D.st;
    st->escalated += (unsigned long long)st->wl_count;
    st->iter++;
    st->go = !halted(st);
    if (st->long_seen && !D.allow_long && st->go) { st->go = 0; st->want_long = 1; }     // hand over to the copy with the full tier
    st->tr_top = st->tr_base;
    st->T = 0; st->S = 0; st->wl_count = 0; st->W = 0;
    unsigned long long used = st->pool_top - st->pool_base;
    if (used > st->pool_peak) st->pool_peak = used;
    if (used > (st->pool_limit - st->pool_base) / 2) {
        // flip regions; stage_compact re-homes every live value before anything allocates again
        unsigned long long b = st->alt_base, l = st->alt_limit;
        st->alt_base = st->pool_base; st->alt_limit = st->pool_limit;
        st->pool_base = b; st->pool_limit = l; st->pool_top = b;
        st->compacting = 1; st->compactions++;
    }
}

// Copy every live escalated Carry element into the fresh pool region (order is irrelevant).
__device__ void stage_compact(const Dev& D, const Grid& g) {
    DState* st = D.st;
    size_t total = (size_t)(st->m + 1) * st->ld;
    Arena ar = persistent(D);
    for (size_t e = g.tid; e < total; e += g.nth) {
        if (D.xd[e] == 0) {
            Val v; v.n = D.xn[e]; v.d = 0;
            Val c = copy_val(v, D.heap, ar);
            D.xn[e] = c.n; D.xd[e] = c.d;
        }
    }
    for (int k = g.tid; k <= st->m; k += g.nth) {
        if (D.zd[k] == 0) {
            Val v; v.n = D.zn[k]; v.d = 0;
            Val c = copy_val(v, D.heap, ar);
            D.zn[k] = c.n; D.zd[k] = c.d;
        }
    }
    if (st->rc_valid) for (int j = g.tid; j < st->ncols; j += g.nth) {
        if (D.rd[j] == 0) {
            Val v; v.n = D.rn[j]; v.d = 0;
            Val c = copy_val(v, D.heap, ar);
            D.rn[j] = c.n; D.rd[j] = c.d;
        }
    }
}

// commit helpers (one thread)
__device__ __forceinline__ void commit_column(DState* st, const Cand& w, int rule) {
    if (w.idx < 0) { st->status = ST_NO_COLUMN; st->last_selected = -1; }
    else { st->j = w.idx; st->rc = w.v; if (rule == RULE_FIRST_MEMORY) st->last_selected = w.idx; }
}
__device__ __forceinline__ void commit_row(DState* st, const Cand& w) {
    if (w.idx < 0) st->status = ST_NO_ROW; else st->r = w.idx;
}

#ifndef RELP_EMUL
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t;
}
#define GRID_SYNC() cooperative_groups::this_grid().sync()
#define TRACE(k) do { if (threadIdx.x == 0 && D.trace) D.trace[(size_t)g.bid * 16 + (k)] = gtime(); } while (0)
#else
static inline unsigned long long gtime() { return 0; }
#define GRID_SYNC() do {} while (0)
#define TRACE(k) do {} while (0)
#endif

#ifndef RELP_EMUL
// ---------------------------------------------------------------------------------------------
// Peer exchange primitives (see XHead).  All of them are called by every thread of ONE CTA.
// ---------------------------------------------------------------------------------------------
#ifndef RELP_PEER_TIMEOUT_NS
#define RELP_PEER_TIMEOUT_NS 2000000000ull
#endif
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// spin until *flag == want; a peer that never answers costs RELP_PEER_TIMEOUT_NS and an error, not a hang
__device__ __forceinline__ bool peer_wait(const unsigned long long* flag, unsigned long long want, DState* st) {
    if (ld_acquire_sys(flag) == want) return true;
    const unsigned long long t0 = gtime();
    unsigned spins = 0;
    while (*(const volatile unsigned long long*)flag != want)
        if ((++spins & 1023u) == 0 && gtime() - t0 > RELP_PEER_TIMEOUT_NS) { atomicOr(&st->err, ERR_PEER); return false; }
    return ld_acquire_sys(flag) == want;       // the acquire that orders the payload reads
}
__device__ __forceinline__ Cand ldcv_cand(const Cand* c) {
    Cand o;
    o.v.n = __ldcv((const long long*)&c->v.n); o.v.d = __ldcv((const long long*)&c->v.d);
    o.key = __ldcv(&c->key); o.idx = __ldcv(&c->idx);
    return o;
}

// All-gather of one ratio-test candidate per rank + (T, flags), reduced identically everywhere.
// Returns the global winner to every thread; Tmax = largest local T, any = OR of the flags.
__device__ Cand peer_exchange_cand(const Dev& D, Cand mine, int T, int flags, int& Tmax, int& any) {
    __shared__ XSlot xs[RELP_MAX_WORLD];
    DState* st = D.st;
    __syncthreads();                                    // previous readers of xs / writers of xseq are done
    const unsigned long long x = *(volatile unsigned long long*)&st->xseq + 1, seq = (D.xepoch << 40) | x;
    const int ring = (int)(x & 3), w = threadIdx.x;
    if (mine.idx >= 0 && mine.v.d == 0) flags |= XF_ERR;          // a multi-limb ratio cannot cross GPUs (handle is rank-local)
    if (st->err) flags |= XF_ERR;
    if (w < D.xworld) {
        XSlot* dst = &D.xp[w].head->slot[ring][D.xrank];
        dst->c = mine; dst->T = T; dst->flags = flags;
        st_release_sys(&dst->seq, seq);            // release: orders this thread's stores above before the flag
        const XSlot* src = &D.xp[D.xrank].head->slot[ring][w];
        XSlot got; got.c = cand_none(); got.T = 0; got.flags = XF_ERR;
        if (peer_wait(&src->seq, seq, st)) { got.c = ldcv_cand(&src->c); got.T = __ldcv(&src->T); got.flags = __ldcv(&src->flags); }
        xs[w] = got;
    }
    __syncthreads();
    if (threadIdx.x == 0) st->xseq = x;
    Cand best = cand_none(); Tmax = 0; any = 0;
    for (int q = 0; q < D.xworld; q++) {
        any |= xs[q].flags;
        if (xs[q].T > Tmax) Tmax = xs[q].T;
        if (!(xs[q].flags & XF_ERR) && cand_better(xs[q].c, best, ARG_RATIO, D.heap)) best = xs[q].c;
    }
    if ((any & XF_ERR) && threadIdx.x == 0) atomicOr(&st->err, (mine.idx >= 0 && mine.v.d == 0) ? ERR_PEER_BIG : ERR_PEER);
    return best;
}

// First launch of a solve: nobody writes into a peer's ring before every rank has arrived in the new solve
// (and so has read everything the previous one sent).
__device__ void peer_hello(const Dev& D) {
    DState* st = D.st;
    const unsigned long long seq = D.xepoch << 40;
    const int w = threadIdx.x, par = (int)(D.xepoch & 1);
    if (w < D.xworld) {
        st_release_sys(&D.xp[w].head->hello[par][D.xrank], seq);
        peer_wait(&D.xp[D.xrank].head->hello[par][w], seq, st);
    }
    __syncthreads();
}

// Pivot-row header: the owner posts (ok, S, W, b_nz) to every peer after the row's entries are on their
// way (every writer has executed a system fence); the others wait for it.  Returns the header.
__device__ XRowHdr peer_row_header(const Dev& D, bool owner, int ok, int S, int W, int b_nz) {
    __shared__ XRowHdr xh;
    DState* st = D.st;
    __syncthreads();
    const unsigned long long x = *(volatile unsigned long long*)&st->xseq + 1, seq = (D.xepoch << 40) | x;
    const int ring = (int)(x & 3), w = threadIdx.x;
    if (owner) {
        if (w < D.xworld && w != D.xrank) {
            XRowHdr* dst = &D.xp[w].head->rowhdr[ring];
            dst->ok = ok; dst->S = S; dst->W = W; dst->b_nz = b_nz;
            st_release_sys(&dst->seq, seq);
        }
        if (w == 0) { xh.ok = ok; xh.S = S; xh.W = W; xh.b_nz = b_nz; }
    } else if (w == 0) {
        const XRowHdr* src = &D.xp[D.xrank].head->rowhdr[ring];
        xh.ok = -1; xh.S = 0; xh.W = 0; xh.b_nz = 0;
        if (peer_wait(&src->seq, seq, st)) { xh.ok = __ldcv(&src->ok); xh.S = __ldcv(&src->S); xh.W = __ldcv(&src->W); xh.b_nz = __ldcv(&src->b_nz); }
    }
    __syncthreads();
    if (threadIdx.x == 0) st->xseq = x;
    return xh;
}
#endif

#ifndef RELP_EMUL
// ---------------------------------------------------------------------------------------------
// Narrow iteration.  While B^-1 is sparse a pivot touches a few dozen Carry elements and a few
// hundred reduced costs: spreading that over 148 CTAs buys nothing and costs a grid-wide barrier
// after every stage.  CTA 0 therefore runs column generation, ratio test, pivot row, Carry update
// and reduced-cost update ALONE, with block barriers in between; the rest of the grid only takes
// part in the selection scan.  Every limit is checked before anything is modified: if the pivot
// turns out to be too big for one CTA, the iteration is handed back (return 1) and redone by the
// whole grid.  Returns 0 = done, 1 = redo wide, 2 = stop (no pivot row / error).
// ---------------------------------------------------------------------------------------------
#define RELP_NARROW_BLOCKS 64       // row blocks (of 32 rows) one CTA generates
#define RELP_NARROW_ENT 1024        // entries of a_j held in shared memory
#define RELP_NARROW_T 512           // rows with d_i != 0
#define RELP_NARROW_S 64            // columns with p_k != 0
#define RELP_NARROW_TS 4096         // Carry elements one CTA updates
#define RELP_NARROW_W 8192          // reduced costs one CTA updates
#define RELP_NARROW_SUMW 256        // words of the row-block summary held in shared memory
#define RELP_NARROW_PBW 2048        // words of the pivot row's bitmap held in shared memory

#if defined(RELP_WTRACE)
#define NT(k) do { if (threadIdx.x == 0 && D.trace) { unsigned long long t_ = gtime0(); D.trace[150 * 16 + (k)] += t_ - tprev_; tprev_ = t_; } } while (0)
#else
#define NT(k) do {} while (0)
#endif

struct NarrowSmem {
    uint32_t sum[RELP_NARROW_SUMW];
    uint32_t pb[RELP_NARROW_PBW];            // p_k != 0 bitmap (copy of the pivot row's row of rowbits)
    int ent[RELP_NARROW_ENT];                // rows of a_j
    int blk[RELP_NARROW_BLOCKS];
    int tl[RELP_NARROW_T]; Val dv[RELP_NARROW_T];       // rows with d_i != 0 and d_i
    int sl[RELP_NARROW_S]; Val pv[RELP_NARROW_S];       // columns with p_k != 0 and p_k
    long long plo[RELP_NARROW_S]; int pre[RELP_NARROW_S + 1];   // CSR row of A behind column k, prefix of lengths
    int nblk, T, S, W, b_nz;
    Val pivot;
};

template <bool PEERS>
__device__ int narrow_iteration(const Dev& D, const Inv& I, const Cand& col, int smem_words) {
    __shared__ NarrowSmem sm;
#if defined(RELP_WTRACE)
    unsigned long long tprev_ = gtime0();
#endif
    DState* st = D.st;
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarps = nt >> 5;
    Grid g0; g0.tid = tid; g0.nth = nt; g0.bid = 0; g0.nb = 1;
    constexpr bool peers = PEERS;          // the same LP on several GPUs: this rank holds rows [row_lo, row_lo + mloc)
    const int j = col.idx;
    const Val rc = col.v;
    const ColumnRef c = column_ref(D, I, j);
    const int nent = c.unit_row >= 0 ? 1 : (int)(c.t1 - c.t0);
    const int wsc = I.wsc;
    // limits that depend on the LP alone: every rank decides the same
    if (wsc > RELP_NARROW_SUMW || nent > RELP_NARROW_ENT || I.wpr > RELP_NARROW_PBW) return 1;
    // (a) which blocks of 32 rows can d = B^-1 a_j touch?   row_idx -> colsum, one entry per thread
    for (int w = tid; w < wsc; w += nt) sm.sum[w] = 0;
    if (tid == 0) { sm.T = 0; sm.S = 0; sm.W = 0; sm.b_nz = 0; }
    __syncthreads();
    for (int e = tid; e < nent; e += nt) sm.ent[e] = c.unit_row >= 0 ? c.unit_row : D.row_idx[c.t0 + e];
    __syncthreads();
    for (int it = tid; it < nent * wsc; it += nt) {
        int e = it / wsc, w = it - e * wsc;
        uint32_t word = D.colsum[(size_t)(1 + sm.ent[e]) * wsc + w];
        if (word) atomicOr(&sm.sum[w], word);
    }
    __syncthreads();
    if (warp == 0) {        // compact the set bits into a list of row blocks
        int base = 0;
        for (int w0 = 0; w0 < wsc; w0 += 32) {
            uint32_t bits = w0 + lane < wsc ? sm.sum[w0 + lane] : 0u;
            int cnt = __popc(bits), incl = cnt;
            for (int d = 1; d < 32; d <<= 1) { int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
            int pos = base + incl - cnt;
            while (bits) { int bb = ctz32(bits); bits &= bits - 1; if (pos < RELP_NARROW_BLOCKS) sm.blk[pos] = (w0 + lane) * 32 + bb; pos++; }
            base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) sm.nblk = base;
    }
    __syncthreads();
    NT(1);
    // limits that depend on this rank's rows: alone, give up now; with peers, say so in the exchange
    bool toobig = sm.nblk > RELP_NARROW_BLOCKS;
    if (toobig && !peers) return 1;
    // (b) d, the rows with d_i != 0 and the ratio test; only scratch (dn, tlist) is written
    Arena tr = transient(D);
    Cand best = cand_none();
    if (!toobig) for (int q = warp; q < sm.nblk; q += nwarps) {
        const int rb = sm.blk[q], li = rb * 32 + lane;
        // b_i and the row's basis column are requested before the column is generated: only rows with
        // d_i > 0 need them, but asking now takes a round trip off the critical path
        Val bv = val_zero(); int key = 0;
        if (li < I.mloc) { bv = ld_val(D.xn, D.xd, li); key = D.basis[I.row_lo + li]; }
        Val acc = gen_rows_block(D, I, c.unit_row, sm.ent, nent, D.an + c.t0, D.ad + c.t0, rb, lane, tr);
        if (li < I.mloc && acc.n != 0) {
            st_val(D.dn, D.dd, li, acc);
            int t = list_append(&sm.T);
            if (t < RELP_NARROW_T) { sm.tl[t] = li; sm.dv[t] = acc; D.tlist[t] = li; }
            if (val_sign(acc, D.heap) > 0) {
                Cand cd; cd.v = val_ratio(bv, acc, D.heap, tr); cd.idx = I.row_lo + li; cd.key = key;
                if (cand_better(cd, best, ARG_RATIO, D.heap)) best = cd;
            }
        }
    }
    NT(2);
    Cand row = block_argmin(best, ARG_RATIO, D.heap);       // (block barriers inside)
    NT(3);
    const int T = sm.T;
    if (!peers) {
        if (row.idx < 0) { if (tid == 0) commit_row(st, row); return 2; }
        if (T > RELP_NARROW_T) return 1;
    }
    // (c) the pivot row into scratch: p_k = X[r][k] / d_r, its bitmap, and the CSR row of A behind each k.
    // With peers every rank does this for ITS OWN best row, speculatively, and sends the result along with the
    // candidate: one exchange per pivot instead of two (the row is a handful of entries here).
    Arena ar = persistent(D);
    const size_t ld = I.ld;
    int ok = 1;                                               // 1 go, 0 too big for one CTA, < 0 error
    const bool have = row.idx >= 0 && !toobig && T <= RELP_NARROW_T;
    int lr = have ? row.idx - I.row_lo : -1;
    if (have) {
        for (int t = tid; t < T; t += nt) if (sm.tl[t] == lr) sm.pivot = sm.dv[t];
        __syncthreads();
        const Val pivot = sm.pivot;
        if (pivot.n == 0) { if (!peers && tid == 0) atomicOr(&st->err, ERR_ZERO_PIVOT); ok = -1; }
        const uint32_t* rbits = D.rowbits + (size_t)lr * I.wpr;
        if (ok > 0) for (int w = tid; w < I.wpr; w += nt) {   // the row's nonzero columns first: nothing is divided for a row that is too long
            uint32_t bits = rbits[w];
            sm.pb[w] = bits;
            while (bits) {
                int k = w * 32 + ctz32(bits); bits &= bits - 1;
                int q = atomicAdd(&sm.S, 1);
                if (q < RELP_NARROW_S) sm.sl[q] = k;
            }
        }
        __syncthreads();
        const int S0 = sm.S;
        if (ok > 0 && (S0 > RELP_NARROW_S || (!peers && (long long)T * S0 > RELP_NARROW_TS))) ok = 0;
        if (ok > 0) for (int q = tid; q < S0; q += nt) {          // one thread per nonzero column
            const int k = sm.sl[q];
            Val x = ld_val(D.xn, D.xd, (size_t)k * ld + lr);
            long long lo = k > 0 ? D.row_ptr[k - 1] : 0, hi = k > 0 ? D.row_ptr[k] : 0;
            sm.pv[q] = val_div(x, pivot, D.heap, peers ? tr : ar);
            sm.plo[q] = lo; sm.pre[q] = (int)(hi - lo);
            atomicAdd(&sm.W, (int)(hi - lo));
            if (k == 0) sm.b_nz = 1;
        }
        __syncthreads();
        if (ok > 0) {
            const long long nnz = D.col_ptr[I.n];
            const bool sparse = S0 <= RELP_RC_SPARSE_S && (long long)sm.W * 32 <= nnz + 1024;
            // with peers the reduced costs must go through the CSR path: the other ranks do not hold this row's bitmap
            if (rc.n != 0 && (sparse ? sm.W > RELP_NARROW_W : (peers || (I.sl_hi - I.sl_lo) > 256))) ok = 0;
        }
    }
    if constexpr (PEERS) {
        __shared__ XSlot xs[RELP_MAX_WORLD];
        __shared__ int xwin;
        __syncthreads();
        const unsigned long long x = *(volatile unsigned long long*)&st->xseq + 1, seq = (D.xepoch << 40) | x;
        const int ring = (int)(x & 3);
        const int S0 = have && ok > 0 ? sm.S : 0;
        for (int it = tid; it < S0 * D.xworld; it += nt) {        // my row's entries into every peer's copy of my lane
            const int w = it / S0, q = it - w * S0;
            if (w == D.xrank) continue;
            if (sm.pv[q].d == 0) ok = -2;                         // a multi-limb p_k cannot cross GPUs
            XNarrowRow* dst = &D.xp[w].head->nrow[ring][D.xrank];
            dst->k[q] = sm.sl[q]; dst->p[q] = sm.pv[q];
        }
        if (__syncthreads_or(ok == -2)) ok = -2;                  // (also orders the entries before the release below)
        if (tid < D.xworld) {
            int flags = (toobig || T > RELP_NARROW_T) ? XF_TOOBIG : 0;
            if (row.idx >= 0 && row.v.d == 0) flags |= XF_ERR;    // a multi-limb ratio cannot cross GPUs either
            if (st->err) flags |= XF_ERR;
            XSlot* dst = &D.xp[tid].head->slot[ring][D.xrank];
            dst->c = row; dst->T = T; dst->flags = flags; dst->ok = have ? ok : 0; dst->S = S0; dst->W = sm.W; dst->b_nz = sm.b_nz;
            st_release_sys(&dst->seq, seq);
            const XSlot* src = &D.xp[D.xrank].head->slot[ring][tid];
            XSlot got; got.c = cand_none(); got.T = 0; got.flags = XF_ERR; got.ok = -1; got.S = 0; got.W = 0; got.b_nz = 0;
            if (peer_wait(&src->seq, seq, st)) {
                got.c = ldcv_cand(&src->c); got.T = __ldcv(&src->T); got.flags = __ldcv(&src->flags);
                got.ok = __ldcv(&src->ok); got.S = __ldcv(&src->S); got.W = __ldcv(&src->W); got.b_nz = __ldcv(&src->b_nz);
            }
            xs[tid] = got;
        }
        __syncthreads();
        if (tid == 0) {
            st->xseq = x;
            int win = -1; Cand best = cand_none();
            for (int q = 0; q < D.xworld; q++)
                if (!(xs[q].flags & XF_ERR) && cand_better(xs[q].c, best, ARG_RATIO, D.heap)) { best = xs[q].c; win = q; }
            xwin = win;
        }
        __syncthreads();
        int any = 0, Tmax = 0;
        for (int q = 0; q < D.xworld; q++) { any |= xs[q].flags; if (xs[q].T > Tmax) Tmax = xs[q].T; }
        if (any & XF_ERR) { if (tid == 0) atomicOr(&st->err, (row.idx >= 0 && row.v.d == 0) ? ERR_PEER_BIG : ERR_PEER); return 2; }
        if (any & XF_TOOBIG) return 1;
        const int win = xwin;
        if (win < 0) { if (tid == 0) commit_row(st, cand_none()); return 2; }
        row = xs[win].c;
        ok = xs[win].ok;
        if (ok == -2 && tid == 0) atomicOr(&st->err, ERR_PEER_BIG);
        if (ok > 0 && (long long)Tmax * xs[win].S > RELP_NARROW_TS) ok = 0;
        lr = row.idx - I.row_lo;
        if (ok > 0 && win != D.xrank) {                          // adopt the winner's row
            __syncthreads();
            if (tid == 0) { sm.S = xs[win].S; sm.W = xs[win].W; sm.b_nz = xs[win].b_nz; }
            const XNarrowRow* src = &D.xp[D.xrank].head->nrow[ring][win];
            for (int q = tid; q < xs[win].S; q += nt) {
                const int k = __ldcv(&src->k[q]);
                Val pq; pq.n = __ldcv((const long long*)&src->p[q].n); pq.d = __ldcv((const long long*)&src->p[q].d);
                long long lo = k > 0 ? D.row_ptr[k - 1] : 0, hi = k > 0 ? D.row_ptr[k] : 0;
                sm.sl[q] = k; sm.pv[q] = pq; sm.plo[q] = lo; sm.pre[q] = (int)(hi - lo);
            }
            __syncthreads();
        }
    }
    if (ok < 0) { if (peers && tid == 0) atomicOr(&st->err, ERR_PEER); return 2; }
    if (ok == 0) return 1;
    const int r = row.idx;
    const bool owner = lr >= 0 && lr < I.mloc;
    const int S = sm.S;
    const int W = sm.W;
    const bool rc_sparse = S <= RELP_RC_SPARSE_S && (long long)W * 32 <= D.col_ptr[I.n] + 1024;
    NT(4);
    // (d) commit: row r in place and in the pivot-row buffer, then everything that depends on p
    if (tid == 0) {
        int acc = 0;
        for (int q = 0; q < S; q++) { int len = sm.pre[q]; sm.pre[q] = acc; acc += len; }
        sm.pre[S] = acc;
        commit_row(st, row);
        st->T = T; st->S = S; st->W = W;
    }
    for (int q = tid; q < S; q += nt) {
        int k = sm.sl[q];
        if (owner) st_val(D.xn, D.xd, (size_t)k * ld + lr, sm.pv[q]);
        st_val(D.pn, D.pd, k, sm.pv[q]);
        D.slist[q] = k;
    }
    __syncthreads();
    NT(5);
    // one item space for the whole CTA: Carry elements, cost row, artificial reduced costs ...
    const int off = I.phase == 1 ? I.nr_art : 0;
    const int n_carry = T * S, n_cost = rc.n != 0 ? S : 0, n_art = (rc.n != 0 && rc_sparse && I.phase == 1) ? S : 0;
    const int n_rc = (rc.n != 0 && rc_sparse) ? W : 0;
    const int total = n_carry + n_cost + n_art;
    if (tid == nt - 1) pivot_bookkeeping(D, j, r, T, S, sm.b_nz, rc, 1);
    for (int w = tid; w < total; w += nt) {
        int it = w;
        if (it < n_carry) {
            int q = it / T, t = it - q * T;
            if (sm.tl[t] != lr) carry_elem_update(D, I, sm.tl[t], sm.sl[q], sm.dv[t], sm.pv[q]);
            continue;
        }
        it -= n_carry;
        if (it < n_cost) { cost_row_update(D, sm.sl[it], rc, sm.pv[it], ar); continue; }
        it -= n_cost;
        int k = sm.sl[it];         // artificial column e_i (cost one): p . e_i = p_k
        int a = k > 0 ? D.row_art[k - 1] : -1;
        if (a >= 0) st_val(D.rn, D.rd, a, val_fma(ld_val(D.rn, D.rd, a), rc, sm.pv[it], -1, D.heap, ar));
    }
    // ... and the structural reduced costs: one thread per entry (k - 1, jj) of the rows of A behind the
    // nonzero p_k, each holding its own term p_k a_kj.  A column met from several rows is taken by one
    // thread per round (ticket exchange on claim[jj]); the others retry in the next round, after the
    // block barrier, so the read-modify-write of rc_jj is never concurrent.  Exact arithmetic makes the
    // order of the terms irrelevant.
    {
        const unsigned base = ((unsigned)st->iter + 1u) << 10;
        unsigned rid = 0;
        for (int w0 = 0; w0 < n_rc; w0 += nt) {
            const int it = w0 + tid;
            bool pending = it < n_rc;
            int jj = 0; Val term = val_zero(), old = val_zero();
            if (pending) {
                int q = 0;
                while (q + 1 < S && sm.pre[q + 1] <= it) q++;
                const int4 ent = D.csr_ent[sm.plo[q] + (it - sm.pre[q])];
                jj = ent.x;
                old = ld_val(D.rn, D.rd, off + jj);
                term = val_fma(val_zero(), sm.pv[q], make_small(D.an[ent.w], I.a_int ? 1 : D.ad[ent.w]), 1, D.heap, ar);
            }
            bool first_round = true;
            while (__syncthreads_or(pending)) {
                if (pending) {
                    const unsigned ticket = base + rid;
                    if (atomicExch(&D.claim[jj], ticket) != ticket) {
                        if (!first_round) old = ld_val(D.rn, D.rd, off + jj);
                        st_val(D.rn, D.rd, off + jj, val_fma(old, rc, term, -1, D.heap, ar));
                        pending = false;
                    }
                }
                first_round = false;
                rid++;
            }
        }
    }
    NT(6);
    if (rc.n != 0 && !rc_sparse) stage_update_rc(D, g0, I, r, rc, smem_words);
    __syncthreads();
    NT(7);
    if (st->wl_count > 0) { stage_update_big(D, g0, I); __syncthreads(); }
    NT(8);
    return 0;
}
#endif

// ---------------------------------------------------------------------------------------------
// The persistent loop: phase_one::primal / phase_two::primal (phase_one.rs:135-169,
// phase_two.rs:32-50) entirely on the device.  Cooperative launch, one CTA per SM.
// ---------------------------------------------------------------------------------------------
// PEERS: the multi-GPU instance (exchanges with the other ranks compiled in); the single-GPU instance carries none of it.
template <bool PEERS>
__global__ void __launch_bounds__(RELP_PTHREADS, 1) k_simplex_loop(Dev D, int rule, int max_iters, int smem_words, int use_narrow) {
    DState* st = D.st;
    const Grid g = this_grid_dims();
    const Inv I = load_inv(st);
    const int mode = rule == RULE_DANTZIG ? ARG_DANTZIG : ARG_FIRST;
    // The reduced-cost row is computed once (full pricing pass) and then kept current by every
    // pivot; `rc_valid` is cleared by whatever replaces the cost row or the basis from outside.
#ifndef RELP_EMUL
    if (PEERS && st->xseq == 0) {
        if (g.bid == 0) peer_hello(D);
        GRID_SYNC();
        if (st->err) return;
    }
#endif
    if (!st->rc_valid) {
        stage_price(D, g, I, mode, smem_words, 1);
        GRID_SYNC();
        if (g.tid == 0) st->rc_valid = 1;
    }
    unsigned long long t0 = gtime();
    for (int it = 0; it < max_iters; it++) {
        TRACE(0);
        // Selection reads only the reduced-cost row; the previous iteration's bookkeeping (one
        // thread: end_of_iteration) overlaps it and is published by the same barrier.
        stage_select_rc(D, g, I, mode);
        TRACE(1);
        GRID_SYNC();
        // `go` and `compacting` were written by one thread before that barrier (or by the host
        // before the launch): every thread reads the same value here
        if (!st->go) break;
        if (st->compacting) {
            stage_compact(D, g);
            GRID_SYNC();
            if (g.tid == 0) st->compacting = 0;
        }
        unsigned long long t1 = gtime();
        TRACE(2);
        Cand col = cand_none();
        bool narrow_failed = false;
        unsigned long long t2 = t1, t3 = t1, t4 = t1;
#ifndef RELP_EMUL
        if (use_narrow && st->narrow_skip == 0) {
            // only CTA 0 needs the entering column now; the others learn it after the next barrier
            if (g.bid == 0) {
                col = reduce_candidates(D.cand, g.nb, mode, D.heap);
                if (g.tid == 0) commit_column(st, col, rule);
                int verdict = col.idx < 0 ? 2 : narrow_iteration<PEERS>(D, I, col, smem_words);
                if (threadIdx.x == 0) {
                    st->narrow_verdict = verdict;
                    if (verdict == 0) { t2 = t3 = t4 = gtime(); }
                }
            }
            TRACE(4);
            GRID_SYNC();
            TRACE(5);
            const int verdict = st->narrow_verdict;
            if (verdict == 2) break;
            if (verdict == 0) {
                if (g.tid == 0) {
                    st->narrow_done++; st->narrow_backoff = 0;
                    end_of_iteration(D);
                    unsigned long long t5 = gtime();
                    st->stage_ns[0] += t1 - t0; st->stage_ns[5] += t4 - t1; st->stage_ns[4] += t5 - t4;
                    t0 = t5;
                }
                continue;
            }
            if (g.bid != 0) { col.idx = st->j; col.v = st->rc; col.key = 0; }
            narrow_failed = true;
            // verdict 1: nothing was modified; the whole grid redoes this iteration
        } else
#endif
        {
            col = reduce_candidates(D.cand, g.nb, mode, D.heap);      // same winner in every CTA
            if (g.tid == 0) commit_column(st, col, rule);
            if (col.idx < 0) break;
        }
        TRACE(3);
        stage_gen_column(D, g, I, col.idx);
        GRID_SYNC();
        t2 = gtime();
        Cand row;
#ifndef RELP_EMUL
        if constexpr (PEERS) {
            // this rank's winner against the other ranks' (CTA 0 talks to the peers, the grid learns the result
            // through gather[0] after one more barrier)
            if (g.bid == 0) {
                Cand mine = reduce_candidates(D.cand2, g.nb, ARG_RATIO, D.heap);
                int Tmax = 0, any = 0;
                Cand win = peer_exchange_cand(D, mine, st->T, 0, Tmax, any);
                if (threadIdx.x == 0) { D.gather[0] = win; st->xflags = any; }
            }
            GRID_SYNC();
            if (st->xflags & XF_ERR) break;
            row = D.gather[0];
        } else
#endif
        row = reduce_candidates(D.cand2, g.nb, ARG_RATIO, D.heap);
        if (g.tid == 0) commit_row(st, row);
        if (row.idx < 0) break;
        TRACE(6);
        stage_pivot_row(D, g, I, row.idx, PEERS ? 2 : 0);
        TRACE(7);
        GRID_SYNC();
#ifndef RELP_EMUL
        if constexpr (PEERS) {
            // the owner's entries are on their way (each writer fenced at system scope before the barrier
            // above); its CTA 0 now posts the header, every other rank waits for it and adopts S and W
            if (g.bid == 0) {
                const int lr = row.idx - I.row_lo;
                const bool owner = lr >= 0 && lr < I.mloc;
                const XRowHdr h = peer_row_header(D, owner, st->err ? -1 : 1, st->S, st->W, 0);
                if (threadIdx.x == 0) {
                    if (h.ok < 0) atomicOr(&st->err, ERR_PEER);
                    else if (!owner) { st->S = h.S; st->W = h.W; }
                }
            }
            GRID_SYNC();
            if (st->err) break;
        }
#endif
        TRACE(8);
        t3 = gtime();
#ifndef RELP_EMUL
        // A dense update of this size is bandwidth work for the whole GPU: leave the loop and let the host
        // run it as k_update (shared-memory ring fed by bulk copies), then come back.  T and S are final
        // after the barrier above, so every thread takes the same branch.
        if ((long long)st->T * st->S >= st->k1_exit && 2 * st->T >= I.mloc) {
            if (g.tid == 0) {
                st->want_wide = 1;
                if (narrow_failed) {
                    int bo = st->narrow_backoff ? st->narrow_backoff * 2 : 2;
                    st->narrow_backoff = bo > 64 ? 64 : bo; st->narrow_skip = st->narrow_backoff;
                } else if (st->narrow_skip > 0) st->narrow_skip--;
            }
            break;
        }
#endif
        stage_update<false>(D, g, I, col.idx, row.idx, col.v, 1, nullptr);
        stage_update_rc(D, g, I, row.idx, col.v, smem_words);
        TRACE(9);
        GRID_SYNC();
        t4 = gtime();
        TRACE(10);
        if (st->wl_count > 0) {                 // final after the barrier: uniform
            stage_update_big(D, g, I);
            GRID_SYNC();
        }
        if (g.tid == 0) {
            if (narrow_failed) {           // too big for one CTA: stay wide for a while, longer each time
                int bo = st->narrow_backoff ? st->narrow_backoff * 2 : 2;
                st->narrow_backoff = bo > 64 ? 64 : bo; st->narrow_skip = st->narrow_backoff;
            } else if (st->narrow_skip > 0) st->narrow_skip--;
            end_of_iteration(D);
            unsigned long long t5 = gtime();
            st->stage_ns[0] += t1 - t0; st->stage_ns[1] += t2 - t1; st->stage_ns[2] += t3 - t2;
            st->stage_ns[3] += t4 - t3; st->stage_ns[4] += t5 - t4;
            t0 = t5;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// One wrapper per stage: step-wise trait API and the multi-GPU path (NCCL between the stages).
// ---------------------------------------------------------------------------------------------
__global__ void k_begin(Dev D) {      // housekeeping that the persistent loop does inline
    DState* st = D.st;
    if (halted(st)) return;
    if (st->compacting) stage_compact(D, this_grid_dims());
}
__global__ void k_begin2(Dev D) { if (!halted(D.st)) D.st->compacting = 0; }
__global__ void k_price(Dev D, int mode, int smem_words) {
    if (halted(D.st)) return;
    stage_price(D, this_grid_dims(), load_inv(D.st), mode, smem_words, 0);
}
// reduce `count` candidates (this rank's blocks, or the all-gather buffer) and either commit the
// decision or post this rank's winner to gather[slot] (multi-GPU stage 1)
__global__ void k_select(Dev D, int mode, int count, int from_gather, int slot, int rule) {
    DState* st = D.st;
    if (halted(st)) return;
    const Cand* src = from_gather ? D.gather : (mode == ARG_RATIO ? D.cand2 : D.cand);
    Cand w = reduce_candidates(src, count, mode, D.heap);
    if (threadIdx.x != 0) return;
    if (slot >= 0) { D.gather[slot] = w; return; }
    if (mode == ARG_RATIO) commit_row(st, w); else commit_column(st, w, rule);
}
__global__ void k_gen_column(Dev D) {
    if (halted(D.st)) return;
    stage_gen_column(D, this_grid_dims(), load_inv(D.st), D.st->j);
}
__global__ void k_pivot_row(Dev D, int multi) {
    if (halted(D.st)) return;
    stage_pivot_row(D, this_grid_dims(), load_inv(D.st), D.st->r, multi);
}
__global__ void k_build_slist(Dev D) {
    if (halted(D.st)) return;
    stage_build_slist(D, this_grid_dims(), load_inv(D.st));
}
__global__ void __launch_bounds__(RELP_UTHREADS, 1) k_update(Dev D, int k1_off) {
    if (halted(D.st)) return;
    stage_update<true>(D, this_grid_dims(), load_inv(D.st), D.st->j, D.st->r, D.st->rc, 0, k1_region(k1_off));
}
__global__ void k_update_big(Dev D) {
    if (halted(D.st)) return;
    stage_update_big(D, this_grid_dims(), load_inv(D.st));
}
__global__ void k_iter_end(Dev D) {
    if (halted(D.st)) return;
    D.st->want_wide = 0;
    end_of_iteration(D);
}

// ---------------------------------------------------------------------------------------------
// Set-up and hand-over kernels (run once per solve)
// ---------------------------------------------------------------------------------------------
// B^-1 = I, b = rhs (local rows).   carry/mod.rs:381-426, basis_inverse_rows.rs:97-101
__global__ void k_init_planes(Dev D, const int64_t* rhs_n, const int64_t* rhs_d) {
    DState* st = D.st;
    size_t ld = st->ld, total = (size_t)(st->m + 1) * ld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int k = (int)(e / ld), li = (int)(e - (size_t)k * ld);
        int64_t n = 0, d = 1;
        if (li < st->mloc) {
            int i = st->row_lo + li;
            if (k == 0) { n = rhs_n[i]; d = rhs_d[i]; }
            else if (k == i + 1) n = 1;
        }
        D.xn[e] = n; D.xd[e] = d;
    }
}

// Nonzero pattern of the planes, rebuilt whenever they are written wholesale (set-up, row removal).
__global__ void k_build_pattern(Dev D) {
    DState* st = D.st;
    const size_t ld = st->ld; const int wpc = st->ld >> 5, wpr = (st->m + 32) >> 5, m = st->m;
    size_t nthreads = (size_t)gridDim.x * blockDim.x, tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (size_t w = tid; w < (size_t)(m + 1) * wpc; w += nthreads) {
        int k = (int)(w / wpc), wb = (int)(w - (size_t)k * wpc);
        uint32_t bits = 0;
        for (int b = 0; b < 32; b++) if (D.xn[(size_t)k * ld + wb * 32 + b] != 0) bits |= 1u << b;
        D.colbits[w] = bits;
    }
    const int wsc = (wpc + 31) >> 5;
    for (size_t w = tid; w < (size_t)(m + 1) * wsc; w += nthreads) {
        int k = (int)(w / wsc), ws = (int)(w - (size_t)k * wsc);
        uint32_t bits = 0;
        for (int b = 0; b < 32; b++) {
            int rb = ws * 32 + b;
            if (rb >= wpc) break;
            bool any = false;
            for (int q = 0; q < 32 && !any; q++) any = D.xn[(size_t)k * ld + rb * 32 + q] != 0;
            if (any) bits |= 1u << b;
        }
        D.colsum[w] = bits;
    }
    for (size_t w = tid; w < (size_t)ld * wpr; w += nthreads) {
        int kw = (int)(w / ld), li = (int)(w - (size_t)kw * ld);      // consecutive threads: consecutive rows
        uint32_t bits = 0;
        for (int b = 0; b < 32; b++) { int k = kw * 32 + b; if (k <= m && D.xn[(size_t)k * ld + li] != 0) bits |= 1u << b; }
        D.rowbits[(size_t)li * wpr + kw] = bits;
    }
}

// Pattern of the initial planes (B^-1 = I, b = rhs) without reading them back: the bitmaps are zeroed by
// the host and this kernel sets one bit per row in each (plus the b column's).
__global__ void k_init_pattern(Dev D, const int64_t* rhs_n) {
    DState* st = D.st;
    const int wpc = st->ld >> 5, wpr = (st->m + 32) >> 5, wsc = (wpc + 31) >> 5;
    for (int li = blockIdx.x * blockDim.x + threadIdx.x; li < st->mloc; li += gridDim.x * blockDim.x) {
        const int k = st->row_lo + li + 1;
        uint32_t rowword0 = 0;
        atomicOr(&D.colbits[(size_t)k * wpc + (li >> 5)], 1u << (li & 31));
        atomicOr(&D.colsum[(size_t)k * wsc + (li >> 10)], 1u << ((li >> 5) & 31));
        if (rhs_n[st->row_lo + li] != 0) {
            atomicOr(&D.colbits[li >> 5], 1u << (li & 31));
            atomicOr(&D.colsum[li >> 10], 1u << ((li >> 5) & 31));
            rowword0 = 1u;
        }
        if ((k >> 5) == 0) D.rowbits[(size_t)li * wpr] = rowword0 | (1u << k);
        else { D.rowbits[(size_t)li * wpr] = rowword0; D.rowbits[(size_t)li * wpr + (k >> 5)] = 1u << (k & 31); }
    }
}

// cost row of the artificial tableau: -pi_i = -1 on artificial rows, -obj = -sum of their b_i
__global__ void k_init_cost_row_artificial(Dev D, const int64_t* rhs_n, const int64_t* rhs_d) {
    DState* st = D.st;
    // -pi: -1 on artificial rows, 0 elsewhere (every thread); -obj: exact sum, one thread
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < st->m; i += gridDim.x * blockDim.x) {
        D.zn[1 + i] = D.row_art[i] >= 0 ? -1 : 0; D.zd[1 + i] = 1;
    }
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    Arena ar = persistent(D);
    Val acc = val_zero();
    Val one = make_small(1, 1);
    for (int a = 0; a < st->nr_art; a++) {
        int i = D.art_row[a];
        acc = val_fma(acc, make_small(rhs_n[i], rhs_d[i]), one, -1, D.heap, ar);
    }
    st_val(D.zn, D.zd, 0, acc);
}

// bitmap of nonzero -pi entries, rebuilt whenever the cost row is replaced wholesale
__global__ void k_build_zbits(Dev D) {
    DState* st = D.st;
    int nwords = (st->m + 31) >> 5;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += gridDim.x * blockDim.x) {
        uint32_t bits = 0;
        for (int b = 0; b < 32; b++) { int i = w * 32 + b; if (i < st->m && D.zn[1 + i] != 0) bits |= 1u << b; }
        D.zbits[w] = bits;
    }
}

// K5: z_k = -sum_i c_B(i) X[i][k] over this rank's rows, for k = 0..m (column 0 gives -obj).
// `cost_rows` lists local rows whose basic column has a nonzero cost.   carry/mod.rs:214-271
__global__ void k_recompute_cost_row(Dev D, const int* cost_rows, int n_cost_rows, int64_t* outn, int64_t* outd) {
    DState* st = D.st;
    Arena ar = persistent(D);
    const size_t ld = st->ld;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= st->m; k += gridDim.x * blockDim.x) {
        Val acc = val_zero();
        for (int q = 0; q < n_cost_rows; q++) {
            int li = cost_rows[q];
            size_t e = (size_t)k * ld + li;
            int64_t xn = D.xn[e];
            if (xn == 0) continue;
            int jb = D.basis[st->row_lo + li];
            Val xv; xv.n = xn; xv.d = D.xd[e];
            acc = val_fma(acc, xv, make_small(D.cn[jb], D.cd[jb]), -1, D.heap, ar);
        }
        st_val(outn, outd, k, acc);
    }
}

// Artificial clean-up search (phase_one.rs:236-243): first non-artificial non-basic column with
// relative cost exactly zero and a nonzero element in `pivot_row` of B^-1 A.
// `any_cost` = 0: the reference's rule (phase_one.rs:239-249: only columns whose relative cost is zero are tried);
// 1: any non-basic column with a nonzero element in the row (the pivot is at zero level either way)
__global__ void k_cleanup_select(Dev D, int pivot_row, int* best_j, int any_cost) {
    DState* st = D.st;
    if (st->err) return;
    Arena ar = transient(D);
    int lr = pivot_row - st->row_lo;
    bool owner = lr >= 0 && lr < st->mloc;
    const size_t ld = st->ld;
    for (int j = st->nr_art + blockIdx.x * blockDim.x + threadIdx.x; j < st->ncols; j += gridDim.x * blockDim.x) {
        if (D.in_basis[j]) continue;
        if (!any_cost) {
            Val rc = price_column_csc(D, st, j, ar);
            if (rc.n != 0) continue;
        }
        if (!owner) continue;
        int jj = j - st->nr_art;
        Val acc = val_zero();
        for (long long t = D.col_ptr[jj]; t < D.col_ptr[jj + 1]; t++) {
            size_t e = (size_t)(1 + D.row_idx[t]) * ld + lr;
            int64_t xn = D.xn[e];
            if (xn == 0) continue;
            Val xv; xv.n = xn; xv.d = D.xd[e];
            acc = val_fma(acc, xv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
        }
        if (acc.n != 0) atomicMin(best_j, j);
    }
}

// one generated element (basis_inverse_rows.rs:157-173)
__global__ void k_generate_element(Dev D, int i, int j, int64_t* outn, int64_t* outd) {
    DState* st = D.st;
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    Arena ar = transient(D);
    int li = i - st->row_lo;
    const size_t ld = st->ld;
    Val acc = val_zero();
    if (st->phase == 1 && j < st->nr_art) acc = ld_val(D.xn, D.xd, (size_t)(1 + D.art_row[j]) * ld + li);
    else {
        int jj = st->phase == 1 ? j - st->nr_art : j;
        for (long long t = D.col_ptr[jj]; t < D.col_ptr[jj + 1]; t++) {
            Val xv = ld_val(D.xn, D.xd, (size_t)(1 + D.row_idx[t]) * ld + li);
            if (xv.n == 0) continue;
            acc = val_fma(acc, xv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
        }
    }
    outn[0] = acc.n; outd[0] = acc.d;
}

__global__ void k_reprice_selected(Dev D) {   // rc of st->j for the step-wise API (relp_generate_column)
    DState* st = D.st;
    Arena ar = transient(D);
    st->rc = price_column_csc(D, st, st->j, ar);
}

// Warm start (Carry::from_basis, carry/mod.rs:428-463): pivot row for the column being brought in = the
// smallest row with d_i != 0 that no earlier column took.  One block.
__global__ void k_pick_unused_row(Dev D, int* taken, int* where, int slot) {
    DState* st = D.st;
    if (halted(st)) return;
    __shared__ int best;
    if (threadIdx.x == 0) best = 0x7fffffff;
    __syncthreads();
    for (int t = threadIdx.x; t < st->T; t += blockDim.x) { int li = D.tlist[t]; if (!taken[li]) atomicMin(&best, li); }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (best == 0x7fffffff) st->status = ST_NO_ROW;       // the columns are linearly dependent
        else { st->r = st->row_lo + best; taken[best] = 1; where[slot] = best; }
    }
}

// Row/column removal (basis_inverse_rows.rs:194-205, carry/mod.rs:650-689): gather the kept rows
// and columns of the column-major planes into fresh planes.
__global__ void k_gather_planes(const int64_t* sn, const int64_t* sd, int sld,
                                int64_t* tn, int64_t* td, int tld, int trows, int tcols,
                                const int* keep_row, const int* keep_col) {
    size_t total = (size_t)tcols * tld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int k = (int)(e / tld), i = (int)(e - (size_t)k * tld);
        int64_t n = 0, d = 1;
        if (i < trows) { size_t s = (size_t)keep_col[k] * sld + keep_row[i]; n = sn[s]; d = sd[s]; }
        tn[e] = n; td[e] = d;
    }
}
// values that would have to cross GPUs must be in the int64 tier: a denominator 0 is a rank-local heap handle
__global__ void k_flag_multi_limb(DState* st, const int64_t* d, long long count) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < count; e += (long long)gridDim.x * blockDim.x)
        if (d[e] == 0) { atomicOr(&st->err, ERR_PEER_BIG); return; }
}
__global__ void k_gather_vec(const int64_t* sn, const int64_t* sd, int64_t* tn, int64_t* td, int count, const int* keep) {
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        tn[k] = sn[keep[k]]; td[k] = sd[keep[k]];
    }
}

// Export a block of rats (element (i, k) at i * rstride + k * cstride): escalated values are
// copied to the end of the transient region so that one D2H copy of [export_base, tr_top) carries
// all limbs; handles become offsets from export_base.  Output is row-major nrows x ncols.
__global__ void k_export(Dev D, const int64_t* sn, const int64_t* sd, long long rstride, long long cstride,
                         int nrows, int ncols, int64_t* on, int64_t* od, unsigned long long export_base) {
    Arena ar = transient(D);
    size_t total = (size_t)nrows * ncols;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int i = (int)(e / ncols), k = (int)(e - (size_t)i * ncols);
        Val v = ld_val(sn, sd, (size_t)((long long)i * rstride + (long long)k * cstride));
        if (v.d == 0) { v = copy_val(v, D.heap, ar); if (v.d == 0) v.n -= (int64_t)export_base; }
        on[e] = v.n; od[e] = v.d;
    }
}

}  // namespace RELP_NS
