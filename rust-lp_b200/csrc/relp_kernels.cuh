// CUDA kernels (sm_100a) of the exact revised-simplex iteration: SURVEY.md section 8(a) rows a3-a10.
//
//   K3 k_price + k_argmin_*     pricing          pivot_rule.rs:38-127, tableau/mod.rs:102-108,
//                                                carry/mod.rs:572-577, dense.rs:81-92
//   K2 k_gen_column             B^-1 a_j         basis_inverse_rows.rs:144-173, sparse.rs:82-106
//   K4 (fused into K2) + argmin ratio test       tableau/mod.rs:221-247
//   K1 k_pivot_row, k_update,   Carry update     carry/mod.rs:283-333, 549-570,
//      k_update_big                              basis_inverse_rows.rs:42-83, sparse.rs:213-301
//   K5 k_recompute_cost_row     phase hand-over  carry/mod.rs:214-271
//
// Control is device resident: every kernel reads (j, r, status, list sizes) from DState, and is a
// no-op once the loop has terminated, so the host can enqueue a batch of iterations (or replay a
// CUDA graph) without a round trip per pivot.
//
// HBM layout.  The Carry matrix is two dense row-major int64 planes xn (numerators) and xd
// (denominators), (mloc + 1) rows x ld columns per GPU: local row li < mloc is
// [b_i | B^-1 row i] for global row i = row_lo + li, local row mloc is the replicated cost row
// [-obj | -pi].  A denominator of 0 tags an element escalated to the multi-limb heap
// (relp_arith.cuh).  Zero (0/1) is the reference's "absent" sparse entry (sparse.rs:235-237).
#pragma once
#ifdef RELP_EMUL
#include "cuda_emul.h"   // tests/native: sequential stand-in, CPU-only CI of the host logic
#else
#include <cuda_runtime.h>
#endif
#include "relp_arith.cuh"

namespace relp {

enum { RULE_FIRST = 0, RULE_FIRST_MEMORY = 1, RULE_DANTZIG = 2 };
enum { ARG_DANTZIG = 0, ARG_FIRST = 1, ARG_RATIO = 2, ARG_CLEANUP = 3 };
enum { ST_RUNNING = 0, ST_NO_COLUMN = 1, ST_NO_ROW = 2 };

struct LogEntry { int tag, j, r, leaving, T, S; };

struct DState {
    int m, mloc, row_lo, ld;       // global rows, local rows, first local row, leading dimension
    int n, nr_art, phase, ncols;   // provider columns, artificials, 1|2, tableau columns
    int col_lo, col_hi;            // tableau columns priced by this rank
    int status, err;
    int j, r, leaving;
    int T, S, wl_count;
    int last_selected;             // FirstProfitableWithMemory (pivot_rule.rs:62-94)
    int tag;                       // log tag of the running loop
    int compacting;
    unsigned long long iter, iter_limit;
    unsigned long long pool_top, pool_limit;     // active persistent region of the limb heap
    unsigned long long pool_base;
    unsigned long long alt_base, alt_limit;      // the other persistent region (compaction target)
    unsigned long long tr_top, tr_base, tr_limit;  // per-iteration transient region
    unsigned long long escalated, sum_T, sum_S, sum_TS, pool_peak, compactions;
    Val rc, pivot;
};

struct Cand { Val v; int key; int idx; };

struct Dev {
    DState* st;
    uint32_t* heap;
    int64_t* xn; int64_t* xd;                 // Carry planes
    const long long* col_ptr; const int* row_idx;
    const int64_t* an; const int64_t* ad;     // CSC values
    const int64_t* cn; const int64_t* cd;     // cost
    int* basis;                               // basis_indices[m], tableau numbering, replicated
    unsigned char* in_basis;                  // per tableau column
    const int* art_row;                       // artificial column -> row (Partially::column_to_row)
    int64_t* rcn; int64_t* rcd;               // reduced cost per tableau column (transient)
    int64_t* dn; int64_t* dd;                 // generated column, local rows; entry mloc = rc
    int64_t* qn; int64_t* qd;                 // ratios b_i / d_i, local rows
    int64_t* pn; int64_t* pd;                 // normalised pivot row (m + 1), replicated
    int* tlist; int* slist;
    int2* worklist;
    Cand* cand;                               // per-block partial argmin results
    Cand* gather;                             // world candidates (multi-GPU all-gather buffer)
    LogEntry* log; long long log_cap;
};

#define RELP_THREADS 256

__device__ __forceinline__ Arena transient(const Dev& D) {
    Arena a; a.top = &D.st->tr_top; a.limit = D.st->tr_limit; a.err = &D.st->err; return a;
}
__device__ __forceinline__ Arena persistent(const Dev& D) {
    Arena a; a.top = &D.st->pool_top; a.limit = D.st->pool_limit; a.err = &D.st->err; return a;
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

// ---------------------------------------------------------------------------------------------
// Start of an iteration: recycle the transient limb region, clear the lists, and decide whether
// the persistent limb pool must be compacted before this iteration allocates more.
// ---------------------------------------------------------------------------------------------
__global__ void k_iter_begin(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    st->tr_top = st->tr_base;
    st->T = 0; st->S = 0; st->wl_count = 0;
    unsigned long long used = st->pool_top - st->pool_base;
    if (used > st->pool_peak) st->pool_peak = used;
    if (used > (st->pool_limit - st->pool_base) / 2) {
        // flip regions; k_compact (next launch) re-homes every live value
        unsigned long long b = st->alt_base, l = st->alt_limit;
        st->alt_base = st->pool_base; st->alt_limit = st->pool_limit;
        st->pool_base = b; st->pool_limit = l; st->pool_top = b;
        st->compacting = 1; st->compactions++;
    } else st->compacting = 0;
}

// Copy every live escalated Carry element into the fresh pool region (order is irrelevant).
__global__ void k_compact(Dev D) {
    DState* st = D.st;
    if (halted(st) || !st->compacting) return;
    size_t total = (size_t)(st->mloc + 1) * st->ld;
    Arena ar = persistent(D);
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        if (D.xd[e] == 0) {
            Val v; v.n = D.xn[e]; v.d = 0;
            Val c = copy_val(v, D.heap, ar);
            D.xn[e] = c.n; D.xd[e] = c.d;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K3 pricing: rc_j = c_j + sum_i (-pi_i) A_ij for every non-basic tableau column of this rank.
// One thread per column; -pi is row mloc of the planes (L2 resident).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ Val price_column(const Dev& D, const DState* st, int j, const Arena& ar) {
    const int64_t* zn = D.xn + (size_t)st->mloc * st->ld + 1;
    const int64_t* zd = D.xd + (size_t)st->mloc * st->ld + 1;
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
        int64_t zden = zd[row];
        if (z == 0 && zden == 1) continue;
        Val zv; zv.n = z; zv.d = zden;
        acc = val_fma(acc, zv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
    }
    return acc;
}

__global__ void k_price(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    Arena ar = transient(D);
    for (int j = st->col_lo + blockIdx.x * blockDim.x + threadIdx.x; j < st->col_hi; j += gridDim.x * blockDim.x) {
        Val rc = D.in_basis[j] ? val_absent() : price_column(D, st, j, ar);
        st_val(D.rcn, D.rcd, j, rc);
    }
}

// ---------------------------------------------------------------------------------------------
// Exact lexicographic argmin used by pricing and the ratio test.
//   ARG_DANTZIG : candidates rc_j < 0, order (rc_j, j)            pivot_rule.rs:111-127
//   ARG_FIRST   : candidates rc_j < 0, order cyclic distance from `start`   :53-57, :78-93
//   ARG_RATIO   : candidates present,  order (ratio_i, basis_indices[i])    tableau/mod.rs:221-247
// ---------------------------------------------------------------------------------------------
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

__device__ Cand block_argmin(Cand best, int mode, const uint32_t* heap) {
    __shared__ Cand sh[RELP_THREADS / 32];
    for (int d = 16; d > 0; d >>= 1) {
        Cand o = cand_shfl(best, d);
        if (cand_better(o, best, mode, heap)) best = o;
    }
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sh[w] = best;
    __syncthreads();
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        if (lane < nw) best = sh[lane]; else best.idx = -1;
        for (int d = 16; d > 0; d >>= 1) {
            Cand o = cand_shfl(best, d);
            if (cand_better(o, best, mode, heap)) best = o;
        }
    }
    __syncthreads();
    return best;   // valid in thread 0
}

// partial: one candidate per block into D.cand[blockIdx.x]
__global__ void k_argmin_partial(Dev D, int mode) {
    DState* st = D.st;
    if (halted(st)) return;
    Cand best; best.idx = -1; best.key = 0; best.v = val_absent();
    if (mode == ARG_RATIO) {
        for (int li = blockIdx.x * blockDim.x + threadIdx.x; li < st->mloc; li += gridDim.x * blockDim.x) {
            Val v = ld_val(D.qn, D.qd, li);
            if (is_absent(v)) continue;
            Cand c; c.v = v; c.idx = st->row_lo + li; c.key = D.basis[c.idx];
            if (cand_better(c, best, mode, D.heap)) best = c;
        }
    } else {
        int ncols = st->ncols;
        int start = (mode == ARG_FIRST && st->last_selected >= 0) ? st->last_selected : 0;
        for (int j = st->col_lo + blockIdx.x * blockDim.x + threadIdx.x; j < st->col_hi; j += gridDim.x * blockDim.x) {
            Val v = ld_val(D.rcn, D.rcd, j);
            if (is_absent(v) || val_sign(v, D.heap) >= 0) continue;
            Cand c; c.v = v; c.idx = j;
            c.key = mode == ARG_FIRST ? (j >= start ? j - start : j - start + ncols) : j;
            if (cand_better(c, best, mode, D.heap)) best = c;
        }
    }
    best = block_argmin(best, mode, D.heap);
    if (threadIdx.x == 0) D.cand[blockIdx.x] = best;
}

// final: reduce `count` candidates (from this rank's blocks, or from the all-gather buffer)
// and commit the decision to DState.  `to_gather`: write the winner to D.gather[rank] instead
// (multi-GPU stage 1).
__global__ void k_argmin_final(Dev D, int mode, int count, int from_gather, int to_gather_slot, int rule) {
    DState* st = D.st;
    if (halted(st)) return;
    const Cand* src = from_gather ? D.gather : D.cand;
    Cand best; best.idx = -1; best.key = 0; best.v = val_absent();
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        Cand c = src[i];
        if (cand_better(c, best, mode, D.heap)) best = c;
    }
    best = block_argmin(best, mode, D.heap);
    if (threadIdx.x != 0) return;
    if (to_gather_slot >= 0) { D.gather[to_gather_slot] = best; return; }
    if (mode == ARG_RATIO) {
        if (best.idx < 0) st->status = ST_NO_ROW; else st->r = best.idx;
    } else {
        if (best.idx < 0) { st->status = ST_NO_COLUMN; st->last_selected = -1; }
        else {
            st->j = best.idx; st->rc = best.v;
            if (rule == RULE_FIRST_MEMORY) st->last_selected = best.idx;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2 + K4: d_i = sum_k Binv[i][k] a_kj for every local row, the list of rows with d_i != 0, and
// the ratio b_i / d_i for rows with d_i > 0.
// ---------------------------------------------------------------------------------------------
__global__ void k_gen_column(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    Arena ar = transient(D);
    int j = st->j;
    int unit_row = -1; long long t0 = 0, t1 = 0;
    if (st->phase == 1 && j < st->nr_art) unit_row = D.art_row[j];
    else { int jj = st->phase == 1 ? j - st->nr_art : j; t0 = D.col_ptr[jj]; t1 = D.col_ptr[jj + 1]; }
    int mloc = st->mloc, ld = st->ld;
    for (int li = blockIdx.x * blockDim.x + threadIdx.x; li <= mloc; li += gridDim.x * blockDim.x) {
        if (li == mloc) {                       // cost-row factor: z -= rc * p   (carry/mod.rs:326-333)
            Val rc = copy_val(st->rc, D.heap, ar);
            st_val(D.dn, D.dd, li, rc);
            if (!val_is_zero(rc)) { int p = atomicAdd(&st->T, 1); D.tlist[p] = li; }
            continue;
        }
        const int64_t* rn = D.xn + (size_t)li * ld + 1;
        const int64_t* rd = D.xd + (size_t)li * ld + 1;
        Val acc = val_zero();
        if (unit_row >= 0) acc = copy_val(ld_val(rn, rd, unit_row), D.heap, ar);
        else for (long long t = t0; t < t1; t++) {
            int k = D.row_idx[t];
            int64_t xn = rn[k], xd = rd[k];
            if (xn == 0 && xd == 1) continue;
            Val xv; xv.n = xn; xv.d = xd;
            acc = val_fma(acc, xv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
        }
        st_val(D.dn, D.dd, li, acc);
        Val q = val_absent();
        if (!val_is_zero(acc)) {
            int p = atomicAdd(&st->T, 1); D.tlist[p] = li;
            if (val_sign(acc, D.heap) > 0) {
                Val b = ld_val(D.xn, D.xd, (size_t)li * ld);
                q = val_div(b, acc, D.heap, ar);
            }
        }
        st_val(D.qn, D.qd, li, q);
    }
}

// ---------------------------------------------------------------------------------------------
// K1 part 1 (owner of row r): p_k = X[r][k] / d_r for k = 0..m, written back to row r and staged
// in the replicated pivot-row buffer; list of columns with p_k != 0; basis bookkeeping.
// basis_inverse_rows.rs:42-51, sparse.rs:291-301, carry/mod.rs:283-313 (b_r), tableau/mod.rs:72-84
// ---------------------------------------------------------------------------------------------
__global__ void k_pivot_row(Dev D, int multi) {
    DState* st = D.st;
    if (halted(st)) return;
    int r = st->r, lr = r - st->row_lo;
    bool owner = lr >= 0 && lr < st->mloc;
    if (!owner) {   // multi-GPU: contribute zeros to the sum all-reduce that replicates the row
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= st->m; k += gridDim.x * blockDim.x) { D.pn[k] = 0; D.pd[k] = 0; }
        return;
    }
    Val pivot = ld_val(D.dn, D.dd, lr);
    if (val_is_zero(pivot)) { if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(&st->err, ERR_ZERO_PIVOT); return; }
    Arena ar = persistent(D);
    int64_t* rn = D.xn + (size_t)lr * st->ld;
    int64_t* rd = D.xd + (size_t)lr * st->ld;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= st->m; k += gridDim.x * blockDim.x) {
        Val x = ld_val(rn, rd, k);
        Val p = x;
        if (!val_is_zero(x)) {
            p = val_div(x, pivot, D.heap, ar);
            st_val(rn, rd, k, p);
            if (!multi) { int s = atomicAdd(&st->S, 1); D.slist[s] = k; }
            else if (p.d == 0) atomicOr(&st->err, ERR_LIMB_CAPACITY);   // escalated pivot rows do not cross GPUs yet
        }
        st_val(D.pn, D.pd, k, p);
    }
}

// multi-GPU: every rank derives the column list from the replicated pivot row
__global__ void k_build_slist(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= st->m; k += gridDim.x * blockDim.x) {
        if (!(D.pn[k] == 0 && D.pd[k] == 1)) { int s = atomicAdd(&st->S, 1); D.slist[s] = k; }
    }
}

// Basis bookkeeping + statistics + log, one thread, after k_pivot_row (every rank).
__global__ void k_bookkeeping(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    int r = st->r, j = st->j;
    int leaving = D.basis[r];
    D.basis[r] = j;
    D.in_basis[leaving] = 0; D.in_basis[j] = 1;
    st->leaving = leaving;
    Val b = ld_val(D.pn, D.pd, 0);
    int costrow = val_is_zero(st->rc) ? 0 : 1;
    int Tl = st->T - 1 - costrow;                     // reference's col.len() - 1 (rows eliminated), this rank
    int Sl = st->S - (val_is_zero(b) ? 0 : 1) + 1;    // nnz(B^-1 row r) + 1
    if ((long long)st->iter < D.log_cap) {
        LogEntry e; e.tag = st->tag; e.j = j; e.r = r; e.leaving = leaving; e.T = Tl; e.S = Sl;
        D.log[st->iter] = e;
    }
    st->sum_T += (unsigned long long)(st->T - 1); st->sum_S += (unsigned long long)st->S;
    st->sum_TS += (unsigned long long)(st->T - 1) * (unsigned long long)st->S;
}

// ---------------------------------------------------------------------------------------------
// K1 part 2: x_ik -= d_i * p_k over the T x S block (local rows incl. the cost row, i != r).
// Fixed-width tier in place; elements whose operands or result leave int64/int64 are queued for
// the multi-limb kernel.   basis_inverse_rows.rs:63-83 -> sparse.rs:213-248; carry/mod.rs:296-312,
// :326-333.
// ---------------------------------------------------------------------------------------------
__global__ void k_update(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    int T = st->T, S = st->S, ld = st->ld;
    int lr = st->r - st->row_lo;
    long long total = (long long)T * S;
    for (long long w = blockIdx.x * (long long)blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
        int t = (int)(w / S), s = (int)(w - (long long)t * S);
        int li = D.tlist[t];
        if (li == lr) continue;
        int k = D.slist[s];
        size_t e = (size_t)li * ld + k;
        Val x = ld_val(D.xn, D.xd, e);
        Val dv = ld_val(D.dn, D.dd, li);
        Val p = ld_val(D.pn, D.pd, k);
        Val out;
        if (x.d > 0 && dv.d > 0 && p.d > 0 && small_fma(x, dv, p, -1, &out)) {
            st_val(D.xn, D.xd, e, out);
        } else {
            int q = atomicAdd(&st->wl_count, 1);
            D.worklist[q] = make_int2(li, k);
        }
    }
}

__global__ void k_update_big(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    int n = st->wl_count, ld = st->ld;
    Arena ar = persistent(D);
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < n; w += gridDim.x * blockDim.x) {
        int2 it = D.worklist[w];
        size_t e = (size_t)it.x * ld + it.y;
        Val x = ld_val(D.xn, D.xd, e);
        Val out = big_fma(x, ld_val(D.dn, D.dd, it.x), ld_val(D.pn, D.pd, it.y), -1, D.heap, ar);
        st_val(D.xn, D.xd, e, out);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) st->escalated += (unsigned long long)n;
}

__global__ void k_iter_end(Dev D) {
    DState* st = D.st;
    if (halted(st)) return;
    st->iter++;
}

// ---------------------------------------------------------------------------------------------
// Set-up and hand-over kernels (run once per solve)
// ---------------------------------------------------------------------------------------------
// B^-1 = I, b = rhs (local rows); cost row cleared.   carry/mod.rs:381-426, basis_inverse_rows.rs:97-101
__global__ void k_init_planes(Dev D, const int64_t* rhs_n, const int64_t* rhs_d) {
    DState* st = D.st;
    size_t ld = st->ld, total = (size_t)(st->mloc + 1) * ld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int li = (int)(e / ld), k = (int)(e - (size_t)li * ld);
        int64_t n = 0, d = 1;
        if (li < st->mloc) {
            int i = st->row_lo + li;
            if (k == 0) { n = rhs_n[i]; d = rhs_d[i]; }
            else if (k == i + 1) n = 1;
        }
        D.xn[e] = n; D.xd[e] = d;
    }
}

// cost row of the artificial tableau: -pi_i = -1 on artificial rows, -obj = -sum of their b_i
__global__ void k_init_cost_row_artificial(Dev D, const int64_t* rhs_n, const int64_t* rhs_d) {
    DState* st = D.st;
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    Arena ar = persistent(D);
    size_t z = (size_t)st->mloc * st->ld;
    Val acc = val_zero();
    Val one = make_small(1, 1);
    for (int a = 0; a < st->nr_art; a++) {
        int i = D.art_row[a];
        acc = val_fma(acc, make_small(rhs_n[i], rhs_d[i]), one, -1, D.heap, ar);
        D.xn[z + 1 + i] = -1; D.xd[z + 1 + i] = 1;
    }
    st_val(D.xn, D.xd, z, acc);
}

// K5: z_k = -sum_i c_B(i) X[i][k] over this rank's rows, for k = 0..m (column 0 gives -obj).
// `cost_rows` lists local rows whose basic column has a nonzero cost.   carry/mod.rs:214-271
__global__ void k_recompute_cost_row(Dev D, const int* cost_rows, int n_cost_rows, int64_t* outn, int64_t* outd) {
    DState* st = D.st;
    Arena ar = persistent(D);
    int ld = st->ld;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= st->m; k += gridDim.x * blockDim.x) {
        Val acc = val_zero();
        for (int q = 0; q < n_cost_rows; q++) {
            int li = cost_rows[q];
            size_t e = (size_t)li * ld + k;
            int64_t xn = D.xn[e], xd = D.xd[e];
            if (xn == 0 && xd == 1) continue;
            int jb = D.basis[st->row_lo + li];
            Val xv; xv.n = xn; xv.d = xd;
            acc = val_fma(acc, xv, make_small(D.cn[jb], D.cd[jb]), -1, D.heap, ar);
        }
        st_val(outn, outd, k, acc);
    }
}

// Artificial clean-up search (phase_one.rs:236-243): first non-artificial non-basic column with
// relative cost exactly zero and a nonzero element in `pivot_row` of B^-1 A.
__global__ void k_cleanup_select(Dev D, int pivot_row, int* best_j) {
    DState* st = D.st;
    if (st->err) return;
    Arena ar = transient(D);
    int lr = pivot_row - st->row_lo;
    bool owner = lr >= 0 && lr < st->mloc;
    for (int j = st->nr_art + blockIdx.x * blockDim.x + threadIdx.x; j < st->ncols; j += gridDim.x * blockDim.x) {
        if (D.in_basis[j]) continue;
        Val rc = price_column(D, st, j, ar);
        if (!val_is_zero(rc)) continue;
        if (!owner) continue;
        int jj = j - st->nr_art;
        const int64_t* rn = D.xn + (size_t)lr * st->ld + 1;
        const int64_t* rd = D.xd + (size_t)lr * st->ld + 1;
        Val acc = val_zero();
        for (long long t = D.col_ptr[jj]; t < D.col_ptr[jj + 1]; t++) {
            int k = D.row_idx[t];
            Val xv = ld_val(rn, rd, k);
            if (val_is_zero(xv)) continue;
            acc = val_fma(acc, xv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
        }
        if (!val_is_zero(acc)) atomicMin(best_j, j);
    }
}

// one generated element (basis_inverse_rows.rs:157-173)
__global__ void k_generate_element(Dev D, int i, int j, int64_t* outn, int64_t* outd) {
    DState* st = D.st;
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    Arena ar = transient(D);
    int li = i - st->row_lo;
    const int64_t* rn = D.xn + (size_t)li * st->ld + 1;
    const int64_t* rd = D.xd + (size_t)li * st->ld + 1;
    Val acc = val_zero();
    if (st->phase == 1 && j < st->nr_art) acc = ld_val(rn, rd, D.art_row[j]);
    else {
        int jj = st->phase == 1 ? j - st->nr_art : j;
        for (long long t = D.col_ptr[jj]; t < D.col_ptr[jj + 1]; t++) {
            Val xv = ld_val(rn, rd, D.row_idx[t]);
            if (val_is_zero(xv)) continue;
            acc = val_fma(acc, xv, make_small(D.an[t], D.ad[t]), 1, D.heap, ar);
        }
    }
    outn[0] = acc.n; outd[0] = acc.d;
}

// Row/column removal (basis_inverse_rows.rs:194-205, carry/mod.rs:650-689): gather the kept rows
// and columns of the planes into fresh planes.  keep_row[i'] / keep_col[k'] give source indices.
__global__ void k_gather_planes(const int64_t* sn, const int64_t* sd, int sld,
                                int64_t* tn, int64_t* td, int tld, int trows, int tcols,
                                const int* keep_row, const int* keep_col) {
    size_t total = (size_t)trows * tld;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int i = (int)(e / tld), k = (int)(e - (size_t)i * tld);
        int64_t n = 0, d = 1;
        if (k < tcols) { size_t s = (size_t)keep_row[i] * sld + keep_col[k]; n = sn[s]; d = sd[s]; }
        tn[e] = n; td[e] = d;
    }
}

// Export a block of rats: escalated values are copied to the end of the transient region so that
// one D2H copy of [export_base, tr_top) carries all limbs; handles become offsets from export_base.
__global__ void k_export(Dev D, const int64_t* sn, const int64_t* sd, int sld, int nrows, int ncols,
                         int64_t* on, int64_t* od, unsigned long long export_base) {
    Arena ar = transient(D);
    size_t total = (size_t)nrows * ncols;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int i = (int)(e / ncols), k = (int)(e - (size_t)i * ncols);
        Val v = ld_val(sn, sd, (size_t)i * sld + k);
        if (v.d == 0) { v = copy_val(v, D.heap, ar); if (v.d == 0) v.n -= (int64_t)export_base; }
        on[e] = v.n; od[e] = v.d;
    }
}

}  // namespace relp
