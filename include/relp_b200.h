/* relp_b200.h -- C ABI of the B200-native exact revised-simplex hot path.
 *
 * The reference (RELP, crate `relp` 0.0.5) has no FFI: its boundary for this path is four generic
 * Rust traits.  Each entry point below names the trait method(s) it replaces, with file:line under
 * /root/reference/src/algorithm/two_phase/.  A Rust shim implements
 * `InverseMaintener` / `BasisInverse` / `PivotRule` on top of exactly these symbols.
 *
 * Conventions
 *   - All functions return a `relp_status` value (>= 0 ok / solver outcome, < 0 error).  No CPU fallback:
 *     without a CUDA device `relp_create` returns NULL and sets relp_last_error().
 *   - Rationals cross the boundary as "rat arrays": parallel arrays num[count], den[count] of
 *     int64 plus one limb buffer.  den[i] >= 1: the value is num[i]/den[i] (canonical: reduced,
 *     zero = 0/1).  den[i] == 0: the value does not fit int64/int64 and num[i] is a word offset
 *     into the limb buffer where [header, numerator limbs, denominator limbs] are stored
 *     (uint32 limbs, little endian; header = nn | nd << 14 | sign << 31).  Always canonical, i.e.
 *     bit-identical to what `RationalBig` (rational/big/mod.rs:17-29) would hold.
 *   - Column indices are tableau indices of the current phase: in phase 1 the `nr_artificial`
 *     artificial columns come first (tableau/kind/artificial/partially.rs:72-88), in phase 2 they
 *     are provider indices.
 *   - Rows: the Carry matrix has m+1 rows and m+1 columns.  Row i < m is [b_i | B^-1 row i];
 *     row m is the cost row [-objective | -pi]  (inverse_maintenance/carry/mod.rs:45-65).
 */
#ifndef RELP_B200_H
#define RELP_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct relp_ctx relp_ctx;

enum relp_status {
    RELP_OK = 0,              /* call succeeded / loop still running (iteration budget reached)     */
    RELP_OPTIMAL = 1,         /* pricing found no column: OptimumResult::FiniteOptimum / phase-1 end */
    RELP_UNBOUNDED = 2,       /* ratio test found no row (phase_two.rs:44-46)                       */
    RELP_INFEASIBLE = 3,      /* phase 1 ended with objective != 0 (phase_one.rs:146-148)           */
    RELP_ERR_CUDA = -1,
    RELP_ERR_ZERO_PIVOT = -2, /* "Pivot value can't be zero" (carry/mod.rs:291-292)                 */
    RELP_ERR_LIMB_CAPACITY = -3,  /* a numerator/denominator exceeded 16 383 limbs (512 Kbit, the record header's length field), or a multi-limb value would have to cross GPUs */
    RELP_ERR_HEAP_EXHAUSTED = -4, /* limb heap full: recreate the context with a larger heap        */
    RELP_ERR_ARGUMENT = -5,
    RELP_ERR_PHASE1_UNBOUNDED = -6, /* "Artificial cost can not be unbounded" (phase_one.rs:143)     */
    RELP_ERR_COMM = -7,
    RELP_ERR_ARTIFICIAL_BASIC = -8 /* an artificial column survived the phase-one clean-up; the reference
                                      underflows `basis_column -= nr_artificial` here (carry/mod.rs:500) */
};

enum relp_rule {              /* strategy/pivot_rule.rs */
    RELP_FIRST_PROFITABLE = 0,             /* :38-58  */
    RELP_FIRST_PROFITABLE_WITH_MEMORY = 1, /* :62-94  (phase-1 default, phase_one.rs:97)   */
    RELP_STEEPEST_DESCENT = 2              /* :97-127 (phase-2 default, two_phase/mod.rs:44) */
};

/* One entry of the pivot log written by relp_iterate / relp_solve_relaxation. */
typedef struct relp_iter_log {
    int32_t tag;      /* '1' phase-1 pivot, '2' phase-2 pivot, 'z' zero-cost artificial clean-up pivot */
    int32_t j;        /* entering column (tableau index of that phase) */
    int32_t r;        /* pivot row */
    int32_t leaving;  /* leaving column */
    int32_t T;        /* nnz(generated column) - 1 : rows eliminated            */
    int32_t S;        /* nnz(B^-1 row r) + 1       : entries per eliminated row */
} relp_iter_log;

typedef struct relp_stats {
    uint64_t pivots;            /* pivots taken since relp_load_matrix                       */
    uint64_t escalated;         /* Carry elements that took the multi-limb path              */
    uint64_t sum_T, sum_S, sum_TS; /* algorithmic work: bytes(K1) = w * (2*sum_TS + sum_S + sum_T) */
    uint64_t heap_peak_words;   /* high-water mark of the persistent limb pool               */
    uint64_t compactions;       /* limb-pool compactions                                     */
    uint64_t kernel_launches;   /* CUDA kernels launched by this context                     */
    int32_t  max_limbs;         /* longest stored numerator/denominator seen at read-back    */
    int32_t  m, n, nr_artificial, phase;
    double   loop_ms;           /* device time (CUDA events) spent inside relp_iterate loops */
    uint64_t stage_ns[8];       /* persistent loop, ns: [0] selection, [1] column+ratio, [2] pivot row, [3] update,
                                   [4] multi-limb + housekeeping, [5] narrow iterations (CTA 0 alone, all stages) */
    uint64_t narrow_pivots;     /* pivots done by one CTA alone (sparse B^-1), the rest by the whole grid   */
    uint64_t wide_handoffs;     /* pivots whose dense Carry update the persistent loop handed to the staged update kernel */
} relp_stats;

/* ---- life cycle ------------------------------------------------------------------------- */
/* One context = one GPU.  rank/world describe the row shard of a multi-GPU solve;
 * world == 1 for a single-GPU solve.  heap_words = size of the limb heap in uint32 words
 * (0 = default).  Returns NULL on failure (see relp_last_error). */
relp_ctx* relp_create(int device, int rank, int world, uint64_t heap_words);
void relp_destroy(relp_ctx*);
const char* relp_last_error(void);
int relp_abi_version(void);
/* options: "persistent" = 1 (default) runs single-GPU loops as one cooperative kernel, 0 as one
 * launch per stage (the path multi-GPU solves use, with NCCL between the stages); "narrow" = 1
 * (default) lets CTA 0 of the persistent kernel do small pivots alone, 0 always uses the whole grid;
 * "k1_exit" / "k1_min" = number of Carry elements (T x S) from which the persistent loop hands a dense
 * update to the staged update kernel / from which that kernel tiles the update (set before the Carry
 * is created; tests lower them so that small problems take the large-problem paths).
 * Multi-limb tier (tests): "coop_from" = limbs from which an operation leaves the fixed-size per-thread routines for
 * the arbitrary-length ones (default 64, the per-thread buffers' size); "long_mode" = 1 runs the copy of the loop that
 * contains that tier from the first pivot instead of handing over when a value beyond 12 limbs is stored;
 * "wslab_words" / "tslab_words" = scratch words per resident warp / per thread for long operations.
 * "cleanup_any" = 1: the artificial clean-up (phase_one.rs:223-260) pivots on ANY nonzero element of the row that holds a
 * zero-level artificial column and removes a row only when it is entirely zero; default 0 = the reference's rule
 * (zero-cost columns only, the artificial's original row), which also removes rows that are not redundant -- DESIGN.md 6. */
int relp_set_option(relp_ctx*, const char* name, int64_t value);
/* the cudaStream_t all work of this context is enqueued on (for CUDA-event timing by the caller) */
void* relp_stream(relp_ctx*);
/* debug (option "trace" = 1): globaltimer of every CTA at 16 points of the last loop iteration */
int relp_debug_trace(relp_ctx*, uint64_t* out /* n_ctas x 16 */, int32_t n_ctas);
/* K1 micro-benchmark: overwrite the Carry planes with dense small integers and a dense generated
 * column with d_r = 1, then time `reps` pivot-row + update kernel pairs (carry/mod.rs:549-570 at
 * T = m - 1 rows, S = m + 1 columns) with CUDA events.  The Carry is invalid afterwards. */
int relp_debug_k1_dense(relp_ctx*, uint64_t seed, int32_t reps, float* ms_pivot_row, float* ms_update,
                        int64_t* T_out, int64_t* S_out);
/* Self-check of that update: the same fabricated block and pivot updated by the staged path (bulk copies into the
 * shared-memory ring) and, from the same fill, by the element-wise path the oracle-parity tests cover; the planes and
 * the pattern bitmap are compared on the device.  *mismatches must be 0.  The Carry is invalid afterwards. */
int relp_debug_k1_verify(relp_ctx*, uint64_t seed, int64_t* mismatches, int64_t* elements);
/* Test hook for the device arithmetic (data/number_types/rational/big/wrapping.rs:47-311 is what it replaces): `count`
 * independent operations, one per thread, through the kernels' own front ends: op 0: x + s a b, 1: x / a, 2: sign(x - a);
 * op 3: x + s a b with every operation done by the 32 lanes of its warp together (the cooperative tier, relp_wcb.cuh).
 * operands = six arrays of `count` int64 (xn xd an ad bn bd), rat-array encoding over limbs_in; results as a rat array. */
int relp_debug_rat_op(relp_ctx*, int op, int32_t count, const int64_t* operands, const uint32_t* limbs_in, size_t limb_words,
                      int s, int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);

/* Multi-GPU only: attach an NCCL communicator (ncclComm_t, created by the caller from
 * torch.distributed's rendezvous) used for the per-pivot pivot-row broadcast and the two
 * candidate all-gathers.  `nccl_lib` is the dlopen path of libnccl (NULL = "libnccl.so.2"). */
int relp_comm_unique_id(void* id128, const char* nccl_lib);
int relp_attach_comm(relp_ctx*, const void* id128, const char* nccl_lib);
/* Peer transport for the multi-GPU solve (one process per GPU, same node): every rank allocates an
 * exchange buffer sized for LPs of up to max_m rows and exports it (CUDA IPC handle, 64 bytes); after
 * the handles have been all-gathered (rank order, 64 bytes each) every rank maps its peers' buffers.
 * The persistent loop kernel then runs on every rank and exchanges the ratio-test candidates and the
 * pivot row itself, by stores into peer memory over NVLink -- no NCCL call and no launch per pivot.
 * The NCCL communicator (or the callbacks) is still used once per solve, at the phase hand-over. */
int relp_peer_export(relp_ctx*, int32_t max_m, void* handle64);
int relp_peer_attach(relp_ctx*, const void* handles /* world x 64 bytes */);
/* Alternative transport: host-staged callbacks (e.g. torch.distributed over gloo).  `send`/`recv`/`buf`
 * are HOST buffers; all-gather concatenates `bytes` from every rank in rank order; all-reduce sums
 * int64 in place.  Return 0 on success.  Used by the world_size-2 CPU tests. */
typedef int (*relp_allgather_fn)(void* user, const void* send, void* recv, size_t bytes);
typedef int (*relp_allreduce_fn)(void* user, int64_t* buf, size_t count);
int relp_attach_comm_callbacks(relp_ctx*, relp_allgather_fn, relp_allreduce_fn, void* user);

/* ---- input: MatrixProvider (matrix_provider/mod.rs:37-136) ---------------------------------- */
/* CSC of ALL provider columns, virtual slack columns materialised (matrix_data.rs:308-348),
 * cost_value(j) with None == 0 (matrix_data.rs:350-357), right_hand_side() >= 0 (:359-371).
 * Host pointers; copied to the device (the e2e timed region of bench.py includes this copy).
 * The arrays are validated (RELP_ERR_ARGUMENT + relp_last_error otherwise): no NULL pointer; col_ptr[0] == 0 and
 * non-decreasing; 0 <= row_idx < m; every value, cost and right-hand side canonical (den >= 1, gcd(num, den) == 1,
 * num != INT64_MIN -- the encoding reserves den == 0 and den < 0, see above); rhs >= 0.
 * The loaded provider is immutable for the caller: a solve that removes redundant rows works on a filtered copy,
 * and the next Carry constructor on the same context puts the loaded provider back (no reload needed). */
int relp_load_matrix(relp_ctx*, int32_t m, int32_t n,
                     const int64_t* col_ptr, const int32_t* row_idx,
                     const int64_t* val_num, const int64_t* val_den,
                     const int64_t* cost_num, const int64_t* cost_den,
                     const int64_t* rhs_num, const int64_t* rhs_den);

/* Replace cost_value(j) of the loaded provider (same canonical-form rules).  Costs are read by phase two only, so this
 * is refused (RELP_ERR_ARGUMENT) while a phase-two tableau is live.  For trait-level callers: RELP's
 * InverseMaintener::create_for_partially_artificial never sees the provider, the real costs arrive with
 * from_artificial / from_artificial_remove_rows (inverse_maintenance/mod.rs:120-151). */
int relp_set_costs(relp_ctx*, int32_t n, const int64_t* cost_num, const int64_t* cost_den);

/* Replace right_hand_side() by a rat array (the encoding every reader of this header returns): den > 0 is num/den as in
 * relp_load_matrix; den == 0 means `num` is the word offset in `limbs` of a record [header | numerator limbs |
 * denominator limbs] (little-endian 32-bit limbs, header = nn | nd << 14, sign bit clear: rhs >= 0).  A value that fits
 * int64/int64 must be passed as such.  Records are trusted to be in lowest terms.  Why: the reference's right-hand side
 * is a RationalBig (MatrixProvider::Rhs, matrix_provider/mod.rs:37-136 with tests/netlib/mod.rs:44-45) and its presolve
 * substitutes solved variables into it -- netlib BORE3D reaches the simplex with 285-bit right-hand sides.  Call after
 * relp_load_matrix (which takes placeholders, e.g. 0/1, for such rows) and before a Carry constructor. */
int relp_set_rhs(relp_ctx*, int32_t m, const int64_t* rhs_num, const int64_t* rhs_den, const uint32_t* limbs, size_t limb_words);

/* ---- InverseMaintener constructors (inverse_maintenance/mod.rs:29-267) --------------------- */
/* Partially::new + Carry::create_for_partially_artificial
 * (tableau/kind/artificial/partially.rs:125-205, carry/mod.rs:381-426).  (piv_rows[k], piv_cols[k])
 * = PartialInitialBasis::pivot_element_indices() (matrix_data.rs:432-452).  n_piv == 0 gives the
 * fully artificial tableau (carry/mod.rs:358-379). */
int relp_carry_create_partially_artificial(relp_ctx*, int32_t n_piv,
                                           const int32_t* piv_rows, const int32_t* piv_cols);
/* NonArtificial::from_artificial + Carry::from_artificial (non_artificial.rs:151-172,
 * carry/mod.rs:484-510, :214-271): drop the artificial columns, recompute -pi and -obj (K5). */
int relp_carry_from_artificial(relp_ctx*);
/* from_artificial_removing_rows (non_artificial.rs:191-220, carry/mod.rs:650-689,
 * filter/generic_wrapper.rs:224-300): additionally remove `rows` (sorted) from the Carry, from
 * every column of A and from b. */
int relp_carry_from_artificial_remove_rows(relp_ctx*, int32_t n_rows, const int32_t* sorted_rows);
/* InverseMaintener::from_basis (inverse_maintenance/mod.rs:91, carry/mod.rs:428-463): the phase-two Carry of
 * a given basis, basis[i] = provider column in row i (from_basis_pivots, carry/mod.rs:465-482, sorts its
 * pairs by row and calls this).  RELP_ERR_ARGUMENT if the columns are linearly dependent. */
int relp_carry_from_basis(relp_ctx*, const int32_t* basis /* m */);

/* ---- per-iteration trait methods (one call per step; results stay on the device) ------------- */
/* PivotRule::select_primal_pivot_column (pivot_rule.rs:21-33) over Tableau::relative_cost
 * (tableau/mod.rs:102-108) = Carry::cost_difference (carry/mod.rs:572-577) + initial cost.  K3.
 * Returns RELP_OK and *j_out, or RELP_OPTIMAL if no column is profitable.  rc (one rat, may be
 * NULL) receives the reduced cost. */
int relp_price(relp_ctx*, int rule, int32_t* j_out,
               int64_t* rc_num, int64_t* rc_den, uint32_t* limbs, size_t limb_cap);
/* Tableau::generate_column (tableau/mod.rs:122-126) -> BasisInverseRows::generate_column
 * (basis_inverse_rows.rs:144-155).  K2.  `rc` of column j is recomputed on the device. */
int relp_generate_column(relp_ctx*, int32_t j);
/* Tableau::select_primal_pivot_row (tableau/mod.rs:221-247) on the generated column.  K4.
 * RELP_OK and *r_out, or RELP_UNBOUNDED. */
int relp_ratio_test(relp_ctx*, int32_t* r_out);
/* Tableau::bring_into_basis (tableau/mod.rs:47-60) -> Carry::change_basis (carry/mod.rs:549-570)
 * with the column generated last.  K1. */
int relp_change_basis(relp_ctx*, int32_t r, int32_t* leaving_out);
/* BasisInverseRows::generate_element (basis_inverse_rows.rs:157-173): one rat. */
int relp_generate_element(relp_ctx*, int32_t i, int32_t j,
                          int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap);

/* ---- fused drivers ---------------------------------------------------------------------------- */
/* phase_one::primal / phase_two::primal loop (phase_one.rs:135-169, phase_two.rs:32-50), device
 * resident: price -> generate column -> ratio test -> change basis, up to max_iters pivots
 * (< 0: until termination).  Appends to the context's pivot log.  Returns RELP_OK (budget
 * reached), RELP_OPTIMAL (pricing found nothing), RELP_UNBOUNDED, or an error. */
int relp_iterate(relp_ctx*, int rule, int64_t max_iters);
/* SolveRelaxation::solve_relaxation (two_phase/mod.rs:30-75): artificial tableau, phase 1, the
 * artificial clean-up (phase_one.rs:223-260), hand-over (with row removal when rank deficient) and
 * phase 2.  max_pivots < 0: run to the end.  Resumable: call again to continue. */
int relp_solve_relaxation(relp_ctx*, int rule_phase1, int rule_phase2,
                          int32_t n_piv, const int32_t* piv_rows, const int32_t* piv_cols,
                          int64_t max_pivots);
int relp_solve_continue(relp_ctx*, int64_t max_pivots);
/* Stop the resumable solve right after the phase hand-over (parity tests of the pinned fixtures). */
int relp_solve_to_handover(relp_ctx*, int rule_phase1, int32_t n_piv,
                           const int32_t* piv_rows, const int32_t* piv_cols);

/* ---- readers (InverseMaintener::current_bfs/b/get_objective_function_value, carry/mod.rs:616-642) */
int relp_dims(relp_ctx*, int32_t* m, int32_t* n, int32_t* nr_artificial, int32_t* phase);
int relp_read_basis(relp_ctx*, int32_t* basis_indices /* m */);
/* Rat array of a rectangular block of the (m+1) x (m+1) Carry, row-major. */
int relp_read_block(relp_ctx*, int32_t row0, int32_t nrows, int32_t col0, int32_t ncols,
                    int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);
int relp_read_carry(relp_ctx*, int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);
int relp_read_b(relp_ctx*, int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);
int relp_read_minus_pi(relp_ctx*, int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);
int relp_read_minus_objective(relp_ctx*, int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);
/* The column generated last (dense, m entries; zeros are the entries the reference drops). */
int relp_read_column(relp_ctx*, int64_t* num, int64_t* den, uint32_t* limbs, size_t limb_cap, size_t* limb_used);
/* Pivot log: number of entries, then a copy of [first, first+count). */
int64_t relp_log_size(relp_ctx*);
int relp_read_log(relp_ctx*, int64_t first, int64_t count, relp_iter_log* out);
int relp_stats_get(relp_ctx*, relp_stats* out);
/* Redundant rows found by the artificial clean-up of the last solve (phase_one.rs:223-260). */
int relp_rows_removed(relp_ctx*, int32_t* out, int32_t cap);

/* ---- allocation-free variant ---------------------------------------------------------------------
 * relp_ratio_test_static has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  It is the K4 ratio test of relp_ratio_test -- argmin over rows with
 * d_i > 0 of b_i / d_i, ties to the smaller basis column, the rule of tableau/mod.rs:221-247 -- as a
 * context-free call that allocates nothing: no cudaMalloc, no cudaFree, no allocating library routine.
 * The per-block candidates of the reduction live in a file-scope `static __device__` array of the library
 * (one candidate per 1024 rows, 24 KB + a 256-byte header; every GPU that loads the library holds it
 * until the process exits), hence the limit m <= RELP_STATIC_MAX_ROWS: a larger m is refused with
 * RELP_ERR_ARGUMENT and a message in relp_last_error().
 *   b_num/b_den, d_num/d_den : DEVICE pointers, m int64 each, canonical rationals of the int64 tier
 *                              (den >= 1; a den <= 0, i.e. a multi-limb handle, gives status
 *                              RELP_ERR_LIMB_CAPACITY in result[1] -- use the context API for those)
 *   basis                    : DEVICE pointer, m int32 (basis column of every row)
 *   result                   : DEVICE pointer, 2 int32: result[0] = pivot row or -1,
 *                              result[1] = RELP_OK / RELP_UNBOUNDED (no row with d_i > 0) / error
 *   stream                   : cudaStream_t the work is enqueued on (NULL = the default stream)
 * The call only enqueues (a 256-byte memset of the scratch header and one kernel) and returns; the caller
 * synchronises the stream before reading `result`.  The current device must be the one the pointers
 * belong to; its copy of the scratch is looked up once per device (cudaGetSymbolAddress), at the first call.
 * ALL calls on one GPU share that scratch array, which keeps the previous call's contents: the header is
 * cleared by every call, but two calls that overlap on the GPU (different streams, or different host
 * threads without ordering between them) race on it and give undefined results.  Serialise them. */
#define RELP_STATIC_MAX_ROWS (1 << 20)
int relp_ratio_test_static(const int64_t* b_num, const int64_t* b_den, const int64_t* d_num, const int64_t* d_den,
                           const int32_t* basis, int32_t m, int32_t* result, void* stream);

#ifdef __cplusplus
}
#endif
#endif
