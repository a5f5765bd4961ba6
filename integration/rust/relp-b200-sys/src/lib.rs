//! Raw FFI of `include/relp_b200.h` (one declaration per exported symbol) and the conversion of the library's
//! "rat arrays" to `num::BigRational`.  Never compiled (no Rust toolchain where this was written).
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)] pub struct relp_ctx { _private: [u8; 0] }

pub const RELP_OK: c_int = 0;
pub const RELP_OPTIMAL: c_int = 1;
pub const RELP_UNBOUNDED: c_int = 2;
pub const RELP_INFEASIBLE: c_int = 3;
pub const RELP_ERR_CUDA: c_int = -1;
pub const RELP_ERR_ZERO_PIVOT: c_int = -2;
pub const RELP_ERR_LIMB_CAPACITY: c_int = -3;
pub const RELP_ERR_HEAP_EXHAUSTED: c_int = -4;
pub const RELP_ERR_ARGUMENT: c_int = -5;
pub const RELP_ERR_PHASE1_UNBOUNDED: c_int = -6;
pub const RELP_ERR_COMM: c_int = -7;
pub const RELP_ERR_ARTIFICIAL_BASIC: c_int = -8;
/// strategy/pivot_rule.rs: FirstProfitable, FirstProfitableWithMemory, SteepestDescent
pub const RELP_FIRST_PROFITABLE: c_int = 0;
pub const RELP_FIRST_PROFITABLE_WITH_MEMORY: c_int = 1;
pub const RELP_STEEPEST_DESCENT: c_int = 2;

#[repr(C)] #[derive(Clone, Copy, Debug, Default)]
pub struct relp_iter_log { pub tag: i32, pub j: i32, pub r: i32, pub leaving: i32, pub t: i32, pub s: i32 }

#[repr(C)] #[derive(Clone, Copy, Debug, Default)]
pub struct relp_stats {
    pub pivots: u64, pub escalated: u64, pub sum_t: u64, pub sum_s: u64, pub sum_ts: u64, pub heap_peak_words: u64,
    pub compactions: u64, pub kernel_launches: u64, pub max_limbs: i32, pub m: i32, pub n: i32, pub nr_artificial: i32,
    pub phase: i32, pub loop_ms: f64, pub stage_ns: [u64; 8], pub narrow_pivots: u64, pub wide_handoffs: u64,
}

pub type relp_allgather_fn = Option<unsafe extern "C" fn(user: *mut c_void, send: *const c_void, recv: *mut c_void, bytes: usize) -> c_int>;
pub type relp_allreduce_fn = Option<unsafe extern "C" fn(user: *mut c_void, buf: *mut i64, count: usize) -> c_int>;

extern "C" {
    pub fn relp_create(device: c_int, rank: c_int, world: c_int, heap_words: u64) -> *mut relp_ctx;
    pub fn relp_destroy(ctx: *mut relp_ctx);
    pub fn relp_last_error() -> *const c_char;
    pub fn relp_abi_version() -> c_int;
    pub fn relp_set_option(ctx: *mut relp_ctx, name: *const c_char, value: i64) -> c_int;
    pub fn relp_stream(ctx: *mut relp_ctx) -> *mut c_void;
    pub fn relp_debug_trace(ctx: *mut relp_ctx, out: *mut u64, n_ctas: i32) -> c_int;
    pub fn relp_debug_k1_dense(ctx: *mut relp_ctx, seed: u64, reps: i32, ms_pivot_row: *mut f32, ms_update: *mut f32,
                               t_out: *mut i64, s_out: *mut i64) -> c_int;
    pub fn relp_debug_k1_verify(ctx: *mut relp_ctx, seed: u64, mismatches: *mut i64, elements: *mut i64) -> c_int;
    pub fn relp_debug_rat_op(ctx: *mut relp_ctx, op: c_int, count: i32, operands: *const i64, limbs_in: *const u32, limb_words: usize,
                             s: c_int, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    pub fn relp_comm_unique_id(id128: *mut c_void, nccl_lib: *const c_char) -> c_int;
    pub fn relp_attach_comm(ctx: *mut relp_ctx, id128: *const c_void, nccl_lib: *const c_char) -> c_int;
    pub fn relp_peer_export(ctx: *mut relp_ctx, max_m: i32, handle64: *mut c_void) -> c_int;
    pub fn relp_peer_attach(ctx: *mut relp_ctx, handles: *const c_void) -> c_int;
    pub fn relp_attach_comm_callbacks(ctx: *mut relp_ctx, ag: relp_allgather_fn, ar: relp_allreduce_fn, user: *mut c_void) -> c_int;
    /// MatrixProvider side: CSC of all provider columns (virtual slacks materialised), costs, right-hand side >= 0
    pub fn relp_load_matrix(ctx: *mut relp_ctx, m: i32, n: i32, col_ptr: *const i64, row_idx: *const i32,
                            val_num: *const i64, val_den: *const i64, cost_num: *const i64, cost_den: *const i64,
                            rhs_num: *const i64, rhs_den: *const i64) -> c_int;
    /// replace cost_value(j) (read by phase two only; refused while a phase-two tableau is live)
    pub fn relp_set_costs(ctx: *mut relp_ctx, n: i32, cost_num: *const i64, cost_den: *const i64) -> c_int;
    /// replace right_hand_side() by a rat array (den == 0: record in `limbs`): the reference's Rhs is a RationalBig
    pub fn relp_set_rhs(ctx: *mut relp_ctx, m: i32, rhs_num: *const i64, rhs_den: *const i64, limbs: *const u32, limb_words: usize) -> c_int;
    /// InverseMaintener::create_for_partially_artificial (carry/mod.rs:381-426)
    pub fn relp_carry_create_partially_artificial(ctx: *mut relp_ctx, n_piv: i32, piv_rows: *const i32, piv_cols: *const i32) -> c_int;
    /// InverseMaintener::from_artificial (carry/mod.rs:484-510)
    pub fn relp_carry_from_artificial(ctx: *mut relp_ctx) -> c_int;
    /// InverseMaintener::from_artificial_remove_rows (carry/mod.rs:650-689)
    pub fn relp_carry_from_artificial_remove_rows(ctx: *mut relp_ctx, n_rows: i32, sorted_rows: *const i32) -> c_int;
    /// InverseMaintener::from_basis (carry/mod.rs:428-463)
    pub fn relp_carry_from_basis(ctx: *mut relp_ctx, basis: *const i32) -> c_int;
    /// PivotRule::select_primal_pivot_column (strategy/pivot_rule.rs:21-33)
    pub fn relp_price(ctx: *mut relp_ctx, rule: c_int, j_out: *mut i32, rc_num: *mut i64, rc_den: *mut i64,
                      limbs: *mut u32, limb_cap: usize) -> c_int;
    /// InverseMaintener::generate_column (carry/mod.rs:579-587); the column stays on the device (relp_read_column)
    pub fn relp_generate_column(ctx: *mut relp_ctx, j: i32) -> c_int;
    /// Tableau::select_primal_pivot_row (tableau/mod.rs:221-247)
    pub fn relp_ratio_test(ctx: *mut relp_ctx, r_out: *mut i32) -> c_int;
    /// InverseMaintener::change_basis (carry/mod.rs:549-570)
    pub fn relp_change_basis(ctx: *mut relp_ctx, r: i32, leaving_out: *mut i32) -> c_int;
    /// InverseMaintener::generate_element (basis_inverse_rows.rs:157-173)
    pub fn relp_generate_element(ctx: *mut relp_ctx, i: i32, j: i32, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize) -> c_int;
    /// phase_one::primal / phase_two::primal on the device (phase_one.rs:125-170, phase_two.rs:22-51)
    pub fn relp_iterate(ctx: *mut relp_ctx, rule: c_int, max_iters: i64) -> c_int;
    /// SolveRelaxation::solve_relaxation (algorithm/two_phase/mod.rs:30-75)
    pub fn relp_solve_relaxation(ctx: *mut relp_ctx, rule_phase1: c_int, rule_phase2: c_int, n_piv: i32,
                                 piv_rows: *const i32, piv_cols: *const i32, max_pivots: i64) -> c_int;
    pub fn relp_solve_continue(ctx: *mut relp_ctx, max_pivots: i64) -> c_int;
    pub fn relp_solve_to_handover(ctx: *mut relp_ctx, rule_phase1: c_int, n_piv: i32, piv_rows: *const i32, piv_cols: *const i32) -> c_int;
    pub fn relp_dims(ctx: *mut relp_ctx, m: *mut i32, n: *mut i32, nr_artificial: *mut i32, phase: *mut i32) -> c_int;
    /// InverseMaintener::basis_column_index_for_row for every row
    pub fn relp_read_basis(ctx: *mut relp_ctx, basis_indices: *mut i32) -> c_int;
    pub fn relp_read_block(ctx: *mut relp_ctx, row0: i32, nrows: i32, col0: i32, ncols: i32, num: *mut i64, den: *mut i64,
                           limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    pub fn relp_read_carry(ctx: *mut relp_ctx, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    /// InverseMaintener::b
    pub fn relp_read_b(ctx: *mut relp_ctx, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    pub fn relp_read_minus_pi(ctx: *mut relp_ctx, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    /// -InverseMaintener::get_objective_function_value
    pub fn relp_read_minus_objective(ctx: *mut relp_ctx, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    /// ColumnComputationInfo::column of the last generated column
    pub fn relp_read_column(ctx: *mut relp_ctx, num: *mut i64, den: *mut i64, limbs: *mut u32, limb_cap: usize, limb_used: *mut usize) -> c_int;
    pub fn relp_log_size(ctx: *mut relp_ctx) -> i64;
    pub fn relp_read_log(ctx: *mut relp_ctx, first: i64, count: i64, out: *mut relp_iter_log) -> c_int;
    pub fn relp_stats_get(ctx: *mut relp_ctx, out: *mut relp_stats) -> c_int;
    pub fn relp_rows_removed(ctx: *mut relp_ctx, out: *mut i32, cap: i32) -> c_int;
    /// No counterpart in the reference: context-free K4 ratio test on DEVICE
    /// pointers that allocates nothing; m <= 2^20; result = [row or -1, status] in device memory, read after a stream sync
    pub fn relp_ratio_test_static(b_num: *const i64, b_den: *const i64, d_num: *const i64, d_den: *const i64,
                                  basis: *const i32, m: i32, result: *mut i32, stream: *mut c_void) -> c_int;
}

/// One entry of a rat array -> the reference's `RationalBig` payload.  `den > 0`: the int64 tier, value `num / den`.
/// `den == 0`: multi-limb, `num` is the word offset into `limbs` of `[header | numerator limbs | denominator limbs]`
/// (little-endian u32 limbs, header = nn | nd << 14 | sign << 31).  Values are canonical: nothing is re-reduced.
pub fn rat_to_big(num: i64, den: i64, limbs: &[u32]) -> num::BigRational {
    use num::bigint::{BigInt, BigUint, Sign};
    if den > 0 {
        return num::BigRational::new_raw(BigInt::from(num), BigInt::from(den));
    }
    let off = num as usize;
    let h = limbs[off];
    let (nn, nd) = ((h & 0x3fff) as usize, ((h >> 14) & 0x3fff) as usize);
    let n = BigUint::from_slice(&limbs[off + 1..off + 1 + nn]);
    let d = BigUint::from_slice(&limbs[off + 1 + nn..off + 1 + nn + nd]);
    let sign = if h >> 31 != 0 { Sign::Minus } else { Sign::Plus };
    num::BigRational::new_raw(BigInt::from_biguint(sign, n), BigInt::from_biguint(Sign::Plus, d))
}
