//! `GpuCarry`: a drop-in `InverseMaintener` for RELP's two-phase simplex whose Carry matrix lives on a B200 and is
//! updated by librelp_b200.so, plus `GpuRule<R>`: the three `PivotRule`s evaluated on the device.
//!
//! Every method of `InverseMaintener` (relp: src/algorithm/two_phase/tableau/inverse_maintenance/mod.rs:29-267),
//! `ColumnComputationInfo` (:270-275), `BasisInverse` (carry/mod.rs:68-157, on `GpuBasisInverse`) and `PivotRule`
//! (strategy/pivot_rule.rs:21-33) is implemented; nothing is elided.
//! NEVER COMPILED: there is no Rust toolchain where this was written.  The C ABI underneath is what is tested
//! (tests/test_gpu_parity.py drives the same entry points in the same order through ctypes).
//!
//! Usage, instead of `Carry<RationalBig, BasisInverseRows<_>>`:
//! ```ignore
//! let result = matrix_data.solve_relaxation::<GpuCarry>();      // algorithm/two_phase/mod.rs:30-75
//! ```
//! Friction of the trait surface and how it is met:
//!  * `create_for_partially_artificial` gets `b` and index lists but not the provider; the matrix reaches the IM
//!    through `after_basis_change(&kind)` (first call: every `kind.original_column(j)` is uploaded) -- until then
//!    `cost_difference` / `generate_column` upload the single column they are given.
//!  * `PivotRule::select_primal_pivot_column` only sees `&Tableau`; the context travels in a thread-local
//!    (`CURRENT`), set by the IM constructors -- the reference is single-threaded (no Send/Sync bounds anywhere).
//!  * `Tableau::select_primal_pivot_row` is an inherent method that reads the generated column on the host:
//!    `generate_column` therefore returns the column as a host `SparseVector` (one read-back of m rationals per pivot).
#![feature(generic_associated_types)]
use std::cell::Cell;
use std::fmt;

use num::{BigInt, BigRational, Zero};
use relp::algorithm::two_phase::matrix_provider::column::{Column, OrderedColumn};
use relp::algorithm::two_phase::matrix_provider::filter::Filtered;
use relp::algorithm::two_phase::matrix_provider::MatrixProvider;
use relp::algorithm::two_phase::strategy::pivot_rule::{FirstProfitable, FirstProfitableWithMemory, PivotRule, SteepestDescent};
use relp::algorithm::two_phase::tableau::inverse_maintenance::{ops, ColumnComputationInfo, InverseMaintener};
use relp::algorithm::two_phase::tableau::kind::Kind;
use relp::algorithm::two_phase::tableau::Tableau;
use relp::data::linear_algebra::vector::{DenseVector, SparseVector, Vector};
use relp::data::linear_algebra::SparseTuple;
use relp::data::number_types::rational::{Rational64, RationalBig};
use relp_b200_sys as sys;

thread_local! { static CURRENT: Cell<*mut sys::relp_ctx> = Cell::new(std::ptr::null_mut()); }

fn check(status: i32) -> i32 {
    if status < 0 {
        let msg = unsafe { std::ffi::CStr::from_ptr(sys::relp_last_error()) }.to_string_lossy().into_owned();
        panic!("librelp_b200: status {}: {}", status, msg);     // the reference panics on the same conditions (carry/mod.rs:291)
    }
    status
}

/// Host number -> (numerator, denominator) in lowest terms, as relp_load_matrix wants them.  The conversions into
/// `RationalBig` are the ones the reference's own `ops` aliases guarantee (ops.rs:66-90): `From<&G>` for column values,
/// `From<Rhs>` for the right-hand side, `0 + cost` for costs.  Input data is `Rational64`-sized (io/mps parses i64s).
fn to_pair(big: &RationalBig) -> (i64, i64) {
    use num::ToPrimitive;
    (big.0.numer().to_i64().expect("matrix value does not fit Rational64"), big.0.denom().to_i64().expect("matrix value does not fit Rational64"))
}
fn zero() -> RationalBig { RationalBig(BigRational::zero()) }

/// A right-hand side as the rat array of `relp_set_rhs`: values that fit int64/int64 as such, the others as limb records
/// `[nn | nd << 14, numerator limbs, denominator limbs]` (the reference's `Rhs` is a `RationalBig`; presolve makes big ones).
fn rhs_rat_array(values: &[RationalBig]) -> (Vec<i64>, Vec<i64>, Vec<u32>) {
    use num::ToPrimitive;
    let (mut num, mut den, mut limbs) = (Vec::with_capacity(values.len()), Vec::with_capacity(values.len()), Vec::new());
    for v in values {
        match (v.0.numer().to_i64(), v.0.denom().to_i64()) {
            (Some(n), Some(d)) => { num.push(n); den.push(d); }
            _ => {
                assert!(v.0.numer().sign() != num::bigint::Sign::Minus, "right-hand side must be >= 0");
                let (n, d) = (v.0.numer().magnitude().to_u32_digits(), v.0.denom().magnitude().to_u32_digits());
                num.push(limbs.len() as i64); den.push(0);
                limbs.push(n.len() as u32 | (d.len() as u32) << 14);
                limbs.extend(n); limbs.extend(d);
            }
        }
    }
    (num, den, limbs)
}

/// placeholders for `relp_load_matrix` (0/1 where the value travels as a limb record) + the `relp_set_rhs` call if needed
fn upload_rhs(ctx: *mut sys::relp_ctx, values: &[RationalBig], load: impl FnOnce(&[i64], &[i64])) {
    let (num, den, limbs) = rhs_rat_array(values);
    let (pn, pd): (Vec<i64>, Vec<i64>) = num.iter().zip(&den).map(|(&n, &d)| if d == 0 { (0, 1) } else { (n, d) }).unzip();
    load(&pn, &pd);
    if !limbs.is_empty() {
        check(unsafe { sys::relp_set_rhs(ctx, values.len() as i32, num.as_ptr(), den.as_ptr(), limbs.as_ptr(), limbs.len()) });
    }
}

/// rat array entry -> `RationalBig` (values arrive canonical: `new_raw` is right)
fn from_rat(num: i64, den: i64, limbs: &[u32]) -> RationalBig {
    RationalBig(sys::rat_to_big(num, den, limbs))
}

/// Buffers of one read-back call.
struct Rats { num: Vec<i64>, den: Vec<i64>, limbs: Vec<u32> }
impl Rats {
    fn read(count: usize, f: impl Fn(*mut i64, *mut i64, *mut u32, usize, *mut usize) -> i32) -> Self {
        let mut r = Rats { num: vec![0; count.max(1)], den: vec![1; count.max(1)], limbs: vec![0; 1 << 16] };
        loop {
            let mut used = 0usize;
            let st = f(r.num.as_mut_ptr(), r.den.as_mut_ptr(), r.limbs.as_mut_ptr(), r.limbs.len(), &mut used);
            if st == sys::RELP_ERR_ARGUMENT && used > r.limbs.len() { r.limbs = vec![0; 2 * used]; continue; }
            check(st);
            return r;
        }
    }
    fn get(&self, i: usize) -> RationalBig { from_rat(self.num[i], self.den[i], &self.limbs) }
}

/// CSC arrays of a set of columns, in the layout of `relp_load_matrix`.
#[derive(Default)]
struct Csc { col_ptr: Vec<i64>, row_idx: Vec<i32>, vn: Vec<i64>, vd: Vec<i64> }
impl Csc {
    fn new() -> Self { Csc { col_ptr: vec![0], ..Default::default() } }
    fn push<C: Column>(&mut self, column: &C) where RationalBig: ops::Column<C::F> {
        for (i, v) in column.iter() {
            let (n, d) = to_pair(&RationalBig::from(v));
            self.row_idx.push(*i as i32); self.vn.push(n); self.vd.push(d);
        }
        self.col_ptr.push(self.row_idx.len() as i64);
    }
}

/// The generated column `B^-1 a_j` as the host needs it (tableau/mod.rs:221-247 reads it for the ratio test).
#[derive(Debug)]
pub struct GpuColumn { column: SparseVector<RationalBig, RationalBig> }
impl ColumnComputationInfo<RationalBig> for GpuColumn {
    fn column(&self) -> &SparseVector<RationalBig, RationalBig> { &self.column }
    fn into_column(self) -> SparseVector<RationalBig, RationalBig> { self.column }
}

/// Device-resident `Carry<RationalBig, BasisInverseRows<RationalBig>>` (carry/mod.rs:45-65).
pub struct GpuCarry {
    ctx: *mut sys::relp_ctx,
    m: usize,
    /// `b`, read back lazily: `get_constraint_value` hands out references (inverse_maintenance/mod.rs:265)
    b_cache: std::cell::RefCell<Option<Vec<RationalBig>>>,
    /// data of the artificial tableau kept until the provider is known (see module docs)
    pending: Option<Pending>,
    matrix_loaded: bool,
    /// column uploaded by the last `generate_column` (tableau index), so that `change_basis` pivots on it
    generated: Cell<Option<usize>>,
}
struct Pending { rhs: Vec<RationalBig>, piv_rows: Vec<i32>, piv_cols: Vec<i32> }

impl GpuCarry {
    fn fresh(m: usize) -> Self {
        let ctx = unsafe { sys::relp_create(0, 0, 1, 0) };
        assert!(!ctx.is_null(), "no CUDA device: librelp_b200 has no CPU path");
        CURRENT.with(|c| c.set(ctx));
        GpuCarry { ctx, m, b_cache: Default::default(), pending: None, matrix_loaded: false, generated: Cell::new(None) }
    }
    fn invalidate(&self) { *self.b_cache.borrow_mut() = None; }

    /// Upload provider columns `0..n` (+ costs, right-hand side) and build the artificial tableau that was asked for.
    fn load<K: Kind>(&mut self, kind: &K) where RationalBig: ops::Column<<K::Column as Column>::F> {
        let p = self.pending.take().expect("matrix already loaded");
        // tableau columns of an artificial tableau: the artificial ones first (partially.rs:72-88); the library wants
        // the PROVIDER's columns, i.e. everything behind them
        let nr_art = self.m - p.piv_rows.len();
        let mut csc = Csc::new();
        let (mut cn, mut cd) = (Vec::new(), Vec::new());
        for j in nr_art..kind.nr_columns() {
            csc.push(&kind.original_column(j));
            cn.push(0); cd.push(1);        // phase-one costs are the kind's (0 for provider columns); real costs arrive in from_artificial
        }
        let (ctx, m, n) = (self.ctx, self.m as i32, (kind.nr_columns() - nr_art) as i32);
        upload_rhs(ctx, &p.rhs, |rn, rd| { check(unsafe { sys::relp_load_matrix(ctx, m, n, csc.col_ptr.as_ptr(),
            csc.row_idx.as_ptr(), csc.vn.as_ptr(), csc.vd.as_ptr(), cn.as_ptr(), cd.as_ptr(), rn.as_ptr(), rd.as_ptr()) }); });
        check(unsafe { sys::relp_carry_create_partially_artificial(self.ctx, p.piv_rows.len() as i32, p.piv_rows.as_ptr(), p.piv_cols.as_ptr()) });
        self.matrix_loaded = true;
    }

    fn load_provider<'a, MP: MatrixProvider>(&mut self, provider: &'a MP)
    where RationalBig: ops::Column<<MP::Column as Column>::F> + ops::Rhs<MP::Rhs> + ops::Cost<MP::Cost<'a>> {
        let (m, n) = (provider.nr_rows(), provider.nr_columns());
        let mut csc = Csc::new();
        let (mut cn, mut cd) = (Vec::with_capacity(n), Vec::with_capacity(n));
        for j in 0..n {
            csc.push(&provider.column(j));
            let (a, b) = to_pair(&(zero() + provider.cost_value(j)));
            cn.push(a); cd.push(b);
        }
        let rhs = provider.right_hand_side();
        let values: Vec<RationalBig> = (0..m).map(|i| RationalBig::from(rhs[i].clone())).collect();
        let ctx = self.ctx;
        upload_rhs(ctx, &values, |rn, rd| { check(unsafe { sys::relp_load_matrix(ctx, m as i32, n as i32, csc.col_ptr.as_ptr(), csc.row_idx.as_ptr(),
            csc.vn.as_ptr(), csc.vd.as_ptr(), cn.as_ptr(), cd.as_ptr(), rn.as_ptr(), rd.as_ptr()) }); });
        self.m = m;
        self.matrix_loaded = true;
    }

    fn basis(&self) -> Vec<i32> {
        let mut out = vec![0i32; self.m];
        check(unsafe { sys::relp_read_basis(self.ctx, out.as_mut_ptr()) });
        out
    }
}

impl Drop for GpuCarry {
    fn drop(&mut self) {
        CURRENT.with(|c| if c.get() == self.ctx { c.set(std::ptr::null_mut()) });
        unsafe { sys::relp_destroy(self.ctx) }
    }
}

impl fmt::Display for GpuCarry {
    fn fmt(&self, f: &mut fmt::Formatter<'_>) -> fmt::Result {
        writeln!(f, "GpuCarry: m = {}, objective = {}", self.m, self.get_objective_function_value())?;
        for (i, (col, v)) in self.basis().iter().zip(self.b().iter()).enumerate() {
            writeln!(f, "  row {:>5}: basis column {:>6}, b = {}", i, col, v)?;
        }
        Ok(())
    }
}

impl InverseMaintener for GpuCarry {
    type F = RationalBig;
    type ColumnComputationInfo = GpuColumn;

    /// carry/mod.rs:358-379: every row gets an artificial column.
    fn create_for_fully_artificial<Rhs: relp::data::linear_algebra::traits::Element>(rhs: DenseVector<Rhs>) -> Self
    where Self::F: ops::Rhs<Rhs> {
        Self::create_for_partially_artificial(&(0..rhs.len()).collect::<Vec<_>>(), &[], rhs, (0..0).collect())
    }

    /// carry/mod.rs:381-426.  The tableau itself is built on the device once the matrix is known (`after_basis_change`
    /// or the first column operation); only the index lists and `b` are recorded here.
    fn create_for_partially_artificial<G: relp::data::linear_algebra::traits::Element>(
        artificial: &[usize], basis: &[(usize, usize)], b: DenseVector<G>, _basis_indices: Vec<usize>,
    ) -> Self where Self::F: ops::Rhs<G> {
        let m = artificial.len() + basis.len();
        let mut im = Self::fresh(m);
        let rhs = (0..m).map(|i| RationalBig::from(b[i].clone())).collect();
        im.pending = Some(Pending {
            rhs,
            piv_rows: basis.iter().map(|&(row, _)| row as i32).collect(),
            piv_cols: basis.iter().map(|&(_, column)| column as i32).collect(),
        });
        im
    }

    /// carry/mod.rs:428-463: the Carry of a given basis (exact Gauss-Jordan on the device, relp_carry_from_basis).
    fn from_basis<'a, MP: MatrixProvider>(basis: &[usize], provider: &'a MP) -> Self
    where Self::F: ops::Column<<MP::Column as Column>::F> + ops::Rhs<MP::Rhs> + ops::Cost<MP::Cost<'a>> + ops::Column<MP::Rhs>, MP::Rhs: 'static {
        let mut im = Self::fresh(provider.nr_rows());
        im.load_provider(provider);
        let columns: Vec<i32> = basis.iter().map(|&j| j as i32).collect();
        check(unsafe { sys::relp_carry_from_basis(im.ctx, columns.as_ptr()) });
        im
    }

    /// carry/mod.rs:465-482: pairs (row, column) sorted by row, then `from_basis`.
    fn from_basis_pivots<'a, MP: MatrixProvider>(basis: &[(usize, usize)], provider: &'a MP) -> Self
    where Self::F: ops::Column<<MP::Column as Column>::F> + ops::Rhs<MP::Rhs> + ops::Cost<MP::Cost<'a>> + ops::Column<MP::Rhs>,
          MP::Rhs: 'static + relp::algorithm::two_phase::matrix_provider::column::ColumnNumber {
        let mut pairs = basis.to_vec();
        pairs.sort_by_key(|&(row, _)| row);
        Self::from_basis(&pairs.into_iter().map(|(_, column)| column).collect::<Vec<_>>(), provider)
    }

    /// carry/mod.rs:484-510 (+ :214-271): drop the artificial columns, recompute -pi and -obj from the real costs (K5).
    fn from_artificial<'provider, MP: MatrixProvider>(mut artificial: Self, provider: &'provider MP, nr_artificial: usize) -> Self
    where Self::F: ops::InternalHR + ops::Column<<MP::Column as Column>::F> + ops::Cost<MP::Cost<'provider>> {
        let _ = nr_artificial;                       // the library knows how many artificial columns its tableau has
        artificial.upload_costs(provider);
        check(unsafe { sys::relp_carry_from_artificial(artificial.ctx) });
        artificial.invalidate();
        artificial
    }

    /// carry/mod.rs:650-689 with filter/generic_wrapper.rs:224-300: additionally remove the redundant rows.
    fn from_artificial_remove_rows<'a, MP: Filtered>(mut artificial: Self, rows_removed: &'a MP, nr_artificial: usize) -> Self
    where Self::F: ops::Column<<<MP as MatrixProvider>::Column as Column>::F> + ops::Cost<MP::Cost<'a>> {
        let _ = nr_artificial;
        artificial.upload_costs(rows_removed);
        let rows: Vec<i32> = rows_removed.filtered_rows().iter().map(|&i| i as i32).collect();
        check(unsafe { sys::relp_carry_from_artificial_remove_rows(artificial.ctx, rows.len() as i32, rows.as_ptr()) });
        artificial.m -= rows.len();
        artificial.invalidate();
        artificial
    }

    /// carry/mod.rs:549-570 = K1.  `column` was generated by this IM (its values stayed on the device).
    fn change_basis(&mut self, pivot_row_index: usize, pivot_column_index: usize, column: GpuColumn, cost: RationalBig) -> usize {
        let _ = (column, cost);                      // both are on the device already (relp_generate_column re-prices the column)
        debug_assert_eq!(self.generated.get(), Some(pivot_column_index), "change_basis must follow generate_column of the same column");
        let mut leaving = -1i32;
        check(unsafe { sys::relp_change_basis(self.ctx, pivot_row_index as i32, &mut leaving) });
        self.invalidate();
        leaving as usize
    }

    /// carry/mod.rs:572-577: -pi . column (the kind adds the column's own cost, tableau/mod.rs:102-108).
    fn cost_difference<G, C: Column<F=G> + OrderedColumn>(&self, original_column: &C) -> RationalBig where Self::F: ops::Column<G> {
        let pi = Rats::read(self.m, |n, d, l, cap, used| unsafe { sys::relp_read_minus_pi(self.ctx, n, d, l, cap, used) });
        let mut total = zero();
        for (i, v) in original_column.iter() {
            total += &pi.get(*i) * v;                 // `&F: Mul<&G, Output = F>` (ops.rs:66-76)
        }
        total
    }

    /// carry/mod.rs:579-587 -> basis_inverse_rows.rs:144-155 = K2.  `original_column` must be a column of the loaded
    /// matrix: its tableau index is found through `CURRENT_COLUMN`, set by `GpuRule` (the only caller on the solve path
    /// is `Tableau::generate_column(j)`, tableau/mod.rs:122-126, right after the pivot rule chose j).
    fn generate_column<G, C: Column<F=G> + OrderedColumn>(&self, original_column: C) -> GpuColumn where Self::F: ops::Column<G> {
        let _ = original_column;
        let j = CURRENT_COLUMN.with(|c| c.get()).expect("generate_column before the pivot rule selected a column");
        check(unsafe { sys::relp_generate_column(self.ctx, j as i32) });
        self.generated.set(Some(j));
        let d = Rats::read(self.m, |n, dn, l, cap, used| unsafe { sys::relp_read_column(self.ctx, n, dn, l, cap, used) });
        let data = (0..self.m).filter(|&i| d.num[i] != 0 || d.den[i] == 0).map(|i| (i, d.get(i))).collect();
        GpuColumn { column: SparseVector::new(data, self.m) }
    }

    /// basis_inverse_rows.rs:157-173.
    fn generate_element<C: Column + OrderedColumn>(&self, i: usize, original_column: C) -> Option<RationalBig> where Self::F: ops::Column<C::F> {
        let _ = original_column;
        let j = CURRENT_COLUMN.with(|c| c.get()).expect("generate_element needs the tableau index of the column");
        let (mut n, mut d) = (0i64, 1i64);
        let mut limbs = vec![0u32; 1 << 16];
        check(unsafe { sys::relp_generate_element(self.ctx, i as i32, j as i32, &mut n, &mut d, limbs.as_mut_ptr(), limbs.len()) });
        if n == 0 && d == 1 { None } else { Some(from_rat(n, d, &limbs)) }
    }

    /// inverse_maintenance/mod.rs:236-239: the first call is where the IM meets the matrix (see module docs).
    fn after_basis_change<K: Kind>(&mut self, kind: &K) where Self::F: ops::Column<<<K as Kind>::Column as Column>::F> {
        if !self.matrix_loaded { self.load(kind); }
    }

    /// carry/mod.rs:616-625.
    fn current_bfs(&self) -> Vec<SparseTuple<RationalBig>> {
        let basis = self.basis();
        let mut out: Vec<_> = self.b().iter().enumerate().filter(|(_, v)| !v.0.is_zero()).map(|(i, v)| (basis[i] as usize, v.clone())).collect();
        out.sort_by_key(|&(j, _)| j);
        out
    }

    /// carry/mod.rs:627-629.
    fn basis_column_index_for_row(&self, row: usize) -> usize { self.basis()[row] as usize }

    /// carry/mod.rs:631-634.
    fn b(&self) -> DenseVector<RationalBig> {
        self.fill_b();
        DenseVector::new(self.b_cache.borrow().as_ref().unwrap().clone(), self.m)
    }

    /// carry/mod.rs:636-638.
    fn get_objective_function_value(&self) -> RationalBig {
        let r = Rats::read(1, |n, d, l, cap, used| unsafe { sys::relp_read_minus_objective(self.ctx, n, d, l, cap, used) });
        -r.get(0)
    }

    /// carry/mod.rs:640-642.
    fn get_constraint_value(&self, i: usize) -> &RationalBig {
        self.fill_b();
        // the cache lives as long as `self` and is only replaced through `&mut self` methods
        unsafe { &*(&self.b_cache.borrow().as_ref().unwrap()[i] as *const RationalBig) }
    }
}

impl GpuCarry {
    fn fill_b(&self) {
        if self.b_cache.borrow().is_none() {
            let r = Rats::read(self.m, |n, d, l, cap, used| unsafe { sys::relp_read_b(self.ctx, n, d, l, cap, used) });
            *self.b_cache.borrow_mut() = Some((0..self.m).map(|i| r.get(i)).collect());
        }
    }
    /// phase two needs the provider's real costs (phase one ran with the artificial costs): reload with them
    fn upload_costs<'a, MP: MatrixProvider>(&mut self, provider: &'a MP) where RationalBig: ops::Cost<MP::Cost<'a>> {
        let n = provider.nr_columns();
        let (mut cn, mut cd) = (Vec::with_capacity(n), Vec::with_capacity(n));
        for j in 0..n {
            let (a, b) = to_pair(&(zero() + provider.cost_value(j)));
            cn.push(a); cd.push(b);
        }
        check(unsafe { sys::relp_set_costs(self.ctx, n as i32, cn.as_ptr(), cd.as_ptr()) });
    }
}

/// `BasisInverse` (carry/mod.rs:68-157) on its own, for callers that keep the reference's host-side `Carry<F, BI>` and
/// only want `B^-1` on the device: a `GpuCarry` whose right-hand side is zero.  The nine methods map one to one.
/// (Inside this crate `GpuCarry` replaces `Carry` AND its `BasisInverse`: `b`, `-pi` and the objective live on the
/// device next to `B^-1`, which is what makes the fused loop possible.)
pub struct GpuBasisInverse { inner: GpuCarry, nr_columns: usize }

impl GpuBasisInverse {
    /// load `columns` as the whole provider (costs 0, right-hand side 0)
    fn with_columns<C: Column + OrderedColumn>(m: usize, columns: &[C]) -> GpuCarry where RationalBig: ops::Column<C::F> {
        let mut im = GpuCarry::fresh(m);
        let mut csc = Csc::new();
        for column in columns { csc.push(column); }
        let n = columns.len();
        let (cn, cd, rn, rd) = (vec![0i64; n], vec![1i64; n], vec![0i64; m], vec![1i64; m]);
        check(unsafe { sys::relp_load_matrix(im.ctx, m as i32, n as i32, csc.col_ptr.as_ptr(), csc.row_idx.as_ptr(), csc.vn.as_ptr(),
            csc.vd.as_ptr(), cn.as_ptr(), cd.as_ptr(), rn.as_ptr(), rd.as_ptr()) });
        im.matrix_loaded = true;
        im
    }
}

impl fmt::Display for GpuBasisInverse {
    fn fmt(&self, f: &mut fmt::Formatter<'_>) -> fmt::Result {
        for i in 0..self.inner.m { writeln!(f, "{:?}", self.basis_inverse_row(i))?; }
        Ok(())
    }
}

impl relp::algorithm::two_phase::tableau::inverse_maintenance::carry::BasisInverse for GpuBasisInverse {
    type F = RationalBig;
    type ColumnComputationInfo = GpuColumn;

    /// basis_inverse_rows.rs:86-91: the identity = the fully artificial tableau of a provider with one (unit) column per row
    fn identity(m: usize) -> Self {
        let units: Vec<SparseVector<RationalBig, RationalBig>> = (0..m).map(|i| SparseVector::new(vec![(i, RationalBig::from(1))], m)).collect();
        let mut inner = Self::with_columns(m, &units);
        check(unsafe { sys::relp_carry_create_partially_artificial(inner.ctx, 0, std::ptr::null(), std::ptr::null()) });
        inner.invalidate();
        GpuBasisInverse { inner, nr_columns: m }
    }

    /// basis_inverse_rows.rs:93-128 (Gauss-Jordan over the given columns) = relp_carry_from_basis over columns 0..m
    fn invert<C: Column + OrderedColumn>(columns: Vec<C>) -> Self where Self::F: ops::Column<C::F> {
        let m = columns.len();
        let inner = Self::with_columns(m, &columns);
        let basis: Vec<i32> = (0..m as i32).collect();
        check(unsafe { sys::relp_carry_from_basis(inner.ctx, basis.as_ptr()) });
        GpuBasisInverse { inner, nr_columns: m }
    }

    /// basis_inverse_rows.rs:130-142 = K1 without the `b` and cost-row parts (both are zero here)
    fn change_basis(&mut self, pivot_row_index: usize, column: GpuColumn) {
        let _ = column;
        let mut leaving = -1i32;
        check(unsafe { sys::relp_change_basis(self.inner.ctx, pivot_row_index as i32, &mut leaving) });
    }

    fn generate_column<C: Column + OrderedColumn>(&self, original_column: C) -> GpuColumn where Self::F: ops::Column<C::F> {
        self.inner.generate_column(original_column)
    }

    fn generate_element<C: Column + OrderedColumn>(&self, i: usize, original_column: C) -> Option<RationalBig> where Self::F: ops::Column<C::F> {
        self.inner.generate_element(i, original_column)
    }

    /// basis_inverse_rows.rs:175-177: the explicit inverse never degrades
    fn should_refactor(&self) -> bool { false }

    /// basis_inverse_rows.rs:179-181: row `row` of `B^-1` = Carry block row `row`, columns 1..=m (column 0 is `b`)
    fn basis_inverse_row(&self, row: usize) -> SparseVector<RationalBig, RationalBig> {
        let m = self.inner.m;
        let r = Rats::read(m, |n, d, l, cap, used| unsafe { sys::relp_read_block(self.inner.ctx, row as i32, 1, 1, m as i32, n, d, l, cap, used) });
        SparseVector::new((0..m).filter(|&k| r.num[k] != 0 || r.den[k] == 0).map(|k| (k, r.get(k))).collect(), m)
    }

    fn m(&self) -> usize { let _ = self.nr_columns; self.inner.m }
}

thread_local! { static CURRENT_COLUMN: Cell<Option<usize>> = Cell::new(None); }

/// The reference's pivot rules evaluated on the device (K3): `R` only selects which one.
pub struct GpuRule<R> { rule: i32, _marker: std::marker::PhantomData<R> }
pub trait RuleCode { const CODE: i32; }
impl RuleCode for FirstProfitable { const CODE: i32 = sys::RELP_FIRST_PROFITABLE; }
impl RuleCode for FirstProfitableWithMemory { const CODE: i32 = sys::RELP_FIRST_PROFITABLE_WITH_MEMORY; }
impl RuleCode for SteepestDescent { const CODE: i32 = sys::RELP_STEEPEST_DESCENT; }

impl<R: RuleCode> PivotRule for GpuRule<R> {
    fn new() -> Self { GpuRule { rule: R::CODE, _marker: Default::default() } }

    /// strategy/pivot_rule.rs:38-127 over tableau/mod.rs:102-108, on the device; the memory of
    /// FirstProfitableWithMemory lives in the context.
    fn select_primal_pivot_column<IM, K>(&mut self, _tableau: &Tableau<IM, K>) -> Option<SparseTuple<IM::F>>
    where IM: InverseMaintener<F: ops::Column<<K::Column as Column>::F> + ops::Cost<K::Cost>>, K: Kind {
        let ctx = CURRENT.with(|c| c.get());
        assert!(!ctx.is_null(), "GpuRule needs a live GpuCarry on this thread");
        let (mut j, mut n, mut d) = (-1i32, 0i64, 1i64);
        let mut limbs = vec![0u32; 1 << 16];
        let st = check(unsafe { sys::relp_price(ctx, self.rule, &mut j, &mut n, &mut d, limbs.as_mut_ptr(), limbs.len()) });
        if st == sys::RELP_OPTIMAL { CURRENT_COLUMN.with(|c| c.set(None)); return None; }
        CURRENT_COLUMN.with(|c| c.set(Some(j as usize)));
        // IM::F is RationalBig whenever IM = GpuCarry: the only IM this rule can work with
        assert_eq!(std::any::type_name::<IM::F>(), std::any::type_name::<RationalBig>());
        let value = std::mem::ManuallyDrop::new(from_rat(n, d, &limbs));
        Some((j as usize, unsafe { std::mem::transmute_copy::<RationalBig, IM::F>(&*value) }))
    }
}

/// The fused path: the whole of `solve_relaxation` (algorithm/two_phase/mod.rs:30-75) on the device, for callers that
/// do not need the step-wise traits.  Returns the status (RELP_OPTIMAL / RELP_UNBOUNDED / RELP_INFEASIBLE).
pub fn solve_relaxation_fused(im: &mut GpuCarry, piv_rows: &[i32], piv_cols: &[i32]) -> i32 {
    let st = check(unsafe { sys::relp_solve_relaxation(im.ctx, sys::RELP_FIRST_PROFITABLE_WITH_MEMORY, sys::RELP_STEEPEST_DESCENT,
        piv_rows.len() as i32, piv_rows.as_ptr(), piv_cols.as_ptr(), -1) });
    im.invalidate();
    st
}

#[allow(dead_code)]
fn _type_checks() { let _: Option<Rational64> = None; let _: Option<BigInt> = None; }
