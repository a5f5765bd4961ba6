"""Synthetic sparse LPs of BASELINE.json config 5.

    min c^T x   s.t.  A x <= b,  x >= 0,     A: m x n, exactly `nnz_col` nonzeros per column

so every row has a `<=` slack that is the initial basis
(matrix_data.rs:432-452: `pivot_element_indices` = the upper-inequality slacks)
and phase 1 starts with zero artificial variables.

Variants
  "tu"      column j = `nnz_col` consecutive ones starting at row0_j (an interval
            matrix, hence totally unimodular: every basis inverse is integral with
            entries in {0, +1, -1}); the int64 tier never overflows.  Roofline config.
  "growth"  same pattern, values in 1..9: overflows int64 quickly; exercises escalation.

Draws come from a counter-based splitmix64 stream z(k), k = 0, 1, 2, ... so that numpy,
C++ and plain Python produce identical data:
  row0_j = z(j)            mod (m - nnz_col + 1)        j = 0..n-1
  b_i    = 1 + z(n + i)    mod 1000                     i = 0..m-1
  c_j    = -(1 + z(n + m + j) mod 100)                  j = 0..n-1
  v_{j,t}= 1 + z(2n + m + j*nnz_col + t) mod 9          growth only
"""
import numpy as np

DEFAULT_SEED = 20260101
_GAMMA = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(seed, k):
    """z(k) for an array (or scalar) of counters k."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + (np.asarray(k, dtype=np.uint64) + np.uint64(1)) * _GAMMA
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


class ProviderArrays:
    """CSC of ALL provider columns (structural, then virtual slacks materialised),
    the arrays `relp_load_matrix` takes (include/relp_b200.h)."""

    def __init__(self, m, n, col_ptr, row_idx, val_num, val_den, cost_num, cost_den,
                 rhs_num, rhs_den, piv_rows, piv_cols, n_normal):
        self.m, self.n = int(m), int(n)
        self.col_ptr = np.ascontiguousarray(col_ptr, dtype=np.int64)
        self.row_idx = np.ascontiguousarray(row_idx, dtype=np.int32)
        self.val_num = np.ascontiguousarray(val_num, dtype=np.int64)
        self.val_den = np.ascontiguousarray(val_den, dtype=np.int64)
        self.cost_num = np.ascontiguousarray(cost_num, dtype=np.int64)
        self.cost_den = np.ascontiguousarray(cost_den, dtype=np.int64)
        self.rhs_num = np.ascontiguousarray(rhs_num, dtype=np.int64)
        self.rhs_den = np.ascontiguousarray(rhs_den, dtype=np.int64)
        self.piv_rows = np.ascontiguousarray(piv_rows, dtype=np.int32)
        self.piv_cols = np.ascontiguousarray(piv_cols, dtype=np.int32)
        self.n_normal = int(n_normal)
        self.rhs_exact = {}      # row -> Fraction for right-hand sides beyond int64/int64 (rhs_num/rhs_den hold 0/1 there): relp_set_rhs

    @property
    def nnz(self):
        return int(self.col_ptr[-1])


def synthetic_lp(m, n, nnz_col=20, variant="tu", seed=DEFAULT_SEED):
    """Build config 5: returns ProviderArrays with n + m columns (n structural + m slacks)."""
    assert m >= nnz_col and variant in ("tu", "growth")
    j = np.arange(n, dtype=np.uint64)
    row0 = (splitmix64(seed, j) % np.uint64(m - nnz_col + 1)).astype(np.int64)
    b = 1 + (splitmix64(seed, np.uint64(n) + np.arange(m, dtype=np.uint64)) % np.uint64(1000)).astype(np.int64)
    c = -(1 + (splitmix64(seed, np.uint64(n + m) + j) % np.uint64(100)).astype(np.int64))
    rows_struct = (row0[:, None] + np.arange(nnz_col, dtype=np.int64)[None, :]).reshape(-1)
    if variant == "tu":
        vals_struct = np.ones(n * nnz_col, dtype=np.int64)
    else:
        k = np.uint64(2 * n + m) + np.arange(n * nnz_col, dtype=np.uint64)
        vals_struct = 1 + (splitmix64(seed, k) % np.uint64(9)).astype(np.int64)
    col_ptr = np.concatenate([np.arange(n + 1, dtype=np.int64) * nnz_col,
                              n * nnz_col + np.arange(1, m + 1, dtype=np.int64)])
    row_idx = np.concatenate([rows_struct, np.arange(m, dtype=np.int64)]).astype(np.int32)
    val_num = np.concatenate([vals_struct, np.ones(m, dtype=np.int64)])
    val_den = np.ones_like(val_num)
    cost_num = np.concatenate([c, np.zeros(m, dtype=np.int64)])
    cost_den = np.ones_like(cost_num)
    return ProviderArrays(m, n + m, col_ptr, row_idx, val_num, val_den, cost_num, cost_den,
                          b, np.ones(m, dtype=np.int64),
                          np.arange(m, dtype=np.int32), n + np.arange(m, dtype=np.int32), n)
