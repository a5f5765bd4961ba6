"""Shared test helpers: conversions between the Python restatement's LP (oracle/pyref.py) and
the array form `relp_load_matrix` takes, plus small random LP generators."""
import random
from fractions import Fraction as Fr

import numpy as np

from oracle import pyref
from rust_lp_b200.synth import ProviderArrays


def pa_from_lp(lp):
    col_ptr = [0]
    rows, vn, vd = [], [], []
    for c in lp.cols:
        for (i, v) in c:
            rows.append(i); vn.append(v.numerator); vd.append(v.denominator)
        col_ptr.append(len(rows))
    pr = [p[0] for p in lp.pivots]
    pc = [p[1] for p in lp.pivots]
    return ProviderArrays(lp.m, lp.n, col_ptr, rows, vn, vd,
                          [c.numerator for c in lp.cost], [c.denominator for c in lp.cost],
                          [c.numerator for c in lp.rhs], [c.denominator for c in lp.rhs],
                          pr, pc, getattr(lp, "n_normal", lp.n))


def lp_from_pa(pa):
    cols = []
    for j in range(pa.n):
        s, e = int(pa.col_ptr[j]), int(pa.col_ptr[j + 1])
        cols.append([(int(pa.row_idx[k]), Fr(int(pa.val_num[k]), int(pa.val_den[k]))) for k in range(s, e)])
    cost = [Fr(int(a), int(b)) for a, b in zip(pa.cost_num, pa.cost_den)]
    rhs = [Fr(int(a), int(b)) for a, b in zip(pa.rhs_num, pa.rhs_den)]
    lp = pyref.LP(pa.m, cols, cost, rhs, list(zip(map(int, pa.piv_rows), map(int, pa.piv_cols))))
    lp.n_normal = pa.n_normal
    return lp


def problem_2_lp():
    """src/tests/problem_2.rs:70-115"""
    cols = pyref.dense_rows_to_columns([[3, 2, 1, 0, 0], [5, 1, 1, 1, 0], [2, 5, 1, 0, 1]], 5)
    return pyref.matrix_data(cols, [1, 3, 4], [], 3, 0, 0, 0, [1] * 5, [None] * 5)


def problem_1_lp():
    """src/tests/problem_1.rs:284-343"""
    cols = pyref.dense_rows_to_columns([[0, -1, 1], [1, 0, 1]], 3)
    return pyref.matrix_data(cols, [6, 10], [], 1, 0, 0, 1, [1, 4, 9], [4, 2, None])


def random_lp(seed, n_eq, n_range, n_le, n_ge, nvars, density=0.5, maxval=5, frac=False, bounds=0.3):
    """Small random LP in MatrixData layout (E, R, L, G rows; b >= 0).  May be infeasible,
    unbounded or rank deficient -- that is the point."""
    rnd = random.Random(seed)
    ncon = n_eq + n_range + n_le + n_ge

    def val():
        v = rnd.randint(1, maxval) * rnd.choice([1, -1])
        return Fr(v, rnd.randint(1, 4)) if frac else Fr(v)

    cols = []
    for j in range(nvars):
        c = [(i, val()) for i in range(ncon) if rnd.random() < density]
        if not c and ncon:
            c = [(rnd.randrange(ncon), val())]
        cols.append(c)
    b = [Fr(rnd.randint(0, 20), rnd.randint(1, 3) if frac else 1) for _ in range(ncon)]
    ranges = [Fr(rnd.randint(1, 10)) for _ in range(n_range)]
    costs = [Fr(rnd.randint(-9, 9), rnd.randint(1, 3) if frac else 1) for _ in range(nvars)]
    ub = [Fr(rnd.randint(1, 15)) if rnd.random() < bounds else None for _ in range(nvars)]
    return pyref.matrix_data(cols, b, ranges, n_eq, n_range, n_le, n_ge, costs, ub)


def log_tuples(log_array):
    return [(chr(int(t)), int(j), int(r), int(l)) for (t, j, r, l, _T, _S) in log_array]


def feasible_lp(seed, n_eq, n_range, n_le, n_ge, nvars, density=0.5, maxval=5, frac=False, bounds=0.3,
                dup_rows=0):
    """Random LP that is feasible by construction (x0 >= 0 satisfies every row and bound).
    `dup_rows` appends copies of equality rows => rank deficient (phase_one.rs:223-260 path)."""
    rnd = random.Random(seed)
    ncon = n_eq + n_range + n_le + n_ge
    x0 = [Fr(rnd.randint(0, 6)) for _ in range(nvars)]

    def val():
        v = rnd.randint(1, maxval) * rnd.choice([1, -1])
        return Fr(v, rnd.randint(1, 4)) if frac else Fr(v)

    rows = []
    for i in range(ncon):
        r = {j: val() for j in range(nvars) if rnd.random() < density}
        if not r:
            r = {rnd.randrange(nvars): val()}
        ax = sum(v * x0[j] for j, v in r.items())
        if ax < 0:
            r = {j: -v for j, v in r.items()}
        rows.append(r)
    for k in range(dup_rows):
        src = rows[rnd.randrange(max(1, n_eq))]
        rows.insert(0, dict(src))
    n_eq += dup_rows
    ncon += dup_rows
    b, ranges = [], []
    for i, r in enumerate(rows):
        ax = sum(v * x0[j] for j, v in r.items())
        if i < n_eq:
            b.append(ax)
        elif i < n_eq + n_range:
            rg = Fr(rnd.randint(1, 10)); s0 = Fr(rnd.randint(0, int(rg)))
            b.append(ax + s0); ranges.append(rg)
        elif i < n_eq + n_range + n_le:
            b.append(ax + rnd.randint(0, 5))
        else:
            b.append(max(Fr(0), ax - rnd.randint(0, 5)))
    cols = [[(i, rows[i][j]) for i in range(ncon) if j in rows[i]] for j in range(nvars)]
    costs = [Fr(rnd.randint(-9, 9), rnd.randint(1, 3) if frac else 1) for _ in range(nvars)]
    ub = [x0[j] + rnd.randint(0, 9) + 1 if rnd.random() < bounds else None for j in range(nvars)]
    return pyref.matrix_data(cols, b, ranges, n_eq, n_range, n_le, n_ge, costs, ub)
