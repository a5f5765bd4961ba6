"""ORACLE (test infrastructure, not product) -- independent Python restatement.

A `fractions.Fraction` restatement of RELP's two-phase revised simplex over
`Carry<RationalBig, BasisInverseRows>`.  It exists to cross-check the C++
oracle (`oracle/simplex.cpp`) pivot-for-pivot on small inputs: two independent
restatements agreeing is the substitute for the un-runnable Rust reference
(no rustc/cargo in this image).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may
import this module.  The product path (`rust-lp_b200/`) never does.

Reference files followed (all under /root/reference/src/algorithm/two_phase/):
  tableau/kind/artificial/partially.rs:125-205   Partially::new
  tableau/inverse_maintenance/carry/mod.rs       Carry (update_b 283-313,
      update_minus_pi_and_obj 326-333, create_for_partially_artificial 381-426,
      from_artificial 484-510, change_basis 549-570, cost_difference 572-577,
      current_bfs 616-625, remove-rows specialisation 650-689)
  tableau/inverse_maintenance/carry/basis_inverse_rows.rs  (normalize 42-51,
      row_reduce 63-83, generate_column 144-155, generate_element 157-173,
      remove_basis_part 194-205)
  strategy/pivot_rule.rs:38-127                  the three PivotRules
  tableau/mod.rs:102-108, 122-126, 221-247       relative_cost, generate_column,
                                                 select_primal_pivot_row
  phase_one.rs:125-170, 223-260                  primal, artificial clean-up
  phase_two.rs:22-51                             primal
  mod.rs:30-75                                   solve_relaxation

Parity status: pinned against the reference's own exact fixtures
(src/tests/problem_1.rs, src/tests/problem_2.rs, tableau/mod.rs tests,
pivot_rule.rs tests) in tests/test_oracle_fixtures.py.
"""
from fractions import Fraction as Fr

FIRST_PROFITABLE = 0
FIRST_PROFITABLE_WITH_MEMORY = 1
STEEPEST_DESCENT = 2


class LP:
    """A `MatrixProvider` view (matrix_provider/mod.rs:37-136): all provider
    columns (structural then virtual slacks, matrix_data.rs:37-52) materialised."""

    def __init__(self, m, cols, cost, rhs, pivots):
        self.m = m
        self.cols = cols          # list of ascending [(row, Fr)]
        self.n = len(cols)
        self.cost = cost          # list of Fr (None in the reference == 0)
        self.rhs = rhs            # list of Fr, >= 0
        self.pivots = pivots      # pivot_element_indices(): [(row, col)] ascending row


class Carry:
    """carry/mod.rs:45-65 with BasisInverseRows rows as dicts {col: Fr}."""

    def __init__(self, minus_obj, minus_pi, b, basis_indices, rows):
        self.minus_obj = minus_obj
        self.minus_pi = minus_pi
        self.b = b
        self.basis_indices = basis_indices
        self.rows = rows

    def m(self):
        return len(self.b)

    # basis_inverse_rows.rs:157-173
    def generate_element(self, i, column):
        row = self.rows[i]
        tot = Fr(0)
        for (k, v) in column:
            a = row.get(k)
            if a is not None:
                tot += a * v
        return tot

    # basis_inverse_rows.rs:144-155 -- ascending rows, exact zeros dropped
    def generate_column(self, column):
        out = []
        for i in range(self.m()):
            e = self.generate_element(i, column)
            if e != 0:
                out.append((i, e))
        return out

    # carry/mod.rs:572-577
    def cost_difference(self, column):
        tot = Fr(0)
        for (i, v) in column:
            tot += self.minus_pi[i] * v
        return tot

    # carry/mod.rs:549-570
    def change_basis(self, r, j, col, rc):
        d = dict(col)
        piv = d[r]
        # update_b (283-313)
        self.b[r] /= piv
        for (i, v) in col:
            if i != r:
                self.b[i] -= v * self.b[r]
        # BasisInverseRows::change_basis (131-142)
        prow = {k: v / piv for k, v in self.rows[r].items()}
        self.rows[r] = prow
        for (i, v) in col:
            if i == r:
                continue
            row = self.rows[i]
            for k, pv in prow.items():
                nv = row.get(k, Fr(0)) - v * pv
                if nv == 0:
                    row.pop(k, None)
                else:
                    row[k] = nv
        # update_minus_pi_and_obj (326-333)
        for k, pv in prow.items():
            self.minus_pi[k] -= rc * pv
        self.minus_obj -= rc * self.b[r]
        leaving = self.basis_indices[r]
        self.basis_indices[r] = j
        return leaving

    # carry/mod.rs:616-625
    def current_bfs(self):
        t = [(self.basis_indices[i], v) for i, v in enumerate(self.b) if v != 0]
        t.sort(key=lambda x: x[0])
        return t


class Tableau:
    """tableau/mod.rs:24-38.  `nr_art` > 0 <=> Partially kind (phase 1)."""

    def __init__(self, lp, carry, basis_columns, column_to_row):
        self.lp = lp
        self.carry = carry
        self.basis = basis_columns
        self.column_to_row = column_to_row  # artificial k -> row (phase 1 only)
        self.phase1 = column_to_row is not None
        self.nr_art = len(column_to_row) if self.phase1 else 0
        self.log = []

    def nr_columns(self):
        return self.nr_art + self.lp.n if self.phase1 else self.lp.n

    # partially.rs:72-80 / non_artificial.rs:58-62
    def original_column(self, j):
        if self.phase1:
            if j < self.nr_art:
                return [(self.column_to_row[j], Fr(1))]
            return self.lp.cols[j - self.nr_art]
        return self.lp.cols[j]

    # partially.rs:52-60 / non_artificial.rs:42-46
    def initial_cost(self, j):
        if self.phase1:
            return Fr(1) if j < self.nr_art else Fr(0)
        return self.lp.cost[j]

    # tableau/mod.rs:102-108
    def relative_cost(self, j):
        return self.carry.cost_difference(self.original_column(j)) + self.initial_cost(j)

    def generate_column(self, j):
        return self.carry.generate_column(self.original_column(j))

    # tableau/mod.rs:221-247
    def select_primal_pivot_row(self, col):
        best = None
        for (row, x) in col:
            if x > 0:
                ratio = self.carry.b[row] / x
                leaving = self.carry.basis_indices[row]
                if best is not None:
                    if ratio == best[1] and leaving < best[2]:
                        best = (row, best[1], leaving)
                    elif ratio < best[1]:
                        best = (row, ratio, leaving)
                else:
                    best = (row, ratio, leaving)
        return None if best is None else best[0]

    # tableau/mod.rs:47-60
    def bring_into_basis(self, j, r, col, rc, tag):
        leaving = self.carry.change_basis(r, j, col, rc)
        self.basis.remove(leaving)
        self.basis.add(j)
        self.log.append((tag, j, r, rc, leaving))


# pivot_rule.rs:38-127
class PivotRule:
    def __init__(self, kind):
        self.kind = kind
        self.last = None

    def _find(self, t, rng):
        for j in rng:
            if j in t.basis:
                continue
            rc = t.relative_cost(j)
            if rc < 0:
                return (j, rc)
        return None

    def select(self, t):
        n = t.nr_columns()
        if self.kind == FIRST_PROFITABLE:
            return self._find(t, range(n))
        if self.kind == FIRST_PROFITABLE_WITH_MEMORY:
            if self.last is None:
                pot = self._find(t, range(n))
            else:
                pot = self._find(t, range(self.last, n)) or self._find(t, range(self.last))
            self.last = None if pot is None else pot[0]
            return pot
        best = None  # SteepestDescent: strict <, ties -> lowest j
        for j in range(n):
            if j in t.basis:
                continue
            rc = t.relative_cost(j)
            if rc < 0 and (best is None or rc < best[1]):
                best = (j, rc)
        return best


# partially.rs:125-205 + carry/mod.rs:381-426
def partially_artificial_tableau(lp):
    m = lp.m
    real = lp.pivots
    real_rows = {r for r, _ in real}
    artificial = [i for i in range(m) if i not in real_rows]
    nr_art = len(artificial)
    real_map = dict(real)
    art_idx = {row: k for k, row in enumerate(artificial)}
    basis_indices = [art_idx[i] if i in art_idx else nr_art + real_map[i] for i in range(m)]
    minus_obj = -sum((lp.rhs[i] for i in artificial), Fr(0))
    minus_pi = [Fr(-1) if i in art_idx else Fr(0) for i in range(m)]
    rows = [{i: Fr(1)} for i in range(m)]
    carry = Carry(minus_obj, minus_pi, list(lp.rhs), basis_indices, rows)
    return Tableau(lp, carry, set(basis_indices), artificial)


# phase_one.rs:223-260
def remove_artificial_basis_variables(t):
    arts = sorted(c for c in t.basis if c < t.nr_art)
    rows_to_remove = []
    for a in arts:
        pivot_row = t.column_to_row[a]
        found = None
        for j in range(t.nr_art, t.nr_columns()):
            if j in t.basis:
                continue
            rc = t.relative_cost(j)
            if rc != 0:
                continue
            if t.carry.generate_element(pivot_row, t.original_column(j)) != 0:
                found = (j, rc)
                break
        if found is not None:
            col = t.generate_column(found[0])
            t.bring_into_basis(found[0], pivot_row, col, found[1], 'z')
        else:
            # the reference pushes `artificial` (phase_one.rs:251); equal to pivot_row whenever the reference survives (else it removes the wrong row and underflows at carry/mod.rs:665)
            rows_to_remove.append(pivot_row)
    return rows_to_remove


# phase_one.rs:125-170
def phase_one(t, rule_kind=FIRST_PROFITABLE_WITH_MEMORY):
    rule = PivotRule(rule_kind)
    while True:
        sel = rule.select(t)
        if sel is None:
            break
        j, rc = sel
        col = t.generate_column(j)
        r = t.select_primal_pivot_row(col)
        if r is None:
            raise RuntimeError("Artificial cost can not be unbounded.")
        t.bring_into_basis(j, r, col, rc, '1')
    if -t.carry.minus_obj != 0:
        return ('infeasible', None)
    rows = []
    if any(c < t.nr_art for c in t.basis):
        rows = remove_artificial_basis_variables(t)
    return ('feasible', rows)


# carry/mod.rs:214-271
def recompute_cost_row(carry, lp):
    m = carry.m()
    pi = [Fr(0)] * m
    for i in range(m):
        c = lp.cost[carry.basis_indices[i]]
        for k, v in carry.rows[i].items():
            pi[k] += v * c
    carry.minus_pi = [-x for x in pi]
    carry.minus_obj = -sum((carry.b[i] * lp.cost[carry.basis_indices[i]] for i in range(m)), Fr(0))


def remove_rows_lp(lp, rows):
    """filter/generic_wrapper.rs:224-300: provider with `rows` (sorted) removed."""
    rs = set(rows)
    newidx = {}
    k = 0
    for i in range(lp.m):
        if i not in rs:
            newidx[i] = k
            k += 1
    cols = [[(newidx[i], v) for (i, v) in c if i not in rs] for c in lp.cols]
    rhs = [v for i, v in enumerate(lp.rhs) if i not in rs]
    return LP(k, cols, lp.cost, rhs, [])


# non_artificial.rs:151-220 + carry/mod.rs:484-510, 650-689
def to_phase_two(t1, rows_to_remove):
    c = t1.carry
    nr_art = t1.nr_art
    lp = t1.lp
    basis = set(t1.basis)
    if rows_to_remove:
        for row in rows_to_remove:
            basis.discard(c.basis_indices[row])
        rs = set(rows_to_remove)
        keep = [i for i in range(c.m()) if i not in rs]
        newidx = {old: new for new, old in enumerate(keep)}
        c.basis_indices = [c.basis_indices[i] for i in keep]
        c.rows = [{newidx[k]: v for k, v in c.rows[i].items() if k not in rs} for i in keep]
        c.b = [c.b[i] for i in keep]
        lp = remove_rows_lp(lp, rows_to_remove)
    c.basis_indices = [x - nr_art for x in c.basis_indices]
    recompute_cost_row(c, lp)
    t2 = Tableau(lp, c, {x - nr_art for x in basis}, None)
    t2.log = t1.log
    return t2


# phase_two.rs:22-51
def phase_two(t, rule_kind=STEEPEST_DESCENT):
    rule = PivotRule(rule_kind)
    while True:
        sel = rule.select(t)
        if sel is None:
            return 'optimal'
        j, rc = sel
        col = t.generate_column(j)
        r = t.select_primal_pivot_row(col)
        if r is None:
            return 'unbounded'
        t.bring_into_basis(j, r, col, rc, '2')


# two_phase/mod.rs:30-75
def solve_relaxation(lp, rule1=FIRST_PROFITABLE_WITH_MEMORY, rule2=STEEPEST_DESCENT):
    t1 = partially_artificial_tableau(lp)
    status, rows = phase_one(t1, rule1)
    if status == 'infeasible':
        return {'status': 'infeasible', 'log': t1.log}
    t2 = to_phase_two(t1, rows)
    st = phase_two(t2, rule2)
    return {'status': st, 'log': t2.log, 'objective': -t2.carry.minus_obj,
            'bfs': t2.carry.current_bfs(), 'tableau': t2, 'rows_removed': rows}


def matrix_data(columns, b, ranges, n_eq, n_range, n_le, n_ge, costs, upper_bounds):
    """matrix_data.rs:37-52, 149-196, 308-371, 432-452: materialise the virtual
    columns/rows of `MatrixData` into an `LP`.

    columns: structural columns [(row, Fr)] over the n_eq+n_range+n_le+n_ge
    constraint rows (already grouped E, R, L, G); upper_bounds: Fr or None."""
    n_con = n_eq + n_range + n_le + n_ge
    nn = len(columns)
    bound_vars = [j for j in range(nn) if upper_bounds[j] is not None]
    n_bnd = len(bound_vars)
    var_to_bound = {j: k for k, j in enumerate(bound_vars)}
    row_start = [0, n_eq, n_eq + n_range, n_eq + n_range + n_le, n_con, n_con + n_bnd]
    col_start = [0, nn, nn + n_range, nn + n_range + n_le, nn + n_range + n_le + n_ge,
                 nn + n_range + n_le + n_ge + n_bnd]
    m = n_con + n_bnd + n_range
    cols = []
    for j in range(nn):
        c = [(i, Fr(v)) for (i, v) in columns[j]]
        if j in var_to_bound:
            c.append((n_con + var_to_bound[j], Fr(1)))
        cols.append(c)
    for k in range(n_range):
        cols.append([(row_start[1] + k, Fr(1)), (row_start[5] + k, Fr(1))])
    for k in range(n_le):
        cols.append([(row_start[2] + k, Fr(1))])
    for k in range(n_ge):
        cols.append([(row_start[3] + k, Fr(-1))])
    for k in range(n_bnd):
        cols.append([(row_start[4] + k, Fr(1))])
    for k in range(n_range):
        cols.append([(row_start[5] + k, Fr(1))])
    cost = [Fr(c) for c in costs] + [Fr(0)] * (len(cols) - nn)
    rhs = [Fr(x) for x in b] + [Fr(upper_bounds[j]) for j in bound_vars] + [Fr(x) for x in ranges]
    pivots = ([(row_start[2] + k, col_start[2] + k) for k in range(n_le)]
              + [(row_start[4] + k, col_start[4] + k) for k in range(n_bnd)]
              + [(row_start[5] + k, col_start[5] + k) for k in range(n_range)])
    lp = LP(m, cols, cost, rhs, pivots)
    lp.n_normal = nn
    return lp


def dense_rows_to_columns(rows, ncols):
    cols = [[] for _ in range(ncols)]
    for i, r in enumerate(rows):
        for j, v in enumerate(r):
            if v != 0:
                cols[j].append((i, Fr(v)))
    return cols
