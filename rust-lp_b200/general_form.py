"""`MPS -> GeneralForm -> MatrixData`, host side, exact rationals.

Restates /root/reference/src/io/mps/convert.rs:28-72, 140-260, 324-480 (bounds with GLPK-style
defaults, ranges table) and data/linear_program/general_form/mod.rs `standardize` (:307-314) =
transform_variables (:479-520: split free variables :522-556, flip upper-only variables, shift lower
bounds to 0) -> make_b_non_negative (:558-590) -> make_minimization_problem (:592-600), then
`derive_matrix_data` (:259-288) = reorder_constraints_by_type (E, Range, L, G; stable, :633-706) +
`MatrixData::new` (matrix_provider/matrix_data.rs:149-196; virtual slack columns and bound rows laid
out as in :37-52, initial basis :432-452).

The presolve rules (general_form/presolve/) are restated in presolve.py and applied by `standardize(presolve=True)`
exactly where the reference applies them (mod.rs:307-314: presolve, then transform_variables, make_b_non_negative,
make_minimization_problem).  `presolve=False` keeps the un-presolved standard form that the round-1 fixtures
(tests/golden/mps/, mps_prefix/) were generated with: same optimum, different rows / columns / pivot sequence.
"""
from fractions import Fraction as Fr

from . import presolve as _presolve
from .synth import ProviderArrays


class Infeasible(Exception):
    pass


class Unbounded(Exception):
    """`Err(LinearProgramType::Unbounded)`: a column that appears in no constraint can improve the objective for ever."""


class SolvedByPresolve(Exception):
    """`Err(LinearProgramType::FiniteOptimum(solution))` of `GeneralForm::presolve` (mod.rs:334-336): every variable was
    eliminated with a value.  .objective / .solution are those of the ORIGINAL problem."""

    def __init__(self, objective, solution):
        super().__init__("linear program solved by presolve")
        self.objective, self.solution = objective, solution


class GeneralForm:
    def __init__(self, mps):
        self.maximize = mps.maximize
        self.n_orig = len(mps.columns)
        self.names = [c[0] for c in mps.columns]
        self.cols = [list(c[1]) for c in mps.columns]
        self.cost = [mps.cost.get(j, Fr(0)) for j in range(self.n_orig)]
        self.orig_cost = list(self.cost)
        self.all_cost = list(self.cost)
        self.objective_constant = Fr(mps.objective_constant)
        self.fixed_cost = Fr(0)              # accumulated by the lower-bound shifts, minimisation form aside
        self.lower, self.upper = self._bounds(mps)
        self.types, self.b = self._constraints(mps)
        # bookkeeping for solution reconstruction: x_orig = sign * (x_active - x_neg_part) - shift ... see below
        self.shift = [Fr(0)] * self.n_orig
        self.flipped = [False] * self.n_orig
        self.neg_part = {}           # original free variable -> index of its negative duplicate
        # presolve bookkeeping (general_form/mod.rs `original_variables`, `from_active_to_original`)
        self.all_names = list(self.names)
        self.active_to_orig = list(range(self.n_orig))
        self.removed = {}            # original index -> ("S", value) | ("F", constant, [(original index, coefficient)])

    # convert.rs:140-260
    @staticmethod
    def _bounds(mps):
        n = len(mps.columns)
        lower, upper = [None] * n, [None] * n
        needs_default, free = [True] * n, [False] * n

        def repl(arr, j, v, greater):
            if arr[j] is None or (v > arr[j] if greater else v < arr[j]):
                arr[j] = v

        for j, kind, v in mps.bounds:
            if kind in ("LO", "LI"):
                repl(lower, j, v, True); needs_default[j] = False
            elif kind in ("UP", "UI"):
                repl(upper, j, v, False)                 # implied zero lower bound unless another is given
            elif kind == "FX":
                repl(lower, j, v, True); repl(upper, j, v, False); needs_default[j] = False
            elif kind == "FR":
                if lower[j] is not None or upper[j] is not None:
                    raise ValueError("Variable can't be bounded and free")
                free[j] = True; needs_default[j] = False
            elif kind == "MI":
                repl(upper, j, Fr(0), False); needs_default[j] = False
            elif kind == "PL":
                repl(lower, j, Fr(0), True); needs_default[j] = False
            elif kind == "BV":
                repl(lower, j, Fr(0), True); repl(upper, j, Fr(1), False); needs_default[j] = False
        for j in range(n):
            if free[j] and (lower[j] is not None or upper[j] is not None):
                raise ValueError("A variable is both free and bounded.")
            if needs_default[j]:
                lower[j] = Fr(0)
        return lower, upper

    # convert.rs:324-480
    @staticmethod
    def _constraints(mps):
        m = len(mps.rows)
        rng = dict(mps.ranges)
        types = []
        for i, (_, kind) in enumerate(mps.rows):
            if i in rng:
                types.append("E" if rng[i] == 0 else ("R", abs(rng[i])))
            else:
                types.append(kind)
        b = [None] * m
        for i, v in mps.rhs:
            kind = mps.rows[i][1]
            if b[i] is None:
                if isinstance(types[i], tuple):
                    r = rng[i]
                    # stored b is the UPPER end of the range; the range length is |r| (mps/mod.rs:236-254)
                    if kind == "G" or (kind == "E" and r >= 0):
                        b[i] = v + abs(r)
                    else:
                        b[i] = v
                else:
                    b[i] = v
            else:
                if kind == "E" and v != b[i]:
                    raise Infeasible(f"a constraint can't equal both {b[i]} and {v}")
                if kind == "G" and v > b[i]:
                    b[i] = v
                if kind == "L" and v < b[i]:
                    b[i] = v
        for i in range(m):
            if b[i] is None:
                # no rhs entry: default 0; a ranged row then spans from its default accordingly
                if isinstance(types[i], tuple):
                    kind, r = mps.rows[i][1], rng[i]
                    b[i] = abs(r) if (kind == "G" or (kind == "E" and r >= 0)) else Fr(0)
                else:
                    b[i] = Fr(0)
        return types, b

    # ---- general_form/mod.rs:479-600 -----------------------------------------------------------
    def presolve(self):
        """`GeneralForm::presolve` (mod.rs:316-366): compute the changes (presolve.py), apply them to what remains
        (:405-435), drop the removed rows and columns (:437-470), resolve removed variables whose dependencies are all
        known (:728-790).  Raises Infeasible / Unbounded / SolvedByPresolve where the reference returns Err."""
        try:
            ch = _presolve.compute_presolve_changes(self)
        except _presolve.PresolveInfeasible:
            raise Infeasible("presolve: the problem is infeasible")
        except _presolve.PresolveUnbounded:
            raise Unbounded("presolve: the problem is unbounded")
        for i, v in ch["b"].items():
            self.b[i] = v
        for i, t in ch["constraints"].items():
            self.types[i] = t
        self.fixed_cost += ch["fixed_cost"]
        for (j, direction), v in ch["bounds"].items():
            if direction == _presolve.LOWER:
                self.lower[j] = v
            else:
                self.upper[j] = v
        for j, sol in ch["removed_variables"]:
            self.removed[self.active_to_orig[j]] = sol
        gone_v = {j for j, _ in ch["removed_variables"]}
        gone_c = set(ch["constraints_marked_removed"])
        new_row, k = {}, 0
        for i in range(len(self.b)):
            if i not in gone_c:
                new_row[i] = k
                k += 1
        keep = [j for j in range(len(self.cols)) if j not in gone_v]
        self.cols = [[(new_row[i], v) for (i, v) in self.cols[j] if i in new_row] for j in keep]
        for name in ("cost", "orig_cost", "lower", "upper", "names", "active_to_orig", "shift", "flipped"):
            setattr(self, name, [getattr(self, name)[j] for j in keep])
        self.n_orig = len(keep)
        self.types = [t for i, t in enumerate(self.types) if i not in gone_c]
        self.b = [v for i, v in enumerate(self.b) if i not in gone_c]
        # compute_solution_where_possible (:728-790): a function of others whose others are all solved is solved
        progress = True
        while progress:
            progress = False
            for j, sol in list(self.removed.items()):
                if sol[0] == "F" and all(self.removed.get(k, ("?",))[0] == "S" for k, _ in sol[2]):
                    self.removed[j] = ("S", sol[1] - sum(c * self.removed[k][1] for k, c in sol[2]))
                    progress = True
        if not keep and all(s[0] == "S" for s in self.removed.values()):
            values = [self.removed[j][1] for j in range(len(self.all_names))]
            objective = sum(c * v for c, v in zip(self.all_cost, values)) + self.objective_constant
            raise SolvedByPresolve(objective, dict(zip(self.all_names, values)))
        return self

    def standardize(self, presolve=True):
        if presolve:
            self.presolve()
        m = len(self.b)
        # split free variables (:522-556)
        n0 = len(self.cols)
        for j in range(n0):
            if self.lower[j] is None and self.upper[j] is None:
                self.neg_part[j] = len(self.cols)
                self.cols.append([(i, -v) for (i, v) in self.cols[j]])
                self.cost.append(-self.cost[j])
                self.lower.append(Fr(0)); self.upper.append(None)
                self.shift.append(Fr(0)); self.flipped.append(False)
                self.lower[j] = Fr(0)
        for j in range(len(self.cols)):
            if self.lower[j] is None and self.upper[j] is not None:       # flip (:490-502)
                self.flipped[j] = not self.flipped[j]
                self.shift[j] = -self.shift[j]
                self.cost[j] = -self.cost[j]
                self.lower[j] = -self.upper[j]; self.upper[j] = None
                self.cols[j] = [(i, -v) for (i, v) in self.cols[j]]
            lo = self.lower[j]
            if lo is not None:                                               # shift (:505-517)
                self.shift[j] -= lo
                if self.upper[j] is not None:
                    self.upper[j] -= lo
                    if self.upper[j] < 0:
                        raise Infeasible(f"variable {j}: upper bound below lower bound")
                self.fixed_cost += lo * self.cost[j]
                for i, v in self.cols[j]:
                    self.b[i] -= v * lo
                self.lower[j] = Fr(0)
        # make b non-negative (:558-590)
        neg = {i for i in range(m) if self.b[i] < 0}
        if neg:
            self.cols = [[(i, -v if i in neg else v) for (i, v) in c] for c in self.cols]
            for i in neg:
                t = self.types[i]
                if t == "L":
                    self.types[i] = "G"; self.b[i] = -self.b[i]
                elif t == "G":
                    self.types[i] = "L"; self.b[i] = -self.b[i]
                elif t == "E":
                    self.b[i] = -self.b[i]
                else:
                    self.b[i] = t[1] - self.b[i]
        # minimisation (:592-600)
        self.min_cost = [-c for c in self.cost] if self.maximize else list(self.cost)
        return self

    # ---- general_form/mod.rs:259-288, :633-706 ; matrix_data.rs:37-52, 149-196, 432-452 ----------
    def derive_matrix_data(self):
        m = len(self.b)
        order = ([i for i in range(m) if self.types[i] == "E"] + [i for i in range(m) if isinstance(self.types[i], tuple)]
                 + [i for i in range(m) if self.types[i] == "L"] + [i for i in range(m) if self.types[i] == "G"])
        n_eq = sum(1 for t in self.types if t == "E")
        n_rg = sum(1 for t in self.types if isinstance(t, tuple))
        n_le = sum(1 for t in self.types if t == "L")
        n_ge = sum(1 for t in self.types if t == "G")
        newpos = {old: new for new, old in enumerate(order)}
        ranges = [self.types[i][1] for i in order if isinstance(self.types[i], tuple)]
        b = [self.b[i] for i in order]
        nn = len(self.cols)
        bound_vars = [j for j in range(nn) if self.upper[j] is not None]
        n_bnd = len(bound_vars)
        vb = {j: k for k, j in enumerate(bound_vars)}
        n_con = m
        rs = [0, n_eq, n_eq + n_rg, n_eq + n_rg + n_le, n_con, n_con + n_bnd]
        cs = [0, nn, nn + n_rg, nn + n_rg + n_le, nn + n_rg + n_le + n_ge, nn + n_rg + n_le + n_ge + n_bnd]
        cols = []
        for j in range(nn):
            c = sorted((newpos[i], v) for (i, v) in self.cols[j] if v != 0)
            if j in vb:
                c.append((n_con + vb[j], Fr(1)))
            cols.append(c)
        cols += [[(rs[1] + k, Fr(1)), (rs[5] + k, Fr(1))] for k in range(n_rg)]
        cols += [[(rs[2] + k, Fr(1))] for k in range(n_le)]
        cols += [[(rs[3] + k, Fr(-1))] for k in range(n_ge)]
        cols += [[(rs[4] + k, Fr(1))] for k in range(n_bnd)]
        cols += [[(rs[5] + k, Fr(1))] for k in range(n_rg)]
        cost = list(self.min_cost) + [Fr(0)] * (len(cols) - nn)
        rhs = b + [self.upper[j] for j in bound_vars] + ranges
        piv = ([(rs[2] + k, cs[2] + k) for k in range(n_le)] + [(rs[4] + k, cs[4] + k) for k in range(n_bnd)]
               + [(rs[5] + k, cs[5] + k) for k in range(n_rg)])
        return provider_arrays(len(rhs), cols, cost, rhs, piv, nn)

    # ---- general_form/mod.rs:817-941 -------------------------------------------------------------
    def recon_state(self):
        """What `reconstruct` needs, as plain JSON-able data (fractions as "n/d" strings)."""
        return {
            "names": list(self.names),
            "cost": [str(c) for c in self.orig_cost],
            "shift": [str(c) for c in self.shift[:self.n_orig]],
            "flipped": [bool(f) for f in self.flipped[:self.n_orig]],
            "neg_part": {str(j): k for j, k in self.neg_part.items()},
            "n_normal": len(self.cols),
            "objective_constant": str(self.objective_constant),
            "maximize": bool(self.maximize),
            # presolve (general_form/mod.rs original_variables): what was eliminated and how to get it back
            "all_names": list(self.all_names),
            "all_cost": [str(c) for c in self.all_cost],
            "active_to_orig": list(self.active_to_orig),
            "removed": {str(j): ([s[0], str(s[1])] if s[0] == "S" else [s[0], str(s[1]), [[k, str(c)] for k, c in s[2]]])
                        for j, s in self.removed.items()},
        }

    def reconstruct(self, bfs):
        return reconstruct(self.recon_state(), bfs)


def reconstruct(state, bfs):
    """(column, value) pairs of the provider solution -> (objective, {name: value}) of the ORIGINAL
    problem: un-split, un-flip, un-shift (general_form/mod.rs:817-941 without the presolve stack).

    The objective is the ORIGINAL objective function (sense as written in the file) evaluated at the
    solution.  The reference reports the objective of its internal minimisation form instead
    (mod.rs:840-870 sums the standardised costs), so for an `OBJSENSE MAX` file its sign differs;
    `minimisation_objective` gives that convention."""
    n_orig = len(state["names"])
    x = {j: v for j, v in bfs if j < state["n_normal"]}
    sol = []
    for j in range(n_orig):
        y = x.get(j, Fr(0))                       # active value, >= 0
        shift = Fr(state["shift"][j])
        # forward: x' = -x if flipped; y = x' + shift   =>   x = +-(y - shift)
        val = (y - shift) if not state["flipped"][j] else -(y - shift)
        k = state["neg_part"].get(str(j))
        if k is not None:                         # free variable: x = y_plus - y_minus
            val -= x.get(k, Fr(0))
        sol.append(val)
    if not state.get("removed"):
        obj = sum(Fr(c) * v for c, v in zip(state["cost"], sol)) + Fr(state["objective_constant"])
        return obj, dict(zip(state["names"], sol))
    # variables eliminated by presolve (compute_solution_value_with_bfs, mod.rs:884-925): fixed values, and functions
    # `constant - sum coefficient * other` of other ORIGINAL variables (possibly eliminated themselves)
    values = {o: v for o, v in zip(state["active_to_orig"], sol)}
    removed = state["removed"]

    def value_of(j):
        if j not in values:
            r = removed[str(j)]
            values[j] = Fr(r[1]) if r[0] == "S" else Fr(r[1]) - sum(Fr(c) * value_of(k) for k, c in r[2])
        return values[j]
    full = [value_of(j) for j in range(len(state["all_names"]))]
    obj = sum(Fr(c) * v for c, v in zip(state["all_cost"], full)) + Fr(state["objective_constant"])
    return obj, dict(zip(state["all_names"], full))


def minimisation_objective(state, objective):
    return -objective if state["maximize"] else objective


def provider_arrays(m, cols, cost, rhs, piv, n_normal):
    """Fraction columns -> the int64 arrays of relp_load_matrix (inputs are Rational64 in the
    reference too: io/mod.rs:66-68)."""
    col_ptr, rows, vn, vd = [0], [], [], []
    for c in cols:
        for i, v in c:
            rows.append(i); vn.append(v.numerator); vd.append(v.denominator)
        col_ptr.append(len(rows))
    lim = 2 ** 63 - 1
    for arr in (vn, vd):
        if any(abs(x) > lim for x in arr):
            raise OverflowError("matrix value does not fit Rational64")
    # the right-hand side alone may leave int64: presolve substitutes solved variables into it (BORE3D: 285 bits); the
    # reference's Rhs is a RationalBig.  Such rows travel separately (ProviderArrays.rhs_exact -> relp_set_rhs).
    big = {i: x for i, x in enumerate(rhs) if abs(x.numerator) > lim or x.denominator > lim}
    pa = ProviderArrays(m, len(cols), col_ptr, rows, vn, vd, [c.numerator for c in cost], [c.denominator for c in cost],
                        [0 if i in big else x.numerator for i, x in enumerate(rhs)], [1 if i in big else x.denominator for i, x in enumerate(rhs)],
                        [p[0] for p in piv], [p[1] for p in piv], n_normal)
    pa.rhs_exact = big
    return pa


def from_mps(mps, presolve=False):
    """MPS -> (GeneralForm standardised, ProviderArrays).  presolve=True is the reference's pipeline (mod.rs:259-314);
    the default keeps the un-presolved form of the committed round-1 fixtures."""
    gf = GeneralForm(mps)
    gf.standardize(presolve=presolve)
    return gf, gf.derive_matrix_data()
