"""The presolve restatement (rust-lp_b200/presolve.py) pinned on the reference's own tests:
/root/reference/src/data/linear_program/general_form/presolve/test/per_rule.rs:16-177 (single rule applications, with
the counters / queues / updates they leave behind), test/changes.rs (whole `compute_presolve_changes` runs) and
test/with_application.rs:27 (a problem presolve solves completely).  Values are copied from those files."""
from fractions import Fraction as Fr

import pytest

from rust_lp_b200 import general_form, presolve
from rust_lp_b200.presolve import LOWER, UPPER

E, L, G = "E", "L", "G"


def R(t):
    return ("R", Fr(t))


def gf(rows, ncols, types, b, variables, maximize=False, fixed_cost=0, names=None):
    """`GeneralForm::new(objective, ColumnMajor::from_test_data(rows, ncols), types, b, variables, names, fixed_cost)`;
    variables = [(cost, lower, upper)]"""
    g = object.__new__(general_form.GeneralForm)
    g.maximize = maximize
    g.cols = [[(i, Fr(r[j])) for i, r in enumerate(rows) if r[j] != 0] for j in range(ncols)]
    g.cost = [Fr(v[0]) for v in variables]
    g.orig_cost, g.all_cost = list(g.cost), list(g.cost)
    g.lower = [None if v[1] is None else Fr(v[1]) for v in variables]
    g.upper = [None if v[2] is None else Fr(v[2]) for v in variables]
    g.types, g.b = list(types), [Fr(x) for x in b]
    g.n_orig = ncols
    g.names = names or ["x"] * ncols
    g.all_names = list(g.names)
    g.active_to_orig = list(range(ncols))
    g.removed, g.neg_part = {}, {}
    g.shift, g.flipped = [Fr(0)] * ncols, [False] * ncols
    g.fixed_cost, g.objective_constant = Fr(0), Fr(fixed_cost)
    return g


def changes(**kw):
    base = {"b": {}, "constraints": {}, "fixed_cost": Fr(0), "bounds": {}, "removed_variables": [], "constraints_marked_removed": []}
    base.update(kw)
    return base


# ---- per_rule.rs ---------------------------------------------------------------------------------------
def test_presolve_fixed_variable_feasible():
    idx = presolve.Index(gf([[1], [2]], 1, [E, G], [1, 1], [(1, 1, 1)], fixed_cost=7))
    idx.presolve_fixed_variable(0)
    assert idx.c_constraint == [0, 0] and idx.c_variable == [0]
    assert idx.constraints_marked_removed == [0, 1]
    assert idx.removed_variables == [(0, ("S", Fr(1)))]
    assert idx.u_b[0] == 0 and idx.u_b[1] == -1
    assert idx.fixed_cost == 1
    assert not idx.are_queues_empty()


def test_presolve_fixed_variable_infeasible():
    idx = presolve.Index(gf([[1], [2]], 1, [E, E], [1, 1], [(1, 1, 1)], fixed_cost=7))
    with pytest.raises(presolve.PresolveInfeasible):
        idx.presolve_fixed_variable(0)


def test_presolve_simple_bound_constraint():
    g = gf([[1], [1]], 1, [E, E], [2, 2], [(1, None, None)])
    idx = presolve.Index(g)
    idx.presolve_bound_constraint(0)
    assert idx.c_constraint == [0, 1] and idx.c_variable == [1]
    assert idx.constraints_marked_removed == [0] and idx.removed_variables == []
    assert idx.u_bounds[(0, LOWER)] == 2 and idx.u_bounds[(0, UPPER)] == 2
    assert not idx.are_queues_empty()
    assert g.b == [2, 2]


def test_presolve_constraint_if_slack_with_suitable_bounds():
    def create(ctype, lower, upper):
        return gf([[2, 2, 2]], 3, [ctype], [3], [(0, lower, upper)] * 3)
    fo = ("F", Fr(3, 2), [(1, Fr(1)), (2, Fr(1))])
    idx = presolve.Index(create(E, 1, 2))
    idx.presolve_slack(0)
    assert idx.c_constraint == [2] and idx.c_variable == [0, 1, 1] and idx.constraints_marked_removed == []
    assert idx.removed_variables == [(0, fo)]
    idx = presolve.Index(create(E, 1, None))
    idx.presolve_slack(0)
    assert idx.c_constraint == [2] and idx.c_variable == [0, 1, 1] and idx.constraints_marked_removed == []
    assert idx.constraint_type(0) == L and idx.removed_variables == [(0, fo)] and idx.b(0) == 3 - 2
    idx = presolve.Index(create(E, None, 1))
    idx.presolve_slack(0)
    assert idx.c_constraint == [2] and idx.c_variable == [0, 1, 1] and idx.constraints_marked_removed == []
    assert idx.constraint_type(0) == G and idx.removed_variables == [(0, fo)] and idx.b(0) == 3 - 2
    idx = presolve.Index(create(G, 1, None))
    idx.presolve_slack(0)
    assert idx.c_constraint == [0] and idx.c_variable == [0, 0, 0] and idx.constraints_marked_removed == [0]
    assert idx.removed_variables == [(1, ("S", Fr(1))), (2, ("S", Fr(1))), (0, fo)]


# ---- changes.rs ------------------------------------------------------------------------------------------
A2 = [[1, 1], [Fr(4, 5), Fr(3, 2)], [Fr(3, 2), Fr(4, 5)]]
A3 = [[1, 1, 1], [Fr(4, 5), Fr(3, 2), 0], [Fr(3, 2), Fr(4, 5), 0]]
B3 = [1, Fr(6, 5), Fr(6, 5)]
BOX = (1, 0, Fr(3, 4))


def test_no_changes():
    assert presolve.compute_presolve_changes(gf(A2, 2, [L, L, L], B3, [BOX] * 2, maximize=True)) == changes()


def test_empty_column():
    g = gf([[1, 1, 0], [Fr(4, 5), Fr(3, 2), 0], [Fr(3, 2), Fr(4, 5), 0]], 3, [L, L, L], B3, [BOX] * 3, maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(fixed_cost=Fr(3, 4), removed_variables=[(2, ("S", Fr(3, 4)))])


def test_remove_slack():
    g = gf(A3, 3, [E, L, L], B3, [BOX, BOX, (0, 0, None)], maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(
        constraints={0: L}, removed_variables=[(2, ("F", Fr(1), [(0, Fr(1)), (1, Fr(1))]))])


def test_remove_range_slack():
    g = gf(A3, 3, [E, L, L], B3, [BOX, BOX, (0, 0, Fr(3, 4))], maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(
        constraints={0: R(Fr(3, 4))}, removed_variables=[(2, ("F", Fr(1), [(0, Fr(1)), (1, Fr(1))]))])


def test_free_variable():
    g = gf(A3, 3, [E, L, L], B3, [BOX, BOX, (0, None, None)], maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(
        removed_variables=[(2, ("F", Fr(1), [(0, Fr(1)), (1, Fr(1))]))], constraints_marked_removed=[0])


def test_range_slack_for_range_bound():
    g = gf(A3, 3, [R(Fr(1, 2)), L, L], B3, [(14, None, None), (14, None, None), (0, 0, Fr(1, 2))], maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(
        constraints={0: R(1)}, removed_variables=[(2, ("F", Fr(1), [(0, Fr(1)), (1, Fr(1))]))])


def test_range_slack_inequality_bound():
    g = gf([[1, 0, 1], [1, 1, 0], [Fr(4, 5), Fr(3, 2), 0], [Fr(3, 2), Fr(4, 5), 0]], 3, [L, L, L, L], [1, 1, Fr(6, 5), Fr(6, 5)],
           [(1, None, None), (1, None, None), (0, 0, 1)], maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(
        bounds={(0, UPPER): Fr(1)}, removed_variables=[(2, ("S", Fr(0)))], constraints_marked_removed=[0])


def test_remove_redundant_constraint():
    g = gf(A2, 2, [L, L, L], [Fr(3, 2), Fr(6, 5), Fr(6, 5)], [BOX] * 2, maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(constraints_marked_removed=[0])


def test_trigger_fixed_unfeasible():
    g = gf(A2, 2, [E, L, L], B3, [(14, Fr(1, 4), None), (14, Fr(3, 4), None)], maximize=True)
    with pytest.raises(presolve.PresolveInfeasible):
        presolve.compute_presolve_changes(g)


def test_trigger_substitution_feasible():
    g = gf(A2, 2, [E, L, L], B3, [(14, Fr(1, 2), None)] * 2, maximize=True)
    assert presolve.compute_presolve_changes(g) == changes(
        fixed_cost=Fr(14), removed_variables=[(0, ("S", Fr(1, 2))), (1, ("S", Fr(1, 2)))], constraints_marked_removed=[0, 1, 2])


def test_trigger_substitution_infeasible():
    g = gf(A2, 2, [L, L, L], B3, [(14, Fr(2, 3), None)] * 2, maximize=True)
    with pytest.raises(presolve.PresolveInfeasible):
        presolve.compute_presolve_changes(g)


# ---- with_application.rs:27 --------------------------------------------------------------------------------
def test_presolve_solves_the_whole_problem():
    rows = [[2, 0, 0, 0, 0, 0], [3, 5, 0, 0, 0, 0], [7, 11, 13, 0, 0, 0], [17, 19, 23, 0, 29, 31]]
    x2 = (Fr(103) - Fr(101, 2) * 3) / 5
    variables = [(211, None, None), (223, x2, None), (227, None, None), (-229, None, 131),
                 (233, Fr(-30736, 65 * 29), 123), (0, 5, None)]
    names = ["XONE", "XTWO", "XTHREE", "XFOUR", "XFIVE", "XSIX"]
    g = gf(rows, 6, [E, L, G, E], [101, 103, 107, 109], variables, names=names, fixed_cost=1)
    with pytest.raises(general_form.SolvedByPresolve) as e:
        g.presolve()
    expected = {"XONE": Fr(101, 2), "XTWO": x2, "XTHREE": (Fr(-3601, 5) + Fr(29 * 30736, 1885)) / 23, "XFOUR": Fr(131),
                "XFIVE": Fr(-30736, 65 * 29), "XSIX": Fr(5)}
    assert e.value.solution == expected
    assert e.value.objective == (1 + Fr(211 * 101, 2) + Fr(223 * -97, 10) + Fr(227 * -699, 65) + Fr(-229 * 131) + Fr(233 * -30736, 1885))
