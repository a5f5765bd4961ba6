"""Pins both CPU restatements (oracle/simplex.cpp via ctypes, oracle/pyref.py) against the
reference's own exact fixtures.  CPU only."""
from fractions import Fraction as Fr

import pytest

from helpers import pa_from_lp, problem_1_lp, problem_2_lp, feasible_lp, log_tuples
from oracle import pyref, coracle


def _carry_py(t):
    c = t.carry
    return c.minus_obj, c.minus_pi, c.b, c.basis_indices, [dict(r) for r in c.rows]


def _dense(rows, m):
    return [[r.get(k, Fr(0)) for k in range(m)] for r in rows]


# ---- src/tests/problem_2.rs:117-174 --------------------------------------------------------
P2_ART = (Fr(-8), [Fr(-1)] * 3, [Fr(1), Fr(3), Fr(4)], [0, 1, 2],
          [[1, 0, 0], [0, 1, 0], [0, 0, 1]])
P2_BFS = (Fr(-9, 2), [Fr(5, 2), Fr(-1), Fr(-1)], [Fr(1, 2), Fr(5, 2), Fr(3, 2)], [1, 3, 4],
          [[Fr(1, 2), 0, 0], [Fr(-1, 2), 1, 0], [Fr(-5, 2), 0, 1]])
# ---- src/tests/problem_1.rs:376-431 --------------------------------------------------------
P1_ART = (Fr(-16), [Fr(-1), Fr(-1), Fr(0), Fr(0)], [Fr(6), Fr(10), Fr(4), Fr(2)], [0, 1, 6, 7],
          [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]])
P1_BFS = (Fr(-58), [Fr(4), Fr(-13), Fr(12), Fr(0)], [Fr(6), Fr(0), Fr(4), Fr(2)], [2, 1, 0, 5],
          [[0, 1, -1, 0], [-1, 1, -1, 0], [0, 0, 1, 0], [1, -1, 1, 1]])


def _check(state, expected):
    mo, mpi, b, basis, rows = state
    emo, empi, eb, ebasis, erows = expected
    assert mo == emo and list(mpi) == empi and list(b) == eb and list(basis) == ebasis
    assert _dense(rows, len(eb)) == [[Fr(x) for x in r] for r in erows]


@pytest.mark.parametrize("make_lp,art,bfs,opt,obj", [
    (problem_2_lp, P2_ART, P2_BFS, [(1, Fr(1, 2)), (3, Fr(5, 2)), (4, Fr(3, 2))], Fr(9, 2)),
    (problem_1_lp, P1_ART, P1_BFS, [(0, Fr(4)), (2, Fr(6)), (5, Fr(2))], Fr(58)),
])
def test_pyref_pipeline(make_lp, art, bfs, opt, obj):
    """conversion_pipeline tests: artificial tableau -> phase 1 (FirstProfitable) -> hand-over
    -> phase 2 (FirstProfitable) -> optimum."""
    t = pyref.partially_artificial_tableau(make_lp())
    _check(_carry_py(t), art)
    status, rows = pyref.phase_one(t, pyref.FIRST_PROFITABLE)
    assert status == "feasible" and rows == []
    t2 = pyref.to_phase_two(t, rows)
    _check(_carry_py(t2), bfs)
    assert pyref.phase_two(t2, pyref.FIRST_PROFITABLE) == "optimal"
    assert t2.carry.current_bfs() == opt
    assert -t2.carry.minus_obj == obj


@pytest.mark.parametrize("make_lp,art,bfs,opt,obj", [
    (problem_2_lp, P2_ART, P2_BFS, [(1, Fr(1, 2)), (3, Fr(5, 2)), (4, Fr(3, 2))], Fr(9, 2)),
    (problem_1_lp, P1_ART, P1_BFS, [(0, Fr(4)), (2, Fr(6)), (5, Fr(2))], Fr(58)),
])
def test_coracle_pipeline(make_lp, art, bfs, opt, obj):
    o = coracle.COracle(pa_from_lp(make_lp()))
    o.begin(0, 0)
    _check(o.carry(), art)
    assert o.advance_to_handover() == coracle.RUNNING and o.stage == 1
    _check(o.carry(), bfs)
    assert o.advance(-1) == coracle.OPTIMAL
    assert o.current_bfs() == opt and o.objective() == obj


def test_tableau_kats():
    """tableau/mod.rs:417-481 and pivot_rule.rs:137-167."""
    lp = problem_2_lp()
    t = pyref.partially_artificial_tableau(lp)
    assert -t.carry.minus_obj == 8
    assert t.relative_cost(0) == 0
    assert t.relative_cost(t.nr_art + 0) == -10
    assert t.generate_column(t.nr_art + 0) == [(0, Fr(3)), (1, Fr(5)), (2, Fr(2))]
    assert pyref.PivotRule(pyref.FIRST_PROFITABLE).select(t)[0] == 3
    assert t.select_primal_pivot_row([(0, Fr(3)), (1, Fr(5)), (2, Fr(2))]) == 0
    assert t.select_primal_pivot_row([(0, Fr(2)), (1, Fr(1)), (2, Fr(5))]) == 0
    col = t.generate_column(3)
    r = t.select_primal_pivot_row(col)
    t.bring_into_basis(3, r, col, t.relative_cost(3), "1")
    assert 3 in t.basis and 0 not in t.basis and -t.carry.minus_obj == Fr(14, 3)
    # non-artificial tableau of tableau/mod.rs:379-404
    carry = pyref.Carry(Fr(-6), [Fr(1), Fr(-1), Fr(-1)], [Fr(1), Fr(2), Fr(3)], [2, 3, 4],
                        [{0: Fr(1)}, {0: Fr(-1), 1: Fr(1)}, {0: Fr(-1), 2: Fr(1)}])
    t2 = pyref.Tableau(lp, carry, {2, 3, 4}, None)
    assert [t2.relative_cost(j) for j in range(3)] == [-3, -3, 0]
    assert t2.generate_column(0) == [(0, Fr(3)), (1, Fr(2)), (2, Fr(-1))]
    assert t2.select_primal_pivot_row([(0, Fr(3)), (1, Fr(2)), (2, Fr(-1))]) == 0
    assert t2.select_primal_pivot_row([(0, Fr(2)), (1, Fr(-1)), (2, Fr(3))]) == 0
    col = t2.generate_column(1)
    t2.bring_into_basis(1, t2.select_primal_pivot_row(col), col, t2.relative_cost(1), "2")
    assert -t2.carry.minus_obj == Fr(9, 2)


@pytest.mark.parametrize("seed", range(60))
def test_two_restatements_agree(seed):
    """Independent C++ and Python restatements produce the same pivot log, status, objective
    and final Carry -- including rank-deficient (rows removed), unbounded and fractional cases."""
    lp = feasible_lp(seed, 3, 1, 3, 2, 10, frac=(seed % 2 == 0), dup_rows=(1 if seed % 5 == 0 else 0))
    o = coracle.COracle(pa_from_lp(lp))
    st = o.solve(1, 2)
    res = pyref.solve_relaxation(lp, 1, 2)
    assert {"optimal": 1, "unbounded": 2, "infeasible": 3}[res["status"]] == st
    assert [(t, j, r, l) for (t, j, r, rc, l) in res["log"]] == log_tuples(o.log())
    for k, e in enumerate(res["log"]):
        assert o.log_rc(k) == e[3]
    t2 = res["tableau"]
    mo, mpi, b, basis, rows = o.carry()
    assert (mo, mpi, b, basis) == (t2.carry.minus_obj, t2.carry.minus_pi, t2.carry.b, t2.carry.basis_indices)
    assert rows == [dict(r) for r in t2.carry.rows]
