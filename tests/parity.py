"""Parity checks shared by the GPU suite (tests/test_gpu_parity.py, the real librelp_b200.so on a
B200) and the CPU-only logic suite (tests/test_emul_logic.py, the same sources compiled against
tests/native/cuda_emul.h).  Every check drives the C ABI and compares with the oracle."""
from fractions import Fraction as Fr

import numpy as np

from helpers import pa_from_lp, log_tuples
from oracle import coracle, pyref
from rust_lp_b200 import capi

STATUS = {coracle.OPTIMAL: capi.OPTIMAL, coracle.UNBOUNDED: capi.UNBOUNDED, coracle.INFEASIBLE: capi.INFEASIBLE}


def oracle_log(o):
    return [(chr(int(t)), int(j), int(r), int(l), int(T), int(S)) for (t, j, r, l, T, S) in o.log()]


def assert_same_carry(c, o, what=""):
    got = c.carry()
    exp = o.carry()
    assert got[3] == exp[3], f"basis differs {what}"
    assert got[0] == exp[0], f"-obj differs {what}: {got[0]} vs {exp[0]}"
    assert got[1] == exp[1], f"-pi differs {what}"
    assert got[2] == exp[2], f"b differs {what}"
    assert got[4] == exp[4], f"B^-1 differs {what}"


def check_full_solve(pa, lib_path, rule1=1, rule2=2, heap_words=1 << 22, expect_escalation=None, persistent=True, options=None):
    """Whole solve through relp_solve_relaxation: identical status, pivot log (entering column,
    row, leaving column, T, S per pivot), objective, basic solution and every Carry entry."""
    o = coracle.COracle(pa)
    st_o = o.solve(rule1, rule2)
    c = capi.Carry(pa, heap_words=heap_words, lib_path=lib_path)
    c.set_option("persistent", persistent)
    for name, value in (options or {}).items():
        c.set_option(name, value)
    st = c.solve_relaxation(rule1, rule2)
    assert st == STATUS[st_o], (st, st_o)
    assert c.log() == oracle_log(o)
    if st_o != coracle.INFEASIBLE:
        assert_same_carry(c, o, "at termination")
        assert c.get_objective_function_value() == o.objective()
        assert c.current_bfs() == o.current_bfs()
    assert c.rows_removed() == [] or o.m < pa.m
    s = c.stats()
    if expect_escalation is True:
        assert s.escalated > 0 and o.max_bits() > 63
    if expect_escalation is False:
        assert s.escalated == 0
    c.close()
    return s, o


def check_resumable(pa, lib_path, chunk=3, rule1=1, rule2=2):
    """Pivot budget: solve in chunks of `chunk` pivots; the Carry equals the oracle's after each."""
    o = coracle.COracle(pa); o.begin(rule1, rule2)
    c = capi.Carry(pa, heap_words=1 << 22, lib_path=lib_path)
    st = c.solve_relaxation(rule1, rule2, max_pivots=chunk)
    st_o = o.advance(chunk)
    for _ in range(10000):
        assert (st == capi.OK) == (st_o == coracle.RUNNING)
        assert c.log() == oracle_log(o)
        if st_o == coracle.INFEASIBLE:
            break
        assert_same_carry(c, o, f"after {o.nr_pivots} pivots")
        if st != capi.OK:
            break
        st = c.solve_continue(chunk); st_o = o.advance(chunk)
    assert st == STATUS.get(st_o, capi.OK)
    c.close()


def check_stepwise(lp, lib_path, rule1=1, rule2=2):
    """Trait-level calls (price / generate_column / ratio test / change_basis), one at a time,
    against the Python restatement: reduced cost, generated column, pivot row, leaving column and
    the full Carry after every pivot.  Phase 1 only up to its end, then hand-over and phase 2."""
    pa = pa_from_lp(lp)
    t = pyref.partially_artificial_tableau(lp)
    c = capi.Carry(pa, heap_words=1 << 22, lib_path=lib_path)
    c.create_for_partially_artificial()

    def same(tab):
        got = c.carry()
        k = tab.carry
        assert got[0] == k.minus_obj and got[1] == list(k.minus_pi) and got[2] == list(k.b)
        assert got[3] == list(k.basis_indices) and got[4] == [dict(r) for r in k.rows]

    def loop(tab, rule, tag):
        pr = pyref.PivotRule(rule)
        for _ in range(10000):
            sel = pr.select(tab)
            got = c.select_primal_pivot_column(rule)
            if sel is None:
                assert got is None
                return "done"
            assert got == (sel[0], sel[1]), (got, sel)
            col = tab.generate_column(sel[0])
            assert c.generate_column(sel[0]) == col
            r = tab.select_primal_pivot_row(col)
            assert c.select_primal_pivot_row() == r
            if r is None:
                return "unbounded"
            leaving = tab.carry.basis_indices[r]
            tab.bring_into_basis(sel[0], r, col, sel[1], tag)
            assert c.change_basis(r) == leaving
            same(tab)
        raise AssertionError("no termination")

    same(t)
    assert loop(t, rule1, "1") == "done"
    if t.carry.minus_obj != 0:
        c.close(); return "infeasible"
    if any(j < t.nr_art for j in t.basis):
        c.close(); return "artificials left"        # covered by check_full_solve
    t2 = pyref.to_phase_two(t, [])
    c.from_artificial()
    same(t2)
    res = loop(t2, rule2, "2")
    c.close()
    return res


def check_fixture(make_lp, art, bfs, opt, obj, lib_path):
    """src/tests/problem_1.rs / problem_2.rs: the pinned Carry states, FirstProfitable in both phases."""
    pa = pa_from_lp(make_lp())
    c = capi.Carry(pa, heap_words=1 << 20, lib_path=lib_path)
    c.create_for_partially_artificial()
    _check_state(c.carry(), art)
    assert c.solve_to_handover(capi.FIRST_PROFITABLE) == capi.OK
    _check_state(c.carry(), bfs)
    assert c.iterate(capi.FIRST_PROFITABLE) == capi.OPTIMAL
    assert c.current_bfs() == opt and c.get_objective_function_value() == obj
    c.close()


def _check_state(state, expected):
    mo, mpi, b, basis, rows = state
    emo, empi, eb, ebasis, erows = expected
    assert mo == emo and list(mpi) == empi and list(b) == eb and list(basis) == ebasis
    m = len(eb)
    assert [[r.get(k, Fr(0)) for k in range(m)] for r in rows] == [[Fr(x) for x in r] for r in erows]


def check_from_basis(pa, lib_path, pivots_before=3, rule2=2):
    """Warm start (`IM::from_basis`): take the basis the oracle has after `pivots_before` phase-two pivots,
    rebuild the Carry from that basis alone on both sides (oracle: Gauss-Jordan on [B | I]; product:
    relp_carry_from_basis) -- every entry must equal the Carry the simplex iterations had produced --
    then solve on from there: identical pivots, objective and solution."""
    o = coracle.COracle(pa); o.begin(1, rule2)
    st = o.advance_to_handover()
    if st != coracle.RUNNING:
        return False
    o.advance(pivots_before)
    reached = o.carry()
    basis = list(reached[3])
    o2 = coracle.COracle(pa)
    assert o2.from_basis(basis, rule2) == 0
    assert o2.carry() == reached, "oracle: from_basis differs from the iterated Carry"
    c = capi.Carry(pa, heap_words=1 << 22, lib_path=lib_path)
    c.from_basis(basis)
    assert_same_carry(c, o2, "after from_basis")
    st_c = c.iterate(rule2)
    st_o = o2.advance(-1)
    assert st_c == STATUS[st_o]
    assert c.log() == oracle_log(o2)
    assert_same_carry(c, o2, "at termination after a warm start")
    assert c.current_bfs() == o2.current_bfs()
    c.close()
    return True


def check_reuse_after_row_removal(pa_deficient, pa_other, lib_path):
    """a context that removed redundant rows (smaller planes, filtered provider) must (a) solve the
    SAME loaded LP again to the same answer without a reload and (b) take a different, larger LP afterwards."""
    o = coracle.COracle(pa_deficient); st_o = o.solve(1, 2)
    c = capi.Carry(pa_deficient, heap_words=1 << 22, lib_path=lib_path)
    for _ in range(2):                       # second round: no reload in between
        assert c.solve_relaxation() == STATUS[st_o]
        assert c.rows_removed() != [] and o.m < pa_deficient.m
        assert c.log() == oracle_log(o)
        assert_same_carry(c, o, "rank-deficient LP")
    o2 = coracle.COracle(pa_other); st2 = o2.solve(1, 2)
    c.load_matrix(pa_other)
    assert c.solve_relaxation() == STATUS[st2]
    assert c.log() == oracle_log(o2)
    assert_same_carry(c, o2, "LP loaded after a rank-deficient one")
    c.close()


def check_load_validation(pa, lib_path):
    """relp_load_matrix refuses what the kernels would misread: bad denominators, non-canonical
    fractions, INT64_MIN, row indices out of range, decreasing col_ptr, negative right-hand side, NULL."""
    import copy
    c = capi.Carry(pa, heap_words=1 << 20, lib_path=lib_path)

    def refused(mutate, what):
        q = copy.deepcopy(pa)
        mutate(q)
        try:
            c.load_matrix(q)
        except capi.RelpError as e:
            assert e.status == capi.ERR_ARGUMENT and what in str(e), str(e)
            return
        raise AssertionError("accepted: " + what)

    def setv(name, idx, v):
        def f(q):
            getattr(q, name)[idx] = v
        return f
    refused(setv("val_den", 0, 0), "denominator")
    refused(setv("val_den", 1, -3), "denominator")
    refused(setv("cost_den", 0, 0), "cost")
    refused(setv("rhs_den", 0, -1), "rhs")
    refused(setv("rhs_num", 0, -1), "rhs < 0")
    refused(setv("val_num", 0, -2 ** 63), "INT64_MIN")
    refused(setv("row_idx", 0, pa.m), "row_idx")
    refused(setv("row_idx", 0, -1), "row_idx")
    refused(setv("col_ptr", 1, int(pa.col_ptr[2]) + 1), "col_ptr")

    def noncanon(q):
        q.val_num[0] *= 2; q.val_den[0] *= 2
    refused(noncanon, "lowest terms")
    c.load_matrix(pa)                        # and the context is still usable
    assert c.solve_relaxation() in (capi.OPTIMAL, capi.UNBOUNDED, capi.INFEASIBLE)
    c.close()


def check_set_costs(pa, lib_path):
    """relp_set_costs (the trait-level caller's route: tableau first, real costs at the phase hand-over): loading with
    zero costs, running phase one, handing the real costs over and then from_artificial + phase two gives the pivot
    log and answer of a load with the costs; non-canonical costs, a wrong length and a live phase-two tableau are refused."""
    import copy

    def run(loaded, late_costs):
        c = capi.Carry(loaded, heap_words=1 << 22, lib_path=lib_path)
        c.create_for_partially_artificial()
        assert c.iterate(capi.FIRST_PROFITABLE_WITH_MEMORY) == capi.OPTIMAL          # phase one
        assert c.get_objective_function_value() == 0 and all(j >= c.dims()[2] for j in c.basis_indices()), "pick a feasible full-rank LP"
        if late_costs:
            for bad_n, bad_d, what in ((pa.cost_num * 2, np.full_like(pa.cost_den, 2), "canonical"),
                                       (pa.cost_num[:-1], pa.cost_den[:-1], "differs")):
                try:
                    c.set_costs(bad_n, bad_d)
                except capi.RelpError as e:
                    assert e.status == capi.ERR_ARGUMENT and what in str(e), str(e)
                else:
                    raise AssertionError("accepted " + what)
            c.set_costs(pa.cost_num, pa.cost_den)
        c.from_artificial()
        st = c.iterate(capi.STEEPEST_DESCENT)
        out = (st, c.log(), c.get_objective_function_value(), c.current_bfs(), c.carry())
        if late_costs:
            try:
                c.set_costs(pa.cost_num, pa.cost_den)
            except capi.RelpError as e:
                assert "phase-two" in str(e)
            else:
                raise AssertionError("set_costs accepted with a phase-two tableau live")
        c.close()
        return out

    blank = copy.deepcopy(pa)
    blank.cost_num = np.zeros_like(pa.cost_num); blank.cost_den = np.ones_like(pa.cost_den)
    assert np.any(pa.cost_num != 0)
    assert run(blank, True) == run(pa, False)


def check_big_rhs(pa, lib_path):
    """relp_set_rhs: right-hand sides beyond int64/int64 (the reference's Rhs is a RationalBig; presolve produces them).
    Whole solves with every row scaled by a 130-bit factor, and with only some rows big, against the oracle; bad records
    are refused."""
    import copy
    scale = Fr(10 ** 40 + 1, 3 ** 5)
    for rows in (range(pa.m), range(0, pa.m, 2)):
        q = copy.deepcopy(pa)
        q.rhs_exact = {i: Fr(int(pa.rhs_num[i]), int(pa.rhs_den[i])) * scale for i in rows if pa.rhs_num[i] != 0}
        for i in q.rhs_exact:
            q.rhs_num[i], q.rhs_den[i] = 0, 1
        s, o = check_full_solve(q, lib_path, heap_words=1 << 22)
        assert s.escalated > 0 and o.max_bits() > 100
    c = capi.Carry(pa, heap_words=1 << 20, lib_path=lib_path)
    m = pa.m
    n = np.zeros(m, np.int64); d = np.ones(m, np.int64)

    def refused(limbs, what, n0=0, d0=0):
        n[0], d[0] = n0, d0
        lin = np.array(limbs, np.uint32)
        st = c.L.relp_set_rhs(c.h, m, n.ctypes.data_as(capi.C.c_void_p), d.ctypes.data_as(capi.C.c_void_p), lin.ctypes.data_as(capi.C.c_void_p), len(limbs))
        assert st == capi.ERR_ARGUMENT and what in c.L.relp_last_error().decode(), c.L.relp_last_error()
    refused([3 | 1 << 14 | 0x80000000, 1, 2, 3, 1], "rhs < 0")
    refused([3 | 1 << 14, 1, 2, 3, 1], "outside", n0=7)
    refused([3 | 1 << 14, 1, 2, 3], "past")
    refused([3 | 1 << 14, 1, 2, 0, 1], "leading zero")
    refused([1 | 1 << 14, 5, 1], "fits int64")
    refused([0], "denominator < 0", n0=1, d0=-2)
    c.close()


def check_cleanup_any(lib_path):
    """The artificial clean-up.  Reference rule (phase_one.rs:239-249): only columns with relative cost zero are tried, and a
    row whose artificial cannot be pivoted out that way is REMOVED -- also when the row is not redundant (here: row 0,
    -x1 - x2 = 0, fixes x1 = x2 = 0; after phase one its only nonzero elements sit in columns with positive relative cost).
    The default mirrors the reference (objective -5 of the relaxed problem); option cleanup_any = 1 pivots on any nonzero
    element of the row and reaches the true optimum 0.  Product and oracle agree under both settings."""
    from rust_lp_b200.synth import ProviderArrays
    #            x1        x2      slack
    pa = ProviderArrays(2, 3, [0, 2, 3, 4], [0, 1, 0, 1], [-1, 1, -1, 1], [1, 1, 1, 1], [-1, 0, 0], [1, 1, 1], [0, 5], [1, 1], [1], [2], 2)
    for any_cost, want in ((0, Fr(-5)), (1, Fr(0))):
        o = coracle.COracle(pa); o.set_cleanup_any(bool(any_cost))
        assert o.solve(1, 2) == coracle.OPTIMAL and o.objective() == want, (any_cost, o.objective())
        c = capi.Carry(pa, heap_words=1 << 20, lib_path=lib_path)
        c.set_option("cleanup_any", any_cost)
        assert c.solve_relaxation() == capi.OPTIMAL
        assert c.get_objective_function_value() == want and c.log() == oracle_log(o) and c.current_bfs() == o.current_bfs()
        assert len(c.rows_removed()) == (0 if any_cost else 1)
        c.close()
