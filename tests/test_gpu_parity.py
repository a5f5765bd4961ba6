"""Parity tests proper: the real librelp_b200.so on a B200, through the C ABI, against the oracle.
Bit-exact (integer/rational work): identical pivot sequence, every Carry entry, objective, solution."""
import os
from fractions import Fraction as Fr

import pytest

import parity
from helpers import pa_from_lp, problem_1_lp, problem_2_lp, feasible_lp, random_lp
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
from test_oracle_fixtures import P1_ART, P1_BFS, P2_ART, P2_BFS

pytestmark = pytest.mark.gpu
LIB = capi.LIB_PATH


def test_library_is_the_cuda_build():
    import torch
    assert torch.cuda.is_available()
    assert os.path.exists(LIB), "librelp_b200.so missing: the product has no fallback"


@pytest.mark.parametrize("make_lp,art,bfs,opt,obj", [
    (problem_2_lp, P2_ART, P2_BFS, [(1, Fr(1, 2)), (3, Fr(5, 2)), (4, Fr(3, 2))], Fr(9, 2)),
    (problem_1_lp, P1_ART, P1_BFS, [(0, Fr(4)), (2, Fr(6)), (5, Fr(2))], Fr(58)),
])
def test_reference_fixtures(make_lp, art, bfs, opt, obj):
    parity.check_fixture(make_lp, art, bfs, opt, obj, LIB)


@pytest.mark.parametrize("seed", range(30))
def test_full_solve_random(seed):
    lp = feasible_lp(seed, 3, 1, 3, 2, 10, frac=(seed % 2 == 0), dup_rows=(1 if seed % 5 == 0 else 0))
    parity.check_full_solve(pa_from_lp(lp), LIB)


@pytest.mark.parametrize("seed", range(100, 110))
def test_full_solve_random_unfiltered(seed):
    lp = random_lp(seed, 2, 1, 3, 2, 8, frac=(seed % 2 == 0))
    parity.check_full_solve(pa_from_lp(lp), LIB)


@pytest.mark.parametrize("seed", range(4))
def test_stepwise_traits(seed):
    lp = feasible_lp(seed + 50, 2, 1, 3, 1, 8, frac=(seed % 2 == 1))
    parity.check_stepwise(lp, LIB)


@pytest.mark.parametrize("seed", range(2))
def test_resumable(seed):
    lp = feasible_lp(seed + 70, 3, 0, 3, 2, 9, frac=True)
    parity.check_resumable(pa_from_lp(lp), LIB)


@pytest.mark.parametrize("seed", range(10))
def test_dense_update_paths(seed):
    """the large-problem paths of the Carry update on small problems: thresholds lowered so that every wide
    pivot with a dense generated column leaves the persistent loop for k_update (shared-memory ring fed by
    bulk copies, two rows per thread) and every other one is tiled there; same pivots, same Carry"""
    lp = feasible_lp(seed + 300, 4, 1, 5, 3, 14, frac=(seed % 2 == 0), dup_rows=(1 if seed % 4 == 0 else 0), density=0.8)
    opts = {"narrow": 0, "k1_exit": 1, "k1_min": 1}
    s, _ = parity.check_full_solve(pa_from_lp(lp), LIB, options=opts)
    assert s.wide_handoffs > 0
    parity.check_full_solve(pa_from_lp(lp), LIB, persistent=False, options=opts)
    if seed == 0:
        s, _ = parity.check_full_solve(pa_from_lp(feasible_lp(0, 6, 2, 6, 4, 24, frac=True, maxval=9, density=0.4)), LIB,
                                       expect_escalation=True, options=opts)
        assert s.wide_handoffs > 0


@pytest.mark.parametrize("seed", range(8))
def test_stage_per_launch_path(seed):
    """the one-launch-per-stage driver (what multi-GPU solves use) gives the same results"""
    lp = feasible_lp(seed + 200, 3, 1, 3, 2, 10, frac=(seed % 2 == 0), dup_rows=(1 if seed % 4 == 0 else 0))
    parity.check_full_solve(pa_from_lp(lp), LIB, persistent=False)
    if seed == 0:
        parity.check_full_solve(pa_from_lp(feasible_lp(0, 6, 2, 6, 4, 24, frac=True, maxval=9, density=0.4)), LIB,
                                expect_escalation=True, persistent=False)


@pytest.mark.parametrize("seed", range(8))
def test_from_basis_warm_start(seed):
    """`IM::from_basis` (carry/mod.rs:428-463): the Carry rebuilt from a basis alone equals the iterated one"""
    lp = feasible_lp(seed + 300, 3, 1, 3, 2, 10, frac=(seed % 2 == 0))
    parity.check_from_basis(pa_from_lp(lp), LIB, pivots_before=seed % 4)
    if seed == 0:      # escalated entries, and a sparse structured basis
        assert parity.check_from_basis(pa_from_lp(feasible_lp(0, 6, 2, 6, 4, 24, frac=True, maxval=9, density=0.4)), LIB, pivots_before=6)
        assert parity.check_from_basis(synthetic_lp(200, 2000, 20, "tu"), LIB, pivots_before=15)


@pytest.mark.parametrize("m,n,k", [(60, 400, 6), (200, 2000, 20), (2000, 20000, 20)])
def test_synthetic_tu(m, n, k):
    """BASELINE config 5a scaled down to sizes the oracle finishes: int64 tier, no escalation."""
    parity.check_full_solve(synthetic_lp(m, n, k, "tu"), LIB, expect_escalation=False, heap_words=1 << 24)


def test_synthetic_growth():
    parity.check_full_solve(synthetic_lp(80, 400, 10, "growth"), LIB, expect_escalation=True)


def test_escalation_to_multi_limb():
    parity.check_full_solve(pa_from_lp(feasible_lp(0, 6, 2, 6, 4, 24, frac=True, maxval=9, density=0.4)), LIB,
                            expect_escalation=True)
    s, o = parity.check_full_solve(pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4)), LIB,
                                   expect_escalation=True)
    assert o.max_bits() > 128 and s.max_limbs >= 5


def test_limb_pool_compaction():
    pa = pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4))
    s, _ = parity.check_full_solve(pa, LIB, heap_words=1 << 17, expect_escalation=True)
    assert s.compactions > 0


def test_heap_exhaustion_is_an_error_not_garbage():
    pa = pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4))
    c = capi.Carry(pa, heap_words=1 << 11)
    with pytest.raises(capi.RelpError) as e:
        c.solve_relaxation()
    assert e.value.status == capi.ERR_HEAP_EXHAUSTED


def test_full_size_properties():
    """BASELINE config 5a at full size (m=20k, n=200k): the oracle cannot finish it, so check
    size-independent properties of the optimum: primal feasibility A x <= b, x >= 0 (exact, int64
    tier), dual feasibility (no negative reduced cost: the loop's own termination), objective
    = c.x, and that a second solve reproduces the identical pivot log (determinism)."""
    import numpy as np
    pa = synthetic_lp(20000, 200000, 20, "tu")
    c = capi.Carry(pa, heap_words=1 << 24)
    assert c.solve_relaxation() == capi.OPTIMAL
    s = c.stats()
    assert s.escalated == 0
    bfs = c.current_bfs()
    x = np.zeros(pa.n, dtype=object)
    for j, v in bfs:
        assert v > 0
        x[j] = v
    # A x = b with slacks included (every row has its slack column): check row sums exactly
    ax = [Fr(0)] * pa.m
    for j, v in bfs:
        for t in range(int(pa.col_ptr[j]), int(pa.col_ptr[j + 1])):
            ax[int(pa.row_idx[t])] += v * Fr(int(pa.val_num[t]), int(pa.val_den[t]))
    assert all(ax[i] == Fr(int(pa.rhs_num[i]), int(pa.rhs_den[i])) for i in range(pa.m))
    obj = sum(v * Fr(int(pa.cost_num[j]), int(pa.cost_den[j])) for j, v in bfs)
    assert obj == c.get_objective_function_value()
    log1 = c.log()
    c.close()
    c2 = capi.Carry(pa, heap_words=1 << 24)
    assert c2.solve_relaxation() == capi.OPTIMAL
    assert c2.log() == log1
    c2.close()


def test_context_reuse_after_row_removal():
    lp = feasible_lp(5, 3, 1, 3, 2, 10, frac=True, dup_rows=2)
    other = feasible_lp(9, 4, 1, 4, 2, 16, frac=False)
    parity.check_reuse_after_row_removal(pa_from_lp(lp), pa_from_lp(other), LIB)


def test_load_matrix_validation():
    parity.check_load_validation(pa_from_lp(feasible_lp(3, 3, 1, 3, 2, 10, frac=True)), LIB)


def test_set_costs_after_phase_one():
    parity.check_set_costs(pa_from_lp(feasible_lp(7, 6, 2, 4, 2, 12, frac=True)), LIB)


def test_artificial_cleanup_rule():
    parity.check_cleanup_any(LIB)


def test_right_hand_side_beyond_int64():
    parity.check_big_rhs(pa_from_lp(feasible_lp(11, 6, 2, 5, 3, 14, frac=True)), LIB)


@pytest.mark.parametrize("m,frac", [(2500, 0), (9000, 0), (9000, 1)])
def test_dense_update_ring_at_scale(m, frac):
    """update_dense_tma where its pipeline turns over: ld = 2528 / 9024 rows = 3 / 9 row tiles with a
    partial last one, S = m + 1 columns => 25 / 273 tiles per CTA through the 6-stage ring (empty-barrier waits, parity
    flips, d_i reloads at every row-tile change), on integers and on general rationals.  The result is compared on the
    device with the element-wise path (the one every oracle-parity test above exercises) run on the same fill."""
    pa = synthetic_lp(m, 2 * m, 20, "tu")
    c = capi.Carry(pa, heap_words=1 << 22)
    c.set_option("k1_bench_frac", frac)
    c.create_for_partially_artificial()
    bad, n = c.debug_k1_verify(seed=20260101 + m)
    assert n == m * (m + 1) and bad == 0, (bad, n)
    c.close()


@pytest.mark.parametrize("persistent,long_mode", [(True, 1), (True, 0), (False, 0)])
def test_cooperative_multi_limb_tier_in_solves(persistent, long_mode):
    """option coop_from = 2: every operation on values beyond two limbs takes the arbitrary-length routines of
    relp_wcb.cuh instead of the fixed-size per-thread ones -- whole warps in the Carry update's multi-limb stage, one
    lane (with its per-thread scratch slab once the solve is in long mode) elsewhere -- same pivots, same Carry as the
    oracle.  long_mode = 1 runs the full copy of the persistent loop from the first pivot (the fast copy has no such tier
    and hands over only when a value beyond 12 limbs is stored)."""
    opts = {"coop_from": 2, "long_mode": long_mode}
    s, o = parity.check_full_solve(pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4)), LIB,
                                   expect_escalation=True, options=opts, persistent=persistent, heap_words=1 << 25)
    assert o.max_bits() > 128
    parity.check_full_solve(synthetic_lp(80, 400, 10, "growth"), LIB, expect_escalation=True, options=opts, persistent=persistent,
                            heap_words=1 << 25)
    parity.check_full_solve(pa_from_lp(feasible_lp(5, 3, 1, 3, 2, 10, frac=True, dup_rows=1)), LIB, options=opts, persistent=persistent)
