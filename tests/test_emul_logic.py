"""CPU-only check of the product's HOST logic and kernel indexing: rust-lp_b200/csrc/*.cu compiled
with g++ against tests/native/cuda_emul.h (every launch = one sequential thread), driven through the
same C ABI and the same parity checks as the GPU suite.  This does not replace the GPU tests (it
exercises no parallelism) and is never used by the product."""
import os
import subprocess

import pytest

import parity
from helpers import pa_from_lp, problem_1_lp, problem_2_lp, feasible_lp, random_lp
from rust_lp_b200.synth import synthetic_lp
from test_oracle_fixtures import P1_ART, P1_BFS, P2_ART, P2_BFS
from fractions import Fraction as Fr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "rust-lp_b200", "csrc")
OUT = os.path.join(ROOT, "tests", "native", "_build", "librelp_emul.so")


@pytest.fixture(scope="module")
def lib():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    deps = [os.path.join(CSRC, f) for f in ("relp_capi.cu", "relp_kernels.cuh", "relp_arith.cuh")]
    deps += [os.path.join(ROOT, "tests", "native", "cuda_emul.h"), os.path.join(ROOT, "include", "relp_b200.h")]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        subprocess.check_call(["g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-DRELP_EMUL",
                               "-fsanitize=undefined", "-fno-sanitize-recover=undefined",
                               "-I", os.path.join(ROOT, "tests", "native"), "-I", CSRC, "-x", "c++",
                               deps[0], "-o", OUT, "-ldl"])
    return OUT


@pytest.mark.parametrize("make_lp,art,bfs,opt,obj", [
    (problem_2_lp, P2_ART, P2_BFS, [(1, Fr(1, 2)), (3, Fr(5, 2)), (4, Fr(3, 2))], Fr(9, 2)),
    (problem_1_lp, P1_ART, P1_BFS, [(0, Fr(4)), (2, Fr(6)), (5, Fr(2))], Fr(58)),
])
def test_reference_fixtures(lib, make_lp, art, bfs, opt, obj):
    parity.check_fixture(make_lp, art, bfs, opt, obj, lib)


@pytest.mark.parametrize("seed", range(40))
def test_full_solve_random(lib, seed):
    lp = feasible_lp(seed, 3, 1, 3, 2, 10, frac=(seed % 2 == 0), dup_rows=(1 if seed % 5 == 0 else 0))
    parity.check_full_solve(pa_from_lp(lp), lib)


@pytest.mark.parametrize("seed", range(100, 112))
def test_full_solve_random_unfiltered(lib, seed):
    """may be infeasible or unbounded"""
    lp = random_lp(seed, 2, 1, 3, 2, 8, frac=(seed % 2 == 0))
    parity.check_full_solve(pa_from_lp(lp), lib)


@pytest.mark.parametrize("seed", range(6))
def test_stepwise_traits(lib, seed):
    lp = feasible_lp(seed + 50, 2, 1, 3, 1, 8, frac=(seed % 2 == 1))
    parity.check_stepwise(lp, lib)


@pytest.mark.parametrize("seed", range(4))
def test_resumable(lib, seed):
    lp = feasible_lp(seed + 70, 3, 0, 3, 2, 9, frac=True)
    parity.check_resumable(pa_from_lp(lp), lib)


@pytest.mark.parametrize("seed", range(8))
def test_stage_per_launch_path(lib, seed):
    """the one-launch-per-stage driver (what multi-GPU solves use) gives the same results"""
    lp = feasible_lp(seed + 200, 3, 1, 3, 2, 10, frac=(seed % 2 == 0), dup_rows=(1 if seed % 4 == 0 else 0))
    parity.check_full_solve(pa_from_lp(lp), lib, persistent=False)
    if seed == 0:
        parity.check_full_solve(pa_from_lp(feasible_lp(0, 6, 2, 6, 4, 24, frac=True, maxval=9, density=0.4)), lib,
                                expect_escalation=True, persistent=False)


def test_synthetic_tu(lib):
    parity.check_full_solve(synthetic_lp(60, 400, 6, "tu"), lib, expect_escalation=False)


def test_escalation_to_multi_limb(lib):
    """entries outgrow int64/int64 (oracle: 95 and 144 bits): per-element escalation, same results"""
    parity.check_full_solve(pa_from_lp(feasible_lp(0, 6, 2, 6, 4, 24, frac=True, maxval=9, density=0.4)), lib,
                            expect_escalation=True)
    s, o = parity.check_full_solve(pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4)), lib,
                                   expect_escalation=True)
    assert o.max_bits() > 128 and s.max_limbs >= 5


def test_limb_pool_compaction(lib):
    """a heap so small that the persistent limb pool must be compacted many times"""
    pa = pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4))
    s, _ = parity.check_full_solve(pa, lib, heap_words=1 << 17, expect_escalation=True)
    assert s.compactions > 0


def test_heap_exhaustion_is_an_error_not_garbage(lib):
    from rust_lp_b200 import capi
    pa = pa_from_lp(feasible_lp(0, 10, 2, 10, 6, 40, frac=True, maxval=9, density=0.4))
    c = capi.Carry(pa, heap_words=1 << 11, lib_path=lib)
    with pytest.raises(capi.RelpError) as e:
        c.solve_relaxation()
    assert e.value.status == capi.ERR_HEAP_EXHAUSTED


def _mps_names():
    from test_mps_golden import NAMES
    # the sequential stand-in needs 25-50 s each for the two longest escalating solves (SHARE1B: 294 bits, ISRAEL: 132):
    # they stay in the GPU suite (tests/test_mps_golden.py -m gpu); SC205, BOEING2, STOCFOR1, KB2 cover escalation here
    return [n for n in NAMES if n not in ("netlib_SCORPION", "netlib_SHARE1B", "netlib_ISRAEL")]


@pytest.mark.parametrize("name", _mps_names())
def test_mps_fixtures(lib, name):
    """Problems read by the MPS reader (tests/golden/mps): bound rows, range rows, free variables, and
    redundant rows whose artificial index differs from its row (VTP-BASE, degenerate_redundant)."""
    from test_mps_golden import load_fixture, BY_NAME
    pa, _ = load_fixture(name)
    s, o = parity.check_full_solve(pa, lib)
    assert pa.m - o.m == BY_NAME[name]["rows_removed"]


def test_artificial_left_in_basis_is_an_error(lib):
    """netlib SCORPION without presolve: an artificial column re-enters the basis in phase one on another
    row and survives the clean-up; the reference underflows there (carry/mod.rs:500).  Status -8 here."""
    from test_mps_golden import load_fixture
    from rust_lp_b200 import capi
    pa, _ = load_fixture("netlib_SCORPION")
    c = capi.Carry(pa, heap_words=1 << 22, lib_path=lib)
    with pytest.raises(capi.RelpError) as ei:
        c.solve_relaxation(1, 2)
    assert ei.value.status == capi.ERR_ARTIFICIAL_BASIC
    c.close()


@pytest.mark.parametrize("seed", range(12))
def test_from_basis_warm_start(lib, seed):
    lp = feasible_lp(seed + 300, 3, 1, 3, 2, 10, frac=(seed % 2 == 0))
    parity.check_from_basis(pa_from_lp(lp), lib, pivots_before=seed % 4)


def test_from_basis_singular(lib):
    from rust_lp_b200 import capi
    lp = feasible_lp(301, 3, 1, 3, 2, 10)
    pa = pa_from_lp(lp)
    c = capi.Carry(pa, heap_words=1 << 22, lib_path=lib)
    # a column that is all zeros in the basis rows cannot happen with a repeated... use a dependent pair: two
    # copies are rejected as a repeated column; take columns until a singular set is found via the oracle
    import itertools
    from oracle import coracle
    m = pa.m
    for comb in itertools.combinations(range(pa.n), m):
        o = coracle.COracle(pa)
        if o.from_basis(list(comb)) != 0:
            with pytest.raises(capi.RelpError):
                c.from_basis(list(comb))
            break
    c.close()


def test_context_reuse_after_row_removal(lib):
    lp = feasible_lp(5, 3, 1, 3, 2, 10, frac=True, dup_rows=2)
    other = feasible_lp(9, 4, 1, 4, 2, 16, frac=False)
    parity.check_reuse_after_row_removal(pa_from_lp(lp), pa_from_lp(other), lib)


def test_load_matrix_validation(lib):
    parity.check_load_validation(pa_from_lp(feasible_lp(3, 3, 1, 3, 2, 10, frac=True)), lib)


def test_set_costs_after_phase_one(lib):
    parity.check_set_costs(pa_from_lp(feasible_lp(7, 6, 2, 4, 2, 12, frac=True)), lib)


def test_artificial_cleanup_rule(lib):
    parity.check_cleanup_any(lib)


def test_right_hand_side_beyond_int64(lib):
    parity.check_big_rhs(pa_from_lp(feasible_lp(11, 6, 2, 5, 3, 14, frac=True)), lib)
