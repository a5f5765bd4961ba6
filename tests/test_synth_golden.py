"""BASELINE.json configs[4] at FULL size (synthetic 5a, m = 20 000, n = 200 000): the oracle's whole solve, recorded
once by tests/golden/make_synth_golden.py (several minutes on one core), against the device loop stopped at the same
pivot counts -- sha256 of the pivot log, of the basis and of `b`, and the exact objective at pivots 500 / 1000 / 1500 /
2000 and at the optimum ("full-size parity is a 150-pivot prefix")."""
import json
import os
import sys

import pytest

from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
from make_prefix_golden import state_digests   # noqa: E402

EXPECTED = json.load(open(os.path.join(GOLD, "synth5a_expected.json")))
KEYS = ("log_sha256", "basis_sha256", "b_sha256", "objective")


def test_golden_is_the_whole_solve():
    cps = EXPECTED["checkpoints"]
    assert [c["pivots"] for c in cps[:-1]] == [500, 1000, 1500, 2000]
    assert cps[-1]["what"] == "optimal" and cps[-1]["pivots"] == EXPECTED["pivots_to_optimal"] == 2373
    assert EXPECTED["max_bits"] <= 63            # the int64 tier never overflows on the totally unimodular variant


@pytest.mark.gpu
def test_full_size_solve_matches_the_oracle_at_every_checkpoint():
    pa = synthetic_lp(20000, 200000, 20, "tu")
    c = capi.Carry(pa, heap_words=1 << 24)
    done = 0
    st = None
    for cp in EXPECTED["checkpoints"]:
        budget = cp["pivots"] - done if cp["what"] == "running" else -1
        st = c.solve_relaxation(1, 2, max_pivots=budget) if st is None else c.solve_continue(budget)
        log = c.log()
        done = len(log)
        assert done == cp["pivots"]
        assert st == (capi.OK if cp["what"] == "running" else capi.OPTIMAL)
        got = state_digests([(ord(x[0]), x[1], x[2]) for x in log], c.basis_indices(), c.b(), c.get_objective_function_value())
        assert {k: got[k] for k in KEYS} == {k: cp[k] for k in KEYS}, cp["pivots"]
    assert c.stats().escalated == 0
    c.close()
