"""BASELINE.json configs[2..3]: netlib 25FV47 / PILOT family and miplib acc-tight4.  The oracle does
not finish these in seconds, so parity is pinned on the first K pivots (tests/golden/
make_prefix_golden.py): identical pivot sequence, basis, `b` and objective after pivot K, with limb
escalation exercised (numbers of several hundred bits).

CPU: the oracle reproduces the committed digests.  GPU: the device loop, stopped after the same K
pivots through relp_solve_relaxation's budget, gives the same digests.
"""
import json
import os
import sys

import pytest

from oracle import coracle
from rust_lp_b200 import capi
from test_mps_golden import GOLD, ProviderArrays, np

sys.path.insert(0, GOLD)
from make_prefix_golden import state_digests   # noqa: E402

EXPECTED = json.load(open(os.path.join(GOLD, "prefix_expected.json")))
BY_NAME = {e["name"]: e for e in EXPECTED}
KEYS = ("log_sha256", "basis_sha256", "b_sha256", "objective")


def load(name):
    z = np.load(os.path.join(GOLD, "mps_prefix", name + ".npz"))
    return ProviderArrays(int(z["m"]), int(z["n"]), z["col_ptr"].tolist(), z["row_idx"].tolist(), z["val_num"].tolist(),
                          z["val_den"].tolist(), z["cost_num"].tolist(), z["cost_den"].tolist(), z["rhs_num"].tolist(),
                          z["rhs_den"].tolist(), z["piv_rows"].tolist(), z["piv_cols"].tolist(), int(z["n_normal"]))


@pytest.mark.parametrize("name", list(BY_NAME))
def test_oracle_prefix(name):
    e = BY_NAME[name]
    o = coracle.COracle(load(name))
    o.begin(1, 2)
    assert o.advance(e["prefix_pivots"]) == 0 and o.nr_pivots == e["prefix_pivots"]
    mo, _, b, basis, _ = o.carry()
    got = state_digests([(int(x[0]), int(x[1]), int(x[2])) for x in o.log()], basis, b, -mo)
    assert {k: got[k] for k in KEYS} == {k: e[k] for k in KEYS}
    assert o.max_bits() == e["max_bits"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(BY_NAME))
def test_gpu_prefix_matches_oracle(name):
    e = BY_NAME[name]
    c = capi.Carry(load(name), heap_words=1 << 26)
    st = c.solve_relaxation(1, 2, max_pivots=e["prefix_pivots"])
    assert st == capi.OK
    log = c.log()
    assert len(log) == e["prefix_pivots"]
    got = state_digests([(ord(x[0]), x[1], x[2]) for x in log], c.basis_indices(), c.b(), c.get_objective_function_value())
    assert {k: got[k] for k in KEYS} == {k: e[k] for k in KEYS}
    s = c.stats()
    assert (s.escalated > 0) == (e["max_bits"] > 63) or e["max_bits"] <= 63
    c.close()
