"""BASELINE.json configs[2], full solve: netlib 25FV47 to its exact optimum (about 10-15 minutes on a B200, so only
with RELP_SLOW=1; the round's run is profiles/big_25fv47_r02.json).  The CPU oracle cannot finish this problem, so the
expected value is pinned three ways: (1) the optimum is a certificate that is checked exactly on the host here -- A x = b,
x >= 0, objective = c.x --, (2) it equals the exact optimum published by Koch, "The final Netlib-LP results", to every
digit that value has, (3) it is the same reduced fraction, pivot count and removed-row count as the committed run.
The first 1 500 pivots of the same solve are compared with the oracle in tests/test_mps_prefix.py."""
import json
import os
from decimal import Decimal, getcontext
from fractions import Fraction as Fr

import pytest

from rust_lp_b200 import capi, general_form
from test_mps_golden import GOLD, np
from test_mps_prefix import load

BIG = json.load(open(os.path.join(GOLD, "big_expected.json")))


def test_committed_optimum_matches_the_published_exact_value():
    num, den = (int(x) for x in BIG["objective_exact"].split("/"))
    getcontext().prec = 60
    got = Decimal(num) / Decimal(den)
    koch = Decimal(BIG["koch_exact_value"])
    assert abs(got - koch) < Decimal("1e-27")                        # 32 significant digits published
    assert abs(float(Fr(num, den)) - 5501.8459) < 2e-5               # the reference test's rounded constant


@pytest.mark.gpu
@pytest.mark.skipif(os.environ.get("RELP_SLOW") != "1", reason="a 10-15 minute exact solve: set RELP_SLOW=1")
def test_25fv47_full_exact_solve():
    name = BIG["name"]
    pa = load(name)
    c = capi.Carry(pa, heap_words=1 << 32)
    assert c.solve_relaxation(1, 2) == capi.OPTIMAL
    bfs = c.current_bfs()
    ax = [Fr(0)] * pa.m
    for j, v in bfs:
        assert v > 0
        for t in range(int(pa.col_ptr[j]), int(pa.col_ptr[j + 1])):
            ax[int(pa.row_idx[t])] += v * Fr(int(pa.val_num[t]), int(pa.val_den[t]))
    assert all(ax[i] == Fr(int(pa.rhs_num[i]), int(pa.rhs_den[i])) for i in range(pa.m))
    assert sum(v * Fr(int(pa.cost_num[j]), int(pa.cost_den[j])) for j, v in bfs) == c.get_objective_function_value()
    z = np.load(os.path.join(GOLD, "mps_prefix", name + ".npz"))
    obj, _ = general_form.reconstruct(json.loads(bytes(z["recon"]).decode()), bfs)
    assert str(obj) == BIG["objective_exact"]
    assert c.stats().pivots == BIG["pivots"] and len(c.rows_removed()) == BIG["rows_removed"]
    c.close()
