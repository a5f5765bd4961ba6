"""Golden for BASELINE.json configs[4] at FULL size (synthetic 5a: m = 20 000, n = 200 000, 20 ones per column, seed
20260101): the oracle (oracle/simplex.cpp, the CPU restatement of the reference path) solves the LP once -- several
minutes on one core -- and this script stores, at pivots 500 / 1000 / 1500 / 2000 and at the optimum, the sha256 of
the pivot log so far, of the basis and of `b`, and the exact objective (tests/golden/synth5a_expected.json).  The GPU
test (tests/test_synth_golden.py) stops the device loop at the same pivot counts and compares; bench.py compares the
hash of the whole log of its timed solve at every N.

    python tests/golden/make_synth_golden.py            # ~10 min, single thread
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from oracle import coracle                                   # noqa: E402
from rust_lp_b200.synth import synthetic_lp                  # noqa: E402
from make_prefix_golden import state_digests                 # noqa: E402

M, N, K = 20000, 200000, 20
CHECKPOINTS = [500, 1000, 1500, 2000]


def main():
    pa = synthetic_lp(M, N, K, "tu")
    o = coracle.COracle(pa)
    o.begin(1, 2)
    out = {"workload": f"synthetic 5a tu m={M} n={N} nnz/col={K} seed=20260101", "rules": [1, 2], "checkpoints": []}
    t0 = time.time()
    done = 0

    def snap(tag):
        mo, _, b, basis, _ = o.carry()
        e = {"pivots": o.nr_pivots, "what": tag, "oracle_seconds": round(time.time() - t0, 1)}
        e.update(state_digests([(int(x[0]), int(x[1]), int(x[2])) for x in o.log()], basis, b, -mo))
        out["checkpoints"].append(e)
        print(e["pivots"], tag, e["objective"], e["oracle_seconds"], flush=True)

    for cp in CHECKPOINTS:
        st = o.advance(cp - done)
        done = o.nr_pivots
        if st != coracle.RUNNING:
            break
        snap("running")
    st = o.advance(-1)
    assert st == coracle.OPTIMAL, st
    snap("optimal")
    out["pivots_to_optimal"] = o.nr_pivots
    out["oracle_seconds_to_optimal"] = round(time.time() - t0, 1)
    out["max_bits"] = o.max_bits()
    with open(os.path.join(HERE, "synth5a_expected.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
