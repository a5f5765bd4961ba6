"""Generates tests/golden/mps_prefix/*.npz + tests/golden/prefix_expected.json: the problem files of
BASELINE.json configs[2..3] that the oracle does NOT finish in seconds (netlib 25FV47 and the PILOT
family, miplib acc-tight4).  Parity on these is pinned on a PREFIX of the solve: the oracle runs the
first K pivots; the fixture stores the sha256 of that pivot log, of the basis and of `b` after pivot
K, the exact objective value at that point, and the widest number seen.  The GPU test stops the
device loop after the same K pivots (relp_solve_relaxation's pivot budget) and compares all of them.

Run HERE (the container that has /root/reference); outputs are committed.

    python tests/golden/make_prefix_golden.py
"""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import coracle                                   # noqa: E402
from rust_lp_b200 import general_form, mps                   # noqa: E402

REF = "/root/reference/tests/"
# (name, file, pivots in the prefix, reference's own expectation for the FULL solve and where it is stated)
CASES = [
    ("netlib_25FV47", "netlib/problem_files/25FV47.SIF", 900, 5501.8459, "tests/netlib/test.rs:152-158 (#[ignore]d there)"),
    ("netlib_PILOT4", "netlib/problem_files/PILOT4.SIF", 550, -2581.1392613, "tests/netlib/problem_files/README (published optimum)"),
    ("netlib_PILOT-WE", "netlib/problem_files/PILOT-WE.SIF", 700, -2720107.5328, "tests/netlib/problem_files/README (published optimum)"),
    ("netlib_PILOT", "netlib/problem_files/PILOT.SIF", 1500, -557.48972928, "tests/netlib/problem_files/README (published optimum)"),
    ("miplib_acc-tight4", "miplib/problem_files/acc-tight4.mps", 1500, 0.0, "tests/miplib/test.rs (LP relaxation)"),
]


def digest(items):
    h = hashlib.sha256()
    for x in items:
        h.update((str(x) + ";").encode())
    return h.hexdigest()


def state_digests(log_triples, basis, b, objective):
    return {"log_sha256": digest("%d,%d,%d" % t for t in log_triples), "basis_sha256": digest(int(x) for x in basis),
            "b_sha256": digest(b), "objective": str(objective)}


def main():
    out = []
    os.makedirs(os.path.join(HERE, "mps_prefix"), exist_ok=True)
    for name, rel, k, reference, cite in CASES:
        gf, pa = general_form.from_mps(mps.import_file(REF + rel, fixed=rel.endswith(".SIF")))
        o = coracle.COracle(pa)
        o.begin(1, 2)
        t0 = time.time()
        st = o.advance(k)
        dt = time.time() - t0
        assert st == 0 and o.nr_pivots == k, (name, st, o.nr_pivots)
        mo, _, b, basis, _ = o.carry()
        e = {"name": name, "source": cite, "m": pa.m, "n": pa.n, "nnz": len(pa.row_idx), "prefix_pivots": k,
             "full_solve_reference": reference, "max_bits": o.max_bits(), "stage": o.stage, "oracle_seconds": round(dt, 2)}
        e.update(state_digests([(int(x[0]), int(x[1]), int(x[2])) for x in o.log()], basis, b, -mo))
        out.append(e)
        np.savez_compressed(
            os.path.join(HERE, "mps_prefix", name + ".npz"), m=pa.m, n=pa.n, n_normal=pa.n_normal,
            col_ptr=np.asarray(pa.col_ptr, np.int64), row_idx=np.asarray(pa.row_idx, np.int32),
            val_num=np.asarray(pa.val_num, np.int64), val_den=np.asarray(pa.val_den, np.int64),
            cost_num=np.asarray(pa.cost_num, np.int64), cost_den=np.asarray(pa.cost_den, np.int64),
            rhs_num=np.asarray(pa.rhs_num, np.int64), rhs_den=np.asarray(pa.rhs_den, np.int64),
            piv_rows=np.asarray(pa.piv_rows, np.int32), piv_cols=np.asarray(pa.piv_cols, np.int32),
            recon=np.frombuffer(json.dumps(gf.recon_state()).encode(), np.uint8))
        print("%-20s m %5d n %5d prefix %5d bits %5d oracle %.1f s" % (name, pa.m, pa.n, k, o.max_bits(), dt), flush=True)
    with open(os.path.join(HERE, "prefix_expected.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
