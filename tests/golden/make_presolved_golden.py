"""Generates tests/golden/mps_presolved/*.npz + tests/golden/mps_presolved_expected.json: the same problem files as
make_mps_golden.py, but through the REFERENCE'S pipeline -- presolve first (rust_lp_b200.presolve, the restatement of
general_form/presolve/), then transform_variables / make_b_non_negative / minimisation (general_form/mod.rs:307-314).
These are the standard forms (rows, columns, initial basis) the reference's simplex actually sees, so the pivot logs
stored here are the ones a build of the reference would take.  Every objective is checked against the value the
reference's own test asserts.  Problems presolve solves completely are recorded as such.

Run HERE (the container that has /root/reference); outputs are committed.

    python tests/golden/make_presolved_golden.py
"""
import hashlib
import json
import os
import sys
import time
from fractions import Fraction as Fr

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from oracle import coracle                                   # noqa: E402
from rust_lp_b200 import general_form, mps                   # noqa: E402
from make_mps_golden import CASES, UNICAMP, REF, RULES, STATUS, log_hash   # noqa: E402

OUT = os.path.join(HERE, "mps_presolved")


def process(name, path, fixed, reference, tol, cite):
    m = mps.import_file(path, fixed=fixed)
    raw = general_form.GeneralForm(m)
    rows0, cols0 = len(raw.b), len(raw.cols)
    entry = {"name": name, "source": cite, "reference": reference, "tolerance": tol, "rows_before": rows0, "columns_before": cols0}
    try:
        gf, pa = general_form.from_mps(m, presolve=True)
    except general_form.SolvedByPresolve as e:
        entry.update({"status": "solved_by_presolve", "objective": str(e.objective)})
        print("%-28s solved by presolve: %s" % (name, e.objective), flush=True)
        return entry
    except general_form.Unbounded:
        entry.update({"status": "unbounded_by_presolve"})
        assert reference == "unbounded", (name, reference)
        print("%-28s unbounded (presolve)" % name, flush=True)
        return entry
    state = gf.recon_state()
    t0 = time.time()
    o = coracle.COracle(pa)
    st = o.solve(*RULES)
    entry.update({"m": pa.m, "n": pa.n, "nnz": len(pa.row_idx), "rows_after": len(gf.b), "columns_after": gf.n_orig,
                  "status": STATUS[st], "pivots": o.nr_pivots, "max_bits": o.max_bits(), "log_sha256": log_hash(o.log()),
                  "rows_removed": pa.m - o.m, "oracle_seconds": round(time.time() - t0, 2)})
    if st == 1:
        obj, sol = general_form.reconstruct(state, o.current_bfs())
        entry["objective"] = str(obj)
        entry["solution_sha256"] = hashlib.sha256(json.dumps({k: str(v) for k, v in sol.items()}, sort_keys=True).encode()).hexdigest()
        ref_obj = general_form.minimisation_objective(state, obj)
        if isinstance(reference, str):
            assert reference != "unbounded" and Fr(reference) == ref_obj, (name, reference, obj)
        elif reference is not None:
            assert abs(float(ref_obj) - reference) < tol, (name, reference, float(obj))
    elif STATUS[st] == "artificial_basic":
        # an artificial column re-entered the basis on another row and survives the clean-up: the reference itself
        # underflows there (carry/mod.rs:500); oracle and product report it (difference 2)
        entry["note"] = "artificial column still basic after phase one on this standard form: no result to compare"
    else:
        assert reference in (STATUS[st], None), (name, reference, STATUS[st])
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"), m=pa.m, n=pa.n, n_normal=pa.n_normal,
        col_ptr=np.asarray(pa.col_ptr, np.int64), row_idx=np.asarray(pa.row_idx, np.int32),
        val_num=np.asarray(pa.val_num, np.int64), val_den=np.asarray(pa.val_den, np.int64),
        cost_num=np.asarray(pa.cost_num, np.int64), cost_den=np.asarray(pa.cost_den, np.int64),
        rhs_num=np.asarray(pa.rhs_num, np.int64), rhs_den=np.asarray(pa.rhs_den, np.int64),
        piv_rows=np.asarray(pa.piv_rows, np.int32), piv_cols=np.asarray(pa.piv_cols, np.int32),
        recon=np.frombuffer(json.dumps(state).encode(), np.uint8),
        # right-hand sides beyond int64/int64 (presolve substitutes solved variables into b): row -> [numerator, denominator]
        rhs_exact=np.frombuffer(json.dumps({str(i): [str(f.numerator), str(f.denominator)] for i, f in pa.rhs_exact.items()}).encode(), np.uint8))
    print("%-28s rows %4d -> %4d cols %4d -> %4d  m %4d n %4d pivots %5d bits %4d %s %s" % (
        name, rows0, len(gf.b), cols0, gf.n_orig, pa.m, pa.n, o.nr_pivots, o.max_bits(), STATUS[st], entry.get("objective", "")[:40]), flush=True)
    return entry


# problems that only this pipeline can pin: un-presolved they are beyond what the CPU oracle finishes in minutes
PRESOLVED_ONLY = [
    ("netlib_BORE3D", "netlib/problem_files/BORE3D.SIF", 0.13730803942084927215581987251301e4, 1e-2, "tests/netlib/test.rs:122-126"),
]


def main():
    only = sys.argv[1:]
    out = []
    for name, rel, reference, tol, cite in CASES + UNICAMP + PRESOLVED_ONLY:
        if only and name not in only:
            continue
        out.append(process(name, REF + rel, rel.endswith(".SIF"), reference, tol, cite))
    f = os.path.join(HERE, "mps_presolved_expected.json")
    if only:                      # regenerate single problems: merge into the committed list
        old = json.load(open(f))
        names = {e["name"] for e in out}
        out = [e for e in old if e["name"] not in names] + out
    with open(f, "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
