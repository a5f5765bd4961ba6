"""Generates tests/golden/mps/*.npz + tests/golden/mps_expected.json.

Run HERE (the container that has /root/reference); the outputs are committed because the GPU box has
no /root/reference.  For every problem file of the reference's integration tests that the oracle
finishes in seconds, this script
  1. reads the file with the product's reader (rust_lp_b200.mps) and standardises it
     (rust_lp_b200.general_form, no presolve),
  2. stores the resulting provider arrays (the input of relp_load_matrix) and the data needed to map
     a basic solution back to the original variables,
  3. stores the value the REFERENCE'S OWN TEST asserts (file:line cited) and its tolerance,
  4. solves with the oracle and stores status, pivot count, exact objective, exact solution, widest
     number and a hash of the pivot log, so that the GPU parity test can compare bit for bit even
     where it does not re-run the oracle.
The hand-written problems under tests/golden/mps_own/ go through the same steps; their independent
check is scipy's HiGHS (float) at generation time.

    python tests/golden/make_mps_golden.py
"""
import hashlib
import json
import os
import sys
from fractions import Fraction as Fr

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import coracle                                   # noqa: E402
from rust_lp_b200 import general_form, mps                   # noqa: E402

REF = "/root/reference/tests/"
RULES = (1, 2)    # FirstProfitableWithMemory in phase one, SteepestDescent in phase two (algorithm/two_phase/mod.rs)

# (name, file, reference assertion, tolerance | None = exact, citation)
CASES = [
    ("burkardt_adlittle", "burkardt/problem_files/adlittle.mps",
     "24975305659811992079614961229/120651674036153428931840", None, "tests/burkardt/test.rs:35-53"),
    ("burkardt_afiro", "burkardt/problem_files/afiro.mps", "-406659/875", None, "tests/burkardt/test.rs:57-110"),
    ("burkardt_maros", "burkardt/problem_files/maros.mps", "385/3", None, "tests/burkardt/test.rs:123-146"),
    ("burkardt_testprob", "burkardt/problem_files/testprob.mps", "54", None, "tests/burkardt/test.rs:161-183"),
    ("burkardt_nazareth", "burkardt/problem_files/nazareth.mps", "unbounded", None, "tests/burkardt/test.rs:150-158"),
    ("netlib_ADLITTLE", "netlib/problem_files/ADLITTLE.SIF", 2.254949632e+05, 1e-3, "tests/netlib/test.rs:10-14"),
    ("netlib_AFIRO", "netlib/problem_files/AFIRO.SIF", -464.75314, 1e-3, "tests/netlib/test.rs:17-21"),
    ("netlib_SC50A", "netlib/problem_files/SC50A.SIF", -6.457507706e+01, 1e-5, "tests/netlib/test.rs:24-28"),
    ("netlib_SC50B", "netlib/problem_files/SC50B.SIF", -70.0, 1e-10, "tests/netlib/test.rs:31-35"),
    ("netlib_KB2", "netlib/problem_files/KB2.SIF", -1.749900130e+03, 1e-3, "tests/netlib/test.rs:38-42"),
    ("netlib_SC105", "netlib/problem_files/SC105.SIF", -5.220206121e+01, 1e-3, "tests/netlib/test.rs:45-49"),
    ("netlib_STOCFOR1", "netlib/problem_files/STOCFOR1.SIF", -4.113197622e+04, 1e-3, "tests/netlib/test.rs:52-56"),
    ("netlib_BLEND", "netlib/problem_files/BLEND.SIF", -30.81215, 1e-3, "tests/netlib/test.rs:59-63"),
    ("netlib_SCAGR7", "netlib/problem_files/SCAGR7.SIF", -2.331389824e+06, 1e-1, "tests/netlib/test.rs:66-70"),
    ("netlib_SC205", "netlib/problem_files/SC205.SIF", -5.220206121e+01, 1e-5, "tests/netlib/test.rs:73-77"),
    ("netlib_SHARE2B", "netlib/problem_files/SHARE2B.SIF", -4.157322407e+02, 1e-5, "tests/netlib/test.rs:80-84"),
    ("netlib_RECIPELP", "netlib/problem_files/RECIPELP.SIF", -0.266616e3, 1e-2, "tests/netlib/test.rs:87-91"),
    ("netlib_LOTFI", "netlib/problem_files/LOTFI.SIF", -0.2526470606188e2, 1e-6, "tests/netlib/test.rs:94-98"),
    ("netlib_VTP-BASE", "netlib/problem_files/VTP-BASE.SIF", 0.1298314624613613657395984384889e6, 1e-2,
     "tests/netlib/test.rs:101-105"),
    ("netlib_SHARE1B", "netlib/problem_files/SHARE1B.SIF", -0.7658931857918568112797274346007e5, 1e-3,
     "tests/netlib/test.rs:108-112"),
    ("netlib_BOEING2", "netlib/problem_files/BOEING2.SIF", -0.31501872801520287870462195913263e3, 1e-3,
     "tests/netlib/test.rs:115-119"),
    ("netlib_ISRAEL", "netlib/problem_files/ISRAEL.SIF", -8.9664482186e+05, 1e-4,
     "tests/netlib/problem_files/README:125 (published optimum; BASELINE.json configs[1])"),
    # the reference ignores this test ("Incorrect optimal value", tests/netlib/test.rs:129-134): an artificial column
    # re-enters the basis in phase one and survives the clean-up -> status -8 here instead of a wrong optimum
    ("netlib_SCORPION", "netlib/problem_files/SCORPION.SIF", "artificial_basic", None, "tests/netlib/test.rs:129-134"),
    ("miplib_50v-10", "miplib/problem_files/50v-10.mps", 2879.065687, 1e-3, "tests/miplib/test.rs:4-6"),
]
UNICAMP = [  # tests/unicamp/test.rs: exact objective and solution
    ("unicamp_model_data_1", "unicamp/problem_files/model_data_1.mps", "123/38", None, "tests/unicamp/test.rs:10-26"),
    ("unicamp_model_data_3_1", "unicamp/problem_files/model_data_3_1.mps", "70", None, "tests/unicamp/test.rs:40-50"),
    ("unicamp_model_data_3_2", "unicamp/problem_files/model_data_3_2.mps", "180", None, "tests/unicamp/test.rs:53-62"),
    ("unicamp_model_data_3_3", "unicamp/problem_files/model_data_3_3.mps", "245", None, "tests/unicamp/test.rs:65-74"),
    ("unicamp_model_data_3_4", "unicamp/problem_files/model_data_3_4.mps", "2250", None, "tests/unicamp/test.rs:77-88"),
    ("unicamp_model_data_4", "unicamp/problem_files/model_data_4.mps", "7", None, "tests/unicamp/test.rs:91-101"),
]
OWN = ["ranges_bounds", "max_negative_rhs", "degenerate_redundant", "infeasible", "unbounded_free"]

STATUS = {1: "optimal", 2: "unbounded", 3: "infeasible", -8: "artificial_basic"}


def log_hash(log):
    h = hashlib.sha256()
    for rec in log:
        h.update(("%d,%d,%d;" % (rec[0], rec[1], rec[2])).encode())
    return h.hexdigest()


def highs_check(m):
    """Independent float solve of the ORIGINAL problem (scipy HiGHS), straight from the MPS sections."""
    from scipy.optimize import linprog
    gf = general_form.GeneralForm(m)
    n, rows = gf.n_orig, len(m.rows)
    A = np.zeros((rows, n))
    for j, col in enumerate(gf.cols):
        for i, v in col:
            A[i, j] = float(v)
    c = np.array([float(x) for x in gf.cost])
    A_ub, b_ub, A_eq, b_eq = [], [], [], []
    for i, t in enumerate(gf.types):
        b = float(gf.b[i])
        if t == "E":
            A_eq.append(A[i]); b_eq.append(b)
        elif t == "L":
            A_ub.append(A[i]); b_ub.append(b)
        elif t == "G":
            A_ub.append(-A[i]); b_ub.append(-b)
        else:                                   # range [b - r, b]
            A_ub.append(A[i]); b_ub.append(b)
            A_ub.append(-A[i]); b_ub.append(-(b - float(t[1])))
    bounds = [(None if lo is None else float(lo), None if up is None else float(up)) for lo, up in zip(gf.lower, gf.upper)]
    res = linprog(-c if m.maximize else c, A_ub=np.array(A_ub) if A_ub else None, b_ub=b_ub or None,
                  A_eq=np.array(A_eq) if A_eq else None, b_eq=b_eq or None, bounds=bounds, method="highs")
    if res.status == 0:
        return "optimal", float(-res.fun if m.maximize else res.fun) + float(m.objective_constant)
    return {2: "infeasible", 3: "unbounded"}.get(res.status, "status%d" % res.status), None


def process(name, path, fixed, reference, tol, cite, own=False):
    m = mps.import_file(path, fixed=fixed)
    gf, pa = general_form.from_mps(m)
    state = gf.recon_state()
    o = coracle.COracle(pa)
    st = o.solve(*RULES)
    entry = {"name": name, "source": cite, "m": pa.m, "n": pa.n, "nnz": len(pa.row_idx),
             "reference": reference, "tolerance": tol, "status": STATUS[st], "pivots": o.nr_pivots,
             "max_bits": o.max_bits(), "log_sha256": log_hash(o.log()), "rows_removed": pa.m - o.m}
    if st == 1:
        obj, sol = general_form.reconstruct(state, o.current_bfs())
        entry["objective"] = str(obj)
        entry["solution_sha256"] = hashlib.sha256(json.dumps({k: str(v) for k, v in sol.items()}, sort_keys=True).encode()).hexdigest()
        ref_obj = general_form.minimisation_objective(state, obj)
        if isinstance(reference, str):
            assert reference != "unbounded" and Fr(reference) == ref_obj, (name, reference, obj)
        elif reference is not None:
            assert abs(float(ref_obj) - reference) < tol, (name, reference, float(obj))
    else:
        assert reference in (STATUS[st], None), (name, reference, STATUS[st])
    if own:
        hs, hv = highs_check(m)
        entry["highs"] = [hs, hv]
        assert hs == STATUS[st], (name, hs, STATUS[st])
        if hv is not None:
            assert abs(hv - float(Fr(entry["objective"]))) < 1e-6 * max(1.0, abs(hv)), (name, hv, entry["objective"])
    os.makedirs(os.path.join(HERE, "mps"), exist_ok=True)
    np.savez_compressed(
        os.path.join(HERE, "mps", name + ".npz"), m=pa.m, n=pa.n, n_normal=pa.n_normal,
        col_ptr=np.asarray(pa.col_ptr, np.int64), row_idx=np.asarray(pa.row_idx, np.int32),
        val_num=np.asarray(pa.val_num, np.int64), val_den=np.asarray(pa.val_den, np.int64),
        cost_num=np.asarray(pa.cost_num, np.int64), cost_den=np.asarray(pa.cost_den, np.int64),
        rhs_num=np.asarray(pa.rhs_num, np.int64), rhs_den=np.asarray(pa.rhs_den, np.int64),
        piv_rows=np.asarray(pa.piv_rows, np.int32), piv_cols=np.asarray(pa.piv_cols, np.int32),
        recon=np.frombuffer(json.dumps(state).encode(), np.uint8))
    print("%-28s m %4d n %4d pivots %5d bits %4d %s %s" % (name, pa.m, pa.n, o.nr_pivots, o.max_bits(), STATUS[st],
                                                          entry.get("objective", "")[:40]), flush=True)
    return entry


def main():
    out = []
    for name, rel, reference, tol, cite in CASES + UNICAMP:
        out.append(process(name, REF + rel, rel.endswith(".SIF"), reference, tol, cite))
    for name in OWN:
        out.append(process("own_" + name, os.path.join(HERE, "mps_own", name + ".mps"), False, None, None,
                           "tests/golden/mps_own/%s.mps (hand-written; checked against scipy HiGHS)" % name, own=True))
    with open(os.path.join(HERE, "mps_expected.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
