"""Full exact solves of the BASELINE.json configs[2..3] problems (tests/golden/mps_prefix) on the GPU, in
chunks of pivots, with a wall-clock limit per problem.  One JSON line per chunk -> gpurun_out/big_times.jsonl.
usage: python scripts/big_time.py [seconds per problem] [chunk] [names...]"""
import json, os, sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from rust_lp_b200 import capi
from test_mps_prefix import load, EXPECTED

limit = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 250
only = sys.argv[3:]
os.makedirs("gpurun_out", exist_ok=True)
out = open(os.path.join("gpurun_out", "big_times.jsonl"), "a")
for e in EXPECTED:
    name = e["name"]
    if only and name not in only: continue
    pa = load(name)
    c = capi.Carry(pa, heap_words=int(os.environ.get("RELP_HEAP_WORDS", str(1 << 28))))
    if os.environ.get("RELP_CLEANUP_ANY") == "1":      # artificial clean-up pivots on any nonzero element (default: the reference's zero-cost columns only)
        c.set_option("cleanup_any", 1)
    t0 = time.time()
    st, first = 0, True
    try:
        while st == 0 and time.time() - t0 < limit:
            t1 = time.time()
            st = c.solve_relaxation(1, 2, max_pivots=chunk) if first else c.solve_continue(chunk)
            first = False
            s = c.stats()
            rec = dict(name=name, m=pa.m, n=pa.n, status=st, pivots=s.pivots, chunk_s=time.time() - t1, total_s=time.time() - t0,
                       loop_ms=s.loop_ms, escalated=s.escalated, max_limbs=s.max_limbs, compactions=s.compactions,
                       narrow=s.narrow_pivots, phase=s.phase, heap_peak_words=s.heap_peak_words)
            print(json.dumps(rec), flush=True); out.write(json.dumps(rec) + "\n"); out.flush()
        if st == capi.OPTIMAL:
            # size-independent checks of the optimum, all exact: A x = b, x >= 0, objective = c x; then the objective of the
            # ORIGINAL problem (solution reconstruction) against the value the reference's own test / the netlib README states
            import numpy as np
            from fractions import Fraction as Fr
            from rust_lp_b200 import general_form
            z = np.load(os.path.join("tests", "golden", "mps_prefix", name + ".npz"))
            state = json.loads(bytes(z["recon"]).decode())
            bfs = c.current_bfs()
            ax = [Fr(0)] * pa.m
            for j, v in bfs:
                assert v > 0
                for t in range(int(pa.col_ptr[j]), int(pa.col_ptr[j + 1])):
                    ax[int(pa.row_idx[t])] += v * Fr(int(pa.val_num[t]), int(pa.val_den[t]))
            feasible = all(ax[i] == Fr(int(pa.rhs_num[i]), int(pa.rhs_den[i])) for i in range(pa.m))
            obj_std = sum(v * Fr(int(pa.cost_num[j]), int(pa.cost_den[j])) for j, v in bfs)
            obj, _sol = general_form.reconstruct(state, bfs)
            s = c.stats()
            sys.path.insert(0, os.path.join("tests", "golden"))
            from make_prefix_golden import state_digests
            dig = state_digests([(ord(x[0]), x[1], x[2]) for x in c.log()], c.basis_indices(), c.b(), c.get_objective_function_value())
            res = dict(name=name, status="optimal", digests={k: dig[k] for k in ("log_sha256", "basis_sha256", "b_sha256")}, pivots=int(s.pivots), total_s=time.time() - t0, loop_ms=s.loop_ms,
                       escalated=int(s.escalated), max_limbs=int(s.max_limbs), objective=float(obj), objective_exact=str(obj),
                       reference=e["full_solve_reference"], abs_diff=abs(float(obj) - e["full_solve_reference"]),
                       exact_Ax_eq_b=bool(feasible), objective_equals_cx=bool(obj_std == c.get_objective_function_value()),
                       rows_removed=len(c.rows_removed()), cleanup_any=os.environ.get("RELP_CLEANUP_ANY") == "1",
                       violated_rows=[i for i in range(pa.m) if ax[i] != Fr(int(pa.rhs_num[i]), int(pa.rhs_den[i]))][:50],
                       removed_rows=[int(i) for i in c.rows_removed()])
            print(json.dumps(res), flush=True)
            with open(os.path.join("gpurun_out", "big_result_" + name + ".json"), "w") as fh:
                json.dump(res, fh, indent=1)
    except capi.RelpError as err:
        print(name, "ERROR", err, flush=True)
        out.write(json.dumps(dict(name=name, error=str(err), total_s=time.time() - t0)) + "\n"); out.flush()
    c.close()
