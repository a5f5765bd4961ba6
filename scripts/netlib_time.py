"""Times the golden MPS problems (tests/golden/mps) on the GPU next to the oracle on one host core.
Output: one JSON line per problem -> gpurun_out/netlib_times.jsonl.  Not a bench.py line: these are the
BASELINE.json configs[0..3] parity cases, timed for context (tiny problems are launch/latency bound)."""
import json, os, sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from rust_lp_b200 import capi
from oracle import coracle
from test_mps_golden import load_fixture, EXPECTED

out = open(os.path.join("gpurun_out", "netlib_times.jsonl"), "w")
only = sys.argv[1:]
for e in EXPECTED:
    name = e["name"]
    if only and name not in only: continue
    if e["status"] == "artificial_basic": continue
    pa, _ = load_fixture(name)
    best = None
    for rep in range(3):
        c = capi.Carry(pa, heap_words=1 << 24)
        t0 = time.time(); st = c.solve_relaxation(1, 2); wall = time.time() - t0
        s = c.stats()
        rec = dict(name=name, m=pa.m, n=pa.n, status=st, pivots=s.pivots, loop_ms=s.loop_ms, wall_ms=1e3 * wall,
                   escalated=s.escalated, launches=s.kernel_launches, compactions=s.compactions)
        c.close()
        if best is None or rec["loop_ms"] < best["loop_ms"]: best = rec
    t0 = time.time(); o = coracle.COracle(pa); o.solve(1, 2); best["oracle_ms"] = 1e3 * o.loop_seconds if hasattr(o, "loop_seconds") else 1e3 * (time.time() - t0)
    best["oracle_wall_ms"] = 1e3 * (time.time() - t0); best["max_bits"] = e["max_bits"]
    best["gpu_pivots_per_s"] = best["pivots"] / (best["loop_ms"] / 1e3) if best["loop_ms"] else None
    best["cpu_pivots_per_s"] = best["pivots"] / (best["oracle_wall_ms"] / 1e3)
    print(json.dumps(best), flush=True); out.write(json.dumps(best) + "\n"); out.flush()
