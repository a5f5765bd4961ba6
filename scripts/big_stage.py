"""Per-stage device time of the persistent loop on one of the escalating problems (tests/golden/mps_prefix).
usage: python scripts/big_stage.py name seconds chunk"""
import json, os, sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from rust_lp_b200 import capi
from test_mps_prefix import load
name, limit, chunk = sys.argv[1], float(sys.argv[2]), int(sys.argv[3])
pa = load(name)
c = capi.Carry(pa, heap_words=1 << 28)
t0 = time.time(); st = 0; first = True; prev = [0] * 8; pp = 0; pe = 0
while st == 0 and time.time() - t0 < limit:
    st = c.solve_relaxation(1, 2, max_pivots=chunk) if first else c.solve_continue(chunk)
    first = False
    s = c.stats()
    ns = list(s.stage_ns); d = [(a - b) / 1e6 for a, b in zip(ns, prev)]; prev = ns
    dp = s.pivots - pp; pp = s.pivots; de = s.escalated - pe; pe = s.escalated
    print(json.dumps(dict(pivots=s.pivots, phase=s.phase, dpiv=dp, escalated=de, narrow=s.narrow_pivots, comp=s.compactions,
                          handoffs=s.wide_handoffs, loop_ms=round(s.loop_ms),
                          ms=dict(zip(("select", "column_ratio", "pivot_row", "update", "big+house", "narrow"), [round(x, 1) for x in d[:6]])))), flush=True)
