"""One exact solve of a golden MPS problem (profiling aid).
usage: python scripts/one_solve.py netlib_SHARE1B [max_pivots (-1)] [persistent (1; 0 = one launch per stage)] [heap_words]"""
import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from rust_lp_b200 import capi
import os
from test_mps_golden import load_fixture, GOLD
if os.path.exists(os.path.join(GOLD, "mps", sys.argv[1] + ".npz")):
    pa, _ = load_fixture(sys.argv[1])
else:                                   # BASELINE configs[2..3]: tests/golden/mps_prefix
    from test_mps_prefix import load
    pa = load(sys.argv[1])
budget = int(sys.argv[2]) if len(sys.argv) > 2 else -1
c = capi.Carry(pa, heap_words=int(sys.argv[4]) if len(sys.argv) > 4 else 1 << 26)
if len(sys.argv) > 3:
    c.set_option("persistent", int(sys.argv[3]))
t0 = time.time(); st = c.solve_relaxation(1, 2, budget); s = c.stats()
print(sys.argv[1], "status", st, "pivots", s.pivots, "loop_ms", round(s.loop_ms, 1), "escalated", s.escalated, "narrow", s.narrow_pivots,
      "stage_ms", [round(x / 1e6, 1) for x in s.stage_ns[:6]])
