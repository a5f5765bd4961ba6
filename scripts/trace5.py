import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
m, n = int(sys.argv[1]), int(sys.argv[2])
pa = synthetic_lp(m, n, 20, "tu")
c = capi.Carry(pa, heap_words=1 << 26, lib_path=sys.argv[3])
if len(sys.argv) > 4: c.set_option("coop_blocks", int(sys.argv[4]))
c.solve_relaxation()
c.set_option("trace", 1)
c.solve_relaxation()
tr = c.debug_trace(4096).astype(np.int64)
piv = len(c.log())
s = c.stats(); print("stage_us/iter", [round(x / 1e3 / max(s.pivots, 1), 2) for x in s.stage_ns[:6]])
names = ["-", "summary", "gen rows", "argmin", "pivot row", "commit", "items(thread0)", "sync/dense rc", "big"]
print("narrow steps, us per pivot (thread 0 of CTA 0):")
for k, nm in enumerate(names): print(f"  {nm:18s} {tr[150, k] / piv / 1e3:6.2f}")
print("  total", tr[150, :9].sum() / piv / 1e3)

print("item classes: slowest thread per pivot (us): carry %.2f cost %.2f rc %.2f bookkeeping %.2f; items/pivot %.1f rc items/pivot %.1f" % (
    tr[151, 0] / piv / 1e3, tr[151, 1] / piv / 1e3, tr[151, 2] / piv / 1e3, tr[151, 3] / piv / 1e3, tr[151, 4] / piv, tr[151, 5] / piv))
