import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
m, n = int(sys.argv[1]), int(sys.argv[2])
pa = synthetic_lp(m, n, 20, "tu")
c = capi.Carry(pa, heap_words=1 << 26, lib_path=sys.argv[3])
c.solve_relaxation()
c.set_option("trace", 1)
c.solve_relaxation()
tr = c.debug_trace(4096).astype(np.int64).reshape(-1)
log = c.log()
it8 = tr[160 * 16:160 * 16 + 8 * len(log)].reshape(-1, 8); it = it8[:, :4]
print('CTA0: reduce1 done, gen done (from t1); reduce2 done, pivot row done (from t2): mean us', it8[:, 4:].mean(0) / 1e3)
np.set_printoptions(linewidth=220)
print("pivots", len(log), "mean us", it.mean(0) / 1e3)
for q, name in enumerate(["select", "column", "pivot_row", "update"]):
    v = it[:, q] / 1e3
    print(name, "percentiles 5/25/50/75/95/99/max:", np.percentile(v, [5, 25, 50, 75, 95, 99, 100]).round(1))
T = np.array([x[4] for x in log]); S = np.array([x[5] for x in log]); tag = np.array([x[0] for x in log])
print("phase1 pivots", int((tag == '1').sum()), "T mean", T.mean(), "S mean", S.mean(), "T max", T.max(), "S max", S.max())
for lo, hi in [(0, 2), (2, 5), (5, 20), (20, 100), (100, 100000)]:
    sel = (T >= lo) & (T < hi)
    if sel.any(): print(f"T in [{lo},{hi}): n={int(sel.sum())} column {it[sel,1].mean()/1e3:.1f} update {it[sel,3].mean()/1e3:.1f}")
for lo, hi in [(0, 2), (2, 5), (5, 20), (20, 100), (100, 100000)]:
    sel = (S >= lo) & (S < hi)
    if sel.any(): print(f"S in [{lo},{hi}): n={int(sel.sum())} pivot_row {it[sel,2].mean()/1e3:.1f} update {it[sel,3].mean()/1e3:.1f}")
ph1 = tag == '1'
print("phase 1 means", it[ph1].mean(0) / 1e3, "phase 2 means", it[~ph1].mean(0) / 1e3)

mx = tr[160 * 16 + 7000 * 8:160 * 16 + 7000 * 8 + len(log)]
dur = (mx >> 8) / 1e3; who = mx & 255
print("slowest CTA gen_column duration per iteration: percentiles", np.percentile(dur, [5, 25, 50, 75, 95, 100]).round(1))
print("first 40 iterations (dur us, cta):", list(zip(dur[:40].round(1), who[:40])))
print("last 20:", list(zip(dur[-20:].round(1), who[-20:])))
rows = np.array([x[2] for x in log]); cols = np.array([x[1] for x in log])
print("row blocks of pivots (r//32 % 148?) first 20:", (rows[:20] // 32))
