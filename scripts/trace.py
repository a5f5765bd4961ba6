import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
m, n = int(sys.argv[1]), int(sys.argv[2])
pa = synthetic_lp(m, n, 20, "tu")
c = capi.Carry(pa, heap_words=1 << 26)
c.solve_relaxation()
c.set_option("trace", 1)
c.solve_relaxation()
tr = c.debug_trace().astype(np.int64)
t0 = tr[:, 3].min()   # start of gen_column in the last complete iteration
np.set_printoptions(linewidth=250)
names = ["price_in", "price_out", "sync1_out", "gen_in", "gen_out", "sync2_out", "piv_in", "piv_out", "sync3_out", "upd_out", "sync4_out"]
print("per-point (ns, relative to earliest gen_in): min / median / max over CTAs, argmax CTA")
for k in range(3, 11):
    v = tr[:, k] - t0
    print(f"{names[k]:10s} min {v.min():7d} med {int(np.median(v)):7d} max {v.max():7d} argmax {int(v.argmax())}")
d = tr[:, 4] - tr[:, 3]
print("gen_column duration per CTA (first 48):", d[:48])
d = tr[:, 7] - tr[:, 6]
print("pivot_row duration per CTA (first 48):", d[:48])
d = tr[:, 1] - tr[:, 0]
print("price duration per CTA (first 64):", d[:64])
print("select duration per CTA (first 32):", (tr[:, 1] - tr[:, 0])[:32])
print("reduce1 (2->3) per CTA (first 32):", (tr[:, 3] - tr[:, 2])[:32])
print("sync after gen (4->5) (first 32):", (tr[:, 5] - tr[:, 4])[:32])
print("reduce2 (5->6) per CTA (first 32):", (tr[:, 6] - tr[:, 5])[:32])
print("sync after pivot row (7->8) (first 32):", (tr[:, 8] - tr[:, 7])[:32])
print("update (8->9) per CTA (first 32):", (tr[:, 9] - tr[:, 8])[:32], "last 8:", (tr[:, 9] - tr[:, 8])[-8:])
print("sync after update (9->10) (first 32):", (tr[:, 10] - tr[:, 9])[:32])
