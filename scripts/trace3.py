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
tr = c.debug_trace().astype(np.int64)
slow = int((tr[:, 4] - tr[:, 3]).argmax())
t0 = tr[slow, 3]
print("slow CTA", slow, "gen_in 0; points 11..15 and gen_out (ns):", [int(tr[slow, k] - t0) for k in (11, 12, 13, 14, 15, 4)])
for cta in range(0, 148, 12):
    print(cta, [int(tr[cta, k] - tr[cta, 3]) for k in (11, 12, 13, 14, 15, 4)])
