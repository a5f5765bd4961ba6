import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
m, n, lib = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
pa = synthetic_lp(m, n, 20, "tu")
c = capi.Carry(pa, heap_words=1 << 26, lib_path=lib)
c.set_option("trace", 1)
c.create_for_partially_artificial()
c.from_artificial()
try:
    c.iterate(2, 3)
except Exception as e:
    print("iterate:", e)
tr = c.debug_trace().astype(np.int64)
d = tr[:, 1] - tr[:, 0]
print(lib, "price duration per CTA: min", d.min(), "median", int(np.median(d)), "max", d.max())
