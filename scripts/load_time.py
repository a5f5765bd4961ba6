"""Where the end-to-end overhead of one solve goes (GPU): relp_load_matrix (host validation overlapped with the H2D copies,
SELL / CSR build), the Carry constructor, the read-back.  usage: python scripts/load_time.py"""
import sys, time
sys.path.insert(0, '.')
import torch
import bench as B
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp

pa = B.pin(synthetic_lp(20000, 200000, 20, "tu"))
c = capi.Carry(pa, heap_words=1 << 24)
for rep in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter(); c.load_matrix(pa); torch.cuda.synchronize(); t1 = time.perf_counter()
    c.create_for_partially_artificial(); torch.cuda.synchronize(); t2 = time.perf_counter()
    st = c.solve_relaxation(); torch.cuda.synchronize(); t3 = time.perf_counter()
    c.get_objective_function_value(); c.current_bfs_arrays(); torch.cuda.synchronize(); t4 = time.perf_counter()
    print(f"load_matrix {1e3*(t1-t0):.2f} ms | constructor alone {1e3*(t2-t1):.2f} ms | solve_relaxation (constructor + loop) {1e3*(t3-t2):.2f} ms | read-back {1e3*(t4-t3):.2f} ms", flush=True)
