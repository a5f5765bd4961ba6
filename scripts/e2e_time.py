import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import torch
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
import bench
pa = bench.pin(synthetic_lp(20000, 200000, 20, "tu"))
c = capi.Carry(pa, heap_words=1 << 24)
for rep in range(4):
    t = [time.perf_counter()]
    c.load_matrix(pa); torch.cuda.synchronize(); t.append(time.perf_counter())
    st = c.solve_relaxation(); torch.cuda.synchronize(); t.append(time.perf_counter())
    obj = c.get_objective_function_value(); t.append(time.perf_counter())
    bfs = c.current_bfs(); t.append(time.perf_counter())
    s = c.stats()
    print("load %.1f ms  solve %.1f ms (loop %.1f)  objective %.1f ms  bfs %.1f ms" % tuple([1e3 * (t[1] - t[0]), 1e3 * (t[2] - t[1]), s.loop_ms, 1e3 * (t[3] - t[2]), 1e3 * (t[4] - t[3])]), flush=True)
