"""torchrun worker: sharded solve with the peer transport against the oracle (debug aid of tests/test_gpu_multi.py)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist
from rust_lp_b200 import capi, multi
from rust_lp_b200.synth import synthetic_lp
from oracle import coracle
import parity
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sizes = [(60, 400, 6), (203, 2000, 20), (2000, 20000, 20)]
for transport, opts in (("peer", {}), ("peer", {"narrow": 0}), ("peer", {"narrow": 0, "k1_exit": 1, "k1_min": 1}), ("nccl", {})):
    for (m, n, k) in sizes:
        pa = synthetic_lp(m, n, k, "tu")
        o = coracle.COracle(pa); st_o = o.solve(1, 2)
        sc = multi.ShardedCarry(pa, device=local, heap_words=1 << 22, transport=transport)
        for name, value in opts.items():
            sc.c.set_option(name, value)
        t0 = time.perf_counter()
        try:
            st = sc.solve_relaxation()
        except Exception as e:
            print("rank", dist.get_rank(), transport, opts, m, "FAILED", e, "pivots", sc.stats().pivots, flush=True)
            raise
        dt = time.perf_counter() - t0
        s = sc.stats()
        print("rank", dist.get_rank(), transport, opts, m, "status", st, "pivots", s.pivots, "narrow", s.narrow_pivots, "handoffs", s.wide_handoffs,
              f"{dt*1e3:.1f} ms", flush=True)
        assert st == parity.STATUS[st_o], (transport, st, st_o)
        assert [x[:4] for x in sc.log()] == [x[:4] for x in parity.oracle_log(o)], transport
        assert sc.get_objective_function_value() == o.objective()
        assert sc.current_bfs() == o.current_bfs()
        sc.close(); del o
print("rank", dist.get_rank(), "ok", flush=True)
dist.barrier(); dist.destroy_process_group(); os._exit(0)
