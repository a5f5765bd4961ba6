"""torchrun worker: one full-size dense pivot of the m x (m + 1) Carry block row-sharded over WORLD_SIZE GPUs
(relp_debug_k1_dense: the owner's pivot row is replicated through the NCCL stage path, every rank updates its rows
with k_update).  Prints one JSON line on rank 0; times are max over ranks.  usage: ... k1_dense_multi.py [m] [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.distributed as dist
from rust_lp_b200 import multi
from rust_lp_b200.synth import synthetic_lp
m = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
world = dist.get_world_size()
pa = synthetic_lp(m, 10 * m, 20, "tu")
sc = multi.ShardedCarry(pa, device=local, heap_words=1 << 24, transport="nccl")
sc.c.create_for_partially_artificial(pa.piv_rows, pa.piv_cols)
ms_row, ms_upd, T, S = sc.c.debug_k1_dense(seed=20260101, reps=reps)
t = torch.tensor([ms_row, ms_upd, ms_row + ms_upd], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if dist.get_rank() == 0:
    alg = 16 * (2 * (m - 1) * S + S + (m - 1))
    print(json.dumps({"n_gpus": world, "m": m, "rows_per_gpu": T, "S": S, "ms_pivot_row_and_exchange": t[0].item(),
                      "ms_update": t[1].item(), "ms_pivot": t[2].item(), "algorithmic_GBps_whole_job": alg / t[2].item() / 1e6}), flush=True)
sc.close()
dist.barrier(); dist.destroy_process_group(); os._exit(0)
