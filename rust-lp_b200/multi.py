"""One process per GPU: row-sharded Carry.

torch.distributed is the plumbing only (rendezvous, exchanging the NCCL unique id and the CUDA IPC
handles of the exchange buffers, gathering the per-rank pieces of the answer).  Transports:
  "peer"       the persistent loop kernel on every rank, per-pivot exchanges by stores into peer memory
               over NVLink (the product path);
  "nccl"       one launch per stage with NCCL collectives in between, issued by librelp_b200.so on its
               own stream (the correctness baseline the peer path is checked against);
  "callbacks"  the same stage-per-launch driver with host-staged collectives over a gloo group (CPU tests).
"""
import ctypes as C
import glob
import json
import os
import time

import numpy as np

from . import capi


def find_nccl():
    """The libnccl torch itself loaded (so that there is one NCCL in the process)."""
    try:
        import torch
        base = os.path.dirname(os.path.dirname(torch.__file__))
        for pat in ("nvidia/nccl/lib/libnccl.so*", "torch/lib/libnccl.so*"):
            hits = sorted(glob.glob(os.path.join(base, pat)))
            if hits:
                return hits[0]
    except Exception:
        pass
    return None


def gloo_callbacks(group=None):
    """Host-staged all-gather / all-reduce over torch.distributed (any backend with CPU tensors)."""
    import torch
    import torch.distributed as dist

    def allgather(_user, send, recv, nbytes):
        world = dist.get_world_size(group)
        src = torch.from_numpy(np.ctypeslib.as_array(C.cast(send, C.POINTER(C.c_uint8)), shape=(nbytes,)).copy())
        outs = [torch.empty(nbytes, dtype=torch.uint8) for _ in range(world)]
        dist.all_gather(outs, src, group=group)
        dst = np.ctypeslib.as_array(C.cast(recv, C.POINTER(C.c_uint8)), shape=(nbytes * world,))
        dst[:] = torch.cat(outs).numpy()
        return 0

    def allreduce(_user, buf, count):
        arr = np.ctypeslib.as_array(buf, shape=(count,))
        t = torch.from_numpy(arr.copy())
        dist.all_reduce(t, group=group)
        arr[:] = t.numpy()
        return 0

    return allgather, allreduce


class ShardedCarry:
    """The same interface as capi.Carry for the calls a sharded solve needs; every rank calls every
    method (they are collective)."""

    def __init__(self, provider, device=None, heap_words=0, lib_path=None, transport="peer"):
        import torch.distributed as dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        dev = device if device is not None else int(os.environ.get("LOCAL_RANK", self.rank))
        self.c = capi.Carry(provider, device=dev, rank=self.rank, world=self.world, heap_words=heap_words,
                            lib_path=lib_path)
        self.transport = transport if self.world > 1 else "single"
        if self.world > 1:
            if transport in ("nccl", "peer"):
                lib = find_nccl()
                box = [self.c.comm_unique_id(lib) if self.rank == 0 else None]
                dist.broadcast_object_list(box, src=0)
                self.c.attach_comm(box[0], lib)
            else:
                self.c.attach_comm_callbacks(*gloo_callbacks())
            if transport == "peer":
                # exchange buffers mapped into every peer (CUDA IPC): the persistent loop kernel runs on every
                # rank and talks to its peers by stores over NVLink; NCCL is only used at the phase hand-over
                handles = [None] * self.world
                dist.all_gather_object(handles, self.c.peer_export(provider.m))
                self.c.peer_attach(handles)
            else:
                self.c.set_option("persistent", 0)

    def solve_relaxation(self, *a, **k):
        return self.c.solve_relaxation(*a, **k)

    def log(self):
        return self.c.log()

    def stats(self):
        return self.c.stats()

    def get_objective_function_value(self):
        return self.c.get_objective_function_value()

    def current_bfs(self):
        """all ranks get the full basic solution"""
        import torch.distributed as dist
        parts = [None] * self.world
        dist.all_gather_object(parts, self.c.current_bfs())
        out = [x for p in parts for x in p]
        out.sort(key=lambda x: x[0])
        return out

    def b(self):
        import torch.distributed as dist
        parts = [None] * self.world
        dist.all_gather_object(parts, self.c.b())
        return [x for p in parts for x in p]

    def close(self):
        self.c.close()


def bench(args, B):
    """`bench.py --gpus N` under torchrun: ONE LP solved cooperatively by N GPUs (strong scaling).  `B` is the bench
    module (workload, config, roofline, parity and clock helpers shared with the N = 1 line)."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pa, gold = B.workload(args)
    sc = ShardedCarry(pa, device=local, heap_words=args.heap_words)
    stream = torch.cuda.ExternalStream(sc.c.stream())

    def solve():
        st = sc.solve_relaxation()
        assert st == capi.OPTIMAL, st
        return sc.stats()

    for _ in range(max(args.warmup, 3)):
        s = solve()
    clocks = B.ClockSampler(local)
    if rank == 0:
        clocks.start()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = s.kernel_launches
    e0.record(stream)
    pivots, loop_ms = 0, 0.0
    for _ in range(args.steps):
        s = solve()
        pivots += s.pivots; loop_ms += s.loop_ms
    e1.record(stream)
    torch.cuda.synchronize(); dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    launches = s.kernel_launches - launches0
    obj = sc.get_objective_function_value()
    # parity on EVERY rank (the pivot log and the objective are replicated): the timed solve against the committed golden
    par = B.parity_block(sc.log(), obj, gold)
    flags = [None] * world
    dist.all_gather_object(flags, (par["log_sha256"], par["objective"], par["identical_to_golden"]))
    # e2e: host buffers in, answer out, wall clock, max over ranks
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sc.c.load_matrix(pa)
        sc.solve_relaxation()
        sc.get_objective_function_value()
        sc.c.current_bfs_arrays()              # this rank's rows of the basic solution (basis and objective are replicated)
    torch.cuda.synchronize(); dist.barrier()
    e2e = torch.tensor([time.perf_counter() - t0], device="cuda")
    dist.all_reduce(e2e, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e.item())
    if rank == 0:
        clk = clocks.stop()
        h2d = 8 * (pa.n + 1) + 20 * pa.nnz + 16 * pa.n + 16 * pa.m
        par["ranks_identical"] = bool(all(f[:2] == flags[0][:2] for f in flags))
        par["ranks_identical_to_golden"] = [f[2] for f in flags]
        roof = B.roofline_block(pa, s, loop_ms / args.steps, world)
        roof["what"] += "  (rank 0's kernel; every rank runs the same loop on its row shard)"
        cpu = {"kind": "port", "cores": 1, "unit": unit_of(B), "value": None,
               "sample": "not re-timed at N > 1 (the CPU arm is timed at N = 1 and by --impl reference); cached full solve below"}
        if gold and gold.get("oracle_seconds_to_optimal"):
            cpu["time_to_optimal_s"] = gold["oracle_seconds_to_optimal"]
            cpu["value"] = gold["pivots"] / gold["oracle_seconds_to_optimal"]
            cpu["sample"] = "whole solve, " + gold["source"] + " -- run once in the build container, not on this box"
        line = {"metric": B.METRIC, "value": pivots / (ms * 1e-3), "unit": B.UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None,
                "dtype": "exact rational, int64/int64 tier (multi-limb escalation armed)", "data": "synthetic" if args.workload == "synth5a" else "netlib/miplib fixture",
                "config": B.config(args, world, pa), "time_to_optimal_ms": ms / args.steps,
                "pivots_per_solve": int(s.pivots), "objective": str(obj),
                "e2e": {"value": pivots / e2e_s, "unit": B.UNIT, "h2d_bytes_per_step": h2d * world,
                        "d2h_bytes_per_step": 16 + 20 * pa.m, "ms_per_step": 1e3 * e2e_s / args.steps},
                "gpu_launches": int(launches), "clocks": clk,
                "roofline": roof, "parity": par, "cpu_baseline": cpu,
                "transport": sc.transport,
                "note": "ONE LP solved cooperatively (strong scaling): Carry rows sharded, cost row / reduced-cost row / basis "
                        "replicated; the persistent loop kernel runs on every rank and exchanges ratio-test candidates and the "
                        "pivot row by stores into peer memory over NVLink (one exchange per narrow pivot, no NCCL call and no "
                        "launch per pivot).  This LP is latency-bound (about 20 Carry rows x 3 columns change per pivot, "
                        "DESIGN.md section 4): sharding removes almost no work and adds one NVLink round per pivot, so "
                        "pivots/s stays just below the single-GPU figure instead of growing with N"}
        print(json.dumps(line), flush=True)
    sc.close()
    dist.destroy_process_group()


def unit_of(B):
    return B.UNIT
