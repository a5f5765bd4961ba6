#!/usr/bin/env python
"""Headline benchmark: exact simplex pivots/sec and time-to-optimal (BASELINE.json's metric).

Default workload = BASELINE.json configs[4]: the synthetic sparse LP 5a (m = 20 000 rows, n = 200 000 structural
columns, 20 nonzeros per column, totally unimodular, `A x <= b`), solved exactly (rational arithmetic) from the
artificial tableau to the optimum.  `--workload NAME` runs one of the committed MPS fixtures instead (BASELINE configs
[0..3]: netlib_ADLITTLE, netlib_SC205, netlib_ISRAEL, miplib_50v-10, ... under tests/golden/mps*/).

A "step" is one complete exact solve through the C ABI (include/relp_b200.h): relp_solve_relaxation = Carry
initialisation + phase 1 + hand-over + phase 2 on the GPU.

  value     pivots/sec, device time (CUDA events on the library's stream), LP already resident in HBM
  e2e       the same metric through the public API with HOST buffers: relp_load_matrix (pinned host -> device), solve,
            read-back of objective + basic solution, wall clock around the calls
  roofline  the persistent loop kernel.  achieved = bytes the kernel as implemented touches / its time (the loop is
            bound by serial latency, so this is a small fraction of the HBM peak and says so); the bytes the REFERENCE
            algorithm would move per SURVEY.md 8(d) are reported separately as reference_equivalent_GBps
  parity    sha256 of the pivot log of the timed solve + exact objective, compared with the committed golden of the
            oracle's full solve (tests/golden/synth5a_expected.json, mps_expected.json) -- on every rank at N > 1
  cpu_baseline  the oracle (CPU restatement of the reference path, 1 core) on a bounded sample, plus its one-off
            full-solve time-to-optimal from the golden

`--impl reference` times the oracle alone (the reference's Rust toolchain does not exist here).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "exact_simplex_pivots_per_sec"
UNIT = "pivots/s"
W_BYTES = 16            # bytes per rational in the int64/int64 tier
GOLD = os.path.join(ROOT, "tests", "golden")


# ---------------------------------------------------------------------------------------------- workloads
def load_npz(path):
    import numpy as np
    from rust_lp_b200.synth import ProviderArrays
    z = np.load(path)
    pa = ProviderArrays(int(z["m"]), int(z["n"]), z["col_ptr"].tolist(), z["row_idx"].tolist(), z["val_num"].tolist(),
                        z["val_den"].tolist(), z["cost_num"].tolist(), z["cost_den"].tolist(), z["rhs_num"].tolist(),
                        z["rhs_den"].tolist(), z["piv_rows"].tolist(), z["piv_cols"].tolist(), int(z["n_normal"]))
    if "rhs_exact" in z.files:          # right-hand sides beyond int64/int64 (relp_set_rhs)
        from fractions import Fraction
        pa.rhs_exact = {int(i): Fraction(int(v[0]), int(v[1])) for i, v in json.loads(bytes(z["rhs_exact"]).decode()).items()}
    return pa


def workload(args):
    """(ProviderArrays, golden dict or None).  golden = {"pivots", "log_sha256", "objective", "source", ...}"""
    if args.workload == "synth5a":
        from rust_lp_b200.synth import synthetic_lp
        pa = synthetic_lp(args.m, args.n, args.nnz_col, args.variant)
        gold = None
        f = os.path.join(GOLD, "synth5a_expected.json")
        if (args.m, args.n, args.nnz_col, args.variant) == (20000, 200000, 20, "tu") and os.path.exists(f):
            g = json.load(open(f))
            last = g["checkpoints"][-1]
            gold = {"pivots": g["pivots_to_optimal"], "log_sha256": last["log_sha256"], "objective": last["objective"],
                    "source": "tests/golden/synth5a_expected.json (oracle full solve, tests/golden/make_synth_golden.py)",
                    "oracle_seconds_to_optimal": g["oracle_seconds_to_optimal"]}
        return pa, gold
    for sub, exp in (("mps", "mps_expected.json"), ("mps_prefix", None)):
        f = os.path.join(GOLD, sub, args.workload + ".npz")
        if os.path.exists(f):
            gold = None
            if exp:
                e = {x["name"]: x for x in json.load(open(os.path.join(GOLD, exp)))}.get(args.workload)
                if e and e["status"] == "optimal":
                    gold = {"pivots": e["pivots"], "log_sha256": e["log_sha256"], "objective_of_standard_form": None,
                            "source": "tests/golden/mps_expected.json (oracle full solve; objective pinned against the reference's "
                                      "test value " + str(e["source"]) + "; no presolve)", "max_bits": e["max_bits"]}
            return load_npz(f), gold
    if args.workload.startswith("presolved:"):       # the reference's own pipeline (presolve first): tests/golden/mps_presolved/
        name = args.workload.split(":", 1)[1]
        f = os.path.join(GOLD, "mps_presolved", name + ".npz")
        if os.path.exists(f):
            e = {x["name"]: x for x in json.load(open(os.path.join(GOLD, "mps_presolved_expected.json")))}.get(name)
            gold = None
            if e and e.get("status") == "optimal":
                gold = {"pivots": e["pivots"], "log_sha256": e["log_sha256"], "objective_of_standard_form": None,
                        "source": "tests/golden/mps_presolved_expected.json (oracle full solve of the presolved standard form; objective pinned "
                                  "against the reference's test value " + str(e["source"]) + ")", "max_bits": e["max_bits"]}
            return load_npz(f), gold
    raise SystemExit(f"unknown workload {args.workload!r}: synth5a, a fixture name under tests/golden/mps*/, or presolved:<fixture>")


def config(args, n_gpus, pa):
    if args.workload == "synth5a":
        what = (f"synthetic sparse LP 5a ({args.variant}): m={args.m} n={args.n} nnz/col={args.nnz_col} (+{args.m} slack columns), "
                "exact solve to optimality, FirstProfitableWithMemory/SteepestDescent")
    else:
        what = (f"{args.workload} (committed MPS fixture, standard form {'through the reference\'s presolve' if args.workload.startswith('presolved:') else 'without presolve'}): m={pa.m} n={pa.n} nnz={pa.nnz}, exact solve to "
                "optimality, FirstProfitableWithMemory/SteepestDescent")
    return {"workload": what, "m": pa.m, "n": pa.n, "nnz": pa.nnz, "seed": 20260101 if args.workload == "synth5a" else None,
            "parallelism": "1 GPU" if n_gpus == 1 else f"ONE LP, Carry rows sharded over {n_gpus} GPUs (strong scaling)",
            "l2": "every step re-initialises the Carry planes (2 x 8 x (m+1) x m bytes = "
                  f"{2 * 8 * (pa.m + 1) * pa.m / 1e9:.2f} GB written)" +
                  (", which flushes the 126 MB L2 between timed steps" if 16 * (pa.m + 1) * pa.m > 126e6 else
                   "; the problem is smaller than L2: the loop is latency-bound, not bandwidth-bound, and nothing is re-read across steps")}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.rows[0][1]) if self.rows and self.rows[0][1].isdigit() else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def pin(pa):
    """Back every input array with pinned host memory (the e2e copy source)."""
    import torch
    for name in ("col_ptr", "row_idx", "val_num", "val_den", "cost_num", "cost_den", "rhs_num", "rhs_den",
                 "piv_rows", "piv_cols"):
        a = getattr(pa, name)
        t = torch.from_numpy(a).pin_memory()
        setattr(pa, name, t.numpy())
        pa.__dict__.setdefault("_pinned", []).append(t)
    return pa


# ---------------------------------------------------------------------------------------------- bytes
def algorithmic_bytes(pa, stats):
    """SURVEY.md 8(d) figures of the REFERENCE algorithm, summed over the pivots of one solve (w = 16 bytes per rational):
    K3 (w+4) nnz(A) + w n + 4 n ; K2 w (m nnz_j + nnz_j + m) ; K4 2 w m + 4 m ; K1 w (2 T S + S + T)."""
    w, m, n, nnz = W_BYTES, pa.m, pa.n, pa.nnz
    pivots = stats.pivots
    k3 = (w + 4) * nnz + w * n + 4 * n
    nnz_j = nnz / n
    k2 = w * (m * nnz_j + nnz_j + m)
    k4 = 2 * w * m + 4 * m
    k1 = w * (2 * stats.sum_TS + stats.sum_S + stats.sum_T)
    # one extra pricing pass per loop detects termination
    return (k3 + k2 + k4) * pivots + 2 * k3 + k1, {"K3_price": k3, "K2_column": k2, "K4_ratio": k4,
                                                   "K1_update_mean": k1 / max(pivots, 1)}


def implemented_bytes(pa, stats):
    """What the kernel as implemented has to touch: per pivot the stored reduced-cost row (w per
    tableau column), a_j with one bitmap word per entry and row block summary, the hit elements of B^-1, the pivot
    row's bitmap row, the T x S update (read + write), and per updated reduced cost one CSR entry, the value and the
    read-modify-write; plus two full pricing passes per solve."""
    w, m, n, nnz = W_BYTES, pa.m, pa.n, pa.nnz
    pivots = stats.pivots
    nnz_j = nnz / n
    row_len = nnz / m
    select = w * n
    column = nnz_j * (4 + 16 + 4 + 4 * ((m // 32 + 31) // 32)) + w * 2 * (stats.sum_T / max(pivots, 1))
    pivot_row = 4 * ((m + 32) // 32) + 2 * w * (stats.sum_S / max(pivots, 1))
    update = 2 * w * (stats.sum_TS / max(pivots, 1))
    rc = (stats.sum_S / max(pivots, 1)) * row_len * (16 + 16 + 2 * w + 4)
    k3 = (w + 4) * nnz + w * n + 4 * n
    return (select + column + pivot_row + update + rc) * pivots + 2 * k3


def roofline_block(pa, s, loop_ms_step, n_gpus=1):
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("hbm_gbs", 6650.0)
    abytes, per_kernel = algorithmic_bytes(pa, s)
    ibytes = implemented_bytes(pa, s)
    achieved = ibytes / (loop_ms_step * 1e-3) / 1e9
    traffic, tsrc = None, None
    for name in ("traffic_r02.json", "traffic_r01.json"):
        tfile = os.path.join(ROOT, "profiles", name)
        if os.path.exists(tfile) and pa.m == 20000:
            traffic = json.load(open(tfile)).get("dram_bytes_per_launch")
            tsrc = f"profiles/{name} (one ncu --set full capture of this kernel on this workload by the builder; NOT measured in this run)"
            break
    return {"bound": "latency", "bound_resource": "hbm",
            "kernel": "k_simplex_loop (persistent cooperative loop: select + column + ratio + pivot row + Carry/reduced-cost update)",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650",
            "what": "achieved = implemented_bytes_per_launch / kernel_ms_per_launch: the bytes this kernel has to touch (almost all of them the "
                    "selection scan of the stored reduced-cost row, served by L2).  The loop is bound by SERIAL LATENCY (a chain of dependent L2 "
                    "round trips and two grid barriers per pivot; ncu: issue slots ~7 % active), so frac is small by construction: a pivot "
                    "changes ~20 rows x 3 columns.  The HBM-bound form of the Carry update is measured separately (roofline_k1_dense).",
            "implemented_bytes_per_launch": ibytes, "kernel_ms_per_launch": loop_ms_step,
            "traffic": traffic, "traffic_source": tsrc,
            "reference_equivalent_GBps": abytes / (loop_ms_step * 1e-3) / 1e9,
            "reference_equivalent_frac": abytes / (loop_ms_step * 1e-3) / 1e9 / peak,
            "reference_equivalent_note": "bytes the REFERENCE algorithm moves per pivot (SURVEY 8(d): full pricing pass, column generation, ratio "
                                         "test, T x S update) / measured time -- a work rate, not a bandwidth: the kernel does not move them",
            "algorithmic_bytes_per_launch": abytes, "algorithmic_bytes_per_pivot": per_kernel,
            "narrow_pivots": int(s.narrow_pivots), "escalated_elements": int(s.escalated), "max_limbs": int(s.max_limbs)}


# ---------------------------------------------------------------------------------------------- parity
def log_sha256(log):
    h = hashlib.sha256()
    for e in log:
        h.update(("%d,%d,%d;" % (ord(e[0]) if isinstance(e[0], str) else int(e[0]), e[1], e[2])).encode())
    return h.hexdigest()


def parity_block(log, objective, gold):
    """sha256 of the pivot log (phase tag, entering column, pivot row per pivot) + exact objective of the standard form,
    compared with the committed golden of the oracle's full solve when there is one."""
    p = {"pivots": len(log), "log_sha256": log_sha256(log), "objective": str(objective)}
    if gold:
        same = p["log_sha256"] == gold["log_sha256"] and p["pivots"] == gold["pivots"]
        if gold.get("objective") is not None:
            same = same and p["objective"] == gold["objective"]
        p.update({"identical_to_golden": bool(same), "golden": gold["source"]})
    else:
        p.update({"identical_to_golden": None, "golden": "none committed for this workload"})
    return p


def cpu_sample(pa, pivots):
    """Oracle (single thread) on the first `pivots` pivots of the same LP."""
    from oracle import coracle
    o = coracle.COracle(pa)
    o.begin(1, 2)
    t0 = time.perf_counter()
    o.advance(pivots)
    dt = time.perf_counter() - t0
    done = o.nr_pivots
    log = [(chr(int(t)), int(j), int(r), int(l), int(T), int(S)) for (t, j, r, l, T, S) in o.log()]
    return done / dt if dt > 0 else 0.0, done, dt, log


def cpu_pivot_budget(args):
    """the committed MPS fixtures finish on the CPU in seconds: time the whole solve there; the synthetic LP takes ten minutes: a prefix"""
    return args.cpu_pivots if args.workload == "synth5a" or args.cpu_pivots != 150 else 10 ** 9


def cpu_baseline_block(pa, args, gold, gpu_log=None):
    cpu_v, cpu_done, cpu_dt, cpu_log = cpu_sample(pa, cpu_pivot_budget(args))
    whole = gpu_log is not None and cpu_done == len(gpu_log)
    b = {"value": cpu_v, "unit": UNIT, "cores": 1, "kind": "port",
         "sample": (f"the WHOLE solve, {cpu_done} pivots in {cpu_dt:.2f} s" if whole else f"first {cpu_done} pivots of the same LP, {cpu_dt:.1f} s") +
                   " (oracle/simplex.cpp, single thread like the reference" +
                   (")" if whole else "; the GPU arm times the WHOLE solve, so the ratio compares pivots/s over different pivot ranges of the same LP)")}
    if whole:
        b["time_to_optimal_s"] = cpu_dt
    if gold and gold.get("oracle_seconds_to_optimal"):
        b["time_to_optimal_s"] = gold["oracle_seconds_to_optimal"]
        b["full_solve_pivots_per_s"] = gold["pivots"] / gold["oracle_seconds_to_optimal"]
        b["time_to_optimal_source"] = gold["source"] + " -- run once in the build container, not on this box"
    if gpu_log is not None:
        b["prefix_identical_to_gpu_log"] = bool(gpu_log[:len(cpu_log)] == cpu_log)
    return b


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pa, gold = workload(args)
    sample = cpu_pivot_budget(args)
    for _ in range(args.warmup):
        cpu_sample(pa, max(2, min(sample, 150) // 8))
    tot_p, tot_t = 0, 0.0
    for _ in range(args.steps):
        _, done, dt, _ = cpu_sample(pa, sample)
        tot_p += done; tot_t += dt
    v = tot_p / tot_t
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "exact rational (arbitrary-precision integers)", "data": "synthetic" if args.workload == "synth5a" else "netlib/miplib fixture",
            "config": config(args, args.gpus, pa),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
                             "sample": (f"first {sample} pivots of the same LP per step" if sample < 10 ** 9 else f"the whole solve ({tot_p // args.steps} pivots) per step") +
                                       " (oracle/simplex.cpp, single thread: the reference is single-threaded Rust and cannot be built here)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if gold and gold.get("oracle_seconds_to_optimal"):
        line["cpu_baseline"]["time_to_optimal_s"] = gold["oracle_seconds_to_optimal"]
        line["cpu_baseline"]["full_solve_pivots_per_s"] = gold["pivots"] / gold["oracle_seconds_to_optimal"]
    print(json.dumps(line), flush=True)


def run_single(args):
    import torch
    from rust_lp_b200 import capi
    import __graft_entry__ as entry
    entry.build_library()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU path); use --impl reference for the CPU arm")
    torch.cuda.set_device(0)
    pa, gold = workload(args)
    pa = pin(pa)
    c = capi.Carry(pa, heap_words=args.heap_words)
    stream = torch.cuda.ExternalStream(c.stream())

    def solve():
        st = c.solve_relaxation()
        assert st == capi.OPTIMAL, st
        return c.stats()

    for _ in range(max(args.warmup, 3)):
        s = solve()
    # ---- value: LP resident in HBM, device time -------------------------------------------------
    clocks = ClockSampler()
    clocks.start()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = s.kernel_launches
    e0.record(stream)
    pivots, loop_ms, stage_ns = 0, 0.0, [0] * 8
    for _ in range(args.steps):
        s = solve()
        pivots += s.pivots; loop_ms += s.loop_ms
        stage_ns = [a + b for a, b in zip(stage_ns, s.stage_ns)]
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = s.kernel_launches - launches0
    clk = clocks.stop()
    objective = c.get_objective_function_value()
    gpu_log = c.log()
    # ---- e2e: host buffers in, result out, wall clock ----------------------------------------------
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(args.steps):
        c.load_matrix(pa)
        st = c.solve_relaxation()
        obj = c.get_objective_function_value()
        bfs = c.current_bfs_arrays()          # basic solution as (column, num, den) arrays: the D2H read of the result
        d2h += 16 + 16 * pa.m + 4 * pa.m
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert obj == objective
    cols, bn, bd, limbs = bfs                  # (outside the timed region) the arrays are the reference's current_bfs()
    assert list(zip(cols.tolist(), capi.decode_rats(bn, bd, limbs))) == c.current_bfs()
    h2d = 8 * (pa.n + 1) + 20 * pa.nnz + 16 * pa.n + 16 * pa.m
    loop_ms_step = loop_ms / args.steps
    roof = roofline_block(pa, s, loop_ms_step)
    roof["stage_us_per_pivot"] = {k: v / 1e3 / max(pivots, 1) for k, v in
                                  zip(("select", "column_ratio_wide", "pivot_row_wide", "update_wide", "housekeeping", "narrow_iteration"), stage_ns)}
    # ---- K1 alone at the dense bound: every element of the Carry block changes (micro-benchmark) -------
    k1 = None
    if not args.no_k1_dense and args.workload == "synth5a":
        peak = roof["peak"]
        c.create_for_partially_artificial(pa.piv_rows, pa.piv_cols)
        ms_row, ms_upd, T, S = c.debug_k1_dense(seed=20260101, reps=6)
        k1_bytes = W_BYTES * (2 * T * S + S + T)
        c.create_for_partially_artificial(pa.piv_rows, pa.piv_cols)
        bad, checked = c.debug_k1_verify(seed=20260101)
        k1 = {"kernel": "k_update (stage_update, staged through shared memory by bulk copies) on a dense Carry block: T = m - 1 rows x "
                        "S = m + 1 columns of small integers, d_r = 1", "T": T, "S": S, "ms_update": ms_upd, "ms_pivot_row": ms_row,
              "algorithmic_bytes": k1_bytes, "achieved": k1_bytes / (ms_upd * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
              "frac": k1_bytes / (ms_upd * 1e-3) / 1e9 / peak, "bound": "hbm",
              "moved_bytes": (8 + 8 + 8) * T * S, "moved_frac": (8 + 8 + 8) * T * S / (ms_upd * 1e-3) / 1e9 / peak,
              "verified": {"elements": checked, "mismatches": bad,
                           "how": "relp_debug_k1_verify: the same fill and pivot updated by this staged path and by the element-wise path "
                                  "(the one the oracle-parity tests cover), planes and pattern bitmap compared on the device"},
              "note": "algorithmic = w (2 T S + S + T), w = 16 (SURVEY 8(d)); the kernel leaves unchanged denominators unwritten (24 B per "
                      "element on integer data): moved_frac is the physical figure"}
        # the same update on general rationals (every element through small_fma_rational): integer-pipe work
        c.set_option("k1_bench_frac", 1)
        c.create_for_partially_artificial(pa.piv_rows, pa.piv_cols)
        _, ms_rat, T, S = c.debug_k1_dense(seed=20260101, reps=3)
        c.set_option("k1_bench_frac", 0)
        k1["general_rationals"] = {"ms_update": ms_rat, "elements_per_s": T * S / (ms_rat * 1e-3), "bound": "integer pipes",
                                   "algorithmic_GBps": W_BYTES * (2 * T * S + S + T) / (ms_rat * 1e-3) / 1e9,
                                   "note": "entries with denominators 2,3,5,7, d_i over 11,13; pipe utilisation (ncu) and the integer "
                                           "peaks measured on this GPU are in profiles/k1_dense_rational_r01.json"}
    line = {"metric": METRIC, "value": pivots / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "exact rational, int64/int64 tier (multi-limb escalation armed, 0 elements escalated)"
            if s.escalated == 0 else f"exact rational, int64/int64 + multi-limb tiers ({int(s.escalated)} element updates escalated)",
            "data": "synthetic" if args.workload == "synth5a" else "netlib/miplib fixture",
            "config": config(args, 1, pa),
            "time_to_optimal_ms": ms / args.steps, "pivots_per_solve": int(s.pivots), "objective": str(objective),
            "e2e": {"value": pivots / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h // args.steps,
                    "ms_per_step": 1e3 * e2e_s / args.steps},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": roof,
            "parity": parity_block(gpu_log, objective, gold),
            "cpu_baseline": cpu_baseline_block(pa, args, gold, gpu_log)}
    if k1:
        line["roofline_k1_dense"] = k1
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="synth5a", help="synth5a (BASELINE configs[4], default) or a fixture under tests/golden/mps*/")
    ap.add_argument("--m", type=int, default=20000)
    ap.add_argument("--n", type=int, default=200000)
    ap.add_argument("--nnz-col", type=int, default=20)
    ap.add_argument("--variant", default="tu")
    ap.add_argument("--heap-words", type=int, default=1 << 24)
    ap.add_argument("--no-k1-dense", action="store_true", help="skip the dense Carry-update micro-benchmark")
    ap.add_argument("--cpu-pivots", type=int, default=150, help="oracle sample: pivots per CPU step")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # started by hand without a launcher: one process per GPU through torch.distributed.run
        import socket
        with socket.socket() as sk:
            sk.bind(("127.0.0.1", 0)); port = sk.getsockname()[1]
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.gpus > 1 or int(os.environ.get("WORLD_SIZE", "1")) > 1:
        from rust_lp_b200 import multi
        return multi.bench(args, sys.modules[__name__])
    run_single(args)


if __name__ == "__main__":
    main()
