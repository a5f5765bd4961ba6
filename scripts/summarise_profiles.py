"""gpurun_out/ (scripts/round_profile.sh) -> profiles/: the bench line, the launch list and its per-kernel shares,
DRAM traffic and pipe utilisation of the two dominant kernels from the ncu reports (read here, no GPU needed)."""
import csv, io, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
R = sys.argv[1] if len(sys.argv) > 1 else "r01"


def raw(rep):
    """`rep` is X.ncu-rep; round 2 exports X_raw.csv on the GPU box (scripts/round_ncu.sh) because the reports are too big to bring back"""
    exported = rep.replace(".ncu-rep", "_raw.csv")
    if os.path.exists(exported):
        out = open(exported).read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return dict(zip(rows[0], rows[2])), dict(zip(rows[0], rows[1]))


def num(d, k):
    try:
        return float(d[k].replace(",", ""))
    except Exception:
        return None


def unit_bytes(v, u):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)


def kernel_summary(rep, extra):
    d, u = raw(rep)
    rd = unit_bytes(num(d, "dram__bytes_read.sum"), u["dram__bytes_read.sum"])
    wr = unit_bytes(num(d, "dram__bytes_write.sum"), u["dram__bytes_write.sum"])
    t = num(d, "gpu__time_duration.sum") * {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(u["gpu__time_duration.sum"], 1)
    s = {"gpu_time_ms_under_ncu": t, "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_launch": rd + wr,
         "dram_GBps_under_ncu": (rd + wr) / t / 1e6,
         "dram_throughput_pct_of_peak": num(d, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
         "warp_instructions": num(d, "smsp__inst_executed.sum"),
         "issue_active_pct": num(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
         "alu_pipe_pct_of_peak_active": num(d, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
         "fma_pipe_pct_of_peak_active (IMAD)": num(d, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
         "lsu_pipe_pct_of_peak_active": num(d, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
         "registers_per_thread": num(d, "launch__registers_per_thread"),
         "warps_active_pct": num(d, "sm__warps_active.avg.pct_of_peak_sustained_active"),
         "xu_pipe_pct_of_peak_active": num(d, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
         "threads_per_warp_instruction": num(d, "smsp__thread_inst_executed_per_inst_executed.ratio"),
         "l2_hit_rate_pct": num(d, "lts__t_sector_hit_rate.pct")}
    s.update(extra)
    return s


os.makedirs(P, exist_ok=True)
bench = json.loads(open(os.path.join(G, "bench.json")).read().strip().splitlines()[-1])
# launch list -> shares
rows = [r for r in csv.reader(open(os.path.join(G, "launches.csv"))) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]; ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = {}
for r in rows[hdr + 1:]:
    name = r[ik].split("(")[0]
    v = float(r[iv].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[iu], 1e-6)
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
shutil.copy(os.path.join(G, "launches.csv"), os.path.join(P, f"launches_{R}.csv"))
with open(os.path.join(P, f"launches_{R}_summary.csv"), "w") as f:
    f.write("kernel,launches,total_ms,share\n")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write(f"{k},{a[0]},{a[1]:.3f},{a[1] / tot:.4f}\n")
piv = bench["pivots_per_solve"]
loop = kernel_summary(os.path.join(G, "loop.ncu-rep"), {
    "kernel": "k_simplex_loop<false>",
    "command": "ncu --set full --clock-control none --import-source on -k regex:k_simplex_loop --launch-skip 7 --launch-count 1 "
               "python bench.py --steps 1 --warmup 3 --cpu-pivots 10 --no-k1-dense",
    "note": f"phase-2 loop of one solve ({piv} pivots, all narrow), m=20000 n=200000"})
loop["dram_bytes_per_pivot"] = loop["dram_bytes_per_launch"] / piv
json.dump(loop, open(os.path.join(P, f"traffic_{R}.json"), "w"), indent=1)
k1 = json.loads(open(os.path.join(G, "k1_dense.json")).read())
k1n = kernel_summary(os.path.join(G, "k1_dense.ncu-rep"), {
    "kernel": "k_update (stage_update<true> -> update_dense_tma)",
    "command": "ncu --set full --clock-control none --import-source on -k regex:k_update$ -s 1 -c 1 python scripts/k1_dense.py 20000 3",
    "note": "one full-size rank-1 update of the Carry block: T = 19999 rows x S = 20001 columns, int64/int64 tier, integers",
    "cuda_event_timing_without_ncu": k1})
json.dump(k1n, open(os.path.join(P, f"k1_dense_{R}.json"), "w"), indent=1)
if os.path.exists(os.path.join(G, "update_big_raw.csv")):
    ub = kernel_summary(os.path.join(G, "update_big.ncu-rep"), {
        "kernel": "k_update_big (stage_update_big: the queued multi-limb elements of one pivot)",
        "command": "ncu --set full --clock-control none --import-source on -k regex:k_update_big --launch-skip 3500 --launch-count 1 "
                   "python scripts/one_solve.py netlib_25FV47 4000 0",
        "note": "one launch about 3 500 pivots into phase one of netlib 25FV47 (staged driver: one launch per stage)",
        "measured_integer_peaks": "profiles/k1_dense_rational_r01.json (scripts/micro/int_peak.cu: IMAD 18.0, IADD3 18.3, LOP3 12.3 T thread-ops/s)"})
    if ub["warp_instructions"] and ub["threads_per_warp_instruction"]:
        ub["thread_instructions_per_s_under_ncu"] = ub["warp_instructions"] * ub["threads_per_warp_instruction"] / (ub["gpu_time_ms_under_ncu"] * 1e-3)
    json.dump(ub, open(os.path.join(P, f"update_big_{R}.json"), "w"), indent=1)
if os.path.exists(os.path.join(G, "k1_frac.ncu-rep")):      # the same update on general rationals (integer-pipe bound) + measured integer peaks
    kf = kernel_summary(os.path.join(G, "k1_frac.ncu-rep"), {
        "kernel": "k_update on general rationals",
        "command": "ncu --set full --clock-control none --import-source on -k regex:k_update$ -s 1 -c 1 python scripts/k1_dense.py 20000 4 \"\" 1",
        "note": "same block, entries with denominators 2,3,5,7 and d_i over 11,13: every element takes small_fma_rational (three 64-bit gcds, 128-bit products and divisions)",
        "cuda_event_timing_without_ncu": json.loads(open(os.path.join(G, "k1_frac_plain.log")).read().strip().splitlines()[-1])})
    d, _ = raw(os.path.join(G, "k1_frac.ncu-rep"))
    kf["xu_pipe_pct_of_peak_active"] = num(d, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active")
    kf["threads_per_warp_instruction"] = num(d, "smsp__thread_inst_executed_per_inst_executed.ratio")
    kf["thread_instructions_per_element"] = kf["warp_instructions"] * kf["threads_per_warp_instruction"] / (20000.0 * 20001.0)
    kf["thread_instructions_per_s_under_ncu"] = kf["warp_instructions"] * kf["threads_per_warp_instruction"] / (kf["gpu_time_ms_under_ncu"] * 1e-3)
    if os.path.exists(os.path.join(G, "int_peak.jsonl")):
        kf["measured_integer_peaks (scripts/micro/int_peak.cu)"] = [json.loads(l) for l in open(os.path.join(G, "int_peak.jsonl"))]
    json.dump(kf, open(os.path.join(P, f"k1_dense_rational_{R}.json"), "w"), indent=1)
if os.path.exists(os.path.join(G, "netlib_times.jsonl")):
    shutil.copy(os.path.join(G, "netlib_times.jsonl"), os.path.join(P, f"netlib_times_{R}.jsonl"))
for n in (2, 4, 8):
    src = os.path.join(G, f"bench_g{n}.json")
    if os.path.exists(src) and os.path.getsize(src) > 10:
        shutil.copy(src, os.path.join(P, f"bench_{R}_gpus{n}.json"))
print(open(os.path.join(P, f"launches_{R}_summary.csv")).read())
print(json.dumps(loop, indent=1)); print(json.dumps(k1n, indent=1))
