"""K1 (Carry update) alone at the dense bound: relp_debug_k1_dense on the bench LP's Carry block.
usage: python scripts/k1_dense.py [m] [reps] [lib or ""] [1 = general rationals instead of integers]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as entry
entry.build_library()
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
m = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
lib = (sys.argv[3] or None) if len(sys.argv) > 3 else None
frac = int(sys.argv[4]) if len(sys.argv) > 4 else 0
pa = synthetic_lp(m, 10 * m, 20, "tu")
c = capi.Carry(pa, heap_words=1 << 24, lib_path=lib)
c.set_option("k1_bench_frac", frac)
c.create_for_partially_artificial(pa.piv_rows, pa.piv_cols)
ms_row, ms_upd, T, S = c.debug_k1_dense(seed=20260101, reps=reps)
alg = 16 * (2 * T * S + S + T)
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
print(json.dumps({"lib": lib, "data": "rationals (den 2,3,5,7; d_i over 11,13)" if frac else "integers", "m": m, "T": T, "S": S, "ms_update": ms_upd, "ms_pivot_row": ms_row, "algorithmic_GBps": alg / ms_upd / 1e6,
                  "moved_GBps": (32 if frac else 24) * T * S / ms_upd / 1e6, "elements_per_s": T * S / ms_upd * 1e3, "frac_of_hbm_peak": alg / ms_upd / 1e6 / peak}))
