import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from rust_lp_b200 import capi
from test_mps_golden import load_fixture
name = sys.argv[1]
opts = dict(kv.split("=") for kv in sys.argv[2:])
pa, _ = load_fixture(name)
for rep in range(2):
    c = capi.Carry(pa, heap_words=1 << 24)
    for k, v in opts.items(): c.set_option(k, int(v))
    st = c.solve_relaxation(1, 2)
    s = c.stats()
    print(name, opts, "status", st, "pivots", s.pivots, "loop_ms %.1f" % s.loop_ms, "narrow", s.narrow_pivots, "esc", s.escalated, "compactions", s.compactions,
          "stage_ms", [round(x / 1e6, 1) for x in s.stage_ns[:6]], flush=True)
    c.close()
