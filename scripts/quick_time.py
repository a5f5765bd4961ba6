import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from rust_lp_b200 import capi
from rust_lp_b200.synth import synthetic_lp
m, n = int(sys.argv[1]), int(sys.argv[2])
variant = sys.argv[3] if len(sys.argv) > 3 else "tu"
lib = sys.argv[4] if len(sys.argv) > 4 else None
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
pa = synthetic_lp(m, n, 20, variant)
for rep in range(reps):
    t0 = time.time()
    c = capi.Carry(pa, heap_words=1 << 26, lib_path=lib)
    t1 = time.time()
    st = c.solve_relaxation()
    t2 = time.time()
    s = c.stats()
    print(f"rep {rep} m={m} n={n} status={st} pivots={s.pivots} loop_ms={s.loop_ms:.2f} wall_solve={1e3*(t2-t1):.1f}ms create+load={1e3*(t1-t0):.1f}ms "
          f"launches={s.kernel_launches} pivots/s(loop)={s.pivots/(s.loop_ms/1e3):.0f} sumTS={s.sum_TS} escal={s.escalated} obj={c.get_objective_function_value()}", flush=True)
    c.close()
    print("   stage_us/iter:", [round(x / 1e3 / max(s.pivots, 1), 2) for x in s.stage_ns[:6]], "narrow", s.narrow_pivots)
