"""Latency and throughput of the device rational arithmetic by operand size (GPU).

    python scripts/arith_time.py [out.jsonl]
    python scripts/arith_time.py --one BITS OP COUNT      (one configuration, three launches: for ncu)

Drives relp_debug_rat_op (the kernels' own front ends val_fma / val_div, one operation per thread, and op 3: one
operation per warp through the cooperative tier) with random canonical operands of a given bit length and reports
the kernel time (CUDA events inside the library, relp_stats.loop_ms): `count = 1` is the latency of ONE dependent
chain -- what a pivot's serial stages (pivot row, cost row, ratio test) wait for --, `count = 32 768` the throughput
the Carry update of escalated elements sees.  Results are checked against fractions.Fraction."""
import json
import os
import random
import sys
from fractions import Fraction as Fr

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rust_lp_b200 import capi                     # noqa: E402
from rust_lp_b200.synth import synthetic_lp       # noqa: E402


def rnd(bits, rng):
    n = rng.getrandbits(bits) | (1 << (bits - 1)) | 1
    d = rng.getrandbits(bits) | (1 << (bits - 1)) | 1
    return Fr(n if rng.random() < 0.5 else -n, d)


def main():
    one = None
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one = (int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 and not one else None
    c = capi.Carry(synthetic_lp(30, 60, 5), heap_words=1 << 28)
    c.create_for_partially_artificial()
    rng = random.Random(7)
    for bits in (40, 96, 128, 256, 512, 1024, 1600, 3200, 6400) if not one else (one[0],):
        for op, name in ((0, "fma, one thread"), (3, "fma, one warp"), (1, "div, one thread")):
            if one and op != one[1]:
                continue
            for count in (1, 32768 if bits <= 1600 else 2048) if not one else (one[2],):
                if op == 3 and count > 1 and not one:
                    count //= 8
                base = [(rnd(bits, rng), rnd(bits, rng), rnd(bits, rng)) for _ in range(min(count, 64))]
                xs = [base[k % len(base)][0] for k in range(count)]
                as_ = [base[k % len(base)][1] for k in range(count)]
                bs = [base[k % len(base)][2] for k in range(count)]
                best = None
                for _ in range(3):
                    got = c.debug_rat_op(op, xs, as_, bs, -1)
                    ms = c.stats().loop_ms
                    best = ms if best is None else min(best, ms)
                for k in range(min(count, 64)):
                    want = xs[k] / as_[k] if op == 1 else xs[k] - as_[k] * bs[k]
                    assert got[k] == want, (bits, op, k)
                rec = {"bits": bits, "op": name, "count": count, "kernel_ms": round(best, 4),
                       "us_per_op_latency" if count == 1 else "Mops_per_s": round(best * 1e3, 2) if count == 1 else round(count / best / 1e3, 3)}
                print(json.dumps(rec), flush=True)
                if out:
                    out.write(json.dumps(rec) + "\n")
    c.close()


if __name__ == "__main__":
    main()
