"""The exact arithmetic as it runs in kernels (relp_arith.cuh + relp_wcb.cuh), against fractions.Fraction: one
operation per GPU thread through val_fma / val_div / val_cmp (per-thread routines up to `coop_from` limbs, the
arbitrary-length routines on one lane beyond), and the same multiply-adds done by whole warps (limbs across lanes,
carries by ballot: what stage_update_big does).  The oracle-sized LPs never drive values beyond ~14 limbs; here
operands go up to 6000 bits (190 limbs: three times what the fixed-size per-thread routines handle)."""
import random
from fractions import Fraction as Fr

import pytest

from helpers import feasible_lp, pa_from_lp
from rust_lp_b200 import capi

pytestmark = pytest.mark.gpu


def _frac(rnd, bits):
    nb = rnd.choice([1, 20, 62, 64, 65, 128, 300, bits, bits])
    db = rnd.choice([0, 0, 1, 40, 64, 100, bits])
    n = rnd.getrandbits(nb) * rnd.choice([1, -1])
    if rnd.random() < 0.06:
        n = 0
    return Fr(n, rnd.getrandbits(db) + 1 if db else 1)


@pytest.fixture(scope="module")
def ctx():
    c = capi.Carry(pa_from_lp(feasible_lp(3, 3, 1, 3, 2, 10)), heap_words=1 << 26)
    c.create_for_partially_artificial()
    yield c
    c.close()


@pytest.mark.parametrize("bits,count,coop_from,seed", [(200, 600, 64, 1), (1500, 400, 64, 2), (3000, 300, 64, 3), (6000, 96, 64, 4),
                                                       (500, 400, 2, 5), (3000, 200, 8, 6)])
def test_device_fma_div_cmp(ctx, bits, count, coop_from, seed):
    rnd = random.Random(seed)
    ctx.set_option("coop_from", coop_from)
    xs, as_, bs = [], [], []
    for _ in range(count):
        x, a, b = _frac(rnd, bits), _frac(rnd, bits), _frac(rnd, bits)
        r = rnd.random()
        if r < 0.15:          # exact cancellation and shared factors
            x = -a * b
        elif r < 0.3:
            x = x * Fr(a.denominator, max(1, b.denominator))
        xs.append(x); as_.append(a); bs.append(b)
    got = ctx.debug_rat_op(0, xs, as_, bs, s=1)
    assert got == [x + a * b for x, a, b in zip(xs, as_, bs)]
    got = ctx.debug_rat_op(0, xs, as_, bs, s=-1)
    assert got == [x - a * b for x, a, b in zip(xs, as_, bs)]
    got = ctx.debug_rat_op(3, xs, as_, bs, s=-1)          # the same operations, each by the 32 lanes of its warp together
    assert got == [x - a * b for x, a, b in zip(xs, as_, bs)]
    nz = [(x, a) for x, a in zip(xs, as_) if a != 0]
    got = ctx.debug_rat_op(1, [p[0] for p in nz], [p[1] for p in nz], None)
    assert got == [x / a for x, a in nz]
    ys = [x if rnd.random() < 0.2 else a for x, a in zip(xs, as_)]
    got = ctx.debug_rat_op(2, xs, ys, None)
    assert got == [(x > y) - (x < y) for x, y in zip(xs, ys)]
    ctx.set_option("coop_from", 64)


def test_long_values_next_to_short_ones(ctx):
    """one 4000-bit operation in a warp of int64 ones, and a warp where every lane has one: the leader loop"""
    rnd = random.Random(9)
    big = lambda: Fr(rnd.getrandbits(4000) | 1, rnd.getrandbits(3900) | 1)
    xs = [Fr(rnd.randint(-9, 9), rnd.randint(1, 5)) for _ in range(64)]
    as_ = [Fr(rnd.randint(-9, 9)) for _ in range(64)]
    bs = [Fr(rnd.randint(1, 9), 7) for _ in range(64)]
    xs[13] = big(); as_[13] = big(); bs[40] = big()
    xs += [big() for _ in range(32)]; as_ += [big() for _ in range(32)]; bs += [big() for _ in range(32)]
    assert ctx.debug_rat_op(0, xs, as_, bs, s=-1) == [x - a * b for x, a, b in zip(xs, as_, bs)]
