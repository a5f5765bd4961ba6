"""The product's __host__ __device__ rational arithmetic (rust-lp_b200/csrc/relp_arith.cuh), built
for the host with g++, against fractions.Fraction.  CPU only: this is how the device number code
is validated before it ever runs in a kernel.  Canonical-form KATs follow
/root/reference/src/data/number_types/rational/mod.rs:14-104."""
import ctypes as C
import math
import os
import random
import subprocess
from fractions import Fraction as Fr

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "arith_host_test.cpp")
INC = os.path.join(ROOT, "rust-lp_b200", "csrc")
OUT = os.path.join(ROOT, "tests", "native", "_build", "libarith_host_test.so")


@pytest.fixture(scope="module")
def L():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    deps = [SRC, os.path.join(INC, "relp_arith.cuh"), os.path.join(INC, "relp_wcb.cuh")]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        subprocess.check_call(["g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-fsanitize=undefined",
                               "-fno-sanitize-recover=undefined", "-I", INC, "-x", "c++", SRC, "-o", OUT, "-lpthread"])
    lib = C.CDLL(OUT)
    lib.t_gcd64.restype = C.c_uint64
    lib.t_gcd64.argtypes = [C.c_uint64, C.c_uint64]
    lib.t_reset()
    return lib


def limbs(x):
    out = []
    while x:
        out.append(x & 0xffffffff); x >>= 32
    return np.array(out if out else [0], dtype=np.uint32), len(out)


def mk(lib, f):
    f = Fr(f)
    n, nn = limbs(abs(f.numerator)); d, nd = limbs(f.denominator)
    out = (C.c_int64 * 2)()
    lib.t_make(-1 if f < 0 else 1, n.ctypes.data_as(C.c_void_p), nn, d.ctypes.data_as(C.c_void_p), nd, out)
    return out


def rd(lib, v):
    if v[1] > 0:
        return Fr(v[0], v[1])
    assert v[1] == 0
    buf = np.zeros(1 + 2 * 16384, dtype=np.uint32)
    w = lib.t_read(v, buf.ctypes.data_as(C.c_void_p))
    h = int(buf[0]); nn = h & 0x3fff; nd = (h >> 14) & 0x3fff
    assert w == 1 + nn + nd
    num = sum(int(buf[1 + i]) << (32 * i) for i in range(nn))
    den = sum(int(buf[1 + nn + i]) << (32 * i) for i in range(nd))
    f = Fr(num, den)
    assert f.numerator == num and f.denominator == den, "stored value not canonical"
    assert buf[nn] != 0 and buf[nn + nd] != 0, "leading zero limb"
    assert num > 2**63 - 1 or den > 2**63 - 1, "big representation of a value that fits small"
    return -f if h >> 31 else f


def rnd_frac(rnd, bits):
    nb = rnd.choice([1, 3, 8, 20, 31, 32, 33, 50, 62, 63, 64, 65, 90, 127, 128, 200, 500])
    nb = min(nb, bits)
    db = rnd.choice([0, 0, 1, 5, 31, 32, 40, 63, 64, 100, 300])
    db = min(db, bits)
    n = rnd.getrandbits(nb) * rnd.choice([1, -1])
    d = rnd.getrandbits(db) + 1 if db else 1
    if rnd.random() < 0.1:
        n = 0
    return Fr(n, d)


def test_gcd64(L):
    rnd = random.Random(1)
    for _ in range(3000):
        a = rnd.getrandbits(rnd.randint(0, 64)); b = rnd.getrandbits(rnd.randint(0, 64))
        assert L.t_gcd64(a, b) == math.gcd(a, b)


def _limbs_arg(x):
    a, n = limbs(x)
    return a, a.ctypes.data_as(C.c_void_p), n


def _from_limbs(buf, n):
    return sum(int(buf[i]) << (32 * i) for i in range(n))


def test_mag_gcd_and_mod(L):
    """Lehmer gcd (relp_arith.cuh mag_gcd) and the Knuth-D remainder behind it, against math.gcd / %:
    equal and very different lengths, shared powers of two, multiples, quotient digits at the
    2^32 - 1 boundary, operands that trigger the add-back step."""
    rnd = random.Random(7)
    out = np.zeros(600, dtype=np.uint32)
    po = out.ctypes.data_as(C.c_void_p)
    for it in range(6000):
        ba = rnd.choice([33, 64, 65, 96, 97, 128, 160, 200, 257, 512, 1000, 2047])
        bb = rnd.choice([33, 64, 65, 96, 97, 128, 160, 200, 257, 512, 1000, 2047])
        a = rnd.getrandbits(ba) | 1 << (ba - 1); b = rnd.getrandbits(bb) | 1 << (bb - 1)
        mode = it % 6
        if mode == 1:
            g = rnd.getrandbits(rnd.choice([1, 31, 32, 64, 100, 300])) + 1; a *= g; b *= g
        elif mode == 2:
            a = b * (rnd.getrandbits(rnd.choice([1, 32, 33, 70])) + 1)
        elif mode == 3:
            a <<= rnd.randint(0, 70); b <<= rnd.randint(0, 70)
        elif mode == 4:            # quotient digits 0xffffffff / add-back candidates
            b = (1 << bb) - rnd.getrandbits(20) - 1; a = b * ((1 << 32) - 1) + rnd.getrandbits(bb - 1)
        elif mode == 5:
            a = (1 << ba) - 1; b = (1 << bb) - 1
        if a.bit_length() > 2048 * 2 or b.bit_length() > 2048 * 2:
            continue
        aa, pa, na = _limbs_arg(a); bbv, pb, nb = _limbs_arg(b)
        if max(na, nb) + 2 > 290:
            continue
        n = L.t_mag_gcd(pa, na, pb, nb, po)
        assert _from_limbs(out, n) == math.gcd(a, b), (hex(a), hex(b))
        if nb >= 2 and na >= nb:
            n = L.t_mag_mod(pa, na, pb, nb, po)
            assert _from_limbs(out, n) == a % b, (hex(a), hex(b))


def test_canonical_form_kats(L):
    """rational/mod.rs:26-60: 3/2 == 6/4, 0/2 == 0/5, 3/2 + 6/4 = 3, 3/2 * 6/4 = 9/4, (3/2)/(6/4) = 1"""
    a = mk(L, Fr(3, 2)); b = mk(L, Fr(6, 4)); z = mk(L, 0); one = mk(L, 1)
    assert (a[0], a[1]) == (b[0], b[1]) == (3, 2)
    assert (z[0], z[1]) == (0, 1)
    out = (C.c_int64 * 2)()
    L.t_fma(a, b, one, 1, out); assert (out[0], out[1]) == (3, 1)
    L.t_fma(z, a, b, 1, out); assert (out[0], out[1]) == (9, 4)
    L.t_div(a, b, out); assert (out[0], out[1]) == (1, 1)
    L.t_fma(a, b, one, -1, out); assert (out[0], out[1]) == (0, 1)


@pytest.mark.parametrize("bits,seed", [(12, 6), (20, 7), (31, 8), (40, 0), (64, 1), (64, 2), (130, 3), (500, 4), (500, 5)])
def test_fma_div_cmp_random(L, bits, seed):
    rnd = random.Random(seed)
    out = (C.c_int64 * 2)()
    for it in range(1500):
        if it % 200 == 0:
            L.t_reset()
        x, a, b = (rnd_frac(rnd, bits) for _ in range(3))
        if rnd.random() < 0.2:       # force cancellations / shared factors
            x = a * b * rnd.choice([1, -1]) if rnd.random() < 0.5 else x * Fr(a.denominator, max(1, b.denominator))
        vx, va, vb = mk(L, x), mk(L, a), mk(L, b)
        assert rd(L, vx) == x and rd(L, va) == a
        s = rnd.choice([1, -1])
        L.t_fma(vx, va, vb, s, out)
        assert rd(L, out) == x + s * a * b, (x, a, b, s)
        if b != 0:
            L.t_div(vx, vb, out)
            assert rd(L, out) == x / b, (x, b)
        assert L.t_cmp(vx, va) == (x > a) - (x < a)
        assert L.t_sign(vx) == (x > 0) - (x < 0)
        assert L.t_err() == 0


def test_no_fixed_limb_capacity(L):
    """RationalBig is unbounded (rational/big/mod.rs:17-29), so a value longer than what ONE thread
    handles (RELP_OPL limbs) is not an error: the cooperative routines take over, through the same front ends."""
    L.t_reset()
    opl = L.t_opl()
    big = Fr(2 ** (32 * opl - 1) + 1, 3)
    v = mk(L, big)
    out = (C.c_int64 * 2)()
    L.t_fma(mk(L, 0), v, v, 1, out)          # the square has 2 * opl limbs
    assert L.t_err() == 0 and rd(L, out) == big * big
    w = mk(L, Fr(7 ** 900, 5 ** 700))        # 2527 / 1626 bits
    L.t_fma(out, w, v, -1, out)
    assert L.t_err() == 0 and rd(L, out) == big * big - Fr(7 ** 900, 5 ** 700) * big
    L.t_div(out, w, out)
    assert rd(L, out) == (big * big - Fr(7 ** 900, 5 ** 700) * big) / Fr(7 ** 900, 5 ** 700)
    val, wf = rd(L, out), Fr(7 ** 900, 5 ** 700)
    assert L.t_cmp(out, w) == (val > wf) - (val < wf) and L.t_cmp(w, out) == (wf > val) - (wf < val) and L.t_cmp(w, w) == 0
    L.t_reset()


# ---- cooperative tier (relp_wcb.cuh): the templates run on S host threads behind a barrier -----------------
W_OPS = {"add": 0, "sub": 1, "mul": 2, "mod": 3, "gcd": 4, "divexact": 5, "mul1": 6, "shl": 7, "shr": 8, "cmp": 9}


def w_op(L, S, op, a, b):
    aa, pa, na = _limbs_arg(a); bb, pb, nb = _limbs_arg(b)
    out = np.zeros(2 * (na + nb) + 16 + (b // 32 if op == "shl" else 0), dtype=np.uint32)
    n = L.t_w_op(S, W_OPS[op], pa, na, pb, nb, out.ctypes.data_as(C.c_void_p))
    return n if op == "cmp" else _from_limbs(out, n)


def _rnd_big(rnd, bits):
    style = rnd.randrange(6)
    if style == 0:
        return (1 << bits) - 1                                   # all ones: carries ripple through every lane
    if style == 1:
        return 1 << (bits - 1)                                   # a single bit: borrows ripple
    if style == 2:
        return (((1 << bits) - 1) ^ (rnd.getrandbits(bits) & rnd.getrandbits(bits) & rnd.getrandbits(bits))) or 1   # mostly ones
    if style == 3:
        return rnd.getrandbits(bits) & ~((1 << rnd.randrange(bits)) - 1) | 1 << (bits - 1)                    # low zeros
    return rnd.getrandbits(bits) | 1 << (bits - 1)


@pytest.mark.parametrize("S", [1, 2, 3, 5, 8, 17, 32])
def test_wcb_primitives(L, S):
    """every pass of relp_wcb.cuh against Python integers, for group sizes that divide the lengths and that do not:
    carry-lookahead across lanes and across chunks, column-sum multiplication, Knuth-D remainder with add-back,
    Lehmer gcd, Jebelean exact division, shifts"""
    rnd = random.Random(100 + S)
    sizes = [1, 31, 32, 33, 63, 64, 65, 96, 127, 128, 129, 255, 256, 300, 511, 777, 1024, 1025, 2049, 3000]
    for it in range(160 if S == 1 else 60):        # S host threads behind a spin barrier: seconds per hundred operations
        ba, bb = rnd.choice(sizes), rnd.choice(sizes)
        a, b = _rnd_big(rnd, ba), _rnd_big(rnd, bb)
        assert w_op(L, S, "add", a, b) == a + b
        hi, lo = max(a, b), min(a, b)
        assert w_op(L, S, "sub", hi, lo) == hi - lo
        assert w_op(L, S, "sub", hi, hi) == 0
        assert w_op(L, S, "mul", a, b) == a * b
        m1 = rnd.choice([1, 2, 0xffffffff, rnd.getrandbits(32) | 1])
        assert w_op(L, S, "mul1", a, m1) == a * m1
        sh = rnd.choice([1, 5, 31, 32, 33, 64, 100])
        assert w_op(L, S, "shl", a, sh) == a << sh
        if sh < ba:
            assert w_op(L, S, "shr", a, sh) == a >> sh
        assert w_op(L, S, "cmp", a, b) == (a > b) - (a < b)
        assert w_op(L, S, "cmp", a, a) == 0
        if a >= b:
            assert w_op(L, S, "mod", a, b) == a % b
        mode = it % 5
        g = 1
        if mode == 1:
            g = rnd.getrandbits(rnd.choice([1, 31, 32, 64, 100, 300])) + 1
        elif mode == 2:
            g = 1 << rnd.randint(1, 70)
        elif mode == 3:
            a = b * (rnd.getrandbits(rnd.choice([1, 32, 33, 70])) + 1)
        elif mode == 4:            # quotient digits at the 2^32 - 1 boundary / add-back candidates
            b = (1 << bb) - rnd.getrandbits(min(20, bb)) - 1 if bb > 1 else 1
            a = b * ((1 << 32) - 1) + rnd.getrandbits(max(bb - 1, 1))
            assert w_op(L, S, "mod", a, b) == a % b
        a2, b2 = a * g, b * g
        assert w_op(L, S, "gcd", a2, b2) == math.gcd(a2, b2), (hex(a2), hex(b2))
        d = math.gcd(a2, b2)
        assert w_op(L, S, "divexact", a2, d) == a2 // d
        q = rnd.getrandbits(rnd.choice([1, 32, 64, 200])) + 1
        assert w_op(L, S, "divexact", a2 * q, q) == a2


@pytest.mark.parametrize("S,bits,seed", [(1, 500, 1), (4, 130, 2), (32, 500, 3), (32, 3000, 4), (7, 3000, 5)])
def test_wcb_rationals(L, S, bits, seed):
    """wcb_fma / wcb_div / wcb_cmp (the cooperative versions of the three rational operations) against Fraction"""
    rnd = random.Random(seed)
    out = (C.c_int64 * 2)()

    def big_frac():
        nb = rnd.choice([1, 40, 64, 65, 200, bits])
        db = rnd.choice([0, 1, 33, 64, 100, bits])
        n = rnd.getrandbits(nb) * rnd.choice([1, -1])
        if rnd.random() < 0.08:
            n = 0
        return Fr(n, rnd.getrandbits(db) + 1 if db else 1)

    for it in range(40 if bits > 1000 else 120):
        if it % 20 == 0:
            L.t_reset()
        x, a, b = big_frac(), big_frac(), big_frac()
        if rnd.random() < 0.3:       # cancellations and shared factors
            x = a * b * rnd.choice([1, -1]) if rnd.random() < 0.5 else x * Fr(a.denominator, max(1, b.denominator))
        vx, va, vb = mk(L, x), mk(L, a), mk(L, b)
        s = rnd.choice([1, -1])
        L.t_w_fma(S, vx, va, vb, s, out)
        assert rd(L, out) == x + s * a * b, (x, a, b, s)
        if b != 0:
            L.t_w_div(S, vx, vb, out)
            assert rd(L, out) == x / b
        assert L.t_w_cmp(S, vx, va) == (x > a) - (x < a)
        assert L.t_w_cmp(S, vx, vx) == 0
        assert L.t_err() == 0


@pytest.mark.parametrize("slab_words", [100000, 700, 150])
def test_wcb_scratch_slab_stack_discipline(L, slab_words):
    """the cooperative operations take their temporaries from a slab and give them back as soon as a partial result is
    final (gcd: four of five buffers, exact quotient: three of four); a slab that is too small makes them spill into the
    arena midway, after which nothing is given back -- results must be the same in every case"""
    rnd = random.Random(slab_words)
    out = (C.c_int64 * 2)()
    L.t_set_slab(slab_words)
    try:
        for it in range(30):
            if it % 10 == 0:
                L.t_reset()
            bits = rnd.choice([70, 200, 500, 1500])
            x, a, b = (Fr(rnd.getrandbits(bits) * rnd.choice([1, -1]), rnd.getrandbits(bits) + 1) for _ in range(3))
            if it % 3 == 0:
                x = x * Fr(a.denominator * 6, 35)         # shared factors: the gcds are not 1
            vx, va, vb = mk(L, x), mk(L, a), mk(L, b)
            for S in (1, 5, 32):
                L.t_w_fma(S, vx, va, vb, -1, out)
                assert rd(L, out) == x - a * b
                if b != 0:
                    L.t_w_div(S, vx, vb, out)
                    assert rd(L, out) == x / b
                assert L.t_w_cmp(S, vx, va) == (x > a) - (x < a)
            assert L.t_err() == 0
    finally:
        L.t_set_slab(0)
