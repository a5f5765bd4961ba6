"""Pins oracle/rational.hpp + bigint.hpp at the `num` boundary: the canonical-form and field-law
KATs of src/data/number_types/rational/mod.rs:14-104 and rational/big/test.rs, plus a fuzz
against Python's fractions.Fraction (multi-limb operands).  CPU only."""
import random
from fractions import Fraction as Fr

from oracle.coracle import rat_op


def test_reference_kats():
    assert rat_op("3/2", "0", "+") == rat_op("6/4", "0", "+") == "3/2"      # 3/2 == 6/4
    assert rat_op("0/2", "0", "+") == rat_op("0/5", "0", "+") == "0"        # 0/2 == 0/5
    assert rat_op("3/2", "6/4", "+") == "3"
    assert rat_op("3/2", "6/4", "*") == "9/4"
    assert rat_op("3/2", "6/4", "/") == "1"
    assert rat_op("3/2", "6/4", "-") == "0"
    assert rat_op("-1/3", "1/6", "+") == "-1/6"
    assert rat_op("1/3", "-2", "/") == "-1/6"
    assert rat_op("1/3", "1/2", "<") == "-1" and rat_op("-1/3", "-1/2", "<") == "1"


def test_fuzz_against_fractions():
    rnd = random.Random(7)

    def big():
        bits = rnd.choice([3, 20, 62, 63, 64, 65, 100, 127, 128, 200, 500, 1500])
        n = rnd.getrandbits(bits) * rnd.choice([1, -1])
        d = rnd.getrandbits(rnd.choice([1, 10, 63, 64, 90, 300])) + 1
        return Fr(n, d)

    for _ in range(1500):
        a, b = big(), big()
        assert Fr(rat_op(a, b, "+")) == a + b
        assert Fr(rat_op(a, b, "-")) == a - b
        assert Fr(rat_op(a, b, "*")) == a * b
        if b != 0:
            q = rat_op(a, b, "/")
            assert Fr(q) == a / b
            # canonical text form: reduced, positive denominator
            assert q == str(a / b)
        assert int(rat_op(a, b, "<")) == (a > b) - (a < b)


def test_int64_edges():
    lo, hi = -(2 ** 63), 2 ** 63 - 1
    for a in (lo, hi, lo + 1, -1, 0, 1):
        for b in (lo, hi, -1, 1, 2):
            fa, fb = Fr(a), Fr(b)
            assert Fr(rat_op(fa, fb, "*")) == fa * fb
            assert Fr(rat_op(fa, fb, "+")) == fa + fb
            assert Fr(rat_op(fa, fb, "-")) == fa - fb
            assert Fr(rat_op(fa, fb, "/")) == fa / fb
