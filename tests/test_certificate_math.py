"""The arithmetic behind csrc/common.cuh certify_quotient (the certificate of the bfb factorisation's speculative
quotients), restated in exact rational arithmetic on the CPU: a quotient q of a/b is accepted iff the exactly computed
remainder a - b*q is below half an ulp of q times |b| (a quarter where q is a power of two, because the spacing below a
power of two is half the spacing above it), with exponents in a range where nothing under- or overflows, or iff a is +0
and q is the zero with b's sign. The claim tested here: accepted  =>  q == RN(a / b); and every correctly rounded quotient
of well-scaled operands IS accepted (the redo path is for rare operands only). The device code is tested against
__ddiv_rn on the GPU (tests/test_gpu_tridiag.py test_pipelined_division_is_ieee); this file pins the reasoning."""
import math
import random
import struct
from fractions import Fraction


def bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def from_bits(b):
    return struct.unpack("<d", struct.pack("<Q", b))[0]


def rn(fr):
    return float(fr)  # CPython rounds a Fraction to the nearest double, ties to even


def fma(a, b, c):
    if not (math.isfinite(a) and math.isfinite(b) and math.isfinite(c)):
        return float("nan")  # (what matters here: the comparison with it is false, as on the device)
    return rn(Fraction(a) * Fraction(b) + Fraction(c))


def certify(a, b, q):
    """certify_quotient(double, double, double) of common.cuh, statement for statement. True = accepted."""
    if math.isnan(q) or math.isinf(q):
        return False
    rem = fma(-b, q, a)
    ha, hb, hq = bits(a) >> 32, bits(b) >> 32, bits(q) >> 32
    ea, eb, eq = (ha >> 20) & 0x7FF, (hb >> 20) & 0x7FF, (hq >> 20) & 0x7FF
    pow2 = ((hq & 0xFFFFF) | (bits(q) & 0xFFFFFFFF)) == 0
    hexp = eq - 53 - (1 if pow2 else 0)
    b_ok = 600 <= eb <= 1446
    a_ok = 600 <= ea <= 1446
    # (for a tiny or zero q the device forms a meaningless h; a_ok and b_ok in range imply eq >= 177, so `small` only
    # counts where h is a proper power of two)
    small = 0 < hexp < 0x7FF and abs(rem) < from_bits(hexp << 52) * abs(b)
    zero = bits(a) == 0 and (bits(q) == (bits(b) & (1 << 63)))
    return b_ok and ((a_ok and small) or zero)


def neighbours(q):
    return [from_bits(bits(q) + 1), from_bits(bits(q) - 1)]


def test_correctly_rounded_quotients_are_accepted_and_neighbours_rejected():
    random.seed(7)
    for _ in range(20000):
        a = (random.random() - 0.5) * 10 ** random.uniform(-30, 30)
        b = (random.random() + 1e-3) * random.choice([1, -1]) * 10 ** random.uniform(-30, 30)
        q = rn(Fraction(a) / Fraction(b))
        assert q == a / b
        assert certify(a, b, q), (a, b)
        for w in neighbours(q):
            assert not certify(a, b, w), (a, b, w)


def test_power_of_two_quotients():
    # exact powers of two: remainder 0 -> accepted; just below a power of two the lower neighbour is only half an ulp away
    for e in (-40, -1, 0, 1, 7, 300):
        b = 3.0
        a = b * 2.0 ** e
        assert certify(a, b, 2.0 ** e)
    random.seed(11)
    for _ in range(20000):
        b = 1.0 + random.random()
        e = random.randint(-60, 60)
        a = rn(Fraction(b) * Fraction(2) ** e * (1 - Fraction(random.randint(1, 1 << 20), 1 << 72)))  # a/b a hair below 2^e
        q = a / b
        for cand in (2.0 ** e, from_bits(bits(2.0 ** e) - 1), q):
            if certify(a, b, cand):
                assert cand == q, (a, b, cand, q)  # whatever is accepted is the correctly rounded quotient


def test_signed_zero_numerators():
    for b in (2.5, -2.5, 1e-100, -1e100):
        assert certify(0.0, b, math.copysign(0.0, b))          # +0 / b = the zero with b's sign: exact on the fast path
        assert not certify(0.0, b, -math.copysign(0.0, b))     # the other zero is wrong
        assert not certify(-0.0, b, 0.0) and not certify(-0.0, b, -0.0)  # -0 numerators always go to the full-range division
    assert not certify(0.0, 1e-300, 0.0) and not certify(0.0, float("inf"), 0.0)  # divisor out of range


def test_out_of_range_operands_are_sent_to_the_redo():
    for a, b in ((1e-200, 3.0), (3.0, 1e-200), (1e200, 3.0), (3.0, 1e200), (float("nan"), 1.0), (1.0, float("nan")),
                 (float("inf"), 2.0), (5e-324, 1.0), (1.0, 0.0)):
        q = float("nan") if (b == 0 or math.isnan(a) or math.isnan(b)) else a / b
        assert not certify(a, b, q), (a, b)
