"""CPU check of the arithmetic K1 uses when sqrt(D) is not a power of two (csrc/corr_build.cu, bf16 epilogue):
x / s as q = x*r, q += (x - q*s)*r with r = RN(1/s), two FMAs.  For every bf16 input of ordinary magnitude this is
the correctly rounded fp32 quotient, i.e. exactly what the reference's division computes (models/core/corr.py:104)."""
from fractions import Fraction

import numpy as np
import pytest


def rn32(fr):
    f = np.float32(float(fr))
    cands = [np.nextafter(f, np.float32(-np.inf)), f, np.nextafter(f, np.float32(np.inf))]
    return np.float32(min(cands, key=lambda c: (abs(Fraction(float(c)) - fr), int(np.float32(c).view(np.uint32)) & 1)))


@pytest.mark.parametrize("D", [128, 192])
def test_two_fma_quotient_equals_ieee_division_on_bf16_inputs(D):
    s = np.float32(np.sqrt(np.float32(D)))
    r = np.float32(np.float32(1.0) / s)
    S, R = Fraction(float(s)), Fraction(float(r))
    checked = 0
    for bits in range(1, 65536, 3):
        e = (bits >> 7) & 0xFF
        if e < 60 or e > 190:                       # no denormal quotients / overflow; |corr| is O(1..1e3)
            continue
        x = np.array([bits << 16], dtype=np.uint32).view(np.float32)[0]
        X = Fraction(float(x))
        q = rn32(X * R)
        err = rn32(X - Fraction(float(q)) * S)                       # fma(-q, s, x)
        q2 = rn32(Fraction(float(q)) + Fraction(float(err)) * R)     # fma(err, r, q)
        assert q2 == rn32(X / S), hex(bits)
        checked += 1
    assert checked > 10000
