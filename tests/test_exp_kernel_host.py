"""The lean kernels' exp (hddm_b200/csrc/like_lean.cu, expv) restated on the host with exact rational arithmetic standing in
for the fused multiply-adds: pins the seven-instruction scheme -- one-constant argument reduction, 2,048-entry table whose
entries carry -j * 2^9 in the high word, exponent put in by one integer multiply-add -- against mpmath.  No GPU."""
import os
import re
import struct
from fractions import Fraction

import mpmath
import numpy as np

mpmath.mp.prec = 200
SRC = os.path.join(os.path.dirname(__file__), "..", "hddm_b200", "csrc", "like_lean.cu")


def _bits(x):
    return struct.unpack("<q", struct.pack("<d", x))[0] & 0xFFFFFFFFFFFFFFFF


def _from_bits(b):
    return struct.unpack("<d", struct.pack("<Q", b & 0xFFFFFFFFFFFFFFFF))[0]


def _fma(a, b, c):
    return float(Fraction(a) * Fraction(b) + Fraction(c))   # one rounding, as the hardware's


def _constants():
    src = open(SRC).read()
    m = re.search(r"__constant__ double kE\[5\] = \{([^}]*)\}", src)
    assert m, "kE[5] not found: the test restates like_lean.cu's exp and must be updated with it"
    parts = [p.strip() for p in m.group(1).split(",")]
    assert parts[2] == "-0.6931471805599453094172321 / 2048.0" and parts[3] == "1.0 / 6.0"
    return float(parts[0]), float(parts[1]), -0.6931471805599453094172321 / 2048.0, 1.0 / 6.0, float(parts[4])


def _table():
    tab = []
    for j in range(2048):
        v = float(mpmath.power(2, mpmath.mpf(j) / 2048))          # correctly rounded, as ensure_exp_table's exp2l
        tab.append((_bits(v) - (j << 41)) & 0xFFFFFFFFFFFFFFFF)    # the pre-biased entry (see expv)
    return tab


def _expv(x, kE, tab):
    t = _fma(x, kE[0], kE[1])
    ki = _bits(t) & 0xFFFFFFFF
    ki = ki - (1 << 32) if ki & 0x80000000 else ki                 # __double2loint: signed low word
    kd = t - kE[1]
    r = _fma(kd, kE[2], x)
    q = _fma(r, kE[3], kE[4])
    p = _fma(q, r * r, r)
    tb = tab[ki & 2047]
    hi = ((ki * 512) + (tb >> 32)) & 0xFFFFFFFF                    # mad.lo.s32 on the high word
    s = _from_bits((hi << 32) | (tb & 0xFFFFFFFF))
    return _fma(s, p, s)


def test_lean_exp_scheme_against_mpmath():
    kE, tab = _constants(), _table()
    rng = np.random.default_rng(7)
    xs = np.concatenate([rng.uniform(-40, 5, 600), rng.uniform(-707, -40, 150), rng.uniform(5, 707, 50),
                         [0.0, -0.0, 1e-300, -1e-17, -707.0, 707.9, np.log(2.0), -np.log(2.0) / 4096]])
    worst = 0.0
    for x in xs:
        got = _expv(float(x), kE, tab)
        want = mpmath.exp(mpmath.mpf(float(x)))
        rel = abs((mpmath.mpf(got) - want) / want)
        # 1.5 ulp of arithmetic + the reduction constant's rounding, which acts like 2^-54 |x| on the argument
        assert rel <= 3.4e-16 + 6e-17 * abs(x), (x, got, float(rel))
        worst = max(worst, float(rel) / (1.0 + abs(x) * 0.17))
    assert worst < 3.4e-16


def test_lean_exp_exponent_add_is_exact():
    """S = 2^e * 2^(j/2048) must come out of the integer multiply-add exactly, for every j and for e at both ends."""
    tab = _table()
    for j in (0, 1, 2, 1023, 1024, 2046, 2047):
        tj = float(mpmath.power(2, mpmath.mpf(j) / 2048))
        for e in (-1020, -1, 0, 1, 1020):
            ki = 2048 * e + j
            hi = ((ki * 512) + (tab[j] >> 32)) & 0xFFFFFFFF
            s = _from_bits((hi << 32) | (tab[j] & 0xFFFFFFFF))
            assert s == float(Fraction(tj) * Fraction(2) ** e), (j, e)
