"""CPU tests of the host-built series-term table (fast variant): for every err, looking tt up in the table must
give exactly the expansion and K that the reference's formulas give (checked against the C oracle, which is bit-
identical to the reference's Cython), including right at the piece boundaries."""
import bisect
import math

import numpy as np
import pytest

from hddm_b200 import _lib


def lookup(tab, tt):
    i = bisect.bisect_right(tab["thr"], tt)
    return tab["desc0"] if i == 0 else tab["desc"][i - 1]


@pytest.mark.parametrize("err", [1.0, 0.1, 1e-2, 1e-3, 1e-4, 1e-6, 1e-8, 1e-10, 1e-13, 1e-15])
def test_table_matches_oracle_decision(oracle, err):
    tab = _lib.ktable(err)
    assert 0 <= tab["n"] <= 32, tab
    assert tab["thr"] == sorted(tab["thr"])
    rng = np.random.default_rng(int(-math.log10(err)) + 1)
    tts = np.concatenate([10 ** rng.uniform(-8, 3, 4000), rng.uniform(0.01, 2.0, 4000)])
    # probe the boundaries themselves and their neighbours
    for th in tab["thr"]:
        tts = np.concatenate([tts, [th, np.nextafter(th, 0), np.nextafter(th, np.inf), th * (1 - 1e-12), th * (1 + 1e-12)]])
    for tt in tts:
        _, K, small = oracle.ftt_01w_info(float(tt), 0.4, err)
        d = lookup(tab, float(tt))
        assert (d >> 8, d & 0xff) == (int(small), K), (err, tt, d, K, small)
        assert _lib.ktable_decision(float(tt), err) == d


def test_table_shape_at_hddm_default():
    tab = _lib.ktable(1e-4)
    # small-time pieces first (K grows with tt), then large-time pieces with K falling to 1
    kinds = [tab["desc0"] >> 8] + [d >> 8 for d in tab["desc"]]
    assert kinds[0] == 1 and kinds[-1] == 0
    assert kinds == sorted(kinds, reverse=True)
    assert tab["desc"][-1] & 0xff == 1
    assert tab["kl_max"] <= 8 and tab["ks_max"] <= 8


def test_exp2_table_constants_are_correctly_rounded():
    """The 64-entry 2^(i/64) table the kernels copy per CTA (csrc/wfpt_fast.cuh, kExp2_64) against exact arithmetic: each entry
    must be the double nearest to 2^(i/64), i.e. the 64th power of its neighbours' midpoints must bracket 2^i."""
    import os
    import re
    from fractions import Fraction

    src = open(os.path.join(os.path.dirname(__file__), "..", "hddm_b200", "csrc", "wfpt_fast.cuh")).read()
    body = src[src.index("kExp2_64[64] = {"):]
    body = body[:body.index("};")]
    vals = [float.fromhex(h) for h in re.findall(r"0x1\.[0-9a-f]+p\+0|0x1p\+0|0x1\.0+p\+0", body)]
    assert len(vals) == 64
    ulp = Fraction(1, 2 ** 52)   # spacing of doubles in [1, 2)
    for i, v in enumerate(vals):
        f = Fraction(v)
        assert 1 <= f < 2
        lo, hi = f - ulp / 2, f + ulp / 2
        assert lo ** 64 <= Fraction(2) ** i <= hi ** 64, i
