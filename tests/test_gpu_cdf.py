"""GPU parity of the diffusion-model CDF (hddm.cdfdif.dmat_cdf_array) against fixtures produced by the reference's own
C code (tests/golden/make_cdf_golden.py) and, where oracle/_ref travels, against that library live.

Tolerances: 1e-9 absolute (values are probabilities in [0,1]) where sz and st are away from zero; 1e-5 absolute where
the wrapper's 1e-10 nudges divide the series by ~1e-10-sized widths -- the reference's own result is only defined to
about 1e-6 there (rounding noise amplified by 1/(sz*st)), whichever libm evaluates it."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "cdf_golden.npz")))


def test_cdf_matches_reference_fixture():
    from hddm_b200 import cdfdif

    x = G["x"]
    for row, want in zip(G["cases"], G["vals"]):
        got = cdfdif.dmat_cdf_array(x, *row)
        tol = 1e-9 if (row[4] > 1e-3 and row[6] > 1e-3) else 1e-5
        assert np.max(np.abs(got - want)) <= tol, (row, np.max(np.abs(got - want)))
        if tol == 1e-9:
            # a CDF in signed-RT convention: non-decreasing, from ~0 to ~1
            assert np.all(np.diff(got[:181]) >= -1e-9)


def test_cdf_live_reference_and_errors():
    from hddm_b200 import cdfdif
    from oracle import cpu
    from oracle.build_ref import CDF_LIB

    with pytest.raises(ValueError, match="out of the support"):
        cdfdif.dmat_cdf_array(np.array([0.5]), 1, 0, -1, 0.5, 0, 0.3, 0, 0, 0.1)
    if not os.path.exists(CDF_LIB):
        pytest.skip("oracle/_ref/libcdfdif_ref.so not present")
    rng = np.random.default_rng(8)
    x = np.where(rng.random(300) < 0.7, 1, -1) * rng.uniform(0.2, 4.0, 300)
    for _ in range(5):
        p = (rng.uniform(-2, 2), rng.uniform(0.2, 1.5), rng.uniform(1, 3), rng.uniform(0.35, 0.65), rng.uniform(0.05, 0.3),
             rng.uniform(0.2, 0.5), rng.uniform(0.05, 0.3), 0.05, 0.1)
        want = cpu.dmat_cdf_array_ref(x, *p)
        got = cdfdif.dmat_cdf_array(x, *p)
        assert np.max(np.abs(got - want)) <= 1e-9, p
