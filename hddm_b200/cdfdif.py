"""Drop-in for the reference's extension module `cdfdif_wrapper` (/root/reference/src/cdfdif_wrapper.pyx), reached in
HDDM as `hddm.cdfdif.dmat_cdf_array` (/root/reference/hddm/likelihoods.py:232-233)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .wfpt import _buffer


def dmat_cdf_array(x, v, sv, a, z, sz, t, st, p_outlier, w_outlier):
    x = _buffer(x, "x")
    v, sv, a, z, sz, t, st = (float(v), float(sv), float(a), float(z), float(sz), float(t), float(st))
    p_outlier, w_outlier = float(p_outlier), float(w_outlier)
    # cdfdif_wrapper.pyx:20-26
    if p_outlier > 0:
        assert np.max(np.abs(x)) < (1.0 / (2 * w_outlier)), ValueError("1. / (2*w_outlier) must be smaller than RT")
    if (sv < 0) or (a <= 0) or (z < 0) or (z > 1) or (sz < 0) or (sz > 1) or (z + sz / 2.0 > 1) or \
            (z - sz / 2.0 < 0) or (t - st / 2.0 < 0) or (t < 0) or (st < 0) or not (0 <= p_outlier <= 1):
        raise ValueError("at least one of the parameters is out of the support")
    y = np.empty(x.shape[0], dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_dmat_cdf_array_host(x.ctypes.data_as(C.c_void_p), x.shape[0], v, sv, a, z, sz, t, st, p_outlier,
                                          w_outlier, y.ctypes.data_as(C.c_void_p)))
    return y
