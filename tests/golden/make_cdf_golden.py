"""Generate tests/golden/cdf_golden.npz from the reference's own C (src/cdfdif.c compiled as it lies into
oracle/_ref/libcdfdif_ref.so) through the restated wrapper arithmetic of src/cdfdif_wrapper.pyx:36-51.

    python tests/golden/make_cdf_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.build_ref import build_cdf_ref  # noqa: E402
from oracle import cpu  # noqa: E402

assert build_cdf_ref()
rng = np.random.default_rng(31)
x = np.concatenate([np.linspace(-4.5, 4.5, 181), [0.31, -0.31, 0.2, -0.2, 1e-3]])
cases, vals = [], []
rows = [
    # v, sv, a, z, sz, t, st, p_outlier, w_outlier
    (0.7, 0.8, 1.6, 0.45, 0.2, 0.25, 0.15, 0.05, 0.1),
    (1.0, 1.0, 2.0, 0.5, 0.2, 0.3, 0.1, 0.0, 0.1),
    (-1.5, 0.5, 1.2, 0.6, 0.3, 0.35, 0.2, 0.05, 0.1),
    (0.0, 1.0, 2.5, 0.5, 0.1, 0.2, 0.3, 0.1, 0.05),
    (1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 0.0, 0.1),      # the wrapper's 1e-10 nudges: ill-conditioned by construction
    (2.0, 0.0, 1.5, 0.4, 0.0, 0.3, 0.1, 0.05, 0.1),
    (0.5, 1.5, 3.0, 0.5, 0.25, 0.5, 0.0, 0.05, 0.1),
]
for r in rows:
    cases.append(r)
    vals.append(cpu.dmat_cdf_array_ref(x, *r))
np.savez_compressed(os.path.join(HERE, "cdf_golden.npz"), x=x, cases=np.array(cases), vals=np.array(vals))
print("wrote cdf_golden.npz", np.array(vals).shape)
