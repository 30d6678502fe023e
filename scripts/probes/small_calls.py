import sys; sys.path.insert(0, "/root/repo")
import numpy as np
from hddm_b200 import wfpt
g = np.load("/root/repo/tests/golden/wfpt_golden.npz")
x1 = np.ascontiguousarray(g["x_cfg1"]); x5 = np.ascontiguousarray(x1[:500]); x2 = np.ascontiguousarray(g["x_cfg2"])
kw2 = dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
for _ in range(3):
    wfpt.wiener_like(x5, 1, 0, 2, .5, 0, .3, 0, 1e-4, **kw2)
    wfpt.wiener_like(x1, 1, 0, 2, .5, 0, .3, 0, 1e-4)
    wfpt.wiener_like(x2, 1, 1, 2, .5, .2, .3, .1, 1e-4, **kw2)
