"""Latency of the reference-shaped calls (host arrays in, float out): configs 1 and 2, and the config-3 call pattern."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from hddm_b200 import wfpt
g = np.load("tests/golden/wfpt_golden.npz")
x1, x2 = np.ascontiguousarray(g["x_cfg1"]), np.ascontiguousarray(g["x_cfg2"])
x5 = np.ascontiguousarray(x1[:500])
def per_call(fn, reps):
    fn(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6
kw2 = dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
for cached in (True, False):
    wfpt.cache_config(cached)
    print("cache", cached, "500 simple %.1f us" % per_call(lambda: wfpt.wiener_like(x5, 1, 0, 2, .5, 0, .3, 0, 1e-4, **kw2), 2000),
          "cfg1 %.1f us" % per_call(lambda: wfpt.wiener_like(x1, 1, 0, 2, .5, 0, .3, 0, 1e-4), 2000),
          "cfg2 %.1f us" % per_call(lambda: wfpt.wiener_like(x2, 1, 1, 2, .5, .2, .3, .1, 1e-4, **kw2), 1000))
wfpt.cache_config(True)
print(bench.leg_cfg3(wfpt, 30, 1)["gpu_per_call"], bench.leg_cfg3(wfpt, 30, 1)["gpu_batched_per_stochastic"])
