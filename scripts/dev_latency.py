"""Dev: latency of single wiener_like calls through the drop-in (host array in, float out)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import wfpt
from hddm_b200 import _lib

g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "wfpt_golden.npz"))
x1, x2 = np.ascontiguousarray(g["x_cfg1"]), np.ascontiguousarray(g["x_cfg2"])

def per_call(fn, reps=300):
    for _ in range(20): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6

print("cfg1 simple 1000 trials      %.1f us" % per_call(lambda: wfpt.wiener_like(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)))
print("cfg2 full adaptive 10000     %.1f us" % per_call(lambda: wfpt.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)))
print("500 trials full adaptive     %.1f us" % per_call(lambda: wfpt.wiener_like(x2[:500], 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)))
print("10 trials simple             %.1f us" % per_call(lambda: wfpt.wiener_like(x1[:10], 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)))
# the C call alone (arguments prepared once)
L = _lib.load()
prm = np.array([[1, 0, 2, 0.5, 0, 0.3, 0, 0.0]], dtype=np.double); res = np.empty(1)
px, pp, pr = x1.ctypes.data, prm.ctypes.data, res.ctypes.data
print("C call alone, cfg1           %.1f us" % per_call(lambda: L.wfpt_like_batch_host(px, 1000, pp, 1, 1e-4, 10, 10, 1, 1e-8, 0.1, 0, pr)))
print("pdf_array 1000               %.1f us" % per_call(lambda: wfpt.pdf_array(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)))
print(wfpt.wiener_like(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4), float(g["cfg1"]) if "cfg1" in g else "")
vv = np.linspace(0.5, 1.5, 1000); tt = np.full(1000, 0.3)
print("wiener_like_multi 1000 (v, t) %.1f us" % per_call(lambda: wfpt.wiener_like_multi(x1, vv, 0, 2, 0.5, 0, tt, 0, 1e-4, multi=["v", "t"], p_outlier=0.05, w_outlier=0.1)))
