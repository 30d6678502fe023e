import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import wfpt
from hddm_b200 import simulators as S
g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "wfpt_golden.npz"))
x1 = np.ascontiguousarray(g["x_cfg1"])
def per_call(fn, reps=200):
    for _ in range(10): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6
rng = np.random.default_rng(0)
n = 1000
resp = rng.integers(0, 2, n).astype(np.int64); fb = rng.integers(0, 2, n).astype(np.float64); split = rng.integers(0, 3, n).astype(np.int64)
print("wiener_like_rlddm 1000       %.1f us" % per_call(lambda: wfpt.wiener_like_rlddm(x1, resp, fb, split, 0.5, 0.1, 100.0, 1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 1e-4, p_outlier=0.05, w_outlier=0.1)))
print("wiener_like_rl 1000          %.1f us" % per_call(lambda: wfpt.wiener_like_rl(resp, fb, split, 0.5, 0.1, 100.0, 1.0, 0.5)))
print("wiener_like_contaminant 1000 %.1f us" % per_call(lambda: wfpt.wiener_like_contaminant(x1, np.zeros(n, dtype=np.int32), 1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 0.1, 5.0, 1e-4)))
print("pdf_array 1000               %.1f us" % per_call(lambda: wfpt.pdf_array(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)))
print("simulator ddm 1000 samples   %.1f us" % per_call(lambda: S.simulator([1.0, 1.5, 0.5, 0.3], model="ddm", n_samples=1000), 50))
print("simulator angle 1e5 samples  %.1f us" % per_call(lambda: S.simulator([1.0, 1.5, 0.5, 0.3, 0.2], model="angle", n_samples=100000), 20))
print("simulator race_3 1000        %.1f us" % per_call(lambda: S.simulator([1.0, 1.2, 0.8, 1.5, 0.5, 0.5, 0.5, 0.3], model="race_3", n_samples=1000), 50))
try:
    from oracle.build_ref import load_ref
    ref = load_ref()[0]
    vv = np.linspace(0.5, 1.5, 1000); tt = np.full(1000, 0.3)
    print("reference Cython on this host:")
    print("  wiener_like_rlddm 1000       %.1f us" % per_call(lambda: ref.wiener_like_rlddm(x1, resp, fb, split, 0.5, 0.1, 100.0, 1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 1e-4, p_outlier=0.05, w_outlier=0.1), 50))
    print("  wiener_like_rl 1000          %.1f us" % per_call(lambda: ref.wiener_like_rl(resp, fb, split, 0.5, 0.1, 100.0, 1.0, 0.5), 50))
    print("  wiener_like_contaminant 1000 %.1f us" % per_call(lambda: ref.wiener_like_contaminant(x1, np.zeros(n, dtype=np.int32), 1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 0.1, 5.0, 1e-4), 50))
    print("  wiener_like_multi 1000 (v,t) %.1f us" % per_call(lambda: ref.wiener_like_multi(x1, vv, 0, 2, 0.5, 0, tt, 0, 1e-4, multi=["v", "t"], p_outlier=0.05, w_outlier=0.1), 50))
    print("  pdf_array 1000               %.1f us" % per_call(lambda: ref.pdf_array(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4), 50))
except Exception as exc:
    print("reference unavailable:", exc)
