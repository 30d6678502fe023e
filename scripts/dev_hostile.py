"""Dev timing: adaptive (HDDM-default) likelihood on the benign grid and on the full param_bounds box of config 4."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import workload
from hddm_b200 import wfpt

Ps = 8192
x, prm = workload(Ps)
hr = np.random.default_rng(4)
hostile = np.stack([hr.uniform(-5, 5, Ps), hr.uniform(0, 4, Ps), hr.uniform(0.1, 5, Ps), hr.uniform(0.3, 0.7, Ps),
                    hr.uniform(0, 0.25, Ps), hr.uniform(0.25, 2.25, Ps), hr.uniform(0, 0.25, Ps), np.full(Ps, 0.05)], axis=1)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()


def timed(dp, **kw):
    for _ in range(2):
        wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, w_outlier=0.1, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, w_outlier=0.1, **kw)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 3


for name, p in (("benign", prm), ("hostile", hostile)):
    dp = torch.from_numpy(np.ascontiguousarray(p)).cuda()
    for ua, se in ((0, 1e-3), (1, 1e-3), (1, 1e-5)):
        ms = timed(dp, use_adaptive=ua, simps_err=se)
        print(f"{name:8s} adaptive={ua} simps_err={se:g}: {ms:8.2f} ms  {len(x) * Ps / ms * 1e3:.3e} evals/s", flush=True)
# parity of the refinement path against the faithful variant on a hostile subsample
sub = torch.from_numpy(np.ascontiguousarray(hostile[:512])).cuda()
for se in (1e-3, 1e-5):
    a = wfpt.wiener_like_batch(dx, sub, 1e-4, 2, 2, 1, se, 0.1).cpu().numpy()
    b = wfpt.wiener_like_batch(dx, sub, 1e-4, 2, 2, 1, se, 0.1, variant=wfpt.FAITHFUL).cpu().numpy()
    fin = np.isfinite(a) & np.isfinite(b)
    print(f"simps_err={se:g}: same non-finite pattern {np.array_equal(np.isfinite(a), np.isfinite(b))}, "
          f"max rel diff {np.max(np.abs(a[fin] - b[fin]) / np.abs(b[fin])):.3e}")
