import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import workload
from hddm_b200 import wfpt
Ps = 8192
x, prm = workload(Ps)
hr = np.random.default_rng(4)
hostile = np.stack([hr.uniform(-5, 5, Ps), hr.uniform(0, 4, Ps), hr.uniform(0.1, 5, Ps), hr.uniform(0.3, 0.7, Ps),
                    hr.uniform(0, 0.25, Ps), hr.uniform(0.25, 2.25, Ps), hr.uniform(0, 0.25, Ps), np.full(Ps, 0.05)], axis=1)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()
def timed(dp, **kw):
    for _ in range(2): wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, w_outlier=0.1, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, w_outlier=0.1, **kw)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 3
dp = torch.from_numpy(np.ascontiguousarray(hostile)).cuda()
for se in (1e3, 1e-1, 1e-2, 1e-3):
    print(f"hostile simps_err={se:g}: {timed(dp, use_adaptive=1, simps_err=se):.2f} ms", flush=True)
# which parameter regions are slow: time subsets sorted by t
order = np.argsort(hostile[:, 5])
for lo in range(0, Ps, 1024):
    sub = torch.from_numpy(np.ascontiguousarray(hostile[order[lo:lo + 1024]])).cuda()
    ms = timed(sub, use_adaptive=1, simps_err=1e-3)
    ms0 = timed(sub, use_adaptive=1, simps_err=1e3)
    tt = hostile[order[lo:lo + 1024], 5]
    print(f"t in [{tt.min():.2f},{tt.max():.2f}]: {ms:.2f} ms (no-refine {ms0:.2f} ms)", flush=True)
