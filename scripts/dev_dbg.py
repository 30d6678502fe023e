import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from hddm_b200 import wfpt
g = bench.load_grid_golden()
x, _ = bench.workload(1)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()
prm = bench.param_bounds_grid(8192)
dp = torch.from_numpy(prm).cuda()
got = wfpt.wiener_like_batch(dx, dp, **bench.ADAPT).cpu().numpy()
want = g["bounds_adaptive"]
fin = np.isfinite(want)
rel = np.where(fin, np.abs(got - want) / np.abs(np.where(fin, want, 1)), 0)
idx = np.argsort(-rel)[:10]
ax = np.abs(x)
for j in idx:
    q = prm[j]
    insup = int(np.sum(ax >= q[5] - q[6] / 2)); inwin = int(np.sum((ax >= q[5] - q[6] / 2) & (ax <= q[5] + q[6] / 2)))
    print(j, rel[j], got[j] - want[j], "in support", insup, "in t window", inwin, [float(f"{v:.3g}") for v in q])
print("sets off:", int(np.sum(rel > 1e-10)))
