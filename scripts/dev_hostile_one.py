import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import workload
from hddm_b200 import wfpt
Ps = 2048
x, prm = workload(Ps)
hr = np.random.default_rng(4)
hostile = np.stack([hr.uniform(-5, 5, Ps), hr.uniform(0, 4, Ps), hr.uniform(0.1, 5, Ps), hr.uniform(0.3, 0.7, Ps),
                    hr.uniform(0, 0.25, Ps), hr.uniform(0.25, 2.25, Ps), hr.uniform(0, 0.25, Ps), np.full(Ps, 0.05)], axis=1)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()
dp = torch.from_numpy(np.ascontiguousarray(hostile)).cuda()
for _ in range(3):
    out = wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, 1, 1e-3, 0.1)
torch.cuda.synchronize()
print(float(out[np.isfinite(out.cpu().numpy())].sum()))
