import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import workload
from hddm_b200 import wfpt
x, prm = workload(4096)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda(); dp = torch.from_numpy(prm).cuda()
for _ in range(3): out = wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, 1, 1e-3, 0.1)
torch.cuda.synchronize(); print(float(out.sum()))
