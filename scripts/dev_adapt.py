import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hddm_b200 import wfpt
from bench import workload
x, prm = workload(16384)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda(); dp = torch.from_numpy(prm).cuda()
for se in (1e3, 1e-1, 1e-2, 1e-3, 1e-4):
    for _ in range(2): wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, 1, se, 0.1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, 1, se, 0.1)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"simps_err={se:g}: {ms:.2f} ms  {len(x)*len(prm)/ms*1e3:.3e} evals/s")
