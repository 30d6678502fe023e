"""Timing of one rank's share of the headline grid at N GPUs (sets r::N of the 65,536), to look at the last-wave effect."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from hddm_b200 import wfpt
g = bench.load_grid_golden()
x, benign = bench.workload(bench.P_FULL)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()
for N in (8, 4, 2):
    prm = np.ascontiguousarray(benign[0::N])
    dp = torch.from_numpy(prm).cuda()
    out = wfpt.wiener_like_batch(dx, dp, **bench.OPTS); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(5):
        e0.record(); out = wfpt.wiener_like_batch(dx, dp, **bench.OPTS); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    rel, mism = bench.compare(out.cpu().numpy(), g["benign_fixed"][0::N])
    print(f"N={N}: {len(prm)} sets {best:.3f} ms -> x{N} = {best*N:.3f} ms-equivalents; rel {rel:.1e} mism {mism}", flush=True)
