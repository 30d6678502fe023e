"""Quick timing of the headline launch (65,536 benign sets x 10,000 RTs, fixed 3x3 grid) with a parity check against the
reference fixture; for kernel experiments.   python scripts/time_fixed.py [adaptive]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from hddm_b200 import wfpt  # noqa: E402

g = bench.load_grid_golden()
x, benign = bench.workload(bench.P_FULL)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()
runs = [("benign_fixed", benign, bench.OPTS)]
if "adaptive" in sys.argv:
    runs = [("benign_adaptive", benign[:8192], bench.ADAPT), ("bounds_adaptive", bench.param_bounds_grid(8192), bench.ADAPT),
            ("bounds_fixed", bench.param_bounds_grid(8192), bench.OPTS)]
for name, prm, opts in runs:
    dp = torch.from_numpy(np.ascontiguousarray(prm)).cuda()
    out = wfpt.wiener_like_batch(dx, dp, **opts)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(4):
        e0.record()
        out = wfpt.wiener_like_batch(dx, dp, **opts)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    rel, mism = bench.compare(out.cpu().numpy(), g[name])
    print(f"{name}: {best:.3f} ms  {len(prm) * len(x) / best * 1e-6:.4g} Gevals/s  max_rel_err {rel:.2e} mismatches {mism}", flush=True)
