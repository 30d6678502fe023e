"""One launch shape per invocation, for ncu captures (scripts/profile_all.sh): prints the units of work of ONE launch.
    python scripts/prof_target.py fixed | fixed_bounds | adaptive_benign | adaptive_bounds | sim_angle"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from hddm_b200 import simulators, wfpt  # noqa: E402

what = sys.argv[1]
if what == "sim_angle":
    theta = torch.from_numpy(bench.cfg5_theta()).cuda()
    for _ in range(3):
        h, s = simulators.simulate_histograms_device(theta, model="angle", n_samples=100_000, nbins=512, random_state=5)
    torch.cuda.synchronize()
    print(int(s.item()))          # Euler steps of one launch
    sys.exit(0)
x, benign = bench.workload(bench.P_FULL)
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda()
prm, opts = {"fixed": (benign, bench.OPTS), "adaptive_benign": (benign[:8192], bench.ADAPT),
             "adaptive_bounds": (bench.param_bounds_grid(8192), bench.ADAPT),
             "fixed_bounds": (bench.param_bounds_grid(8192), bench.OPTS)}[what]
dp = torch.from_numpy(np.ascontiguousarray(prm)).cuda()
for _ in range(4):
    out = wfpt.wiener_like_batch(dx, dp, **opts)
torch.cuda.synchronize()
print(len(prm) * len(x))          # evals of one launch
