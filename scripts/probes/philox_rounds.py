"""Philox4x32-7 against -10 (run with WFPT_B200_LIB pointing at a -7 build to compare): throughput of the config-5 launch and
the first moments / lag-1 correlation of what the streams produce (angle model, raw output)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench
from hddm_b200 import simulators
theta = torch.from_numpy(bench.cfg5_theta()).cuda()
n = 1_000_000
for _ in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); h, s = simulators.simulate_histograms_device(theta, model="angle", n_samples=n, nbins=512, random_state=5); e1.record()
    torch.cuda.synchronize()
t = e0.elapsed_time(e1) * 1e-3
print("euler steps/s", int(s.item()) / t, "particles/s", 1024 * n / t)
rts, ch, _ = simulators.simulator(np.array([0.0, 1.0, 0.5, 0.0, 0.0], dtype=np.float32), model="angle", n_samples=2_000_000, random_state=11)
x = (rts.reshape(-1) * ch.reshape(-1)).astype(np.float64)
print("P(upper)", float(np.mean(ch == 1)), "mean|rt|", float(np.abs(x).mean()), "var", float(np.abs(x).var()),
      "lag-1 corr of consecutive particles", float(np.corrcoef(np.abs(x[:-1]), np.abs(x[1:]))[0, 1]))
