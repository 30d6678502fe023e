"""Where the end-to-end step of one rank at N = 8 goes: the host-buffer call on 8,192 sets against the kernel alone."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from hddm_b200 import wfpt
x, benign = bench.workload(bench.P_FULL)
my = np.ascontiguousarray(benign[0::8])
hx = torch.from_numpy(x).pin_memory().numpy(); hp = torch.from_numpy(my).pin_memory().numpy()
hout_t = torch.empty(len(my), dtype=torch.float64).pin_memory(); hout = hout_t.numpy()
dx = torch.from_numpy(wfpt.sort_rts(x)).cuda(); dp = torch.from_numpy(my).cuda(); dout = torch.empty(len(my), dtype=torch.float64, device="cuda")
for _ in range(3):
    wfpt.wiener_like_batch(hx, hp, out=hout, **bench.OPTS); wfpt.wiener_like_batch(dx, dp, out=dout, **bench.OPTS)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): wfpt.wiener_like_batch(hx, hp, out=hout, **bench.OPTS)
t_host = (time.perf_counter() - t0) / 20
t0 = time.perf_counter()
for _ in range(20):
    wfpt.wiener_like_batch(dx, dp, out=dout, **bench.OPTS)
torch.cuda.synchronize()
t_dev = (time.perf_counter() - t0) / 20
print(f"host-buffer call {t_host*1e3:.3f} ms, device-resident {t_dev*1e3:.3f} ms, difference {1e6*(t_host-t_dev):.0f} us")
