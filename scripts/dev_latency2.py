import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hddm_b200 import _lib, wfpt
g = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "wfpt_golden.npz"))
x1 = np.ascontiguousarray(g["x_cfg1"]); x2 = np.ascontiguousarray(g["x_cfg2"])
L = _lib.load()
for name, x, prm, opts in (("cfg1", x1, [1, 0, 2, .5, 0, .3, 0, 0.0], (1e-4, 10, 10, 1, 1e-8)),
                           ("cfg2", x2, [1, 1, 2, .5, .2, .3, .1, 0.05], (1e-4, 2, 2, 1, 1e-3))):
    dx = torch.from_numpy(x).cuda(); dp = torch.tensor([prm], dtype=torch.float64).cuda(); out = torch.empty(1, dtype=torch.float64).cuda()
    s = torch.cuda.current_stream().cuda_stream
    call = lambda: L.wfpt_like_batch(dx.data_ptr(), x.shape[0], dp.data_ptr(), 1, *opts, 0.1, 0, out.data_ptr(), s)
    for _ in range(20): call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200): call()
    e1.record(); torch.cuda.synchronize()
    print(name, "device time per call (back-to-back launches): %.2f us" % (e0.elapsed_time(e1) / 200 * 1e3))
    t0 = time.perf_counter()
    for _ in range(200):
        call(); torch.cuda.synchronize()
    print(name, "launch + sync wall: %.2f us" % ((time.perf_counter() - t0) / 200 * 1e6))
    t0 = time.perf_counter()
    for _ in range(200): call()
    t1 = time.perf_counter(); torch.cuda.synchronize()
    print(name, "enqueue only: %.2f us" % ((t1 - t0) / 200 * 1e6))
    res = np.empty(1); pr = np.array([prm], dtype=np.double)
    h = lambda: L.wfpt_like_batch_host(x.ctypes.data, x.shape[0], pr.ctypes.data, 1, *opts, 0.1, 0, res.ctypes.data)
    for _ in range(20): h()
    t0 = time.perf_counter()
    for _ in range(300): h()
    print(name, "host call: %.2f us" % ((time.perf_counter() - t0) / 300 * 1e6))
