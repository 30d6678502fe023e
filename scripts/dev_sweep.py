import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import workload
from hddm_b200 import wfpt
x, params = workload(256)
rng = np.random.default_rng(3)
for nn, ntr in ((60, 500), (60, 100), (12, 2000), (300, 500)):
    nodes = [np.ascontiguousarray(rng.permutation(x)[:ntr]) for _ in range(nn)]
    prm = params[:nn].copy() if nn <= 256 else np.tile(params, (2, 1))[:nn].copy()
    rd = wfpt.ResidentData(nodes)
    dprm = torch.from_numpy(prm).cuda()
    kw = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=float(os.environ.get("SE", "1e-3")), w_outlier=0.1)
    for _ in range(5): rd.logp(dprm, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): rd.logp(dprm, **kw)
    e1.record(); torch.cuda.synchronize()
    dev = e0.elapsed_time(e1) / 50 * 1e3
    t0 = time.perf_counter()
    for _ in range(50): rd.logp(prm, **kw)
    e2e = (time.perf_counter() - t0) / 50 * 1e6
    ref = np.array([wfpt.wiener_like(nodes[i], *prm[i, :7], 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=float(os.environ.get("SE", "1e-3")), p_outlier=prm[i, 7], w_outlier=0.1) for i in range(nn)])
    got = rd.logp(prm, **kw)
    print(f"{nn} nodes x {ntr} trials: device {dev:.1f} us, e2e {e2e:.1f} us, max rel diff vs per-node calls {np.max(np.abs(got - ref) / np.abs(ref)):.2e}")
