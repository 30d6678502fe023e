import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from bench import workload
from hddm_b200 import wfpt
from oracle import cpu as oracle
x, params = workload(256)
rng = np.random.default_rng(3)
nodes = [np.ascontiguousarray(rng.permutation(x)[:500]) for _ in range(60)]
rows = []
for k in range(60):
    dx = torch.from_numpy(wfpt.sort_rts(nodes[k])).cuda(); dp = torch.from_numpy(params[k:k + 1].copy()).cuda()
    f = lambda: wfpt.wiener_like_batch(dx, dp, 1e-4, 2, 2, 1, 1e-3, 0.1)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    oracle.nodes_reset()
    p = params[k]
    oracle.wiener_like(nodes[k], *p[:7], 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=p[7], w_outlier=0.1)
    per = oracle.nodes_get() / 500
    # count refined trials
    cnt = 0
    for xv in nodes[k][:200]:
        oracle.nodes_reset(); oracle.full_pdf(xv, *p[:7], 1e-4, 2, 2, 1, 1e-3); cnt += oracle.nodes_get() > 25
    rows.append((us, k, per, cnt / 200, np.round(p[:7], 3)))
rows.sort(key=lambda r: -r[0])
for r in rows[:12]: print("%.1f us set %d nodes/trial %.1f refined frac %.2f" % r[:4], r[4])
print("median us", np.median([r[0] for r in rows]))
