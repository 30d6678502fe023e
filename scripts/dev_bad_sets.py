import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from hddm_b200 import wfpt
from oracle import cpu as oracle
rng = np.random.default_rng(21)
n, P = 3001, 257
x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.3, 4.0, n)
prm = np.stack([rng.uniform(-4, 4, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 3, P), rng.uniform(0.5, 4, P),
                rng.uniform(0.2, 0.8, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                rng.uniform(0.05, 0.5, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                rng.choice([0.0, 0.05, 1.0, 1.5], P)], axis=1)
want = oracle.wiener_like_batch(x, prm, 1e-4, 2, 2, 1, 1e-3, 0.1)
got = wfpt.wiener_like_batch(x, prm, 1e-4, 2, 2, 1, 1e-3, 0.1)
fin = np.isfinite(want)
rel = np.where(fin, np.abs(got - want) / np.abs(np.where(fin, want, 1)), 0)
for j in np.argsort(-rel)[:8]:
    print(j, rel[j], got[j], want[j], [float(f"{v:.4g}") for v in prm[j]])
j = int(np.argmax(rel))
q = prm[j]
d_want = oracle.pdf_array(x, *q[:7], 1e-4, 0, 2, 2, 1, 1e-3, 0.0, 0.0)
# per-trial through the batch kernel: one trial at a time is slow; use chunks of single trials via P=1 calls on 1-element arrays
bad = []
for i in range(0, n):
    g1 = wfpt.wiener_like_batch(x[i:i+1], q[None, :] * np.array([1,1,1,1,1,1,1,0.0]) , 1e-4, 2, 2, 1, 1e-3, 0.1)[0]
    w1 = np.log(d_want[i]) if d_want[i] > 0 else -np.inf
    if not (g1 == w1 or abs(g1 - w1) <= 1e-9 * abs(w1)):
        bad.append((i, x[i], g1, w1))
print(len(bad), bad[:10])
