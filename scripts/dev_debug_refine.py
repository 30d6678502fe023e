import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import cpu as oracle
from hddm_b200 import wfpt
rng = np.random.default_rng(33)
P, n = 96, 700
prm = np.stack([rng.uniform(-5, 5, P), rng.uniform(0, 4, P), rng.uniform(0.1, 5, P), rng.uniform(0.3, 0.7, P),
                rng.uniform(0, 0.25, P), rng.uniform(0.25, 2.25, P), rng.uniform(0, 0.25, P), np.full(P, 0.05)], axis=1)
prm[::7, 4] = 0.0
prm[3::7, 6] = 0.0
x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.3, 3.0, n)
for k, row in enumerate(prm[:40]):
    want = oracle.pdf_array(x, *row[:7], 1e-4, 0, 2, 2, 1, 1e-3, 0.0, 0.0)
    got = wfpt.pdf_array(x, *row[:7], 1e-4, 0, 2, 2, 1, 1e-3, 0.0, 0.0)
    wfpt.set_variant(wfpt.FAITHFUL); gf = wfpt.pdf_array(x, *row[:7], 1e-4, 0, 2, 2, 1, 1e-3, 0.0, 0.0); wfpt.set_variant(wfpt.FAST)
    nz = want != 0
    rel = np.zeros(n); rel[nz] = np.abs(got[nz] - want[nz]) / np.abs(want[nz])
    bad = np.flatnonzero(rel > 1e-9)
    for i in bad[:3]:
        oracle.nodes_reset(); oracle.full_pdf(x[i], *row[:7], 1e-4, 2, 2, 1, 1e-3); nn = oracle.nodes_get()
        print(k, i, repr(x[i]), [repr(float(v)) for v in row[:7]], "want %r got %r faithful %r rel %.2e nodes %d" % (want[i], got[i], gf[i], rel[i], nn))
