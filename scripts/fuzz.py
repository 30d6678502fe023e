"""Dev: random single calls / small batches against the oracle over option combinations (new launch plans and refinement
routines), reporting the worst disagreement."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import cpu as oracle
from hddm_b200 import wfpt
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
worst = 0.0; bad = 0; cases = 0
for it in range(400):
    n = int(rng.choice([1, 3, 17, 128, 129, 300, 1000, 2500]))
    x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.25, 3.5, n)
    P = int(rng.choice([1, 1, 2, 7, 33]))
    prm = np.stack([rng.uniform(-4, 4, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 3, P), rng.uniform(0.4, 4, P),
                    rng.uniform(0.25, 0.75, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                    rng.uniform(0.05, 1.5, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                    rng.choice([0.0, 0.05], P)], axis=1)
    ad = int(rng.integers(0, 2))
    n_st, n_sz = (int(rng.choice([2, 4, 6])), int(rng.choice([2, 4]))) if not ad else (int(rng.integers(0, 5)), int(rng.integers(0, 5)))
    se = float(rng.choice([1e-2, 1e-3, 1e-5, 1e-8]))
    err = float(rng.choice([1e-3, 1e-4, 1e-6]))
    want = oracle.wiener_like_batch(x, prm, err, n_st, n_sz, ad, se, 0.1)
    got = wfpt.wiener_like_batch(x, prm, err, n_st, n_sz, ad, se, 0.1, presort=bool(rng.integers(0, 2)))
    cases += P
    same = np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(np.isinf(got), np.isinf(want))
    fin = np.isfinite(want) & np.isfinite(got)
    rel = np.max(np.abs(got[fin] - want[fin]) / np.maximum(np.abs(want[fin]), 1e-300)) if fin.any() else 0.0
    if not same or rel > 1e-10:
        bad += 1
        print("MISMATCH", dict(n=n, P=P, ad=ad, n_st=n_st, n_sz=n_sz, se=se, err=err), "rel", rel, "same-nonfinite", same)
        k = int(np.argmax(np.where(fin, np.abs(got - want) / np.maximum(np.abs(want), 1e-300), 0))) if fin.any() else 0
        print("   row", prm[k].tolist(), got[k], want[k])
    worst = max(worst, rel)
print(f"{cases} parameter sets in 400 calls, mismatches {bad}, worst relative error {worst:.2e}")
