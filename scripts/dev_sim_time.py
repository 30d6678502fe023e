"""Developer probe: simulator throughput (particles/s, Euler steps/s), histogram mode, device timed via wall clock."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from hddm_b200 import simulators as S
rng = np.random.default_rng(5)
n_sets = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
theta = np.stack([rng.uniform(-3, 3, n_sets), rng.uniform(.3, 3, n_sets), rng.uniform(.1, .9, n_sets), rng.uniform(1e-3, 2, n_sets), rng.uniform(-.1, 1.3, n_sets)], 1).astype(np.float32)
for model, th in (("angle", theta), ("ddm", theta[:, :4]), ("ddm_hddm_base", theta[:, :4]), ("weibull", np.concatenate([theta[:, :4], rng.uniform(.31, 4.99, (n_sets, 1)), rng.uniform(.31, 6.99, (n_sets, 1))], 1).astype(np.float32))):
    S.simulate_histograms(th[:4], model=model, n_samples=1000, random_state=1)
    t0 = time.perf_counter()
    h, steps = S.simulate_histograms(th, model=model, n_samples=n, nbins=512, random_state=5)
    dt = time.perf_counter() - t0
    print(f"{model:14s} sets={n_sets} n={n}: {dt*1e3:8.1f} ms  {n_sets*n/dt:.3e} particles/s  {steps/dt:.3e} steps/s  mean steps {steps/(n_sets*n):.0f}", flush=True)
