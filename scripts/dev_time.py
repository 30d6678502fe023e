"""Developer timing probe (not the bench contract): device-resident batched likelihood, CUDA-event timed."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from hddm_b200 import _lib, wfpt

def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "wfpt_golden.npz")))
    x2 = g["x_cfg2"]
    rng = np.random.default_rng(4)
    prm = np.stack([rng.uniform(-5, 5, P), rng.uniform(0, 4, P), rng.uniform(1, 3, P), rng.uniform(0.3, 0.7, P),
                    rng.uniform(0, 0.25, P), rng.uniform(0.1, 0.3, P), rng.uniform(0, 0.2, P), np.full(P, 0.05)], axis=1)
    simple = prm.copy(); simple[:, [1, 4, 6]] = 0
    print(_lib.device_info())
    pk = _lib.fp64_peak(16384)
    print("fp64 probe: %.3e DFMA/s = %.2f TFLOP/s (%.2f ms)" % (pk["dfma_per_s"], 2e-12 * pk["dfma_per_s"], pk["ms"]))
    dx = torch.from_numpy(x2).cuda()
    dxs = torch.from_numpy(x2[np.argsort(np.abs(x2))]).cuda()
    for name, pp, ad in (("simple", simple, 0), ("full-fixed9", prm, 0), ("full-adaptive25", prm, 1)):
        dp = torch.from_numpy(pp).cuda()
        for var in (0, 1):
            for xs, tag in ((dx, "unsorted"), (dxs, "sorted")):
                for _ in range(2):
                    out = wfpt.wiener_like_batch(xs, dp, 1e-4, 2, 2, ad, 1e-3, 0.1, variant=var)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                reps = 3
                for _ in range(reps):
                    out = wfpt.wiener_like_batch(xs, dp, 1e-4, 2, 2, ad, 1e-3, 0.1, variant=var)
                e1.record(); torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                print(f"{name:16s} variant={var} {tag:8s} P={P} n={len(x2)}  {ms:9.3f} ms  {P*len(x2)/ms*1e3:.3e} evals/s", flush=True)

main()
