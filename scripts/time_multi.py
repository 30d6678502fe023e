"""Device-resident timing of wfpt_like_multi (per-trial v, a, t; sz, st above the thresholds; the reference's defaults of
wiener_like_multi: adaptive, depth 10, simps_err 1e-3) through the C ABI, both variants."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from hddm_b200 import _lib  # noqa: E402


def time_multi(nm=1_000_000, sz=0.1, st=0.1, variant=0, opts=(1e-4, 10, 10, 1, 1e-3), reps=3):
    L = _lib.load()
    rng = np.random.default_rng(6)
    xm = np.where(rng.random(nm) < 0.8, 1.0, -1.0) * rng.uniform(0.5, 3.0, nm)
    cols = [rng.normal(1, 0.3, nm), None, rng.uniform(1.5, 2.5, nm), None, None, rng.uniform(0.1, 0.3, nm), None]
    dx = torch.from_numpy(xm).cuda()
    dcols = [torch.from_numpy(c).cuda() if c is not None else None for c in cols]
    ptrs = (C.c_void_p * 7)(*[c.data_ptr() if c is not None else None for c in dcols])
    scal = (C.c_double * 7)(0.0, 0.5, 0.0, 0.5, sz, 0.0, st)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    st_ = torch.cuda.current_stream().cuda_stream
    call = lambda: _lib.check(L.wfpt_like_multi(dx.data_ptr(), nm, ptrs, scal, *opts, 0.05, 0.1, variant, out.data_ptr(), st_))
    call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        call()
    e1.record()
    torch.cuda.synchronize()
    return nm / (e0.elapsed_time(e1) / reps * 1e-3), float(out.item())


if __name__ == "__main__":
    for tag, sz, st in (("sz, st = 0.1", 0.1, 0.1), ("sz = st = 0", 0.0, 0.0)):
        for var in (0, 1):
            r, val = time_multi(sz=sz, st=st, variant=var)
            print(f"{tag}, variant {var}: {r:.4g} evals/s  (sum {val:.6f})")
