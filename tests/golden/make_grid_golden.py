"""Reference answers for the WHOLE benchmark grids (run where oracle/_ref has been built, i.e. with /root/reference):

    python tests/golden/make_grid_golden.py

Writes tests/golden/grid_golden.npz with the outputs of the reference's own Cython `wiener_like`
(/root/reference/src/wfpt.pyx:54-76, compiled unmodified into oracle/_ref) on the cfg-2 RTs for

  benign_fixed      float64[65536]  bench.workload(65536), fixed 3x3 Simpson            -- the headline grid
  benign_adaptive   float64[8192]   first 8,192 of the same sets, HDDM-default adaptive integration
  bounds_fixed      float64[8192]   bench.param_bounds_grid(8192) (whole param_bounds box), fixed grid
  bounds_adaptive   float64[8192]   same sets, adaptive
  special_params    float64[10, 8]  hand-picked sets on the same RTs: outside the support with and without the outlier mix
                                    (N*log(w*p) / -inf), t above some RTs, p_outlier out of range, a == 0 (NaN), simple
                                    DDM, sz/st under the 1e-3 thresholds, one-axis integration
  special_fixed / special_adaptive  float64[10]  their answers in both integration modes
  *_sha             the SHA-256 of the parameter matrix each vector belongs to (the grids are regenerated from their
                    seeds by bench.py and the tests; the digest proves they are the same numbers)

bench.py compares every set of every step's output with these at every GPU count; tests/test_gpu_parity.py does the
same through the C ABI.  About three minutes on eight cores.
"""
from __future__ import annotations

import hashlib
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from oracle.build_ref import build_ref, load_ref  # noqa: E402

_W = {}


def _init():
    _W["like"] = load_ref()[0].wiener_like
    _W["x"] = bench.workload(1)[0]


def _chunk(job):
    rows, adaptive = job
    like, x = _W["like"], _W["x"]
    out = np.empty(len(rows))
    for i, q in enumerate(rows):
        out[i] = like(x, q[0], q[1], q[2], q[3], q[4], q[5], q[6], 1e-4, n_st=2, n_sz=2, use_adaptive=adaptive,
                      simps_err=1e-3, p_outlier=q[7], w_outlier=0.1)
    return out


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def special_params():
    # v, sv, a, z, sz, t, st, p_outlier
    return np.array([
        [1.0, 0.0, 2.0, 1.5, 0.0, 0.3, 0.0, 0.05],      # z > 1 under the HDDM outlier mix: N * log(w_outlier * p_outlier)
        [1.0, 0.0, 2.0, 1.5, 0.0, 0.3, 0.0, 0.0],       # the same without the mix: -inf
        [1.0, 1.0, 2.0, 0.5, 0.2, 0.6, 0.1, 0.0],       # t - st/2 above some |rt|: zero density there, -inf
        [1.0, 1.0, 2.0, 0.5, 0.2, 0.6, 0.1, 0.05],      # the same with the mix: finite
        [1.0, 1.0, 2.0, 0.5, 0.2, 0.3, 0.1, 1.5],       # p_outlier outside [0, 1]: -inf
        [1.0, 1.0, 0.0, 0.5, 0.2, 0.3, 0.1, 0.05],      # a == 0 passes the validity test (pdf.pxi:111) and yields NaN
        [1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 0.05],      # simple DDM
        [0.5, 0.7, 1.5, 0.45, 9e-4, 0.25, 9e-4, 0.05],  # sz, st under the 1e-3 thresholds: zeroed after the validity test
        [-0.8, 1.2, 2.2, 0.55, 0.0, 0.2, 0.15, 0.05],   # t-axis integration only
        [-0.8, 1.2, 2.2, 0.55, 0.15, 0.2, 0.0, 0.05],   # z-axis integration only
    ])


def main():
    assert build_ref(), "reference build failed"
    cores = os.cpu_count() or 1
    pool = mp.get_context("fork").Pool(cores, initializer=_init)

    def run(params, adaptive):
        t0 = time.time()
        jobs = [(c, adaptive) for c in np.array_split(params, cores * 16) if len(c)]
        res = np.concatenate(pool.map(_chunk, jobs, chunksize=1))
        print(f"{len(params)} sets, adaptive={adaptive}: {time.time() - t0:.1f} s", flush=True)
        return res

    if "--only-special" in sys.argv:   # keep the grids of an existing fixture, (re)compute the hand-picked sets only
        out = dict(np.load(os.path.join(HERE, "grid_golden.npz")))
        sp = special_params()
        out.update(special_params=sp, special_fixed=run(sp, 0), special_adaptive=run(sp, 1))
        pool.close()
        np.savez_compressed(os.path.join(HERE, "grid_golden.npz"), **out)
        print(out["special_fixed"], out["special_adaptive"])
        return
    _, benign = bench.workload(bench.P_FULL)
    bounds = bench.param_bounds_grid(8192)
    out = {
        "benign_fixed": run(benign, 0), "benign_sha": sha(benign),
        "benign_adaptive": run(benign[:8192], 1), "benign8192_sha": sha(benign[:8192]),
        "bounds_fixed": run(bounds, 0), "bounds_adaptive": run(bounds, 1), "bounds_sha": sha(bounds),
    }
    sp = special_params()
    out.update(special_params=sp, special_fixed=run(sp, 0), special_adaptive=run(sp, 1))
    pool.close()
    np.savez_compressed(os.path.join(HERE, "grid_golden.npz"), **out)
    for k, v in out.items():
        print(k, v.shape, v.dtype)


if __name__ == "__main__":
    main()
