"""Generate tests/golden/*.npz from the reference itself (run where /root/reference exists).

    python tests/golden/make_golden.py

Sources of the numbers written here:
  * `matlab`      : the 20 known-answer tuples of /root/reference/hddm/tests/matlab_values.py:3-204
                    (v, t, a, z, z_nonorm, rt, err, matlab_wfpt), read by importing that file in place.
  * everything else: outputs of the reference's own Cython (src/wfpt.pyx, src/pdf.pxi, src/integrate.pxi) compiled
                    UNMODIFIED into oracle/_ref by oracle/build_ref.py, evaluated on seeded inputs.
The fixtures travel to the GPU box (where /root/reference does not exist); tests compare both the C oracle and
the CUDA path with them.
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.build_ref import build_ref, load_ref  # noqa: E402


def bounds_params(rng, P, benign=False):
    """cfg-4 parameter sets: uniform in model_config['full_ddm_hddm_base']['param_bounds']
    (/root/reference/hddm/model_config.py:31-34); columns v,sv,a,z,sz,t,st,p_outlier."""
    v = rng.uniform(-5, 5, P)
    a = rng.uniform(1, 3, P) if benign else rng.uniform(0.1, 5, P)
    z = rng.uniform(0.3, 0.7, P)
    t = rng.uniform(0.1, 0.3, P) if benign else rng.uniform(0.25, 2.25, P)
    sz = rng.uniform(0, 0.25, P)
    sv = rng.uniform(0, 4, P)
    st = rng.uniform(0, 0.25, P)
    if benign:
        st = np.minimum(st, 2 * t)  # keep t - st/2 >= 0
    po = np.full(P, 0.05)
    return np.stack([v, sv, a, z, sz, t, st, po], axis=1)


def main():
    assert build_ref(), "reference build failed"
    wfpt, _ = load_ref()
    out = {}

    spec = importlib.util.spec_from_file_location("matlab_values", "/root/reference/hddm/tests/matlab_values.py")
    mv = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mv)
    out["matlab"] = np.array(mv.vals, dtype=np.float64)

    # --- benchmark inputs ---------------------------------------------------------------------------
    np.random.seed(1)
    x1 = wfpt.gen_rts_from_cdf(1, 0, 2, 0.5, 0, 0.3, 0, 1000, -6, 6, 1e-3)
    np.random.seed(2)
    x2 = wfpt.gen_rts_from_cdf(1, 1, 2, 0.5, 0.2, 0.3, 0.1, 10000, -6, 6, 1e-3)
    out["x_cfg1"] = x1
    out["x_cfg2"] = x2
    out["cfg1_logp"] = np.array(wfpt.wiener_like(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4))
    kw = dict(p_outlier=0.05, w_outlier=0.1)
    out["cfg2a_logp"] = np.array(wfpt.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=0, simps_err=1e-3, **kw))
    out["cfg2b_logp"] = np.array(wfpt.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, **kw))
    out["cfg2c_logp"] = np.array(wfpt.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8, **kw))
    out["cfg2b_pdf"] = wfpt.pdf_array(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, 0, 2, 2, 1, 1e-3, 0.05, 0.1)

    # --- cfg-4 slices: 48 hostile + 48 benign parameter sets on the cfg-2 RTs, fixed grid and HDDM-adaptive -----
    rng = np.random.default_rng(4)
    P4 = np.concatenate([bounds_params(rng, 48, False), bounds_params(rng, 48, True)], axis=0)
    out["cfg4_params"] = P4
    for tag, ad in (("fixed", 0), ("adaptive", 1)):
        res = np.empty(len(P4))
        for j, q in enumerate(P4):
            res[j] = wfpt.wiener_like(x2, q[0], q[1], q[2], q[3], q[4], q[5], q[6], 1e-4, n_st=2, n_sz=2,
                                      use_adaptive=ad, simps_err=1e-3, p_outlier=q[7], w_outlier=0.1)
        out["cfg4_" + tag] = res

    # --- random single-trial full_pdf cases, every dispatch branch ---------------------------------------------
    rng = np.random.default_rng(11)
    n = 6000
    cases = np.empty((n, 13))
    vals = np.empty(n)
    for i in range(n):
        v = rng.uniform(-5, 5)
        a = rng.uniform(0.1, 5)
        z = rng.uniform(0.05, 0.95)
        t = rng.uniform(0, 1)
        sv = 0.0 if rng.random() < 0.4 else rng.uniform(0, 4)
        sz = 0.0 if rng.random() < 0.4 else rng.uniform(0, 0.3)
        st = 0.0 if rng.random() < 0.4 else rng.uniform(0, 0.3)
        if rng.random() < 0.05:  # thresholds: values below 1e-3 are zeroed (pdf.pxi:122-125)
            st = rng.uniform(0, 2e-3)
            sz = rng.uniform(0, 2e-3)
        x = rng.uniform(-4, 4)
        err = 10 ** rng.uniform(-13, 0)
        ad = int(rng.integers(0, 2))
        nn = int(rng.choice([2, 4, 6, 10]))
        se = 10 ** rng.uniform(-9, -2)
        if ad and nn == 10 and se < 1e-6:
            se = 1e-6  # keep the generator fast
        cases[i] = (x, v, sv, a, z, sz, t, st, err, nn, nn, ad, se)
        vals[i] = wfpt.full_pdf(x, v, sv, a, z, sz, t, st, err, nn, nn, ad, se)
    out["full_pdf_cases"] = cases
    out["full_pdf_vals"] = vals

    # --- wiener_like on a short array: valid, invalid, outlier mixes, early exit -----------------------------------
    xs = np.array([0.8, -0.9, 1.3, -0.5, 0.45])
    wl_cases = []
    wl_vals = []
    rows = [
        # v, sv, a, z, sz, t, st, err, n_st, n_sz, adaptive, simps_err, p_outlier, w_outlier
        (1, 0, 2, 0.5, 0, 0.3, 0, 1e-4, 10, 10, 1, 1e-8, 0, 0.1),
        (1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, 2, 2, 1, 1e-3, 0.05, 0.1),
        (1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, 2, 2, 0, 1e-3, 0.05, 0.1),
        (1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, 10, 10, 1, 1e-8, 0, 0.1),
        (1, 0, 2, 1.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, 0.05, 0.1),   # z>1 invalid, outlier mix -> N*log(.005)
        (1, 0, 2, 1.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, 0, 0.1),      # invalid, no mix -> -inf
        (1, 0, 2, 0.5, 0, 0.6, 0, 1e-4, 2, 2, 1, 1e-3, 0, 0.1),      # |rt|<t for some -> -inf
        (1, 0, 2, 0.5, 0, 0.6, 0, 1e-4, 2, 2, 1, 1e-3, 0.05, 0.1),
        (1, 0, 0, 0.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, 0, 0.1),      # a=0 -> NaN
        (1, 0, 2, 0.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, 1.5, 0.1),    # p_outlier out of range -> -inf
        (1, 0, 2, 0.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, -0.1, 0.1),
        (0, 0.5, 1.5, 0.4, 0.3, 0.2, 0, 1e-6, 4, 4, 0, 1e-3, 0.1, 0.2),
        (-2, 0, 1.2, 0.6, 0, 0.25, 0.2, 1e-9, 6, 6, 1, 1e-5, 0, 0.1),
        (1, 0, 2, 0.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, 1.0, 0.1),    # p_outlier = 1
        (1, 0, 2, 0.5, 0, 0.3, 0, 1e-4, 2, 2, 1, 1e-3, 1.0, 0.0),    # p=0 for all -> -inf
    ]
    for r in rows:
        wl_cases.append(r)
        wl_vals.append(wfpt.wiener_like(xs, *r[:8], n_st=r[8], n_sz=r[9], use_adaptive=r[10], simps_err=r[11],
                                        p_outlier=r[12], w_outlier=r[13]))
    out["wl_x"] = xs
    out["wl_cases"] = np.array(wl_cases, dtype=np.float64)
    out["wl_vals"] = np.array(wl_vals)
    out["wl_rt0"] = np.array(wfpt.wiener_like(np.array([1.0, 2, 3, 0]), 1, 0, 2, 0.5, 0, 0, 0, 1e-4))  # test_likelihoods.py:122-127
    out["wl_empty"] = np.array(wfpt.wiener_like(np.empty(0), 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4))

    # --- wiener_like_multi: per-trial arrays, +-999 coding -----------------------------------------------------------
    rng = np.random.default_rng(12)
    n = 400
    xm = np.where(rng.random(n) < 0.8, 1, -1) * rng.uniform(0.35, 3.0, n)
    xm[5] = 999.0
    xm[17] = -999.0
    vm = rng.normal(1, 0.5, n)
    am = rng.uniform(1, 3, n)
    tm = rng.uniform(0.1, 0.3, n)
    zm = rng.uniform(0.3, 0.7, n)
    out["multi_x"] = xm
    out["multi_v"] = vm
    out["multi_a"] = am
    out["multi_t"] = tm
    out["multi_z"] = zm
    out["multi_v_only"] = np.array(wfpt.wiener_like_multi(xm, vm, 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.05, w_outlier=0.1))
    out["multi_vat"] = np.array(wfpt.wiener_like_multi(xm, vm, 1.0, am, 0.5, 0.1, tm, 0.1, 1e-4, ["v", "a", "t"], p_outlier=0.05, w_outlier=0.1))
    out["multi_vazt_fixed"] = np.array(wfpt.wiener_like_multi(xm, vm, 0.5, am, zm, 0.1, tm, 0.05, 1e-4, ["v", "a", "z", "t"], n_st=2, n_sz=2, use_adaptive=0, p_outlier=0.0, w_outlier=0.0))
    xm2 = xm.copy()
    xm2[3] = 0.05  # |rt| < t: density 0, log(0) flows through the sum without early exit (wfpt.pyx:385)
    out["multi_neginf"] = np.array(wfpt.wiener_like_multi(xm2, vm, 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.0, w_outlier=0.1))
    out["multi_small"] = np.array(wfpt.wiener_like_multi(xs, np.linspace(0.5, 1.5, 5), 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.05, w_outlier=0.1))
    out["multi_999"] = np.array(wfpt.wiener_like_multi(np.array([0.8, 999.0, -999.0]), np.ones(3), 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"]))

    # --- RL flavours -------------------------------------------------------------------------------------------------------
    rng = np.random.default_rng(13)
    n = 300
    split = rng.integers(0, 3, n).astype(np.int64)
    resp = rng.integers(0, 2, n).astype(np.int64)
    fb = rng.integers(0, 2, n).astype(np.float64)
    xr = np.where(resp == 1, 1, -1) * rng.uniform(0.4, 2.5, n)
    out["rl_x"] = xr
    out["rl_response"] = resp
    out["rl_feedback"] = fb
    out["rl_split"] = split
    out["rlddm_a"] = np.array(wfpt.wiener_like_rlddm(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.3, 0, 1e-4, p_outlier=0.05, w_outlier=0.1))
    out["rlddm_b"] = np.array(wfpt.wiener_like_rlddm(xr, resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.5, 1.8, 0.5, 0.1, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1))
    out["rlddm_neginf"] = np.array(wfpt.wiener_like_rlddm(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.9, 0, 1e-4))
    out["rl_a"] = np.array(wfpt.wiener_like_rl(resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0.5, p_outlier=0.05, w_outlier=0.1))
    out["rl_b"] = np.array(wfpt.wiener_like_rl(resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.45))
    alpha_arr = rng.normal(-1, 0.3, n)
    ssort = np.sort(split)
    out["rl_split_sorted"] = ssort
    out["rl_alpha_arr"] = alpha_arr
    v_arr = rng.normal(2, 0.3, n)
    out["rl_v_arr"] = v_arr
    out["multi_rlddm"] = np.array(wfpt.wiener_like_multi_rlddm(xr, resp, fb, ssort, 0.5, v_arr, 0, 1.8, 0.5, 0, 0.3, 0, alpha_arr, 1e-4, ["v", "alpha"], p_outlier=0.05, w_outlier=0.1))

    # --- contaminant -------------------------------------------------------------------------------------------------------
    cont = (rng.random(n) < 0.1).astype(np.int32)
    out["cont_x"] = cont
    out["contaminant"] = np.array(wfpt.wiener_like_contaminant(xr, cont, 1.0, 0.5, 1.8, 0.5, 0.1, 0.3, 0.1, 0.1, 5.0, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3))

    # --- pdf_array on a grid (plots / gen_cdf_using_pdf input, wfpt.pyx:598) --------------------------------------
    grid = np.linspace(-5, 5, 1001)
    out["pdf_grid_x"] = grid
    out["pdf_grid_full"] = wfpt.pdf_array(grid, 0.7, 0.8, 1.6, 0.45, 0.2, 0.25, 0.15, 1e-4, 0, 2, 2, 1, 1e-3, 0.05, 0.1)
    out["pdf_grid_simple_log"] = wfpt.pdf_array(grid, 0.7, 0, 1.6, 0.45, 0, 0.25, 0, 1e-4, 1, 2, 2, 1, 1e-3, 0.05, 0.1)

    path = os.path.join(HERE, "wfpt_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
