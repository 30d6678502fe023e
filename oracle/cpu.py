"""ctypes front for oracle/wfpt_oracle.c and oracle/sim_oracle.c.  TEST INFRASTRUCTURE ONLY.

Importers allowed: tests/, bench.py (cpu_baseline / --impl reference legs) and __graft_entry__.smoke().
The product package hddm_b200/ never imports this module (tests/test_boundary.py greps for it).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")
LIB = os.path.join(BUILD, "liboracle.so")
SRCS = [os.path.join(HERE, "wfpt_oracle.c"), os.path.join(HERE, "sim_oracle.c")]

_dp = C.POINTER(C.c_double)
_i64p = C.POINTER(C.c_int64)


def build_oracle(force: bool = False) -> str:
    srcs = [s for s in SRCS if os.path.exists(s)]
    if (not force) and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in srcs):
        return LIB
    os.makedirs(BUILD, exist_ok=True)
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-std=c11", "-o", LIB] + srcs + ["-lm"]
    subprocess.run(cmd, check=True)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_oracle())
        L = _lib
        L.orc_ftt_01w.restype = C.c_double
        L.orc_ftt_01w.argtypes = [C.c_double] * 3
        L.orc_ftt_01w_info.restype = C.c_double
        L.orc_ftt_01w_info.argtypes = [C.c_double] * 3 + [C.POINTER(C.c_int)] * 2
        L.orc_prob_ub.restype = C.c_double
        L.orc_prob_ub.argtypes = [C.c_double] * 3
        L.orc_pdf_sv.restype = C.c_double
        L.orc_pdf_sv.argtypes = [C.c_double] * 6
        L.orc_full_pdf.restype = C.c_double
        L.orc_full_pdf.argtypes = [C.c_double] * 9 + [C.c_int] * 3 + [C.c_double]
        L.orc_pdf_array.restype = None
        L.orc_pdf_array.argtypes = (
            [_dp, C.c_ssize_t, C.c_ssize_t] + [C.c_double] * 8 + [C.c_int] * 4 + [C.c_double] * 3 + [_dp]
        )
        L.orc_wiener_like.restype = C.c_double
        L.orc_wiener_like.argtypes = [_dp, C.c_ssize_t, C.c_ssize_t] + [C.c_double] * 8 + [C.c_int] * 3 + [C.c_double] * 3
        L.orc_wiener_like_batch.restype = None
        L.orc_wiener_like_batch.argtypes = [_dp, C.c_ssize_t, _dp, C.c_ssize_t, C.c_double, C.c_int, C.c_int, C.c_int,
                                            C.c_double, C.c_double, _dp]
        L.orc_wiener_like_multi.restype = C.c_double
        L.orc_wiener_like_multi.argtypes = [_dp, C.c_ssize_t, C.POINTER(_dp), _dp, C.c_double, C.c_int, C.c_int,
                                            C.c_int, C.c_double, C.c_double, C.c_double]
        L.orc_wiener_like_rlddm.restype = C.c_double
        L.orc_wiener_like_rlddm.argtypes = ([_dp, _i64p, _dp, _i64p, _i64p, C.c_ssize_t] + [C.c_double] * 11 +
                                            [C.c_int] * 3 + [C.c_double] * 3)
        L.orc_wiener_like_rl.restype = C.c_double
        L.orc_wiener_like_rl.argtypes = [_i64p, _dp, _i64p, _i64p, C.c_ssize_t] + [C.c_double] * 7
        L.orc_wiener_like_multi_rlddm.restype = C.c_double
        L.orc_wiener_like_multi_rlddm.argtypes = [_dp, _i64p, _dp, _i64p, C.c_ssize_t, C.c_double, C.POINTER(_dp), _dp,
                                                  C.c_double, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                                                  C.c_double]
        L.orc_wiener_like_contaminant.restype = C.c_double
        L.orc_wiener_like_contaminant.argtypes = ([_dp, C.POINTER(C.c_int32), C.c_ssize_t] + [C.c_double] * 10 +
                                                  [C.c_int] * 3 + [C.c_double])
        L.orc_cdf_grid.restype = None
        L.orc_cdf_grid.argtypes = [_dp, C.c_ssize_t] + [C.c_double] * 5 + [_dp]
        L.orc_nodes_get.restype = C.c_longlong
        L.orc_nodes_reset.restype = None
        if hasattr(L, "orc_sim_run"):
            L.orc_sim_run.restype = C.c_int
            L.orc_sim_run.argtypes = [C.c_int, C.POINTER(C.c_float), C.c_int, C.c_int, C.c_int64, C.c_float,
                                      C.c_float, C.c_float, C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
    return _lib


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_dp)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(_i64p)


# ---- reference-shaped callables (same positional order and defaults as src/wfpt.pyx) -------------------------------

def ftt_01w(tt, w, err):
    return lib().orc_ftt_01w(tt, w, err)


def ftt_01w_info(tt, w, err):
    K = C.c_int(0)
    small = C.c_int(0)
    p = lib().orc_ftt_01w_info(tt, w, err, C.byref(K), C.byref(small))
    return p, K.value, bool(small.value)


def prob_ub(v, a, z):
    return lib().orc_prob_ub(v, a, z)


def full_pdf(x, v, sv, a, z, sz, t, st, err, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
    return lib().orc_full_pdf(x, v, sv, a, z, sz, t, st, err, n_st, n_sz, int(use_adaptive), simps_err)


def pdf_array(x, v, sv, a, z, sz, t, st, err=1e-4, logp=0, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
              p_outlier=0, w_outlier=0):
    x = np.asarray(x, dtype=np.float64)
    xs = x if x.flags.c_contiguous else np.ascontiguousarray(x)
    y = np.empty(xs.shape[0], dtype=np.float64)
    with np.errstate(all="ignore"):
        lib().orc_pdf_array(xs.ctypes.data_as(_dp), xs.shape[0], 1, v, sv, a, z, sz, t, st, err, int(logp), n_st, n_sz,
                            int(use_adaptive), simps_err, p_outlier, w_outlier, y.ctypes.data_as(_dp))
    return y


def wiener_like(x, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0,
                w_outlier=0.1):
    xs, xp = _d(x)
    return lib().orc_wiener_like(xp, xs.shape[0], 1, v, sv, a, z, sz, t, st, err, n_st, n_sz, int(use_adaptive),
                                 simps_err, p_outlier, w_outlier)


def wiener_like_batch(x, params, err, n_st, n_sz, use_adaptive, simps_err, w_outlier):
    xs, xp = _d(x)
    ps, pp = _d(params)
    assert ps.ndim == 2 and ps.shape[1] == 8
    out = np.empty(ps.shape[0], dtype=np.float64)
    lib().orc_wiener_like_batch(xp, xs.shape[0], pp, ps.shape[0], err, n_st, n_sz, int(use_adaptive), simps_err,
                                w_outlier, out.ctypes.data_as(_dp))
    return out


_NAMES = ("v", "sv", "a", "z", "sz", "t", "st")


def _pack_multi(names, values, multi, n):
    keep = []
    ptrs = (_dp * len(names))()
    scal = (C.c_double * len(names))()
    for k, name in enumerate(names):
        if name in multi:
            arr = np.ascontiguousarray(values[name], dtype=np.float64)
            assert arr.shape == (n,)
            keep.append(arr)
            ptrs[k] = arr.ctypes.data_as(_dp)
            scal[k] = 0.0
        else:
            ptrs[k] = None
            scal[k] = float(values[name])
    return keep, ptrs, scal


def wiener_like_multi(x, v, sv, a, z, sz, t, st, err, multi=None, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3,
                      p_outlier=0, w_outlier=0):
    xs, xp = _d(x)
    vals = dict(v=v, sv=sv, a=a, z=z, sz=sz, t=t, st=st)
    keep, ptrs, scal = _pack_multi(_NAMES, vals, multi, xs.shape[0])
    return lib().orc_wiener_like_multi(xp, xs.shape[0], ptrs, scal, err, n_st, n_sz, int(use_adaptive), simps_err,
                                       p_outlier, w_outlier)


def group_by_split(split_by):
    """np.unique + boolean masks of the reference (wfpt.pyx:99,113-116) as one stable grouping."""
    split_by = np.asarray(split_by, dtype=np.int64)
    order = np.argsort(split_by, kind="stable").astype(np.int64)
    uniq, counts = np.unique(split_by, return_counts=True)
    starts = np.zeros(len(uniq) + 1, dtype=np.int64)
    np.cumsum(counts, out=starts[1:])
    return order, starts


def wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha, v, sv, a, z, sz, t, st, err, n_st=10,
                      n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    xs, xp = _d(x)
    rs, rp = _i(response)
    fs, fp = _d(feedback)
    order, starts = group_by_split(split_by)
    return lib().orc_wiener_like_rlddm(xp, rp, fp, order.ctypes.data_as(_i64p), starts.ctypes.data_as(_i64p),
                                       len(starts) - 1, q, alpha, pos_alpha, v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                       int(use_adaptive), simps_err, p_outlier, w_outlier)


def wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, v, z, err=1e-4, n_st=10, n_sz=10,
                   use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    rs, rp = _i(response)
    fs, fp = _d(feedback)
    order, starts = group_by_split(split_by)
    return lib().orc_wiener_like_rl(rp, fp, order.ctypes.data_as(_i64p), starts.ctypes.data_as(_i64p), len(starts) - 1,
                                    q, alpha, pos_alpha, v, z, p_outlier, w_outlier)


def wiener_like_multi_rlddm(x, response, feedback, split_by, q, v, sv, a, z, sz, t, st, alpha, err, multi=None,
                            n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3, p_outlier=0, w_outlier=0):
    xs, xp = _d(x)
    rs, rp = _i(response)
    fs, fp = _d(feedback)
    ss, sp = _i(split_by)
    vals = dict(v=v, sv=sv, a=a, z=z, sz=sz, t=t, st=st, alpha=alpha)
    keep, ptrs, scal = _pack_multi(_NAMES + ("alpha",), vals, multi, xs.shape[0])
    return lib().orc_wiener_like_multi_rlddm(xp, rp, fp, sp, xs.shape[0], q, ptrs, scal, err, n_st, n_sz,
                                             int(use_adaptive), simps_err, p_outlier, w_outlier)


def wiener_like_contaminant(x, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err, n_st=10, n_sz=10, use_adaptive=1,
                            simps_err=1e-8):
    xs, xp = _d(x)
    cs = np.ascontiguousarray(cont_x, dtype=np.int32)
    return lib().orc_wiener_like_contaminant(xp, cs.ctypes.data_as(C.POINTER(C.c_int32)), xs.shape[0], v, sv, a, z, sz,
                                             t, st, t_min, t_max, err, n_st, n_sz, int(use_adaptive), simps_err)


def cdf_grid(x, v, sv, a, z, sz):
    xs, xp = _d(x)
    out = np.empty_like(xs)
    lib().orc_cdf_grid(xp, xs.shape[0], v, sv, a, z, sz, out.ctypes.data_as(_dp))
    return out


def gen_rts_from_cdf(v, sv, a, z, sz, t, st, samples=1000, cdf_lb=-6, cdf_ub=6, dt=1e-2):
    """wfpt.pyx:517-548 with the same global-NumPy-RNG call sequence (rand(samples) [, rand(samples)])."""
    x = np.arange(cdf_lb, cdf_ub, dt)
    l_cdf = cdf_grid(x, v, sv, a, z, sz)
    f = np.random.rand(samples)
    if st != 0:
        delay = np.random.rand(samples) * st + (t - st / 2.0)
    idx = np.searchsorted(l_cdf, f)
    rt = x[idx]
    if st == 0:
        return rt + np.sign(rt) * t
    return rt + np.sign(rt) * delay


def sim_run(model_id, theta, n_samples, delta_t=0.001, max_t=20.0, s=1.0, seed=0):
    """(rts float32[n_samples, n_sets], choices int32[n_samples, n_sets]) from oracle/sim_oracle.c."""
    theta = np.ascontiguousarray(np.atleast_2d(np.asarray(theta, dtype=np.float32)))
    n_sets, n_par = theta.shape
    rts = np.empty((n_samples, n_sets), dtype=np.float32)
    ch = np.empty((n_samples, n_sets), dtype=np.int32)
    rc = lib().orc_sim_run(model_id, theta.ctypes.data_as(C.POINTER(C.c_float)), n_par, n_sets, n_samples, delta_t, max_t,
                           s, seed, rts.ctypes.data_as(C.POINTER(C.c_float)), ch.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc:
        raise ValueError(f"orc_sim_run failed ({rc})")
    return rts, ch


def sim_race_run(lca, n_particles, angle, theta, n_samples, delta_t=0.001, max_t=20.0, s=1.0, seed=0):
    """race_model / lca from oracle/sim_oracle.c; theta rows in the canonical layout v[np], a, z[np], g, b, t, angle."""
    theta = np.ascontiguousarray(np.atleast_2d(np.asarray(theta, dtype=np.float32)))
    n_sets = theta.shape[0]
    assert theta.shape[1] == 2 * n_particles + 5
    rts = np.empty((n_samples, n_sets), dtype=np.float32)
    ch = np.empty((n_samples, n_sets), dtype=np.int32)
    f = lib().orc_sim_race_run
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_int, C.c_int64, C.c_float, C.c_float, C.c_float,
                  C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
    rc = f(int(lca), n_particles, int(angle), theta.ctypes.data_as(C.POINTER(C.c_float)), n_sets, n_samples, delta_t, max_t, s,
           seed, rts.ctypes.data_as(C.POINTER(C.c_float)), ch.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc:
        raise ValueError(f"orc_sim_race_run failed ({rc})")
    return rts, ch


def sim_par2_run(bound_kind, theta, n_samples, delta_t=0.001, max_t=20.0, s=1.0, seed=0):
    """ddm_flexbound_par2 from oracle/sim_oracle.c; theta rows: v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1."""
    theta = np.ascontiguousarray(np.atleast_2d(np.asarray(theta, dtype=np.float32)))
    n_sets = theta.shape[0]
    assert theta.shape[1] == 10
    rts = np.empty((n_samples, n_sets), dtype=np.float32)
    ch = np.empty((n_samples, n_sets), dtype=np.int32)
    f = lib().orc_sim_par2_run
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.POINTER(C.c_float), C.c_int, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_uint64,
                  C.POINTER(C.c_float), C.POINTER(C.c_int32)]
    rc = f(int(bound_kind), theta.ctypes.data_as(C.POINTER(C.c_float)), n_sets, n_samples, delta_t, max_t, s, seed,
           rts.ctypes.data_as(C.POINTER(C.c_float)), ch.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc:
        raise ValueError(f"orc_sim_par2_run failed ({rc})")
    return rts, ch


def sim_hier_run(kind, bound_kind, theta, n_samples, delta_t=0.001, max_t=20.0, s=1.0, seed=0):
    """ddm_flexbound_seq2 / _mic2_adj / _tradeoff (kind 1 / 2 / 3) from oracle/sim_oracle.c; theta rows:
    v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1, d, b2, b3 (bound_kind 3 = conflict-gamma bound)."""
    theta = np.ascontiguousarray(np.atleast_2d(np.asarray(theta, dtype=np.float32)))
    n_sets = theta.shape[0]
    assert theta.shape[1] == 13
    rts = np.empty((n_samples, n_sets), dtype=np.float32)
    ch = np.empty((n_samples, n_sets), dtype=np.int32)
    f = lib().orc_sim_hier_run
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_float), C.c_int, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_uint64,
                  C.POINTER(C.c_float), C.POINTER(C.c_int32)]
    rc = f(int(kind), int(bound_kind), theta.ctypes.data_as(C.POINTER(C.c_float)), n_sets, n_samples, delta_t, max_t, s,
           seed, rts.ctypes.data_as(C.POINTER(C.c_float)), ch.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc:
        raise ValueError(f"orc_sim_hier_run failed ({rc})")
    return rts, ch


def dmat_cdf_array_ref(x, v, sv, a, z, sz, t, st, p_outlier, w_outlier):
    """hddm.cdfdif.dmat_cdf_array with the reference's own C (oracle/_ref/libcdfdif_ref.so, built from
    /root/reference/src/cdfdif.c as it lies) and the wrapper's arithmetic (src/cdfdif_wrapper.pyx:36-51) restated."""
    from oracle.build_ref import load_cdf_ref

    L = load_cdf_ref()
    x = np.asarray(x, dtype=np.float64)
    epsi = 1e-10
    params = (C.c_double * 7)(a / 10.0, t, sv / 10.0 + epsi, z * (a / 10.0), sz * (a / 10.0) + epsi, st + epsi, v / 10.0)
    y = np.empty(len(x))
    pb = C.c_double(0.0)
    for i, xi in enumerate(x):
        yi = L.cdfdif(abs(float(xi)), int(xi > 0), params, C.byref(pb))
        yi = (1 - pb.value) + float(np.sign(xi)) * yi
        y[i] = yi * (1 - p_outlier) + (xi + (1.0 / (2 * w_outlier))) * w_outlier * p_outlier
    return y


def nodes_reset():
    lib().orc_nodes_reset()


def nodes_get():
    return int(lib().orc_nodes_get())
