"""The likelihood adapter HDDM puts between its PyMC nodes and the `wfpt` extension, on top of the GPU drop-in
(/root/reference/hddm/likelihoods.py:83-243).

`generate_wfpt_stochastic_class(wiener_params, ...)` returns a class with the closure semantics of the reference's
(`wfpt_like` with the `|rt| >= 999` no-response branch, `pdf`, `cdf`, `cdf_vec`, `random`).  The reference builds the
class with kabuki's `stochastic_from_dist`, i.e. as a PyMC 2 stochastic; neither package exists in this image, so the
class returned here is a plain Python object with the attributes the sampler reads (`value`, `parents`, `logp`) -- where
kabuki is importable the same `wfpt_like` is handed to `stochastic_from_dist` and the real thing comes back.

What the GPU adds is below the call: `wfpt.wiener_like` keeps every node's RT array resident (keyed by address, verified
byte for byte per call), so a logp evaluation moves 64 bytes; and `batch_logp` evaluates MANY nodes (all children of one
stochastic, several proposals per node) in one launch through `wfpt.ResidentData`.
"""
from __future__ import annotations

import numpy as np

from . import wfpt as _wfpt

DEFAULT_WIENER_PARAMS = {"err": 1e-4, "n_st": 2, "n_sz": 2, "use_adaptive": 1, "simps_err": 1e-3, "w_outlier": 0.1}
_PARAM_ORDER = ("v", "sv", "a", "z", "sz", "t", "st")


def _rt_column(x):
    """x["rt"] as a float64 ndarray without copying when it already is one (pandas: `.values` of a float column is a
    view, so the resident cache sees the same address on every call)."""
    col = x["rt"]
    vals = getattr(col, "values", col)
    return vals if (type(vals) is np.ndarray and vals.dtype == np.float64) else np.asarray(vals, dtype=np.float64)


def make_wfpt_like(wiener_params=None):
    """The closure of likelihoods.py:115-149: `wfpt_like(x, v, sv, a, z, sz, t, st, p_outlier=0)` on a frame with an
    "rt" column (signed RTs; |rt| >= 999 marks a missing response, upper / lower by sign)."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)

    def wfpt_like(x, v, sv, a, z, sz, t, st, p_outlier=0):
        rt = _rt_column(x)
        if rt.shape[0] == 0 or np.abs(rt).max() < 998:
            return _wfpt.wiener_like(rt, v, sv, a, z, sz, t, st, p_outlier=p_outlier, **wp)
        # missing RTs (likelihoods.py:122-149): the observed ones as usual, the others through the choice probability
        noresponse = np.abs(rt) >= 999
        logp_resp = _wfpt.wiener_like(np.ascontiguousarray(rt[~noresponse]), v, sv, a, z, sz, t, st, p_outlier=p_outlier, **wp)
        n_noresponse = int(noresponse.sum())
        k_upper = int((rt[noresponse] > 0).sum())
        if v == 0:
            p_upper = z
        else:
            p_upper = (np.exp(-2 * a * z * v) - 1) / (np.exp(-2 * a * v) - 1)
        from scipy import stats

        return logp_resp + stats.binom.logpmf(k_upper, n_noresponse, p_upper)

    return wfpt_like


def batch_logp(nodes, params, wiener_params=None, resident=None):
    """log-likelihoods of many nodes in ONE launch: `nodes` is a sequence of RT arrays (or frames with an "rt" column),
    `params` float64[len(nodes), 8] rows (v, sv, a, z, sz, t, st, p_outlier).  Pass the `ResidentData` returned by an
    earlier call (second return value) as `resident` to skip the upload: that is the sampler's loop.  Equals
    [wfpt_like(node, *row) for node, row in zip(nodes, params)] for nodes without missing responses."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)
    if resident is None:
        arrays = [n if isinstance(n, np.ndarray) else _rt_column(n) for n in nodes]
        resident = _wfpt.ResidentData(arrays)
    return resident.logp(np.ascontiguousarray(params, dtype=np.float64), **wp), resident


def generate_wfpt_stochastic_class(wiener_params=None, sampling_method="cssm", cdf_range=(-5, 5), sampling_dt=1e-4):
    """likelihoods.py:83-243.  Returns the `wfpt` class: `wfpt(value=frame, v=..., sv=..., a=..., z=..., sz=..., t=...,
    st=..., p_outlier=...)` with `.logp`, `.pdf(x)`, `.cdf(x)`, `.cdf_vec()`, `.random()`."""
    wp = dict(DEFAULT_WIENER_PARAMS if wiener_params is None else wiener_params)
    wfpt_like = make_wfpt_like(wp)
    try:  # the real PyMC stochastic where the reference's stack exists
        from kabuki.utils import stochastic_from_dist  # type: ignore

        base = stochastic_from_dist("wfpt", wfpt_like)
    except Exception:
        class base:  # noqa: N801  (the reference names the class `wfpt`)
            def __init__(self, name="wfpt", value=None, observed=True, **parents):
                self.__name__ = name
                self.value = value
                self.observed = observed
                self.parents = dict(parents)

            @property
            def shape(self):
                return (len(self.value),) if self.value is not None else (0,)

            @property
            def logp(self):
                return wfpt_like(self.value, **self.parents)

    def _parents(self):
        p = self.parents
        return dict(getattr(p, "value", p))

    def pdf(self, x):
        p = _parents(self)
        return _wfpt.pdf_array(np.asarray(x, dtype=np.float64), *(p[k] for k in _PARAM_ORDER), p_outlier=p.get("p_outlier", 0))

    def cdf(self, x):
        from . import cdfdif

        p = _parents(self)
        return cdfdif.dmat_cdf_array(np.asarray(x, dtype=np.float64), w_outlier=wp["w_outlier"], **p)

    def cdf_vec(self):
        p = _parents(self)
        return _wfpt.gen_cdf_using_pdf(time=cdf_range[1], **dict(list(p.items()) + list(wp.items())))

    def random(self, keep_negative_responses=True, add_model=False, add_outliers=False, add_model_parameters=False,
               keep_subj_idx=False):
        assert sampling_method in ["cdf", "drift", "cssm"], "Sampling method is invalid!"
        p = _parents(self)
        if sampling_method == "cdf":
            rts = _wfpt.gen_rts_from_cdf(*(p[k] for k in _PARAM_ORDER), samples=self.shape[0], cdf_lb=cdf_range[0],
                                         cdf_ub=cdf_range[1], dt=sampling_dt)
            return rts
        if sampling_method == "drift":
            raise NotImplementedError("gen_rts(method='drift') lives in hddm.generate (host Python), outside this package")
        from . import simulators

        theta = np.array([p[k] for k in ("v", "a", "z", "t", "sz", "sv", "st")], dtype=np.float32)
        sim_out = simulators.simulator(theta=theta, model="full_ddm_hddm_base", n_samples=self.shape[0], max_t=20)
        if add_outliers and p.get("p_outlier", 0) > 0.0:
            sim_out = simulators._add_outliers(sim_out=sim_out, p_outlier=p["p_outlier"], max_rt_outlier=1 / wp["w_outlier"])
        out = simulators.hddm_preprocess(sim_out, keep_negative_responses=keep_negative_responses, keep_subj_idx=keep_subj_idx,
                                         add_model_parameters=add_model_parameters)
        if add_model:
            out["model"] = "ddm_hddm_base" if (p["sz"] == 0 and p["sv"] == 0 and p["st"] == 0) else "full_ddm_hddm_base"
        if not np.any(out["rt"] < 0):                                      # hddm.utils.flip_errors (utils.py:489-517)
            out = out.copy()
            idx = out["response"] != 1
            out.loc[idx, "rt"] = -out.loc[idx, "rt"]
        return out

    wfpt = type("wfpt", (base,), {"pdf": pdf, "cdf": cdf, "cdf_vec": cdf_vec, "random": random,
                                  "wfpt_like": staticmethod(wfpt_like)})
    return wfpt
