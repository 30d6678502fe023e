"""Network-likelihood (LAN) wrappers of the reference `wfpt` module, host side only.

The reference functions (`src/wfpt.pyx:156-267, 436-514, 633-753`) contain no arithmetic of their own beyond
assembling the float32 input matrix, running the Q-learning recurrence that produces the per-trial drift column,
and mixing the network's log-density with the outlier density; the network call (`network.predict_on_batch`)
is the caller's object.  SURVEY.md section 8(b) keeps them out of the CUDA scope; they are provided so that
`import wfpt` exposes every name HDDM's model classes look up.  Nothing here touches the GPU library.
"""
from __future__ import annotations

import numpy as np

LL_MIN = np.float32(-16.11809)  # floor applied to the network output (`wfpt.pyx:643`)
_RL_BASE = 2.718281828459       # truncated e of the reference's learning-rate transform (`wfpt.pyx:222`)


def _net_logp(network, data, p_outlier, w_outlier):
    """Per-row log density: floor at LL_MIN, then the outlier mixture when p_outlier != 0 (`wfpt.pyx:650-653`)."""
    lp = np.maximum(network.predict_on_batch(data), LL_MIN)
    if p_outlier == 0:
        return lp
    return np.log(np.exp(lp) * (1.0 - p_outlier) + (w_outlier * p_outlier))


def _as_f32_scalar(v):
    # the reference stores the sum in a C float before returning it
    return float(np.float32(v))


def _trial_matrix(rt, response, params):
    n, k = rt.shape[0], params.shape[0]
    data = np.empty((n, k + 2), dtype=np.float32)
    data[:, :k] = params.astype(np.float32)[None, :]
    data[:, k] = rt
    data[:, k + 1] = response
    return data


def _check_f32_1d(**arrs):
    for name, a in arrs.items():
        if not isinstance(a, np.ndarray):
            raise TypeError(f"Argument '{name}' has incorrect type (expected numpy.ndarray, got {type(a).__name__})")
        if a.ndim != 1:
            raise ValueError(f"Buffer has wrong number of dimensions (expected 1, got {a.ndim})")
        if a.dtype != np.float32:
            raise ValueError("Buffer dtype mismatch, expected 'float' but got '%s'" %
                             ("double" if a.dtype == np.float64 else a.dtype.name))


def wiener_like_nn_mlp(rt, response, params, p_outlier=0, w_outlier=0, network=None):
    """`wfpt.pyx:633-655`."""
    _check_f32_1d(rt=rt, response=response, params=params)
    return _as_f32_scalar(np.sum(_net_logp(network, _trial_matrix(rt, response, params), p_outlier, w_outlier)))


def wiener_like_nn_mlp_info(rt, response, params, upper_bounds, lower_bounds, p_outlier=0, w_outlier=0,
                            network=None):
    """`wfpt.pyx:658-690`: as above, -inf when a parameter leaves [lower_bounds, upper_bounds]."""
    _check_f32_1d(rt=rt, response=response, params=params, upper_bounds=upper_bounds, lower_bounds=lower_bounds)
    k = params.shape[0]
    if np.any(params > upper_bounds[:k]) or np.any(params < lower_bounds[:k]):
        return -np.inf
    return _as_f32_scalar(np.sum(_net_logp(network, _trial_matrix(rt, response, params), p_outlier, w_outlier)))


def wiener_like_nn_mlp_pdf(rt, response, params, p_outlier=0, w_outlier=0, logp=0, network=None):
    """`wfpt.pyx:692-716`: per-trial (log-)density."""
    _check_f32_1d(rt=rt, response=response, params=params)
    lp = np.squeeze(_net_logp(network, _trial_matrix(rt, response, params), p_outlier, w_outlier))
    return lp if logp else np.exp(lp)


def wiener_like_multi_nn_mlp(data, p_outlier=0, w_outlier=0, network=None):
    """`wfpt.pyx:721-735`: regression style, the caller supplies the assembled matrix."""
    return _as_f32_scalar(np.sum(_net_logp(network, data, p_outlier, w_outlier)))


def wiener_like_multi_nn_mlp_pdf(data, p_outlier=0, w_outlier=0, network=None):
    """`wfpt.pyx:737-752`.  The reference assigns the array to a C float and raises for more than one row; this
    returns the per-row log-densities (what `wiener_like_nn_mlp_pdf(..., logp=1)` returns)."""
    return np.squeeze(_net_logp(network, data, p_outlier, w_outlier))


def _learning_rate(alpha):
    g = _RL_BASE ** alpha
    return g / (1 + g)


def _groups(split_by):
    """Row indices of each condition, conditions in `np.unique` order (`wfpt.pyx:184, 209-214`)."""
    return [np.flatnonzero(split_by == s) for s in np.unique(split_by)]


def wiener_like_rlssm_nn(model, x, response, feedback, split_by, q, params_ssm, params_rl, params_bnds,
                         p_outlier=0, w_outlier=0, network=None):
    """`wfpt.pyx:156-267`.  The per-trial drift `(Q_upper - Q_lower) * v` is written into column 0 in
    condition-grouped row order while the rt / response columns keep the caller's order, as in the reference."""
    if not (0 <= p_outlier <= 1):
        return -np.inf
    k = params_ssm.shape[0]
    lo, hi = params_bnds[0], params_bnds[1]
    for i in range(1, k):
        if params_ssm[i] < lo[i] or params_ssm[i] > hi[i]:
            return -np.inf
    v = float(params_ssm[0])
    a_neg = _learning_rate(float(params_rl[0]))
    a_pos = _learning_rate(float(params_rl[1])) if len(params_rl) == 2 else a_neg

    n = x.shape[0]
    data = np.zeros((n, k + 2), dtype=np.float32)
    row = 0
    for idx in _groups(split_by):
        fb = feedback[idx]
        ch = np.where(response[idx] == -1, 0, response[idx])
        qs = [q, q]
        for i in range(idx.shape[0]):
            if i > 0:
                data[row + i, 0] = (qs[1] - qs[0]) * v
                if data[row + i, 0] < lo[0] or data[row + i, 0] > hi[0]:
                    return -np.inf
            c = int(ch[i])
            rate = a_pos if fb[i] > qs[c] else a_neg
            qs[c] = qs[c] + rate * (fb[i] - qs[c])
        row += idx.shape[0]
    data[:, 1:k] = params_ssm[1:].astype(np.float32)[None, :]
    data[:, k] = x
    data[:, k + 1] = response
    return float(np.sum(_net_logp(network, data, p_outlier, w_outlier)))


def wiener_like_rlssm_nn_reg(data, rl_arr, x, response, feedback, split_by, q, params_bnds, p_outlier=0,
                             w_outlier=0, network=None):
    """`wfpt.pyx:436-514`.  Column 0 of `data` holds the drift scaling on entry and is overwritten IN PLACE with the
    per-trial drift (the reference aliases its `data_copy` to `data`), so the caller rebuilds `data` per call."""
    if not (0 <= p_outlier <= 1):
        return -np.inf
    lo, hi = params_bnds[0], params_bnds[1]
    for i in range(1, data.shape[1] - 2):
        if data[:, i].min() < lo[i] or data[:, i].max() > hi[i]:
            return -np.inf
    row = 0
    for idx in _groups(split_by):
        fb = feedback[idx]
        ch = np.where(response[idx] == -1, 0, response[idx])
        qs = [q, q]
        for i in range(idx.shape[0]):
            scale = float(data[row + i, 0])
            if scale < 0:
                return -np.inf
            data[row + i, 0] = (qs[1] - qs[0]) * scale
            if data[row + i, 0] < lo[0] or data[row + i, 0] > hi[0]:
                return -np.inf
            rate = _learning_rate(float(rl_arr[row + i, 0]))
            c = int(ch[i])
            qs[c] = qs[c] + rate * (fb[i] - qs[c])
        row += idx.shape[0]
    return float(np.sum(_net_logp(network, data, p_outlier, w_outlier)))


__all__ = ["wiener_like_nn_mlp", "wiener_like_nn_mlp_info", "wiener_like_nn_mlp_pdf", "wiener_like_multi_nn_mlp",
           "wiener_like_multi_nn_mlp_pdf", "wiener_like_rlssm_nn", "wiener_like_rlssm_nn_reg"]
