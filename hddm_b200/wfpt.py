"""Drop-in for the reference's Cython extension module `wfpt` (/root/reference/src/wfpt.pyx), backed by
libwfpt_b200.so.  Same callables, positional orders, defaults, return types and argument errors; the arithmetic
runs on the GPU (FP64, sm_100a kernels).  There is no CPU path: without the built library / a GPU every call raises.

A top-level module named `wfpt` (repo root `wfpt.py`) re-exports this one, because the reference imports it as
`import wfpt` (hddm/__init__.py:29) and `from wfpt import wiener_like_rlddm` (hddm/models/hddm_rl.py:12).

Host work kept here is the host work the reference itself does in NumPy: argument coercion, np.unique grouping
(wfpt.pyx:99,113-116), cumsum / searchsorted / RNG of gen_rts_from_cdf (wfpt.pyx:533-547), cumulative_trapezoid of
gen_cdf_using_pdf (wfpt.pyx:602).
"""
from __future__ import annotations

import ctypes as C
from copy import copy  # noqa: F401  (the reference module exposes `copy`)

import numpy as np

from . import _lib
from ._lib import FAITHFUL, FAST, WfptError  # noqa: F401
from .lan import *  # noqa: F401,F403  (network-likelihood wrappers, host-only; `src/wfpt.pyx:156-267,436-514,633-753`)
from .lan import __all__ as _lan_all

__all__ = [
    "pdf_array", "wiener_like", "wiener_like_rlddm", "wiener_like_rl", "wiener_like_multi", "wiener_like_multi_rlddm",
    "wiener_like_contaminant", "gen_rts_from_cdf", "gen_cdf_using_pdf", "split_cdf", "full_pdf", "wiener_like_batch", "sort_rts",
    "set_variant", "get_variant", "ResidentData", "wiener_like_segments", "cache_config", "cache_clear", "cache_stats",
] + list(_lan_all)

_variant = FAST


def set_variant(v: int) -> None:
    """Select the device evaluator for the reference-shaped calls: FAST (default) or FAITHFUL."""
    global _variant
    if v not in (FAST, FAITHFUL):
        raise ValueError("variant must be hddm_b200.wfpt.FAST or FAITHFUL")
    _variant = v


def get_variant() -> int:
    return _variant


# ---- Cython-buffer argument semantics (errors to reproduce) ---------------------------------------------
_CNAME = {"float64": "double", "float32": "float", "int64": "long", "int32": "int", "int16": "short", "int8": "signed char",
          "uint8": "unsigned char", "bool": "bool", "complex128": "double complex", "float16": "half",
          "uint64": "unsigned long", "uint32": "unsigned int", "object": "Python object"}


def _buffer(arg, name, dtype=np.float64):
    """np.ndarray[dtype, ndim=1] argument of a Cython def: type / ndim / dtype checks, strides allowed."""
    if not isinstance(arg, np.ndarray):
        raise TypeError(
            f"Argument '{name}' has incorrect type (expected numpy.ndarray, got {type(arg).__name__})"
        )
    if arg.ndim != 1:
        raise ValueError(f"Buffer has wrong number of dimensions (expected 1, got {arg.ndim})")
    want = np.dtype(dtype)
    if arg.dtype != want:
        raise ValueError(
            f"Buffer dtype mismatch, expected '{_CNAME.get(want.name, want.name)}' but got "
            f"'{_CNAME.get(arg.dtype.name, arg.dtype.name)}'"
        )
    return np.ascontiguousarray(arg)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _opts(err, n_st, n_sz, use_adaptive, simps_err):
    return float(err), int(n_st), int(n_sz), 1 if use_adaptive else 0, float(simps_err)


# ---- src/wfpt.pyx:32-48 --------------------------------------------------------------------------------------------------
def pdf_array(x, v, sv, a, z, sz, t, st, err=1e-4, logp=0, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
              p_outlier=0, w_outlier=0):
    x = _buffer(x, "x")
    prm = (C.c_double * 8)(float(v), float(sv), float(a), float(z), float(sz), float(t), float(st), float(p_outlier))
    y = np.empty(x.shape[0], dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_pdf_array_host(_ptr(x), x.shape[0], prm, *_opts(err, n_st, n_sz, use_adaptive, simps_err),
                                     float(w_outlier), 1 if logp == 1 else 0, _variant, _ptr(y)))
    return y


# ---- src/pdf.pxi:104-146 (cpdef, also called from Python: hddm/tests/test_likelihoods.py:56) --------------------------
def full_pdf(x, v, sv, a, z, sz, t, st, err, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
    xa = np.array([float(x)], dtype=np.double)
    prm = (C.c_double * 8)(float(v), float(sv), float(a), float(z), float(sz), float(t), float(st), 0.0)
    y = np.empty(1, dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_pdf_array_host(_ptr(xa), 1, prm, *_opts(err, n_st, n_sz, use_adaptive, simps_err), 0.0, 0,
                                     _variant, _ptr(y)))
    return float(y[0])


# ---- batched form (no reference counterpart with P > 1; P == 1 is wiener_like) -----------------------------------------
def sort_rts(x):
    """RTs reordered by |rt| (stable).  tt = (|rt| - t)/a^2 is then monotone along the thread index for every
    parameter set, so neighbouring lanes take the same series expansion and term count; the trial sum only changes
    by summation order (~1e-16 relative).  Works on NumPy arrays and torch tensors."""
    if isinstance(x, np.ndarray):
        return x[np.argsort(np.abs(x), kind="stable")]
    import torch

    return x[torch.argsort(x.abs(), stable=True)]


def wiener_like_batch(x, params, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, w_outlier=0.1,
                      variant=None, out=None, stream=None, presort=None):
    """out[j] = wiener_like(x, *params[j, :7], err, ..., p_outlier=params[j, 7], w_outlier) for every row j.

    `x` (float64[n]) and `params` (float64[P, 8]: v, sv, a, z, sz, t, st, p_outlier) are either both NumPy arrays
    (host buffers; copies happen inside the call and a NumPy array is returned) or both torch CUDA tensors (device
    resident; the kernels are enqueued on torch's current stream and a torch tensor is returned without a sync).

    `presort`: reorder the RTs by |rt| before the launch (see sort_rts).  Default: on for host buffers when the grid
    is large enough to gain from it (P >= 64; done inside the library, on the device), off for device tensors (sort
    resident data once yourself).
    """
    var = _variant if variant is None else variant
    L = _lib.load()
    o = _opts(err, n_st, n_sz, use_adaptive, simps_err)
    if isinstance(x, np.ndarray):
        xs = _buffer(x, "x")
        ps = np.ascontiguousarray(params, dtype=np.double)
        if ps.ndim != 2 or ps.shape[1] != 8:
            raise ValueError("params must have shape (P, 8): v, sv, a, z, sz, t, st, p_outlier")
        res = np.empty(ps.shape[0], dtype=np.double) if out is None else out
        # the reordering happens inside the library (on the device for up to 32,768 RTs): no host argsort per call
        fn = L.wfpt_like_batch_host_presort if (presort or (presort is None and ps.shape[0] >= 64)) else L.wfpt_like_batch_host
        _lib.check(fn(_ptr(xs), xs.shape[0], _ptr(ps), ps.shape[0], *o, float(w_outlier), var, _ptr(res)))
        return res
    import torch  # the batching boundary: device tensors in, device tensor out

    if not (isinstance(x, torch.Tensor) and isinstance(params, torch.Tensor)):
        raise TypeError("x and params must both be numpy arrays or both be torch CUDA tensors")
    if not (x.is_cuda and params.is_cuda and x.device == params.device):
        raise ValueError("x and params must live on the same CUDA device")
    if x.dtype != torch.float64 or params.dtype != torch.float64:
        raise ValueError("x and params must be float64")
    if x.dim() != 1 or params.dim() != 2 or params.shape[1] != 8:
        raise ValueError("x must be 1-D and params (P, 8)")
    x = x.contiguous()
    if presort:
        x = sort_rts(x)
    params = params.contiguous()
    res = torch.empty(params.shape[0], dtype=torch.float64, device=x.device) if out is None else out
    with torch.cuda.device(x.device):
        s = torch.cuda.current_stream().cuda_stream if stream is None else stream
        _lib.check(L.wfpt_like_batch(x.data_ptr(), x.shape[0], params.data_ptr(), params.shape[0], *o, float(w_outlier),
                                     var, res.data_ptr(), s))
    return res


# ---- resident cache behind wiener_like ------------------------------------------------------------
_cache_on = True


def cache_config(enabled=True, max_entries=65536, max_bytes=2 << 30):
    """The resident-array cache behind `wiener_like`: on by default.  `enabled=False` makes every call upload its RTs."""
    global _cache_on
    _cache_on = bool(enabled)
    _lib.check(_lib.load().wfpt_cache_config(int(max_entries) if enabled else 0, int(max_bytes)))


def cache_clear():
    _lib.check(_lib.load().wfpt_cache_clear())


def cache_stats() -> dict:
    v = [C.c_int64() for _ in range(5)]
    _lib.check(_lib.load().wfpt_cache_stats(*[C.byref(k) for k in v]))
    return dict(zip(("entries", "bytes", "hits", "misses", "stale"), (k.value for k in v)))


# ---- src/wfpt.pyx:54-76 --------------------------------------------------------------------------------------------------
_F64 = np.dtype(np.float64)
_Prm8 = C.c_double * 8


try:  # CPython binding of the hot call (built beside the library; ctypes below does the same, a few microseconds slower)
    from . import _wfptcall as _fastcall
except ImportError:
    _fastcall = None


def wiener_like(x, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0,
                w_outlier=0.1):
    # The call a sampler makes millions of times: keep the Python side to a handful of microseconds.
    if _fastcall is not None:
        try:
            r = _fastcall.wiener_like(x, v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive, simps_err, p_outlier, w_outlier,
                                      _variant, _cache_on)
        except RuntimeError as exc:
            raise WfptError(str(exc)) from None
        if r is not None:
            return r
    if type(x) is not np.ndarray or x.dtype != _F64 or x.ndim != 1 or not x.flags.c_contiguous:
        x = _buffer(x, "x")                      # the reference's TypeError / ValueError, or a contiguous copy
    try:
        prm = _Prm8(v, sv, a, z, sz, t, st, p_outlier)
    except TypeError:                            # 0-d arrays and other things float() accepts
        prm = _Prm8(float(v), float(sv), float(a), float(z), float(sz), float(t), float(st), float(p_outlier))
    res = C.c_double()
    # the same array comes back thousands of times under MCMC: the library keeps it on the device (keyed by address and
    # length, every call compared byte for byte with the copy it holds) and only the parameters travel
    L = _lib.load()
    fn = L.wfpt_like_cached_host if _cache_on else L.wfpt_like_batch_host
    try:
        rc = fn(x.ctypes.data, x.shape[0], prm, 1, err, n_st, n_sz, 1 if use_adaptive else 0, simps_err, w_outlier,
                _variant, C.byref(res))
    except C.ArgumentError:                      # options given as 0-d arrays etc.
        rc = fn(x.ctypes.data, x.shape[0], prm, 1, *_opts(err, n_st, n_sz, use_adaptive, simps_err), float(w_outlier),
                _variant, C.byref(res))
    if rc:
        _lib.check(rc)
    return res.value


# ---- src/wfpt.pyx:357-387 ------------------------------------------------------------------------------------------------
_NAMES = ("v", "sv", "a", "z", "sz", "t", "st")


def _pack_multi(names, values, multi, n):
    keep, ptrs, scal = [], (C.c_void_p * len(names))(), (C.c_double * len(names))()
    for k, name in enumerate(names):
        if name in multi:
            val = values[name]
            arr = np.ascontiguousarray(np.asarray(val), dtype=np.double)
            if arr.ndim != 1 or arr.shape[0] < n:
                raise IndexError(f"per-trial parameter '{name}' must be a 1-D array with at least {n} entries")
            keep.append(arr)
            ptrs[k] = arr.ctypes.data
            scal[k] = 0.0
        else:
            ptrs[k] = None
            scal[k] = float(values[name])
    return keep, ptrs, scal


def wiener_like_multi(x, v, sv, a, z, sz, t, st, err, multi=None, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3,
                      p_outlier=0, w_outlier=0):
    x = _buffer(x, "x")
    if multi is None:
        # wfpt.pyx:366-367 hands the whole array to the scalar full_pdf, which cannot convert it
        raise TypeError("only size-1 arrays can be converted to Python scalars")
    vals = dict(v=v, sv=sv, a=a, z=z, sz=sz, t=t, st=st)
    for name in multi:
        if name not in vals:
            raise KeyError(name)
    n = x.shape[0]
    keep, ptrs, scal = _pack_multi(_NAMES, vals, multi, n)
    res = np.empty(1, dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_like_multi_host(_ptr(x), n, ptrs, scal, *_opts(err, n_st, n_sz, use_adaptive, simps_err),
                                      float(p_outlier), float(w_outlier), _variant, _ptr(res)))
    return float(res[0])


# ---- src/wfpt.pyx:78-153 -------------------------------------------------------------------------------------------------
def _group_by_split(split_by):
    """The grouping `for s in np.unique(split_by): mask = split_by == s` (wfpt.pyx:99,109-116) as index lists (the
    library's host utility: a counting sort, a few microseconds instead of NumPy's argsort + unique)."""
    n = split_by.shape[0]
    order = np.empty(n, dtype=np.int64)
    starts = np.empty(n + 1, dtype=np.int64)
    ng = C.c_int64()
    _lib.check(_lib.load().wfpt_group_by_split(split_by.ctypes.data, n, order.ctypes.data, starts.ctypes.data, C.byref(ng)))
    return order, starts[:ng.value + 1]


def wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha, v, sv, a, z, sz, t, st, err, n_st=10,
                      n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    x = _buffer(x, "x")
    response = _buffer(response, "response", np.int64)
    feedback = _buffer(feedback, "feedback")
    split_by = _buffer(split_by, "split_by", np.int64)
    n = x.shape[0]
    order, starts = _group_by_split(split_by)
    res = np.empty(1, dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_like_rlddm_host(_ptr(x), _ptr(response), _ptr(feedback), _ptr(order), _ptr(starts),
                                      len(starts) - 1, n, float(q), float(alpha), float(pos_alpha), float(v), float(sv),
                                      float(a), float(z), float(sz), float(t), float(st),
                                      *_opts(err, n_st, n_sz, use_adaptive, simps_err), float(p_outlier),
                                      float(w_outlier), _variant, _ptr(res)))
    return float(res[0])


# ---- src/wfpt.pyx:270-354 ------------------------------------------------------------------------------------------------
def wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, v, z, err=1e-4, n_st=10, n_sz=10,
                   use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0):
    response = _buffer(response, "response", np.int64)
    feedback = _buffer(feedback, "feedback")
    split_by = _buffer(split_by, "split_by", np.int64)
    order, starts = _group_by_split(split_by)
    res = np.empty(1, dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_like_rl_host(_ptr(response), _ptr(feedback), _ptr(order), _ptr(starts), len(starts) - 1,
                                   response.shape[0], float(q), float(alpha), float(pos_alpha), float(v), float(z),
                                   float(p_outlier), float(w_outlier), _ptr(res)))
    return float(res[0])


# ---- src/wfpt.pyx:390-433 ------------------------------------------------------------------------------------------------
def wiener_like_multi_rlddm(x, response, feedback, split_by, q, v, sv, a, z, sz, t, st, alpha, err, multi=None,
                            n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-3, p_outlier=0, w_outlier=0):
    x = _buffer(x, "x")
    response = _buffer(response, "response", np.int64)
    feedback = _buffer(feedback, "feedback")
    split_by = _buffer(split_by, "split_by", np.int64)
    if multi is None:
        raise TypeError("only size-1 arrays can be converted to Python scalars")  # wfpt.pyx:403-404, as in wiener_like_multi
    vals = dict(v=v, sv=sv, a=a, z=z, sz=sz, t=t, st=st, alpha=alpha)
    for name in multi:
        if name not in vals:
            raise KeyError(name)
    n = x.shape[0]
    keep, ptrs, scal = _pack_multi(_NAMES + ("alpha",), vals, multi, n)
    res = np.empty(1, dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_like_multi_rlddm_host(_ptr(x), _ptr(response), _ptr(feedback), _ptr(split_by), n, float(q), ptrs,
                                            scal, *_opts(err, n_st, n_sz, use_adaptive, simps_err), float(p_outlier),
                                            float(w_outlier), _variant, _ptr(res)))
    return float(res[0])


# ---- src/wfpt.pyx:551-580 ------------------------------------------------------------------------------------------------
def wiener_like_contaminant(x, cont_x, v, sv, a, z, sz, t, st, t_min, t_max, err, n_st=10, n_sz=10, use_adaptive=1,
                            simps_err=1e-8):
    x = _buffer(x, "x")
    cont_x = _buffer(cont_x, "cont_x", np.int32)
    res = np.empty(1, dtype=np.double)
    L = _lib.load()
    _lib.check(L.wfpt_like_contaminant_host(_ptr(x), _ptr(cont_x), x.shape[0], float(v), float(sv), float(a), float(z),
                                            float(sz), float(t), float(st), float(t_min), float(t_max),
                                            *_opts(err, n_st, n_sz, use_adaptive, simps_err), _variant, _ptr(res)))
    return float(res[0])


# ---- one launch for all likelihood nodes of a hierarchical model -----------------------------------
class ResidentData:
    """RT arrays of many likelihood nodes (one per subject x condition, /root/reference/hddm/models/base.py:1488-1493)
    kept resident on the GPU, so that an MCMC sweep sends only parameters.

    The observed RTs never change while sampling; only the 8 parameters of each node do.  `ResidentData(arrays)`
    uploads the arrays once (each |rt|-sorted, packed back to back on even offsets); `logp(params)` evaluates
    wiener_like for every node -- params[i] belongs to arrays[i] -- in ONE kernel launch and returns float64[len(arrays)].
    `logp(params, nodes=idx)` evaluates a subset (params[i] belongs to arrays[idx[i]]; a node may appear many times,
    e.g. several proposals for the same node).  The device memory belongs to the library (`wfpt_resident_*`): NumPy in,
    NumPy out needs nothing but ctypes; torch CUDA tensors as `params` are accepted too (result stays on the device,
    enqueued on torch's current stream)."""

    def __init__(self, arrays, device=None):
        arrs = [sort_rts(_buffer(np.asarray(a), "x")) for a in arrays]
        self.lengths = np.array([len(a) for a in arrs], dtype=np.int64)
        starts = np.zeros(len(arrs) + 1, dtype=np.int64)
        pos = 0
        for i, n in enumerate(self.lengths):
            starts[i] = pos
            pos += int(n) + (int(n) & 1)
        starts[-1] = pos
        packed = np.zeros(max(pos, 1), dtype=np.double)
        for a, s0 in zip(arrs, starts[:-1]):
            packed[s0:s0 + len(a)] = a
        self.starts = starts[:-1].copy()
        self.n_nodes = len(arrs)
        seg = np.ascontiguousarray(np.stack([self.starts, self.starts + self.lengths], axis=1), dtype=np.int64) \
            if arrs else np.zeros((0, 2), dtype=np.int64)
        self._seg_all = seg
        self.mean_len = int(max(1, self.lengths.mean())) if len(arrs) else 1
        self._device = None
        if device is not None:
            import torch

            self._device = torch.device(device)
        L = _lib.load()
        h = C.c_uint64()
        with self._on_device():
            _lib.check(L.wfpt_resident_create(packed.ctypes.data, packed.shape[0], seg.ctypes.data, self.n_nodes, C.byref(h)))
        self._h = h
        dx, dseg = C.c_void_p(), C.c_void_p()
        _lib.check(L.wfpt_resident_pointers(h, C.byref(dx), C.byref(dseg), None, None))
        self._x_ptr, self._dseg_ptr = dx.value, dseg.value

    def _on_device(self):
        """Context for library calls: torch's device guard when a device was named, nothing otherwise (the library
        then works on the CUDA runtime's current device, as every other entry point does)."""
        if self._device is None:
            import contextlib

            return contextlib.nullcontext()
        import torch

        return torch.cuda.device(self._device)

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                _lib.load().wfpt_resident_release(h)
            except Exception:  # interpreter shutdown
                pass

    __del__ = close

    def logp(self, params, nodes=None, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, w_outlier=0.1,
             variant=None):
        var = _variant if variant is None else variant
        L = _lib.load()
        o = _opts(err, n_st, n_sz, use_adaptive, simps_err)
        idx = None if nodes is None else np.ascontiguousarray(nodes, dtype=np.int64)
        P = self.n_nodes if idx is None else idx.shape[0]
        if isinstance(params, np.ndarray):
            ps = params if (params.dtype == _F64 and params.flags.c_contiguous) else np.ascontiguousarray(params, dtype=np.double)
            if ps.shape != (P, 8):
                raise ValueError(f"params must have shape ({P}, 8)")
            res = np.empty(P, dtype=np.double)
            if P:
                args = (self._h, None if idx is None else idx.ctypes.data, ps.ctypes.data, P, *o, float(w_outlier), var,
                        res.ctypes.data)
                if self._device is None:
                    rc = L.wfpt_resident_logp(*args)
                else:
                    with self._on_device():
                        rc = L.wfpt_resident_logp(*args)
                if rc:
                    _lib.check(rc)
            return res
        import torch  # device-resident parameters: torch owns them and the stream

        ps = params.contiguous()
        if ps.shape != (P, 8):
            raise ValueError(f"params must have shape ({P}, 8)")
        out = torch.empty(P, dtype=torch.float64, device=ps.device)
        if P == 0:
            return out
        if idx is None:
            dseg_ptr, mean_len, keep = self._dseg_ptr, self.mean_len, None
        else:
            keep = torch.from_numpy(self._seg_all[idx].reshape(-1)).to(ps.device)
            dseg_ptr, mean_len = keep.data_ptr(), int(max(1, self.lengths[idx].mean()))
        with torch.cuda.device(ps.device):
            s = torch.cuda.current_stream().cuda_stream
            _lib.check(L.wfpt_like_segments(self._x_ptr, dseg_ptr, mean_len, ps.data_ptr(), P, *o, float(w_outlier), var,
                                            out.data_ptr(), s))
        return out


def wiener_like_segments(arrays, params, **kw):
    """One-shot form of ResidentData: out[i] = wiener_like(arrays[i], *params[i, :7], ..., p_outlier=params[i, 7])."""
    return ResidentData(arrays).logp(np.asarray(params, dtype=np.double), **kw)


# ---- src/wfpt.pyx:517-548 ------------------------------------------------------------------------------------------------
def gen_rts_from_cdf(v, sv, a, z, sz, t, st, samples=1000, cdf_lb=-6, cdf_ub=6, dt=1e-2):
    v, sv, a, z, sz, t, st = (float(v), float(sv), float(a), float(z), float(sz), float(t), float(st))
    samples = int(samples)
    x = np.arange(float(cdf_lb), float(cdf_ub), float(dt))
    l_cdf = np.empty(x.shape[0], dtype=np.double)
    l_cdf[0] = 0
    # full_pdf(x[i], v, sv, a, z, sz, 0, 0, 1e-4) with full_pdf's own defaults (n=2, adaptive, 1e-3), summed in order
    l_cdf[1:] = np.cumsum(pdf_array(x[1:], v, sv, a, z, sz, 0.0, 0.0, 1e-4, 0, 2, 2, 1, 1e-3, 0, 0))
    l_cdf /= l_cdf[x.shape[0] - 1]
    f = np.random.rand(samples)
    if st != 0:
        delay = np.random.rand(samples) * st + (t - st / 2.0)
    idx = np.searchsorted(l_cdf, f)
    rt = x[idx]
    if st == 0:
        return rt + np.sign(rt) * t
    return rt + np.sign(rt) * delay


# ---- src/wfpt.pyx:582-607 ------------------------------------------------------------------------------------------------
def gen_cdf_using_pdf(v, sv, a, z, sz, t, st, err, N=500, time=5.0, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                      p_outlier=0, w_outlier=0):
    v, sv, a, z, sz, t, st = (float(v), float(sv), float(a), float(z), float(sz), float(t), float(st))
    p_outlier = float(p_outlier)
    if (sv < 0) or (a <= 0) or (z < 0) or (z > 1) or (sz < 0) or (sz > 1) or (z + sz / 2.0 > 1) or \
            (z - sz / 2.0 < 0) or (t - st / 2.0 < 0) or (t < 0) or (st < 0) or not (0 <= p_outlier <= 1):
        raise ValueError("at least one of the parameters is out of the support")
    N = int(N)
    x = np.linspace(-float(time), float(time), 2 * N + 1)
    cdf_array = pdf_array(x, v, sv, a, z, sz, t, st, err, 0, n_st, n_sz, use_adaptive, simps_err, p_outlier, w_outlier)
    # integrate.cumtrapz(y) == cumulative trapezoid with unit spacing (removed from SciPy >= 1.14; same arithmetic)
    cdf_array[1:] = np.cumsum((cdf_array[1:] + cdf_array[:-1]) / 2.0)
    cdf_array /= cdf_array[x.shape[0] - 1]
    return x, cdf_array


# ---- src/wfpt.pyx:610-626 ------------------------------------------------------------------------------------------------
def split_cdf(x, data):
    x = _buffer(x, "x")
    data = _buffer(data, "data")
    N = (len(data) - 1) // 2
    x_lb = -x[:N][::-1]
    lb = data[:N][::-1]
    lb = np.cumsum(np.concatenate([np.array([0]), -np.diff(lb)]))
    x_ub = x[N + 1:]
    ub = data[N + 1:]
    ub -= ub[0]  # in place on a view of `data`, as the reference does
    return (x_lb, lb, x_ub, ub)
