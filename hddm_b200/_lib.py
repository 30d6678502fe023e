"""ctypes binding of libwfpt_b200.so (C ABI: include/wfpt_b200.h).  Lazy, loud, no fallback."""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("WFPT_B200_LIB") or os.path.join(HERE, "libwfpt_b200.so")   # (override: A/B runs of two builds)

c_double_p = C.POINTER(C.c_double)
c_int64_p = C.POINTER(C.c_int64)
c_float_p = C.POINTER(C.c_float)
c_int32_p = C.POINTER(C.c_int32)

FAST = 0
FAITHFUL = 1

_lock = threading.Lock()
_lib = None


class WfptError(RuntimeError):
    pass


def _declare(L):
    d, i, i64, vp = C.c_double, C.c_int, C.c_int64, C.c_void_p
    L.wfpt_last_error.restype = C.c_char_p
    L.wfpt_last_error.argtypes = []
    L.wfpt_version.restype = i
    L.wfpt_launch_count.restype = i64
    L.wfpt_device_info.argtypes = [C.c_char_p, i, C.POINTER(i), C.POINTER(i), C.POINTER(i64)]
    L.wfpt_release_scratch.argtypes = []
    L.wfpt_fp64_peak.argtypes = [i, c_double_p, c_double_p]
    opts = [d, i, i, i, d]  # err, n_st, n_sz, use_adaptive, simps_err
    L.wfpt_like_batch.argtypes = [vp, i64, vp, i64] + opts + [d, i, vp, vp]
    L.wfpt_like_batch_host.argtypes = [vp, i64, vp, i64] + opts + [d, i, vp]
    L.wfpt_like_batch_host_presort.argtypes = [vp, i64, vp, i64] + opts + [d, i, vp]
    L.wfpt_like_cached_host.argtypes = [vp, i64, vp, i64] + opts + [d, i, vp]
    L.wfpt_sort_rts.argtypes = [vp, i64, vp, vp]
    L.wfpt_cache_config.argtypes = [i64, i64]
    L.wfpt_cache_clear.argtypes = []
    L.wfpt_cache_stats.argtypes = [vp, vp, vp, vp, vp]
    for name in ("wfpt_like_batch_host_presort", "wfpt_like_cached_host", "wfpt_sort_rts", "wfpt_cache_config",
                 "wfpt_cache_clear", "wfpt_cache_stats"):
        getattr(L, name).restype = i
    L.wfpt_like_batch_noalloc.argtypes = [vp, i64, vp, i64] + opts + [d, i, vp, vp]
    L.wfpt_like_batch_noalloc.restype = i
    L.wfpt_pdf_array.argtypes = [vp, i64, c_double_p] + opts + [d, i, i, vp, vp]
    L.wfpt_pdf_array_host.argtypes = [vp, i64, c_double_p] + opts + [d, i, i, vp]
    L.wfpt_like_multi.argtypes = [vp, i64, C.POINTER(vp), c_double_p] + opts + [d, d, i, vp, vp]
    L.wfpt_like_multi_host.argtypes = [vp, i64, C.POINTER(vp), c_double_p] + opts + [d, d, i, vp]
    L.wfpt_like_rlddm_host.argtypes = [vp, vp, vp, vp, vp, i64, i64] + [d] * 10 + opts + [d, d, i, vp]
    L.wfpt_ktable.argtypes = [d, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(i), c_double_p, C.POINTER(i), i]
    L.wfpt_ktable.restype = i
    L.wfpt_ktable_decision.argtypes = [d, d]
    L.wfpt_ktable_decision.restype = i
    L.wfpt_ktable_decision_array.argtypes = [vp, i64, d, vp]
    L.wfpt_ktable_decision_array.restype = i
    L.wfpt_kflip_census.argtypes = [vp, i64, vp, i64, d, i, i, i, vp, vp]
    L.wfpt_kflip_census.restype = i
    for name in ("wfpt_device_info", "wfpt_release_scratch", "wfpt_fp64_peak", "wfpt_like_batch",
                 "wfpt_like_batch_host", "wfpt_pdf_array", "wfpt_pdf_array_host", "wfpt_like_multi",
                 "wfpt_like_multi_host", "wfpt_like_rlddm_host"):
        getattr(L, name).restype = i
    L.wfpt_like_segments.argtypes = [vp, vp, i64, vp, i64] + opts + [d, i, vp, vp]
    L.wfpt_like_segments_host.argtypes = [vp, vp, vp, i64] + opts + [d, i, vp]
    L.wfpt_like_segments_resident.argtypes = [vp, vp, i64, vp, i64] + opts + [d, i, vp]
    L.wfpt_resident_create.argtypes = [vp, i64, vp, i64, vp]
    L.wfpt_resident_logp.argtypes = [C.c_uint64, vp, vp, i64] + opts + [d, i, vp]
    L.wfpt_resident_pointers.argtypes = [C.c_uint64, vp, vp, vp, vp]
    L.wfpt_resident_release.argtypes = [C.c_uint64]
    for name in ("wfpt_resident_create", "wfpt_resident_logp", "wfpt_resident_pointers", "wfpt_resident_release"):
        getattr(L, name).restype = i
    L.wfpt_like_contaminant_host.argtypes = [vp, vp, i64] + [d] * 9 + opts + [i, vp]
    L.wfpt_like_rl_host.argtypes = [vp, vp, vp, vp, i64, i64] + [d] * 7 + [vp]
    L.wfpt_group_by_split.argtypes = [vp, i64, vp, vp, vp]
    L.wfpt_group_by_split.restype = i
    L.wfpt_like_multi_rlddm_host.argtypes = [vp, vp, vp, vp, i64, d, C.POINTER(vp), c_double_p] + opts + [d, d, i, vp]
    for name in ("wfpt_like_segments", "wfpt_like_segments_host", "wfpt_like_segments_resident", "wfpt_like_contaminant_host", "wfpt_like_rl_host",
                 "wfpt_like_multi_rlddm_host"):
        getattr(L, name).restype = i
    L.wfpt_dmat_cdf_array_host.argtypes = [vp, i64] + [d] * 9 + [vp]
    L.wfpt_dmat_cdf_array_host.restype = i
    f = C.c_float
    sim_tail = [i, i, i64, i64, f, f, f, C.c_uint64, vp, vp, vp, i, f, vp, vp, i]
    L.wfpt_sim_run.argtypes = [i, vp] + sim_tail + [vp]
    L.wfpt_sim_run.restype = i
    L.wfpt_sim_run_sharded.argtypes = [i, vp, i, i, i, i, i64, i64, f, f, f, C.c_uint64, vp, vp, vp, i, f, vp, vp, i, vp]
    L.wfpt_sim_run_sharded.restype = i
    L.wfpt_sim_run_host.argtypes = [i, vp] + sim_tail
    L.wfpt_sim_run_host.restype = i
    L.wfpt_sim_race_run_host.argtypes = [i, i, i, vp, i, i64, i64, f, f, f, C.c_uint64, vp, vp, vp, i, f, vp]
    L.wfpt_sim_race_run_host.restype = i
    L.wfpt_sim_par2_run_host.argtypes = [i, vp, i, i64, i64, f, f, f, C.c_uint64, vp, vp, vp, i, f, vp]
    L.wfpt_sim_par2_run_host.restype = i
    L.wfpt_sim_hier_run_host.argtypes = [i, i, vp, i, i64, i64, f, f, f, C.c_uint64, vp, vp, vp, i, f, vp]
    L.wfpt_sim_hier_run_host.restype = i
    L.wfpt_sim_n_params.argtypes = [i]
    L.wfpt_sim_n_params.restype = i


def load():
    """Returns the loaded library; raises WfptError when it has not been built (python -m hddm_b200.build)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise WfptError(
                        f"{LIB_PATH} is missing: build it with `python -m hddm_b200.build` (needs nvcc). "
                        "hddm_b200 has no CPU fallback."
                    )
                L = C.CDLL(LIB_PATH)
                _declare(L)
                _lib = L
    return _lib


def check(rc: int):
    if rc != 0:
        msg = load().wfpt_last_error()
        raise WfptError(msg.decode("utf-8", "replace") if msg else f"libwfpt_b200 call failed (rc={rc})")


def launch_count() -> int:
    return int(load().wfpt_launch_count())


def device_info() -> dict:
    L = load()
    name = C.create_string_buffer(256)
    sm, khz, mem = C.c_int(0), C.c_int(0), C.c_int64(0)
    check(L.wfpt_device_info(name, 256, C.byref(sm), C.byref(khz), C.byref(mem)))
    return {"name": name.value.decode(), "sm_count": sm.value, "clock_khz": khz.value, "mem_bytes": mem.value}


def ktable(err: float) -> dict:
    """The fast variant's series-term table for `err` (host computation; works without a GPU)."""
    L = load()
    n, klm, ksm, d0 = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
    thr = (C.c_double * 32)()
    desc = (C.c_int * 32)()
    check(L.wfpt_ktable(float(err), C.byref(n), C.byref(klm), C.byref(ksm), C.byref(d0), thr, desc, 32))
    m = max(n.value, 0)
    return {"n": n.value, "kl_max": klm.value, "ks_max": ksm.value, "desc0": d0.value,
            "thr": [thr[k] for k in range(m)], "desc": [desc[k] for k in range(m)]}


def ktable_decision(tt: float, err: float) -> int:
    return int(load().wfpt_ktable_decision(float(tt), float(err)))


def ktable_decision_array(tt, err: float):
    """The reference's series-term decision (host libm, src/pdf.pxi:35-60) for an array of tt; int32 descriptors."""
    import numpy as np

    tt = np.ascontiguousarray(tt, dtype=np.float64)
    out = np.empty(tt.shape[0], dtype=np.int32)
    check(load().wfpt_ktable_decision_array(tt.ctypes.data, tt.shape[0], float(err), out.ctypes.data))
    return out


def fp64_peak(iters: int = 8192) -> dict:
    L = load()
    rate, ms = C.c_double(0), C.c_double(0)
    check(L.wfpt_fp64_peak(iters, C.byref(rate), C.byref(ms)))
    return {"dfma_per_s": rate.value, "ms": ms.value}
