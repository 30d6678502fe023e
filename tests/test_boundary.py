"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/*.h declares, the
product never touches the oracle, and the reference-shaped Python callables reproduce the Cython argument errors
(SURVEY 8(b)).  No compute call is made here (no GPU)."""
import ctypes
import glob
import inspect
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    syms = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        txt = open(h).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        syms += re.findall(r"\b(wfpt_[a-z0-9_]+)\s*\(", txt)
    return sorted(set(syms))


def test_library_exports_every_declared_symbol():
    from hddm_b200 import _lib

    assert os.path.exists(_lib.LIB_PATH), "run python -m hddm_b200.build"
    L = ctypes.CDLL(_lib.LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/ but not exported"
    L.wfpt_version.restype = ctypes.c_int
    assert L.wfpt_version() >= 100
    _lib.load()  # argtypes declaration must not fail either


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "hddm_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "liboracle" not in txt and "oracle/" not in txt, f
    assert "oracle" not in open(os.path.join(ROOT, "wfpt.py")).read()
    assert "oracle" not in open(os.path.join(ROOT, "cdfdif_wrapper.py")).read()


def test_top_level_module_name_and_callables():
    import wfpt

    for name in ("wiener_like", "wiener_like_multi", "pdf_array", "wiener_like_rlddm", "full_pdf", "gen_rts_from_cdf",
                 "gen_cdf_using_pdf", "split_cdf", "wiener_like_contaminant", "wiener_like_rl",
                 "wiener_like_multi_rlddm", "wiener_like_nn_mlp", "wiener_like_nn_mlp_info", "wiener_like_nn_mlp_pdf",
                 "wiener_like_multi_nn_mlp", "wiener_like_multi_nn_mlp_pdf", "wiener_like_rlssm_nn",
                 "wiener_like_rlssm_nn_reg"):
        assert callable(getattr(wfpt, name)), name


def test_signatures_match_reference():
    """Positional order and defaults of /root/reference/src/wfpt.pyx:32-34,54-56,78-85,357-359 and pdf.pxi:104-107."""
    import wfpt

    def sig(f):
        return [(p.name, p.default) for p in inspect.signature(f).parameters.values()]

    E = inspect.Parameter.empty
    assert sig(wfpt.wiener_like) == [("x", E), ("v", E), ("sv", E), ("a", E), ("z", E), ("sz", E), ("t", E), ("st", E),
                                     ("err", E), ("n_st", 10), ("n_sz", 10), ("use_adaptive", 1), ("simps_err", 1e-8),
                                     ("p_outlier", 0), ("w_outlier", 0.1)]
    assert sig(wfpt.pdf_array) == [("x", E), ("v", E), ("sv", E), ("a", E), ("z", E), ("sz", E), ("t", E), ("st", E),
                                   ("err", 1e-4), ("logp", 0), ("n_st", 2), ("n_sz", 2), ("use_adaptive", 1),
                                   ("simps_err", 1e-3), ("p_outlier", 0), ("w_outlier", 0)]
    assert sig(wfpt.wiener_like_multi) == [("x", E), ("v", E), ("sv", E), ("a", E), ("z", E), ("sz", E), ("t", E),
                                           ("st", E), ("err", E), ("multi", None), ("n_st", 10), ("n_sz", 10),
                                           ("use_adaptive", 1), ("simps_err", 1e-3), ("p_outlier", 0), ("w_outlier", 0)]
    assert sig(wfpt.wiener_like_rlddm) == [("x", E), ("response", E), ("feedback", E), ("split_by", E), ("q", E),
                                           ("alpha", E), ("pos_alpha", E), ("v", E), ("sv", E), ("a", E), ("z", E),
                                           ("sz", E), ("t", E), ("st", E), ("err", E), ("n_st", 10), ("n_sz", 10),
                                           ("use_adaptive", 1), ("simps_err", 1e-8), ("p_outlier", 0), ("w_outlier", 0)]
    assert sig(wfpt.full_pdf) == [("x", E), ("v", E), ("sv", E), ("a", E), ("z", E), ("sz", E), ("t", E), ("st", E),
                                  ("err", E), ("n_st", 2), ("n_sz", 2), ("use_adaptive", 1), ("simps_err", 1e-3)]
    assert sig(wfpt.gen_rts_from_cdf)[7:] == [("samples", 1000), ("cdf_lb", -6), ("cdf_ub", 6), ("dt", 1e-2)]


def test_buffer_argument_errors():
    """Cython buffer semantics: the checks fire before any device work, so they are testable without a GPU."""
    import wfpt

    with pytest.raises(TypeError, match="Argument 'x' has incorrect type"):
        wfpt.wiener_like([0.5, 0.6], 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)
    with pytest.raises(ValueError, match="Buffer dtype mismatch, expected 'double' but got 'float'"):
        wfpt.wiener_like(np.ones(3, dtype=np.float32), 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)
    with pytest.raises(ValueError, match=r"Buffer has wrong number of dimensions \(expected 1, got 2\)"):
        wfpt.pdf_array(np.ones((2, 2)), 1, 0, 2, 0.5, 0, 0.3, 0)
    with pytest.raises(ValueError, match="expected 'long' but got 'double'"):
        wfpt.wiener_like_rlddm(np.ones(3), np.ones(3), np.ones(3), np.ones(3, dtype=np.int64), 0.5, 0, 100.0, 1, 0, 2,
                               0.5, 0, 0.3, 0, 1e-4)
    with pytest.raises(TypeError):
        wfpt.wiener_like_multi(np.ones(3), 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)  # multi=None is broken in the reference too
    with pytest.raises(ValueError, match="out of the support"):
        wfpt.gen_cdf_using_pdf(1, 0, -2, 0.5, 0, 0.3, 0, 1e-4)


def test_split_cdf_host_logic():
    import wfpt

    x = np.linspace(-2, 2, 9)
    data = np.linspace(0, 1, 9) ** 2
    want_ub = data[5:] - data[5]
    x_lb, lb, x_ub, ub = wfpt.split_cdf(x, data.copy())
    assert np.allclose(x_lb, -x[:4][::-1]) and np.allclose(x_ub, x[5:])
    assert lb[0] == 0 and np.all(np.diff(lb) >= 0)
    assert np.allclose(ub, want_ub)


def test_import_is_lazy():
    """Import must not load CUDA or the shared library (models get pickled / re-imported, base.py:1394-1403)."""
    import subprocess
    import sys

    code = "import sys; sys.path.insert(0, %r); import wfpt, hddm_b200._lib as L; assert L._lib is None; print('ok')" % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr


def test_group_by_split_matches_numpy():
    """wfpt_group_by_split (host utility, no device work) == np.argsort(kind='stable') + np.unique, the grouping of
    /root/reference/src/wfpt.pyx:99,109-116."""
    from hddm_b200 import wfpt as W

    rng = np.random.default_rng(0)
    cases = [rng.integers(0, 3, 1000), rng.integers(-5, 7, 517), rng.integers(0, 2**40, 300) * 3, np.zeros(5),
             np.array([]), np.array([7, 7, 2, 9, 2])]
    for split in cases:
        split = np.ascontiguousarray(split, dtype=np.int64)
        order, starts = W._group_by_split(split)
        uniq, counts = np.unique(split, return_counts=True)
        assert np.array_equal(order, np.argsort(split, kind="stable"))
        assert np.array_equal(starts, np.concatenate([[0], np.cumsum(counts)]))
