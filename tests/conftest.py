import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    return dict(np.load(os.path.join(GOLDEN, "wfpt_golden.npz")))


@pytest.fixture(scope="session")
def grid_golden():
    """Reference answers for the whole benchmark grids (tests/golden/make_grid_golden.py) with the grids themselves,
    regenerated from their seeds and checked against the digests stored beside the answers."""
    import hashlib

    import bench

    g = dict(np.load(os.path.join(GOLDEN, "grid_golden.npz")))
    x, benign = bench.workload(bench.P_FULL)
    bounds = bench.param_bounds_grid(8192)
    sha = lambda a: np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)
    assert np.array_equal(sha(benign), g["benign_sha"]) and np.array_equal(sha(benign[:8192]), g["benign8192_sha"])
    assert np.array_equal(sha(bounds), g["bounds_sha"])
    g.update(x=x, benign=benign, bounds=bounds)
    return g


@pytest.fixture(scope="session")
def oracle():
    from oracle import cpu

    cpu.lib()
    return cpu


@pytest.fixture(scope="session")
def ref_wfpt():
    """The reference's own Cython build (oracle/_ref); skipped where it was not built."""
    from oracle.build_ref import load_ref, ref_built

    if not ref_built():
        pytest.skip("oracle/_ref not built on this machine")
    return load_ref()[0]


@pytest.fixture(scope="session")
def wfpt_gpu():
    """The product module, on a machine with a GPU and the built library."""
    from hddm_b200 import _lib, wfpt

    _lib.load()
    return wfpt


def rel_err(got, want):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    same = (got == want) | (np.isnan(got) & np.isnan(want))
    with np.errstate(all="ignore"):
        r = np.abs(got - want) / np.abs(want)
    r = np.where(same, 0.0, r)
    return float(np.max(r)) if r.size else 0.0
