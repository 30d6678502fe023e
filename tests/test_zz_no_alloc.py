"""GPU test of wfpt_like_batch_noalloc, the batched likelihood for callers that cannot allocate device memory
(their GPU memory belongs to another pool, or they cannot afford cudaMalloc's device-wide synchronisation, or they
run where allocation calls are not allowed).  Its partial sums live in a fixed array inside the library instead of
cudaMalloc'ed scratch, so the array is shared by all calls and still holds the previous call's data on entry.

The test calls the entry point through the C ABI with device pointers, twice in a row with different inputs and
different launch geometries, and compares each result with wfpt_like_batch on the same inputs (tolerance 1e-10
relative, the bar of the other parity tests; -inf / NaN must agree exactly).  It also checks that a batch needing
more scratch slots than the array holds is refused with a clear message instead of being truncated.

Collected last on purpose (file name), and independent of the other tests: it passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_SLOTS = 16384  # WFPT_NOALLOC_MAX_SLOTS in include/wfpt_b200.h
OPTS = (1e-4, 2, 2, 0, 1e-3)  # err, n_st, n_sz, use_adaptive, simps_err


def _inputs(seed, n, P):
    rng = np.random.default_rng(seed)
    x = np.where(rng.random(n) < 0.75, 1.0, -1.0) * rng.uniform(0.3, 3.5, n)
    prm = np.stack([rng.uniform(-3, 3, P), rng.uniform(0, 2, P), rng.uniform(0.8, 3, P), rng.uniform(0.3, 0.7, P),
                    rng.uniform(0, 0.2, P), rng.uniform(0.05, 0.4, P), rng.uniform(0, 0.2, P),
                    rng.choice([0.0, 0.05], P)], axis=1)
    return x, prm


def _both(L, check, torch, x, prm, variant):
    dx = torch.from_numpy(x).cuda()
    dp = torch.from_numpy(np.ascontiguousarray(prm)).cuda()
    ref = torch.empty(len(prm), dtype=torch.float64, device="cuda")
    got = torch.full((len(prm),), 123.0, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    check(L.wfpt_like_batch(dx.data_ptr(), len(x), dp.data_ptr(), len(prm), *OPTS, 0.1, variant, ref.data_ptr(), stream))
    check(L.wfpt_like_batch_noalloc(dx.data_ptr(), len(x), dp.data_ptr(), len(prm), *OPTS, 0.1, variant, got.data_ptr(), stream))
    torch.cuda.synchronize()
    return got.cpu().numpy(), ref.cpu().numpy()


def _agree(got, ref):
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    assert np.array_equal(np.isinf(got), np.isinf(ref))
    fin = np.isfinite(ref)
    assert np.array_equal(got[~fin & ~np.isnan(ref)], ref[~fin & ~np.isnan(ref)])
    assert np.all(np.abs(got[fin] - ref[fin]) <= 1e-10 * np.abs(ref[fin]))


@pytest.mark.parametrize("variant", [0, 1], ids=["fast", "faithful"])
def test_noalloc_matches_like_batch_on_consecutive_calls(variant):
    import torch

    from hddm_b200 import _lib

    L = _lib.load()
    # call 1: few sets, many trials -> many trial slices per set, scratch holds (partial, flag) pairs
    x1, p1 = _inputs(1, 6001, 3)
    got, ref = _both(L, _lib.check, torch, x1, p1, variant)
    assert np.all(np.isfinite(ref) | np.isneginf(ref))
    _agree(got, ref)
    # call 2, straight after, different data and a different geometry (one slice per set): whatever call 1 left in
    # the shared array must not leak into these sums
    x2, p2 = _inputs(2, 1500, 2000)
    p2[7, 3] = 1.4      # one set outside the support with p_outlier = 0 -> -inf
    p2[7, 7] = 0.0
    got, ref = _both(L, _lib.check, torch, x2, p2, variant)
    assert np.isneginf(ref[7])
    _agree(got, ref)
    # and the first geometry again, now on top of call 2's leftovers
    got, ref = _both(L, _lib.check, torch, x1, p1, variant)
    _agree(got, ref)


def test_noalloc_leaves_free_memory_unchanged():
    """The point of the entry point: a call of a shape the process has never run before must not take device memory.
    cudaMemGetInfo before and after (through torch, which only reports the driver's numbers) must agree."""
    import torch

    from hddm_b200 import _lib

    L = _lib.load()
    x, prm = _inputs(5, 7001, 5)            # 5 sets x many trial slices: scratch needed
    dx = torch.from_numpy(x).cuda()
    dp = torch.from_numpy(np.ascontiguousarray(prm)).cuda()
    out = torch.empty(len(prm), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    # first call on this device (symbol look-up) outside the measurement, as the header asks
    _lib.check(L.wfpt_like_batch_noalloc(dx.data_ptr(), 16, dp.data_ptr(), 1, *OPTS, 0.1, 0, out.data_ptr(), stream))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for variant in (0, 1):
        _lib.check(L.wfpt_like_batch_noalloc(dx.data_ptr(), len(x), dp.data_ptr(), len(prm), *OPTS, 0.1, variant,
                                             out.data_ptr(), stream))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)


def test_noalloc_rejects_batches_beyond_its_scratch():
    import torch

    from hddm_b200 import _lib

    L = _lib.load()
    x, prm = _inputs(3, 64, MAX_SLOTS + 1)
    dx = torch.from_numpy(x).cuda()
    dp = torch.from_numpy(np.ascontiguousarray(prm)).cuda()
    out = torch.zeros(len(prm), dtype=torch.float64, device="cuda")
    rc = L.wfpt_like_batch_noalloc(dx.data_ptr(), len(x), dp.data_ptr(), len(prm), *OPTS, 0.1, 0, out.data_ptr(), None)
    assert rc != 0
    msg = L.wfpt_last_error().decode()
    assert "scratch" in msg and str(MAX_SLOTS) in msg, msg
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0  # nothing was launched
    # exactly at the limit it runs
    x, prm = _inputs(4, 64, MAX_SLOTS)
    got, ref = _both(L, _lib.check, torch, x, prm, 0)
    _agree(got, ref)
