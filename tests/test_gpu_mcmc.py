"""GPU tests of the sampler-facing pieces: the resident-array cache behind the reference-shaped `wiener_like` call
(SURVEY 8(f) rank 1), the |rt| reordering done inside the library, and the likelihood adapter on the CUDA path."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu
KW = dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)


def _rts(seed, n):
    rng = np.random.default_rng(seed)
    return np.where(rng.random(n) < 0.8, 1.0, -1.0) * rng.uniform(0.35, 3.0, n)


def test_cached_call_is_bit_identical_and_notices_mutation(oracle, wfpt_gpu):
    w = wfpt_gpu
    w.cache_config(True)
    w.cache_clear()
    s0 = w.cache_stats()
    for n in (1, 37, 500, 1000, 4099, 10000):
        x = _rts(n, n)
        args = (0.9, 0.8, 1.9, 0.48, 0.15, 0.3, 0.1, 1e-4)
        w.cache_config(False)
        plain = w.wiener_like(x, *args, **KW)
        w.cache_config(True)
        first = w.wiener_like(x, *args, **KW)           # miss: uploads
        again = w.wiener_like(x, *args, **KW)           # hit: parameters only
        assert plain == first == again, n
        other = w.wiener_like(x, 1.1, *args[1:], **KW)  # hit with other parameters
        assert other == pytest.approx(oracle.wiener_like(x, 1.1, *args[1:], **KW), rel=1e-10)
        # in-place mutation of ONE element (not the first, not the last): same address, same length, new contents
        x[n // 2] = -x[n // 2] if n > 2 else x[n // 2] + 0.25
        mutated = w.wiener_like(x, *args, **KW)
        w.cache_config(False)
        assert mutated == w.wiener_like(x, *args, **KW), n
        w.cache_config(True)
        if n > 2:
            assert mutated != plain
        # a view into the same buffer with another length is another key
        if n > 40:
            assert w.wiener_like(x[:n - 3], *args, **KW) == pytest.approx(oracle.wiener_like(x[:n - 3], *args, **KW), rel=1e-10)
    s1 = w.cache_stats()
    assert s1["hits"] - s0["hits"] >= 12 and s1["stale"] - s0["stale"] == 6 and s1["entries"] >= 6
    # eviction: a two-entry cache cycles without error and keeps answering correctly
    w.cache_config(True, max_entries=2)
    xs = [_rts(100 + k, 300) for k in range(5)]
    for rep in range(2):
        for x in xs:
            assert w.wiener_like(x, 1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 1e-4) == pytest.approx(
                oracle.wiener_like(x, 1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 1e-4), rel=1e-10)
    assert w.cache_stats()["entries"] <= 2
    w.cache_config(True)
    w.cache_clear()
    assert w.cache_stats()["entries"] == 0


def test_library_side_reordering(wfpt_gpu):
    """wfpt_sort_rts / the presort entry point against np.argsort(|x|, stable): same order (ties, zeros, signs), same sums."""
    import torch

    from hddm_b200 import _lib

    L = _lib.load()
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 63, 64, 65, 255, 256, 257, 1023, 1025, 10000, 32768):
        x = _rts(n, n)
        x[rng.integers(0, n, max(1, n // 10))] = x[rng.integers(0, n, max(1, n // 10))]   # ties
        if n > 4:
            x[1], x[3] = 0.0, -0.0
        dx = torch.from_numpy(x).cuda()
        out = torch.empty_like(dx)
        _lib.check(L.wfpt_sort_rts(dx.data_ptr(), n, out.data_ptr(), None))
        torch.cuda.synchronize()
        want = x[np.argsort(np.abs(x), kind="stable")]
        got = out.cpu().numpy()
        assert np.array_equal(got, want) and np.array_equal(np.signbit(got), np.signbit(want)), n
    x = _rts(9, 10000)
    prm = np.stack([rng.uniform(-3, 3, 200), rng.uniform(0, 2, 200), rng.uniform(1, 3, 200), rng.uniform(0.3, 0.7, 200),
                    rng.uniform(0, 0.2, 200), rng.uniform(0.1, 0.3, 200), rng.uniform(0, 0.2, 200), np.full(200, 0.05)], axis=1)
    a = wfpt_gpu.wiener_like_batch(x, prm, 1e-4, 2, 2, 0, 1e-3, 0.1)                                   # sorted inside the library
    b = wfpt_gpu.wiener_like_batch(wfpt_gpu.sort_rts(x), prm, 1e-4, 2, 2, 0, 1e-3, 0.1, presort=False)  # sorted by NumPy
    assert np.array_equal(a, b)
    big = _rts(10, 40000)                                                                              # host-sorted inside the library
    a = wfpt_gpu.wiener_like_batch(big, prm, 1e-4, 2, 2, 0, 1e-3, 0.1)
    b = wfpt_gpu.wiener_like_batch(wfpt_gpu.sort_rts(big), prm, 1e-4, 2, 2, 0, 1e-3, 0.1, presort=False)
    assert np.array_equal(a, b)
    with pytest.raises(wfpt_gpu.WfptError):
        _lib.check(L.wfpt_sort_rts(dx.data_ptr(), 40000, out.data_ptr(), None))


def test_adapter_on_the_cuda_path(oracle, wfpt_gpu):
    import pandas as pd

    from hddm_b200 import likelihoods

    wfpt_gpu.cache_config(True)
    rt = _rts(3, 500)
    frame = pd.DataFrame({"rt": rt})
    cls = likelihoods.generate_wfpt_stochastic_class()
    node = cls(value=frame, v=1.0, sv=0.5, a=2.0, z=0.5, sz=0.1, t=0.3, st=0.1, p_outlier=0.05)
    h0 = wfpt_gpu.cache_stats()["hits"]
    vals = []
    for v in (1.0, 1.1, 0.9, 1.05):
        node.parents["v"] = v
        vals.append(node.logp)
    want = [oracle.wiener_like(rt, v, 0.5, 2.0, 0.5, 0.1, 0.3, 0.1, 1e-4, **KW) for v in (1.0, 1.1, 0.9, 1.05)]
    assert rel_err(np.array(vals), np.array(want)) <= 1e-10
    assert wfpt_gpu.cache_stats()["hits"] - h0 >= 3          # the frame's column keeps its address between calls
    # batch_logp: many nodes, one launch, same numbers as the per-node calls
    nodes = [pd.DataFrame({"rt": _rts(50 + k, 200 + 13 * k)}) for k in range(12)]
    rng = np.random.default_rng(8)
    prm = np.stack([rng.normal(1, 0.3, 12), rng.uniform(0, 1, 12), rng.uniform(1.5, 2.5, 12), rng.uniform(0.4, 0.6, 12),
                    rng.uniform(0, 0.2, 12), rng.uniform(0.1, 0.3, 12), rng.uniform(0, 0.15, 12), np.full(12, 0.05)], axis=1)
    got, resident = likelihoods.batch_logp(nodes, prm)
    like = cls.wfpt_like
    per_node = np.array([like(nd, *q[:7], p_outlier=q[7]) for nd, q in zip(nodes, prm)])
    assert rel_err(got, per_node) <= 1e-12
    got2, _ = likelihoods.batch_logp(nodes, prm * 1.0, resident=resident)
    assert np.array_equal(got, got2)
    assert len(node.pdf(np.array([0.5, -0.7]))) == 2 and len(node.cdf(np.array([0.5, -0.7]))) == 2
    out = node.random()
    assert len(out) == 500 and (out["rt"] < 0).any()
