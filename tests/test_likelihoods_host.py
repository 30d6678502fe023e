"""CPU tests of the likelihood adapter (hddm_b200/likelihoods.py): the closure semantics of the reference's
generate_wfpt_stochastic_class (/root/reference/hddm/likelihoods.py:83-149) with the device call replaced by the C
oracle -- no GPU here; tests/test_gpu_mcmc.py runs the same adapter on the CUDA path."""
import numpy as np
import pandas as pd
import pytest
from scipy import stats


@pytest.fixture()
def like_with_oracle(monkeypatch, oracle):
    from hddm_b200 import likelihoods

    calls = []

    def fake(x, v, sv, a, z, sz, t, st, err, n_st=10, n_sz=10, use_adaptive=1, simps_err=1e-8, p_outlier=0, w_outlier=0.1):
        calls.append(x)
        return oracle.wiener_like(x, v, sv, a, z, sz, t, st, err, n_st=n_st, n_sz=n_sz, use_adaptive=use_adaptive,
                                  simps_err=simps_err, p_outlier=p_outlier, w_outlier=w_outlier)

    monkeypatch.setattr(likelihoods._wfpt, "wiener_like", fake)
    return likelihoods, calls


def test_wfpt_like_closure_matches_reference_semantics(like_with_oracle, oracle):
    likelihoods, calls = like_with_oracle
    rng = np.random.default_rng(0)
    rt = np.where(rng.random(200) < 0.8, 1.0, -1.0) * rng.uniform(0.4, 2.5, 200)
    frame = pd.DataFrame({"rt": rt, "response": (rt > 0).astype(int)})
    wfpt_like = likelihoods.make_wfpt_like()
    got = wfpt_like(frame, 1.0, 0.5, 2.0, 0.5, 0.1, 0.3, 0.1, p_outlier=0.05)
    want = oracle.wiener_like(rt, 1.0, 0.5, 2.0, 0.5, 0.1, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                              p_outlier=0.05, w_outlier=0.1)
    assert got == want
    # the array handed down is the frame's own column (same address every call: what the resident cache keys on)
    assert calls[-1].ctypes.data == frame["rt"].values.ctypes.data
    wfpt_like(frame, 1.1, 0.5, 2.0, 0.5, 0.1, 0.3, 0.1, p_outlier=0.05)
    assert calls[-1].ctypes.data == calls[-2].ctypes.data
    # missing responses (likelihoods.py:122-149): |rt| >= 999 rows go through the binomial on the choice probability
    rt2 = np.concatenate([rt, [999.0, 999.0, -999.0]])
    frame2 = pd.DataFrame({"rt": rt2})
    for v in (0.7, 0.0):
        got = wfpt_like(frame2, v, 0.0, 2.0, 0.45, 0.0, 0.3, 0.0, p_outlier=0.05)
        base = oracle.wiener_like(rt, v, 0.0, 2.0, 0.45, 0.0, 0.3, 0.0, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                                  p_outlier=0.05, w_outlier=0.1)
        p_up = 0.45 if v == 0 else (np.exp(-2 * 2.0 * 0.45 * v) - 1) / (np.exp(-2 * 2.0 * v) - 1)
        assert got == pytest.approx(base + stats.binom.logpmf(2, 3, p_up), rel=1e-14)
    # custom wiener_params travel into the call
    like2 = likelihoods.make_wfpt_like({"err": 1e-6, "n_st": 4, "n_sz": 4, "use_adaptive": 0, "simps_err": 1e-3, "w_outlier": 0.2})
    got = like2(frame, 1.0, 0.5, 2.0, 0.5, 0.1, 0.3, 0.1, p_outlier=0.1)
    want = oracle.wiener_like(rt, 1.0, 0.5, 2.0, 0.5, 0.1, 0.3, 0.1, 1e-6, n_st=4, n_sz=4, use_adaptive=0, simps_err=1e-3,
                              p_outlier=0.1, w_outlier=0.2)
    assert got == want


def test_stochastic_class_shape(like_with_oracle):
    likelihoods, _ = like_with_oracle
    cls = likelihoods.generate_wfpt_stochastic_class()
    rt = np.array([0.8, -0.9, 1.3, 0.6])
    node = cls(value=pd.DataFrame({"rt": rt}), v=1.0, sv=0.0, a=2.0, z=0.5, sz=0.0, t=0.3, st=0.0, p_outlier=0.05)
    assert np.isfinite(node.logp) and node.shape == (4,)
    node.parents["v"] = 1.2
    lp2 = node.logp
    assert np.isfinite(lp2) and lp2 != cls(value=node.value, v=1.0, sv=0.0, a=2.0, z=0.5, sz=0.0, t=0.3, st=0.0, p_outlier=0.05).logp
    for name in ("pdf", "cdf", "cdf_vec", "random", "wfpt_like"):
        assert hasattr(cls, name)
