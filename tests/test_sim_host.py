"""CPU tests of the simulator drop-in's host logic (no GPU): theta coercion, model table, canonical race layout,
pointwise binning against a literal restatement of the reference loop."""
import numpy as np
import pytest

from hddm_b200 import simulators as S


def test_pointwise_binning_matches_reference_loop():
    # /root/reference/graveyard/basic_simulator_old.py:27-67
    rng = np.random.default_rng(0)
    rt = rng.uniform(0, 25, (400, 1)).astype(np.float32)
    rt[3, 0] = 0.0
    ch = rng.choice([-1, 1], (400, 1)).astype(np.int32)
    got = S.bin_simulator_output_pointwise((rt, ch, {"max_t": 20.0}), nbins=64)
    bins = np.zeros(65)
    bins[:64] = np.linspace(0, 20, 64)
    bins[64] = np.inf
    want = rt.copy()
    for i in range(rt.shape[0]):
        for j in range(1, bins.shape[0]):
            if rt[i] > bins[j - 1] and rt[i] < bins[j]:
                want[i] = j - 1
    c2 = ch.copy()
    c2[c2 == -1] = 0
    want = np.concatenate([want, c2.astype(np.float32)], axis=-1).astype(np.int32)
    assert np.array_equal(got, want)


def test_histogram_binning_rule():
    # basic_simulator_old.py:70-119 on hand-made data
    rt = np.array([[0.0], [0.1], [19.99], [20.0], [25.0]], dtype=np.float32)
    ch = np.array([[1], [-1], [1], [1], [-1]], dtype=np.int32)
    h = S.bin_simulator_output((rt, ch, {"max_t": 20.0, "possible_choices": [-1, 1], "n_samples": 5}), nbins=5, freq_cnt=True)
    assert h.shape == (5, 2) and h.sum() == 5
    assert h[4, 0] == 1 and h[4, 1] == 1 and h[0, 1] == 1 and h[0, 0] == 1 and h[3, 1] == 1


def test_theta_coercion_and_model_table():
    th = S._coerce_theta([1.0, 1.0, 0.5, 0.3, 0.2], "angle")
    assert th.shape == (1, 5) and th.dtype == np.float32
    with pytest.raises(ValueError):
        S._coerce_theta(np.zeros((3, 4)), "angle")
    assert S._resolve("weibull_cdf") == "weibull" and S._resolve("full_ddm_legacy") == "full_ddm_hddm_base"
    assert S._resolve("ddm_seq2_no_bias") == "ddm_seq2_no_bias" and len(S.model_config) == 42
    assert S._par2_canonical("tradeoff_conflict_gamma_no_bias", np.arange(10, dtype=np.float32)[None, :]).tolist() == \
        [[0, 1, 2, 5, 0.5, 0.5, 0.5, 4, 6, 7, 3, 8, 9]]
    assert S._par2_canonical("ddm_mic2_adj", np.arange(9, dtype=np.float32)[None, :]).tolist() == \
        [[0, 1, 2, 3, 4, 5, 6, 8, 0, 0, 7, 0, 0]]
    assert S._par2_canonical("ddm_par2_weibull_no_bias", np.arange(7, dtype=np.float32)[None, :]).tolist() == \
        [[0, 1, 2, 3, 0.5, 0.5, 0.5, 4, 5, 6]]
    with pytest.raises(ValueError):
        S._resolve("no_such_model")
    # parameter lists follow /root/reference/hddm/model_config.py
    assert S.model_config["angle"]["params"] == ["v", "a", "z", "t", "theta"]
    assert S.model_config["lca_no_bias_angle_4"]["params"] == ["v0", "v1", "v2", "v3", "a", "z", "g", "b", "t", "theta"]
    assert S.model_config["race_no_bias_3"]["choices"] == [0, 1, 2]
    c = S._race_canonical("lca_no_bias_angle_4", np.arange(10, dtype=np.float32)[None, :])
    assert c.tolist() == [[0, 1, 2, 3, 4, 5, 5, 5, 5, 6, 7, 8, 9]]


def test_preprocess_and_outliers():
    """`hddm_dataset_generators.py:12-101`: frame layout, -1 -> 0, outliers replace the tail using NumPy's global RNG."""
    n = 500
    rts = np.linspace(0.3, 2.0, n, dtype=np.float32)[:, None]
    ch = np.where(np.arange(n) % 3 == 0, -1, 1).astype(np.int32)[:, None]
    meta = {"model": "angle", "possible_choices": [-1, 1], "max_t": 20.0, "n_samples": n,
            "v": np.array([1.0], np.float32), "a": np.array([1.5], np.float32), "z": np.array([0.5], np.float32),
            "t": np.array([0.3], np.float32), "theta": np.array([0.2], np.float32)}
    df = S.hddm_preprocess((rts, ch, meta), subj_id=3, add_model_parameters=True)
    assert list(df.columns) == ["rt", "response", "subj_idx", "v", "a", "z", "t", "theta"]
    assert df["rt"].dtype == np.float64 and set(df["response"]) == {0, 1} and (df["subj_idx"] == 3).all()
    assert np.isclose(df["theta"].iloc[0], 0.2)
    keep = S.hddm_preprocess((rts, ch, meta), keep_negative_responses=True, keep_subj_idx=False)
    assert set(keep["response"]) == {-1, 1} and "subj_idx" not in keep

    assert S._add_outliers((rts, ch, meta), p_outlier=0) [0] is rts
    np.random.seed(11)
    k = np.random.binomial(n=n, p=0.1)
    want_rt = np.random.uniform(0.0, 10.0, k)
    want_ch = np.random.choice([-1, 1], size=k)
    r2, c2 = rts.copy(), ch.copy()
    np.random.seed(11)
    out = S._add_outliers((r2, c2, meta), p_outlier=0.1, max_rt_outlier=10.0)
    assert out[0] is r2 and k > 0
    assert np.array_equal(r2[-k:, 0], want_rt.astype(np.float32)) and np.array_equal(c2[-k:, 0], want_ch)
    assert np.array_equal(r2[:-k], rts[:-k]) and np.array_equal(c2[:-k], ch[:-k])
