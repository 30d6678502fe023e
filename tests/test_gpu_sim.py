"""GPU tests of the simulators (run with -m gpu): the CUDA kernels, through the C ABI / the simulator() drop-in,
against the CPU oracle port (distributional: two-sample KS on RTs, overlapping choice-probability intervals), the
analytic choice probability, the reference's own shape/metadata test, and exact properties of the counter-based
streams (geometry / sharding independence) and of the shared-memory histograms (== np.histogram of the raw output)."""
import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu

CASES = {
    "ddm_hddm_base": (0, [1.0, 2.0, 0.5, 0.3]),
    "full_ddm_hddm_base": (1, [0.8, 1.8, 0.5, 0.4, 0.1, 0.7, 0.15]),
    "ddm": (2, [0.6, 1.1, 0.45, 0.3]),
    "angle": (3, [1.0, 1.0, 0.5, 0.3, 0.2]),
    "weibull": (4, [-0.5, 1.5, 0.4, 0.2, 2.0, 1.5]),
    "full_ddm": (5, [0.8, 1.2, 0.5, 0.4, 0.1, 0.7, 0.15]),
    "ddm_sdv": (6, [0.8, 1.2, 0.5, 0.3, 1.0]),
    "ornstein": (7, [0.5, 1.2, 0.5, 0.6, 0.3]),
    "levy": (8, [0.5, 1.2, 0.5, 1.6, 0.3]),
    "gamma_drift": (9, [0.3, 1.2, 0.5, 0.3, 3.0, 0.3, 1.5]),
    "gamma_drift_angle": (10, [0.3, 1.2, 0.5, 0.3, 0.15, 3.0, 0.3, 1.5]),
    "ds_conflict_drift": (11, [1.3, 0.5, 0.3, 0.5, 2.0, 2.0, 3.0, 1.5, 1.0, -1.0]),
    "ds_conflict_drift_angle": (12, [1.3, 0.5, 0.3, 0.5, 2.0, 2.0, 3.0, 1.5, 1.0, -1.0, 0.2]),
    "ddm_cov": (13, [1.0, 2.0, 0.5, 0.3]),
}


def _ci(k, n, zc=4.0):
    p = k / n
    h = zc * np.sqrt(max(p * (1 - p), 1e-12) / n)
    return p - h, p + h


@pytest.fixture(scope="module")
def sims():
    from hddm_b200 import _lib, simulators

    _lib.load()
    return simulators


@pytest.mark.parametrize("model", list(CASES))
def test_distribution_matches_oracle(model, sims, oracle):
    mid, theta = CASES[model]
    n_gpu, n_cpu = 400000, 100000
    rts, ch, meta = sims.simulator(theta, model=model, n_samples=n_gpu, random_state=100 + mid)
    ort, och = oracle.sim_run(mid, theta, n_cpu, seed=200 + mid)
    assert rts.shape == (n_gpu, 1) and rts.dtype == np.float32 and ch.dtype == np.int32
    ks = stats.ks_2samp(rts[:, 0], ort[:, 0])
    assert ks.pvalue > 1e-3, (model, ks)
    upper = 1
    la, ha = _ci(int(np.sum(ch == upper)), n_gpu)
    lb, hb = _ci(int(np.sum(och == upper)), n_cpu)
    assert la <= hb and lb <= ha, (model, (la, ha), (lb, hb))
    # conditional RT distributions per boundary as well
    for c in meta["possible_choices"]:
        a, b = rts[ch == c], ort[och == c]
        if min(len(a), len(b)) > 2000:
            assert stats.ks_2samp(a, b).pvalue > 1e-3, (model, c)
    assert set(np.unique(ch)) <= set(meta["possible_choices"]) | {0}


RACE_CASES = {
    "race_no_bias_3": [1.0, 0.5, 0.2, 2.0, 0.3, 0.3],
    "race_no_bias_angle_4": [1.0, 0.8, 0.5, 0.2, 2.0, 0.3, 0.3, 0.2],
    "race_3": [1.0, 0.5, 0.2, 2.0, 0.2, 0.3, 0.4, 0.3],
    "lca_no_bias_3": [1.0, 0.5, 0.2, 2.0, 0.3, 0.5, 0.3, 0.3],
    "lca_no_bias_angle_4": [1.0, 0.8, 0.5, 0.2, 2.0, 0.3, 0.5, 0.3, 0.3, 0.2],
    "lca_4": [1.0, 0.8, 0.5, 0.2, 2.0, 0.2, 0.3, 0.3, 0.4, 0.5, 0.3, 0.3],
}


@pytest.mark.parametrize("model", list(RACE_CASES))
def test_race_and_lca_match_oracle(model, sims, oracle):
    theta = np.asarray(RACE_CASES[model], dtype=np.float32)
    lca, npar, ang, cols = sims._RACE[model]
    n_gpu, n_cpu = 300000, 60000
    rts, ch, meta = sims.simulator(theta, model=model, n_samples=n_gpu, random_state=41)
    canon = sims._race_canonical(model, theta[None, :])
    ort, och = oracle.sim_race_run(lca, npar, ang, canon, n_cpu, seed=42)
    assert rts.shape == (n_gpu, 1) and meta["possible_choices"] == list(range(npar))
    assert stats.ks_2samp(rts[:, 0], ort[:, 0]).pvalue > 1e-3
    for c in range(npar):
        la, ha = _ci(int(np.sum(ch == c)), n_gpu)
        lb, hb = _ci(int(np.sum(och == c)), n_cpu)
        assert la <= hb and lb <= ha, (model, c)
    for p_name in cols:
        assert p_name in meta
    # histogram mode == np.histogram of the raw output, per accumulator
    r2, c2, hist, steps = sims._run_race(model, theta[None, :], 20000, 0.001, 20.0, 1.0, 7, True, 128)
    out = (r2[:, 0, :], c2[:, 0, :], {"max_t": 20.0, "possible_choices": list(range(npar)), "n_samples": 20000})
    assert np.array_equal(hist[0].astype(np.float64), sims.bin_simulator_output(out, nbins=128, freq_cnt=True))


PAR2_CASES = {
    "ddm_par2": [1.0, 0.5, -0.5, 1.2, 0.55, 0.45, 0.5, 0.3],
    "ddm_par2_no_bias": [1.0, 0.5, -0.5, 1.2, 0.3],
    "ddm_par2_angle_no_bias": [1.0, 0.5, -0.5, 1.2, 0.3, 0.2],
    "ddm_par2_weibull_no_bias": [1.0, 0.5, -0.5, 1.2, 0.3, 2.0, 1.5],
}


@pytest.mark.parametrize("model", list(PAR2_CASES))
def test_par2_matches_oracle(model, sims, oracle):
    theta = np.asarray(PAR2_CASES[model], dtype=np.float32)
    kind, cols = sims._PAR2[model]
    n_gpu, n_cpu = 300000, 60000
    rts, ch, meta = sims.simulator(theta, model=model, n_samples=n_gpu, random_state=51)
    ort, och = oracle.sim_par2_run(kind, sims._par2_canonical(model, theta[None, :]), n_cpu, seed=52)
    assert rts.shape == (n_gpu, 1) and meta["possible_choices"] == [0, 1, 2, 3]
    assert stats.ks_2samp(rts[:, 0], ort[:, 0]).pvalue > 1e-3
    for c in range(4):
        la, ha = _ci(int(np.sum(ch == c)), n_gpu)
        lb, hb = _ci(int(np.sum(och == c)), n_cpu)
        assert la <= hb and lb <= ha, (model, c)
    for p_name in cols + ["zh", "zl1", "zl2"]:
        assert p_name in meta
    r2, c2, hist, steps = sims._run_par2(model, theta[None, :], 20000, 0.001, 20.0, 1.0, 7, True, 128)
    out = (r2[:, 0, :], c2[:, 0, :], {"max_t": 20.0, "possible_choices": [0, 1, 2, 3], "n_samples": 20000})
    assert np.array_equal(hist[0].astype(np.float64), sims.bin_simulator_output(out, nbins=128, freq_cnt=True))


HIER_CASES = {
    "ddm_seq2": [1.0, 0.5, -0.5, 1.2, 0.55, 0.45, 0.5, 0.3],
    "ddm_seq2_angle_no_bias": [0.4, 0.5, -0.5, 1.0, 0.3, 0.6],        # bound collapses early: the carried-over y_l matters
    "ddm_seq2_weibull_no_bias": [1.0, 0.5, -0.5, 1.2, 0.3, 2.0, 1.5],
    "ddm_mic2_adj": [1.0, 0.8, -0.8, 1.2, 0.55, 0.45, 0.5, 0.6, 0.3],
    "ddm_mic2_adj_angle_no_bias": [0.6, 0.8, -0.8, 1.0, 0.4, 0.3, 0.4],
    "tradeoff_no_bias": [1.0, 0.8, -0.8, 1.2, 0.6, 0.3],
    "tradeoff_weibull_no_bias": [1.0, 0.8, -0.8, 1.2, 0.6, 0.3, 2.0, 1.5],
    "tradeoff_conflict_gamma_no_bias": [1.0, 0.8, -0.8, 0.6, 0.3, 0.8, 0.3, 1.5, 2.0, 0.6],
}


@pytest.mark.parametrize("model", list(HIER_CASES))
def test_hier_matches_oracle(model, sims, oracle):
    """seq2 / mic2_adj / tradeoff against the serial CPU restatement (which keeps the reference's carried-over state)."""
    theta = np.asarray(HIER_CASES[model], dtype=np.float32)
    kind, bk, cols = sims._HIER[model]
    n_gpu, n_cpu = 300000, 60000
    rts, ch, meta = sims.simulator(theta, model=model, n_samples=n_gpu, random_state=61)
    ort, och = oracle.sim_hier_run(kind, bk, sims._par2_canonical(model, theta[None, :]), n_cpu, seed=62)
    assert rts.shape == (n_gpu, 1) and meta["possible_choices"] == [0, 1, 2, 3]
    assert stats.ks_2samp(rts[:, 0], ort[:, 0]).pvalue > 1e-3
    for c in range(4):
        la, ha = _ci(int(np.sum(ch == c)), n_gpu)
        lb, hb = _ci(int(np.sum(och == c)), n_cpu)
        assert la <= hb and lb <= ha, (model, c)
        sel_g, sel_c = rts[:, 0][ch[:, 0] == c], ort[:, 0][och[:, 0] == c]
        if min(sel_g.size, sel_c.size) > 2000:
            assert stats.ks_2samp(sel_g, sel_c).pvalue > 1e-4, (model, c)
    for p_name in cols + ["zh", "zl1", "zl2"]:
        assert p_name in meta
    # same particles for any sharding of the particle axis (the predecessor replay of seq2 is keyed by global index)
    full = sims._run_par2(model, theta[None, :], 6000, 0.001, 20.0, 1.0, 9, True, 0)
    tail = sims._run_par2(model, theta[None, :], 2000, 0.001, 20.0, 1.0, 9, True, 0, sample_offset=4000)
    assert np.array_equal(full[0][4000:], tail[0]) and np.array_equal(full[1][4000:], tail[1])


def test_shapes_and_metadata_like_reference_test(sims):
    """/root/reference/hddm/tests/test_simulators.py:65-87."""
    n_samples, n_trials = 100, 7
    for model, (mid, theta) in CASES.items():
        out = sims.simulator(theta=np.tile(theta, reps=(n_trials, 1)), model=model, n_samples=n_samples)
        assert out[0].shape == out[1].shape == (n_samples, n_trials, 1)
        assert type(out[2]) is dict
        for param in sims.model_config[model]["params"]:
            assert param in out[2].keys()
        for key in ("possible_choices", "max_t", "n_samples", "delta_t", "s", "model", "trajectory"):
            assert key in out[2]
    # squeeze rules (graveyard/basic_simulator_old.py:1041-1045)
    assert sims.simulator([1.0, 1.0, 0.5, 0.3, 0.2], model="angle", n_samples=50)[0].shape == (50, 1)
    assert sims.simulator(np.tile([1.0, 1.0, 0.5, 0.3, 0.2], (9, 1)), model="angle", n_samples=1)[0].shape == (9, 1)
    with pytest.raises(ValueError):
        sims.simulator([1.0, 1.0, 0.5], model="angle")
    with pytest.raises(ValueError):
        sims.simulator([0.0] * 8, model="ddm_seq3")
    # ddm_cov codes the lower boundary as +1 (cddm_data_simulation.pyx:437)
    r0, c0, _ = sims.simulator([2.0, 2.0, 0.5, 0.3], model="ddm_hddm_base", n_samples=4000, random_state=3)
    r1, c1, _ = sims.simulator([2.0, 2.0, 0.5, 0.3], model="ddm_cov", n_samples=4000, random_state=3)
    assert np.array_equal(r0, r1) and np.array_equal(c1, np.where(c0 == 1, -1, 1))


def test_streams_do_not_depend_on_geometry_or_sharding(sims):
    """Same seed -> the same particles, whether simulated in one launch, in two shards with global offsets (what two
    ranks do), next to other parameter sets, or while a histogram is also being built."""
    theta = np.array([[1.0, 1.0, 0.5, 0.3, 0.2], [-0.7, 1.4, 0.6, 0.2, 0.1]], dtype=np.float32)
    n = 5000
    full = sims._run("angle", theta, n, 0.001, 20.0, 1.0, 77, True, 0)
    a = sims._run("angle", theta, 2000, 0.001, 20.0, 1.0, 77, True, 0, sample_offset=0)
    b = sims._run("angle", theta, 3000, 0.001, 20.0, 1.0, 77, True, 0, sample_offset=2000)
    assert np.array_equal(np.concatenate([a[0], b[0]]), full[0])
    assert np.array_equal(np.concatenate([a[1], b[1]]), full[1])
    both = sims._run("angle", theta, n, 0.001, 20.0, 1.0, 77, True, 64)
    assert np.array_equal(both[0], full[0]) and np.array_equal(both[1], full[1])
    other = sims._run("angle", theta, n, 0.001, 20.0, 1.0, 78, True, 0)
    assert not np.array_equal(other[0], full[0])
    # first particle's trajectory ends where its walk ended
    rts, ch, meta = sims.simulator(theta[0], model="angle", n_samples=10, random_state=5)
    traj = meta["trajectory"][:, 0]
    k = int(np.sum(traj > -999)) - 1
    assert abs((k * 0.001 + 0.3) - rts[0, 0]) < 2e-3


@pytest.mark.parametrize("model,nbins", [("angle", 512), ("ddm_hddm_base", 64)])
def test_histograms_equal_host_binning_of_raw_output(model, nbins, sims):
    mid, theta = CASES[model]
    theta = np.tile(np.asarray(theta, dtype=np.float32), (3, 1))
    theta[1, 0] = -theta[1, 0]
    n = 50000
    rts, ch, hist, _, steps = sims._run(model, theta, n, 0.001, 20.0, 1.0, 9, True, nbins)
    assert int(hist.sum()) + int(np.sum(ch == 0) if model == "angle" else 0) == 3 * n
    choices = sims.model_config[model]["choices"]
    for k in range(3):
        out = (rts[:, k, :], ch[:, k, :], {"max_t": 20.0, "possible_choices": choices, "n_samples": n})
        want = sims.bin_simulator_output(out, nbins=nbins, freq_cnt=True)
        assert np.array_equal(hist[k].astype(np.float64), want), (model, k)
    # the reference-shaped binned call: proportions for one parameter set
    h, meta = sims.simulator(theta[0], model=model, n_samples=n, bin_dim=nbins, random_state=9)
    assert h.shape == (nbins, 2) and abs(h.sum() - 1.0) < 1e-3
    assert steps > n


def test_choice_probability_at_1e7_particles(sims, oracle):
    """north_star: exact choice-probability interval at 1e7 particles.  Constant-bound DDM, P(upper) analytic; the
    Euler scheme (dt = 1e-3) has a known O(sqrt(dt)) bias, so the analytic value is compared at a finer dt as well."""
    v, a, z, t = 0.7, 1.6, 0.45, 0.25
    n = 10_000_000
    hist, steps = sims.simulate_histograms([v, a, z, t], model="ddm_hddm_base", n_samples=n, nbins=512, random_state=3)
    assert int(hist.sum()) == n
    p_gpu = hist[0, :, 1].sum() / n
    p_true = (np.exp(-2 * v * a * z) - 1) / (np.exp(-2 * v * a) - 1)
    assert abs(p_gpu - p_true) < 0.01          # Euler bias at dt = 1e-3 (about 0.006 here), same in the reference
    ort, och = oracle.sim_run(0, [v, a, z, t], 300000, seed=31)
    lo, hi = _ci(int(np.sum(och == 1)), 300000)
    glo, ghi = _ci(int(hist[0, :, 1].sum()), n)
    assert glo <= hi and lo <= ghi, ((glo, ghi), (lo, hi))   # interval overlap with the CPU port of the reference
    hist2, _ = sims.simulate_histograms([v, a, z, t], model="ddm_hddm_base", n_samples=2_000_000, nbins=512, delta_t=1e-4,
                                        random_state=4)
    p_fine = hist2[0, :, 1].sum() / 2_000_000
    assert abs(p_fine - p_true) < abs(p_gpu - p_true) + 4 * np.sqrt(p_true * (1 - p_true) / 2e6)
    assert abs(p_fine - p_true) < 0.003
    # two independent 1e7 runs agree within their binomial interval, bin by bin in aggregate
    hist3, _ = sims.simulate_histograms([v, a, z, t], model="ddm_hddm_base", n_samples=n, nbins=512, random_state=5)
    p3 = hist3[0, :, 1].sum() / n
    assert abs(p3 - p_gpu) < 5 * np.sqrt(2 * p_true * (1 - p_true) / n)
    ks = np.max(np.abs(np.cumsum(hist[0].sum(1).astype(np.int64)) - np.cumsum(hist3[0].sum(1).astype(np.int64)))) / n
    assert ks < 1.95 * np.sqrt(2 / n)  # two-sample KS at alpha = 0.001 on the binned RTs


def test_every_model_name_runs(sims):
    """All 42 names of the simulator boundary: shapes, finite positive RTs, choices inside `possible_choices`, and every
    parameter name in the metadata (hddm/tests/test_simulators.py:78-87 does this for the reference's list)."""
    defaults = {"v": 0.8, "a": 1.3, "z": 0.5, "t": 0.3, "theta": 0.3, "alpha": 1.8, "beta": 2.0, "sz": 0.1, "sv": 0.5, "st": 0.1,
                "g": 0.3, "b": 0.2, "shape": 2.0, "scale": 1.5, "c": 1.0, "tinit": 0.5, "dinit": 0.5, "tslope": 1.0,
                "dslope": 1.0, "tfixedp": 1.0, "tcoh": 0.5, "dcoh": 0.5, "angle": 0.3, "vh": 0.8, "vl1": 0.5, "vl2": -0.5,
                "zh": 0.5, "zl1": 0.5, "zl2": 0.5, "d": 0.5, "alphagamma": 2.0, "scalegamma": 0.6}
    assert len(sims.model_config) == 42
    for name, cfg in sims.model_config.items():
        th = []
        for p_name in cfg["params"]:
            if p_name in defaults:
                th.append(defaults[p_name])
            elif p_name[0] == "v":
                th.append(0.5 + 0.3 * int(p_name[1:]))
            else:
                assert p_name[0] == "z", p_name
                th.append(0.3)
        rts, ch, meta = sims.simulator(th, model=name, n_samples=1500, random_state=1)
        assert rts.shape == (1500, 1) and ch.shape == (1500, 1), name
        assert np.isfinite(rts).all() and (rts > 0).all(), name
        assert set(np.unique(ch).tolist()) <= set(meta["possible_choices"]), name
        assert meta["model"] == name and all(p_name in meta for p_name in cfg["params"]), name
