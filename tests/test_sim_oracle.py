"""CPU tests of the simulator oracle (oracle/sim_oracle.c): distributional agreement with the reference's own
Cython simulators (compiled unmodified into oracle/_ref, skipped where that build is absent) and with the analytic
WFPT for the constant-bound DDM.  The reference's tests pin no simulator numerics (parity unpinned), so the bar is the
one the north_star names: two-sample KS on RTs and overlapping choice-probability intervals."""
import numpy as np
import pytest
from scipy import stats


def _ci(k, n, zc=4.0):
    p = k / n
    h = zc * np.sqrt(max(p * (1 - p), 1e-12) / n)
    return p - h, p + h


def _agree(rt_a, ch_a, rt_b, ch_b, upper=1):
    # RT grid values are shared (same float32 time accumulation); KS handles ties conservatively
    ks = stats.ks_2samp(rt_a, rt_b)
    assert ks.pvalue > 1e-3, ks
    la, ha = _ci(int(np.sum(ch_a == upper)), len(ch_a))
    lb, hb = _ci(int(np.sum(ch_b == upper)), len(ch_b))
    assert la <= hb and lb <= ha, ((la, ha), (lb, hb))


@pytest.fixture(scope="module")
def ref_sim():
    from oracle.build_ref import load_ref, ref_built

    if not ref_built():
        pytest.skip("oracle/_ref not built on this machine")
    return load_ref()[1]


def constant(t=0):
    return 1


def angle(t=1, theta=1):
    return np.multiply(t, (-np.sin(theta) / np.cos(theta)))


def weibull_cdf(t=1, alpha=1, beta=1):
    return np.exp(-np.power(np.divide(t, beta), alpha))


N = 40000


def _f32(*xs):
    return [np.array([x], dtype=np.float32) for x in xs]


def test_ddm_vs_reference(oracle, ref_sim):
    v, a, z, t = _f32(1.0, 2.0, 0.5, 0.3)
    r = ref_sim.ddm(v, a, z, t, n_samples=N, n_trials=1)
    rt, ch = oracle.sim_run(0, [1.0, 2.0, 0.5, 0.3], N, seed=11)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])


def test_angle_and_weibull_vs_reference(oracle, ref_sim):
    v, a, z, t = _f32(1.0, 1.0, 0.5, 0.3)
    r = ref_sim.ddm_flexbound(v, a, z, t, n_samples=N, n_trials=1, boundary_fun=angle, boundary_multiplicative=False,
                              boundary_params={"theta": np.array([0.2], dtype=np.float32)})
    rt, ch = oracle.sim_run(3, [1.0, 1.0, 0.5, 0.3, 0.2], N, seed=12)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])
    v, a, z, t = _f32(-0.5, 1.5, 0.4, 0.2)
    r = ref_sim.ddm_flexbound(v, a, z, t, n_samples=N, n_trials=1, boundary_fun=weibull_cdf, boundary_multiplicative=True,
                              boundary_params={"alpha": np.array([2.0], dtype=np.float32), "beta": np.array([1.5], dtype=np.float32)})
    rt, ch = oracle.sim_run(4, [-0.5, 1.5, 0.4, 0.2, 2.0, 1.5], N, seed=13)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])


def test_full_ddm_family_vs_reference(oracle, ref_sim):
    th = [0.8, 1.8, 0.5, 0.4, 0.1, 0.7, 0.15]  # v, a, z, t, sz, sv, st
    arrs = _f32(*th)
    r = ref_sim.full_ddm_hddm_base(*arrs, n_samples=N, n_trials=1)
    rt, ch = oracle.sim_run(1, th, N, seed=14)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])
    r = ref_sim.full_ddm(*arrs, n_samples=N, n_trials=1, boundary_fun=constant, boundary_multiplicative=True, boundary_params={})
    rt, ch = oracle.sim_run(5, th, N, seed=15)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])
    v, a, z, t, sv = _f32(0.8, 1.2, 0.5, 0.3, 1.0)
    r = ref_sim.ddm_sdv(v, a, z, t, sv, n_samples=N, n_trials=1, boundary_fun=constant, boundary_multiplicative=True, boundary_params={})
    rt, ch = oracle.sim_run(6, [0.8, 1.2, 0.5, 0.3, 1.0], N, seed=16)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])


def test_ornstein_and_levy_vs_reference(oracle, ref_sim):
    v, a, z, g, t = _f32(0.5, 1.2, 0.5, 0.6, 0.3)
    r = ref_sim.ornstein_uhlenbeck(v, a, z, g, t, n_samples=N, n_trials=1, boundary_fun=constant, boundary_multiplicative=True, boundary_params={})
    rt, ch = oracle.sim_run(7, [0.5, 1.2, 0.5, 0.6, 0.3], N, seed=17)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])
    v, a, z, al, t = _f32(0.5, 1.2, 0.5, 1.6, 0.3)
    r = ref_sim.levy_flexbound(v, a, z, al, t, n_samples=N, n_trials=1, boundary_fun=constant, boundary_multiplicative=True, boundary_params={})
    rt, ch = oracle.sim_run(8, [0.5, 1.2, 0.5, 1.6, 0.3], N, seed=18)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])


def gamma_drift(t, shape=2, scale=0.01, c=1.5):
    # /root/reference/graveyard/drift_functions.py:11-36
    num_ = np.power(t, shape - 1) * np.exp(np.divide(-t, scale))
    div_ = np.power(shape - 1, shape - 1) * np.power(scale, shape - 1) * np.exp(-(shape - 1))
    return c * np.divide(num_, div_)


def ds_conflict_drift(t, init_p_t=0, init_p_d=0, slope_t=1, slope_d=1, fixed_p_t=1, coherence_t=1.5, coherence_d=1.5):
    # drift_functions.py:39-103 (the drift timecourse itself)
    w_t = (init_p_t - fixed_p_t) * np.exp(-(slope_t * t)) + fixed_p_t
    w_d = (init_p_d - 0) * np.exp(-(slope_d * t)) + 0
    return (w_t * coherence_t) + (w_d * coherence_d)


def test_ddm_flex_and_ddm_cov_vs_reference(oracle, ref_sim):
    v, a, z, t = _f32(0.3, 1.2, 0.5, 0.3)
    dp = {"shape": np.array([3.0], dtype=np.float32), "scale": np.array([0.3], dtype=np.float32), "c": np.array([1.5], dtype=np.float32)}
    r = ref_sim.ddm_flex(v, a, z, t, n_samples=N, n_trials=1, boundary_fun=angle, drift_fun=gamma_drift, boundary_multiplicative=False,
                         boundary_params={"theta": np.array([0.15], dtype=np.float32)}, drift_params=dp)
    rt, ch = oracle.sim_run(10, [0.3, 1.2, 0.5, 0.3, 0.15, 3.0, 0.3, 1.5], N, seed=19)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])
    th = [1.3, 0.5, 0.3, 0.5, 2.0, 2.0, 3.0, 1.5, 1.0, -1.0]
    names = ["init_p_t", "init_p_d", "slope_t", "slope_d", "fixed_p_t", "coherence_t", "coherence_d"]
    dp = {k: np.array([x], dtype=np.float32) for k, x in zip(names, th[3:])}
    v0, a, z, t = _f32(0.0, th[0], th[1], th[2])
    r = ref_sim.ddm_flex(v0, a, z, t, n_samples=N, n_trials=1, boundary_fun=constant, drift_fun=ds_conflict_drift,
                         boundary_multiplicative=True, boundary_params={}, drift_params=dp)
    rt, ch = oracle.sim_run(11, th, N, seed=20)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])
    v, a, z, t = _f32(1.0, 2.0, 0.5, 0.3)
    r = ref_sim.ddm_cov(v, a, z, t, n_samples=N, n_trials=1)
    rt, ch = oracle.sim_run(13, [1.0, 2.0, 0.5, 0.3], N, seed=21)
    _agree(rt[:, 0], ch[:, 0], r[0][:, 0, 0], r[1][:, 0, 0])


def test_race_and_lca_vs_reference(oracle, ref_sim):
    n = 20000
    s3 = np.ones((1, 3), dtype=np.float32)
    v = np.array([[1.0, 0.5, 0.2]], dtype=np.float32)
    a = np.array([[2.0]], dtype=np.float32)
    z = np.array([[0.3, 0.3, 0.3]], dtype=np.float32)
    t = np.array([[0.3]], dtype=np.float32)
    r = ref_sim.race_model(v=v, a=a, z=z, t=t, s=s3, n_samples=n, n_trials=1, boundary_fun=constant,
                           boundary_multiplicative=True, boundary_params={})
    rt, ch = oracle.sim_race_run(0, 3, 0, [1.0, 0.5, 0.2, 2.0, 0.3, 0.3, 0.3, 0, 0, 0.3, 0], n, seed=22)
    assert stats.ks_2samp(rt[:, 0], r[0][:, 0, 0]).pvalue > 1e-3
    for c in range(3):
        la, ha = _ci(int(np.sum(ch == c)), n)
        lb, hb = _ci(int(np.sum(r[1] == c)), n)
        assert la <= hb and lb <= ha
    g = np.array([[0.5]], dtype=np.float32)
    b = np.array([[0.3]], dtype=np.float32)
    r = ref_sim.lca(v=v, a=a, z=z, g=g, b=b, t=t, s=s3, n_samples=n, n_trials=1, boundary_fun=angle,
                    boundary_multiplicative=False, boundary_params={"theta": np.array([0.2], dtype=np.float32)})
    rt, ch = oracle.sim_race_run(1, 3, 1, [1.0, 0.5, 0.2, 2.0, 0.3, 0.3, 0.3, 0.5, 0.3, 0.3, 0.2], n, seed=23)
    assert stats.ks_2samp(rt[:, 0], r[0][:, 0, 0]).pvalue > 1e-3
    for c in range(3):
        la, ha = _ci(int(np.sum(ch == c)), n)
        lb, hb = _ci(int(np.sum(r[1] == c)), n)
        assert la <= hb and lb <= ha


def test_par2_vs_reference(oracle, ref_sim):
    n = 20000
    th = [1.0, 0.5, -0.5, 1.2, 0.55, 0.45, 0.5, 0.3]
    arrs = _f32(*th)
    r = ref_sim.ddm_flexbound_par2(*arrs, n_samples=n, n_trials=1, boundary_fun=angle, boundary_multiplicative=False,
                                   boundary_params={"theta": np.array([0.2], dtype=np.float32)})
    rt, ch = oracle.sim_par2_run(1, th + [0.2, 0.0], n, seed=24)
    assert stats.ks_2samp(rt[:, 0], r[0][:, 0, 0]).pvalue > 1e-3
    for c in range(4):
        la, ha = _ci(int(np.sum(ch == c)), n)
        lb, hb = _ci(int(np.sum(r[1] == c)), n)
        assert la <= hb and lb <= ha, c


def _conflict_gamma_bound(t=0, a=0.5, theta=0.5, scale=1, alphagamma=1.01, scalegamma=0.3):
    # graveyard/boundary_functions.py:66-97
    return np.maximum(a + scale * stats.gamma.pdf(t, a=alphagamma, loc=0, scale=scalegamma)
                      + np.multiply(t, (-np.sin(theta) / np.cos(theta))), 0)


def _constant(t=0):
    return 1


@pytest.mark.parametrize("kind,fn", [(1, "ddm_flexbound_seq2"), (2, "ddm_flexbound_mic2_adj"), (3, "ddm_flexbound_tradeoff")])
@pytest.mark.parametrize("bk", [0, 1, 3])
def test_hier_vs_reference(oracle, ref_sim, kind, fn, bk):
    n = 20000
    th = [0.6, 0.5, -0.5, 1.1, 0.55, 0.45, 0.5, 0.3]
    d = 0.6
    arrs = _f32(*th[:7]) + (_f32(d) if kind > 1 else []) + _f32(th[7])
    a_arg = th[3]
    if bk == 0:
        bnd, canon_b = dict(boundary_fun=_constant, boundary_multiplicative=True, boundary_params={}), [0, 0, 0, 0]
    elif bk == 1:
        bnd = dict(boundary_fun=angle, boundary_multiplicative=False, boundary_params={"theta": np.array([0.5], dtype=np.float32)})
        canon_b = [0.5, 0, 0, 0]
    else:
        bp = {"a": [th[3]], "theta": [0.3], "scale": [1.5], "alphagamma": [2.0], "scalegamma": [0.6]}
        bnd = dict(boundary_fun=_conflict_gamma_bound, boundary_multiplicative=False,
                   boundary_params={k: np.array(v, dtype=np.float32) for k, v in bp.items()})
        canon_b = [0.3, 1.5, 2.0, 0.6]
        arrs[3] = np.zeros(1, dtype=np.float32)      # the dispatch passes a zero vector; the bound carries its own a
    r = getattr(ref_sim, fn)(*arrs, n_samples=n, n_trials=1, **bnd)
    r_rt, r_ch = (r["rts"], r["choices"]) if isinstance(r, dict) else (r[0], r[1])
    canon = th + [canon_b[0], canon_b[1], d, canon_b[2], canon_b[3]]
    rt, ch = oracle.sim_hier_run(kind, bk, canon, n, seed=31 + kind)
    assert stats.ks_2samp(rt[:, 0], r_rt[:, 0, 0]).pvalue > 1e-3
    for c in range(4):
        la, ha = _ci(int(np.sum(ch == c)), n)
        lb, hb = _ci(int(np.sum(r_ch == c)), n)
        assert la <= hb and lb <= ha, c


def test_ddm_vs_analytic_wfpt(oracle):
    """Constant bounds: P(upper) from prob_ub (pdf.pxi:67-72) and the RT density from full_pdf, as
    /root/reference/hddm/tests/test_generate.py:33-50 validates its own random walk."""
    v, a, z, t = 0.7, 1.6, 0.45, 0.25
    n = 60000
    rt, ch = oracle.sim_run(0, [v, a, z, t], n, seed=21)
    p_up = (np.exp(-2 * v * a * z) - 1) / (np.exp(-2 * v * a) - 1) if v != 0 else z
    lo, hi = _ci(int(np.sum(ch == 1)), n)
    assert lo - 0.004 <= p_up <= hi + 0.004  # Euler discretisation (dt=1e-3) biases the hit probability slightly
    # CDF of upper-boundary RTs against the analytic density integrated on a fine grid
    grid = np.linspace(t + 1e-4, 8.0, 8000)
    dens = oracle.pdf_array(grid, v, 0, a, z, 0, t, 0, 1e-10)
    cdf = np.concatenate([[0], np.cumsum((dens[1:] + dens[:-1]) / 2 * np.diff(grid))])
    cdf /= cdf[-1]
    up = np.sort(rt[ch[:, 0] == 1, 0].astype(np.float64))
    emp = np.arange(1, len(up) + 1) / len(up)
    D = np.max(np.abs(emp - np.interp(up, grid, cdf)))
    assert D < 0.04, D  # the Euler walk detects crossings late by O(sqrt(dt)): a small, known shift (~0.024 here)
