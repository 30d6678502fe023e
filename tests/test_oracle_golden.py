"""CPU tests (no GPU): pin the C oracle against the reference's own known answers and against fixtures produced
by the reference's Cython (tests/golden/make_golden.py), and - where oracle/_ref exists - bit for bit against it."""
import math

import numpy as np
import pytest

from conftest import rel_err


def test_matlab_known_answers(golden, oracle):
    # /root/reference/hddm/tests/test_likelihoods.py:51-58 (test_pdf_no_matlab): 9 decimals
    for v, t, a, z, z_nonorm, rt, err, matlab_wfpt in golden["matlab"]:
        p = oracle.full_pdf(-rt, v, 0, a, z, 0, t, 0, err, 0)
        assert p == pytest.approx(matlab_wfpt, abs=5e-10)


def test_rt_zero_gives_minus_inf(golden, oracle):
    # test_likelihoods.py:122-127
    assert oracle.wiener_like(np.array([1.0, 2, 3, 0]), 1, 0, 2, 0.5, 0, 0, 0, 1e-4) == -np.inf
    assert golden["wl_rt0"] == -np.inf
    assert oracle.wiener_like(np.empty(0), 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4) == 0.0 == float(golden["wl_empty"])


def test_failure_modes(oracle):
    # test_likelihoods.py:405-548 (test_failure_mode): density exactly 0 outside the support
    f = oracle.full_pdf
    assert f(0.5, 1, 0, 2, 1.1, 0, 0.3, 0, 1e-4) == 0      # z > 1
    assert f(0.5, 1, 0, 2, -0.1, 0, 0.3, 0, 1e-4) == 0     # z < 0
    assert f(0.5, 1, 0, 2, 0.9, 0.3, 0.3, 0, 1e-4) == 0    # z + sz/2 > 1
    assert f(0.5, 1, 0, 2, 0.1, 0.3, 0.3, 0, 1e-4) == 0    # z - sz/2 < 0
    assert f(0.5, 1, 0, -2, 0.5, 0, 0.3, 0, 1e-4) == 0     # a < 0
    assert f(0.2, 1, 0, 2, 0.5, 0, 0.3, 0.1, 1e-4) == 0    # |rt| < t - st/2
    assert f(0.5, 1, 0, 2, 0.5, 0, -0.3, 0, 1e-4) == 0     # t < 0
    assert f(0.5, 1, 0, 2, 0.5, 0, 0.1, 0.4, 1e-4) == 0    # t - st/2 < 0
    assert math.isnan(f(0.5, 1, 0, 0, 0.5, 0, 0.3, 0, 1e-4))  # a == 0 is not masked (Appendix A.5)


def test_threshold_zeroing(oracle):
    # pdf.pxi:122-125: st, sz below 1e-3 are treated as 0 after the validity test
    base = oracle.full_pdf(0.9, 1, 0.5, 2, 0.5, 0, 0.3, 0, 1e-4)
    assert oracle.full_pdf(0.9, 1, 0.5, 2, 0.5, 9e-4, 0.3, 9e-4, 1e-4) == base


def test_pdf_sv_against_quadrature(oracle):
    # test_likelihoods.py:129-153 (test_pdf_sv): closed-form sv versus numeric integration over the drift
    from scipy.integrate import quad
    from scipy.stats import norm

    rng = np.random.default_rng(5)
    for _ in range(20):
        sv = rng.uniform(0.1, 1.0)
        v = rng.normal(0, 1)
        a = rng.uniform(1, 2.5)
        z = rng.uniform(0.3, 0.7)
        x = rng.uniform(0.4, 2.0)
        closed = oracle.full_pdf(x, v, sv, a, z, 0, 0.2, 0, 1e-10, 0, 0, 0)
        num, _ = quad(lambda vi: oracle.full_pdf(x, vi, 0, a, z, 0, 0.2, 0, 1e-10, 0, 0, 0) * norm.pdf(vi, v, sv),
                      v - 10 * sv, v + 10 * sv, epsabs=1e-12, epsrel=1e-10)
        assert closed == pytest.approx(num, rel=1e-6, abs=1e-9)


def test_density_integrates_to_one(oracle):
    # test_likelihoods.py:215-231
    xs = np.linspace(-8, 8, 16001)
    y = oracle.pdf_array(xs, 0.6, 0.4, 1.7, 0.45, 0.1, 0.2, 0.1, 1e-6, 0, 2, 2, 1, 1e-4, 0, 0)
    assert np.trapezoid(y, xs) == pytest.approx(1.0, abs=1e-3)


def test_sz_st_integration_meaning(oracle):
    # test_likelihoods.py:233-403: sz / st are uniform widths, the integral is normalised by the width
    from scipy.integrate import simpson

    x, v, sv, a, z, t, err = 1.1, 0.8, 0.5, 1.9, 0.5, 0.3, 1e-8
    sz, st = 0.3, 0.2
    zs = np.linspace(z - sz / 2, z + sz / 2, 61)
    ts = np.linspace(t - st / 2, t + st / 2, 61)
    want_z = simpson([oracle.full_pdf(x, v, sv, a, zi, 0, t, 0, err) for zi in zs], x=zs) / sz
    got_z = oracle.full_pdf(x, v, sv, a, z, sz, t, 0, err, 60, 60, 0, 1e-3)
    assert got_z == pytest.approx(want_z, rel=1e-7)
    want_t = simpson([oracle.full_pdf(x, v, sv, a, z, 0, ti, 0, err) for ti in ts], x=ts) / st
    got_t = oracle.full_pdf(x, v, sv, a, z, 0, t, st, err, 60, 60, 0, 1e-3)
    assert got_t == pytest.approx(want_t, rel=1e-7)
    adapt = oracle.full_pdf(x, v, sv, a, z, sz, t, st, err, 10, 10, 1, 1e-8)
    fixed = oracle.full_pdf(x, v, sv, a, z, sz, t, st, err, 60, 60, 0, 1e-8)
    assert adapt == pytest.approx(fixed, rel=1e-5)


def test_full_pdf_cases_bit_exact(golden, oracle):
    cases, vals = golden["full_pdf_cases"], golden["full_pdf_vals"]
    got = np.array([oracle.full_pdf(*c[:9], int(c[9]), int(c[10]), int(c[11]), c[12]) for c in cases])
    same = (got == vals) | (np.isnan(got) & np.isnan(vals))
    assert same.all(), f"{(~same).sum()} of {len(vals)} differ from the reference's Cython"


def test_wiener_like_cases_bit_exact(golden, oracle):
    x = golden["wl_x"]
    for c, want in zip(golden["wl_cases"], golden["wl_vals"]):
        got = oracle.wiener_like(x, *c[:8], n_st=int(c[8]), n_sz=int(c[9]), use_adaptive=int(c[10]), simps_err=c[11],
                                 p_outlier=c[12], w_outlier=c[13])
        assert got == want or (math.isnan(got) and math.isnan(want)), (c, got, want)
    # with the HDDM outlier mix an invalid parameter set yields exactly N*log(w*p)
    assert golden["wl_vals"][4] == pytest.approx(5 * math.log(0.005), rel=1e-15)


def test_benchmark_inputs_bit_exact(golden, oracle):
    x1, x2 = golden["x_cfg1"], golden["x_cfg2"]
    assert oracle.wiener_like(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4) == float(golden["cfg1_logp"]) == -927.6658171227821
    kw = dict(p_outlier=0.05, w_outlier=0.1)
    assert oracle.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=0, simps_err=1e-3,
                              **kw) == float(golden["cfg2a_logp"]) == -10872.759152153509
    assert oracle.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                              **kw) == float(golden["cfg2b_logp"]) == -10872.29410744754
    got = oracle.pdf_array(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, 0, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert np.array_equal(got, golden["cfg2b_pdf"])


def test_cfg4_slices_bit_exact(golden, oracle):
    x2, P4 = golden["x_cfg2"], golden["cfg4_params"]
    for tag, ad in (("fixed", 0), ("adaptive", 1)):
        got = oracle.wiener_like_batch(x2, P4, 1e-4, 2, 2, ad, 1e-3, 0.1)
        want = golden["cfg4_" + tag]
        assert np.array_equal(got, want, equal_nan=True)


def test_adaptive_node_census(golden, oracle):
    # at simps_err=1e-3 every cfg-2 trial stops at the first level -> exactly 25 nodes per trial
    x2 = golden["x_cfg2"][:2000]
    oracle.nodes_reset()
    oracle.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3,
                       p_outlier=0.05, w_outlier=0.1)
    assert oracle.nodes_get() == 25 * len(x2)
    oracle.nodes_reset()
    oracle.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=0, simps_err=1e-3,
                       p_outlier=0.05, w_outlier=0.1)
    assert oracle.nodes_get() == 9 * len(x2)


def test_multi_bit_exact(golden, oracle):
    g = golden
    xm, vm, am, tm, zm = g["multi_x"], g["multi_v"], g["multi_a"], g["multi_t"], g["multi_z"]
    assert oracle.wiener_like_multi(xm, vm, 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.05, w_outlier=0.1) == float(g["multi_v_only"])
    assert oracle.wiener_like_multi(xm, vm, 1.0, am, 0.5, 0.1, tm, 0.1, 1e-4, ["v", "a", "t"], p_outlier=0.05, w_outlier=0.1) == float(g["multi_vat"])
    assert oracle.wiener_like_multi(xm, vm, 0.5, am, zm, 0.1, tm, 0.05, 1e-4, ["v", "a", "z", "t"], n_st=2, n_sz=2, use_adaptive=0, p_outlier=0.0, w_outlier=0.0) == float(g["multi_vazt_fixed"])
    xm2 = xm.copy()
    xm2[3] = 0.05
    assert oracle.wiener_like_multi(xm2, vm, 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.0, w_outlier=0.1) == -np.inf == float(g["multi_neginf"])
    assert oracle.wiener_like_multi(g["wl_x"], np.linspace(0.5, 1.5, 5), 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.05, w_outlier=0.1) == float(g["multi_small"])
    assert oracle.wiener_like_multi(np.array([0.8, 999.0, -999.0]), np.ones(3), 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"]) == float(g["multi_999"])


def test_rl_flavours_bit_exact(golden, oracle):
    g = golden
    xr, resp, fb, split = g["rl_x"], g["rl_response"], g["rl_feedback"], g["rl_split"]
    assert oracle.wiener_like_rlddm(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.3, 0, 1e-4, p_outlier=0.05, w_outlier=0.1) == float(g["rlddm_a"])
    assert oracle.wiener_like_rlddm(xr, resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.5, 1.8, 0.5, 0.1, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1) == float(g["rlddm_b"])
    assert oracle.wiener_like_rlddm(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.9, 0, 1e-4) == -np.inf == float(g["rlddm_neginf"])
    assert oracle.wiener_like_rl(resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0.5, p_outlier=0.05, w_outlier=0.1) == float(g["rl_a"])
    assert oracle.wiener_like_rl(resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.45) == float(g["rl_b"])
    assert oracle.wiener_like_multi_rlddm(xr, resp, fb, g["rl_split_sorted"], 0.5, g["rl_v_arr"], 0, 1.8, 0.5, 0, 0.3, 0, g["rl_alpha_arr"], 1e-4, ["v", "alpha"], p_outlier=0.05, w_outlier=0.1) == float(g["multi_rlddm"])
    assert oracle.wiener_like_contaminant(xr, g["cont_x"], 1.0, 0.5, 1.8, 0.5, 0.1, 0.3, 0.1, 0.1, 5.0, 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3) == float(g["contaminant"])


def test_pdf_grid(golden, oracle):
    g = golden
    got = oracle.pdf_array(g["pdf_grid_x"], 0.7, 0.8, 1.6, 0.45, 0.2, 0.25, 0.15, 1e-4, 0, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert np.array_equal(got, g["pdf_grid_full"])
    got = oracle.pdf_array(g["pdf_grid_x"], 0.7, 0, 1.6, 0.45, 0, 0.25, 0, 1e-4, 1, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert rel_err(got, g["pdf_grid_simple_log"]) < 1e-15  # np.log (NumPy SIMD) vs libm log: <= 1 ulp


def test_gen_rts_from_cdf_matches_reference_inputs(golden, oracle):
    np.random.seed(1)
    x1 = oracle.gen_rts_from_cdf(1, 0, 2, 0.5, 0, 0.3, 0, 1000, -6, 6, 1e-3)
    assert np.array_equal(x1, golden["x_cfg1"])


def test_against_live_reference(ref_wfpt, oracle):
    """Where the reference's Cython is built (this container): random bit-for-bit comparison."""
    rng = np.random.default_rng(99)
    for _ in range(3000):
        v, a, z, t = rng.uniform(-5, 5), rng.uniform(0.1, 5), rng.uniform(0.05, 0.95), rng.uniform(0, 1)
        sv = rng.choice([0.0, rng.uniform(0, 4)])
        sz = rng.choice([0.0, rng.uniform(0, 0.3)])
        st = rng.choice([0.0, rng.uniform(0, 0.3)])
        x = rng.uniform(-4, 4)
        err = 10 ** rng.uniform(-12, 0)
        ad, n, se = int(rng.integers(0, 2)), int(rng.choice([2, 4, 6])), 10 ** rng.uniform(-7, -2)
        p1 = oracle.full_pdf(x, v, sv, a, z, sz, t, st, err, n, n, ad, se)
        p2 = ref_wfpt.full_pdf(x, v, sv, a, z, sz, t, st, err, n, n, ad, se)
        assert p1 == p2 or (math.isnan(p1) and math.isnan(p2))


def test_cdf_fixture_comes_from_the_reference_library():
    """tests/golden/cdf_golden.npz is what the reference's own cdfdif.c (oracle/_ref) returns, bit for bit."""
    import os

    from oracle import cpu
    from oracle.build_ref import CDF_LIB

    if not os.path.exists(CDF_LIB):
        pytest.skip("oracle/_ref/libcdfdif_ref.so not present")
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "cdf_golden.npz")))
    for row, want in zip(g["cases"], g["vals"]):
        assert np.array_equal(cpu.dmat_cdf_array_ref(g["x"], *row), want)
    # sanity of the quantity itself: in signed-RT convention the CDF rises from 0 to 1 (outlier-free row)
    y = g["vals"][1]
    assert abs(y[0]) < 5e-3 and abs(y[180] - 1) < 5e-3 and np.all(np.diff(y[:181]) >= -1e-12)


def test_oracle_matches_whole_grid_fixture_on_a_sample(grid_golden, oracle):
    """The C oracle against the reference's answers for the benchmark grids (every 211th benign set, every 61st of the
    whole-box grid, both integration modes, and the hand-picked special sets): bit for bit."""
    g = grid_golden
    x = g["x"]
    idx = np.arange(0, 65536, 211)
    got = oracle.wiener_like_batch(x, g["benign"][idx], 1e-4, 2, 2, 0, 1e-3, 0.1)
    assert np.array_equal(got, g["benign_fixed"][idx])
    idx = np.arange(0, 8192, 61)
    for tag, ad in (("fixed", 0), ("adaptive", 1)):
        got = oracle.wiener_like_batch(x, g["bounds"][idx], 1e-4, 2, 2, ad, 1e-3, 0.1)
        assert np.array_equal(got, g["bounds_" + tag][idx], equal_nan=True)
        got = oracle.wiener_like_batch(x, g["special_params"], 1e-4, 2, 2, ad, 1e-3, 0.1)
        assert np.array_equal(got, g["special_" + tag], equal_nan=True)
    got = oracle.wiener_like_batch(x, g["benign"][idx], 1e-4, 2, 2, 1, 1e-3, 0.1)
    assert np.array_equal(got, g["benign_adaptive"][idx])
    # the special sets say what they are meant to say
    sp = g["special_fixed"]
    assert sp[0] == pytest.approx(len(x) * math.log(0.1 * 0.05), rel=1e-12) and np.isneginf(sp[[1, 2, 4]]).all()
    assert np.isnan(sp[5]) and np.isfinite(sp[[3, 6, 7, 8, 9]]).all()


def test_fixed_grid_with_odd_interval_counts(oracle, ref_wfpt):
    """integrate.pxi:30-45 with an odd number of intervals (the "n even" asserts are commented out): the last node ends up
    with weight 3.  The oracle follows the reference bit for bit there too."""
    rng = np.random.default_rng(12)
    x = np.where(rng.random(40) < 0.7, 1.0, -1.0) * rng.uniform(0.4, 3.0, 40)
    for n_st, n_sz in ((1, 1), (3, 3), (5, 2), (2, 5), (1, 4), (3, 0), (0, 3)):
        for sz, st in ((0.2, 0.1), (0.0, 0.1), (0.2, 0.0)):
            if (sz and not n_sz) or (st and not n_st):
                continue
            want = ref_wfpt.wiener_like(x, 0.8, 0.9, 1.9, 0.45, sz, 0.3, st, 1e-4, n_st=n_st, n_sz=n_sz, use_adaptive=0,
                                        simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
            got = oracle.wiener_like(x, 0.8, 0.9, 1.9, 0.45, sz, 0.3, st, 1e-4, n_st=n_st, n_sz=n_sz, use_adaptive=0,
                                     simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
            assert got == want, (n_st, n_sz, sz, st)
