"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against the CPU oracle
on the same seeded inputs, against the committed fixtures produced by the reference's own Cython, and through
size-independent properties at full size.  Tolerance for summed log-likelihoods: 1e-10 relative (north_star);
bit-exact for the outlier mix of masked trials, the -inf masking and the p_outlier range check."""
import math

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10
VARIANTS = [0, 1]  # FAST, FAITHFUL


@pytest.fixture(params=VARIANTS, ids=["fast", "faithful"])
def variant(request, wfpt_gpu):
    old = wfpt_gpu.get_variant()
    wfpt_gpu.set_variant(request.param)
    yield request.param
    wfpt_gpu.set_variant(old)


def test_matlab_known_answers(golden, wfpt_gpu, variant):
    # /root/reference/hddm/tests/test_likelihoods.py:51-58
    for v, t, a, z, z_nonorm, rt, err, matlab_wfpt in golden["matlab"]:
        p = wfpt_gpu.full_pdf(-rt, v, 0, a, z, 0, t, 0, err, 0)
        assert p == pytest.approx(matlab_wfpt, abs=5e-10)


def test_full_pdf_cases(golden, wfpt_gpu, variant):
    """6000 single-trial densities over every dispatch branch (sv/sz/st on/off, fixed/adaptive, err 1e-13..1)."""
    cases, vals = golden["full_pdf_cases"], golden["full_pdf_vals"]
    # group by option tuple so that each group is one pdf_array-style launch per case row (n=1 calls are slow)
    got = np.array([wfpt_gpu.full_pdf(*c[:9], int(c[9]), int(c[10]), int(c[11]), c[12]) for c in cases])
    want = vals
    assert np.array_equal(got == 0, want == 0)                 # validity mask: exact zeros in the same places
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = np.isfinite(want) & (np.abs(want) > 1e-300)
    # a K flip (1-ulp libm difference at an integer ks/kl) would show up here as an error of order `err`
    assert rel_err(got[ok], want[ok]) < 1e-9
    tiny = np.isfinite(want) & (want != 0) & ~ok     # subnormal results carry only a few digits in either implementation
    assert np.all(np.abs(got[tiny] - want[tiny]) < 1e-300)


def test_wiener_like_cases(golden, wfpt_gpu, variant):
    x = golden["wl_x"]
    for c, want in zip(golden["wl_cases"], golden["wl_vals"]):
        got = wfpt_gpu.wiener_like(x, *c[:8], n_st=int(c[8]), n_sz=int(c[9]), use_adaptive=int(c[10]), simps_err=c[11],
                                   p_outlier=c[12], w_outlier=c[13])
        if math.isnan(want):
            assert math.isnan(got), c
        elif math.isinf(want):
            assert got == want, c
        else:
            assert abs(got - want) <= TOL * abs(want), (c, got, want)
    # bit-exact pieces: invalid set under the HDDM outlier mix == N*log(w*p) to the last bit
    c = golden["wl_cases"][4]
    got = wfpt_gpu.wiener_like(x, *c[:8], n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=c[12], w_outlier=c[13])
    assert abs(got - golden["wl_vals"][4]) <= 4e-15 * abs(got)
    assert wfpt_gpu.wiener_like(np.array([1.0, 2, 3, 0]), 1, 0, 2, 0.5, 0, 0, 0, 1e-4) == -np.inf
    assert wfpt_gpu.wiener_like(np.empty(0), 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4) == 0.0


def test_benchmark_configs(golden, wfpt_gpu, variant):
    x1, x2 = golden["x_cfg1"], golden["x_cfg2"]
    got = wfpt_gpu.wiener_like(x1, 1, 0, 2, 0.5, 0, 0.3, 0, 1e-4)
    assert abs(got - float(golden["cfg1_logp"])) <= TOL * abs(got)
    kw = dict(p_outlier=0.05, w_outlier=0.1)
    for tag, ad, n, se in (("cfg2a_logp", 0, 2, 1e-3), ("cfg2b_logp", 1, 2, 1e-3), ("cfg2c_logp", 1, 10, 1e-8)):
        got = wfpt_gpu.wiener_like(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, n_st=n, n_sz=n, use_adaptive=ad, simps_err=se, **kw)
        want = float(golden[tag])
        assert abs(got - want) <= TOL * abs(want), (tag, got, want)
    y = wfpt_gpu.pdf_array(x2, 1, 1, 2, 0.5, 0.2, 0.3, 0.1, 1e-4, 0, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert rel_err(y, golden["cfg2b_pdf"]) < 1e-11


def test_cfg4_slices_vs_reference_fixture(golden, wfpt_gpu, variant):
    x2, P4 = golden["x_cfg2"], golden["cfg4_params"]
    for tag, ad in (("fixed", 0), ("adaptive", 1)):
        got = wfpt_gpu.wiener_like_batch(x2, P4, 1e-4, 2, 2, ad, 1e-3, 0.1, variant=variant)
        want = golden["cfg4_" + tag]
        assert np.array_equal(np.isinf(got), np.isinf(want)) and np.array_equal(np.isnan(got), np.isnan(want))
        assert rel_err(got, want) <= TOL


def test_batch_vs_oracle_random(oracle, wfpt_gpu, variant):
    rng = np.random.default_rng(21)
    n, P = 3001, 257  # odd sizes: head/tail paths of the vectorised loads, partial slices
    x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.3, 4.0, n)
    prm = np.stack([rng.uniform(-4, 4, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 3, P), rng.uniform(0.5, 4, P),
                    rng.uniform(0.2, 0.8, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                    rng.uniform(0.05, 0.5, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                    rng.choice([0.0, 0.05, 1.0, 1.5], P)], axis=1)
    for ad, nn, se in ((0, 2, 1e-3), (0, 4, 1e-3), (1, 2, 1e-3), (1, 4, 1e-5)):
        want = oracle.wiener_like_batch(x, prm, 1e-4, nn, nn, ad, se, 0.1)
        got = wfpt_gpu.wiener_like_batch(x, prm, 1e-4, nn, nn, ad, se, 0.1, variant=variant)
        assert np.array_equal(np.isneginf(got), np.isneginf(want))
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert rel_err(got, want) <= TOL, (ad, nn, se)
    # unaligned view of x (8-byte aligned only) and P=1 with many slices
    want = oracle.wiener_like_batch(x[1:], prm[:3], 1e-4, 2, 2, 0, 1e-3, 0.1)
    got = wfpt_gpu.wiener_like_batch(x[1:], prm[:3], 1e-4, 2, 2, 0, 1e-3, 0.1, variant=variant)
    assert rel_err(got, want) <= TOL


def test_batch_device_tensors_match_host_buffers(golden, wfpt_gpu, variant):
    import torch

    x2, P4 = golden["x_cfg2"], golden["cfg4_params"]
    host = wfpt_gpu.wiener_like_batch(x2, P4, 1e-4, 2, 2, 0, 1e-3, 0.1, variant=variant, presort=False)
    dev = wfpt_gpu.wiener_like_batch(torch.from_numpy(x2).cuda(), torch.from_numpy(P4).cuda(), 1e-4, 2, 2, 0, 1e-3, 0.1,
                                     variant=variant)
    torch.cuda.synchronize()
    assert np.array_equal(host, dev.cpu().numpy(), equal_nan=True)


def test_multi(golden, wfpt_gpu, variant):
    g = golden
    xm, vm, am, tm, zm = g["multi_x"], g["multi_v"], g["multi_a"], g["multi_t"], g["multi_z"]
    f = wfpt_gpu.wiener_like_multi
    for got, want in (
        (f(xm, vm, 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.05, w_outlier=0.1), g["multi_v_only"]),
        (f(xm, vm, 1.0, am, 0.5, 0.1, tm, 0.1, 1e-4, ["v", "a", "t"], p_outlier=0.05, w_outlier=0.1), g["multi_vat"]),
        (f(xm, vm, 0.5, am, zm, 0.1, tm, 0.05, 1e-4, ["v", "a", "z", "t"], n_st=2, n_sz=2, use_adaptive=0, p_outlier=0.0, w_outlier=0.0), g["multi_vazt_fixed"]),
        (f(g["wl_x"], np.linspace(0.5, 1.5, 5), 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.05, w_outlier=0.1), g["multi_small"]),
        (f(np.array([0.8, 999.0, -999.0]), np.ones(3), 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"]), g["multi_999"]),
    ):
        assert abs(got - float(want)) <= TOL * abs(float(want)), (got, want)
    xm2 = xm.copy()
    xm2[3] = 0.05
    assert f(xm2, vm, 0, 2, 0.5, 0, 0.3, 0, 1e-4, ["v"], p_outlier=0.0, w_outlier=0.1) == -np.inf


def test_rlddm(golden, wfpt_gpu, variant):
    g = golden
    xr, resp, fb, split = g["rl_x"], g["rl_response"], g["rl_feedback"], g["rl_split"]
    f = wfpt_gpu.wiener_like_rlddm
    got = f(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.3, 0, 1e-4, p_outlier=0.05, w_outlier=0.1)
    assert abs(got - float(g["rlddm_a"])) <= TOL * abs(got)
    got = f(xr, resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.5, 1.8, 0.5, 0.1, 0.3, 0.1, 1e-4, n_st=2, n_sz=2, use_adaptive=1,
            simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
    assert abs(got - float(g["rlddm_b"])) <= TOL * abs(got)
    assert f(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.9, 0, 1e-4) == -np.inf
    assert f(xr, resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0, 1.8, 0.5, 0, 0.3, 0, 1e-4, p_outlier=2.0) == -np.inf


def test_rl_multi_rlddm_contaminant(golden, wfpt_gpu, variant):
    g = golden
    xr, resp, fb, split = g["rl_x"], g["rl_response"], g["rl_feedback"], g["rl_split"]
    got = wfpt_gpu.wiener_like_rl(resp, fb, split, 0.5, -1.0, 100.0, 2.0, 0.5, p_outlier=0.05, w_outlier=0.1)
    assert abs(got - float(g["rl_a"])) <= TOL * abs(got)
    got = wfpt_gpu.wiener_like_rl(resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.45)
    assert abs(got - float(g["rl_b"])) <= TOL * abs(got)
    assert wfpt_gpu.wiener_like_rl(resp, fb, split, 0.5, -0.5, 0.7, 3.0, 0.45, p_outlier=-1) == -np.inf
    got = wfpt_gpu.wiener_like_multi_rlddm(xr, resp, fb, g["rl_split_sorted"], 0.5, g["rl_v_arr"], 0, 1.8, 0.5, 0, 0.3, 0,
                                           g["rl_alpha_arr"], 1e-4, ["v", "alpha"], p_outlier=0.05, w_outlier=0.1)
    assert abs(got - float(g["multi_rlddm"])) <= TOL * abs(got)
    got = wfpt_gpu.wiener_like_contaminant(xr, g["cont_x"], 1.0, 0.5, 1.8, 0.5, 0.1, 0.3, 0.1, 0.1, 5.0, 1e-4, n_st=2, n_sz=2,
                                           use_adaptive=1, simps_err=1e-3)
    assert abs(got - float(g["contaminant"])) <= TOL * abs(got)
    # a non-contaminant trial with zero density -> -inf, contaminant term or not
    assert wfpt_gpu.wiener_like_contaminant(xr, g["cont_x"], 1.0, 0.5, 1.8, 0.5, 0.1, 0.9, 0.1, 0.1, 5.0, 1e-4) == -np.inf


def test_segments_resident_data(oracle, wfpt_gpu, variant):
    """60 nodes of ragged length, one parameter set each (config 3's sweep), in one launch; subsets and repeats."""
    rng = np.random.default_rng(33)
    arrays = [np.where(rng.random(n) < 0.8, 1.0, -1.0) * rng.uniform(0.35, 3.0, n)
              for n in rng.integers(1, 700, 60)]
    arrays[7] = np.empty(0)          # an empty node contributes 0.0
    P = len(arrays)
    prm = np.stack([rng.normal(1, 0.5, P), rng.uniform(0, 1, P), rng.uniform(1.5, 2.5, P), rng.uniform(0.4, 0.6, P),
                    rng.uniform(0, 0.2, P), rng.uniform(0.1, 0.3, P), rng.uniform(0, 0.15, P), np.full(P, 0.05)], axis=1)
    kw = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, w_outlier=0.1)
    want = np.array([oracle.wiener_like(a, *q[:7], 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=q[7],
                                        w_outlier=0.1) for a, q in zip(arrays, prm)])
    rd = wfpt_gpu.ResidentData(arrays)
    got = rd.logp(prm, variant=variant, **kw)
    assert want[7] == 0.0 and got[7] == 0.0
    assert rel_err(got, want) <= TOL
    idx = np.array([3, 3, 59, 0, 3])
    got = rd.logp(prm[idx], nodes=idx, variant=variant, **kw)
    assert rel_err(got, want[idx]) <= TOL
    # device-resident parameters (torch tensors) through the same handle's device pointers
    import torch

    got_t = rd.logp(torch.from_numpy(prm).cuda(), variant=variant, **kw)
    assert got_t.is_cuda and np.array_equal(got_t.cpu().numpy(), rd.logp(prm, variant=variant, **kw))
    got_t = rd.logp(torch.from_numpy(prm[idx]).cuda(), nodes=idx, variant=variant, **kw)
    assert rel_err(got_t.cpu().numpy(), want[idx]) <= TOL
    with pytest.raises(wfpt_gpu.WfptError):
        rd.logp(prm[:2], nodes=np.array([0, 60]), variant=variant, **kw)
    rd.close()
    got = wfpt_gpu.wiener_like_segments(arrays[:5], prm[:5], variant=variant, **kw)
    assert rel_err(got, want[:5]) <= TOL
    # one long node: the segment is cut into many slices
    big = [np.where(rng.random(9001) < 0.8, 1.0, -1.0) * rng.uniform(0.35, 3.0, 9001), arrays[1]]
    got = wfpt_gpu.ResidentData(big).logp(prm[:2], variant=variant, **kw)
    w2 = [oracle.wiener_like(a, *q[:7], 1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=q[7], w_outlier=0.1)
          for a, q in zip(big, prm[:2])]
    assert rel_err(got, np.array(w2)) <= TOL


def test_pdf_array_and_cdf_helpers(golden, oracle, wfpt_gpu, variant):
    g = golden
    got = wfpt_gpu.pdf_array(g["pdf_grid_x"], 0.7, 0.8, 1.6, 0.45, 0.2, 0.25, 0.15, 1e-4, 0, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert rel_err(got, g["pdf_grid_full"]) < 1e-11
    # masked trials carry exactly w_outlier*p_outlier (bit-exact outlier mix)
    masked = np.abs(g["pdf_grid_x"]) < 0.25 - 0.15 / 2
    assert masked.any() and np.all(got[masked] == 0.1 * 0.05)
    got = wfpt_gpu.pdf_array(g["pdf_grid_x"], 0.7, 0, 1.6, 0.45, 0, 0.25, 0, 1e-4, 1, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert rel_err(got, g["pdf_grid_simple_log"]) < 1e-12
    # gen_rts_from_cdf reproduces the benchmark inputs generated by the reference (same global-RNG call sequence);
    # a sample sitting exactly on a CDF step may move by one grid cell under a 1e-16 perturbation
    np.random.seed(1)
    x1 = wfpt_gpu.gen_rts_from_cdf(1, 0, 2, 0.5, 0, 0.3, 0, 1000, -6, 6, 1e-3)
    assert np.mean(x1 == g["x_cfg1"]) > 0.995 and np.max(np.abs(x1 - g["x_cfg1"])) <= 1.5e-3
    xs, cdf = wfpt_gpu.gen_cdf_using_pdf(0.7, 0.8, 1.6, 0.45, 0.2, 0.25, 0.15, 1e-4, N=300, time=5.0)
    assert 0 <= cdf[0] < 1e-4 and cdf[-1] == 1 and np.all(np.diff(cdf) >= 0) and len(xs) == 601
    # numbers, not only shape (wfpt.pyx:582-607): the oracle's pdf_array on the same grid + scipy's unit-spacing cumulative
    # trapezoid (the `cumtrapz` the reference calls, under its current name) + the same normalisation
    from scipy.integrate import cumulative_trapezoid

    for args, kw in (((0.7, 0.8, 1.6, 0.45, 0.2, 0.25, 0.15, 1e-4), dict(N=300, time=5.0)),
                     ((-1.1, 0.0, 2.1, 0.6, 0.0, 0.3, 0.0, 1e-4), dict(N=500, time=5.0, p_outlier=0.05, w_outlier=0.1)),
                     ((0.3, 1.5, 1.2, 0.5, 0.1, 0.2, 0.2, 1e-6), dict(N=64, time=3.0, n_st=4, n_sz=4, use_adaptive=0))):
        xs, cdf = wfpt_gpu.gen_cdf_using_pdf(*args, **kw)
        N, time = kw["N"], kw["time"]
        grid = np.linspace(-time, time, 2 * N + 1)
        assert np.array_equal(xs, grid)
        dens = oracle.pdf_array(grid, *args, 0, kw.get("n_st", 2), kw.get("n_sz", 2), kw.get("use_adaptive", 1),
                                kw.get("simps_err", 1e-3), kw.get("p_outlier", 0), kw.get("w_outlier", 0))
        want = np.empty_like(dens)
        want[0] = dens[0]
        want[1:] = cumulative_trapezoid(dens)
        want /= want[-1]
        assert np.max(np.abs(cdf - want)) < 1e-12, args
        # split_cdf (wfpt.pyx:610-626) on the GPU-made CDF against the same slicing done by hand on the oracle's
        x_lb, lb, x_ub, ub = wfpt_gpu.split_cdf(xs, cdf.copy())
        w_lb = np.cumsum(np.concatenate([[0.0], -np.diff(want[:N][::-1])]))
        w_ub = want[N + 1:] - want[N + 1]
        assert np.array_equal(x_lb, -grid[:N][::-1]) and np.array_equal(x_ub, grid[N + 1:])
        assert np.max(np.abs(lb - w_lb)) < 1e-12 and np.max(np.abs(ub - w_ub)) < 1e-12
    with pytest.raises(ValueError):
        wfpt_gpu.gen_cdf_using_pdf(0.7, 0.8, 1.6, 1.2, 0.2, 0.25, 0.15, 1e-4)


def test_density_integrates_to_one(wfpt_gpu, variant):
    # test_likelihoods.py:215-231, at a size the CPU oracle would not want: 400k-point quadrature
    xs = np.linspace(-8, 8, 400001)
    y = wfpt_gpu.pdf_array(xs, 0.6, 0.4, 1.7, 0.45, 0.1, 0.2, 0.1, 1e-6, 0, 2, 2, 1, 1e-4, 0, 0)
    assert np.trapezoid(y, xs) == pytest.approx(1.0, abs=1e-3)


def test_full_size_properties(golden, wfpt_gpu):
    """cfg-4 sized launch (4096 sets x 10,000 trials here; the bench runs 65,536): (i) permutation invariance of
    the trial sum within rounding, (ii) additivity over a split of the trials, (iii) fast == faithful."""
    rng = np.random.default_rng(4)
    x2 = golden["x_cfg2"]
    P = 4096
    prm = np.stack([rng.uniform(-5, 5, P), rng.uniform(0, 4, P), rng.uniform(1, 3, P), rng.uniform(0.3, 0.7, P),
                    rng.uniform(0, 0.25, P), rng.uniform(0.1, 0.3, P), rng.uniform(0, 0.2, P), np.full(P, 0.05)], axis=1)
    kw = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=0, simps_err=1e-3, w_outlier=0.1)
    full = wfpt_gpu.wiener_like_batch(x2, prm, **kw)
    assert np.all(np.isfinite(full))
    perm = wfpt_gpu.wiener_like_batch(x2[rng.permutation(len(x2))], prm, **kw)
    assert rel_err(perm, full) < 1e-12
    a = wfpt_gpu.wiener_like_batch(x2[:4000], prm, **kw)
    b = wfpt_gpu.wiener_like_batch(x2[4000:], prm, **kw)
    assert rel_err(a + b, full) < 1e-12
    faithful = wfpt_gpu.wiener_like_batch(x2, prm, variant=1, **kw)
    assert rel_err(full, faithful) <= TOL


def test_refinement_paths_vs_oracle(oracle, wfpt_gpu):
    """Adaptive recursion below the first level (config 4 over the whole param_bounds box: RTs inside the t +- st/2
    window switch the integrand on inside the interval, which is what makes the reference refine): per-trial densities
    and summed log-likelihoods against the oracle, for the HDDM budget (2, 2) and for deeper / asymmetric budgets."""
    rng = np.random.default_rng(33)
    P, n = 96, 700
    prm = np.stack([rng.uniform(-5, 5, P), rng.uniform(0, 4, P), rng.uniform(0.1, 5, P), rng.uniform(0.3, 0.7, P),
                    rng.uniform(0, 0.25, P), rng.uniform(0.25, 2.25, P), rng.uniform(0, 0.25, P), np.full(P, 0.05)], axis=1)
    prm[::7, 4] = 0.0    # t-only integration
    prm[3::7, 6] = 0.0   # z-only integration
    x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.3, 3.0, n)
    for n_st, n_sz, se in ((2, 2, 1e-3), (2, 2, 1e-6), (2, 3, 1e-4), (3, 2, 1e-4)):
        want = oracle.wiener_like_batch(x, prm, 1e-4, n_st, n_sz, 1, se, 0.1)
        got = wfpt_gpu.wiener_like_batch(x, prm, 1e-4, n_st, n_sz, 1, se, 0.1, variant=wfpt_gpu.FAST)
        assert np.array_equal(np.isfinite(got), np.isfinite(want)), (n_st, n_sz, se)
        ok = np.isfinite(want)
        assert rel_err(got[ok], want[ok]) < TOL, (n_st, n_sz, se)
    # per-trial values for a few sets (a wrong node anywhere shows up undiluted)
    refined = 0
    for row in prm[:12]:
        oracle.nodes_reset()
        want = oracle.pdf_array(x, *row[:7], 1e-4, 0, 2, 2, 1, 1e-3, 0.0, 0.0)
        first_level = (5 if row[4] >= 1e-3 else 1) * (5 if row[6] >= 1e-3 else 1)
        refined += oracle.nodes_get() > first_level * int(np.sum(np.abs(x) - (row[5] - row[6] / 2) >= 0))
        got = wfpt_gpu.pdf_array(x, *row[:7], 1e-4, 0, 2, 2, 1, 1e-3, 0.0, 0.0)
        assert np.array_equal(got == 0, want == 0)
        nz = np.abs(want) > 1e-290      # (subnormal results carry only a few digits in either implementation)
        assert np.max(np.abs(got[nz] - want[nz]) / np.abs(want[nz])) < 1e-9
        assert np.all(np.abs(got[~nz] - want[~nz]) < 1e-290)
    assert refined >= 6   # the case really exercises the levels below the first


@pytest.mark.parametrize("n", [1, 2, 127, 128, 129, 255, 256, 257, 1000, 4099, 20001])
def test_single_call_sizes(oracle, wfpt_gpu, n, variant):
    """One wiener_like call at sizes around the launch-geometry switches (one RT per thread for small grids, pairs
    otherwise; pinned staging), odd lengths and unaligned views included."""
    rng = np.random.default_rng(n)
    buf = np.where(rng.random(n + 1) < 0.8, 1.0, -1.0) * rng.uniform(0.35, 4.0, n + 1)
    for x in (np.ascontiguousarray(buf[:n]), buf[1:]):      # the second view starts 8 bytes off a 16-byte boundary
        for args, kw in (((1.0, 0.0, 2.0, 0.5, 0.0, 0.3, 0.0, 1e-4), {}),
                         ((0.7, 1.1, 1.8, 0.45, 0.2, 0.3, 0.1, 1e-4),
                          dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)),
                         ((0.7, 1.1, 1.8, 0.45, 0.2, 0.3, 0.1, 1e-4), dict(n_st=2, n_sz=2, use_adaptive=0, simps_err=1e-3))):
            want = oracle.wiener_like(x, *args, **kw)
            got = wfpt_gpu.wiener_like(x, *args, **kw)
            assert abs(got - want) <= TOL * abs(want), (n, args)


def test_batch_host_sizes_through_both_staging_paths(oracle, wfpt_gpu):
    """P on both sides of the pinned-staging limits (results block 8,192 sets, input block 1 MiB)."""
    rng = np.random.default_rng(9)
    x = np.where(rng.random(64) < 0.8, 1.0, -1.0) * rng.uniform(0.4, 3.0, 64)
    for P in (1, 511, 513, 8192, 8193, 20000):
        prm = np.stack([rng.uniform(-2, 2, P), rng.uniform(0, 2, P), rng.uniform(1, 3, P), rng.uniform(0.3, 0.7, P),
                        rng.uniform(0, 0.2, P), rng.uniform(0.1, 0.3, P), rng.uniform(0, 0.2, P), np.full(P, 0.05)], axis=1)
        got = wfpt_gpu.wiener_like_batch(x, prm, 1e-4, 2, 2, 1, 1e-3, 0.1)
        idx = rng.choice(P, size=min(P, 40), replace=False)
        want = oracle.wiener_like_batch(x, prm[idx], 1e-4, 2, 2, 1, 1e-3, 0.1)
        assert rel_err(got[idx], want) < TOL, P
        assert np.all(np.isfinite(got))


def test_random_option_sweep_vs_oracle(oracle, wfpt_gpu):
    """Random calls over the option space (sizes around the launch-plan switches, fixed / adaptive integration with
    budgets 0..4 and grids 2..6, err, simps_err, sv/sz/st on or off, presorted or not): same NaN / inf pattern as the
    oracle, finite values within TOL."""
    rng = np.random.default_rng(77)
    for _ in range(150):
        n = int(rng.choice([1, 3, 17, 128, 129, 300, 1000, 2500]))
        x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.25, 3.5, n)
        P = int(rng.choice([1, 1, 2, 7, 33]))
        prm = np.stack([rng.uniform(-4, 4, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 3, P), rng.uniform(0.4, 4, P),
                        rng.uniform(0.25, 0.75, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                        rng.uniform(0.05, 1.5, P), rng.choice([0.0, 1.0], P) * rng.uniform(0, 0.3, P),
                        rng.choice([0.0, 0.05], P)], axis=1)
        ad = int(rng.integers(0, 2))
        if ad:
            n_st, n_sz = int(rng.integers(0, 5)), int(rng.integers(0, 5))
        else:
            n_st, n_sz = int(rng.choice([2, 4, 6])), int(rng.choice([2, 4]))
        se, err = float(rng.choice([1e-2, 1e-3, 1e-5, 1e-8])), float(rng.choice([1e-3, 1e-4, 1e-6]))
        want = oracle.wiener_like_batch(x, prm, err, n_st, n_sz, ad, se, 0.1)
        got = wfpt_gpu.wiener_like_batch(x, prm, err, n_st, n_sz, ad, se, 0.1, presort=bool(rng.integers(0, 2)))
        ctx = dict(n=n, P=P, ad=ad, n_st=n_st, n_sz=n_sz, se=se, err=err)
        assert np.array_equal(np.isnan(got), np.isnan(want)), ctx
        assert np.array_equal(np.isinf(got), np.isinf(want)), ctx
        fin = np.isfinite(want)
        if fin.any():
            assert rel_err(got[fin], want[fin]) < TOL, ctx


def test_whole_benchmark_grids_vs_reference_fixture(grid_golden, wfpt_gpu):
    """Every set of the benchmark grids against the reference's own answers (tests/golden/grid_golden.npz): the 65,536
    benign sets on the fixed grid (the headline workload), the first 8,192 of them with adaptive integration, the 8,192
    sets over the whole param_bounds box in both modes, and the hand-picked special sets."""
    import torch

    g = grid_golden
    dx = torch.from_numpy(wfpt_gpu.sort_rts(g["x"])).cuda()
    runs = (("benign", "benign_fixed", 0, slice(None)), ("benign", "benign_adaptive", 1, slice(0, 8192)),
            ("bounds", "bounds_fixed", 0, slice(None)), ("bounds", "bounds_adaptive", 1, slice(None)),
            ("special_params", "special_fixed", 0, slice(None)), ("special_params", "special_adaptive", 1, slice(None)))
    for pname, wname, ad, sl in runs:
        prm = torch.from_numpy(np.ascontiguousarray(g[pname][sl])).cuda()
        got = wfpt_gpu.wiener_like_batch(dx, prm, 1e-4, 2, 2, ad, 1e-3, 0.1).cpu().numpy()
        want = g[wname]
        assert got.shape == want.shape
        assert np.array_equal(np.isnan(got), np.isnan(want)), wname
        assert np.array_equal(np.isneginf(got), np.isneginf(want)) and not np.isposinf(got).any(), wname
        assert rel_err(got, want) <= TOL, (wname, rel_err(got, want))


def test_fixed_grid_with_odd_interval_counts(oracle, wfpt_gpu, variant):
    """integrate.pxi:30-45 does not insist on an even number of intervals: with an odd one the last node carries weight 3."""
    rng = np.random.default_rng(12)
    x = np.where(rng.random(500) < 0.7, 1.0, -1.0) * rng.uniform(0.4, 3.0, 500)
    P = 24
    prm = np.stack([rng.uniform(-3, 3, P), rng.uniform(0, 2, P), rng.uniform(1, 3, P), rng.uniform(0.3, 0.7, P),
                    rng.uniform(0.01, 0.25, P), rng.uniform(0.1, 0.3, P), rng.uniform(0.01, 0.2, P), np.full(P, 0.05)], axis=1)
    prm[::5, 4] = 0.0
    prm[1::5, 6] = 0.0
    for n_st, n_sz in ((1, 1), (3, 3), (5, 2), (2, 5), (1, 4), (7, 3)):
        want = oracle.wiener_like_batch(x, prm, 1e-4, n_st, n_sz, 0, 1e-3, 0.1)
        got = wfpt_gpu.wiener_like_batch(x, prm, 1e-4, n_st, n_sz, 0, 1e-3, 0.1, variant=variant)
        assert rel_err(got, want) <= TOL, (n_st, n_sz)


def test_series_term_census_has_no_flips(grid_golden, wfpt_gpu):
    """The fast kernels decide the expansion and K on a tt formed from a reciprocal; the reference on x'/pow(a,2).  Over
    8,192 benign sets x 10,000 RTs x 3 t-nodes (fixed grid) and 5 t-nodes (adaptive first level): how often the two tt
    differ at all, and how often that changes the table's answer (must be never).  The table itself is then held against
    the reference's formulas (host libm) at tt values of these very inputs."""
    import ctypes as C

    import torch

    from hddm_b200 import _lib

    g = grid_golden
    L = _lib.load()
    dx = torch.from_numpy(g["x"]).cuda()
    prm = np.ascontiguousarray(np.concatenate([g["benign"][:8192], g["bounds"]]))
    dp = torch.from_numpy(prm).cuda()
    for ad in (0, 1):
        counts = torch.zeros(3, dtype=torch.int64, device="cuda")
        _lib.check(L.wfpt_kflip_census(dx.data_ptr(), len(g["x"]), dp.data_ptr(), len(prm), 1e-4, 2, 2, ad, counts.data_ptr(), None))
        torch.cuda.synchronize()
        seen, tt_diff, flips = (int(v) for v in counts.cpu())
        assert seen > 1e8 and flips == 0, (ad, seen, tt_diff, flips)
        assert tt_diff < 1e-3 * seen   # the reciprocal form is correctly rounded nearly always
    import math

    # (the compiled reference forms a^2 as a*a: gcc folds Cython's pow(a, 2.0) into a multiplication, which is what the
    # bit-for-bit agreement of oracle/wfpt_oracle.c with oracle/_ref rests on; libm's pow(a, 2.0) itself is not always a*a)
    kt = _lib.ktable(1e-4)
    thr, desc = np.array(kt["thr"]), np.array([kt["desc0"]] + kt["desc"])
    for q in prm[::257]:
        t_nodes = (q[5] - q[6] / 2) + (q[6] / 2) * np.arange(3)
        xp = (np.abs(g["x"])[:, None] - t_nodes[None, :]).ravel()
        tt = xp[xp > 0] / (q[2] * q[2])
        assert np.array_equal(desc[np.searchsorted(thr, tt, side="right")], _lib.ktable_decision_array(tt, 1e-4))


def test_multi_with_integration_vs_oracle(oracle, wfpt_gpu, variant):
    """wiener_like_multi with per-trial v, a, t (and z) while sz / st are above the 1e-3 thresholds: every trial integrates
    with its own parameter set (what HDDMRegressor asks for, /root/reference/hddm/models/hddm_regression.py:48-61, with the
    reference's depth-10 adaptive defaults, src/wfpt.pyx:357-387), plus fixed grids and the one-axis cases."""
    rng = np.random.default_rng(41)
    n = 1500
    x = np.where(rng.random(n) < 0.75, 1.0, -1.0) * rng.uniform(0.3, 3.5, n)
    x[5], x[9] = 999.0, -999.0          # choice-only trials (wfpt.pyx:370-379)
    v, a, t, z = rng.normal(0.8, 0.5, n), rng.uniform(1.2, 2.8, n), rng.uniform(0.1, 0.4, n), rng.uniform(0.35, 0.65, n)
    cases = [
        (dict(sv=0.6, sz=0.15, st=0.12), ["v", "a", "t"], dict()),                                   # defaults: adaptive, depth 10
        (dict(sv=0.0, sz=0.2, st=0.0), ["v", "a", "t", "z"], dict()),                                # z axis only
        (dict(sv=1.1, sz=0.0, st=0.2), ["v", "t"], dict(n_st=3, n_sz=3, simps_err=1e-5)),            # t axis only
        (dict(sv=0.6, sz=0.15, st=0.12), ["v", "a", "t"], dict(use_adaptive=0, n_st=2, n_sz=2)),      # fixed 3 x 3
        (dict(sv=0.3, sz=0.1, st=0.15), ["a"], dict(use_adaptive=0, n_st=5, n_sz=4)),                 # odd grid
        (dict(sv=0.6, sz=5e-4, st=0.12), ["v", "a", "t"], dict(use_adaptive=0, n_st=4, n_sz=0)),      # sz under the threshold
    ]
    for fixed, multi, kw in cases:
        vals = dict(v=v if "v" in multi else 0.9, a=a if "a" in multi else 2.0, t=t if "t" in multi else 0.25,
                    z=z if "z" in multi else 0.5, **fixed)
        args = (vals["v"], vals["sv"], vals["a"], vals["z"], vals["sz"], vals["t"], vals["st"], 1e-4)
        want = oracle.wiener_like_multi(x, *args, multi, p_outlier=0.05, w_outlier=0.1, **kw)
        got = wfpt_gpu.wiener_like_multi(x, *args, multi, p_outlier=0.05, w_outlier=0.1, **kw)
        assert np.isfinite(want) and abs(got - want) <= TOL * abs(want), (fixed, multi, kw, got, want)


def test_adaptive_two_kernel_lists_overflow_and_shrink(oracle, wfpt_gpu):
    """Adaptive launches keep, per (set, slice) item, a list of the trials whose recursion goes below the first level for
    the refinement kernel.  Here nearly every trial lands in it (all RTs inside the t +- st/2 window), on a grid large
    enough that the library shrinks the per-item regions to stay under its 2 GiB budget: lists overflow and the
    overflow is refined inside the first kernel.  Same answers as the one-kernel general path and as the oracle."""
    import os

    import torch

    rng = np.random.default_rng(77)
    n, P = 2100, 140000
    x = np.where(rng.random(n) < 0.7, 1.0, -1.0) * rng.uniform(0.42, 0.6, n)
    prm = np.stack([rng.uniform(-2, 2, P), rng.uniform(0, 2, P), rng.uniform(0.8, 2.0, P), rng.uniform(0.35, 0.65, P),
                    rng.uniform(0.05, 0.25, P), rng.uniform(0.48, 0.52, P), rng.uniform(0.2, 0.25, P), np.full(P, 0.05)], axis=1)
    dx = torch.from_numpy(wfpt_gpu.sort_rts(x)).cuda()
    dp = torch.from_numpy(prm).cuda()
    kw = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, w_outlier=0.1)
    got = wfpt_gpu.wiener_like_batch(dx, dp, **kw).cpu().numpy()
    os.environ["WFPT_NO_LEAN"] = "1"
    try:
        sub = slice(0, 2048)
        general = wfpt_gpu.wiener_like_batch(dx, dp[sub].contiguous(), **kw).cpu().numpy()
    finally:
        del os.environ["WFPT_NO_LEAN"]
    assert np.all(np.isfinite(got)) and rel_err(got[sub], general) < 1e-12
    idx = rng.choice(P, 24, replace=False)
    want = oracle.wiener_like_batch(x, prm[idx], 1e-4, 2, 2, 1, 1e-3, 0.1)
    assert rel_err(got[idx], want) <= TOL
