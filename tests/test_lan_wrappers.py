"""Host-only network-likelihood wrappers vs the reference Cython module (stub network)."""
import numpy as np
import pytest

pytestmark = pytest.mark.filterwarnings("ignore:numpy.core is deprecated")

from oracle import build_ref


class _Net:
    """Deterministic stand-in for a LAN: any smooth function of the input rows will do."""

    def predict_on_batch(self, data):
        d = np.asarray(data, dtype=np.float32)
        w = np.linspace(-0.7, 0.9, d.shape[1]).astype(np.float32)
        return (np.tanh(d @ w) * 12.0 - 8.0).astype(np.float32)[:, None]


@pytest.fixture(scope="module")
def ref():
    try:
        return build_ref.load_ref()[0]
    except Exception as e:  # reference tree absent (GPU box)
        pytest.skip(f"reference module unavailable: {e}")


@pytest.fixture(scope="module")
def wfpt():
    import wfpt as m
    return m


def _trial_data(n, seed):
    rng = np.random.default_rng(seed)
    rt = rng.uniform(0.3, 3.0, n).astype(np.float32)
    resp = rng.choice([-1.0, 1.0], n).astype(np.float32)
    return rt, resp


@pytest.mark.parametrize("p_outlier,w_outlier", [(0.0, 0.0), (0.05, 0.1)])
def test_mlp_family(ref, wfpt, p_outlier, w_outlier):
    net = _Net()
    rt, resp = _trial_data(257, 1)
    params = np.array([0.8, 1.4, 0.45, 0.3], dtype=np.float32)
    kw = dict(p_outlier=p_outlier, w_outlier=w_outlier, network=net)
    assert wfpt.wiener_like_nn_mlp(rt, resp, params, **kw) == ref.wiener_like_nn_mlp(rt, resp, params, **kw)
    ub = np.array([3, 2.5, 0.9, 2], dtype=np.float32)
    lb = np.array([-3, 0.3, 0.1, 0], dtype=np.float32)
    assert wfpt.wiener_like_nn_mlp_info(rt, resp, params, ub, lb, **kw) == \
        ref.wiener_like_nn_mlp_info(rt, resp, params, ub, lb, **kw)
    bad = params.copy()
    bad[1] = 2.6
    assert wfpt.wiener_like_nn_mlp_info(rt, resp, bad, ub, lb, **kw) == -np.inf
    for logp in (0, 1):
        assert np.array_equal(wfpt.wiener_like_nn_mlp_pdf(rt, resp, params, logp=logp, **kw),
                              ref.wiener_like_nn_mlp_pdf(rt, resp, params, logp=logp, **kw))
    data = np.column_stack([np.tile(params, (rt.size, 1)), rt, resp]).astype(np.float32)
    assert wfpt.wiener_like_multi_nn_mlp(data, **kw) == ref.wiener_like_multi_nn_mlp(data, **kw)
    # the reference's version assigns the array to a C float and raises for more than one row (`wfpt.pyx:743,750`);
    # the replacement returns the per-row values the single-parameter variant returns
    with pytest.raises(TypeError):
        ref.wiener_like_multi_nn_mlp_pdf(data, **kw)
    assert np.array_equal(wfpt.wiener_like_multi_nn_mlp_pdf(data, **kw),
                          wfpt.wiener_like_nn_mlp_pdf(rt, resp, params, logp=1, **kw))
    with pytest.raises(ValueError):
        wfpt.wiener_like_nn_mlp(rt.astype(np.float64), resp, params, **kw)


def _rl_data(n, seed):
    rng = np.random.default_rng(seed)
    x = rng.uniform(0.3, 3.0, n)
    response = rng.integers(0, 2, n).astype(np.int64)
    feedback = rng.integers(0, 2, n).astype(np.float64)
    split_by = rng.integers(0, 3, n).astype(np.int64)
    return x, response, feedback, split_by


@pytest.mark.parametrize("rl", [[0.3], [0.3, -0.4]])
@pytest.mark.parametrize("p_outlier", [0.0, 0.05])
def test_rlssm_nn(ref, wfpt, rl, p_outlier):
    net = _Net()
    x, response, feedback, split_by = _rl_data(200, 2)
    ssm = np.array([1.5, 1.4, 0.5, 0.3])
    bnds = np.array([[-3, 0.3, 0.1, 0.0], [3, 2.5, 0.9, 2.0]])
    args = ("ddm", x, response, feedback, split_by, 0.5, ssm, np.array(rl), bnds)
    kw = dict(p_outlier=p_outlier, w_outlier=0.1, network=net)
    assert wfpt.wiener_like_rlssm_nn(*args, **kw) == ref.wiener_like_rlssm_nn(*args, **kw)
    tight = bnds.copy()
    tight[1, 0] = 0.2   # the per-trial drift leaves the bound somewhere along the recurrence
    args_t = args[:-1] + (tight,)
    assert wfpt.wiener_like_rlssm_nn(*args_t, **kw) == ref.wiener_like_rlssm_nn(*args_t, **kw) == -np.inf
    assert wfpt.wiener_like_rlssm_nn(*args, p_outlier=1.5, network=net) == -np.inf


def test_rlssm_nn_reg(ref, wfpt):
    net = _Net()
    n = 150
    x, response, feedback, split_by = _rl_data(n, 3)
    rng = np.random.default_rng(4)
    data = np.column_stack([rng.uniform(0.5, 2.0, n), np.full(n, 1.4), np.full(n, 0.5), np.full(n, 0.3),
                            x, response]).astype(np.float32)
    rl_arr = rng.normal(0, 1, (n, 1)).astype(np.float32)
    bnds = np.array([[-3, 0.3, 0.1, 0.0], [3, 2.5, 0.9, 2.0]])
    d1, d2 = data.copy(), data.copy()
    a = wfpt.wiener_like_rlssm_nn_reg(d1, rl_arr, x, response, feedback, split_by, 0.5, bnds, p_outlier=0.05,
                                      w_outlier=0.1, network=net)
    b = ref.wiener_like_rlssm_nn_reg(d2, rl_arr, x, response, feedback, split_by, 0.5, bnds, p_outlier=0.05,
                                     w_outlier=0.1, network=net)
    assert a == b and np.array_equal(d1, d2) and not np.array_equal(d1, data)
    neg = data.copy()
    neg[5, 0] = -1.0
    assert wfpt.wiener_like_rlssm_nn_reg(neg, rl_arr, x, response, feedback, split_by, 0.5, bnds, network=net) == -np.inf
