"""world_size-2 and -3 gloo tests (CPU) of the multi-GPU host logic: round-robin sharding of parameter sets,
padded all-gather, un-permutation, histogram merge.  The evaluator is injected (the C oracle) because there is no
GPU here; on the box the same function runs with the CUDA evaluator under NCCL (bench.py --gpus N)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, P, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from hddm_b200 import distributed as D
    from oracle import cpu as oracle

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    rng = np.random.default_rng(7)
    x = np.where(rng.random(300) < 0.8, 1.0, -1.0) * rng.uniform(0.4, 3.0, 300)
    prm = np.stack([rng.uniform(-3, 3, P), rng.uniform(0, 2, P), rng.uniform(1, 3, P), rng.uniform(0.3, 0.7, P),
                    rng.uniform(0, 0.2, P), rng.uniform(0.1, 0.3, P), rng.uniform(0, 0.2, P), np.full(P, 0.05)], axis=1)
    calls = []

    def evaluate(xt, pt, **opts):
        calls.append(pt.shape[0])
        return torch.from_numpy(oracle.wiener_like_batch(xt.numpy(), pt.numpy(), 1e-4, 2, 2, 0, 1e-3, 0.1))

    full = D.wiener_like_batch_sharded(torch.from_numpy(x), torch.from_numpy(prm), evaluate=evaluate)
    want = oracle.wiener_like_batch(x, prm, 1e-4, 2, 2, 0, 1e-3, 0.1)
    # trial-axis sharding: all-reduce of partial sums; a zero-density trial on one rank makes the total -inf
    prm2 = prm[:5].copy()
    prm2[4, 5] = 0.9    # t above some |rt| -> density 0 for those trials; with p_outlier = 0 the sum is -inf
    prm2[4, 6] = 0.0
    prm2[4, 7] = 0.0
    tot = D.wiener_like_batch_trial_sharded(torch.from_numpy(x), torch.from_numpy(prm2), evaluate=evaluate)
    want2 = oracle.wiener_like_batch(x, prm2, 1e-4, 2, 2, 0, 1e-3, 0.1)
    ok_trial = bool(np.isneginf(tot.numpy()[4]) and np.isneginf(want2[4]) and
                    np.allclose(tot.numpy()[:4], want2[:4], rtol=1e-12, atol=0))
    # -inf on one rank meeting NaN / +inf on another: the zero-density flag wins (wfpt.pyx:71-72), a plain sum would say NaN
    def fake(xt, pt, **opts):
        return torch.tensor([float("-inf") if rank == 0 else float("nan"), float("-inf") if rank == 0 else float("inf"),
                             float("nan") if rank == 0 else 1.0, 2.0], dtype=torch.float64)

    mixed = D.wiener_like_batch_trial_sharded(torch.from_numpy(x), torch.from_numpy(prm[:4]), evaluate=fake).numpy()
    ok_trial = ok_trial and bool(np.isneginf(mixed[0]) and np.isneginf(mixed[1]) and np.isnan(mixed[2]) and mixed[3] == 2.0 * world)
    # simulator histograms sharded by sets and by particles, merged by collectives: equal to the unsharded counts.  The
    # injected stand-in counts, per global (set, particle), into a bin derived from those indices alone (what the device
    # kernel's counter-based streams guarantee).
    def fake_sim(theta, model="angle", n_samples=0, nbins=8, sample_offset=0, set_offset=0, set_stride=1, hist=None, steps=None):
        n_sets = theta.shape[0]
        hist = torch.zeros((n_sets, nbins, 2), dtype=torch.int32) if hist is None else hist
        steps = torch.zeros(1, dtype=torch.int64) if steps is None else steps
        for k in range(n_sets):
            gs = set_offset + k * set_stride
            assert float(theta[k, 0]) == float(gs)          # the shard really holds the rows it claims
            for p in range(sample_offset, sample_offset + n_samples):
                h = (gs * 7919 + p * 104729) % (2 * nbins)
                hist[k, h // 2, h % 2] += 1
                steps += gs + p
        return hist, steps

    n_sets_h = 7
    theta = torch.arange(n_sets_h, dtype=torch.float32).reshape(-1, 1).repeat(1, 5)
    want_h, want_s = fake_sim(theta, n_samples=23)
    ok_hist = True
    for mode in ("sets", "particles"):
        got_h, got_s = D.simulate_histograms_sharded(theta, n_samples=23, nbins=8, shard=mode, simulate=fake_sim)
        ok_hist = ok_hist and bool(torch.equal(got_h, want_h)) and int(got_s) == int(want_s)
    ok_trial = ok_trial and ok_hist
    hist = torch.full((4, 8, 2), rank + 1, dtype=torch.int64)
    D.merge_histograms(hist)
    ok = bool(np.array_equal(full.numpy(), want)) and calls[0] == len(D.shard_rows(P, rank, world)) and \
        bool((hist == sum(range(1, world + 1))).all()) and ok_trial
    q.put((rank, ok))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,P", [(2, 64), (2, 37), (3, 10)])
def test_sharded_batch_matches_single_process(world, P):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, P, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]


def test_shard_helpers():
    from hddm_b200 import distributed as D

    for P, W in [(1, 1), (8, 2), (37, 2), (10, 3), (65536, 8), (5, 8)]:
        rows = [D.shard_rows(P, r, W) for r in range(W)]
        assert sorted(np.concatenate(rows).tolist()) == list(range(P))
        per = -(-P // W)
        gathered = np.full(per * W, -1, dtype=np.int64)
        for r in range(W):
            gathered[r * per: r * per + len(rows[r])] = rows[r]
        assert np.array_equal(gathered[D.unshard_order(P, W)], np.arange(P))
