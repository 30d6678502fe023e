"""Sharding of the batched likelihood and of simulator histograms over the GPUs of one box.

One process per GPU (torchrun); `torch.distributed` (NCCL on GPUs, gloo in the CPU tests) is plumbing only.
Parameter sets are independent, so the likelihood needs no data-path collective: rank r evaluates rows r, r+W, ...
(round-robin balances cheap and expensive sets) and the per-set sums are all-gathered -- no cross-rank arithmetic,
hence results are bit-identical for every world size.  Simulator histograms are integer counts: all-reduce(sum).
"""
from __future__ import annotations

import numpy as np


def shard_rows(P: int, rank: int, world: int) -> np.ndarray:
    """Row indices owned by `rank`: rank, rank+world, ...  (requires nothing of P)."""
    return np.arange(rank, P, world, dtype=np.int64)


def unshard_order(P: int, world: int) -> np.ndarray:
    """Permutation that maps the rank-major concatenation of padded shards back to row order.

    Every rank pads its shard to ceil(P/world) rows (all_gather needs equal sizes); position k of rank r holds row
    r + k*world.  Returns idx with full[row] = gathered[idx[row]]."""
    per = -(-P // world)
    rows = np.arange(P, dtype=np.int64)
    return (rows % world) * per + rows // world


def wiener_like_batch_sharded(x, params, group=None, evaluate=None, **opts):
    """All ranks call this with the same (x, params); every rank returns the full float64[P] vector.

    x, params: torch tensors on this rank's device (CUDA for the product path).  `evaluate(x, shard_params, **opts)`
    defaults to hddm_b200.wfpt.wiener_like_batch; the CPU tests inject their own."""
    import torch
    import torch.distributed as dist

    if evaluate is None:
        from . import wfpt

        evaluate = wfpt.wiener_like_batch
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    P = params.shape[0]
    if world == 1:
        return evaluate(x, params, **opts)
    rows = torch.from_numpy(shard_rows(P, rank, world)).to(params.device)
    mine = evaluate(x, params.index_select(0, rows).contiguous(), **opts)
    per = -(-P // world)
    padded = torch.zeros(per, dtype=torch.float64, device=params.device)
    padded[: mine.shape[0]] = mine
    gathered = torch.empty(per * world, dtype=torch.float64, device=params.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    idx = torch.from_numpy(unshard_order(P, world)).to(params.device)
    return gathered.index_select(0, idx)


def wiener_like_batch_trial_sharded(x, params, group=None, evaluate=None, **opts):
    """Large-N / small-P layout: every rank holds ALL parameter sets and evaluates them on its slice of the trials
    (rank r takes x[r::W]); the per-set partial sums are combined with an all-reduce(sum) and a "some rank saw a zero
    density" flag with an all-reduce(max): a zero-density trial on any rank makes the total -inf even when another rank
    produced NaN, as in the reference and in the single-GPU kernel; the summation order
    differs from the single-GPU run by rounding (~1e-16 relative).  All ranks return the full float64[P] vector."""
    import torch.distributed as dist

    if evaluate is None:
        from . import wfpt

        evaluate = wfpt.wiener_like_batch
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1:
        return evaluate(x, params, **opts)
    import torch

    part = evaluate(x[rank::world].contiguous(), params, **opts).clone()
    # wfpt.pyx:71-72: the first zero density returns -inf whatever the other trials give.  A plain sum would turn
    # "-inf on one rank, NaN or +inf on another" into NaN, so the "saw a zero" flag travels with the sums (max over ranks)
    # and wins, as it does inside the single-GPU kernel.
    zero = torch.isneginf(part).to(torch.float64)
    dist.all_reduce(part, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(zero, op=dist.ReduceOp.MAX, group=group)
    return torch.where(zero > 0, torch.full_like(part, float("-inf")), part)


def merge_histograms(counts, group=None):
    """Sum integer histogram counts over ranks in place (particles sharded over ranks) and return them."""
    import torch.distributed as dist

    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts


def simulate_histograms_sharded(theta, model="angle", n_samples=100000, nbins=512, shard="sets", group=None, simulate=None,
                                **kw):
    """Config 5 across GPUs: RT x choice histograms of `theta` (float32 CUDA tensor [n_sets, n_par], the same on every
    rank) with the work split over the ranks of `group` and merged by NCCL on the stream the kernels run on -- no host
    round trip.  shard="sets": rank r simulates sets r, r+W, ... and the per-set histograms are all-gathered;
    shard="particles": every rank simulates n_samples/W particles of every set (global particle offsets) and the counts
    are all-reduced.  Either way the random streams are keyed by GLOBAL set and particle indices, so the merged
    histogram equals the single-GPU one bit for bit.  Returns (hist int32[n_sets, nbins, 2], steps int64[1]) on every rank.
    (Histogram semantics: /root/reference/graveyard/basic_simulator_old.py:70-119.)"""
    import torch
    import torch.distributed as dist

    if simulate is None:
        from . import simulators

        simulate = simulators.simulate_histograms_device
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_sets = theta.shape[0]
    if world == 1:
        return simulate(theta, model=model, n_samples=n_samples, nbins=nbins, **kw)
    if shard == "particles":
        per = -(-int(n_samples) // world)
        lo = min(rank * per, int(n_samples))
        mine = min(per, int(n_samples) - lo)
        hist = torch.zeros((n_sets, int(nbins), 2), dtype=torch.int32, device=theta.device)
        steps = torch.zeros(1, dtype=torch.int64, device=theta.device)
        if mine > 0:
            simulate(theta, model=model, n_samples=mine, nbins=nbins, sample_offset=lo, hist=hist, steps=steps, **kw)
        dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(steps, op=dist.ReduceOp.SUM, group=group)
        return hist, steps
    if shard != "sets":
        raise ValueError("shard must be 'sets' or 'particles'")
    rows = torch.from_numpy(shard_rows(n_sets, rank, world)).to(theta.device)
    per = -(-n_sets // world)
    padded = torch.zeros((per, int(nbins), 2), dtype=torch.int32, device=theta.device)
    steps = torch.zeros(1, dtype=torch.int64, device=theta.device)
    if rows.numel():
        simulate(theta.index_select(0, rows).contiguous(), model=model, n_samples=n_samples, nbins=nbins, set_offset=rank,
                 set_stride=world, hist=padded[: rows.numel()], steps=steps, **kw)
    gathered = torch.empty((per * world, int(nbins), 2), dtype=torch.int32, device=theta.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    dist.all_reduce(steps, op=dist.ReduceOp.SUM, group=group)
    idx = torch.from_numpy(unshard_order(n_sets, world)).to(theta.device)
    return gathered.index_select(0, idx), steps
