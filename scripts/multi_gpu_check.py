"""Multi-GPU correctness check (run under torchrun, NCCL): (1) parameter-set sharding + all-gather reproduces the
single-GPU log-likelihoods bit for bit; (2) trial sharding + all-reduce(sum) agrees to rounding; (3) simulator particles
sharded over ranks (global particle offsets) + NCCL all-reduce of the histograms == the unsharded histogram, bit for bit.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/multi_gpu_check.py"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from hddm_b200 import wfpt, simulators, distributed as D
from bench import workload

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", local)
x, prm = workload(4096)
dx, dp = torch.from_numpy(wfpt.sort_rts(x)).to(dev), torch.from_numpy(prm).to(dev)
kw = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=0, simps_err=1e-3, w_outlier=0.1)
single = wfpt.wiener_like_batch(dx, dp, **kw)
sharded = D.wiener_like_batch_sharded(dx, dp, **kw)
trial = D.wiener_like_batch_trial_sharded(dx, dp[:256].contiguous(), **kw)
torch.cuda.synchronize()
res = {"world": world, "param_shard_bit_identical": bool(torch.equal(single, sharded)),
       "trial_shard_max_rel_err": float(((trial - single[:256]).abs() / single[:256].abs()).max())}
theta = np.array([[1.0, 1.0, 0.5, 0.3, 0.2], [-0.5, 1.5, 0.6, 0.2, 0.1]], dtype=np.float32)
n = 200000
per = n // world
h_part, _ = simulators.simulate_histograms(theta, model="angle", n_samples=per, nbins=256, random_state=11, sample_offset=rank * per)
merged = D.merge_histograms(torch.from_numpy(h_part.astype(np.int64)).to(dev))
h_full, _ = simulators.simulate_histograms(theta, model="angle", n_samples=per * world, nbins=256, random_state=11)
res["histogram_allreduce_bit_identical"] = bool(np.array_equal(merged.cpu().numpy(), h_full.astype(np.int64)))
if rank == 0:
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
