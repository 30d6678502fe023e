"""Digests of the ONE-GPU config-5 histograms (1,024 angle parameter sets, seed 5, 512 bins): run on a GPU box,

    python tests/golden/make_cfg5_sha.py > gpurun_out/cfg5_hist_sha.json     # then copy to tests/golden/

bench.py --gpus N compares the histogram it merged over N GPUs (NCCL all-gather / all-reduce) with these: counter-based
random streams keyed by global set and particle indices make the merged counts independent of the sharding, bit for bit."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch  # noqa: E402

import bench  # noqa: E402
from hddm_b200 import simulators  # noqa: E402

theta = torch.from_numpy(bench.cfg5_theta()).cuda()
out = {}
for n in (100_000, 10_000_000):
    h, _ = simulators.simulate_histograms_device(theta, model="angle", n_samples=n, nbins=512, random_state=5, max_t=20.0, delta_t=1e-3)
    torch.cuda.synchronize()
    out[str(n)] = bench.sha(h.cpu().numpy())
print(json.dumps(out))
