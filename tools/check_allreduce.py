#!/usr/bin/env python
"""Two-or-more-GPU check of sharding.allreduce_mean_grads over NCCL (both paths: grouped in-place launch and flat buckets):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_allreduce.py

Every rank fills its gradients with a rank-dependent pattern; afterwards every rank must hold the mean over ranks.  The CPU
suite covers the same logic over gloo (tests/test_multirank_cpu.py); this is the NCCL leg, which needs more than one GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import pkgload

pkgload.load()
from monodepth_estimation_b200.sharding import allreduce_mean_grads

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
shapes = {"w.big": (3072, 768), "w.bias": (768,), "pos": (1, 197, 768), "scalar3": (3,), "conv": (128, 256, 3, 3)}
for mode in ("coalesced", "bucket"):
    os.environ["MDE_ALLREDUCE"] = mode
    base = {k: torch.randn(*s, generator=torch.Generator().manual_seed(len(k))).to(dev) for k, s in shapes.items()}
    grads = {k: v * (rank + 1) for k, v in base.items()}
    n = allreduce_mean_grads(grads, dist, bucket_bytes=4 << 20)
    torch.cuda.synchronize()
    factor = sum(r + 1 for r in range(world)) / world
    worst = max(float((grads[k] - base[k] * factor).abs().max()) for k in grads)
    assert worst < 1e-5, (mode, worst)
    if rank == 0:
        print(f"allreduce_mean_grads[{mode}] over {world} ranks: {n} collective launch(es), max abs error {worst:.2e}", flush=True)
dist.destroy_process_group()
