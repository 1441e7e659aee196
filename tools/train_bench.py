#!/usr/bin/env python
"""BASELINE.json configs[4]: the fine-tuning step (forward + backward + gradient all-reduce + Adam), batch 8 per GPU at 518 px.

    python tools/train_bench.py [--steps K] [--warmup W] [--batch 8] [--size 518] [--cpu-steps N]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/train_bench.py ...

One process per GPU, each with its own replica and its own batch; the only collective is the bucketed gradient all-reduce
(NCCL) between backward and Adam.  Timed with CUDA events on the launch stream, barrier + synchronize on both sides, max over
ranks.  Rank 0 prints one JSON line; with one rank it also times the CPU oracle of the same step (the reference's own
arithmetic, oracle/train_oracle.py) on the host cores as the baseline.  Not the headline metric: bench.py is.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import pkgload

pkgload.load()
import monodepth_estimation_b200 as mde
from monodepth_estimation_b200.sharding import allreduce_mean_grads
from monodepth_estimation_b200.synth import synth_images, synth_state_dict
from monodepth_estimation_b200.training import DepthTrainer


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--size", type=int, default=518)
    ap.add_argument("--cpu-steps", type=int, default=1)
    ap.add_argument("--no-graph", action="store_true", help="eager launches every step (for a per-kernel launch list under ncu)")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    B, S, md = a.batch, a.size, {"source": "rover"}
    m = mde.DepthAnything(pretrained=False)
    sd0 = synth_state_dict(m.state_dict(), 0)
    m.load_state_dict(sd0)
    m = m.to(dev)
    tr = DepthTrainer(m, lr=1e-4, graph=not a.no_graph)
    x = synth_images(B, S, S, 77 + rank).to(dev)
    rnd = torch.rand(B, 1, S, S, generator=torch.Generator().manual_seed(5 + rank)).to(dev)
    hook = (lambda g: allreduce_mean_grads(g, dist)) if dist is not None else None

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        losses = tr.step(x, md, rnd, hook)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        losses = tr.step(x, md, rnd, hook)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t = float(t.item())
    line = {"metric": "train_images_per_sec", "value": round(world * B * a.steps / t, 1), "unit": "images/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(t / a.steps * 1e3, 3), "higher_is_better": True, "scaling": "weak", "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"fine-tuning step (forward + backward + gradient all-reduce + Adam), batch {B}/GPU, 3x{S}x{S} (BASELINE.json configs[4])",
                       "global_batch": B * world, "collective": "fp32 gradient all-reduce over NCCL (one grouped in-place launch per backward stage, overlapped with the next stage)" if world > 1 else "none (one rank)",
                       "launch": "eager" if a.no_graph else "two CUDA graphs per step (forward+backward, Adam) around the all-reduce"},
            "losses_last_step": [round(float(v), 6) for v in losses.cpu()]}
    if rank == 0 and world == 1 and a.cpu_steps > 0:
        from oracle import train_oracle as T

        torch.set_num_threads(os.cpu_count() or 1)
        xc, rc = x.cpu(), rnd.cpu()
        ts = []
        for i in range(a.cpu_steps + 1):
            t0 = time.perf_counter()
            _, grads, _ = T.loss_and_grads(sd0, xc, md, rc)
            T.adam_reference({k: sd0[k] for k in grads}, grads, {}, lambda k: 1e-5 if k.startswith("encoder.") else 1e-4, 1)
            if i > 0:
                ts.append(time.perf_counter() - t0)
        line["cpu_baseline"] = {"value": round(B * len(ts) / sum(ts), 3), "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
                                "sample": f"{len(ts)} step(s) of the same batch after 1 warm-up (oracle/train_oracle.py: fp32 autograd + Adam)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
