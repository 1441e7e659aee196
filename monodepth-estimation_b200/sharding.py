"""Image-batch sharding across ranks: images are independent, so a global batch is cut into contiguous
per-rank slices and no collective touches the data path.  The only communication is the timing reduction of the
benchmark (max over ranks) — `torch.distributed` with NCCL on the GPU box, gloo in the CPU tests."""
from __future__ import annotations


def shard_bounds(global_batch: int, world: int, rank: int) -> tuple[int, int]:
    """[lo, hi) of rank's contiguous slice; the first `global_batch % world` ranks get one extra image."""
    if world < 1 or not 0 <= rank < world or global_batch < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(global_batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def whole_job_rate(local_images: int, local_seconds: float, dist=None, device=None) -> tuple[float, float, int]:
    """(images/s of the whole job, max seconds over ranks, total images): every rank's image count is summed and divided by
    the SLOWEST rank's time, as the bench contract asks.  `dist` is torch.distributed (initialised) or None for one rank."""
    import torch

    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_images / local_seconds, local_seconds, local_images
    t = torch.tensor([local_seconds], dtype=torch.float64, device=device)
    n = torch.tensor([local_images], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return float(n.item()) / float(t.item()), float(t.item()), int(n.item())
