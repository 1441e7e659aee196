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


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_cpus(device_index: int, sysfs: str = "/sys") -> tuple[int, set[int]] | None:
    """(NUMA node, its CPUs) of the PCIe slot of CUDA device `device_index`, or None when the platform does not say
    (single-node box, container without sysfs, numa_node = -1)."""
    import torch

    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"{sysfs}/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = _parse_cpulist(open(f"{sysfs}/devices/system/node/node{node}/cpulist").read())
        return (node, cpus) if cpus else None
    except Exception:
        return None


def bind_host_to_gpu(device_index: int):
    """One process per GPU: run this rank's host thread on the CPUs of its GPU's NUMA node, so that the pinned staging buffers
    it allocates afterwards (first touch) sit behind the same PCIe root as the GPU — with 8 ranks copying 400 MB per step each,
    cross-socket traffic is what the end-to-end rate is bound by.  Returns (node, previous affinity) or None if nothing was
    changed; undo with ``os.sched_setaffinity(0, previous)``."""
    import os

    info = gpu_numa_cpus(device_index)
    if info is None or not hasattr(os, "sched_setaffinity"):
        return None
    node, cpus = info
    prev = os.sched_getaffinity(0)
    use = cpus & prev
    if not use or use == prev:
        return None
    os.sched_setaffinity(0, use)
    return node, prev


def allreduce_mean_grads(grads: dict, dist=None, bucket_bytes: int = 64 << 20):
    """Data-parallel gradient averaging for the training step (BASELINE config 5; no counterpart in the reference, whose loop is
    single-device: train/mars_adapter.py:138-140).  Gradients are packed in a fixed (sorted-key) order into flat fp32 buckets of
    about `bucket_bytes`, each bucket is all-reduced once (NCCL over NVLink on the GPU box, gloo in the CPU tests) and divided by
    the world size, then copied back.  Few large collectives instead of ~170 small ones: on NVSwitch the cost is launch latency,
    not link bandwidth.  On NCCL the default is instead ONE grouped launch over the gradient tensors in place (no packing copies;
    ``MDE_ALLREDUCE=bucket`` selects the bucket path).  Every rank must hold the same keys.  In place; returns the number of
    collectives issued."""
    import torch

    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return 0
    world = dist.get_world_size()
    keys = sorted(grads)
    import os

    if (dist.get_backend() == "nccl" and os.environ.get("MDE_ALLREDUCE", "coalesced") == "coalesced"
            and all(g.is_cuda and g.is_contiguous() and g.dtype == torch.float32 for g in grads.values())):
        # NCCL: one grouped launch over the gradients where they lie (ncclGroupStart/End around per-tensor all-reduces, averaging
        # inside the collective): no packing copies, nothing but the collective competes with the backward for HBM
        try:
            from torch.distributed.distributed_c10d import _coalescing_manager

            with _coalescing_manager(device=grads[keys[0]].device):
                for k in keys:
                    dist.all_reduce(grads[k], op=dist.ReduceOp.AVG)
            return 1
        except (ImportError, TypeError):
            pass
    calls, i = 0, 0
    while i < len(keys):
        j, size = i, 0
        while j < len(keys) and (j == i or (size + grads[keys[j]].numel()) * 4 <= bucket_bytes):
            size += grads[keys[j]].numel()
            j += 1
        group = [grads[k] for k in keys[i:j]]
        flat = torch.cat([g.reshape(-1).to(torch.float32) for g in group])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(world)
        off = 0
        for g in group:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        calls += 1
        i = j
    return calls
