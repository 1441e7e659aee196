"""CPU, world_size 2, gloo: the host-side multi-rank logic (batch sharding and the whole-job throughput reduction that
bench.py uses under torchrun).  The data path itself has no collective."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pkgload  # noqa: E402  (spawned workers re-import this module without conftest.py)

pkgload.load()

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from monodepth_estimation_b200.sharding import shard_bounds, whole_job_rate


def test_shard_bounds_partition_every_batch():
    for B in (0, 1, 7, 16, 32, 33):
        for world in (1, 2, 3, 8):
            cuts = [shard_bounds(B, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_bounds(33, world, rank)
    x = torch.arange(33 * 4, dtype=torch.float32).view(33, 4)[lo:hi]          # this rank's slice of the global batch
    secs = 0.5 + 0.25 * rank                                                    # rank 1 is the slow one
    rate, tmax, total = whole_job_rate(x.shape[0], secs, dist)
    got = [None] * world
    dist.all_gather_object(got, (lo, hi, float(x.sum())))
    if rank == 0:
        out.put((rate, tmax, total, got))
    dist.destroy_process_group()


def test_world2_gloo_rate_is_total_images_over_slowest_rank():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    rate, tmax, total, got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert total == 33 and tmax == 0.75 and abs(rate - 33 / 0.75) < 1e-9
    assert [g[:2] for g in got] == [(0, 17), (17, 33)]
    full = torch.arange(33 * 4, dtype=torch.float32).view(33, 4)
    assert abs(sum(g[2] for g in got) - float(full.sum())) < 1e-3  # the two slices cover the batch exactly once


def test_numa_binding_helpers():
    """Host side of one-process-per-GPU: cpulist parsing, and no CUDA device / no sysfs entry means 'leave placement alone'."""
    import os

    from monodepth_estimation_b200 import sharding as S

    assert S._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert S._parse_cpulist("") == set()
    before = os.sched_getaffinity(0)
    assert S.gpu_numa_cpus(0, sysfs="/nonexistent") is None
    if not __import__("torch").cuda.is_available():
        assert S.bind_host_to_gpu(0) is None
        assert os.sched_getaffinity(0) == before


def _allreduce_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from monodepth_estimation_b200.sharding import allreduce_mean_grads

    g = torch.Generator().manual_seed(100 + rank)
    grads = {"b.weight": torch.randn(300, 7, generator=g), "a.bias": torch.randn(5, generator=g), "c": torch.randn(1000, generator=g)}
    calls = allreduce_mean_grads(grads, dist, bucket_bytes=4096)  # small buckets: several collectives
    q.put((rank, calls, {k: v.clone() for k, v in grads.items()}))
    dist.destroy_process_group()


def test_gradient_allreduce_averages_over_ranks():
    """world_size 2, gloo: the bucketed gradient all-reduce of the training step leaves the mean of the ranks' gradients everywhere."""
    world, port = 2, 29631
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_allreduce_worker, args=(r, world, port, q)) for r in range(world)]
    for p_ in procs:
        p_.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p_ in procs:
        p_.join(60)
    want = {}
    for k, shape in (("b.weight", (300, 7)), ("a.bias", (5,)), ("c", (1000,))):
        per_rank = []
        for r in range(world):
            g = torch.Generator().manual_seed(100 + r)
            d = {"b.weight": torch.randn(300, 7, generator=g), "a.bias": torch.randn(5, generator=g), "c": torch.randn(1000, generator=g)}
            per_rank.append(d[k])
        want[k] = sum(per_rank) / world
    for rank, calls, got in res:
        assert calls >= 2
        for k in want:
            assert torch.allclose(got[k], want[k], atol=1e-6), (rank, k)
