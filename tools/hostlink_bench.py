#!/usr/bin/env python
"""What can the host link of this box do?  Plain pinned-memory copies from N processes (one per GPU) at once.

    python tools/hostlink_bench.py [--ranks 1,2,4,8] [--mb 256] [--iters 20] [--out gpurun_out/hostlink.json]

For every N in --ranks (capped at the visible GPU count) N worker processes each own one GPU, a pinned host buffer and a device
buffer of --mb MiB, and run three legs — H2D only, D2H only, both directions at once on two streams — of --iters copies each
(ONE cudaMemcpyAsync per copy).  All ranks start a leg together (process barrier), each leg is timed on the device with CUDA
events on the copy streams, and the aggregate is N * bytes / max-over-ranks time.  `aggregate_bidir_gbs` is the roofline that
bench.py's `e2e.host_link.frac` divides by: the end-to-end path cannot move its bytes faster than this, whatever the kernels do.
Also reports the NUMA node each GPU sits behind (sysfs) so that a flat curve can be read against the topology.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(rank, n, mb, iters, barrier, result, bind):
    import torch

    torch.cuda.set_device(rank)
    node = None
    if bind:
        import pkgload

        pkgload.load()
        from monodepth_estimation_b200.sharding import bind_host_to_gpu

        b = bind_host_to_gpu(rank)
        node = None if b is None else b[0]
    nbytes = mb << 20
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_in.fill_(rank + 1)  # first touch on this process's CPUs
    h_out.zero_()
    d_in = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_out = torch.full((nbytes,), 7, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    out = {}
    for leg in ("h2d", "d2h", "bidir"):
        for timed in (False, True):
            reps = iters if timed else 2
            torch.cuda.synchronize()
            barrier.wait()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record(s1)
            e[2].record(s2)
            for _ in range(reps):
                if leg in ("h2d", "bidir"):
                    with torch.cuda.stream(s1):
                        d_in.copy_(h_in, non_blocking=True)
                if leg in ("d2h", "bidir"):
                    with torch.cuda.stream(s2):
                        h_out.copy_(d_out, non_blocking=True)
            e[1].record(s1)
            e[3].record(s2)
            torch.cuda.synchronize()
            if timed:
                t1, t2 = e[0].elapsed_time(e[1]) * 1e-3, e[2].elapsed_time(e[3]) * 1e-3
                out[leg] = {"h2d": t1, "d2h": t2}[leg] if leg != "bidir" else max(t1, t2)
    assert int(d_in[12345]) == rank + 1 and int(h_out[54321]) == 7
    result[rank] = (out, node)
    barrier.wait()


def numa_of_gpus(n):
    import torch

    nodes = []
    for i in range(n):
        try:
            p = torch.cuda.get_device_properties(i)
            bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
            nodes.append(int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read()))
        except Exception:
            nodes.append(None)
    return nodes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", default="1,2,4,8")
    ap.add_argument("--mb", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-bind", action="store_true", help="leave CPU placement of the ranks to the OS")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "hostlink.json"))
    a = ap.parse_args()
    import torch
    import torch.multiprocessing as mp

    ngpu = torch.cuda.device_count()
    ctx = mp.get_context("spawn")
    runs = []
    for n in [int(v) for v in a.ranks.split(",") if int(v) <= ngpu]:
        barrier = ctx.Barrier(n)
        result = ctx.Manager().dict()
        procs = [ctx.Process(target=worker, args=(r, n, a.mb, a.iters, barrier, result, not a.no_bind)) for r in range(n)]
        for p in procs:
            p.start()
        for p in procs:
            p.join()
        if len(result) != n:
            print(f"N={n}: {n - len(result)} worker(s) failed", file=sys.stderr)
            continue
        nbytes = (a.mb << 20) * a.iters
        agg = lambda leg, mult: round(mult * n * nbytes / max(result[r][0][leg] for r in range(n)) / 1e9, 2)
        run = {"ranks": n, "h2d_gbs": agg("h2d", 1), "d2h_gbs": agg("d2h", 1), "aggregate_bidir_gbs": agg("bidir", 2),
               "per_rank_bidir_gbs": [round(2 * nbytes / result[r][0]["bidir"] / 1e9, 2) for r in range(n)],
               "host_numa_node_of_rank": [result[r][1] for r in range(n)]}
        print(json.dumps(run), flush=True)
        runs.append(run)
    doc = {"what": "plain pinned cudaMemcpyAsync (one per copy), N processes = N GPUs at once; GB/s = N * bytes / max-over-ranks CUDA-event time",
           "mb_per_copy": a.mb, "copies_per_leg": a.iters, "gpus_visible": ngpu, "gpu_numa_nodes": numa_of_gpus(ngpu),
           "host_cpus": os.cpu_count(), "bound_to_gpu_numa_node": not a.no_bind, "runs": runs}
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(doc, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
