#!/usr/bin/env python
"""Headline benchmark: images/sec of the batched depth-inference forward at 1024 px.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--model standard|enhanced] [--batch B] [--size S]
    python bench.py --impl reference ...      # the CPU arm (oracle port of the reference on host cores)

A step = one forward of one synthetic batch (BASELINE.json configs[1]: standard model, batch 32,
3x1024x1024 fp32 images, bf16 compute on one B200).  For N > 1 the driver launches one rank per GPU
with torchrun; every rank runs its own replica on its own batch (images are independent: no
collective on the data path), so scaling is weak and `value` is N * B * K / t_max.

One JSON line is printed by rank 0 (see the keys at the bottom).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GFLOP_PER_IMAGE = {"standard": 38.43, "enhanced": 39.11}  # SURVEY.md §8(d), algorithmic


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="standard", choices=["standard", "enhanced"])
    ap.add_argument("--batch", type=int, default=32, help="images per GPU per step")
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--no-extras", action="store_true", help="skip the e2e / roofline / cpu_baseline / enhanced legs")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"], "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------ reference arm (CPU)
def cpu_forward_rate(kind, size, batch, steps, warmup, threads=None):
    """images/sec of the oracle port (same ATen CPU ops as the reference classes) on the host cores."""
    import torch

    import pkgload

    pkgload.load()
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.synth import synth_images, synth_state_dict
    from oracle import depth_oracle as O

    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    cls = mde.DepthAnything if kind == "standard" else mde.EnhancedDepthAnything
    m = cls(pretrained=False)
    sd = synth_state_dict(m.state_dict(), 0)
    x = synth_images(batch, size, size, 1234)
    fwd = O.standard_forward if kind == "standard" else O.enhanced_forward
    ts = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            fwd(sd, x, {"source": "rover"})
            if i >= warmup:
                ts.append(time.perf_counter() - t0)
    total = sum(ts)
    return batch * len(ts) / total, total / len(ts) * 1e3, threads


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = 2
    steps, warmup = max(1, min(a.steps, 8)), max(1, min(a.warmup, 2))
    v, ms, threads = cpu_forward_rate(a.model, a.size, batch, steps, warmup)
    sample = f"{a.model} forward, batch {batch} x 3x{a.size}x{a.size} fp32, {steps} timed steps after {warmup} warm-up (bounded sample of the batch-{a.batch} workload)"
    line = {"impl": "reference", "metric": "images_per_sec", "value": round(v, 3), "unit": "images/s", "n_gpus": a.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": round(ms, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": f"{a.model} model forward, batch {a.batch}/GPU, 3x{a.size}x{a.size} (BASELINE.json configs[1])", "device": "host CPU"},
            "cpu_baseline": {"value": round(v, 3), "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": round(v, 3), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------ B200 arm
def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)

    import torch

    import pkgload

    pkgload.load()
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200 import ops
    from monodepth_estimation_b200.sharding import whole_job_rate
    from monodepth_estimation_b200.synth import bench_batch, synth_images, synth_state_dict

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # host side of one-process-per-GPU: pinned staging buffers on the GPU's own NUMA node (MDE_NUMA_BIND=0 leaves placement to the OS)
    from monodepth_estimation_b200.sharding import bind_host_to_gpu
    numa = bind_host_to_gpu(local) if os.environ.get("MDE_NUMA_BIND", "1") != "0" else None
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    B, S, md = a.batch, a.size, {"source": "rover"}

    def make(kind):
        cls = mde.DepthAnything if kind == "standard" else mde.EnhancedDepthAnything
        m = cls(pretrained=False)
        m.load_state_dict(synth_state_dict(m.state_dict(), 0))
        return m.eval().to(dev)

    # one image set per rank (seeded by rank); a batch is 403 MB, i.e. > the 126 MB L2, so no step can be served from cache
    x_host = bench_batch(B, S, rank).pin_memory()
    x_dev = x_host.to(dev, non_blocking=True)

    def timed(model, steps, warmup, feed_host):
        """returns (seconds for `steps` steps, max over ranks), CUDA-event timed on the launch stream."""
        out_host = torch.empty(B, 1, S, S, dtype=torch.float32).pin_memory() if feed_host else None
        xd = x_dev
        stream = torch.cuda.current_stream()

        def step():
            nonlocal xd
            if feed_host:
                xd = x_host.to(dev, non_blocking=True)
            out = model(xd, md)
            if feed_host:
                out_host.copy_(out[("disp", 0)], non_blocking=True)  # what the caller reads back (run_mars_depth.py:145)
            return out

        for _ in range(warmup):
            step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        barrier()
        return whole_job_rate(B * steps, e0.elapsed_time(e1) / 1e3, dist, dev)[1]  # max seconds over ranks

    model = make(a.model)
    sampler = ClockSampler(local) if rank == 0 else None
    t = timed(model, a.steps, a.warmup, feed_host=False)
    clocks = sampler.stop() if sampler else None
    launches = model._get_engine().launches_per_forward * a.steps
    value = world * B * a.steps / t
    pk = peaks()

    line = {"metric": "images_per_sec", "value": round(value, 1), "unit": "images/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": round(t / a.steps * 1e3, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": f"{a.model} model forward, batch {B}/GPU, 3x{S}x{S} fp32 in -> {4 if a.model == 'standard' else 10} fp32 maps out "
                                   f"(BASELINE.json configs[{1 if a.model == 'standard' else 2}])",
                       "global_batch": B * world, "image_size": S, "parallelism": f"image-sharded x{world}, no collective",
                       "l2": "inputs (403 MB/batch) and outputs (>=537 MB/batch) exceed the 126 MB L2", "weights": "synthetic random-init, seed 0",
                       "host_numa_node": None if numa is None else numa[0]},
            "clocks": clocks, "gpu_launches": launches,
            "tensor_frac_of_step": round(value / world * GFLOP_PER_IMAGE[a.model] / 1e3 / pk["bf16_tflops_sustained"], 4)}

    if not a.no_extras:
        # ---- e2e: pinned host input -> H2D -> forward -> D2H of ("disp",0), every step inside the timed region, through the
        # package's public batching API (BatchPipeline: copy-in / compute / copy-out on three streams, ring of 3 buffers)
        def timed_pipeline(host_batch, steps, warmup):
            pipe = mde.BatchPipeline(model, keys=(("disp", 0),), depth=3)
            pipe.run([host_batch] * warmup, md)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(torch.cuda.current_stream())
            pipe.run([host_batch] * steps, md)  # returns once the last result is in host memory
            e1.record(torch.cuda.current_stream())
            barrier()
            return whole_job_rate(B * steps, e0.elapsed_time(e1) / 1e3, dist, dev)[1]

        te = timed_pipeline(x_host, a.steps, a.warmup)
        line["e2e"] = {"value": round(world * B * a.steps / te, 1), "unit": "images/s", "h2d_bytes_per_step": x_host.numel() * 4,
                       "d2h_bytes_per_step": B * S * S * 4, "ms_per_step": round(te / a.steps * 1e3, 4),
                       "how": "mde.BatchPipeline: float32 [B,3,H,W] pinned host batches in, (disp,0) float32 to pinned host out; H2D(i+1), forward(i), D2H(i-1) overlap"}
        ts = timed(model, max(2, a.steps // 2), a.warmup, feed_host=True)
        line["e2e_serial"] = {"value": round(world * B * max(2, a.steps // 2) / ts, 1), "unit": "images/s", "ms_per_step": round(ts / max(2, a.steps // 2) * 1e3, 4),
                              "how": "same copies issued serially on one stream around model(x), as the reference callers do"}
        x_u8 = (x_host.permute(0, 2, 3, 1) * 255.0).round().to(torch.uint8).contiguous().pin_memory()
        tu = timed_pipeline(x_u8, a.steps, a.warmup)
        line["e2e_u8"] = {"value": round(world * B * a.steps / tu, 1), "unit": "images/s", "h2d_bytes_per_step": x_u8.numel(),
                          "d2h_bytes_per_step": B * S * S * 4, "ms_per_step": round(tu / a.steps * 1e3, 4),
                          "how": "mde.BatchPipeline fed uint8 [B,H,W,3] (decoder output); float()/255 + HWC->CHW run on the device"}
        del x_u8

        # ---- roofline of the dominant kernel: per-launch CUDA-event timing of every tensor-core GEMM launch (eager, same stream)
        if rank == 0:
            line["roofline"] = roofline_leg(model, x_dev, md, ops, pk)

        # ---- the other half of the headline metric: enhanced model (BASELINE.json configs[2]) on the same inputs
        if a.model == "standard":
            del model
            torch.cuda.empty_cache()
            enh = make("enhanced")
            t2 = timed(enh, max(2, a.steps // 2), a.warmup, feed_host=False)
            line["enhanced"] = {"value": round(world * B * max(2, a.steps // 2) / t2, 1), "unit": "images/s", "ms_per_step": round(t2 / max(2, a.steps // 2) * 1e3, 4),
                                "config": "enhanced model, same batch/size (BASELINE.json configs[2])"}
            del enh

        # ---- CPU baseline (rank 0, N=1 only): the oracle port of the reference on the host cores, bounded sample
        if rank == 0 and world == 1:
            if numa is not None:
                os.sched_setaffinity(0, numa[1])  # the CPU leg uses every host core again
            v, ms, threads = cpu_forward_rate(a.model, S, 2, 3, 1)
            line["cpu_baseline"] = {"value": round(v, 3), "unit": "images/s", "cores": threads, "kind": "port",
                                    "sample": f"{a.model} forward, batch 2 x 3x{S}x{S} fp32, 3 timed steps after 1 warm-up"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def ncu_traffic(M, N, K):
    """DRAM bytes per launch of a GEMM class from the committed `ncu --set full` capture (profiles/ncu_traffic.json, written by
    tools/ncu_summary.py --traffic); None when no capture of that shape is on file.  Not measured live: ncu cannot run inside a bench."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    for e in json.load(open(p)).get("gemm", []):
        if (e["M"], e["N"], e["K"]) == (M, N, K):
            return e["dram_bytes"]
    return None


def roofline_leg(model, x_dev, md, ops, pk):
    """Roofline of the tensor-core GEMM class with the largest total time.

    Pass 1 (which class, share of the step): one eager forward with a CUDA-event pair around every tc-GEMM launch on the launching
    stream.  An event pair around ONE 20-40 us launch also brackets the launch gap that programmatic dependent launch and the CUDA
    graph remove in the real step, so these per-launch times are upper bounds; they are reported as `eager_event_us`.
    Pass 2 (`achieved`): every class re-launched 20x back to back with its recorded operands between two events on the same
    stream, synchronize on both sides — the kernel's own duration including its prologue/epilogue, without the event gaps."""
    import torch

    eng = model._get_engine()
    records = []
    orig = ops.gemm

    def timed_gemm(a, w, bias, epilogue, out, aux0=None, aux1=None, M=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(torch.cuda.current_stream())
        r = orig(a, w, bias, epilogue, out, aux0, aux1, M)
        e1.record(torch.cuda.current_stream())
        records.append(((a.shape[0] if M is None else M, w.shape[0], a.shape[1], epilogue), e0, e1, (a, w, bias, epilogue, out, aux0, aux1, M)))
        return r

    was = eng.use_graph
    eng.use_graph = False
    ops.gemm = timed_gemm
    try:
        for _ in range(3):
            records.clear()
            # park the GPU for ~4 ms so the host (ctypes call + tensor-map encodes + event records, ~15 us per launch) gets
            # ahead: every event pair then brackets device time only, not a host gap in front of the launch
            torch.cuda._sleep(int(4e-3 * 1.9e9))
            model(x_dev, md)
        torch.cuda.synchronize()
    finally:
        ops.gemm = orig
        eng.use_graph = was
    agg, args = {}, {}
    for key, e0, e1, call in records:
        agg.setdefault(key, []).append(e0.elapsed_time(e1) * 1e-3)
        args.setdefault(key, call)
    # pass 2: back-to-back launches of each class (operands are the live workspace buffers of the last forward)
    reps, b2b = 20, {}
    stream = torch.cuda.current_stream()
    for key, call in args.items():
        for _ in range(3):
            orig(*call)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            orig(*call)
        e1.record(stream)
        torch.cuda.synchronize()
        b2b[key] = e0.elapsed_time(e1) * 1e-3 / reps
    fl = lambda k: 2.0 * k[0] * k[1] * k[2]
    key = max(agg, key=lambda k: b2b[k] * len(agg[k]))
    M, N, K, epi = key
    total_b2b = sum(b2b[k] * len(v) for k, v in agg.items())
    total_eager = sum(sum(v) for v in agg.values())
    ach = fl(key) / b2b[key] / 1e12
    allf = sum(fl(k) * len(v) for k, v in agg.items())
    vit = [k for k in agg if k[0] == M and k[3] in (1, 2, 5)]  # the four per-layer ViT linears (QKV, proj, fc1, fc2)
    return {"bound": "tensor", "kernel": f"tc_gemm2_kernel (cta_group::2, 256x256x64 tiles) M={M} N={N} K={K} epilogue={epi}",
            "achieved": round(ach, 1), "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s", "frac": round(ach / pk["bf16_tflops_sustained"], 4),
            "peak_source": f"{pk['source']} (sustained cuBLAS bf16: kernel timed inside a long step)",
            "how": f"{reps} back-to-back launches with the step's own operands between two CUDA events on the launch stream; "
                   "eager_event_us = one event pair per launch inside an eager step (upper bound: includes the launch gap)",
            "traffic": ncu_traffic(M, N, K), "traffic_source": "profiles/ncu_traffic.json (ncu --set full, dram__bytes_read+write per launch)",
            "algorithmic_flops_per_launch": fl(key), "launches_per_step": len(agg[key]), "avg_us": round(b2b[key] * 1e6, 2),
            "eager_event_us": round(sum(agg[key]) / len(agg[key]) * 1e6, 2),
            "share_of_gemm_time": round(b2b[key] * len(agg[key]) / total_b2b, 4),
            "all_gemms_tflops": round(allf / total_b2b / 1e12, 1), "all_gemms_ms_per_step": round(total_b2b * 1e3, 4),
            "all_gemms_ms_per_step_eager_events": round(total_eager * 1e3, 4),
            "vit_linear_tflops": round(sum(fl(k) * len(agg[k]) for k in vit) / sum(b2b[k] * len(agg[k]) for k in vit) / 1e12, 1) if vit else None,
            "gemm_classes": [{"M": k[0], "N": k[1], "K": k[2], "epi": k[3], "n": len(v), "avg_us": round(b2b[k] * 1e6, 2),
                              "eager_event_us": round(sum(v) / len(v) * 1e6, 2), "tflops": round(fl(k) / b2b[k] / 1e12, 1),
                              "frac": round(fl(k) / b2b[k] / 1e12 / pk["bf16_tflops_sustained"], 4)}
                             for k, v in sorted(agg.items(), key=lambda kv: -b2b[kv[0]] * len(kv[1]))]}


if __name__ == "__main__":
    main()
