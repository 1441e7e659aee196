#!/usr/bin/env python
"""Headline benchmark: images/sec of the batched depth-inference forward at 1024 px.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--model standard|enhanced|config4|train] [--batch B] [--size S]
    python bench.py --impl reference ...      # the CPU arm (oracle port of the reference on host cores), same config

Default (BASELINE.json configs[1]): standard model, batch 32 per GPU, 3x1024x1024 fp32 images, bf16 compute.  A step = one
forward of one synthetic batch.  Other workloads of BASELINE.json, each with its own `metric` name:
    --model enhanced            configs[2]  (batch 32/GPU; `--batch 4` at N=8 is the strong-scaled global-32 split of SURVEY §8d)
    --model config4             configs[3]  enhanced + MultiScaleDepthFusion + TextureAwareRefinementModule (batch 16/GPU; --batch 2 = 16 global at N=8)
    --model train               configs[4]  fine-tuning step (forward + backward + gradient all-reduce + Adam), batch 8/GPU at 518 px
For N > 1 the driver launches one rank per GPU with torchrun; every rank runs its own replica on its own batch (images are
independent: no collective on the forward path; the training step all-reduces gradients over NCCL), scaling is
weak and `value` is N * B * K / t_max.  One JSON line is printed by rank 0 (keys: see `base_line` and the legs below it).
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

GFLOP_PER_IMAGE = {"standard": 38.43, "enhanced": 39.11, "config4": 392.1}  # SURVEY.md §8(d), algorithmic (2*MAC), per image
ENCODER_GFLOP_PER_IMAGE = 35.12       # ViT linears + attention + patch embedding: the "encoder GEMM/attention" figure of north_star
ATTN_FLOP_PER_IMAGE_LAYER = 4.0 * 197 * 197 * 64 * 12  # QK^T + PV, 12 heads (0.119 GFLOP)
DEFAULT_BATCH = {"standard": 32, "enhanced": 32, "config4": 16, "train": 8}
DEFAULT_SIZE = {"standard": 1024, "enhanced": 1024, "config4": 1024, "train": 518}
METRIC = {"standard": "images_per_sec", "enhanced": "images_per_sec", "config4": "config4_images_per_sec", "train": "train_images_per_sec"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="standard", choices=["standard", "enhanced", "config4", "train"])
    ap.add_argument("--batch", type=int, default=None, help="images per GPU per step (default: the BASELINE.json config's)")
    ap.add_argument("--size", type=int, default=None)
    ap.add_argument("--no-extras", action="store_true", help="skip the e2e / roofline / sustained / cpu_baseline / enhanced legs")
    ap.add_argument("--sustain-seconds", type=float, default=3.0, help="length of the steady-state companion run of the same step")
    a = ap.parse_args()
    a.batch = a.batch or DEFAULT_BATCH[a.model]
    a.size = a.size or DEFAULT_SIZE[a.model]
    return a


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"], "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


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
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                pw.append(float(r[2]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "power_w_max": max(pw) if pw else None, "samples": len(sm)}


def workload_text(a):
    B, S = a.batch, a.size
    return {"standard": f"standard model forward, batch {B}/GPU, 3x{S}x{S} fp32 in -> 4 fp32 maps out (BASELINE.json configs[1])",
            "enhanced": f"enhanced model forward, batch {B}/GPU, 3x{S}x{S} fp32 in -> 10 fp32 maps out (BASELINE.json configs[2])",
            "config4": f"enhanced + MultiScaleDepthFusion + TextureAwareRefinementModule, batch {B}/GPU, 3x{S}x{S} (BASELINE.json configs[3])",
            "train": f"fine-tuning step (forward + backward + gradient all-reduce + Adam), batch {B}/GPU, 3x{S}x{S} (BASELINE.json configs[4])"}[a.model]


# ------------------------------------------------------------------------------------ reference arm (CPU)
def _cpu_models(kind):
    import pkgload

    pkgload.load()
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.synth import synth_state_dict

    mk = lambda cls, seed: synth_state_dict(cls().state_dict(), seed)
    if kind == "config4":
        enh = synth_state_dict(mde.EnhancedDepthAnything(pretrained=False).state_dict(), 0)
        return enh, mk(mde.MultiScaleDepthFusion, 100), mk(mde.TextureAwareRefinementModule, 200)
    cls = mde.EnhancedDepthAnything if kind == "enhanced" else mde.DepthAnything
    return (synth_state_dict(cls(pretrained=False).state_dict(), 0),)


def cpu_rate(kind, size, batch, steps, warmup, threads=None, budget_s=240.0):
    """(images/sec, ms/step, threads, timed steps) of the oracle port — the reference's own ATen CPU arithmetic — on the host cores.
    Timed steps stop early once `budget_s` of timed work is spent (the count actually run is returned)."""
    sds = _cpu_models(kind)
    import torch

    from monodepth_estimation_b200.synth import bench_batch, synth_images
    from oracle import depth_oracle as O

    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    md = {"source": "rover"}
    if kind == "train":
        from oracle import train_oracle as T

        x = synth_images(batch, size, size, 77)
        rnd = torch.rand(batch, 1, size, size, generator=torch.Generator().manual_seed(5))

        def step():
            _, grads, _ = T.loss_and_grads(sds[0], x, md, rnd)
            T.adam_reference({k: sds[0][k].clone() for k in grads}, grads, {}, lambda k: 1e-5 if k.startswith("encoder.") else 1e-4, 1)
    else:
        x = bench_batch(batch, size, 0)
        fwd = {"standard": O.standard_forward, "enhanced": O.enhanced_forward}.get(kind)

        def step():
            with torch.no_grad():
                if kind == "config4":
                    O.config4_forward(sds[0], sds[1], sds[2], x, md)
                else:
                    fwd(sds[0], x, md)
    ts = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        step()
        if i >= warmup:
            ts.append(time.perf_counter() - t0)
            if sum(ts) > budget_s:
                break
    total = sum(ts)
    return batch * len(ts) / total, total / len(ts) * 1e3, threads, len(ts)


def cpu_config1(threads=None):
    """BASELINE.json configs[0] exactly: standard model, batch 1, 3x518x518, fp32 on the host cores (3 warm-up, median of 10)."""
    sd = _cpu_models("standard")[0]
    import torch

    from monodepth_estimation_b200.synth import synth_images
    from oracle import depth_oracle as O

    torch.set_num_threads(threads or os.cpu_count() or 1)
    x = synth_images(1, 518, 518, 12)
    ts = []
    with torch.no_grad():
        for i in range(13):
            t0 = time.perf_counter()
            O.standard_forward(sd, x, {"source": "rover"})
            if i >= 3:
                ts.append(time.perf_counter() - t0)
    med = statistics.median(ts)
    return {"value": round(1.0 / med, 3), "unit": "images/s", "ms_median": round(med * 1e3, 2), "cores": torch.get_num_threads(),
            "config": "standard model, batch 1, 3x518x518 fp32 (BASELINE.json configs[0]), median of 10 after 3 warm-up"}


def run_reference(a):
    """The CPU arm on THIS arm's config: same model, batch, size, steps and warm-up (a step of batch 32 at 1024 px is ~1 s of host
    work); only a wall-clock budget may cut the timed steps short, and the line says how many ran."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = a.batch
    note = ""
    if a.model == "config4" and batch > 2:  # ~10 s of host work per image: a bounded sample of the batch
        batch, note = 2, f" (bounded sample: batch 2 of the batch-{a.batch} workload; the refinement chain costs ~10 s/image on host cores)"
    v, ms, threads, done = cpu_rate(a.model, a.size, batch, max(1, a.steps), max(0, a.warmup))
    sample = f"{a.model}, batch {batch} x 3x{a.size}x{a.size} fp32, {done} timed steps after {a.warmup} warm-up{note}"
    line = {"impl": "reference", "metric": METRIC[a.model], "value": round(v, 3), "unit": "images/s", "n_gpus": a.gpus, "steps": done, "warmup": a.warmup,
            "ms_per_step": round(ms, 2), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": workload_text(a), "device": "host CPU", "batch_timed": batch},
            "cpu_baseline": {"value": round(v, 3), "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": round(v, 3), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if a.model == "standard":
        line["cpu_baseline"]["config1"] = cpu_config1(threads)
    emit(line)


# ------------------------------------------------------------------------------------ B200 arm
class Env:
    """rank / device / process group of one bench process."""

    def __init__(self):
        import torch

        self.torch = torch
        self.rank, self.world, self.local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        from monodepth_estimation_b200.sharding import bind_host_to_gpu

        # pinned staging buffers on the GPU's own NUMA node where the platform exposes one (MDE_NUMA_BIND=0 leaves placement to the OS)
        self.numa = bind_host_to_gpu(self.local) if os.environ.get("MDE_NUMA_BIND", "1") != "0" else None
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_seconds(self, local_images, seconds):
        from monodepth_estimation_b200.sharding import whole_job_rate

        return whole_job_rate(local_images, seconds, self.dist, self.dev)[1]

    def timed(self, step, steps, warmup, images_per_step):
        """`warmup` untimed + exactly `steps` timed calls of step(), CUDA events on the launch stream, barrier + synchronize on both
        sides, max over ranks.  Returns seconds."""
        torch = self.torch
        stream = torch.cuda.current_stream()
        for _ in range(warmup):
            step()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        self.barrier()
        return self.max_seconds(images_per_step * steps, e0.elapsed_time(e1) / 1e3)

    def sustained(self, step, seconds, est_step_s, images_per_step):
        """The same step run back to back for >= `seconds` (power-capped steady state), with its own clock record."""
        n = max(3, int(seconds / max(est_step_s, 1e-5)) + 1)
        sampler = ClockSampler(self.local) if self.rank == 0 else None
        t = self.timed(step, n, 1, images_per_step)
        clocks = sampler.stop() if sampler else None
        return n, t, clocks


def b2b_seconds(torch, fn, reps=20, warm=3):
    """Average duration of fn() launched `reps` times back to back between two CUDA events (synchronize on both sides)."""
    stream = torch.cuda.current_stream()
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


_REAL_STDOUT = None


def _quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner under torchrun): from here on
    file descriptor 1 points at stderr, and emit() writes the line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def main():
    a = parse()
    _quiet_stdout()
    if a.impl == "reference":
        return run_reference(a)

    import pkgload

    pkgload.load()
    env = Env()
    line = bench_train(a, env) if a.model == "train" else bench_forward(a, env)
    if env.rank == 0:
        emit(line)
    if env.dist is not None:
        env.dist.destroy_process_group()


def base_line(a, env, value, seconds, steps, clocks, launches, extra_config=None):
    cfg = {"workload": workload_text(a), "global_batch": a.batch * env.world, "image_size": a.size,
           "parallelism": f"image-sharded x{env.world}" + (", gradient all-reduce over NCCL" if a.model == "train" else ", no collective"),
           "l2": "inputs and outputs of a step exceed the 126 MB L2 (12.6 MB in / >= 16.8 MB out per image at 1024 px)" if a.model != "train"
                 else "activations + weights + Adam state of a step (> 2 GB) exceed the 126 MB L2",
           "weights": "synthetic random-init, seed 0", "host_numa_node": None if env.numa is None else env.numa[0]}
    cfg.update(extra_config or {})
    return {"metric": METRIC[a.model], "value": round(value, 1), "unit": "images/s", "n_gpus": env.world, "steps": steps, "warmup": a.warmup,
            "ms_per_step": round(seconds / steps * 1e3, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": launches}


def bench_forward(a, env):
    torch = env.torch
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200 import ops
    from monodepth_estimation_b200.synth import bench_batch, synth_state_dict

    dev, rank, world = env.dev, env.rank, env.world
    B, S, md = a.batch, a.size, {"source": "rover"}
    pk = peaks()

    def make(kind):
        def load(mod, seed):
            mod.load_state_dict(synth_state_dict(mod.state_dict(), seed))
            return mod.eval().to(dev)

        if kind == "config4":
            return mde.EnhancedWithRefinement(load(mde.EnhancedDepthAnything(pretrained=False), 0), load(mde.MultiScaleDepthFusion(), 100),
                                              load(mde.TextureAwareRefinementModule(), 200))
        return load((mde.DepthAnything if kind == "standard" else mde.EnhancedDepthAnything)(pretrained=False), 0)

    def launches_of(model):
        if a.model == "config4":  # + depth_fusion + the refinement chain (pack, 5 convs, output kernel)
            return model.model._get_engine().launches_per_forward + 8
        return model._get_engine().launches_per_forward

    # one image set per rank (seeded by rank); a batch is far larger than the 126 MB L2, so no step is served from cache
    x_host = bench_batch(B, S, rank).pin_memory()
    x_dev = x_host.to(dev, non_blocking=True)
    out_key = ("disp_refined", 0) if a.model == "config4" else ("disp", 0)

    model = make(a.model)
    step = lambda: model(x_dev, md)
    sampler = ClockSampler(env.local) if rank == 0 else None
    t = env.timed(step, a.steps, a.warmup, B)
    clocks = sampler.stop() if sampler else None
    value = world * B * a.steps / t
    line = base_line(a, env, value, t, a.steps, clocks, launches_of(model) * a.steps)
    gf = GFLOP_PER_IMAGE[a.model]
    line["step_tflops"] = round(value / world * gf / 1e3, 1)
    line["tensor_frac_of_step"] = round(value / world * gf / 1e3 / pk["bf16_tflops"], 4)
    line["tensor_frac_of_step_note"] = "algorithmic GFLOP of the whole step / time / BURST bf16 peak (the timed region is a short burst; see `sustained`)"
    if a.no_extras:
        return line

    # ---- steady state: the same step for >= 3 s (the headline above is a ~50 ms burst), own clock record, sustained peak
    n_s, t_s, clocks_s = env.sustained(step, a.sustain_seconds, t / a.steps, B)
    v_s = world * B * n_s / t_s
    line["sustained"] = {"value": round(v_s, 1), "unit": "images/s", "steps": n_s, "seconds": round(t_s, 3), "ms_per_step": round(t_s / n_s * 1e3, 4),
                         "clocks": clocks_s, "tensor_frac_of_step": round(v_s / world * gf / 1e3 / pk["bf16_tflops_sustained"], 4),
                         "peak": pk["bf16_tflops_sustained"], "peak_source": pk["source"] + ", sustained cuBLAS figure"}

    # ---- e2e through the public batching API (mde.BatchPipeline: H2D(i+1), forward(i), D2H(i-1) on three streams), host buffers
    # pinned, copies inside the timed region every step.  Declared feed: what the reference's caller holds after decoding an image
    # (uint8 HWC, run_mars_depth.py:94-100; float()/255 + HWC->CHW run on the device), result read back as fp16.
    def timed_pipeline(host_batch, out_dtype):
        pipe = mde.BatchPipeline(model, keys=(out_key,), depth=3, out_dtype=out_dtype)
        pipe.run([host_batch] * a.warmup, md)
        env.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(torch.cuda.current_stream())
        pipe.run([host_batch] * a.steps, md)  # returns once the last result is in host memory
        e1.record(torch.cuda.current_stream())
        env.barrier()
        return env.max_seconds(B * a.steps, e0.elapsed_time(e1) / 1e3)

    x_u8 = (x_host.permute(0, 2, 3, 1) * 255.0).round().to(torch.uint8).contiguous().pin_memory()
    te = timed_pipeline(x_u8, torch.float16)
    line["e2e"] = {"value": round(world * B * a.steps / te, 1), "unit": "images/s", "h2d_bytes_per_step": x_u8.numel(), "d2h_bytes_per_step": B * S * S * 2,
                   "ms_per_step": round(te / a.steps * 1e3, 4),
                   "how": f"mde.BatchPipeline: uint8 [B,H,W,3] pinned host batches in (the decoder's output, run_mars_depth.py:94-100), {out_key} read back "
                          "as fp16 into pinned host memory; H2D(i+1), forward(i), D2H(i-1) overlap on three streams.  The pipeline calls "
                          "model.forward_keys: the uint8 image goes straight into the input kernel (pixel = float(u8)/255 there, bit-identical "
                          "to converting first), the whole encoder + decoder runs as in `value`, and only the requested map is resized to full "
                          "resolution, written as fp16 by the output kernel (config 4: the composed callable takes the fp32 image and returns every map)"}
    tf = timed_pipeline(x_host, torch.float32)
    line["e2e_f32"] = {"value": round(world * B * a.steps / tf, 1), "unit": "images/s", "h2d_bytes_per_step": x_host.numel() * 4,
                       "d2h_bytes_per_step": B * S * S * 4, "ms_per_step": round(tf / a.steps * 1e3, 4),
                       "how": "same pipeline fed float32 [B,3,H,W] and read back as float32: byte for byte what the reference's callers move"}
    if a.model != "config4":
        out_host = torch.empty(B, 1, S, S, dtype=torch.float32).pin_memory()

        def serial():
            out = model(x_host.to(dev, non_blocking=True), md)
            out_host.copy_(out[out_key], non_blocking=True)

        ns = max(2, a.steps // 2)
        ts = env.timed(serial, ns, a.warmup, B)
        line["e2e_serial"] = {"value": round(world * B * ns / ts, 1), "unit": "images/s", "ms_per_step": round(ts / ns * 1e3, 4),
                              "how": "float32 copies issued serially on one stream around model(x), as the reference callers do"}
    hl = hostlink_reference(world)
    if hl is not None:
        bytes_per_image = (x_u8.numel() + B * S * S * 2) / B
        line["e2e"]["host_link"] = {"aggregate_gbs": hl, "bound_images_per_s": round(hl * 1e9 / bytes_per_image, 1),
                                    "frac": round(line["e2e"]["value"] * bytes_per_image / (hl * 1e9), 4),
                                    "source": "profiles/hostlink_r02.json (plain pinned cudaMemcpyAsync H2D+D2H from N ranks at once)"}
    del x_u8

    # ---- rooflines (rank 0): the dominant tensor-core GEMM class, attention, and the two HBM-bound ends of the step
    if rank == 0 and a.model != "config4":
        line["roofline"] = roofline_leg(model, x_dev, md, ops, pk, B)
    elif rank == 0:
        line["roofline"] = {"bound": "tensor", "kernel": "conv_halo_kernel (refinement.py chain, fp16 operands)", "achieved": line["step_tflops"],
                            "peak": pk["bf16_tflops"], "unit": "TFLOP/s", "frac": line["tensor_frac_of_step"], "traffic": None,
                            "how": "whole-step algorithmic FLOPs / step time (352.9 of the 392.1 GFLOP/image are the chain's 3x3 convs)"}

    # ---- the other half of the headline metric: enhanced model (BASELINE.json configs[2]) on the same inputs
    if a.model == "standard":
        del model
        torch.cuda.empty_cache()
        enh = make("enhanced")
        n2 = max(2, a.steps // 2)
        t2 = env.timed(lambda: enh(x_dev, md), n2, a.warmup, B)
        line["enhanced"] = {"value": round(world * B * n2 / t2, 1), "unit": "images/s", "ms_per_step": round(t2 / n2 * 1e3, 4),
                            "config": "enhanced model, same batch/size (BASELINE.json configs[2])"}
        del enh

    # ---- CPU baseline (rank 0, N=1 only): the oracle port of the reference on the host cores, bounded sample of the same workload
    if rank == 0 and world == 1:
        if env.numa is not None:
            os.sched_setaffinity(0, env.numa[1])  # the CPU leg uses every host core again
        cb = min(B, 2 if a.model == "config4" else 8)
        v, ms, threads, done = cpu_rate(a.model, S, cb, 3, 1, budget_s=60.0)
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "images/s", "cores": threads, "kind": "port",
                                "sample": f"{a.model} forward, batch {cb} x 3x{S}x{S} fp32, {done} timed steps after 1 warm-up ({ms:.0f} ms/step)"}
        if a.model == "standard":
            line["cpu_baseline"]["config1"] = cpu_config1(threads)
    return line


def hostlink_reference(world):
    """Aggregate H2D+D2H GB/s of plain pinned copies from `world` ranks at once (tools/hostlink_bench.py), if on file."""
    p = os.path.join(ROOT, "profiles", "hostlink_r02.json")
    if not os.path.exists(p):
        return None
    try:
        for e in json.load(open(p))["runs"]:
            if e["ranks"] == world:
                return e["aggregate_bidir_gbs"]
    except Exception:
        pass
    return None


def _traffic_file():
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    return json.load(open(p)) if os.path.exists(p) else {}


def ncu_traffic(M, N, K):
    """DRAM bytes per launch of a GEMM class from the committed `ncu --set full` capture (profiles/ncu_traffic.json, written by
    tools/ncu_summary.py --traffic); None when no capture of that shape is on file.  Not measured live: ncu cannot run inside a bench."""
    for e in _traffic_file().get("gemm", []):
        if (e["M"], e["N"], e["K"]) == (M, N, K):
            return e["dram_bytes"]
    return None


def ncu_traffic_named(name):
    return _traffic_file().get("kernels", {}).get(name)


def roofline_leg(model, x_dev, md, ops, pk, B):
    """Roofline of the tensor-core GEMM class with the largest total time, plus attention and the HBM-bound input / output kernels.

    Pass 1 (which class, share of the step): one eager forward with a CUDA-event pair around every tc-GEMM launch on the launching
    stream.  An event pair around ONE 20-40 us launch also brackets the launch gap that programmatic dependent launch and the CUDA
    graph remove in the real step, so these per-launch times are upper bounds; they are reported as `eager_event_us`.
    Pass 2 (`achieved`): every class re-launched 20x back to back with its recorded operands between two events on the same
    stream, synchronize on both sides — 20 launches are < 1 ms at full clocks with no power cap, i.e. a kernel timed in isolation,
    so every fraction here is against the BURST peaks of MEASURED_PEAKS.json (bf16_tflops / hbm_gbs)."""
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
    reps = 20
    b2b = {key: b2b_seconds(torch, lambda c=call: orig(*c), reps) for key, call in args.items()}
    fl = lambda k: 2.0 * k[0] * k[1] * k[2]
    key = max(agg, key=lambda k: b2b[k] * len(agg[k]))
    M, N, K, epi = key
    total_b2b = sum(b2b[k] * len(v) for k, v in agg.items())
    total_eager = sum(sum(v) for v in agg.values())
    ach = fl(key) / b2b[key] / 1e12
    allf = sum(fl(k) * len(v) for k, v in agg.items())
    vit = [k for k in agg if k[0] == M and k[3] in (1, 2, 5)]  # the four per-layer ViT linears (QKV, proj, fc1, fc2)
    peak = pk["bf16_tflops"]
    out = {"bound": "tensor", "kernel": f"ViT linear GEMM (tcgen05 cta_group::2) M={M} N={N} K={K} epilogue={epi}",
           "achieved": round(ach, 1), "peak": peak, "unit": "TFLOP/s", "frac": round(ach / peak, 4),
           "peak_source": f"{pk['source']}, BURST cuBLAS bf16 figure: the kernel is timed alone, {reps} back-to-back launches (< 1 ms)",
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
                             "frac": round(fl(k) / b2b[k] / 1e12 / peak, 4)}
                            for k, v in sorted(agg.items(), key=lambda kv: -b2b[kv[0]] * len(kv[1]))]}

    # ---- attention (tensor roofline, tiny FLOPs: latency-bound) and the two HBM-bound ends of the step, same back-to-back timing
    ws = eng.workspace(x_dev.device, x_dev.shape[0], x_dev.shape[2], x_dev.shape[3])
    w = eng.weights()
    others = []
    t_attn = b2b_seconds(torch, lambda: ops.vit_attention(ws.q, ws.k, ws.vt, B, ws.attn), reps)
    fa = ATTN_FLOP_PER_IMAGE_LAYER * B
    others.append({"kernel": "vit_attention (S = QK^T, softmax, PV fused; 12 launches/step)", "bound": "tensor", "achieved": round(fa / t_attn / 1e12, 1),
                   "peak": peak, "unit": "TFLOP/s", "frac": round(fa / t_attn / 1e12 / peak, 4), "avg_us": round(t_attn * 1e6, 2),
                   "algorithmic_flops_per_launch": fa, "traffic": ncu_traffic_named("vit_attention"),
                   "note": "0.3 % of the step's FLOPs; bounded by TMA latency + the exponentials (MUFU), not by the tensor pipe"})
    enc_flops = ENCODER_GFLOP_PER_IMAGE * 1e9 * B
    enc_t = sum(b2b[k] * len(agg[k]) for k in agg if k[0] in (M, B * 196) and k[1] in (768, 2304, 3072)) + 12 * t_attn
    out["encoder_gemm_attention"] = {"tflops": round(enc_flops / enc_t / 1e12, 1), "frac_of_burst": round(enc_flops / enc_t / 1e12 / peak, 4),
                                     "frac_of_sustained": round(enc_flops / enc_t / 1e12 / pk["bf16_tflops_sustained"], 4),
                                     "ms_per_step": round(enc_t * 1e3, 4), "target": "north_star: >= 0.60 of the bf16 tensor-core peak"}
    t_in = b2b_seconds(torch, lambda: eng._pre(ws, w, x_dev), reps)
    bytes_in = x_dev.numel() * 4 + B * 3 * 224 * 224 * 4 + B * 196 * 768 * 2
    others.append({"kernel": "input stage: input_scan_resample_kernel + range_finalize + patchify224 (one pass over the input)", "bound": "hbm",
                   "achieved": round(bytes_in / t_in / 1e9, 1), "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(bytes_in / t_in / 1e9 / pk["hbm_gbs"], 4),
                   "avg_us": round(t_in * 1e6, 2), "algorithmic_bytes_per_launch": bytes_in, "traffic": ncu_traffic_named("input_stage"),
                   "note": "12.58 MB/image fp32 read once (SURVEY §8d) + 0.9 MB/image of 224 px raster and patch rows written"})
    t_out = b2b_seconds(torch, lambda: eng._post(ws, 1.0), reps)
    n_maps = 10 if eng.enhanced else 4
    bytes_out = n_maps * B * x_dev.shape[2] * x_dev.shape[3] * 4
    others.append({"kernel": "output stage: resize_maps kernel (scale, bilinear resize to the input size, disp->depth; every map in one launch)", "bound": "hbm",
                   "achieved": round(bytes_out / t_out / 1e9, 1), "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": round(bytes_out / t_out / 1e9 / pk["hbm_gbs"], 4),
                   "avg_us": round(t_out * 1e6, 2), "algorithmic_bytes_per_launch": bytes_out, "traffic": ncu_traffic_named("resize_maps"),
                   "note": f"{n_maps} fp32 maps x 4.19 MB/image written once (SURVEY §8d); reads are <= 0.3 MB/image"})
    out["others"] = others
    return out


# ------------------------------------------------------------------------------------ training step (configs[4])
def bench_train(a, env):
    torch = env.torch
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.sharding import allreduce_mean_grads
    from monodepth_estimation_b200.synth import synth_images, synth_state_dict
    from monodepth_estimation_b200.training import DepthTrainer

    dev, rank, world = env.dev, env.rank, env.world
    B, S, md = a.batch, a.size, {"source": "rover"}
    m = mde.DepthAnything(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), 0))
    m = m.to(dev)
    tr = DepthTrainer(m, lr=1e-4)
    x_host = synth_images(B, S, S, 77 + rank).pin_memory()
    rnd_host = torch.rand(B, 1, S, S, generator=torch.Generator().manual_seed(5 + rank)).pin_memory()
    x, rnd = x_host.to(dev), rnd_host.to(dev)
    hook = (lambda g: allreduce_mean_grads(g, env.dist)) if env.dist is not None else None
    step = lambda: tr.step(x, md, rnd, hook)
    sampler = ClockSampler(env.local) if rank == 0 else None
    t = env.timed(step, a.steps, max(a.warmup, 3), B)  # step 1 eager, step 2 captures the graphs, step 3+ replay
    clocks = sampler.stop() if sampler else None
    value = world * B * a.steps / t
    line = base_line(a, env, value, t, a.steps, clocks, getattr(tr, "launches_per_step", 0) * a.steps,
                     {"collective": "fp32 gradient all-reduce over NCCL (one grouped in-place launch per backward stage, overlapped with the next stage)"
                      if world > 1 else "none (one rank)", "launch": "CUDA graphs per backward stage + one multi-tensor Adam launch"})
    line["step_tflops"] = round(value / world * 3 * 38.43 / 1e3, 1)
    if a.no_extras:
        return line
    loss_host = torch.empty(3, dtype=torch.float32).pin_memory()

    def e2e_step():
        xd, rd = x_host.to(dev, non_blocking=True), rnd_host.to(dev, non_blocking=True)
        loss_host.copy_(tr.step(xd, md, rd, hook), non_blocking=True)

    te = env.timed(e2e_step, a.steps, 2, B)
    line["e2e"] = {"value": round(world * B * a.steps / te, 1), "unit": "images/s", "h2d_bytes_per_step": (x_host.numel() + rnd_host.numel()) * 4,
                   "d2h_bytes_per_step": 12, "ms_per_step": round(te / a.steps * 1e3, 4),
                   "how": "DepthTrainer.step on a pinned host batch (images + the random target factors) copied in every step, the three losses read back"}
    n_s, t_s, clocks_s = env.sustained(step, a.sustain_seconds, t / a.steps, B)
    line["sustained"] = {"value": round(world * B * n_s / t_s, 1), "unit": "images/s", "steps": n_s, "seconds": round(t_s, 3),
                         "ms_per_step": round(t_s / n_s * 1e3, 4), "clocks": clocks_s}
    pk = peaks()
    line["roofline"] = {"bound": "tensor", "kernel": "whole training step (forward + backward GEMMs on the tcgen05 GEMM kernels)", "achieved": line["step_tflops"],
                        "peak": pk["bf16_tflops"], "unit": "TFLOP/s", "frac": round(line["step_tflops"] / pk["bf16_tflops"], 4), "traffic": None,
                        "how": "3 x 38.43 GFLOP/image (SURVEY §8 config table) / step time, against the burst bf16 peak"}
    if rank == 0 and world == 1:
        v, ms, threads, done = cpu_rate("train", S, B, 1, 1, budget_s=60.0)
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "images/s", "cores": threads, "kind": "port",
                                "sample": f"{done} step(s) of the same batch after 1 warm-up (oracle/train_oracle.py: fp32 autograd + Adam), {ms:.0f} ms/step"}
    return line


if __name__ == "__main__":
    main()
