#!/usr/bin/env python
"""Marginal cost of each launch class inside the real CUDA-graph step:  python tools/ablate.py [B] [model]

A kernel's duration under programmatic dependent launch overlaps its neighbours, so neither the ncu launch list
(serialised, cold-cache) nor a CUPTI trace says what a kernel COSTS the step.  This tool re-captures the graph with one
launch class removed (outputs are then garbage — timing only, nothing here is a bench value) and prints the step time
saved.  Guidance for where to spend optimisation effort.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import pkgload

pkgload.load()
import monodepth_estimation_b200 as mde
from monodepth_estimation_b200 import _lib as L
from monodepth_estimation_b200 import ops
from monodepth_estimation_b200.synth import synth_images, synth_state_dict

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
KIND = sys.argv[2] if len(sys.argv) > 2 else "standard"
S = 1024
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
cls = mde.DepthAnything if KIND == "standard" else mde.EnhancedDepthAnything
base = synth_images(4, S, S, 1234)
x = base.repeat((B + 3) // 4, 1, 1, 1)[:B].contiguous().to(dev)
md = {"source": "rover"}
sd = None


def step_ms(patch):
    """patch: dict ops-attribute -> replacement, applied while a fresh model (fresh graph) is built and timed."""
    global sd
    saved = {k: getattr(ops, k) for k in patch}
    for k, v in patch.items():
        setattr(ops, k, v)
    try:
        m = cls(pretrained=False)
        if sd is None:
            sd = synth_state_dict(m.state_dict(), 0)
        m.load_state_dict(sd)
        m = m.eval().to(dev)
        for _ in range(4):
            m(x, md)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            m(x, md)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 20
    finally:
        for k, v in saved.items():
            setattr(ops, k, v)
        del m
        torch.cuda.empty_cache()


def skip(*a, **k):
    return None


_gemm = ops.gemm
_conv = ops.conv_nhwc


def gemm_without(epis, n_only=None):
    def f(a, w, bias, epilogue, out, aux0=None, aux1=None, M=None):
        if epilogue in epis and (n_only is None or w.shape[0] == n_only):
            return out
        return _gemm(a, w, bias, epilogue, out, aux0, aux1, M)
    return f


def conv_without(pred):
    def f(inp, h, w, cin, wpacked, bias, cout, *a, **k):
        if pred(h, w, cin, cout):
            return None
        return _conv(inp, h, w, cin, wpacked, bias, cout, *a, **k)
    return f


cases = [
    ("layernorm x24", {"layernorm": skip}),
    ("vit_attention x12", {"vit_attention": skip}),
    ("gemm qkv x12", {"gemm": gemm_without({L.EPI_QKV_SPLIT})}),
    ("gemm fc1 x12", {"gemm": gemm_without({L.EPI_BIAS_GELU_BF16})}),
    ("gemm proj x12 (resid, K=768)", {"gemm": None}),
    ("gemm fc2 x12 (resid, K=3072)", {"gemm": None}),
    ("gemm reduce x4", {"gemm": gemm_without({L.EPI_BIAS_F32})}),
    ("tokens_to_grid x4", {"tokens_to_grid": skip}),
    ("instnorm x4", {"instnorm_nhwc": skip}),
    ("upsample2x_concat x4", {"upsample2x_concat": skip}),
    ("conv: all", {"conv_nhwc": skip}),
    ("conv: 224x240 maps (16ch x3 + head)", {"conv_nhwc": conv_without(lambda h, w, ci, co: h >= 200)}),
    ("conv: 112x120 maps", {"conv_nhwc": conv_without(lambda h, w, ci, co: 100 <= h < 200)}),
    ("conv: 56x60 maps", {"conv_nhwc": conv_without(lambda h, w, ci, co: 50 <= h < 100)}),
    ("conv: 28x30 maps", {"conv_nhwc": conv_without(lambda h, w, ci, co: 20 <= h < 50)}),
    ("conv: 14x15 maps", {"conv_nhwc": conv_without(lambda h, w, ci, co: h < 20)}),
    ("input_stage", {"input_stage": skip}),
    ("resize_maps (output)", {"resize_maps": skip}),
]
if KIND == "enhanced":
    cases += [("spatial_attention x2", {"spatial_attention": skip}), ("enh_fuse", {"enh_fuse": skip}), ("rgb_to_grid", {"rgb_to_grid": skip})]


def gemm_resid(kdim):
    def f(a, w, bias, epilogue, out, aux0=None, aux1=None, M=None):
        if epilogue == L.EPI_BIAS_RESID_F32 and a.shape[1] == kdim:
            return out
        return _gemm(a, w, bias, epilogue, out, aux0, aux1, M)
    return f


cases[4] = (cases[4][0], {"gemm": gemm_resid(768)})
cases[5] = (cases[5][0], {"gemm": gemm_resid(3072)})

full = step_ms({})
full2 = step_ms({})
print(f"{KIND} B={B} {S}px: full step {full:.4f} ms (repeat {full2:.4f})")
print(f"{'removed':44s} {'step ms':>9s} {'saved us':>9s} {'share':>7s}")
tot = 0.0
for name, patch in cases:
    t = step_ms(patch)
    saved = (full - t) * 1e3
    if not name.startswith("conv: all"):
        tot += saved
    print(f"{name:44s} {t:9.4f} {saved:9.1f} {saved / (full * 1e3) * 100:6.1f}%")
print(f"sum of marginal savings (conv: all excluded) {tot:.1f} us of {full * 1e3:.1f} us")
