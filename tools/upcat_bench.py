"""The four upsample2x_concat launches of one standard forward (batch 32), each replayed from a CUDA graph (device time only).
usage: python tools/upcat_bench.py [batch] [reps]    MDE_UPCAT_ROWS=0 selects the row-per-block kernel (guidance, not a bench value)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
from monodepth_estimation_b200 import ops  # noqa: E402

B, reps = ([int(v) for v in sys.argv[1:]] + [32, 20][len(sys.argv) - 1:])[:2]
BF16 = torch.bfloat16
tot = 0.0
for lvl, (h, w, ca, cs) in enumerate([(14, 15, 128, 128), (28, 30, 64, 64), (56, 60, 32, 32), (112, 120, 16, 0)]):
    a = torch.randn(B, h + 2, w + 2, ca, device="cuda").to(BF16)
    s = torch.randn(B, 16, 17, cs, device="cuda").to(BF16) if cs else None
    out = torch.empty(B, 2 * h + 2, 2 * w + 2, ca + cs, device="cuda", dtype=BF16)
    run = lambda: ops.upsample2x_concat(a, h, w, s, 14, 15, out)
    run()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            run()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    byt = out.numel() * 2 + a.numel() * 2 + (s.numel() * 2 if cs else 0)
    tot += us
    print(f"level {lvl}: {h}x{w} ca={ca} cs={cs}: {us:.1f} us, {byt / us * 1e-3:.0f} GB/s")
print(f"sum {tot:.1f} us")
