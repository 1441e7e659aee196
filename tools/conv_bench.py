"""One 3x3 conv layer launched back to back between two CUDA events (guidance for conv_halo.cu work, not a bench value).
usage: python tools/conv_bench.py [cin] [cout] [h] [w] [batch] [reps] [f32head]
MDE_HALO_DBG=<mask> removes parts of the kernel (1 epilogue, 2 MMAs, 4 stores, 8 halo loads) to see what the rest costs."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
from monodepth_estimation_b200 import _lib as L, ops  # noqa: E402
from monodepth_estimation_b200.packing import pack_conv  # noqa: E402


def main():
    a = [int(v) for v in sys.argv[1:]]
    cin, cout, h, w, B, reps, head = (a + [16, 16, 224, 224, 32, 20, 0][len(a):])[:7]
    torch.manual_seed(0)
    x = torch.randn(B, h + 2, w + 2, cin, device="cuda").to(torch.bfloat16)
    wt, b = pack_conv(torch.randn(cout, cin, 3, 3, device="cuda") * 0.1, torch.zeros(cout, device="cuda"))
    if head:
        out32, out16 = torch.empty(B, h, w, cout, device="cuda"), None
    else:
        out32, out16 = None, torch.empty(B, h + 2, w + 2, max(cout, 16), device="cuda", dtype=torch.bfloat16)
    run = lambda: ops.conv_nhwc(x, h, w, cin, wt, b, cout, 3, L.PAD_ZERO if head else L.PAD_REFLECT, L.ACT_SIGMOID if head else L.ACT_ELU, out_bf16=out16, out_f32=out32)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()  # the Python + ctypes + tensor-map encode of one call is ~15 us: replay a graph so the device is timed
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
    px = B * h * w
    print(f"conv {cin}->{cout} {h}x{w} B={B} dbg={os.environ.get('MDE_HALO_DBG', '0')}: {us:.1f} us/launch, "
          f"{2 * 9 * cin * cout * px / us * 1e-6:.1f} TFLOP/s, {px * 2 * (cin + (4 if head else 2) * cout / 2) / us * 1e-3:.0f} GB/s algorithmic")


if __name__ == "__main__":
    main()
