"""Last decoder level: the fused kernel (mde_decoder_tail16) against the four-launch path it replaces, back to back.
usage: python tools/tail_bench.py [batch] [reps] [fused|unfused|both]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
from monodepth_estimation_b200 import _lib as L  # noqa: E402
from monodepth_estimation_b200 import ops  # noqa: E402
from monodepth_estimation_b200.packing import pack_conv  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    which = sys.argv[3] if len(sys.argv) > 3 else "both"
    h, w = 112, 120
    g = torch.Generator(device="cuda").manual_seed(0)
    rnd = lambda *s, sc=1.0: torch.randn(*s, device="cuda", generator=g) * sc
    xp = rnd(B, h + 2, w + 2, 16).to(torch.bfloat16)
    packed = [pack_conv(rnd(c, 16, 3, 3, sc=0.1), rnd(c, sc=0.1)) for c in (16, 16, 1)]
    out = torch.empty(B, 2 * h, 2 * w, device="cuda")
    cat = torch.empty(B, 2 * h + 2, 2 * w + 2, 16, device="cuda", dtype=torch.bfloat16)
    a1, a2 = torch.empty_like(cat), torch.empty_like(cat)

    def fused():
        ops.decoder_tail16(xp, h, w, packed[0], packed[1], packed[2], out)

    def unfused():
        ops.upsample2x_concat(xp, h, w, None, 0, 0, cat)
        ops.conv_nhwc(cat, 2 * h, 2 * w, 16, packed[0][0], packed[0][1], 16, 3, L.PAD_REFLECT, L.ACT_ELU, out_bf16=a1)
        ops.conv_nhwc(a1, 2 * h, 2 * w, 16, packed[1][0], packed[1][1], 16, 3, L.PAD_REFLECT, L.ACT_ELU, out_bf16=a2)
        ops.conv_nhwc(a2, 2 * h, 2 * w, 16, packed[2][0], packed[2][1], 1, 3, L.PAD_ZERO, L.ACT_SIGMOID, out_f32=out)

    for name, fn in (("fused", fused), ("unfused", unfused)):
        if which not in (name, "both"):
            continue
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{name:8s} B={B}: {e0.elapsed_time(e1) * 1e3 / reps:8.2f} us per call")


if __name__ == "__main__":
    main()
