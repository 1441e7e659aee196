"""ViT attention through the C-ABI, launched `reps` times back to back between two CUDA events, and checked against torch.
usage: python tools/attn_bench.py [batch] [reps]      (MDE_ATTN=ps selects the persistent kernel)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
from monodepth_estimation_b200 import _lib as L  # noqa: E402
from monodepth_estimation_b200 import ops  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    q, k, v = (torch.randn(B, 12, 197, 64, device=dev, generator=g).to(torch.bfloat16) for _ in range(3))
    qs = (q.float() * 0.125).to(torch.bfloat16)
    vt = torch.zeros(B, 12, 64, L.VT_PITCH, device=dev, dtype=torch.bfloat16)
    vt[..., :197] = v.transpose(-1, -2)
    tm = lambda t: t.permute(0, 2, 1, 3).contiguous().view(-1)
    qf, kf, vf = tm(qs), tm(k), vt.view(-1)
    out = ops.vit_attention(qf, kf, vf, B)
    att = torch.softmax(qs.float() @ k.float().transpose(-1, -2), dim=-1)
    ref = (att @ v.float()).transpose(1, 2).reshape(B * 197, 768)
    err = float((out.float() - ref).norm() / ref.norm())
    for _ in range(5):
        ops.vit_attention(qf, kf, vf, B, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ops.vit_attention(qf, kf, vf, B, out)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    flops = 4.0 * B * 12 * 197 * 197 * 64
    which = os.environ.get('MDE_ATTN', 'tp')
    print(f"attention[{which}] B={B}: {us:7.2f} us  {flops / us / 1e6:6.1f} TFLOP/s (algorithmic)  rel-L2 vs torch fp32 {err:.2e}")


if __name__ == "__main__":
    main()
