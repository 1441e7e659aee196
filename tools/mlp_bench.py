"""fc1 + fc2 of one ViT layer at batch B: the fused launch (mde_mlp_bf16) against the two GEMM launches, each replayed from a CUDA
graph of `reps` launches.  usage: python tools/mlp_bench.py [batch] [reps]   (guidance, not a bench value)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
from monodepth_estimation_b200 import _lib as L, ops  # noqa: E402

B, reps = ([int(v) for v in sys.argv[1:]] + [32, 20][len(sys.argv) - 1:])[:2]
M, D, Hd = B * 197, 768, 3072
BF16 = torch.bfloat16
torch.manual_seed(0)
a = torch.randn(M, D, device="cuda").to(BF16)
w1, b1 = (torch.randn(Hd, D, device="cuda") * D ** -0.5).to(BF16), torch.randn(Hd, device="cuda") * 0.1
w2, b2 = (torch.randn(D, Hd, device="cuda") * Hd ** -0.5).to(BF16), torch.randn(D, device="cuda") * 0.1
resid, out = torch.randn(M, D, device="cuda"), torch.empty(M, D, device="cuda")
hid = torch.empty(M, Hd, device="cuda", dtype=BF16)
sync = torch.zeros(128, device="cuda", dtype=torch.int32)


def timed(fn):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def two():
    ops.gemm(a, w1, b1, L.EPI_BIAS_GELU_BF16, hid)
    ops.gemm(hid, w2, b2, L.EPI_BIAS_RESID_F32, out, resid)


t2 = timed(two)
t1 = timed(lambda: ops.mlp(a, w1, b1, hid, w2, b2, resid, out, sync))
fl = 4.0 * M * D * Hd
print(f"B={B} M={M}: two GEMMs {t2:.2f} us ({fl / t2 * 1e-6:.0f} TFLOP/s), fused MLP {t1:.2f} us ({fl / t1 * 1e-6:.0f} TFLOP/s); sync clean: {int(sync.abs().sum()) == 0}")

if os.environ.get("MDE_GEMM_PROF") == "1":  # in-kernel time stamps of ONE fused launch (build with MDE_BUILD_PROF=1)
    prof = torch.zeros(148 * 8 + 74 * 16, device="cuda", dtype=torch.int64)
    L.call("mde_debug_gemm_prof", prof.data_ptr())
    ops.mlp(a, w1, b1, hid, w2, b2, resid, out, sync)
    torch.cuda.synchronize()
    L.call("mde_debug_gemm_prof", None)
    hist = prof[148 * 8:].view(74, 16).double().cpu() / 1.9e3
    pr = prof[:148 * 8].view(148, 8).double().cpu() / 1.9e3
    print("   fc1 operand wait by k-block index (us per cluster, summed over its fc1 tiles):", [round(float(v), 2) for v in hist.mean(0)[:12]])
    names = ["fc1: MMA waits accumulator", "producer at first fc2", "row block complete", "MMA starts first fc2", "last MMA issued", "epilogue done", "last fc1 signalled", "fc1: MMA waits operands"]
    lead = pr[0::2]
    for i, nme in enumerate(names):
        c = lead[:, i]
        c = c[c > 0]
        if len(c):
            print(f"   {nme:24s} mean {c.mean():7.2f}  min {c.min():7.2f}  max {c.max():7.2f} us ({len(c)} leader CTAs)")
    print("   per cluster [prod@fc2, ready, mma fc2 start, last mma, epi done, fc1 signalled]:")
    for c in (0, 1, 7, 8, 36, 72, 73):
        print("    cluster", c, [round(float(v), 1) for v in lead[c, 1:7]])
