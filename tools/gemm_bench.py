"""Per-class timing of the ViT linear layers through the C-ABI GEMM: each class is launched `reps` times back to back
(GPU-bound: one launch is 10-40 us of device time against ~5 us of host time) between two CUDA events.
usage: python tools/gemm_bench.py [batch] [reps]
MDE_BENCH_LN=1 times the LayerNorm-fused forms (mde_gemm_ln_bf16: proj/fc2 as producers, qkv/fc1 as consumers) next to the plain ones."""
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
    M = B * 197
    bf, f32 = torch.bfloat16, torch.float32
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s, dt=bf: (torch.randn(*s, device=dev, generator=g) * 0.05).to(dt)
    q, k, vt = (torch.empty(B * 12 * 197 * 64, device=dev, dtype=bf), torch.empty(B * 12 * 197 * 64, device=dev, dtype=bf),
                torch.zeros(B * 12 * 64 * L.VT_PITCH, device=dev, dtype=bf))
    cases = [("qkv", 2304, 768, L.EPI_QKV_SPLIT), ("proj", 768, 768, L.EPI_BIAS_RESID_F32), ("fc1", 3072, 768, L.EPI_BIAS_GELU_BF16),
             ("fc2", 768, 3072, L.EPI_BIAS_RESID_F32)]
    total = 0.0
    for name, N, K, epi in cases:
        a, w, bias = rnd(M, K), rnd(N, K), rnd(N, dt=f32)
        if epi == L.EPI_QKV_SPLIT:
            args = (q, k, vt)
        elif epi == L.EPI_BIAS_RESID_F32:
            args = (torch.empty(M, N, device=dev, dtype=f32), rnd(M, N, dt=f32))
        else:
            args = (torch.empty(M, N, device=dev, dtype=bf),)
        for _ in range(5):
            ops.gemm(a, w, bias, epi, *args)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            ops.gemm(a, w, bias, epi, *args)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        total += us
        print(f"{name:5s} M={M} N={N} K={K}: {us:7.2f} us  {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s")
        if os.environ.get("MDE_BENCH_LN") == "1":
            st = torch.zeros(24, M, 2, device=dev, dtype=f32)
            if epi == L.EPI_BIAS_RESID_F32:
                kw = dict(xb_out=torch.empty(M, N, device=dev, dtype=bf), stats_out=st)
            else:
                st[..., 1] = 1.0
                kw = dict(stats_in=st, colsum=rnd(N, dt=f32))
            for _ in range(5):
                ops.gemm_ln(a, w, bias, epi, *args, **kw)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(reps):
                ops.gemm_ln(a, w, bias, epi, *args, **kw)
            e1.record()
            torch.cuda.synchronize()
            us2 = e0.elapsed_time(e1) * 1e3 / reps
            print(f"      LayerNorm-fused ({'producer' if 'xb_out' in kw else 'consumer'}): {us2:7.2f} us  ({us2 - us:+.2f})")
        if os.environ.get("MDE_GEMM_PROF") == "1":  # in-kernel clock64 counters of ONE launch (mde_debug_gemm_prof)
            prof = torch.zeros(148, 8, device=dev, dtype=torch.int64)
            L.call("mde_debug_gemm_prof", prof.data_ptr())
            ops.gemm(a, w, bias, epi, *args)
            torch.cuda.synchronize()
            L.call("mde_debug_gemm_prof", None)
            pr = prof.double().cpu() / 1.9e3  # us at ~1.9 GHz
            lead, both = pr[0::2], pr
            print(f"      per CTA (us, mean/max): MMA waits accumulator {lead[:, 0].mean():.2f}/{lead[:, 0].max():.2f}, MMA waits operands "
                  f"{lead[:, 1].mean():.2f}/{lead[:, 1].max():.2f}, epilogue waits MMA {both[:, 2].mean():.2f}/{both[:, 2].max():.2f}, epilogue works "
                  f"{both[:, 3].mean():.2f}/{both[:, 3].max():.2f} over {both[:, 4].mean():.2f} tiles, CTA lifetime {both[:, 5].mean():.2f}/{both[:, 5].max():.2f}, "
                  f"prologue {both[:, 6].mean():.2f}/{both[:, 6].max():.2f}, last MMA issued at {lead[:, 7].mean():.2f}/{lead[:, 7].max():.2f}")
            print("      lifetime per cluster (us):", " ".join(f"{v:.1f}" for v in lead[:, 5].tolist()[:12]), "...", " ".join(f"{v:.1f}" for v in lead[:, 5].tolist()[-4:]))
    print(f"per ViT layer (4 GEMMs): {total:.1f} us; x12 = {total * 12 / 1e3:.3f} ms")


if __name__ == "__main__":
    main()
