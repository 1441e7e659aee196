"""Input stage and output stage of the standard forward, each replayed from a CUDA graph of `reps` launches (device time only).
usage: python tools/io_bench.py [batch] [size] [reps]      (guidance for pointwise.cu work, not a bench value)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
import monodepth_estimation_b200 as mde  # noqa: E402
from monodepth_estimation_b200.synth import bench_batch, synth_state_dict  # noqa: E402


def timed(fn, reps):
    for _ in range(2):
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
    return e0.elapsed_time(e1) * 1e-3 / reps


def main():
    a = [int(v) for v in sys.argv[1:]]
    B, S, reps = (a + [32, 1024, 10][len(a):])[:3]
    kind = os.environ.get("IO_MODEL", "standard")
    m = (mde.EnhancedDepthAnything if kind == "enhanced" else mde.DepthAnything)(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), 0))
    m = m.eval().cuda()
    x = bench_batch(B, S, 0).cuda()
    with torch.no_grad():
        m(x, {"source": "rover"})
    eng = m._engine
    ws, w = eng.workspace(x.device, B, S, S), eng.weights()
    t_in = timed(lambda: eng._pre(ws, w, x), reps)
    keep = []
    t_out = timed(lambda: keep.append(eng._post(ws, 1.0)), reps)
    n_maps = 10 if eng.enhanced else 4
    b_in, b_out = x.numel() * 4, n_maps * B * S * S * 4
    print(f"{kind} B={B} {S}px: input stage {t_in * 1e6:.1f} us = {b_in / t_in / 1e9:.0f} GB/s;  output stage {t_out * 1e6:.1f} us = {b_out / t_out / 1e9:.0f} GB/s")


if __name__ == "__main__":
    main()
