"""BASELINE.json config 4 on one GPU: enhanced forward -> MultiScaleDepthFusion -> TextureAwareRefinementModule
(composition as stated in oracle.config4_forward).  Prints images/s and the warm per-kernel split of the two modules.
usage: python tools/config4_bench.py [batch] [size] [steps]"""
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pkgload  # noqa: E402

pkgload.load()
import monodepth_estimation_b200 as mde  # noqa: E402
from monodepth_estimation_b200.synth import synth_images, synth_state_dict  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    S = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    mk = lambda cls, seed: (lambda m: (m.load_state_dict(synth_state_dict(m.state_dict(), seed)), m.eval().cuda())[1])(cls())
    m = mde.EnhancedDepthAnything(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), 0))
    m = m.eval().cuda()
    fu, rf = mk(mde.MultiScaleDepthFusion, 100), mk(mde.TextureAwareRefinementModule, 200)
    x = synth_images(min(B, 2), S, S, 1234).repeat((B + 1) // 2, 1, 1, 1)[:B].contiguous().cuda()
    md = {"source": "rover"}

    def step():
        out = m(x, md)
        fused = fu(out)
        out[("disp_msfused", 0)] = fused
        out[("disp_refined", 0)] = rf(fused, x)
        return out

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    gflop = B * (39.11 + 352.90 + 0.084)
    print(f"config4 B={B} {S}px: {ms:.3f} ms/step, {B / ms * 1e3:.1f} images/s, {gflop / ms:.1f} TFLOP/s algorithmic")
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            step()
        torch.cuda.synchronize()
    agg = collections.OrderedDict()
    for ev in prof.events():
        if ev.device_type != torch.autograd.DeviceType.CUDA or "mem" in ev.name.lower()[:8]:
            continue
        name = ev.name.split("(")[0].replace("void ", "").replace("mde::", "")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ev.device_time_total
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print(f"{t / 2:10.1f} us/step {c / 2:6.1f} n/step {t / c:9.2f} avg us  {n[:80]}")


if __name__ == "__main__":
    main()
