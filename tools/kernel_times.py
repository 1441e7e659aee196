"""Warm, in-situ kernel durations of one forward (CUDA-graph replay included) from the CUPTI activity trace
(torch.profiler).  Guidance only: numbers taken under a profiler are never reported as bench values.
usage: python tools/kernel_times.py [standard|enhanced] [batch] [size]"""
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
    kind = sys.argv[1] if len(sys.argv) > 1 else "standard"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    S = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    cls = mde.DepthAnything if kind == "standard" else mde.EnhancedDepthAnything
    m = cls(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), 0))
    m = m.eval().cuda()
    x = synth_images(min(B, 4), S, S, 1234).repeat((B + 3) // 4, 1, 1, 1)[:B].contiguous().cuda()
    md = {"source": "rover"}
    for _ in range(3):
        m(x, md)
    torch.cuda.synchronize()
    steps = 3
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        for _ in range(steps):
            m(x, md)
        torch.cuda.synchronize()
    agg = collections.OrderedDict()
    tot = 0.0
    for ev in prof.events():
        if ev.device_type != torch.autograd.DeviceType.CUDA or "memcpy" in ev.name.lower() or "memset" in ev.name.lower():
            continue
        name = ev.name.split("(")[0].replace("void ", "").replace("mde::", "")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
        tot += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
    print(f"{kind} B={B} {S}px: kernel time {tot / steps:.1f} us/forward (sum of kernels, warm)")
    print(f"{'us/fwd':>9} {'share':>6} {'n/fwd':>6} {'avg us':>8}  kernel")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t / steps:9.1f} {t / tot:6.1%} {c / steps:6.1f} {t / c:8.2f}  {n[:90]}")


if __name__ == "__main__":
    main()
