"""ORACLE — test infrastructure.  Generates ``tests/golden/*.npz`` by running the REAL
reference classes imported from ``/root/reference`` (this container only; the path does
not exist on the GPU box) on the deterministic synthetic weights/inputs of
``monodepth-estimation_b200/synth.py``, and cross-checks the restatement in
``oracle/depth_oracle.py`` against them while doing so.

    python oracle/make_golden.py            # writes tests/golden, prints max deviations
    python oracle/make_golden.py --only cfg4_1024_b1_rover   # (re)generate the named model cases only

Each .npz holds, per output map, a strided subsample (``cases.stride_for``) plus the
full-map mean/std/min/max in float64, and the torch/torchvision versions used.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import pkgload  # noqa: E402

pkg = pkgload.load()
from monodepth_estimation_b200.synth import synth_images, synth_state_dict  # noqa: E402
from oracle import depth_oracle as O  # noqa: E402
from oracle.cases import CASES, FUSION_CASES, REFINE_CASES, key_name, stride_for  # noqa: E402

from models.enhanced_model import EnhancedDepthAnything as RefEnhanced  # noqa: E402
from models.model import DepthAnything as RefStandard  # noqa: E402
from models.refinement import MultiScaleDepthFusion as RefFusion  # noqa: E402
from models.refinement import TextureAwareRefinementModule as RefRefine  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def pack(maps: dict) -> dict:
    import torchvision

    rec = {"torch_version": np.array(torch.__version__), "torchvision_version": np.array(torchvision.__version__)}
    for k, v in maps.items():
        a = v.detach().double().numpy()
        sy, sx = stride_for(a.shape[2]), stride_for(a.shape[3])
        n = key_name(k) if isinstance(k, tuple) else k
        rec[n] = a[:, :, ::sy, ::sx].astype(np.float32)
        rec[n + "__stats"] = np.array([a.mean(), a.std(), a.min(), a.max()], dtype=np.float64)
        rec[n + "__shape"] = np.array(a.shape)
    return rec


def dev(a, b):
    return float((a - b).abs().max()), float((a - b).norm() / b.norm().clamp_min(1e-30))


def ref_refine_live(m, depth, rgb):
    """The reference module's own sub-modules called in refinement.py's order, skipping
    the crashing dead lines :87-90."""
    import torch.nn.functional as F

    if rgb.shape[2:] != depth.shape[2:]:
        rgb = F.interpolate(rgb, size=depth.shape[2:], mode="bilinear", align_corners=False)
    rf = m.activation(m.rgb_conv2(m.activation(m.rgb_conv1(rgb))))
    df = m.activation(m.depth_conv2(m.activation(m.depth_conv1(depth))))
    edge = torch.sqrt(m.edge_detect_x(rgb).pow(2) + m.edge_detect_y(rgb).pow(2))
    cf = m.activation(m.combined_conv1(torch.cat([rf, df], dim=1)))
    cf = m.activation(m.combined_conv2(cf))
    return depth + m.output_conv(cf) * torch.sigmoid(edge * 5.0)


def main():
    only = set(sys.argv[sys.argv.index("--only") + 1:]) if "--only" in sys.argv else None
    os.makedirs(GOLD, exist_ok=True)
    torch.manual_seed(0)
    models = {}
    worst = 0.0
    with torch.no_grad():
        for name, (kind, B, H, W, ikind, md, wseed, iseed) in CASES.items():
            if only is not None and name not in only:
                continue
            base = "enhanced" if kind == "config4" else kind
            mk = (base, wseed)
            if mk not in models:
                m = (RefEnhanced if base == "enhanced" else RefStandard)(pretrained=False)
                m.load_state_dict(synth_state_dict(m.state_dict(), wseed))
                models[mk] = m.eval()
            m = models[mk]
            sd = m.state_dict()
            x = synth_images(B, H, W, iseed, ikind)
            out = m(x.clone(), md)
            if kind == "config4":
                fu, rf = RefFusion().eval(), RefRefine().eval()
                fu.load_state_dict(synth_state_dict(fu.state_dict(), wseed + 100))
                rf.load_state_dict(synth_state_dict(rf.state_dict(), wseed + 200))
                fused = fu(out)
                out[("disp_msfused", 0)] = fused
                out[("disp_refined", 0)] = ref_refine_live(rf, fused, x)
                mine = O.config4_forward(sd, fu.state_dict(), rf.state_dict(), x.clone(), md)
            elif kind == "enhanced":
                mine = O.enhanced_forward(sd, x.clone(), md)
            else:
                mine = O.standard_forward(sd, x.clone(), md)
            assert set(mine) == set(out), (set(mine) ^ set(out))
            for k in out:
                ma, rl = dev(mine[k], out[k])
                worst = max(worst, rl)
                print(f"{name:28s} {str(k):24s} oracle-vs-reference max_abs {ma:.3e} rel_l2 {rl:.3e}")
            np.savez_compressed(os.path.join(GOLD, name + ".npz"), **pack(out))

        for name, (H, W, rh, rw, seed) in ({} if only is not None else REFINE_CASES).items():
            rf = RefRefine().eval()
            rf.load_state_dict(synth_state_dict(rf.state_dict(), seed))
            depth = synth_images(1, H, W, seed, "unit")[:, :1].contiguous()
            rgb = synth_images(1, rh, rw, seed + 1, "unit")
            try:
                rf(depth, rgb)
                raise SystemExit("reference TextureAwareRefinementModule.forward unexpectedly ran")
            except RuntimeError as e:
                print(f"{name}: reference forward raises as documented: {str(e)[:80]}")
            want = ref_refine_live(rf, depth, rgb)
            got = O.texture_refine_module(rf.state_dict(), depth, rgb)
            print(f"{name:28s} oracle-vs-reference-submodules {dev(got, want)}")
            worst = max(worst, dev(got, want)[1])
            np.savez_compressed(os.path.join(GOLD, name + ".npz"), **pack({"refined": want}))

        for name, (H, W, seed) in ({} if only is not None else FUSION_CASES).items():
            fu = RefFusion().eval()
            fu.load_state_dict(synth_state_dict(fu.state_dict(), seed))
            maps = {("disp", s): synth_images(1, H >> s, W >> s, seed + s, "unit")[:, :1].contiguous() for s in range(4)}
            want = fu(maps)
            got = O.multiscale_depth_fusion(fu.state_dict(), maps)
            print(f"{name:28s} oracle-vs-reference {dev(got, want)}")
            worst = max(worst, dev(got, want)[1])
            np.savez_compressed(os.path.join(GOLD, name + ".npz"), **pack({"fused": want}))
    print(f"worst oracle-vs-reference rel_l2 over all cases: {worst:.3e}")


if __name__ == "__main__":
    main()
