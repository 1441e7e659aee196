"""ORACLE — test infrastructure.  Pins ``oracle/terrain_oracle.py`` to the REAL reference classes
(``utils/terrain_reconstruction_fixed.py``, ``models/refinement.py:DepthQualityEnhancement``) imported from
``/root/reference`` in the build container, and writes ``tests/golden/terrain_*.npz``.

    python oracle/make_golden_terrain.py

The reference module imports matplotlib at the top (not installed, unused on this path): empty stand-in modules are
registered for the import only.  ``cv2.ximgproc`` is not installed, so the joint-bilateral branch cannot be run: unpinned.
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

for name in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["mpl_toolkits.mplot3d"].Axes3D = object

from oracle import terrain_oracle as T  # noqa: E402
from oracle.terrain_cases import (BILATERAL_CASES, GUIDED_CASES, HOLE_CASES, PC_CASES, TEMPORAL_CASES, guided_inputs, hole_inputs,  # noqa: E402
                                  pc_inputs, temporal_inputs)

from models.refinement import DepthQualityEnhancement as RefDQE  # noqa: E402
from utils.terrain_reconstruction_fixed import MarsTerrainMapper as RefMapper  # noqa: E402
from utils.terrain_reconstruction_fixed import TerrainReconstructor as RefReconstructor  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


class _FixedDisp(torch.nn.Module):
    """Stands in for the depth model so that infer_depth's own arithmetic is what gets pinned."""

    def __init__(self, disp):
        super().__init__()
        self.disp = disp
        self.p = torch.nn.Parameter(torch.zeros(1))

    def forward(self, image, metadata=None):
        return {("disp", 0): self.disp}


def same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


def main():
    rec = {"numpy_version": np.array(np.__version__)}
    for name, case in PC_CASES.items():
        depth, k, rgb = pc_inputs(case)
        r = RefReconstructor(min_depth=case["min_depth"], max_depth=case["max_depth"])
        d_in = torch.from_numpy(depth)[None] if case.get("as_tensor") else depth
        if rgb is None:
            want, wantc = r.create_point_cloud(d_in, k), None
        else:
            want, wantc = r.reconstruct_terrain(depth, rgb, intrinsics=k)
        got, valid = T.create_point_cloud(depth, k, case["min_depth"], case["max_depth"])
        assert same(got, want), name
        rec[f"{name}__points"] = want
        if rgb is not None:
            assert same(T.reconstruct_colors(rgb, valid), wantc), name
            rec[f"{name}__colors"] = wantc
        print(f"{name:24s} points {want.shape} {want.dtype}: oracle bit-identical to the reference")

    # infer_depth: the reference method with a stand-in model (CPU)
    disp = torch.from_numpy(np.random.Generator(np.random.PCG64(7)).random((2, 1, 24, 40), dtype=np.float32))
    m = RefMapper(_FixedDisp(disp), min_depth=0.1, max_depth=100.0, use_cuda=False)
    want = m.infer_depth(torch.zeros(2, 3, 24, 40)).numpy()
    got = T.infer_depth_from_disp(disp, 0.1, 100.0).numpy()
    assert same(got, want)
    m2 = RefMapper(_FixedDisp(disp), min_depth=0.5, max_depth=30.0, use_cuda=False)
    want2 = m2.infer_depth(torch.zeros(2, 3, 24, 40)).numpy()
    assert same(T.infer_depth_from_disp(disp, 0.5, 30.0).numpy(), want2)
    rec["infer_depth__disp"], rec["infer_depth__0.1_100"], rec["infer_depth__0.5_30"] = disp.numpy(), want, want2
    print("infer_depth              oracle bit-identical to the reference")

    for name, case in TEMPORAL_CASES.items():
        maps, w = temporal_inputs(case)
        want = RefDQE.temporal_consistency_filter(maps, w)
        assert same(T.temporal_consistency_filter(maps, w), want), name
        rec[f"{name}__out"] = want
        print(f"{name:24s} {want.shape} {want.dtype}: oracle bit-identical to the reference")

    for name, case in BILATERAL_CASES.items():
        depth = T.synth_depth(case["H"], case["W"], case["seed"], holes=False) * case.get("mul", 1.0)
        want = RefDQE.bilateral_depth_filter(depth, None, **case.get("kw", {}))
        got = T.bilateral_depth_filter(depth, None, **case.get("kw", {}))
        span = float(depth.max() - depth.min())
        err = float(np.abs(got - want).max()) / span
        print(f"{name:24s} oracle vs reference (cv2 {__import__('cv2').__version__}): max abs / range {err:.2e}")
        assert err < 2e-6, (name, err)
        rec[f"{name}__out"] = want.astype(np.float32)
    for name, case in HOLE_CASES.items():
        depth = hole_inputs(case)
        with np.errstate(invalid="ignore"):
            want = RefDQE.fill_depth_holes(depth, case["max_hole_size"])
            got = T.fill_depth_holes(depth, case["max_hole_size"])
            got2 = T.fill_depth_holes_numpy(depth, case["max_hole_size"])
        assert same(got, want), name
        ok = T.nan_free_windows(T.fill_depth_holes(depth, case["max_hole_size"], return_prefilter=True)[1])
        assert case.get("nan_ring") or ok.all(), name
        assert same(np.where(ok, got2, 0), np.where(ok, want, 0)), name  # the library-free restatement computes the same thing
        n_bad = int((~np.isfinite(depth) | (depth <= 0)).sum())
        print(f"{name:24s} {n_bad} invalid pixels: oracle (and its NumPy-only form) bit-identical to the reference")
        rec[f"{name}__out"] = want
    # edge_preserving_smoothing needs cv2.ximgproc (opencv-contrib): absent, the reference method cannot run -> parity unpinned.
    # What CAN be pinned is its grey conversion: cv2.cvtColor is in the installed cv2.
    import cv2

    for name, case in GUIDED_CASES.items():
        depth, rgb = guided_inputs(case)
        try:
            RefDQE.edge_preserving_smoothing(depth, rgb, case["lam"])
            raise SystemExit("cv2.ximgproc unexpectedly present: pin the guided filter against it")
        except AttributeError as e:
            print(f"{name:24s} reference raises as documented ({str(e)[:60]}): unpinned")
        if rgb is not None:
            assert np.array_equal(T.rgb_to_gray_u8(rgb), cv2.cvtColor(rgb, cv2.COLOR_RGB2GRAY)), name
        rec[f"{name}__oracle"] = T.edge_preserving_smoothing(depth, rgb, case["lam"])  # the oracle's own output (regression vector, not a pin)
    np.savez_compressed(os.path.join(GOLD, "terrain_golden.npz"), **rec)
    print("wrote", os.path.join(GOLD, "terrain_golden.npz"))


if __name__ == "__main__":
    main()
