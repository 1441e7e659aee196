"""GPU: the point-cloud / disp->depth / temporal / bilateral kernels (csrc/terrain.cu) through the C-ABI and the
reference-shaped mirror classes, against the golden vectors of the real reference and the oracle.  Integer / index work
(which pixels survive, their order) and the IEEE float arithmetic are checked BIT-EXACT; the bilateral filter (cv2's SIMD
summation order is not reproducible) to 5e-6 of the depth range."""
import numpy as np
import pytest
import torch

from helpers import cached_model, load_golden, synth_images
from oracle import terrain_oracle as T
from oracle.terrain_cases import (BILATERAL_CASES, GUIDED_CASES, HOLE_CASES, PC_CASES, TEMPORAL_CASES, guided_inputs, hole_inputs, pc_inputs,
                                  temporal_inputs)

import monodepth_estimation_b200 as mde
from monodepth_estimation_b200 import ops

pytestmark = pytest.mark.gpu
GOLD = load_golden("terrain_golden")


def same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("name", list(PC_CASES))
def test_point_cloud_vs_reference_golden(name):
    case = PC_CASES[name]
    depth, k, rgb = pc_inputs(case)
    r = mde.TerrainReconstructor(min_depth=case["min_depth"], max_depth=case["max_depth"])
    d_in = torch.from_numpy(depth)[None] if case.get("as_tensor") else depth
    if rgb is None:
        pts = r.create_point_cloud(d_in, k)
    else:
        pts, cols = r.reconstruct_terrain(depth, rgb, intrinsics=k)
        assert same(cols, GOLD[f"{name}__colors"])
        pts2, cols2 = r.reconstruct_terrain(torch.from_numpy(depth), torch.from_numpy(rgb).permute(2, 0, 1), intrinsics=k)  # CHW tensor
        assert same(pts2, pts) and same(cols2, cols)
    assert same(pts, GOLD[f"{name}__points"])


@pytest.mark.parametrize("H,W,B,f64", [(1, 1, 1, True), (1, 2049, 2, True), (45, 91, 3, False), (3, 4096, 1, False), (257, 263, 2, True),
                                       (1024, 1024, 2, True), (1024, 1024, 2, False)])
def test_point_cloud_ragged_and_batched_vs_oracle(H, W, B, f64):
    """tile boundaries (2048 px), H*W % 4 != 0, per-image restart of the scan, both point dtypes"""
    depth = np.stack([T.synth_depth(H, W, 900 + 7 * b + H) for b in range(B)])
    k = T.default_intrinsics(H, W).astype(np.float64 if f64 else np.float32)
    rng = np.random.Generator(np.random.PCG64(H * W))
    rgb = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    r = mde.TerrainReconstructor()
    out = r.point_clouds(torch.from_numpy(depth)[:, None], k, rgb=rgb)
    assert len(out) == B
    for b in range(B):
        want, valid = T.create_point_cloud(depth[b], k)
        assert same(out[b][0].cpu().numpy(), want), (b, out[b][0].shape, want.shape)
        assert same(out[b][1].cpu().numpy(), T.reconstruct_colors(rgb[b], valid))


def test_point_cloud_all_and_none_valid():
    z = torch.full((2, 1, 50, 70), 5.0)
    z[1] = 0.0
    (p0, _), (p1, _) = mde.TerrainReconstructor().point_clouds(z)
    assert p0.shape == (3500, 3) and p1.shape == (0, 3)
    assert torch.equal(p0[:, 2].cpu(), torch.full((3500,), 5.0, dtype=torch.float64))


def test_point_cloud_full_size_properties():
    """BASELINE size (32 x 1024^2): survivors counted and ordered as a torch masked select; x, y invert to the pixel grid"""
    B, H, W = 32, 1024, 1024
    g = torch.Generator(device="cuda").manual_seed(5)
    z = torch.rand(B, H, W, device="cuda", generator=g) * 120.0 - 10.0  # ~8 % below min_depth, ~17 % above max_depth
    pts, _, counts = ops.point_cloud(z, (0.7 * W, 0.7 * W, W / 2, H / 2), 0.1, 100.0, f64=False)
    valid = (z > 0.1) & (z < 100.0)
    assert torch.equal(counts.long(), valid.flatten(1).sum(1))
    n = counts.cpu().tolist()
    for b in (0, 13, 31):
        zz = pts[b, : n[b], 2]
        assert torch.equal(zz, z[b][valid[b]])
        idx = valid[b].flatten().nonzero().squeeze(1)
        u = (pts[b, : n[b], 0].double() * (0.7 * W) / zz.double() + W / 2).round().long()
        v = (pts[b, : n[b], 1].double() * (0.7 * W) / zz.double() + H / 2).round().long()
        assert torch.equal(v * W + u, idx)


def test_disp_to_depth_and_fused_point_cloud():
    disp = torch.from_numpy(GOLD["infer_depth__disp"]).cuda()
    assert same(ops.disp_to_depth(disp, 0.1, 100.0).cpu().numpy(), GOLD["infer_depth__0.1_100"])
    assert same(ops.disp_to_depth(disp, 0.5, 30.0).cpu().numpy(), GOLD["infer_depth__0.5_30"])
    # disparity straight into the point-cloud kernel == depth map first, then the point cloud
    r = mde.TerrainReconstructor(min_depth=0.5, max_depth=30.0)
    fused = r.point_clouds(disp, from_disp=True)
    two_step = r.point_clouds(ops.disp_to_depth(disp, 0.5, 30.0))
    for (a, _), (b, _) in zip(fused, two_step):
        assert torch.equal(a, b) and a.shape[0] > 0
    want, _ = T.create_point_cloud(GOLD["infer_depth__0.5_30"][1, 0], None, 0.5, 30.0)
    assert same(fused[1][0].cpu().numpy(), want)


def test_mars_terrain_mapper_end_to_end():
    m = cached_model("standard", 0, "cuda")
    x = synth_images(2, 224, 224, 5)
    mapper = mde.MarsTerrainMapper(m, min_depth=0.1, max_depth=100.0)
    depth = mapper.infer_depth(x, {"source": "rover"})
    disp = m(x.cuda(), {"source": "rover"})[("disp", 0)]
    assert depth.shape == (2, 1, 224, 224)
    assert same(depth.cpu().numpy(), T.infer_depth_from_disp(disp.cpu(), 0.1, 100.0).numpy())
    # the NumPy entry with the /255 heuristic (terrain_reconstruction_fixed.py:152-163)
    depth_np = mapper.infer_depth((x[0].numpy() * 255.0), {"source": "rover"})
    assert depth_np.shape == (1, 1, 224, 224)
    clouds = mapper.infer_point_clouds(x, {"source": "rover"}, with_colors=True)
    want, valid = T.create_point_cloud(depth[0, 0].cpu().numpy())
    assert same(clouds[0][0].cpu().numpy(), want)
    assert same(clouds[0][1].cpu().numpy(), T.reconstruct_colors(x[0].permute(1, 2, 0).numpy(), valid))


@pytest.mark.parametrize("name", list(TEMPORAL_CASES))
def test_temporal_vs_reference_golden(name):
    maps, w = temporal_inputs(TEMPORAL_CASES[name])
    assert same(mde.DepthQualityEnhancement.temporal_consistency_filter(maps, w), GOLD[f"{name}__out"])


@pytest.mark.parametrize("name", list(BILATERAL_CASES))
def test_bilateral_vs_reference_golden(name):
    case = BILATERAL_CASES[name]
    depth = T.synth_depth(case["H"], case["W"], case["seed"], holes=False) * case.get("mul", 1.0)
    got = mde.DepthQualityEnhancement.bilateral_depth_filter(depth, None, **case.get("kw", {}))
    span = float(depth.max() - depth.min())
    assert got.dtype == np.float32 and got.shape == depth.shape
    assert float(np.abs(got - GOLD[f"{name}__out"]).max()) / span < 5e-6


@pytest.mark.parametrize("H,W", [(97, 131), (1024, 1024)])
def test_bilateral_vs_oracle_sizes_and_constant_map(H, W):
    depth = T.synth_depth(H, W, 77, holes=False)
    got = mde.DepthQualityEnhancement.bilateral_depth_filter(depth)
    want = T.bilateral_depth_filter(depth)
    assert float(np.abs(got - want).max()) / float(depth.max() - depth.min()) < 5e-6
    const = np.full((H, W), 2.25, np.float32)
    assert np.array_equal(mde.DepthQualityEnhancement.bilateral_depth_filter(const), const)


def test_joint_bilateral_vs_oracle():
    """cv2.ximgproc is not installed: this branch is checked against the oracle's restatement only (parity unpinned)"""
    H, W = 70, 90
    depth = T.synth_depth(H, W, 81, holes=False)
    rng = np.random.Generator(np.random.PCG64(82))
    rgb = (rng.integers(0, 4, (H, W, 3)) * 60).astype(np.uint8)  # few distinct colours: sigma_color = 0.05 only averages equal ones
    got = mde.DepthQualityEnhancement.bilateral_depth_filter(depth, rgb)
    want = T.bilateral_depth_filter(depth, rgb)
    assert float(np.abs(got - want).max()) / float(depth.max() - depth.min()) < 5e-6
    got2 = mde.DepthQualityEnhancement.bilateral_depth_filter(depth, rgb.astype(np.float32) / 255.0, sigma_color=40.0)
    want2 = T.bilateral_depth_filter(depth, rgb.astype(np.float32) / 255.0, sigma_color=40.0)
    assert float(np.abs(got2 - want2).max()) / float(depth.max() - depth.min()) < 5e-6
    assert float(np.abs(got2 - got).max()) > 1e-3  # the wide colour kernel really smooths across colours


def test_terrain_errors_are_loud():
    from monodepth_estimation_b200._lib import MdeError

    r = mde.TerrainReconstructor()
    with pytest.raises(ValueError):
        r.reconstruct_terrain(np.ones((8, 8), np.float32), np.zeros((4, 4, 3), np.uint8))  # the cv2.resize branch is not built
    with pytest.raises(ValueError):
        r.create_point_cloud(np.ones((2, 8, 8), np.float32))
    with pytest.raises(MdeError):
        mde.TerrainReconstructor(min_depth=5.0, max_depth=1.0).create_point_cloud(np.ones((8, 8), np.float32))
    with pytest.raises(ValueError):
        mde.DepthQualityEnhancement.bilateral_depth_filter(np.ones((8, 8), np.float32), np.zeros((8, 9, 3), np.uint8))


# --------------------------------------------------------------------------- fill_depth_holes / edge_preserving_smoothing
@pytest.mark.parametrize("name", list(HOLE_CASES))
def test_fill_depth_holes_vs_reference_golden(name):
    """DepthQualityEnhancement.fill_depth_holes on the device (union-find labelling, ring median, 3x3 median) against the vectors
    of the REAL reference method: bit-exact (only selections and one fp32 mean of two values).  Where a NaN reaches the 3x3 blur
    window cv2's result is not specified (holes_80x80_nanring): those pixels are left out, except an all-NaN window."""
    case = HOLE_CASES[name]
    depth = hole_inputs(case)
    got = mde.DepthQualityEnhancement.fill_depth_holes(depth, case["max_hole_size"])
    want = GOLD[f"{name}__out"]
    assert got.dtype == np.float32 and got.shape == want.shape
    with np.errstate(invalid="ignore"):
        pre = T.fill_depth_holes(depth, case["max_hole_size"], return_prefilter=True)[1]
    ok = T.nan_free_windows(pre)
    assert case.get("nan_ring") or ok.all()
    assert np.array_equal(np.where(ok, got, 0), np.where(ok, want, 0), equal_nan=True)
    if case.get("nan_ring"):  # the 3x3 zero blob next to the NaN blob got a NaN fill value (np.median semantics): its centre's window is all NaN
        y, x = case["H"] - 10, case["W"] - 12
        assert np.isnan(pre[y:y + 3, x + 4:x + 7]).all() and np.isnan(got[y + 1, x + 5]) and np.isnan(want[y + 1, x + 5])


@pytest.mark.parametrize("H,W,frac,mhs", [(257, 263, 0.02, 10), (48, 512, 0.3, 3), (33, 47, 0.0, 10), (40, 40, 1.0, 10), (1024, 1024, 0.0003, 10), (50, 60, 0.45, 13)])
def test_fill_depth_holes_vs_oracle_random(H, W, frac, mhs):
    """Random salt holes (many components, merged blobs, percolating masks at 45 %), no hole at all, an all-invalid map, the
    full 1024 px size: against the oracle (scipy + cv2, as the reference calls them)."""
    rng = np.random.Generator(np.random.PCG64(H * 7 + W))
    d = T.synth_depth(H, W, H + W, holes=False)
    r = rng.random((H, W))
    d[r < frac] = 0.0
    d[(r >= frac) & (r < frac * 1.2)] = -2.0
    want = T.fill_depth_holes(d, mhs)
    got = ops.fill_depth_holes(torch.from_numpy(d).cuda(), mhs).cpu().numpy()
    assert np.array_equal(got, want)


def test_fill_depth_holes_rejects_unbuilt_sizes():
    with pytest.raises(RuntimeError):
        ops.fill_depth_holes(torch.ones(8, 8).cuda(), 14)


@pytest.mark.parametrize("name", list(GUIDED_CASES))
def test_guided_filter_vs_oracle(name):
    """edge_preserving_smoothing against the restated guided filter (float64 NumPy; cv2.ximgproc is absent: parity unpinned).
    Tolerance 2e-5 of the map's range: the device keeps a, b as fp32 planes like the oracle and sums in double."""
    case = GUIDED_CASES[name]
    depth, rgb = guided_inputs(case)
    got = mde.DepthQualityEnhancement.edge_preserving_smoothing(depth, rgb, case["lam"])
    want = T.edge_preserving_smoothing(depth, rgb, case["lam"])
    assert got.dtype == np.float32 and got.shape == want.shape
    span = float(depth.max() - depth.min())
    assert float(np.abs(got - want).max()) <= 2e-5 * span, float(np.abs(got - want).max()) / span
    if rgb is not None:  # float [0,1] guide takes the (rgb * 255).astype(uint8) branch (refinement.py:268-269)
        got_f = mde.DepthQualityEnhancement.edge_preserving_smoothing(depth, (rgb / 255.0).astype(np.float32), case["lam"])
        want_f = T.edge_preserving_smoothing(depth, (rgb / 255.0).astype(np.float32), case["lam"])
        assert float(np.abs(got_f - want_f).max()) <= 2e-5 * span


def test_guided_filter_full_size_and_grey_guide():
    d = T.synth_depth(512, 640, 7, holes=False)
    want = T.edge_preserving_smoothing(d)
    got = mde.DepthQualityEnhancement.edge_preserving_smoothing(d)
    assert float(np.abs(got - want).max()) <= 2e-5 * float(d.max() - d.min())
    g = (np.random.Generator(np.random.PCG64(3)).integers(0, 256, d.shape)).astype(np.uint8)
    want = T.edge_preserving_smoothing(d, g)
    got = mde.DepthQualityEnhancement.edge_preserving_smoothing(d, g)
    assert float(np.abs(got - want).max()) <= 2e-5 * float(d.max() - d.min())
