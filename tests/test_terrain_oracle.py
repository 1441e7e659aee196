"""CPU: oracle/terrain_oracle.py against the golden vectors produced by the REAL reference classes
(oracle/make_golden_terrain.py, run where /root/reference exists), plus the host-side argument logic of the mirror."""
import numpy as np
import pytest
import torch

from helpers import load_golden
from oracle import terrain_oracle as T
from oracle.terrain_cases import BILATERAL_CASES, PC_CASES, TEMPORAL_CASES, pc_inputs, temporal_inputs

GOLD = load_golden("terrain_golden")


def same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("name", list(PC_CASES))
def test_point_cloud_oracle_is_bit_identical_to_reference(name):
    case = PC_CASES[name]
    depth, k, rgb = pc_inputs(case)
    pts, valid = T.create_point_cloud(depth, k, case["min_depth"], case["max_depth"])
    assert same(pts, GOLD[f"{name}__points"])
    if rgb is not None:
        assert same(T.reconstruct_colors(rgb, valid), GOLD[f"{name}__colors"])


def test_infer_depth_oracle_is_bit_identical_to_reference():
    disp = torch.from_numpy(GOLD["infer_depth__disp"])
    assert same(T.infer_depth_from_disp(disp, 0.1, 100.0).numpy(), GOLD["infer_depth__0.1_100"])
    assert same(T.infer_depth_from_disp(disp, 0.5, 30.0).numpy(), GOLD["infer_depth__0.5_30"])


@pytest.mark.parametrize("name", list(TEMPORAL_CASES))
def test_temporal_oracle_is_bit_identical_to_reference(name):
    maps, w = temporal_inputs(TEMPORAL_CASES[name])
    assert same(T.temporal_consistency_filter(maps, w), GOLD[f"{name}__out"])


@pytest.mark.parametrize("name", list(BILATERAL_CASES))
def test_bilateral_oracle_matches_reference_cv2(name):
    case = BILATERAL_CASES[name]
    depth = T.synth_depth(case["H"], case["W"], case["seed"], holes=False) * case.get("mul", 1.0)
    got = T.bilateral_depth_filter(depth, None, **case.get("kw", {}))
    span = float(depth.max() - depth.min())
    assert float(np.abs(got - GOLD[f"{name}__out"]).max()) / span < 2e-6


def test_bilateral_constant_map_is_returned_unchanged():
    depth = np.full((20, 30), 3.5, np.float32)
    assert np.array_equal(T.bilateral_depth_filter(depth), depth)


def test_mirror_host_logic():
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.refinement import _bilateral_tables
    from monodepth_estimation_b200.terrain import _intrinsics4

    (fx, fy, cx, cy), f64 = _intrinsics4(None, 30, 40)
    assert (fx, fy, cx, cy, f64) == (0.7 * 40, 0.7 * 40, 20.0, 15.0, True)
    assert _intrinsics4(np.eye(3, dtype=np.float32), 3, 3)[1] is False
    assert _intrinsics4(np.eye(3, dtype=np.int64), 3, 3)[1] is True  # NumPy promotes float32 - int64 scalar to float64
    with pytest.raises(ValueError):
        _intrinsics4(np.eye(4), 3, 3)
    space, lut = _bilateral_tables(15, 0.05, 15, joint=True)
    assert space.shape == (15, 15) and space[7, 7] == 1.0 and space[0, 0] == 0.0 and space[0, 7] > 0 and lut.shape == (768,) and lut[0] == 1.0
    assert int((space > 0).sum()) == len(T._window(15, 15)[1])
    with pytest.raises(ValueError):
        _bilateral_tables(31, 0.05, 15, joint=False)
    if not torch.cuda.is_available():  # no CPU path: the filters fail loudly without a device
        with pytest.raises(RuntimeError):
            mde.DepthQualityEnhancement.fill_depth_holes(np.zeros((4, 4)))
        with pytest.raises(RuntimeError):
            mde.DepthQualityEnhancement.edge_preserving_smoothing(np.zeros((4, 4)))
    assert mde.DepthQualityEnhancement.temporal_consistency_filter([]) is None
    one = np.ones((3, 3), np.float32)
    assert mde.DepthQualityEnhancement.temporal_consistency_filter([one]) is one


@pytest.mark.parametrize("name", list(__import__("oracle.terrain_cases", fromlist=["HOLE_CASES"]).HOLE_CASES))
def test_fill_depth_holes_oracle_vs_reference_golden(name):
    """oracle.fill_depth_holes (the reference's lines with its own scipy / cv2 calls) and its library-free NumPy form against the
    vectors the REAL DepthQualityEnhancement.fill_depth_holes produced (oracle/make_golden_terrain.py): bit-exact."""
    from oracle.terrain_cases import HOLE_CASES, hole_inputs

    case = HOLE_CASES[name]
    depth = hole_inputs(case)
    want = GOLD[f"{name}__out"]
    with np.errstate(invalid="ignore"):
        got, pre = T.fill_depth_holes(depth, case["max_hole_size"], return_prefilter=True)
        got2 = T.fill_depth_holes_numpy(depth, case["max_hole_size"])
    assert np.array_equal(got, want, equal_nan=True)
    ok = T.nan_free_windows(pre)
    assert case.get("nan_ring") or ok.all()
    assert np.array_equal(np.where(ok, got2, 0), np.where(ok, want, 0), equal_nan=True)
    invalid = ~np.isfinite(depth) | (depth <= 0)
    assert np.array_equal(want[~invalid], depth[~invalid])  # valid pixels pass through


def test_guided_filter_oracle_properties():
    """cv2.ximgproc is absent (parity unpinned): the restated guided filter is checked through what the algorithm guarantees — a
    constant map is a fixed point, eps -> infinity gives the box mean, eps = 0 self-guided returns the input — and its grey
    conversion against the installed cv2."""
    import cv2
    from oracle.terrain_cases import GUIDED_CASES, guided_inputs

    rng = np.random.Generator(np.random.PCG64(5))
    p = rng.random((40, 56)).astype(np.float32)
    assert np.abs(T.guided_filter_gray(p, p, 3, 0.0) - p).max() < 1e-5
    mean = T._box_mean_reflect(p, 3)
    assert np.abs(T.guided_filter_gray(p, p, 3, 1e9) - T._box_mean_reflect(mean.astype(np.float32), 3)).max() < 1e-5
    assert np.abs(T._box_mean_reflect(p, 2) - cv2.boxFilter(p, -1, (5, 5), borderType=cv2.BORDER_REFLECT)).max() < 1e-6
    const = np.full((33, 47), 3.25, np.float32)
    assert np.abs(T.edge_preserving_smoothing(const) - 3.25).max() < 1e-6
    rgb = rng.integers(0, 256, (50, 70, 3), dtype=np.uint8)
    assert np.array_equal(T.rgb_to_gray_u8(rgb), cv2.cvtColor(rgb, cv2.COLOR_RGB2GRAY))
    gold = GOLD
    for name, case in GUIDED_CASES.items():
        depth, g = guided_inputs(case)
        out = T.edge_preserving_smoothing(depth, g, case["lam"])
        assert out.shape == depth.shape and out.dtype == np.float32
        assert np.abs(out - gold[f"{name}__oracle"]).max() < 1e-5 * float(depth.max() - depth.min())
