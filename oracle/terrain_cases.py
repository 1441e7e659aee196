"""ORACLE — test infrastructure.  Case list shared by ``oracle/make_golden_terrain.py`` (runs the real reference) and the
tests (re-create the same inputs and compare against the stored results)."""
import numpy as np

from oracle.terrain_oracle import synth_depth, synth_holes

PC_CASES = {
    "pc_default_37x53":   dict(H=37, W=53, seed=51, min_depth=0.1, max_depth=100.0, K=None),
    "pc_f64K_64x96":      dict(H=64, W=96, seed=52, min_depth=0.1, max_depth=100.0, K="f64", as_tensor=True),
    "pc_f32K_48x80":      dict(H=48, W=80, seed=53, min_depth=0.5, max_depth=6.0, K="f32"),
    "pc_rgb_u8_40x72":    dict(H=40, W=72, seed=54, min_depth=0.1, max_depth=100.0, K=None, rgb="u8"),
    "pc_rgb_f32_33x47":   dict(H=33, W=47, seed=55, min_depth=0.1, max_depth=100.0, K="f32", rgb="f32"),
    "pc_none_valid_16x24": dict(H=16, W=24, seed=56, min_depth=50.0, max_depth=60.0, K=None),
}
TEMPORAL_CASES = {
    "temporal_3_weighted": dict(H=30, W=44, n=3, seed=61, weights=[1.0, 2.0, 3.0]),
    "temporal_5_equal":    dict(H=17, W=29, n=5, seed=62, weights=None),
}
BILATERAL_CASES = {
    "bilateral_64x80":   dict(H=64, W=80, seed=71),
    "bilateral_45x33_x10": dict(H=45, W=33, seed=72, mul=10.0, kw=dict(sigma_space=9, sigma_depth=0.1)),
}


HOLE_CASES = {
    "holes_96x128":        dict(H=96, W=128, seed=81, max_hole_size=10),
    "holes_70x90_max4":    dict(H=70, W=90, seed=82, max_hole_size=4),
    "holes_64x64_nonan":   dict(H=64, W=64, seed=83, max_hole_size=10, nan_small=False),
    "holes_80x80_nanring": dict(H=80, W=80, seed=84, max_hole_size=10, nan_ring=True),  # compared where the blur window is NaN-free
}
GUIDED_CASES = {
    "guided_self_80x112":  dict(H=80, W=112, seed=91, rgb=False, lam=0.1),
    "guided_rgb_64x96":    dict(H=64, W=96, seed=92, rgb=True, lam=0.1),
    "guided_self_48x48":   dict(H=48, W=48, seed=93, rgb=False, lam=0.01),
}


def hole_inputs(case):
    return synth_holes(case["H"], case["W"], case["seed"], case.get("nan_small", True), case.get("nan_ring", False))


def guided_inputs(case):
    depth = synth_depth(case["H"], case["W"], case["seed"], holes=False)
    rgb = None
    if case["rgb"]:
        rng = np.random.Generator(np.random.PCG64(case["seed"] + 1000))
        yy, xx = np.mgrid[0:case["H"], 0:case["W"]]
        base = (96 + 64 * np.sin(xx / 9.0) + 48 * (xx > case["W"] // 2))[..., None]
        rgb = np.clip(base + rng.integers(-20, 20, (case["H"], case["W"], 3)), 0, 255).astype(np.uint8)
    return depth, rgb


def pc_inputs(case):
    depth = synth_depth(case["H"], case["W"], case["seed"])
    k = None
    if case["K"] is not None:
        dt = np.float64 if case["K"] == "f64" else np.float32
        k = np.array([[0.83 * case["W"], 0, case["W"] / 2 - 3.25], [0, 0.79 * case["W"], case["H"] / 2 + 1.5], [0, 0, 1]], dtype=dt)
    rgb = None
    if case.get("rgb"):
        rng = np.random.Generator(np.random.PCG64(case["seed"] + 1000))
        rgb = rng.integers(0, 256, (case["H"], case["W"], 3), dtype=np.uint8)
        if case["rgb"] == "f32":
            rgb = (rgb / 255.0).astype(np.float32)
    return depth, k, rgb


def temporal_inputs(case):
    maps = [synth_depth(case["H"], case["W"], case["seed"] + i, holes=False) for i in range(case["n"])]
    return maps, case["weights"]
