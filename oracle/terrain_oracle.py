"""ORACLE — test infrastructure (never imported by the product).  CPU restatement of the callers on either side
of the forward path, NumPy / torch-CPU only:

* ``create_point_cloud`` / ``reconstruct_colors`` — ``utils/terrain_reconstruction_fixed.py:18-108``
* ``infer_depth_from_disp``                      — ``utils/terrain_reconstruction_fixed.py:168-171`` + ``layers.py:16-25``
* ``temporal_consistency_filter``                — ``models/refinement.py:354-388``
* ``bilateral_depth_filter``                     — ``models/refinement.py:190-243``; the plain branch restates
  OpenCV's ``bilateralFilter`` float32 algorithm (opencv-python 4.13.0 in the build image: circular window of radius d/2,
  4096-bin linearly interpolated colour table over the range of the input, BORDER_REFLECT_101, centre weight 1);
  the joint branch restates ``cv2.ximgproc.jointBilateralFilter`` (opencv-contrib, NOT installed here).

Pinning (``oracle/make_golden_terrain.py``, outputs in ``tests/golden/terrain_*.npz``): the point cloud, colour, depth and
temporal functions are bit-identical to the real reference classes run in the build container (NumPy 2.3); the plain
bilateral branch is pinned against the reference's ``DepthQualityEnhancement.bilateral_depth_filter`` (i.e. the installed
cv2) to 2e-6 of the normalised range.  **The joint bilateral branch is parity unpinned** (no ximgproc to run).
"""
import numpy as np
import torch


def default_intrinsics(h, w):
    f = 0.7 * w  # terrain_reconstruction_fixed.py:40-48
    return np.array([[f, 0, w / 2], [0, f, h / 2], [0, 0, 1]])


def create_point_cloud(depth_map, intrinsics=None, min_depth=0.1, max_depth=100.0):
    """terrain_reconstruction_fixed.py:18-70 -> (points [N,3], valid mask [H*W])"""
    depth_map = np.asarray(depth_map)
    if depth_map.ndim == 3:
        depth_map = depth_map.squeeze(0)
    h, w = depth_map.shape
    k = default_intrinsics(h, w) if intrinsics is None else intrinsics
    u, v = np.meshgrid(np.arange(w), np.arange(h))
    u = u.reshape(-1).astype(np.float32)
    v = v.reshape(-1).astype(np.float32)
    z = depth_map.reshape(-1).astype(np.float32)
    valid = (z > min_depth) & (z < max_depth)
    u, v, z = u[valid], v[valid], z[valid]
    fx, fy, cx, cy = k[0, 0], k[1, 1], k[0, 2], k[1, 2]
    x = (u - cx) * z / fx
    y = (v - cy) * z / fy
    return np.stack((x, y, z), axis=1), valid


def reconstruct_colors(rgb_hwc, valid):
    """terrain_reconstruction_fixed.py:99-104"""
    return np.asarray(rgb_hwc).reshape(-1, 3)[valid]


def infer_depth_from_disp(disp, min_depth=0.1, max_depth=100.0):
    """layers.py:16-25 on a float32 torch CPU tensor, as terrain_reconstruction_fixed.py:168-171 calls it"""
    min_disp = 1 / max_depth
    max_disp = 1 / min_depth
    scaled = min_disp + (max_disp - min_disp) * disp
    return 1 / scaled


def temporal_consistency_filter(depth_maps, weights=None):
    """models/refinement.py:354-388"""
    maps = [np.squeeze(np.asarray(d)) for d in depth_maps]
    w = np.ones(len(maps)) / len(maps) if weights is None else np.array(weights) / np.sum(weights)
    acc = maps[0] * w[0]
    for m, wi in zip(maps[1:], w[1:]):  # np.sum over the stacking axis adds frame by frame
        acc = acc + m * wi
    return acc


def _window(d, sigma_space):
    sigma_space = 1.0 if sigma_space <= 0 else float(sigma_space)
    radius = max(int(d) // 2 if d > 0 else int(round(sigma_space * 1.5)), 1)
    taps = []
    for i in range(-radius, radius + 1):
        for j in range(-radius, radius + 1):
            r = np.sqrt(float(i * i + j * j))
            if r > radius:
                continue
            taps.append((i, j, np.float32(np.exp(r * r * (-0.5 / (sigma_space * sigma_space))))))
    return radius, taps


def _bilateral_f32(src, d, sigma_color, sigma_space):
    """OpenCV bilateralFilter, CV_32FC1 branch"""
    src = src.astype(np.float32)
    sigma_color = 1.0 if sigma_color <= 0 else float(sigma_color)
    radius, taps = _window(d, sigma_space)
    lo, hi = float(src.min()), float(src.max())
    if abs(lo - hi) < np.finfo(np.float32).eps:
        return src.copy()
    bins = 1 << 12
    scale_index = np.float32(bins) / np.float32(hi - lo)
    idx = np.arange(bins + 2, dtype=np.float64)
    val = idx / np.float64(scale_index)
    lut = np.exp(val * val * (-0.5 / (sigma_color * sigma_color))).astype(np.float32)
    pad = np.pad(src, radius, mode="reflect")  # BORDER_REFLECT_101
    H, W = src.shape
    total = src.copy()
    wsum = np.ones_like(src)
    for i, j, sw in taps:
        if i == 0 and j == 0:
            continue
        nb = pad[radius + i: radius + i + H, radius + j: radius + j + W]
        alpha = np.abs(nb - src) * scale_index
        k = np.floor(alpha).astype(np.int32)
        alpha = alpha - k.astype(np.float32)
        wgt = sw * (lut[k] + alpha * (lut[k + 1] - lut[k]))
        total = total + nb * wgt
        wsum = wsum + wgt
    return (total / wsum).astype(np.float32)


def _joint_bilateral_u8guide(guide, src, d, sigma_color, sigma_space):
    """cv2.ximgproc.jointBilateralFilter, 8-bit 3-channel guide, float32 source (published algorithm; unpinned)"""
    src = src.astype(np.float32)
    sigma_color = 1.0 if sigma_color <= 0 else float(sigma_color)
    radius, taps = _window(d, sigma_space)
    k = np.arange(3 * 256, dtype=np.float64)
    lut = np.exp(k * k * (-0.5 / (sigma_color * sigma_color))).astype(np.float32)
    pad = np.pad(src, radius, mode="reflect")
    gpad = np.pad(guide.astype(np.int32), ((radius, radius), (radius, radius), (0, 0)), mode="reflect")
    H, W = src.shape
    g0 = guide.astype(np.int32)
    total = np.zeros_like(src)
    wsum = np.zeros_like(src)
    for i, j, sw in taps:
        nb = pad[radius + i: radius + i + H, radius + j: radius + j + W]
        gd = np.abs(gpad[radius + i: radius + i + H, radius + j: radius + j + W] - g0).sum(axis=2)
        wgt = sw * lut[gd]
        total = total + nb * wgt
        wsum = wsum + wgt
    return (total / wsum).astype(np.float32)


def bilateral_depth_filter(depth_map, rgb_image=None, sigma_space=15, sigma_color=0.05, sigma_depth=0.05):
    """models/refinement.py:190-243"""
    depth_map = np.squeeze(depth_map).astype(np.float32)
    depth_min = np.min(depth_map)
    depth_max = np.max(depth_map)
    normalized = (depth_map - depth_min) / (depth_max - depth_min + 1e-8)
    if rgb_image is not None:
        rgb = np.squeeze(rgb_image)
        if rgb.dtype != np.uint8:
            rgb = (rgb * 255).astype(np.uint8)
        filtered = _joint_bilateral_u8guide(rgb, normalized, sigma_space, sigma_color, sigma_space)
    else:
        filtered = _bilateral_f32(normalized, sigma_space, sigma_depth, sigma_space)
    return filtered * (depth_max - depth_min) + depth_min


def synth_depth(h, w, seed, holes=True):
    """Deterministic test depth map: smooth ramp + texture + a step edge, with out-of-range pixels (<= min_depth, >= max_depth,
    NaN) sprinkled in so the validity filter has something to reject."""
    rng = np.random.Generator(np.random.PCG64(seed))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    z = 2.0 + 0.02 * yy + 0.5 * np.sin(xx / 7.0) + rng.random((h, w), dtype=np.float32)
    z[:, w // 2:] += 3.0
    if holes:
        r = rng.random((h, w))
        z[r < 0.05] = 0.0
        z[(r >= 0.05) & (r < 0.08)] = 150.0
        z[(r >= 0.08) & (r < 0.09)] = np.nan
        z[(r >= 0.09) & (r < 0.10)] = 0.1  # == min_depth: rejected (strict compare)
    return z.astype(np.float32)
