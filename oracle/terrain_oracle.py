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


def fill_depth_holes(depth_map, max_hole_size=10, return_prefilter=False):
    """models/refinement.py:301-352, line by line, with the reference's own third-party calls (scipy.ndimage.label /
    binary_dilation with their default cross-shaped structuring element, cv2.medianBlur ksize 3) — both are installed in the build
    image and on the GPU box.  Pinned bit-exactly against the reference method (oracle/make_golden_terrain.py)."""
    import cv2
    import scipy.ndimage as ndimage

    depth_map = np.squeeze(depth_map).astype(np.float32)
    invalid_mask = ~np.isfinite(depth_map) | (depth_map <= 0)
    filled = depth_map.copy()
    labeled, num = ndimage.label(invalid_mask)
    for i in range(1, num + 1):
        component = labeled == i
        if np.sum(component) > max_hole_size * max_hole_size:
            continue
        neighbor_mask = ndimage.binary_dilation(component, iterations=3) & ~component
        if np.any(neighbor_mask):
            filled[component] = np.median(depth_map[neighbor_mask])
    prefilter = filled.copy()
    final = cv2.medianBlur(filled.copy(), 3)
    filled[invalid_mask] = final[invalid_mask]
    return (filled, prefilter) if return_prefilter else filled


def nan_free_windows(prefilter):
    """Pixels whose 3x3 window (replicated border) of the pre-blur map holds no NaN: where cv2.medianBlur's result is defined
    (with a NaN in the window it depends on OpenCV's SIMD min/max operand order, which nothing specifies)."""
    H, W = prefilter.shape
    pad = np.pad(np.isnan(prefilter), 1, mode="edge")
    bad = np.zeros((H, W), bool)
    for dy in range(3):
        for dx in range(3):
            bad |= pad[dy:dy + H, dx:dx + W]
    return ~bad


def fill_depth_holes_numpy(depth_map, max_hole_size=10):
    """The same function WITHOUT scipy / cv2 (plain NumPy: flood-fill labelling, Manhattan-distance-3 ring, np.median, 3x3 median
    with replicated border): states what the two library calls compute, and is what the device kernels implement.  Checked against
    ``fill_depth_holes`` in tests/test_terrain_oracle.py."""
    d = np.squeeze(depth_map).astype(np.float32)
    H, W = d.shape
    invalid = ~np.isfinite(d) | (d <= 0)
    label = np.zeros((H, W), np.int32)
    comps = []
    for y in range(H):
        for x in range(W):
            if invalid[y, x] and label[y, x] == 0:
                comps.append([])
                stack = [(y, x)]
                label[y, x] = len(comps)
                while stack:
                    cy, cx = stack.pop()
                    comps[-1].append((cy, cx))
                    for ny, nx in ((cy - 1, cx), (cy + 1, cx), (cy, cx - 1), (cy, cx + 1)):
                        if 0 <= ny < H and 0 <= nx < W and invalid[ny, nx] and label[ny, nx] == 0:
                            label[ny, nx] = len(comps)
                            stack.append((ny, nx))
    filled = d.copy()
    for ci, px in enumerate(comps, 1):
        if len(px) > max_hole_size * max_hole_size:
            continue
        ring = set()
        for cy, cx in px:
            for dy in range(-3, 4):
                for dx in range(-(3 - abs(dy)), 3 - abs(dy) + 1):
                    ny, nx = cy + dy, cx + dx
                    if 0 <= ny < H and 0 <= nx < W and label[ny, nx] != ci:
                        ring.add((ny, nx))
        if ring:
            with np.errstate(invalid="ignore"):
                med = np.median(np.array([d[p] for p in sorted(ring)], dtype=np.float32))
            for cy, cx in px:
                filled[cy, cx] = med
    pad = np.pad(filled, 1, mode="edge")
    win = np.stack([pad[dy:dy + H, dx:dx + W] for dy in range(3) for dx in range(3)], axis=0)
    blur = np.sort(win, axis=0)[4]  # NaN-free windows only: callers compare where no NaN reaches the window
    out = filled.copy()
    out[invalid] = blur[invalid]
    return out


def rgb_to_gray_u8(rgb):
    """cv2.cvtColor(rgb, cv2.COLOR_RGB2GRAY) for uint8 (opencv-python 4.13): 15-bit fixed point, (9798 R + 19235 G + 3735 B + 16384) >> 15
    (checked against cv2 in tests/test_terrain_oracle.py)."""
    r = rgb.astype(np.int64)
    return ((9798 * r[..., 0] + 19235 * r[..., 1] + 3735 * r[..., 2] + 16384) >> 15).astype(np.uint8)


def _box_mean_reflect(a, r):
    """cv2.boxFilter(ksize 2r+1, normalised, BORDER_REFLECT) in float64."""
    if r == 0:
        return a.astype(np.float64)
    p = np.pad(a.astype(np.float64), r, mode="symmetric")  # numpy 'symmetric' == cv2.BORDER_REFLECT (fedcba|abcdef|fedcba)
    c = np.cumsum(np.cumsum(np.pad(p, ((1, 0), (1, 0))), axis=0), axis=1)
    k = 2 * r + 1
    H, W = a.shape
    s = c[k:k + H, k:k + W] - c[0:H, k:k + W] - c[k:k + H, 0:W] + c[0:H, 0:W]
    return s / float(k * k)


def guided_filter_gray(guide, src, radius, eps):
    """cv2.ximgproc.guidedFilter, one-channel guide (He, Sun, Tang: "Guided Image Filtering", box-filter form): opencv-contrib is
    NOT installed, so this restates the published algorithm in float64 — parity unpinned."""
    I, p = guide.astype(np.float64), src.astype(np.float64)
    mI, mp = _box_mean_reflect(I, radius), _box_mean_reflect(p, radius)
    var = _box_mean_reflect(I * I, radius) - mI * mI
    cov = _box_mean_reflect(I * p, radius) - mI * mp
    a = cov / (var + eps)
    b = mp - a * mI
    # the device keeps a and b as float32 planes between the two box passes, as OpenCV's CV_32F implementation does
    a, b = a.astype(np.float32), b.astype(np.float32)
    return (_box_mean_reflect(a, radius) * I + _box_mean_reflect(b, radius)).astype(np.float32)


def edge_preserving_smoothing(depth_map, rgb_image=None, lambda_param=0.1):
    """models/refinement.py:244-299 with guided_filter_gray in place of cv2.ximgproc.guidedFilter."""
    depth_map = np.squeeze(depth_map).astype(np.float32)
    depth_min = np.min(depth_map)
    depth_max = np.max(depth_map)
    normalized = (depth_map - depth_min) / (depth_max - depth_min + 1e-8)
    radius = min(depth_map.shape) // 16
    if rgb_image is not None:
        rgb = np.squeeze(rgb_image)
        if rgb.dtype != np.uint8:
            rgb = (rgb * 255).astype(np.uint8)
        guide = rgb_to_gray_u8(rgb) if rgb.ndim == 3 else rgb
        smoothed = guided_filter_gray(guide.astype(np.float32), normalized, radius, lambda_param)
    else:
        smoothed = guided_filter_gray(normalized, normalized, radius, lambda_param)
    return smoothed * (depth_max - depth_min) + depth_min


def synth_holes(h, w, seed, nan_small=True, nan_ring=False):
    """Depth map with holes of every kind fill_depth_holes distinguishes: single invalid pixels and small blobs (0, negative, +inf,
    and — nan_small — NaN), each isolated so that no other hole lies within its ring, two blobs larger than 10 x 10 (kept as holes,
    value 0 so that the 3x3 median around them stays NaN-free), a long thin blob, a blob in a corner and one on the bottom border.
    nan_ring adds a 3x3 zero blob two pixels away from a NaN blob: its ring then holds NaN and np.median makes the fill value NaN."""
    rng = np.random.Generator(np.random.PCG64(seed))
    z = synth_depth(h, w, seed, holes=False)
    special = [(h // 2, w // 2, 12, 11, 0.0),      # 132 px > 100: stays a hole
               (2, max(0, w - 30), 3, 26, -1.0),    # 78 px, long and thin: filled (at max_hole_size 10), wide window
               (0, 0, 2, 3, 0.0),                   # corner
               (h - 1, w // 3, 1, 4, 0.0)]          # bottom border
    taken = np.zeros((h, w), bool)
    for y, x, hh, ww, v in special:
        z[y:y + hh, x:x + ww] = v
        taken[max(0, y - 8):y + hh + 8, max(0, x - 8):x + ww + 8] = True
    vals = [0.0, -1.5, np.inf] + ([np.nan] if nan_small else [])
    k = 0
    for y in range(6, h - 12, 14):
        for x in range(6, w - 12, 14):
            hh, ww = int(rng.integers(1, 5)), int(rng.integers(1, 5))
            if taken[y:y + hh + 1, x:x + ww].any():
                continue
            z[y:y + hh, x:x + ww] = vals[k % len(vals)]
            if k % 5 == 0:
                z[y + hh, x] = vals[k % len(vals)]  # an L-shaped component
            k += 1
    if nan_ring:
        y, x = h - 10, w - 12
        z[y:y + 3, x:x + 2] = np.nan
        z[y:y + 3, x + 4:x + 7] = 0.0  # 3x3: the window of its centre pixel is all-NaN after the fill, whatever the blur's NaN order
    return z.astype(np.float32)


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
