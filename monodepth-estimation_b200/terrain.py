"""Drop-in for the depth -> terrain callers of ``utils/terrain_reconstruction_fixed.py``.

``TerrainReconstructor.create_point_cloud`` / ``reconstruct_terrain`` (``:18-108``) and
``MarsTerrainMapper.infer_depth`` (``:132-179``) with the reference's signatures and return types
(NumPy arrays / a depth tensor), evaluated by ``csrc/terrain.cu`` on the device: one pass that
filters ``min_depth < z < max_depth``, back-projects and compacts in raster order.  The batched,
device-resident forms (``point_clouds``, ``infer_point_clouds``) skip the host round trips and the
intermediate depth map; mesh plotting (``:181+``, matplotlib / open3d) is out of scope.

Result dtype follows NumPy >= 2 (the version the oracle was pinned with): the reference's pixel grid and
depth are float32, its default intrinsics float64, so ``(u - cx) * z / fx`` and the stacked points are
float64; float32 intrinsics give float32 points.  Both are reproduced bit-exactly.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops


def _intrinsics4(intrinsics, h, w):
    """-> ((fx, fy, cx, cy) python floats, points are float64?)  terrain_reconstruction_fixed.py:39-48,61-62"""
    if intrinsics is None:
        f = 0.7 * w
        return (f, f, w / 2, h / 2), True
    if isinstance(intrinsics, torch.Tensor):
        intrinsics = intrinsics.detach().cpu().numpy()
    k = np.asarray(intrinsics)
    if k.shape != (3, 3):
        raise ValueError(f"intrinsics must be a 3x3 matrix, got shape {k.shape}")
    return (float(k[0, 0]), float(k[1, 1]), float(k[0, 2]), float(k[1, 2])), k.dtype != np.float32


def _as_device_maps(depth_map, device):
    """[H,W] / [1,H,W] / [B,1,H,W] / [B,H,W] array or tensor -> contiguous float32 CUDA tensor [B,H,W]"""
    t = depth_map if isinstance(depth_map, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(depth_map))
    t = t.detach()
    if t.dim() == 4:
        if t.shape[1] != 1:
            raise ValueError(f"expected depth maps [B,1,H,W], got {tuple(t.shape)}")
        t = t[:, 0]
    elif t.dim() == 2:
        t = t[None]
    elif t.dim() != 3:
        raise ValueError(f"expected a depth map [H,W], [1,H,W] or a batch [B,1,H,W], got {tuple(t.shape)}")
    return t.to(device=device, dtype=torch.float32).contiguous()


class TerrainReconstructor:
    """terrain_reconstruction_fixed.py:10-108"""

    def __init__(self, min_depth=0.1, max_depth=100.0, device="cuda"):
        self.min_depth = min_depth
        self.max_depth = max_depth
        self.device = torch.device(device)

    # ---- batched, device-resident form --------------------------------------------------------------------------
    def point_clouds(self, maps, intrinsics=None, rgb=None, from_disp=False):
        """maps: depth (or, with ``from_disp``, the model's disparity) [B,1,H,W] / [B,H,W] -> list of B
        ``(points [N_b,3], colors [N_b,3] or None)`` CUDA tensors (views into one buffer)."""
        z = _as_device_maps(maps, self.device)
        B, H, W = z.shape
        k4, f64 = _intrinsics4(intrinsics, H, W)
        if rgb is not None:
            rgb = rgb if isinstance(rgb, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(rgb))
            if rgb.dim() == 3:
                rgb = rgb[None]
            if rgb.shape[-1] != 3 and rgb.shape[1] == 3:
                rgb = rgb.permute(0, 2, 3, 1)
            if tuple(rgb.shape) != (B, H, W, 3):
                raise ValueError(f"rgb of shape {tuple(rgb.shape)} does not match depth maps {B}x{H}x{W} "
                                 "(the reference's cv2.resize branch is not built)")
            if rgb.dtype not in (torch.uint8, torch.float32):
                raise ValueError(f"rgb must be uint8 or float32, got {rgb.dtype}")
            rgb = rgb.to(self.device).contiguous()
        pts, cols, counts = ops.point_cloud(z, k4, self.min_depth, self.max_depth, from_disp=from_disp, f64=f64, rgb=rgb)
        n = counts.cpu().tolist()  # the one host sync: the sizes of the result
        return [(pts[b, : n[b]], None if cols is None else cols[b, : n[b]]) for b in range(B)]

    # ---- the reference's signatures --------------------------------------------------------------------------------
    def create_point_cloud(self, depth_map, intrinsics=None):
        """depth map [H,W] or [1,H,W] -> NumPy points [N,3] (``:18-70``)"""
        z = depth_map if isinstance(depth_map, torch.Tensor) else np.asarray(depth_map)
        if z.ndim == 3:
            z = z.squeeze(0)
        if z.ndim != 2:
            raise ValueError(f"expected a depth map [H,W] or [1,H,W], got {tuple(depth_map.shape)}")
        (pts, _), = self.point_clouds(z, intrinsics)
        return pts.cpu().numpy()

    def reconstruct_terrain(self, depth_map, rgb_image=None, mask=None, intrinsics=None):
        """-> (points [N,3], colors [N,3] or None) as NumPy arrays (``:72-108``; ``mask`` is unused there too)"""
        if rgb_image is None:
            return self.create_point_cloud(depth_map, intrinsics), None
        z = depth_map if isinstance(depth_map, torch.Tensor) else np.asarray(depth_map)
        if z.ndim == 3:
            z = z.squeeze(0)
        rgb = rgb_image.detach().cpu().numpy() if isinstance(rgb_image, torch.Tensor) else np.asarray(rgb_image)
        if rgb.shape[0] == 3:
            rgb = np.moveaxis(rgb, 0, -1)
        (pts, cols), = self.point_clouds(z, intrinsics, rgb=np.ascontiguousarray(rgb))
        return pts.cpu().numpy(), cols.cpu().numpy()


class MarsTerrainMapper:
    """terrain_reconstruction_fixed.py:110-179 (``infer_depth``); the plotting / mesh methods are out of scope."""

    def __init__(self, model, min_depth=0.1, max_depth=100.0, use_cuda=True):
        if not (use_cuda and torch.cuda.is_available()):
            raise RuntimeError("MarsTerrainMapper needs a CUDA device (there is no CPU path)")
        self.model = model
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.model.to(self.device)
        self.reconstructor = TerrainReconstructor(min_depth=min_depth, max_depth=max_depth, device=self.device)

    def _prepare(self, image):
        if isinstance(image, torch.Tensor):
            return image.to(self.device)
        image = torch.from_numpy(image).float().to(self.device)  # :152-163
        if image.dim() == 3:
            image = image.unsqueeze(0)
        if image.max() > 1.0:
            image = image / 255.0
        return image

    @torch.no_grad()
    def infer_depth(self, image, metadata=None):
        """image [B,3,H,W] (tensor, or NumPy with the reference's /255 heuristic) -> depth [B,1,H,W] (``:132-179``).
        The model's outputs already have the input size, so the reference's final resize (``:174-177``) never runs."""
        self.model.eval()
        image = self._prepare(image)
        disp = self.model(image.contiguous(), metadata)[("disp", 0)]
        return ops.disp_to_depth(disp.contiguous(), self.reconstructor.min_depth, self.reconstructor.max_depth)

    @torch.no_grad()
    def infer_point_clouds(self, image, metadata=None, intrinsics=None, with_colors=False):
        """``infer_depth`` + ``create_point_cloud`` per image without materialising the depth maps: the disparity maps
        go straight into the point-cloud kernel.  -> list of B (points, colors) CUDA tensors."""
        self.model.eval()
        image = self._prepare(image)
        disp = self.model(image.contiguous(), metadata)[("disp", 0)]
        rgb = image.permute(0, 2, 3, 1).contiguous().float() if with_colors else None
        return self.reconstructor.point_clouds(disp, intrinsics, rgb=rgb, from_disp=True)
