"""Drop-in for the two GPU modules of ``models/refinement.py``.

``TextureAwareRefinementModule`` (``refinement.py:8-107``) and ``MultiScaleDepthFusion``
(``refinement.py:109-183``) with the reference's constructors and ``state_dict`` keys.
The reference's ``TextureAwareRefinementModule.forward`` raises on every valid input
(``refinement.py:88`` convolves a 1-channel map with a 3-channel Sobel weight; the
value it would build is never used), so this implementation computes the live
arithmetic, ``refinement.py:68-85`` and ``:92-107``.  ``DepthQualityEnhancement``
(``refinement.py:185-388``, OpenCV/NumPy CPU post-processing with no caller in the reference;
SURVEY.md §8(f)#4): all four filters run on the device (``csrc/terrain.cu``, ``csrc/depth_filters.cu``).
``edge_preserving_smoothing`` and the joint branch of ``bilateral_depth_filter`` call ``cv2.ximgproc`` in the
reference (opencv-contrib, absent from the build image): they follow the published algorithms, parity unpinned.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn


class TextureAwareRefinementModule(nn.Module):
    def __init__(self, input_channels=3, depth_channels=1, features=64):
        super().__init__()
        if (input_channels, depth_channels, features) != (3, 1, 64):
            raise ValueError("only the default configuration (3, 1, 64) is built")
        self.rgb_conv1 = nn.Conv2d(input_channels, features, kernel_size=3, padding=1)
        self.rgb_conv2 = nn.Conv2d(features, features, kernel_size=3, padding=1)
        self.depth_conv1 = nn.Conv2d(depth_channels, features, kernel_size=3, padding=1)
        self.depth_conv2 = nn.Conv2d(features, features, kernel_size=3, padding=1)
        self.combined_conv1 = nn.Conv2d(features * 2, features, kernel_size=3, padding=1)
        self.combined_conv2 = nn.Conv2d(features, features // 2, kernel_size=3, padding=1)
        self.output_conv = nn.Conv2d(features // 2, depth_channels, kernel_size=1)
        self.edge_detect_x = nn.Conv2d(input_channels, 1, kernel_size=3, padding=1, bias=False)
        self.edge_detect_y = nn.Conv2d(input_channels, 1, kernel_size=3, padding=1, bias=False)
        with torch.no_grad():  # Sobel / input_channels, frozen (refinement.py:40-53)
            gx = torch.tensor([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]])
            self.edge_detect_x.weight.copy_((gx / input_channels).expand(1, input_channels, 3, 3))
            self.edge_detect_y.weight.copy_((gx.t() / input_channels).expand(1, input_channels, 3, 3))
        self.edge_detect_x.requires_grad_(False)
        self.edge_detect_y.requires_grad_(False)
        self.activation = nn.ELU()

    @torch.no_grad()
    def forward(self, depth_map, rgb_image):
        from .engine import texture_refine_forward

        return texture_refine_forward(self, depth_map, rgb_image)


class MultiScaleDepthFusion(nn.Module):
    def __init__(self, scales=[0, 1, 2, 3], num_channels=1):
        super().__init__()
        if num_channels != 1:
            raise ValueError("only num_channels=1 is built")
        self.scales = scales
        self.num_scales = len(scales)
        self.confidence_conv = nn.ModuleList(
            [nn.Conv2d(num_channels, num_channels, kernel_size=3, padding=1) for _ in range(self.num_scales)]
        )
        self.fusion_conv = nn.Conv2d(num_channels * self.num_scales, num_channels, kernel_size=1)
        self.activation = nn.Sigmoid()

    @torch.no_grad()
    def forward(self, depth_scales):
        from .engine import depth_fusion_forward

        return depth_fusion_forward(self, depth_scales)


def _bilateral_tables(d, sigma_color, sigma_space, joint):
    """cv2's window: radius d/2 (or round(1.5 sigma_space) for d <= 0), clipped to a circle; weights in double, stored float32."""
    sigma_color = 1.0 if sigma_color <= 0 else float(sigma_color)
    sigma_space = 1.0 if sigma_space <= 0 else float(sigma_space)
    radius = max(int(d) // 2 if d > 0 else int(round(sigma_space * 1.5)), 1)
    if radius > 7:
        raise ValueError(f"bilateral window radius {radius} > 7 is not built (the reference uses d = 15)")
    i, j = np.mgrid[-7:8, -7:8].astype(np.float64)
    r = np.sqrt(i * i + j * j)
    space = np.where(r <= radius, np.exp(r * r * (-0.5 / (sigma_space * sigma_space))), 0.0).astype(np.float32)
    lut = None
    if joint:
        k = np.arange(3 * 256, dtype=np.float64)
        lut = np.exp(k * k * (-0.5 / (sigma_color * sigma_color))).astype(np.float32)
    return space, lut


class DepthQualityEnhancement:
    """``refinement.py:185-388``: NumPy in, NumPy out like the reference; the arithmetic runs in ``csrc/terrain.cu``."""

    @staticmethod
    def _device_map(depth_map):
        a = depth_map.detach().cpu().numpy() if isinstance(depth_map, torch.Tensor) else np.asarray(depth_map)
        a = np.squeeze(a).astype(np.float32)
        if a.ndim != 2:
            raise ValueError(f"expected one depth map, got shape {a.shape}")
        if not torch.cuda.is_available():
            raise RuntimeError("DepthQualityEnhancement needs a CUDA device (there is no CPU path)")
        return torch.from_numpy(np.ascontiguousarray(a)).cuda()

    @staticmethod
    def bilateral_depth_filter(depth_map, rgb_image=None, sigma_space=15, sigma_color=0.05, sigma_depth=0.05):
        """``refinement.py:190-243``: min/max normalise, bilateral (``rgb_image`` None) or joint bilateral filter, rescale."""
        from . import ops

        z = DepthQualityEnhancement._device_map(depth_map)
        guide = None
        if rgb_image is not None:
            rgb = np.squeeze(rgb_image.detach().cpu().numpy() if isinstance(rgb_image, torch.Tensor) else np.asarray(rgb_image))
            if rgb.dtype != np.uint8:
                rgb = (rgb * 255).astype(np.uint8)  # :215-217
            if rgb.shape != (*z.shape, 3):
                raise ValueError(f"guide image must be [H,W,3] matching the depth map, got {rgb.shape}")
            guide = torch.from_numpy(np.ascontiguousarray(rgb)).cuda()
        space, lut = _bilateral_tables(sigma_space, sigma_color if guide is not None else sigma_depth, sigma_space, guide is not None)
        out = ops.bilateral_depth(z, sigma_color if guide is not None else sigma_depth, torch.from_numpy(space).cuda(), guide,
                                  None if lut is None else torch.from_numpy(lut).cuda())
        return out.cpu().numpy()

    @staticmethod
    def edge_preserving_smoothing(depth_map, rgb_image=None, lambda_param=0.1):
        """``refinement.py:244-299``: min/max normalise, guided filter (guide = grey image, or the map itself), radius
        ``min(H, W) // 16``, rescale."""
        from . import ops

        z = DepthQualityEnhancement._device_map(depth_map)
        guide = None
        if rgb_image is not None:
            rgb = np.squeeze(rgb_image.detach().cpu().numpy() if isinstance(rgb_image, torch.Tensor) else np.asarray(rgb_image))
            if rgb.dtype != np.uint8:
                rgb = (rgb * 255).astype(np.uint8)  # :268-269
            if rgb.ndim == 2:  # :272-275: a one-channel guide is used as it is
                rgb = np.repeat(rgb[..., None], 3, axis=2)  # (9798 + 19235 + 3735) g + 16384 >> 15 == g: the grey conversion is the identity
            if rgb.shape != (*z.shape, 3):
                raise ValueError(f"guide image must be [H,W,3] or [H,W] matching the depth map, got {rgb.shape}")
            guide = torch.from_numpy(np.ascontiguousarray(rgb)).cuda()
        return ops.guided_filter_depth(z, min(z.shape) // 16, lambda_param, guide).cpu().numpy()

    @staticmethod
    def fill_depth_holes(depth_map, max_hole_size=10):
        """``refinement.py:301-352``: fill every 4-connected hole (non-finite or <= 0) of at most ``max_hole_size``^2 pixels with the
        median of its ring, then a 3x3 median over what was invalid."""
        from . import ops

        z = DepthQualityEnhancement._device_map(depth_map)
        return ops.fill_depth_holes(z, max_hole_size).cpu().numpy()

    @staticmethod
    def temporal_consistency_filter(depth_maps, weights=None):
        """``refinement.py:354-388``: weighted mean over frames, float64 like ``np.sum(stacked * weights[:, None, None], axis=0)``."""
        from . import ops

        if len(depth_maps) == 0:
            return None
        if len(depth_maps) == 1:
            return depth_maps[0]
        maps = [DepthQualityEnhancement._device_map(d) for d in depth_maps]
        if any(m.shape != maps[0].shape for m in maps):
            raise ValueError("all depth maps must have the same shape")
        if weights is None:
            w = np.ones(len(maps)) / len(maps)
        else:
            w = np.array(weights) / np.sum(weights)
        return ops.temporal_mean(maps, w.astype(np.float64).tolist()).cpu().numpy()
