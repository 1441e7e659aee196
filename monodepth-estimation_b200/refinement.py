"""Drop-in for the two GPU modules of ``models/refinement.py``.

``TextureAwareRefinementModule`` (``refinement.py:8-107``) and ``MultiScaleDepthFusion``
(``refinement.py:109-183``) with the reference's constructors and ``state_dict`` keys.
The reference's ``TextureAwareRefinementModule.forward`` raises on every valid input
(``refinement.py:88`` convolves a 1-channel map with a 3-channel Sobel weight; the
value it would build is never used), so this implementation computes the live
arithmetic, ``refinement.py:68-85`` and ``:92-107``.  ``DepthQualityEnhancement``
(OpenCV/NumPy CPU post-processing, ``refinement.py:185-388``) is out of scope.
"""
from __future__ import annotations

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
