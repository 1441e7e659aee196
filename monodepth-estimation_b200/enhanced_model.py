"""Drop-in for ``models/enhanced_model.py``: ``EnhancedDepthAnything`` with spatial
self-attention on encoder features 1/2, the uncertainty head, texture refinement and
3-scale softmax fusion (``enhanced_model.py:116-279``).

``forward(x, metadata=None)`` returns the reference's 11 maps, all ``[B,1,H,W]`` fp32:
``("disp",0..3)``, ``("uncertainty",0)``, ``("disp_unrefined",0)``, ``("disp_fused",0)``,
``("depth",0..3)``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .decoder import DepthAnythingDecoder
from .encoder import DepthAnythingEncoder
from .model import _EngineMixin, source_scale


class SelfAttentionBlock(nn.Module):
    """enhanced_model.py:11-19 (container): 1x1 q/k (C/8), 1x1 v (C), scalar gamma."""

    def __init__(self, in_channels):
        super().__init__()
        in_channels = int(in_channels)
        self.query = nn.Conv2d(in_channels, in_channels // 8, kernel_size=1)
        self.key = nn.Conv2d(in_channels, in_channels // 8, kernel_size=1)
        self.value = nn.Conv2d(in_channels, in_channels, kernel_size=1)
        self.gamma = nn.Parameter(torch.zeros(1))


class UncertaintyModule(nn.Module):
    """enhanced_model.py:41-48 (container)."""

    def __init__(self, in_channels):
        super().__init__()
        in_channels = int(in_channels)
        self.conv1 = nn.Conv2d(in_channels, in_channels // 2, kernel_size=3, padding=1)
        self.conv2 = nn.Conv2d(in_channels // 2, in_channels // 2, kernel_size=3, padding=1)
        self.conv3 = nn.Conv2d(in_channels // 2, 1, kernel_size=1)
        self.activation = nn.ELU()


class MultiScaleFeatureFusion(nn.Module):
    """enhanced_model.py:56-61 (container): one learnable weight per scale."""

    def __init__(self, scales=[0, 1, 2, 3]):
        super().__init__()
        self.scales = scales
        self.weights = nn.Parameter(torch.ones(len(scales)) / len(scales))


class TextureAwareRefinement(nn.Module):
    """enhanced_model.py:84-91 (container)."""

    def __init__(self, in_channels):
        super().__init__()
        in_channels = int(in_channels)
        self.conv1 = nn.Conv2d(in_channels + 3, in_channels, kernel_size=3, padding=1)
        self.conv2 = nn.Conv2d(in_channels, in_channels // 2, kernel_size=3, padding=1)
        self.conv3 = nn.Conv2d(in_channels // 2, 1, kernel_size=1)
        self.activation = nn.ELU()


class EnhancedDepthAnything(_EngineMixin, nn.Module):
    _engine_kind = "enhanced"

    def __init__(self, pretrained=True, mars_finetuned=False):
        super().__init__()
        self.encoder = DepthAnythingEncoder(pretrained=pretrained)
        self.decoder = DepthAnythingDecoder(
            num_ch_enc=self.encoder.num_ch_enc, scales=range(4), num_output_channels=1, use_skips=True
        )
        self.attention_blocks = nn.ModuleDict(
            {
                "1": SelfAttentionBlock(self.encoder.num_ch_enc[1]),
                "2": SelfAttentionBlock(self.encoder.num_ch_enc[2]),
            }
        )
        self.feature_fusion = MultiScaleFeatureFusion(scales=[0, 1, 2])
        self.uncertainty_module = UncertaintyModule(self.encoder.num_ch_enc[0])
        self.refinement = TextureAwareRefinement(self.encoder.num_ch_enc[0])
        self.mars_finetuned = mars_finetuned
        self.input_size_warning_shown = False

    @torch.no_grad()
    def forward(self, x, metadata=None):
        self._check_input(x)
        return self._get_engine().forward(x, source_scale(metadata))
