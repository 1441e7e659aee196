"""Host-side mirror of the reference encoder (``models/encoder.py``).

The classes here are *parameter containers*: they own fp32 ``nn.Parameter`` /
buffer tensors under exactly the module tree (and therefore exactly the
``state_dict`` keys) of the reference, and are built from the same ``torch.nn``
primitives in the same order as ``torchvision.models.vision_transformer`` so a
fixed ``torch.manual_seed`` yields the reference's random init.  None of the
``nn`` sub-modules is ever *called*: ``forward`` hands packed bf16 weights to
the sm_100a kernels in ``csrc/`` through the C-ABI (see ``engine.py``).

Reference: ``/root/reference/models/encoder.py:16-58`` (constructor) and
torchvision ``vision_transformer.py:83-107,160-252`` (ViT-B/16 layout).
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

VIT_HIDDEN = 768
VIT_HEADS = 12
VIT_MLP = 3072
VIT_LAYERS = 12
VIT_PATCH = 16
VIT_IMAGE = 224
VIT_SEQ = (VIT_IMAGE // VIT_PATCH) ** 2 + 1  # 197
HOOK_LAYERS = (2, 5, 8, 11)  # encoder.py:44


class _MLPBlock(nn.Sequential):
    """Key layout of torchvision's MLPBlock: ``0`` = fc1, ``3`` = fc2 (vision_transformer.py:38-54)."""

    def __init__(self, in_dim: int, mlp_dim: int):
        super().__init__(
            nn.Linear(in_dim, mlp_dim), nn.GELU(), nn.Dropout(0.0), nn.Linear(mlp_dim, in_dim), nn.Dropout(0.0)
        )
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    nn.init.normal_(m.bias, std=1e-6)


class _EncoderBlock(nn.Module):
    """vision_transformer.py:83-107 — container only."""

    def __init__(self):
        super().__init__()
        self.num_heads = VIT_HEADS
        self.ln_1 = nn.LayerNorm(VIT_HIDDEN, eps=1e-6)
        self.self_attention = nn.MultiheadAttention(VIT_HIDDEN, VIT_HEADS, dropout=0.0, batch_first=True)
        self.dropout = nn.Dropout(0.0)
        self.ln_2 = nn.LayerNorm(VIT_HIDDEN, eps=1e-6)
        self.mlp = _MLPBlock(VIT_HIDDEN, VIT_MLP)


class _Encoder(nn.Module):
    """vision_transformer.py:123-157 — ``pos_embedding``, ``layers.encoder_layer_i``, ``ln``."""

    def __init__(self):
        super().__init__()
        self.pos_embedding = nn.Parameter(torch.empty(1, VIT_SEQ, VIT_HIDDEN).normal_(std=0.02))
        self.dropout = nn.Dropout(0.0)
        layers = OrderedDict()
        for i in range(VIT_LAYERS):
            layers[f"encoder_layer_{i}"] = _EncoderBlock()
        self.layers = nn.Sequential(layers)
        self.ln = nn.LayerNorm(VIT_HIDDEN, eps=1e-6)  # output is discarded by the reference (encoder.py:38,142)


class _ViT(nn.Module):
    """vision_transformer.py:160-252 with ``heads`` replaced by Identity (encoder.py:38)."""

    def __init__(self):
        super().__init__()
        self.image_size = VIT_IMAGE
        self.patch_size = VIT_PATCH
        self.hidden_dim = VIT_HIDDEN
        self.conv_proj = nn.Conv2d(3, VIT_HIDDEN, kernel_size=VIT_PATCH, stride=VIT_PATCH)
        self.class_token = nn.Parameter(torch.zeros(1, 1, VIT_HIDDEN))
        self.encoder = _Encoder()
        # torchvision builds (and the reference then discards) a 768->1000 head; building
        # it here keeps the RNG stream aligned with the reference's constructor.
        nn.Linear(VIT_HIDDEN, 1000)
        self.heads = nn.Identity()
        fan_in = 3 * VIT_PATCH * VIT_PATCH
        nn.init.trunc_normal_(self.conv_proj.weight, std=math.sqrt(1 / fan_in))
        nn.init.zeros_(self.conv_proj.bias)


def _imagenet_vit_state_dict():
    """The ImageNet ViT-B/16 weights the reference loads at ``encoder.py:32-33`` (``vit_b_16(weights=ViT_B_16_Weights.IMAGENET1K_V1)``, ``encoder.py:32``),
    read from torchvision's LOCAL checkpoint cache only (``torch.hub`` directory) — this build never downloads.  None if absent."""
    import os

    try:
        from torchvision.models import ViT_B_16_Weights

        url = ViT_B_16_Weights.IMAGENET1K_V1.url
        path = os.path.join(torch.hub.get_dir(), "checkpoints", os.path.basename(url))
    except Exception:
        return None
    if not os.path.exists(path):
        return None
    return torch.load(path, map_location="cpu", weights_only=True)


class DepthAnythingEncoder(nn.Module):
    """Mirror of ``models/encoder.py:16`` — ViT-B/16 taps at blocks 2/5/8/11,
    1x1 768->256 channel reduction and instance norm.  The arithmetic lives in
    ``engine.DepthEngine``; this class owns the parameters."""

    def _load_imagenet_vit(self):
        """``pretrained=True`` (the reference's default, ``encoder.py:21,32``): copy the ImageNet weights into the mirror ViT.  The
        reference downloads them; here they must already sit in torchvision's local cache — otherwise fail loudly."""
        sd = _imagenet_vit_state_dict()
        if sd is None:
            raise RuntimeError(
                "pretrained=True needs the ImageNet ViT-B/16 checkpoint (vit_b_16-c867db91.pth) in torch.hub's checkpoint cache: the "
                "reference downloads it (encoder.py:32), this build never uses the network — construct with pretrained=False and load a state_dict"
            )
        own = self.encoder.state_dict()
        missing = [k for k in own if k not in sd]
        if missing:
            raise RuntimeError(f"cached ViT checkpoint lacks {missing[:3]}...")
        self.encoder.load_state_dict({k: sd[k] for k in own})  # the 'heads.*' entries are dropped (heads = Identity, encoder.py:38)

    def __init__(self, pretrained: bool = True):
        super().__init__()
        self.num_ch_enc = np.array([256, 256, 256, 256])
        self.encoder = _ViT()
        if pretrained:
            self._load_imagenet_vit()
        self.hook_layer_indices = list(HOOK_LAYERS)
        self.channel_reduce = nn.ModuleList([nn.Conv2d(VIT_HIDDEN, 256, 1) for _ in range(4)])
        self.normalize = nn.BatchNorm2d(3)
