"""BASELINE.json config 4 as one callable: enhanced model -> ``MultiScaleDepthFusion`` -> ``TextureAwareRefinementModule``.

The reference never composes these modules (``models/refinement.py`` has no caller); the composition is the one
stated once in ``oracle.depth_oracle.config4_forward``, following the module docstrings (``refinement.py:61-66,135-140``):
the enhanced forward's ``("disp", 0..3)`` at input size are fused, and the fused map is refined against the input image.
Two extra keys (names are ours) are added to the reference's 11: ``("disp_msfused", 0)`` and ``("disp_refined", 0)``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .enhanced_model import EnhancedDepthAnything
from .refinement import MultiScaleDepthFusion, TextureAwareRefinementModule


class EnhancedWithRefinement(nn.Module):
    """``forward(x, metadata=None)`` like the model classes, so ``BatchPipeline`` and ``bench.py --model config4`` can drive it."""

    def __init__(self, model: EnhancedDepthAnything | None = None, fusion: MultiScaleDepthFusion | None = None,
                 refine: TextureAwareRefinementModule | None = None):
        super().__init__()
        self.model = model if model is not None else EnhancedDepthAnything(pretrained=False)
        self.fusion = fusion if fusion is not None else MultiScaleDepthFusion()
        self.refine = refine if refine is not None else TextureAwareRefinementModule()

    @torch.no_grad()
    def forward(self, x, metadata=None):
        out = self.model(x, metadata)
        fused = self.fusion(out)
        out[("disp_msfused", 0)] = fused
        out[("disp_refined", 0)] = self.refine(fused, x)
        return out
