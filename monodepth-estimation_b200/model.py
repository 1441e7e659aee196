"""Drop-in for ``models/model.py``: ``DepthAnything(pretrained, mars_finetuned)`` with
``forward(x, metadata=None) -> {("disp", 0..3): [B,1,H,W] fp32}``.

Same constructor, forward signature, sub-module names (``encoder``, ``decoder``) and
``state_dict`` keys as the reference (``model.py:18-32``); the computation runs on
hand-written sm_100a kernels through ``libmde.so`` (no PyTorch op on the hot path).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .decoder import DepthAnythingDecoder
from .encoder import DepthAnythingEncoder


def source_scale(metadata):
    """Source-specific disparity scale, ``model.py:63-87`` == ``enhanced_model.py:218-242``.

    Returns None when the reference applies no scaling at all (metadata missing or
    without ``'source'``), else the python float it multiplies every ("disp", k) by."""
    if metadata is None or "source" not in metadata:
        return None
    src = metadata["source"]
    if src == "rover":
        scale = 1.0
        if "height" in metadata:
            scale = 2.0 / max(metadata["height"], 0.1)
    elif src == "ingenuity":
        scale = 10.0 / max(metadata["altitude"], 1.0) if "altitude" in metadata else 0.5
    elif src == "satellite":
        scale = 0.01
    else:
        scale = 1.0
    return float(scale)


class _EngineMixin:
    """Shared plumbing: lazily (re)packs the fp32 parameters into kernel layouts."""

    _engine_kind = "standard"

    def _get_engine(self):
        from .engine import DepthEngine

        eng = self.__dict__.get("_engine")
        if eng is None:
            eng = DepthEngine(self, self._engine_kind)
            self.__dict__["_engine"] = eng
        return eng

    @torch.no_grad()
    def forward_keys(self, x, metadata=None, keys=(("disp", 0),), out_dtype=torch.float32, out_buffers=None):
        """The forward for callers that know which outputs they read (no counterpart in the reference, whose ``forward`` always
        returns every map; used by ``BatchPipeline`` / ``run_depth_estimation``, where only ``("disp", 0)`` and the uncertainty
        cross the host link).  Same values as ``forward(x, metadata)[key]`` for every requested key.

        ``x``: float32 ``[B,3,H,W]`` as in ``forward``, or the decoder's uint8 ``[B,H,W,3]`` image — a pixel then is
        ``float(u8) / 255`` exactly as ``run_mars_depth.py:100-103`` computes it, evaluated inside the input kernel (3 bytes read
        per pixel, no fp32 copy of the image).  ``out_dtype=torch.float16`` returns the maps rounded to IEEE fp16 by the output
        kernel itself; ``out_buffers`` ({key: tensor}) makes it write into caller-owned tensors."""
        if out_dtype not in (torch.float32, torch.float16):
            raise ValueError("out_dtype must be torch.float32 or torch.float16")
        if x.dtype == torch.uint8:
            if self.training:
                self._check_input(torch.empty(1, 3, 224, 224))  # raises the train-mode error
            if not (x.dim() == 4 and x.shape[-1] == 3):
                raise ValueError(f"expected a uint8 image batch [B,H,W,3], got {tuple(x.shape)}")
        else:
            self._check_input(x)
        return self._get_engine().forward(x, source_scale(metadata), None, tuple(keys), out_dtype == torch.float16, out_buffers)

    def _check_input(self, x):
        if self.training:
            raise RuntimeError(
                "model(x) is the inference forward (BatchNorm running statistics): call model.eval() first.  The training step "
                "(train/mars_adapter.py:117-153) runs through monodepth_estimation_b200.training.DepthTrainer(model).step(...)"
            )
        if not (isinstance(x, torch.Tensor) and x.dim() == 4 and x.shape[1] == 3):
            raise ValueError(f"expected an image batch [B,3,H,W], got {getattr(x, 'shape', type(x))}")
        if min(x.shape[2:]) < 160 and not self.input_size_warning_shown:
            # model.py:45-47
            print("Warning: Input image is very small. For best results, use images of at least 224x224 pixels.")
            self.input_size_warning_shown = True


class DepthAnything(_EngineMixin, nn.Module):
    def __init__(self, pretrained=True, mars_finetuned=False):
        super().__init__()
        self.encoder = DepthAnythingEncoder(pretrained=pretrained)
        self.decoder = DepthAnythingDecoder(
            num_ch_enc=self.encoder.num_ch_enc, scales=range(4), num_output_channels=1, use_skips=True
        )
        self.mars_finetuned = mars_finetuned
        self.input_size_warning_shown = False

    @torch.no_grad()
    def forward(self, x, metadata=None):
        self._check_input(x)
        return self._get_engine().forward(x, source_scale(metadata))
