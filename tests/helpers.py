"""Shared test helpers: error metrics, layout converters, model builders."""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

import monodepth_estimation_b200 as mde  # noqa: E402
from monodepth_estimation_b200.synth import synth_images, synth_state_dict  # noqa: E402

# north_star tolerances
DEPTH_REL_L2 = 1e-2
UNCERTAINTY_MAX_ABS = 1e-2


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def max_abs(a, b):
    return float((a.double() - b.double()).abs().max())


def to_padded_nhwc(x_nchw):
    """fp32 NCHW -> bf16 padded NHWC with reflected border."""
    return F.pad(x_nchw, (1, 1, 1, 1), mode="reflect").permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def from_padded_nhwc(t, with_border=False):
    t = t.float().permute(0, 3, 1, 2)
    return t if with_border else t[:, :, 1:-1, 1:-1]


def build_model(kind, seed=0, device="cpu"):
    cls = {"standard": mde.DepthAnything, "enhanced": mde.EnhancedDepthAnything}[kind]
    m = cls(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), seed))
    return m.eval().to(device)


_MODEL_CACHE = {}


def cached_model(kind, seed=0, device="cpu"):
    key = (kind, seed, str(device))
    if key not in _MODEL_CACHE:
        _MODEL_CACHE[key] = build_model(kind, seed, device)
    return _MODEL_CACHE[key]


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))
