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


# ---------------------------------------------------------------------------------------- parity ledger
# Every end-to-end comparison appends {case, key, metric, value, bar} here; tests/conftest.py writes the list to
# gpurun_out/parity_ledger.json at session end (the GPU box only brings gpurun_out/ back) and prints the worst entry.
# A committed copy per round lives under profiles/ (profiles/parity_r02.json).
LEDGER = []


def ledger_add(case, key, metric, value, bar, stage=None):
    LEDGER.append({"case": case, "key": "/".join(str(p) for p in key) if isinstance(key, tuple) else str(key), "metric": metric,
                   "value": float(value), "bar": float(bar), "frac_of_bar": float(value) / float(bar), **({"stage": stage} if stage else {})})


def check_outputs(got, ref, where, subsample=False):
    """All output maps of one forward against the oracle (or, subsample=True, a golden file's strided samples): the north_star bars,
    every number recorded in the ledger."""
    from oracle.cases import key_name, stride_for

    if not subsample:
        assert set(got) == set(ref), (set(got) ^ set(ref))
    worst = 0.0
    for k, v in got.items():
        if subsample:
            r = torch.from_numpy(ref[key_name(k)])
            sy, sx = stride_for(v.shape[2]), stride_for(v.shape[3])
            g = v[:, :, ::sy, ::sx].float().cpu()
        else:
            g, r = v.float().cpu(), ref[k]
        assert g.shape == r.shape, (k, g.shape, r.shape)
        assert torch.isfinite(g).all(), k
        if k[0] == "uncertainty":
            e = max_abs(g, r)
            ledger_add(where, k, "max_abs", e, UNCERTAINTY_MAX_ABS)
            assert e <= UNCERTAINTY_MAX_ABS, (where, k, e)
        else:
            e = rel_l2(g, r)
            ledger_add(where, k, "rel_l2", e, DEPTH_REL_L2)
            assert e <= DEPTH_REL_L2, (where, k, e)
        worst = max(worst, e)
        if k[0] == "disp":  # what the callers compute, run_mars_depth.py:137
            e = rel_l2(1 / g.clamp(min=1e-6), 1 / r.clamp(min=1e-6))
            ledger_add(where, (k[0] + "^-1", k[1]), "rel_l2", e, DEPTH_REL_L2)
            assert e <= DEPTH_REL_L2, (where, k, "1/disp", e)
    return worst
