"""Deterministic synthetic weights and inputs for parity tests and benchmarks.

There is no network for checkpoints or datasets, so every test / bench run uses
"random-init weights of that architecture" and synthetic images.  Values come from
NumPy's PCG64 stream (bit-reproducible across machines, unlike the order of
``torch.nn`` constructor RNG calls), keyed by the ``state_dict`` key, so the *same*
tensors can be loaded into the reference classes (when generating
``tests/golden``), into the oracle and into this package's classes.

The fill rules keep activations O(1) and — per SURVEY.md §0.9 — de-degenerate the
two code paths a default init hides: ``SelfAttentionBlock.gamma`` (0 at init) is set
to 0.5 and every BatchNorm except the 3-channel input ``normalize`` gets random
running statistics and affine terms.
"""
from __future__ import annotations

import zlib

import numpy as np
import torch


def _rng(seed: int, key: str) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64([seed, zlib.crc32(key.encode())]))


def synth_tensor(key: str, ref: torch.Tensor, seed: int) -> torch.Tensor:
    """Value for one ``state_dict`` entry, chosen from its key and shape."""
    shape = tuple(ref.shape)
    g = _rng(seed, key)
    leaf = key.rsplit(".", 1)[-1]
    if leaf == "num_batches_tracked":
        return torch.zeros(shape, dtype=ref.dtype)
    if "edge_detect" in key:  # frozen Sobel filters keep their constructor values (refinement.py:40-53)
        return ref.detach().clone().cpu()
    if ".normalize." in key or key.startswith("normalize."):  # input BN(3): default stats
        return ref.detach().clone().cpu()
    if leaf == "running_mean":
        v = g.normal(0.0, 0.1, shape)
    elif leaf == "running_var":
        v = g.uniform(0.5, 1.5, shape)
    elif leaf == "gamma":
        v = np.full(shape, 0.5)
    elif key.endswith("feature_fusion.weights"):
        v = g.uniform(0.0, 1.0, shape)
    elif leaf in ("class_token", "pos_embedding"):
        v = g.normal(0.0, 0.02, shape)
    elif leaf in ("bias", "in_proj_bias"):
        v = g.normal(0.0, 0.05, shape)
    elif len(shape) == 1:  # LayerNorm / BatchNorm scale
        v = g.uniform(0.5, 1.5, shape)
    else:  # conv / linear weight: N(0, 1/fan_in)
        fan_in = int(np.prod(shape[1:]))
        v = g.normal(0.0, 1.0 / np.sqrt(fan_in), shape)
    return torch.from_numpy(np.asarray(v, dtype=np.float32)).to(ref.dtype)


def synth_state_dict(template: dict, seed: int = 0) -> dict:
    """Synthetic replacement for every entry of ``template`` (a ``state_dict``)."""
    return {k: synth_tensor(k, v, seed) for k, v in template.items()}


def synth_images(batch: int, height: int, width: int, seed: int = 1234, kind: str = "unit") -> torch.Tensor:
    """Synthetic image batch ``[B,3,H,W]`` fp32.

    ``unit``: smooth blobs + noise in [0,1] (what callers feed: uint8/255,
    ``run_mars_depth.py:103``); ``byte``: the same scaled to [0,255] (exercises the
    ``/255`` branch, ``encoder.py:132-133``); ``signed``: shifted to [-0.5,0.5]
    (no ``2x-1``, ``encoder.py:134-135``)."""
    g = _rng(seed, f"img{batch}x{height}x{width}")
    yy = np.linspace(0.0, 1.0, height, dtype=np.float32)[None, None, :, None]
    xx = np.linspace(0.0, 1.0, width, dtype=np.float32)[None, None, None, :]
    ph = g.uniform(0.0, 6.28, (batch, 3, 1, 1)).astype(np.float32)
    fr = g.uniform(2.0, 9.0, (batch, 3, 1, 1)).astype(np.float32)
    img = 0.5 + 0.25 * np.sin(fr * 6.28 * yy + ph) * np.cos(fr * 4.1 * xx - ph)
    img = img + 0.25 * g.random((batch, 3, height, width), dtype=np.float32)
    img = np.clip(img, 0.0, 1.0).astype(np.float32)
    if kind == "byte":
        img = np.round(img * 255.0).astype(np.float32)
    elif kind == "signed":
        img = img - 0.5
    elif kind != "unit":
        raise ValueError(kind)
    return torch.from_numpy(img)


def bench_batch(batch: int, size: int, rank: int = 0) -> torch.Tensor:
    """The benchmark's input batch (bench.py and the parity tests at the benchmarked configurations share it): up to four
    distinct synthetic images per rank, repeated to `batch` — at 1024 px a batch of 32 is 403 MB, well past the 126 MB L2."""
    base = synth_images(min(batch, 4), size, size, 1234 + rank)
    return base.repeat((batch + base.shape[0] - 1) // base.shape[0], 1, 1, 1)[:batch].contiguous()
