"""CPU: the training-step oracle (oracle/train_oracle.py) pinned against the REAL reference classes in train() mode — runs only
where /root/reference exists (the build container).  Everything the device path is later compared with (losses, every parameter
gradient, BatchNorm running statistics after the step) is checked here first."""
import os
import sys

import pytest
import torch

from helpers import rel_l2
from monodepth_estimation_b200.synth import synth_images, synth_state_dict

REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sources only exist in the build container")
def test_train_oracle_matches_reference_train_mode():
    from oracle import train_oracle as T

    sys.path.insert(0, REF)
    try:
        from layers import get_smooth_loss
        from models.model import DepthAnything as RS
    finally:
        sys.path.remove(REF)
    torch.manual_seed(0)
    ref = RS(pretrained=False)
    sd = synth_state_dict(ref.state_dict(), 0)
    ref.load_state_dict(sd)
    ref.train()
    B, H, W = 2, 176, 240
    x = synth_images(B, H, W, 5)
    md = {"source": "rover", "height": 1.6}
    rnd = torch.rand(B, 1, H, W, generator=torch.Generator().manual_seed(9))
    # the reference step (train/mars_adapter.py:131-140,209-247) with the rand draw replaced by `rnd`
    out = ref(x, md)
    disp = out[("disp", 0)]
    smooth = 0.001 * get_smooth_loss(disp, x)
    depth = 1.0 * torch.abs(disp - disp.detach() * (0.9 + 0.2 * rnd)).mean()
    (depth + smooth).backward()

    losses, grads, buffers = T.loss_and_grads(sd, x, md, rnd)
    assert abs(losses["depth"] - float(depth)) <= 1e-6 * abs(float(depth))
    assert abs(losses["smoothness"] - float(smooth)) <= 1e-5 * abs(float(smooth))
    n_checked = 0
    for k, p in ref.named_parameters():
        if p.grad is None:
            assert k not in grads, k
            continue
        assert k in grads, k
        # biases in front of a normalisation (channel_reduce.*.bias, conv biases before BatchNorm) have an analytically zero
        # gradient: both sides hold rounding noise there, compared absolutely
        e = rel_l2(grads[k], p.grad)
        assert e < 2e-4 or float((grads[k] - p.grad).abs().max()) < 1e-7, (k, e)
        n_checked += 1
    assert n_checked > 150
    for k, b in ref.named_buffers():
        if k.endswith(("running_mean", "running_var")):
            assert rel_l2(buffers[k], b) < 1e-5, k
        elif k.endswith("num_batches_tracked"):
            assert int(buffers[k]) == int(b) == 1, k
    assert sum(k.endswith("num_batches_tracked") for k in buffers) == 14
