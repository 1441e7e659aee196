"""CPU/torch restatement of the reference's training step — TEST INFRASTRUCTURE ONLY (imported by tests/ and bench legs that time
the CPU baseline; the product path never touches it).

Follows ``train/mars_adapter.py``: ``train_one_epoch`` (:117-153: ``model.train()``, forward, ``compute_losses``, ``zero_grad``,
``backward``, ``optimizer.step``), ``compute_losses`` (:209-247) with ``layers.get_smooth_loss`` (layers.py:202-215) and the Adam
set-up of ``__init__`` (:47-55: encoder parameters at lr/10, decoder parameters at lr).  The one deviation is explicit: the
``torch.rand_like`` draw of :238 is an argument (``rnd``), so that the device path can be fed the same numbers.

Pinned against the real classes (``models.model.DepthAnything`` in ``train()`` mode + the functions above, imported from
``/root/reference`` in the build container) by ``tests/test_train_oracle.py``.
"""
from __future__ import annotations

import torch

from . import depth_oracle as O


def smooth_loss(disp, img):
    """layers.py:202-215 get_smooth_loss."""
    gdx = torch.abs(disp[:, :, :, :-1] - disp[:, :, :, 1:])
    gdy = torch.abs(disp[:, :, :-1, :] - disp[:, :, 1:, :])
    gix = torch.mean(torch.abs(img[:, :, :, :-1] - img[:, :, :, 1:]), 1, keepdim=True)
    giy = torch.mean(torch.abs(img[:, :, :-1, :] - img[:, :, 1:, :]), 1, keepdim=True)
    return (gdx * torch.exp(-gix)).mean() + (gdy * torch.exp(-giy)).mean()


def compute_losses(images, outputs, rnd, w_depth=1.0, w_smooth=0.001):
    """train/mars_adapter.py:209-247 with the random target factor passed in (rnd ~ U[0,1), shape of disp)."""
    disp = outputs[("disp", 0)]
    smooth = w_smooth * smooth_loss(disp, images)
    target = disp.detach() * (0.9 + 0.2 * rnd)
    depth = w_depth * torch.abs(disp - target).mean()
    return {"depth": depth, "smoothness": smooth, "total": depth + smooth}


def loss_and_grads(sd, images, metadata, rnd, w_depth=1.0, w_smooth=0.001):
    """One forward in train mode + backward.  ``sd``: the model's state_dict (fp32); parameters are cloned into autograd leaves,
    buffers (BatchNorm running statistics) are cloned and updated as ``model.train()`` would.  Returns (losses as floats,
    {parameter name: gradient} for every parameter autograd reaches, the updated buffers)."""
    is_buffer = lambda k: k.endswith(("running_mean", "running_var", "num_batches_tracked"))
    leaves = {k: v.detach().clone().requires_grad_(not is_buffer(k) and v.is_floating_point()) for k, v in sd.items()}
    out = O.standard_forward(leaves, images, metadata, train=True)
    losses = compute_losses(images, out, rnd, w_depth, w_smooth)
    losses["total"].backward()
    grads = {k: v.grad for k, v in leaves.items() if v.requires_grad and v.grad is not None}
    buffers = {k: v.detach() for k, v in leaves.items() if is_buffer(k)}
    for k in buffers:  # nn.BatchNorm2d.forward in train mode: every BatchNorm of the model runs once per step (encoder.normalize + 13 ConvBlocks)
        if k.endswith("num_batches_tracked"):
            buffers[k] = buffers[k] + 1
    return {k: float(v) for k, v in losses.items()}, grads, buffers


def adam_reference(params, grads, state, lr_of, step, betas=(0.9, 0.999), eps=1e-8):
    """torch.optim.Adam's update (defaults of train/mars_adapter.py:52-55) applied per tensor; ``lr_of(name)`` gives the group's
    learning rate (encoder.* -> lr/10, decoder.* -> lr).  Parameters without a gradient are skipped, as Adam skips .grad None."""
    b1, b2 = betas
    for k, g in grads.items():
        m, v = state.setdefault(k, (torch.zeros_like(g), torch.zeros_like(g)))
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
        denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
        params[k].addcdiv_(m, denom, value=-lr_of(k) / bc1)
