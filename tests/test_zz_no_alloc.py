"""GPU: ``mde_input_stage_noalloc`` — the allocation-free variant of ``mde_input_stage`` (scratch in a device array built into
libmde.so; no counterpart in the reference) — called through the C-ABI and compared
with the existing entry point on two calls in a row.  Both entry points launch the same kernels with the same schedule, so the
comparison is bit-exact (stricter than the 1e-2 / 5e-3 bars of the other tests).  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from monodepth_estimation_b200 import _lib as L
from monodepth_estimation_b200.synth import synth_images

pytestmark = pytest.mark.gpu

F32, BF16 = torch.float32, torch.bfloat16


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _both(x, bn):
    """(affine6, raw224, patches) from the existing entry point and from the allocation-free one, same inputs."""
    B, _, H, W = x.shape
    outs = []
    for noalloc in (False, True):
        affine = torch.full((6,), float("nan"), device="cuda", dtype=F32)
        raw = torch.full((B, 3, 224, 224), float("nan"), device="cuda", dtype=F32)
        patches = torch.zeros(B * 196, 768, device="cuda", dtype=BF16)
        if noalloc:
            L.call("mde_input_stage_noalloc", _p(x), B, H, W, _p(bn), 1e-5, _p(affine), _p(raw), _p(patches), _stream())
        else:
            scratch = torch.empty(L.RANGE_SCRATCH_FLOATS, device="cuda", dtype=F32)
            L.call("mde_input_stage", _p(x), B, H, W, _p(bn), 1e-5, _p(scratch), _p(affine), _p(raw), _p(patches), _stream())
        torch.cuda.synchronize()
        outs.append((affine, raw, patches))
    return outs


def test_input_stage_noalloc_matches_input_stage_on_two_calls_in_a_row():
    bn = torch.tensor([1.0, 1.1, 0.9, 0.0, 0.01, -0.01, 0.0, 0.02, -0.02, 1.0, 1.2, 0.8], device="cuda", dtype=F32)
    # first call: the library's maximum batch, byte-range inputs (the /255 branch); second call: a smaller batch of another size and
    # range right after it — the shared array still holds the first call's records, none of which the second call may read
    cases = [(L.INPUT_STAGE_NOALLOC_MAX_BATCH, 256, 320, "byte", 401), (3, 518 + 10, 528, "unit", 402)]
    for B, H, W, kind, seed in cases:
        x = synth_images(B, H, W, seed, kind).cuda()
        (a0, r0, p0), (a1, r1, p1) = _both(x, bn)
        assert torch.isfinite(a1).all() and torch.isfinite(r1).all()
        assert torch.equal(a0, a1), (a0, a1)
        assert torch.equal(r0, r1)
        assert torch.equal(p0, p1)
    # sanity against a plain torch statement of encoder.py:129-144 for the last case (bf16 patch rows: 4e-3)
    a, b = a1[:3].view(1, 3, 1, 1), a1[3:].view(1, 3, 1, 1)
    ref = torch.nn.functional.interpolate(x, size=(224, 224), mode="bilinear", align_corners=False) * a + b
    rows = ref.view(B, 3, 14, 16, 14, 16).permute(0, 2, 4, 1, 3, 5).reshape(B * 196, 768)
    assert float((p1.float() - rows).norm() / rows.norm()) < 4e-3


def test_input_stage_noalloc_rejects_batches_beyond_its_scratch():
    B = L.INPUT_STAGE_NOALLOC_MAX_BATCH + 1
    dummy = torch.zeros(64, device="cuda", dtype=F32)  # argument checks come before any launch: nothing is read or written
    with pytest.raises(L.MdeError, match="exceeds the built-in scratch"):
        L.call("mde_input_stage_noalloc", _p(dummy), B, 224, 224, _p(dummy), 1e-5, _p(dummy), _p(dummy), _p(dummy), _stream())
    torch.cuda.synchronize()
