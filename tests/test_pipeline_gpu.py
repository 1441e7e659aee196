"""GPU: the caller-side batching pipeline (pipeline.py) returns, batch for batch, exactly what a plain
``model(x.to(device))`` call returns, for float32 CHW and uint8 HWC host batches."""
import pytest
import torch

from helpers import cached_model, synth_images

import monodepth_estimation_b200 as mde
from monodepth_estimation_b200 import ops

pytestmark = pytest.mark.gpu


def test_u8hwc_to_f32chw_is_bit_exact():
    g = torch.Generator().manual_seed(3)
    u8 = torch.randint(0, 256, (3, 40, 52, 3), dtype=torch.uint8, generator=g)
    want = u8.float().permute(0, 3, 1, 2) / 255.0  # run_mars_depth.py:100-103
    got = ops.u8hwc_to_f32chw(u8.cuda())
    assert torch.equal(got.cpu(), want.contiguous())


@pytest.mark.parametrize("fmt", ["f32", "u8"])
def test_pipeline_matches_direct_calls(fmt):
    m = cached_model("enhanced", 0, "cuda")
    md = {"source": "rover"}
    batches = []
    for i in range(5):
        x = synth_images(2, 192, 256, 300 + i)
        if fmt == "u8":
            x = (x.permute(0, 2, 3, 1) * 255.0).round().to(torch.uint8).contiguous()
        batches.append(x.pin_memory())
    keys = (("disp", 0), ("uncertainty", 0), ("depth", 2))
    got = {}

    def consume(i, host):
        got[i] = {k: v.clone() for k, v in host.items()}

    pipe = mde.BatchPipeline(m, keys=keys, depth=2)
    assert pipe.run(batches, md, consume) == 5
    assert sorted(got) == list(range(5))
    for i, hb in enumerate(batches):
        x = hb
        if fmt == "u8":  # on the HOST, as run_mars_depth.py:100-103 does (torch's CUDA div-by-scalar multiplies by 1/255: not the same bits)
            x = (x.float().permute(0, 3, 1, 2) / 255.0).contiguous()
        ref = m(x.cuda(), md)
        for k in keys:
            assert torch.equal(got[i][k], ref[k].cpu()), (i, k)


def test_pipeline_rejects_bad_input():
    m = cached_model("standard", 0, "cuda")
    pipe = mde.BatchPipeline(m)
    with pytest.raises(ValueError):
        pipe.run([torch.zeros(1, 3, 64, 64, dtype=torch.float16).pin_memory()])
    with pytest.raises(ValueError):
        mde.BatchPipeline(m, depth=1)


def test_pipeline_fp16_output_is_the_rounded_fp32_map():
    """out_dtype=float16: the maps come back as round-to-nearest fp16 of exactly what model(x) returns (half the D2H bytes)."""
    m = cached_model("standard", 0, "cuda")
    md = {"source": "rover"}
    batches = [(synth_images(2, 192, 256, 310 + i).permute(0, 2, 3, 1) * 255.0).round().to(torch.uint8).contiguous().pin_memory() for i in range(4)]
    got = {}
    pipe = mde.BatchPipeline(m, keys=(("disp", 0), ("disp", 3)), depth=3, out_dtype=torch.float16)
    pipe.run(batches, md, lambda i, host: got.__setitem__(i, {k: v.clone() for k, v in host.items()}))
    for i, hb in enumerate(batches):
        ref = m((hb.float().permute(0, 3, 1, 2) / 255.0).contiguous().cuda(), md)
        for k in (("disp", 0), ("disp", 3)):
            assert got[i][k].dtype == torch.float16
            assert torch.equal(got[i][k], ref[k].cpu().to(torch.float16)), (i, k)


def test_f32_to_f16_ragged_and_saturating():
    x = torch.randn(100003, generator=torch.Generator().manual_seed(4)).cuda() * 10
    x[:4] = torch.tensor([1e6, -1e6, 65504.0, 1e-9]).cuda()
    got = ops.f32_to_f16(x)
    want = x.clamp(-65504, 65504).to(torch.float16)
    assert torch.equal(got, want)


def test_config4_composition_through_the_pipeline():
    """mde.EnhancedWithRefinement (BASELINE config 4 as one callable) under BatchPipeline == the three modules called by hand."""
    from helpers import synth_state_dict

    m = cached_model("enhanced", 0, "cuda")
    mk = lambda cls, seed: (lambda mod: (mod.load_state_dict(synth_state_dict(mod.state_dict(), seed)), mod.eval().cuda())[1])(cls())
    fu, rf = mk(mde.MultiScaleDepthFusion, 100), mk(mde.TextureAwareRefinementModule, 200)
    comp = mde.EnhancedWithRefinement(m, fu, rf)
    x = synth_images(2, 128, 192, 320)
    out = m(x.cuda(), None)
    want = rf(fu(out), x.cuda())
    got = {}
    mde.BatchPipeline(comp, keys=(("disp_refined", 0),)).run([x.pin_memory()], None, lambda i, h: got.update({k: v.clone() for k, v in h.items()}))
    assert torch.equal(got[("disp_refined", 0)], want.cpu())
