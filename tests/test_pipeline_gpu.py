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
