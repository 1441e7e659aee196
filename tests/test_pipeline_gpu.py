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


@pytest.mark.parametrize("kind,batch_size", [("standard", 3), ("enhanced", 2), ("standard", 1)])
def test_run_depth_estimation_matches_the_reference_loop(tmp_path, kind, batch_size):
    """runner.run_depth_estimation (decode pool -> buckets -> BatchPipeline) against the reference's one-image-at-a-time loop
    (run_mars_depth.py:90-147) written out with this package's model: same maps bit for bit, same files on disk."""
    import os

    import numpy as np
    from PIL import Image

    rng = np.random.Generator(np.random.PCG64(21))
    specs = [("rover_a.png", (256, 192)), ("rover_b.png", (256, 192)), ("aerial_c.png", (256, 192)), ("rover_d.png", (600, 400)), ("rover_e.png", (256, 192)),
             ("satellite_f.png", (224, 224)), ("rover_g.png", (256, 192)), ("rover_h.png", (256, 192))]
    paths = []
    for name, (w, h) in specs:
        p = os.path.join(str(tmp_path), name)
        Image.fromarray(rng.integers(0, 256, (h, w, 3), dtype=np.uint8)).save(p)
        paths.append(p)
    m = cached_model(kind, 0, "cuda")
    out_dir = os.path.join(str(tmp_path), "out")
    res = mde.run_depth_estimation(m, paths, out_dir, batch_size=batch_size, max_size=256, workers=4)
    assert [r["path"] for r in res] == paths
    for r, p in zip(res, paths):
        img = Image.open(p).convert("RGB")  # run_mars_depth.py:94-104
        if max(img.size) > 256:
            scale = 256 / max(img.size)
            img = img.resize((int(img.size[0] * scale), int(img.size[1] * scale)), Image.LANCZOS)
        x = (torch.from_numpy(np.array(img).transpose((2, 0, 1))).float() / 255.0).unsqueeze(0).cuda()
        src = "ingenuity" if "aerial" in p else ("satellite" if "satellite" in p else "rover")
        out = m(x, {"source": src})
        disp = out[("disp", 0)]
        depth = 1.0 / disp.clamp(min=1e-6)
        assert r["source"] == src
        assert np.array_equal(r["disp"], disp.cpu().squeeze().numpy())
        assert np.array_equal(r["depth"], depth.cpu().squeeze().numpy())
        if kind == "enhanced":
            assert np.array_equal(r["uncertainty"], out[("uncertainty", 0)].cpu().squeeze().numpy())
        else:
            assert r["uncertainty"] is None
        name = os.path.splitext(os.path.basename(p))[0]
        assert np.array_equal(np.load(os.path.join(out_dir, name, f"{name}_depth.npy")), r["depth"])


@pytest.mark.parametrize("B,H,W", [(3, 320, 256), (2, 1024, 1024), (1, 224, 228)])
def test_input_stage_u8_is_bit_identical_to_convert_then_fp32_stage(B, H, W):
    """mde_input_stage_u8 (the decoder's HWC bytes, pixel = float(u8)/255 inside the kernel) against mde_u8hwc_to_f32chw followed by
    mde_input_stage: the affine map of the range heuristic, the raw 224 px raster and the bf16 patch rows, bit for bit."""
    from monodepth_estimation_b200 import _lib as L

    g = torch.Generator().manual_seed(H + W)
    u8 = torch.randint(3, 250, (B, H, W, 3), dtype=torch.uint8, generator=g)
    u8[0, 5, 7, 1], u8[B - 1, H - 1, W - 1, 2] = 255, 0  # the extremes sit in single pixels
    u8 = u8.cuda()
    bn = torch.tensor([1.1, 0.9, 1.0, 0.01, -0.02, 0.0, 0.05, 0.0, -0.05, 1.2, 0.8, 1.0], device="cuda")
    outs = []
    for fused in (False, True):
        scratch = torch.empty(L.RANGE_SCRATCH_FLOATS, device="cuda")
        affine, raw = torch.empty(6, device="cuda"), torch.empty(B, 3, 224, 224, device="cuda")
        patches = torch.empty(B * 196, 768, device="cuda", dtype=torch.bfloat16)
        if fused:
            ops.input_stage_u8(u8, bn, 1e-5, scratch, affine, raw, patches)
        else:
            ops.input_stage(ops.u8hwc_to_f32chw(u8), bn, 1e-5, scratch, affine, raw, patches)
        outs.append((affine, raw, patches))
    for a, b in zip(*outs):
        assert torch.equal(a, b)


def test_forward_keys_subsets_dtypes_and_fallbacks():
    """model.forward_keys: a subset of the maps, uint8 input (also at a size that needs the pad-to-/16 resize, where the image is
    converted first), fp16 output written by the output kernel (disparities) or converted (depth), caller-owned buffers."""
    m = cached_model("enhanced", 0, "cuda")
    md = {"source": "ingenuity", "height": 12.0}
    for H, W in ((192, 256), (200, 250)):
        u8 = (synth_images(2, H, W, 77).permute(0, 2, 3, 1) * 255.0).round().to(torch.uint8).contiguous()
        ref = m((u8.float().permute(0, 3, 1, 2) / 255.0).contiguous().cuda(), md)
        keys = (("disp", 0), ("disp_fused", 0), ("uncertainty", 0), ("depth", 1), ("disp_unrefined", 0))
        got = m.forward_keys(u8.cuda(), md, keys)
        assert set(got) == set(keys)
        for k in keys:
            assert torch.equal(got[k], ref[k]), (H, W, k)
        buf = {("disp", 0): torch.empty(2, 1, H, W, device="cuda", dtype=torch.float16)}
        half = m.forward_keys(u8.cuda(), md, (("disp", 0), ("depth", 0)), torch.float16, buf)
        assert half[("disp", 0)] is buf[("disp", 0)]
        for k in half:
            assert half[k].dtype == torch.float16 and torch.equal(half[k], ref[k].to(torch.float16)), (H, W, k)
    with pytest.raises(KeyError):
        m.forward_keys(u8.cuda(), md, (("disp", 7),))
    with pytest.raises(ValueError):
        m.forward_keys(u8.cuda(), md, (("disp", 0),), torch.float16, {("disp", 0): torch.empty(2, 1, 8, 8, device="cuda", dtype=torch.float16)})
