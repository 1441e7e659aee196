"""GPU: the drop-in model classes (C-ABI kernels underneath) against
 (a) the CPU fp32 oracle on the same synthetic weights/inputs, and
 (b) the committed golden vectors produced by the real reference.
Tolerances are the north_star's: depth rel-L2 <= 1e-2, uncertainty max-abs <= 1e-2 (bf16 storage,
fp32 accumulate, vs the fp32 reference)."""
import pytest
import torch

from helpers import DEPTH_REL_L2, UNCERTAINTY_MAX_ABS, cached_model, check_outputs, load_golden, max_abs, rel_l2, synth_images
from oracle import depth_oracle as O
from oracle.cases import CASES, key_name, stride_for

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["std_224_b2_none", "std_256x320_b1_ingalt", "std_256_b1_byte_other", "std_256_b1_signed_none",
                                  "std_160x200_b2_ing", "enh_224_b2_none", "enh_320x256_b1_sat", "std_518_b1_rover", "enh_518_b1_roverh"])
def test_model_vs_oracle(name):
    kind, B, H, W, ikind, md, wseed, iseed = CASES[name]
    x = synth_images(B, H, W, iseed, ikind)
    m_cpu = cached_model(kind, wseed)
    with torch.no_grad():
        fwd = O.standard_forward if kind == "standard" else O.enhanced_forward
        ref = fwd(m_cpu.state_dict(), x, md)
    m = cached_model(kind, wseed, "cuda")
    got = m(x.cuda(), md)
    check_outputs(got, ref, name)
    for v in got.values():
        assert v.shape == (B, 1, H, W) and v.dtype == torch.float32 and v.is_cuda


@pytest.mark.parametrize("name", ["std_1024_b1_rover", "enh_1024_b1_rover", "std_518_b1_rover", "enh_518_b1_roverh", "enh_224_b2_none"])
def test_model_vs_reference_golden(name):
    kind, B, H, W, ikind, md, wseed, iseed = CASES[name]
    gold = load_golden(name)
    m = cached_model(kind, wseed, "cuda")
    got = m(synth_images(B, H, W, iseed, ikind).cuda(), md)
    check_outputs(got, gold, "golden:" + name, subsample=True)


@pytest.mark.parametrize("kind", ["standard", "enhanced"])
def test_model_batch8_vs_oracle(kind):
    """B=8 -> M = 1576 GEMM rows: the ViT linear layers run on the CTA-pair kernel (tc_gemm2.cu)."""
    x = synth_images(8, 224, 256, 55)
    md = {"source": "ingenuity", "altitude": 3.0}
    with torch.no_grad():
        ref = (O.standard_forward if kind == "standard" else O.enhanced_forward)(cached_model(kind, 0).state_dict(), x, md)
    got = cached_model(kind, 0, "cuda")(x.cuda(), md)
    check_outputs(got, ref, f"{kind}_b8")


def test_batch_shard_invariance():
    """Images are independent: forward(B=4) == cat(forward(2), forward(2)) bit for bit."""
    m = cached_model("enhanced", 0, "cuda")
    x = synth_images(4, 256, 256, 77).cuda()
    md = {"source": "rover"}
    full = m(x, md)
    a, b = m(x[:2].contiguous(), md), m(x[2:].contiguous(), md)
    for k in full:
        assert torch.equal(full[k], torch.cat([a[k], b[k]])), k


def test_layernorm_fusion_matches_separate_kernels(monkeypatch):
    """MDE_LN_FUSION=1 (LayerNorm folded into the neighbouring GEMM epilogues, CTA-pair kernel, B >= 6) against the default path
    and the fp32 oracle; 12 + 8 images also covers the bit-exact shard invariance of the fused statistics."""
    import monodepth_estimation_b200.engine as E

    m = cached_model("standard", 0, "cuda")
    x = synth_images(12, 224, 224, 79)
    md = {"source": "rover"}
    eng = m._get_engine()
    base = {k: v.clone() for k, v in m(x.cuda(), md).items()}
    monkeypatch.setattr(E, "_LN_FUSION", True)
    eng._ws.clear()  # captured graphs hold the unfused launch sequence
    try:
        fused = {k: v.clone() for k, v in m(x.cuda(), md).items()}
        part = m(x[:8].contiguous().cuda(), md)
    finally:
        monkeypatch.setattr(E, "_LN_FUSION", False)
        eng._ws.clear()
    with torch.no_grad():
        ref = O.standard_forward(cached_model("standard", 0).state_dict(), x, md)
    for k in base:
        assert rel_l2(fused[k].float().cpu(), base[k].float().cpu()) < 8e-3, k  # two bf16 paths, each within 1e-2 of the oracle
        assert torch.equal(fused[k][:8], part[k]), k
    check_outputs(fused, ref, "ln_fusion")


def test_graph_and_eager_agree_and_repeat():
    import monodepth_estimation_b200.engine as E

    m = cached_model("standard", 0, "cuda")
    x = synth_images(2, 320, 320, 78).cuda()
    eng = m._get_engine()
    eng.use_graph = False
    eager = m(x, None)
    eng.use_graph = True
    g1 = m(x, None)
    g2 = m(x, None)  # graph replay
    for k in eager:
        assert torch.equal(eager[k], g1[k]) and torch.equal(g1[k], g2[k]), k
    assert eng.launches_per_forward > 100


def test_metadata_scales_and_reload():
    """Source scaling is applied after the sigmoid (model.py:89-92); load_state_dict re-packs the weights."""
    m = cached_model("standard", 0, "cuda")
    x = synth_images(1, 224, 224, 79).cuda()
    base = m(x, None)
    sat = m(x, {"source": "satellite"})
    assert rel_l2(sat[("disp", 0)], base[("disp", 0)] * 0.01) < 1e-6
    from helpers import build_model

    m2 = build_model("standard", 3, "cuda")
    o3 = m2(x, None)
    m2.load_state_dict(cached_model("standard", 0).state_dict())
    o0 = m2(x, None)
    assert torch.equal(o0[("disp", 0)], base[("disp", 0)]) and not torch.equal(o3[("disp", 0)], o0[("disp", 0)])


def test_errors_are_loud():
    from helpers import build_model

    m = cached_model("standard", 0, "cuda")
    with pytest.raises(RuntimeError):
        m(torch.rand(1, 3, 224, 224), None)  # CPU tensor: no fallback
    m_cpu = build_model("standard", 0)
    with pytest.raises(RuntimeError):
        m_cpu(torch.rand(1, 3, 224, 224), None)
    with pytest.raises(ValueError):
        m(torch.rand(1, 1, 224, 224).cuda(), None)
