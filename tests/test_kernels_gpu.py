"""GPU: every C-ABI kernel against a plain PyTorch fp32 statement of the same op, on seeded inputs.

The kernels compute in bf16 storage / fp32 accumulate, so tolerances are bf16-level
(2^-8 relative per rounding); each test states its own."""
import math

import pytest
import torch
import torch.nn.functional as F

from helpers import from_padded_nhwc, max_abs, rel_l2, to_padded_nhwc

pytestmark = pytest.mark.gpu

BF16 = torch.bfloat16


def _mods():
    from monodepth_estimation_b200 import _lib as L
    from monodepth_estimation_b200 import ops

    return L, ops


def _rand(*shape, seed=0, scale=1.0, dtype=torch.float32):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).cuda()


# --------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 256, 128), (1000, 768, 768), (197 * 3, 2304, 768), (333, 384, 256), (260, 768, 3072)])
def test_gemm_bias_f32(M, N, K):
    L, ops = _mods()
    a, w, b = _rand(M, K, seed=1, dtype=BF16), _rand(N, K, seed=2, scale=K ** -0.5, dtype=BF16), _rand(N, seed=3)
    out = torch.full((M, N), float("nan"), device="cuda")
    ops.gemm(a, w, b, L.EPI_BIAS_F32, out)
    ref = a.float() @ w.float().t() + b
    assert rel_l2(out, ref) < 1e-5, rel_l2(out, ref)


# M >= 1024 and N % 256 == 0 route to the CTA-pair kernel (tc_gemm2.cu); the shapes below hit full waves,
# the four-way split tail items, partial row pairs and the k=3072 pipeline wrap-around.
@pytest.mark.parametrize("M,N,K,epi", [(1576, 768, 768, "resid"), (1576, 2304, 768, "f32"), (6304, 768, 3072, "resid"), (6304, 3072, 768, "gelu"),
                                       (1024, 256, 64, "f32"), (6304, 2304, 768, "bf16"), (19000, 768, 768, "resid"), (1100, 256, 768, "f32")])
def test_gemm_pair_kernel(M, N, K, epi):
    L, ops = _mods()
    a, w, b = _rand(M, K, seed=101, dtype=BF16), _rand(N, K, seed=102, scale=K ** -0.5, dtype=BF16), _rand(N, seed=103)
    ref = a.float() @ w.float().t() + b
    if epi == "f32":
        out = torch.full((M, N), float("nan"), device="cuda")
        ops.gemm(a, w, b, L.EPI_BIAS_F32, out)
        assert rel_l2(out, ref) < 1e-5, rel_l2(out, ref)
    elif epi == "resid":
        r = _rand(M, N, seed=104)
        out = torch.full((M, N), float("nan"), device="cuda")
        ops.gemm(a, w, b, L.EPI_BIAS_RESID_F32, out, r)
        assert rel_l2(out, ref + r) < 1e-5, rel_l2(out, ref + r)
    else:
        out = torch.full((M, N), float("nan"), device="cuda", dtype=BF16)
        ops.gemm(a, w, b, L.EPI_BIAS_GELU_BF16 if epi == "gelu" else L.EPI_BIAS_BF16, out)
        want = F.gelu(ref) if epi == "gelu" else ref
        assert torch.isfinite(out.float()).all()
        assert rel_l2(out.float(), want) < 4e-3, rel_l2(out.float(), want)


def test_gemm_pair_kernel_qkv_split():
    L, ops = _mods()
    B = 8
    M = B * 197
    a, w, b = _rand(M, 768, seed=111, dtype=BF16), _rand(2304, 768, seed=112, scale=768 ** -0.5, dtype=BF16), _rand(2304, seed=113)
    q = torch.empty(B * 12 * 197 * 64, device="cuda", dtype=BF16)
    k = torch.empty_like(q)
    vt = torch.zeros(B * 12 * 64 * L.VT_PITCH, device="cuda", dtype=BF16)
    ops.gemm(a, w, b, L.EPI_QKV_SPLIT, q, k, vt)
    ref = (a.float() @ w.float().t() + b).view(B, 197, 3, 12, 64).permute(2, 0, 3, 1, 4)
    assert rel_l2(q.view(B, 197, 12, 64).permute(0, 2, 1, 3).float(), ref[0] * 0.125) < 4e-3  # q, k are token-major
    assert rel_l2(k.view(B, 197, 12, 64).permute(0, 2, 1, 3).float(), ref[1]) < 4e-3
    assert rel_l2(vt.view(B, 12, 64, L.VT_PITCH)[..., :197].float().transpose(-1, -2), ref[2]) < 4e-3


# --------------------------------------------------------------------------- LayerNorm folded into the neighbouring GEMMs
def _ln_slots(x, ops):
    """The statistics layout of mde_gemm_ln_bf16 (include/mde.h) for an fp32 matrix x [M][N]: [N/256][8][M][2].  A producer tile
    that ran whole leaves two partials (even / odd 32-column chunks), a tile of the split tail one partial per chunk."""
    M, N = x.shape
    nt_n = N // 256
    clusters = ops.sm_count() // 2
    tiles = (M + 255) // 256 * nt_n
    full = tiles // clusters * clusters
    ch = x.double().view(M, nt_n, 8, 32)
    part = torch.stack([ch.sum(-1), (ch * ch).sum(-1)], -1)  # [M][nt][8][2]
    st = torch.full((nt_n, 8, M, 2), float("nan"), dtype=torch.float64, device=x.device)
    for mt in range((M + 255) // 256):
        r = slice(mt * 256, min(M, mt * 256 + 256))
        for nt in range(nt_n):
            if mt * nt_n + nt < full:
                st[nt, 0, r] = part[r, nt, 0::2].sum(1)
                st[nt, 1, r] = part[r, nt, 1::2].sum(1)
            else:
                st[nt, :, r] = part[r, nt].transpose(0, 1)
    return st.view(nt_n * 8, M, 2)


@pytest.mark.parametrize("M,K,epi", [(6304, 768, "resid"), (1576, 3072, "resid"), (1100, 768, "f32"), (19 * 256, 768, "resid")])
def test_gemm_ln_producer(M, K, epi):
    """The producer epilogues also emit bf16(x) and per-row (sum, sum of squares) partials of the fp32 output."""
    L, ops = _mods()
    N = 768
    a, w, b = _rand(M, K, seed=201, dtype=BF16), _rand(N, K, seed=202, scale=K ** -0.5, dtype=BF16), _rand(N, seed=203)
    rows = M
    out = torch.full((rows, N), float("nan"), device="cuda")
    xb = torch.full((rows + 2, N), 5.0, device="cuda", dtype=BF16)
    st = torch.full((N // 32, rows, 2), float("nan"), device="cuda")  # [N/256][8][M][2]
    if epi == "resid":
        r = _rand(M, N, seed=206)
        ops.gemm_ln(a, w, b, L.EPI_BIAS_RESID_F32, out, r, xb_out=xb, stats_out=st)
        ref = torch.empty_like(out)
        ops.gemm(a, w, b, L.EPI_BIAS_RESID_F32, ref, r)
    else:
        ops.gemm_ln(a, w, b, L.EPI_BIAS_F32, out, xb_out=xb, stats_out=st)
        ref = torch.empty_like(out)
        ops.gemm(a, w, b, L.EPI_BIAS_F32, ref)
    assert torch.equal(out, ref)  # the fp32 output itself is unchanged
    assert torch.equal(xb[:rows], ref.to(BF16))
    assert float((xb[rows:].float() - 5.0).abs().max()) == 0.0  # rows beyond M untouched
    want = _ln_slots(ref, ops)
    used = ~torch.isnan(want)
    assert torch.equal(torch.isnan(st), ~used)  # exactly the documented slots are written
    assert max_abs(st[used[..., 0]][:, 0].double(), want[used[..., 0]][:, 0]) < 2e-4
    assert rel_l2(st[used[..., 1]][:, 1].double(), want[used[..., 1]][:, 1]) < 1e-6


def test_gemm_ln_patch_embed_is_an_error():
    L, ops = _mods()
    M = 8 * 196
    a, w, b = _rand(M, 768, seed=1, dtype=BF16), _rand(768, 768, seed=2, dtype=BF16), _rand(768, seed=3)
    out = torch.empty(8 * 197, 768, device="cuda")
    pos, cls = _rand(197, 768, seed=204), _rand(768, seed=205)
    with pytest.raises(RuntimeError, match="fp32 TMA epilogue"):
        ops.gemm_ln(a, w, b, L.EPI_PATCH_EMBED, out, pos, cls, xb_out=torch.empty(8 * 197, 768, device="cuda", dtype=BF16),
                    stats_out=torch.empty(24, 8 * 197, 2, device="cuda"))


@pytest.mark.parametrize("M,N,epi", [(6304, 3072, "gelu"), (1576, 2304, "qkv"), (1100, 768, "bf16")])
def test_gemm_ln_consumer(M, N, epi):
    """rstd * (bf16(x) (gamma.W)^T - mean * colsum) + (b + W beta)  ==  LN(x) W^T + b, with x carrying a per-row offset."""
    L, ops = _mods()
    from monodepth_estimation_b200.packing import fold_layernorm

    K = 768
    x = _rand(M, K, seed=211) * (0.5 + torch.rand(M, 1, device="cuda")) + 0.3 * _rand(M, 1, seed=212)
    w, b = _rand(N, K, seed=213, scale=K ** -0.5), _rand(N, seed=214)
    gamma, beta = 1.0 + 0.2 * _rand(K, seed=215), 0.1 * _rand(K, seed=216)
    wf, bf, cs = fold_layernorm(w, b, gamma, beta)
    st = _ln_slots(x, ops).float().contiguous()  # [K/256][8][M][2]; unused slots hold NaN and must not be read
    ref = F.layer_norm(x, (K,), gamma, beta, 1e-6) @ w.t() + b
    xb = x.to(BF16)
    if epi == "qkv":
        B = M // 197
        q = torch.empty(B * 12 * 197 * 64, device="cuda", dtype=BF16)
        k = torch.empty_like(q)
        vt = torch.zeros(B * 12 * 64 * L.VT_PITCH, device="cuda", dtype=BF16)
        ops.gemm_ln(xb, wf, bf, L.EPI_QKV_SPLIT, q, k, vt, stats_in=st, colsum=cs)
        r = ref.view(B, 197, 3, 12, 64).permute(2, 0, 3, 1, 4)
        assert rel_l2(q.view(B, 197, 12, 64).permute(0, 2, 1, 3).float(), r[0] * 0.125) < 5e-3
        assert rel_l2(k.view(B, 197, 12, 64).permute(0, 2, 1, 3).float(), r[1]) < 5e-3
        assert rel_l2(vt.view(B, 12, 64, L.VT_PITCH)[..., :197].float().transpose(-1, -2), r[2]) < 5e-3
        return
    out = torch.full((M, N), float("nan"), device="cuda", dtype=BF16)
    ops.gemm_ln(xb, wf, bf, L.EPI_BIAS_GELU_BF16 if epi == "gelu" else L.EPI_BIAS_BF16, out, stats_in=st, colsum=cs)
    want = F.gelu(ref) if epi == "gelu" else ref
    assert torch.isfinite(out.float()).all()
    assert rel_l2(out.float(), want) < 5e-3, rel_l2(out.float(), want)
    # against the unfused pair of kernels: same error level as the separate LayerNorm
    h = ops.layernorm(x, gamma, beta, 1e-6)
    out2 = torch.empty_like(out)
    ops.gemm(h, w.to(BF16).contiguous(), b, L.EPI_BIAS_GELU_BF16 if epi == "gelu" else L.EPI_BIAS_BF16, out2)
    assert rel_l2(out.float(), want) < 1.5 * rel_l2(out2.float(), want) + 1e-4


def test_gemm_ln_small_m_is_an_error():
    L, ops = _mods()
    a, w, b = _rand(512, 768, seed=1, dtype=BF16), _rand(768, 768, seed=2, dtype=BF16), _rand(768, seed=3)
    out = torch.empty(512, 768, device="cuda")
    with pytest.raises(RuntimeError, match="CTA-pair"):
        ops.gemm_ln(a, w, b, L.EPI_BIAS_F32, out, xb_out=torch.empty(512, 768, device="cuda", dtype=BF16), stats_out=torch.empty(24, 512, 2, device="cuda"))


def test_gemm_bias_bf16_and_gelu():
    L, ops = _mods()
    M, N, K = 500, 3072, 768
    a, w, b = _rand(M, K, seed=4, dtype=BF16), _rand(N, K, seed=5, scale=K ** -0.5, dtype=BF16), _rand(N, seed=6)
    ref = a.float() @ w.float().t() + b
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    ops.gemm(a, w, b, L.EPI_BIAS_BF16, out)
    assert rel_l2(out.float(), ref) < 4e-3
    ops.gemm(a, w, b, L.EPI_BIAS_GELU_BF16, out)
    assert rel_l2(out.float(), F.gelu(ref)) < 4e-3
    assert max_abs(out.float(), F.gelu(ref)) < 3e-2


def test_gemm_residual_f32_and_row_tail():
    L, ops = _mods()
    M, N, K = 197 * 2 + 5, 768, 768  # M not a multiple of 128: TMA zero-fills, epilogue masks
    a, w, b = _rand(M, K, seed=7, dtype=BF16), _rand(N, K, seed=8, scale=K ** -0.5, dtype=BF16), _rand(N, seed=9)
    r = _rand(M, N, seed=10)
    out = torch.full((M + 3, N), 7.0, device="cuda")
    ops.gemm(a, w, b, L.EPI_BIAS_RESID_F32, out, r, M=M)
    ref = a.float() @ w.float().t() + b + r
    assert rel_l2(out[:M], ref) < 1e-5
    assert float((out[M:] - 7.0).abs().max()) == 0.0  # rows beyond M untouched


@pytest.mark.parametrize("B", [3, 8, 32])  # 1-CTA kernel; CTA-pair kernel with the TMA epilogue (split tail items only; whole tiles + tail)
def test_gemm_patch_embed_epilogue(B):
    L, ops = _mods()
    a, w, b = _rand(B * 196, 768, seed=11, dtype=BF16), _rand(768, 768, seed=12, scale=768 ** -0.5, dtype=BF16), _rand(768, seed=13)
    pos, cls = _rand(197, 768, seed=14, scale=0.02), _rand(768, seed=15, scale=0.02)
    out = torch.full((B * 197 + 3, 768), 7.0, device="cuda")
    ops.gemm(a, w, b, L.EPI_PATCH_EMBED, out, pos, cls)
    tok = (a.float() @ w.float().t() + b).view(B, 196, 768)
    ref = torch.cat([cls.view(1, 1, 768).expand(B, -1, -1), tok], 1) + pos
    assert rel_l2(out[:B * 197].view(B, 197, 768), ref) < 1e-5
    assert max_abs(out[:B * 197].view(B, 197, 768), ref) < 1e-3  # every token row of every image (a missing row would keep its 7.0)
    assert float((out[B * 197:] - 7.0).abs().max()) == 0.0  # nothing past the last image


def test_gemm_qkv_split_epilogue():
    L, ops = _mods()
    B = 2
    M = B * 197
    a, w, b = _rand(M, 768, seed=16, dtype=BF16), _rand(2304, 768, seed=17, scale=768 ** -0.5, dtype=BF16), _rand(2304, seed=18)
    q = torch.empty(B * 12 * 197 * 64, device="cuda", dtype=BF16)
    k = torch.empty_like(q)
    vt = torch.zeros(B * 12 * 64 * L.VT_PITCH, device="cuda", dtype=BF16)
    ops.gemm(a, w, b, L.EPI_QKV_SPLIT, q, k, vt)
    ref = (a.float() @ w.float().t() + b).view(B, 197, 3, 12, 64).permute(2, 0, 3, 1, 4)  # [3,B,12,197,64]
    assert rel_l2(q.view(B, 197, 12, 64).permute(0, 2, 1, 3).float(), ref[0] * 0.125) < 4e-3  # q, k are token-major
    assert rel_l2(k.view(B, 197, 12, 64).permute(0, 2, 1, 3).float(), ref[1]) < 4e-3
    v = vt.view(B, 12, 64, L.VT_PITCH)[..., :197].float().transpose(-1, -2)
    assert rel_l2(v, ref[2]) < 4e-3


# --------------------------------------------------------------------------- LayerNorm / attention
def test_layernorm():
    _, ops = _mods()
    x, g, b = _rand(197 * 3 + 1, 768, seed=20, scale=2.0) + 0.3, _rand(768, seed=21) + 1.0, _rand(768, seed=22)
    out = ops.layernorm(x, g, b, 1e-6)
    ref = F.layer_norm(x, (768,), g, b, 1e-6)
    assert rel_l2(out.float(), ref) < 3e-3
    assert max_abs(out.float(), ref) < 5e-2


@pytest.mark.parametrize("B", [1, 3])
def test_vit_attention(B):
    L, ops = _mods()
    q, k, v = (_rand(B, 12, 197, 64, seed=30 + i, dtype=BF16) for i in range(3))
    qs = (q.float() * 0.125).to(BF16)
    vt = torch.zeros(B, 12, 64, L.VT_PITCH, device="cuda", dtype=BF16)
    vt[..., :197] = v.transpose(-1, -2)
    tm = lambda t: t.permute(0, 2, 1, 3).contiguous().view(-1)  # [B,12,197,64] -> token-major [B*197][768]
    out = ops.vit_attention(tm(qs), tm(k), vt.view(-1), B)
    att = torch.softmax(qs.float() @ k.float().transpose(-1, -2), dim=-1)
    ref = (att @ v.float()).transpose(1, 2).reshape(B * 197, 768)
    assert rel_l2(out.float(), ref) < 1e-2, rel_l2(out.float(), ref)
    assert max_abs(out.float(), ref) < 3e-2


# --------------------------------------------------------------------------- conv / upsample / norm
CONV_CASES = [
    # B, h, w, cin, cout, ksize, pad, act
    (2, 14, 15, 256, 256, 3, "reflect", "elu"),
    (2, 14, 15, 256, 128, 3, "zero", "elu"),
    (1, 28, 30, 256, 128, 3, "reflect", "elu"),
    (2, 28, 30, 128, 64, 3, "reflect", "elu"),
    (1, 56, 60, 128, 64, 3, "reflect", "elu"),
    (1, 56, 60, 64, 32, 3, "reflect", "elu"),
    (1, 112, 120, 64, 32, 3, "reflect", "none"),
    (1, 112, 120, 32, 16, 3, "reflect", "elu"),
    (1, 224, 240, 16, 16, 3, "reflect", "elu"),
    (2, 14, 14, 256, 64, 3, "reflect", "sigmoid"),
    (1, 14, 15, 320, 256, 3, "zero", "elu"),
    (2, 14, 15, 128, 16, 1, "zero", "none"),
    (40, 14, 15, 128, 64, 3, "reflect", "elu"),   # conv_small.cu: more tiles than SMs (persistent loop), two chunks
    (1, 44, 40, 64, 64, 3, "reflect", "elu"),     # conv_small.cu: several row tiles per image, one chunk
    (3, 30, 28, 192, 96, 3, "zero", "none"),      # conv_small.cu: 32-column tiles (cout % 64 != 0), three chunks
]


@pytest.mark.parametrize("B,h,w,cin,cout,ks,pad,act", CONV_CASES)
def test_conv_bf16_out(B, h, w, cin, cout, ks, pad, act):
    L, ops = _mods()
    from monodepth_estimation_b200.packing import pack_conv

    x = _rand(B, cin, h, w, seed=40)
    wt = _rand(cout, cin, ks, ks, seed=41, scale=(cin * ks * ks) ** -0.5)
    bs = _rand(cout, seed=42, scale=0.1)
    xp = to_padded_nhwc(x)
    wp, bp = pack_conv(wt, bs)
    out = torch.full((B, h + 2, w + 2, cout), float("nan"), device="cuda", dtype=BF16)
    ops.conv_nhwc(xp, h, w, cin, wp, bp, cout, ks, L.PAD_REFLECT if pad == "reflect" else L.PAD_ZERO,
                  {"none": L.ACT_NONE, "elu": L.ACT_ELU, "sigmoid": L.ACT_SIGMOID}[act], out_bf16=out)
    xr, wr = xp.float().permute(0, 3, 1, 2)[:, :, 1:-1, 1:-1], wp[:cout].float().view(cout, ks, ks, cin).permute(0, 3, 1, 2)
    if ks == 3:
        xr = F.pad(xr, (1, 1, 1, 1), mode="reflect") if pad == "reflect" else F.pad(xr, (1, 1, 1, 1))
    ref = F.conv2d(xr, wr, bs)
    ref = {"none": lambda t: t, "elu": F.elu, "sigmoid": torch.sigmoid}[act](ref)
    got = from_padded_nhwc(out, with_border=True)
    assert rel_l2(got[:, :, 1:-1, 1:-1], ref) < 5e-3, rel_l2(got[:, :, 1:-1, 1:-1], ref)
    # the border must hold the reflection of the interior
    assert max_abs(got, F.pad(got[:, :, 1:-1, 1:-1], (1, 1, 1, 1), mode="reflect")) == 0.0


@pytest.mark.parametrize("B,h,w,cin", [(2, 28, 30, 128), (1, 56, 60, 64), (1, 112, 120, 32), (1, 224, 240, 16), (3, 14, 15, 128)])
def test_conv_head_f32_out(B, h, w, cin):
    L, ops = _mods()
    from monodepth_estimation_b200.packing import pack_conv

    x = _rand(B, cin, h, w, seed=50)
    wt, bs = _rand(1, cin, 3, 3, seed=51, scale=(cin * 9) ** -0.5), _rand(1, seed=52)
    xp = to_padded_nhwc(x)
    wp, bp = pack_conv(wt, bs)
    out = torch.full((B, h, w), float("nan"), device="cuda")
    ops.conv_nhwc(xp, h, w, cin, wp, bp, 1, 3, L.PAD_ZERO, L.ACT_SIGMOID, out_f32=out)
    ref = torch.sigmoid(F.conv2d(xp.float().permute(0, 3, 1, 2)[:, :, 1:-1, 1:-1], wp[:1].float().view(1, 3, 3, cin).permute(0, 3, 1, 2), bs, padding=1))
    assert max_abs(out, ref[:, 0]) < 2e-5, max_abs(out, ref[:, 0])


@pytest.mark.parametrize("B,h,w", [(1, 112, 120), (2, 112, 112), (3, 20, 27), (1, 8, 16), (2, 9, 5)])
def test_decoder_tail16_vs_torch_and_unfused_path(B, h, w):
    """mde_decoder_tail16 (decoder.py:139-168, level 3: upsample x2, upconv_3_1, mars_adapt, dispconv_3 + sigmoid in one kernel)
    against (a) a plain fp32 torch statement with the same bf16 rounding points and (b) the four-launch path it replaces.
    Shapes: the real level (112x120 -> 224x240), the 14x14-grid level, sizes that are not tile multiples, maps smaller than a tile."""
    L, ops = _mods()
    from monodepth_estimation_b200.packing import pack_conv

    x = _rand(B, 16, h, w, seed=60)
    ws_, bs_ = [], []
    for k, cout in enumerate((16, 16, 1)):
        ws_.append(_rand(cout, 16, 3, 3, seed=61 + k, scale=(16 * 9) ** -0.5 * 1.5))
        bs_.append(_rand(cout, seed=64 + k, scale=0.1))
    packed = [pack_conv(wt, bs) for wt, bs in zip(ws_, bs_)]
    xp = to_padded_nhwc(x)
    out = torch.full((B, 2 * h, 2 * w), float("nan"), device="cuda")
    ops.decoder_tail16(xp, h, w, packed[0], packed[1], packed[2], out)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    # (a) torch fp32 with bf16-rounded operands / activations
    r = lambda t: t.to(BF16).float()
    wr = [p[0][:c].float().view(c, 3, 3, 16).permute(0, 3, 1, 2) for p, c in zip(packed, (16, 16, 1))]
    t = r(F.interpolate(r(x), size=(2 * h, 2 * w), mode="bilinear", align_corners=True))
    t = r(F.elu(F.conv2d(F.pad(t, (1, 1, 1, 1), mode="reflect"), wr[0], bs_[0])))
    t = r(F.elu(F.conv2d(F.pad(t, (1, 1, 1, 1), mode="reflect"), wr[1], bs_[1])))
    ref = torch.sigmoid(F.conv2d(t, wr[2], bs_[2], padding=1))[:, 0]
    assert max_abs(out, ref) < 4e-3, max_abs(out, ref)  # bf16 re-rounding of two intermediate maps (1 ulp flips), sigmoid slope <= 1/4
    assert rel_l2(out, ref) < 1.5e-3, rel_l2(out, ref)
    # (b) the unfused path: upsample2x_concat + three conv launches
    cat = torch.empty(B, 2 * h + 2, 2 * w + 2, 16, device="cuda", dtype=BF16)
    ops.upsample2x_concat(xp, h, w, None, 0, 0, cat)
    a1, a2 = torch.empty_like(cat), torch.empty_like(cat)
    ops.conv_nhwc(cat, 2 * h, 2 * w, 16, packed[0][0], packed[0][1], 16, 3, L.PAD_REFLECT, L.ACT_ELU, out_bf16=a1)
    ops.conv_nhwc(a1, 2 * h, 2 * w, 16, packed[1][0], packed[1][1], 16, 3, L.PAD_REFLECT, L.ACT_ELU, out_bf16=a2)
    old = torch.empty(B, 2 * h, 2 * w, device="cuda")
    ops.conv_nhwc(a2, 2 * h, 2 * w, 16, packed[2][0], packed[2][1], 1, 3, L.PAD_ZERO, L.ACT_SIGMOID, out_f32=old)
    assert max_abs(out, old) < 4e-3, max_abs(out, old)
    assert rel_l2(out, old) < 1.5e-3, rel_l2(out, old)


def test_conv_folded_batchnorm_and_channel_offset():
    L, ops = _mods()
    from monodepth_estimation_b200.packing import pack_conv

    B, h, w, cin, cout = 2, 14, 15, 256, 128
    x, wt, bs = _rand(B, cin, h, w, seed=60), _rand(cout, cin, 3, 3, seed=61, scale=(cin * 9) ** -0.5), _rand(cout, seed=62, scale=0.1)
    g, beta, mean, var = _rand(cout, seed=63) * 0.2 + 1, _rand(cout, seed=64, scale=0.1), _rand(cout, seed=65, scale=0.1), torch.rand(cout).cuda() + 0.5
    wp, bp = pack_conv(wt, bs, (g, beta, mean, var))
    xp = to_padded_nhwc(x)
    out = torch.zeros(B, h + 2, w + 2, 320, device="cuda", dtype=BF16)
    ops.conv_nhwc(xp, h, w, cin, wp, bp, cout, 3, L.PAD_REFLECT, L.ACT_ELU, out_bf16=out, out_coff=64)
    y = F.conv2d(F.pad(xp.float().permute(0, 3, 1, 2)[:, :, 1:-1, 1:-1], (1, 1, 1, 1), mode="reflect"), wt, bs)
    ref = F.elu(F.batch_norm(y, mean, var, g, beta, False, 0.0, 1e-5))
    got = from_padded_nhwc(out)
    assert rel_l2(got[:, 64:192], ref) < 8e-3
    assert float(got[:, :64].abs().max()) == 0.0 and float(got[:, 192:].abs().max()) == 0.0


@pytest.mark.parametrize("B,h,w,ca,cs", [(2, 14, 15, 128, 128), (1, 28, 30, 64, 64), (1, 112, 120, 16, 0), (1, 14, 14, 32, 32), (3, 56, 60, 32, 32),
                                           (1, 14, 15, 16, 8)])  # the last one (3 channel groups) takes the row-per-block kernel
def test_upsample_concat(B, h, w, ca, cs):
    _, ops = _mods()
    a = _rand(B, ca, h, w, seed=70)
    ap = to_padded_nhwc(a)
    sp, s = None, None
    if cs:
        s = _rand(B, cs, 14, 15, seed=71)
        sp = to_padded_nhwc(s)
    out = torch.full((B, 2 * h + 2, 2 * w + 2, ca + cs), float("nan"), device="cuda", dtype=BF16)
    ops.upsample2x_concat(ap, h, w, sp, 14, 15, out)
    ref = F.interpolate(ap.float().permute(0, 3, 1, 2)[:, :, 1:-1, 1:-1], size=(2 * h, 2 * w), mode="bilinear", align_corners=True)
    if cs:
        rs = F.interpolate(sp.float().permute(0, 3, 1, 2)[:, :, 1:-1, 1:-1], size=(2 * h, 2 * w), mode="bilinear", align_corners=True)
        ref = torch.cat([ref, rs], 1)
    got = from_padded_nhwc(out, with_border=True)
    assert rel_l2(got[:, :, 1:-1, 1:-1], ref) < 3e-3
    assert max_abs(got, F.pad(got[:, :, 1:-1, 1:-1], (1, 1, 1, 1), mode="reflect")) == 0.0


@pytest.mark.parametrize("gh,gw", [(14, 15), (14, 14)])
def test_tokens_to_grid_and_instnorm(gh, gw):
    _, ops = _mods()
    B, G = 2, gh * gw
    strip = G == 196
    tok = _rand(B, 197, 768, seed=80)
    grid = ops.tokens_to_grid(tok.view(-1, 768), B, strip, G)
    ref = tok[:, 1:] if strip else torch.cat([tok, tok.new_zeros(B, G - 197, 768)], 1)
    assert max_abs(grid.view(B, G, 768).float(), ref.to(BF16).float()) == 0.0
    x = _rand(B, G, 256, seed=81, scale=3.0) + 1.5
    outp = torch.zeros(B, gh + 2, gw + 2, 320, device="cuda", dtype=BF16)
    flat = torch.empty(B * G, 256, device="cuda", dtype=BF16)
    ops.instnorm_nhwc(x.view(-1, 256), B, gh, gw, 256, 1e-5, outp, 0, flat)
    refn = F.instance_norm(x.view(B, gh, gw, 256).permute(0, 3, 1, 2))
    got = from_padded_nhwc(outp, with_border=True)[:, :256]
    assert rel_l2(got[:, :, 1:-1, 1:-1], refn) < 3e-3
    assert max_abs(got, F.pad(got[:, :, 1:-1, 1:-1], (1, 1, 1, 1), mode="reflect")) == 0.0
    assert max_abs(flat.view(B, gh, gw, 256).permute(0, 3, 1, 2).float(), got[:, :, 1:-1, 1:-1]) == 0.0


# --------------------------------------------------------------------------- input / output stages
@pytest.mark.parametrize("kind,H,W", [("unit", 256, 320), ("byte", 224, 224), ("signed", 528, 528), ("unit", 1024, 1024)])
def test_input_range_and_patchify(kind, H, W):
    _, ops = _mods()
    from helpers import synth_images
    from oracle import depth_oracle as O

    x = synth_images(2, H, W, 5, kind).cuda()
    sd = {"encoder.normalize.weight": torch.tensor([1.0, 0.9, 1.1]).cuda(), "encoder.normalize.bias": torch.tensor([0.0, 0.01, -0.01]).cuda(),
          "encoder.normalize.running_mean": torch.tensor([0.0, 0.02, -0.02]).cuda(), "encoder.normalize.running_var": torch.tensor([1.0, 1.1, 0.9]).cuda()}
    if kind == "unit":  # default BN statistics: the common 2x-1 branch
        sd = {k: (torch.ones(3) if k.endswith(("weight", "running_var")) else torch.zeros(3)).cuda() for k in sd}
    bn = torch.cat([sd["encoder.normalize.weight"], sd["encoder.normalize.bias"], sd["encoder.normalize.running_mean"], sd["encoder.normalize.running_var"]])
    aff = ops.input_range(x, bn, 1e-5)
    y = O.input_range_map(sd, "encoder.", x)
    if (H, W) != (224, 224):
        y = F.interpolate(y, size=(224, 224), mode="bilinear", align_corners=False)
    ref = y.unfold(2, 16, 16).unfold(3, 16, 16).permute(0, 2, 3, 1, 4, 5).reshape(2 * 196, 768)
    got = ops.preprocess_patchify(x, aff)
    assert max_abs(got.float(), ref) < 1.2e-2 * max(1.0, float(ref.abs().max()))
    assert rel_l2(got.float(), ref) < 4e-3


@pytest.mark.parametrize("kind,H,W,B", [("unit", 256, 320, 2), ("byte", 224, 224, 2), ("signed", 528, 528, 3), ("unit", 1024, 1024, 2), ("byte", 100, 148, 1),
                                        ("unit", 1024, 1024, 33)])
def test_input_stage_fused(kind, H, W, B):
    """The one-pass input stage (scan + raw resample, finalize, patchify) against the two-pass kernels and the oracle."""
    L, ops = _mods()
    from helpers import synth_images

    x = synth_images(min(B, 3), H, W, 7, kind).repeat((B + 2) // 3, 1, 1, 1)[:B].contiguous().cuda()
    x[B - 1, 2, H - 1, W - 4:] = x.max() + 0.25  # extremes in the last rows / first row: every source row must be scanned once
    x[0, 0, 0, :4] = x.min() - 0.25
    bn = torch.tensor([1.0, 0.9, 1.1, 0.0, 0.01, -0.01, 0.0, 0.02, -0.02, 1.0, 1.1, 0.9], device="cuda")
    aff_ref = ops.input_range(x, bn, 1e-5)
    ref = ops.preprocess_patchify(x, aff_ref).float()
    scratch = torch.empty(L.RANGE_SCRATCH_FLOATS, device="cuda")
    aff = torch.empty(6, device="cuda")
    raw = torch.empty(B, 3, 224, 224, device="cuda")
    got = ops.input_stage(x, bn, 1e-5, scratch, aff, raw, torch.empty(B * 196, 768, device="cuda", dtype=BF16)).float()
    assert torch.equal(aff, aff_ref)
    want_raw = x if (H, W) == (224, 224) else F.interpolate(x, size=(224, 224), mode="bilinear", align_corners=False)
    assert max_abs(raw, want_raw) < 2e-5 * max(1.0, float(x.abs().max()))
    assert max_abs(got, ref) <= 1.6e-2 * max(1.0, float(ref.abs().max()))  # one bf16 ulp where the fp32 values straddle a rounding boundary
    assert rel_l2(got, ref) < 1e-3


@pytest.mark.parametrize("H,W", [(518, 518), (1024, 1024), (250, 333)])
def test_resize_maps_and_depth(H, W):
    _, ops = _mods()
    B = 2
    srcs = [torch.rand(B, 28 << i, 30 << i, device="cuda") for i in range(3)]
    disp = [torch.empty(B, 1, H, W, device="cuda") for _ in srcs]
    depth = [torch.empty(B, 1, H, W, device="cuda") for _ in srcs]
    ops.resize_maps([(s, 0.5, d, z) for s, d, z in zip(srcs, disp, depth)], B, H, W, 1 / 100.0, 1 / 0.1)
    for s, d, z in zip(srcs, disp, depth):
        ref = F.interpolate((s * 0.5).unsqueeze(1), size=(H, W), mode="bilinear", align_corners=False)
        assert max_abs(d, ref) < 2e-6
        assert rel_l2(z, 1 / (0.01 + 9.99 * ref)) < 1e-6


def test_resize_bilinear_nchw():
    _, ops = _mods()
    x = torch.rand(2, 3, 518, 518, device="cuda")
    out = ops.resize_bilinear_nchw(x, 528, 528)
    assert max_abs(out, F.interpolate(x, size=(528, 528), mode="bilinear", align_corners=False)) < 2e-6


# --------------------------------------------------------------------------- enhanced-model pieces
def test_spatial_attention():
    L, ops = _mods()
    from oracle import depth_oracle as O

    B, gh, gw = 2, 14, 15
    G = gh * gw
    x = _rand(B, 256, gh, gw, seed=90)
    sd = {"a.query.weight": _rand(32, 256, 1, 1, seed=91, scale=1 / 16), "a.query.bias": _rand(32, seed=92, scale=0.05),
          "a.key.weight": _rand(32, 256, 1, 1, seed=93, scale=1 / 16), "a.key.bias": _rand(32, seed=94, scale=0.05),
          "a.value.weight": _rand(256, 256, 1, 1, seed=95, scale=1 / 16), "a.value.bias": _rand(256, seed=96, scale=0.05),
          "a.gamma": torch.tensor([0.5]).cuda()}
    xb = x.to(BF16)
    ref = O.spatial_attention(sd, "a.", xb.float())
    flat = xb.permute(0, 2, 3, 1).reshape(B * G, 256).contiguous()
    wq = torch.cat([sd["a.query.weight"], sd["a.key.weight"], sd["a.value.weight"]]).view(320, 256)
    bq = torch.cat([sd["a.query.bias"], sd["a.key.bias"], sd["a.value.bias"]])
    from monodepth_estimation_b200.packing import pack_linear

    pw, pb = pack_linear(wq, bq)
    qkv = torch.empty(B * G, 384, device="cuda")
    ops.gemm(flat, pw, pb, L.EPI_BIAS_F32, qkv)
    out = torch.empty(B, gh + 2, gw + 2, 256, device="cuda", dtype=BF16)
    ops.spatial_attention(flat, qkv, B, gh, gw, sd["a.gamma"], out)
    got = from_padded_nhwc(out)
    assert rel_l2(got, ref) < 1e-2, rel_l2(got, ref)


def test_enh_fuse():
    _, ops = _mods()
    B, gh, gw = 2, 14, 15
    d = [torch.rand(B, (2 * gh) << i, (2 * gw) << i, device="cuda") for i in range(3)]
    refn = torch.randn(B, gh, gw, device="cuda") * 0.1
    fw = torch.tensor([0.3, 0.9, 0.5]).cuda()
    fused = torch.empty(B, 2 * gh, 2 * gw, device="cuda")
    ops.enh_fuse(d[0], d[1], d[2], refn, B, 2 * gh, 2 * gw, gh, gw, 0.7, fw, fused)
    r = F.interpolate(refn.unsqueeze(1), size=(2 * gh, 2 * gw), mode="bilinear", align_corners=False)
    s = [(d[0].unsqueeze(1) + r) * 0.7, d[1].unsqueeze(1) * 0.7, d[2].unsqueeze(1) * 0.7]
    wts = torch.softmax(fw, 0)
    ref = sum(wts[i] * (s[i] if i == 0 else F.interpolate(s[i], size=(2 * gh, 2 * gw), mode="bilinear", align_corners=False)) for i in range(3))
    assert max_abs(fused.unsqueeze(1), ref) < 2e-6


@pytest.mark.parametrize("M", [6304, 3152, 1576, 1024])
def test_fused_mlp_equals_the_two_gemms(M):
    """mde_mlp_bf16 (fc1 + GELU + fc2 + residual in one persistent launch, fc2 tiles released per finished 256-row block) against
    mde_gemm_bf16 x 2: the same MMAs in the same order and the same epilogue arithmetic, so the results are bit-identical; the
    row-block counters are zero again afterwards, also after back-to-back launches."""
    L, ops = _mods()
    D, Hd = 768, 3072
    a = _rand(M, D, seed=90).to(BF16)
    w1, b1 = _rand(Hd, D, seed=91, scale=D ** -0.5).to(BF16), _rand(Hd, seed=92, scale=0.1)
    w2, b2 = _rand(D, Hd, seed=93, scale=Hd ** -0.5).to(BF16), _rand(D, seed=94, scale=0.1)
    resid = _rand(M, D, seed=95)
    hid_ref = torch.empty(M, Hd, device="cuda", dtype=BF16)
    out_ref = torch.empty(M, D, device="cuda")
    ops.gemm(a, w1, b1, L.EPI_BIAS_GELU_BF16, hid_ref)
    ops.gemm(hid_ref, w2, b2, L.EPI_BIAS_RESID_F32, out_ref, resid)
    sync = torch.zeros(128, device="cuda", dtype=torch.int32)
    for rep in range(3):
        hid = torch.full((M, Hd), float("nan"), device="cuda", dtype=BF16)
        out = torch.full((M, D), float("nan"), device="cuda")
        ops.mlp(a, w1, b1, hid, w2, b2, resid, out, sync)
        torch.cuda.synchronize()
        assert int(sync.abs().sum()) == 0, rep
        assert torch.equal(hid, hid_ref), rep
        assert torch.equal(out, out_ref), (rep, rel_l2(out, out_ref))
    # and against torch (bf16 operands, fp32 accumulation, the exact-erf GELU the epilogue approximates to 2.5e-5)
    ref = resid + torch.nn.functional.gelu(a.float() @ w1.float().t() + b1).to(BF16).float() @ w2.float().t() + b2
    assert rel_l2(out, ref) < 2e-3, rel_l2(out, ref)
