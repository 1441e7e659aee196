"""fp32 ``state_dict`` -> kernel-ready weights.

Everything the kernels read is derived here once (and again whenever a parameter changes):
bf16 K-major GEMM weights, conv weights re-laid as ``[cout_pad][(ky*3+kx)*cin + c]`` with the
eval-BatchNorm scale folded in (fp32 fold, then one bf16 rounding), fp32 biases with the BN
shift folded in, and the small fp32 vectors (LayerNorm, positional embedding, ...).
All arithmetic here is one-time setup on the device with torch ops; none of it is on the
per-forward path.
"""
from __future__ import annotations

import torch

BF16, F16, F32 = torch.bfloat16, torch.float16, torch.float32


def _pad_rows(w2d: torch.Tensor, mult: int) -> torch.Tensor:
    n = w2d.shape[0]
    n_pad = (n + mult - 1) // mult * mult
    if n_pad == n:
        return w2d.contiguous()
    out = w2d.new_zeros((n_pad,) + tuple(w2d.shape[1:]))
    out[:n] = w2d
    return out


def pack_conv(weight, bias, bn=None, cin_pad=None, eps=1e-5, dtype=BF16):
    """[cout,cin,k,k] conv (+ optional eval BN) -> (bf16 [cout_pad, k*k*cin_pad], f32 bias [cout_pad])."""
    w = weight.detach().to(F32)
    cout, cin, kh, kw = w.shape
    b = bias.detach().to(F32) if bias is not None else w.new_zeros(cout)
    if bn is not None:
        g, beta, mean, var = (t.detach().to(F32) for t in bn)
        scale = g / torch.sqrt(var + eps)
        w = w * scale.view(-1, 1, 1, 1)
        b = (b - mean) * scale + beta
    w = w.permute(0, 2, 3, 1)  # [cout, ky, kx, cin]
    if cin_pad is not None and cin_pad != cin:
        wp = w.new_zeros(cout, kh, kw, cin_pad)
        wp[..., :cin] = w
        w = wp
    w2 = _pad_rows(w.reshape(cout, -1), 16).to(dtype).contiguous()
    b2 = _pad_rows(b.view(-1, 1), 16).view(-1).contiguous()
    return w2, b2


def pack_linear(weight, bias, n_mult=128):
    w = _pad_rows(weight.detach().to(F32).reshape(weight.shape[0], -1), n_mult).to(BF16).contiguous()
    b = _pad_rows(bias.detach().to(F32).view(-1, 1), n_mult).view(-1).contiguous()
    return w, b


def fold_layernorm(weight, bias, gamma, beta, n_mult=128):
    """LayerNorm folded into the linear layer that follows it (``mde_gemm_ln_bf16`` consumer side):
    LN(x) W^T + b = rstd * (x (gamma.W)^T - mean * colsum) + (b + W beta).
    Returns (bf16 gamma.W [N,K], f32 b + W beta [N], f32 colsum [N]); colsum is taken over the ROUNDED weight so that the
    mean term cancels exactly what the tensor cores accumulate."""
    w = weight.detach().to(F32).reshape(weight.shape[0], -1)
    g, be = gamma.detach().to(F32), beta.detach().to(F32)
    wf = _pad_rows(w * g.view(1, -1), n_mult).to(BF16).contiguous()
    b = _pad_rows((bias.detach().to(F32) + w @ be).view(-1, 1), n_mult).view(-1).contiguous()
    cs = wf.to(torch.float64).sum(1).to(F32).contiguous()
    return wf, b, cs


def pack_encoder(sd, p="encoder."):
    """ViT-B/16 + channel_reduce + input BN (encoder.py:16-58)."""
    f = lambda k: sd[p + k].detach().to(F32).contiguous()
    out = {}
    out["bn4x3"] = torch.cat([f("normalize.weight"), f("normalize.bias"), f("normalize.running_mean"), f("normalize.running_var")]).contiguous()
    out["patch_w"], out["patch_b"] = pack_linear(sd[p + "encoder.conv_proj.weight"], sd[p + "encoder.conv_proj.bias"])
    out["pos"] = f("encoder.encoder.pos_embedding").view(197, 768).contiguous()
    out["cls"] = f("encoder.class_token").view(768).contiguous()
    layers = []
    for i in range(12):
        q = f"encoder.encoder.layers.encoder_layer_{i}."
        ly = {"ln1_w": f(q + "ln_1.weight"), "ln1_b": f(q + "ln_1.bias"), "ln2_w": f(q + "ln_2.weight"), "ln2_b": f(q + "ln_2.bias")}
        ly["qkv_w"], ly["qkv_b"] = pack_linear(sd[p + q + "self_attention.in_proj_weight"], sd[p + q + "self_attention.in_proj_bias"])
        ly["proj_w"], ly["proj_b"] = pack_linear(sd[p + q + "self_attention.out_proj.weight"], sd[p + q + "self_attention.out_proj.bias"])
        ly["fc1_w"], ly["fc1_b"] = pack_linear(sd[p + q + "mlp.0.weight"], sd[p + q + "mlp.0.bias"])
        ly["fc2_w"], ly["fc2_b"] = pack_linear(sd[p + q + "mlp.3.weight"], sd[p + q + "mlp.3.bias"])
        # the same two layers with ln_1 / ln_2 folded in (used when the batch is large enough for the CTA-pair GEMM)
        ly["qkv_ln"] = fold_layernorm(sd[p + q + "self_attention.in_proj_weight"], sd[p + q + "self_attention.in_proj_bias"], ly["ln1_w"], ly["ln1_b"])
        ly["fc1_ln"] = fold_layernorm(sd[p + q + "mlp.0.weight"], sd[p + q + "mlp.0.bias"], ly["ln2_w"], ly["ln2_b"])
        layers.append(ly)
    out["layers"] = layers
    out["reduce"] = [pack_linear(sd[f"{p}channel_reduce.{i}.weight"], sd[f"{p}channel_reduce.{i}.bias"]) for i in range(4)]
    return out


def pack_decoder(sd, p="decoder.convs."):
    """decoder.py:41-88: every ConvBlock with its BatchNorm folded; heads padded to 16 output rows."""
    out = {}
    names = ["init_conv", "mars_adapt"] + [f"upconv_{i}_{j}" for i in range(4) for j in range(2)] + [f"skip_{i}" for i in range(3)]
    for n in names:
        q = p + n + "."
        bn = (sd[q + "conv.1.weight"], sd[q + "conv.1.bias"], sd[q + "conv.1.running_mean"], sd[q + "conv.1.running_var"])
        out[n] = pack_conv(sd[q + "conv.0.conv.1.weight"], sd[q + "conv.0.conv.1.bias"], bn)
    for i in range(4):
        out[f"dispconv_{i}"] = pack_conv(sd[f"{p}dispconv_{i}.weight"], sd[f"{p}dispconv_{i}.bias"])
    return out


def pack_enhanced(sd):
    """enhanced_model.py:11-91 extras: attention projections (q|k|v stacked, padded to 384 rows),
    uncertainty / refinement convs (259 input channels padded to 320), fusion weights."""
    out = {"attn": {}}
    for i in ("1", "2"):
        q = f"attention_blocks.{i}."
        w = torch.cat([sd[q + "query.weight"], sd[q + "key.weight"], sd[q + "value.weight"]]).detach().to(F32).view(320, 256)
        b = torch.cat([sd[q + "query.bias"], sd[q + "key.bias"], sd[q + "value.bias"]]).detach().to(F32)
        pw, pb = pack_linear(w, b)
        out["attn"][i] = (pw, pb, sd[q + "gamma"].detach().to(F32).contiguous())
    u, r = "uncertainty_module.", "refinement."
    out["unc1"] = pack_conv(sd[u + "conv1.weight"], sd[u + "conv1.bias"])
    out["unc2"] = pack_conv(sd[u + "conv2.weight"], sd[u + "conv2.bias"])
    out["unc3"] = pack_conv(sd[u + "conv3.weight"], sd[u + "conv3.bias"])
    out["ref1"] = pack_conv(sd[r + "conv1.weight"], sd[r + "conv1.bias"], cin_pad=320)
    out["ref2"] = pack_conv(sd[r + "conv2.weight"], sd[r + "conv2.bias"])
    out["ref3"] = pack_conv(sd[r + "conv3.weight"], sd[r + "conv3.bias"])
    out["fusion_w"] = sd["feature_fusion.weights"].detach().to(F32).contiguous()
    return out


def pack_texture_refine(sd, p=""):
    """refinement.py:19-36 TextureAwareRefinementModule: rgb_conv1 (3->64) and depth_conv1 (1->64) are merged into ONE
    16->128 conv over the packed input {r,g,b,depth,0..} (block-structured weights: rows 0..63 see r,g,b, rows 64..127 see
    depth), so the first layer is a single implicit GEMM; operands are fp16 (see mde_conv_nhwc_f16); the 1x1 output conv stays fp32 for the output kernel."""
    f = lambda k: sd[p + k].detach().to(F32)
    w1 = f("rgb_conv1.weight").new_zeros(128, 16, 3, 3)
    w1[:64, 0:3] = f("rgb_conv1.weight")
    w1[64:, 3:4] = f("depth_conv1.weight")
    b1 = torch.cat([f("rgb_conv1.bias"), f("depth_conv1.bias")])
    out = {"l1": pack_conv(w1, b1, dtype=F16)}
    for n in ("rgb_conv2", "depth_conv2", "combined_conv1", "combined_conv2"):
        out[n] = pack_conv(sd[p + n + ".weight"], sd[p + n + ".bias"], dtype=F16)
    out["out_w"] = f("output_conv.weight").reshape(32).contiguous()
    out["out_b"] = f("output_conv.bias").reshape(1).contiguous()
    return out


def pack_depth_fusion(sd, n, p=""):
    """refinement.py:122-129 MultiScaleDepthFusion: the n 3x3 1->1 confidence convs and the 1x1 n->1 fusion conv as flat fp32."""
    f = lambda k: sd[p + k].detach().to(F32)
    return {"conf_w": torch.stack([f(f"confidence_conv.{i}.weight").reshape(9) for i in range(n)]).contiguous(),
            "conf_b": torch.cat([f(f"confidence_conv.{i}.bias").reshape(1) for i in range(n)]).contiguous(),
            "fuse_w": f("fusion_conv.weight").reshape(n).contiguous(), "fuse_b": f("fusion_conv.bias").reshape(1).contiguous()}
