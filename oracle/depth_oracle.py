"""ORACLE — test infrastructure, NOT product code.

A CPU fp32 restatement of the reference's batched depth-inference forward path,
written as plain functions over a ``state_dict`` (no ``nn.Module``, no torchvision,
no import of the reference).  It exists so that parity tests, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline leg have a checker that travels to the GPU box
(``/root/reference`` does not).  Nothing under ``monodepth-estimation_b200/`` may import it.

Pinning: ``oracle/make_golden.py`` runs the REAL reference classes (imported from
``/root/reference``) on the synthetic weights/inputs of ``synth.py`` and commits their
outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks every function
here against those vectors (fp32 agreement ~1e-6).  One exception, stated in DESIGN.md:
``texture_refine_module`` restates ``refinement.py:68-85,92-107`` because the reference's
own ``forward`` raises at ``refinement.py:88`` on every valid input; its golden vector
is produced by calling the reference module's own sub-modules in the reference's order.

Each function cites the reference lines it follows (paths relative to the reference
root; ``tv:`` = torchvision 0.26.0 ``models/vision_transformer.py``).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

HOOK_LAYERS = (2, 5, 8, 11)  # encoder.py:44


# --------------------------------------------------------------------------- ViT
def vit_block(sd, p, x, heads=12):
    """tv:110-120 EncoderBlock.forward — LN, MHA, +res, LN, MLP(erf-GELU), +res."""
    B, S, D = x.shape
    h = F.layer_norm(x, (D,), sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], 1e-6)
    qkv = F.linear(h, sd[p + "self_attention.in_proj_weight"], sd[p + "self_attention.in_proj_bias"])
    q, k, v = qkv.view(B, S, 3, heads, D // heads).permute(2, 0, 3, 1, 4)
    att = torch.softmax((q @ k.transpose(-1, -2)) / math.sqrt(D // heads), dim=-1)
    ctx = (att @ v).transpose(1, 2).reshape(B, S, D)
    x = x + F.linear(ctx, sd[p + "self_attention.out_proj.weight"], sd[p + "self_attention.out_proj.bias"])
    h = F.layer_norm(x, (D,), sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], 1e-6)
    h = F.gelu(F.linear(h, sd[p + "mlp.0.weight"], sd[p + "mlp.0.bias"]))
    return x + F.linear(h, sd[p + "mlp.3.weight"], sd[p + "mlp.3.bias"])


def vit_taps(sd, p, x224):
    """tv:268-306 + tv:156 up to the last hooked block; the final ``ln`` and the CLS
    read-out are computed-then-discarded by the reference (encoder.py:38,142) and are
    omitted.  Returns the residual stream after blocks 2,5,8,11 (encoder.py:44-50,60-67)."""
    B = x224.shape[0]
    t = F.conv2d(x224, sd[p + "conv_proj.weight"], sd[p + "conv_proj.bias"], stride=16)
    t = t.reshape(B, 768, -1).permute(0, 2, 1)
    t = torch.cat([sd[p + "class_token"].expand(B, -1, -1), t], dim=1) + sd[p + "encoder.pos_embedding"]
    taps = []
    for i in range(12):
        t = vit_block(sd, f"{p}encoder.layers.encoder_layer_{i}.", t)
        if i in HOOK_LAYERS:
            taps.append(t)
    return taps


# ----------------------------------------------------------------------- encoder
def tokens_to_grid(tok, h, w):
    """encoder.py:69-110 — strip CLS only when tokens == h*w+1, else keep it and
    zero-pad to int(sqrt(T)) x ceil(T / int(sqrt(T))) (14x15 for T=197)."""
    B, T, C = tok.shape
    if T == h * w + 1:
        tok, T = tok[:, 1:], T - 1
    if T != h * w:
        nh = int(math.sqrt(T))
        nw = T // nh
        if nh * nw < T:
            nw += 1
        if nh * nw > T:
            tok = torch.cat([tok, tok.new_zeros(B, nh * nw - T, C)], dim=1)
        h, w = nh, nw
    return tok.reshape(B, h, w, C).permute(0, 3, 1, 2)


def input_range_map(sd, p, x, train=False):
    """encoder.py:129-135 — BatchNorm2d(3) (eval: running statistics; train: batch statistics, running statistics updated in
    place as nn.BatchNorm2d does) then the whole-batch range heuristic."""
    x = F.batch_norm(x, sd[p + "normalize.running_mean"], sd[p + "normalize.running_var"],
                     sd[p + "normalize.weight"], sd[p + "normalize.bias"], train, 0.1 if train else 0.0, 1e-5)
    if x.max() > 1.0:
        x = x / 255.0
    if x.min() >= 0 and x.max() <= 1.0:
        x = 2.0 * x - 1.0
    return x


def encoder_forward(sd, x, p="encoder.", want=None, train=False):
    """encoder.py:112-187 — returns 4 x [B,256,gh,gw]."""
    H, W = x.shape[2:]
    x = input_range_map(sd, p, x, train)
    if (H, W) != (224, 224):
        x = F.interpolate(x, size=(224, 224), mode="bilinear", align_corners=False)
    taps = vit_taps(sd, p + "encoder.", x)
    if want is not None:
        want["vit_in"] = x
        want["taps"] = taps
    feats = []
    for i, t in enumerate(taps):
        g = tokens_to_grid(t, H // 16, W // 16)
        g = F.conv2d(g, sd[f"{p}channel_reduce.{i}.weight"], sd[f"{p}channel_reduce.{i}.bias"])
        feats.append(F.instance_norm(g))  # eps 1e-5, no affine (encoder.py:167)
    return feats


# ----------------------------------------------------------------------- decoder
def conv_block(sd, p, x, train=False):
    """decoder.py:8-34 — ReflectionPad2d(1), 3x3 conv, BatchNorm2d (eval, or train: batch statistics), ELU."""
    x = F.conv2d(F.pad(x, (1, 1, 1, 1), mode="reflect"), sd[p + "conv.0.conv.1.weight"], sd[p + "conv.0.conv.1.bias"])
    x = F.batch_norm(x, sd[p + "conv.1.running_mean"], sd[p + "conv.1.running_var"],
                     sd[p + "conv.1.weight"], sd[p + "conv.1.bias"], train, 0.1 if train else 0.0, 1e-5)
    return F.elu(x)


def decoder_forward(sd, feats, p="decoder.convs.", want=None, train=False):
    """decoder.py:90-173 — returns {("disp", i): [B,1,28*2^i,30*2^i]} (level sizes for a 14x15 grid)."""
    out = {}
    x = conv_block(sd, p + "init_conv.", feats[-1], train)
    for i in range(4):
        x = conv_block(sd, f"{p}upconv_{i}_0.", x, train)
        x = F.interpolate(x, size=(x.shape[2] * 2, x.shape[3] * 2), mode="bilinear", align_corners=True)
        if i < 3:
            s = conv_block(sd, f"{p}skip_{i}.", feats[-(i + 2)], train)
            if s.shape[2:] != x.shape[2:]:
                s = F.interpolate(s, size=x.shape[2:], mode="bilinear", align_corners=True)
            x = torch.cat([x, s], dim=1)
        x = conv_block(sd, f"{p}upconv_{i}_1.", x, train)
        if i == 3:
            x = conv_block(sd, p + "mars_adapt.", x, train)
        if want is not None:
            want[f"dec_x{i}"] = x
        out[("disp", i)] = torch.sigmoid(F.conv2d(x, sd[f"{p}dispconv_{i}.weight"], sd[f"{p}dispconv_{i}.bias"], padding=1))
    return out


# ------------------------------------------------------------------- full models
def _pad16(x):
    """model.py:51-54 == enhanced_model.py:174-177."""
    if x.shape[2] % 16 != 0 or x.shape[3] % 16 != 0:
        h, w = (x.shape[2] // 16 + 1) * 16, (x.shape[3] // 16 + 1) * 16
        x = F.interpolate(x, size=(h, w), mode="bilinear", align_corners=False)
    return x


def source_scale(metadata):
    """model.py:63-87 == enhanced_model.py:218-242; None = no scaling applied."""
    if metadata is None or "source" not in metadata:
        return None
    s = metadata["source"]
    if s == "rover":
        return 2.0 / max(metadata["height"], 0.1) if "height" in metadata else 1.0
    if s == "ingenuity":
        return 10.0 / max(metadata["altitude"], 1.0) if "altitude" in metadata else 0.5
    if s == "satellite":
        return 0.01
    return 1.0


def standard_forward(sd, x, metadata=None, want=None, train=False):
    """model.py:32-102 DepthAnything.forward (train=True: the BatchNorms use and update batch statistics, as under model.train())."""
    size = x.shape[2:]
    feats = encoder_forward(sd, _pad16(x), want=want, train=train)
    out = decoder_forward(sd, feats, want=want, train=train)
    if want is not None:
        want["features"] = feats
        want["disp_lowres"] = dict(out)
    sc = source_scale(metadata)
    for k in out:
        if sc is not None:
            out[k] = out[k] * sc
        if out[k].shape[2:] != size:
            out[k] = F.interpolate(out[k], size=size, mode="bilinear", align_corners=False)
    return out


def spatial_attention(sd, p, x):
    """enhanced_model.py:20-39 SelfAttentionBlock.forward (no 1/sqrt(d) scale)."""
    B, C, H, W = x.shape
    q = F.conv2d(x, sd[p + "query.weight"], sd[p + "query.bias"]).view(B, -1, H * W).permute(0, 2, 1)
    k = F.conv2d(x, sd[p + "key.weight"], sd[p + "key.bias"]).view(B, -1, H * W)
    a = torch.softmax(torch.bmm(q, k), dim=2)
    v = F.conv2d(x, sd[p + "value.weight"], sd[p + "value.bias"]).view(B, -1, H * W)
    o = torch.bmm(v, a.permute(0, 2, 1)).view(B, C, H, W)
    return sd[p + "gamma"] * o + x


def uncertainty_head(sd, p, x):
    """enhanced_model.py:50-54."""
    x = F.elu(F.conv2d(x, sd[p + "conv1.weight"], sd[p + "conv1.bias"], padding=1))
    x = F.elu(F.conv2d(x, sd[p + "conv2.weight"], sd[p + "conv2.bias"], padding=1))
    return torch.sigmoid(F.conv2d(x, sd[p + "conv3.weight"], sd[p + "conv3.bias"]))


def texture_refinement(sd, p, feat, rgb):
    """enhanced_model.py:92-113 (``target_size`` is never set, so :106-111 is dead)."""
    if rgb.shape[2:] != feat.shape[2:]:
        rgb = F.interpolate(rgb, size=feat.shape[2:], mode="bilinear", align_corners=False)
    x = torch.cat([feat, rgb], dim=1)
    x = F.elu(F.conv2d(x, sd[p + "conv1.weight"], sd[p + "conv1.bias"], padding=1))
    x = F.elu(F.conv2d(x, sd[p + "conv2.weight"], sd[p + "conv2.bias"], padding=1))
    return F.conv2d(x, sd[p + "conv3.weight"], sd[p + "conv3.bias"])


def enhanced_forward(sd, x, metadata=None, want=None):
    """enhanced_model.py:152-279 EnhancedDepthAnything.forward — 11 maps [B,1,H,W]."""
    size = x.shape[2:]
    x = _pad16(x)
    feats = encoder_forward(sd, x, want=want)
    enh = list(feats)
    for i in (1, 2):
        enh[i] = spatial_attention(sd, f"attention_blocks.{i}.", feats[i])
    out = decoder_forward(sd, enh, want=want)
    out[("uncertainty", 0)] = uncertainty_head(sd, "uncertainty_module.", feats[0])
    ref = texture_refinement(sd, "refinement.", feats[0], x)
    if want is not None:
        want["features"] = feats
        want["enh_features"] = enh
        want["refinement"] = ref
        want["uncertainty_lowres"] = out[("uncertainty", 0)]
    if ref.shape[2:] != out[("disp", 0)].shape[2:]:
        ref = F.interpolate(ref, size=out[("disp", 0)].shape[2:], mode="bilinear", align_corners=False)
    out[("disp_unrefined", 0)] = out[("disp", 0)].clone()
    out[("disp", 0)] = out[("disp", 0)] + ref
    sc = source_scale(metadata)
    if sc is not None:
        for k in out:
            if k[0] == "disp":
                out[k] = out[k] * sc
    # enhanced_model.py:63-82,250-263 — scales 1,2 are resized DOWN to scale 0's size
    wts = torch.softmax(sd["feature_fusion.weights"], dim=0)
    tgt = out[("disp", 0)].shape[2:]
    fused = 0
    for i in range(3):
        f = out[("disp", i)]
        if f.shape[2:] != tgt:
            f = F.interpolate(f, size=tgt, mode="bilinear", align_corners=False)
        fused = fused + wts[i] * f
    out[("disp_fused", 0)] = fused
    out[("disp", 0)] = fused
    if want is not None:
        want["fused_lowres"] = fused
    for k in out:
        if out[k].shape[2:] != size:
            out[k] = F.interpolate(out[k], size=size, mode="bilinear", align_corners=False)
    for k in list(out):
        if k[0] == "disp":  # layers.py:16-25 with (0.1, 100.0)
            out[("depth", k[1])] = 1 / (1 / 100.0 + (1 / 0.1 - 1 / 100.0) * out[k])
    return out


# --------------------------------------------------- refinement.py GPU modules
def texture_refine_module(sd, depth, rgb, p=""):
    """refinement.py:57-107 TextureAwareRefinementModule.forward, live lines only
    (:68-85 and :92-107).  Lines :87-90 (Sobel of the depth map) raise in the reference
    and their result is never used; they are omitted."""
    if rgb.shape[2:] != depth.shape[2:]:
        rgb = F.interpolate(rgb, size=depth.shape[2:], mode="bilinear", align_corners=False)
    c = lambda n, t: F.conv2d(t, sd[f"{p}{n}.weight"], sd.get(f"{p}{n}.bias"), padding=sd[f"{p}{n}.weight"].shape[-1] // 2)
    rf = F.elu(c("rgb_conv2", F.elu(c("rgb_conv1", rgb))))
    df = F.elu(c("depth_conv2", F.elu(c("depth_conv1", depth))))
    ex, ey = c("edge_detect_x", rgb), c("edge_detect_y", rgb)
    edge = torch.sqrt(ex.pow(2) + ey.pow(2))
    cf = F.elu(c("combined_conv1", torch.cat([rf, df], dim=1)))
    cf = F.elu(c("combined_conv2", cf))
    res = c("output_conv", cf)
    return depth + res * torch.sigmoid(edge * 5.0)


def multiscale_depth_fusion(sd, maps, scales=(0, 1, 2, 3), p=""):
    """refinement.py:131-183 MultiScaleDepthFusion.forward."""
    key = lambda s: ("disp", s) if ("disp", s) in maps else ("depth", s)
    tgt = maps[key(min(scales))].shape[2:]
    feats, confs = [], []
    for i, s in enumerate(scales):
        if key(s) not in maps:
            continue
        f = maps[key(s)]
        if f.shape[2:] != tgt:
            f = F.interpolate(f, size=tgt, mode="bilinear", align_corners=False)
        confs.append(torch.sigmoid(F.conv2d(f, sd[f"{p}confidence_conv.{i}.weight"], sd[f"{p}confidence_conv.{i}.bias"], padding=1)))
        feats.append(f)
    feats, confs = torch.cat(feats, 1), torch.cat(confs, 1)
    w = confs / (confs.sum(dim=1, keepdim=True) + 1e-8)
    return F.conv2d(feats * w, sd[p + "fusion_conv.weight"], sd[p + "fusion_conv.bias"])


def config4_forward(sd_enh, sd_fuse, sd_refine, x, metadata=None):
    """BASELINE.json config 4.  The reference never composes these modules (they are
    orphaned); the composition is stated once here, following the module
    docstrings (refinement.py:61-66,135-140): enhanced forward -> MultiScaleDepthFusion over
    ("disp",0..3) at input size -> TextureAwareRefinementModule(fused, x).  Extra keys
    ("disp_msfused",0) and ("disp_refined",0) are ours; the 11 reference keys are untouched."""
    out = enhanced_forward(sd_enh, x, metadata)
    fused = multiscale_depth_fusion(sd_fuse, out)
    out[("disp_msfused", 0)] = fused
    out[("disp_refined", 0)] = texture_refine_module(sd_refine, fused, x)
    return out
