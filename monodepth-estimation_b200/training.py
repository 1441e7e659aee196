"""The reference's training step on the device (BASELINE config 5).

The reference's training step (``train/mars_adapter.py:131-140``) is ``model(images)`` -> ``compute_losses`` ->
``losses['total'].backward()`` -> ``optimizer.step()``; autograd walks torchvision's ``EncoderBlock`` (tv:110-120) twelve
times.  This module states that walk explicitly over the C-ABI: ``ViTBlockTrainer.forward`` keeps what the backward needs
(the LayerNorm inputs, q / k / v^T, the attention output, the pre-GELU activation), ``backward`` returns the gradient of the
block input and of the block's twelve parameter tensors under the reference's ``state_dict`` names.

GEMM-shaped gradients run on the tcgen05 GEMM (``mde_gemm_bf16``) with transposed operands (``csrc/backward.cu`` has the
transposing casts and everything that is not a GEMM).  Operands are bf16, accumulation and every parameter gradient fp32.
``DepthTrainer`` (bottom of the file) composes the whole step: train-mode forward of the ``('disp', 0)`` path, ``compute_losses``,
backward through decoder level 0 / features / ViT, Adam.  ``sharding.allreduce_mean_grads`` is the data-parallel hook.
``DepthAnything.forward`` itself still raises in ``train()`` mode: training goes through ``DepthTrainer.step``.
"""
from __future__ import annotations

import torch

from . import _lib as L
from . import ops

BF16, F32 = torch.bfloat16, torch.float32
_NAMES = ("ln_1.weight", "ln_1.bias", "self_attention.in_proj_weight", "self_attention.in_proj_bias", "self_attention.out_proj.weight",
          "self_attention.out_proj.bias", "ln_2.weight", "ln_2.bias", "mlp.0.weight", "mlp.0.bias", "mlp.3.weight", "mlp.3.bias")


def _pad_rows(t, mult=128):
    n = t.shape[0]
    if n % mult == 0:
        return t.contiguous()
    out = torch.zeros((n + mult - 1) // mult * mult, *t.shape[1:], device=t.device, dtype=t.dtype)
    out[:n] = t
    return out


class ViTBlockTrainer:
    """One torchvision ``EncoderBlock`` (tv:84-120; ViT-B/16: 768 wide, 12 heads, MLP 3072, LayerNorm eps 1e-6, no dropout).

    ``params``: dict with the twelve tensors of the block under the names they have inside
    ``encoder.encoder.encoder.layers.encoder_layer_i.`` (``ln_1.weight`` ... ``mlp.3.bias``), fp32, on a CUDA device.
    """

    def __init__(self, params: dict):
        self.p = {k: params[k].detach().to(F32).contiguous() for k in _NAMES}
        dev = self.p["ln_1.weight"].device
        if dev.type != "cuda":
            raise RuntimeError("ViTBlockTrainer needs CUDA tensors (there is no CPU path)")
        self.dev = dev
        self.repack()

    def repack(self):
        """bf16 operand copies of the weights, in both orientations; call again after an optimizer step changed ``self.p``."""
        p = self.p
        self.w = {}
        for name, key in (("qkv", "self_attention.in_proj_weight"), ("proj", "self_attention.out_proj.weight"), ("fc1", "mlp.0.weight"), ("fc2", "mlp.3.weight")):
            w = p[key]
            self.w[name] = _pad_rows(w.to(BF16))                      # [N][K]: forward operand
            wt, _ = ops.transpose_bf16(w, ldt=w.shape[0])            # [K][N]: dX operand (N is a multiple of 64 for every layer)
            self.w[name + "_t"] = _pad_rows(wt)
        self.zero = {n: torch.zeros(n, device=self.dev, dtype=F32) for n in (768, 2304, 3072)}

    # ------------------------------------------------------------------ forward (keeps activations)
    def forward(self, x):
        """x f32 [B*197, 768] -> block output f32 [B*197, 768]; activations for ``backward`` are kept on ``self``."""
        p, w = self.p, self.w
        M = x.shape[0]
        B = M // L.VIT_SEQ
        if M != B * L.VIT_SEQ or x.shape[1] != L.VIT_DIM or x.dtype != F32:
            raise ValueError("x must be float32 [B*197, 768]")
        e = lambda *s, dt=BF16: torch.empty(*s, device=self.dev, dtype=dt)
        s = self.saved = {"x_in": x.contiguous(), "B": B}
        s["h1"] = ops.layernorm(s["x_in"], p["ln_1.weight"], p["ln_1.bias"], 1e-6)
        s["q"], s["k"] = e(M, 768), e(M, 768)
        s["vt"] = torch.zeros(B * 12 * 64 * L.VT_PITCH, device=self.dev, dtype=BF16)
        ops.gemm(s["h1"], w["qkv"], p["self_attention.in_proj_bias"], L.EPI_QKV_SPLIT, s["q"], s["k"], s["vt"])
        s["attn"] = ops.vit_attention(s["q"], s["k"], s["vt"], B)
        s["x_mid"] = e(M, 768, dt=F32)
        ops.gemm(s["attn"], w["proj"], p["self_attention.out_proj.bias"], L.EPI_BIAS_RESID_F32, s["x_mid"], s["x_in"])
        s["h2"] = ops.layernorm(s["x_mid"], p["ln_2.weight"], p["ln_2.bias"], 1e-6)
        s["u"] = e(M, 3072)
        ops.gemm(s["h2"], w["fc1"], p["mlp.0.bias"], L.EPI_BIAS_BF16, s["u"])  # pre-activation kept: GELU' needs it
        s["mlp"] = ops.gelu_bf16(s["u"])
        out = e(M, 768, dt=F32)
        ops.gemm(s["mlp"], w["fc2"], p["mlp.3.bias"], L.EPI_BIAS_RESID_F32, out, s["x_mid"])
        return out

    # ------------------------------------------------------------------ backward
    def _linear_bwd(self, dy_rm, dy_t, x_act, w_t, n_in, want_dx=True):
        """dy_rm bf16 [M][N], dy_t bf16 [N][Mp], x_act bf16 [M][K] -> (dX bf16 [M][K] | None, dW f32 [N][K], db f32 [N])."""
        M, N = dy_rm.shape
        Mp = dy_t.shape[1]
        x_t, _ = ops.transpose_bf16(x_act, ldt=Mp)                                   # [K][Mp]
        dw = torch.empty(N, n_in, device=self.dev, dtype=F32)
        ops.gemm(dy_t, _pad_rows(x_t), self.zero[n_in], L.EPI_BIAS_F32, dw)
        db = ops.colsum_bf16(dy_rm)
        dx = None
        if want_dx:
            dx = torch.empty(M, n_in, device=self.dev, dtype=BF16)
            ops.gemm(dy_rm, w_t, self.zero[n_in], L.EPI_BIAS_BF16, dx)
        return dx, dw, db

    def backward(self, dout):
        """dout f32 [B*197, 768] = gradient of the block output -> (gradient of the block input f32, {name: fp32 gradient})."""
        s, p, w = self.saved, self.p, self.w
        B = s["B"]
        M = dout.shape[0]
        Mp = (M + 63) // 64 * 64
        g = {}
        # x_out = x_mid + fc2(gelu(fc1(ln_2(x_mid))))
        d2_t, d2 = ops.transpose_bf16(dout.contiguous(), ldt=Mp, want_rm=True)                       # bf16 dY of mlp.3, both orientations
        dmlp, g["mlp.3.weight"], g["mlp.3.bias"] = self._linear_bwd(d2, d2_t, s["mlp"], w["fc2_t"], 3072)
        du = ops.gelu_bwd_bf16(s["u"], dmlp)
        du_t, _ = ops.transpose_bf16(du, ldt=Mp)
        dh2, g["mlp.0.weight"], g["mlp.0.bias"] = self._linear_bwd(du, du_t, s["h2"], w["fc1_t"], 768)
        dx_mid, g["ln_2.weight"], g["ln_2.bias"] = ops.layernorm_bwd(s["x_mid"], dh2, p["ln_2.weight"], 1e-6, dres=dout.contiguous())
        # x_mid = x_in + out_proj(attention(in_proj(ln_1(x_in))))
        d1_t, d1 = ops.transpose_bf16(dx_mid, ldt=Mp, want_rm=True)
        dattn, g["self_attention.out_proj.weight"], g["self_attention.out_proj.bias"] = self._linear_bwd(d1, d1_t, s["attn"], w["proj_t"], 768)
        dqkv = ops.vit_attention_bwd(s["q"], s["k"], s["vt"], dattn, B)
        dqkv_t, _ = ops.transpose_bf16(dqkv, ldt=Mp)
        dh1, g["self_attention.in_proj_weight"], g["self_attention.in_proj_bias"] = self._linear_bwd(dqkv, dqkv_t, s["h1"], w["qkv_t"], 768)
        dx_in, g["ln_1.weight"], g["ln_1.bias"] = ops.layernorm_bwd(s["x_in"], dh1, p["ln_1.weight"], 1e-6, dres=dx_mid)
        return dx_in, g


HOOKS = (2, 5, 8, 11)  # encoder.py:44-50: forward hooks on these blocks feed the decoder


class ViTEncoderTrainer:
    """torchvision ``VisionTransformer`` as the reference uses it (``encoder.py:29-50,139-147``): patch embedding + class token +
    position embedding (tv:268-300,156), twelve ``EncoderBlock`` s, the outputs of blocks 2 / 5 / 8 / 11 tapped by forward hooks.
    The final ``encoder.ln`` and ``heads`` produce a value the reference discards (``encoder.py:38,142``): no gradient, not run.

    ``params``: the VisionTransformer's own ``named_parameters()`` names (``class_token``, ``conv_proj.weight``, ``conv_proj.bias``,
    ``encoder.pos_embedding``, ``encoder.layers.encoder_layer_<i>.<...>``), i.e. the reference's ``state_dict`` keys without the
    leading ``encoder.encoder.``.
    """

    def __init__(self, params: dict):
        f = lambda k: params[k].detach().to(F32).contiguous()
        self.p = {k: f(k) for k in ("class_token", "conv_proj.weight", "conv_proj.bias", "encoder.pos_embedding")}
        self.dev = self.p["class_token"].device
        if self.dev.type != "cuda":
            raise RuntimeError("ViTEncoderTrainer needs CUDA tensors (there is no CPU path)")
        self.blocks = []
        for i in range(12):
            pre = f"encoder.layers.encoder_layer_{i}."
            self.blocks.append(ViTBlockTrainer({n: params[pre + n] for n in _NAMES}))
        self.repack()

    def repack(self):
        p = self.p
        self.patch_w = p["conv_proj.weight"].reshape(768, 768).to(BF16).contiguous()
        self.patch_w_t, _ = ops.transpose_bf16(p["conv_proj.weight"].reshape(768, 768), ldt=768)  # [in][out]: dX operand
        self.pos = p["encoder.pos_embedding"].reshape(197, 768).contiguous()
        self.cls = p["class_token"].reshape(768).contiguous()
        self.zero768 = torch.zeros(768, device=self.dev, dtype=F32)

    def forward(self, patches):
        """patches bf16 [B*196, 768] (row = (image, patch), columns (c, ky, kx): what ``mde_input_stage`` writes) -> the four
        tapped residual streams, f32 [B*197, 768] each."""
        if patches.dtype != BF16 or patches.shape[1] != 768 or patches.shape[0] % 196:
            raise ValueError("patches must be bfloat16 [B*196, 768]")
        B = patches.shape[0] // 196
        self.B, self.patches = B, patches.contiguous()
        x = torch.empty(B * L.VIT_SEQ, 768, device=self.dev, dtype=F32)
        ops.gemm(self.patches, self.patch_w, self.p["conv_proj.bias"], L.EPI_PATCH_EMBED, x, self.pos, self.cls)
        taps = []
        for i, blk in enumerate(self.blocks):
            x = blk.forward(x)
            if i in HOOKS:
                taps.append(x)
        return taps

    def backward(self, dtaps):
        """dtaps: four entries (f32 [B*197, 768] or None) = the gradients of the tapped outputs.  Returns {name: fp32 gradient}
        under the VisionTransformer's parameter names (parameters that received no gradient are absent, as autograd leaves
        their ``.grad`` at None: blocks above the last tap that matters)."""
        grads = {}
        for _, g in self.backward_iter(dtaps):
            grads.update(g)
        return grads

    def backward_iter(self, dtaps):
        """The same walk as a generator: yields (block index, that block's gradients) from block 11 down, then (-1, the gradients
        of the embedding parameters).  A data-parallel step all-reduces what it has while the rest is still being computed."""
        B, M = self.B, self.B * L.VIT_SEQ
        dx = None
        for i in range(11, -1, -1):
            if i in HOOKS and dtaps[HOOKS.index(i)] is not None:
                d = dtaps[HOOKS.index(i)]
                dx = d.to(F32).clone() if dx is None else ops.axpy_f32(dx, d.contiguous())
            if dx is None:
                continue
            dx, g = self.blocks[i].backward(dx)
            yield i, {f"encoder.layers.encoder_layer_{i}.{n}": v for n, v in g.items()}
        if dx is None:
            return
        grads = {}
        # tokens = [class_token ; conv_proj(patches)] + pos_embedding  (tv:288-300, 156)
        dpos = ops.batch_sum_f32(dx.view(B, L.VIT_SEQ * 768))
        grads["encoder.pos_embedding"] = dpos.view(1, L.VIT_SEQ, 768)
        grads["class_token"] = dpos.view(L.VIT_SEQ, 768)[0].clone().view(1, 1, 768)
        Mp = (B * 196 + 63) // 64 * 64
        dy_t = torch.empty(768, Mp, device=self.dev, dtype=BF16)
        dy = torch.empty(B * 196, 768, device=self.dev, dtype=BF16)
        for b in range(B):  # the patch rows of image b (its class-token row is skipped), placed side by side
            rows = dx[b * L.VIT_SEQ + 1:(b + 1) * L.VIT_SEQ]
            last = b == B - 1
            ops.transpose_bf16(rows, dst_t=dy_t, dst_rm=dy, col0=b * 196, cols_t=(Mp - b * 196) if last else 196, row0_rm=b * 196)
        p_t, _ = ops.transpose_bf16(self.patches, ldt=Mp)
        dw = torch.empty(768, 768, device=self.dev, dtype=F32)
        ops.gemm(dy_t, p_t, self.zero768, L.EPI_BIAS_F32, dw)
        grads["conv_proj.weight"] = dw.view(768, 3, 16, 16)
        grads["conv_proj.bias"] = ops.colsum_bf16(dy)
        # gradient of the patch rows themselves (f32 [B*196, 768], columns (c, ky, kx)): the input-stage backward (encoder.normalize) reads it
        self.dpatches = torch.empty(B * 196, 768, device=self.dev, dtype=F32)
        ops.gemm(dy, self.patch_w_t, self.zero768, L.EPI_BIAS_F32, self.dpatches)
        yield -1, grads


# ====================================================================== the rest of the step: features, decoder level 0, loss, Adam
def _linear_grads(dy_f32, x_t, w_t, zero_in, dev):
    """dy f32 [N, Co] (gradient of x W^T + b), x_t bf16 [Ci, Np] (the layer input, transposed, zero padded), w_t bf16 [Ci, Co]
    -> (dx f32 [N, Ci], dW f32 [Co, Ci], db f32 [Co])."""
    N, Co = dy_f32.shape
    Np, Ci = x_t.shape[1], x_t.shape[0]
    dy_t, dy_rm = ops.transpose_bf16(dy_f32, ldt=Np, want_rm=True)
    dw = torch.empty(Co, Ci, device=dev, dtype=F32)
    ops.gemm(dy_t, x_t, zero_in, L.EPI_BIAS_F32, dw)
    db = ops.colsum_bf16(dy_rm)
    dx = torch.empty(N, Ci, device=dev, dtype=F32)
    ops.gemm(dy_rm, w_t, zero_in, L.EPI_BIAS_F32, dx)
    return dx, dw, db


class _FeatureT:
    """encoder.py:156-168 for one tap: token -> grid rows, 1x1 conv 768 -> 256 (channel_reduce[i]), instance norm."""

    def __init__(self, weight, bias, dev):
        self.weight, self.bias, self.dev = weight, bias, dev  # views of the model's parameters
        self.zero = torch.zeros(768, device=dev, dtype=F32)
        self.repack()

    def repack(self):
        w = self.weight.detach().reshape(256, 768)
        self.w = w.to(BF16).contiguous()
        self.w_t = _pad_rows(ops.transpose_bf16(w.contiguous(), ldt=256)[0])  # [768][256]

    def forward(self, tap, B, strip, G):
        self.B, self.strip, self.G = B, strip, G
        self.grid_a = ops.tokens_to_grid(tap, B, strip, G)                      # bf16 [B*G, 768]
        self.reduced = torch.empty(B * G, 256, device=self.dev, dtype=F32)
        ops.gemm(self.grid_a, self.w, self.bias.detach(), L.EPI_BIAS_F32, self.reduced)
        return ops.instnorm_f32(self.reduced.view(B, G, 256))                    # f32 [B, G, 256] = NHWC [B, gh, gw, 256]

    def backward(self, dfeat):
        B, G = self.B, self.G
        dred = ops.instnorm_bwd_f32(self.reduced.view(B, G, 256), dfeat.reshape(B, G, 256).contiguous()).view(B * G, 256)
        Np = (B * G + 63) // 64 * 64
        x_t = _pad_rows(ops.transpose_bf16(self.grid_a, ldt=Np)[0])
        dgrid, dw, db = _linear_grads(dred, x_t, self.w_t, self.zero, self.dev)
        return ops.tokens_to_grid_bwd(dgrid, B, self.strip, G), {"weight": dw.view(256, 768, 1, 1), "bias": db}


class _ConvBlockT:
    """decoder.py:22-34 in train mode: ReflectionPad2d(1) + 3x3 conv + BatchNorm2d (batch statistics) + ELU, as im2col + GEMM."""

    def __init__(self, sd, prefix, dev):
        g = lambda k: sd[prefix + k]
        self.cw, self.cb = g("conv.0.conv.1.weight"), g("conv.0.conv.1.bias")
        self.bw, self.bb, self.rm, self.rv = g("conv.1.weight"), g("conv.1.bias"), g("conv.1.running_mean"), g("conv.1.running_var")
        self.co, self.ci = self.cw.shape[0], self.cw.shape[1]
        self.dev = dev
        self.zero = torch.zeros(9 * self.ci, device=dev, dtype=F32)
        self.repack()

    def repack(self):
        wm = self.cw.detach().permute(0, 2, 3, 1).reshape(self.co, 9 * self.ci).contiguous()  # k = (ky*3+kx)*Ci + c
        self.wm = _pad_rows(wm.to(BF16))
        self.wm_t = _pad_rows(ops.transpose_bf16(wm, ldt=self.co)[0])                          # [9Ci][Co]

    def forward(self, x):
        B, h, w, _ = x.shape
        self.shape = (B, h, w)
        col, self.colT = ops.im2col3x3(x.contiguous(), L.PAD_REFLECT)
        self.z = torch.empty(B * h * w, self.co, device=self.dev, dtype=F32)
        ops.gemm(col, self.wm, self.cb.detach(), L.EPI_BIAS_F32, self.z)
        y, self.mean, self.rstd = ops.bn_elu_train_fwd(self.z, self.co, self.bw.detach(), self.bb.detach(), run_mean=self.rm, run_var=self.rv)
        return y.view(B, h, w, self.co)

    def backward(self, dy):
        B, h, w = self.shape
        dz, dg, db = ops.bn_elu_train_bwd(self.z, self.co, self.mean, self.rstd, self.bw.detach(), self.bb.detach(), dy.reshape(B * h * w, self.co).contiguous())
        dxcol, dw, dcb = _linear_grads(dz, _pad_rows(self.colT), self.wm_t, self.zero, self.dev)
        dx = ops.col2im3x3(dxcol, B, h, w, self.ci, L.PAD_REFLECT)
        dwt = dw.view(self.co, 3, 3, self.ci).permute(0, 3, 1, 2).contiguous()
        return dx, {"conv.0.conv.1.weight": dwt, "conv.0.conv.1.bias": dcb, "conv.1.weight": dg, "conv.1.bias": db}


def _to_padded_bf16(y_nhwc):
    """fp32 NHWC -> bf16 padded NHWC with the reflected border the inference conv kernels read (layout glue, torch ops)."""
    import torch.nn.functional as F

    B, h, w, Cc = y_nhwc.shape  # reflection over (h, w) with the channel axis innermost: a 3-D reflection pad with no padding on the last axis
    return F.pad(y_nhwc.reshape(1, B, h, w, Cc), (0, 0, 1, 1, 1, 1), mode="reflect").view(B, h + 2, w + 2, Cc).to(BF16)


class _ConvBlockStatsT:
    """decoder.py:22-34 in train mode, FORWARD ONLY, for the decoder levels whose outputs the loss never sees (levels 1-3): the
    reference still runs them under ``model.train()``, so their BatchNorm running statistics advance every step.  The 3x3
    convolution runs on the inference kernels (``mde_conv_nhwc``, BatchNorm NOT folded, no activation, bf16 out), the batch
    statistics / running-statistic update / normalisation / ELU on ``mde_bn_elu_train_fwd``."""

    def __init__(self, sd, prefix, dev):
        g = lambda k: sd[prefix + k]
        self.cw, self.cb = g("conv.0.conv.1.weight"), g("conv.0.conv.1.bias")
        self.bw, self.bb, self.rm, self.rv = g("conv.1.weight"), g("conv.1.bias"), g("conv.1.running_mean"), g("conv.1.running_var")
        self.co, self.ci = self.cw.shape[0], self.cw.shape[1]
        self.dev = dev
        self.repack()

    def repack(self):
        from .packing import pack_conv

        self.wp, self.bp = pack_conv(self.cw.detach(), self.cb.detach())

    def forward(self, xp, h, w):
        """xp: bf16 padded NHWC [B, h+2, w+2, ci] -> bf16 padded NHWC [B, h+2, w+2, co]."""
        B = xp.shape[0]
        z = torch.empty(B, h + 2, w + 2, self.co, device=self.dev, dtype=BF16)
        ops.conv_nhwc(xp, h, w, self.ci, self.wp, self.bp, self.co, 3, L.PAD_REFLECT, L.ACT_NONE, out_bf16=z)
        zf = z[:, 1:-1, 1:-1, :].float().reshape(B * h * w, self.co)
        y, _, _ = ops.bn_elu_train_fwd(zf, self.co, self.bw.detach(), self.bb.detach(), run_mean=self.rm, run_var=self.rv)
        return _to_padded_bf16(y.view(B, h, w, self.co))


class _DispHeadT:
    """decoder.py:86-88,166: Conv2d(C, 1, 3, padding=1) + sigmoid, times the source scale (model.py:63-92)."""

    def __init__(self, weight, bias, dev):
        self.weight, self.bias, self.dev = weight, bias, dev
        self.ci = weight.shape[1]
        self.zero = torch.zeros(9 * self.ci, device=dev, dtype=F32)
        self.repack()

    def repack(self):
        wm = self.weight.detach().permute(0, 2, 3, 1).reshape(1, 9 * self.ci)
        self.wm = _pad_rows(wm.to(BF16))                                     # [128][9Ci], row 0 real
        wm64 = torch.zeros(64, 9 * self.ci, device=self.dev, dtype=F32)
        wm64[0] = wm[0]
        self.wm_t = _pad_rows(ops.transpose_bf16(wm64, ldt=64)[0])           # [9Ci][64], column 0 real
        self.bias128 = torch.zeros(128, device=self.dev, dtype=F32)
        self.bias128[:1] = self.bias.detach()

    def forward(self, x, mul):
        B, h, w, _ = x.shape
        self.shape, self.mul = (B, h, w), mul
        col, self.colT = ops.im2col3x3(x.contiguous(), L.PAD_ZERO)
        self.z = torch.empty(B * h * w, 128, device=self.dev, dtype=F32)
        ops.gemm(col, self.wm, self.bias128, L.EPI_BIAS_F32, self.z)
        return ops.sigmoid_head(self.z, B * h * w, mul).view(B, h, w, 1)

    def backward(self, ddisp):
        B, h, w = self.shape
        N = B * h * w
        dz64 = torch.zeros(N, 64, device=self.dev, dtype=F32)
        ops.sigmoid_head_bwd(self.z, ddisp.reshape(N).contiguous(), self.mul, out=dz64)
        dxcol, dw, db = _linear_grads(dz64, _pad_rows(self.colT), self.wm_t, self.zero, self.dev)
        dx = ops.col2im3x3(dxcol, B, h, w, self.ci, L.PAD_ZERO)
        return dx, {"weight": dw[0].view(1, 3, 3, self.ci).permute(0, 3, 1, 2).contiguous(), "bias": db[:1].clone()}


class DepthTrainer:
    """One training step of the reference's fine-tuning loop (``train/mars_adapter.py:117-153``) for ``DepthAnything`` on the device:
    train-mode forward of everything ``('disp', 0)`` depends on, ``compute_losses`` (:209-247), backward, Adam with the
    reference's two learning-rate groups (:47-55).  The model's fp32 parameters and BatchNorm buffers are updated in place.

    Decoder levels 1-3 do not influence the loss; they are run forward-only in train mode (``full_decoder=True``, the default) so that
    every BatchNorm buffer — running statistics and ``num_batches_tracked`` — ends the step as ``model.train(); model(x)`` leaves it
    in the reference.  Deviations, stated once: the ``torch.rand_like`` draw of :238 is an argument; only ``('disp', 0)`` is returned.
    """

    def __init__(self, model, lr=1e-4, w_depth=1.0, w_smooth=0.001, betas=(0.9, 0.999), eps=1e-8, graph=True, full_decoder=True):
        self.model, self.lr, self.w_depth, self.w_smooth, self.betas, self.eps = model, lr, w_depth, w_smooth, betas, eps
        self.graph, self._graphs = graph, {}
        self.launches_per_step = 0  # set when the step's graphs are captured
        # ViT blocks after which a data-parallel step starts reducing what it has (see step()); MDE_TRAIN_SPLITS="9,6,3" overrides
        import os
        env = os.environ.get("MDE_TRAIN_SPLITS")
        self.splits = tuple(int(v) for v in env.split(",") if v != "") if env is not None else (8, 5, 2)
        self.dev = next(model.parameters()).device
        if self.dev.type != "cuda":
            raise RuntimeError("the model must live on a CUDA (sm_100) device: there is no CPU path")
        self.sd = dict(model.state_dict(keep_vars=True))  # parameters and buffers themselves: updates land in the model
        d = lambda k: self.sd[k].detach()
        vit = {k[len("encoder.encoder."):]: d(k) for k in self.sd if k.startswith("encoder.encoder.")}
        self.enc = ViTEncoderTrainer(vit)
        self.full_decoder = full_decoder
        self.feat = {i: _FeatureT(d(f"encoder.channel_reduce.{i}.weight"), d(f"encoder.channel_reduce.{i}.bias"), self.dev)
                     for i in ((0, 1, 2, 3) if full_decoder else (2, 3))}
        det = {k: d(k) for k in self.sd}
        self.blocks = {n: _ConvBlockT(det, f"decoder.convs.{n}.", self.dev) for n in ("init_conv", "upconv_0_0", "skip_0", "upconv_0_1")}
        self.head = _DispHeadT(d("decoder.convs.dispconv_0.weight"), d("decoder.convs.dispconv_0.bias"), self.dev)
        # decoder levels 1-3 (decoder.py:106-168): forward only, for their BatchNorm running statistics
        self._stats_names = ["upconv_1_0", "skip_1", "upconv_1_1", "upconv_2_0", "skip_2", "upconv_2_1", "upconv_3_0", "upconv_3_1", "mars_adapt"] if full_decoder else []
        self.stats_blocks = {n: _ConvBlockStatsT(det, f"decoder.convs.{n}.", self.dev) for n in self._stats_names}
        self.state, self.t = {}, 0
        self.step_dev = torch.zeros(1, device=self.dev, dtype=torch.int32)  # Adam's step count where a replayed graph can read it
        # the two learning rates (train/mars_adapter.py:47-55: encoder.* at lr / 10, the rest at lr) live on the device too: a
        # captured Adam graph reads them at run time, so set_lr() / a StepLR schedule (:57-61) takes effect on the next step
        self.lr_dev = torch.tensor([lr / 10, lr], device=self.dev, dtype=F32)
        # BatchNorm layers that run in train mode here: their num_batches_tracked advances once per step, as nn.BatchNorm2d does
        self.nbt = [self.sd[k] for k in (["encoder.normalize.num_batches_tracked"] +
                                         [f"decoder.convs.{n}.conv.1.num_batches_tracked" for n in ["init_conv", "upconv_0_0", "skip_0", "upconv_0_1"] + self._stats_names])]
        self.scratch = torch.empty(L.RANGE_SCRATCH_FLOATS, device=self.dev, dtype=F32)
        self.affine = torch.empty(6, device=self.dev, dtype=F32)

    def _repack(self):
        for b in self.enc.blocks:
            b.repack()
        self.enc.repack()
        for m in list(self.feat.values()) + list(self.blocks.values()) + [self.head] + list(self.stats_blocks.values()):
            m.repack()

    def forward_backward(self, images, metadata, rnd):
        """-> (losses f32 [3] on the device: total, depth, smoothness; {state_dict key: fp32 gradient})"""
        grads = {}
        for losses, part in self._stages(images, metadata, rnd, ()):
            grads.update(part)
        return losses, grads

    def _stages(self, images, metadata, rnd, splits):
        """forward_backward as a generator over len(splits) + 1 stages: yields (losses, the gradients finished since the last
        yield) after the ViT blocks named in ``splits`` (descending block indices) and at the end.  A data-parallel step captures
        one CUDA graph per stage and all-reduces a stage's gradients while the next stage runs."""
        from .engine import grid_shape
        from .model import source_scale

        sd, dev = self.sd, self.dev
        self._repack()  # bf16 operand copies of the current fp32 master parameters
        x = images.to(F32).contiguous()
        B, _, H0, W0 = x.shape
        xin = x
        if H0 % 16 or W0 % 16:  # model.py:51-54
            xin = ops.resize_bilinear_nchw(x, (H0 // 16 + 1) * 16, (W0 // 16 + 1) * 16)
        Hp, Wp = xin.shape[2:]
        if Wp % 4:
            raise ValueError("unsupported input width")
        gh, gw, strip = grid_shape(Hp, Wp)
        G = gh * gw
        # encoder.py:129-144 with BatchNorm2d(3) in train mode: batch statistics, then the same one-pass input stage
        mean, var = ops.channel_stats_nchw(xin, 0.1, sd["encoder.normalize.running_mean"].detach(), sd["encoder.normalize.running_var"].detach())
        bn = torch.cat([sd["encoder.normalize.weight"].detach(), sd["encoder.normalize.bias"].detach(), mean, var]).contiguous()
        raw224 = torch.empty(B, 3, 224, 224, device=dev, dtype=F32)
        patches = torch.empty(B * 196, 768, device=dev, dtype=BF16)
        ops.input_stage(xin, bn, 1e-5, self.scratch, self.affine, raw224, patches)
        torch._foreach_add_(self.nbt, 1)  # nn.BatchNorm2d.forward in train mode: num_batches_tracked += 1, once per forward
        taps = self.enc.forward(patches)
        f3 = self.feat[3].forward(taps[3], B, strip, G).view(B, gh, gw, 256)
        f2 = self.feat[2].forward(taps[2], B, strip, G).view(B, gh, gw, 256)
        blk = self.blocks
        x0 = blk["init_conv"].forward(f3)                                   # decoder.py:103
        x1 = blk["upconv_0_0"].forward(x0)                                  # :113
        s0 = blk["skip_0"].forward(f2)                                      # :133 (features[-2])
        cat = torch.empty(B, 2 * gh, 2 * gw, 256, device=dev, dtype=F32)
        ops.bilinear_nhwc(x1, 2 * gh, 2 * gw, True, dst=cat, coff=0)        # :118-120
        ops.bilinear_nhwc(s0, 2 * gh, 2 * gw, True, dst=cat, coff=128)      # :136-145
        x2 = blk["upconv_0_1"].forward(cat)                                 # :152
        if self.full_decoder:
            self._decoder_stats_forward(x2, taps, B, strip, gh, gw)
        sc = source_scale(metadata)
        disp0 = self.head.forward(x2, 1.0 if sc is None else sc)            # :166 + model.py:63-92
        disp = ops.bilinear_nhwc(disp0, H0, W0, False).view(B, 1, H0, W0)   # model.py:95-100 (C = 1: NHWC == NCHW)
        losses, ddisp = ops.depth_loss(disp, x, rnd.to(F32).contiguous(), self.w_depth, self.w_smooth)
        self.outputs = {("disp", 0): disp}
        # ---------------- backward
        g = {}
        ddisp0 = ops.bilinear_nhwc_bwd(ddisp.view(B, H0, W0, 1), 2 * gh, 2 * gw, False)
        dx2, gh_ = self.head.backward(ddisp0)
        for k, v in gh_.items():
            g[f"decoder.convs.dispconv_0.{k}"] = v
        def back(name, dy):
            dx, gg = blk[name].backward(dy)
            for k, v in gg.items():
                g[f"decoder.convs.{name}.{k}"] = v
            return dx
        dcat = back("upconv_0_1", dx2)
        dx1 = ops.bilinear_nhwc_bwd(dcat, gh, gw, True, Cc=128, coff=0)
        ds0 = ops.bilinear_nhwc_bwd(dcat, gh, gw, True, Cc=128, coff=128)
        df3 = back("init_conv", back("upconv_0_0", dx1))
        df2 = back("skip_0", ds0)
        dt3, g3 = self.feat[3].backward(df3)
        dt2, g2 = self.feat[2].backward(df2)
        for i, gg in ((3, g3), (2, g2)):
            for k, v in gg.items():
                g[f"encoder.channel_reduce.{i}.{k}"] = v
        for i, gi in self.enc.backward_iter([None, None, dt2, dt3]):
            for k, v in gi.items():
                g["encoder.encoder." + k] = v
            if i in splits:
                yield losses, g
                g = {}
        dn = ops.input_norm_bwd(self.enc.dpatches, raw224, self.affine, bn, 1e-5)                    # encoder.py:129-137
        g["encoder.normalize.weight"], g["encoder.normalize.bias"] = dn[:3].clone(), dn[3:].clone()
        yield losses, g

    def _decoder_stats_forward(self, x2, taps, B, strip, gh, gw):
        """decoder.py:106-168 for levels 1-3 in train mode: their outputs do not reach the loss (only ('disp', 0) does,
        train/mars_adapter.py:219), but ``model.train(); model(x)`` runs them, so every BatchNorm there sees one batch per step."""
        G = gh * gw
        sb = self.stats_blocks
        feats = {i: _to_padded_bf16(self.feat[i].forward(taps[i], B, strip, G).view(B, gh, gw, 256)) for i in (0, 1)}
        xp = _to_padded_bf16(x2)
        h, w = 2 * gh, 2 * gw
        for i in (1, 2, 3):
            a = sb[f"upconv_{i}_0"].forward(xp, h, w)                                    # :113
            s = sb[f"skip_{i}"].forward(feats[2 - i], gh, gw) if i < 3 else None           # :126-137 (features[-(i+2)])
            cat = torch.empty(B, 2 * h + 2, 2 * w + 2, a.shape[-1] * (2 if i < 3 else 1), device=self.dev, dtype=BF16)
            ops.upsample2x_concat(a, h, w, s, gh, gw, cat)                                 # :118-145
            h, w = 2 * h, 2 * w
            xp = sb[f"upconv_{i}_1"].forward(cat, h, w)                                   # :152
            if i == 3:
                xp = sb["mars_adapt"].forward(xp, h, w)                                   # :157-158

    def set_lr(self, lr: float):
        """Change the base learning rate (group 0 = encoder.* gets lr / 10).  Safe between replays of the captured step."""
        self.lr = float(lr)
        self.lr_dev.copy_(torch.tensor([self.lr / 10, self.lr], dtype=F32), non_blocking=False)

    def lr_scheduler_step(self, epoch: int, step_size: int = 5, gamma: float = 0.5, base_lr: float | None = None):
        """torch.optim.lr_scheduler.StepLR as the reference configures it (train/mars_adapter.py:57-61): lr = base * gamma^(epoch // step_size)."""
        if base_lr is not None:
            self._base_lr = float(base_lr)
        base = self.__dict__.setdefault("_base_lr", self.lr)
        self.set_lr(base * gamma ** (epoch // step_size))

    def _adam(self, grads, plan=None):
        """torch.optim.Adam over the reference's two groups (train/mars_adapter.py:47-55): encoder.* at lr/10, decoder.* at lr.
        One multi-tensor launch (``mde_adam_multi``); ``plan`` caches its device tables when the gradient buffers are stable."""
        self.step_dev += 1
        if plan is None:
            plan = self._adam_plan(grads)
        plan.step(self.lr_dev, self.step_dev, self.betas[0], self.betas[1], self.eps)
        return plan

    def _adam_plan(self, grads):
        entries = []
        for k, gk in grads.items():
            p = self.sd[k].detach()
            m, v = self.state[k]
            if not gk.is_contiguous() or gk.numel() != p.numel():
                raise RuntimeError(f"gradient of {k} must be a contiguous tensor of the parameter's size (its pointer is baked into the Adam table)")
            entries.append((p, gk.view(p.shape), m, v, 0 if k.startswith("encoder.") else 1))
        return ops.AdamPlan(entries)

    def _adam_state(self, grads):
        for k in grads:
            if k not in self.state:
                p = self.sd[k].detach()
                self.state[k] = (torch.zeros_like(p), torch.zeros_like(p))

    def step(self, images, metadata, rnd, grads_hook=None):
        """forward + backward + Adam (encoder.* at lr/10, decoder.* at lr).  ``grads_hook(grads)`` runs between backward and the
        update: the place for the gradient all-reduce of a data-parallel job; it may be called several times per step, each time
        with a disjoint part of the gradients.  Returns the three losses (device tensor; with ``graph=True`` a buffer the next
        step overwrites).

        ``graph=True`` (default): the first step of an input shape runs eagerly (one-time initialisation inside the library must
        not happen under capture); the second captures the CUDA graphs — [re-pack bf16 operands, forward, loss, backward] and
        [Adam] — and every step from then on replays them around the hook.  When a hook is present at capture time the backward
        is cut after the ViT blocks in ``self.splits`` into one graph per stage: the hook of a stage runs on a side stream while
        the next stage's graph computes (the all-reduce overlaps the rest of the backward)."""
        from .model import source_scale

        self.t += 1
        key = (tuple(images.shape), source_scale(metadata))
        st = self._graphs.get(key) if self.graph else None
        if st is None:
            losses, grads = self.forward_backward(images, metadata, rnd)
            if grads_hook is not None:
                grads_hook(grads)
            self._adam_state(grads)
            self._adam(grads)
            self._invalidate_inference()
            if self.graph:
                self._graphs[key] = {}
            return losses
        if not st:
            st["x"], st["rnd"] = torch.empty_like(images, dtype=F32), torch.empty_like(rnd, dtype=F32)
            st["x"].copy_(images)
            st["rnd"].copy_(rnd)
            torch.cuda.synchronize(self.dev)
            splits = tuple(self.splits) if grads_hook is not None else ()
            calls0 = L.N_CALLS[0]
            gen = self._stages(st["x"], metadata, st["rnd"], splits)
            st["fb"], st["parts"], pool = [], [], None
            for _ in range(len(splits) + 1):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, pool=pool):
                    st["losses"], part = next(gen)
                pool = g.pool()
                st["fb"].append(g)
                st["parts"].append(part)
            st["grads"] = {k: v for part in st["parts"] for k, v in part.items()}
            self._adam_state(st["grads"])
            st["adam_plan"] = self._adam_plan(st["grads"])  # gradient buffers live in the graphs' pool: pointers are stable
            st["adam"] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(st["adam"]):
                self._adam(st["grads"], st["adam_plan"])
            st["side"] = torch.cuda.Stream(self.dev) if len(st["fb"]) > 1 else None
            # library entry points captured into the step's graphs (each launches >= 1 kernel): what one replayed step launches
            self.launches_per_step = L.N_CALLS[0] - calls0
        else:
            st["x"].copy_(images)
            st["rnd"].copy_(rnd)
        cur, side = torch.cuda.current_stream(self.dev), st["side"]
        for g, part in zip(st["fb"], st["parts"]):
            g.replay()
            if grads_hook is None:
                continue
            if side is None:
                grads_hook(part)
            else:  # this stage's gradients are final: reduce them next to the following stage's graph
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    grads_hook(part)
        if side is not None and grads_hook is not None:
            cur.wait_stream(side)
        st["adam"].replay()
        self._invalidate_inference()
        return st["losses"]

    def _invalidate_inference(self):
        """Parameters and BatchNorm buffers were just rewritten through raw pointers (tensor._version did not move): bump the
        module's generation so that the inference engine re-packs its bf16 weights and drops its CUDA graphs on the next
        ``model.eval(); model(x)`` (engine._fingerprint)."""
        d = self.model.__dict__
        d["_mde_gen"] = d.get("_mde_gen", 0) + 1
