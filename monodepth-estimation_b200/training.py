"""Backward of the ViT encoder on the device (first part).

The reference's training step (``train/mars_adapter.py:131-140``) is ``model(images)`` -> ``compute_losses`` ->
``losses['total'].backward()`` -> ``optimizer.step()``; autograd walks torchvision's ``EncoderBlock`` (tv:110-120) twelve
times.  This module states that walk explicitly over the C-ABI: ``ViTBlockTrainer.forward`` keeps what the backward needs
(the LayerNorm inputs, q / k / v^T, the attention output, the pre-GELU activation), ``backward`` returns the gradient of the
block input and of the block's twelve parameter tensors under the reference's ``state_dict`` names.

GEMM-shaped gradients run on the tcgen05 GEMM (``mde_gemm_bf16``) with transposed operands (``csrc/backward.cu`` has the
transposing casts and everything that is not a GEMM).  Operands are bf16, accumulation and every parameter gradient fp32.
Scope so far: the encoder blocks.  The decoder / loss backward, Adam and the NCCL gradient all-reduce of config 5 are the next
pieces; ``DepthAnything.forward`` still raises in ``train()`` mode.
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
        B, M = self.B, self.B * L.VIT_SEQ
        grads = {}
        dx = None
        for i in range(11, -1, -1):
            if i in HOOKS and dtaps[HOOKS.index(i)] is not None:
                d = dtaps[HOOKS.index(i)]
                dx = d.to(F32).clone() if dx is None else ops.axpy_f32(dx, d.contiguous())
            if dx is None:
                continue
            dx, g = self.blocks[i].backward(dx)
            for n, v in g.items():
                grads[f"encoder.layers.encoder_layer_{i}.{n}"] = v
        if dx is None:
            return grads
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
        self.dpatch_rows = dy  # bf16 [B*196, 768]: dY of conv_proj, kept for the input-stage backward (encoder.normalize)
        return grads
