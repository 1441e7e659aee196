"""Forward-path orchestration: one ``DepthEngine`` per model instance.

The engine owns (a) the packed kernel weights (re-derived whenever a parameter tensor
changes or the module moves), (b) one pre-allocated ``Workspace`` per input geometry and
(c) optionally a CUDA graph of the whole ViT + decoder section per geometry, so a forward is
``range/patchify`` -> graph replay -> output kernel, with no host synchronisation anywhere
(the reference's ``x.max() > 1.0`` host branch, encoder.py:132-135, is evaluated on device).

Reference call stack being replaced: ``model.py:32-102`` / ``enhanced_model.py:152-279`` ->
``encoder.py:112-187`` -> torchvision ViT -> ``decoder.py:90-173``.
"""
from __future__ import annotations

import os

import torch

from . import _lib as L
from . import ops
from .packing import pack_conv, pack_decoder, pack_encoder, pack_enhanced

BF16, F16, F32 = torch.bfloat16, torch.float16, torch.float32
HOOKS = (2, 5, 8, 11)
DEC_CH = (256, 128, 64, 32, 16)


def grid_shape(Hp: int, Wp: int):
    """encoder.py:69-110 for T=197 tokens: (gh, gw, strip_cls)."""
    h, w = Hp // 16, Wp // 16
    if h * w == 196:
        return h, w, True
    if h * w == 197:
        raise ValueError("inputs whose /16 grid has exactly 197 cells (a 1x197 strip) are not supported")
    return 14, 15, False  # int(sqrt(197)) x ceil(197/14), CLS kept, 13 zero cells appended


def _fingerprint(module):
    """Identity of the weights the packed copies were made from.  ``_mde_gen`` is bumped by code that rewrites parameters through
    raw pointers (training.DepthTrainer), which ``tensor._version`` cannot see."""
    return (module.__dict__.get("_mde_gen", 0),) + tuple((t.data_ptr(), t._version, t.device.index) for t in list(module.parameters()) + list(module.buffers()))


_LN_FUSION = os.environ.get("MDE_LN_FUSION") == "1"  # opt-in, see DepthEngine._vit
# MDE_FUSED_TAIL=1: last decoder level as one kernel (csrc/decoder_tail.cu).  Opt-in: correct (tests/test_kernels_gpu.py) but slower than
# the four tcgen05 launches it replaces (202 vs 141 us at batch 32: warp-level mma.sync runs at 1/8 of the tcgen05 rate, profiles/r02_experiments.md)
_FUSED_TAIL = os.environ.get("MDE_FUSED_TAIL", "0") == "1"
# the 1-channel heads of decoder levels 0-2 run on the side stream (MDE_SIDE_HEADS=0: in line, for A/B timing)
_SIDE_HEADS = os.environ.get("MDE_SIDE_HEADS", "1") == "1"
# MDE_FUSED_MLP=1: fc1 + fc2 of a ViT layer in one persistent launch (csrc/tc_mlp.cu).  Opt-in: bit-identical to the two GEMMs
# (tests/test_kernels_gpu.py) and equally fast back to back (60 us per layer at batch 32), but 1.3 us per layer SLOWER inside the
# captured step (2.173 vs 2.157 ms): profiles/r02_experiments.md has the in-kernel time line and why the overlap does not pay
_FUSED_MLP = os.environ.get("MDE_FUSED_MLP", "0") == "1"

class Workspace:
    """All intermediates of one forward for a fixed (B, Hp, Wp); allocated once."""

    def __init__(self, dev, B, H0, W0, enhanced):
        Hp, Wp = H0, W0
        if H0 % 16 or W0 % 16:  # model.py:51-54
            Hp, Wp = (H0 // 16 + 1) * 16, (W0 // 16 + 1) * 16
        if (Hp * Wp) % 4:
            raise ValueError("unsupported input size")
        self.B, self.H0, self.W0, self.Hp, self.Wp = B, H0, W0, Hp, Wp
        self.gh, self.gw, self.strip = grid_shape(Hp, Wp)
        gh, gw = self.gh, self.gw
        G = gh * gw
        M = B * L.VIT_SEQ
        self.nbytes = 0  # sum of the buffers below, each rounded up to 256 bytes == mde_workspace_bytes() (tests/test_dropin_cpu.py)

        def e(*s, dt=BF16, zero=False):
            t = (torch.zeros if zero else torch.empty)(*s, device=dev, dtype=dt)
            self.nbytes += (t.numel() * t.element_size() + 255) // 256 * 256
            return t

        z = lambda *s, dt=BF16: e(*s, dt=dt, zero=True)
        self.xp = e(B, 3, Hp, Wp, dt=F32) if (Hp, Wp) != (H0, W0) else None
        self.scratch = e(L.RANGE_SCRATCH_FLOATS, dt=F32)
        self.affine = e(6, dt=F32)
        self.patches = e(B * 196, 768)
        self.raw224 = e(B, 3, 224, 224, dt=F32)
        self.xa, self.xb = e(M, 768, dt=F32), e(M, 768, dt=F32)
        self.taps = [e(M, 768, dt=F32) for _ in range(4)]
        self.h = e(M, 768)
        self.lnstats = e(24, M, 2, dt=F32)  # per-row (sum, sum of squares) partials of the residual stream (LayerNorm fused into the GEMMs)
        self.q, self.k = e(B * 12 * L.VIT_SEQ * 64), e(B * 12 * L.VIT_SEQ * 64)
        self.vt = z(B * 12 * 64 * L.VT_PITCH)
        self.attn = e(M, 768)
        self.mlp = e(M, 3072)
        self.mlp_sync = e(128, dt=torch.int32, zero=True)  # row-block counters of the fused MLP kernel (all zero between launches)
        self.grid_a = [e(B * G, 768) for _ in range(4)]  # per tap: the side stream works on tap i while the ViT runs on
        self.reduced = [e(B * G, 256, dt=F32) for _ in range(4)]
        c0 = 320 if enhanced else 256  # feature 0 carries the 3 rgb + 61 zero channels of the refinement conv
        self.feats = [z(B, gh + 2, gw + 2, c0)] + [e(B, gh + 2, gw + 2, 256) for _ in range(3)]
        # decoder
        self.d_init = e(B, gh + 2, gw + 2, 256)
        self.d_a, self.d_s, self.d_cat, self.d_x, self.disp = [], [], [], [], []
        h, w = gh, gw
        for i in range(4):
            co = DEC_CH[i + 1]
            self.d_a.append(e(B, h + 2, w + 2, co))
            self.d_s.append(e(B, gh + 2, gw + 2, co) if i < 3 else None)
            h, w = 2 * h, 2 * w
            self.d_cat.append(e(B, h + 2, w + 2, co * (2 if i < 3 else 1)))
            self.d_x.append(e(B, h + 2, w + 2, co))
            self.disp.append(e(B, h, w, dt=F32))
        self.d_x3b = e(B, h + 2, w + 2, 16)  # mars_adapt output
        if enhanced:
            self.flat = [None, e(B * G, 256), e(B * G, 256), None]
            self.qkv = [None, e(B * G, 384, dt=F32), e(B * G, 384, dt=F32), None]
            self.enh = [None, e(B, gh + 2, gw + 2, 256), e(B, gh + 2, gw + 2, 256), None]
            self.u1, self.u2 = e(B, gh + 2, gw + 2, 128), e(B, gh + 2, gw + 2, 128)
            self.r1, self.r2 = e(B, gh + 2, gw + 2, 256), e(B, gh + 2, gw + 2, 128)
            self.unc = e(B, gh, gw, dt=F32)
            self.refn = e(B, gh, gw, dt=F32)
            self.fused = e(B, 2 * gh, 2 * gw, dt=F32)
        self.graph = None
        self.graph_scale = None


class DepthEngine:
    def __init__(self, module, kind: str):
        self.module = module
        self.kind = kind
        self.enhanced = kind == "enhanced"
        self._fp = None
        self._w = None
        self._ws = {}
        # CUDA graphs remove ~150 launch latencies per forward; MDE_NO_GRAPH=1 runs eagerly (debugging)
        self.use_graph = os.environ.get("MDE_NO_GRAPH", "0") != "1"
        self.launches_per_forward = 0
        # MDE_NO_OVERLAP=1 keeps every launch on one stream (A/B testing)
        self.overlap = os.environ.get("MDE_NO_OVERLAP", "0") != "1"
        self.side = None

    # ------------------------------------------------------------------ weights
    def weights(self):
        fp = _fingerprint(self.module)
        if fp != self._fp:
            dev = next(self.module.parameters()).device
            if dev.type != "cuda":
                raise RuntimeError("the model must live on a CUDA (sm_100) device: there is no CPU path — call model.to('cuda')")
            with torch.cuda.device(dev):
                ops.device_check()
                sd = self.module.state_dict()
                w = {"enc": pack_encoder(sd), "dec": pack_decoder(sd)}
                if self.enhanced:
                    w["enh"] = pack_enhanced(sd)
            self._w, self._fp = w, fp
            for ws in self._ws.values():
                ws.graph = None
        return self._w

    def workspace(self, dev, B, H0, W0):
        key = (dev.index, B, H0, W0)
        ws = self._ws.get(key)
        if ws is None:
            if len(self._ws) >= 4:  # bound the cache
                self._ws.pop(next(iter(self._ws)))
            ws = self._ws[key] = Workspace(dev, B, H0, W0, self.enhanced)
        return ws

    # ------------------------------------------------------------------ stages
    def _pre(self, ws, w, x):
        """pad-to-/16 resize (model.py:51-54), range heuristic (encoder.py:129-135), resize->224 + patchify."""
        if x.dtype == torch.uint8:  # the decoder's [B,H,W,3] image: one pass at 3 bytes per pixel, no fp32 copy (forward() checked H, W % 16 == 0)
            ops.input_stage_u8(x, w["enc"]["bn4x3"], 1e-5, ws.scratch, ws.affine, ws.raw224, ws.patches)
            if self.enhanced:
                ops.rgb_to_grid_u8(x, ws.gh, ws.gw, ws.feats[0], 256, 64)
            return 3 + self.enhanced
        xin = x
        if ws.xp is not None:
            xin = ops.resize_bilinear_nchw(x, ws.Hp, ws.Wp, out=ws.xp)
        if ws.Wp % 4 == 0:  # one pass over the input (scan + raw resample), then finalize + patchify on 0.6 MB/image
            ops.input_stage(xin, w["enc"]["bn4x3"], 1e-5, ws.scratch, ws.affine, ws.raw224, ws.patches)
        else:
            ops.input_range(xin, w["enc"]["bn4x3"], 1e-5, ws.scratch, ws.affine)
            ops.preprocess_patchify(xin, ws.affine, ws.patches)
        n = 3 + (ws.xp is not None)
        if self.enhanced:
            ops.rgb_to_grid(xin, ws.gh, ws.gw, ws.feats[0], 256, 64)  # enhanced_model.py:95-98
            n += 1
        return n

    def _vit(self, ws, w, want=None, on_tap=None):
        enc, B = w["enc"], ws.B
        n = 0
        # LayerNorm folded into the GEMMs around it when the CTA-pair kernel runs (M >= 1024; MDE_LN_FUSION=0 switches it off): the
        # producer of the residual stream (out-proj, fc2) also writes bf16(x) into ws.h and per-row partial sums, the QKV / fc1
        # epilogues normalise.  23 launches fewer per forward (the LayerNorm after the patch embedding stays a kernel).
        fused = ws.B * 197 >= L.GEMM_LN_MIN_ROWS and _LN_FUSION
        prod = dict(xb_out=ws.h, stats_out=ws.lnstats) if fused else {}
        ops.gemm(ws.patches, enc["patch_w"], enc["patch_b"], L.EPI_PATCH_EMBED, ws.xa, enc["pos"], enc["cls"])
        n += 1
        # residual stream: x_in --(attention)--> x_mid --(MLP)--> x_out.  x_in is dead once the out-proj
        # epilogue has read it, so two working buffers suffice; hooked layers (encoder.py:44-50) write
        # x_out straight into their tap buffer, which then serves as the next layer's x_in (no copy).
        work = (ws.xa, ws.xb)
        x_in = ws.xa
        for i, ly in enumerate(enc["layers"]):
            x_mid = work[1] if x_in is work[0] else work[0]
            x_out = ws.taps[HOOKS.index(i)] if i in HOOKS else (work[0] if x_mid is work[1] else work[1])
            if fused:
                qw, qb, qcs = ly["qkv_ln"]
                fw, fb, fcs = ly["fc1_ln"]
                if i == 0:
                    ops.layernorm(x_in, ly["ln1_w"], ly["ln1_b"], 1e-6, ws.h)
                    ops.gemm(ws.h, ly["qkv_w"], ly["qkv_b"], L.EPI_QKV_SPLIT, ws.q, ws.k, ws.vt)
                    n += 1
                else:
                    ops.gemm_ln(ws.h, qw, qb, L.EPI_QKV_SPLIT, ws.q, ws.k, ws.vt, stats_in=ws.lnstats, colsum=qcs)
                ops.vit_attention(ws.q, ws.k, ws.vt, B, ws.attn)
                ops.gemm_ln(ws.attn, ly["proj_w"], ly["proj_b"], L.EPI_BIAS_RESID_F32, x_mid, x_in, **prod)
                ops.gemm_ln(ws.h, fw, fb, L.EPI_BIAS_GELU_BF16, ws.mlp, stats_in=ws.lnstats, colsum=fcs)
                ops.gemm_ln(ws.mlp, ly["fc2_w"], ly["fc2_b"], L.EPI_BIAS_RESID_F32, x_out, x_mid, **(prod if i < 11 else {}))  # nothing normalises the last output (tv:157 is discarded)
                n += 5
            else:
                ops.layernorm(x_in, ly["ln1_w"], ly["ln1_b"], 1e-6, ws.h)
                ops.gemm(ws.h, ly["qkv_w"], ly["qkv_b"], L.EPI_QKV_SPLIT, ws.q, ws.k, ws.vt)
                ops.vit_attention(ws.q, ws.k, ws.vt, B, ws.attn)
                ops.gemm(ws.attn, ly["proj_w"], ly["proj_b"], L.EPI_BIAS_RESID_F32, x_mid, x_in)
                ops.layernorm(x_mid, ly["ln2_w"], ly["ln2_b"], 1e-6, ws.h)
                if _FUSED_MLP and ws.h.shape[0] >= L.GEMM_LN_MIN_ROWS:
                    # fc1 + GELU + fc2 + residual as ONE persistent kernel: fc2 tiles start per finished 256-row block (csrc/tc_mlp.cu)
                    ops.mlp(ws.h, ly["fc1_w"], ly["fc1_b"], ws.mlp, ly["fc2_w"], ly["fc2_b"], x_mid, x_out, ws.mlp_sync)
                    n -= 1
                else:
                    ops.gemm(ws.h, ly["fc1_w"], ly["fc1_b"], L.EPI_BIAS_GELU_BF16, ws.mlp)
                    ops.gemm(ws.mlp, ly["fc2_w"], ly["fc2_b"], L.EPI_BIAS_RESID_F32, x_out, x_mid)
                n += 7
            if want is not None and i == 0:
                want["block0_q"] = ws.q.clone()
                want["block0_attn"] = ws.attn.clone()
                want["block0_out"] = x_out.clone()
            if on_tap is not None and i in HOOKS:
                n += on_tap(HOOKS.index(i))
            x_in = x_out
        return n

    def _feature(self, ws, w, i):
        """encoder.py:147-168 for tap i: token->grid rows, 1x1 reduce, instance norm; enhanced: + the per-feature extras
        (attention block on features 1/2, uncertainty + refinement heads on feature 0) and the decoder skip conv that
        consumes this feature (decoder.py:126-137) — everything that depends on this tap only."""
        enc, B, gh, gw = w["enc"], ws.B, ws.gh, ws.gw
        G = gh * gw
        ops.tokens_to_grid(ws.taps[i], B, ws.strip, G, ws.grid_a[i])
        rw, rb = enc["reduce"][i]
        ops.gemm(ws.grid_a[i], rw, rb, L.EPI_BIAS_F32, ws.reduced[i])
        flat = ws.flat[i] if self.enhanced else None
        ops.instnorm_nhwc(ws.reduced[i], B, gh, gw, 256, 1e-5, ws.feats[i], 0, flat)
        n = 3
        feat = ws.feats[i]
        if self.enhanced:
            n += self._enhanced_feature(ws, w, i)
            if i in (1, 2):
                feat = ws.enh[i]
        if i < 3:
            lvl = 2 - i  # decoder level whose skip branch reads feature i
            self._conv(feat, gh, gw, 256, w["dec"][f"skip_{lvl}"], DEC_CH[lvl + 1], L.ACT_ELU, out=ws.d_s[lvl])
            n += 1
        return n

    def _conv(self, inp, h, w_, cin, wb, cout, act, out=None, out_f32=None, ksize=3, pad=L.PAD_REFLECT):
        ops.conv_nhwc(inp, h, w_, cin, wb[0], wb[1], cout, ksize, pad, act, out_bf16=out, out_f32=out_f32)

    def _enhanced_feature(self, ws, w, i):
        """enhanced_model.py:183-196: attention on features 1,2; uncertainty + refinement heads on feature 0."""
        enh, B, gh, gw = w["enh"], ws.B, ws.gh, ws.gw
        if i in (1, 2):
            pw, pb, gamma = enh["attn"][str(i)]
            ops.gemm(ws.flat[i], pw, pb, L.EPI_BIAS_F32, ws.qkv[i])
            ops.spatial_attention(ws.flat[i], ws.qkv[i], B, gh, gw, gamma, ws.enh[i])
            return 2
        if i != 0:
            return 0
        f0 = ws.feats[0]
        self._conv(f0, gh, gw, 256, enh["unc1"], 128, L.ACT_ELU, out=ws.u1, pad=L.PAD_ZERO)
        self._conv(ws.u1, gh, gw, 128, enh["unc2"], 128, L.ACT_ELU, out=ws.u2, pad=L.PAD_ZERO)
        self._conv(ws.u2, gh, gw, 128, enh["unc3"], 1, L.ACT_SIGMOID, out_f32=ws.unc, ksize=1)
        self._conv(f0, gh, gw, 320, enh["ref1"], 256, L.ACT_ELU, out=ws.r1, pad=L.PAD_ZERO)
        self._conv(ws.r1, gh, gw, 256, enh["ref2"], 128, L.ACT_ELU, out=ws.r2, pad=L.PAD_ZERO)
        self._conv(ws.r2, gh, gw, 128, enh["ref3"], 1, L.ACT_NONE, out_f32=ws.refn, ksize=1)
        return 6

    def _decoder(self, ws, w, feats):
        """decoder.py:90-173."""
        dec, gh, gw = w["dec"], ws.gh, ws.gw
        n = 0
        main = torch.cuda.current_stream()
        side = self.side if _SIDE_HEADS else None
        self._conv(feats[3], gh, gw, 256, dec["init_conv"], 256, L.ACT_ELU, out=ws.d_init)
        n += 1
        x, h, w_ = ws.d_init, gh, gw
        for i in range(4):
            ci, co = DEC_CH[i], DEC_CH[i + 1]
            self._conv(x, h, w_, ci, dec[f"upconv_{i}_0"], co, L.ACT_ELU, out=ws.d_a[i])
            if i == 3 and _FUSED_TAIL:
                # upsample x2 + upconv_3_1 + mars_adapt + dispconv_3 + sigmoid: one kernel, intermediates stay in shared memory
                ops.decoder_tail16(ws.d_a[3], h, w_, dec["upconv_3_1"], dec["mars_adapt"], dec["dispconv_3"], ws.disp[3])
                n += 2
                break
            s = ws.d_s[i] if i < 3 else None  # skip branch: computed with its feature (_feature)
            ops.upsample2x_concat(ws.d_a[i], h, w_, s, gh, gw, ws.d_cat[i])
            h, w_ = 2 * h, 2 * w_
            self._conv(ws.d_cat[i], h, w_, ws.d_cat[i].shape[-1], dec[f"upconv_{i}_1"], co, L.ACT_ELU, out=ws.d_x[i])
            x = ws.d_x[i]
            n += 3
            if i == 3:
                self._conv(x, h, w_, 16, dec["mars_adapt"], 16, L.ACT_ELU, out=ws.d_x3b)
                x = ws.d_x3b
                n += 1
            if i < 3 and side is not None:
                # the heads of levels 0-2 feed only the output stage: off the critical path, next to the next level's convs
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    self._conv(x, h, w_, co, dec[f"dispconv_{i}"], 1, L.ACT_SIGMOID, out_f32=ws.disp[i], pad=L.PAD_ZERO)
            else:
                self._conv(x, h, w_, co, dec[f"dispconv_{i}"], 1, L.ACT_SIGMOID, out_f32=ws.disp[i], pad=L.PAD_ZERO)
            n += 1
        if side is not None:
            main.wait_stream(side)
        return n

    def _body(self, ws, w, scale, want=None):
        """Everything between the input kernels and the output kernel (graph-captured).  The work that depends on one tap
        only (features 0..2 and their consumers) runs on a side stream, forked after the hooked ViT layer and joined before
        the decoder: those launches are small (210 positions per image) and fill the tail of the persistent GEMMs."""
        main = torch.cuda.current_stream()
        counts = []

        def on_tap(i):
            if i == 3 or self.side is None:
                return self._feature(ws, w, i)
            self.side.wait_stream(main)  # fork (first use) / order after the layer that produced tap i
            with torch.cuda.stream(self.side):
                counts.append(self._feature(ws, w, i))
            return 0

        n = self._vit(ws, w, want, on_tap)
        if self.side is not None:
            main.wait_stream(self.side)  # join
        n += sum(counts)
        feats = ws.feats
        if self.enhanced:
            feats = [ws.feats[0], ws.enh[1], ws.enh[2], ws.feats[3]]
        n += self._decoder(ws, w, feats)
        if self.enhanced:
            mul = 1.0 if scale is None else scale
            ops.enh_fuse(ws.disp[0], ws.disp[1], ws.disp[2], ws.refn, ws.B, 2 * ws.gh, 2 * ws.gw, ws.gh, ws.gw, mul, w["enh"]["fusion_w"], ws.fused)
            n += 1
        return n

    def _post(self, ws, scale, keys=None, half=False, out_buffers=None):
        """model.py:89-100 / enhanced_model.py:266-277 + layers.py:16-25 — one kernel writes every output map.

        ``keys`` (no reference counterpart; the caller-side pipeline's entry): produce only these outputs — the low-resolution maps
        are all computed, but a full-resolution map nobody reads is 4 MB per image of HBM writes.  ``half``: disparity-type keys are
        written as IEEE fp16 by the same kernel (round to nearest of the fp32 value, i.e. identical to ``ops.f32_to_f16`` of the
        fp32 map); ``out_buffers``: {key: preallocated tensor of the right dtype} to write into."""
        B, H0, W0 = ws.B, ws.H0, ws.W0
        dev = ws.scratch.device
        mul = 1.0 if scale is None else scale
        if not self.enhanced:
            sources = [(ws.disp[i], mul, (("disp", i),), None) for i in range(4)]
            rng = (0.01, 10.0)
        else:
            # ("disp_fused",0) holds the same values as ("disp",0) (enhanced_model.py:259-263); one buffer, two keys
            sources = [(ws.fused, 1.0, (("disp", 0), ("disp_fused", 0)), ("depth", 0))]
            sources += [(ws.disp[i], mul, (("disp", i),), ("depth", i)) for i in (1, 2, 3)]
            sources += [(ws.unc, 1.0, (("uncertainty", 0),), None), (ws.disp[0], 1.0, (("disp_unrefined", 0),), None)]
            rng = (1 / 100.0, 1 / 0.1)
        if keys is not None:
            known = {k for _, _, dk, zk in sources for k in dk + ((zk,) if zk else ())}
            bad = [k for k in keys if k not in known]
            if bad:
                raise KeyError(f"unknown output key(s) {bad}; this model produces {sorted(known)}")
        bufs = out_buffers or {}

        def new(key, dt):
            t = bufs.get(key)
            if t is None:
                return torch.empty(B, 1, H0, W0, device=dev, dtype=dt)
            if t.dtype != dt or tuple(t.shape) != (B, 1, H0, W0) or t.device != dev or not t.is_contiguous():
                raise ValueError(f"out_buffers[{key}] must be a contiguous {dt} tensor [{B},1,{H0},{W0}] on {dev}")
            return t

        out, maps, to_half = {}, [], []
        for src, m, disp_keys, depth_key in sources:
            wd = [k for k in disp_keys if keys is None or k in keys]
            wz = depth_key is not None and (keys is None or depth_key in keys)
            if not wd and not wz:
                continue
            dd = dh = dz = None
            if wd:
                if half:
                    dh = new(wd[0], F16)
                else:
                    dd = new(wd[0], F32)
                for k in wd:
                    out[k] = dh if half else dd
            if wz:
                dz = new(depth_key, F32) if not half else torch.empty(B, 1, H0, W0, device=dev, dtype=F32)
                if half:
                    to_half.append((depth_key, dz))
                else:
                    out[depth_key] = dz
            maps.append((src, m, dd, dz, dh))
        n = 0
        for i in range(0, len(maps), L.MAX_MAPS):
            ops.resize_maps(maps[i:i + L.MAX_MAPS], B, H0, W0, *rng)
            n += 1
        for k, dz in to_half:  # depth maps have no fused fp16 form: convert (rare: the pipeline's default keys are disparities)
            out[k] = ops.f32_to_f16(dz, new(k, F16))
            n += 1
        return out, n

    # ------------------------------------------------------------------ forward
    def forward(self, x, scale, want=None, keys=None, half=False, out_buffers=None):
        w = self.weights()
        dev = next(self.module.parameters()).device
        if not x.is_cuda or x.device != dev:
            raise RuntimeError(f"input is on {x.device} but the model is on {dev}")
        if x.dtype == torch.uint8:  # [B,H,W,3] as decoded (run_mars_depth.py:97-100); a pixel is float(u8)/255
            if x.dim() != 4 or x.shape[-1] != 3:
                raise ValueError("uint8 input must be [B,H,W,3]")
            x = x.contiguous()
            B, H0, W0, _ = x.shape
            if H0 % 16 or W0 % 16:  # the pad-to-/16 resize (model.py:51-54) works on the fp32 image
                with torch.cuda.device(dev):
                    x = ops.u8hwc_to_f32chw(x, torch.empty(B, 3, H0, W0, device=dev, dtype=F32))
        else:
            if x.dtype != F32 or not x.is_contiguous():
                x = x.to(F32).contiguous()
            B, _, H0, W0 = x.shape
        with torch.cuda.device(dev):
            ws = self.workspace(dev, B, H0, W0)
            if self.overlap and (self.side is None or self.side.device != dev):
                self.side = torch.cuda.Stream(device=dev)
            n = self._pre(ws, w, x)
            if want is not None or not self.use_graph:
                n += self._body(ws, w, scale, want)
            else:
                if ws.graph is None or ws.graph_scale != scale:
                    ws.body_launches = self._body(ws, w, scale)  # warm-up run (also the first result)
                    torch.cuda.current_stream().synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        self._body(ws, w, scale)
                    ws.graph, ws.graph_scale = g, scale
                ws.graph.replay()
                n += ws.body_launches
            out, k = self._post(ws, scale, keys, half, out_buffers)
            self.launches_per_forward = n + k
            if want is not None:
                want["ws"] = ws
        return out


# ---------------------------------------------------------------------- refinement.py modules
class _ModuleCache:
    """Packed weights (re-derived when a parameter changes or moves) + per-geometry buffers of a refinement.py module."""

    def __init__(self, module, pack):
        self.module, self.pack = module, pack
        self._fp, self._w, self.bufs = None, None, {}

    def weights(self):
        fp = _fingerprint(self.module)
        if fp != self._fp:
            dev = next(self.module.parameters()).device
            if dev.type != "cuda":
                raise RuntimeError("the module must live on a CUDA (sm_100) device: there is no CPU path — call module.to('cuda')")
            with torch.cuda.device(dev):
                ops.device_check()
                self._w = self.pack(self.module.state_dict())
            self._fp = fp
        return self._w

    def buffers(self, key, make):
        b = self.bufs.get(key)
        if b is None:
            if len(self.bufs) >= 2:
                self.bufs.pop(next(iter(self.bufs)))
            b = self.bufs[key] = make()
        return b


def _cache(module, pack):
    c = module.__dict__.get("_mde_cache")
    if c is None:
        c = module.__dict__["_mde_cache"] = _ModuleCache(module, pack)
    return c


def _as_f32(t, dev):
    if not t.is_cuda or t.device != dev:
        raise RuntimeError(f"input is on {t.device} but the module is on {dev}")
    return t if (t.dtype == F32 and t.is_contiguous()) else t.to(F32).contiguous()


def texture_refine_forward(module, depth_map, rgb_image):
    """refinement.py:57-107 (live lines :68-85,:92-107): pack -> merged 16->128 conv -> rgb_conv2 | depth_conv2 (written side by
    side = the channel concat of :93) -> combined_conv1 -> combined_conv2 -> output kernel (1x1 conv, edge gate, residual add)."""
    from .packing import pack_texture_refine

    c = _cache(module, pack_texture_refine)
    w = c.weights()
    dev = next(module.parameters()).device
    depth_map, rgb_image = _as_f32(depth_map, dev), _as_f32(rgb_image, dev)
    B, cd, H, W = depth_map.shape
    if cd != 1 or rgb_image.shape[0] != B or rgb_image.shape[1] != 3:
        raise ValueError("expected depth_map [B,1,H,W] and rgb_image [B,3,h,w]")
    with torch.cuda.device(dev):
        def make():
            e = lambda ch: torch.empty(B, H + 2, W + 2, ch, device=dev, dtype=F16)
            return {"in16": e(16), "f128": e(128), "cat": e(128), "gate": torch.empty(B, H, W, device=dev, dtype=F32),
                    "rgb": torch.empty(B, 3, H, W, device=dev, dtype=F32)}

        bf = c.buffers((B, H, W), make)
        if rgb_image.shape[2:] != depth_map.shape[2:]:  # refinement.py:71-72
            rgb_image = ops.resize_bilinear_nchw(rgb_image, H, W, out=bf["rgb"])
        ops.refine_pack_inputs(rgb_image, depth_map, bf["in16"], bf["gate"])
        conv = lambda inp, cin, wb, cout, out, in_coff=0, out_coff=0: ops.conv_nhwc(
            inp, H, W, cin, wb[0], wb[1], cout, 3, L.PAD_ZERO, L.ACT_ELU, out_bf16=out, out_coff=out_coff, in_coff=in_coff)
        conv(bf["in16"], 16, w["l1"], 128, bf["f128"])
        conv(bf["f128"], 64, w["rgb_conv2"], 64, bf["cat"], in_coff=0, out_coff=0)
        conv(bf["f128"], 64, w["depth_conv2"], 64, bf["cat"], in_coff=64, out_coff=64)
        conv(bf["cat"], 128, w["combined_conv1"], 64, bf["f128"], out_coff=0)      # f128[..., 0:64]  <- combined_conv1
        conv(bf["f128"], 64, w["combined_conv2"], 32, bf["f128"], in_coff=0, out_coff=64)  # f128[..., 64:96] <- combined_conv2
        out = torch.empty(B, 1, H, W, device=dev, dtype=F32)
        feat = bf["f128"]
        L.call("mde_refine_output", ops.C.c_void_p(feat.data_ptr() + 2 * 64), feat.shape[-1], ops._ptr(w["out_w"], F32), ops._ptr(w["out_b"], F32),
               ops._ptr(depth_map, F32), ops._ptr(bf["gate"], F32), B, H, W, ops._ptr(out, F32), ops._stream())
    return out


def depth_fusion_forward(module, depth_scales):
    """refinement.py:131-183: key lookup and target size as the reference, then one fused kernel."""
    from .packing import pack_depth_fusion

    scales = list(module.scales)
    key = lambda s: ("disp", s) if ("disp", s) in depth_scales else ("depth", s)
    maps = []
    for s in scales:
        if key(s) not in depth_scales:
            # the reference `continue`s and then fails in fusion_conv (channel mismatch, refinement.py:152-153,179)
            raise RuntimeError(f"MultiScaleDepthFusion: scale {s} missing from depth_scales (the reference fails in fusion_conv)")
        maps.append(depth_scales[key(s)])
    n = len(scales)
    if n > L.MAX_FUSION_SCALES:
        raise ValueError(f"at most {L.MAX_FUSION_SCALES} scales are built")
    c = _cache(module, lambda sd: pack_depth_fusion(sd, n))
    w = c.weights()
    dev = next(module.parameters()).device
    maps = [_as_f32(m, dev) for m in maps]
    tgt = depth_scales[key(min(scales))]
    H, W = tgt.shape[2:]
    for m in maps:
        if m.shape[0] != tgt.shape[0] or m.shape[1] != 1:
            raise ValueError("expected [B,1,h,w] maps of one batch size")
    with torch.cuda.device(dev):
        return ops.depth_fusion(maps, H, W, w["conf_w"], w["conf_b"], w["fuse_w"], w["fuse_b"])
