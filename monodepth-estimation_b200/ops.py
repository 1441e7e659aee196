"""Tensor-level wrappers over the C-ABI (``include/mde.h``).

Each function takes CUDA torch tensors, checks dtype/contiguity, and enqueues the kernel on
``torch.cuda.current_stream()``.  PyTorch is used only to own device memory and streams; no
torch operator runs here.  Outputs are allocated by the caller (``engine.Workspace``) or, for
test convenience, by the wrapper when ``out`` is omitted.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L

BF16, F16, F32 = torch.bfloat16, torch.float16, torch.float32


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t, dtype=None):
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda:
        raise ValueError("mde ops need CUDA tensors (there is no CPU path)")
    if not t.is_contiguous():
        raise ValueError("mde ops need contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise ValueError(f"expected {dtype}, got {t.dtype}")
    return C.c_void_p(t.data_ptr())


def device_check():
    L.call("mde_device_check")


def sm_count() -> int:
    return int(L.load().mde_sm_count())


def workspace_bytes(batch: int, height: int, width: int, enhanced: bool) -> int:
    """``mde_workspace_bytes``: device scratch of one forward (what ``engine.Workspace`` allocates).  No device work."""
    cfg = L.WorkspaceCfg(int(batch), int(height), int(width), int(bool(enhanced)))
    n = int(L.load().mde_workspace_bytes(C.byref(cfg)))
    if n < 0:
        L.check(n, "mde_workspace_bytes")
    return n


def f32_to_f16(src, out=None):
    out = out if out is not None else torch.empty(src.shape, device=src.device, dtype=F16)
    L.call("mde_f32_to_f16", _ptr(src, F32), src.numel(), _ptr(out, F16), _stream())
    return out


def input_range(x, bn4x3, eps, scratch=None, affine=None):
    B, _, H, W = x.shape
    scratch = scratch if scratch is not None else torch.empty(L.RANGE_SCRATCH_FLOATS, device=x.device, dtype=F32)
    affine = affine if affine is not None else torch.empty(6, device=x.device, dtype=F32)
    L.call("mde_input_range", _ptr(x, F32), B, H, W, _ptr(bn4x3, F32), float(eps), _ptr(scratch, F32), _ptr(affine, F32), _stream())
    return affine


def resize_bilinear_nchw(src, Hd, Wd, out=None):
    B, Cc, Hs, Ws = src.shape
    out = out if out is not None else torch.empty(B, Cc, Hd, Wd, device=src.device, dtype=F32)
    L.call("mde_resize_bilinear_nchw", _ptr(src, F32), B, Cc, Hs, Ws, _ptr(out, F32), Hd, Wd, _stream())
    return out


def preprocess_patchify(x, affine, out=None):
    B, _, H, W = x.shape
    out = out if out is not None else torch.empty(B * 196, 768, device=x.device, dtype=BF16)
    L.call("mde_preprocess_patchify", _ptr(x, F32), B, H, W, _ptr(affine, F32), _ptr(out, BF16), _stream())
    return out


def gemm(a, w, bias, epilogue, out, aux0=None, aux1=None, M=None):
    """out = epilogue(a[M,K] @ w[N,K]^T); see ``mde_gemm_bf16`` for the aux operands."""
    M = a.shape[0] if M is None else M
    K = a.shape[1]
    N = w.shape[0]
    ldo = out.shape[-1] if epilogue <= L.EPI_BIAS_F32 else 0
    L.call("mde_gemm_bf16", _ptr(a, BF16), a.stride(0), _ptr(w, BF16), w.stride(0), M, N, K, epilogue, _ptr(bias, F32), _ptr(out), ldo,
           _ptr(aux0), _ptr(aux1), _stream())
    return out


def gemm_ln(a, w, bias, epilogue, out, aux0=None, aux1=None, stats_in=None, colsum=None, eps=1e-6, xb_out=None, stats_out=None):
    """``gemm`` with the neighbouring LayerNorm folded in (``mde_gemm_ln_bf16``): a producer passes ``xb_out`` + ``stats_out``,
    a consumer ``stats_in`` + ``colsum`` (and ``a`` = the producer's ``xb_out``, ``w``/``bias`` folded by ``packing.fold_layernorm``)."""
    M, K = a.shape
    N = w.shape[0]
    ldo = out.shape[-1] if epilogue <= L.EPI_BIAS_F32 else 0
    L.call("mde_gemm_ln_bf16", _ptr(a, BF16), a.stride(0), _ptr(w, BF16), w.stride(0), M, N, K, epilogue, _ptr(bias, F32), _ptr(out), ldo,
           _ptr(aux0), _ptr(aux1), _ptr(stats_in, F32), _ptr(colsum, F32), float(eps), _ptr(xb_out, BF16), _ptr(stats_out, F32), _stream())
    return out


def layernorm(x, gamma, beta, eps, out=None):
    rows, dim = x.shape
    out = out if out is not None else torch.empty(rows, dim, device=x.device, dtype=BF16)
    L.call("mde_layernorm_bf16", _ptr(x, F32), rows, dim, _ptr(gamma, F32), _ptr(beta, F32), float(eps), _ptr(out, BF16), _stream())
    return out


def vit_attention(q, k, vt, B, out=None):
    out = out if out is not None else torch.empty(B * L.VIT_SEQ, L.VIT_DIM, device=q.device, dtype=BF16)
    L.call("mde_vit_attention", _ptr(q, BF16), _ptr(k, BF16), _ptr(vt, BF16), B, _ptr(out, BF16), _stream())
    return out


def tokens_to_grid(tokens, B, strip_cls, G, out=None):
    out = out if out is not None else torch.empty(B * G, L.VIT_DIM, device=tokens.device, dtype=BF16)
    L.call("mde_tokens_to_grid_bf16", _ptr(tokens, F32), B, int(strip_cls), G, _ptr(out, BF16), _stream())
    return out


def instnorm_nhwc(x, B, gh, gw, Cc, eps, out_padded, out_coff=0, out_flat=None):
    L.call("mde_instnorm_nhwc", _ptr(x, F32), B, gh, gw, Cc, float(eps), _ptr(out_padded, BF16), out_padded.shape[-1], out_coff,
           _ptr(out_flat, BF16), _stream())
    return out_padded


def conv_nhwc(inp, h, w, cin, wpacked, bias, cout, ksize, padding, act, out_bf16=None, out_coff=0, out_f32=None, in_coff=0):
    """``in_coff``: first input channel read (a multiple of 8), so a conv can consume a channel slice of a wider tensor.
    The operand dtype (bf16, or fp16 -> ``mde_conv_nhwc_f16``) is taken from ``inp``; weights and output must match."""
    B = inp.shape[0]
    dt = inp.dtype
    if dt not in (BF16, F16):
        raise ValueError("conv operands must be bf16 or fp16")
    if in_coff % 8 or in_coff + cin > inp.shape[-1]:
        raise ValueError("in_coff must be a multiple of 8 and in_coff + cin <= channels")
    src = C.c_void_p(_ptr(inp, dt).value + 2 * in_coff)
    L.call("mde_conv_nhwc" if dt == BF16 else "mde_conv_nhwc_f16", src, B, h, w, cin, inp.shape[-1], _ptr(wpacked, dt), _ptr(bias, F32), cout, ksize,
           padding, act, _ptr(out_bf16, dt), 0 if out_bf16 is None else out_bf16.shape[-1], out_coff, _ptr(out_f32, F32), _stream())
    return out_bf16 if out_bf16 is not None else out_f32


def mlp(a, w1, b1, hidden, w2, b2, resid, out, sync):
    """tv:117-118 + the residual add (tv:119) in one launch: out = resid + (gelu(a @ w1.T + b1) @ w2.T + b2).  ``hidden`` [M, Hd] bf16 is
    scratch (it holds the GELU output afterwards), ``sync`` an int32 [128] buffer that is all zero before and after the call."""
    M, D = a.shape
    Hd = w1.shape[0]
    if w1.shape != (Hd, D) or w2.shape != (D, Hd) or hidden.shape != (M, Hd) or resid.shape != (M, D) or out.shape != (M, D) or sync.numel() < 128:
        raise ValueError("mlp: shape mismatch")
    L.call("mde_mlp_bf16", _ptr(a, BF16), a.stride(0), _ptr(w1, BF16), _ptr(b1, F32), _ptr(hidden, BF16), hidden.stride(0), _ptr(w2, BF16), _ptr(b2, F32),
           _ptr(resid, F32), _ptr(out, F32), out.stride(0), M, D, Hd, _ptr(sync, torch.int32), _stream())
    return out


def upsample2x_concat(a, h, w, s, hs, ws, out):
    B, ca = a.shape[0], a.shape[-1]
    cs = 0 if s is None else s.shape[-1]
    L.call("mde_upsample2x_concat_nhwc", _ptr(a, BF16), B, h, w, ca, _ptr(s, BF16), hs, ws, cs, _ptr(out, BF16), _stream())
    return out


def decoder_tail16(src_padded, h, w, wb1, wb2, wb3, out):
    """decoder.py:139-168, level 3: upsample x2 + upconv_3_1 + mars_adapt + dispconv_3 + sigmoid in one kernel."""
    B = src_padded.shape[0]
    if src_padded.shape[-1] != 16 or any(t[0].shape != (16, 144) for t in (wb1, wb2, wb3)):
        raise ValueError("decoder_tail16 serves the 16-channel last decoder level only")
    L.call("mde_decoder_tail16", _ptr(src_padded, BF16), B, h, w, _ptr(wb1[0], BF16), _ptr(wb1[1], F32), _ptr(wb2[0], BF16), _ptr(wb2[1], F32),
           _ptr(wb3[0], BF16), _ptr(wb3[1], F32), _ptr(out, F32), _stream())
    return out


def rgb_to_grid(x, gh, gw, out_padded, out_coff, cfill):
    B, _, H, W = x.shape
    L.call("mde_rgb_to_grid_nhwc", _ptr(x, F32), B, H, W, gh, gw, _ptr(out_padded, BF16), out_padded.shape[-1], out_coff, cfill, _stream())
    return out_padded


def resize_maps(maps, B, H, W, min_disp=0.01, max_disp=10.0):
    """maps: list of (src [B,hs,ws] f32, mul, dst_disp|None, dst_depth|None[, dst_half|None]); dst_half is an fp16 tensor that
    receives the resized disparity rounded to nearest (the host-link format of BatchPipeline)."""
    arr = (L.MapDesc * len(maps))()
    for i, m in enumerate(maps):
        src, mul, dd, dz = m[:4]
        dh = m[4] if len(m) > 4 else None
        arr[i].src = src.data_ptr()
        arr[i].hs, arr[i].ws = src.shape[-2], src.shape[-1]
        arr[i].mul = float(mul)
        arr[i].dst_disp = dd.data_ptr() if dd is not None else None
        arr[i].dst_disp2 = None
        arr[i].dst_depth = dz.data_ptr() if dz is not None else None
        arr[i].dst_half = _ptr(dh, F16).value if dh is not None else None
    L.call("mde_resize_maps", arr, len(maps), B, H, W, float(min_disp), float(max_disp), _stream())


def spatial_attention(x_flat, qkv, B, gh, gw, gamma, out_padded):
    L.call("mde_spatial_attention", _ptr(x_flat, BF16), _ptr(qkv, F32), B, gh, gw, _ptr(gamma, F32), _ptr(out_padded, BF16), _stream())
    return out_padded


def enh_fuse(d0, d1, d2, refinement, B, h0, w0, gh, gw, mul, fusion_w, fused):
    L.call("mde_enh_fuse", _ptr(d0, F32), _ptr(d1, F32), _ptr(d2, F32), _ptr(refinement, F32), B, h0, w0, gh, gw, float(mul),
           _ptr(fusion_w, F32), _ptr(fused, F32), _stream())
    return fused


def depth_fusion(maps, H, W, conf_w, conf_b, fuse_w, fuse_b, out=None):
    """maps: list of f32 [B,1,h,w] (or [B,h,w]) tensors in scale order; out f32 [B,1,H,W]."""
    n, B = len(maps), maps[0].shape[0]
    ptrs = (C.c_void_p * n)(*[_ptr(m, F32).value for m in maps])
    hs = (C.c_int * n)(*[m.shape[-2] for m in maps])
    ws = (C.c_int * n)(*[m.shape[-1] for m in maps])
    out = out if out is not None else torch.empty(B, 1, H, W, device=maps[0].device, dtype=F32)
    L.call("mde_depth_fusion", ptrs, hs, ws, n, B, H, W, _ptr(conf_w, F32), _ptr(conf_b, F32), _ptr(fuse_w, F32), _ptr(fuse_b, F32),
           _ptr(out, F32), _stream())
    return out


def refine_pack_inputs(rgb, depth, in16_padded, gate):
    B, _, H, W = rgb.shape
    L.call("mde_refine_pack_inputs", _ptr(rgb, F32), _ptr(depth, F32), B, H, W, _ptr(in16_padded, F16), _ptr(gate, F32), _stream())


def refine_output(feat_padded, w32, bias1, depth, gate, out):
    B, H, W = gate.shape
    L.call("mde_refine_output", _ptr(feat_padded, F16), feat_padded.shape[-1], _ptr(w32, F32), _ptr(bias1, F32), _ptr(depth, F32),
           _ptr(gate, F32), B, H, W, _ptr(out, F32), _stream())
    return out


def u8hwc_to_f32chw(src, out=None):
    """uint8 [B,H,W,3] -> float32 [B,3,H,W] / 255 (run_mars_depth.py:100-103 on the device)."""
    B, H, W, _ = src.shape
    out = out if out is not None else torch.empty(B, 3, H, W, device=src.device, dtype=F32)
    L.call("mde_u8hwc_to_f32chw", _ptr(src, torch.uint8), B, H, W, _ptr(out, F32), _stream())
    return out


def input_stage(x, bn4x3, eps, scratch, affine, raw224, patches):
    """encoder.py:129-144 in one pass: min/max scan + raw 224x224 resample, device-side range heuristic, affine + patchify."""
    B, _, H, W = x.shape
    L.call("mde_input_stage", _ptr(x, F32), B, H, W, _ptr(bn4x3, F32), float(eps), _ptr(scratch, F32), _ptr(affine, F32), _ptr(raw224, F32),
           _ptr(patches, BF16), _stream())
    return patches


def input_stage_u8(x_hwc, bn4x3, eps, scratch, affine, raw224, patches):
    """mde_input_stage from the uint8 [B,H,W,3] image (pixel = float(u8)/255): bit-identical outputs, 3 bytes read per pixel."""
    B, H, W, _ = x_hwc.shape
    L.call("mde_input_stage_u8", _ptr(x_hwc, torch.uint8), B, H, W, _ptr(bn4x3, F32), float(eps), _ptr(scratch, F32), _ptr(affine, F32), _ptr(raw224, F32),
           _ptr(patches, BF16), _stream())
    return patches


def rgb_to_grid_u8(x_hwc, gh, gw, out_padded, out_coff, cfill):
    B, H, W, _ = x_hwc.shape
    L.call("mde_rgb_to_grid_nhwc_u8", _ptr(x_hwc, torch.uint8), B, H, W, gh, gw, _ptr(out_padded, BF16), out_padded.shape[-1], out_coff, cfill, _stream())
    return out_padded


def point_cloud(src, intrinsics4, min_depth, max_depth, from_disp=False, f64=True, rgb=None):
    """terrain_reconstruction_fixed.py:18-108 for a batch: src f32 [B,H,W] (depth, or disparity with ``from_disp``) ->
    (points [B,H*W,3] f32/f64 capacity buffer, colors or None, counts int32 [B]); image b's points are points[b, :counts[b]]."""
    B, H, W = src.shape
    dev = src.device
    points = torch.empty(B, H * W, 3, device=dev, dtype=torch.float64 if f64 else F32)
    counts = torch.empty(B, device=dev, dtype=torch.int32)
    scratch = torch.empty(int(L.load().mde_point_cloud_scratch_bytes(B, H, W)), device=dev, dtype=torch.uint8)
    colors = None
    if rgb is not None:
        if rgb.shape != (B, H, W, 3) or rgb.dtype not in (torch.uint8, F32):
            raise ValueError(f"rgb must be uint8 or float32 [B,H,W,3] matching the depth maps, got {tuple(rgb.shape)} {rgb.dtype}")
        colors = torch.empty(B, H * W, 3, device=dev, dtype=rgb.dtype)
    k4 = (C.c_double * 4)(*[float(v) for v in intrinsics4])
    L.call("mde_point_cloud", _ptr(src, F32), B, H, W, int(bool(from_disp)), float(min_depth), float(max_depth), k4, int(bool(f64)),
           _ptr(points), _ptr(rgb), int(rgb is not None and rgb.dtype == F32), _ptr(colors), _ptr(counts, torch.int32), _ptr(scratch), _stream())
    return points, colors, counts


def temporal_mean(maps, weights, out=None):
    """refinement.py:354-388: maps = list of f32 tensors of one shape, weights = normalised python floats -> f64 tensor."""
    n = len(maps)
    ptrs = (C.c_void_p * n)(*[_ptr(m, F32).value for m in maps])
    w = (C.c_double * n)(*[float(v) for v in weights])
    out = out if out is not None else torch.empty(maps[0].shape, device=maps[0].device, dtype=torch.float64)
    L.call("mde_temporal_mean", ptrs, w, n, maps[0].numel(), _ptr(out, torch.float64), _stream())
    return out


def bilateral_depth(depth, sigma_color, space_w, guide=None, joint_lut=None, out=None):
    """refinement.py:190-243 on the device: depth f32 [H,W]; guide uint8 [H,W,3] selects the joint form."""
    H, W = depth.shape
    scratch = torch.empty(L.BILATERAL_SCRATCH_FLOATS, device=depth.device, dtype=F32)
    out = out if out is not None else torch.empty_like(depth)
    L.call("mde_bilateral_depth", _ptr(depth, F32), H, W, _ptr(guide, torch.uint8) if guide is not None else _ptr(None), float(sigma_color),
           _ptr(space_w, F32), _ptr(joint_lut, F32) if joint_lut is not None else _ptr(None), _ptr(scratch, F32), _ptr(out, F32), _stream())
    return out


def fill_depth_holes(depth, max_hole_size=10, out=None):
    """refinement.py:301-352 on the device: depth f32 [H,W] -> filled f32 [H,W]."""
    H, W = depth.shape
    n = int(L.load().mde_fill_depth_holes_scratch_bytes(H, W))
    scratch = torch.empty(n, device=depth.device, dtype=torch.uint8)
    out = out if out is not None else torch.empty_like(depth)
    L.call("mde_fill_depth_holes", _ptr(depth, F32), H, W, int(max_hole_size), _ptr(scratch, torch.uint8), _ptr(out, F32), _stream())
    return out


def guided_filter_depth(depth, radius, eps, rgb=None, out=None):
    """refinement.py:244-299 on the device: depth f32 [H,W], rgb uint8 [H,W,3] or None (self-guided)."""
    H, W = depth.shape
    n = int(L.load().mde_guided_filter_scratch_bytes(H, W))
    scratch = torch.empty(n, device=depth.device, dtype=torch.uint8)
    out = out if out is not None else torch.empty_like(depth)
    L.call("mde_guided_filter_depth", _ptr(depth, F32), H, W, _ptr(rgb, torch.uint8) if rgb is not None else _ptr(None), int(radius), float(eps),
           _ptr(scratch, torch.uint8), _ptr(out, F32), _stream())
    return out


def disp_to_depth(disp, min_depth, max_depth, out=None):
    """layers.py:16-25 (depth only) on the device, bit-identical to the torch CPU expression."""
    out = out if out is not None else torch.empty_like(disp)
    L.call("mde_disp_to_depth", _ptr(disp, F32), disp.numel(), float(min_depth), float(max_depth), _ptr(out, F32), _stream())
    return out


# ---------------------------------------------------------------------- backward of the ViT encoder block (csrc/backward.cu)
def transpose_bf16(src, ldt=None, want_t=True, want_rm=False, dst_t=None, dst_rm=None, col0=0, cols_t=None, row0_rm=0):
    """src [R,C] (f32 or bf16, rows contiguous) -> (dst_t bf16 [C, ldt] | None, dst_rm bf16 [R,C] | None).  Without ``dst_t`` a new
    buffer is made and zero-filled beyond R.  With ``dst_t`` given, the block lands at columns [col0, col0+R) and columns up to
    col0+cols_t are zero-filled (``cols_t`` defaults to R: nothing else is touched), so row blocks can be placed side by side;
    ``dst_rm`` likewise receives its rows at ``row0_rm``."""
    R, Cc = src.shape
    if src.dtype not in (F32, BF16) or src.stride(1) != 1:
        raise ValueError("transpose_bf16 takes float32 or bfloat16 with contiguous rows")
    if dst_t is None and want_t:
        ldt = ldt if ldt is not None else (R + 63) // 64 * 64
        dst_t = torch.empty(Cc, ldt, device=src.device, dtype=BF16)
        cols_t = ldt
    elif dst_t is not None:
        ldt = dst_t.shape[1]
        cols_t = R if cols_t is None else cols_t
    if dst_rm is None and want_rm:
        dst_rm = torch.empty(R, Cc, device=src.device, dtype=BF16)
    pt = C.c_void_p(_ptr(dst_t, BF16).value + 2 * col0) if dst_t is not None else _ptr(None)
    pr = C.c_void_p(_ptr(dst_rm, BF16).value + 2 * row0_rm * Cc) if dst_rm is not None else _ptr(None)
    L.call("mde_transpose_bf16", C.c_void_p(src.data_ptr()), int(src.dtype == F32), R, Cc, src.stride(0), pt, ldt or 0, cols_t or 0, pr, Cc, _stream())
    return dst_t, dst_rm


def batch_sum_f32(src, out=None, accumulate=False):
    """src f32 [B, ...] -> sum over the first dimension, fixed order."""
    B = src.shape[0]
    n = src.numel() // B
    out = out if out is not None else torch.empty(src.shape[1:], device=src.device, dtype=F32)
    L.call("mde_batch_sum_f32", _ptr(src, F32), B, n, _ptr(out, F32), int(accumulate), _stream())
    return out


def axpy_f32(y, x, a=1.0):
    L.call("mde_axpy_f32", _ptr(y, F32), _ptr(x, F32), float(a), y.numel(), _stream())
    return y


def _reduce_scratch(Cc, dev):
    """per-chunk partial sums of the two-stage column reductions (8-byte aligned: float64 storage)"""
    return torch.empty(int(L.load().mde_reduce_scratch_bytes(int(Cc))) // 8, device=dev, dtype=torch.float64)


def colsum_bf16(src, out=None, accumulate=False):
    R, Cc = src.shape
    out = out if out is not None else torch.empty(Cc, device=src.device, dtype=F32)
    L.call("mde_colsum_bf16", _ptr(src, BF16), R, Cc, src.stride(0), _ptr(out, F32), int(accumulate), _ptr(_reduce_scratch(Cc, src.device)), _stream())
    return out


def gelu_bf16(u, out=None):
    out = out if out is not None else torch.empty_like(u)
    L.call("mde_gelu_bf16", _ptr(u, BF16), u.numel(), _ptr(out, BF16), _stream())
    return out


def gelu_bwd_bf16(u, dy, out=None):
    out = out if out is not None else torch.empty_like(u)
    L.call("mde_gelu_bwd_bf16", _ptr(u, BF16), _ptr(dy, BF16), u.numel(), _ptr(out, BF16), _stream())
    return out


def layernorm_bwd(x, dy, gamma, eps, dres=None, dgamma=None, dbeta=None, accumulate=False):
    """-> (dx f32 [rows,dim] (+ dres), dgamma f32 [dim], dbeta f32 [dim])"""
    rows, dim = x.shape
    dx = torch.empty(rows, dim, device=x.device, dtype=F32)
    dgamma = dgamma if dgamma is not None else torch.empty(dim, device=x.device, dtype=F32)
    dbeta = dbeta if dbeta is not None else torch.empty(dim, device=x.device, dtype=F32)
    scratch = torch.empty(int(L.load().mde_layernorm_bwd_scratch_floats(rows)), device=x.device, dtype=F32)
    L.call("mde_layernorm_bwd", _ptr(x, F32), _ptr(dy, BF16), _ptr(gamma, F32), rows, dim, float(eps), _ptr(dres, F32) if dres is not None else _ptr(None),
           _ptr(dx, F32), _ptr(dgamma, F32), _ptr(dbeta, F32), int(accumulate), _ptr(scratch, F32), _stream())
    return dx, dgamma, dbeta


def vit_attention_bwd(q, k, vt, dout, B, out=None):
    out = out if out is not None else torch.empty(B * L.VIT_SEQ, 3 * L.VIT_DIM, device=q.device, dtype=BF16)
    scratch = torch.empty(int(L.load().mde_vit_attention_bwd_scratch_floats(B)), device=q.device, dtype=F32)
    L.call("mde_vit_attention_bwd", _ptr(q, BF16), _ptr(k, BF16), _ptr(vt, BF16), _ptr(dout, BF16), B, _ptr(out, BF16), _ptr(scratch, F32), _stream())
    return out


# ---------------------------------------------------------------------- the rest of the training step (csrc/train_ops.cu)
def _f32(*shape, dev):
    return torch.empty(*shape, device=dev, dtype=F32)


def im2col3x3(x, padding, want_col=True, want_colT=True):
    """x f32 [B,h,w,C] -> (col bf16 [B*h*w, 9C] | None, colT bf16 [9C, ldt] | None) with ldt = B*h*w rounded up to 64."""
    B, h, w, Cc = x.shape
    n = B * h * w
    ldt = (n + 63) // 64 * 64
    col = torch.empty(n, 9 * Cc, device=x.device, dtype=BF16) if want_col else None
    colT = torch.empty(9 * Cc, ldt, device=x.device, dtype=BF16) if want_colT else None
    L.call("mde_im2col3x3_bf16", _ptr(x, F32), B, h, w, Cc, padding, _ptr(col), _ptr(colT), ldt, _stream())
    return col, colT


def col2im3x3(dxcol, B, h, w, Cc, padding):
    dx = _f32(B, h, w, Cc, dev=dxcol.device)
    L.call("mde_col2im3x3_f32", _ptr(dxcol, F32), dxcol.shape[1], B, h, w, Cc, padding, _ptr(dx, F32), _stream())
    return dx


def bn_elu_train_fwd(z, Cc, gamma, beta, eps=1e-5, momentum=0.1, run_mean=None, run_var=None):
    """z f32 [N, ldz>=C] -> (y f32 [N,C], mean [C], rstd [C]); running statistics updated in place when given."""
    N, ldz = z.shape
    mean, rstd, y = _f32(Cc, dev=z.device), _f32(Cc, dev=z.device), _f32(N, Cc, dev=z.device)
    L.call("mde_bn_elu_train_fwd", _ptr(z, F32), N, Cc, ldz, _ptr(gamma, F32), _ptr(beta, F32), float(eps), float(momentum),
           _ptr(run_mean, F32) if run_mean is not None else _ptr(None), _ptr(run_var, F32) if run_var is not None else _ptr(None),
           _ptr(mean, F32), _ptr(rstd, F32), _ptr(y, F32), _ptr(_reduce_scratch(Cc, z.device)), _stream())
    return y, mean, rstd


def bn_elu_train_bwd(z, Cc, mean, rstd, gamma, beta, dy):
    N, ldz = z.shape
    dz, dg, db = _f32(N, Cc, dev=z.device), _f32(Cc, dev=z.device), _f32(Cc, dev=z.device)
    L.call("mde_bn_elu_train_bwd", _ptr(z, F32), N, Cc, ldz, _ptr(mean, F32), _ptr(rstd, F32), _ptr(gamma, F32), _ptr(beta, F32), _ptr(dy, F32),
           _ptr(dz, F32), _ptr(dg, F32), _ptr(db, F32), _ptr(_reduce_scratch(Cc, z.device)), _stream())
    return dz, dg, db


def bilinear_nhwc(src, hd, wd, align_corners, dst=None, coff=0):
    B, hs, ws, Cc = src.shape
    dst = dst if dst is not None else _f32(B, hd, wd, Cc, dev=src.device)
    L.call("mde_bilinear_nhwc_f32", _ptr(src, F32), B, hs, ws, Cc, _ptr(dst, F32), hd, wd, dst.shape[-1], coff, int(align_corners), _stream())
    return dst


def bilinear_nhwc_bwd(ddst, hs, ws, align_corners, Cc=None, coff=0):
    B, hd, wd, ldd = ddst.shape
    Cc = ldd if Cc is None else Cc
    dsrc = _f32(B, hs, ws, Cc, dev=ddst.device)
    scratch = _f32(B * hd * ws * Cc, dev=ddst.device) if hd >= 4 * hs and wd >= 4 * ws else None  # large magnification: separable form
    L.call("mde_bilinear_nhwc_bwd_f32", _ptr(ddst, F32), B, hd, wd, Cc, ldd, coff, _ptr(dsrc, F32), hs, ws, int(align_corners),
           _ptr(scratch, F32) if scratch is not None else _ptr(None), _stream())
    return dsrc


def instnorm_f32(x, eps=1e-5):
    B, G, Cc = x.shape
    y = torch.empty_like(x)
    L.call("mde_instnorm_f32", _ptr(x, F32), B, G, Cc, float(eps), _ptr(y, F32), _stream())
    return y


def instnorm_bwd_f32(x, dy, eps=1e-5):
    B, G, Cc = x.shape
    dx = torch.empty_like(x)
    L.call("mde_instnorm_bwd_f32", _ptr(x, F32), _ptr(dy, F32), B, G, Cc, float(eps), _ptr(dx, F32), _stream())
    return dx


def tokens_to_grid_bwd(dgrid, B, strip_cls, G):
    dtok = _f32(B * L.VIT_SEQ, L.VIT_DIM, dev=dgrid.device)
    L.call("mde_tokens_to_grid_bwd_f32", _ptr(dgrid, F32), B, int(strip_cls), G, _ptr(dtok, F32), _stream())
    return dtok


def sigmoid_head(z, n, mul):
    d = _f32(n, dev=z.device)
    L.call("mde_sigmoid_head_f32", _ptr(z, F32), z.shape[-1], n, float(mul), _ptr(d, F32), _stream())
    return d


def sigmoid_head_bwd(z, ddisp, mul, out=None):
    """dz[i] = ddisp[i] * mul * s(1-s); ``out`` f32 [n, ldo] receives it in column 0 (its other columns are left alone)."""
    n = ddisp.numel()
    out = out if out is not None else _f32(n, 1, dev=z.device)
    L.call("mde_sigmoid_head_bwd_f32", _ptr(z, F32), z.shape[-1], _ptr(ddisp, F32), n, float(mul), _ptr(out, F32), out.shape[-1], _stream())
    return out


def channel_stats_nchw(x, momentum=0.1, run_mean=None, run_var=None):
    """x f32 [B,C,H,W] -> (mean [C], biased var [C]); running statistics updated in place when given (nn.BatchNorm2d.train())."""
    B, Cc, H, W = x.shape
    mean, var = _f32(Cc, dev=x.device), _f32(Cc, dev=x.device)
    L.call("mde_channel_stats_nchw", _ptr(x, F32), B, Cc, H * W, float(momentum), _ptr(run_mean, F32) if run_mean is not None else _ptr(None),
           _ptr(run_var, F32) if run_var is not None else _ptr(None), _ptr(mean, F32), _ptr(var, F32), _ptr(_reduce_scratch(Cc, x.device)), _stream())
    return mean, var


def input_norm_bwd(dpatches, raw224, affine6, bn4x3, eps=1e-5):
    """dpatches f32 [B*196, 768], raw224 f32 [B,3,224,224], affine6 / bn4x3 as ``input_stage`` saw them -> f32 [6] = (dweight[3], dbias[3]) of
    ``encoder.normalize``."""
    B = raw224.shape[0]
    scratch = _f32(int(L.load().mde_input_norm_bwd_scratch_floats(B)), dev=raw224.device)
    out = _f32(6, dev=raw224.device)
    L.call("mde_input_norm_bwd", _ptr(dpatches, F32), _ptr(raw224, F32), B, _ptr(affine6, F32), _ptr(bn4x3, F32), float(eps), _ptr(scratch, F32),
           _ptr(out, F32), _stream())
    return out


def depth_loss(disp, img, rnd, w_depth=1.0, w_smooth=0.001):
    """disp f32 [B,1,H,W] or [B,H,W], img f32 [B,3,H,W], rnd f32 like disp -> (losses f32 [3] = total, depth, smoothness; ddisp like disp)."""
    B, H, W = disp.shape[0], disp.shape[-2], disp.shape[-1]
    scratch = _f32(int(L.load().mde_depth_loss_scratch_floats()), dev=disp.device)
    losses, dd = _f32(3, dev=disp.device), torch.empty_like(disp)
    L.call("mde_depth_loss", _ptr(disp, F32), _ptr(img, F32), _ptr(rnd, F32), B, H, W, float(w_depth), float(w_smooth), _ptr(scratch, F32),
           _ptr(losses, F32), _ptr(dd, F32), _stream())
    return losses, dd


def adam_step(p, g, m, v, lr, step, beta1=0.9, beta2=0.999, eps=1e-8):
    """``step``: int, or an int32 device tensor holding the step count (read at run time: CUDA-graph replays)."""
    if torch.is_tensor(step):
        L.call("mde_adam_step_dev", _ptr(p, F32), _ptr(g, F32), _ptr(m, F32), _ptr(v, F32), p.numel(), float(lr), float(beta1), float(beta2), float(eps),
               _ptr(step, torch.int32), _stream())
    else:
        L.call("mde_adam_step", _ptr(p, F32), _ptr(g, F32), _ptr(m, F32), _ptr(v, F32), p.numel(), float(lr), float(beta1), float(beta2), float(eps), int(step),
               _stream())
    return p


class AdamPlan:
    """Device tables of ``mde_adam_multi`` for a fixed set of (parameter, gradient, exp_avg, exp_avg_sq, group) tensors: built once
    (the pointers are baked in, as a captured CUDA graph needs anyway), launched once per step."""

    def __init__(self, entries):
        chunk = int(L.load().mde_adam_chunk_elems())
        rows, starts, total = [], [0], 0
        self.keep = []
        for p, g, m, v, group in entries:
            for t in (p, g, m, v):
                if not (t.is_cuda and t.is_contiguous() and t.dtype == F32 and t.numel() == p.numel()):
                    raise ValueError("Adam tensors must be contiguous fp32 CUDA tensors of one size")
            rows.append([p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), int(group)])
            total += (p.numel() + chunk - 1) // chunk
            starts.append(total)
            self.keep.append((p, g, m, v))
        dev = entries[0][0].device
        self.table = torch.tensor(rows, dtype=torch.int64).to(dev)
        self.starts = torch.tensor(starts, dtype=torch.int32).to(dev)
        self.n, self.chunks = len(rows), total

    def step(self, lr_groups, step_dev, beta1=0.9, beta2=0.999, eps=1e-8):
        L.call("mde_adam_multi", _ptr(self.table, torch.int64), self.n, _ptr(self.starts, torch.int32), self.chunks, _ptr(lr_groups, F32),
               float(beta1), float(beta2), float(eps), _ptr(step_dev, torch.int32), _stream())
