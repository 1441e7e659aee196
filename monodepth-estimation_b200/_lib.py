"""ctypes binding of ``libmde.so`` (the C-ABI declared in ``include/mde.h``).

The product path has no fallback: if the library is missing or the device is not
sm_100, calls raise.  ``lib()`` loads (and, if the sources are newer or the .so is
absent and nvcc exists, builds) the shared object once per process.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("MDE_LIB_PATH") or os.path.join(HERE, "libmde.so")  # MDE_LIB_PATH: A/B builds during kernel work

MDE_OK = 0
EPI_BIAS_BF16, EPI_BIAS_GELU_BF16, EPI_BIAS_RESID_F32, EPI_BIAS_F32, EPI_PATCH_EMBED, EPI_QKV_SPLIT = range(6)
ACT_NONE, ACT_ELU, ACT_SIGMOID = range(3)
PAD_REFLECT, PAD_ZERO = range(2)
VIT_SEQ, VIT_DIM, VIT_HEADS, VT_PITCH = 197, 768, 12, 208
RANGE_SCRATCH_FLOATS = 6 * 8192
INPUT_STAGE_NOALLOC_MAX_BATCH = 32  # MDE_INPUT_STAGE_NOALLOC_MAX_BATCH
GEMM_LN_MIN_ROWS = 1024  # MDE_GEMM_LN_MIN_ROWS
MAX_MAPS = 8
MAX_FUSION_SCALES = 4
MAX_TEMPORAL_MAPS = 16
BILATERAL_SCRATCH_FLOATS = 4128

_p, _i, _f, _d, _ll = C.c_void_p, C.c_int, C.c_float, C.c_double, C.c_longlong


class WorkspaceCfg(C.Structure):
    _fields_ = [("batch", _i), ("height", _i), ("width", _i), ("enhanced", _i)]


class MapDesc(C.Structure):
    _fields_ = [("src", _p), ("hs", _i), ("ws", _i), ("mul", _f), ("dst_disp", _p), ("dst_disp2", _p), ("dst_depth", _p), ("dst_half", _p)]


# name -> argtypes (every function returns int status unless listed in _RESTYPES)
SIGNATURES = {
    "mde_version": [],
    "mde_last_error": [],
    "mde_device_check": [],
    "mde_sm_count": [],
    "mde_input_range": [_p, _i, _i, _i, _p, _f, _p, _p, _p],
    "mde_resize_bilinear_nchw": [_p, _i, _i, _i, _i, _p, _i, _i, _p],
    "mde_u8hwc_to_f32chw": [_p, _i, _i, _i, _p, _p],
    "mde_f32_to_f16": [_p, _ll, _p, _p],
    "mde_workspace_bytes": [_p],
    "mde_preprocess_patchify": [_p, _i, _i, _i, _p, _p, _p],
    "mde_input_stage": [_p, _i, _i, _i, _p, _f, _p, _p, _p, _p, _p],
    "mde_input_stage_u8": [_p, _i, _i, _i, _p, _f, _p, _p, _p, _p, _p],
    "mde_input_stage_noalloc": [_p, _i, _i, _i, _p, _f, _p, _p, _p, _p],
    "mde_gemm_bf16": [_p, _i, _p, _i, _i, _i, _i, _i, _p, _p, _i, _p, _p, _p],
    "mde_mlp_bf16": [_p, _i, _p, _p, _p, _i, _p, _p, _p, _p, _i, _i, _i, _i, _p, _p],
    "mde_debug_gemm_prof": [_p],
    "mde_gemm_ln_bf16": [_p, _i, _p, _i, _i, _i, _i, _i, _p, _p, _i, _p, _p, _p, _p, _f, _p, _p, _p],
    "mde_layernorm_bf16": [_p, _i, _i, _p, _p, _f, _p, _p],
    "mde_vit_attention": [_p, _p, _p, _i, _p, _p],
    "mde_tokens_to_grid_bf16": [_p, _i, _i, _i, _p, _p],
    "mde_instnorm_nhwc": [_p, _i, _i, _i, _i, _f, _p, _i, _i, _p, _p],
    "mde_conv_nhwc": [_p, _i, _i, _i, _i, _i, _p, _p, _i, _i, _i, _i, _p, _i, _i, _p, _p],
    "mde_conv_nhwc_f16": [_p, _i, _i, _i, _i, _i, _p, _p, _i, _i, _i, _i, _p, _i, _i, _p, _p],
    "mde_upsample2x_concat_nhwc": [_p, _i, _i, _i, _i, _p, _i, _i, _i, _p, _p],
    "mde_decoder_tail16": [_p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p],
    "mde_rgb_to_grid_nhwc": [_p, _i, _i, _i, _i, _i, _p, _i, _i, _i, _p],
    "mde_rgb_to_grid_nhwc_u8": [_p, _i, _i, _i, _i, _i, _p, _i, _i, _i, _p],
    "mde_resize_maps": [C.POINTER(MapDesc), _i, _i, _i, _i, _f, _f, _p],
    "mde_spatial_attention": [_p, _p, _i, _i, _i, _p, _p, _p],
    "mde_enh_fuse": [_p, _p, _p, _p, _i, _i, _i, _i, _i, _f, _p, _p, _p],
    "mde_depth_fusion": [C.POINTER(_p), C.POINTER(_i), C.POINTER(_i), _i, _i, _i, _i, _p, _p, _p, _p, _p, _p],
    "mde_refine_pack_inputs": [_p, _p, _i, _i, _i, _p, _p, _p],
    "mde_refine_output": [_p, _i, _p, _p, _p, _p, _i, _i, _i, _p, _p],
    "mde_point_cloud_scratch_bytes": [_i, _i, _i],
    "mde_point_cloud": [_p, _i, _i, _i, _i, _d, _d, C.POINTER(_d), _i, _p, _p, _i, _p, _p, _p, _p],
    "mde_disp_to_depth": [_p, _ll, _d, _d, _p, _p],
    "mde_temporal_mean": [C.POINTER(_p), C.POINTER(_d), _i, _ll, _p, _p],
    "mde_bilateral_depth": [_p, _i, _i, _p, _f, _p, _p, _p, _p, _p],
    "mde_fill_depth_holes_scratch_bytes": [_i, _i],
    "mde_fill_depth_holes": [_p, _i, _i, _i, _p, _p, _p],
    "mde_guided_filter_scratch_bytes": [_i, _i],
    "mde_guided_filter_depth": [_p, _i, _i, _p, _i, _f, _p, _p, _p],
    "mde_transpose_bf16": [_p, _i, _i, _i, _i, _p, _i, _i, _p, _i, _p],
    "mde_batch_sum_f32": [_p, _i, _ll, _p, _i, _p],
    "mde_axpy_f32": [_p, _p, _f, _ll, _p],
    "mde_colsum_bf16": [_p, _i, _i, _i, _p, _i, _p, _p],
    "mde_reduce_scratch_bytes": [_i],
    "mde_gelu_bf16": [_p, _ll, _p, _p],
    "mde_gelu_bwd_bf16": [_p, _p, _ll, _p, _p],
    "mde_layernorm_bwd_scratch_floats": [_i],
    "mde_layernorm_bwd": [_p, _p, _p, _i, _i, _f, _p, _p, _p, _p, _i, _p, _p],
    "mde_vit_attention_bwd": [_p, _p, _p, _p, _i, _p, _p, _p],
    "mde_vit_attention_bwd_scratch_floats": [_i],
    "mde_im2col3x3_bf16": [_p, _i, _i, _i, _i, _i, _p, _p, _i, _p],
    "mde_col2im3x3_f32": [_p, _i, _i, _i, _i, _i, _i, _p, _p],
    "mde_bn_elu_train_fwd": [_p, _i, _i, _i, _p, _p, _f, _f, _p, _p, _p, _p, _p, _p, _p],
    "mde_bn_elu_train_bwd": [_p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "mde_bilinear_nhwc_f32": [_p, _i, _i, _i, _i, _p, _i, _i, _i, _i, _i, _p],
    "mde_bilinear_nhwc_bwd_f32": [_p, _i, _i, _i, _i, _i, _i, _p, _i, _i, _i, _p, _p],
    "mde_instnorm_f32": [_p, _i, _i, _i, _f, _p, _p],
    "mde_instnorm_bwd_f32": [_p, _p, _i, _i, _i, _f, _p, _p],
    "mde_tokens_to_grid_bwd_f32": [_p, _i, _i, _i, _p, _p],
    "mde_sigmoid_head_f32": [_p, _i, _ll, _f, _p, _p],
    "mde_sigmoid_head_bwd_f32": [_p, _i, _p, _ll, _f, _p, _i, _p],
    "mde_channel_stats_nchw": [_p, _i, _i, _ll, _f, _p, _p, _p, _p, _p, _p],
    "mde_depth_loss_scratch_floats": [],
    "mde_input_norm_bwd_scratch_floats": [_i],
    "mde_input_norm_bwd": [_p, _p, _i, _p, _p, _f, _p, _p, _p],
    "mde_depth_loss": [_p, _p, _p, _i, _i, _i, _f, _f, _p, _p, _p, _p],
    "mde_adam_step": [_p, _p, _p, _p, _ll, _f, _f, _f, _f, _i, _p],
    "mde_adam_step_dev": [_p, _p, _p, _p, _ll, _f, _f, _f, _f, _p, _p],
    "mde_adam_chunk_elems": [],
    "mde_adam_multi": [_p, _i, _p, _i, _p, _f, _f, _f, _p, _p],
}
_RESTYPES = {"mde_last_error": C.c_char_p, "mde_fill_depth_holes_scratch_bytes": C.c_longlong, "mde_guided_filter_scratch_bytes": C.c_longlong, "mde_workspace_bytes": C.c_longlong, "mde_point_cloud_scratch_bytes": C.c_longlong, "mde_layernorm_bwd_scratch_floats": C.c_longlong,
             "mde_depth_loss_scratch_floats": C.c_longlong, "mde_input_norm_bwd_scratch_floats": C.c_longlong, "mde_reduce_scratch_bytes": C.c_longlong, "mde_vit_attention_bwd_scratch_floats": C.c_longlong}

_lock = threading.Lock()
_lib = None


class MdeError(RuntimeError):
    pass


def load(build_if_needed: bool = True):
    """Return the loaded CDLL with prototypes set (no device work)."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if build_if_needed:
            from .build import build

            try:
                build()
            except Exception as e:
                # Falling back to an existing libmde.so is only sound when it was built from the sources as they are now: a stale
                # binary would make every parity result describe code that is no longer checked in.
                from .build import OUT, is_stale

                if SO_PATH != OUT:  # MDE_LIB_PATH names an A/B build: the in-tree build is irrelevant
                    pass
                elif not os.path.exists(SO_PATH) or is_stale():
                    raise MdeError(f"building libmde.so failed and no up-to-date binary exists ({SO_PATH}): {e}") from e
        if not os.path.exists(SO_PATH):
            raise MdeError(f"{SO_PATH} is missing: run `python monodepth-estimation_b200/build.py` (needs nvcc)")
        lib = C.CDLL(SO_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here == header/library mismatch
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, C.c_int)
        _lib = lib
        return lib


def check(status: int, what: str = ""):
    if status != MDE_OK:
        msg = load().mde_last_error()
        raise MdeError(f"{what} failed with status {status}: {msg.decode() if msg else '?'}")


# Number of compute entry points called so far (every ``call`` launches at least one kernel of libmde.so): callers take
# differences around a region — e.g. the trainer around its graph capture — to report how many launches a step replays.
N_CALLS = [0]


def call(name: str, *args):
    N_CALLS[0] += 1
    check(getattr(load(), name)(*args), name)
