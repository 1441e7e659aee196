/* libmde.so — C-ABI of the B200-native depth-inference forward path.
 *
 * The reference (buzzpranav/Monodepth-Estimation) is pure Python/PyTorch and has no FFI
 * of its own; its boundary for this path is the nn.Module surface
 * (models/model.py:18,32; models/enhanced_model.py:120,152).  This header is the op layer a
 * Python (ctypes) or C++ host binds instead of the ATen/cuDNN/cuBLAS calls listed in
 * SURVEY.md §2a.  Each entry point names the reference lines it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer (cudaMalloc'ed by the caller; PyTorch owns all
 *     buffers); the kernels never allocate and never synchronise the host;
 *   - every call enqueues work on `stream` (a cudaStream_t) and returns immediately;
 *   - return value: MDE_OK or a negative mde_status; mde_last_error() gives the message
 *     of the last failure on the calling thread.  Unsupported shapes are errors — there is
 *     no fallback path;
 *   - "bf16" arguments are raw bfloat16 bit patterns (uint16_t);
 *   - "padded NHWC" = bf16 tensor [B][h+2][w+2][C] whose 1-pixel border holds the
 *     REFLECTION of the interior (so a reflect-pad 3x3 conv reads it directly and a
 *     zero-pad conv addresses the interior through TMA out-of-bounds zero fill).
 */
#ifndef MDE_H_
#define MDE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* mde_stream_t; /* == cudaStream_t */
typedef uint16_t mde_bf16;
typedef uint16_t mde_f16; /* IEEE binary16 bit pattern */

enum mde_status {
  MDE_OK = 0,
  MDE_ERR_BAD_ARGUMENT = -1, /* shape / alignment / null pointer */
  MDE_ERR_UNSUPPORTED = -2,  /* configuration not built */
  MDE_ERR_CUDA = -3,         /* a CUDA runtime/driver call failed */
  MDE_ERR_DEVICE = -4        /* current device is not sm_100 */
};

/* Epilogues of mde_gemm_bf16 (what is fused behind the tensor-core main loop). */
enum mde_gemm_epilogue {
  MDE_EPI_BIAS_BF16 = 0,      /* out bf16[M,ldo]  = acc + bias                              */
  MDE_EPI_BIAS_GELU_BF16 = 1, /* out bf16[M,ldo]  = gelu_erf(acc + bias)      (tv:MLPBlock) */
  MDE_EPI_BIAS_RESID_F32 = 2, /* out f32[M,ldo]   = acc + bias + resid[M,ldo] (tv:115,120)  */
  MDE_EPI_BIAS_F32 = 3,       /* out f32[M,ldo]   = acc + bias                              */
  MDE_EPI_PATCH_EMBED = 4,    /* rows (b,p) -> token row b*197+1+p, + bias + pos_emb; also writes the CLS rows */
  MDE_EPI_QKV_SPLIT = 5       /* q*(1/8), k -> token-major [B*197][768]; v -> transposed [B*12][64][MDE_VT_PITCH] */
};

enum mde_activation { MDE_ACT_NONE = 0, MDE_ACT_ELU = 1, MDE_ACT_SIGMOID = 2 };
enum mde_padding { MDE_PAD_REFLECT = 0, MDE_PAD_ZERO = 1 };

#define MDE_VIT_SEQ 197        /* tokens per image: 14*14 patches + CLS                      */
#define MDE_VIT_DIM 768
#define MDE_VIT_HEADS 12
#define MDE_VT_PITCH 208       /* row pitch (elements) of the transposed V buffer            */
#define MDE_RANGE_SCRATCH_FLOATS (6 * 8192)

int mde_version(void);
const char* mde_last_error(void);

/* Device scratch one forward of `batch` images of height x width needs (kernels never allocate; a host binding
 * this header sizes its workspace with this call).  The sum of every intermediate of the path — ViT residual stream and taps,
 * q/k/v, MLP hidden, grid features, decoder activations, low-resolution heads (+ the enhanced extras) — each rounded up to 256
 * bytes; the Python host (engine.Workspace) allocates exactly these buffers.  Returns a negative mde_status for a size the path
 * does not serve. */
typedef struct mde_workspace_cfg {
  int batch, height, width; /* the INPUT size (the pad-to-/16 resize of model.py:51-54 is applied inside) */
  int enhanced;             /* 0: models/model.py, 1: models/enhanced_model.py */
} mde_workspace_cfg;
long long mde_workspace_bytes(const mde_workspace_cfg* cfg);
/* MDE_OK iff the current CUDA device is compute capability 10.x. */
int mde_device_check(void);
/* Number of SMs of the current device (148 on B200). */
int mde_sm_count(void);

/* ---- encoder.py:129-135: eval BatchNorm2d(3) + whole-batch range heuristic ------------
 * One pass over x[B,3,H,W] (per-channel min/max), then a 1-block kernel that evaluates the
 * reference's two data-dependent branches ON DEVICE (no host sync) and emits the combined
 * per-channel affine  y = a_c * x + b_c  as affine6 = {a0,a1,a2,b0,b1,b2}.
 * bn4x3 = {weight[3], bias[3], running_mean[3], running_var[3]}.  scratch: MDE_RANGE_SCRATCH_FLOATS floats. */
int mde_input_range(const float* x, int B, int H, int W, const float* bn4x3, float eps,
                    float* scratch, float* affine6, mde_stream_t stream);

/* ---- F.interpolate(mode='bilinear', align_corners=False) on fp32 NCHW ------------------
 * model.py:51-54 (pad-to-/16 resize), enhanced_model.py:95-96 (rgb -> 14x15). */
int mde_resize_bilinear_nchw(const float* src, int B, int C, int Hs, int Ws, float* dst, int Hd, int Wd,
                             mde_stream_t stream);

/* ---- run_mars_depth.py:97-104 (caller side): uint8 HWC image batch [B][H][W][3] -> float32 CHW [B][3][H][W] / 255,
 * i.e. `torch.from_numpy(np.array(img)).float().permute(2,0,1) / 255.0` evaluated on the device (IEEE fp32 division,
 * bit-identical to the host expression) so the host->device copy moves 1 byte per sample instead of 4. H*W % 4 == 0. */
int mde_u8hwc_to_f32chw(const uint8_t* src, int B, int H, int W, float* dst, mde_stream_t stream);

/* ---- caller side, output direction (run_mars_depth.py:131-147 reads outputs[("disp",0)] back with .cpu()): fp32 -> IEEE fp16,
 * round-to-nearest, saturating — halves the device->host copy of a result map; n values, any n. */
int mde_f32_to_f16(const float* src, long long n, mde_f16* dst, mde_stream_t stream);

/* ---- encoder.py:139-144 + tv:_process_input: resize to 224^2 (bilinear, align_corners=False),
 * apply affine6, and lay the result out as the patch-embedding GEMM's A operand:
 * patches bf16 [B*196][768], column = c*256 + ky*16 + kx. */
int mde_preprocess_patchify(const float* x, int B, int H, int W, const float* affine6, mde_bf16* patches,
                            mde_stream_t stream);

/* ---- encoder.py:129-144 in ONE pass over x (the fused form of mde_input_range + mde_preprocess_patchify) -------------
 * resize(a*x + b) == a*resize(x) + b for the per-channel affine map of BN + range heuristic, so the input is read once:
 * the scan kernel takes per-channel min/max AND resamples the raw image to raw224 f32 [B][3][224][224]; the finalize
 * kernel evaluates the heuristic on device (affine6 as above); the patchify kernel applies it and writes
 * patches bf16 [B*196][768].  W % 4 == 0.  scratch: MDE_RANGE_SCRATCH_FLOATS floats. */
int mde_input_stage(const float* x, int B, int H, int W, const float* bn4x3, float eps, float* scratch, float* affine6,
                    float* raw224, mde_bf16* patches, mde_stream_t stream);

/* mde_input_stage fed with the image decoder's output: x_hwc uint8 [B][H][W][3] (run_mars_depth.py:94-100, np.array(PIL image)).
 * A pixel is float(u8) / 255 (the caller-side expression of :100-103, same IEEE division), so affine6, raw224 and patches are
 * bit-identical to mde_u8hwc_to_f32chw followed by mde_input_stage, while the image is read once at 3 bytes per pixel and no
 * fp32 copy of it is ever written.  W % 4 == 0, x_hwc 4-byte aligned. */
int mde_input_stage_u8(const uint8_t* x_hwc, int B, int H, int W, const float* bn4x3, float eps, float* scratch, float* affine6,
                       float* raw224, mde_bf16* patches, mde_stream_t stream);

/* mde_input_stage for callers that cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).  No
 * counterpart in the reference.  Same kernels, same outputs (bit for bit) as
 * mde_input_stage, but the partial-record scratch is a device array built into the library (126 KB per GPU that loads it, held until
 * the process exits), so the caller passes inputs and outputs only and nothing here calls cudaMalloc / cudaFree or any allocating
 * routine.  Limit: B <= MDE_INPUT_STAGE_NOALLOC_MAX_BATCH images (any H, W with W % 4 == 0); larger batches are an error
 * (MDE_ERR_BAD_ARGUMENT) — use mde_input_stage with caller scratch.  Nothing in the array must start at zero (each call writes
 * every record it later reads), but ALL calls on a device share it: two calls that overlap on the GPU (different streams, or
 * graph branches) race — keep them on one stream or order them with events. */
#define MDE_INPUT_STAGE_NOALLOC_MAX_BATCH 32
int mde_input_stage_noalloc(const float* x, int B, int H, int W, const float* bn4x3, float eps, float* affine6, float* raw224,
                            mde_bf16* patches, mde_stream_t stream);

/* ---- tcgen05 GEMM: D[M,N] = epilogue(A[M,K] * W[N,K]^T), bf16 in, fp32 accumulate --------
 * Replaces conv_proj (tv:277), in_proj / out_proj (tv:113), mlp.0 / mlp.3 (tv:118) and
 * channel_reduce (encoder.py:164).  A row-major [M][lda], W row-major [N][ldw] (K contiguous).
 * Requirements: K % 64 == 0, N % 128 == 0, lda/ldw % 8 == 0, 16-byte aligned pointers.
 * aux0/aux1/aux2 depend on the epilogue:
 *   RESID_F32 : aux0 = resid f32[M][ldo]
 *   PATCH_EMBED: aux0 = pos_embedding f32[197][768], aux1 = class_token f32[768]; out f32[B*197][768]
 *   QKV_SPLIT : out = q bf16, aux0 = k bf16, aux1 = vt bf16 (layouts above); M = B*197 */
int mde_gemm_bf16(const mde_bf16* A, int lda, const mde_bf16* W, int ldw, int M, int N, int K, int epilogue,
                  const float* bias, void* out, int ldo, const void* aux0, const void* aux1,
                  mde_stream_t stream);

/* ---- the same GEMM with the LayerNorm between two linear layers folded into their epilogues (tv:112-113, 117-118) -----
 * y = LN(x) W^T + b  ==  rstd * (bf16(x) (gamma.W)^T - mean * colsum) + (b + W beta)   with colsum[n] = sum_k (gamma.W)[n][k].
 * PRODUCER side (the GEMM that writes the fp32 residual stream x: RESID_F32, BIAS_F32 or PATCH_EMBED epilogue):
 *   xb_out bf16 [rows][N] receives bf16(x), ln_stats_out f32 [N/32][rows][2] (chunk-major, so warps access it coalesced) the (sum, sum of squares) of every
 *   32-column chunk of every row (fp32 values, before rounding).  For PATCH_EMBED rows = B*197.
 * CONSUMER side (the GEMM that follows the LayerNorm: BIAS_BF16, BIAS_GELU_BF16 or QKV_SPLIT epilogue): A = xb,
 *   W = bf16(gamma . W), bias = b + W beta, ln_stats_in = the producer's buffer (K/32 chunks per row, summed in a fixed
 *   order: deterministic), ln_colsum f32 [N] (computed from the ROUNDED folded weight), ln_eps.
 * Either side may be used alone (pass NULL for the other side's pointers).  Only the CTA-pair kernel has these epilogues:
 * M >= MDE_GEMM_LN_MIN_ROWS and N % 256 == 0, otherwise MDE_ERR_UNSUPPORTED (use mde_layernorm_bf16 + mde_gemm_bf16). */
#define MDE_GEMM_LN_MIN_ROWS 1024
int mde_gemm_ln_bf16(const mde_bf16* A, int lda, const mde_bf16* W, int ldw, int M, int N, int K, int epilogue,
                     const float* bias, void* out, int ldo, const void* aux0, const void* aux1,
                     const float* ln_stats_in, const float* ln_colsum, float ln_eps, mde_bf16* xb_out, float* ln_stats_out,
                     mde_stream_t stream);

/* ---- torchvision MLPBlock + residual (tv:117-119) as ONE launch -------------------------------------------------------------
 * out[M][D] (f32) = resid + gelu(a[M][D] * w1[Hd][D]^T + b1) * w2[D][Hd]^T + b2, bf16 operands, fp32 accumulation; the GELU output
 * is written to `hidden` [M][Hd] bf16 (caller scratch) and read back by the second product, whose tiles start as soon as the 256-row
 * block they read is complete (csrc/tc_mlp.cu).  Same arithmetic as mde_gemm_bf16(MDE_EPI_BIAS_GELU_BF16) followed by
 * mde_gemm_bf16(MDE_EPI_BIAS_RESID_F32): bit-identical results.  M >= 256 (at most 64 row blocks of 256), D and Hd multiples of 256,
 * Hd a multiple of D.  sync: 128 uint32 of device memory, ALL ZERO on entry; zero again when the kernel ends.  Two calls that
 * overlap on the GPU must not share a sync buffer. */
int mde_mlp_bf16(const mde_bf16* a, int lda, const mde_bf16* w1, const float* b1, mde_bf16* hidden, int ldh, const mde_bf16* w2,
                 const float* b2, const float* resid, float* out, int ldo, int M, int D, int Hd, unsigned* sync, mde_stream_t stream);

/* Development hook (tools/gemm_prof.py): while `counters` (device, 8 x int64 per CTA, zeroed by the caller) is set, the
 * CTA-pair GEMM adds clock64 deltas per CTA: [0] MMA thread waiting for a free accumulator, [1] MMA thread waiting for
 * operands, [2] epilogue warp waiting for an accumulator, [3] epilogue warp working, [4] tiles, [5] CTA lifetime.
 * Pass NULL to switch it off. */
int mde_debug_gemm_prof(long long* counters);

/* ---- LayerNorm over the last dim (tv:112,117; eps 1e-6), fp32 in, bf16 out ------------- */
int mde_layernorm_bf16(const float* x, int rows, int dim, const float* gamma, const float* beta, float eps,
                       mde_bf16* out, mde_stream_t stream);

/* ---- ViT self-attention core (tv:113, nn.MultiheadAttention, 12 heads x 64, S=197) -----
 * softmax(q k^T) v per (image, head) with q pre-scaled by 1/8; tcgen05 for both products,
 * softmax fused in between.  q,k: token-major [B*197][768] (head h in columns 64h..64h+63); vt: [B*12][64][MDE_VT_PITCH];
 * out bf16 [B*197][768] (token-major, head h in columns 64h..64h+63). */
int mde_vit_attention(const mde_bf16* q, const mde_bf16* k, const mde_bf16* vt, int B, mde_bf16* out,
                      mde_stream_t stream);

/* ---- encoder.py:69-110: token sequence -> grid rows, as the channel_reduce GEMM's A operand.
 * tokens f32 [B][197][768].  strip_cls=1: rows = tokens 1..196 (G=196, exact 224^2 input);
 * strip_cls=0: rows = tokens 0..196 then zero rows up to G (G=210, the 14x15 grid). out bf16 [B*G][768]. */
int mde_tokens_to_grid_bf16(const float* tokens, int B, int strip_cls, int G, mde_bf16* out, mde_stream_t stream);

/* ---- encoder.py:167 F.instance_norm (eps 1e-5, biased variance, no affine) over the G = gh*gw
 * positions of x f32 [B][G][C]; writes channels [out_coff, out_coff+C) of the padded NHWC bf16
 * tensor [B][gh+2][gw+2][out_cstride] and, if non-null, the flat copy bf16 [B*G][C]. */
int mde_instnorm_nhwc(const float* x, int B, int gh, int gw, int C, float eps, mde_bf16* out_padded,
                      int out_cstride, int out_coff, mde_bf16* out_flat, mde_stream_t stream);

/* ---- 3x3 (or 1x1) convolution as a tcgen05 implicit GEMM over padded NHWC ----------------
 * decoder.py:8-34 (reflect pad + conv + folded eval-BN + ELU), decoder.py:86-88,166 (zero-pad
 * heads + sigmoid), enhanced_model.py:45-54,87-103 (zero-pad convs + ELU / 1x1 heads).
 *   in      : padded NHWC bf16 [B][h+2][w+2][in_cstride], channels [0,cin) are read
 *             (cin a multiple of 64, or 32, or 16)
 *   wpacked : bf16 [cout_pad][ksize*ksize*cin], k = (ky*ksize+kx)*cin + c, BN scale folded in;
 *             cout_pad = cout rounded up to a multiple of 16
 *   bias    : f32 [cout_pad] (conv bias with BN shift folded in)
 *   out_bf16: padded NHWC bf16 [B][h+2][w+2][out_cstride], written at channels
 *             [out_coff, out_coff+cout) including the reflected border, or NULL
 *   out_f32 : f32 [B][h][w][cout] (unpadded), or NULL          — exactly one of the two */
int mde_conv_nhwc(const mde_bf16* in, int B, int h, int w, int cin, int in_cstride, const mde_bf16* wpacked, const float* bias,
                  int cout, int ksize, int padding, int activation, mde_bf16* out_bf16, int out_cstride,
                  int out_coff, float* out_f32, mde_stream_t stream);

/* Same kernel with IEEE fp16 operands and fp16 padded-NHWC output (10 mantissa bits instead of 8, fp32 accumulate;
 * values saturate at +-65504).  Used by the refinement.py conv chain, whose output is a small difference of larger
 * terms and needs the extra bits to stay inside the 1e-2 bar. */
int mde_conv_nhwc_f16(const mde_f16* in, int B, int h, int w, int cin, int in_cstride, const mde_f16* wpacked, const float* bias,
                      int cout, int ksize, int padding, int activation, mde_f16* out_f16, int out_cstride,
                      int out_coff, float* out_f32, mde_stream_t stream);

/* ---- decoder.py:118-145: x2 bilinear upsample (align_corners=True) of `a`, optional bilinear
 * resize (align_corners=True) of the skip branch `s` to the same size, channel concat.
 * a: padded NHWC [B][h+2][w+2][ca]; s: padded NHWC [B][hs+2][ws+2][cs] or NULL (cs=0);
 * out: padded NHWC [B][2h+2][2w+2][ca+cs] (border reflected). */
int mde_upsample2x_concat_nhwc(const mde_bf16* a, int B, int h, int w, int ca, const mde_bf16* s, int hs, int ws,
                               int cs, mde_bf16* out, mde_stream_t stream);

/* ---- decoder.py:139-168 for the last level (i = 3) as one kernel: x2 bilinear upsample (align_corners=True) of src (padded NHWC
 * bf16 [B][h+2][w+2][16], the output of upconv_3_0), upconv_3_1 and mars_adapt (3x3, reflection pad, eval BatchNorm folded, ELU),
 * dispconv_3 (3x3, zero pad) + sigmoid -> out f32 [B][2h][2w].  w1, w2, w3: packed [16][(ky*3+kx)*16 + c] bf16 (w3: row 0 = the
 * single output channel), b1, b2, b3: f32 [16].  Intermediates never leave shared memory. */
int mde_decoder_tail16(const mde_bf16* src_padded, int B, int h, int w, const mde_bf16* w1, const float* b1, const mde_bf16* w2, const float* b2,
                       const mde_bf16* w3, const float* b3, float* out, mde_stream_t stream);

/* ---- enhanced_model.py:95-98: rgb image f32 [B,3,H,W] -> bilinear (align_corners=False) to the gh x gw
 * grid, written as channels [out_coff, out_coff+3) of a padded NHWC bf16 tensor; channels up to
 * out_coff+cfill are zero-filled (K padding of the 259-channel refinement conv).  cfill, out_coff and out_cstride
 * are multiples of 8 channels (16-byte stores). */
int mde_rgb_to_grid_nhwc(const float* x, int B, int H, int W, int gh, int gw, mde_bf16* out_padded, int out_cstride,
                         int out_coff, int cfill, mde_stream_t stream);
/* The same from the 8-bit HWC image (pixel = float(u8) / 255, bit-identical to the fp32 form after mde_u8hwc_to_f32chw). */
int mde_rgb_to_grid_nhwc_u8(const uint8_t* x_hwc, int B, int H, int W, int gh, int gw, mde_bf16* out_padded, int out_cstride,
                            int out_coff, int cfill, mde_stream_t stream);

/* ---- output stage: model.py:89-100, enhanced_model.py:266-277, layers.py:16-25 ----------
 * For each of n_maps low-resolution fp32 maps [B][hs][ws]: multiply by `mul`, resize to
 * [B][H][W] (bilinear, align_corners=False) into dst_disp and, where dst_depth != NULL, also
 * write 1/(min_disp + (max_disp-min_disp)*disp) of the RESIZED value. */
typedef struct mde_map_desc {
  const float* src;
  int hs, ws;
  float mul;
  float* dst_disp;  /* may be NULL */
  float* dst_disp2; /* optional second copy (("disp_fused",0) aliases ("disp",0)), may be NULL */
  float* dst_depth; /* may be NULL */
  mde_f16* dst_half; /* optional IEEE fp16 copy of the resized disparity (round to nearest, saturating), may be NULL: what
                        BatchPipeline(out_dtype=float16) sends over the host link — no reference counterpart */
} mde_map_desc;
#define MDE_MAX_MAPS 8
int mde_resize_maps(const mde_map_desc* maps, int n_maps, int B, int H, int W, float min_disp, float max_disp,
                    mde_stream_t stream);

/* ---- enhanced_model.py:20-39 SelfAttentionBlock on a 14x15 (or 14x14) grid -----------------
 * x: flat bf16 [B*G][256] (instance-normed features); qkv: f32 [B*G][384] = the 1x1 projections
 * q(32)|k(32)|v(256)|pad(64) incl. bias (one mde_gemm_bf16 with MDE_EPI_BIAS_F32);
 * out = gamma * (v softmax(q^T k)^T) + x, written as padded NHWC bf16 [B][gh+2][gw+2][256].
 * gamma is read from device memory (f32[1]). */
int mde_spatial_attention(const mde_bf16* x, const float* qkv, int B, int gh, int gw, const float* gamma,
                          mde_bf16* out_padded, mde_stream_t stream);

/* ---- enhanced_model.py:199-215,244-263: low-resolution tail of the enhanced forward --------
 * refinement f32 [B][gh][gw] -> bilinear (align False) to disp0's size, added to disp0;
 * disp_unrefined = disp0 (unscaled); scaled maps d_k*mul; fused = sum_k softmax(w)_k *
 * resize(d_k*mul -> disp0 size).  d0,d1,d2: f32 [B][h0<<k][w0<<k].
 * Outputs (f32 [B][h0][w0]): fused. */
int mde_enh_fuse(const float* d0, const float* d1, const float* d2, const float* refinement, int B, int h0,
                 int w0, int gh, int gw, float mul, const float* fusion_w3, float* fused, mde_stream_t stream);

/* ---- models/refinement.py:131-183 MultiScaleDepthFusion.forward, one fused stencil kernel ---------
 * maps[i]: f32 [B][hs[i]][ws[i]] device pointers (host array of n_maps pointers, scale order); every map
 * is resampled to H x W (bilinear, align_corners=False; identity when already H x W), then
 *   conf_i = sigmoid(conv3x3_i(f_i) + conf_b[i])  (zero padding),  w_i = conf_i / (sum_j conf_j + 1e-8),
 *   out    = sum_i fuse_w[i] * f_i * w_i + fuse_b[0].
 * conf_w f32 [n_maps][9], conf_b f32 [n_maps], fuse_w f32 [n_maps], fuse_b f32 [1] (device); out f32 [B][H][W]. */
#define MDE_MAX_FUSION_SCALES 4
int mde_depth_fusion(const float* const* maps, const int* hs, const int* ws, int n_maps, int B, int H, int W,
                     const float* conf_w, const float* conf_b, const float* fuse_w, const float* fuse_b, float* out,
                     mde_stream_t stream);

/* ---- models/refinement.py:57-107 TextureAwareRefinementModule.forward (live lines :68-85,:92-107) ----
 * The module runs as  mde_refine_pack_inputs -> 5 x mde_conv_nhwc_f16 (zero pad, ELU) -> mde_refine_output:
 *   pack  : rgb f32 [B,3,H,W] + depth f32 [B,1,H,W] -> in16 fp16 padded NHWC [B][H+2][W+2][16] with channels
 *           {r,g,b,depth,0...}  (operand of the merged rgb_conv1|depth_conv1 implicit GEMM, 16 -> 128) and
 *           gate f32 [B][H][W] = sigmoid(5*sqrt(ex^2+ey^2)), ex/ey the frozen Sobel/3 convs of :41-49,:83-85;
 *   output: out f32 [B][H][W] = depth + (w32 . feat[...,0:32] + bias1[0]) * gate  (:98-105),
 *           feat = combined_conv2 output, padded NHWC fp16 [B][H+2][W+2][cstride]. */
int mde_refine_pack_inputs(const float* rgb, const float* depth, int B, int H, int W, mde_f16* in16_padded, float* gate,
                           mde_stream_t stream);
int mde_refine_output(const mde_f16* feat_padded, int cstride, const float* w32, const float* bias1, const float* depth,
                      const float* gate, int B, int H, int W, float* out, mde_stream_t stream);

/* ======================= callers on either side of the forward path =======================

 * ---- utils/terrain_reconstruction_fixed.py:18-70 TerrainReconstructor.create_point_cloud, :72-108 reconstruct_terrain,
 *      :168-171 MarsTerrainMapper.infer_depth's disp_to_depth (layers.py:16-25) -----------------------------------------
 * src f32 [B][H][W]: depth maps, or (from_disp != 0) the model's disparity maps, converted first as
 *   depth = 1 / (1/max_depth + (1/min_depth - 1/max_depth) * disp)   (float32 steps, as torch evaluates layers.py:21-24).
 * For every image, the pixels with min_depth < z < max_depth (float32 compare) are back-projected,
 *   x = (u - cx) * z / fx,  y = (v - cy) * z / fy,  intrinsics4 = {fx, fy, cx, cy} (HOST pointer),
 * and written COMPACTED IN RASTER ORDER to points[b][0..counts[b])[3]; image b's run starts at element b*H*W*3 of `points`
 * (capacity B*H*W*3 elements).  points_f64 != 0: float64 points — what NumPy >= 2 produces for the reference's default
 * (float64) intrinsics, u/v/z being float32; points_f64 == 0: float32 points, every step rounded to float32 (float32
 * intrinsics).  Both are bit-identical to the NumPy expressions.  rgb (optional, [B][H][W][3] uint8, or float32 when
 * rgb_f32 != 0) is compacted the same way into colors (:99-104).  counts: int32 [B] on the device.  scratch:
 * mde_point_cloud_scratch_bytes(B,H,W) bytes.  One pass: 4 B read per pixel, 12/24 B written per surviving point. */
long long mde_point_cloud_scratch_bytes(int B, int H, int W);
int mde_point_cloud(const float* src, int B, int H, int W, int from_disp, double min_depth, double max_depth,
                    const double* intrinsics4, int points_f64, void* points, const void* rgb, int rgb_f32, void* colors,
                    int* counts, void* scratch, mde_stream_t stream);

/* ---- layers.py:16-25 disp_to_depth as MarsTerrainMapper.infer_depth calls it (terrain_reconstruction_fixed.py:168-171) ----
 * depth[i] = 1 / (1/max_depth + (1/min_depth - 1/max_depth) * disp[i]); the two scalars are formed in double and rounded to
 * float32, the multiply, add and divide round to float32 one by one (what torch does for a float32 tensor and Python-float
 * scalars): bit-identical to the reference on CPU.  16-byte aligned pointers. */
int mde_disp_to_depth(const float* disp, long long count, double min_depth, double max_depth, float* depth, mde_stream_t stream);

/* ---- models/refinement.py:354-388 DepthQualityEnhancement.temporal_consistency_filter ----
 * out f64 [count] = sum_n weights[n] * maps[n][i], float32 maps times float64 weights added map by map in float64
 * (np.sum(stacked * weights[:, None, None], axis=0)); weights are already normalised by the caller (:378-381).
 * maps / weights are HOST arrays (of device pointers / doubles), n <= MDE_MAX_TEMPORAL_MAPS. */
#define MDE_MAX_TEMPORAL_MAPS 16
int mde_temporal_mean(const float* const* maps, const double* weights, int n, long long count, double* out,
                      mde_stream_t stream);

/* ---- models/refinement.py:190-243 DepthQualityEnhancement.bilateral_depth_filter ----
 * depth f32 [H][W] -> out f32 [H][W]:  n = (depth - min) / (max - min + 1e-8);  f = bilateral(n);  out = f * (max - min) + min,
 * min/max taken on the device (no host round trip).  guide == NULL: cv2.bilateralFilter's float32 algorithm (window of
 * radius <= 7 clipped to a circle, colour weight from a 4096-bin linearly interpolated table of exp(-0.5 (dv/sigma_color)^2)
 * built from the range of n, BORDER_REFLECT_101; the input is returned when max(n) - min(n) < FLT_EPSILON).  guide != NULL
 * (uint8 [H][W][3]): cv2.ximgproc.jointBilateralFilter — colour weight joint_lut[|dr|+|dg|+|db|] of the guide, centre
 * pixel included (joint_lut: f32 [768] = exp(-0.5 (i/sigma_color)^2); ximgproc is not installed in the build image: this
 * branch restates the published algorithm, parity unpinned).  space_w: f32 [15][15], exp(-0.5 (r/sigma_space)^2) for
 * r = sqrt(dy^2+dx^2) <= radius, 0 elsewhere.  scratch: MDE_BILATERAL_SCRATCH_FLOATS floats.  out != depth. */
#define MDE_BILATERAL_SCRATCH_FLOATS 4128
int mde_bilateral_depth(const float* depth, int H, int W, const uint8_t* guide, float sigma_color, const float* space_w,
                        const float* joint_lut, float* scratch, float* out, mde_stream_t stream);

/* DepthQualityEnhancement.fill_depth_holes (models/refinement.py:301-352) on one map depth f32 [H][W]: pixels that are not finite
 * or <= 0 are holes; every 4-connected hole of at most max_hole_size^2 pixels is filled with the median of the ORIGINAL map over its
 * ring (3 cross-shaped binary dilations of the hole minus the hole: np.median semantics, NaN if the ring holds a NaN), then every
 * originally invalid pixel takes the 3x3 median (replicated border, cv2.medianBlur) of that result.  max_hole_size 0..13 (the
 * reference's default is 10).  scratch: mde_fill_depth_holes_scratch_bytes(H, W) bytes, 256-byte aligned.  out may not alias depth. */
long long mde_fill_depth_holes_scratch_bytes(int H, int W);
int mde_fill_depth_holes(const float* depth, int H, int W, int max_hole_size, void* scratch, float* out, mde_stream_t stream);

/* DepthQualityEnhancement.edge_preserving_smoothing (models/refinement.py:244-299): min/max-normalise depth f32 [H][W], guided filter
 * (cv2.ximgproc.guidedFilter: box-filter form of He et al., window 2*radius+1, BORDER_REFLECT) with guide = the normalised map
 * (rgb == NULL) or the 8-bit grey image of rgb u8 [H][W][3] (cv2.cvtColor RGB2GRAY), regulariser eps, then back to the map's range.
 * The reference passes radius = min(H, W) / 16 and eps = lambda_param.  opencv-contrib is absent from the build image: the filter
 * follows the published algorithm ("parity unpinned").  scratch: mde_guided_filter_scratch_bytes(H, W) bytes, 256-byte aligned. */
long long mde_guided_filter_scratch_bytes(int H, int W);
int mde_guided_filter_depth(const float* depth, int H, int W, const uint8_t* rgb, int radius, float eps, void* scratch, float* out,
                            mde_stream_t stream);

/* ==== backward of the ViT encoder block (train/mars_adapter.py:131-140 calls loss.backward() through
 * torchvision's EncoderBlock, tv:110-120).  GEMM-shaped gradients reuse mde_gemm_bf16 on transposed operands:
 *   dX = dY W    : A = dY [M][N],     W operand = W^T [K][N]
 *   dW = dY^T X  : A = dY^T [N][Mp],  W operand = X^T [K][Mp],  Mp = M rounded up to a multiple of 64 (zero filled)
 * Everything below is deterministic (fixed-order reductions, no atomics). ==== */

/* src [R][lds] (f32 when src_is_f32, else bf16) -> dst_t bf16 [C][ldt]: dst_t[c][r] = src[r][c], zero for R <= r < cols_t (the K
 * padding of a dW GEMM; R <= cols_t <= ldt, columns beyond cols_t are left alone so that row blocks can be transposed side by
 * side); and/or dst_rm bf16 [R][ldr] = cast(src).  Either destination may be NULL. */
int mde_transpose_bf16(const void* src, int src_is_f32, int R, int C, int lds, mde_bf16* dst_t, int ldt, int cols_t, mde_bf16* dst_rm,
                       int ldr, mde_stream_t stream);

/* Scratch of the column reductions below (mde_colsum_bf16, mde_bn_elu_train_fwd / _bwd, mde_channel_stats_nchw): per-chunk
 * partial sums that a second kernel adds in a fixed order.  8-byte aligned, mde_reduce_scratch_bytes(C) bytes for C columns. */
long long mde_reduce_scratch_bytes(int C);

/* bias gradient of a Linear (tv:113,118): out[c] (+)= sum_r src[r][c]. */
int mde_colsum_bf16(const mde_bf16* src, int R, int C, int lds, float* out, int accumulate, void* scratch, mde_stream_t stream);

/* nn.GELU (exact erf, tv:118) on a bf16 buffer, and its backward du = dy * (Phi(u) + u phi(u)).  count % 8 == 0. */
int mde_gelu_bf16(const mde_bf16* u, long long count, mde_bf16* y, mde_stream_t stream);
int mde_gelu_bwd_bf16(const mde_bf16* u, const mde_bf16* dy, long long count, mde_bf16* du, mde_stream_t stream);

/* LayerNorm backward (tv:112,117; dim = 768): x f32 [rows][dim] (the layer's input), dy bf16 [rows][dim].
 * dx[r] = rstd (g - mean(g) - xhat mean(g xhat)) with g = dy gamma, plus dres[r] when dres != NULL (the gradient that reaches x
 * through the residual connection); dgamma / dbeta (+)= column sums of dy xhat / dy.  scratch: mde_layernorm_bwd_scratch_floats(rows). */
long long mde_layernorm_bwd_scratch_floats(int rows);
int mde_layernorm_bwd(const float* x, const mde_bf16* dy, const float* gamma, int rows, int dim, float eps, const float* dres, float* dx,
                      float* dgamma, float* dbeta, int accumulate, float* scratch, mde_stream_t stream);

/* out[i] (+)= sum over b of src[b][i] (gradients of pos_embedding / class_token, tv:156,296: shared by the images of a batch), and
 * y += a x (the gradient of a hooked block's output is the sum of what its two consumers send back, encoder.py:60-67). */
int mde_batch_sum_f32(const float* src, int B, long long n, float* out, int accumulate, mde_stream_t stream);
int mde_axpy_f32(float* y, const float* x, float a, long long n, mde_stream_t stream);

/* Backward of mde_vit_attention (tv:113): q (pre-scaled by 1/8), k token-major bf16 [B*197][768], vt bf16 [B*12][64][208] as the
 * forward wrote them, dout bf16 [B*197][768] = gradient of the attention output.  dqkv bf16 [B*197][2304] = (dq | dk | dv) per
 * token with respect to in_proj's output (dq includes the 1/8), i.e. the dY of the in_proj gradient GEMMs.  q, k, vt, dout 16-byte
 * aligned; scratch: mde_vit_attention_bwd_scratch_floats(B) floats (the softmax row statistics between the two passes). */
long long mde_vit_attention_bwd_scratch_floats(int B);
int mde_vit_attention_bwd(const mde_bf16* q, const mde_bf16* k, const mde_bf16* vt, const mde_bf16* dout, int B, mde_bf16* dqkv, float* scratch,
                          mde_stream_t stream);

/* ==== the rest of the training step (train/mars_adapter.py:117-153,209-247), train-mode decoder path of ('disp', 0) ====
 * Train-mode activations are plain fp32 NHWC [B][h][w][C]; a 3x3 convolution is im2col + mde_gemm_bf16 in both directions. */

/* decoder.py:8-19 (pad 1 + 3x3): x f32 [B][h][w][C] -> col bf16 [B*h*w][9C] (k = (ky*3+kx)*C + c) and/or colT bf16 [9C][ldt]
 * (zero beyond B*h*w: the K padding of the weight-gradient GEMM).  padding = MDE_PAD_REFLECT | MDE_PAD_ZERO. */
int mde_im2col3x3_bf16(const float* x, int B, int h, int w, int C, int padding, mde_bf16* col, mde_bf16* colT, int ldt, mde_stream_t stream);
/* the adjoint: dxcol f32 [B*h*w][ldc] -> dx f32 [B][h][w][C] (reflected border rows/columns fold back onto rows 1 and h-2). */
int mde_col2im3x3_f32(const float* dxcol, int ldc, int B, int h, int w, int C, int padding, float* dx, mde_stream_t stream);

/* nn.BatchNorm2d in train mode + nn.ELU (decoder.py:22-34) on z f32 [N][ldz] (N = B*h*w rows, C channels): batch mean / biased
 * variance -> mean, rstd (kept for the backward); y f32 [N][C] = ELU(gamma zhat + beta); running_mean / running_var updated
 * with `momentum` and the unbiased variance when given.  Backward: dy f32 [N][C] -> dz f32 [N][C], dgamma, dbeta.
 * scratch: mde_reduce_scratch_bytes(C). */
int mde_bn_elu_train_fwd(const float* z, int N, int C, int ldz, const float* gamma, const float* beta, float eps, float momentum, float* run_mean,
                         float* run_var, float* mean, float* rstd, float* y, void* scratch, mde_stream_t stream);
int mde_bn_elu_train_bwd(const float* z, int N, int C, int ldz, const float* mean, const float* rstd, const float* gamma, const float* beta,
                         const float* dy, float* dz, float* dgamma, float* dbeta, void* scratch, mde_stream_t stream);

/* F.interpolate(mode='bilinear') on fp32 NHWC and its adjoint (decoder.py:118-142 align_corners=True; model.py:95-100
 * align_corners=False).  The destination may be a channel slice [coff, coff+C) of rows ldd wide (the concat of decoder.py:145). */
int mde_bilinear_nhwc_f32(const float* src, int B, int hs, int ws, int C, float* dst, int hd, int wd, int ldd, int coff, int align_corners,
                          mde_stream_t stream);
/* scratch: NULL (one gather kernel), or B*hd*ws*C floats: two separable gathers (along x, then y; same order of additions), which
 * is what a large magnification with few source pixels wants (the disparity map back from the input size to level 0). */
int mde_bilinear_nhwc_bwd_f32(const float* ddst, int B, int hd, int wd, int C, int ldd, int coff, float* dsrc, int hs, int ws, int align_corners,
                              float* scratch, mde_stream_t stream);

/* F.instance_norm (encoder.py:167: no affine, biased variance) on x f32 [B][G][C] and its backward; the adjoint of
 * mde_tokens_to_grid_bf16 (encoder.py:69-110): dgrid f32 [B][G][768] -> dtokens f32 [B][197][768]. */
int mde_instnorm_f32(const float* x, int B, int G, int C, float eps, float* y, mde_stream_t stream);
int mde_instnorm_bwd_f32(const float* x, const float* dy, int B, int G, int C, float eps, float* dx, mde_stream_t stream);
int mde_tokens_to_grid_bwd_f32(const float* dgrid, int B, int strip_cls, int G, float* dtokens, mde_stream_t stream);

/* decoder.py:166 + model.py:63-92: disp = mul * sigmoid(z[i*ldz]) and dz = ddisp * mul * s (1 - s). */
int mde_sigmoid_head_f32(const float* z, int ldz, long long n, float mul, float* disp, mde_stream_t stream);
int mde_sigmoid_head_bwd_f32(const float* z, int ldz, const float* ddisp, long long n, float mul, float* dz, int ldo, mde_stream_t stream);

/* encoder.py:129 in train mode: per-channel batch mean and BIASED variance of x f32 [B][C][hw] (what nn.BatchNorm2d normalises with),
 * running statistics updated with `momentum` and the unbiased variance when given.  Passing (weight, bias, mean, var) as the
 * bn4x3 block of mde_input_stage then evaluates the train-mode forward. */
int mde_channel_stats_nchw(const float* x, int B, int C, long long hw, float momentum, float* run_mean, float* run_var, float* mean, float* var,
                           void* scratch, mde_stream_t stream);

/* Gradient of encoder.normalize.{weight,bias} (encoder.py:129-137; the six scalars in front of the patch embedding).  dpatches f32
 * [B*196][768] = gradient of the patch rows mde_input_stage wrote (columns (c, ky, kx): dY of conv_proj times its weight), raw224 /
 * affine6 / bn4x3 / eps as mde_input_stage saw them.  The range heuristic's scale is a Python branch in the reference (no gradient
 * through the choice); which one was taken is read back from affine6.  dgamma_dbeta6 = (dweight[3], dbias[3]).
 * scratch: mde_input_norm_bwd_scratch_floats(B) floats. */
long long mde_input_norm_bwd_scratch_floats(int B);
int mde_input_norm_bwd(const float* dpatches, const float* raw224, int B, const float* affine6, const float* bn4x3, float eps, float* scratch,
                       float* dgamma_dbeta6, mde_stream_t stream);

/* MarsDomainAdapter.compute_losses (train/mars_adapter.py:209-247) with get_smooth_loss (layers.py:202-215): disp f32 [B][H][W] (=
 * outputs[('disp',0)]), img f32 [B][3][H][W], rnd f32 [B][H][W] = the torch.rand_like draw of :238 (passed in so that both sides
 * see the same numbers).  losses3 = (total, depth, smoothness), ddisp = d total / d disp.  scratch: mde_depth_loss_scratch_floats(). */
long long mde_depth_loss_scratch_floats(void);
int mde_depth_loss(const float* disp, const float* img, const float* rnd, int B, int H, int W, float w_depth, float w_smooth, float* scratch,
                   float* losses3, float* ddisp, mde_stream_t stream);

/* torch.optim.Adam step (train/mars_adapter.py:52-55,139; defaults: no weight decay, no amsgrad) on n contiguous fp32 values. */
int mde_adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2, float eps, int step,
                  mde_stream_t stream);
/* The same update with the step count (>= 1) read from device memory at run time: one captured launch serves every step of a
 * CUDA graph. */
int mde_adam_step_dev(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2, float eps, const int* step_dev,
                      mde_stream_t stream);

/* Multi-tensor Adam: one launch for every parameter tensor.  table_dev[t] (DEVICE memory) names tensor t's parameter, gradient and
 * the two moment buffers (fp32, contiguous), its length and its learning-rate group; chunk_start_dev[0..n_tensors] is the prefix sum
 * of ceil(n / mde_adam_chunk_elems()) (n_chunks = its last entry).  lr_groups_dev[group] is read on the device at run time, so a
 * learning-rate schedule (train/mars_adapter.py:47-61: two groups, StepLR) takes effect between replays of a captured graph. */
typedef struct mde_adam_tensor {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  long long group;
} mde_adam_tensor;
int mde_adam_chunk_elems(void);
int mde_adam_multi(const mde_adam_tensor* table_dev, int n_tensors, const int* chunk_start_dev, int n_chunks, const float* lr_groups_dev,
                   float beta1, float beta2, float eps, const int* step_dev, mde_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MDE_H_ */
