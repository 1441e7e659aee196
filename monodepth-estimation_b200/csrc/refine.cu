// models/refinement.py on the device.
//  * depth_fusion_kernel  — MultiScaleDepthFusion.forward (refinement.py:131-183): one fused stencil.
//    Every scale is resampled to the target size on the fly (bilinear, align_corners=False), staged with a
//    1-pixel zero halo in shared memory, then per pixel: 3x3 confidence conv + sigmoid per scale, normalise
//    over scales, 1x1 fusion conv.  HBM traffic = the input maps once + one fp32 output map.
//  * refine_pack_kernel / refine_output_kernel — the memory-bound ends of TextureAwareRefinementModule.forward
//    (refinement.py:68-107, live lines): pack rgb+depth into the 16-channel NHWC bf16 operand of the first
//    implicit-GEMM conv and evaluate the Sobel edge gate in fp32; apply the 1x1 output conv, the gate and the
//    residual add.  The 3x3 convs in between run on the tensor-core conv kernel (tc_gemm.cu / mde_conv_nhwc).
#include <cuda_fp16.h>

#include "common.h"
#include "ptx.cuh"

namespace mde {

__device__ __forceinline__ void src_index_acf2(float scale, int dst, int in_size, int& i0, int& i1, float& l1) {
  float s = scale * (dst + 0.5f) - 0.5f;
  s = s < 0.0f ? 0.0f : s;
  i0 = (int)s;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = s - (float)i0;
}

__device__ __forceinline__ float sample_acf(const float* __restrict__ p, int hs, int ws, int Hd, int Wd, int Y, int X) {
  if (hs == Hd && ws == Wd) return __ldg(p + (size_t)Y * ws + X);
  int y0, y1, x0, x1;
  float ly, lx;
  src_index_acf2((float)hs / (float)Hd, Y, hs, y0, y1, ly);
  src_index_acf2((float)ws / (float)Wd, X, ws, x0, x1, lx);
  const float v00 = __ldg(p + (size_t)y0 * ws + x0), v01 = __ldg(p + (size_t)y0 * ws + x1), v10 = __ldg(p + (size_t)y1 * ws + x0),
              v11 = __ldg(p + (size_t)y1 * ws + x1);
  return (1.0f - ly) * ((1.0f - lx) * v00 + lx * v01) + ly * ((1.0f - lx) * v10 + lx * v11);
}

// ------------------------------------------------------------------ MultiScaleDepthFusion (refinement.py:131-183)
constexpr int kFuTW = 64, kFuTH = 16;
struct FusionParams {
  const float* src[MDE_MAX_FUSION_SCALES];
  int hs[MDE_MAX_FUSION_SCALES], ws[MDE_MAX_FUSION_SCALES];
  int n, B, H, W;
  const float* conf_w;  // [n][9]
  const float* conf_b;  // [n]
  const float* fuse_w;  // [n]
  const float* fuse_b;  // [1]
  float* out;           // [B][H][W]
};

__global__ void __launch_bounds__(256) depth_fusion_kernel(const __grid_constant__ FusionParams p) {
  pdl_launch_dependents();
  pdl_grid_wait();
  __shared__ float tile[MDE_MAX_FUSION_SCALES][kFuTH + 2][kFuTW + 2];
  __shared__ float sw[MDE_MAX_FUSION_SCALES][12];
  const int b = blockIdx.z, X0 = blockIdx.x * kFuTW, Y0 = blockIdx.y * kFuTH;
  const int tid = threadIdx.x;
  if (tid < p.n * 12) {
    const int s = tid / 12, j = tid - s * 12;
    sw[s][j] = j < 9 ? p.conf_w[s * 9 + j] : (j == 9 ? p.conf_b[s] : (j == 10 ? p.fuse_w[s] : 0.0f));
  }
  constexpr int HALO = (kFuTH + 2) * (kFuTW + 2);
  for (int s = 0; s < p.n; ++s) {
    const float* src = p.src[s] + (size_t)b * p.hs[s] * p.ws[s];
    for (int i = tid; i < HALO; i += 256) {
      const int ty = i / (kFuTW + 2), tx = i - ty * (kFuTW + 2);
      const int Y = Y0 + ty - 1, X = X0 + tx - 1;
      float v = 0.0f;  // zero padding of the confidence conv (padding=1 on the RESIZED map)
      if (Y >= 0 && Y < p.H && X >= 0 && X < p.W) v = sample_acf(src, p.hs[s], p.ws[s], p.H, p.W, Y, X);
      tile[s][ty][tx] = v;
    }
  }
  __syncthreads();
  const int tx = tid & (kFuTW - 1);
  const int X = X0 + tx;
  if (X >= p.W) return;
  const float fb = __ldg(p.fuse_b);
#pragma unroll 1
  for (int ty = tid >> 6; ty < kFuTH; ty += 4) {
    const int Y = Y0 + ty;
    if (Y >= p.H) break;
    float conf[MDE_MAX_FUSION_SCALES], f[MDE_MAX_FUSION_SCALES];
    float csum = 0.0f;
#pragma unroll
    for (int s = 0; s < MDE_MAX_FUSION_SCALES; ++s) {
      conf[s] = 0.0f;
      f[s] = 0.0f;
      if (s < p.n) {
        float a = sw[s][9];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) a = fmaf(sw[s][ky * 3 + kx], tile[s][ty + ky][tx + kx], a);
        conf[s] = 1.0f / (1.0f + expf(-a));
        f[s] = tile[s][ty + 1][tx + 1];
        csum += conf[s];
      }
    }
    const float inv = 1.0f / (csum + 1e-8f);
    float o = fb;
#pragma unroll
    for (int s = 0; s < MDE_MAX_FUSION_SCALES; ++s)
      if (s < p.n) o = fmaf(sw[s][10], f[s] * (conf[s] * inv), o);
    __stcs(p.out + ((size_t)b * p.H + Y) * p.W + X, o);
  }
}

// ------------------------------------------------------------------ TextureAwareRefinementModule ends (refinement.py:68-107)
// One thread per pixel.  rgb f32 [B,3,H,W], depth f32 [B,1,H,W] ->
//   in16  fp16 padded NHWC [B][H+2][W+2][16]: channels {r,g,b,depth,0..0} (interior only; the convs that read it are
//         zero-padded and address the interior through TMA out-of-bounds fill, so the border is never read)
//   gate  f32 [B][H][W] = sigmoid(5 * sqrt(ex^2 + ey^2)), ex/ey = Sobel/3 summed over the 3 channels, zero padded
__global__ void __launch_bounds__(256) refine_pack_kernel(const float* __restrict__ rgb, const float* __restrict__ depth, int B, int H, int W,
                                                          __half* __restrict__ in16, float* __restrict__ gate) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const long long total = (long long)B * H * W;
  const size_t plane = (size_t)H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W), y = (int)((i / W) % H), b = (int)(i / plane);
    const float* base = rgb + (size_t)b * 3 * plane;
    float ex = 0.0f, ey = 0.0f, c[3];
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      const float* pl = base + ch * plane;
      float v[3][3];
#pragma unroll
      for (int dy = 0; dy < 3; ++dy)
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
          const int yy = y + dy - 1, xx = x + dx - 1;
          v[dy][dx] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(pl + (size_t)yy * W + xx) : 0.0f;
        }
      c[ch] = v[1][1];
      ex += (v[0][2] - v[0][0]) + 2.0f * (v[1][2] - v[1][0]) + (v[2][2] - v[2][0]);
      ey += (v[2][0] - v[0][0]) + 2.0f * (v[2][1] - v[0][1]) + (v[2][2] - v[0][2]);
    }
    ex *= (1.0f / 3.0f);
    ey *= (1.0f / 3.0f);
    const float mag = sqrtf(ex * ex + ey * ey);
    gate[i] = 1.0f / (1.0f + expf(-5.0f * mag));
    const float d = __ldg(depth + i);
    uint4* dst = reinterpret_cast<uint4*>(in16 + (((size_t)b * (H + 2) + y + 1) * (W + 2) + x + 1) * 16);
    dst[0] = make_uint4(pack_f16x2(c[0], c[1]), pack_f16x2(c[2], d), 0u, 0u);
    dst[1] = make_uint4(0u, 0u, 0u, 0u);
  }
}

// out = depth + (w . feat + b) * gate   (refinement.py:98-105); feat = padded NHWC fp16 [B][H+2][W+2][cstride], 32 channels read
__global__ void __launch_bounds__(256) refine_output_kernel(const __half* __restrict__ feat, int cstride, const float* __restrict__ w32,
                                                            const float* __restrict__ bias1, const float* __restrict__ depth,
                                                            const float* __restrict__ gate, int B, int H, int W, float* __restrict__ out) {
  pdl_launch_dependents();
  pdl_grid_wait();
  __shared__ float sw[32];
  if (threadIdx.x < 32) sw[threadIdx.x] = w32[threadIdx.x];
  __syncthreads();
  const float b0 = __ldg(bias1);
  const long long total = (long long)B * H * W;
  const size_t plane = (size_t)H * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W), y = (int)((i / W) % H), b = (int)(i / plane);
    const uint4* src = reinterpret_cast<const uint4*>(feat + (((size_t)b * (H + 2) + y + 1) * (W + 2) + x + 1) * cstride);
    float acc = b0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint4 u = __ldg(src + j);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[k]));
        acc = fmaf(sw[8 * j + 2 * k], f.x, acc);
        acc = fmaf(sw[8 * j + 2 * k + 1], f.y, acc);
      }
    }
    __stcs(out + i, __ldg(depth + i) + acc * __ldg(gate + i));
  }
}

static inline int grid_cap(long long work_items, int threads, int per_sm) {
  long long blocks = (work_items + threads - 1) / threads;
  const long long cap = (long long)sm_count() * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace mde

using namespace mde;

extern "C" int mde_depth_fusion(const float* const* maps, const int* hs, const int* ws, int n_maps, int B, int H, int W,
                                const float* conf_w, const float* conf_b, const float* fuse_w, const float* fuse_b, float* out,
                                mde_stream_t stream) {
  MDE_CHECK_ARG(maps && hs && ws && conf_w && conf_b && fuse_w && fuse_b && out, "mde_depth_fusion: null pointer");
  MDE_CHECK_ARG(n_maps >= 1 && n_maps <= MDE_MAX_FUSION_SCALES, "mde_depth_fusion: n_maps=%d not in [1,%d]", n_maps, MDE_MAX_FUSION_SCALES);
  MDE_CHECK_ARG(B > 0 && B <= 65535 && H > 0 && W > 0, "mde_depth_fusion: bad geometry B=%d H=%d W=%d", B, H, W);
  FusionParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < n_maps; ++i) {
    MDE_CHECK_ARG(maps[i] && hs[i] > 0 && ws[i] > 0, "mde_depth_fusion: map %d is empty", i);
    p.src[i] = maps[i];
    p.hs[i] = hs[i];
    p.ws[i] = ws[i];
  }
  p.n = n_maps;
  p.B = B;
  p.H = H;
  p.W = W;
  p.conf_w = conf_w;
  p.conf_b = conf_b;
  p.fuse_w = fuse_w;
  p.fuse_b = fuse_b;
  p.out = out;
  MDE_CHECK_ARG(ceil_div(H, kFuTH) <= 65535, "mde_depth_fusion: H too large");
  launch_k(depth_fusion_kernel, dim3(ceil_div(W, kFuTW), ceil_div(H, kFuTH), B), dim3(256), 0, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_refine_pack_inputs(const float* rgb, const float* depth, int B, int H, int W, mde_f16* in16_padded, float* gate,
                                      mde_stream_t stream) {
  MDE_CHECK_ARG(rgb && depth && in16_padded && gate && B > 0 && H >= 3 && W >= 3, "mde_refine_pack_inputs: bad arguments");
  MDE_CHECK_ARG((uintptr_t)in16_padded % 16 == 0, "mde_refine_pack_inputs: in16_padded must be 16-byte aligned");
  launch_k(refine_pack_kernel, dim3(grid_cap((long long)B * H * W, 256, 16)), dim3(256), 0, stream, rgb, depth, B, H, W, reinterpret_cast<__half*>(in16_padded), gate);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_refine_output(const mde_f16* feat_padded, int cstride, const float* w32, const float* bias1, const float* depth,
                                 const float* gate, int B, int H, int W, float* out, mde_stream_t stream) {
  MDE_CHECK_ARG(feat_padded && w32 && bias1 && depth && gate && out && B > 0 && H >= 3 && W >= 3, "mde_refine_output: bad arguments");
  MDE_CHECK_ARG(cstride >= 32 && cstride % 8 == 0 && (uintptr_t)feat_padded % 16 == 0, "mde_refine_output: feat must have >= 32 channels, 16-byte aligned");
  launch_k(refine_output_kernel, dim3(grid_cap((long long)B * H * W, 256, 16)), dim3(256), 0, stream, reinterpret_cast<const __half*>(feat_padded), cstride, w32, bias1, depth,
                                                                                    gate, B, H, W, out);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
