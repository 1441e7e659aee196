// HBM-bound kernels of the path: input range scan, resampling, patchify, LayerNorm, instance norm,
// upsample+concat, and the output stage.  All are coalesced / 16-byte vectorised, grid-stride over
// a multiple of the SM count, fp32 arithmetic, no host synchronisation.
#include <float.h>

#include <mutex>

#include "common.h"
#include "ptx.cuh"

namespace mde {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// torch's area_pixel_compute_source_index for bilinear, align_corners=False (UpSample.h)
__device__ __forceinline__ void src_index_acf(float scale, int dst, int in_size, int& i0, int& i1, float& l1) {
  float s = scale * (dst + 0.5f) - 0.5f;
  s = s < 0.0f ? 0.0f : s;
  i0 = (int)s;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = s - (float)i0;
}
// align_corners=True
__device__ __forceinline__ void src_index_act(float scale, int dst, int in_size, int& i0, int& i1, float& l1) {
  const float s = scale * dst;
  i0 = (int)s;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = s - (float)i0;
}

// ------------------------------------------------------------------ input range (encoder.py:129-135)
__global__ void __launch_bounds__(256) minmax_kernel(const float4* __restrict__ x4, long long n4, int hw4, float* __restrict__ partials) {
  pdl_launch_dependents();
  pdl_grid_wait();
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  const long long per = (n4 + gridDim.x - 1) / gridDim.x;
  const long long beg = per * blockIdx.x;
  long long end = beg + per;
  if (end > n4) end = n4;
  for (long long i0 = beg + threadIdx.x; i0 < end; i0 += 4 * blockDim.x) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {  // 4 independent 16-byte loads in flight per thread
      const long long i = i0 + u * blockDim.x;
      v[u] = i < end ? __ldg(x4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * blockDim.x;
      if (i >= end) break;
      const int c = (int)((i / hw4) % 3);
      const float lo = fminf(fminf(v[u].x, v[u].y), fminf(v[u].z, v[u].w));
      const float hi = fmaxf(fmaxf(v[u].x, v[u].y), fmaxf(v[u].z, v[u].w));
#pragma unroll
      for (int k = 0; k < 3; ++k)
        if (c == k) {
          mn[k] = fminf(mn[k], lo);
          mx[k] = fmaxf(mx[k], hi);
        }
    }
  }
  __shared__ float red[8][6];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float a = warp_min(mn[k]), b = warp_max(mx[k]);
    if (lane == 0) {
      red[warp][k] = a;
      red[warp][3 + k] = b;
    }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float r = red[0][threadIdx.x];
    for (int w = 1; w < 8; ++w) r = threadIdx.x < 3 ? fminf(r, red[w][threadIdx.x]) : fmaxf(r, red[w][threadIdx.x]);
    partials[blockIdx.x * 6 + threadIdx.x] = r;
  }
}

__global__ void __launch_bounds__(768) range_finalize_kernel(const float* __restrict__ partials, int nparts, const float* __restrict__ bn, float eps,
                                      float* __restrict__ affine6) {
  pdl_launch_dependents();
  // One block on the critical path of every forward (input scan -> this -> patchify): the BatchNorm parameters are requested before
  // the wait, every thread keeps 8 partials in flight (a single block cannot hide the L2 latency of a one-load-at-a-time loop: 56
  // round trips before), and the 128 threads of a statistic are combined by shuffles + ONE block-wide sync (the shared-memory tree
  // took eight).  min / max are order-independent: the result is bit-identical.
  __shared__ float s_bn[12];
  __shared__ float s_w[24];
  float bnv = 0.0f;
  if (threadIdx.x < 12) bnv = ld_param_now_f32(bn + threadIdx.x);
  pdl_grid_wait();
  {
    const int k = threadIdx.x >> 7, j0 = threadIdx.x & 127;  // statistic (0..2 minima, 3..5 maxima), first partial
    const bool is_min = k < 3;
    float r = is_min ? FLT_MAX : -FLT_MAX;
    for (int j = j0; j < nparts; j += 128 * 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int jj = j + 128 * u;
        v[u] = jj < nparts ? ld_after_wait_f32(partials + (size_t)jj * 6 + k) : r;  // not a read-only load: see ptx.cuh
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) r = is_min ? fminf(r, v[u]) : fmaxf(r, v[u]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float t = __shfl_xor_sync(0xffffffffu, r, o);
      r = is_min ? fminf(r, t) : fmaxf(r, t);
    }
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = r;  // warps 4k .. 4k+3 hold statistic k
    if (threadIdx.x < 12) s_bn[threadIdx.x] = bnv;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float st[6];
    for (int k = 0; k < 6; ++k) {
      const float u0 = s_w[4 * k], u1 = s_w[4 * k + 1], u2 = s_w[4 * k + 2], u3 = s_w[4 * k + 3];
      st[k] = k < 3 ? fminf(fminf(u0, u1), fminf(u2, u3)) : fmaxf(fmaxf(u0, u1), fmaxf(u2, u3));
    }
    const float* bn = s_bn;
    float a[3], b[3], gmin = FLT_MAX, gmax = -FLT_MAX;
    for (int c = 0; c < 3; ++c) {
      const float w = bn[c], bias = bn[3 + c], mean = bn[6 + c], var = bn[9 + c];
      const float invstd = 1.0f / sqrtf(var + eps);
      a[c] = w * invstd;
      b[c] = bias - mean * a[c];
      const float y0 = a[c] * st[c] + b[c], y1 = a[c] * st[3 + c] + b[c];
      gmin = fminf(gmin, fminf(y0, y1));
      gmax = fmaxf(gmax, fmaxf(y0, y1));
    }
    if (gmax > 1.0f) {  // encoder.py:132-133
      for (int c = 0; c < 3; ++c) {
        a[c] /= 255.0f;
        b[c] /= 255.0f;
      }
      gmin /= 255.0f;
      gmax /= 255.0f;
    }
    if (gmin >= 0.0f && gmax <= 1.0f) {  // encoder.py:134-135
      for (int c = 0; c < 3; ++c) {
        a[c] *= 2.0f;
        b[c] = 2.0f * b[c] - 1.0f;
      }
    }
    for (int c = 0; c < 3; ++c) {
      affine6[c] = a[c];
      affine6[3 + c] = b[c];
    }
  }
}

// ------------------------------------------------------------------ generic fp32 NCHW bilinear (align_corners=False)
__global__ void __launch_bounds__(256) resize_nchw_kernel(const float* __restrict__ src, int planes, int Hs, int Ws, float* __restrict__ dst,
                                                          int Hd, int Wd) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const float sh = (float)Hs / (float)Hd, sw = (float)Ws / (float)Wd;
  const long long total = (long long)planes * Hd * Wd;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int X = (int)(i % Wd);
    const int Y = (int)((i / Wd) % Hd);
    const long long pl = i / ((long long)Wd * Hd);
    int y0, y1, x0, x1;
    float ly, lx;
    src_index_acf(sh, Y, Hs, y0, y1, ly);
    src_index_acf(sw, X, Ws, x0, x1, lx);
    const float* p = src + pl * (long long)Hs * Ws;
    const float v00 = __ldg(p + (long long)y0 * Ws + x0), v01 = __ldg(p + (long long)y0 * Ws + x1);
    const float v10 = __ldg(p + (long long)y1 * Ws + x0), v11 = __ldg(p + (long long)y1 * Ws + x1);
    dst[i] = (1.0f - ly) * ((1.0f - lx) * v00 + lx * v01) + ly * ((1.0f - lx) * v10 + lx * v11);
  }
}

// ------------------------------------------------------------------ resize->224^2 + affine + patchify (encoder.py:139-144, tv:277)
// one thread = 8 consecutive kx of one (b, patch, c, ky): one 16-byte store of the GEMM A operand
__global__ void __launch_bounds__(256) patchify_kernel(const float* __restrict__ x, int B, int H, int W, const float* __restrict__ affine6,
                                                       bf16* __restrict__ patches) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const float sh = (float)H / 224.0f, sw = (float)W / 224.0f;
  const bool identity = (H == 224 && W == 224);
  const long long total = (long long)B * 196 * 96;  // 768 / 8 groups per patch row
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int grp = (int)(i % 96);
    const long long row = i / 96;  // b*196 + patch
    const int patch = (int)(row % 196);
    const int b = (int)(row / 196);
    const int c = grp >> 5, ky = (grp >> 1) & 15, kx0 = (grp & 1) * 8;
    const int Y = (patch / 14) * 16 + ky, X0 = (patch % 14) * 16 + kx0;
    const float a = __ldg(affine6 + c), bb = __ldg(affine6 + 3 + c);
    const float* p = x + ((long long)b * 3 + c) * H * W;
    float v[8];
    if (identity) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = __ldg(p + (long long)Y * W + X0 + j);
    } else {
      int y0, y1;
      float ly;
      src_index_acf(sh, Y, H, y0, y1, ly);
      const float* r0 = p + (long long)y0 * W;
      const float* r1 = p + (long long)y1 * W;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        int x0, x1;
        float lx;
        src_index_acf(sw, X0 + j, W, x0, x1, lx);
        v[j] = (1.0f - ly) * ((1.0f - lx) * __ldg(r0 + x0) + lx * __ldg(r0 + x1)) + ly * ((1.0f - lx) * __ldg(r1 + x0) + lx * __ldg(r1 + x1));
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaf(a, v[j], bb);
    *reinterpret_cast<uint4*>(patches + row * 768 + grp * 8) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}

// ------------------------------------------------------------------ fused input stage (encoder.py:129-144), one pass over x
// Bilinear resampling is linear with weights summing to 1 and BN + the range heuristic are one per-channel affine map, so
//   resize(a*x + b) == a*resize(x) + b
// and the 224x224 image can be resampled from the RAW input in the same pass that scans it for min/max; the affine map is
// applied afterwards to the 0.6 MB/image result.  A block owns `rpb` consecutive output rows of one (image, channel)
// plane: it scans the source rows [y0(first), y0(last+1)) — a contiguous range; the ranges of all blocks tile the plane
// exactly once — with four independent 16-byte loads in flight per thread, writes its min/max partial, then produces its
// output rows from taps that are now in L1/L2.  The separate patchify pass read 44 % of the input a second time from DRAM.
__device__ __forceinline__ int scan_row_start(float sh, int oy, int H) {  // first source row owned by output row oy
  if (oy <= 0) return 0;
  if (oy >= 224) return H;
  int y0, y1;
  float l;
  src_index_acf(sh, oy, H, y0, y1, l);
  return y0;
}
__global__ void __launch_bounds__(256) input_scan_resample_kernel(const float* __restrict__ x, int H, int W, int rpb, int groups,
                                                                  float* __restrict__ partials, float* __restrict__ raw224) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const int grp = blockIdx.x % groups, plane = blockIdx.x / groups;  // plane = b*3 + c
  const int c = plane % 3;
  const int oy0 = grp * rpb, oy1 = min(224, oy0 + rpb);
  const float sh = (float)H / 224.0f, sw = (float)W / 224.0f;
  const float* pl = x + (size_t)plane * H * W;
  const int lo = scan_row_start(sh, oy0, H), hi = scan_row_start(sh, oy1, H);
  float mn = FLT_MAX, mx = -FLT_MAX;
  {
    const float4* p4 = reinterpret_cast<const float4*>(pl + (size_t)lo * W);  // host checked W % 4 == 0 and 16-byte alignment
    const int n4 = (hi - lo) * (W >> 2);
#ifndef MDE_SCAN_UNROLL
#define MDE_SCAN_UNROLL 8
#endif
    constexpr int U = MDE_SCAN_UNROLL;  // independent 16-byte loads in flight per thread (r02, 32 x 3 x 1024^2: 2 / 4 / 8 -> 85.6 / 85.9 / 79.0 us)
    for (int i0 = threadIdx.x; i0 < n4; i0 += U * 256) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * 256;
        v[u] = __ldg(p4 + min(i, n4 - 1));  // past the end: re-read the last vector (neutral for min and max)
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        mn = fminf(mn, fminf(fminf(v[u].x, v[u].y), fminf(v[u].z, v[u].w)));
        mx = fmaxf(mx, fmaxf(fmaxf(v[u].x, v[u].y), fmaxf(v[u].z, v[u].w)));
      }
    }
  }
  __shared__ float red[8][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  mn = warp_min(mn);
  mx = warp_max(mx);
  if (lane == 0) {
    red[warp][0] = mn;
    red[warp][1] = mx;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    const int k = threadIdx.x;
    float r = k < 3 ? FLT_MAX : -FLT_MAX;
    if (k == c || k == 3 + c)
      for (int w = 0; w < 8; ++w) r = k < 3 ? fminf(r, red[w][0]) : fmaxf(r, red[w][1]);
    partials[(size_t)blockIdx.x * 6 + k] = r;
  }
  // resample this block's output rows (torch bilinear, align_corners=False)
  const int ox = threadIdx.x;
  if (ox >= 224) return;
  int x0, x1;
  float lx;
  src_index_acf(sw, ox, W, x0, x1, lx);
  float* dst = raw224 + ((size_t)plane * 224 + oy0) * 224 + ox;
  for (int oy = oy0; oy < oy1; ++oy, dst += 224) {
    int y0, y1;
    float ly;
    src_index_acf(sh, oy, H, y0, y1, ly);
    const float* r0 = pl + (size_t)y0 * W;
    const float* r1 = pl + (size_t)y1 * W;
    *dst = (1.0f - ly) * ((1.0f - lx) * __ldg(r0 + x0) + lx * __ldg(r0 + x1)) + ly * ((1.0f - lx) * __ldg(r1 + x0) + lx * __ldg(r1 + x1));
  }
}

// raw224 f32 [B][3][224][224] -> affine -> patch-embedding A operand bf16 [B*196][768], column = c*256 + ky*16 + kx
__global__ void __launch_bounds__(256) patchify224_kernel(const float* __restrict__ raw, int B, const float* __restrict__ affine6,
                                                          bf16* __restrict__ patches) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const int total = B * 196 * 96;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int grp = i % 96, row = i / 96;  // row = b*196 + patch
    const int patch = row % 196, b = row / 196;
    const int c = grp >> 5, ky = (grp >> 1) & 15, kx0 = (grp & 1) * 8;
    const int Y = (patch / 14) * 16 + ky, X0 = (patch % 14) * 16 + kx0;
    const float a = __ldg(affine6 + c), bb = __ldg(affine6 + 3 + c);
    const float4* src = reinterpret_cast<const float4*>(raw + (((size_t)b * 3 + c) * 224 + Y) * 224 + X0);
    const float4 u = __ldg(src), w = __ldg(src + 1);
    *reinterpret_cast<uint4*>(patches + (size_t)row * 768 + grp * 8) =
        make_uint4(pack_bf16x2(fmaf(a, u.x, bb), fmaf(a, u.y, bb)), pack_bf16x2(fmaf(a, u.z, bb), fmaf(a, u.w, bb)),
                   pack_bf16x2(fmaf(a, w.x, bb), fmaf(a, w.y, bb)), pack_bf16x2(fmaf(a, w.z, bb), fmaf(a, w.w, bb)));
  }
}

// ------------------------------------------------------------------ LayerNorm (tv:112,117) fp32 -> bf16, one warp per R rows
// R rows per warp with all R*24 loads of a lane in flight before the first reduction: at B = 32 (6304 rows) one row per warp
// needs 788 blocks for 592 resident slots (4 blocks/SM), i.e. a second, one-third-full wave; R = 2 fits one wave.
template <int DIM, int R>
__global__ void __launch_bounds__(256, R == 1 ? 4 : 3) layernorm_kernel(const float* __restrict__ x, int rows, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, float eps, bf16* __restrict__ out) {
  pdl_launch_dependents();
  constexpr int PER = DIM / 256;  // 8-element groups per lane
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // gamma / beta are parameters, not written by the previous kernel: they go to shared memory while this kernel still waits for it
  // (read straight from global memory after the reductions, their L2 round trip sat on every block's critical path)
  __shared__ float4 s_gb[2][DIM / 4];
  float4 pg = make_float4(0.0f, 0.0f, 0.0f, 0.0f), pb = pg;
  if (threadIdx.x < DIM / 4) {  // in flight across the wait and the row loads; stored to shared memory once those are issued
    pg = ld_param_now_f4(reinterpret_cast<const float4*>(gamma) + threadIdx.x);
    pb = ld_param_now_f4(reinterpret_cast<const float4*>(beta) + threadIdx.x);
  }
  pdl_grid_wait();
  const int row0 = min((blockIdx.x * 8 + warp) * R, rows - 1);  // a warp past the end computes the last row again and stores nothing
  const bool live0 = (blockIdx.x * 8 + warp) * R < rows;
  float v[R][PER * 8];
  float s[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int row = min(row0 + r, rows - 1);  // a clamped duplicate row is computed and not stored
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * DIM);
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const float4 a = ld_after_wait_f4(xr + (j * 32 + lane) * 2), b = ld_after_wait_f4(xr + (j * 32 + lane) * 2 + 1);  // not __ldg: see ptx.cuh
      v[r][8 * j + 0] = a.x; v[r][8 * j + 1] = a.y; v[r][8 * j + 2] = a.z; v[r][8 * j + 3] = a.w;
      v[r][8 * j + 4] = b.x; v[r][8 * j + 5] = b.y; v[r][8 * j + 6] = b.z; v[r][8 * j + 7] = b.w;
    }
  }
  if (threadIdx.x < DIM / 4) {
    s_gb[0][threadIdx.x] = pg;
    s_gb[1][threadIdx.x] = pb;
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    s[r] = 0.0f;
#pragma unroll
    for (int j = 0; j < PER; ++j)
#pragma unroll
      for (int k = 0; k < 8; ++k) s[r] += v[r][8 * j + k];
  }
  float mean[R], rstd[R];
#pragma unroll
  for (int r = 0; r < R; ++r) mean[r] = warp_sum(s[r]) * (1.0f / DIM);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float q = 0.0f;
#pragma unroll
    for (int k = 0; k < PER * 8; ++k) {
      const float d = v[r][k] - mean[r];
      q = fmaf(d, d, q);
    }
    s[r] = q;
  }
#pragma unroll
  for (int r = 0; r < R; ++r) rstd[r] = 1.0f / sqrtf(warp_sum(s[r]) * (1.0f / DIM) + eps);
  __syncthreads();  // s_gb is complete (its loads were issued before everything else)
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int g = (j * 32 + lane) * 2;
    const float4 ga = s_gb[0][g], gb = s_gb[0][g + 1], ba = s_gb[1][g], bb = s_gb[1][g + 1];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (!live0 || row0 + r >= rows) break;
      const float m = mean[r], t = rstd[r];
      const float y0 = (v[r][8 * j + 0] - m) * t * ga.x + ba.x, y1 = (v[r][8 * j + 1] - m) * t * ga.y + ba.y;
      const float y2 = (v[r][8 * j + 2] - m) * t * ga.z + ba.z, y3 = (v[r][8 * j + 3] - m) * t * ga.w + ba.w;
      const float y4 = (v[r][8 * j + 4] - m) * t * gb.x + bb.x, y5 = (v[r][8 * j + 5] - m) * t * gb.y + bb.y;
      const float y6 = (v[r][8 * j + 6] - m) * t * gb.z + bb.z, y7 = (v[r][8 * j + 7] - m) * t * gb.w + bb.w;
      reinterpret_cast<uint4*>(out + (size_t)(row0 + r) * DIM)[j * 32 + lane] =
          make_uint4(pack_bf16x2(y0, y1), pack_bf16x2(y2, y3), pack_bf16x2(y4, y5), pack_bf16x2(y6, y7));
    }
  }
}

// ------------------------------------------------------------------ tokens -> grid rows (encoder.py:69-110), fp32 -> bf16
__global__ void __launch_bounds__(256) tokens_to_grid_kernel(const float* __restrict__ tok, int B, int strip, int G, bf16* __restrict__ out) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const int T = MDE_VIT_SEQ - strip;  // tokens copied per image
  const long long total = (long long)B * G * 96;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int grp = (int)(i % 96);
    const long long row = i / 96;
    const int g = (int)(row % G), b = (int)(row / G);
    uint4 o = make_uint4(0, 0, 0, 0);
    if (g < T) {
      const float4* s = reinterpret_cast<const float4*>(tok + ((long long)b * MDE_VIT_SEQ + g + strip) * MDE_VIT_DIM + grp * 8);
      const float4 a = __ldg(s), c = __ldg(s + 1);
      o = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(c.x, c.y), pack_bf16x2(c.z, c.w));
    }
    *reinterpret_cast<uint4*>(out + row * MDE_VIT_DIM + grp * 8) = o;
  }
}

// ------------------------------------------------------------------ instance norm (encoder.py:167) -> padded NHWC bf16
// grid (C/16, B) so the 4 x 32-image launch fills the machine; block 256 = 4 channel quads x 64 position lanes.  A thread keeps its
// <= 4 positions x 4 channels in registers (float4 loads, one pass over the data), the position lanes of a warp are summed by
// shuffles and the 8 warps through shared memory, and every position is written as one 8-byte store plus its reflected-border
// copies as straight-line predicated stores (the first version: one channel per thread, 14 positions each with an integer division
// and a 3 x 3 candidate loop — issue-bound at ~11 us for 7 MB).
constexpr int kInLanes = 64, kInMaxPer = 4;
__device__ __forceinline__ float4 quad_sum_positions(float4 s) {  // over the 8 lanes of the warp that share a channel quad
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) {
    s.x += __shfl_xor_sync(0xffffffffu, s.x, o);
    s.y += __shfl_xor_sync(0xffffffffu, s.y, o);
    s.z += __shfl_xor_sync(0xffffffffu, s.z, o);
    s.w += __shfl_xor_sync(0xffffffffu, s.w, o);
  }
  return s;
}
__global__ void __launch_bounds__(256) instnorm_kernel(const float* __restrict__ x, int gh, int gw, int C, float eps, bf16* __restrict__ outp,
                                                       int cstride, int coff, bf16* __restrict__ outf) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const int G = gh * gw;
  const int quad = threadIdx.x & 3, pl = threadIdx.x >> 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x * 16 + quad * 4, b = blockIdx.y;
  const float* xb = x + (size_t)b * G * C + c;
  __shared__ float4 red[2][8][4];
  float4 v[kInMaxPer];
  float4 s = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
  for (int i = 0; i < kInMaxPer; ++i) {
    const int g = pl + kInLanes * i;
    v[i] = g < G ? __ldg(reinterpret_cast<const float4*>(xb + (size_t)g * C)) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    s.x += v[i].x; s.y += v[i].y; s.z += v[i].z; s.w += v[i].w;
  }
  s = quad_sum_positions(s);
  if (lane < 4) red[0][warp][lane] = s;
  __syncthreads();
  float4 mean = red[0][0][quad];
#pragma unroll
  for (int w = 1; w < 8; ++w) {
    const float4 t = red[0][w][quad];
    mean.x += t.x; mean.y += t.y; mean.z += t.z; mean.w += t.w;
  }
  const float fG = (float)G;
  mean.x /= fG; mean.y /= fG; mean.z /= fG; mean.w /= fG;
  float4 q = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
  for (int i = 0; i < kInMaxPer; ++i) {
    if (pl + kInLanes * i < G) {
      const float dx = v[i].x - mean.x, dy = v[i].y - mean.y, dz = v[i].z - mean.z, dw = v[i].w - mean.w;
      q.x = fmaf(dx, dx, q.x); q.y = fmaf(dy, dy, q.y); q.z = fmaf(dz, dz, q.z); q.w = fmaf(dw, dw, q.w);
    }
  }
  q = quad_sum_positions(q);
  if (lane < 4) red[1][warp][lane] = q;
  __syncthreads();
  float4 var = red[1][0][quad];
#pragma unroll
  for (int w = 1; w < 8; ++w) {
    const float4 t = red[1][w][quad];
    var.x += t.x; var.y += t.y; var.z += t.z; var.w += t.w;
  }
  // biased variance, like F.instance_norm
  const float4 rstd = make_float4(1.0f / sqrtf(var.x / fG + eps), 1.0f / sqrtf(var.y / fG + eps), 1.0f / sqrtf(var.z / fG + eps), 1.0f / sqrtf(var.w / fG + eps));
  const int W2 = gw + 2;
  const long long rowp = (long long)W2 * cstride;  // one padded row, in elements
#pragma unroll
  for (int i = 0; i < kInMaxPer; ++i) {
    const int g = pl + kInLanes * i;
    if (g < G) {
      const uint2 o = make_uint2(pack_bf16x2((v[i].x - mean.x) * rstd.x, (v[i].y - mean.y) * rstd.y), pack_bf16x2((v[i].z - mean.z) * rstd.z, (v[i].w - mean.w) * rstd.w));
      if (outf) *reinterpret_cast<uint2*>(outf + ((size_t)b * G + g) * C + c) = o;
      const int y = g / gw, xx = g - y * gw;
      bf16* d = outp + (((size_t)b * (gh + 2) + y + 1) * W2 + xx + 1) * cstride + coff + c;
      // ReflectionPad2d(1): padded row / column 0 repeats source 1, padded row / column n + 1 repeats source n - 2
      const bool xl = xx == 1, xr = xx == gw - 2, yt = y == 1, yb = y == gh - 2;
      const long long cl = -2LL * cstride, cr = 2LL * cstride, rt = -2 * rowp, rb = 2 * rowp;
      *reinterpret_cast<uint2*>(d) = o;
      if (xl) *reinterpret_cast<uint2*>(d + cl) = o;
      if (xr) *reinterpret_cast<uint2*>(d + cr) = o;
      if (yt) {
        *reinterpret_cast<uint2*>(d + rt) = o;
        if (xl) *reinterpret_cast<uint2*>(d + rt + cl) = o;
        if (xr) *reinterpret_cast<uint2*>(d + rt + cr) = o;
      }
      if (yb) {
        *reinterpret_cast<uint2*>(d + rb) = o;
        if (xl) *reinterpret_cast<uint2*>(d + rb + cl) = o;
        if (xr) *reinterpret_cast<uint2*>(d + rb + cr) = o;
      }
    }
  }
}

__device__ __forceinline__ void bf16x8_to_float(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 t = __bfloat1622float2(h[k]);
    f[2 * k] = t.x;
    f[2 * k + 1] = t.y;
  }
}
__device__ __forceinline__ int reflect_idx(int i, int n) {  // ReflectionPad2d(1) source index for padded coordinate i-1
  return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
}

// One block per padded output row (b, Yp): the vertical taps/weights of both sources are block constants, the item loop is
// 32-bit (the first version spent its time in 64-bit index divisions: issue-bound at 4x the HBM time, r01b profile).
__global__ void __launch_bounds__(256) upsample_concat_kernel(const bf16* __restrict__ a, int B, int h, int w, int ca, const bf16* __restrict__ s,
                                                              int hs, int ws, int cs, bf16* __restrict__ out) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const int Ho = 2 * h, Wo = 2 * w, ct = ca + cs, groups = ct >> 3, ga = ca >> 3;
  const int b = blockIdx.x / (Ho + 2), Yp = blockIdx.x - b * (Ho + 2);
  const int oy = reflect_idx(Yp - 1, Ho);
  const float sxa = (float)(w - 1) / (float)(Wo - 1);
  const float sxs = cs ? (float)(ws - 1) / (float)(Wo - 1) : 0.0f;
  int ya0, ya1, ys0 = 0, ys1 = 0;
  float lya, lys = 0.0f;
  src_index_act((float)(h - 1) / (float)(Ho - 1), oy, h, ya0, ya1, lya);
  if (cs) src_index_act((float)(hs - 1) / (float)(Ho - 1), oy, hs, ys0, ys1, lys);
  const bf16* a0 = a + ((size_t)b * (h + 2) + ya0 + 1) * (w + 2) * ca;
  const bf16* a1 = a + ((size_t)b * (h + 2) + ya1 + 1) * (w + 2) * ca;
  const bf16* s0 = cs ? s + ((size_t)b * (hs + 2) + ys0 + 1) * (ws + 2) * cs : nullptr;
  const bf16* s1 = cs ? s + ((size_t)b * (hs + 2) + ys1 + 1) * (ws + 2) * cs : nullptr;
  bf16* orow = out + ((size_t)b * (Ho + 2) + Yp) * (Wo + 2) * ct;
  const int items = (Wo + 2) * groups;
  for (int i = threadIdx.x; i < items; i += 256) {
    const int Xp = i / groups, grp = i - Xp * groups;
    const int ox = reflect_idx(Xp - 1, Wo);
    const bf16 *r0, *r1;
    int x0, x1, cc, cstr;
    float ly, lx;
    if (grp < ga) {
      src_index_act(sxa, ox, w, x0, x1, lx);
      r0 = a0; r1 = a1; ly = lya; cc = grp << 3; cstr = ca;
    } else {
      src_index_act(sxs, ox, ws, x0, x1, lx);
      r0 = s0; r1 = s1; ly = lys; cc = (grp - ga) << 3; cstr = cs;
    }
    const uint4 u00 = __ldg(reinterpret_cast<const uint4*>(r0 + (x0 + 1) * cstr + cc));
    const uint4 u01 = __ldg(reinterpret_cast<const uint4*>(r0 + (x1 + 1) * cstr + cc));
    const uint4 u10 = __ldg(reinterpret_cast<const uint4*>(r1 + (x0 + 1) * cstr + cc));
    const uint4 u11 = __ldg(reinterpret_cast<const uint4*>(r1 + (x1 + 1) * cstr + cc));
    float f00[8], f01[8], f10[8], f11[8], r[8];
    bf16x8_to_float(u00, f00);
    bf16x8_to_float(u01, f01);
    bf16x8_to_float(u10, f10);
    bf16x8_to_float(u11, f11);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      r[k] = (1.0f - ly) * ((1.0f - lx) * f00[k] + lx * f01[k]) + ly * ((1.0f - lx) * f10[k] + lx * f11[k]);
    *reinterpret_cast<uint4*>(orow + (size_t)i * 8) =
        make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
  }
}

// Faster form for channel counts whose 8-channel groups are a power of two (every decoder level: 16 / 64 / 128 / 256 channels).
// A block owns R padded output rows of one image.  Phase 1 blends the two source rows of each output row VERTICALLY into shared
// memory as fp32 (both sources, at source width: 7.7 KB + <= 7.7 KB per row at every level), phase 2 produces each 8-channel
// output group from two shared-memory pixels with one lerp: the horizontal index arithmetic is done once per thread and reused
// for the R rows, and the four 16-byte global loads + 32 conversions per output group of the row-per-block kernel become four
// shared-memory reads.  (r02 ablation: the four launches cost 77 us of the 2.17 ms step for 35 us of traffic.)
template <int R>
__global__ void __launch_bounds__(256) upsample_concat_rows_kernel(const bf16* __restrict__ a, int B, int h, int w, int ca, const bf16* __restrict__ s,
                                                                   int hs, int ws, int cs, bf16* __restrict__ out, int lg_groups) {
  pdl_launch_dependents();
  extern __shared__ float4 ucr_smem[];
  const int Ho = 2 * h, Wo = 2 * w, ct = ca + cs, groups = 1 << lg_groups, ga = ca >> 3;
  const int rows_total = Ho + 2, bpi = (rows_total + R - 1) / R;
  const int b = blockIdx.x / bpi, Y0 = (blockIdx.x - b * bpi) * R;
  const int nrows = min(R, rows_total - Y0);
  const int na8 = (w * ca) >> 3, ns8 = (ws * cs) >> 3, row_f4 = 2 * (na8 + ns8);
  const float sya = (float)(h - 1) / (float)(Ho - 1), sys = cs ? (float)(hs - 1) / (float)(Ho - 1) : 0.0f;
  pdl_grid_wait();
  // per-row source rows and weights first (R is a compile-time constant: these live in registers), then ONE pass over the source
  // elements with the R rows innermost, so a thread has 2R independent 16-byte loads in flight instead of two
  const uint4 *a0[R], *a1[R], *s0[R], *s1[R];
  float lya[R], lys[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int oy = reflect_idx(min(Y0 + r, rows_total - 1) - 1, Ho);
    int y0, y1;
    src_index_act(sya, oy, h, y0, y1, lya[r]);
    // interior of a padded source row is contiguous: element x * C + c sits at (x + 1) * C + c of the row
    a0[r] = reinterpret_cast<const uint4*>(a + (((size_t)b * (h + 2) + y0 + 1) * (w + 2) + 1) * ca);
    a1[r] = reinterpret_cast<const uint4*>(a + (((size_t)b * (h + 2) + y1 + 1) * (w + 2) + 1) * ca);
    s0[r] = s1[r] = nullptr;
    lys[r] = 0.0f;
    if (cs) {
      src_index_act(sys, oy, hs, y0, y1, lys[r]);
      s0[r] = reinterpret_cast<const uint4*>(s + (((size_t)b * (hs + 2) + y0 + 1) * (ws + 2) + 1) * cs);
      s1[r] = reinterpret_cast<const uint4*>(s + (((size_t)b * (hs + 2) + y1 + 1) * (ws + 2) + 1) * cs);
    }
  }
  for (int i = threadIdx.x; i < na8 + ns8; i += 256) {
    const bool isa = i < na8;
    const int k = isa ? i : i - na8;
    uint4 u0[R], u1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      u0[r] = __ldg((isa ? a0[r] : s0[r]) + k);
      u1[r] = __ldg((isa ? a1[r] : s1[r]) + k);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float ly = isa ? lya[r] : lys[r];
      float f0[8], f1[8];
      bf16x8_to_float(u0[r], f0);
      bf16x8_to_float(u1[r], f1);
#pragma unroll
      for (int c = 0; c < 8; ++c) f0[c] = (1.0f - ly) * f0[c] + ly * f1[c];
      float4* dst = ucr_smem + (size_t)r * row_f4;
      dst[2 * i] = make_float4(f0[0], f0[1], f0[2], f0[3]);
      dst[2 * i + 1] = make_float4(f0[4], f0[5], f0[6], f0[7]);
    }
  }
  __syncthreads();
  const float sxa = (float)(w - 1) / (float)(Wo - 1), sxs = cs ? (float)(ws - 1) / (float)(Wo - 1) : 0.0f;
  const int items = (Wo + 2) << lg_groups;
  bf16* orow0 = out + ((size_t)b * rows_total + Y0) * (Wo + 2) * ct;
  const size_t row_elems = (size_t)(Wo + 2) * ct;
  for (int i = threadIdx.x; i < items; i += 256) {
    const int Xp = i >> lg_groups, grp = i & (groups - 1);
    const int ox = reflect_idx(Xp - 1, Wo);
    int x0, x1, i0, i1;
    float lx;
    if (grp < ga) {
      src_index_act(sxa, ox, w, x0, x1, lx);
      i0 = (x0 * ca + (grp << 3)) >> 2;
      i1 = (x1 * ca + (grp << 3)) >> 2;
    } else {
      src_index_act(sxs, ox, ws, x0, x1, lx);
      i0 = 2 * na8 + ((x0 * cs + ((grp - ga) << 3)) >> 2);
      i1 = 2 * na8 + ((x1 * cs + ((grp - ga) << 3)) >> 2);
    }
    const float l0 = 1.0f - lx;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (r < nrows) {
        const float4* row = ucr_smem + (size_t)r * row_f4;
        const float4 p0 = row[i0], p1 = row[i0 + 1], q0 = row[i1], q1 = row[i1 + 1];
        *reinterpret_cast<uint4*>(orow0 + r * row_elems + (size_t)i * 8) =
            make_uint4(pack_bf16x2(l0 * p0.x + lx * q0.x, l0 * p0.y + lx * q0.y), pack_bf16x2(l0 * p0.z + lx * q0.z, l0 * p0.w + lx * q0.w),
                       pack_bf16x2(l0 * p1.x + lx * q1.x, l0 * p1.y + lx * q1.y), pack_bf16x2(l0 * p1.z + lx * q1.z, l0 * p1.w + lx * q1.w));
      }
    }
  }
}

// The same stage fed with the decoder's 8-bit HWC image (run_mars_depth.py:94-103: np.array(PIL) -> float()/255 -> permute): min/max
// are taken on the packed bytes (monotonic under the division, so float(min)/255 is the min of the floats), the resampled 224 px
// raster is computed from float(u8)/255 with the SAME IEEE division and the same expression as the fp32 kernel, i.e. the outputs
// are bit-identical to mde_u8hwc_to_f32chw + mde_input_stage while reading 3 bytes per pixel instead of writing 12 and reading 12.
// A block owns `rpb` output rows of ALL THREE channels of one image.
__global__ void __launch_bounds__(256) input_scan_resample_u8_kernel(const uint8_t* __restrict__ x, int H, int W, int rpb, int groups,
                                                                     float* __restrict__ partials, float* __restrict__ raw224) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const int grp = blockIdx.x % groups, b = blockIdx.x / groups;
  const int oy0 = grp * rpb, oy1 = min(224, oy0 + rpb);
  const float sh = (float)H / 224.0f, sw = (float)W / 224.0f;
  const uint8_t* img = x + (size_t)b * H * W * 3;
  const int lo = scan_row_start(sh, oy0, H), hi = scan_row_start(sh, oy1, H);
  // bytes of 4 pixels: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3 — one packed min and max per word position
  uint32_t mn0 = 0xffffffffu, mn1 = 0xffffffffu, mn2 = 0xffffffffu, mx0 = 0, mx1 = 0, mx2 = 0;
  {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(img + (size_t)lo * W * 3);  // host checked W % 4 == 0 and 4-byte alignment
    const int n = (hi - lo) * (W >> 2);                                               // groups of 4 pixels = 3 words
    for (int i0 = threadIdx.x; i0 < n; i0 += 4 * 256) {
      uint32_t w[4][3];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = min(i0 + u * 256, n - 1);  // past the end: re-read the last group (neutral)
        w[u][0] = __ldg(p + 3 * i);
        w[u][1] = __ldg(p + 3 * i + 1);
        w[u][2] = __ldg(p + 3 * i + 2);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        mn0 = __vminu4(mn0, w[u][0]); mx0 = __vmaxu4(mx0, w[u][0]);
        mn1 = __vminu4(mn1, w[u][1]); mx1 = __vmaxu4(mx1, w[u][1]);
        mn2 = __vminu4(mn2, w[u][2]); mx2 = __vmaxu4(mx2, w[u][2]);
      }
    }
  }
  auto byte_of = [](uint32_t v, int k) { return (v >> (8 * k)) & 255u; };
  uint32_t cmn[3], cmx[3];
  cmn[0] = min(min(byte_of(mn0, 0), byte_of(mn0, 3)), min(byte_of(mn1, 2), byte_of(mn2, 1)));
  cmn[1] = min(min(byte_of(mn0, 1), byte_of(mn1, 0)), min(byte_of(mn1, 3), byte_of(mn2, 2)));
  cmn[2] = min(min(byte_of(mn0, 2), byte_of(mn1, 1)), min(byte_of(mn2, 0), byte_of(mn2, 3)));
  cmx[0] = max(max(byte_of(mx0, 0), byte_of(mx0, 3)), max(byte_of(mx1, 2), byte_of(mx2, 1)));
  cmx[1] = max(max(byte_of(mx0, 1), byte_of(mx1, 0)), max(byte_of(mx1, 3), byte_of(mx2, 2)));
  cmx[2] = max(max(byte_of(mx0, 2), byte_of(mx1, 1)), max(byte_of(mx2, 0), byte_of(mx2, 3)));
  __shared__ uint32_t red[8][6];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    cmn[c] = __reduce_min_sync(0xffffffffu, cmn[c]);
    cmx[c] = __reduce_max_sync(0xffffffffu, cmx[c]);
  }
  if (lane == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) red[warp][c] = cmn[c], red[warp][3 + c] = cmx[c];
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    const int k = threadIdx.x;
    uint32_t r = red[0][k];
    for (int w = 1; w < 8; ++w) r = k < 3 ? min(r, red[w][k]) : max(r, red[w][k]);
    partials[(size_t)blockIdx.x * 6 + k] = __fdiv_rn((float)r, 255.0f);
  }
  // resample this block's output rows (torch bilinear, align_corners=False) from float(u8) / 255
  const int ox = threadIdx.x;
  if (ox >= 224) return;
  int x0, x1;
  float lx;
  src_index_acf(sw, ox, W, x0, x1, lx);
  for (int oy = oy0; oy < oy1; ++oy) {
    int y0, y1;
    float ly;
    src_index_acf(sh, oy, H, y0, y1, ly);
    const uint8_t* r0 = img + (size_t)y0 * W * 3;
    const uint8_t* r1 = img + (size_t)y1 * W * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v00 = __fdiv_rn((float)__ldg(r0 + 3 * x0 + c), 255.0f), v01 = __fdiv_rn((float)__ldg(r0 + 3 * x1 + c), 255.0f);
      const float v10 = __fdiv_rn((float)__ldg(r1 + 3 * x0 + c), 255.0f), v11 = __fdiv_rn((float)__ldg(r1 + 3 * x1 + c), 255.0f);
      raw224[(((size_t)b * 3 + c) * 224 + oy) * 224 + ox] = (1.0f - ly) * ((1.0f - lx) * v00 + lx * v01) + ly * ((1.0f - lx) * v10 + lx * v11);
    }
  }
}

// ------------------------------------------------------------------ rgb -> grid channels of a padded NHWC buffer (enhanced_model.py:95-98)
// U8 = true: x is the 8-bit HWC image and a pixel is float(u8) / 255 (the caller-side expression, same IEEE division)
template <bool U8>
__global__ void __launch_bounds__(256) rgb_to_grid_kernel(const void* __restrict__ xv, int B, int H, int W, int gh, int gw, bf16* __restrict__ out,
                                                          int cstride, int coff, int cfill) {
  pdl_launch_dependents();
  pdl_grid_wait();
  // thread = (grid position, group of 8 output channels): group 0 takes the three bilinear samples (12 independent loads),
  // the other groups write the zero fill — one 16-byte store per thread instead of a 64-iteration 2-byte store loop per position
  const float sh = (float)H / (float)gh, sw = (float)W / (float)gw;
  const int groups = cfill >> 3;  // host checked cfill % 8 == 0
  const int total = B * gh * gw * groups;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
    const int grp = t % groups, i = t / groups;
    const int xx = i % gw, y = (i / gw) % gh, b = i / (gw * gh);
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (grp == 0) {
      int y0, y1, x0, x1;
      float ly, lx;
      src_index_acf(sh, y, H, y0, y1, ly);
      src_index_acf(sw, xx, W, x0, x1, lx);
      float v[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float v00, v01, v10, v11;
        if constexpr (U8) {
          const uint8_t* p = reinterpret_cast<const uint8_t*>(xv) + (size_t)b * H * W * 3 + c;
          v00 = __fdiv_rn((float)__ldg(p + ((size_t)y0 * W + x0) * 3), 255.0f);
          v01 = __fdiv_rn((float)__ldg(p + ((size_t)y0 * W + x1) * 3), 255.0f);
          v10 = __fdiv_rn((float)__ldg(p + ((size_t)y1 * W + x0) * 3), 255.0f);
          v11 = __fdiv_rn((float)__ldg(p + ((size_t)y1 * W + x1) * 3), 255.0f);
        } else {
          const float* p = reinterpret_cast<const float*>(xv) + ((size_t)b * 3 + c) * H * W;
          v00 = __ldg(p + (size_t)y0 * W + x0);
          v01 = __ldg(p + (size_t)y0 * W + x1);
          v10 = __ldg(p + (size_t)y1 * W + x0);
          v11 = __ldg(p + (size_t)y1 * W + x1);
        }
        v[c] = (1.0f - ly) * ((1.0f - lx) * v00 + lx * v01) + ly * ((1.0f - lx) * v10 + lx * v11);
      }
      o.x = pack_bf16x2(v[0], v[1]);
      o.y = pack_bf16x2(v[2], 0.0f);
    }
    bf16* dst = out + (((size_t)b * (gh + 2) + y + 1) * (gw + 2) + xx + 1) * cstride + coff + grp * 8;
    *reinterpret_cast<uint4*>(dst) = o;
  }
}

// ------------------------------------------------------------------ output stage (model.py:89-100, enhanced_model.py:266-277, layers.py:16-25)
struct MapTable {
  mde_map_desc m[MDE_MAX_MAPS];
  int n;
};

// grid = (ceil(W/PX/256), ceil(H/kRowsPerBlock), n_maps*B).  The block first blends its kRowsPerBlock output rows VERTICALLY
// at source resolution into shared memory (<= 1024 floats per row, `mul` folded in, one duplicated element at the end so
// the right tap is always x0+1), then every thread produces PX consecutive pixels per row with two shared-memory reads
// and one lerp each: ~25 instructions per 16-byte store instead of ~190 (r01b: the 4-tap-per-pixel version was issue-
// bound at 2.4 TB/s).  DRAM traffic is the output stream only.
constexpr int kMaxSrcW = 1024;
constexpr int kRowsPerBlock = 8;
template <bool VEC4>
__global__ void __launch_bounds__(256) resize_maps_kernel(const __grid_constant__ MapTable tab, int B, int H, int W, float min_disp, float disp_range) {
  pdl_launch_dependents();
  pdl_grid_wait();
  constexpr int PX = VEC4 ? 4 : 1;
  __shared__ float srow[kRowsPerBlock][kMaxSrcW + 1];
  const int mi = blockIdx.z / B, b = blockIdx.z - mi * B;
  const mde_map_desc& d = tab.m[mi];
  const int ws = d.ws, hs = d.hs;
  const float sh = (float)hs / (float)H;
  const float mul = d.mul;
  const float* src = d.src + (size_t)b * hs * ws;
  const int Y0 = blockIdx.y * kRowsPerBlock;
  const int nrows = min(kRowsPerBlock, H - Y0);
  for (int r = 0; r < nrows; ++r) {
    int y0, y1;
    float ly;
    src_index_acf(sh, Y0 + r, hs, y0, y1, ly);
    const float* r0 = src + (size_t)y0 * ws;
    const float* r1 = src + (size_t)y1 * ws;
    const float w0 = (1.0f - ly) * mul, w1 = ly * mul;
    for (int x = threadIdx.x; x <= ws; x += 256) {
      const int xc = min(x, ws - 1);
      srow[r][x] = w0 * __ldg(r0 + xc) + w1 * __ldg(r1 + xc);
    }
  }
  __syncthreads();
  const int xq = blockIdx.x * 256 + threadIdx.x;
  if (xq * PX >= W) return;
  const float sw = (float)ws / (float)W;
  int x0[PX];
  float lx[PX];
#pragma unroll
  for (int j = 0; j < PX; ++j) {
    int x1;
    src_index_acf(sw, xq * PX + j, ws, x0[j], x1, lx[j]);  // x1 == x0 only at the clamped right edge, where srow[x0+1] == srow[x0]
  }
  size_t off = ((size_t)b * H + Y0) * W + (size_t)xq * PX;
#pragma unroll 2
  for (int r = 0; r < nrows; ++r, off += W) {
    float o[PX], dep[PX];
#pragma unroll
    for (int j = 0; j < PX; ++j) {
      const float a = srow[r][x0[j]], c = srow[r][x0[j] + 1];
      o[j] = fmaf(lx[j], c - a, a);
      dep[j] = d.dst_depth ? fast_rcp(fmaf(disp_range, o[j], min_disp)) : 0.0f;
    }
    if constexpr (VEC4) {
      if (d.dst_half) __stcs(reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(d.dst_half) + off), make_uint2(pack_f16x2(o[0], o[1]), pack_f16x2(o[2], o[3])));
      if (d.dst_disp) __stcs(reinterpret_cast<float4*>(d.dst_disp + off), make_float4(o[0], o[1], o[2], o[3]));
      if (d.dst_disp2) __stcs(reinterpret_cast<float4*>(d.dst_disp2 + off), make_float4(o[0], o[1], o[2], o[3]));
      if (d.dst_depth) __stcs(reinterpret_cast<float4*>(d.dst_depth + off), make_float4(dep[0], dep[1], dep[2], dep[3]));
    } else {
      if (d.dst_half) reinterpret_cast<uint16_t*>(d.dst_half)[off] = static_cast<uint16_t>(pack_f16x2(o[0], 0.0f) & 0xFFFFu);
      if (d.dst_disp) d.dst_disp[off] = o[0];
      if (d.dst_disp2) d.dst_disp2[off] = o[0];
      if (d.dst_depth) d.dst_depth[off] = dep[0];
    }
  }
}

// Fast variant for >= 3x horizontal upscaling with W % 4 == 0 (every 1024-px configuration: sources are <= 240 wide).
// The four consecutive pixels of a thread then fall into at most two adjacent source intervals, so each pixel is
// wa*s[i] + wb*s[i+1] + wc*s[i+2] with per-thread constant weights (one of wa, wc is zero): 3 shared-memory reads and 12
// FP instructions per 16-byte store, rows fully unrolled so every address is base + immediate.
#ifndef MDE_UP3_ROWS
#define MDE_UP3_ROWS 16
#endif
constexpr int kUp3Rows = MDE_UP3_ROWS;  // output rows per block (r02, 32 images x 4 maps at 1024 px: 4 rows 120 us, 8 rows 101, 16 rows 92.5, 32 rows 103)
constexpr int kUp3MaxSrcW = 344;        // 3 * ws < W: covers outputs up to 1032 px wide (wider ones take the generic kernel)
__global__ void __launch_bounds__(256) resize_maps_up3_kernel(const __grid_constant__ MapTable tab, int B, int H, int W, float min_disp, float disp_range) {
  pdl_launch_dependents();
  pdl_grid_wait();
  __shared__ float srow[kUp3Rows][kUp3MaxSrcW + 2];
  const int mi = blockIdx.z / B, b = blockIdx.z - mi * B;
  const mde_map_desc& d = tab.m[mi];
  const int ws = d.ws, hs = d.hs;
  const float sh = (float)hs / (float)H;
  const float mul = d.mul;
  const float* src = d.src + (size_t)b * hs * ws;
  const int Y0 = blockIdx.y * kUp3Rows;
  const int nrows = min(kUp3Rows, H - Y0);
  for (int r = 0; r < nrows; ++r) {
    int y0, y1;
    float ly;
    src_index_acf(sh, Y0 + r, hs, y0, y1, ly);
    const float* r0 = src + (size_t)y0 * ws;
    const float* r1 = src + (size_t)y1 * ws;
    const float w0 = (1.0f - ly) * mul, w1 = ly * mul;
    for (int x = threadIdx.x; x < ws + 2; x += 256) {
      const int xc = min(x, ws - 1);
      srow[r][x] = w0 * __ldg(r0 + xc) + w1 * __ldg(r1 + xc);
    }
  }
  __syncthreads();
  const int xq = blockIdx.x * 256 + threadIdx.x;
  if (xq * 4 >= W) return;
  const float sw = (float)ws / (float)W;
  int xb, x1;
  float l0;
  src_index_acf(sw, xq * 4, ws, xb, x1, l0);
  float wa[4], wb[4], wc[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int xj;
    float lj;
    src_index_acf(sw, xq * 4 + j, ws, xj, x1, lj);
    const bool same = xj == xb;  // else xj == xb + 1 (3 * sw < 1)
    wa[j] = same ? 1.0f - lj : 0.0f;
    wb[j] = same ? lj : 1.0f - lj;
    wc[j] = same ? 0.0f : lj;
  }
  const float* sp = &srow[0][xb];
  const size_t off0 = ((size_t)b * H + Y0) * W + (size_t)xq * 4;
  float* dd = d.dst_disp ? d.dst_disp + off0 : nullptr;
  float* d2 = d.dst_disp2 ? d.dst_disp2 + off0 : nullptr;
  float* dz = d.dst_depth ? d.dst_depth + off0 : nullptr;
  uint16_t* dh = d.dst_half ? reinterpret_cast<uint16_t*>(d.dst_half) + off0 : nullptr;
#pragma unroll
  for (int r = 0; r < kUp3Rows; ++r) {
    if (r < nrows) {
      const float a = sp[r * (kUp3MaxSrcW + 2)], bb = sp[r * (kUp3MaxSrcW + 2) + 1], c = sp[r * (kUp3MaxSrcW + 2) + 2];
      float o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] = fmaf(wc[j], c, fmaf(wb[j], bb, wa[j] * a));
      if (dh) __stcs(reinterpret_cast<uint2*>(dh + (size_t)r * W), make_uint2(pack_f16x2(o[0], o[1]), pack_f16x2(o[2], o[3])));
      if (dd) __stcs(reinterpret_cast<float4*>(dd + (size_t)r * W), make_float4(o[0], o[1], o[2], o[3]));
      if (d2) __stcs(reinterpret_cast<float4*>(d2 + (size_t)r * W), make_float4(o[0], o[1], o[2], o[3]));
      if (dz)
        __stcs(reinterpret_cast<float4*>(dz + (size_t)r * W),
               make_float4(fast_rcp(fmaf(disp_range, o[0], min_disp)), fast_rcp(fmaf(disp_range, o[1], min_disp)),
                           fast_rcp(fmaf(disp_range, o[2], min_disp)), fast_rcp(fmaf(disp_range, o[3], min_disp))));
    }
  }
}

// ------------------------------------------------------------------ enhanced low-res tail (enhanced_model.py:199-215,244-263)
__device__ __forceinline__ float bilinear_acf(const float* p, int hs, int ws, int Hd, int Wd, int Y, int X, float mul) {
  int y0, y1, x0, x1;
  float ly, lx;
  src_index_acf((float)hs / (float)Hd, Y, hs, y0, y1, ly);
  src_index_acf((float)ws / (float)Wd, X, ws, x0, x1, lx);
  const float v00 = p[(size_t)y0 * ws + x0] * mul, v01 = p[(size_t)y0 * ws + x1] * mul, v10 = p[(size_t)y1 * ws + x0] * mul,
              v11 = p[(size_t)y1 * ws + x1] * mul;
  return (1.0f - ly) * ((1.0f - lx) * v00 + lx * v01) + ly * ((1.0f - lx) * v10 + lx * v11);
}

__global__ void __launch_bounds__(256) enh_fuse_kernel(const float* __restrict__ d0, const float* __restrict__ d1, const float* __restrict__ d2,
                                                       const float* __restrict__ refn, int B, int h0, int w0, int gh, int gw, float mul,
                                                       const float* __restrict__ fw, float* __restrict__ fused) {
  pdl_launch_dependents();
  pdl_grid_wait();
  // softmax over the three fusion weights (enhanced_model.py:70)
  const float a0 = fw[0], a1 = fw[1], a2 = fw[2];
  const float m = fmaxf(a0, fmaxf(a1, a2));
  const float e0 = expf(a0 - m), e1 = expf(a1 - m), e2 = expf(a2 - m);
  const float inv = 1.0f / (e0 + e1 + e2);
  const float w_0 = e0 * inv, w_1 = e1 * inv, w_2 = e2 * inv;
  const int total = B * h0 * w0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int X = i % w0, Y = (i / w0) % h0, b = i / (w0 * h0);
    const float r = (gh != h0 || gw != w0) ? bilinear_acf(refn + (size_t)b * gh * gw, gh, gw, h0, w0, Y, X, 1.0f)
                                           : refn[((size_t)b * gh + Y) * gw + X];
    const float s0 = (d0[i] + r) * mul;
    const float s1 = bilinear_acf(d1 + (size_t)b * 4 * h0 * w0, 2 * h0, 2 * w0, h0, w0, Y, X, mul);
    const float s2 = bilinear_acf(d2 + (size_t)b * 16 * h0 * w0, 4 * h0, 4 * w0, h0, w0, Y, X, mul);
    fused[i] = w_0 * s0 + w_1 * s1 + w_2 * s2;
  }
}

// ------------------------------------------------------------------ caller-side input conversion (run_mars_depth.py:97-104)
// uint8 HWC (decoder output) -> float32 CHW / 255: `torch.from_numpy(arr).float().permute(2,0,1) / 255.0` on the device.
// One thread converts 4 pixels: three aligned 4-byte loads, three 16-byte streaming-friendly stores.
__global__ void __launch_bounds__(256) u8hwc_to_f32chw_kernel(const uint32_t* __restrict__ src, long long quads, int HW4, float4* __restrict__ dst) {
  pdl_launch_dependents();
  pdl_grid_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (long long)gridDim.x * blockDim.x) {
    const long long b = i / HW4;
    const int q = (int)(i - b * HW4);
    const uint32_t w0 = __ldg(src + 3 * i), w1 = __ldg(src + 3 * i + 1), w2 = __ldg(src + 3 * i + 2);
    // bytes: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
    const float r[4] = {(float)(w0 & 255u), (float)(w0 >> 24), (float)((w1 >> 16) & 255u), (float)((w2 >> 8) & 255u)};
    const float g[4] = {(float)((w0 >> 8) & 255u), (float)(w1 & 255u), (float)(w1 >> 24), (float)((w2 >> 16) & 255u)};
    const float bl[4] = {(float)((w0 >> 16) & 255u), (float)((w1 >> 8) & 255u), (float)(w2 & 255u), (float)(w2 >> 24)};
    float4* o = dst + b * 3 * HW4 + q;
    o[0] = make_float4(__fdiv_rn(r[0], 255.0f), __fdiv_rn(r[1], 255.0f), __fdiv_rn(r[2], 255.0f), __fdiv_rn(r[3], 255.0f));
    o[HW4] = make_float4(__fdiv_rn(g[0], 255.0f), __fdiv_rn(g[1], 255.0f), __fdiv_rn(g[2], 255.0f), __fdiv_rn(g[3], 255.0f));
    o[2 * (size_t)HW4] = make_float4(__fdiv_rn(bl[0], 255.0f), __fdiv_rn(bl[1], 255.0f), __fdiv_rn(bl[2], 255.0f), __fdiv_rn(bl[3], 255.0f));
  }
}

static inline int grid_for(long long work_items, int threads) {
  long long blocks = (work_items + threads - 1) / threads;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace mde

using namespace mde;

extern "C" int mde_input_range(const float* x, int B, int H, int W, const float* bn4x3, float eps, float* scratch, float* affine6,
                               mde_stream_t stream) {
  MDE_CHECK_ARG(x && bn4x3 && scratch && affine6, "mde_input_range: null pointer");
  MDE_CHECK_ARG(B > 0 && H > 0 && W > 0 && ((long long)H * W) % 4 == 0, "mde_input_range: H*W must be a positive multiple of 4 (H=%d W=%d)", H, W);
  MDE_CHECK_ARG((uintptr_t)x % 16 == 0, "mde_input_range: x must be 16-byte aligned");
  const long long n4 = (long long)B * 3 * H * W / 4;
  int parts = sm_count() * 4;
  if (parts > 1024) parts = 1024;
  if (parts > n4) parts = (int)n4;
  launch_k(minmax_kernel, dim3(parts), dim3(256), 0, stream, reinterpret_cast<const float4*>(x), n4, H * W / 4, scratch);
  MDE_CHECK_LAUNCH();
  launch_k(range_finalize_kernel, dim3(1), dim3(768), 0, stream, scratch, parts, bn4x3, eps, affine6);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_resize_bilinear_nchw(const float* src, int B, int C, int Hs, int Ws, float* dst, int Hd, int Wd, mde_stream_t stream) {
  MDE_CHECK_ARG(src && dst && B > 0 && C > 0 && Hs > 0 && Ws > 0 && Hd > 0 && Wd > 0, "mde_resize_bilinear_nchw: bad arguments");
  const long long total = (long long)B * C * Hd * Wd;
  launch_k(resize_nchw_kernel, dim3(grid_for(total, 256)), dim3(256), 0, stream, src, B * C, Hs, Ws, dst, Hd, Wd);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_preprocess_patchify(const float* x, int B, int H, int W, const float* affine6, mde_bf16* patches, mde_stream_t stream) {
  MDE_CHECK_ARG(x && affine6 && patches && B > 0 && H > 0 && W > 0, "mde_preprocess_patchify: bad arguments");
  MDE_CHECK_ARG((uintptr_t)patches % 16 == 0, "mde_preprocess_patchify: patches must be 16-byte aligned");
  const long long total = (long long)B * 196 * 96;
  launch_k(patchify_kernel, dim3(grid_for(total, 256)), dim3(256), 0, stream, x, B, H, W, affine6, reinterpret_cast<bf16*>(patches));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_layernorm_bf16(const float* x, int rows, int dim, const float* gamma, const float* beta, float eps, mde_bf16* out,
                                  mde_stream_t stream) {
  MDE_CHECK_ARG(x && gamma && beta && out && rows > 0, "mde_layernorm_bf16: bad arguments");
  MDE_CHECK_ARG(((uintptr_t)x | (uintptr_t)out | (uintptr_t)gamma | (uintptr_t)beta) % 16 == 0, "mde_layernorm_bf16: pointers must be 16-byte aligned");
  if (dim != 768) {
    set_error("mde_layernorm_bf16: dim=%d not built (768 only)", dim);
    return MDE_ERR_UNSUPPORTED;
  }
  // one resident wave: 4 blocks of 8 warps per SM (register-limited); rows per warp chosen so that the grid fits it
  static const int r_env = getenv("MDE_LN_ROWS") ? atoi(getenv("MDE_LN_ROWS")) : 0;
  const int slots = 4 * sm_count();
  const int R = r_env > 0 ? r_env : (ceil_div(rows, 8) <= slots ? 1 : 2);
  if (R == 1) launch_k(layernorm_kernel<768, 1>, dim3(ceil_div(rows, 8)), dim3(256), 0, stream, x, rows, gamma, beta, eps, reinterpret_cast<bf16*>(out));
  else launch_k(layernorm_kernel<768, 2>, dim3(ceil_div(rows, 16)), dim3(256), 0, stream, x, rows, gamma, beta, eps, reinterpret_cast<bf16*>(out));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_tokens_to_grid_bf16(const float* tokens, int B, int strip_cls, int G, mde_bf16* out, mde_stream_t stream) {
  MDE_CHECK_ARG(tokens && out && B > 0, "mde_tokens_to_grid_bf16: bad arguments");
  MDE_CHECK_ARG((strip_cls == 0 || strip_cls == 1) && G >= MDE_VIT_SEQ - strip_cls, "mde_tokens_to_grid_bf16: G=%d too small", G);
  const long long total = (long long)B * G * 96;
  launch_k(tokens_to_grid_kernel, dim3(grid_for(total, 256)), dim3(256), 0, stream, tokens, B, strip_cls, G, reinterpret_cast<bf16*>(out));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_instnorm_nhwc(const float* x, int B, int gh, int gw, int C, float eps, mde_bf16* out_padded, int out_cstride, int out_coff,
                                 mde_bf16* out_flat, mde_stream_t stream) {
  MDE_CHECK_ARG(x && out_padded && B > 0 && gh >= 3 && gw >= 3, "mde_instnorm_nhwc: bad arguments");
  MDE_CHECK_ARG(C % 16 == 0 && out_coff + C <= out_cstride, "mde_instnorm_nhwc: C must be a multiple of 16 and fit the output stride");
  MDE_CHECK_ARG(gh * gw <= kInLanes * kInMaxPer, "mde_instnorm_nhwc: grid %dx%d exceeds %d positions", gh, gw, kInLanes * kInMaxPer);
  MDE_CHECK_ARG(out_cstride % 4 == 0 && out_coff % 4 == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)out_padded % 8 == 0 && (uintptr_t)out_flat % 8 == 0,
                "mde_instnorm_nhwc: channel stride / offset must be multiples of 4 and the buffers 16 / 8-byte aligned");
  launch_k(instnorm_kernel, dim3(C / 16, B), dim3(256), 0, stream, x, gh, gw, C, eps, reinterpret_cast<bf16*>(out_padded), out_cstride, out_coff,
                                                       reinterpret_cast<bf16*>(out_flat));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_upsample2x_concat_nhwc(const mde_bf16* a, int B, int h, int w, int ca, const mde_bf16* s, int hs, int ws, int cs,
                                          mde_bf16* out, mde_stream_t stream) {
  MDE_CHECK_ARG(a && out && B > 0 && h >= 2 && w >= 2 && ca > 0 && ca % 8 == 0, "mde_upsample2x_concat_nhwc: bad arguments");
  MDE_CHECK_ARG((s == nullptr) == (cs == 0) && cs % 8 == 0, "mde_upsample2x_concat_nhwc: skip tensor / cs mismatch");
  if (s) MDE_CHECK_ARG(hs >= 2 && ws >= 2, "mde_upsample2x_concat_nhwc: bad skip size");
  MDE_CHECK_ARG((long long)B * (2 * h + 2) < (1ll << 31), "mde_upsample2x_concat_nhwc: too many rows");
  {
    const int groups = (ca + cs) >> 3;
    static const bool rows_off = getenv("MDE_UPCAT_ROWS") != nullptr && getenv("MDE_UPCAT_ROWS")[0] == '0';
    // rows per block: as many as still leave ~4 blocks per SM (the low-resolution levels have few rows: 960 at 28 x 30 x 32 images)
    int R = 4;
    while (R > 1 && (long long)B * ceil_div(2 * h + 2, R) < 4ll * sm_count()) R >>= 1;
    const size_t smem = (size_t)R * ((size_t)w * ca + (size_t)ws * cs) * sizeof(float);
    if (!rows_off && groups > 0 && (groups & (groups - 1)) == 0 && smem <= 100 * 1024) {
      static bool attr = false;
      if (!attr) {
        MDE_CHECK_CUDA(cudaFuncSetAttribute(upsample_concat_rows_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        MDE_CHECK_CUDA(cudaFuncSetAttribute(upsample_concat_rows_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        MDE_CHECK_CUDA(cudaFuncSetAttribute(upsample_concat_rows_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        attr = true;
      }
      int lg = 0;
      while ((1 << lg) < groups) ++lg;
      const dim3 grid(B * ceil_div(2 * h + 2, R));
#define MDE_UPCAT_CASE(RR)                                                                                                                        \
  if (R == RR)                                                                                                                                    \
    launch_k(upsample_concat_rows_kernel<RR>, grid, dim3(256), smem, stream, reinterpret_cast<const bf16*>(a), B, h, w, ca, reinterpret_cast<const bf16*>(s), \
             hs, ws, cs, reinterpret_cast<bf16*>(out), lg);
      MDE_UPCAT_CASE(4) MDE_UPCAT_CASE(2) MDE_UPCAT_CASE(1)
#undef MDE_UPCAT_CASE
      MDE_CHECK_LAUNCH();
      return MDE_OK;
    }
  }
  launch_k(upsample_concat_kernel, dim3(B * (2 * h + 2)), dim3(256), 0, stream, reinterpret_cast<const bf16*>(a), B, h, w, ca, reinterpret_cast<const bf16*>(s),
                                                                   hs, ws, cs, reinterpret_cast<bf16*>(out));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_rgb_to_grid_nhwc(const float* x, int B, int H, int W, int gh, int gw, mde_bf16* out_padded, int out_cstride, int out_coff,
                                    int cfill, mde_stream_t stream) {
  MDE_CHECK_ARG(x && out_padded && B > 0 && cfill >= 8 && out_coff + cfill <= out_cstride, "mde_rgb_to_grid_nhwc: bad arguments");
  MDE_CHECK_ARG(cfill % 8 == 0 && out_coff % 8 == 0 && out_cstride % 8 == 0 && (uintptr_t)out_padded % 16 == 0,
                "mde_rgb_to_grid_nhwc: cfill, out_coff and out_cstride must be multiples of 8 channels (16-byte stores)");
  launch_k(rgb_to_grid_kernel<false>, dim3(grid_for((long long)B * gh * gw * (cfill / 8), 256)), dim3(256), 0, stream, static_cast<const void*>(x), B, H, W, gh, gw,
           reinterpret_cast<bf16*>(out_padded), out_cstride, out_coff, cfill);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_rgb_to_grid_nhwc_u8(const uint8_t* x_hwc, int B, int H, int W, int gh, int gw, mde_bf16* out_padded, int out_cstride, int out_coff,
                                       int cfill, mde_stream_t stream) {
  MDE_CHECK_ARG(x_hwc && out_padded && B > 0 && cfill >= 8 && out_coff + cfill <= out_cstride, "mde_rgb_to_grid_nhwc_u8: bad arguments");
  MDE_CHECK_ARG(cfill % 8 == 0 && out_coff % 8 == 0 && out_cstride % 8 == 0 && (uintptr_t)out_padded % 16 == 0,
                "mde_rgb_to_grid_nhwc_u8: cfill, out_coff and out_cstride must be multiples of 8 channels (16-byte stores)");
  launch_k(rgb_to_grid_kernel<true>, dim3(grid_for((long long)B * gh * gw * (cfill / 8), 256)), dim3(256), 0, stream, static_cast<const void*>(x_hwc), B, H, W, gh, gw,
           reinterpret_cast<bf16*>(out_padded), out_cstride, out_coff, cfill);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_resize_maps(const mde_map_desc* maps, int n_maps, int B, int H, int W, float min_disp, float max_disp, mde_stream_t stream) {
  MDE_CHECK_ARG(maps && n_maps >= 1 && n_maps <= MDE_MAX_MAPS && B > 0 && H > 0 && W > 0, "mde_resize_maps: bad arguments");
  MapTable tab;
  tab.n = n_maps;
  bool vec = (W % 4 == 0);
  for (int i = 0; i < n_maps; ++i) {
    tab.m[i] = maps[i];
    MDE_CHECK_ARG(maps[i].src && maps[i].hs > 0 && maps[i].ws > 0 && maps[i].ws <= kMaxSrcW, "mde_resize_maps: map %d has no source or is wider than %d", i, kMaxSrcW);
    const uintptr_t al = (uintptr_t)maps[i].dst_disp | (uintptr_t)maps[i].dst_disp2 | (uintptr_t)maps[i].dst_depth;
    if (al % 16 != 0 || (uintptr_t)maps[i].dst_half % 8 != 0) vec = false;
  }
  const float range = max_disp - min_disp;
  MDE_CHECK_ARG(H <= 65535 && (long long)n_maps * B <= 65535, "mde_resize_maps: H and n_maps*B must be <= 65535");
  if (!vec) {
    launch_k(resize_maps_kernel<false>, dim3(ceil_div(W, 256), ceil_div(H, kRowsPerBlock), n_maps * B), dim3(256), 0, stream, tab, B, H, W, min_disp, range);
    MDE_CHECK_LAUNCH();
    return MDE_OK;
  }
  // The kernel is chosen PER MAP (>= 3x horizontal upscaling takes the 3-tap form), never per call: the two forms round
  // differently, and a map must come out the same whichever other maps are requested with it (model.forward_keys).
  MapTable fast, slow;
  fast.n = slow.n = 0;
  for (int i = 0; i < n_maps; ++i) {
    const bool up3 = 3 * maps[i].ws < W && maps[i].ws <= kUp3MaxSrcW;
    MapTable& t = up3 ? fast : slow;
    t.m[t.n++] = maps[i];
  }
  if (fast.n) {
    launch_k(resize_maps_up3_kernel, dim3(ceil_div(W / 4, 256), ceil_div(H, kUp3Rows), fast.n * B), dim3(256), 0, stream, fast, B, H, W, min_disp, range);
    MDE_CHECK_LAUNCH();
  }
  if (slow.n) {
    launch_k(resize_maps_kernel<true>, dim3(ceil_div(W / 4, 256), ceil_div(H, kRowsPerBlock), slow.n * B), dim3(256), 0, stream, slow, B, H, W, min_disp, range);
    MDE_CHECK_LAUNCH();
  }
  return MDE_OK;
}

extern "C" int mde_enh_fuse(const float* d0, const float* d1, const float* d2, const float* refinement, int B, int h0, int w0, int gh, int gw,
                            float mul, const float* fusion_w3, float* fused, mde_stream_t stream) {
  MDE_CHECK_ARG(d0 && d1 && d2 && refinement && fusion_w3 && fused && B > 0 && h0 > 0 && w0 > 0, "mde_enh_fuse: bad arguments");
  launch_k(enh_fuse_kernel, dim3(grid_for((long long)B * h0 * w0, 256)), dim3(256), 0, stream, d0, d1, d2, refinement, B, h0, w0, gh, gw, mul, fusion_w3, fused);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_u8hwc_to_f32chw(const uint8_t* src, int B, int H, int W, float* dst, mde_stream_t stream) {
  MDE_CHECK_ARG(src && dst && B > 0 && H > 0 && W > 0, "mde_u8hwc_to_f32chw: bad arguments");
  MDE_CHECK_ARG(((long long)H * W) % 4 == 0 && (uintptr_t)src % 4 == 0 && (uintptr_t)dst % 16 == 0,
                "mde_u8hwc_to_f32chw: H*W must be a multiple of 4, src 4-byte and dst 16-byte aligned");
  const int hw4 = (int)(((long long)H * W) / 4);
  const long long quads = (long long)B * hw4;
  launch_k(u8hwc_to_f32chw_kernel, dim3(grid_for(quads, 256)), dim3(256), 0, stream, reinterpret_cast<const uint32_t*>(src), quads, hw4, reinterpret_cast<float4*>(dst));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

__global__ void __launch_bounds__(256) f32_to_f16_kernel(const float* __restrict__ src, long long n, uint16_t* __restrict__ dst) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const long long n8 = n >> 3;
  const float4* s4 = reinterpret_cast<const float4*>(src);
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const float4 a = __ldcs(s4 + 2 * i), b = __ldcs(s4 + 2 * i + 1);
    __stcs(d4 + i, make_uint4(pack_f16x2(a.x, a.y), pack_f16x2(a.z, a.w), pack_f16x2(b.x, b.y), pack_f16x2(b.z, b.w)));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
    const long long i = (n8 << 3) + threadIdx.x;
    dst[i] = static_cast<uint16_t>(pack_f16x2(src[i], 0.0f) & 0xFFFFu);
  }
}

extern "C" int mde_f32_to_f16(const float* src, long long n, mde_f16* dst, mde_stream_t stream) {
  MDE_CHECK_ARG(src && dst && n > 0, "mde_f32_to_f16: bad arguments");
  MDE_CHECK_ARG((uintptr_t)src % 16 == 0 && (uintptr_t)dst % 16 == 0, "mde_f32_to_f16: pointers must be 16-byte aligned");
  launch_k(f32_to_f16_kernel, dim3(grid_for(n >> 3, 256)), dim3(256), 0, stream, src, n, dst);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_input_stage(const float* x, int B, int H, int W, const float* bn4x3, float eps, float* scratch, float* affine6,
                               float* raw224, mde_bf16* patches, mde_stream_t stream) {
  MDE_CHECK_ARG(x && bn4x3 && scratch && affine6 && raw224 && patches, "mde_input_stage: null pointer");
  MDE_CHECK_ARG(B > 0 && H > 0 && W > 0 && W % 4 == 0 && (uintptr_t)x % 16 == 0, "mde_input_stage: W must be a multiple of 4 and x 16-byte aligned (H=%d W=%d)", H, W);
  MDE_CHECK_ARG((uintptr_t)raw224 % 16 == 0 && (uintptr_t)patches % 16 == 0, "mde_input_stage: raw224 / patches must be 16-byte aligned");
  const int max_parts = MDE_RANGE_SCRATCH_FLOATS / 6;
  int rpb = ceil_div(224 * 3 * B, max_parts);  // output rows per block: as few as the partials buffer allows, at least kMinRows
  static const int kMinRows = getenv("MDE_SCAN_ROWS") ? atoi(getenv("MDE_SCAN_ROWS")) : 3;  // r02: 3 rows 77.0 us, 4 rows 79.0, 6 rows 92.7 (with 8 loads in flight)
  if (rpb < kMinRows) rpb = kMinRows;
  MDE_CHECK_ARG(rpb <= 224, "mde_input_stage: batch %d too large for the partials buffer", B);
  const int groups = ceil_div(224, rpb);
  const int parts = groups * 3 * B;
  launch_k(input_scan_resample_kernel, dim3(parts), dim3(256), 0, stream, x, H, W, rpb, groups, scratch, raw224);
  MDE_CHECK_LAUNCH();
  launch_k(range_finalize_kernel, dim3(1), dim3(768), 0, stream, scratch, parts, bn4x3, eps, affine6);
  MDE_CHECK_LAUNCH();
  launch_k(patchify224_kernel, dim3(grid_for((long long)B * 196 * 96, 256)), dim3(256), 0, stream, raw224, B, affine6, reinterpret_cast<bf16*>(patches));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_input_stage_u8(const uint8_t* x_hwc, int B, int H, int W, const float* bn4x3, float eps, float* scratch, float* affine6,
                                  float* raw224, mde_bf16* patches, mde_stream_t stream) {
  MDE_CHECK_ARG(x_hwc && bn4x3 && scratch && affine6 && raw224 && patches, "mde_input_stage_u8: null pointer");
  MDE_CHECK_ARG(B > 0 && H > 0 && W > 0 && W % 4 == 0 && (uintptr_t)x_hwc % 4 == 0, "mde_input_stage_u8: W must be a multiple of 4 and x 4-byte aligned (H=%d W=%d)", H, W);
  MDE_CHECK_ARG((uintptr_t)raw224 % 16 == 0 && (uintptr_t)patches % 16 == 0, "mde_input_stage_u8: raw224 / patches must be 16-byte aligned");
  const int max_parts = MDE_RANGE_SCRATCH_FLOATS / 6;
  int rpb = ceil_div(224 * B, max_parts);  // output rows per block (all three channels): as few as the partials buffer allows, at least 2
  if (rpb < 2) rpb = 2;
  MDE_CHECK_ARG(rpb <= 224, "mde_input_stage_u8: batch %d too large for the partials buffer", B);
  const int groups = ceil_div(224, rpb);
  const int parts = groups * B;
  launch_k(input_scan_resample_u8_kernel, dim3(parts), dim3(256), 0, stream, x_hwc, H, W, rpb, groups, scratch, raw224);
  MDE_CHECK_LAUNCH();
  launch_k(range_finalize_kernel, dim3(1), dim3(768), 0, stream, scratch, parts, bn4x3, eps, affine6);
  MDE_CHECK_LAUNCH();
  launch_k(patchify224_kernel, dim3(grid_for((long long)B * 196 * 96, 256)), dim3(256), 0, stream, raw224, B, affine6, reinterpret_cast<bf16*>(patches));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

// ------------------------------------------------------------------ mde_input_stage without caller scratch (no allocation anywhere)
// Added for callers that cannot cudaMalloc (code under CUDA-graph capture, tight serving
// loops, harnesses that forbid it); it has no counterpart in the reference.  The partial-record scratch of the scan lives in a
// file-scope device array: MDE_INPUT_STAGE_NOALLOC_MAX_BATCH images x 3 planes x 56 row groups x 6 floats = 126 KB, held by every
// GPU that loads the library until the process exits.  In HOST code the name `g_input_stage_scratch` is only a stand-in for the
// symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kNoallocRowsPerBlock = 4;                                   // the schedule mde_input_stage uses up to batch 48
constexpr int kNoallocGroups = 224 / kNoallocRowsPerBlock;                // 56
constexpr int kNoallocMaxParts = MDE_INPUT_STAGE_NOALLOC_MAX_BATCH * 3 * kNoallocGroups;
static __device__ __align__(256) float g_input_stage_scratch[kNoallocMaxParts * 6];

static int noalloc_scratch_address(float** out) {
  static std::mutex mu;
  static float* addr[64] = {nullptr};
  int dev = 0;
  MDE_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("mde_input_stage_noalloc: device index %d out of range", dev);
    return MDE_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    MDE_CHECK_CUDA(cudaGetSymbolAddress(&p, g_input_stage_scratch));  // the CURRENT device's copy of the array
    addr[dev] = static_cast<float*>(p);
  }
  *out = addr[dev];
  return MDE_OK;
}

extern "C" int mde_input_stage_noalloc(const float* x, int B, int H, int W, const float* bn4x3, float eps, float* affine6, float* raw224,
                                       mde_bf16* patches, mde_stream_t stream) {
  MDE_CHECK_ARG(x && bn4x3 && affine6 && raw224 && patches, "mde_input_stage_noalloc: null pointer");
  MDE_CHECK_ARG(B > 0 && H > 0 && W > 0 && W % 4 == 0 && (uintptr_t)x % 16 == 0,
                "mde_input_stage_noalloc: W must be a multiple of 4 and x 16-byte aligned (H=%d W=%d)", H, W);
  MDE_CHECK_ARG((uintptr_t)raw224 % 16 == 0 && (uintptr_t)patches % 16 == 0, "mde_input_stage_noalloc: raw224 / patches must be 16-byte aligned");
  MDE_CHECK_ARG(B <= MDE_INPUT_STAGE_NOALLOC_MAX_BATCH,
                "mde_input_stage_noalloc: batch %d exceeds the built-in scratch (MDE_INPUT_STAGE_NOALLOC_MAX_BATCH = %d images); use "
                "mde_input_stage with caller scratch", B, MDE_INPUT_STAGE_NOALLOC_MAX_BATCH);
  float* scratch = nullptr;
  const int rc = noalloc_scratch_address(&scratch);
  if (rc != MDE_OK) return rc;
  const int parts = kNoallocGroups * 3 * B;  // <= kNoallocMaxParts; every record the finalize kernel reads is written by the scan first
  launch_k(input_scan_resample_kernel, dim3(parts), dim3(256), 0, stream, x, H, W, kNoallocRowsPerBlock, kNoallocGroups, scratch, raw224);
  MDE_CHECK_LAUNCH();
  launch_k(range_finalize_kernel, dim3(1), dim3(768), 0, stream, scratch, parts, bn4x3, eps, affine6);
  MDE_CHECK_LAUNCH();
  launch_k(patchify224_kernel, dim3(grid_for((long long)B * 196 * 96, 256)), dim3(256), 0, stream, raw224, B, affine6, reinterpret_cast<bf16*>(patches));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
