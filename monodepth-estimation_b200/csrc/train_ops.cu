// Training-step kernels around the encoder backward (train/mars_adapter.py:117-153,209-247): the decoder
// path that ('disp', 0) depends on (decoder.py:97-166, level 0 only: init_conv, upconv_0_0, skip_0, upconv_0_1, dispconv_0) in
// train mode — BatchNorm with batch statistics — the loss (layers.py:202-215 + the L1 term of mars_adapter.py:238-240) and Adam.
//
// Train-mode activations are plain fp32 NHWC [B][h][w][C] (210 .. 840 positions per image: small), convolutions are
// im2col + the tcgen05 GEMM in both directions.  Everything is deterministic: reductions run in a fixed order, no atomics.
#include <math.h>

#include "common.h"
#include "ptx.cuh"

namespace mde {

static inline int grid1(long long items) {
  long long g = (items + 255) / 256;
  const long long cap = (long long)sm_count() * 32;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
#define MDE_GRID_STRIDE(i, n) for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (long long)gridDim.x * blockDim.x)

__device__ __forceinline__ int reflect1(int i, int n) { return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i); }  // ReflectionPad2d(1)

// ------------------------------------------------------------------ im2col (decoder.py:8-19: pad 1, 3x3)
// x f32 [B][h][w][C] -> col bf16 [B*h*w][9*C] (k = (ky*3+kx)*C + c) and/or colT bf16 [9*C][ldt] (zero beyond B*h*w)
__global__ void __launch_bounds__(256) im2col_kernel(const float* __restrict__ x, int B, int h, int w, int C, int reflect, bf16* __restrict__ col,
                                                     bf16* __restrict__ colT, int ldt) {
  const long long npx = (long long)B * h * w, K = 9LL * C;
  const long long total = (colT ? (long long)ldt : npx) * K;
  MDE_GRID_STRIDE(i, total) {
    const long long p = i / K;
    const int k = (int)(i - p * K);
    float v = 0.0f;
    if (p < npx) {
      const int t = k / C, c = k - t * C, ky = t / 3, kx = t - 3 * ky;
      const int xx = (int)(p % w), y = (int)((p / w) % h), b = (int)(p / ((long long)w * h));
      int sy = y + ky - 1, sx = xx + kx - 1;
      bool ok = true;
      if (reflect) {
        sy = reflect1(sy, h);
        sx = reflect1(sx, w);
      } else {
        ok = sy >= 0 && sy < h && sx >= 0 && sx < w;
      }
      if (ok) v = x[(((size_t)b * h + sy) * w + sx) * C + c];
      if (col) col[(size_t)p * K + k] = __float2bfloat16_rn(v);
    }
    if (colT) colT[(size_t)k * ldt + p] = __float2bfloat16_rn(v);
  }
}

// dxcol f32 [B*h*w][ldc] (columns k = (ky*3+kx)*C + c) -> dx f32 [B][h][w][C]: the adjoint of im2col.  Gather form: padded
// position (Y,X) collects the taps of the outputs that read it; with reflection the border rows/columns fold onto rows 1 / h-2.
__global__ void __launch_bounds__(256) col2im_kernel(const float* __restrict__ dxcol, int ldc, int B, int h, int w, int C, int reflect, float* __restrict__ dx) {
  const long long total = (long long)B * h * w * C;
  MDE_GRID_STRIDE(i, total) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int xx = (int)(p % w), y = (int)((p / w) % h), b = (int)(p / ((long long)w * h));
    int Ys[2], Xs[2], ny = 0, nx = 0;  // padded coordinates (interior pixel (y,x) sits at (y+1, x+1))
    Ys[ny++] = y + 1;
    Xs[nx++] = xx + 1;
    if (reflect) {
      if (y == 1) Ys[ny++] = 0;
      if (y == h - 2) Ys[ny++] = h + 1;
      if (xx == 1) Xs[nx++] = 0;
      if (xx == w - 2) Xs[nx++] = w + 1;
    }
    float acc = 0.0f;
    for (int a = 0; a < ny; ++a)
      for (int e = 0; e < nx; ++e)
        for (int ky = 0; ky < 3; ++ky) {
          const int yo = Ys[a] - ky;  // output row whose tap ky reads padded row Ys[a]
          if (yo < 0 || yo >= h) continue;
          for (int kx = 0; kx < 3; ++kx) {
            const int xo = Xs[e] - kx;
            if (xo < 0 || xo >= w) continue;
            acc += dxcol[(((size_t)b * h + yo) * w + xo) * ldc + (ky * 3 + kx) * C + c];
          }
        }
    dx[i] = acc;
  }
}

// ------------------------------------------------------------------ BatchNorm2d in train mode + ELU (decoder.py:22-34)
// z f32 [N][C].  stats: per channel mean and 1/sqrt(biased var + eps) over the N rows; running statistics updated as
// nn.BatchNorm2d does (momentum, UNBIASED variance).  Grid (C/32, row chunks): every thread adds its rows in double, the block's
// eight warps are added in warp order, the finalize kernel adds the chunks in chunk order (deterministic).
constexpr int kRedChunks = 64;
static inline int red_chunks(long long rows, int rows_per_chunk_min) {
  long long c = (rows + rows_per_chunk_min - 1) / rows_per_chunk_min;
  return (int)(c < 1 ? 1 : (c > kRedChunks ? kRedChunks : c));
}
__global__ void __launch_bounds__(256) bn_stats_partial_kernel(const float* __restrict__ z, int N, int C, int ldz, double* __restrict__ part) {
  __shared__ double p1[8][32], p2[8][32];
  const int l = threadIdx.x & 31, wq = threadIdx.x >> 5, c = blockIdx.x * 32 + l;
  const int per = (N + gridDim.y - 1) / gridDim.y, r0 = blockIdx.y * per, r1 = min(N, r0 + per);
  double a = 0.0, q = 0.0;
  if (c < C)
    for (int r = r0 + wq; r < r1; r += 8) {
      const double v = z[(size_t)r * ldz + c];
      a += v;
      q += v * v;
    }
  p1[wq][l] = a;
  p2[wq][l] = q;
  __syncthreads();
  if (wq == 0 && c < C) {
    double ta = 0.0, tq = 0.0;
    for (int k = 0; k < 8; ++k) {
      ta += p1[k][l];
      tq += p2[k][l];
    }
    part[((size_t)blockIdx.y * C + c) * 2] = ta;
    part[((size_t)blockIdx.y * C + c) * 2 + 1] = tq;
  }
}
__global__ void __launch_bounds__(256) bn_stats_finalize_kernel(const double* __restrict__ part, int chunks, int N, int C, float eps, float* __restrict__ mean,
                                                                float* __restrict__ rstd, float* __restrict__ run_mean, float* __restrict__ run_var, float momentum) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  double a = 0.0, q = 0.0;
  for (int k = 0; k < chunks; ++k) {
    a += part[((size_t)k * C + c) * 2];
    q += part[((size_t)k * C + c) * 2 + 1];
  }
  const double m = a / N, var = fmax(q / N - m * m, 0.0);
  mean[c] = (float)m;
  rstd[c] = (float)(1.0 / sqrt(var + (double)eps));
  if (run_mean) {
    run_mean[c] = (1.0f - momentum) * run_mean[c] + momentum * (float)m;
    run_var[c] = (1.0f - momentum) * run_var[c] + momentum * (float)(N > 1 ? var * N / (N - 1.0) : var);
  }
}
__global__ void __launch_bounds__(256) bn_elu_apply_kernel(const float* __restrict__ z, long long N, int C, int ldz, const float* __restrict__ mean,
                                                           const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           float* __restrict__ y) {
  MDE_GRID_STRIDE(i, N * C) {
    const int c = (int)(i % C);
    const long long r = i / C;
    const float a = fmaf((z[(size_t)r * ldz + c] - mean[c]) * rstd[c], gamma[c], beta[c]);
    y[i] = a > 0.0f ? a : expm1f(a);
  }
}
// backward: da = dy * ELU'(a); dbeta = sum da; dgamma = sum da zhat; dz = gamma rstd (da - dbeta/N - zhat dgamma/N)
__global__ void __launch_bounds__(256) bn_elu_bwd_reduce_kernel(const float* __restrict__ z, int N, int C, int ldz, const float* __restrict__ mean,
                                                                const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                const float* __restrict__ dy, double* __restrict__ part) {
  __shared__ float pg[8][32], pb[8][32];
  const int l = threadIdx.x & 31, wq = threadIdx.x >> 5, c = blockIdx.x * 32 + l;
  const int per = (N + gridDim.y - 1) / gridDim.y, r0 = blockIdx.y * per, r1 = min(N, r0 + per);
  float sg = 0.0f, sb = 0.0f;
  if (c < C) {
    const float m = mean[c], rs = rstd[c], g = gamma[c], be = beta[c];
    for (int r = r0 + wq; r < r1; r += 8) {
      const float zh = (z[(size_t)r * ldz + c] - m) * rs;
      const float a = fmaf(zh, g, be);
      const float da = dy[(size_t)r * C + c] * (a > 0.0f ? 1.0f : expf(a));
      sg = fmaf(da, zh, sg);
      sb += da;
    }
  }
  pg[wq][l] = sg;
  pb[wq][l] = sb;
  __syncthreads();
  if (wq == 0 && c < C) {
    double tg = 0.0, tb = 0.0;
    for (int k = 0; k < 8; ++k) {
      tg += pg[k][l];
      tb += pb[k][l];
    }
    part[((size_t)blockIdx.y * C + c) * 2] = tg;
    part[((size_t)blockIdx.y * C + c) * 2 + 1] = tb;
  }
}
__global__ void __launch_bounds__(256) bn_elu_bwd_finalize_kernel(const double* __restrict__ part, int chunks, int C, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  double g = 0.0, b = 0.0;
  for (int k = 0; k < chunks; ++k) {
    g += part[((size_t)k * C + c) * 2];
    b += part[((size_t)k * C + c) * 2 + 1];
  }
  dgamma[c] = (float)g;
  dbeta[c] = (float)b;
}
__global__ void __launch_bounds__(256) bn_elu_bwd_apply_kernel(const float* __restrict__ z, long long N, int C, int ldz, const float* __restrict__ mean,
                                                               const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                                                               const float* __restrict__ dy, const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                                                               float* __restrict__ dz) {
  const float invn = 1.0f / (float)N;
  MDE_GRID_STRIDE(i, N * C) {
    const int c = (int)(i % C);
    const long long r = i / C;
    const float zh = (z[(size_t)r * ldz + c] - mean[c]) * rstd[c];
    const float a = fmaf(zh, gamma[c], beta[c]);
    const float da = dy[i] * (a > 0.0f ? 1.0f : expf(a));
    dz[i] = gamma[c] * rstd[c] * (da - dbeta[c] * invn - zh * dgamma[c] * invn);
  }
}

// ------------------------------------------------------------------ bilinear resize, NHWC f32, forward and adjoint
__device__ __forceinline__ void src_idx(int dst, int in, int out, int align, int& i0, int& i1, float& l1) {
  float s;
  if (align) s = out > 1 ? (float)dst * ((float)(in - 1) / (float)(out - 1)) : 0.0f;  // F.interpolate(align_corners=True)
  else {
    s = ((float)in / (float)out) * ((float)dst + 0.5f) - 0.5f;                          // area_pixel_compute_source_index
    s = s < 0.0f ? 0.0f : s;
  }
  i0 = min((int)s, in - 1);
  i1 = min(i0 + 1, in - 1);
  l1 = s - (float)i0;
}
__global__ void __launch_bounds__(256) bilinear_fwd_kernel(const float* __restrict__ src, int B, int hs, int ws, int C, float* __restrict__ dst, int hd, int wd,
                                                           int align, int ldd, int coff) {
  MDE_GRID_STRIDE(i, (long long)B * hd * wd * C) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int X = (int)(p % wd), Y = (int)((p / wd) % hd), b = (int)(p / ((long long)wd * hd));
    int y0, y1, x0, x1;
    float ly, lx;
    src_idx(Y, hs, hd, align, y0, y1, ly);
    src_idx(X, ws, wd, align, x0, x1, lx);
    const float* s = src + (size_t)b * hs * ws * C + c;
    const float v = (1.0f - ly) * ((1.0f - lx) * s[((size_t)y0 * ws + x0) * C] + lx * s[((size_t)y0 * ws + x1) * C]) +
                    ly * ((1.0f - lx) * s[((size_t)y1 * ws + x0) * C] + lx * s[((size_t)y1 * ws + x1) * C]);
    dst[(size_t)p * ldd + coff + c] = v;
  }
}
// adjoint, gather form: source pixel (y,x) visits the destination pixels whose taps can touch it
__global__ void __launch_bounds__(256) bilinear_bwd_kernel(const float* __restrict__ dd, int B, int hd, int wd, int C, int ldd, int coff, float* __restrict__ ds, int hs,
                                                           int ws, int align) {
  const float ry = (float)hd / (float)hs, rx = (float)wd / (float)ws;  // destination pixels per source pixel
  MDE_GRID_STRIDE(i, (long long)B * hs * ws * C) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int x = (int)(p % ws), y = (int)((p / ws) % hs), b = (int)(p / ((long long)ws * hs));
    const int Y0 = max(0, (int)floorf((y - 1) * ry) - 2), Y1 = min(hd - 1, (int)ceilf((y + 2) * ry) + 2);
    const int X0 = max(0, (int)floorf((x - 1) * rx) - 2), X1 = min(wd - 1, (int)ceilf((x + 2) * rx) + 2);
    float acc = 0.0f;
    for (int Y = Y0; Y <= Y1; ++Y) {
      int a0, a1;
      float ly;
      src_idx(Y, hs, hd, align, a0, a1, ly);
      const float wy = (a0 == y ? 1.0f - ly : 0.0f) + (a1 == y ? ly : 0.0f);
      if (wy == 0.0f) continue;
      float row = 0.0f;
      for (int X = X0; X <= X1; ++X) {
        int e0, e1;
        float lx;
        src_idx(X, ws, wd, align, e0, e1, lx);
        const float wx = (e0 == x ? 1.0f - lx : 0.0f) + (e1 == x ? lx : 0.0f);
        if (wx != 0.0f) row = fmaf(wx, dd[(((size_t)b * hd + Y) * wd + X) * ldd + coff + c], row);
      }
      acc = fmaf(wy, row, acc);
    }
    ds[i] = acc;
  }
}

// The same adjoint in two separable steps for large magnifications (the disparity map back from the input size to the decoder's
// level-0 grid: every source pixel collects ~(2 hd/hs)^2 destination pixels and there are few source pixels): first along x into
// tmp [B][hd][ws][C], then along y.  The order of the additions is the direct kernel's (row sums, then the sum over rows).
__global__ void __launch_bounds__(256) bilinear_bwd_x_kernel(const float* __restrict__ dd, int B, int hd, int wd, int C, int ldd, int coff, float* __restrict__ tmp, int ws,
                                                             int align) {
  const float rx = (float)wd / (float)ws;
  MDE_GRID_STRIDE(i, (long long)B * hd * ws * C) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int x = (int)(p % ws);
    const long long bY = p / ws;  // b * hd + Y
    const int X0 = max(0, (int)floorf((x - 1) * rx) - 2), X1 = min(wd - 1, (int)ceilf((x + 2) * rx) + 2);
    float row = 0.0f;
    for (int X = X0; X <= X1; ++X) {
      int e0, e1;
      float lx;
      src_idx(X, ws, wd, align, e0, e1, lx);
      const float wx = (e0 == x ? 1.0f - lx : 0.0f) + (e1 == x ? lx : 0.0f);
      if (wx != 0.0f) row = fmaf(wx, dd[((size_t)bY * wd + X) * ldd + coff + c], row);
    }
    tmp[i] = row;
  }
}
__global__ void __launch_bounds__(256) bilinear_bwd_y_kernel(const float* __restrict__ tmp, int B, int hd, int C, float* __restrict__ ds, int hs, int ws, int align) {
  const float ry = (float)hd / (float)hs;
  MDE_GRID_STRIDE(i, (long long)B * hs * ws * C) {
    const int c = (int)(i % C);
    const long long p = i / C;
    const int x = (int)(p % ws), y = (int)((p / ws) % hs), b = (int)(p / ((long long)ws * hs));
    const int Y0 = max(0, (int)floorf((y - 1) * ry) - 2), Y1 = min(hd - 1, (int)ceilf((y + 2) * ry) + 2);
    float acc = 0.0f;
    for (int Y = Y0; Y <= Y1; ++Y) {
      int a0, a1;
      float ly;
      src_idx(Y, hs, hd, align, a0, a1, ly);
      const float wy = (a0 == y ? 1.0f - ly : 0.0f) + (a1 == y ? ly : 0.0f);
      if (wy == 0.0f) continue;
      acc = fmaf(wy, tmp[(((size_t)b * hd + Y) * ws + x) * C + c], acc);
    }
    ds[i] = acc;
  }
}

// ------------------------------------------------------------------ instance norm (encoder.py:167) and its backward, x f32 [B][G][C], no affine
// One block per (32 channels, image): lane = channel, the eight warps stride the G grid positions; the eight partial sums are added
// in warp order.  sum8 returns the total to every thread of the block.
__device__ __forceinline__ float instnorm_sum8(float v, float (*buf)[32], int w, int l) {
  __syncthreads();  // the previous round's readers are done
  buf[w][l] = v;
  __syncthreads();
  float t = 0.0f;
#pragma unroll
  for (int k = 0; k < 8; ++k) t += buf[k][l];
  return t;
}
__global__ void __launch_bounds__(256) instnorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, int G, int C, float eps, float* __restrict__ dx) {
  __shared__ float buf[8][32];
  const int l = threadIdx.x & 31, w = threadIdx.x >> 5, c = blockIdx.x * 32 + l, b = blockIdx.y;
  const bool ok = c < C;
  const float* xb = x + (size_t)b * G * C + (ok ? c : 0);
  const float* db = dy + (size_t)b * G * C + (ok ? c : 0);
  float s = 0.0f;
  if (ok)
    for (int g = w; g < G; g += 8) s += xb[(size_t)g * C];
  const float m = instnorm_sum8(s, buf, w, l) / (float)G;
  float q = 0.0f;
  if (ok)
    for (int g = w; g < G; g += 8) {
      const float d = xb[(size_t)g * C] - m;
      q = fmaf(d, d, q);
    }
  const float rs = 1.0f / sqrtf(instnorm_sum8(q, buf, w, l) / (float)G + eps);
  float sd = 0.0f, sdy = 0.0f;
  if (ok)
    for (int g = w; g < G; g += 8) {
      const float yh = (xb[(size_t)g * C] - m) * rs, d = db[(size_t)g * C];
      sd += d;
      sdy = fmaf(d, yh, sdy);
    }
  sd = instnorm_sum8(sd, buf, w, l) / (float)G;
  sdy = instnorm_sum8(sdy, buf, w, l) / (float)G;
  if (!ok) return;
  float* ob = dx + (size_t)b * G * C + c;
  for (int g = w; g < G; g += 8) {
    const float yh = (xb[(size_t)g * C] - m) * rs;
    ob[(size_t)g * C] = rs * (db[(size_t)g * C] - sd - yh * sdy);
  }
}
__global__ void __launch_bounds__(256) instnorm_fwd_kernel(const float* __restrict__ x, int G, int C, float eps, float* __restrict__ y) {
  __shared__ float buf[8][32];
  const int l = threadIdx.x & 31, w = threadIdx.x >> 5, c = blockIdx.x * 32 + l, b = blockIdx.y;
  const bool ok = c < C;
  const float* xb = x + (size_t)b * G * C + (ok ? c : 0);
  float s = 0.0f;
  if (ok)
    for (int g = w; g < G; g += 8) s += xb[(size_t)g * C];
  const float m = instnorm_sum8(s, buf, w, l) / (float)G;
  float q = 0.0f;
  if (ok)
    for (int g = w; g < G; g += 8) {
      const float d = xb[(size_t)g * C] - m;
      q = fmaf(d, d, q);
    }
  const float rs = 1.0f / sqrtf(instnorm_sum8(q, buf, w, l) / (float)G + eps);
  if (!ok) return;
  float* ob = y + (size_t)b * G * C + c;
  for (int g = w; g < G; g += 8) ob[(size_t)g * C] = (xb[(size_t)g * C] - m) * rs;
}
// adjoint of mde_tokens_to_grid_bf16 (encoder.py:69-110): grid rows g < T go back to token g + strip, zero-padded cells and (when
// stripped) the class token receive nothing.  dgrid f32 [B][G][768] -> dtok f32 [B][197][768]
__global__ void __launch_bounds__(256) grid_to_tokens_bwd_kernel(const float* __restrict__ dgrid, int B, int strip, int G, float* __restrict__ dtok) {
  const int T = MDE_VIT_SEQ - strip;
  MDE_GRID_STRIDE(i, (long long)B * MDE_VIT_SEQ * MDE_VIT_DIM) {
    const int c = (int)(i % MDE_VIT_DIM);
    const long long r = i / MDE_VIT_DIM;
    const int t = (int)(r % MDE_VIT_SEQ), b = (int)(r / MDE_VIT_SEQ);
    const int g = t - strip;
    dtok[i] = (g >= 0 && g < T && g < G) ? dgrid[((size_t)b * G + g) * MDE_VIT_DIM + c] : 0.0f;
  }
}

// ------------------------------------------------------------------ sigmoid head: disp = mul * sigmoid(z) and its backward
__global__ void __launch_bounds__(256) sigmoid_fwd_kernel(const float* __restrict__ z, int ldz, long long n, float mul, float* __restrict__ d) {
  MDE_GRID_STRIDE(i, n) d[i] = mul / (1.0f + expf(-z[(size_t)i * ldz]));
}
__global__ void __launch_bounds__(256) sigmoid_bwd_kernel(const float* __restrict__ z, int ldz, const float* __restrict__ dd, long long n, float mul, float* __restrict__ dz,
                                                          int ldo) {
  MDE_GRID_STRIDE(i, n) {
    const float s = 1.0f / (1.0f + expf(-z[(size_t)i * ldz]));
    dz[(size_t)i * ldo] = dd[i] * mul * s * (1.0f - s);
  }
}

// ------------------------------------------------------------------ BatchNorm2d(3) of encoder.py:129 in train mode: batch statistics of an NCHW image batch
// Grid (C, chunks of the plane): every thread accumulates in double over a fixed stride, the 256 partials of a block are added in
// thread order, the finalize kernel adds the chunks in chunk order.
__global__ void __launch_bounds__(256) channel_stats_partial_kernel(const float* __restrict__ x, int B, int C, long long hw, double* __restrict__ part) {
  __shared__ double s1[256], s2[256];
  const int c = blockIdx.x;
  const long long per = (hw + gridDim.y - 1) / gridDim.y, lo = blockIdx.y * per, hi = lo + per < hw ? lo + per : hw;
  double a = 0.0, q = 0.0;
  for (int b = 0; b < B; ++b) {
    const float* p = x + ((size_t)b * C + c) * hw;
    for (long long i = lo + threadIdx.x; i < hi; i += 256) {
      const double v = p[i];
      a += v;
      q += v * v;
    }
  }
  s1[threadIdx.x] = a;
  s2[threadIdx.x] = q;
  __syncthreads();
  if (threadIdx.x == 0) {
    double ta = 0.0, tq = 0.0;
    for (int k = 0; k < 256; ++k) {
      ta += s1[k];
      tq += s2[k];
    }
    part[((size_t)c * gridDim.y + blockIdx.y) * 2] = ta;
    part[((size_t)c * gridDim.y + blockIdx.y) * 2 + 1] = tq;
  }
}
__global__ void channel_stats_finalize_kernel(const double* __restrict__ part, int chunks, int B, int C, long long hw, float* __restrict__ mean, float* __restrict__ var,
                                              float* __restrict__ run_mean, float* __restrict__ run_var, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double ta = 0.0, tq = 0.0;
  for (int k = 0; k < chunks; ++k) {
    ta += part[((size_t)c * chunks + k) * 2];
    tq += part[((size_t)c * chunks + k) * 2 + 1];
  }
  const double n = (double)B * (double)hw, m = ta / n;
  const double vb = fmax(tq / n - m * m, 0.0);
  mean[c] = (float)m;
  var[c] = (float)vb;
  if (run_mean) {
    run_mean[c] = (1.0f - momentum) * run_mean[c] + momentum * (float)m;
    run_var[c] = (1.0f - momentum) * run_var[c] + momentum * (float)(n > 1.0 ? vb * n / (n - 1.0) : vb);
  }
}

// ------------------------------------------------------------------ gradient of encoder.normalize.{weight,bias} (encoder.py:129-137)
// The input stage is x224[c] = a_c raw224[c] + b_c with a_c = s gamma_c rstd_c, b_c = s (beta_c - mu_c gamma_c rstd_c) + t, where
// (s, t) is one of (1,0), (1/255,0), (2,-1), (2/255,-1) chosen by the range heuristic (a Python branch in the reference: no
// gradient through the choice) and mu / rstd are batch statistics of the INPUT (independent of the parameters).  So
//   dgamma_c = s rstd_c (S2_c - mu_c S1_c),  dbeta_c = s S1_c,  S1_c = sum dX224[c],  S2_c = sum dX224[c] raw224[c].
// dX224 arrives in patch layout (row = (image, patch), column = (c, ky, kx)): the output of the conv_proj dX GEMM.
// One block per (image, patch row, channel) writes its two partial sums; the finalize block adds them in a fixed order.
__global__ void __launch_bounds__(256) input_norm_bwd_kernel(const float* __restrict__ dpatches, const float* __restrict__ raw224, float* __restrict__ part) {
  __shared__ float r1[256], r2[256];
  const int c = blockIdx.x % 3, py = (blockIdx.x / 3) % 14, b = blockIdx.x / 42;
  float s1 = 0.0f, s2 = 0.0f;
  for (int i = threadIdx.x; i < 16 * 224; i += 256) {
    const int ky = i / 224, X = i % 224;
    const float d = dpatches[((size_t)b * 196 + py * 14 + (X >> 4)) * 768 + c * 256 + ky * 16 + (X & 15)];
    const float v = raw224[(((size_t)b * 3 + c) * 224 + py * 16 + ky) * 224 + X];
    s1 += d;
    s2 = fmaf(d, v, s2);
  }
  r1[threadIdx.x] = s1;
  r2[threadIdx.x] = s2;
  for (int s = 128; s > 0; s >>= 1) {
    __syncthreads();
    if ((int)threadIdx.x < s) {
      r1[threadIdx.x] += r1[threadIdx.x + s];
      r2[threadIdx.x] += r2[threadIdx.x + s];
    }
  }
  if (threadIdx.x == 0) {
    part[blockIdx.x * 2] = r1[0];
    part[blockIdx.x * 2 + 1] = r2[0];
  }
}

__global__ void input_norm_bwd_finalize_kernel(const float* __restrict__ part, int nrows, const float* __restrict__ affine6, const float* __restrict__ bn,
                                               float eps, float* __restrict__ dgamma_dbeta6) {
  __shared__ double S[3][2];
  __shared__ float rstd[3];
  if (threadIdx.x < 3) {
    const int c = threadIdx.x;
    double a = 0.0, q = 0.0;
    for (int r = 0; r < nrows; ++r) {
      a += part[(r * 3 + c) * 2];
      q += part[(r * 3 + c) * 2 + 1];
    }
    S[c][0] = a;
    S[c][1] = q;
    rstd[c] = 1.0f / sqrtf(bn[9 + c] + eps);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // which (s, t) did the forward pick?  the candidate that reproduces affine6 (no division: gamma may be 0)
    const float cs[4] = {1.0f, 1.0f / 255.0f, 2.0f, 2.0f / 255.0f}, ct[4] = {0.0f, 0.0f, -1.0f, -1.0f};
    int best = 0;
    float best_err = 3.0e38f;
    for (int k = 0; k < 4; ++k) {
      float err = 0.0f;
      for (int c = 0; c < 3; ++c) {
        const float a0 = bn[c] * rstd[c], b0 = bn[3 + c] - bn[6 + c] * a0;
        err += fabsf(cs[k] * a0 - affine6[c]) + fabsf(cs[k] * b0 + ct[k] - affine6[3 + c]);
      }
      if (err < best_err) {
        best_err = err;
        best = k;
      }
    }
    const double s = cs[best];
    for (int c = 0; c < 3; ++c) {
      dgamma_dbeta6[c] = (float)(s * rstd[c] * (S[c][1] - (double)bn[6 + c] * S[c][0]));
      dgamma_dbeta6[3 + c] = (float)(s * S[c][0]);
    }
  }
}

// ------------------------------------------------------------------ loss (mars_adapter.py:209-247, layers.py:202-215)
// depth term: mean |disp - target|, target = disp.detach() * (0.9 + 0.2 rnd); smoothness: mean_x(|dx disp| e^{-mean_c |dx img|}) + the
// same along y.  One pass writes per-block partial sums (fixed order) and the gradient with respect to disp.
__global__ void __launch_bounds__(256) depth_loss_kernel(const float* __restrict__ disp, const float* __restrict__ img, const float* __restrict__ rnd, int B, int H, int W,
                                                         float w_depth, float w_smooth, float* __restrict__ part, float* __restrict__ ddisp) {
  __shared__ float red[3][256];
  const long long n = (long long)B * H * W;
  const float inv_n = 1.0f / (float)n, inv_nx = 1.0f / (float)((long long)B * H * (W - 1)), inv_ny = 1.0f / (float)((long long)B * (H - 1) * W);
  float s_d = 0.0f, s_x = 0.0f, s_y = 0.0f;
  auto sgn = [](float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); };
  MDE_GRID_STRIDE(i, n) {
    const int x = (int)(i % W), y = (int)((i / W) % H), b = (int)(i / ((long long)W * H));
    const float d = disp[i];
    const float diff = d - d * (0.9f + 0.2f * rnd[i]);
    s_d += fabsf(diff);
    float g = w_depth * sgn(diff) * inv_n;
    const float* ib = img + (size_t)b * 3 * H * W + (size_t)y * W + x;
    const size_t pl = (size_t)H * W;
    auto wgt = [&](long long off) {  // exp(-mean_c |I(p) - I(p + off)|)
      const float a = fabsf(ib[0] - ib[off]) + fabsf(ib[pl] - ib[pl + off]) + fabsf(ib[2 * pl] - ib[2 * pl + off]);
      return expf(-a * (1.0f / 3.0f));
    };
    if (x < W - 1) {
      const float e = d - disp[i + 1], wq = wgt(1);
      s_x = fmaf(fabsf(e), wq, s_x);
      g += w_smooth * sgn(e) * wq * inv_nx;
    }
    if (x > 0) g -= w_smooth * sgn(disp[i - 1] - d) * wgt(-1) * inv_nx;
    if (y < H - 1) {
      const float e = d - disp[i + W], wq = wgt(W);
      s_y = fmaf(fabsf(e), wq, s_y);
      g += w_smooth * sgn(e) * wq * inv_ny;
    }
    if (y > 0) g -= w_smooth * sgn(disp[i - W] - d) * wgt(-(long long)W) * inv_ny;
    ddisp[i] = g;
  }
  red[0][threadIdx.x] = s_d;
  red[1][threadIdx.x] = s_x;
  red[2][threadIdx.x] = s_y;
  __syncthreads();
  if (threadIdx.x < 3) {
    float t = 0.0f;
    for (int k = 0; k < 256; ++k) t += red[threadIdx.x][k];
    part[(size_t)blockIdx.x * 3 + threadIdx.x] = t;
  }
}
__global__ void __launch_bounds__(256) depth_loss_finalize_kernel(const float* __restrict__ part, int nblocks, int B, int H, int W, float w_depth, float w_smooth,
                                                                  float* __restrict__ out) {
  __shared__ double r[3][256];
  double d = 0.0, sx = 0.0, sy = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += 256) {  // fixed assignment, then the 256 partials in thread order
    d += part[(size_t)i * 3];
    sx += part[(size_t)i * 3 + 1];
    sy += part[(size_t)i * 3 + 2];
  }
  r[0][threadIdx.x] = d;
  r[1][threadIdx.x] = sx;
  r[2][threadIdx.x] = sy;
  __syncthreads();
  if (threadIdx.x != 0) return;
  d = sx = sy = 0.0;
  for (int k = 0; k < 256; ++k) {
    d += r[0][k];
    sx += r[1][k];
    sy += r[2][k];
  }
  const double depth = w_depth * d / ((double)B * H * W);
  const double smooth = w_smooth * (sx / ((double)B * H * (W - 1)) + sy / ((double)B * (H - 1) * W));
  out[0] = (float)(depth + smooth);
  out[1] = (float)depth;
  out[2] = (float)smooth;
}

// ------------------------------------------------------------------ Adam (torch.optim.Adam defaults: no weight decay, no amsgrad)
// step_dev != nullptr: the step count lives on the device (a CUDA graph replays the same launch every step) and the bias
// corrections are evaluated here; otherwise they arrive as arguments.
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long long n,
                                                   float lr, float b1, float b2, float eps, float bc1, float bc2_sqrt, const int* __restrict__ step_dev, int vec4) {
  if (step_dev) {
    const float t = (float)*step_dev;
    bc1 = 1.0f - powf(b1, t);
    bc2_sqrt = sqrtf(1.0f - powf(b2, t));
  }
  auto upd = [&](float& pi, float gi, float& mi, float& vi) {
    mi = fmaf(b1, mi, (1.0f - b1) * gi);   // exp_avg.lerp_(grad, 1 - beta1)
    vi = fmaf(b2, vi, (1.0f - b2) * gi * gi);
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    pi -= (lr / bc1) * (mi / denom);
  };
  if (vec4) {  // n % 4 == 0 and 16-byte aligned pointers: 128-bit accesses
    float4* p4 = reinterpret_cast<float4*>(p);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    MDE_GRID_STRIDE(i, n / 4) {
      float4 pp = p4[i], mm = m4[i], vv = v4[i];
      const float4 gg = g4[i];
      upd(pp.x, gg.x, mm.x, vv.x);
      upd(pp.y, gg.y, mm.y, vv.y);
      upd(pp.z, gg.z, mm.z, vv.z);
      upd(pp.w, gg.w, mm.w, vv.w);
      p4[i] = pp;
      m4[i] = mm;
      v4[i] = vv;
    }
    return;
  }
  MDE_GRID_STRIDE(i, n) {
    float pi = p[i], mi = m[i], vi = v[i];
    upd(pi, g[i], mi, vi);
    p[i] = pi;
    m[i] = mi;
    v[i] = vi;
  }
}

// Multi-tensor Adam: ONE launch updates every parameter tensor of the model (91 single-tensor launches before, 355 us of launch
// and tail cost per step).  table[t] = {p, g, m, v, n, group}; chunk_start[t] = first chunk of tensor t in a global list of
// kAdamChunk-element chunks (prefix sum, chunk_start[n_tensors] = total).  The learning rate of a group is read from device
// memory, so a schedule (train/mars_adapter.py:57-61 StepLR) changes it between replays of a captured graph.
constexpr int kAdamChunk = 4096;  // elements per block-iteration: 256 threads x 4 x float4
struct AdamTensor {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  long long group;
};
__global__ void __launch_bounds__(256) adam_multi_kernel(const AdamTensor* __restrict__ table, int n_tensors, const int* __restrict__ chunk_start,
                                                         const float* __restrict__ lr_groups, float b1, float b2, float eps,
                                                         const int* __restrict__ step_dev) {
  const float t = (float)*step_dev;
  const float bc1 = 1.0f - powf(b1, t), bc2_sqrt = sqrtf(1.0f - powf(b2, t));
  const int total = chunk_start[n_tensors];
  for (int c = blockIdx.x; c < total; c += gridDim.x) {
    int lo = 0, hi = n_tensors - 1;  // last tensor whose first chunk is <= c
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (chunk_start[mid] <= c) lo = mid; else hi = mid - 1;
    }
    const AdamTensor T = table[lo];
    const float step_size = lr_groups[T.group] / bc1;
    const long long off = (long long)(c - chunk_start[lo]) * kAdamChunk;
    const long long len = min((long long)kAdamChunk, T.n - off);
    auto upd = [&](float& pi, float gi, float& mi, float& vi) {
      mi = fmaf(b1, mi, (1.0f - b1) * gi);
      vi = fmaf(b2, vi, (1.0f - b2) * gi * gi);
      pi -= step_size * (mi / (sqrtf(vi) / bc2_sqrt + eps));
    };
    const bool vec = ((uintptr_t)(T.p + off) | (uintptr_t)(T.g + off) | (uintptr_t)(T.m + off) | (uintptr_t)(T.v + off)) % 16 == 0;
    if (vec) {
      float4* p4 = reinterpret_cast<float4*>(T.p + off);
      float4* m4 = reinterpret_cast<float4*>(T.m + off);
      float4* v4 = reinterpret_cast<float4*>(T.v + off);
      const float4* g4 = reinterpret_cast<const float4*>(T.g + off);
      for (long long i = threadIdx.x; i < len / 4; i += 256) {
        float4 pp = p4[i], mm = m4[i], vv = v4[i];
        const float4 gg = g4[i];
        upd(pp.x, gg.x, mm.x, vv.x);
        upd(pp.y, gg.y, mm.y, vv.y);
        upd(pp.z, gg.z, mm.z, vv.z);
        upd(pp.w, gg.w, mm.w, vv.w);
        p4[i] = pp;
        m4[i] = mm;
        v4[i] = vv;
      }
    }
    for (long long i = (vec ? (len / 4) * 4 : 0) + threadIdx.x; i < len; i += 256) {
      float pi = T.p[off + i], mi = T.m[off + i], vi = T.v[off + i];
      upd(pi, T.g[off + i], mi, vi);
      T.p[off + i] = pi;
      T.m[off + i] = mi;
      T.v[off + i] = vi;
    }
  }
}

}  // namespace mde

using namespace mde;

extern "C" int mde_adam_chunk_elems(void) { return kAdamChunk; }

extern "C" int mde_adam_multi(const mde_adam_tensor* table_dev, int n_tensors, const int* chunk_start_dev, int n_chunks, const float* lr_groups_dev,
                              float beta1, float beta2, float eps, const int* step_dev, mde_stream_t stream) {
  static_assert(sizeof(AdamTensor) == sizeof(mde_adam_tensor), "table layout");
  MDE_CHECK_ARG(table_dev && chunk_start_dev && lr_groups_dev && step_dev && n_tensors > 0 && n_chunks > 0, "mde_adam_multi: bad arguments");
  const int cap = sm_count() * 8;
  adam_multi_kernel<<<n_chunks < cap ? n_chunks : cap, 256, 0, stream>>>(reinterpret_cast<const AdamTensor*>(table_dev), n_tensors, chunk_start_dev,
                                                                         lr_groups_dev, beta1, beta2, eps, step_dev);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_im2col3x3_bf16(const float* x, int B, int h, int w, int C, int padding, mde_bf16* col, mde_bf16* colT, int ldt, mde_stream_t stream) {
  MDE_CHECK_ARG(x && (col || colT) && B > 0 && h >= 3 && w >= 3 && C > 0, "mde_im2col3x3_bf16: bad arguments");
  MDE_CHECK_ARG(padding == MDE_PAD_REFLECT || padding == MDE_PAD_ZERO, "mde_im2col3x3_bf16: bad padding mode");
  MDE_CHECK_ARG(!colT || ldt >= B * h * w, "mde_im2col3x3_bf16: ldt too small");
  const long long total = (colT ? (long long)ldt : (long long)B * h * w) * 9 * C;
  im2col_kernel<<<grid1(total), 256, 0, stream>>>(x, B, h, w, C, padding == MDE_PAD_REFLECT, reinterpret_cast<bf16*>(col), reinterpret_cast<bf16*>(colT), ldt);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_col2im3x3_f32(const float* dxcol, int ldc, int B, int h, int w, int C, int padding, float* dx, mde_stream_t stream) {
  MDE_CHECK_ARG(dxcol && dx && B > 0 && h >= 3 && w >= 3 && C > 0 && ldc >= 9 * C, "mde_col2im3x3_f32: bad arguments");
  col2im_kernel<<<grid1((long long)B * h * w * C), 256, 0, stream>>>(dxcol, ldc, B, h, w, C, padding == MDE_PAD_REFLECT, dx);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_reduce_scratch_bytes(int C) { return (long long)kRedChunks * (C > 0 ? C : 1) * 2 * (long long)sizeof(double); }

extern "C" int mde_bn_elu_train_fwd(const float* z, int N, int C, int ldz, const float* gamma, const float* beta, float eps, float momentum, float* run_mean,
                                    float* run_var, float* mean, float* rstd, float* y, void* scratch, mde_stream_t stream) {
  MDE_CHECK_ARG(z && gamma && beta && mean && rstd && y && scratch && N > 0 && C > 0 && ldz >= C, "mde_bn_elu_train_fwd: bad arguments");
  MDE_CHECK_ARG((run_mean == nullptr) == (run_var == nullptr), "mde_bn_elu_train_fwd: running_mean and running_var go together");
  MDE_CHECK_ARG((uintptr_t)scratch % 8 == 0, "mde_bn_elu_train_fwd: scratch must be 8-byte aligned");
  const int chunks = red_chunks(N, 64);
  bn_stats_partial_kernel<<<dim3(ceil_div(C, 32), chunks), 256, 0, stream>>>(z, N, C, ldz, static_cast<double*>(scratch));
  MDE_CHECK_LAUNCH();
  bn_stats_finalize_kernel<<<ceil_div(C, 256), 256, 0, stream>>>(static_cast<const double*>(scratch), chunks, N, C, eps, mean, rstd, run_mean, run_var, momentum);
  MDE_CHECK_LAUNCH();
  bn_elu_apply_kernel<<<grid1((long long)N * C), 256, 0, stream>>>(z, N, C, ldz, mean, rstd, gamma, beta, y);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_bn_elu_train_bwd(const float* z, int N, int C, int ldz, const float* mean, const float* rstd, const float* gamma, const float* beta,
                                    const float* dy, float* dz, float* dgamma, float* dbeta, void* scratch, mde_stream_t stream) {
  MDE_CHECK_ARG(z && mean && rstd && gamma && beta && dy && dz && dgamma && dbeta && scratch && N > 0 && C > 0 && ldz >= C, "mde_bn_elu_train_bwd: bad arguments");
  MDE_CHECK_ARG((uintptr_t)scratch % 8 == 0, "mde_bn_elu_train_bwd: scratch must be 8-byte aligned");
  const int chunks = red_chunks(N, 64);
  bn_elu_bwd_reduce_kernel<<<dim3(ceil_div(C, 32), chunks), 256, 0, stream>>>(z, N, C, ldz, mean, rstd, gamma, beta, dy, static_cast<double*>(scratch));
  MDE_CHECK_LAUNCH();
  bn_elu_bwd_finalize_kernel<<<ceil_div(C, 256), 256, 0, stream>>>(static_cast<const double*>(scratch), chunks, C, dgamma, dbeta);
  MDE_CHECK_LAUNCH();
  bn_elu_bwd_apply_kernel<<<grid1((long long)N * C), 256, 0, stream>>>(z, N, C, ldz, mean, rstd, gamma, beta, dy, dgamma, dbeta, dz);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_bilinear_nhwc_f32(const float* src, int B, int hs, int ws, int C, float* dst, int hd, int wd, int ldd, int coff, int align_corners,
                                     mde_stream_t stream) {
  MDE_CHECK_ARG(src && dst && B > 0 && hs > 0 && ws > 0 && hd > 0 && wd > 0 && C > 0 && coff >= 0 && coff + C <= ldd, "mde_bilinear_nhwc_f32: bad arguments");
  bilinear_fwd_kernel<<<grid1((long long)B * hd * wd * C), 256, 0, stream>>>(src, B, hs, ws, C, dst, hd, wd, align_corners, ldd, coff);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_bilinear_nhwc_bwd_f32(const float* ddst, int B, int hd, int wd, int C, int ldd, int coff, float* dsrc, int hs, int ws, int align_corners,
                                         float* scratch, mde_stream_t stream) {
  MDE_CHECK_ARG(ddst && dsrc && B > 0 && hs > 0 && ws > 0 && hd > 0 && wd > 0 && C > 0 && coff >= 0 && coff + C <= ldd, "mde_bilinear_nhwc_bwd_f32: bad arguments");
  if (scratch) {  // separable: B*hd*ws*C floats of scratch
    bilinear_bwd_x_kernel<<<grid1((long long)B * hd * ws * C), 256, 0, stream>>>(ddst, B, hd, wd, C, ldd, coff, scratch, ws, align_corners);
    MDE_CHECK_LAUNCH();
    bilinear_bwd_y_kernel<<<grid1((long long)B * hs * ws * C), 256, 0, stream>>>(scratch, B, hd, C, dsrc, hs, ws, align_corners);
    MDE_CHECK_LAUNCH();
    return MDE_OK;
  }
  bilinear_bwd_kernel<<<grid1((long long)B * hs * ws * C), 256, 0, stream>>>(ddst, B, hd, wd, C, ldd, coff, dsrc, hs, ws, align_corners);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_instnorm_f32(const float* x, int B, int G, int C, float eps, float* y, mde_stream_t stream) {
  MDE_CHECK_ARG(x && y && B > 0 && G > 0 && C > 0, "mde_instnorm_f32: bad arguments");
  instnorm_fwd_kernel<<<dim3(ceil_div(C, 32), B), 256, 0, stream>>>(x, G, C, eps, y);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_instnorm_bwd_f32(const float* x, const float* dy, int B, int G, int C, float eps, float* dx, mde_stream_t stream) {
  MDE_CHECK_ARG(x && dy && dx && B > 0 && G > 0 && C > 0, "mde_instnorm_bwd_f32: bad arguments");
  instnorm_bwd_kernel<<<dim3(ceil_div(C, 32), B), 256, 0, stream>>>(x, dy, G, C, eps, dx);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_tokens_to_grid_bwd_f32(const float* dgrid, int B, int strip_cls, int G, float* dtokens, mde_stream_t stream) {
  MDE_CHECK_ARG(dgrid && dtokens && B > 0 && (strip_cls == 0 || strip_cls == 1) && G >= MDE_VIT_SEQ - strip_cls, "mde_tokens_to_grid_bwd_f32: bad arguments");
  grid_to_tokens_bwd_kernel<<<grid1((long long)B * MDE_VIT_SEQ * MDE_VIT_DIM), 256, 0, stream>>>(dgrid, B, strip_cls, G, dtokens);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_sigmoid_head_f32(const float* z, int ldz, long long n, float mul, float* disp, mde_stream_t stream) {
  MDE_CHECK_ARG(z && disp && n > 0 && ldz >= 1, "mde_sigmoid_head_f32: bad arguments");
  sigmoid_fwd_kernel<<<grid1(n), 256, 0, stream>>>(z, ldz, n, mul, disp);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_channel_stats_nchw(const float* x, int B, int C, long long hw, float momentum, float* run_mean, float* run_var, float* mean, float* var,
                                      void* scratch, mde_stream_t stream) {
  MDE_CHECK_ARG(x && mean && var && scratch && B > 0 && C > 0 && hw > 0, "mde_channel_stats_nchw: bad arguments");
  MDE_CHECK_ARG((run_mean == nullptr) == (run_var == nullptr), "mde_channel_stats_nchw: running_mean and running_var go together");
  MDE_CHECK_ARG((uintptr_t)scratch % 8 == 0, "mde_channel_stats_nchw: scratch must be 8-byte aligned");
  const int chunks = red_chunks(hw, 4096);
  channel_stats_partial_kernel<<<dim3(C, chunks), 256, 0, stream>>>(x, B, C, hw, static_cast<double*>(scratch));
  MDE_CHECK_LAUNCH();
  channel_stats_finalize_kernel<<<ceil_div(C, 64), 64, 0, stream>>>(static_cast<const double*>(scratch), chunks, B, C, hw, mean, var, run_mean, run_var, momentum);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_input_norm_bwd_scratch_floats(int B) { return (long long)B * 14 * 3 * 2; }

extern "C" int mde_input_norm_bwd(const float* dpatches, const float* raw224, int B, const float* affine6, const float* bn4x3, float eps, float* scratch,
                                  float* dgamma_dbeta6, mde_stream_t stream) {
  MDE_CHECK_ARG(dpatches && raw224 && affine6 && bn4x3 && scratch && dgamma_dbeta6 && B > 0, "mde_input_norm_bwd: bad arguments");
  input_norm_bwd_kernel<<<B * 42, 256, 0, stream>>>(dpatches, raw224, scratch);
  MDE_CHECK_LAUNCH();
  input_norm_bwd_finalize_kernel<<<1, 32, 0, stream>>>(scratch, B * 14, affine6, bn4x3, eps, dgamma_dbeta6);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_sigmoid_head_bwd_f32(const float* z, int ldz, const float* ddisp, long long n, float mul, float* dz, int ldo, mde_stream_t stream) {
  MDE_CHECK_ARG(z && ddisp && dz && n > 0 && ldz >= 1 && ldo >= 1, "mde_sigmoid_head_bwd_f32: bad arguments");
  sigmoid_bwd_kernel<<<grid1(n), 256, 0, stream>>>(z, ldz, ddisp, n, mul, dz, ldo);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_depth_loss_scratch_floats(void) { return 3LL * sm_count() * 32; }

extern "C" int mde_depth_loss(const float* disp, const float* img, const float* rnd, int B, int H, int W, float w_depth, float w_smooth, float* scratch,
                              float* losses3, float* ddisp, mde_stream_t stream) {
  MDE_CHECK_ARG(disp && img && rnd && scratch && losses3 && ddisp && B > 0 && H > 1 && W > 1, "mde_depth_loss: bad arguments");
  const int nb = grid1((long long)B * H * W);
  depth_loss_kernel<<<nb, 256, 0, stream>>>(disp, img, rnd, B, H, W, w_depth, w_smooth, scratch, ddisp);
  MDE_CHECK_LAUNCH();
  depth_loss_finalize_kernel<<<1, 256, 0, stream>>>(scratch, nb, B, H, W, w_depth, w_smooth, losses3);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

static inline int adam_vec4(const float* p, const float* g, const float* m, const float* v, long long n) {
  return n % 4 == 0 && ((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0;
}

extern "C" int mde_adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2, float eps, int step,
                             mde_stream_t stream) {
  MDE_CHECK_ARG(p && g && m && v && n > 0 && step >= 1, "mde_adam_step: bad arguments");
  const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
  const int vec4 = adam_vec4(p, g, m, v, n);
  adam_kernel<<<grid1(vec4 ? n / 4 : n), 256, 0, stream>>>(p, g, m, v, n, lr, beta1, beta2, eps, bc1, sqrtf(bc2), nullptr, vec4);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_adam_step_dev(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2, float eps, const int* step_dev,
                                 mde_stream_t stream) {
  MDE_CHECK_ARG(p && g && m && v && n > 0 && step_dev, "mde_adam_step_dev: bad arguments");
  const int vec4 = adam_vec4(p, g, m, v, n);
  adam_kernel<<<grid1(vec4 ? n / 4 : n), 256, 0, stream>>>(p, g, m, v, n, lr, beta1, beta2, eps, 1.0f, 1.0f, step_dev, vec4);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
