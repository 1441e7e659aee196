// Backward kernels of the ViT encoder block (the training step of train/mars_adapter.py:131-140 runs
// loss.backward() through torchvision's EncoderBlock, tv:110-120).  The GEMM-shaped gradients reuse mde_gemm_bf16:
//     dX = dY W          -> A = dY [M][N],        W operand = W^T [K][N]   (mde_transpose_bf16 of the weight, once per step)
//     dW = dY^T X        -> A = dY^T [N][Mp],     W operand = X^T [K][Mp]  (mde_transpose_bf16 of both, Mp = M rounded up to 64, zero filled)
// and this file holds what is not a GEMM: the transposing casts, bias gradients (column sums), GELU, LayerNorm backward and
// the attention backward.  First version: correctness and determinism (no atomics anywhere: every reduction has a fixed
// order) before speed — the attention backward runs on CUDA cores in fp32 from shared memory.
#include <math.h>

#include "common.h"
#include "ptx.cuh"

namespace mde {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------ transpose (+ cast, + zero-filled K padding)
// src [R][lds] (float or bf16) -> dst_t bf16 [C][ldt] with dst_t[c][r] = src[r][c] for r < R and 0 for R <= r < cols_t (<= ldt);
// optionally also dst_rm bf16 [R][ldr] = cast(src).  32x32 tiles through shared memory, both sides coalesced.
template <class TIn>
__global__ void __launch_bounds__(256) transpose_pad_kernel(const TIn* __restrict__ src, int R, int C, int lds, bf16* __restrict__ dst_t, int ldt, int cols_t,
                                                            bf16* __restrict__ dst_rm, int ldr) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = r0 + ty + 8 * k, c = c0 + tx;
    float v = 0.0f;
    if (r < R && c < C) {
      if constexpr (sizeof(TIn) == 4) v = src[(size_t)r * lds + c];
      else v = __bfloat162float(src[(size_t)r * lds + c]);
      if (dst_rm) dst_rm[(size_t)r * ldr + c] = __float2bfloat16_rn(v);
    }
    tile[ty + 8 * k][tx] = v;
  }
  __syncthreads();
  if (dst_t) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = c0 + ty + 8 * k, r = r0 + tx;
      if (c < C && r < cols_t) dst_t[(size_t)c * ldt + r] = __float2bfloat16_rn(tile[tx][ty + 8 * k]);
    }
  }
}

// ------------------------------------------------------------------ bias gradient: out[c] = sum_r src[r][c]
// Grid (C/32, row chunks); warp w of a block adds rows w, w+8, ... of its chunk (fixed order), the eight partials are added in
// warp order, the finalize kernel adds the chunks in chunk order.
__global__ void __launch_bounds__(256) colsum_kernel(const bf16* __restrict__ src, int R, int C, int lds, float* __restrict__ part) {
  __shared__ float sp[8][32];
  const int c = blockIdx.x * 32 + (threadIdx.x & 31), w = threadIdx.x >> 5;
  const int per = (R + gridDim.y - 1) / gridDim.y, r0 = blockIdx.y * per, r1 = min(R, r0 + per);
  float s = 0.0f;
  if (c < C)
    for (int r = r0 + w; r < r1; r += 8) s += __bfloat162float(src[(size_t)r * lds + c]);
  sp[w][threadIdx.x & 31] = s;
  __syncthreads();
  if (w == 0 && c < C) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sp[k][threadIdx.x & 31];
    part[(size_t)blockIdx.y * C + c] = t;
  }
}
__global__ void __launch_bounds__(256) colsum_finalize_kernel(const float* __restrict__ part, int chunks, int C, float* __restrict__ out, int accumulate) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  float t = 0.0f;
  for (int k = 0; k < chunks; ++k) t += part[(size_t)k * C + c];
  out[c] = accumulate ? out[c] + t : t;
}

// ------------------------------------------------------------------ exact-erf GELU (tv:118 nn.GELU) forward / backward, bf16
__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
  return fmaf(x, pdf, cdf);
}
__global__ void __launch_bounds__(256) gelu_fwd_kernel(const bf16* __restrict__ u, long long n8, bf16* __restrict__ y) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 a = reinterpret_cast<const uint4*>(u)[i];
    const uint32_t in[4] = {a.x, a.y, a.z, a.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&in[k]);
      o[k] = pack_bf16x2(gelu_exact(__low2float(h)), gelu_exact(__high2float(h)));
    }
    reinterpret_cast<uint4*>(y)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
__global__ void __launch_bounds__(256) gelu_bwd_kernel(const bf16* __restrict__ u, const bf16* __restrict__ dy, long long n8, bf16* __restrict__ du) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 a = reinterpret_cast<const uint4*>(u)[i], g = reinterpret_cast<const uint4*>(dy)[i];
    const uint32_t in[4] = {a.x, a.y, a.z, a.w}, gi[4] = {g.x, g.y, g.z, g.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&in[k]);
      const __nv_bfloat162 d = *reinterpret_cast<const __nv_bfloat162*>(&gi[k]);
      o[k] = pack_bf16x2(__low2float(d) * gelu_grad(__low2float(h)), __high2float(d) * gelu_grad(__high2float(h)));
    }
    reinterpret_cast<uint4*>(du)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ------------------------------------------------------------------ LayerNorm backward (tv:112,117), DIM = 768
// y = xhat * gamma + beta, xhat = (x - mean) * rstd.  With g = dy * gamma:
//   dx = rstd * (g - mean(g) - xhat * mean(g * xhat))  [+ dres: the gradient arriving through the residual connection]
//   dgamma[c] = sum_rows dy * xhat,  dbeta[c] = sum_rows dy
// One warp per row; a block keeps per-column partial sums of its rows in registers (thread = 24 fixed columns x its warp's
// rows), reduces the eight warps through shared memory in warp order and writes ONE partial row per block; ln_bwd_finalize
// adds the block partials in block order.  Deterministic.
constexpr int kLnDim = 768;
constexpr int kLnRowsPerBlock = 64;  // 8 warps x 8 rows
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ x, const bf16* __restrict__ dy, const float* __restrict__ gamma, int rows,
                                                     float eps, const float* __restrict__ dres, float* __restrict__ dx, float* __restrict__ part) {
  __shared__ float red[8][kLnDim];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float ag[24], ab[24];  // this lane's columns: c = 32 * k + lane ... no: c = (k * 32 + lane) (stride-32 mapping, coalesced)
#pragma unroll
  for (int k = 0; k < 24; ++k) ag[k] = ab[k] = 0.0f;
  float gm[24];
#pragma unroll
  for (int k = 0; k < 24; ++k) gm[k] = gamma[k * 32 + lane];
  for (int rr = 0; rr < 8; ++rr) {
    const int row = blockIdx.x * kLnRowsPerBlock + rr * 8 + warp;
    if (row >= rows) break;
    const float* xr = x + (size_t)row * kLnDim;
    const bf16* dr = dy + (size_t)row * kLnDim;
    float v[24], d[24];
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 24; ++k) {
      v[k] = xr[k * 32 + lane];
      d[k] = __bfloat162float(dr[k * 32 + lane]);
      s += v[k];
    }
    const float mean = warp_sum(s) * (1.0f / kLnDim);
    float q = 0.0f;
#pragma unroll
    for (int k = 0; k < 24; ++k) {
      v[k] -= mean;
      q = fmaf(v[k], v[k], q);
    }
    const float rstd = 1.0f / sqrtf(warp_sum(q) * (1.0f / kLnDim) + eps);
    float sg = 0.0f, sgx = 0.0f;
#pragma unroll
    for (int k = 0; k < 24; ++k) {
      v[k] *= rstd;  // xhat
      const float g = d[k] * gm[k];
      sg += g;
      sgx = fmaf(g, v[k], sgx);
      ag[k] = fmaf(d[k], v[k], ag[k]);
      ab[k] += d[k];
    }
    const float mg = warp_sum(sg) * (1.0f / kLnDim), mgx = warp_sum(sgx) * (1.0f / kLnDim);
    float* o = dx + (size_t)row * kLnDim;
    const float* rs = dres ? dres + (size_t)row * kLnDim : nullptr;
#pragma unroll
    for (int k = 0; k < 24; ++k) {
      float r = rstd * (d[k] * gm[k] - mg - v[k] * mgx);
      if (rs) r += rs[k * 32 + lane];
      o[k * 32 + lane] = r;
    }
  }
  // block partial of dgamma, then of dbeta (two rounds through the same shared buffer)
  for (int pass = 0; pass < 2; ++pass) {
#pragma unroll
    for (int k = 0; k < 24; ++k) red[warp][k * 32 + lane] = pass == 0 ? ag[k] : ab[k];
    __syncthreads();
    for (int c = threadIdx.x; c < kLnDim; c += 256) {
      float t = 0.0f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w][c];
      part[((size_t)blockIdx.x * 2 + pass) * kLnDim + c] = t;
    }
    __syncthreads();
  }
}
__global__ void __launch_bounds__(256) ln_bwd_finalize_kernel(const float* __restrict__ part, int nblocks, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                              int accumulate) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= kLnDim) return;
  float g = 0.0f, b = 0.0f;
  for (int i = 0; i < nblocks; ++i) {
    g += part[((size_t)i * 2) * kLnDim + c];
    b += part[((size_t)i * 2 + 1) * kLnDim + c];
  }
  dgamma[c] = accumulate ? dgamma[c] + g : g;
  dbeta[c] = accumulate ? dbeta[c] + b : b;
}

// ------------------------------------------------------------------ ViT attention backward (tv:113, 12 heads x 64, S = 197)
// q (pre-scaled by 1/8, as the QKV epilogue stores it), k: token-major [B*197][768]; vt: [B*12][64][208]; dout: gradient of the
// attention output, token-major [B*197][768].  Writes dqkv [B*197][2304] = (dq | dk | dv) per token, dq with respect to the
// UNscaled q (what in_proj produced), so it is the A operand of the in_proj gradient GEMMs as is.
//   P = softmax(q_s k^T);  dV = P^T dO;  dP = dO V^T;  dS = P o (dP - rowsum(P o dP));  dq = dS k / 8;  dk = dS^T q_s
// Two launches over a grid (image x head, row chunk).  Pass 1, warp per query row (K and V of the slab in shared memory): row
// statistics (max, 1/sum, D = rowsum(P o dP)) -> `stats`, and dq.  Pass 2, warp per key row (Q and dO in shared memory): P and
// dS of that column recomputed from the statistics, dk and dv.  Shared-memory rows are 33 words apart: lane = row reads are
// conflict-free, lane = word reads are contiguous.  A row's arithmetic does not depend on the chunking.
constexpr int kAbS = MDE_VIT_SEQ;
constexpr int kAbLd = 33;  // words (bf16 pairs) per row: 32 + 1 pad
constexpr int kAbT = (kAbS + 31) / 32;  // 7 row groups per lane
constexpr int kAbSmem = 2 * kAbS * kAbLd * 4 + 3 * kAbS * 4 + 8 * 2 * 224 * 4;  // 69 KB; registers allow two blocks per SM

__device__ __forceinline__ float dot2(uint32_t a, uint32_t b, float acc) {
  const __nv_bfloat162 x = *reinterpret_cast<const __nv_bfloat162*>(&a), y = *reinterpret_cast<const __nv_bfloat162*>(&b);
  return fmaf(__high2float(x), __high2float(y), fmaf(__low2float(x), __low2float(y), acc));
}
// one 64-wide bf16 row (128 bytes, 16-byte aligned) into 32 registers of every lane (broadcast loads)
__device__ __forceinline__ void load_row32(const bf16* __restrict__ p, uint32_t (&r)[32]) {
  const uint4* v = reinterpret_cast<const uint4*>(p);
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const uint4 t = __ldg(v + w);
    r[4 * w] = t.x;
    r[4 * w + 1] = t.y;
    r[4 * w + 2] = t.z;
    r[4 * w + 3] = t.w;
  }
}

template <int PASS>
__global__ void __launch_bounds__(256, 2) vit_attn_bwd_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k, const bf16* __restrict__ vt,
                                                              const bf16* __restrict__ dout, float* __restrict__ stats, bf16* __restrict__ dqkv) {
  extern __shared__ uint32_t sm[];
  uint32_t* sA = sm;                   // pass 1: K          pass 2: Q
  uint32_t* sB = sA + kAbS * kAbLd;    // pass 1: V          pass 2: dO
  float* sMax = reinterpret_cast<float*>(sB + kAbS * kAbLd);
  float* sInv = sMax + kAbS;
  float* sD = sInv + kAbS;
  float* sRow = sD + kAbS;  // [8 warps][2][224]
  const int slab = blockIdx.x, b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const size_t tok0 = (size_t)b * kAbS;
  const int per = (kAbS + gridDim.y - 1) / gridDim.y, lo = blockIdx.y * per, hi = min(kAbS, lo + per);
  const bf16* vb = vt + (size_t)slab * 64 * MDE_VT_PITCH;
  float* st = stats + (size_t)slab * 3 * kAbS;
  {
    const bf16* a = PASS == 1 ? k : q;
    for (int i = tid; i < kAbS * 32; i += 256) {
      const int r = i >> 5, w = i & 31;
      const size_t off = (tok0 + r) * MDE_VIT_DIM + hd * 64 + 2 * w;
      sA[r * kAbLd + w] = *reinterpret_cast<const uint32_t*>(a + off);
      if (PASS == 2) sB[r * kAbLd + w] = *reinterpret_cast<const uint32_t*>(dout + off);
    }
    if (PASS == 1) {  // V^T [64][208] -> sB[token][d]
      bf16* sVh = reinterpret_cast<bf16*>(sB);
      for (int i = tid; i < 64 * kAbS; i += 256) {
        const int d = i / kAbS, t = i - d * kAbS;
        sVh[t * (2 * kAbLd) + d] = vb[(size_t)d * MDE_VT_PITCH + t];
      }
    } else {
      for (int i = tid; i < kAbS; i += 256) {
        sMax[i] = st[i];
        sInv[i] = st[kAbS + i];
        sD[i] = st[2 * kAbS + i];
      }
    }
  }
  __syncthreads();
  float* rowA = sRow + warp * 2 * 224;
  float* rowB = rowA + 224;

  if (PASS == 1) {  // ---- query rows
    const uint32_t* sK = sA;
    const uint32_t* sV = sB;
    for (int i = lo + warp; i < hi; i += 8) {
      uint32_t qi[32], gi[32];
      load_row32(q + (tok0 + i) * MDE_VIT_DIM + hd * 64, qi);
      load_row32(dout + (tok0 + i) * MDE_VIT_DIM + hd * 64, gi);
      float s[kAbT], dp[kAbT];
      float mx = -INFINITY;
#pragma unroll
      for (int t = 0; t < kAbT; ++t) {
        const int j = lane + 32 * t;
        s[t] = -INFINITY;
        dp[t] = 0.0f;
        if (j < kAbS) {
          float a = 0.0f, c = 0.0f;
#pragma unroll
          for (int w = 0; w < 32; ++w) {
            a = dot2(qi[w], sK[j * kAbLd + w], a);
            c = dot2(gi[w], sV[j * kAbLd + w], c);
          }
          s[t] = a;
          dp[t] = c;
          mx = fmaxf(mx, a);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      float sum = 0.0f;
#pragma unroll
      for (int t = 0; t < kAbT; ++t) {
        s[t] = (lane + 32 * t < kAbS) ? __expf(s[t] - mx) : 0.0f;
        sum += s[t];
      }
      const float inv = 1.0f / warp_sum(sum);
      float dd = 0.0f;
#pragma unroll
      for (int t = 0; t < kAbT; ++t) {
        s[t] *= inv;
        dd = fmaf(s[t], dp[t], dd);
      }
      dd = warp_sum(dd);
      if (lane == 0) {
        st[i] = mx;
        st[kAbS + i] = inv;
        st[2 * kAbS + i] = dd;
      }
#pragma unroll
      for (int t = 0; t < kAbT; ++t)
        if (lane + 32 * t < kAbS) rowA[lane + 32 * t] = s[t] * (dp[t] - dd);  // dS
      __syncwarp();
      float a0 = 0.0f, a1 = 0.0f;  // dq[2*lane], dq[2*lane+1]
      for (int j = 0; j < kAbS; ++j) {
        const float ds = rowA[j];
        const uint32_t kw = sK[j * kAbLd + lane];
        const __nv_bfloat162 kk = *reinterpret_cast<const __nv_bfloat162*>(&kw);
        a0 = fmaf(ds, __low2float(kk), a0);
        a1 = fmaf(ds, __high2float(kk), a1);
      }
      *reinterpret_cast<uint32_t*>(dqkv + (tok0 + i) * (3 * MDE_VIT_DIM) + hd * 64 + 2 * lane) = pack_bf16x2(a0 * 0.125f, a1 * 0.125f);
      __syncwarp();
    }
  } else {  // ---- key rows
    const uint32_t* sQ = sA;
    const uint32_t* sG = sB;
    for (int j = lo + warp; j < hi; j += 8) {
      uint32_t kj[32], vj[32];
      load_row32(k + (tok0 + j) * MDE_VIT_DIM + hd * 64, kj);
      {  // V row j out of the transposed slab: lane l fetches dims 2l, 2l+1, then 32 broadcasts
        const uint32_t mine = (uint32_t)__bfloat16_as_ushort(vb[(size_t)(2 * lane) * MDE_VT_PITCH + j]) |
                              ((uint32_t)__bfloat16_as_ushort(vb[(size_t)(2 * lane + 1) * MDE_VT_PITCH + j]) << 16);
#pragma unroll
        for (int w = 0; w < 32; ++w) vj[w] = __shfl_sync(0xffffffffu, mine, w);
      }
#pragma unroll
      for (int t = 0; t < kAbT; ++t) {
        const int i = lane + 32 * t;
        if (i < kAbS) {
          float a = 0.0f, c = 0.0f;
#pragma unroll
          for (int w = 0; w < 32; ++w) {
            a = dot2(kj[w], sQ[i * kAbLd + w], a);
            c = dot2(vj[w], sG[i * kAbLd + w], c);
          }
          const float p = __expf(a - sMax[i]) * sInv[i];
          rowA[i] = p;
          rowB[i] = p * (c - sD[i]);
        }
      }
      __syncwarp();
      float k0 = 0.0f, k1 = 0.0f, v0 = 0.0f, v1 = 0.0f;
      for (int i = 0; i < kAbS; ++i) {
        const float p = rowA[i], ds = rowB[i];
        const uint32_t qw = sQ[i * kAbLd + lane], gw = sG[i * kAbLd + lane];
        const __nv_bfloat162 qq = *reinterpret_cast<const __nv_bfloat162*>(&qw), gg = *reinterpret_cast<const __nv_bfloat162*>(&gw);
        k0 = fmaf(ds, __low2float(qq), k0);
        k1 = fmaf(ds, __high2float(qq), k1);
        v0 = fmaf(p, __low2float(gg), v0);
        v1 = fmaf(p, __high2float(gg), v1);
      }
      bf16* dst = dqkv + (tok0 + j) * (3 * MDE_VIT_DIM) + hd * 64 + 2 * lane;
      *reinterpret_cast<uint32_t*>(dst + MDE_VIT_DIM) = pack_bf16x2(k0, k1);
      *reinterpret_cast<uint32_t*>(dst + 2 * MDE_VIT_DIM) = pack_bf16x2(v0, v1);
      __syncwarp();
    }
  }
}

// out[i] (+)= sum_b src[b][i]  (pos_embedding / class_token gradients: the sum over the images of a batch), fixed order
__global__ void __launch_bounds__(256) batch_sum_kernel(const float* __restrict__ src, int B, long long n, float* __restrict__ out, int accumulate) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float t = 0.0f;
    for (int b = 0; b < B; ++b) t += src[(size_t)b * n + i];
    out[i] = accumulate ? out[i] + t : t;
  }
}
__global__ void __launch_bounds__(256) axpy_kernel(float* __restrict__ y, const float* __restrict__ x, float a, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) y[i] = fmaf(a, x[i], y[i]);
}

static inline int grid_1d(long long items, int threads) {
  long long g = (items + threads - 1) / threads;
  const long long cap = (long long)sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace mde

using namespace mde;

extern "C" int mde_transpose_bf16(const void* src, int src_is_f32, int R, int C, int lds, mde_bf16* dst_t, int ldt, int cols_t, mde_bf16* dst_rm, int ldr,
                                  mde_stream_t stream) {
  MDE_CHECK_ARG(src && R > 0 && C > 0 && lds >= C, "mde_transpose_bf16: bad source geometry");
  MDE_CHECK_ARG(dst_t || dst_rm, "mde_transpose_bf16: no destination");
  MDE_CHECK_ARG(!dst_t || (cols_t >= R && ldt >= cols_t), "mde_transpose_bf16: need R <= cols_t <= ldt (R=%d cols_t=%d ldt=%d)", R, cols_t, ldt);
  MDE_CHECK_ARG(!dst_rm || ldr >= C, "mde_transpose_bf16: ldr=%d < C=%d", ldr, C);
  const dim3 grid(ceil_div(C, 32), ceil_div(dst_t ? cols_t : R, 32));  // tiles also cover the zero-filled tail R..cols_t of the transposed rows
  if (src_is_f32)
    transpose_pad_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(src), R, C, lds, reinterpret_cast<bf16*>(dst_t), ldt, cols_t,
                                                          reinterpret_cast<bf16*>(dst_rm), ldr);
  else
    transpose_pad_kernel<bf16><<<grid, 256, 0, stream>>>(static_cast<const bf16*>(src), R, C, lds, reinterpret_cast<bf16*>(dst_t), ldt, cols_t,
                                                         reinterpret_cast<bf16*>(dst_rm), ldr);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_colsum_bf16(const mde_bf16* src, int R, int C, int lds, float* out, int accumulate, void* scratch, mde_stream_t stream) {
  MDE_CHECK_ARG(src && out && scratch && R > 0 && C > 0 && lds >= C, "mde_colsum_bf16: bad arguments");
  int chunks = ceil_div(R, 128);
  if (chunks > 64) chunks = 64;  // mde_reduce_scratch_bytes(C) holds 64 chunks
  colsum_kernel<<<dim3(ceil_div(C, 32), chunks), 256, 0, stream>>>(reinterpret_cast<const bf16*>(src), R, C, lds, static_cast<float*>(scratch));
  MDE_CHECK_LAUNCH();
  colsum_finalize_kernel<<<ceil_div(C, 256), 256, 0, stream>>>(static_cast<const float*>(scratch), chunks, C, out, accumulate);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_gelu_bf16(const mde_bf16* u, long long count, mde_bf16* y, mde_stream_t stream) {
  MDE_CHECK_ARG(u && y && count > 0 && count % 8 == 0 && ((uintptr_t)u | (uintptr_t)y) % 16 == 0, "mde_gelu_bf16: count %% 8 == 0, 16-byte aligned pointers");
  gelu_fwd_kernel<<<grid_1d(count / 8, 256), 256, 0, stream>>>(reinterpret_cast<const bf16*>(u), count / 8, reinterpret_cast<bf16*>(y));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_gelu_bwd_bf16(const mde_bf16* u, const mde_bf16* dy, long long count, mde_bf16* du, mde_stream_t stream) {
  MDE_CHECK_ARG(u && dy && du && count > 0 && count % 8 == 0 && ((uintptr_t)u | (uintptr_t)dy | (uintptr_t)du) % 16 == 0,
                "mde_gelu_bwd_bf16: count %% 8 == 0, 16-byte aligned pointers");
  gelu_bwd_kernel<<<grid_1d(count / 8, 256), 256, 0, stream>>>(reinterpret_cast<const bf16*>(u), reinterpret_cast<const bf16*>(dy), count / 8,
                                                               reinterpret_cast<bf16*>(du));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_layernorm_bwd_scratch_floats(int rows) { return (long long)ceil_div(rows, kLnRowsPerBlock) * 2 * kLnDim; }

extern "C" int mde_layernorm_bwd(const float* x, const mde_bf16* dy, const float* gamma, int rows, int dim, float eps, const float* dres, float* dx,
                                 float* dgamma, float* dbeta, int accumulate, float* scratch, mde_stream_t stream) {
  MDE_CHECK_ARG(x && dy && gamma && dx && dgamma && dbeta && scratch && rows > 0, "mde_layernorm_bwd: null pointer");
  if (dim != kLnDim) {
    set_error("mde_layernorm_bwd: dim=%d not built (768 only)", dim);
    return MDE_ERR_UNSUPPORTED;
  }
  const int nb = ceil_div(rows, kLnRowsPerBlock);
  ln_bwd_kernel<<<nb, 256, 0, stream>>>(x, reinterpret_cast<const bf16*>(dy), gamma, rows, eps, dres, dx, scratch);
  MDE_CHECK_LAUNCH();
  ln_bwd_finalize_kernel<<<ceil_div(kLnDim, 256), 256, 0, stream>>>(scratch, nb, dgamma, dbeta, accumulate);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_vit_attention_bwd_scratch_floats(int B) { return (long long)(B > 0 ? B : 0) * MDE_VIT_HEADS * 3 * kAbS; }

extern "C" int mde_vit_attention_bwd(const mde_bf16* q, const mde_bf16* k, const mde_bf16* vt, const mde_bf16* dout, int B, mde_bf16* dqkv, float* scratch,
                                     mde_stream_t stream) {
  MDE_CHECK_ARG(q && k && vt && dout && dqkv && scratch && B > 0, "mde_vit_attention_bwd: bad arguments");
  MDE_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)dout) % 16 == 0 && (uintptr_t)dqkv % 4 == 0, "mde_vit_attention_bwd: q, k, dout must be 16-byte aligned");
  static bool attr = false;
  if (!attr) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    attr = true;
  }
  // row chunks so that the grid is at most one resident wave (two blocks per SM), at least 16 rows (two per warp) each
  const int slabs = B * MDE_VIT_HEADS;
  int chunks = 2 * sm_count() / slabs;
  if (chunks > 12) chunks = 12;
  if (chunks < 1) chunks = 1;
  const dim3 grid(slabs, chunks);
  const bf16 *qq = reinterpret_cast<const bf16*>(q), *kk = reinterpret_cast<const bf16*>(k), *vv = reinterpret_cast<const bf16*>(vt),
             *gg = reinterpret_cast<const bf16*>(dout);
  vit_attn_bwd_kernel<1><<<grid, 256, kAbSmem, stream>>>(qq, kk, vv, gg, scratch, reinterpret_cast<bf16*>(dqkv));
  MDE_CHECK_LAUNCH();
  vit_attn_bwd_kernel<2><<<grid, 256, kAbSmem, stream>>>(qq, kk, vv, gg, scratch, reinterpret_cast<bf16*>(dqkv));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_batch_sum_f32(const float* src, int B, long long n, float* out, int accumulate, mde_stream_t stream) {
  MDE_CHECK_ARG(src && out && B > 0 && n > 0, "mde_batch_sum_f32: bad arguments");
  batch_sum_kernel<<<grid_1d(n, 256), 256, 0, stream>>>(src, B, n, out, accumulate);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_axpy_f32(float* y, const float* x, float a, long long n, mde_stream_t stream) {
  MDE_CHECK_ARG(y && x && n > 0, "mde_axpy_f32: bad arguments");
  axpy_kernel<<<grid_1d(n, 256), 256, 0, stream>>>(y, x, a, n);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
