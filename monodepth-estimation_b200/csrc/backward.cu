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
constexpr int kLnRowsPerBlock = 16;  // 8 warps x 2 rows: ~100 blocks for the 1576 token rows of a batch of 8
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
  for (int rr = 0; rr < kLnRowsPerBlock / 8; ++rr) {
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
// one block per 32 columns: warp w adds the partials w, w + 8, ... of its lane's column, the eight sums are added in warp order
__global__ void __launch_bounds__(256) ln_bwd_finalize_kernel(const float* __restrict__ part, int nblocks, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                              int accumulate) {
  __shared__ float sg[8][32], sb[8][32];
  const int l = threadIdx.x & 31, w = threadIdx.x >> 5, c = blockIdx.x * 32 + l;
  float g = 0.0f, b = 0.0f;
  for (int i = w; i < nblocks; i += 8) {
    g += part[((size_t)i * 2) * kLnDim + c];
    b += part[((size_t)i * 2 + 1) * kLnDim + c];
  }
  sg[w][l] = g;
  sb[w][l] = b;
  __syncthreads();
  if (w != 0) return;
  g = b = 0.0f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    g += sg[k][l];
    b += sb[k][l];
  }
  dgamma[c] = accumulate ? dgamma[c] + g : g;
  dbeta[c] = accumulate ? dbeta[c] + b : b;
}

// ------------------------------------------------------------------ ViT attention backward (tv:113, 12 heads x 64, S = 197)
// q (pre-scaled by 1/8, as the QKV epilogue stores it), k: token-major [B*197][768]; vt: [B*12][64][208]; dout: gradient of the
// attention output, token-major [B*197][768].  Writes dqkv [B*197][2304] = (dq | dk | dv) per token, dq with respect to the
// UNscaled q (what in_proj produced), so it is the A operand of the in_proj gradient GEMMs as is.
//   P = softmax(q_s k^T);  dV = P^T dO;  dP = dO V^T;  dS = P o (dP - rowsum(P o dP));  dq = dS k / 8;  dk = dS^T q_s
// Warp-level tensor-core MMAs (mma.sync m16n8k16, bf16 operands, fp32 accumulation; 197 tokens do not fill a tcgen05 tile and
// the products chain through registers).  Two launches over a grid (image x head, 4 row tiles of 16):
//   pass 1, a warp owns 16 query rows (K and V^T of the slab in shared memory): S, the softmax statistics (max, 1/sum) and
//           D = rowsum(P o dP) -> `stats`; dS as the A operand straight out of the accumulators; dq = dS K.
//   pass 2, a warp owns 16 key rows (Q and dO in shared memory): S^T and dP^T per 16-query step, P^T and dS^T rebuilt from the
//           statistics, dv += P^T dO, dk += dS^T Q.
// Shared rows are 72 bf16 (36 words) apart: the 32-bit B-fragment reads (bank 4g + t) and the ldmatrix rows are conflict-free.
constexpr int kAbS = MDE_VIT_SEQ;
constexpr int kAbSp = 208;                 // 13 tiles of 16
constexpr int kAbTiles = kAbSp / 16;
constexpr int kAbPw = 36;                  // row pitch in 32-bit words
constexpr int kAbVw = 108;                 // pitch of a V^T row (208 keys + 8 pad) in words: ldmatrix rows 12 banks apart
constexpr int kAbSmem = 2 * kAbSp * kAbPw * 4 + 3 * kAbSp * 4;  // 62.4 KB
static_assert(MDE_VT_PITCH == kAbSp, "the V^T slab rows are copied 16 bytes at a time");

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// four transposed 8x8 bf16 matrices: with rows = k and columns = n in shared memory this yields the B fragments (b0, b1) of two
// adjacent n-tiles (lane address: row k0 + (lane & 15), column n0 + 8 (lane >> 4))
__device__ __forceinline__ void ldsm_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(a));
}
// A fragments (4 k-steps of 16) of a 16 x 64 row-major bf16 tile in global memory; rows >= kAbS read as zero
__device__ __forceinline__ void load_a_tile(const bf16* __restrict__ base, int row_lo, int t, uint32_t (&a)[4][4]) {
  const bool vlo = row_lo < kAbS, vhi = row_lo + 8 < kAbS;
  const uint32_t* plo = reinterpret_cast<const uint32_t*>(base + (size_t)row_lo * MDE_VIT_DIM);
  const uint32_t* phi = reinterpret_cast<const uint32_t*>(base + (size_t)(row_lo + 8) * MDE_VIT_DIM);
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    a[kk][0] = vlo ? __ldg(plo + kk * 8 + t) : 0u;
    a[kk][1] = vhi ? __ldg(phi + kk * 8 + t) : 0u;
    a[kk][2] = vlo ? __ldg(plo + kk * 8 + 4 + t) : 0u;
    a[kk][3] = vhi ? __ldg(phi + kk * 8 + 4 + t) : 0u;
  }
}

template <int PASS>
__global__ void __launch_bounds__(128) vit_attn_bwd_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k, const bf16* __restrict__ vt,
                                                           const bf16* __restrict__ dout, float* __restrict__ stats, bf16* __restrict__ dqkv) {
  extern __shared__ __align__(16) uint32_t sm[];
  uint32_t* sA = sm;                     // pass 1: K [key][d]                  pass 2: Q [query][d]
  uint32_t* sB = sA + kAbSp * kAbPw;     // pass 1: V^T [d][key], pitch kAbVw   pass 2: dO [query][d]
  float* sMax = reinterpret_cast<float*>(sB + kAbSp * kAbPw);
  float* sInv = sMax + kAbSp;
  float* sD = sInv + kAbSp;
  const int slab = blockIdx.x, b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const size_t tok0 = (size_t)b * kAbS;
  const bf16* vb = vt + (size_t)slab * 64 * MDE_VT_PITCH;
  float* st = stats + (size_t)slab * 3 * kAbS;
  const bf16* qh = q + tok0 * MDE_VIT_DIM + hd * 64;
  const bf16* kh = k + tok0 * MDE_VIT_DIM + hd * 64;
  const bf16* gh = dout + tok0 * MDE_VIT_DIM + hd * 64;
  {
    const bf16* a = PASS == 1 ? kh : qh;
    for (int i = tid; i < kAbSp * 8; i += 128) {  // 16-byte pieces; rows 197..207 are zero
      const int r = i >> 3, c = i & 7;
      uint4 va = make_uint4(0u, 0u, 0u, 0u), vg = va;
      if (r < kAbS) {
        va = __ldg(reinterpret_cast<const uint4*>(a + (size_t)r * MDE_VIT_DIM) + c);
        if (PASS == 2) vg = __ldg(reinterpret_cast<const uint4*>(gh + (size_t)r * MDE_VIT_DIM) + c);
      }
      *reinterpret_cast<uint4*>(sA + r * kAbPw + c * 4) = va;
      if (PASS == 2) *reinterpret_cast<uint4*>(sB + r * kAbPw + c * 4) = vg;
    }
    if (PASS == 1) {  // V^T [64][208] as stored (rows = d: ldmatrix.trans turns them into B fragments); keys 197..207 zeroed
      for (int i = tid; i < 64 * (kAbSp / 8); i += 128) {
        const int d = i / (kAbSp / 8), c = i - d * (kAbSp / 8);
        *reinterpret_cast<uint4*>(sB + d * kAbVw + c * 4) = __ldg(reinterpret_cast<const uint4*>(vb + (size_t)d * MDE_VT_PITCH) + c);
      }
      __syncthreads();
      bf16* sVh = reinterpret_cast<bf16*>(sB);
      for (int i = tid; i < 64 * (kAbSp - kAbS); i += 128) {
        const int d = i / (kAbSp - kAbS), c = i - d * (kAbSp - kAbS);
        sVh[d * (2 * kAbVw) + kAbS + c] = __float2bfloat16_rn(0.0f);
      }
    } else {
      for (int i = tid; i < kAbSp; i += 128) {  // padded queries: 1/sum = 0 switches them off
        const bool ok = i < kAbS;
        sMax[i] = ok ? st[i] : 0.0f;
        sInv[i] = ok ? st[kAbS + i] : 0.0f;
        sD[i] = ok ? st[2 * kAbS + i] : 0.0f;
      }
    }
  }
  __syncthreads();
  const int mt = blockIdx.y * 4 + warp;
  if (mt >= kAbTiles) return;
  const int r_lo = mt * 16 + g, r_hi = r_lo + 8;

  if (PASS == 1) {  // ------------------------------------------------ 16 query rows
    uint32_t ga[4][4];
    float P[kAbSp / 8][4];
    {
      uint32_t qa[4][4];
      load_a_tile(qh, r_lo, t, qa);
#pragma unroll
      for (int j = 0; j < kAbSp / 8; ++j) {
        P[j][0] = P[j][1] = P[j][2] = P[j][3] = 0.0f;
        const uint32_t* kr = sA + (j * 8 + g) * kAbPw + t;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) mma_bf16_16816(P[j], qa[kk], kr[kk * 8], kr[kk * 8 + 4]);
      }
    }
    load_a_tile(gh, r_lo, t, ga);
    float mlo = -INFINITY, mhi = -INFINITY;
#pragma unroll
    for (int j = 0; j < kAbSp / 8; ++j) {
      const int key = j * 8 + 2 * t;
      if (key >= kAbS) P[j][0] = P[j][2] = -INFINITY;
      if (key + 1 >= kAbS) P[j][1] = P[j][3] = -INFINITY;
      mlo = fmaxf(mlo, fmaxf(P[j][0], P[j][1]));
      mhi = fmaxf(mhi, fmaxf(P[j][2], P[j][3]));
    }
    mlo = fmaxf(mlo, __shfl_xor_sync(0xffffffffu, mlo, 1));
    mlo = fmaxf(mlo, __shfl_xor_sync(0xffffffffu, mlo, 2));
    mhi = fmaxf(mhi, __shfl_xor_sync(0xffffffffu, mhi, 1));
    mhi = fmaxf(mhi, __shfl_xor_sync(0xffffffffu, mhi, 2));
    float slo = 0.0f, shi = 0.0f;
#pragma unroll
    for (int j = 0; j < kAbSp / 8; ++j) {
      P[j][0] = __expf(P[j][0] - mlo);
      P[j][1] = __expf(P[j][1] - mlo);
      P[j][2] = __expf(P[j][2] - mhi);
      P[j][3] = __expf(P[j][3] - mhi);
      slo += P[j][0] + P[j][1];
      shi += P[j][2] + P[j][3];
    }
    slo += __shfl_xor_sync(0xffffffffu, slo, 1);
    slo += __shfl_xor_sync(0xffffffffu, slo, 2);
    shi += __shfl_xor_sync(0xffffffffu, shi, 1);
    shi += __shfl_xor_sync(0xffffffffu, shi, 2);
    const float ilo = 1.0f / slo, ihi = 1.0f / shi;
    float dlo = 0.0f, dhi = 0.0f;  // D = rowsum(P o dP)
    const uint32_t* vrow = sB + (lane & 15) * kAbVw + (lane >> 4) * 4;  // + kk * 16 rows, + kp * 8 words
#pragma unroll
    for (int kp = 0; kp < kAbTiles; ++kp) {
      float dp[2][4] = {{0.0f, 0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f, 0.0f}};
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_trans(b0, b1, b2, b3, vrow + kk * 16 * kAbVw + kp * 8);
        mma_bf16_16816(dp[0], ga[kk], b0, b1);
        mma_bf16_16816(dp[1], ga[kk], b2, b3);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = 2 * kp + h;
        P[j][0] *= ilo;
        P[j][1] *= ilo;
        P[j][2] *= ihi;
        P[j][3] *= ihi;
        dlo = fmaf(P[j][0], dp[h][0], fmaf(P[j][1], dp[h][1], dlo));
        dhi = fmaf(P[j][2], dp[h][2], fmaf(P[j][3], dp[h][3], dhi));
      }
    }
    dlo += __shfl_xor_sync(0xffffffffu, dlo, 1);
    dlo += __shfl_xor_sync(0xffffffffu, dlo, 2);
    dhi += __shfl_xor_sync(0xffffffffu, dhi, 1);
    dhi += __shfl_xor_sync(0xffffffffu, dhi, 2);
    if (t == 0) {
      if (r_lo < kAbS) {
        st[r_lo] = mlo;
        st[kAbS + r_lo] = ilo;
        st[2 * kAbS + r_lo] = dlo;
      }
      if (r_hi < kAbS) {
        st[r_hi] = mhi;
        st[kAbS + r_hi] = ihi;
        st[2 * kAbS + r_hi] = dhi;
      }
    }
    float dq[8][4];
#pragma unroll
    for (int n = 0; n < 8; ++n) dq[n][0] = dq[n][1] = dq[n][2] = dq[n][3] = 0.0f;
#pragma unroll
    for (int kp = 0; kp < kAbTiles; ++kp) {  // 16 keys per step: dP again (cheaper than keeping it), dS, dq += dS K
      uint32_t a[4];
      float dp[2][4] = {{0.0f, 0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f, 0.0f}};
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_trans(b0, b1, b2, b3, vrow + kk * 16 * kAbVw + kp * 8);
        mma_bf16_16816(dp[0], ga[kk], b0, b1);
        mma_bf16_16816(dp[1], ga[kk], b2, b3);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = 2 * kp + h;
        a[2 * h] = pack_bf16x2(P[j][0] * (dp[h][0] - dlo), P[j][1] * (dp[h][1] - dlo));
        a[2 * h + 1] = pack_bf16x2(P[j][2] * (dp[h][2] - dhi), P[j][3] * (dp[h][3] - dhi));
      }
      const uint32_t* krow = sA + (kp * 16 + (lane & 15)) * kAbPw + (lane >> 4) * 4;
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_trans(b0, b1, b2, b3, krow + np * 8);
        mma_bf16_16816(dq[2 * np], a, b0, b1);
        mma_bf16_16816(dq[2 * np + 1], a, b2, b3);
      }
    }
    bf16* olo = dqkv + (tok0 + r_lo) * (3 * MDE_VIT_DIM) + hd * 64 + 2 * t;
    bf16* ohi = dqkv + (tok0 + r_hi) * (3 * MDE_VIT_DIM) + hd * 64 + 2 * t;
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      if (r_lo < kAbS) *reinterpret_cast<uint32_t*>(olo + n * 8) = pack_bf16x2(dq[n][0] * 0.125f, dq[n][1] * 0.125f);
      if (r_hi < kAbS) *reinterpret_cast<uint32_t*>(ohi + n * 8) = pack_bf16x2(dq[n][2] * 0.125f, dq[n][3] * 0.125f);
    }
  } else {  // -------------------------------------------------------- 16 key rows
    uint32_t ka[4][4], va[4][4];
    load_a_tile(kh, r_lo, t, ka);
    {
      const bool vlo = r_lo < kAbS, vhi = r_hi < kAbS;
      const unsigned short* vs = reinterpret_cast<const unsigned short*>(vb);
      auto ld2 = [&](int d, int key, bool ok) -> uint32_t {
        return ok ? ((uint32_t)vs[(size_t)d * MDE_VT_PITCH + key] | ((uint32_t)vs[(size_t)(d + 1) * MDE_VT_PITCH + key] << 16)) : 0u;
      };
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        va[kk][0] = ld2(kk * 16 + 2 * t, r_lo, vlo);
        va[kk][1] = ld2(kk * 16 + 2 * t, r_hi, vhi);
        va[kk][2] = ld2(kk * 16 + 8 + 2 * t, r_lo, vlo);
        va[kk][3] = ld2(kk * 16 + 8 + 2 * t, r_hi, vhi);
      }
    }
    float dk[8][4], dv[8][4];
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      dk[n][0] = dk[n][1] = dk[n][2] = dk[n][3] = 0.0f;
      dv[n][0] = dv[n][1] = dv[n][2] = dv[n][3] = 0.0f;
    }
#pragma unroll 1
    for (int qp = 0; qp < kAbTiles; ++qp) {  // 16 queries per step
      uint32_t pa[4], da[4];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = 2 * qp + h;
        float sT[4] = {0.0f, 0.0f, 0.0f, 0.0f}, dp[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        const uint32_t* qr = sA + (j * 8 + g) * kAbPw + t;
        const uint32_t* gr = sB + (j * 8 + g) * kAbPw + t;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          mma_bf16_16816(sT, ka[kk], qr[kk * 8], qr[kk * 8 + 4]);
          mma_bf16_16816(dp, va[kk], gr[kk * 8], gr[kk * 8 + 4]);
        }
        const int q0 = j * 8 + 2 * t;
        const float m0 = sMax[q0], m1 = sMax[q0 + 1], i0 = sInv[q0], i1 = sInv[q0 + 1], d0 = sD[q0], d1 = sD[q0 + 1];
        const float p0 = __expf(sT[0] - m0) * i0, p1 = __expf(sT[1] - m1) * i1, p2 = __expf(sT[2] - m0) * i0, p3 = __expf(sT[3] - m1) * i1;
        pa[2 * h] = pack_bf16x2(p0, p1);
        pa[2 * h + 1] = pack_bf16x2(p2, p3);
        da[2 * h] = pack_bf16x2(p0 * (dp[0] - d0), p1 * (dp[1] - d1));
        da[2 * h + 1] = pack_bf16x2(p2 * (dp[2] - d0), p3 * (dp[3] - d1));
      }
      const int roff = (qp * 16 + (lane & 15)) * kAbPw + (lane >> 4) * 4;
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_trans(b0, b1, b2, b3, sB + roff + np * 8);  // dO [k = query][n = d]
        mma_bf16_16816(dv[2 * np], pa, b0, b1);
        mma_bf16_16816(dv[2 * np + 1], pa, b2, b3);
        ldsm_x4_trans(b0, b1, b2, b3, sA + roff + np * 8);  // Q
        mma_bf16_16816(dk[2 * np], da, b0, b1);
        mma_bf16_16816(dk[2 * np + 1], da, b2, b3);
      }
    }
    bf16* olo = dqkv + (tok0 + r_lo) * (3 * MDE_VIT_DIM) + MDE_VIT_DIM + hd * 64 + 2 * t;
    bf16* ohi = dqkv + (tok0 + r_hi) * (3 * MDE_VIT_DIM) + MDE_VIT_DIM + hd * 64 + 2 * t;
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      if (r_lo < kAbS) {
        *reinterpret_cast<uint32_t*>(olo + n * 8) = pack_bf16x2(dk[n][0], dk[n][1]);
        *reinterpret_cast<uint32_t*>(olo + MDE_VIT_DIM + n * 8) = pack_bf16x2(dv[n][0], dv[n][1]);
      }
      if (r_hi < kAbS) {
        *reinterpret_cast<uint32_t*>(ohi + n * 8) = pack_bf16x2(dk[n][2], dk[n][3]);
        *reinterpret_cast<uint32_t*>(ohi + MDE_VIT_DIM + n * 8) = pack_bf16x2(dv[n][2], dv[n][3]);
      }
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
  ln_bwd_finalize_kernel<<<kLnDim / 32, 256, 0, stream>>>(scratch, nb, dgamma, dbeta, accumulate);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_vit_attention_bwd_scratch_floats(int B) { return (long long)(B > 0 ? B : 0) * MDE_VIT_HEADS * 3 * kAbS; }

extern "C" int mde_vit_attention_bwd(const mde_bf16* q, const mde_bf16* k, const mde_bf16* vt, const mde_bf16* dout, int B, mde_bf16* dqkv, float* scratch,
                                     mde_stream_t stream) {
  MDE_CHECK_ARG(q && k && vt && dout && dqkv && scratch && B > 0, "mde_vit_attention_bwd: bad arguments");
  MDE_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)vt | (uintptr_t)dout) % 16 == 0 && (uintptr_t)dqkv % 4 == 0,
                "mde_vit_attention_bwd: q, k, vt, dout must be 16-byte aligned");
  static bool attr = false;
  if (!attr) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    attr = true;
  }
  const dim3 grid(B * MDE_VIT_HEADS, (kAbTiles + 3) / 4);
  const bf16 *qq = reinterpret_cast<const bf16*>(q), *kk = reinterpret_cast<const bf16*>(k), *vv = reinterpret_cast<const bf16*>(vt),
             *gg = reinterpret_cast<const bf16*>(dout);
  vit_attn_bwd_kernel<1><<<grid, 128, kAbSmem, stream>>>(qq, kk, vv, gg, scratch, reinterpret_cast<bf16*>(dqkv));
  MDE_CHECK_LAUNCH();
  vit_attn_bwd_kernel<2><<<grid, 128, kAbSmem, stream>>>(qq, kk, vv, gg, scratch, reinterpret_cast<bf16*>(dqkv));
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
