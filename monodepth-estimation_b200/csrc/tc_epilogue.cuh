// Fused epilogues shared by the tensor-core kernels (tc_gemm.cu: 1-CTA GEMM + implicit-GEMM conv,
// tc_gemm2.cu: 2-CTA GEMM).  One thread owns one accumulator row (one TMEM lane) and NCOL consecutive
// fp32 columns per call; `P` is the kernel's parameter block (fields epi, act, bias, out, ldo, aux0,
// aux1, cout, h, w, out_cstride, out_coff, f16, st32).
#pragma once
#include "common.h"
#include "ptx.cuh"

namespace mde {

enum { EPI_CONV_BF16 = 16, EPI_CONV_F32 = 17 };

struct RowInfo {
  bool valid;
  int m;        // GEMM row
  int b, y, x;  // conv pixel
};

#ifdef MDE_STORE_EXPERIMENT  // tools/build_variant.sh only: what would fully coalesced stores of the same bytes cost?
template <class P>
__device__ __forceinline__ auto dbg_of(const P& p, int) -> decltype(p.dbg) { return p.dbg; }
template <class P>
__device__ __forceinline__ int dbg_of(const P&, long) { return 0; }
#endif

// conv epilogue proper (after the bias add): activation, then a bf16/fp16 padded-NHWC store with the reflected border copies
// the pixel owns, or an fp32 NHWC head.  conv_halo.cu calls it directly (no GEMM cases in its instruction stream).
template <int NCOL, class P>
__device__ __forceinline__ void epilogue_conv(const P& p, float (&v)[NCOL], int n0, const RowInfo& ri) {
  const int nv = min(NCOL, p.cout - n0);  // valid output channels in this chunk
  if (nv <= 0) return;
  if (p.epi == EPI_CONV_F32 && nv == 1) {  // the 1-channel heads (dispconv, uncertainty): one activation, one store
    float r = v[0];
    r = p.act == MDE_ACT_ELU ? elu1(r) : p.act == MDE_ACT_SIGMOID ? sigmoidf_(r) : r;
    reinterpret_cast<float*>(p.out)[(((size_t)ri.b * p.h + ri.y) * p.w + ri.x) * p.cout + n0] = r;
    return;
  }
  if (p.act == MDE_ACT_ELU) {
#pragma unroll
    for (int j = 0; j < NCOL; ++j) v[j] = elu1(v[j]);
  } else if (p.act == MDE_ACT_SIGMOID) {
#pragma unroll
    for (int j = 0; j < NCOL; ++j) v[j] = sigmoidf_(v[j]);
  }
  if (p.epi == EPI_CONV_F32) {
    float* dst = reinterpret_cast<float*>(p.out) + (((size_t)ri.b * p.h + ri.y) * p.w + ri.x) * p.cout + n0;
#pragma unroll
    for (int j = 0; j < NCOL; ++j)
      if (j < nv) dst[j] = v[j];
    return;
  }
  uint32_t pk[NCOL / 2];
#pragma unroll
  for (int j = 0; j < NCOL / 2; ++j) pk[j] = p.f16 ? pack_f16x2(v[2 * j], v[2 * j + 1]) : pack_bf16x2(v[2 * j], v[2 * j + 1]);
  const int W2 = p.w + 2;
  // interior pixel (y, x) lives at padded (y + 1, x + 1)
  bf16* px = reinterpret_cast<bf16*>(p.out) + p.out_coff + n0 + (((size_t)ri.b * (p.h + 2) + ri.y + 1) * W2 + ri.x + 1) * p.out_cstride;
  auto store_at = [&](int dpix) {  // this pixel's channels, dpix padded pixels away
    uint4* dst = reinterpret_cast<uint4*>(px + (ptrdiff_t)dpix * p.out_cstride);
#ifdef MDE_STORE_EXPERIMENT
    if (dbg_of(p, 0) & 64) {  // WRONG RESULTS, timing only: the warp's stores land on one contiguous KB per instruction
      const unsigned long long b0 = __shfl_sync(__activemask(), (unsigned long long)dst, __ffs(__activemask()) - 1);
      uint4* d2 = reinterpret_cast<uint4*>(b0) + (threadIdx.x & 31) * 2;
#pragma unroll
      for (int j = 0; j < NCOL / 16; ++j)
        st_global_v8(d2 + 64 * j, make_uint4(pk[8 * j], pk[8 * j + 1], pk[8 * j + 2], pk[8 * j + 3]),
                     make_uint4(pk[8 * j + 4], pk[8 * j + 5], pk[8 * j + 6], pk[8 * j + 7]));
      return;
    }
#endif
    if (p.st32) {  // 32-byte stores (host checked alignment and cout % 16 == 0)
#pragma unroll
      for (int j = 0; j < NCOL / 16; ++j)
        if (16 * j < nv)
          st_global_v8(dst + 2 * j, make_uint4(pk[8 * j], pk[8 * j + 1], pk[8 * j + 2], pk[8 * j + 3]),
                       make_uint4(pk[8 * j + 4], pk[8 * j + 5], pk[8 * j + 6], pk[8 * j + 7]));
    } else {
#pragma unroll
      for (int j = 0; j < NCOL / 8; ++j)
        if (8 * j < nv) dst[j] = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    }
  };
  store_at(0);
  // the reflected border copies this pixel owns: column 1 -> padded column 0, column w-2 -> padded column w+1, same for rows, and
  // the corner copies where both hold.  Straight-line code: a loop over the 3x3 candidates cost the border tiles a third more time.
  const bool xl = ri.x == 1, xr = ri.x == p.w - 2, yt = ri.y == 1, yb = ri.y == p.h - 2;
  if (xl | xr | yt | yb) {
    if (xl) store_at(-2);
    if (xr) store_at(2);
    if (yt) {
      store_at(-2 * W2);
      if (xl) store_at(-2 * W2 - 2);
      if (xr) store_at(-2 * W2 + 2);
    }
    if (yb) {
      store_at(2 * W2);
      if (xl) store_at(2 * W2 - 2);
      if (xr) store_at(2 * W2 + 2);
    }
  }
}

template <int NCOL, class P>
__device__ __forceinline__ void add_bias(const P& p, float (&v)[NCOL], int n0) {
  // same address for every lane: served as an L1 broadcast
  const float4* bp = reinterpret_cast<const float4*>(p.bias + n0);
#pragma unroll
  for (int j = 0; j < NCOL / 4; ++j) {
    const float4 b4 = __ldg(bp + j);
    v[4 * j + 0] += b4.x;
    v[4 * j + 1] += b4.y;
    v[4 * j + 2] += b4.z;
    v[4 * j + 3] += b4.w;
  }
}

template <int NCOL, class P>
__device__ __forceinline__ void epilogue_chunk(const P& p, float (&v)[NCOL], int n0, const RowInfo& ri) {
  add_bias<NCOL, P>(p, v, n0);
  if (!ri.valid) return;

  switch (p.epi) {
    case MDE_EPI_BIAS_GELU_BF16:
#pragma unroll
      for (int j = 0; j < NCOL; ++j) v[j] = gelu_fast(v[j]);
      // fallthrough
    case MDE_EPI_BIAS_BF16: {
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.out) + (size_t)ri.m * p.ldo + n0);
#pragma unroll
      for (int j = 0; j < NCOL / 8; ++j)
        dst[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                            pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
      break;
    }
    case MDE_EPI_BIAS_RESID_F32: {
      const float4* rs = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux0) + (size_t)ri.m * p.ldo + n0);
      float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (size_t)ri.m * p.ldo + n0);
#pragma unroll
      for (int j = 0; j < NCOL / 4; ++j) {
        const float4 r4 = __ldg(rs + j);
        dst[j] = make_float4(v[4 * j] + r4.x, v[4 * j + 1] + r4.y, v[4 * j + 2] + r4.z, v[4 * j + 3] + r4.w);
      }
      break;
    }
    case MDE_EPI_BIAS_F32: {
      float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (size_t)ri.m * p.ldo + n0);
#pragma unroll
      for (int j = 0; j < NCOL / 4; ++j) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      break;
    }
    case MDE_EPI_PATCH_EMBED: {
      // A row (b, patch) -> token row b*197 + 1 + patch; + pos_embedding (tv:156); CLS row written by patch 0.
      const int b = ri.m / 196, pp = ri.m - b * 196;
      const float* pos = reinterpret_cast<const float*>(p.aux0);
      const float4* ps = reinterpret_cast<const float4*>(pos + (size_t)(1 + pp) * MDE_VIT_DIM + n0);
      const size_t trow = (size_t)b * MDE_VIT_SEQ + 1 + pp;
      float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + trow * MDE_VIT_DIM + n0);
#pragma unroll
      for (int j = 0; j < NCOL / 4; ++j) {
        const float4 q4 = __ldg(ps + j);
        v[4 * j] += q4.x;
        v[4 * j + 1] += q4.y;
        v[4 * j + 2] += q4.z;
        v[4 * j + 3] += q4.w;
        dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
      if (pp == 0) {
        const float4* cls = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux1) + n0);
        const float4* p0 = reinterpret_cast<const float4*>(pos + n0);
        const size_t crow = (size_t)b * MDE_VIT_SEQ;
        float4* d0 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + crow * MDE_VIT_DIM + n0);
        float c[NCOL];
#pragma unroll
        for (int j = 0; j < NCOL / 4; ++j) {
          const float4 c4 = __ldg(cls + j), q4 = __ldg(p0 + j);
          c[4 * j] = c4.x + q4.x;
          c[4 * j + 1] = c4.y + q4.y;
          c[4 * j + 2] = c4.z + q4.z;
          c[4 * j + 3] = c4.w + q4.w;
          d0[j] = make_float4(c[4 * j], c[4 * j + 1], c[4 * j + 2], c[4 * j + 3]);
        }
      }
      break;
    }
    case MDE_EPI_QKV_SPLIT: {
      const int b = ri.m / MDE_VIT_SEQ, t = ri.m - b * MDE_VIT_SEQ;
      const int which = n0 / MDE_VIT_DIM;  // 0 q, 1 k, 2 v (uniform per chunk: 768 % 32 == 0)
      const int c = n0 - which * MDE_VIT_DIM;
      const int hd = c >> 6, d0 = c & 63;
      const size_t slab = (size_t)b * MDE_VIT_HEADS + hd;
      if (which < 2) {
        const float sc = which == 0 ? 0.125f : 1.0f;  // 1/sqrt(64) folded into q (exact in bf16)
        bf16* base = reinterpret_cast<bf16*>(which == 0 ? p.out : const_cast<void*>(p.aux0));
        uint4* dst = reinterpret_cast<uint4*>(base + (size_t)ri.m * MDE_VIT_DIM + c);  // token-major [B*197][768]
#pragma unroll
        for (int j = 0; j < NCOL / 8; ++j)
          dst[j] = make_uint4(pack_bf16x2(v[8 * j] * sc, v[8 * j + 1] * sc), pack_bf16x2(v[8 * j + 2] * sc, v[8 * j + 3] * sc),
                              pack_bf16x2(v[8 * j + 4] * sc, v[8 * j + 5] * sc), pack_bf16x2(v[8 * j + 6] * sc, v[8 * j + 7] * sc));
      } else {
        // V transposed: vt[slab][d][t]; consecutive lanes hold consecutive tokens -> coalesced 2-byte stores
        bf16* dst = reinterpret_cast<bf16*>(const_cast<void*>(p.aux1)) + (slab * 64 + d0) * MDE_VT_PITCH + t;
#pragma unroll
        for (int j = 0; j < NCOL; ++j) dst[(size_t)j * MDE_VT_PITCH] = __float2bfloat16_rn(v[j]);
      }
      break;
    }
    case EPI_CONV_BF16:
    case EPI_CONV_F32:
      epilogue_conv<NCOL, P>(p, v, n0, ri);
      break;
    default:
      break;
  }
}

}  // namespace mde
