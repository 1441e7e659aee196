// Callers on either side of the forward path, on the device.
//  * point_cloud_kernel   — TerrainReconstructor.create_point_cloud / reconstruct_terrain
//    (utils/terrain_reconstruction_fixed.py:18-108), optionally with MarsTerrainMapper.infer_depth's disp -> depth
//    (:168-171, layers.py:16-25) in front: ONE pass over the map that filters min_depth < z < max_depth, back-projects the
//    surviving pixels and writes them compacted in raster order.  The compaction is a single-pass chained scan (tile status
//    words with decoupled look-back, tiles handed out by a ticket counter so every predecessor of a tile is already
//    running); a tile stages its surviving points in shared memory and streams them out as one contiguous run.
//    HBM traffic = 4 B read per pixel + 12 B (fp32) or 24 B (fp64) written per surviving point.
//  * temporal_mean_kernel — DepthQualityEnhancement.temporal_consistency_filter (models/refinement.py:354-388).
//  * range / LUT / bilateral kernels — DepthQualityEnhancement.bilateral_depth_filter (models/refinement.py:190-243):
//    min/max normalisation, cv2.bilateralFilter's float32 algorithm (circular window, 4096-bin interpolated colour LUT,
//    reflect-101 border) or the joint form guided by a uint8 RGB image, and the rescale back, with no host round trip.
#include <math.h>

#include "common.h"
#include "ptx.cuh"

namespace mde {

// ------------------------------------------------------------------ point cloud (terrain_reconstruction_fixed.py:18-108)
constexpr int kPcThreads = 256;
constexpr int kPcPer = 8;
constexpr int kPcTile = kPcThreads * kPcPer;  // 2048 pixels per tile
constexpr unsigned long long kPcAggregate = 1ull << 62, kPcPrefix = 2ull << 62, kPcValueMask = 0xffffffffull;

struct PcParams {
  const float* src;            // [B][H*W]: depth, or disparity when from_disp
  const void* rgb;             // optional [B][H*W][3] uint8 or fp32
  void* points;                // [B][H*W][3] fp32 or fp64 (capacity); image b starts at b*H*W*3
  void* colors;                // [B][H*W][3], same element type as rgb
  int* counts;                 // [B]
  unsigned long long* state;   // [B*T] tile status words, zeroed before the launch
  unsigned int* ticket;        // zeroed before the launch
  int B, H, W, T;
  int from_disp;
  float min_disp, disp_range;  // depth = 1 / (min_disp + disp_range * disp), each step rounded to fp32 (layers.py:21-24)
  float min_depth, max_depth;
  double fx, fy, cx, cy;
  int rgb_f32;
};

template <typename T>
__global__ void __launch_bounds__(kPcThreads) point_cloud_kernel(const __grid_constant__ PcParams p) {
  pdl_launch_dependents();
  pdl_grid_wait();
  extern __shared__ __align__(16) uint8_t pc_smem[];
  __shared__ int s_tile, s_prefix, s_total;
  __shared__ int s_warp[kPcThreads / 32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) s_tile = (int)atomicAdd(p.ticket, 1u);
  __syncthreads();
  const int tile = s_tile;
  const int b = tile / p.T, t = tile - b * p.T;
  const int HW = p.H * p.W;
  const int px0 = t * kPcTile + tid * kPcPer;
  const float* src = p.src + (size_t)b * HW;

  float z[kPcPer];
  if (px0 + kPcPer <= HW && (HW & 3) == 0) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src + px0)), c = __ldg(reinterpret_cast<const float4*>(src + px0 + 4));
    z[0] = a.x, z[1] = a.y, z[2] = a.z, z[3] = a.w, z[4] = c.x, z[5] = c.y, z[6] = c.z, z[7] = c.w;
  } else {
#pragma unroll
    for (int k = 0; k < kPcPer; ++k) z[k] = px0 + k < HW ? __ldg(src + px0 + k) : 0.0f;
  }
  unsigned valid = 0;
#pragma unroll
  for (int k = 0; k < kPcPer; ++k) {
    if (p.from_disp) z[k] = __fdiv_rn(1.0f, __fadd_rn(p.min_disp, __fmul_rn(p.disp_range, z[k])));
    if (px0 + k < HW && z[k] > p.min_depth && z[k] < p.max_depth) valid |= 1u << k;
  }

  // block-wide exclusive scan of the per-thread survivor counts
  const int cnt = __popc(valid);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kPcThreads / 32; ++w) {
    const int v = s_warp[w];
    woff += w < warp ? v : 0;
    total += v;
  }
  const int excl = woff + incl - cnt;

  // chained scan over the tiles of this image: publish the aggregate, look back for the prefix, publish the inclusive prefix
  if (warp == 0) {
    volatile unsigned long long* st = p.state + (size_t)b * p.T;
    int prefix = 0;
    if (t == 0) {
      if (lane == 0) st[0] = kPcPrefix | (unsigned long long)total;
    } else {
      if (lane == 0) st[t] = kPcAggregate | (unsigned long long)total;
      int look = t - 1;
      const long long t0 = clock64();
      while (true) {
        const int idx = look - lane;
        unsigned long long w = 2ull << 62;
        if (idx >= 0) w = st[idx];  // before the first tile: an inclusive prefix of 0
        const unsigned flag = (unsigned)(w >> 62);
        if (__ballot_sync(0xffffffffu, flag != 0) != 0xffffffffu) {
          if (clock64() - t0 > 4000000000LL) {
            if (lane == 0) printf("mde: point_cloud_kernel look-back timed out (tile %d)\n", tile);
            __trap();
          }
          continue;
        }
        const unsigned pmask = __ballot_sync(0xffffffffu, flag == 2);
        const int first = pmask ? __ffs(pmask) - 1 : 32;
        int v = lane <= first ? (int)(w & kPcValueMask) : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        prefix += v;
        if (pmask) break;
        look -= 32;
      }
      if (lane == 0) st[t] = kPcPrefix | (unsigned long long)(prefix + total);
    }
    if (lane == 0) {
      s_prefix = prefix;
      s_total = total;
      if (t == p.T - 1) p.counts[b] = prefix + total;
    }
  }

  // back-project the survivors into shared memory at their rank inside the tile
  T* sp = reinterpret_cast<T*>(pc_smem);
  {
    int v = px0 / p.W, u = px0 - v * p.W, r = excl;
#pragma unroll
    for (int k = 0; k < kPcPer; ++k) {
      if (valid & (1u << k)) {
        T x, y;
        if (sizeof(T) == 8) {  // float32 pixel grid, float64 intrinsics: NumPy promotes (u - cx) * z / fx to float64
          x = (T)(((double)(float)u - p.cx) * (double)z[k] / p.fx);
          y = (T)(((double)(float)v - p.cy) * (double)z[k] / p.fy);
        } else {               // float32 intrinsics: every step rounds to float32
          x = (T)__fdiv_rn(__fmul_rn(__fsub_rn((float)u, (float)p.cx), z[k]), (float)p.fx);
          y = (T)__fdiv_rn(__fmul_rn(__fsub_rn((float)v, (float)p.cy), z[k]), (float)p.fy);
        }
        sp[3 * r] = x, sp[3 * r + 1] = y, sp[3 * r + 2] = (T)z[k];
        ++r;
      }
      if (++u == p.W) u = 0, ++v;
    }
  }
  __syncthreads();
  const int prefix = s_prefix, n3 = 3 * s_total;
  {
    T* dst = reinterpret_cast<T*>(p.points) + ((size_t)b * HW + prefix) * 3;
    for (int i = tid; i < n3; i += kPcThreads) dst[i] = sp[i];
  }
  if (p.rgb != nullptr) {  // reconstruct_terrain: the colours of the same pixels, same order (:99-104)
    __syncthreads();
    if (p.rgb_f32) {
      float* sc = reinterpret_cast<float*>(pc_smem);
      const float* rgb = reinterpret_cast<const float*>(p.rgb) + ((size_t)b * HW + px0) * 3;
      int r = excl;
#pragma unroll
      for (int k = 0; k < kPcPer; ++k)
        if (valid & (1u << k)) {
          sc[3 * r] = __ldg(rgb + 3 * k), sc[3 * r + 1] = __ldg(rgb + 3 * k + 1), sc[3 * r + 2] = __ldg(rgb + 3 * k + 2);
          ++r;
        }
      __syncthreads();
      float* dst = reinterpret_cast<float*>(p.colors) + ((size_t)b * HW + prefix) * 3;
      for (int i = tid; i < n3; i += kPcThreads) dst[i] = sc[i];
    } else {
      uint8_t* sc = pc_smem;
      const uint8_t* rgb = reinterpret_cast<const uint8_t*>(p.rgb) + ((size_t)b * HW + px0) * 3;
      int r = excl;
#pragma unroll
      for (int k = 0; k < kPcPer; ++k)
        if (valid & (1u << k)) {
          sc[3 * r] = __ldg(rgb + 3 * k), sc[3 * r + 1] = __ldg(rgb + 3 * k + 1), sc[3 * r + 2] = __ldg(rgb + 3 * k + 2);
          ++r;
        }
      __syncthreads();
      uint8_t* dst = reinterpret_cast<uint8_t*>(p.colors) + ((size_t)b * HW + prefix) * 3;
      // byte head up to a 4-byte boundary, words, byte tail
      const int head = min(n3, (int)((4 - (reinterpret_cast<uintptr_t>(dst) & 3)) & 3));
      if (tid < head) dst[tid] = sc[tid];
      const int nw = (n3 - head) >> 2;
      for (int i = tid; i < nw; i += kPcThreads) {
        const uint8_t* s = sc + head + 4 * i;
        reinterpret_cast<uint32_t*>(dst + head)[i] = (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | ((uint32_t)s[3] << 24);
      }
      const int done = head + 4 * nw;
      if (tid < n3 - done) dst[done + tid] = sc[done + tid];
    }
  }
}

// ------------------------------------------------------------------ disp -> depth (layers.py:16-25, terrain_reconstruction_fixed.py:168-171)
__global__ void __launch_bounds__(256) disp_to_depth_kernel(const float* __restrict__ disp, long long count, float min_disp, float disp_range,
                                                            float* __restrict__ depth) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const long long stride = (long long)gridDim.x * blockDim.x * 4;
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if ((count & 3) == 0) {
    for (; i < count; i += stride) {
      float4 v = __ldg(reinterpret_cast<const float4*>(disp + i));
      v.x = __fdiv_rn(1.0f, __fadd_rn(min_disp, __fmul_rn(disp_range, v.x)));
      v.y = __fdiv_rn(1.0f, __fadd_rn(min_disp, __fmul_rn(disp_range, v.y)));
      v.z = __fdiv_rn(1.0f, __fadd_rn(min_disp, __fmul_rn(disp_range, v.z)));
      v.w = __fdiv_rn(1.0f, __fadd_rn(min_disp, __fmul_rn(disp_range, v.w)));
      *reinterpret_cast<float4*>(depth + i) = v;
    }
  } else {
    for (; i < count; i += stride)
      for (int k = 0; k < 4 && i + k < count; ++k) depth[i + k] = __fdiv_rn(1.0f, __fadd_rn(min_disp, __fmul_rn(disp_range, __ldg(disp + i + k))));
  }
}

// ------------------------------------------------------------------ temporal mean (refinement.py:354-388)
struct TemporalParams {
  const float* maps[MDE_MAX_TEMPORAL_MAPS];
  double w[MDE_MAX_TEMPORAL_MAPS];
  int n;
  long long count;
  double* out;
};

// np.sum(stacked * weights[:, None, None], axis=0): float32 maps times float64 weights, added map by map in float64
__global__ void __launch_bounds__(256) temporal_mean_kernel(const __grid_constant__ TemporalParams p) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.count; i += stride) {
    double acc = __dmul_rn((double)__ldg(p.maps[0] + i), p.w[0]);
    for (int n = 1; n < p.n; ++n) acc = __dadd_rn(acc, __dmul_rn((double)__ldg(p.maps[n] + i), p.w[n]));
    p.out[i] = acc;
  }
}

// ------------------------------------------------------------------ bilateral depth filter (refinement.py:190-243)
constexpr int kBfBins = 4096;            // cv2: kExpNumBinsPerChannel
constexpr int kBfRadius = 7;             // d = 15
constexpr int kBfTile = 32;              // 32 x 32 outputs per block
constexpr int kBfSpan = kBfTile + 2 * kBfRadius;

__device__ __forceinline__ unsigned f32_ordered(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float f32_unordered(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

// range[0] = min, range[1] = max as order-preserving unsigned keys (initialised to 0xffffffff / 0 by the host side)
__global__ void __launch_bounds__(256) depth_range_kernel(const float* __restrict__ x, long long count, unsigned* __restrict__ range) {
  pdl_launch_dependents();
  pdl_grid_wait();
  float lo = INFINITY, hi = -INFINITY;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    const float v = __ldg(x + i);
    lo = fminf(lo, v), hi = fmaxf(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  __shared__ float slo[8], shi[8];
  if ((threadIdx.x & 31) == 0) slo[threadIdx.x >> 5] = lo, shi[threadIdx.x >> 5] = hi;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) lo = fminf(lo, slo[w]), hi = fmaxf(hi, shi[w]);
    atomicMin(range, f32_ordered(lo));
    atomicMax(range + 1, f32_ordered(hi));
  }
}

// The colour-weight table of cv2.bilateralFilter for float32 input, built from the range of the NORMALISED map:
// lut[i] = (float)exp((i / scale_index)^2 * -0.5 / sigma_color^2), scale_index = 4096 / (max_n - min_n); ctl = {scale_index,
// degenerate flag (cv2 returns the input when max_n - min_n < FLT_EPSILON), depth min, depth max - min + 1e-8, depth max - min}.
// lut == nullptr (joint form: the table depends on sigma_color only and comes from the host): ctl only.
__global__ void __launch_bounds__(256) bilateral_lut_kernel(const unsigned* __restrict__ range, float sigma_color, float* __restrict__ lut,
                                                            float* __restrict__ ctl) {
  pdl_launch_dependents();
  pdl_grid_wait();
  const float dmin = f32_unordered(range[0]), dmax = f32_unordered(range[1]);
  const float den = __fadd_rn(__fsub_rn(dmax, dmin), 1e-8f);
  const float nmin = __fdiv_rn(__fsub_rn(dmin, dmin), den), nmax = __fdiv_rn(__fsub_rn(dmax, dmin), den);
  const double span = (double)nmax - (double)nmin;
  const float len = (float)span;
  const float scale_index = (float)kBfBins / len;
  const double sc = sigma_color <= 0.0f ? 1.0 : (double)sigma_color;
  const double coeff = -0.5 / (sc * sc);
  if (lut != nullptr)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kBfBins + 2; i += gridDim.x * blockDim.x) {
      const double val = (double)i / (double)scale_index;
      lut[i] = (float)exp(val * val * coeff);
    }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    ctl[0] = scale_index;
    ctl[1] = fabs(span) < 1.1920928955078125e-07 ? 1.0f : 0.0f;
    ctl[2] = dmin;
    ctl[3] = den;
    ctl[4] = __fsub_rn(dmax, dmin);
  }
}

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
  return i;
}

struct BilateralParams {
  const float* depth;    // [H][W]
  const uint8_t* guide;  // optional [H][W][3] uint8: joint form
  const float* lut;      // kBfBins + 2 (plain form) or 3*256 entries (joint form)
  const float* ctl;
  const float* space_w;  // [15*15], 0 outside the circle
  float* out;
  int H, W;
};

// JOINT = false: cv2.bilateralFilter(depth_n, 15, sigma_depth, 15) on the normalised map (refinement.py:229-235)
// JOINT = true : cv2.ximgproc.jointBilateralFilter(guide=rgb uint8, src=depth_n, 15, sigma_color, 15) (:220-227): colour weight
//                indexed by |dr|+|dg|+|db| of the guide, centre pixel included
template <bool JOINT>
__global__ void __launch_bounds__(256) bilateral_kernel(const __grid_constant__ BilateralParams p) {
  pdl_launch_dependents();
  pdl_grid_wait();
  __shared__ float tile[kBfSpan][kBfSpan + 1];
  __shared__ uint32_t gtile[JOINT ? kBfSpan : 1][JOINT ? kBfSpan + 1 : 1];
  __shared__ float slut[JOINT ? 3 * 256 : kBfBins + 2];
  __shared__ float ssw[15 * 15];
  const int tid = threadIdx.x;
  const int X0 = blockIdx.x * kBfTile, Y0 = blockIdx.y * kBfTile;
  const float scale_index = p.ctl[0], dmin = p.ctl[2], den = p.ctl[3], span = p.ctl[4];
  const bool degenerate = p.ctl[1] != 0.0f;
  for (int i = tid; i < (JOINT ? 3 * 256 : kBfBins + 2); i += 256) slut[i] = p.lut[i];
  if (tid < 225) ssw[tid] = p.space_w[tid];
  for (int i = tid; i < kBfSpan * kBfSpan; i += 256) {
    const int ty = i / kBfSpan, tx = i - ty * kBfSpan;
    const int y = reflect101(Y0 + ty - kBfRadius, p.H), x = reflect101(X0 + tx - kBfRadius, p.W);
    tile[ty][tx] = __fdiv_rn(__fsub_rn(__ldg(p.depth + (size_t)y * p.W + x), dmin), den);
    if (JOINT) {
      const uint8_t* g = p.guide + ((size_t)y * p.W + x) * 3;
      gtile[ty][tx] = (uint32_t)__ldg(g) | ((uint32_t)__ldg(g + 1) << 8) | ((uint32_t)__ldg(g + 2) << 16);
    }
  }
  __syncthreads();
  const int lx = tid & 31;
#pragma unroll 1
  for (int q = 0; q < 4; ++q) {
    const int ly = (tid >> 5) + 8 * q;
    const int X = X0 + lx, Y = Y0 + ly;
    if (X >= p.W || Y >= p.H) continue;
    const float v0 = tile[ly + kBfRadius][lx + kBfRadius];
    float res = v0;
    if (JOINT) {
      const uint32_t g0 = gtile[ly + kBfRadius][lx + kBfRadius];
      const int r0 = g0 & 255, gg0 = (g0 >> 8) & 255, b0 = (g0 >> 16) & 255;
      float sum = 0.0f, wsum = 0.0f;
      for (int dy = -kBfRadius; dy <= kBfRadius; ++dy)
#pragma unroll
        for (int dx = -kBfRadius; dx <= kBfRadius; ++dx) {
          const float sw = ssw[(dy + kBfRadius) * 15 + dx + kBfRadius];
          if (sw == 0.0f) continue;
          const uint32_t g = gtile[ly + kBfRadius + dy][lx + kBfRadius + dx];
          const int idx = abs((int)(g & 255) - r0) + abs((int)((g >> 8) & 255) - gg0) + abs((int)((g >> 16) & 255) - b0);
          const float w = sw * slut[idx];
          sum = fmaf(tile[ly + kBfRadius + dy][lx + kBfRadius + dx], w, sum);
          wsum += w;
        }
      res = sum / wsum;
    } else if (!degenerate) {
      float sum = v0, wsum = 1.0f;  // centre pixel: both weights are 1
      for (int dy = -kBfRadius; dy <= kBfRadius; ++dy)
#pragma unroll
        for (int dx = -kBfRadius; dx <= kBfRadius; ++dx) {
          const float sw = ssw[(dy + kBfRadius) * 15 + dx + kBfRadius];
          if (sw == 0.0f || (dy == 0 && dx == 0)) continue;
          const float val = tile[ly + kBfRadius + dy][lx + kBfRadius + dx];
          float alpha = fabsf(val - v0) * scale_index;
          const int idx = (int)alpha;  // alpha >= 0: floor
          alpha -= (float)idx;
          const float w = sw * (slut[idx] + alpha * (slut[idx + 1] - slut[idx]));
          sum = fmaf(val, w, sum);
          wsum += w;
        }
      res = sum / wsum;
    }
    p.out[(size_t)Y * p.W + X] = __fadd_rn(__fmul_rn(res, span), dmin);  // filtered * (max - min) + min in float32 (:240)
  }
}

}  // namespace mde

using namespace mde;

extern "C" long long mde_point_cloud_scratch_bytes(int B, int H, int W) {
  if (B <= 0 || H <= 0 || W <= 0) return 0;
  const long long T = ((long long)H * W + kPcTile - 1) / kPcTile;
  return (B * T + 2) * 8;
}

extern "C" int mde_point_cloud(const float* src, int B, int H, int W, int from_disp, double min_depth, double max_depth, const double* intrinsics4,
                               int points_f64, void* points, const void* rgb, int rgb_f32, void* colors, int* counts, void* scratch,
                               mde_stream_t stream) {
  MDE_CHECK_ARG(src && points && counts && scratch && intrinsics4, "mde_point_cloud: null pointer");
  MDE_CHECK_ARG(B > 0 && H > 0 && W > 0 && (long long)H * W < (1ll << 30), "mde_point_cloud: bad shape B=%d H=%d W=%d", B, H, W);
  MDE_CHECK_ARG((rgb == nullptr) == (colors == nullptr), "mde_point_cloud: rgb and colors must be given together");
  MDE_CHECK_ARG(min_depth > 0.0 && max_depth > min_depth, "mde_point_cloud: need 0 < min_depth < max_depth");
  MDE_CHECK_ARG(intrinsics4[0] != 0.0 && intrinsics4[1] != 0.0, "mde_point_cloud: zero focal length");
  static bool attr_set = false;
  if (!attr_set) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(point_cloud_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPcTile * 3 * 8));
    attr_set = true;
  }
  PcParams p;
  memset(&p, 0, sizeof(p));
  p.src = src, p.rgb = rgb, p.points = points, p.colors = colors, p.counts = counts;
  p.B = B, p.H = H, p.W = W, p.T = (int)(((long long)H * W + kPcTile - 1) / kPcTile);
  p.state = reinterpret_cast<unsigned long long*>(scratch);
  p.ticket = reinterpret_cast<unsigned int*>(p.state + (size_t)B * p.T);
  p.from_disp = from_disp;
  // layers.py:21-24 with Python-float scalars: min_disp = 1/max_depth, max_disp = 1/min_depth in double, the products with a
  // float32 tensor round the scalar to float32 first
  const double min_disp = 1.0 / max_depth, max_disp = 1.0 / min_depth;
  p.min_disp = (float)min_disp, p.disp_range = (float)(max_disp - min_disp);
  p.min_depth = (float)min_depth, p.max_depth = (float)max_depth;  // z > min_depth on a float32 array compares in float32
  p.fx = intrinsics4[0], p.fy = intrinsics4[1], p.cx = intrinsics4[2], p.cy = intrinsics4[3];
  p.rgb_f32 = rgb_f32;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  MDE_CHECK_CUDA(cudaMemsetAsync(scratch, 0, (size_t)mde_point_cloud_scratch_bytes(B, H, W), s));
  const dim3 grid((unsigned)(B * p.T));
  if (points_f64) launch_k(point_cloud_kernel<double>, grid, dim3(kPcThreads), (size_t)kPcTile * 3 * 8, s, p);
  else launch_k(point_cloud_kernel<float>, grid, dim3(kPcThreads), (size_t)kPcTile * 3 * 4, s, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_temporal_mean(const float* const* maps, const double* weights, int n, long long count, double* out, mde_stream_t stream) {
  MDE_CHECK_ARG(maps && weights && out, "mde_temporal_mean: null pointer");
  MDE_CHECK_ARG(n >= 1 && n <= MDE_MAX_TEMPORAL_MAPS && count > 0, "mde_temporal_mean: need 1 <= n <= %d maps (got %d)", MDE_MAX_TEMPORAL_MAPS, n);
  TemporalParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < n; ++i) {
    MDE_CHECK_ARG(maps[i] != nullptr, "mde_temporal_mean: map %d is null", i);
    p.maps[i] = maps[i], p.w[i] = weights[i];
  }
  p.n = n, p.count = count, p.out = out;
  const long long blocks = (count + 255) / 256, cap = 8ll * sm_count();
  launch_k(temporal_mean_kernel, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_bilateral_depth(const float* depth, int H, int W, const uint8_t* guide, float sigma_color, const float* space_w,
                                   const float* joint_lut, float* scratch, float* out, mde_stream_t stream) {
  MDE_CHECK_ARG(depth && space_w && scratch && out, "mde_bilateral_depth: null pointer");
  MDE_CHECK_ARG(H > 0 && W > 0 && (long long)H * W < (1ll << 31), "mde_bilateral_depth: bad shape %dx%d", H, W);
  MDE_CHECK_ARG((guide == nullptr) == (joint_lut == nullptr), "mde_bilateral_depth: guide and joint_lut must be given together");
  MDE_CHECK_ARG(depth != out, "mde_bilateral_depth: in-place filtering is not supported");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  unsigned* range = reinterpret_cast<unsigned*>(scratch);
  float* ctl = scratch + 8;
  float* lut = scratch + 16;
  MDE_CHECK_CUDA(cudaMemsetAsync(range, 0xff, 4, s));
  MDE_CHECK_CUDA(cudaMemsetAsync(range + 1, 0, 4, s));
  const long long count = (long long)H * W, blocks = (count + 1023) / 1024, cap = 8ll * sm_count();
  launch_k(depth_range_kernel, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, s, depth, count, range);
  launch_k(bilateral_lut_kernel, dim3(guide ? 1 : 17), dim3(256), 0, s, (const unsigned*)range, sigma_color, guide ? (float*)nullptr : lut, ctl);
  BilateralParams p;
  memset(&p, 0, sizeof(p));
  p.depth = depth, p.guide = guide, p.lut = guide ? joint_lut : lut, p.ctl = ctl, p.space_w = space_w, p.out = out, p.H = H, p.W = W;
  const dim3 grid((unsigned)ceil_div(W, kBfTile), (unsigned)ceil_div(H, kBfTile));
  if (guide) launch_k(bilateral_kernel<true>, grid, dim3(256), 0, s, p);
  else launch_k(bilateral_kernel<false>, grid, dim3(256), 0, s, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_disp_to_depth(const float* disp, long long count, double min_depth, double max_depth, float* depth, mde_stream_t stream) {
  MDE_CHECK_ARG(disp && depth && count > 0, "mde_disp_to_depth: null pointer or empty map");
  MDE_CHECK_ARG(min_depth > 0.0 && max_depth > min_depth, "mde_disp_to_depth: need 0 < min_depth < max_depth");
  MDE_CHECK_ARG(((reinterpret_cast<uintptr_t>(disp) | reinterpret_cast<uintptr_t>(depth)) & 15) == 0, "mde_disp_to_depth: pointers must be 16-byte aligned");
  const double min_disp = 1.0 / max_depth, max_disp = 1.0 / min_depth;
  const long long blocks = (count + 1023) / 1024, cap = 16ll * sm_count();
  launch_k(disp_to_depth_kernel, dim3((unsigned)(blocks < cap ? blocks : cap)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), disp, count,
           (float)min_disp, (float)(max_disp - min_disp), depth);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
