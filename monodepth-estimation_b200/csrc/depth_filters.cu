// DepthQualityEnhancement.fill_depth_holes (models/refinement.py:301-352) and .edge_preserving_smoothing (:244-299) on the device.
//
// fill_depth_holes: the reference labels the 4-connected components of the invalid mask (scipy.ndimage.label), and for every
// component of at most max_hole_size^2 pixels takes the median of the ORIGINAL map over the ring "3 binary dilations of the component
// minus the component" (a Manhattan-distance-3 neighbourhood; it may contain pixels of OTHER holes, whose NaN then makes the median
// NaN, as np.median does) and writes it into the component; a 3x3 cv2.medianBlur (replicated border) of the result replaces every
// originally invalid pixel.  Here: union-find labelling with atomics (root = smallest linear index of the component), per-root size
// and bounding box, a compacted list of the small components, one block per component that gathers the ring into shared memory and
// bitonic-sorts it, and one stencil kernel for the blur + merge.  No host synchronisation anywhere.
//
// edge_preserving_smoothing: cv2.ximgproc.guidedFilter (He et al., box-filter form) on the min/max-normalised map with either the
// map itself or the grey image (cv2.cvtColor RGB2GRAY, 8-bit fixed point) as guide, radius min(H, W) / 16, eps = lambda.
// opencv-contrib is not installed in the build image, so this restates the published algorithm (oracle/terrain_oracle.py):
// "parity unpinned".  Window sums are accumulated in double from exact fp32 products.
#include <float.h>
#include <limits.h>
#include <math.h>

#include "common.h"

namespace mde {

#define MDE_GS(i, n) for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (long long)gridDim.x * blockDim.x)

static inline int grid_px(long long n) {
  long long g = (n + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

__device__ __forceinline__ bool hole_px(float v) { return !isfinite(v) || v <= 0.0f; }  // refinement.py:320

// ------------------------------------------------------------------ connected components (4-connectivity)
__device__ __forceinline__ int uf_find(const int* L, int i) {
  int p = L[i];
  while (p != i) {
    i = p;
    p = L[i];
  }
  return i;
}
__device__ __forceinline__ void uf_union(int* L, int a, int b) {
  while (true) {
    a = uf_find(L, a);
    b = uf_find(L, b);
    if (a == b) return;
    if (a < b) {
      const int t = a;
      a = b;
      b = t;
    }
    const int old = atomicMin(&L[a], b);  // a > b: hang the larger root under the smaller one
    if (old == a) return;
    a = old;
  }
}

struct HoleScratch {
  int* label;     // [H*W] root index of the pixel's component, -1 for valid pixels
  int* count;     // [H*W] at roots: component size
  int* bbox;      // [H*W][4] at roots: ymin, ymax, xmin, xmax
  int* list;      // [H*W] roots of the components to fill
  int* n_list;    // [1]
  float* filled;  // [H*W]
};

__global__ void __launch_bounds__(256) holes_init_kernel(const float* __restrict__ depth, long long n, HoleScratch s) {
  MDE_GS(i, n) {
    const float v = depth[i];
    s.filled[i] = v;
    s.label[i] = hole_px(v) ? (int)i : -1;
    s.count[i] = 0;
    s.bbox[4 * i] = INT_MAX;
    s.bbox[4 * i + 1] = -1;
    s.bbox[4 * i + 2] = INT_MAX;
    s.bbox[4 * i + 3] = -1;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *s.n_list = 0;
}

__global__ void __launch_bounds__(256) holes_merge_kernel(int H, int W, HoleScratch s) {
  MDE_GS(i, (long long)H * W) {
    if (s.label[i] < 0) continue;
    const int x = (int)(i % W), y = (int)(i / W);
    if (x > 0 && s.label[i - 1] >= 0) uf_union(s.label, (int)i, (int)i - 1);
    if (y > 0 && s.label[i - W] >= 0) uf_union(s.label, (int)i, (int)i - W);
  }
}

__global__ void __launch_bounds__(256) holes_stats_kernel(int H, int W, HoleScratch s) {
  MDE_GS(i, (long long)H * W) {
    if (s.label[i] < 0) continue;
    const int r = uf_find(s.label, (int)i);
    s.label[i] = r;  // parents only ever move towards the root: concurrent finds stay valid
    const int x = (int)(i % W), y = (int)(i / W);
    atomicAdd(&s.count[r], 1);
    atomicMin(&s.bbox[4 * r], y);
    atomicMax(&s.bbox[4 * r + 1], y);
    atomicMin(&s.bbox[4 * r + 2], x);
    atomicMax(&s.bbox[4 * r + 3], x);
  }
}

__global__ void __launch_bounds__(256) holes_select_kernel(long long n, int max_px, HoleScratch s) {
  MDE_GS(i, n) {
    if (s.label[i] == (int)i && s.count[i] <= max_px) s.list[atomicAdd(s.n_list, 1)] = (int)i;
  }
}

// One block per small component: ring = pixels within Manhattan distance 3 of the component that are not part of it
// (3 iterations of scipy's binary_dilation with the default cross structuring element, minus the component).
constexpr int kRingCap = 4096;       // >= 24 * max_px for max_hole_size <= 13 (each pixel's diamond has 24 other cells)
constexpr int kMaxHoleSize = 13;
__global__ void __launch_bounds__(256) holes_fill_kernel(const float* __restrict__ depth, int H, int W, HoleScratch s) {
  __shared__ float ring[kRingCap];
  __shared__ int n_ring, has_nan;
  const int n_list = *s.n_list;
  for (int e = blockIdx.x; e < n_list; e += gridDim.x) {
    const int root = s.list[e];
    const int y0 = s.bbox[4 * root], y1 = s.bbox[4 * root + 1], x0 = s.bbox[4 * root + 2], x1 = s.bbox[4 * root + 3];
    const int wy0 = max(0, y0 - 3), wy1 = min(H - 1, y1 + 3), wx0 = max(0, x0 - 3), wx1 = min(W - 1, x1 + 3);
    const int ww = wx1 - wx0 + 1, wn = ww * (wy1 - wy0 + 1);
    if (threadIdx.x == 0) {
      n_ring = 0;
      has_nan = 0;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < wn; t += blockDim.x) {
      const int y = wy0 + t / ww, x = wx0 + t % ww;
      if (s.label[(size_t)y * W + x] == root) continue;
      bool near = false;
      for (int dy = -3; dy <= 3 && !near; ++dy) {
        const int yy = y + dy;
        if (yy < 0 || yy >= H) continue;
        const int rem = 3 - abs(dy);
        for (int dx = -rem; dx <= rem; ++dx) {
          const int xx = x + dx;
          if (xx < 0 || xx >= W) continue;
          if (s.label[(size_t)yy * W + xx] == root) {
            near = true;
            break;
          }
        }
      }
      if (!near) continue;
      const float v = depth[(size_t)y * W + x];
      const int k = atomicAdd(&n_ring, 1);
      if (isnan(v)) has_nan = 1;
      if (k < kRingCap) ring[k] = isnan(v) ? INFINITY : v;
    }
    __syncthreads();
    const int n = min(n_ring, kRingCap);
    if (n > 0) {  // np.any(neighbor_mask), refinement.py:340
      int np2 = 1;
      while (np2 < n) np2 <<= 1;
      for (int t = n + threadIdx.x; t < np2; t += blockDim.x) ring[t] = INFINITY;
      __syncthreads();
      for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
          for (int t = threadIdx.x; t < np2; t += blockDim.x) {
            const int o = t ^ j;
            if (o > t) {
              const float a = ring[t], b = ring[o];
              const bool up = (t & k) == 0;
              if ((a > b) == up) {
                ring[t] = b;
                ring[o] = a;
              }
            }
          }
          __syncthreads();
        }
      // np.median: NaN if any NaN; else the middle element, or the fp32 mean of the two middle elements
      float med;
      if (has_nan) med = NAN;
      else if (n & 1) med = ring[n >> 1];
      else med = (ring[(n >> 1) - 1] + ring[n >> 1]) * 0.5f;
      const int bw = x1 - x0 + 1, bn = bw * (y1 - y0 + 1);
      for (int t = threadIdx.x; t < bn; t += blockDim.x) {
        const size_t p = (size_t)(y0 + t / bw) * W + x0 + t % bw;
        if (s.label[p] == root) s.filled[p] = med;
      }
    }
    __syncthreads();
  }
}

// cv2.medianBlur(ksize 3, float32): median of the 3x3 window with replicated border, selected (no arithmetic) — written where the
// ORIGINAL map was invalid (refinement.py:344-349); every other pixel keeps its value.  Windows that still hold a NaN (a hole too
// large to fill, or a NaN median) follow this min/max network's operand order, which OpenCV does not specify either: unpinned.
__device__ __forceinline__ void cswap(float& a, float& b) {
  const float lo = fminf(a, b), hi = fmaxf(a, b);
  a = lo;
  b = hi;
}
__global__ void __launch_bounds__(256) holes_blur_merge_kernel(const float* __restrict__ depth, const float* __restrict__ filled, int H, int W,
                                                               float* __restrict__ out) {
  MDE_GS(i, (long long)H * W) {
    const float d = depth[i];
    if (!hole_px(d)) {
      out[i] = d;
      continue;
    }
    const int x = (int)(i % W), y = (int)(i / W);
    float p[9];
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        const int yy = min(max(y + dy, 0), H - 1), xx = min(max(x + dx, 0), W - 1);
        p[(dy + 1) * 3 + dx + 1] = filled[(size_t)yy * W + xx];
      }
    cswap(p[1], p[2]); cswap(p[4], p[5]); cswap(p[7], p[8]); cswap(p[0], p[1]);
    cswap(p[3], p[4]); cswap(p[6], p[7]); cswap(p[1], p[2]); cswap(p[4], p[5]);
    cswap(p[7], p[8]); cswap(p[0], p[3]); cswap(p[5], p[8]); cswap(p[4], p[7]);
    cswap(p[3], p[6]); cswap(p[1], p[4]); cswap(p[2], p[5]); cswap(p[4], p[7]);
    cswap(p[4], p[2]); cswap(p[6], p[4]); cswap(p[4], p[2]);
    out[i] = p[4];
  }
}

// ------------------------------------------------------------------ guided filter (refinement.py:244-299)
// min/max of the map (refinement.py:262-264): per-block partials, then every consumer block reduces the <= kMmParts partials itself
constexpr int kMmParts = 512;
__global__ void __launch_bounds__(256) gf_minmax_kernel(const float* __restrict__ d, long long n, float* __restrict__ parts) {
  float mn = FLT_MAX, mx = -FLT_MAX;
  MDE_GS(i, n) {
    const float v = d[i];
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
  __shared__ float smn[256], smx[256];
  smn[threadIdx.x] = mn;
  smx[threadIdx.x] = mx;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if ((int)threadIdx.x < s) {
      smn[threadIdx.x] = fminf(smn[threadIdx.x], smn[threadIdx.x + s]);
      smx[threadIdx.x] = fmaxf(smx[threadIdx.x], smx[threadIdx.x + s]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    parts[2 * blockIdx.x] = smn[0];
    parts[2 * blockIdx.x + 1] = smx[0];
  }
}
__global__ void gf_minmax_final_kernel(const float* __restrict__ parts, int nparts, float* __restrict__ mm) {
  float mn = FLT_MAX, mx = -FLT_MAX;
  for (int i = threadIdx.x; i < nparts; i += 32) {
    mn = fminf(mn, parts[2 * i]);
    mx = fmaxf(mx, parts[2 * i + 1]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (threadIdx.x == 0) {
    mm[0] = mn;
    mm[1] = mx;
  }
}

// p = (d - min) / (max - min + 1e-8) in fp32 as NumPy evaluates it; guide I = p (self-guided) or the 8-bit grey value of the RGB
// image: cv2.cvtColor(RGB2GRAY) = (9798 R + 19235 G + 3735 B + 16384) >> 15 (opencv 4.13; pinned against cv2 in the tests).
__global__ void __launch_bounds__(256) gf_prepare_kernel(const float* __restrict__ d, const uint8_t* __restrict__ rgb, long long n, const float* __restrict__ mm,
                                                         float* __restrict__ p, float* __restrict__ I) {
  const float mn = mm[0], den = (mm[1] - mm[0]) + 1e-8f;
  MDE_GS(i, n) {
    const float v = __fdiv_rn(d[i] - mn, den);
    p[i] = v;
    if (rgb) {
      const int g = (9798 * rgb[3 * i] + 19235 * rgb[3 * i + 1] + 3735 * rgb[3 * i + 2] + 16384) >> 15;
      I[i] = (float)g;
    } else {
      I[i] = v;
    }
  }
}

__device__ __forceinline__ int reflect_idx(int i, int n) {  // cv2.BORDER_REFLECT: fedcba|abcdefgh|hgfedcb (radius < n is checked on the host)
  return i < 0 ? -i - 1 : (i >= n ? 2 * n - 1 - i : i);
}
// Separable box sums, window 2r+1, double accumulators.  Pass H (horizontal) of NQ quantities derived from (I, p) or read from a
// planar buffer; pass V (vertical) finishes the window mean and evaluates the point-wise step.
// step 1: quantities (I, p, I*I, I*p) -> a = cov / (var + eps), b = mean_p - a * mean_I;  step 2: quantities (a, b) -> q = mean_a * I + mean_b
__global__ void __launch_bounds__(256) gf_hsum4_kernel(const float* __restrict__ I, const float* __restrict__ p, int H, int W, int r, double* __restrict__ hs) {
  const long long n = (long long)H * W;
  MDE_GS(i, n) {
    const int x = (int)(i % W);
    const float* Ir = I + (i - x);
    const float* pr = p + (i - x);
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int dx = -r; dx <= r; ++dx) {
      const int xx = reflect_idx(x + dx, W);
      const double a = Ir[xx], b = pr[xx];
      s0 += a;
      s1 += b;
      s2 += a * a;
      s3 += a * b;
    }
    hs[i] = s0;
    hs[n + i] = s1;
    hs[2 * n + i] = s2;
    hs[3 * n + i] = s3;
  }
}
__global__ void __launch_bounds__(256) gf_vsum4_ab_kernel(const double* __restrict__ hs, int H, int W, int r, float eps, float* __restrict__ a_out,
                                                          float* __restrict__ b_out) {
  const long long n = (long long)H * W;
  const double inv = 1.0 / ((double)(2 * r + 1) * (2 * r + 1));
  MDE_GS(i, n) {
    const int x = (int)(i % W), y = (int)(i / W);
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int dy = -r; dy <= r; ++dy) {
      const size_t j = (size_t)reflect_idx(y + dy, H) * W + x;
      s0 += hs[j];
      s1 += hs[n + j];
      s2 += hs[2 * n + j];
      s3 += hs[3 * n + j];
    }
    const double mI = s0 * inv, mp = s1 * inv, var = s2 * inv - mI * mI, cov = s3 * inv - mI * mp;
    const double a = cov / (var + (double)eps);
    a_out[i] = (float)a;
    b_out[i] = (float)(mp - a * mI);
  }
}
__global__ void __launch_bounds__(256) gf_hsum2_kernel(const float* __restrict__ a, const float* __restrict__ b, int H, int W, int r, double* __restrict__ hs) {
  const long long n = (long long)H * W;
  MDE_GS(i, n) {
    const int x = (int)(i % W);
    const float* ar = a + (i - x);
    const float* br = b + (i - x);
    double s0 = 0, s1 = 0;
    for (int dx = -r; dx <= r; ++dx) {
      const int xx = reflect_idx(x + dx, W);
      s0 += ar[xx];
      s1 += br[xx];
    }
    hs[i] = s0;
    hs[n + i] = s1;
  }
}
// q = mean_a * I + mean_b, then back to the map's range: q * (max - min) + min in fp32 (refinement.py:297)
__global__ void __launch_bounds__(256) gf_vsum2_out_kernel(const double* __restrict__ hs, const float* __restrict__ I, int H, int W, int r,
                                                           const float* __restrict__ mm, float* __restrict__ out) {
  const long long n = (long long)H * W;
  const double inv = 1.0 / ((double)(2 * r + 1) * (2 * r + 1));
  const float mn = mm[0], range = mm[1] - mm[0];
  MDE_GS(i, n) {
    const int x = (int)(i % W), y = (int)(i / W);
    double s0 = 0, s1 = 0;
    for (int dy = -r; dy <= r; ++dy) {
      const size_t j = (size_t)reflect_idx(y + dy, H) * W + x;
      s0 += hs[j];
      s1 += hs[n + j];
    }
    const float q = (float)(s0 * inv * (double)I[i] + s1 * inv);
    out[i] = __fadd_rn(__fmul_rn(q, range), mn);
  }
}

}  // namespace mde

using namespace mde;

static inline long long align256(long long b) { return (b + 255) / 256 * 256; }

extern "C" long long mde_fill_depth_holes_scratch_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return MDE_ERR_BAD_ARGUMENT;
  const long long n = (long long)H * W;
  return align256(n * 4) * 4 + align256(n * 16) + 256;  // label, count, list, filled; bbox; n_list
}

extern "C" int mde_fill_depth_holes(const float* depth, int H, int W, int max_hole_size, void* scratch, float* out, mde_stream_t stream) {
  MDE_CHECK_ARG(depth && scratch && out && H > 0 && W > 0 && (long long)H * W < INT_MAX, "mde_fill_depth_holes: bad arguments");
  MDE_CHECK_ARG((uintptr_t)scratch % 256 == 0, "mde_fill_depth_holes: scratch must be 256-byte aligned");
  if (max_hole_size < 0 || max_hole_size > kMaxHoleSize) {
    set_error("mde_fill_depth_holes: max_hole_size %d not built (0..%d: the ring of a component is sorted in shared memory)", max_hole_size, kMaxHoleSize);
    return MDE_ERR_UNSUPPORTED;
  }
  const long long n = (long long)H * W;
  uint8_t* b = reinterpret_cast<uint8_t*>(scratch);
  HoleScratch s;
  s.label = reinterpret_cast<int*>(b);
  b += align256(n * 4);
  s.count = reinterpret_cast<int*>(b);
  b += align256(n * 4);
  s.list = reinterpret_cast<int*>(b);
  b += align256(n * 4);
  s.filled = reinterpret_cast<float*>(b);
  b += align256(n * 4);
  s.bbox = reinterpret_cast<int*>(b);
  b += align256(n * 16);
  s.n_list = reinterpret_cast<int*>(b);
  const int g = grid_px(n);
  holes_init_kernel<<<g, 256, 0, stream>>>(depth, n, s);
  MDE_CHECK_LAUNCH();
  holes_merge_kernel<<<g, 256, 0, stream>>>(H, W, s);
  MDE_CHECK_LAUNCH();
  holes_stats_kernel<<<g, 256, 0, stream>>>(H, W, s);
  MDE_CHECK_LAUNCH();
  holes_select_kernel<<<g, 256, 0, stream>>>(n, max_hole_size * max_hole_size, s);
  MDE_CHECK_LAUNCH();
  holes_fill_kernel<<<sm_count() * 8, 256, 0, stream>>>(depth, H, W, s);
  MDE_CHECK_LAUNCH();
  holes_blur_merge_kernel<<<g, 256, 0, stream>>>(depth, s.filled, H, W, out);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" long long mde_guided_filter_scratch_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return MDE_ERR_BAD_ARGUMENT;
  const long long n = (long long)H * W;
  return align256(n * 4) * 4 + align256(n * 8 * 4) + align256(kMmParts * 8) + 256;  // p, I, a, b; hs (4 double planes); partials; (min, max)
}

extern "C" int mde_guided_filter_depth(const float* depth, int H, int W, const uint8_t* rgb, int radius, float eps, void* scratch, float* out,
                                       mde_stream_t stream) {
  MDE_CHECK_ARG(depth && scratch && out && H > 0 && W > 0 && (long long)H * W < INT_MAX, "mde_guided_filter_depth: bad arguments");
  MDE_CHECK_ARG((uintptr_t)scratch % 256 == 0, "mde_guided_filter_depth: scratch must be 256-byte aligned");
  MDE_CHECK_ARG(radius >= 0 && radius < H && radius < W && eps >= 0.0f, "mde_guided_filter_depth: need 0 <= radius < min(H, W) and eps >= 0 (radius %d)", radius);
  const long long n = (long long)H * W;
  uint8_t* bp = reinterpret_cast<uint8_t*>(scratch);
  float* p = reinterpret_cast<float*>(bp);
  bp += align256(n * 4);
  float* I = reinterpret_cast<float*>(bp);
  bp += align256(n * 4);
  float* a = reinterpret_cast<float*>(bp);
  bp += align256(n * 4);
  float* b = reinterpret_cast<float*>(bp);
  bp += align256(n * 4);
  double* hs = reinterpret_cast<double*>(bp);
  bp += align256(n * 8 * 4);
  float* parts = reinterpret_cast<float*>(bp);
  bp += align256(kMmParts * 8);
  float* mm = reinterpret_cast<float*>(bp);
  const int g = grid_px(n);
  const int nparts = g < kMmParts ? g : kMmParts;
  gf_minmax_kernel<<<nparts, 256, 0, stream>>>(depth, n, parts);
  MDE_CHECK_LAUNCH();
  gf_minmax_final_kernel<<<1, 32, 0, stream>>>(parts, nparts, mm);
  MDE_CHECK_LAUNCH();
  gf_prepare_kernel<<<g, 256, 0, stream>>>(depth, rgb, n, mm, p, I);
  MDE_CHECK_LAUNCH();
  gf_hsum4_kernel<<<g, 256, 0, stream>>>(I, p, H, W, radius, hs);
  MDE_CHECK_LAUNCH();
  gf_vsum4_ab_kernel<<<g, 256, 0, stream>>>(hs, H, W, radius, eps, a, b);
  MDE_CHECK_LAUNCH();
  gf_hsum2_kernel<<<g, 256, 0, stream>>>(a, b, H, W, radius, hs);
  MDE_CHECK_LAUNCH();
  gf_vsum2_out_kernel<<<g, 256, 0, stream>>>(hs, I, H, W, radius, mm, out);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
