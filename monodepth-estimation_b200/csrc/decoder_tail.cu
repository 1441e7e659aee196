// The last decoder level as ONE kernel (decoder.py:139-168 for i = 3): bilinear x2 upsample (align_corners=True) of the 16-channel
// map, upconv_3_1 (3x3 reflect, BN folded, ELU), mars_adapt (3x3 reflect, BN folded, ELU), dispconv_3 (3x3 zero pad, sigmoid).
//
// Why: at 224x240x16 channels per image these were four launches and four HBM round trips of a 55 MB (batch 32) activation —
// 188 us of a 2.25 ms step (r01s ablation) for 0.5 GFLOP per image — and a tcgen05 MMA with N = 16 reads a full 128 x 32 B A tile
// per 16 output columns (tensor pipe 11-14 % busy, r01s ncu).  Here a CTA owns a 16 x 32 output tile: it reads the low-resolution
// source once (L2), builds the upsampled tile with a 3-pixel halo in shared memory, runs both 16 -> 16 convolutions as implicit
// GEMMs on warp-level tensor-core MMAs (mma.sync m16n8k16, bf16 in, fp32 accumulate: M = 16 consecutive frame positions,
// K = 16 channels per tap, N = 16) from shared memory to shared memory, and writes only the 1-channel head.  DRAM traffic per
// image: 0.43 MB in (the 112x120x16 map) + 0.215 MB out instead of ~10 MB.
//
// Frame: the CTA's (TH + 6) x (TW + 6) window in image coordinates, addressed flat (f = fy * P + fx) so that tap (dy, dx) of any
// position is the position f + dy * P + dx; positions whose tap wraps around a row end are halo positions whose results are
// never used.  ReflectionPad2d(1) of each intermediate is restored between the stages by copying the reflected ring in place
// (zero ring for the zero-padded head).  Activations are rounded to bf16 at the same points as the unfused path.
#include "common.h"
#include "ptx.cuh"

namespace mde {

constexpr int kDtTH = 16, kDtTW = 32;
constexpr int kDtRows = kDtTH + 6, kDtP = kDtTW + 6;         // 22 x 38 frame
constexpr int kDtF = kDtRows * kDtP;                          // 836 frame positions
constexpr int kDtBlocks = (kDtF + 15) / 16;                   // 53 M-blocks of 16 positions
constexpr int kDtGuard = kDtP + 2;                            // positions of slack on either side (a tap reaches P + 1 away)
constexpr int kDtBufPos = kDtBlocks * 16 + 2 * kDtGuard;      // 928 positions x 32 B
constexpr int kDtThreads = 256;

// byte offset of channel-half `hf` (8 channels = 16 B) of frame position `pos` (may be negative down to -kDtGuard).  No swizzle: a tap
// shift is then a compile-time byte offset on one base address (the swizzled form spent ~10 instructions per tap on address
// arithmetic — the kernel was issue-bound, r02d ncu: 124.5 M warp instructions, HMMA pipe 24 %); an ldmatrix 8x8 tile sees a
// 2-way bank conflict instead.
__device__ __forceinline__ uint32_t dt_off(int pos, int hf) { return static_cast<uint32_t>(pos + kDtGuard) * 32u + static_cast<uint32_t>(hf << 4); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// One 3x3, 16 -> 16 convolution over the frame: src -> dst (shared memory), + bias, ELU, bf16.  Only the M-blocks [blk0, blk1) are
// computed: the ones that hold a position some later stage reads.  `w` (bf16 [16][144]) and `bias` are in SHARED memory (staged once
// at kernel start: fetching the B fragments from global at the start of each convolution cost an L2 round trip with every warp
// waiting, 23 % of the stall samples in r02d); a warp works on two M-blocks at a time = four independent accumulator chains.
__device__ __forceinline__ void dt_conv16(const uint8_t* src, uint8_t* dst, const bf16* w, const float* bias, int blk0, int blk1, int warp, int lane) {
  // B fragments of all nine taps: W[n][tap * 16 + k] (k contiguous) -> b0b1 = k (lane % 4) * 2, +1; b2b3 = k + 8, for n = nb * 8 + lane / 4
  uint32_t bf[9][2][2];
  const int n = lane >> 2, kq = (lane & 3) * 2;
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int nb = 0; nb < 2; ++nb) {
      const bf16* wr = w + (nb * 8 + n) * 144 + t * 16 + kq;
      bf[t][nb][0] = *reinterpret_cast<const uint32_t*>(wr);
      bf[t][nb][1] = *reinterpret_cast<const uint32_t*>(wr + 8);
    }
  float bv[2][2];
#pragma unroll
  for (int nb = 0; nb < 2; ++nb) {
    bv[nb][0] = bias[nb * 8 + kq];
    bv[nb][1] = bias[nb * 8 + kq + 1];
  }
  // ldmatrix.x4: lane l supplies the address of row (l % 8) + 8 * ((l / 8) & 1) of channel-half l / 16
  const uint32_t lbase = smem_u32(src) + dt_off((lane & 7) + ((lane >> 3) & 1) * 8, lane >> 4);
  constexpr int kWarps = kDtThreads / 32;
  for (int blk = blk0 + warp; blk < blk1; blk += 2 * kWarps) {
    const bool two = blk + kWarps < blk1;  // warp-uniform
    float acc[2][2][4];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int nb = 0; nb < 2; ++nb)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[u][nb][e] = 0.0f;
    const uint32_t abase0 = lbase + static_cast<uint32_t>(blk) * 512u;
    const uint32_t abase1 = two ? abase0 + kWarps * 512u : abase0;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int shift = ((t / 3 - 1) * kDtP + (t % 3 - 1)) * 32;
      uint32_t a0[4], a1[4];
      ldmatrix_x4(abase0 + shift, a0);
      ldmatrix_x4(abase1 + shift, a1);
      mma_bf16_16816(acc[0][0], a0, bf[t][0][0], bf[t][0][1]);
      mma_bf16_16816(acc[0][1], a0, bf[t][1][0], bf[t][1][1]);
      mma_bf16_16816(acc[1][0], a1, bf[t][0][0], bf[t][0][1]);
      mma_bf16_16816(acc[1][1], a1, bf[t][1][0], bf[t][1][1]);
    }
    // accumulator rows lane / 4 and lane / 4 + 8, channels nb * 8 + (lane % 4) * 2, +1
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int pos = (blk + u * kWarps) * 16 + (lane >> 2) + half * 8;
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
          const float v0 = elu1(acc[u][nb][2 * half] + bv[nb][0]), v1 = elu1(acc[u][nb][2 * half + 1] + bv[nb][1]);
          *reinterpret_cast<uint32_t*>(dst + dt_off(pos, nb) + kq * 2) = pack_bf16x2(v0, v1);
        }
      }
    }
  }
}

// Ring of width 1 around the image inside the frame: reflection of the interior (ReflectionPad2d(1)) or zeros (zero padding).
// Tiles whose frame lies inside the image (two thirds of them) have no ring: the caller skips the call (CTA-uniform test).
__device__ __forceinline__ void dt_fix_ring(uint8_t* buf, int Y0, int X0, int H, int W, bool zero, int tid) {
  for (int i = tid; i < kDtF * 2; i += kDtThreads) {
    const int f = i >> 1, hf = i & 1;
    const int fy = f / kDtP, fx = f - fy * kDtP;
    const int Y = Y0 - 3 + fy, X = X0 - 3 + fx;
    const bool in = Y >= 0 && Y < H && X >= 0 && X < W;
    const bool ring = !in && Y >= -1 && Y <= H && X >= -1 && X <= W;
    if (!ring) continue;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (!zero) {
      const int sy = Y < 0 ? 1 : (Y >= H ? H - 2 : Y), sx = X < 0 ? 1 : (X >= W ? W - 2 : X);
      const int sfy = sy - (Y0 - 3), sfx = sx - (X0 - 3);
      if (sfy < 0 || sfy >= kDtRows || sfx < 0 || sfx >= kDtP) continue;  // the source lies outside this CTA's frame: so does every reader of this ring position
      v = *reinterpret_cast<const uint4*>(buf + dt_off(sfy * kDtP + sfx, hf));
    }
    *reinterpret_cast<uint4*>(buf + dt_off(f, hf)) = v;
  }
}

struct DtParams {
  const bf16* src;  // padded NHWC [B][h + 2][w + 2][16], interior at (+1, +1)
  int h, w;
  const bf16 *w1, *w2, *w3;  // [16][144] each (BN folded); w3: row 0 is the head's single output channel
  const float *b1, *b2, *b3;
  float* out;  // [B][2h][2w]
  int tiles_x, tiles_y;
};

#ifdef MDE_GEMM_PROF
#define DT_T(i) do { if (prof_on) tp[i] = clock64(); } while (0)
#else
#define DT_T(i) do { } while (0)
#endif
__global__ void __launch_bounds__(kDtThreads, 3) decoder_tail_kernel(const DtParams p) {
#ifdef MDE_GEMM_PROF
  long long tp[10] = {0};
  const bool prof_on = threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == 500 || blockIdx.x == 2000 || blockIdx.x == 3500);
  DT_T(0);
#endif
  pdl_launch_dependents();
  extern __shared__ __align__(128) uint8_t dt_smem[];
  uint8_t* bufA = dt_smem;                    // upsampled input, later the second convolution's output
  uint8_t* bufB = dt_smem + kDtBufPos * 32;   // first convolution's output
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x % (p.tiles_x * p.tiles_y), b = blockIdx.x / (p.tiles_x * p.tiles_y);
  const int Y0 = (tile / p.tiles_x) * kDtTH, X0 = (tile % p.tiles_x) * kDtTW;
  const int H = 2 * p.h, W = 2 * p.w;
  // weights and biases of the three layers -> shared memory (3 x 4.6 KB), once
  __shared__ __align__(16) bf16 s_w[3][16 * 144];
  __shared__ float s_b[3][16];
  {
    const bf16* wsrc[3] = {p.w1, p.w2, p.w3};
    const float* bsrc[3] = {p.b1, p.b2, p.b3};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      for (int i = tid; i < 16 * 144 / 8; i += kDtThreads) reinterpret_cast<uint4*>(s_w[k])[i] = __ldg(reinterpret_cast<const uint4*>(wsrc[k]) + i);
      if (tid < 16) s_b[k][tid] = __ldg(bsrc[k] + tid);
    }
  }
  // the guard bands and the block padding are read by taps of halo positions: keep them finite
  for (int i = tid; i < kDtBufPos * 2; i += kDtThreads) {
    const int q = i >> 1;
    if (q < kDtGuard || q >= kDtGuard + kDtF) {
      *reinterpret_cast<uint4*>(bufA + (size_t)i * 16) = make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(bufB + (size_t)i * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  // per frame row / column: source taps and weights of the x2 upsample (src_index_act of pointwise.cu, align_corners=True), with
  // ring 1 reflected; rows / columns further outside the image are marked invalid (zero-filled)
  __shared__ int s_y0[kDtRows], s_y1[kDtRows], s_x0[kDtP], s_x1[kDtP];
  __shared__ float s_ly[kDtRows], s_lx[kDtP];
  if (tid < kDtRows + kDtP) {
    const bool isy = tid < kDtRows;
    const int fi = isy ? tid : tid - kDtRows;
    const int n_out = isy ? H : W, n_in = isy ? p.h : p.w;
    int c = (isy ? Y0 : X0) - 3 + fi;
    int i0 = -1, i1 = 0;
    float l = 0.0f;
    if (c >= -1 && c <= n_out) {
      c = c < 0 ? 1 : (c >= n_out ? n_out - 2 : c);
      const float sc = (float)(n_in - 1) / (float)(n_out - 1) * (float)c;
      i0 = (int)sc;
      if (i0 > n_in - 1) i0 = n_in - 1;
      i1 = i0 + (i0 < n_in - 1 ? 1 : 0);
      l = sc - (float)i0;
    }
    if (isy) {
      s_y0[fi] = i0;
      s_y1[fi] = i1;
      s_ly[fi] = l;
    } else {
      s_x0[fi] = i0;
      s_x1[fi] = i1;
      s_lx[fi] = l;
    }
  }
  __syncthreads();
  DT_T(1);
  pdl_grid_wait();
  DT_T(2);
  // ---- stage 0: x = upsample2x(src) (decoder.py:139-142) on the frame
  {
    const bf16* sb = p.src + ((size_t)b * (p.h + 2) * (p.w + 2) + (p.w + 2) + 1) * 16;  // interior origin
    const int rowp = (p.w + 2) * 16;
    for (int i0 = tid; i0 < kDtF * 2; i0 += 2 * kDtThreads) {
      // two items per iteration, their eight 16-byte loads issued before the first use (L2 latency: r02d, 23 % of the stall samples)
      uint4 u[2][4];
      float wgt[2][4];
      bool live[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int i = i0 + e * kDtThreads;
        const int f = min(i, kDtF * 2 - 1) >> 1, hf = i & 1;
        const int fy = f / kDtP, fx = f - fy * kDtP;
        const int y0 = s_y0[fy], x0 = s_x0[fx];
        live[e] = i < kDtF * 2 && y0 >= 0 && x0 >= 0;
        const int y1 = s_y1[fy], x1 = s_x1[fx];
        const float ly = s_ly[fy], lx = s_lx[fx];
        wgt[e][0] = (1.0f - ly) * (1.0f - lx);
        wgt[e][1] = (1.0f - ly) * lx;
        wgt[e][2] = ly * (1.0f - lx);
        wgt[e][3] = ly * lx;
        const bf16* r0 = sb + (size_t)max(y0, 0) * rowp + hf * 8;
        const bf16* r1 = sb + (size_t)y1 * rowp + hf * 8;
        if (live[e]) {
          u[e][0] = __ldg(reinterpret_cast<const uint4*>(r0 + x0 * 16));
          u[e][1] = __ldg(reinterpret_cast<const uint4*>(r0 + x1 * 16));
          u[e][2] = __ldg(reinterpret_cast<const uint4*>(r1 + x0 * 16));
          u[e][3] = __ldg(reinterpret_cast<const uint4*>(r1 + x1 * 16));
        }
      }
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int i = i0 + e * kDtThreads;
        if (i >= kDtF * 2) break;
        uint4 o = make_uint4(0u, 0u, 0u, 0u);
        if (live[e]) {
          const uint32_t a00[4] = {u[e][0].x, u[e][0].y, u[e][0].z, u[e][0].w}, a01[4] = {u[e][1].x, u[e][1].y, u[e][1].z, u[e][1].w};
          const uint32_t a10[4] = {u[e][2].x, u[e][2].y, u[e][2].z, u[e][2].w}, a11[4] = {u[e][3].x, u[e][3].y, u[e][3].z, u[e][3].w};
          uint32_t r[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // a bf16 pair -> two floats: the low half shifted up, the high half masked (exact)
            const float vx = wgt[e][0] * __uint_as_float(a00[k] << 16) + wgt[e][1] * __uint_as_float(a01[k] << 16) +
                             wgt[e][2] * __uint_as_float(a10[k] << 16) + wgt[e][3] * __uint_as_float(a11[k] << 16);
            const float vy = wgt[e][0] * __uint_as_float(a00[k] & 0xffff0000u) + wgt[e][1] * __uint_as_float(a01[k] & 0xffff0000u) +
                             wgt[e][2] * __uint_as_float(a10[k] & 0xffff0000u) + wgt[e][3] * __uint_as_float(a11[k] & 0xffff0000u);
            r[k] = pack_bf16x2(vx, vy);
          }
          o = make_uint4(r[0], r[1], r[2], r[3]);
        }
        *reinterpret_cast<uint4*>(bufA + dt_off(i >> 1, i & 1)) = o;
      }
    }
  }
  __syncthreads();
  DT_T(3);
  // does the frame reach outside the image?  (CTA-uniform; only then is there a padding ring to restore)
  const bool border = Y0 - 3 < 0 || X0 - 3 < 0 || Y0 + kDtTH + 3 > H || X0 + kDtTW + 3 > W;
  // ---- upconv_3_1 (decoder.py:152) -> bufB, frame rows 1..20 (blocks 2..49); its reflected ring
  dt_conv16(bufA, bufB, s_w[0], s_b[0], (kDtP + 1) / 16, (20 * kDtP + kDtP - 2) / 16 + 1, warp, lane);
  __syncthreads();
  if (border) {
    dt_fix_ring(bufB, Y0, X0, H, W, false, tid);
    __syncthreads();
  }
  DT_T(4);
  // ---- mars_adapt (decoder.py:157-158) -> bufA, frame rows 2..19; zero ring for the zero-padded head (decoder.py:86-88)
  dt_conv16(bufB, bufA, s_w[1], s_b[1], (2 * kDtP + 2) / 16, (19 * kDtP + kDtP - 3) / 16 + 1, warp, lane);
  __syncthreads();
  if (border) {
    dt_fix_ring(bufA, Y0, X0, H, W, true, tid);
    __syncthreads();
  }
  DT_T(5);
  // ---- dispconv_3 + sigmoid (decoder.py:166): 16 -> 1 as an N = 8 MMA whose columns 1..7 are zero; frame rows 3..18
  {
    uint32_t bf[9][2];
    const int n = lane >> 2, kq = (lane & 3) * 2;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      bf[t][0] = n == 0 ? *reinterpret_cast<const uint32_t*>(s_w[2] + t * 16 + kq) : 0u;
      bf[t][1] = n == 0 ? *reinterpret_cast<const uint32_t*>(s_w[2] + t * 16 + kq + 8) : 0u;
    }
    const float b3 = s_b[2][0];
    const uint32_t lbase = smem_u32(bufA) + dt_off((lane & 7) + ((lane >> 3) & 1) * 8, lane >> 4);
    constexpr int kB0 = (3 * kDtP + 3) / 16, kB1 = (18 * kDtP + kDtP - 4) / 16 + 1;
    for (int blk = kB0 + warp; blk < kB1; blk += kDtThreads / 32) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const uint32_t abase = lbase + static_cast<uint32_t>(blk) * 512u;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        uint32_t a[4];
        ldmatrix_x4(abase + ((t / 3 - 1) * kDtP + (t % 3 - 1)) * 32, a);
        mma_bf16_16816(acc, a, bf[t][0], bf[t][1]);
      }
      if ((lane & 3) == 0) {  // column 0 of rows lane / 4 (acc[0]) and lane / 4 + 8 (acc[2])
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int f = blk * 16 + (lane >> 2) + half * 8;
          const int fy = f / kDtP, fx = f - fy * kDtP;
          const int ty = fy - 3, tx = fx - 3;
          if (ty >= 0 && ty < kDtTH && tx >= 0 && tx < kDtTW && Y0 + ty < H && X0 + tx < W)
            p.out[((size_t)b * H + Y0 + ty) * W + X0 + tx] = sigmoidf_(acc[2 * half] + b3);
        }
      }
    }
  }
#ifdef MDE_GEMM_PROF
  if (prof_on) {
    DT_T(6);
    printf("tail blk %d: setup %lld pdlwait %lld upsample %lld conv1+ring %lld conv2+ring %lld head %lld total %lld cycles\n", blockIdx.x, tp[1] - tp[0], tp[2] - tp[1],
           tp[3] - tp[2], tp[4] - tp[3], tp[5] - tp[4], tp[6] - tp[5], tp[6] - tp[0]);
  }
#endif
}

}  // namespace mde

using namespace mde;

extern "C" int mde_decoder_tail16(const mde_bf16* src_padded, int B, int h, int w, const mde_bf16* w1, const float* b1, const mde_bf16* w2, const float* b2,
                                  const mde_bf16* w3, const float* b3, float* out, mde_stream_t stream) {
  MDE_CHECK_ARG(src_padded && w1 && b1 && w2 && b2 && w3 && b3 && out && B > 0 && h >= 2 && w >= 2, "mde_decoder_tail16: bad arguments");
  MDE_CHECK_ARG(((uintptr_t)src_padded | (uintptr_t)w1 | (uintptr_t)w2 | (uintptr_t)w3) % 16 == 0, "mde_decoder_tail16: pointers must be 16-byte aligned");
  DtParams p;
  p.src = reinterpret_cast<const bf16*>(src_padded);
  p.h = h;
  p.w = w;
  p.w1 = reinterpret_cast<const bf16*>(w1);
  p.w2 = reinterpret_cast<const bf16*>(w2);
  p.w3 = reinterpret_cast<const bf16*>(w3);
  p.b1 = b1;
  p.b2 = b2;
  p.b3 = b3;
  p.out = out;
  p.tiles_x = ceil_div(2 * w, kDtTW);
  p.tiles_y = ceil_div(2 * h, kDtTH);
  const long long blocks = (long long)B * p.tiles_x * p.tiles_y;
  MDE_CHECK_ARG(blocks < (1ll << 31), "mde_decoder_tail16: too many tiles");
  const int smem = 2 * kDtBufPos * 32;
  static bool attr_set = false;
  if (!attr_set) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(decoder_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(decoder_tail_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  launch_k(decoder_tail_kernel, dim3((unsigned)blocks), dim3(kDtThreads), smem, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
