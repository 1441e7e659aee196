// 3x3 convolution on SMALL maps (width < 48: the 14x15 and 28x30 levels of decoder.py:97-145 and the 14x15 heads of
// enhanced_model.py:45-54,87-103) with wide channel counts (cin a multiple of 64, cout a multiple of 32), bf16 in / bf16 out.
//
// The tap-shifted implicit GEMM of tc_gemm.cu re-reads every input pixel 9 x (cout / 64) times and every weight once per 128
// pixels: 221 MB of L2 reads for a 256 -> 256 layer at 14x15x32 whose operands are 5.6 MB, and each CTA streams its 36 k-blocks
// at the per-SM L2 rate whatever the tile width (r02: 16.3 / 17.7 / 20.7 us with 64 / 128 / 256-column tiles) — 16 us for 5 us of
// MMAs.  Here a CTA owns (image rows, BLOCK_N output channels): per 64-channel chunk it loads the halo tile of its rows ONCE (one
// 4-D TMA box, pixels addressed flat with pitch w + 2 so that tap (dy,dx) is the same tile viewed from row dy*pitch + dx, as in
// conv_halo.cu) together with the chunk's nine BLOCK_N x 64 weight slices (one 3-D TMA box over [cin][cout][tap]), and runs
// 9 taps x 4 k-steps per 128-pixel MMA block on them: 55 MB of L2 reads for the same layer.  conv_halo.cu keeps ALL weights
// resident, which does not fit 256 input channels (9 x 4 chunks x 64 x 128 B = 295 KB); here they stream with the activations.
//
// Warp roles (320 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2..9 epilogue (two per TMEM lane
// quarter taking alternate MMA blocks).  One accumulator set (a CTA normally has ONE tile: <= 148 tiles at batch 32); the loop over
// tiles is there for larger problems.  The epilogue is conv_halo.cu's (bias, folded BN, ELU, padded-NHWC store with border copies).
#include <stdlib.h>

#include <algorithm>

#include "common.h"
#include "ptx.cuh"
#include "tc_epilogue.cuh"

namespace mde {

constexpr int kSmThreads = 320;
constexpr int kSmEpiWarps = 8;
constexpr int kSmMaxStages = 4;
constexpr int kSmMaxBlk = 4;

struct alignas(64) SmallParams {
  CUtensorMap tmA;  // 4-D box {64 channels, pitch px, th + 2 rows, 1 image} over the padded NHWC activation
  CUtensorMap tmW;  // 3-D box {64 channels, BLOCK_N rows, 9 taps} over the packed weight viewed as [cin][cout_pad][tap]
  int B, h, w, th, pitch, tiles_y, n_tiles, total_tiles;
  int chunks, nblk, stages, coord_off, halo_bytes, halo_slot, stage_bytes;
  // epilogue (field names shared with tc_epilogue.cuh)
  int epi, act;
  const float* bias;
  void* out;
  int ldo;
  const void* aux0;
  const void* aux1;
  int cout, out_cstride, out_coff, f16, st32;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kSmThreads, 1) conv_small_kernel(const __grid_constant__ SmallParams p) {
  pdl_launch_dependents();
  constexpr int WT = BLOCK_N * 128;  // one tap's weight slice: BLOCK_N rows x 64 channels
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * p.stage_bytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(p.nblk * BLOCK_N)) tmem_cols <<= 1;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmW);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tfull_bar, 1);
      mbar_init(tempty_bar, kSmEpiWarps);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, tmem_cols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_grid_wait();  // the set-up above overlaps the previous kernel's tail

  const int per_img = p.tiles_y * p.n_tiles;
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int b = tile / per_img, r = tile - b * per_img;
        const int y0 = (r / p.n_tiles) * p.th, n0 = (r % p.n_tiles) * BLOCK_N;
        for (int c = 0; c < p.chunks; ++c) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* st = smem + (size_t)stage * p.stage_bytes;
          mbar_expect_tx(&full_bar[stage], (uint32_t)(p.halo_bytes + 9 * WT));
          tma_load_4d(st, &p.tmA, &full_bar[stage], c * 64, p.coord_off, y0 + p.coord_off, b);
          tma_load_3d(st + p.halo_slot, &p.tmW, &full_bar[stage], c * 64, n0, 0);
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      const uint32_t idesc = p.f16 ? umma_idesc_f16(128, BLOCK_N) : umma_idesc_bf16(128, BLOCK_N);
      // descriptors differ only in the start-address field (address >> 4, no carry out of it below 256 KB): offsets are added
      const uint64_t d0 = umma_smem_desc_kmajor<128>(smem_u32(smem));
      const uint32_t stage_step = (uint32_t)p.stage_bytes >> 4, w_off = (uint32_t)p.halo_slot >> 4;
      int stage = 0;
      uint32_t phase = 0, tphase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        mbar_wait(tempty_bar, tphase ^ 1);  // the epilogue has drained the previous tile (passes at once for the first)
        tc_fence_after();
        for (int c = 0; c < p.chunks; ++c) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t da_s = d0 + (uint32_t)stage * stage_step;
          const uint64_t dw_s = da_s + w_off;
          for (int j = 0; j < p.nblk; ++j) {
            const uint32_t d_tmem = tmem_base + (uint32_t)(j * BLOCK_N);
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              const int dy = t / 3, dx = t - 3 * dy;
              const uint64_t da = da_s + (uint32_t)(((128 * j + dy * p.pitch + dx) * 128) >> 4);
              const uint64_t dw = dw_s + (uint32_t)((t * WT) >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k) umma_bf16_ss(d_tmem, da + 2 * k, dw + 2 * k, idesc, (c | t | k) != 0);
            }
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(tfull_bar);
        tphase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;  // the two warps of a lane quarter take alternate MMA blocks
    uint32_t tphase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const int b = tile / per_img, r = tile - b * per_img;
      const int y0 = (r / p.n_tiles) * p.th, n0 = (r % p.n_tiles) * BLOCK_N;
      mbar_wait(tfull_bar, tphase);
      tc_fence_after();
      for (int j = half; j < p.nblk; j += 2) {
        const int o = 128 * j + q * 32 + lane;
        const int yy = o / p.pitch, xx = o - yy * p.pitch;
        RowInfo ri;
        ri.m = 0;
        ri.b = b;
        ri.y = y0 + yy;
        ri.x = xx;
        ri.valid = xx < p.w && yy < p.th && ri.y < p.h;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + (uint32_t)(j * BLOCK_N);
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 32; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32b_x32(taddr + c * 32, raw);
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) v[jj] = __uint_as_float(raw[jj]);
          add_bias<32, SmallParams>(p, v, n0 + c * 32);
          if (ri.valid) epilogue_conv<32, SmallParams>(p, v, n0 + c * 32, ri);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar);
      tphase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ------------------------------------------------------------------------------------ host side
template <int BLOCK_N>
static int launch_small(SmallParams& p, int smem_bytes, cudaStream_t stream) {
  static bool attr = false;
  if (!attr) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(conv_small_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr = true;
  }
  const int grid = std::min(sm_count(), p.total_tiles);
  launch_k(conv_small_kernel<BLOCK_N>, dim3(grid), dim3(kSmThreads), smem_bytes, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

// Returns MDE_OK after launching, MDE_ERR_UNSUPPORTED (without setting an error) if this geometry is not served here.
int launch_conv_small(int f16, const void* in, int B, int h, int w, int cin, int in_cstride, const void* wpacked, const float* bias, int cout,
                      int padding, int activation, void* out_bf16, int out_cstride, int out_coff, cudaStream_t stream) {
  static const bool disabled = getenv("MDE_CONV_SMALL") != nullptr && getenv("MDE_CONV_SMALL")[0] == '0';
  if (disabled || !out_bf16 || w >= 48 || w < 3 || cin % 64 != 0 || cout % 32 != 0) return MDE_ERR_UNSUPPORTED;
  const int pitch = w + 2, chunks = cin / 64;
  // Tile shape: BLOCK_N = 64 output channels where cout allows it and the grid still covers most SMs, else 32; the fewest row
  // tiles whose flat extent th * pitch fits kSmMaxBlk MMA blocks of 128 pixels AND whose stage (halo slot + nine weight slices)
  // fits shared memory at least twice (two stages: the next chunk loads while this one is multiplied).
  static const int bn_env = getenv("MDE_CONV_SMALL_BN") ? atoi(getenv("MDE_CONV_SMALL_BN")) : 0;
  int block_n = 0, th = 0, tiles_y = 0, nblk = 0, halo_bytes = 0, halo_slot = 0, stage_bytes = 0, stages = 0;
  for (int bn = 64; bn >= 32 && block_n == 0; bn >>= 1) {
    if (cout % bn != 0 || (bn_env && bn != bn_env)) continue;
    for (int ty = ceil_div(h * pitch, kSmMaxBlk * 128); ty <= h; ++ty) {
      const int t = ceil_div(h, ty), nb = ceil_div(t * pitch, 128);
      if (nb > kSmMaxBlk) continue;
      const int hb = (t + 2) * pitch * 128;
      // the tap views of the last block read up to row 128 * nblk + 2 * pitch + 2 (garbage rows feed discarded outputs)
      const int slot = (std::max(hb, (128 * nb + 2 * pitch + 2) * 128) + 1023) & ~1023;
      const int sb = slot + 9 * bn * 128;
      const int st = std::min((227 * 1024 - 1024 - 256) / sb, std::min(kSmMaxStages, chunks));
      if (st < std::min(2, chunks)) continue;
      if (bn == 64 && !bn_env && cout % 32 == 0 && (long long)B * ty * (cout / 64) * 2 <= sm_count()) break;  // too few tiles: take 32 columns
      block_n = bn, th = t, tiles_y = ty, nblk = nb, halo_bytes = hb, halo_slot = slot, stage_bytes = sb, stages = st;
      break;
    }
  }
  if (block_n == 0) return MDE_ERR_UNSUPPORTED;
  // r02 (batch 32, back to back): 14x15 maps 256->256 16.4 -> 14.6 us, 256->128 16.2 -> 11.9, 256->64 15.5 -> 11.5; 28x30 maps (two or
  // four row tiles per image, each re-loading the weights) 29.3 -> 29.8 and 11.3 -> 12.1: only whole-image tiles are served here
  static const bool all_sizes = getenv("MDE_CONV_SMALL_ALL") != nullptr;
  if (tiles_y != 1 && !all_sizes) return MDE_ERR_UNSUPPORTED;

  SmallParams p;
  memset(&p, 0, sizeof(p));
  const bool zero = padding == MDE_PAD_ZERO;
  {
    const int H2 = h + 2, W2 = w + 2;
    const uint64_t strides[3] = {(uint64_t)in_cstride * 2, (uint64_t)W2 * in_cstride * 2, (uint64_t)H2 * W2 * in_cstride * 2};
    const uint32_t box[4] = {64, (uint32_t)pitch, (uint32_t)(th + 2), 1};
    int rc;
    if (zero) {  // view of the interior: out-of-bounds halo pixels read TMA zero fill
      const uint64_t dims[4] = {(uint64_t)cin, (uint64_t)w, (uint64_t)h, (uint64_t)B};
      rc = encode_tmap_bf16(&p.tmA, reinterpret_cast<const uint16_t*>(in) + ((size_t)W2 + 1) * in_cstride, 4, dims, strides, box, 128);
    } else {
      const uint64_t dims[4] = {(uint64_t)cin, (uint64_t)W2, (uint64_t)H2, (uint64_t)B};
      rc = encode_tmap_bf16(&p.tmA, in, 4, dims, strides, box, 128);
    }
    if (rc != MDE_OK) return rc;
  }
  {
    const int cout_pad = (cout + 15) & ~15;
    const uint64_t dims[3] = {(uint64_t)cin, (uint64_t)cout_pad, 9};
    const uint64_t strides[2] = {(uint64_t)9 * cin * 2, (uint64_t)cin * 2};
    const uint32_t box[3] = {64, (uint32_t)block_n, 9};
    int rc = encode_tmap_bf16(&p.tmW, wpacked, 3, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
  }
  p.B = B;
  p.h = h;
  p.w = w;
  p.th = th;
  p.pitch = pitch;
  p.tiles_y = tiles_y;
  p.n_tiles = cout / block_n;
  p.total_tiles = B * tiles_y * p.n_tiles;
  p.chunks = chunks;
  p.nblk = nblk;
  p.stages = stages;
  p.coord_off = zero ? -1 : 0;
  p.halo_bytes = halo_bytes;
  p.halo_slot = halo_slot;
  p.stage_bytes = stage_bytes;
  p.epi = EPI_CONV_BF16;
  p.act = activation;
  p.bias = bias;
  p.out = out_bf16;
  p.cout = cout;
  p.out_cstride = out_cstride;
  p.out_coff = out_coff;
  p.f16 = f16;
  p.st32 = cout % 16 == 0 && ((uintptr_t)out_bf16 % 32 == 0) && (out_cstride * 2) % 32 == 0 && (out_coff * 2) % 32 == 0;
  const int smem_bytes = 1024 + stages * stage_bytes + 256;
  if (block_n == 64) return launch_small<64>(p, smem_bytes, stream);
  return launch_small<32>(p, smem_bytes, stream);
}

}  // namespace mde
