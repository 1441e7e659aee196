// Halo-resident implicit-GEMM 3x3 convolution (tcgen05 + TMA), the fast path of mde_conv_nhwc for maps >= 48 px wide.
//
// The tap-shifted kernel in tc_gemm.cu pulls every input pixel from L2 nine times (one TMA box per tap); for a
// 64-channel layer that is 147 KB per 128-pixel tile against ~1 us of MMA time, i.e. L2->SM bound at ~1/3 of tensor
// peak, and for 16-channel layers it is issue-bound on 18 tiny TMA loads per tile (r01b: 157 us for a 17 us layer).
// Here a CTA loads a (TH+2) x 64-pixel halo tile of the padded NHWC activation ONCE per 64-channel chunk with a single
// TMA box and keeps the whole 3x3 weight set resident in shared memory.  The pixels of the tile are addressed FLAT,
// o = y*64 + x (62 real outputs per row + 2 discarded), so the A operand of tap (dy,dx) for the 128-row MMA block j is
// simply the same shared-memory tile viewed from row 128*j + dy*64 + dx: nine descriptors with shifted start
// addresses, no data movement.  The hardware swizzle is a function of the shared-memory address bits, so a view that
// starts at any 16-byte-aligned row of a TMA-written tile stays consistent with how TMA laid the tile out.
//
// Warp roles: warp 0 TMA producer (weights once, then one halo box per tile and channel chunk), warp 1
// TMEM allocator + MMA issuer, warps 2.. epilogue — two or four warps per TMEM lane quarter dealing out the MMA blocks and column chunks, and for
// the low-channel layers (small shared-memory footprint) several CTAs per SM: with one warp per scheduler the
// dependent-issue latency of the ELU/pack/store chain, not HBM, bounded the 16-channel layers (r01g: 113 us for 17 us of traffic).  The epilogue itself is shared with tc_gemm.cu: bias,
// folded BN, ELU/sigmoid, bf16/fp16 padded-NHWC store with reflected border, or fp32 head.  Two accumulator sets in TMEM overlap the epilogue of tile
// i with the MMAs of tile i+1.  (A finer hand-off, one tcgen05.commit per 128-pixel MMA block, measured 8-12 % SLOWER on every
// layer shape, r02: profiles/r02_experiments.md.)
#include <stdlib.h>

#include <algorithm>

#include "common.h"
#include "ptx.cuh"
#include "tc_epilogue.cuh"

namespace mde {

constexpr int kHaloPitch = 64;  // halo tile width in pixels
constexpr int kHaloTW = 62;     // real outputs per tile row
// Epilogue warps per CTA.  The 128-channel layers run ONE CTA per SM (TMEM: 2 x 2 blocks x 128 columns), and with 8 epilogue
// warps that is two warps per scheduler for ~9 instructions (one of them MUFU) per output element: 16 warps, four per TMEM lane
// quarter, took the 16->128 layer at 512x512x4 from 131 to 111 us (r02).  N = 64 measured no better with 12 or 16 (89.6 / 89.0 / 91.7 us).
__host__ __device__ constexpr int halo_epi_warps(int block_n) {
#ifdef MDE_HALO_EPIW
  return block_n >= 128 ? MDE_HALO_EPIW : 8;
#else
  return block_n >= 128 ? 16 : 8;
#endif
}
__host__ __device__ constexpr int halo_threads(int block_n) { return 64 + 32 * halo_epi_warps(block_n); }
constexpr int kHaloMaxStages = 4;

struct alignas(64) HaloParams {
  CUtensorMap tmA;  // 4-D box {KC channels, 64 px, TH+2 rows, 1 image}
  CUtensorMap tmB;  // 2-D box {KC, BLOCK_N} over wpacked [cout_pad][9*cin]
  int B, h, w, th, tiles_x, tiles_y, total_tiles;
  int chunks, nblk, stages, coord_off, halo_bytes, stage_bytes;
  // epilogue (field names shared with tc_epilogue.cuh)
  int epi, act;
  const float* bias;
  void* out;
  int ldo;
  const void* aux0;
  const void* aux1;
  int cout, out_cstride, out_coff, f16, st32;
  int dbg;  // MDE_HALO_DBG bit mask, tools/conv_bench.py only: 1 no epilogue, 2 no MMAs, 4 no activation/stores, 8 no halo loads, 16 stores to one pixel
};

template <int BLOCK_N, int ROWB>
__global__ void __launch_bounds__(halo_threads(BLOCK_N), BLOCK_N <= 16 ? 3 : BLOCK_N <= 32 ? 2 : 1) conv_halo_kernel(const __grid_constant__ HaloParams p) {
  pdl_launch_dependents();
  constexpr int KC = ROWB / 2;
  constexpr int WB_RAW = BLOCK_N * ROWB;
  constexpr int WB = (WB_RAW + 1023) & ~1023;
  constexpr int NCOL = BLOCK_N >= 32 ? 32 : 16;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int wtiles = 9 * p.chunks;
  const int stages = p.stages;
  uint8_t* sW = smem;
  uint8_t* sH = smem + (size_t)wtiles * WB;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sH + (size_t)stages * p.stage_bytes + 1024);  // 1 KB pad: tap views overrun by 2 pixels
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* w_bar = tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_img = p.tiles_x * p.tiles_y;
  uint32_t tmem_cols = 32;
  while (tmem_cols < static_cast<uint32_t>(2 * p.nblk * BLOCK_N)) tmem_cols <<= 1;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tfull_bar[a], 1);
        mbar_init(&tempty_bar[a], halo_epi_warps(BLOCK_N));
      }
      mbar_init(w_bar, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, tmem_cols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // The 3x3 weight set is a parameter, not the previous kernel's output: it is requested BEFORE the grid-dependency wait, so a CTA
  // that became resident while the previous kernel drains has its weights in shared memory when the wait ends.
  if (warp == 0 && elect_one()) {
    mbar_expect_tx(w_bar, (uint32_t)(wtiles * WB_RAW));
    for (int t = 0; t < wtiles; ++t) tma_load_2d(sW + (size_t)t * WB, &p.tmB, w_bar, t * KC, 0);
  }
  pdl_grid_wait();  // the set-up above overlaps the previous kernel's tail

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int b = tile / per_img, r = tile - b * per_img;
        const int y0 = (r / p.tiles_x) * p.th, x0 = (r % p.tiles_x) * kHaloTW;
        for (int c = 0; c < p.chunks; ++c) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (p.dbg & 8) {
            mbar_arrive(&full_bar[stage]);
          } else {
            mbar_expect_tx(&full_bar[stage], (uint32_t)p.halo_bytes);
            tma_load_4d(sH + (size_t)stage * p.stage_bytes, &p.tmA, &full_bar[stage], c * KC, x0 + p.coord_off, y0 + p.coord_off, b);
          }
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
      // Shared-memory descriptors differ only in their start-address field (bits 0..13, address >> 4; every address here is below
      // 256 KB, so sums never carry out of the field): build one per operand and ADD the stage / block / tap / k-step offsets
      // instead of rebuilding a descriptor (~15 instructions) per MMA.
      const uint64_t da0 = umma_smem_desc_kmajor<ROWB>(smem_u32(sH));
      const uint64_t db0 = umma_smem_desc_kmajor<ROWB>(smem_u32(sW));
      const uint32_t stage_step = (uint32_t)p.stage_bytes >> 4;
      const uint32_t wchunk_step = (uint32_t)WB >> 4, wtap_step = (uint32_t)(p.chunks * WB) >> 4;
      constexpr uint32_t kBlkStep = (128 * ROWB) >> 4;
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      mbar_wait(w_bar, 0);
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        for (int c = 0; c < p.chunks; ++c) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t da_s = da0 + (uint32_t)stage * stage_step;
          const uint64_t db_c = db0 + (uint32_t)c * wchunk_step;
          for (int j = 0; j < ((p.dbg & 2) ? 0 : p.nblk); ++j) {
            const uint32_t d_tmem = tmem_base + (uint32_t)((acc * p.nblk + j) * BLOCK_N);
            const uint64_t da_b = da_s + (uint32_t)j * kBlkStep;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              const int dy = t / 3, dx = t - 3 * dy;
              const uint64_t da = da_b + (uint32_t)(((dy * kHaloPitch + dx) * ROWB) >> 4);
              const uint64_t db = db_c + (uint32_t)t * wtap_step;
#pragma unroll
              for (int k = 0; k < KC / 16; ++k) umma_bf16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (c | t | k) != 0);
            }
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tfull_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int q = warp & 3;
    const int sub = (warp - 2) >> 2;  // the warps of a lane quarter deal out the (MMA block, column chunk) items of a tile
    constexpr int kSubs = halo_epi_warps(BLOCK_N) / 4, kChunks = BLOCK_N / NCOL;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const int b = tile / per_img, r = tile - b * per_img;
      const int y0 = (r / p.tiles_x) * p.th, x0 = (r % p.tiles_x) * kHaloTW;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      for (int i = sub; i < ((p.dbg & 1) ? 0 : p.nblk * kChunks); i += kSubs) {
        const int j = i / kChunks, c = i - j * kChunks;
        const int o = 128 * j + q * 32 + lane;
        const int yy = o >> 6, xx = o & 63;
        RowInfo ri;
        ri.m = 0;
        ri.b = b;
        ri.y = y0 + yy;
        ri.x = x0 + xx;
        ri.valid = xx < kHaloTW && ri.y < p.h && ri.x < p.w && !(p.dbg & 4);
        if (p.dbg & 16) ri.b = 0, ri.y = 2, ri.x = 2;  // every lane stores to one pixel: same instructions, no store traffic
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + (uint32_t)((acc * p.nblk + j) * BLOCK_N);
        uint32_t raw[NCOL];
        if constexpr (NCOL == 32) tmem_ld_32x32b_x32(taddr + c * NCOL, raw);
        else tmem_ld_32x32b_x16(taddr + c * NCOL, raw);
        tmem_ld_wait();
        float v[NCOL];
#pragma unroll
        for (int jj = 0; jj < NCOL; ++jj) v[jj] = __uint_as_float(raw[jj]);
        add_bias<NCOL, HaloParams>(p, v, c * NCOL);
        if (ri.valid) epilogue_conv<NCOL, HaloParams>(p, v, c * NCOL, ri);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ------------------------------------------------------------------------------------ host side
template <int BLOCK_N, int ROWB>
static int launch_halo(HaloParams& p, int smem_bytes, cudaStream_t stream) {
  static int smem_attr = 0;
  if (smem_attr == 0) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_kernel<BLOCK_N, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(conv_halo_kernel<BLOCK_N, ROWB>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    smem_attr = 1;
  }
  // persistent CTAs: as many per SM as registers / shared memory / TMEM columns allow (1 for the 64-channel layers).  Computed
  // here rather than with cudaOccupancyMaxActiveBlocksPerMultiprocessor: that call answered 1 for the first launch of an
  // instantiation (before the carveout preference had taken effect) and 2 afterwards, so eager and captured runs differed.
  static int regs_limit = 0;
  if (regs_limit == 0) {
    cudaFuncAttributes fa;
    MDE_CHECK_CUDA(cudaFuncGetAttributes(&fa, conv_halo_kernel<BLOCK_N, ROWB>));
    const int per_thread = (fa.numRegs + 7) & ~7;
    regs_limit = std::max(1, 65536 / (per_thread * halo_threads(BLOCK_N)));
  }
  int tmem_cols = 32;
  while (tmem_cols < 2 * p.nblk * BLOCK_N) tmem_cols <<= 1;
  int occ = std::min(regs_limit, (227 * 1024) / (smem_bytes + 1024));
  occ = std::max(1, std::min(occ, 512 / tmem_cols));
  int grid = sm_count() * occ;
  if (grid > p.total_tiles) grid = p.total_tiles;
  launch_k(conv_halo_kernel<BLOCK_N, ROWB>, dim3(grid), dim3(halo_threads(BLOCK_N)), smem_bytes, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

// Returns MDE_OK after launching, MDE_ERR_UNSUPPORTED (without setting an error) if this geometry is not served here.
int launch_conv_halo(int f16, const void* in, int B, int h, int w, int cin, int in_cstride, const void* wpacked, const float* bias, int cout,
                     int padding, int activation, void* out_bf16, int out_cstride, int out_coff, float* out_f32, cudaStream_t stream) {
  static const bool disabled = getenv("MDE_CONV_HALO") != nullptr && getenv("MDE_CONV_HALO")[0] == '0';
  if (disabled || w < 48) return MDE_ERR_UNSUPPORTED;
  int rowb;
  if (cin % 64 == 0) rowb = 128;
  else if (cin == 32) rowb = 64;
  else if (cin == 16) rowb = 32;
  else return MDE_ERR_UNSUPPORTED;
  const int kc = rowb / 2;
  const int cout_pad = (cout + 15) & ~15;
  if (cout_pad != 16 && cout_pad != 32 && cout_pad != 64 && cout_pad != 128) return MDE_ERR_UNSUPPORTED;
  const int block_n = cout_pad;
  const int chunks = cin / kc;
  const int wb = (block_n * rowb + 1023) & ~1023;
  const int weights_bytes = 9 * chunks * wb;
  int th = 0, stages = 0, stage_bytes = 0;
  // Low-channel layers are bound by the latency of the epilogue's dependent instruction chain, so resident warps matter more
  // than pipeline depth: give up halo stages (down to 2) to fit 3 CTAs per SM (N = 16) or 2 (N = 32).  MDE_HALO_OCC overrides.
  static const int occ_env = getenv("MDE_HALO_OCC") ? atoi(getenv("MDE_HALO_OCC")) : 0;
  const int occ_target = occ_env > 0 ? occ_env : (block_n <= 16 ? 3 : block_n <= 32 ? 2 : 1);
  for (int t = 8; t >= 2 && th == 0; t -= 2) {
    const int nblk = t / 2;
    if (2 * nblk * block_n > 512) continue;
    const int sb = ((t + 2) * kHaloPitch * rowb + 1023) & ~1023;
    for (int oc = occ_target; oc >= 1; --oc) {
      const int per_cta = (227 * 1024) / oc - 1024 /*per-CTA reservation*/ - 1024 /*align*/ - 1024 /*view overrun pad*/ - 256 /*barriers*/ - weights_bytes;
      int st = per_cta / sb;
      if (st > kHaloMaxStages) st = kHaloMaxStages;
      if (st >= 2) {
        th = t;
        stages = st;
        stage_bytes = sb;
        break;
      }
    }
  }
  if (th == 0) return MDE_ERR_UNSUPPORTED;

  HaloParams p;
  memset(&p, 0, sizeof(p));
  const bool zero = padding == MDE_PAD_ZERO;
  {
    const int H2 = h + 2, W2 = w + 2;
    const uint64_t strides[3] = {(uint64_t)in_cstride * 2, (uint64_t)W2 * in_cstride * 2, (uint64_t)H2 * W2 * in_cstride * 2};
    const uint32_t box[4] = {(uint32_t)kc, (uint32_t)kHaloPitch, (uint32_t)(th + 2), 1};
    int rc;
    if (zero) {  // view of the interior: out-of-bounds halo pixels read TMA zero fill
      const uint64_t dims[4] = {(uint64_t)cin, (uint64_t)w, (uint64_t)h, (uint64_t)B};
      rc = encode_tmap_bf16(&p.tmA, reinterpret_cast<const uint16_t*>(in) + ((size_t)W2 + 1) * in_cstride, 4, dims, strides, box, rowb);
    } else {
      const uint64_t dims[4] = {(uint64_t)cin, (uint64_t)W2, (uint64_t)H2, (uint64_t)B};
      rc = encode_tmap_bf16(&p.tmA, in, 4, dims, strides, box, rowb);
    }
    if (rc != MDE_OK) return rc;
  }
  {
    const int ktot = 9 * cin;
    const uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)cout_pad};
    const uint64_t strides[1] = {(uint64_t)ktot * 2};
    const uint32_t box[2] = {(uint32_t)kc, (uint32_t)block_n};
    int rc = encode_tmap_bf16(&p.tmB, wpacked, 2, dims, strides, box, rowb);
    if (rc != MDE_OK) return rc;
  }
  p.B = B;
  p.h = h;
  p.w = w;
  p.th = th;
  p.tiles_x = ceil_div(w, kHaloTW);
  p.tiles_y = ceil_div(h, th);
  p.total_tiles = B * p.tiles_x * p.tiles_y;
  p.chunks = chunks;
  p.nblk = th / 2;
  p.stages = stages;
  p.coord_off = zero ? -1 : 0;
  p.halo_bytes = (th + 2) * kHaloPitch * rowb;
  p.stage_bytes = stage_bytes;
  p.epi = out_bf16 ? EPI_CONV_BF16 : EPI_CONV_F32;
  p.act = activation;
  p.bias = bias;
  p.out = out_bf16 ? out_bf16 : (void*)out_f32;
  p.cout = cout;
  p.out_cstride = out_cstride;
  p.out_coff = out_coff;
  p.f16 = f16;
  static const int dbg = getenv("MDE_HALO_DBG") ? atoi(getenv("MDE_HALO_DBG")) : 0;
  p.dbg = dbg;
  p.st32 = out_bf16 && cout % 16 == 0 && ((uintptr_t)out_bf16 % 32 == 0) && (out_cstride * 2) % 32 == 0 && (out_coff * 2) % 32 == 0;
  const int smem_bytes = 1024 + weights_bytes + stages * stage_bytes + 1024 + 256;
#define MDE_HALO_CASE(BN, RB) \
  if (block_n == BN && rowb == RB) return launch_halo<BN, RB>(p, smem_bytes, stream);
  MDE_HALO_CASE(16, 128) MDE_HALO_CASE(32, 128) MDE_HALO_CASE(64, 128) MDE_HALO_CASE(128, 128)
  MDE_HALO_CASE(16, 64) MDE_HALO_CASE(32, 64) MDE_HALO_CASE(64, 64) MDE_HALO_CASE(128, 64)
  MDE_HALO_CASE(16, 32) MDE_HALO_CASE(32, 32) MDE_HALO_CASE(64, 32) MDE_HALO_CASE(128, 32)
#undef MDE_HALO_CASE
  return MDE_ERR_UNSUPPORTED;
}

}  // namespace mde
