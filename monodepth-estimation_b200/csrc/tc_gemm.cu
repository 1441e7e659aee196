// Tensor-core main loop of the path: ONE warp-specialised, persistent tcgen05 kernel that serves
//   * the ViT linear layers   D[M,N] = epi(A[M,K] W[N,K]^T)            (mde_gemm_bf16)
//   * the decoder/head convs  as implicit GEMM over padded NHWC          (mde_conv_nhwc)
// Warp roles (192 threads): warp 0 = TMA producer (one elected lane), warp 1 = TMEM allocator + MMA
// issuer (one elected lane), warps 2..5 = epilogue (TMEM -> registers -> fused epilogue -> global).
// Pipelines: smem ring full/empty mbarriers (TMA <-> MMA) and a 2-deep TMEM accumulator ring
// tmem_full/tmem_empty (MMA <-> epilogue), so the epilogue of tile i overlaps the MMAs of tile i+1.
//
// Operands are K-major bf16 tiles written by TMA with hardware swizzle (128/64/32-byte rows for
// 64/32/16 input channels per k-block); for convolutions the A tile of tap (dy,dx) is a 4-D TMA box
// [channels, tw, th, 1] of the padded NHWC activation shifted by the tap, so no im2col buffer exists
// and zero padding falls out of TMA out-of-bounds fill.
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "common.h"
#include "ptx.cuh"
#include "tc_epilogue.cuh"

namespace mde {

constexpr int kBlockM = 128;
constexpr int kThreads = 192;
constexpr int kMaxStages = 12;

struct alignas(64) TcParams {
  CUtensorMap tmA;
  CUtensorMap tmB;
  int M, N;
  int num_m_tiles, num_n_tiles, k_blocks, stages;
  // implicit-GEMM geometry
  int is_conv, ksize, chunks, coord_off;
  int B, h, w, tw, th, tiles_x, tiles_y;
  // epilogue
  int epi, act;
  const float* bias;
  void* out;
  int ldo;
  const void* aux0;
  const void* aux1;
  int cout, out_cstride, out_coff;
  int f16;   // operands (and bf16-slot outputs) are fp16 instead of bf16
  int st32;  // conv bf16/fp16 output: 32-byte stores are legal (alignment, cout % 16 == 0)
  // LayerNorm fusion (pair kernel only; always null here, the shared patch-embed epilogue refers to them)
  void* xb;
  float* ln_out;
  int ln_rows;
};

// co-resident CTAs per SM the register budget is compiled for (small tiles hide latency with more CTAs)
template <int BLOCK_N>
constexpr int min_ctas() {
  return BLOCK_N >= 128 ? 1 : (BLOCK_N == 64 ? 2 : (BLOCK_N == 32 ? 3 : 4));
}

template <int BLOCK_N, int ROWB>
__global__ void __launch_bounds__(kThreads, min_ctas<BLOCK_N>()) tc_gemm_kernel(const __grid_constant__ TcParams p) {
  pdl_launch_dependents();
  constexpr int KCHUNK = ROWB / 2;
  constexpr int A_BYTES = kBlockM * ROWB;
  constexpr int B_BYTES_RAW = BLOCK_N * ROWB;
  constexpr int B_BYTES = (B_BYTES_RAW + 1023) & ~1023;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = (2 * BLOCK_N < 32) ? 32 : 2 * BLOCK_N;
  constexpr int NCOL = BLOCK_N >= 32 ? 32 : 16;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.num_m_tiles * p.num_n_tiles;

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
        mbar_init(&tempty_bar[a], 4);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_grid_wait();  // the set-up above overlaps the previous kernel's tail

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const int per_img = p.tiles_x * p.tiles_y;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_blk = tile % p.num_n_tiles, m_blk = tile / p.num_n_tiles;
        int b = 0, y0 = 0, x0 = 0;
        if (p.is_conv) {
          b = m_blk / per_img;
          const int r = m_blk - b * per_img;
          y0 = (r / p.tiles_x) * p.th;
          x0 = (r % p.tiles_x) * p.tw;
        }
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          mbar_expect_tx(&full_bar[stage], A_BYTES + B_BYTES_RAW);
          if (!p.is_conv) {
            tma_load_2d(sa, &p.tmA, &full_bar[stage], kb * KCHUNK, m_blk * kBlockM);
          } else {
            const int tap = kb / p.chunks, cc = kb - tap * p.chunks;
            const int dy = p.ksize == 3 ? tap / 3 : 1, dx = p.ksize == 3 ? tap % 3 : 1;
            tma_load_4d(sa, &p.tmA, &full_bar[stage], cc * KCHUNK, x0 + dx + p.coord_off, y0 + dy + p.coord_off, b);
          }
          tma_load_2d(sb, &p.tmB, &full_bar[stage], kb * KCHUNK, n_blk * BLOCK_N);
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
      const uint32_t idesc = p.f16 ? umma_idesc_f16(kBlockM, BLOCK_N) : umma_idesc_bf16(kBlockM, BLOCK_N);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * STAGE_BYTES);
          const uint64_t da = umma_smem_desc_kmajor<ROWB>(sa);
          const uint64_t db = umma_smem_desc_kmajor<ROWB>(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < KCHUNK / 16; ++k)  // +32 bytes per UMMA_K step inside the swizzle span
            umma_bf16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tfull_bar[acc]);  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int q = warp & 3;  // TMEM lane quarter this warp can address
    const int r = q * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    const int per_img = p.tiles_x * p.tiles_y;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int n_blk = tile % p.num_n_tiles, m_blk = tile / p.num_n_tiles;
      RowInfo ri;
      ri.m = m_blk * kBlockM + r;
      ri.b = ri.y = ri.x = 0;
      if (p.is_conv) {
        ri.b = m_blk / per_img;
        const int t = m_blk - ri.b * per_img;
        ri.y = (t / p.tiles_x) * p.th + r / p.tw;
        ri.x = (t % p.tiles_x) * p.tw + r % p.tw;
        ri.valid = ri.y < p.h && ri.x < p.w;
      } else {
        ri.valid = ri.m < p.M;
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / NCOL; ++c) {
        uint32_t raw[NCOL];
        if constexpr (NCOL == 32) tmem_ld_32x32b_x32(taddr + c * NCOL, raw);
        else tmem_ld_32x32b_x16(taddr + c * NCOL, raw);
        tmem_ld_wait();
        float v[NCOL];
#pragma unroll
        for (int j = 0; j < NCOL; ++j) v[j] = __uint_as_float(raw[j]);
        epilogue_chunk<NCOL, TcParams>(p, v, n_blk * BLOCK_N + c * NCOL, ri);
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
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------ host side
template <int BLOCK_N, int ROWB>
static int launch_tc(TcParams& p, cudaStream_t stream) {
  constexpr int A_BYTES = kBlockM * ROWB;
  constexpr int B_BYTES = (BLOCK_N * ROWB + 1023) & ~1023;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // Large-N tiles own the SM (deep ring); small-N tiles (low-channel convs) leave room for several CTAs per SM
  // because their per-tile MMA time is tiny and latency hiding comes from co-resident CTAs.
  const int budget = BLOCK_N >= 128 ? 200 * 1024 : (220 * 1024) / min_ctas<BLOCK_N>() - 1024;
  int stages = (budget - 2048) / STAGE_BYTES;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) stages = 2;
  const int smem_bytes = stages * STAGE_BYTES + 1024 /*alignment slack*/ + (2 * kMaxStages + 8) * 8;
  static int ctas_per_sm = 0;  // per instantiation: what the hardware can actually keep resident
  if (ctas_per_sm == 0) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BLOCK_N, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BLOCK_N, ROWB>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    cudaFuncAttributes fa;
    MDE_CHECK_CUDA(cudaFuncGetAttributes(&fa, tc_gemm_kernel<BLOCK_N, ROWB>));
    // resident CTAs per SM from the four per-SM budgets: registers (64 K), shared memory (227 KB usable,
    // 1 KB reserved per CTA), threads (2048) and TMEM columns (512)
    const int regs_per_cta = ((fa.numRegs + 7) & ~7) * kThreads;
    const int tmem_cols = 2 * BLOCK_N < 32 ? 32 : 2 * BLOCK_N;
    int occ = 65536 / regs_per_cta;
    occ = std::min(occ, (227 * 1024) / (smem_bytes + 1024));
    occ = std::min(occ, 2048 / kThreads);
    occ = std::min(occ, 512 / tmem_cols);
    ctas_per_sm = occ < 1 ? 1 : occ;
  }
  p.stages = stages;
  const int total = p.num_m_tiles * p.num_n_tiles;
  int grid = sm_count() * ctas_per_sm;
  if (grid > total) grid = total;
  launch_k(tc_gemm_kernel<BLOCK_N, ROWB>, dim3(grid), dim3(kThreads), smem_bytes, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

static int dispatch_tc(TcParams& p, int block_n, int rowb, cudaStream_t stream) {
#define MDE_TC_CASE(BN, RB) \
  if (block_n == BN && rowb == RB) return launch_tc<BN, RB>(p, stream);
  MDE_TC_CASE(256, 128) MDE_TC_CASE(128, 128) MDE_TC_CASE(64, 128) MDE_TC_CASE(32, 128) MDE_TC_CASE(16, 128)
  MDE_TC_CASE(64, 64) MDE_TC_CASE(32, 64) MDE_TC_CASE(16, 64)
  MDE_TC_CASE(64, 32) MDE_TC_CASE(32, 32) MDE_TC_CASE(16, 32)
#undef MDE_TC_CASE
  set_error("tc kernel: no instantiation for BLOCK_N=%d row_bytes=%d", block_n, rowb);
  return MDE_ERR_UNSUPPORTED;
}

}  // namespace mde

using namespace mde;

static int gemm_impl(const mde_bf16* A, int lda, const mde_bf16* W, int ldw, int M, int N, int K, int epilogue, const float* bias, void* out, int ldo,
                     const void* aux0, const void* aux1, const LnFuse* ln, mde_stream_t stream) {
  MDE_CHECK_ARG(A && W && bias && out, "mde_gemm_bf16: null pointer");
  MDE_CHECK_ARG(M > 0 && N > 0 && K > 0, "mde_gemm_bf16: empty problem M=%d N=%d K=%d", M, N, K);
  MDE_CHECK_ARG(K % 64 == 0 && N % 128 == 0, "mde_gemm_bf16: need K %% 64 == 0 and N %% 128 == 0 (K=%d N=%d)", K, N);
  MDE_CHECK_ARG(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K, "mde_gemm_bf16: bad leading dims lda=%d ldw=%d", lda, ldw);
  MDE_CHECK_ARG(((uintptr_t)A | (uintptr_t)W | (uintptr_t)out | (uintptr_t)bias) % 16 == 0, "mde_gemm_bf16: pointers must be 16-byte aligned");
  MDE_CHECK_ARG(epilogue >= MDE_EPI_BIAS_BF16 && epilogue <= MDE_EPI_QKV_SPLIT, "mde_gemm_bf16: unknown epilogue %d", epilogue);
  if (epilogue == MDE_EPI_BIAS_RESID_F32) MDE_CHECK_ARG(aux0 != nullptr, "mde_gemm_bf16: RESID_F32 needs aux0 = residual");
  if (epilogue == MDE_EPI_PATCH_EMBED) {
    MDE_CHECK_ARG(aux0 && aux1, "mde_gemm_bf16: PATCH_EMBED needs aux0 = pos_embedding, aux1 = class_token");
    MDE_CHECK_ARG(N == MDE_VIT_DIM && M % 196 == 0, "mde_gemm_bf16: PATCH_EMBED expects N=768 and M = B*196");
  }
  if (epilogue == MDE_EPI_QKV_SPLIT) {
    MDE_CHECK_ARG(aux0 && aux1, "mde_gemm_bf16: QKV_SPLIT needs aux0 = k, aux1 = vt");
    MDE_CHECK_ARG(N == 3 * MDE_VIT_DIM && M % MDE_VIT_SEQ == 0, "mde_gemm_bf16: QKV_SPLIT expects N=2304 and M = B*197");
  }
  if (epilogue <= MDE_EPI_BIAS_F32) MDE_CHECK_ARG(ldo >= N && ldo % 8 == 0, "mde_gemm_bf16: bad ldo=%d", ldo);

  // large problems run on CTA pairs (tc_gemm2.cu); MDE_GEMM_1CTA=1 forces the single-CTA kernel (A/B testing)
  static const bool force_1cta = getenv("MDE_GEMM_1CTA") != nullptr && getenv("MDE_GEMM_1CTA")[0] == '1';
  if (!force_1cta && N % 256 == 0 && M >= MDE_GEMM_LN_MIN_ROWS) return launch_gemm_pair(A, lda, W, ldw, M, N, K, epilogue, bias, out, ldo, aux0, aux1, ln, stream);
  if (ln) {
    set_error("mde_gemm_ln_bf16: the LayerNorm-fused epilogues exist in the CTA-pair kernel only (M >= %d, N %% 256 == 0; M=%d N=%d)", MDE_GEMM_LN_MIN_ROWS, M, N);
    return MDE_ERR_UNSUPPORTED;
  }

  TcParams p;
  memset(&p, 0, sizeof(p));
  // Tile width: 256 (or 128) columns when that already gives every SM a tile; otherwise halve it down to 64 — at batch 4 the
  // N = 768 layers were 21 CTAs of 128 x 256 on 148 SMs, 31 us per launch and 44 % of the forward (profiles/r02e_launch_summary_b4.txt).
  // A is re-read once per column tile, from L2.  MDE_GEMM_BLOCK_N overrides (A/B timing).
  int block_n = (N % 256 == 0) ? 256 : 128;
  static const int bn_env = getenv("MDE_GEMM_BLOCK_N") ? atoi(getenv("MDE_GEMM_BLOCK_N")) : 0;
  if (bn_env > 0 && N % bn_env == 0) {
    block_n = bn_env;
  } else {
    const int want = (sm_count() * 4) / 5;
    while (block_n > 64 && N % (block_n / 2) == 0 && ceil_div(M, kBlockM) * (N / block_n) < want) block_n /= 2;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
    const uint64_t strides[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, (uint32_t)kBlockM};
    int rc = encode_tmap_bf16(&p.tmA, A, 2, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
    const uint64_t strides[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {64, (uint32_t)block_n};
    int rc = encode_tmap_bf16(&p.tmB, W, 2, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
  }
  p.M = M;
  p.N = N;
  p.num_m_tiles = ceil_div(M, kBlockM);
  p.num_n_tiles = N / block_n;
  p.k_blocks = K / 64;
  p.is_conv = 0;
  p.tiles_x = p.tiles_y = 1;
  p.tw = 128;
  p.th = 1;
  p.epi = epilogue;
  p.bias = bias;
  p.out = out;
  p.ldo = ldo;
  p.aux0 = aux0;
  p.aux1 = aux1;
  return dispatch_tc(p, block_n, 128, stream);
}

extern "C" int mde_gemm_bf16(const mde_bf16* A, int lda, const mde_bf16* W, int ldw, int M, int N, int K, int epilogue,
                             const float* bias, void* out, int ldo, const void* aux0, const void* aux1,
                             mde_stream_t stream) {
  return gemm_impl(A, lda, W, ldw, M, N, K, epilogue, bias, out, ldo, aux0, aux1, nullptr, stream);
}

extern "C" int mde_gemm_ln_bf16(const mde_bf16* A, int lda, const mde_bf16* W, int ldw, int M, int N, int K, int epilogue,
                                const float* bias, void* out, int ldo, const void* aux0, const void* aux1, const float* ln_stats_in,
                                const float* ln_colsum, float ln_eps, mde_bf16* xb_out, float* ln_stats_out, mde_stream_t stream) {
  LnFuse ln;
  ln.stats_in = ln_stats_in;
  ln.colsum = ln_colsum;
  ln.eps = ln_eps;
  ln.xb_out = xb_out;
  ln.stats_out = ln_stats_out;
  const bool any = ln_stats_in || ln_colsum || xb_out || ln_stats_out;
  return gemm_impl(A, lda, W, ldw, M, N, K, epilogue, bias, out, ldo, aux0, aux1, any ? &ln : nullptr, stream);
}

static int conv_nhwc_impl(int f16, const mde_bf16* in, int B, int h, int w, int cin, int in_cstride, const mde_bf16* wpacked, const float* bias,
                             int cout, int ksize, int padding, int activation, mde_bf16* out_bf16, int out_cstride,
                             int out_coff, float* out_f32, mde_stream_t stream) {
  MDE_CHECK_ARG(in && wpacked && bias, "mde_conv_nhwc: null pointer");
  MDE_CHECK_ARG((out_bf16 != nullptr) != (out_f32 != nullptr), "mde_conv_nhwc: exactly one of out_bf16 / out_f32 must be given");
  MDE_CHECK_ARG(B > 0 && h >= 3 && w >= 3, "mde_conv_nhwc: bad geometry B=%d h=%d w=%d", B, h, w);
  MDE_CHECK_ARG(ksize == 3 || ksize == 1, "mde_conv_nhwc: ksize must be 1 or 3");
  MDE_CHECK_ARG(padding == MDE_PAD_REFLECT || padding == MDE_PAD_ZERO, "mde_conv_nhwc: bad padding mode");
  MDE_CHECK_ARG(cout >= 1, "mde_conv_nhwc: cout");
  MDE_CHECK_ARG(in_cstride >= cin && in_cstride % 8 == 0, "mde_conv_nhwc: in_cstride=%d must be >= cin and a multiple of 8", in_cstride);
  int rowb;
  if (cin % 64 == 0) rowb = 128;
  else if (cin == 32) rowb = 64;
  else if (cin == 16) rowb = 32;
  else {
    set_error("mde_conv_nhwc: cin=%d not supported (multiple of 64, or 32, or 16)", cin);
    return MDE_ERR_UNSUPPORTED;
  }
  const int kchunk = rowb / 2;
  const int cout_pad = (cout + 15) & ~15;
  if (out_bf16) {
    MDE_CHECK_ARG(cout % 8 == 0 && out_cstride % 8 == 0 && out_coff % 8 == 0 && out_coff + cout <= out_cstride,
                  "mde_conv_nhwc: bf16 output needs cout/out_cstride/out_coff multiples of 8 (cout=%d cstride=%d coff=%d)", cout,
                  out_cstride, out_coff);
  }
  MDE_CHECK_ARG(((uintptr_t)in | (uintptr_t)wpacked | (uintptr_t)bias) % 16 == 0, "mde_conv_nhwc: pointers must be 16-byte aligned");

  if (ksize == 3) {  // wide maps: halo-resident kernel (conv_halo.cu); it declines geometries it does not serve
    const int rc = launch_conv_halo(f16, in, B, h, w, cin, in_cstride, wpacked, bias, cout, padding, activation, out_bf16, out_cstride, out_coff,
                                    out_f32, stream);
    if (rc != MDE_ERR_UNSUPPORTED) return rc;
    if (out_bf16) {  // small maps, wide channels: halo tile + streamed weights (conv_small.cu)
      const int rc2 = launch_conv_small(f16, in, B, h, w, cin, in_cstride, wpacked, bias, cout, padding, activation, out_bf16, out_cstride, out_coff, stream);
      if (rc2 != MDE_ERR_UNSUPPORTED) return rc2;
    }
  }

  // spatial tile: 128 output pixels = th rows x tw columns
  int tw = 16;
  while (tw < w && tw < 128) tw <<= 1;
  const int th = 128 / tw;

  TcParams p;
  memset(&p, 0, sizeof(p));
  const bool zero = (padding == MDE_PAD_ZERO) && ksize == 3;
  {
    const int H2 = h + 2, W2 = w + 2;
    const uint64_t strides[3] = {(uint64_t)in_cstride * 2, (uint64_t)W2 * in_cstride * 2, (uint64_t)H2 * W2 * in_cstride * 2};
    const uint32_t box[4] = {(uint32_t)kchunk, (uint32_t)tw, (uint32_t)th, 1};
    int rc;
    if (zero) {  // view of the interior: out-of-bounds taps read TMA zero fill
      const uint64_t dims[4] = {(uint64_t)cin, (uint64_t)w, (uint64_t)h, (uint64_t)B};
      rc = encode_tmap_bf16(&p.tmA, in + ((size_t)W2 + 1) * in_cstride, 4, dims, strides, box, rowb);
    } else {
      const uint64_t dims[4] = {(uint64_t)cin, (uint64_t)W2, (uint64_t)H2, (uint64_t)B};
      rc = encode_tmap_bf16(&p.tmA, in, 4, dims, strides, box, rowb);
    }
    if (rc != MDE_OK) return rc;
  }
  // block_n: largest tile that divides cout_pad, shrunk while the grid would leave SMs idle
  int block_n = 16;
  for (int bn = (rowb == 128 ? 256 : 64); bn >= 16; bn >>= 1)  // narrow-K tiles (16/32 input channels) are built for N <= 64
    if (cout_pad % bn == 0) {
      block_n = bn;
      break;
    }
  const int m_tiles = B * ceil_div(h, th) * ceil_div(w, tw);
  while (block_n > 64 && m_tiles * (cout_pad / block_n) < sm_count()) block_n >>= 1;  // (N = 32 tiles measured slower: r01r)
  const int ktot = ksize * ksize * cin;
  {
    const uint64_t dims[2] = {(uint64_t)ktot, (uint64_t)cout_pad};
    const uint64_t strides[1] = {(uint64_t)ktot * 2};
    const uint32_t box[2] = {(uint32_t)kchunk, (uint32_t)block_n};
    int rc = encode_tmap_bf16(&p.tmB, wpacked, 2, dims, strides, box, rowb);
    if (rc != MDE_OK) return rc;
  }
  p.M = m_tiles * kBlockM;
  p.N = cout_pad;
  p.num_m_tiles = m_tiles;
  p.num_n_tiles = cout_pad / block_n;
  p.chunks = cin / kchunk;
  p.k_blocks = ksize * ksize * p.chunks;
  p.is_conv = 1;
  p.ksize = ksize;
  p.coord_off = zero ? -1 : 0;
  p.B = B;
  p.h = h;
  p.w = w;
  p.tw = tw;
  p.th = th;
  p.tiles_x = ceil_div(w, tw);
  p.tiles_y = ceil_div(h, th);
  p.epi = out_bf16 ? EPI_CONV_BF16 : EPI_CONV_F32;
  p.act = activation;
  p.bias = bias;
  p.out = out_bf16 ? (void*)out_bf16 : (void*)out_f32;
  p.cout = cout;
  p.out_cstride = out_cstride;
  p.out_coff = out_coff;
  p.f16 = f16;
  p.st32 = out_bf16 && cout % 16 == 0 && ((uintptr_t)out_bf16 % 32 == 0) && (out_cstride * 2) % 32 == 0 && (out_coff * 2) % 32 == 0;
  return dispatch_tc(p, block_n, rowb, stream);
}

extern "C" int mde_conv_nhwc(const mde_bf16* in, int B, int h, int w, int cin, int in_cstride, const mde_bf16* wpacked, const float* bias,
                             int cout, int ksize, int padding, int activation, mde_bf16* out_bf16, int out_cstride,
                             int out_coff, float* out_f32, mde_stream_t stream) {
  return conv_nhwc_impl(0, in, B, h, w, cin, in_cstride, wpacked, bias, cout, ksize, padding, activation, out_bf16, out_cstride, out_coff, out_f32, stream);
}

extern "C" int mde_conv_nhwc_f16(const mde_f16* in, int B, int h, int w, int cin, int in_cstride, const mde_f16* wpacked, const float* bias,
                                 int cout, int ksize, int padding, int activation, mde_f16* out_f16, int out_cstride,
                                 int out_coff, float* out_f32, mde_stream_t stream) {
  return conv_nhwc_impl(1, in, B, h, w, cin, in_cstride, wpacked, bias, cout, ksize, padding, activation, out_f16, out_cstride, out_coff, out_f32, stream);
}
