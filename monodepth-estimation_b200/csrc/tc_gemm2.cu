// 2-CTA (cta_group::2) tcgen05 GEMM for the large ViT linear layers.
//
// Why a pair kernel: with one CTA per 128x256 tile every k-block pulls 48 KB from L2 per SM, i.e. 85 FLOP
// per L2 byte, and the measured ~8 TB/s of L2->SM bandwidth caps the GEMM near 0.65 PFLOP/s (r01a profile).
// A CTA pair computes a 256x256 tile with ONE tcgen05.mma.cta_group::2 per K=16 step: each CTA loads only
// its own 128 rows of A and HALF of B (32 KB per k-block, 128 FLOP/byte), the tensor cores of both SMs read
// the two B halves from both shared memories, and each CTA keeps its 128 accumulator rows in its own TMEM.
//
// Schedule: persistent clusters over a static item list.  Items are 256x256 tiles; the tiles left over after
// the last full wave (tiles mod clusters) are split into four 256x64 items so the tail costs a quarter of a
// tile time instead of a whole one (M = B*197 = 6304 gives 75 / 225 / 300 tiles for 74 clusters).
//
// Warp roles per CTA (320 threads): warp 0 TMA producer (loads its CTA's A rows and B half, credits the
// leader's full barrier), warp 1 TMEM allocator + (leader only) MMA issuer, warps 2..9 epilogue: two warps
// per TMEM lane quarter, each taking every other 128-byte output chunk.
//
// Epilogue data path (bias | bias+GELU | bias+residual, i.e. every ViT linear layer except QKV / patch embed): a thread owns
// one accumulator row, so direct global stores would touch 32 different cache lines per warp instruction — measured at
// ~20 us per 256x256 tile against 3 us of MMA time (r01e).  Instead each epilogue warp stages its 32-row x 128-byte
// sub-tile in a private, 128B-swizzled shared-memory buffer and moves it with TMA: the residual arrives by TMA LOAD
// one chunk ahead, bias/activation/residual are applied in place, and a TMA STORE (bulk group) writes it out while
// the warp is already on its next chunk.  No LSU global traffic, full-line writes, rows >= M clipped by the tensor map.
//
// Residual layers (out-proj, fc2): the residual and the bias are loaded INTO the accumulator before the tile's MMAs start
// (TMA load -> shared -> tcgen05.st by the epilogue warps, while the previous tile's MMAs run; the first MMA then accumulates),
// so the epilogue of such a tile is TMEM -> shared -> TMA store only.  With the residual fetched inside the epilogue every
// 32-column chunk waited ~1 us for its load: 5.3 us per tile on the un-overlapped tail of proj / fc2 (in-kernel counters, r02b).
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "common.h"
#include "ptx.cuh"
#include "tc_epilogue.cuh"

namespace mde {

constexpr int k2Threads = 320;
constexpr int k2MaxStages = 8;
constexpr int k2ABytes = 128 * 128;       // 128 rows x 64 bf16
constexpr int k2BBytes = 128 * 128;       // B half of a wide item (128 rows); narrow items use the first 4 KB
constexpr int k2StageBytes = k2ABytes + k2BBytes;
constexpr uint32_t k2TmemCols = 512;      // 2 accumulator buffers x 256 columns
constexpr int k2EpiWarps = 8;
constexpr int k2EpiBufBytes = 32 * 128;   // one epilogue warp's staging buffer: 32 rows x 128 bytes (SWIZZLE_128B)
constexpr int k2StagingBytes = k2EpiWarps * 2 * k2EpiBufBytes;
constexpr int k2BarrierBytes = 512;

struct alignas(64) Tc2Params {
  CUtensorMap tmA;   // box [64, 128]
  CUtensorMap tmB;   // box [64, 128]: wide items (pair N = 256)
  CUtensorMap tmBn;  // box [64, 32] : narrow items (pair N = 64)
  CUtensorMap tmOut; // box [128 bytes of columns, 32 rows] over the output (TMA-store epilogue)
  CUtensorMap tmRes; // same geometry over the residual input
  int tma_epi;       // 1: TMA-store epilogue (out_elt = 4: fp32, 2: bf16); 0: direct-store epilogue (QKV split, patch embed)
  int out_elt;
  int M, N;
  int k_blocks, stages;
  int n_wide;        // N / 256
  int full_items;    // items [0, full_items) are whole 256x256 tiles
  int total_items;   // full_items + 4 * (tiles - full_items)
  // epilogue (field names shared with tc_epilogue.cuh)
  int epi, act;
  const float* bias;
  void* out;
  int ldo;
  const void* aux0;
  const void* aux1;
  int cout, h, w, out_cstride, out_coff, f16, st32;
  // LayerNorm fusion (see LnFuse in common.h)
  void* xb;             // producer: bf16 copy of the fp32 output, row pitch xb_ld
  float* ln_out;        // producer: [N/256][8][M][2] partial (sum, sum of squares), slot-major (coalesced); see ln_slot()
  int ln_rows, xb_ld;
  const float* ln_in;   // consumer: the producer's buffer, [ln_ktiles = K/256][8][M][2]
  const float* ln_cs;   // consumer: [N] column sums of the folded weight
  int ln_ktiles;        // consumer: the producer's N / 256
  int ln_full;          // consumer: the producer's full_items (its tiles >= ln_full were split into four 64-column items)
  float ln_eps, ln_invk;
  long long* prof;      // debug (mde_debug_gemm_prof): per CTA 8 counters of cycles spent waiting / working, or null
};

// In-kernel counters (mde_debug_gemm_prof) exist only in builds with -DMDE_GEMM_PROF (MDE_BUILD_PROF=1 python build.py):
// even a never-taken branch per k-block in the MMA-issuing thread costs ~0.7 us per launch.
#ifdef MDE_GEMM_PROF
#define MDE_PROF(...) __VA_ARGS__
#else
#define MDE_PROF(...)
#endif
__device__ __forceinline__ long long clk() { return clock64(); }

struct Item {
  int m0, n0, bn;
};
__device__ __forceinline__ Item decode_item(const Tc2Params& p, int i) {
  Item it;
  int tile, q = 0;
  if (i < p.full_items) {
    tile = i;
    it.bn = 256;
  } else {
    const int j = i - p.full_items;
    tile = p.full_items + (j >> 2);
    q = j & 3;
    it.bn = 64;
  }
  it.m0 = (tile / p.n_wide) * 256;
  it.n0 = (tile % p.n_wide) * 256 + q * 64;
  return it;
}

// Bias values of one epilogue warp's columns, one column per lane: the warp owns the chunks par, par + 2, ... of the item (`my` of
// them); b[jj] = bias of column `lane` of its jj-th 32-column chunk (fp32 epilogues), or b[2 jj + h] = column 32 h + lane of its
// jj-th 64-column chunk (bf16 epilogues).  Four coalesced loads per item instead of 8-16 float4 loads per chunk and lane.
__device__ __forceinline__ void lane_bias32(const float* bias, int n0, int par, int my, int lane, float (&b)[4]) {
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) b[jj] = jj < my ? __ldg(bias + n0 + (par + 2 * jj) * 32 + lane) : 0.0f;
}
__device__ __forceinline__ void lane_bias64(const float* bias, int n0, int par, int my, int lane, float (&b)[4]) {
#pragma unroll
  for (int jj = 0; jj < 2; ++jj) {
    b[2 * jj] = jj < my ? __ldg(bias + n0 + (par + 2 * jj) * 64 + lane) : 0.0f;
    b[2 * jj + 1] = jj < my ? __ldg(bias + n0 + (par + 2 * jj) * 64 + 32 + lane) : 0.0f;
  }
}

// LNF = 1 adds the LayerNorm-fusion code (producer: bf16 copy + row sums, consumer: normalising epilogue); the default
// instantiation carries none of it (the extra registers / instructions cost ~0.6 us per launch even when never executed).
// EC = epilogue class: 0 fp32 through TMA (bias / bias + residual), 1 bf16 through TMA (bias / GELU / QKV split), 2 direct stores
// (patch embed with the LayerNorm-fusion producer), 3 patch embed through TMA (fp32 + position embedding, token rows b*197 + 1 + patch
// addressed through a 3-D tensor map whose bounds clip a 32-row chunk at the end of its image): each instantiation carries only its own
// epilogue code.
template <int LNF, int EC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(k2Threads, 1) tc_gemm2_kernel(const __grid_constant__ Tc2Params p) {
  pdl_launch_dependents();
  MDE_PROF(const long long t_start = clk();)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int stages = p.stages;
  uint8_t* staging = smem + (size_t)stages * k2StageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + k2StagingBytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* res_bar = tempty_bar + 2;  // [k2EpiWarps][4] residual-chunk arrival barriers (2 staging buffers + 2 borrowed slots)
  uint64_t* init_bar = res_bar + 4 * k2EpiWarps;  // the epilogue warps no longer need the pipeline stages they borrowed (count 8)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(init_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmBn);
    if (EC != 2) {
      tma_prefetch_desc(&p.tmOut);
      tma_prefetch_desc(&p.tmRes);
    }
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);   // leader's expect_tx arrive; bytes from both CTAs' TMA
        mbar_init(&empty_bar[s], 1);  // pair commit (multicast) from the leader's MMA thread
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tfull_bar[a], 1);    // pair commit (multicast)
        mbar_init(&tempty_bar[a], 16);  // 8 epilogue warps x 2 CTAs (only the leader's copy is waited on)
      }
      for (int a = 0; a < 4 * k2EpiWarps; ++a) mbar_init(&res_bar[a], 1);
      mbar_init(init_bar, k2EpiWarps);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, k2TmemCols);
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_grid_wait();  // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail
  MDE_PROF(if (p.prof && threadIdx.x == 0) p.prof[blockIdx.x * 8 + 6] = clk() - t_start;)  // prologue incl. the wait for the previous kernel

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      bool borrow = EC == 0 && p.epi == MDE_EPI_BIAS_RESID_F32 && stages >= 4;
      for (int i = cluster_id; i < p.total_items; i += num_clusters) {
        const Item it = decode_item(p, i);
        const int half = it.bn >> 1;
        const CUtensorMap* tb = it.bn == 256 ? &p.tmB : &p.tmBn;
        const uint32_t tx = 2u * (k2ABytes + half * 128);
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (borrow && stage == stages - 2) {  // first pass only: stages - 2 and stages - 1 hold residual chunks of the first item
            mbar_wait(init_bar, 0);
            borrow = false;
          }
          uint8_t* sa = smem + (size_t)stage * k2StageBytes;
          if (leader) mbar_expect_tx(&full_bar[stage], tx);
          tma_load_2d_pair(sa, &p.tmA, &full_bar[stage], kb * 64, it.m0 + (int)rank * 128);
          tma_load_2d_pair(sa + k2ABytes, tb, &full_bar[stage], kb * 64, it.n0 + (int)rank * half);
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader && elect_one()) {
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      const bool pre_acc = EC == 0 && p.epi == MDE_EPI_BIAS_RESID_F32;  // accumulator pre-loaded with residual + bias
      for (int i = cluster_id; i < p.total_items; i += num_clusters) {
        const Item it = decode_item(p, i);
        const uint32_t idesc = umma_idesc_bf16(256, it.bn);
        MDE_PROF(long long t0 = 0; if (p.prof) t0 = clk();)
        mbar_wait(&tempty_bar[acc], pre_acc ? acc_phase : (acc_phase ^ 1));  // pre_acc: the epilogue warps arrive after initialising the buffer
        MDE_PROF(if (p.prof) p.prof[blockIdx.x * 8 + 0] += clk() - t0;)  // MMA thread waiting for a free accumulator (epilogue behind)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          MDE_PROF(if (p.prof) t0 = clk();)
          mbar_wait(&full_bar[stage], phase);
          MDE_PROF(if (p.prof) p.prof[blockIdx.x * 8 + 1] += clk() - t0;)  // MMA thread waiting for operands (TMA behind)
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * k2StageBytes);
          const uint64_t da = umma_smem_desc_kmajor<128>(sa);
          const uint64_t db = umma_smem_desc_kmajor<128>(sa + k2ABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, pre_acc || (kb | k) != 0);
          umma_commit_pair(&empty_bar[stage]);
          if (++stage == stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(&tfull_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      MDE_PROF(if (p.prof) p.prof[blockIdx.x * 8 + 7] = clk() - t_start;)  // last MMA issued (since kernel start)
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs)
    const int q = warp & 3;          // TMEM lane quarter
    const int par = (warp - 2) >> 2; // which half of the output chunks this warp takes
    const int r = q * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    if constexpr (EC != 2) {
      const int ew = warp - 2;
      uint8_t* stg = staging + ew * 2 * k2EpiBufBytes;
      uint64_t* rbar = res_bar + ew * 4;
      const bool resid = EC == 0 && p.epi == MDE_EPI_BIAS_RESID_F32;
      constexpr bool f32out = EC == 0 || EC == 3;
      const int cw = f32out ? 32 : 64;  // columns per 128-byte chunk
      const uint32_t sw = static_cast<uint32_t>(lane & 7);
      uint32_t g = 0;   // output chunks this warp has staged: staging buffer = g & 1
      uint32_t rph = 0;  // bit b: parity of the next completion of residual barrier b (staging buffer b) to wait for
      // Residual layers: (residual + bias) of item `it2` -> accumulator buffer `acc2`, then arrive on tempty (the MMA thread
      // starts the item when all 16 epilogue warps of the pair have).  Two TMA loads in flight per warp.
      auto init_acc = [&](const Item& it2, int acc2) {
        const int my = ((it2.bn >> 5) - par + 1) >> 1;  // this warp's 32-column chunks: par, par + 2, ...
        const int row02 = it2.m0 + (int)rank * 128 + q * 32;
        const uint32_t ta = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc2 * 256;
        auto load = [&](int j) {  // lane 0: chunk j of this warp -> staging buffer j & 1
          uint64_t* b = &rbar[j & 1];
          mbar_expect_tx(b, k2EpiBufBytes);
          tma_load_2d(stg + (j & 1) * k2EpiBufBytes, &p.tmRes, b, it2.n0 + (par + 2 * j) * 32, row02);
        };
        if (lane == 0) {
          bulk_wait_read<0>();  // the stores of the previous drain no longer read the staging buffers
          if (my > 0) load(0);
          if (my > 1) load(1);
        }
        float bq[4];
        lane_bias32(p.bias, it2.n0, par, my, lane, bq);  // in flight together with the residual loads
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j >= my) break;
          const int c = par + 2 * j, b = j & 1;
          mbar_wait(&rbar[b], (rph >> b) & 1);
          rph ^= 1u << b;
          const uint8_t* row = stg + b * k2EpiBufBytes + lane * 128;
          uint32_t v[32];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const float4 r4 = *reinterpret_cast<const float4*>(row + ((static_cast<uint32_t>(u) ^ sw) << 4));
            v[4 * u] = __float_as_uint(r4.x + __shfl_sync(0xffffffffu, bq[j], 4 * u));
            v[4 * u + 1] = __float_as_uint(r4.y + __shfl_sync(0xffffffffu, bq[j], 4 * u + 1));
            v[4 * u + 2] = __float_as_uint(r4.z + __shfl_sync(0xffffffffu, bq[j], 4 * u + 2));
            v[4 * u + 3] = __float_as_uint(r4.w + __shfl_sync(0xffffffffu, bq[j], 4 * u + 3));
          }
          tmem_st_32x32b_x32(ta + c * 32, v);
          __syncwarp();  // every lane has read the buffer
          if (lane == 0 && j + 2 < my) load(j + 2);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc2], 0);
      };
      // The first item's residual: all (up to four) chunks of the warp in flight at once — two in the warp's staging buffers, two in
      // 4 KB slots of pipeline stages `stages - 2` and `stages - 1`, which the producer leaves alone until init_bar completes
      // (the MMAs cannot start before this buffer is initialised anyway).  One TMA round trip instead of two on the critical path.
      auto init_first = [&](const Item& it2) {
        const int my = ((it2.bn >> 5) - par + 1) >> 1;
        const int row02 = it2.m0 + (int)rank * 128 + q * 32;
        const uint32_t ta = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        uint8_t* bufs[4] = {stg, stg + k2EpiBufBytes, smem + (size_t)(stages - 2) * k2StageBytes + ew * k2EpiBufBytes,
                            smem + (size_t)(stages - 1) * k2StageBytes + ew * k2EpiBufBytes};
        if (lane == 0) {
          for (int j = 0; j < my; ++j) {
            // chunks 0, 1 -> staging buffers (barriers 0, 1), chunks 2, 3 -> borrowed slots (barriers 2, 3)
            mbar_expect_tx(&rbar[j], k2EpiBufBytes);
            tma_load_2d(bufs[j], &p.tmRes, &rbar[j], it2.n0 + (par + 2 * j) * 32, row02);
          }
        }
        float bq[4];
        lane_bias32(p.bias, it2.n0, par, my, lane, bq);  // in flight together with the residual loads
        __syncwarp();
        // consume the borrowed slots first, so that the producer gets its stages back as early as possible
#pragma unroll
        for (int o = 0; o < 4; ++o) {
          const int j = (o + 2) & 3;  // 2, 3, 0, 1
          if (j < my) {
            const int c = par + 2 * j;
            mbar_wait(&rbar[j], 0);
            if (j < 2) rph ^= 1u << j;
            const uint8_t* row = bufs[j] + lane * 128;
            uint32_t v[32];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const float4 r4 = *reinterpret_cast<const float4*>(row + ((static_cast<uint32_t>(u) ^ sw) << 4));
              v[4 * u] = __float_as_uint(r4.x + __shfl_sync(0xffffffffu, bq[j], 4 * u));
              v[4 * u + 1] = __float_as_uint(r4.y + __shfl_sync(0xffffffffu, bq[j], 4 * u + 1));
              v[4 * u + 2] = __float_as_uint(r4.z + __shfl_sync(0xffffffffu, bq[j], 4 * u + 2));
              v[4 * u + 3] = __float_as_uint(r4.w + __shfl_sync(0xffffffffu, bq[j], 4 * u + 3));
            }
            tmem_st_32x32b_x32(ta + c * 32, v);
          }
          if (o == 1) {  // both borrowed slots have been read (or were never used)
            __syncwarp();
            if (lane == 0) mbar_arrive(init_bar);
          }
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&tempty_bar[0], 0);
      };
      int i = cluster_id;
      Item it = decode_item(p, i < p.total_items ? i : 0);
      if (resid) {  // the first two items' buffers (the MMA thread waits for them)
        if (i < p.total_items) {
          if (stages >= 4) init_first(it);
          else init_acc(it, 0);
        }
        if (i + num_clusters < p.total_items) init_acc(decode_item(p, i + num_clusters), 1);
      }
      while (i < p.total_items) {
        const int inext = i + num_clusters;
        const bool has_next = inext < p.total_items;
        const Item itn = decode_item(p, has_next ? inext : i);
        const int nch = it.bn / cw;
        const int row0 = it.m0 + (int)rank * 128 + q * 32;
        [[maybe_unused]] const float4* pe_row = nullptr;  // EC == 3: this thread's position-embedding row
        [[maybe_unused]] int pe_patch = -1, pe_img = 0;
        if constexpr (EC == 3) {
          const int m = min(row0 + lane, p.M - 1);  // rows past M are computed on a clamped row and clipped by the store
          pe_img = m / 196;
          pe_patch = row0 + lane < p.M ? m - pe_img * 196 : -1;
          pe_row = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux0) + (size_t)(1 + m - pe_img * 196) * MDE_VIT_DIM);
        }
        // LayerNorm of this thread's A row from the producer's partial sums.  Slot layout per 256-column producer tile: a whole
        // tile leaves two partials (epilogue-warp parity 0 / 1 = even / odd 32-column chunks, each summed chunk by chunk), a
        // tile of the split tail leaves one partial per chunk (slot = chunk).  Both are combined in the SAME order here —
        // ((c0 + c2) + c4) + c6 plus ((c1 + c3) + c5) + c7 — so the statistics do not depend on how the producer was tiled.
        float mu = 0.0f, rstd = 1.0f;
        if (LNF && EC == 1 && p.ln_in && row0 + lane < p.M) {
          const float2* st = reinterpret_cast<const float2*>(p.ln_in) + (row0 + lane);
          const int mt = it.m0 >> 8;
          float2 a0[4], a1[4];
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {  // host checked ln_ktiles <= 4; all loads in flight together
            a0[nt] = a1[nt] = make_float2(0.0f, 0.0f);
            if (nt < p.ln_ktiles) {
              a0[nt] = __ldg(st + (size_t)(nt * 8) * p.M);
              a1[nt] = __ldg(st + (size_t)(nt * 8 + 1) * p.M);
            }
          }
          float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {
            if (nt >= p.ln_ktiles) break;
            float2 w0 = a0[nt], w1 = a1[nt];
            if (mt * p.ln_ktiles + nt >= p.ln_full) {  // split tile (uniform per warp, at most one tile row of the producer)
              float2 e[6];
#pragma unroll
              for (int u = 0; u < 6; ++u) e[u] = __ldg(st + (size_t)(nt * 8 + 2 + u) * p.M);
              w0.x = ((w0.x + e[0].x) + e[2].x) + e[4].x;
              w0.y = ((w0.y + e[0].y) + e[2].y) + e[4].y;
              w1.x = ((w1.x + e[1].x) + e[3].x) + e[5].x;
              w1.y = ((w1.y + e[1].y) + e[3].y) + e[5].y;
            }
            s1 += w0.x + w1.x;
            s2 += w0.y + w1.y;
          }
          mu = s1 * p.ln_invk;
          rstd = rsqrtf(fmaxf(fmaf(-mu, mu, s2 * p.ln_invk), 0.0f) + p.ln_eps);
        }
        const float nmr = -mu * rstd;
        MDE_PROF(long long t0 = 0; const bool prof = p.prof && warp == 2 && lane == 0; if (prof) t0 = clk();)
        // this warp's bias values, one column per lane and 32-column group (shuffled to the row owners below): loaded once per item
        // and BEFORE the wait, so that no chunk pays an L2 round trip for them (r02c: ~0.5 us per chunk, serialised behind the waits)
        float bq[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (!resid) {
          if (f32out) lane_bias32(p.bias, it.n0, par, (nch - par + 1) >> 1, lane, bq);
          else lane_bias64(p.bias, it.n0, par, (nch - par + 1) >> 1, lane, bq);
        }
        mbar_wait(&tfull_bar[acc], acc_phase);
        MDE_PROF(if (prof) {
          const long long t1 = clk();
          p.prof[blockIdx.x * 8 + 2] += t1 - t0;  // epilogue warp 2 waiting for the accumulator (MMA behind)
          t0 = t1;
        })
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 256;
        float ls1 = 0.0f, ls2 = 0.0f;  // LayerNorm producer: this row's (sum, sum of squares) over the chunks this warp handles
#pragma unroll 1
        for (int c = par; c < nch; c += 2, ++g) {
          uint8_t* buf = stg + (g & 1) * k2EpiBufBytes;
          uint8_t* row = buf + lane * 128;
          const int n0 = it.n0 + c * cw;
          if constexpr (f32out) {
            uint32_t raw[32];
            tmem_ld_32x32b_x32(taddr + c * 32, raw);
            if (lane == 0) bulk_wait_read<1>();  // this buffer's previous store (chunk g-2) has finished reading it
            __syncwarp();
            tmem_ld_wait();
            const int jj = (c - par) >> 1;
            const float bsel = jj == 0 ? bq[0] : (jj == 1 ? bq[1] : (jj == 2 ? bq[2] : bq[3]));
            float c1 = 0.0f, c2 = 0.0f;  // this chunk's partial, started from zero (see the consumer: tiling-independent sums)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float4* sp = reinterpret_cast<float4*>(row + ((static_cast<uint32_t>(j) ^ sw) << 4));
              float4 o = make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]), __uint_as_float(raw[4 * j + 2]),
                                     __uint_as_float(raw[4 * j + 3]));
              if (!resid) {  // residual layers: bias and residual are already in the accumulator (init_acc)
                o.x += __shfl_sync(0xffffffffu, bsel, 4 * j);
                o.y += __shfl_sync(0xffffffffu, bsel, 4 * j + 1);
                o.z += __shfl_sync(0xffffffffu, bsel, 4 * j + 2);
                o.w += __shfl_sync(0xffffffffu, bsel, 4 * j + 3);
              }
              if constexpr (EC == 3) {  // + pos_embedding[1 + patch] (tv:156); a parameter, 128 contiguous bytes per row and chunk
                const float4 q4 = __ldg(pe_row + (n0 >> 2) + j);
                o.x += q4.x; o.y += q4.y; o.z += q4.z; o.w += q4.w;
              }
              *sp = o;
              if constexpr (LNF) {
                c1 += (o.x + o.y) + (o.z + o.w);
                c2 = fmaf(o.x, o.x, fmaf(o.y, o.y, fmaf(o.z, o.z, fmaf(o.w, o.w, c2))));
              }
            }
            if constexpr (LNF) {
              ls1 += c1;
              ls2 += c2;
            }
          } else {
            uint32_t raw[64];
            tmem_ld_32x32b_x32(taddr + c * 64, *reinterpret_cast<uint32_t(*)[32]>(&raw[0]));
            tmem_ld_32x32b_x32(taddr + c * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&raw[32]));
            if (lane == 0) bulk_wait_read<1>();
            __syncwarp();
            tmem_ld_wait();
            const int jj = (c - par) >> 1;
            const float blo = jj == 0 ? bq[0] : bq[2], bhi = jj == 0 ? bq[1] : bq[3];  // bias of columns [0,32) / [32,64) of this chunk, by lane
            const bool lnc = LNF && p.ln_in != nullptr;  // LayerNorm consumer: (acc - mean * colsum) * rstd + bias
            const float4* csp = reinterpret_cast<const float4*>(p.ln_cs + n0);
            const bool gelu = p.epi == MDE_EPI_BIAS_GELU_BF16;
            const int which = p.epi == MDE_EPI_QKV_SPLIT ? n0 / MDE_VIT_DIM : 0;  // 0 q, 1 k, 2 v (uniform per 64-column chunk)
            if (which == 2) {
              // V goes out transposed, vt[b*12+head][d][token]: consecutive lanes hold consecutive tokens -> coalesced 2-byte stores
              const int m = row0 + lane;
              const bool live = m < p.M;  // the shuffles below need every lane: only the stores are guarded
              {
                const int mm = live ? m : 0;
                const int bb = mm / MDE_VIT_SEQ, t = mm - bb * MDE_VIT_SEQ;
                const int cc = n0 - 2 * MDE_VIT_DIM;  // = head * 64
                bf16* dst = reinterpret_cast<bf16*>(const_cast<void*>(p.aux1)) + ((size_t)bb * MDE_VIT_DIM + cc) * MDE_VT_PITCH + t;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const float bsrc = j < 8 ? blo : bhi;
                  const float4 b4 = make_float4(__shfl_sync(0xffffffffu, bsrc, (4 * j) & 31), __shfl_sync(0xffffffffu, bsrc, (4 * j + 1) & 31),
                                                __shfl_sync(0xffffffffu, bsrc, (4 * j + 2) & 31), __shfl_sync(0xffffffffu, bsrc, (4 * j + 3) & 31));
                  float4 a4 = make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]), __uint_as_float(raw[4 * j + 2]), __uint_as_float(raw[4 * j + 3]));
                  if (lnc) {  // (acc - mu * colsum) * rstd + bias  ==  acc * rstd + (nmr * colsum + bias),  nmr = -mu * rstd
                    const float4 c4 = __ldg(csp + j);
                    a4 = make_float4(fmaf(a4.x, rstd, fmaf(nmr, c4.x, b4.x)), fmaf(a4.y, rstd, fmaf(nmr, c4.y, b4.y)),
                                     fmaf(a4.z, rstd, fmaf(nmr, c4.z, b4.z)), fmaf(a4.w, rstd, fmaf(nmr, c4.w, b4.w)));
                  } else {
                    a4 = make_float4(a4.x + b4.x, a4.y + b4.y, a4.z + b4.z, a4.w + b4.w);
                  }
                  if (live) {
                    dst[(size_t)(4 * j) * MDE_VT_PITCH] = __float2bfloat16_rn(a4.x);
                    dst[(size_t)(4 * j + 1) * MDE_VT_PITCH] = __float2bfloat16_rn(a4.y);
                    dst[(size_t)(4 * j + 2) * MDE_VT_PITCH] = __float2bfloat16_rn(a4.z);
                    dst[(size_t)(4 * j + 3) * MDE_VT_PITCH] = __float2bfloat16_rn(a4.w);
                  }
                }
              }
              continue;  // nothing staged, no TMA store for this chunk (g is advanced by the loop header)
            }
            const float qs = (p.epi == MDE_EPI_QKV_SPLIT && which == 0) ? 0.125f : 1.0f;  // 1/sqrt(64) folded into q
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float v[8];
#pragma unroll
              for (int t = 0; t < 8; ++t) v[t] = __uint_as_float(raw[8 * j + t]);
              {
                float bb[8];
#pragma unroll
                for (int t = 0; t < 8; ++t) bb[t] = __shfl_sync(0xffffffffu, j < 4 ? blo : bhi, (8 * j + t) & 31);
                if (lnc) {  // (acc - mu * colsum) * rstd + bias  ==  acc * rstd + (nmr * colsum + bias),  nmr = -mu * rstd
                  const float4 c0 = __ldg(csp + 2 * j), c1 = __ldg(csp + 2 * j + 1);
                  const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
                  for (int t = 0; t < 8; ++t) v[t] = fmaf(v[t], rstd, fmaf(nmr, cs[t], bb[t]));
                } else {
#pragma unroll
                  for (int t = 0; t < 8; ++t) v[t] += bb[t];
                }
              }
              if (gelu) {
#pragma unroll
                for (int t = 0; t < 8; ++t) v[t] = gelu_fast(v[t]);
              }
              if (qs != 1.0f) {
#pragma unroll
                for (int t = 0; t < 8; ++t) v[t] *= qs;
              }
              *reinterpret_cast<uint4*>(row + ((static_cast<uint32_t>(j) ^ sw) << 4)) =
                  make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
            }
          }
          fence_proxy_async_smem();  // generic-proxy writes -> visible to the TMA engine
          __syncwarp();
          if constexpr (EC == 3) {
            if (lane == 0) {
              // token rows of image b start at patch 0 -> coordinate 0 of a map over tokens 1..196: a chunk that crosses into the next
              // image is stored twice, each store clipped by the map's bounds to its own image's rows
              const int b0 = row0 / 196, t0 = row0 - b0 * 196;
              tma_store_3d(&p.tmOut, buf, n0, t0, b0);  // rows past patch 195 of image b0 are clipped by the map
              // the rest of a chunk that crosses into image b0 + 1 (a multiple of 4 rows, 196 = 4 x 49): 4-row boxes from the matching
              // rows of the staging buffer (tmRes = the same tensor with a {32, 4, 1} box).  A single store at the negative row
              // coordinate t0 - 196 would be the obvious way — it raises "illegal instruction" (measured; bring-up notes in r02_experiments.md)
              for (int r = 196 - t0; r < 32; r += 4) tma_store_3d(&p.tmRes, buf + r * 128, n0, r - (196 - t0), b0 + 1);
              bulk_commit();
            }
            if (pe_patch == 0) {  // the image's CLS row = class_token + pos_embedding[0] (tv:287-289), these 32 columns
              const float4* cls = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux1) + n0);
              const float4* p0 = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.aux0) + n0);
              float4* d0 = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (size_t)pe_img * MDE_VIT_SEQ * MDE_VIT_DIM + n0);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 c4 = __ldg(cls + j), q4 = __ldg(p0 + j);
                d0[j] = make_float4(c4.x + q4.x, c4.y + q4.y, c4.z + q4.z, c4.w + q4.w);
              }
            }
            continue;
          }
          if (lane == 0) {
            if (p.epi == MDE_EPI_QKV_SPLIT) {  // q -> tmOut, k -> tmRes, both token-major [B*197][768]
              const int which = n0 / MDE_VIT_DIM;
              tma_store_2d(which == 0 ? &p.tmOut : &p.tmRes, buf, n0 - which * MDE_VIT_DIM, row0);
            } else {
              tma_store_2d(&p.tmOut, buf, n0, row0);
            }
            bulk_commit();
          }
          if (LNF && EC == 0 && p.xb) {
            // LayerNorm producer: bf16 copy of the staged fp32 chunk, read back with lane = four columns so that a warp store
            // covers four 64-byte row segments (a row-per-thread store would touch 32 lines per instruction).  The staged
            // tile is only read here and by the TMA store.
            bf16* xb = reinterpret_cast<bf16*>(p.xb);
            const int cc = (lane & 7) * 4;  // lane = four columns: a warp store covers four 64-byte row segments
#pragma unroll
            for (int t = 0; t < 8; ++t) {
              const int rr = 4 * t + (lane >> 3);
              const float4 f = *reinterpret_cast<const float4*>(buf + rr * 128 + ((static_cast<uint32_t>(lane & 7) ^ (rr & 7)) << 4));
              if (row0 + rr < p.M)
                *reinterpret_cast<uint2*>(xb + (size_t)(row0 + rr) * p.xb_ld + n0 + cc) = make_uint2(pack_bf16x2(f.x, f.y), pack_bf16x2(f.z, f.w));
            }
            __syncwarp();  // every lane has read the tile before lane 0 may recycle the buffer
          }
        }
        if (LNF && EC == 0 && p.xb && row0 + lane < p.M) {
          // whole tile: slot = parity (even / odd chunks); split tail item (64 columns, one chunk per warp): slot = chunk index
          const int slot = (it.n0 >> 8) * 8 + (it.bn == 256 ? par : ((it.n0 & 255) >> 5) + par);
          *reinterpret_cast<float2*>(p.ln_out + ((size_t)slot * p.M + (row0 + lane)) * 2) = make_float2(ls1, ls2);
        }
        tc_fence_before();
        __syncwarp();
        if (resid) {  // drained: load the residual of the item that will use this buffer next (two items ahead), then arrive
          if (inext + num_clusters < p.total_items) init_acc(decode_item(p, inext + num_clusters), acc);
        } else if (lane == 0) {
          mbar_arrive_cluster(&tempty_bar[acc], 0);  // the leader's MMA thread owns the accumulator ring
        }
        MDE_PROF(if (prof) {
          p.prof[blockIdx.x * 8 + 3] += clk() - t0;  // epilogue warp 2 working on the tile
          p.prof[blockIdx.x * 8 + 4] += 1;           // tiles
        })
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
        i = inext;
        it = itn;
      }
      if (lane == 0) bulk_wait_exit<0>();  // every store has left shared memory
      __syncwarp();
    } else {
      for (int i = cluster_id; i < p.total_items; i += num_clusters) {
        const Item it = decode_item(p, i);
        RowInfo ri;
        ri.m = it.m0 + (int)rank * 128 + r;
        ri.b = ri.y = ri.x = 0;
        ri.valid = ri.m < p.M;
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 256;
        const int nchunks = it.bn >> 5;
#pragma unroll 1
        for (int c = par; c < nchunks; c += 2) {
          uint32_t raw[32];
          tmem_ld_32x32b_x32(taddr + c * 32, raw);
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
          epilogue_chunk<32, Tc2Params>(p, v, it.n0 + c * 32, ri);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], 0);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();  // both CTAs are done with each other's shared memory / TMEM
  tc_fence_after();
  if (warp == 1) tmem_dealloc_pair(tmem_base, k2TmemCols);
  MDE_PROF(if (p.prof && threadIdx.x == 0) p.prof[blockIdx.x * 8 + 5] = clk() - t_start;)  // CTA lifetime
}

// ------------------------------------------------------------------------------------ host side
long long* g_gemm_prof = nullptr;  // set by mde_debug_gemm_prof (tools/gemm_prof.py)

int launch_gemm_pair(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int epilogue, const float* bias, void* out, int ldo,
                     const void* aux0, const void* aux1, const LnFuse* ln, cudaStream_t stream) {
  Tc2Params p;
  memset(&p, 0, sizeof(p));
  if (ln) {
    const bool f32res = epilogue == MDE_EPI_BIAS_RESID_F32 || epilogue == MDE_EPI_BIAS_F32 || epilogue == MDE_EPI_PATCH_EMBED;
    if (ln->xb_out || ln->stats_out) {  // producer
      if (!(ln->xb_out && ln->stats_out) || !f32res || epilogue == MDE_EPI_PATCH_EMBED || N % 256 != 0 || (uintptr_t)ln->xb_out % 16 || (uintptr_t)ln->stats_out % 16) {
        set_error("mde_gemm_ln_bf16: xb_out + stats_out need an fp32 TMA epilogue (RESID_F32 / BIAS_F32), N %% 256 == 0, 16-byte alignment");
        return MDE_ERR_BAD_ARGUMENT;
      }
      p.xb = ln->xb_out;
      p.ln_out = ln->stats_out;
      p.ln_rows = epilogue == MDE_EPI_PATCH_EMBED ? M / 196 * MDE_VIT_SEQ : M;
      p.xb_ld = N;
    }
    if (ln->stats_in || ln->colsum) {  // consumer
      const bool b16 = epilogue == MDE_EPI_BIAS_BF16 || epilogue == MDE_EPI_BIAS_GELU_BF16 || epilogue == MDE_EPI_QKV_SPLIT;
      if (!(ln->stats_in && ln->colsum) || !b16 || K % 256 != 0 || K > 1024 || (uintptr_t)ln->stats_in % 16 || (uintptr_t)ln->colsum % 16) {
        set_error("mde_gemm_ln_bf16: stats_in + colsum need a bf16 epilogue (BIAS_BF16 / BIAS_GELU_BF16 / QKV_SPLIT), K %% 256 == 0, K <= 1024, 16-byte alignment");
        return MDE_ERR_BAD_ARGUMENT;
      }
      p.ln_in = ln->stats_in;
      p.ln_cs = ln->colsum;
      p.ln_ktiles = K / 256;
      {  // the producer (same M, N = this K) ran launch_gemm_pair's schedule: whole waves of 256x256 tiles, the rest split four ways
        const int ptiles = ceil_div(M, 256) * p.ln_ktiles, pclusters = sm_count() / 2;
        p.ln_full = (ptiles / pclusters) * pclusters;
      }
      p.ln_eps = ln->eps;
      p.ln_invk = 1.0f / (float)K;
    }
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
    const uint64_t strides[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, 128};
    int rc = encode_tmap_bf16(&p.tmA, A, 2, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
    const uint64_t strides[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {64, 128};
    const uint32_t boxn[2] = {64, 32};
    int rc = encode_tmap_bf16(&p.tmB, W, 2, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
    rc = encode_tmap_bf16(&p.tmBn, W, 2, dims, strides, boxn, 128);
    if (rc != MDE_OK) return rc;
  }
  const int clusters = sm_count() / 2;
  const int tiles = ceil_div(M, 256) * (N / 256);
  p.M = M;
  p.N = N;
  p.k_blocks = K / 64;
  p.n_wide = N / 256;
  p.full_items = (tiles / clusters) * clusters;           // whole waves run as 256x256 tiles
  p.total_items = p.full_items + 4 * (tiles - p.full_items);  // the remainder is split four ways along N
  p.epi = epilogue;
  p.bias = bias;
  p.out = out;
  p.ldo = ldo;
  p.aux0 = aux0;
  p.aux1 = aux1;
  p.tma_epi = epilogue <= MDE_EPI_BIAS_F32 || epilogue == MDE_EPI_QKV_SPLIT;
  p.out_elt = (epilogue == MDE_EPI_BIAS_RESID_F32 || epilogue == MDE_EPI_BIAS_F32) ? 4 : 2;
  static const bool pe_direct = getenv("MDE_PATCH_EMBED_DIRECT") != nullptr && getenv("MDE_PATCH_EMBED_DIRECT")[0] == '1';  // A/B timing
  const bool pe_tma = epilogue == MDE_EPI_PATCH_EMBED && !p.xb && !p.ln_in && !pe_direct && M % 196 == 0 && N == MDE_VIT_DIM &&  // (the output pitch of this epilogue is MDE_VIT_DIM by definition: ldo is not used)
                      (uintptr_t)out % 16 == 0 && (uintptr_t)aux0 % 16 == 0 && (uintptr_t)aux1 % 16 == 0;
  if (pe_tma) {
    // tokens 1..196 of every image: [768 columns][196 patches][B images], image pitch 197 rows — rows that do not exist are clipped
    const uint64_t dims[3] = {(uint64_t)MDE_VIT_DIM, 196, (uint64_t)(M / 196)};
    const uint64_t strides[2] = {(uint64_t)MDE_VIT_DIM * 4, (uint64_t)MDE_VIT_SEQ * MDE_VIT_DIM * 4};
    const uint32_t box[3] = {32, 32, 1};
    int rc = encode_tmap(&p.tmOut, 4, reinterpret_cast<const float*>(out) + MDE_VIT_DIM, 3, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
    const uint32_t box4[3] = {32, 4, 1};  // the part of a chunk that belongs to the NEXT image goes out in 4-row pieces
    rc = encode_tmap(&p.tmRes, 4, reinterpret_cast<const float*>(out) + MDE_VIT_DIM, 3, dims, strides, box4, 128);
    if (rc != MDE_OK) return rc;
    p.out_elt = 4;
  }
  if (p.tma_epi) {
    const bool qkv = epilogue == MDE_EPI_QKV_SPLIT;
    const int out_cols = qkv ? MDE_VIT_DIM : N, out_ld = qkv ? MDE_VIT_DIM : ldo;
    const void* second = epilogue == MDE_EPI_BIAS_RESID_F32 ? aux0 : (qkv ? aux0 : out);  // residual in / k out / unused
    if (((uintptr_t)out % 16) || ((uintptr_t)second % 16) || (out_ld * p.out_elt) % 16) {
      set_error("mde_gemm_bf16: output/residual must be 16-byte aligned with a 16-byte multiple row pitch");
      return MDE_ERR_BAD_ARGUMENT;
    }
    const uint64_t dims[2] = {(uint64_t)out_cols, (uint64_t)M};
    const uint64_t strides[1] = {(uint64_t)out_ld * p.out_elt};
    const uint32_t box[2] = {(uint32_t)(128 / p.out_elt), 32};
    int rc = encode_tmap(&p.tmOut, p.out_elt, out, 2, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
    rc = encode_tmap(&p.tmRes, p.out_elt, second, 2, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
  }
  p.prof = g_gemm_prof;
  const int smem_max = 227 * 1024;
  int stages = (smem_max - 1024 - k2StagingBytes - k2BarrierBytes) / k2StageBytes;
  if (stages > k2MaxStages) stages = k2MaxStages;
  static const int stages_env = getenv("MDE_GEMM_STAGES") ? atoi(getenv("MDE_GEMM_STAGES")) : 0;  // experiments: fewer pipeline stages
  if (stages_env > 0 && stages_env < stages) stages = stages_env;
  p.stages = stages;
  const int smem_bytes = stages * k2StageBytes + k2StagingBytes + 1024 + k2BarrierBytes;
  static bool attr_set = false;
  if (!attr_set) {
#define MDE_G2_ATTR(L_, E_) MDE_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm2_kernel<L_, E_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    MDE_G2_ATTR(0, 0) MDE_G2_ATTR(0, 1) MDE_G2_ATTR(0, 2) MDE_G2_ATTR(0, 3) MDE_G2_ATTR(1, 0) MDE_G2_ATTR(1, 1) MDE_G2_ATTR(1, 2)
#undef MDE_G2_ATTR
    attr_set = true;
  }
  const int grid = 2 * std::min(clusters, p.total_items);
  const int ec = pe_tma ? 3 : (!p.tma_epi ? 2 : (p.out_elt == 4 ? 0 : 1));
  const int lnf = (p.xb || p.ln_in) ? 1 : 0;
#define MDE_G2_CASE(L_, E_) \
  if (lnf == L_ && ec == E_) launch_k(tc_gemm2_kernel<L_, E_>, dim3(grid), dim3(k2Threads), smem_bytes, stream, p);
  MDE_G2_CASE(0, 0) MDE_G2_CASE(0, 1) MDE_G2_CASE(0, 2) MDE_G2_CASE(0, 3) MDE_G2_CASE(1, 0) MDE_G2_CASE(1, 1) MDE_G2_CASE(1, 2)
#undef MDE_G2_CASE
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

}  // namespace mde

extern "C" int mde_debug_gemm_prof(long long* counters) {
#ifdef MDE_GEMM_PROF
  mde::g_gemm_prof = counters;  // device buffer of 8 counters per CTA (zeroed by the caller), or null to switch the counters off
  return MDE_OK;
#else
  (void)counters;
  mde::set_error("mde_debug_gemm_prof: this build has no counters (rebuild with MDE_BUILD_PROF=1 python monodepth-estimation_b200/build.py)");
  return MDE_ERR_UNSUPPORTED;
#endif
}
