// The ViT MLP (torchvision MLPBlock, tv:117-118: Linear(768,3072) -> GELU -> Linear(3072,768), + residual) as ONE persistent
// kernel of CTA pairs: both GEMMs of tc_gemm2.cu in a single launch, with the second one's tiles released per 256-row block.
//
// Why: as two launches at batch 32 (M = 6304 = 24.6 row blocks of 256) fc1 is 300 tiles on 74 clusters (4.05 waves) and fc2 is
// 75 tiles on 74 clusters: fc2 pays a second, quarter-filled wave (its 75th tile, split four ways, is load-bound) — 28.8 us for
// 18.4 us of tile time — and each launch has its own prologue, pipeline fill and un-overlapped last epilogue.  Here a cluster
// works through a static list: its share of the 300 fc1 tiles (cost 1 each: K = 768), then its fc2 tile(s) (cost 4 each:
// K = 3072).  The one cluster that gets TWO fc2 tiles gets correspondingly fewer fc1 tiles, so every cluster carries 8-9 units
// instead of 4-5 + 4-8, and fc2's pipeline starts while the last fc1 tiles of OTHER clusters are still in flight.
//
// Dependency: an fc2 tile of row block m reads the bf16 hidden activations of ALL 12 fc1 tiles of row block m, written by up to 12
// other clusters with TMA stores.  Every epilogue warp signals `done[m]` (release, after its bulk stores have completed) once per
// fc1 tile; the TMA producer of an fc2 tile spins (acquire, bounded) until done[m] holds 12 tiles x 16 warps, then crosses to the
// async proxy with a fence before its first load.  No cluster ever waits before it has finished ALL its fc1 tiles and fc1 tiles
// wait for nothing, so the wait cannot deadlock however the CTAs are scheduled.  The last reader of a row block resets its
// counters: the buffer is all zeros again when the kernel ends (and must be on entry).
//
// Everything else is tc_gemm2.cu's machinery: 256x256x64 tcgen05.mma.cta_group::2 tiles, TMA-fed 5-stage ring, two TMEM
// accumulator buffers, 8 epilogue warps per CTA staging 32-row sub-tiles in swizzled shared memory for TMA stores, GELU in the
// fc1 epilogue, residual + bias pre-loaded into the fc2 accumulator (tcgen05.st) so its epilogue is TMEM -> shared -> TMA store.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.h"
#include "ptx.cuh"
#include "tc_epilogue.cuh"

namespace mde {

constexpr int kMlpEpiWarps = 16;  // four per TMEM lane quarter: warp `sub` of a quarter owns columns [64 sub, 64 sub + 64) of every 256-column tile
constexpr int kMlpThreads = 64 + 32 * kMlpEpiWarps;
constexpr int kMlpMaxStages = 8;
constexpr int kMlpABytes = 128 * 128;  // 128 rows x 64 bf16
constexpr int kMlpBBytes = 128 * 128;  // this CTA's half of the 256-row B tile
constexpr int kMlpStageBytes = kMlpABytes + kMlpBBytes;
constexpr uint32_t kMlpTmemCols = 512;  // 2 accumulator buffers x 256 columns
constexpr int kMlpEpiBufBytes = 32 * 128;  // one epilogue warp's staging buffer: 32 rows x 128 bytes (SWIZZLE_128B)
constexpr int kMlpStagingBytes = kMlpEpiWarps * kMlpEpiBufBytes;  // ONE buffer per warp: it stores one or two chunks per tile
constexpr int kMlpBarrierBytes = 512;
constexpr int kMlpMaxClusters = 80;
constexpr int kMlpMaxItems = 14;          // per cluster
constexpr uint16_t kMlpEnd = 0xFFFFu;
constexpr uint16_t kMlpFc2 = 0x8000u;     // item code: bit 15 = fc2, bit 14 = signal this fc1 tile without delay, low bits = tile index
constexpr uint16_t kMlpPrompt = 0x4000u;  // (row block * n_wide + column block)

struct MlpProb {
  CUtensorMap tmA;    // box [64, 128] over the A operand (fc1: LayerNorm output, fc2: hidden activations)
  CUtensorMap tmB;    // box [64, 128] over the weight [N][K]
  CUtensorMap tmOut;  // box [128 bytes of columns, 32 rows] over the output (fc1: bf16 hidden, fc2: fp32 stream)
  const float* bias;
  int k_blocks, n_wide;
};

struct alignas(64) MlpParams {
  MlpProb pr[2];      // 0 = fc1, 1 = fc2
  CUtensorMap tmRes;  // fc2: box [32 fp32 columns, 32 rows] over the residual input
  int M, stages;
  unsigned* sync;     // [2][64]: done[m] (fc1 epilogue-warp signals of row block m), readers[m]; all zero between launches
  unsigned done_target, readers_target;
  uint16_t items[kMlpMaxClusters][kMlpMaxItems];  // per cluster, in order, kMlpEnd-terminated
  long long* prof;    // MDE_GEMM_PROF builds: 8 time stamps per CTA (mde_debug_gemm_prof), or null
  int dbg;            // MDE_MLP_DBG (tools/mlp_bench.py only, WRONG RESULTS): 1 no GELU, 2 no fc1 staging / store, 4 no fc1 TMEM load either
};
#ifdef MDE_GEMM_PROF
#define MLP_PROF(...) __VA_ARGS__
extern long long* g_gemm_prof;
#else
#define MLP_PROF(...)
#endif

struct MlpItem {
  int fc2, m0, n0, mblk;
};
__device__ __forceinline__ MlpItem mlp_decode(const MlpParams& p, uint16_t code) {
  MlpItem it;
  it.fc2 = (code & kMlpFc2) ? 1 : 0;
  const int tile = code & 0x3FFF, nw = p.pr[it.fc2].n_wide;
  it.mblk = tile / nw;
  it.m0 = it.mblk * 256;
  it.n0 = (tile - it.mblk * nw) * 256;
  return it;
}

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Column ownership inside a 256-column tile is the SAME for both GEMMs: warp `sub` of a lane quarter owns columns [64 sub, 64 sub + 64)
// = one 64-column bf16 chunk of an fc1 tile = the 32-column fp32 chunks 2 sub, 2 sub + 1 of an fc2 tile and of its residual pre-load.
// It has to be: a warp initialises an accumulator buffer for the item two ahead right after draining ITS columns of it, while the
// other warps of the quarter may still be reading theirs.
// Bias values, one column per lane: b[h] = bias of column 64 sub + 32 h + lane (shuffled to the row owners in the epilogue).
__device__ __forceinline__ void lane_bias_mlp(const float* bias, int n0, int sub, int lane, float (&b)[2]) {
  b[0] = __ldg(bias + n0 + sub * 64 + lane);
  b[1] = __ldg(bias + n0 + sub * 64 + 32 + lane);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kMlpThreads, 1) tc_mlp_kernel(const __grid_constant__ MlpParams p) {
  pdl_launch_dependents();
  MLP_PROF(const long long t_start = clock64();)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int stages = p.stages;
  uint8_t* staging = smem + (size_t)stages * kMlpStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + kMlpStagingBytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* res_bar = tempty_bar + 2;  // [kMlpEpiWarps] residual-chunk arrival barriers
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + kMlpEpiWarps);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const uint16_t* my = p.items[cluster_id];

  if (warp == 0 && elect_one()) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      tma_prefetch_desc(&p.pr[k].tmA);
      tma_prefetch_desc(&p.pr[k].tmB);
      tma_prefetch_desc(&p.pr[k].tmOut);
    }
    tma_prefetch_desc(&p.tmRes);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full_bar[s], 1);   // leader's expect_tx arrive; bytes from both CTAs' TMA
        mbar_init(&empty_bar[s], 1);  // pair commit (multicast) from the leader's MMA thread
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tfull_bar[a], 1);    // pair commit (multicast)
        mbar_init(&tempty_bar[a], 2 * kMlpEpiWarps);  // every epilogue warp of both CTAs (only the leader's copy is waited on)
      }
      for (int a = 0; a < kMlpEpiWarps; ++a) mbar_init(&res_bar[a], 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, kMlpTmemCols);
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_grid_wait();  // everything above overlaps the previous kernel's tail
  MLP_PROF(long long* pf = p.prof ? p.prof + blockIdx.x * 8 : nullptr;)

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int idx = 0; idx < kMlpMaxItems; ++idx) {
        const uint16_t code = my[idx];
        if (code == kMlpEnd) break;
        const MlpItem it = mlp_decode(p, code);
        const MlpProb& pr = p.pr[it.fc2];
        if (it.fc2) {
          // every fc1 tile of this row block has been stored (see the header); bounded spin: a protocol bug is a launch error
          const unsigned* done = p.sync + it.mblk;
          const long long t0 = clock64();
          MLP_PROF(if (pf && pf[1] == 0) pf[1] = t0 - t_start;)  // producer reaches its first fc2 tile
          while (ld_acquire_gpu(done) < p.done_target) {
            __nanosleep(64);
            if (clock64() - t0 > 4000000000LL) {
              printf("mde: tc_mlp_kernel: row block %d never completed (block %d, have %u of %u)\n", it.mblk, blockIdx.x, ld_acquire_gpu(done), p.done_target);
              __trap();
            }
          }
          MLP_PROF(if (pf && pf[2] == 0) pf[2] = clock64() - t_start;)  // its row block is complete
          unsigned* readers = p.sync + 64 + it.mblk;
          if (atomicAdd(readers, 1u) == p.readers_target - 1) {  // last reader of this row block: leave the counters zero for the next launch
            *const_cast<volatile unsigned*>(p.sync + it.mblk) = 0u;
            *const_cast<volatile unsigned*>(readers) = 0u;
            __threadfence();
          }
          fence_proxy_async_all();  // the acquire above orders generic accesses; the loads below go through the async proxy
        }
        const uint32_t tx = 2u * (kMlpABytes + kMlpBBytes);
        for (int kb = 0; kb < pr.k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * kMlpStageBytes;
          if (leader) mbar_expect_tx(&full_bar[stage], tx);
          tma_load_2d_pair(sa, &pr.tmA, &full_bar[stage], kb * 64, it.m0 + (int)rank * 128);
          tma_load_2d_pair(sa + kMlpABytes, &pr.tmB, &full_bar[stage], kb * 64, it.n0 + (int)rank * 128);
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
      constexpr uint32_t idesc = umma_idesc_bf16(256, 256);
      for (int idx = 0; idx < kMlpMaxItems; ++idx) {
        const uint16_t code = my[idx];
        if (code == kMlpEnd) break;
        const MlpItem it = mlp_decode(p, code);
        const int kbs = p.pr[it.fc2].k_blocks;
        // the epilogue warps arrive once per use of a buffer, the first two included: fc1 -> free, fc2 -> initialised with residual + bias
        MLP_PROF(long long w0 = clock64();)
        mbar_wait(&tempty_bar[acc], acc_phase);
        tc_fence_after();
        MLP_PROF(if (pf && !it.fc2) pf[0] += clock64() - w0;)  // fc1 tiles: MMA thread waiting for a free accumulator
        MLP_PROF(if (pf && it.fc2 && pf[3] == 0) pf[3] = clock64() - t_start;)  // MMA thread starts its first fc2 tile (last fc1 tile issued)
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = 0; kb < kbs; ++kb) {
          MLP_PROF(w0 = clock64();)
          mbar_wait(&full_bar[stage], phase);
          MLP_PROF(if (pf && !it.fc2) { const long long dw = clock64() - w0; pf[7] += dw; p.prof[148 * 8 + cluster_id * 16 + kb] += dw; })  // fc1 tiles: MMA thread waiting for operands (and per k-block index)
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * kMlpStageBytes);
          const uint64_t da = umma_smem_desc_kmajor<128>(sa);
          const uint64_t db = umma_smem_desc_kmajor<128>(sa + kMlpABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, it.fc2 | kb | k);
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
      MLP_PROF(if (pf) pf[4] = clock64() - t_start;)  // last MMA issued
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs)
    const int q = warp & 3;           // TMEM lane quarter
    const int sub = (warp - 2) >> 2;  // which 64 columns of a tile
    const int ew = warp - 2;
    uint8_t* buf = staging + ew * kMlpEpiBufBytes;
    uint8_t* row = buf + lane * 128;
    uint64_t* rbar = res_bar + ew;
    const uint32_t sw = static_cast<uint32_t>(lane & 7);
    uint32_t rph = 0;  // parity of the next completion of this warp's residual barrier
    // fc2: (residual + bias) of item `it2` -> this warp's columns of accumulator buffer `acc2`, then arrive on tempty
    auto init_acc = [&](const MlpItem& it2, int acc2) {
      const int row02 = it2.m0 + (int)rank * 128 + q * 32;
      const uint32_t ta = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc2 * 256 + sub * 64;
      float bq[2];
      lane_bias_mlp(p.pr[1].bias, it2.n0, sub, lane, bq);  // in flight together with the residual load
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if (lane == 0) {
          bulk_wait_read<0>();  // earlier stores no longer read the staging buffer
          mbar_expect_tx(rbar, kMlpEpiBufBytes);
          tma_load_2d(buf, &p.tmRes, rbar, it2.n0 + sub * 64 + j * 32, row02);
        }
        __syncwarp();
        mbar_wait(rbar, rph);
        rph ^= 1u;
        uint32_t v[32];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float4 r4 = *reinterpret_cast<const float4*>(row + ((static_cast<uint32_t>(u) ^ sw) << 4));
          v[4 * u] = __float_as_uint(r4.x + __shfl_sync(0xffffffffu, bq[j], 4 * u));
          v[4 * u + 1] = __float_as_uint(r4.y + __shfl_sync(0xffffffffu, bq[j], 4 * u + 1));
          v[4 * u + 2] = __float_as_uint(r4.z + __shfl_sync(0xffffffffu, bq[j], 4 * u + 2));
          v[4 * u + 3] = __float_as_uint(r4.w + __shfl_sync(0xffffffffu, bq[j], 4 * u + 3));
        }
        tmem_st_32x32b_x32(ta + j * 32, v);
        __syncwarp();  // every lane has read the buffer
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc2], 0);
    };
    // this warp's bulk stores of an fc1 tile are complete -> count it for the tile's row block (lane 0 only)
    auto signal = [&](int mblk) {
      fence_proxy_async_all();
      __threadfence();
      red_release_gpu_add(p.sync + mblk, 1u);
    };
    // first use of the two accumulator buffers
#pragma unroll 1
    for (int j = 0; j < 2; ++j) {
      const uint16_t code = my[j];
      if (code == kMlpEnd) break;
      if (code & kMlpFc2) init_acc(mlp_decode(p, code), j);
      else if (lane == 0) mbar_arrive_cluster(&tempty_bar[j], 0);
    }
    int acc = 0;
    uint32_t acc_phase = 0;
    int sig_m = -1;  // row block of the previous fc1 tile: signalled one tile later, when its store is known complete without waiting
#pragma unroll 1
    for (int idx = 0; idx < kMlpMaxItems; ++idx) {
      const uint16_t code = my[idx];
      if (code == kMlpEnd) break;
      const MlpItem it = mlp_decode(p, code);
      const uint16_t code1 = idx + 1 < kMlpMaxItems ? my[idx + 1] : kMlpEnd;
      const uint16_t code2 = idx + 2 < kMlpMaxItems ? my[idx + 2] : kMlpEnd;
      const int row0 = it.m0 + (int)rank * 128 + q * 32;
      float bq[2] = {0.0f, 0.0f};
      if (!it.fc2) lane_bias_mlp(p.pr[0].bias, it.n0, sub, lane, bq);  // before the wait: the chunk pays no L2 round trip for it
#ifdef MDE_MLP_POLL1
      if (lane == 0) mbar_wait(&tfull_bar[acc], acc_phase);  // one poller per warp
      __syncwarp();
#else
      mbar_wait(&tfull_bar[acc], acc_phase);
#endif
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 256 + sub * 64;
      if (!it.fc2) {
        // fc1: bias + GELU -> bf16; this warp's 64 columns as two 32-column halves of one 128-byte staging row
        if (lane == 0) bulk_wait_read<0>();  // the previous tile's store has finished reading the buffer (long ago)
        __syncwarp();
        if (!(p.dbg & 4)) {
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t raw[32];
            tmem_ld_32x32b_x32(taddr + hh * 32, raw);
            tmem_ld_wait();
            if (p.dbg & 2) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float v[8];
#pragma unroll
              for (int t = 0; t < 8; ++t) {
                v[t] = __uint_as_float(raw[8 * j + t]) + __shfl_sync(0xffffffffu, bq[hh], 8 * j + t);
                if (!(p.dbg & 1)) v[t] = gelu_fast(v[t]);
              }
              *reinterpret_cast<uint4*>(row + ((static_cast<uint32_t>(4 * hh + j) ^ sw) << 4)) =
                  make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
            }
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && !(p.dbg & 6)) {
          tma_store_2d(&p.pr[0].tmOut, buf, it.n0 + sub * 64, row0);
          bulk_commit();
        }
      } else {
        // fc2: residual and bias are already in the accumulator; two 32-column fp32 chunks through the one staging buffer
#pragma unroll 1
        for (int j = 0; j < 2; ++j) {
          uint32_t raw[32];
          tmem_ld_32x32b_x32(taddr + j * 32, raw);
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
          tmem_ld_wait();
#pragma unroll
          for (int u = 0; u < 8; ++u)
            *reinterpret_cast<uint4*>(row + ((static_cast<uint32_t>(u) ^ sw) << 4)) = make_uint4(raw[4 * u], raw[4 * u + 1], raw[4 * u + 2], raw[4 * u + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&p.pr[1].tmOut, buf, it.n0 + sub * 64 + j * 32, row0);
            bulk_commit();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      // hand the accumulator buffer on to the item two ahead (before the signalling below: the MMA thread waits for this)
      if (code2 != kMlpEnd && (code2 & kMlpFc2)) init_acc(mlp_decode(p, code2), acc);
      else if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], 0);
      if (!it.fc2 && lane == 0) {
        if ((code & kMlpPrompt) || code1 == kMlpEnd || (code1 & kMlpFc2)) {
          // a row block somebody is about to need, or the cluster's last fc1 tile: wait for the store now (~1 us, once or twice per cluster)
          bulk_wait_all<0>();
          if (sig_m >= 0) signal(sig_m);
          signal(it.mblk);
          sig_m = -1;
        } else {
          if (sig_m >= 0) {  // the previous tile's store was committed a whole tile ago: only the group just committed may be pending
            bulk_wait_all<1>();
            signal(sig_m);
          }
          sig_m = it.mblk;
        }
        MLP_PROF(if (pf && warp == 2 && (code1 == kMlpEnd || (code1 & kMlpFc2))) pf[6] = clock64() - t_start;)  // last fc1 tile of this cluster signalled
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (lane == 0) bulk_wait_all<0>();  // every store has left shared memory and is globally performed
    __syncwarp();
    MLP_PROF(if (pf && warp == 2 && lane == 0) pf[5] = clock64() - t_start;)  // epilogue warp 2 done
  }

  tc_fence_before();
  cluster_sync_all();  // both CTAs are done with each other's shared memory / TMEM
  tc_fence_after();
  if (warp == 1) tmem_dealloc_pair(tmem_base, kMlpTmemCols);
}

// ------------------------------------------------------------------------------------ host side
// Static schedule, found by simulating the kernel on the host (costs in microseconds, measured with the in-kernel time stamps of
// tools/mlp_bench.py at batch 32: an fc1 tile 5.2, an fc2 tile 15, ~3 between the last MMA of a row block's fc1 tiles and the moment
// its consumers see it complete).  Event-driven list scheduling: the cluster that is free first takes the next fc1 tile (row-block
// major) — or, if a row block is complete, there are fc2 tiles left and fewer than `max_switched` clusters have switched yet,
// it switches to fc2 tiles for good (fc1 tiles never wait, so all of them complete whatever the clusters that switched are
// waiting for: no deadlock).  When the fc1 tiles run out everybody switches.  The few clusters that switch early work off the
// first row blocks' fc2 tiles while the others finish fc1, so that what is left for the end fits ONE wave: the kernel then
// ends one fc2 tile after the last fc1 tile instead of two (75 fc2 tiles on 74 clusters).  `max_switched` is chosen by makespan.
struct MlpSim {
  double makespan;
  std::vector<std::vector<uint16_t>> items;
};
static MlpSim simulate_mlp(int G, int mt, int nw1, int nw2, double c1, double c2, double lag, int max_switched) {
  const int T1 = mt * nw1, T2 = mt * nw2;
  std::vector<double> t(G, 1.5), finish(T1, 0.0), ready(mt, -1.0);
  std::vector<int> mode(G, 0), done(mt, 0);
  std::vector<char> active(G, 1);
  MlpSim r;
  r.items.assign(G, {});
  int next1 = 0, next2 = 0, switched = 0, left = G;
  while (left > 0) {
    int c = -1;
    for (int k = 0; k < G; ++k)
      if (active[k] && (c < 0 || t[k] < t[c])) c = k;
    if (mode[c] == 0) {
      const bool can2 = next2 < T2 && ready[next2 / nw2] >= 0.0 && ready[next2 / nw2] <= t[c];
      bool sw = can2 && switched < max_switched;
      if (next1 >= T1) sw = next2 < T2;
      if (sw) {
        mode[c] = 1;
        ++switched;
      } else if (next1 < T1) {
        const int tile = next1++, m = tile / nw1;
        t[c] += c1;
        finish[tile] = t[c];
        r.items[c].push_back((uint16_t)tile);
        if (++done[m] == nw1) {
          double f = 0.0;
          for (int k = 0; k < nw1; ++k) f = std::max(f, finish[m * nw1 + k]);
          ready[m] = f + lag;
        }
        continue;
      } else {
        active[c] = 0;
        --left;
        continue;
      }
    }
    if (next2 >= T2) {
      active[c] = 0;
      --left;
      continue;
    }
    const int m2 = next2 / nw2;
    if (ready[m2] < 0.0) {  // its row block is not even scheduled completely yet: let the fc1 clusters advance
      double o = 1e30;
      for (int k = 0; k < G; ++k)
        if (active[k] && k != c && mode[k] == 0) o = std::min(o, t[k]);
      if (o > 1e29) {  // cannot happen (fc1 tiles left implies an fc1 cluster is active); bail out with a huge makespan
        r.makespan = 1e30;
        return r;
      }
      t[c] = std::max(t[c], o) + 1e-3;
      continue;
    }
    t[c] = std::max(t[c], ready[m2]) + c2;
    r.items[c].push_back((uint16_t)(kMlpFc2 | next2++));
  }
  r.makespan = 0.0;
  for (int k = 0; k < G; ++k) r.makespan = std::max(r.makespan, t[k]);
  return r;
}

static int build_mlp_schedule(MlpParams& p, int clusters, int mt, int nw1, int nw2, int cost2) {
  const double c1 = 5.2, c2 = 3.75 * cost2, lag = 3.0;
  static const int force = getenv("MDE_MLP_SWITCH") ? atoi(getenv("MDE_MLP_SWITCH")) : -1;  // experiments: fixed number of early switchers
  MlpSim best;
  best.makespan = 1e30;
  for (int ms = 0; ms <= clusters / 2; ++ms) {
    if (force >= 0 && ms != force) continue;
    MlpSim r = simulate_mlp(clusters, mt, nw1, nw2, c1, c2, lag, ms);
    if (r.makespan < best.makespan - 1e-9) best = std::move(r);
  }
  if (best.makespan > 1e29) return MDE_ERR_UNSUPPORTED;
  static const bool defer = getenv("MDE_MLP_DEFER") != nullptr && getenv("MDE_MLP_DEFER")[0] == '1';  // experiments: signal one tile late
  for (int c = 0; c < clusters; ++c) {
    if ((int)best.items[c].size() > kMlpMaxItems - 1) return MDE_ERR_UNSUPPORTED;
    for (int k = 0; k < kMlpMaxItems; ++k) p.items[c][k] = kMlpEnd;
    for (size_t k = 0; k < best.items[c].size(); ++k) {
      uint16_t code = best.items[c][k];
      if (!(code & kMlpFc2) && !defer) code |= kMlpPrompt;  // the epilogue warps have slack (16 of them): every tile is signalled as soon as its store is out
      p.items[c][k] = code;
    }
  }
  return MDE_OK;
}

}  // namespace mde

using namespace mde;

extern "C" int mde_mlp_bf16(const mde_bf16* a, int lda, const mde_bf16* w1, const float* b1, mde_bf16* hidden, int ldh, const mde_bf16* w2, const float* b2,
                            const float* resid, float* out, int ldo, int M, int D, int Hd, unsigned* sync, mde_stream_t stream) {
  MDE_CHECK_ARG(a && w1 && b1 && hidden && w2 && b2 && resid && out && sync, "mde_mlp_bf16: null pointer");
  MDE_CHECK_ARG(M >= 256 && D % 256 == 0 && Hd % 256 == 0 && D >= 256 && Hd >= D && Hd % D == 0, "mde_mlp_bf16: need M >= 256, D and Hd multiples of 256, Hd a multiple of D (M=%d D=%d Hd=%d)", M, D, Hd);
  MDE_CHECK_ARG(lda >= D && ldh >= Hd && ldo >= D && lda % 8 == 0 && ldh % 8 == 0 && ldo % 4 == 0, "mde_mlp_bf16: bad leading dimensions");
  MDE_CHECK_ARG((((uintptr_t)a | (uintptr_t)w1 | (uintptr_t)hidden | (uintptr_t)w2 | (uintptr_t)resid | (uintptr_t)out | (uintptr_t)b1 | (uintptr_t)b2) % 16) == 0,
                "mde_mlp_bf16: pointers must be 16-byte aligned");
  const int mt = ceil_div(M, 256);
  MDE_CHECK_ARG(mt <= 64 && mt * (Hd / 256) < 0x4000, "mde_mlp_bf16: at most 64 row blocks of 256 (M=%d)", M);
  MlpParams p;  // ~3.3 KB: filled per call, passed by value at launch
  memset(&p, 0, sizeof(p));
  auto map_ab = [&](CUtensorMap* tm, const void* ptr, int rows, int k, int ld) {
    const uint64_t dims[2] = {(uint64_t)k, (uint64_t)rows};
    const uint64_t strides[1] = {(uint64_t)ld * 2};
    const uint32_t box[2] = {64, 128};
    return encode_tmap_bf16(tm, ptr, 2, dims, strides, box, 128);
  };
  auto map_out = [&](CUtensorMap* tm, int elt, const void* ptr, int cols, int ld) {
    const uint64_t dims[2] = {(uint64_t)cols, (uint64_t)M};
    const uint64_t strides[1] = {(uint64_t)ld * elt};
    const uint32_t box[2] = {(uint32_t)(128 / elt), 32};
    return encode_tmap(tm, elt, ptr, 2, dims, strides, box, 128);
  };
  int rc;
  if ((rc = map_ab(&p.pr[0].tmA, a, M, D, lda)) != MDE_OK) return rc;
  if ((rc = map_ab(&p.pr[0].tmB, w1, Hd, D, D)) != MDE_OK) return rc;
  if ((rc = map_out(&p.pr[0].tmOut, 2, hidden, Hd, ldh)) != MDE_OK) return rc;
  if ((rc = map_ab(&p.pr[1].tmA, hidden, M, Hd, ldh)) != MDE_OK) return rc;
  if ((rc = map_ab(&p.pr[1].tmB, w2, D, Hd, Hd)) != MDE_OK) return rc;
  if ((rc = map_out(&p.pr[1].tmOut, 4, out, D, ldo)) != MDE_OK) return rc;
  if ((rc = map_out(&p.tmRes, 4, resid, D, ldo)) != MDE_OK) return rc;
  p.pr[0].bias = b1;
  p.pr[0].k_blocks = D / 64;
  p.pr[0].n_wide = Hd / 256;
  p.pr[1].bias = b2;
  p.pr[1].k_blocks = Hd / 64;
  p.pr[1].n_wide = D / 256;
  p.M = M;
  p.sync = sync;
  MLP_PROF(p.prof = g_gemm_prof;)
  static const int dbg = getenv("MDE_MLP_DBG") ? atoi(getenv("MDE_MLP_DBG")) : 0;
  p.dbg = dbg;
  p.done_target = (unsigned)(p.pr[0].n_wide * 2 * kMlpEpiWarps);  // every epilogue warp of both CTAs, once per fc1 tile of the row block
  p.readers_target = (unsigned)(p.pr[1].n_wide * 2);
  const int clusters = std::min(sm_count() / 2, kMlpMaxClusters);
  if ((rc = build_mlp_schedule(p, clusters, mt, p.pr[0].n_wide, p.pr[1].n_wide, Hd / D)) != MDE_OK) {
    set_error("mde_mlp_bf16: schedule does not fit (M=%d on %d clusters)", M, clusters);
    return rc;
  }
  const int smem_max = 227 * 1024;
  int stages = (smem_max - 1024 - kMlpStagingBytes - kMlpBarrierBytes) / kMlpStageBytes;
  if (stages > kMlpMaxStages) stages = kMlpMaxStages;
  p.stages = stages;
  const int smem_bytes = stages * kMlpStageBytes + kMlpStagingBytes + 1024 + kMlpBarrierBytes;
  static bool attr_set = false;
  if (!attr_set) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(tc_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    attr_set = true;
  }
  launch_k(tc_mlp_kernel, dim3(2 * clusters), dim3(kMlpThreads), smem_bytes, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
