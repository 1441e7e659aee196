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
// per TMEM lane quarter, each taking every other 32-column chunk, residual loads software-pipelined.
#include <stdio.h>

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

struct alignas(64) Tc2Params {
  CUtensorMap tmA;   // box [64, 128]
  CUtensorMap tmB;   // box [64, 128]: wide items (pair N = 256)
  CUtensorMap tmBn;  // box [64, 32] : narrow items (pair N = 64)
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
  int cout, h, w, out_cstride, out_coff, f16;
};

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

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(k2Threads, 1) tc_gemm2_kernel(const __grid_constant__ Tc2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int stages = p.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)stages * k2StageBytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmBn);
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
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_slot, k2TmemCols);
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = cluster_id; i < p.total_items; i += num_clusters) {
        const Item it = decode_item(p, i);
        const int half = it.bn >> 1;
        const CUtensorMap* tb = it.bn == 256 ? &p.tmB : &p.tmBn;
        const uint32_t tx = 2u * (k2ABytes + half * 128);
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
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
      for (int i = cluster_id; i < p.total_items; i += num_clusters) {
        const Item it = decode_item(p, i);
        const uint32_t idesc = umma_idesc_bf16(256, it.bn);
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * k2StageBytes);
          const uint64_t da = umma_smem_desc_kmajor<128>(sa);
          const uint64_t db = umma_smem_desc_kmajor<128>(sa + k2ABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
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
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs)
    const int q = warp & 3;          // TMEM lane quarter
    const int par = (warp - 2) >> 2; // which half of the 32-column chunks this warp takes
    const int r = q * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
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
      if (p.epi == MDE_EPI_BIAS_RESID_F32) {
        // residual rows are read one chunk ahead of the accumulator they are added to
        const float* rs_row = reinterpret_cast<const float*>(p.aux0) + (size_t)ri.m * p.ldo + it.n0;
        float* out_row = reinterpret_cast<float*>(p.out) + (size_t)ri.m * p.ldo + it.n0;
        float4 nxt[8];
        int c = par;
        if (ri.valid && c < nchunks) {
#pragma unroll
          for (int j = 0; j < 8; ++j) nxt[j] = __ldg(reinterpret_cast<const float4*>(rs_row + c * 32) + j);
        }
#pragma unroll 1
        for (; c < nchunks; c += 2) {
          float4 cur[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) cur[j] = nxt[j];
          if (ri.valid && c + 2 < nchunks) {
#pragma unroll
            for (int j = 0; j < 8; ++j) nxt[j] = __ldg(reinterpret_cast<const float4*>(rs_row + (c + 2) * 32) + j);
          }
          uint32_t raw[32];
          tmem_ld_32x32b_x32(taddr + c * 32, raw);
          tmem_ld_wait();
          if (ri.valid) {
            const float4* bp = reinterpret_cast<const float4*>(p.bias + it.n0 + c * 32);
            float4* dst = reinterpret_cast<float4*>(out_row + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b4 = __ldg(bp + j);
              dst[j] = make_float4(__uint_as_float(raw[4 * j]) + b4.x + cur[j].x, __uint_as_float(raw[4 * j + 1]) + b4.y + cur[j].y,
                                   __uint_as_float(raw[4 * j + 2]) + b4.z + cur[j].z, __uint_as_float(raw[4 * j + 3]) + b4.w + cur[j].w);
            }
          }
        }
      } else {
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
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], 0);  // the leader's MMA thread owns the accumulator ring
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  cluster_sync_all();  // both CTAs are done with each other's shared memory / TMEM
  tc_fence_after();
  if (warp == 1) tmem_dealloc_pair(tmem_base, k2TmemCols);
}

// ------------------------------------------------------------------------------------ host side
int launch_gemm_pair(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int epilogue, const float* bias, void* out, int ldo,
                     const void* aux0, const void* aux1, cudaStream_t stream) {
  Tc2Params p;
  memset(&p, 0, sizeof(p));
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
  int stages = (220 * 1024 - 2048) / k2StageBytes;
  if (stages > k2MaxStages) stages = k2MaxStages;
  p.stages = stages;
  const int smem_bytes = stages * k2StageBytes + 1024 + (2 * k2MaxStages + 8) * 8;
  static bool attr_set = false;
  if (!attr_set) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    attr_set = true;
  }
  const int grid = 2 * std::min(clusters, p.total_items);
  tc_gemm2_kernel<<<grid, k2Threads, smem_bytes, stream>>>(p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

}  // namespace mde
