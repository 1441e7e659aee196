// Attention kernels.
//  * vit_attn_kernel   — tv:113 nn.MultiheadAttention core for ViT-B/16 (S=197, 12 heads x 64):
//    one CTA per (image, head); Q/K (token-major, 4-D box) and V^T tiles arrive by TMA (out-of-bounds rows zero-filled, so
//    the 197-token sequence needs no padded copy), S = Q K^T and O = P V run on tcgen05 with
//    the accumulators in TMEM, and the softmax is fused in between: each of the 128 threads owns
//    one query row (one TMEM lane), reads its scores with tcgen05.ld, and writes the
//    un-normalised bf16 probabilities straight into the swizzled K-major smem tile that the
//    second MMA consumes.  1/rowsum is applied in the output epilogue.
//  * spatial_attn_kernel — enhanced_model.py:20-39 on the 14x15 grid (G<=224 positions,
//    32-dim q/k, 256-dim v): tiny (25 MFLOP/image), fp32 CUDA-core kernel.
#include <stdlib.h>

#include "common.h"
#include "ptx.cuh"

namespace mde {

constexpr int kS = MDE_VIT_SEQ;   // 197
constexpr int kKeys = 208;        // keys padded to a multiple of 16 (UMMA N)
constexpr int kQBytes = 128 * 128;            // one 128-row Q tile, 64 bf16 per row
constexpr int kKBytes = kKeys * 128;          // 26624
constexpr int kKBytesPad = 27 * 1024;
constexpr int kVBytes = 64 * 128;             // one k-block of V^T: 64 d-rows x 64 keys
constexpr int kPBytes = 128 * 128;            // one k-block of P: 128 rows x 64 keys
constexpr int kOffK = 2 * kQBytes;
constexpr int kOffV = kOffK + kKBytesPad;
constexpr int kOffP = kOffV + 4 * kVBytes;
constexpr int kOffBar = kOffP + 2 * 4 * kPBytes;  // one P tile (4 k-blocks) per query tile
constexpr int kAttnSmem = kOffBar + 64 + 1024;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemTile = 256;  // column stride between the two query tiles; O reuses the first 64 columns of its tile's S

struct alignas(64) AttnParams {
  CUtensorMap tmQ, tmK, tmV;
  CUtensorMap tmO;  // the output viewed like q: box {64 d, 1 head, 128 tokens, 1 image}; tokens >= 197 of a box are clipped by the store
  bf16* out;
  int B;
};

// 256 threads = two warpgroups; warpgroup g owns query tile g (rows 128g .. 128g+127) end to end, so the two tiles'
// S = Q K^T MMAs are issued back to back and the softmax of one tile overlaps the MMAs / TMEM traffic of the other.
__global__ void __launch_bounds__(256, 1) vit_attn_kernel(const __grid_constant__ AttnParams p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kOffK;
  uint8_t* sV = smem + kOffV;
  uint64_t* bar_load = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* bar_s = bar_load + 1;  // [2]
  uint64_t* bar_o = bar_load + 3;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_load + 5);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int wg = tid >> 7, wt = tid & 127;  // warpgroup = query tile, row within the tile
  const int slab = blockIdx.x;              // b * 12 + head
  const int b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;
  uint8_t* sP = smem + kOffP + wg * 4 * kPBytes;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&p.tmQ);
      tma_prefetch_desc(&p.tmK);
      tma_prefetch_desc(&p.tmV);
      mbar_init(bar_load, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&bar_s[i], 1);
        mbar_init(&bar_o[i], 1);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_grid_wait();  // the set-up above overlaps the previous kernel's tail

  if (tid == 0) {
    mbar_expect_tx(bar_load, 2 * kQBytes + kKBytes + 4 * kVBytes);
    tma_load_4d(sQ, &p.tmQ, bar_load, 0, hd, 0, b);  // q, k are token-major [B][197][12][64]: box = {64 d, 1 head, rows, 1 image}
    tma_load_4d(sQ + kQBytes, &p.tmQ, bar_load, 0, hd, 128, b);
    tma_load_4d(sK, &p.tmK, bar_load, 0, hd, 0, b);
#pragma unroll
    for (int j = 0; j < 4; ++j) tma_load_3d(sV + j * kVBytes, &p.tmV, bar_load, 64 * j, 0, slab);
  }
  mbar_wait(bar_load, 0);

  const uint32_t t_tile = tmem_base + wg * kTmemTile;                                   // this tile's S (and later O) columns
  const uint32_t t_row = t_tile + (static_cast<uint32_t>((warp & 3) * 32) << 16);       // this thread's TMEM lane
  const uint32_t sw = static_cast<uint32_t>(wt & 7);                                    // swizzle phase of row wt
  uint8_t* p_row = sP + wt * 128;
  constexpr float kLog2e = 1.4426950408889634f;

  // ---- S = Q_g K^T  (128 x 208, K = 64)
  if (wt == 0) {
    tc_fence_after();
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, kKeys);
    const uint64_t da = umma_smem_desc_kmajor<128>(smem_u32(sQ + wg * kQBytes));
    const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sK));
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_bf16_ss(t_tile, da + 2 * k, db + 2 * k, idesc_s, k != 0);
    umma_commit(&bar_s[wg]);
  }
  mbar_wait(&bar_s[wg], 0);
  tc_fence_after();

  // ---- row softmax over the 197 real keys (columns >= 197 are zero-filled K rows: masked)
  float mx = -INFINITY;
#pragma unroll 1
  for (int c = 0; c < 6; ++c) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(t_row + c * 32, r);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(r[j]));
  }
  {
    uint32_t r[16];
    tmem_ld_32x32b_x16(t_row + 192, r);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < kS - 192; ++j) mx = fmaxf(mx, __uint_as_float(r[j]));
  }
  const float mxs = mx * kLog2e;
  float sum = 0.0f;
#pragma unroll 1
  for (int c = 0; c < 6; ++c) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(t_row + c * 32, r);
    tmem_ld_wait();
    float e[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      e[j] = fast_ex2(fmaf(__uint_as_float(r[j]), kLog2e, -mxs));
      sum += e[j];
    }
    uint8_t* blk = p_row + (c >> 1) * kPBytes;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint32_t chunk = static_cast<uint32_t>((c & 1) * 4 + g);
      *reinterpret_cast<uint4*>(blk + ((chunk ^ sw) << 4)) =
          make_uint4(pack_bf16x2(e[8 * g], e[8 * g + 1]), pack_bf16x2(e[8 * g + 2], e[8 * g + 3]),
                     pack_bf16x2(e[8 * g + 4], e[8 * g + 5]), pack_bf16x2(e[8 * g + 6], e[8 * g + 7]));
    }
  }
  {
    uint32_t r[16];
    tmem_ld_32x32b_x16(t_row + 192, r);
    tmem_ld_wait();
    float e[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      e[j] = (j < kS - 192) ? fast_ex2(fmaf(__uint_as_float(r[j]), kLog2e, -mxs)) : 0.0f;
      sum += e[j];
    }
    uint8_t* blk = p_row + 3 * kPBytes;
#pragma unroll
    for (int g = 0; g < 2; ++g)
      *reinterpret_cast<uint4*>(blk + ((static_cast<uint32_t>(g) ^ sw) << 4)) =
          make_uint4(pack_bf16x2(e[8 * g], e[8 * g + 1]), pack_bf16x2(e[8 * g + 2], e[8 * g + 3]),
                     pack_bf16x2(e[8 * g + 4], e[8 * g + 5]), pack_bf16x2(e[8 * g + 6], e[8 * g + 7]));
  }
  const float inv = 1.0f / sum;
  fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
  tc_fence_before();
  asm volatile("bar.sync %0, 128;" ::"r"(1 + wg) : "memory");  // this warpgroup only: P complete, S columns dead

  // ---- O = P V  (128 x 64, K = 208 = 3 x 64 + 16), accumulated over the first 64 columns of this tile's S
  if (wt == 0) {
    tc_fence_after();
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64);
#pragma unroll
    for (int kb = 0; kb < 4; ++kb) {
      const uint64_t da = umma_smem_desc_kmajor<128>(smem_u32(sP + kb * kPBytes));
      const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sV + kb * kVBytes));
      const int nk = kb < 3 ? 4 : 1;
      for (int k = 0; k < nk; ++k) umma_bf16_ss(t_tile, da + 2 * k, db + 2 * k, idesc_o, (kb | k) != 0);
    }
    umma_commit(&bar_o[wg]);
  }
  mbar_wait(&bar_o[wg], 0);
  tc_fence_after();

  const int t = wg * 128 + wt;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(t_row + c * 32, r);
    tmem_ld_wait();
    if (t < kS) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + ((size_t)b * kS + t) * MDE_VIT_DIM + hd * 64 + c * 32);
#pragma unroll
      for (int g = 0; g < 4; ++g)
        dst[g] = make_uint4(pack_bf16x2(__uint_as_float(r[8 * g]) * inv, __uint_as_float(r[8 * g + 1]) * inv),
                            pack_bf16x2(__uint_as_float(r[8 * g + 2]) * inv, __uint_as_float(r[8 * g + 3]) * inv),
                            pack_bf16x2(__uint_as_float(r[8 * g + 4]) * inv, __uint_as_float(r[8 * g + 5]) * inv),
                            pack_bf16x2(__uint_as_float(r[8 * g + 6]) * inv, __uint_as_float(r[8 * g + 7]) * inv));
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------ ViT attention, probabilities kept in TMEM
// One CTA per (image, head, 128-row query tile), 128 threads, 75 KB of shared memory and 256 TMEM columns, so TWO CTAs
// share an SM and one's softmax overlaps the other's loads and MMAs (the shared-memory-P kernel above needs 225 KB and
// runs alone).  S = Q K^T lands in TMEM columns [0,208); each thread (= query row = TMEM lane) reads its row twice (max,
// then exp), and writes the un-normalised bf16 probabilities back into TMEM columns [104,208) — two per 32-bit column,
// processed from the last chunk to the first so a chunk's probabilities only ever overwrite scores that are already in
// registers.  O = P V then runs with the A operand read from TMEM (tcgen05.mma [d], [a], b-desc) and accumulates into
// columns [0,64), which the scores no longer need.  No P tile in shared memory, no proxy fence.
constexpr int kTpOffK = kQBytes;
constexpr int kTpOffV = kTpOffK + kKBytesPad;
constexpr int kTpOffBar = kTpOffV + 4 * kVBytes;
constexpr int kTpSmem = kTpOffBar + 64 + 1024;
constexpr uint32_t kTpTmemCols = 256;
constexpr uint32_t kTpP = 104;  // first TMEM column of P (208 keys x bf16 = 104 columns)

#ifdef MDE_ATTN_PROF
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define PS_T(i) do { if (prof_on) tp[i] = gtime(); } while (0)
#else
#define PS_T(i) do { } while (0)
#endif

__global__ void __launch_bounds__(128, 2) vit_attn_tp_kernel(const __grid_constant__ AttnParams p) {
#ifdef MDE_ATTN_PROF
  unsigned long long tp[12] = {0};
  const bool prof_on = threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == 295 || blockIdx.x == 300 || blockIdx.x == 700);
  PS_T(0);
#endif
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kTpOffK;
  uint8_t* sV = smem + kTpOffV;
  uint64_t* bar_load = reinterpret_cast<uint64_t*>(smem + kTpOffBar);  // Q and K have landed
  uint64_t* bar_s = bar_load + 1;
  uint64_t* bar_o = bar_load + 2;
  uint64_t* bar_v = bar_load + 3;  // V^T has landed (needed only by P V: S = Q K^T and the softmax start without it)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_load + 4);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int slab = blockIdx.x >> 1, mt = blockIdx.x & 1;  // (image, head), query tile
  const int b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&p.tmQ);
      tma_prefetch_desc(&p.tmK);
      tma_prefetch_desc(&p.tmV);
      mbar_init(bar_load, 1);
      mbar_init(bar_s, 1);
      mbar_init(bar_o, 1);
      mbar_init(bar_v, 1);
      fence_barrier_init();
      // the loads go out before the TMEM allocation and the block-wide sync below: their round trip is the head of the CTA's chain
      pdl_grid_wait();
      mbar_expect_tx(bar_load, kQBytes + kKBytes);
      tma_load_4d(sQ, &p.tmQ, bar_load, 0, hd, mt * 128, b);
      tma_load_4d(sK, &p.tmK, bar_load, 0, hd, 0, b);
      mbar_expect_tx(bar_v, 4 * kVBytes);
#pragma unroll
      for (int j = 0; j < 4; ++j) tma_load_3d(sV + j * kVBytes, &p.tmV, bar_v, 64 * j, 0, slab);
    }
    __syncwarp();
    tmem_alloc(tmem_slot, kTpTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  PS_T(1);
  pdl_grid_wait();  // the set-up above overlaps the previous kernel's tail
  PS_T(2);

  if (tid == 0) {
    mbar_wait(bar_load, 0);
    PS_T(3);
    tc_fence_after();
    // ---- S = Q K^T  (128 x 208, K = 64)
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, kKeys);
    const uint64_t da = umma_smem_desc_kmajor<128>(smem_u32(sQ));
    const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sK));
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem_base, da + 2 * k, db + 2 * k, idesc_s, k != 0);
    umma_commit(bar_s);
  }
  mbar_wait(bar_s, 0);
  tc_fence_after();
  PS_T(4);

  const uint32_t t_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);  // this thread's TMEM lane
  constexpr float kLog2e = 1.4426950408889634f;

  // ---- pass 1: row maximum over the 197 real keys (four independent chains: a thread is alone with its row, so the
  // dependent-issue latency of one 197-long fmax chain was ~0.4 us of every CTA)
  // Both passes keep the TMEM load of the NEXT 32-column chunk in flight while the current one is reduced / exponentiated (a thread
  // is alone with its row: an un-overlapped tcgen05.ld round trip per chunk was a third of the second pass).
  float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
  uint32_t tail[16];
  {
    uint32_t r[2][32];
    tmem_ld_32x32b_x32(t_row, r[0]);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      tmem_ld_wait();
      if (c + 1 < 6) tmem_ld_32x32b_x32(t_row + (c + 1) * 32, r[(c + 1) & 1]);
      else tmem_ld_32x32b_x16(t_row + 192, tail);
#pragma unroll
      for (int j = 0; j < 32; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(r[c & 1][j]));
    }
    tmem_ld_wait();
  }
#pragma unroll
  for (int j = 0; j < kS - 192; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(tail[j]));
  const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
  const float mxs = mx * kLog2e;
  PS_T(5);

  // ---- pass 2, last chunk first: exp, row sum, bf16 probabilities back into TMEM
  float s4[4] = {0.0f, 0.0f, 0.0f, 0.0f};  // partial row sums (fixed assignment of columns to partials: deterministic)
  {
    uint32_t r[2][32];
    tmem_ld_32x32b_x32(t_row + 5 * 32, r[1]);  // chunk 5 on its way while the tail is exponentiated
    uint32_t pk[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float e0 = (2 * j < kS - 192) ? fast_ex2(fmaf(__uint_as_float(tail[2 * j]), kLog2e, -mxs)) : 0.0f;
      const float e1 = (2 * j + 1 < kS - 192) ? fast_ex2(fmaf(__uint_as_float(tail[2 * j + 1]), kLog2e, -mxs)) : 0.0f;
      s4[j & 3] += e0 + e1;
      pk[j] = pack_bf16x2(e0, e1);
    }
    tmem_st_32x32b_x8(t_row + kTpP + 96, pk);
#pragma unroll
    for (int c = 5; c >= 0; --c) {
      tmem_ld_wait();
      // the next chunk's scores (columns 32 (c-1) ..) lie below everything written so far (P columns >= 104 + 16 c)
      if (c > 0) tmem_ld_32x32b_x32(t_row + (c - 1) * 32, r[(c - 1) & 1]);
      uint32_t pc[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float e0 = fast_ex2(fmaf(__uint_as_float(r[c & 1][2 * j]), kLog2e, -mxs));
        const float e1 = fast_ex2(fmaf(__uint_as_float(r[c & 1][2 * j + 1]), kLog2e, -mxs));
        s4[j & 3] += e0 + e1;
        pc[j] = pack_bf16x2(e0, e1);
      }
      tmem_st_32x32b_x16(t_row + kTpP + 16 * c, pc);  // columns >= 104 + 16c: scores there (>= 32c) are already in registers
    }
  }
  tmem_st_wait();
  const float inv = 1.0f / ((s4[0] + s4[1]) + (s4[2] + s4[3]));
  tc_fence_before();
  __syncthreads();
  PS_T(6);

  // ---- O = P V  (128 x 64, K = 208 = 13 x 16): A from TMEM, B = V^T tiles in shared memory
  if (tid == 0) {
    mbar_wait(bar_v, 0);
    tc_fence_after();
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64);
#pragma unroll
    for (int kk = 0; kk < 13; ++kk) {
      const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sV + (kk >> 2) * kVBytes)) + 2 * (kk & 3);
      umma_bf16_ts(tmem_base, tmem_base + kTpP + 8 * kk, db, idesc_o, kk != 0);
    }
    umma_commit(bar_o);
  }
  mbar_wait(bar_o, 0);
  tc_fence_after();
  PS_T(7);

  // O * (1 / rowsum) -> bf16 into the Q tile's shared memory (Q was consumed by S = Q K^T), in the 128-byte swizzle the store
  // expects; ONE TMA store writes the 128 x 64 tile as full lines and clips the rows past token 196 (the per-thread stores it
  // replaces touched 32 different lines per instruction).
  {
    uint8_t* orow = sQ + tid * 128;
    const uint32_t sw = static_cast<uint32_t>(tid & 7);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(t_row + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g)
        *reinterpret_cast<uint4*>(orow + ((static_cast<uint32_t>(4 * c + g) ^ sw) << 4)) =
            make_uint4(pack_bf16x2(__uint_as_float(r[8 * g]) * inv, __uint_as_float(r[8 * g + 1]) * inv),
                       pack_bf16x2(__uint_as_float(r[8 * g + 2]) * inv, __uint_as_float(r[8 * g + 3]) * inv),
                       pack_bf16x2(__uint_as_float(r[8 * g + 4]) * inv, __uint_as_float(r[8 * g + 5]) * inv),
                       pack_bf16x2(__uint_as_float(r[8 * g + 6]) * inv, __uint_as_float(r[8 * g + 7]) * inv));
    }
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tma_store_4d(&p.tmO, sQ, 0, hd, mt * 128, b);
    bulk_commit();
    bulk_wait_exit<0>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) tmem_dealloc(tmem_base, kTpTmemCols);
#ifdef MDE_ATTN_PROF
  if (prof_on) {
    PS_T(8);
    printf("tp blk %d: setup %llu pdlwait %llu load %llu S %llu pass1 %llu pass2 %llu PV %llu epi+dealloc %llu total %llu ns (start %llu)\n", blockIdx.x, tp[1] - tp[0],
           tp[2] - tp[1], tp[3] - tp[2], tp[4] - tp[3], tp[5] - tp[4], tp[6] - tp[5], tp[7] - tp[6], tp[8] - tp[7], tp[8] - tp[0], tp[0] % 1000000ull);
  }
#endif
}

// ------------------------------------------------------------------ ViT attention, probabilities in TMEM, TWO threads per query row
// vit_attn_tp_kernel with 256 threads: warps w and w + 4 share TMEM lane quarter w & 3 (a warp may only touch the lanes of its
// quarter), and each takes half of the row's 208 score columns — half 0 keys [0,96), half 1 keys [96,208).  A thread alone with its
// row is a ~400-instruction dependent chain per pass; two threads halve it and give every scheduler four softmax warps per CTA pair
// instead of two.  The halves exchange the row maximum (one named barrier per lane quarter) and the row sum through shared memory.
// Probabilities: half 0 writes keys [0,96) into the 48 unused columns [208,256); half 1 writes keys [96,208) into [152,208), over
// scores that only half 1 reads and has already taken (last chunk first), so the halves never wait for each other inside a pass.
constexpr int kT2OffX = kTpOffBar + 64;                   // float [2][2][128]: row maximum, row sum per half
constexpr int kT2Smem = kT2OffX + 2048 + 1024;
constexpr uint32_t kT2P0 = 208, kT2P1 = 152;             // first TMEM column of the probabilities of half 0 / half 1

__global__ void __launch_bounds__(256, 2) vit_attn_tp2_kernel(const __grid_constant__ AttnParams p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kTpOffK;
  uint8_t* sV = smem + kTpOffV;
  uint64_t* bar_load = reinterpret_cast<uint64_t*>(smem + kTpOffBar);  // Q and K have landed
  uint64_t* bar_s = bar_load + 1;
  uint64_t* bar_o = bar_load + 2;
  uint64_t* bar_v = bar_load + 3;  // V^T has landed (needed only by P V)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_load + 4);
  float* xmax = reinterpret_cast<float*>(smem + kT2OffX);  // [2][128]
  float* xsum = xmax + 256;                                // [2][128]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qt = warp & 3, half = warp >> 2;  // TMEM lane quarter, column half (warp-uniform)
  const int row = qt * 32 + lane;             // query row of the tile = TMEM lane
  const int slab = blockIdx.x >> 1, mt = blockIdx.x & 1;
  const int b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&p.tmQ);
      tma_prefetch_desc(&p.tmK);
      tma_prefetch_desc(&p.tmV);
      mbar_init(bar_load, 1);
      mbar_init(bar_s, 1);
      mbar_init(bar_o, 1);
      mbar_init(bar_v, 1);
      fence_barrier_init();
      // the loads go out before the TMEM allocation and the block-wide sync below: their round trip is the head of the CTA's chain
      pdl_grid_wait();
      mbar_expect_tx(bar_load, kQBytes + kKBytes);
      tma_load_4d(sQ, &p.tmQ, bar_load, 0, hd, mt * 128, b);
      tma_load_4d(sK, &p.tmK, bar_load, 0, hd, 0, b);
      mbar_expect_tx(bar_v, 4 * kVBytes);
#pragma unroll
      for (int j = 0; j < 4; ++j) tma_load_3d(sV + j * kVBytes, &p.tmV, bar_v, 64 * j, 0, slab);
    }
    __syncwarp();
    tmem_alloc(tmem_slot, kTpTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_grid_wait();

  if (tid == 0) {
    mbar_wait(bar_load, 0);
    tc_fence_after();
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, kKeys);
    const uint64_t da = umma_smem_desc_kmajor<128>(smem_u32(sQ));
    const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sK));
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem_base, da + 2 * k, db + 2 * k, idesc_s, k != 0);
    umma_commit(bar_s);
  }
  mbar_wait(bar_s, 0);
  tc_fence_after();

  const uint32_t t_row = tmem_base + (static_cast<uint32_t>(qt * 32) << 16);
  const uint32_t t_s = t_row + half * 96;  // this half's three 32-column score chunks; half 1 also owns the tail [192,197)
  constexpr float kLog2e = 1.4426950408889634f;
  constexpr int kTail = kS - 192;  // 5 real keys in the last chunk

  // ---- pass 1: maximum over this half's keys, next chunk's tcgen05.ld in flight while the current one is reduced
  float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
  {
    uint32_t r[2][32];
    tmem_ld_32x32b_x32(t_s, r[0]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      tmem_ld_wait();
      if (c + 1 < 3) tmem_ld_32x32b_x32(t_s + (c + 1) * 32, r[(c + 1) & 1]);
#pragma unroll
      for (int j = 0; j < 32; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(r[c & 1][j]));
    }
    if (half) {
      uint32_t tl[8];
      tmem_ld_32x32b_x8(t_row + 192, tl);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < kTail; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(tl[j]));
    }
  }
  float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
  xmax[half * 128 + row] = mx;
  switch (qt) {  // the two warps of this lane quarter meet on their own named barrier (immediate ids: 5 barriers per CTA, not 16)
    case 0: named_bar_sync<1, 64>(); break;
    case 1: named_bar_sync<2, 64>(); break;
    case 2: named_bar_sync<3, 64>(); break;
    default: named_bar_sync<4, 64>(); break;
  }
  mx = fmaxf(mx, xmax[(half ^ 1) * 128 + row]);
  const float mxs = mx * kLog2e;

  // ---- pass 2, last chunk first: exp, partial row sum, bf16 probabilities into this half's TMEM columns
  float s4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  {
    const uint32_t t_p = t_row + (half ? kT2P1 : kT2P0);
    uint32_t r[2][32];
    tmem_ld_32x32b_x32(t_s + 64, r[0]);
    if (half) {  // keys [192,208): five real ones, the rest zero — P columns [200,208), over scores nobody needs
      uint32_t tl[8];
      tmem_ld_32x32b_x8(t_row + 192, tl);
      tmem_ld_wait();  // (also completes the chunk load above)
      uint32_t pk[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float e0 = (2 * j < kTail) ? fast_ex2(fmaf(__uint_as_float(tl[(2 * j) & 7]), kLog2e, -mxs)) : 0.0f;
        const float e1 = (2 * j + 1 < kTail) ? fast_ex2(fmaf(__uint_as_float(tl[(2 * j + 1) & 7]), kLog2e, -mxs)) : 0.0f;
        s4[j & 3] += e0 + e1;
        pk[j] = pack_bf16x2(e0, e1);
      }
      tmem_st_32x32b_x8(t_row + kT2P1 + 48, pk);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int c = 2 - i;
      tmem_ld_wait();
      if (c > 0) tmem_ld_32x32b_x32(t_s + (c - 1) * 32, r[(i + 1) & 1]);
      uint32_t pc[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float e0 = fast_ex2(fmaf(__uint_as_float(r[i & 1][2 * j]), kLog2e, -mxs));
        const float e1 = fast_ex2(fmaf(__uint_as_float(r[i & 1][2 * j + 1]), kLog2e, -mxs));
        s4[j & 3] += e0 + e1;
        pc[j] = pack_bf16x2(e0, e1);
      }
      // half 1: columns [152 + 16c, +16) hold scores of its chunks c.. (already in registers) or of the tail (done above)
      tmem_st_32x32b_x16(t_p + 16 * c, pc);
    }
  }
  tmem_st_wait();
  xsum[half * 128 + row] = (s4[0] + s4[1]) + (s4[2] + s4[3]);
  tc_fence_before();
  __syncthreads();

  // ---- O = P V  (128 x 64, K = 208 = 13 x 16): A from TMEM (k-steps 0..5 at [208,256), 6..12 at [152,208)), B = V^T in shared memory
  if (tid == 0) {
    mbar_wait(bar_v, 0);
    tc_fence_after();
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64);
#pragma unroll
    for (int kk = 0; kk < 13; ++kk) {
      const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sV + (kk >> 2) * kVBytes)) + 2 * (kk & 3);
      const uint32_t a_col = kk < 6 ? kT2P0 + 8 * kk : kT2P1 + 8 * (kk - 6);
      umma_bf16_ts(tmem_base, tmem_base + a_col, db, idesc_o, kk != 0);
    }
    umma_commit(bar_o);
  }
  const float inv = 1.0f / (xsum[row] + xsum[128 + row]);  // half 0 + half 1 in both threads of the row: same bits
  mbar_wait(bar_o, 0);
  tc_fence_after();

  // O * (1 / row sum) -> bf16 into the Q tile's shared memory (128-byte swizzle), this half's 32 of the 64 columns; one TMA store
  {
    uint8_t* orow = sQ + row * 128;
    const uint32_t sw = static_cast<uint32_t>(row & 7);
    uint32_t r[32];
    tmem_ld_32x32b_x32(t_row + half * 32, r);
    tmem_ld_wait();
#pragma unroll
    for (int g = 0; g < 4; ++g)
      *reinterpret_cast<uint4*>(orow + ((static_cast<uint32_t>(4 * half + g) ^ sw) << 4)) =
          make_uint4(pack_bf16x2(__uint_as_float(r[8 * g]) * inv, __uint_as_float(r[8 * g + 1]) * inv),
                     pack_bf16x2(__uint_as_float(r[8 * g + 2]) * inv, __uint_as_float(r[8 * g + 3]) * inv),
                     pack_bf16x2(__uint_as_float(r[8 * g + 4]) * inv, __uint_as_float(r[8 * g + 5]) * inv),
                     pack_bf16x2(__uint_as_float(r[8 * g + 6]) * inv, __uint_as_float(r[8 * g + 7]) * inv));
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tma_store_4d(&p.tmO, sQ, 0, hd, mt * 128, b);
    bulk_commit();
    bulk_wait_exit<0>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) tmem_dealloc(tmem_base, kTpTmemCols);
}

// ------------------------------------------------------------------ ViT attention, persistent and pipelined
// One CTA per SM loops over (image, head) pairs.  320 threads: warpgroup t (warps 4t..4t+3) owns query tile t (rows 128t..)
// and TMEM columns [256t, 256t+208); warp 8 is the TMA producer (Q tiles, K and V^T of the NEXT pair land in the other
// shared-memory stage while this one is being used: K and V are loaded once for both query tiles); one thread of warp 9
// issues every MMA.  Each tile runs the vit_attn_tp_kernel scheme (S in TMEM, un-normalised bf16 P written back into TMEM
// columns [104,208) of the tile, O = P V as a TS-mode MMA into columns [0,64)).  The MMA thread is event-driven: it polls
// the two tiles' barriers and issues whichever step (S of the tile's next pair, or P V) has become possible, so the two
// warpgroups drift out of phase and one tile's MMA / TMEM round trips hide behind the other tile's exponentials.
constexpr int kPsStage = 2 * kQBytes + kKBytesPad + 4 * kVBytes;   // 93184
constexpr int kPsOffBar = 2 * kPsStage;
constexpr int kPsSmem = kPsOffBar + 256 + 1024;
constexpr uint32_t kPsTile = 256;  // TMEM column stride between the two query tiles

__device__ __forceinline__ void ps_poll_timeout(long long t0) {
  if (clock64() - t0 > 4000000000LL) {
    printf("mde: vit_attn_ps_kernel MMA scheduler timed out (block %d)\n", blockIdx.x);
    __trap();
  }
}


__global__ void __launch_bounds__(320, 1) vit_attn_ps_kernel(const __grid_constant__ AttnParams p) {
#ifdef MDE_ATTN_PROF
  unsigned long long tp[12] = {0};
  const bool prof_on = (threadIdx.x == 0 || threadIdx.x == 128) && (blockIdx.x == 0 || blockIdx.x == 100);
  PS_T(0);
#endif
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kPsOffBar);
  uint64_t* bar_full = bars;        // [2] stage loaded (TMA transaction bytes)
  uint64_t* bar_empty = bars + 2;   // [2] both tiles' MMAs of the pair in this stage have completed
  uint64_t* bar_s = bars + 4;       // [2] per tile: S = Q K^T is in TMEM
  uint64_t* bar_p = bars + 6;       // [2] per tile: P written (4 warp arrivals)
  uint64_t* bar_o = bars + 8;       // [2] per tile: O = P V is in TMEM
  uint64_t* bar_free = bars + 10;   // [2] per tile: O read out, the tile's columns may be overwritten (4 warp arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slabs = p.B * MDE_VIT_HEADS;
  const int n_items = (slabs - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 9) {
    if (elect_one()) {
      tma_prefetch_desc(&p.tmQ);
      tma_prefetch_desc(&p.tmK);
      tma_prefetch_desc(&p.tmV);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&bar_full[i], 1);
        mbar_init(&bar_empty[i], 1);
        mbar_init(&bar_s[i], 1);
        mbar_init(&bar_p[i], 4);
        mbar_init(&bar_o[i], 1);
        mbar_init(&bar_free[i], 4);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  PS_T(1);
  pdl_grid_wait();  // the set-up above overlaps the previous kernel's tail
  PS_T(2);

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int it = 0; it < n_items; ++it) {
        const int s = it & 1;
        if (it >= 2) mbar_wait(&bar_empty[s], ((it >> 1) - 1) & 1);
        const int slab = blockIdx.x + it * gridDim.x;
        const int b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;
        uint8_t* st = smem + s * kPsStage;
        mbar_expect_tx(&bar_full[s], 2 * kQBytes + kKBytes + 4 * kVBytes);
        tma_load_4d(st, &p.tmQ, &bar_full[s], 0, hd, 0, b);
        tma_load_4d(st + kQBytes, &p.tmQ, &bar_full[s], 0, hd, 128, b);
        tma_load_4d(st + 2 * kQBytes, &p.tmK, &bar_full[s], 0, hd, 0, b);
#pragma unroll
        for (int j = 0; j < 4; ++j) tma_load_3d(st + 2 * kQBytes + kKBytesPad + j * kVBytes, &p.tmV, &bar_full[s], 64 * j, 0, slab);
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer (event-driven over the two tiles)
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, kKeys);
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64);
      int item[2] = {0, 0};
      int need_pv[2] = {0, 0};
      int released = 0;  // stages handed back to the producer so far (pairs whose P V of both tiles has been issued)
      const long long t0 = clock64();
      while (item[0] < n_items || item[1] < n_items) {
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const int i = item[t];
          if (i >= n_items) continue;
          const uint32_t t_tile = tmem_base + t * kPsTile;
          uint8_t* st = smem + (i & 1) * kPsStage;
          if (!need_pv[t]) {
            if (!mbar_try_wait(&bar_full[i & 1], (i >> 1) & 1)) continue;
            if (i > 0 && !mbar_try_wait(&bar_free[t], (i - 1) & 1)) continue;
            tc_fence_after();
            const uint64_t da = umma_smem_desc_kmajor<128>(smem_u32(st + t * kQBytes));
            const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(st + 2 * kQBytes));
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16_ss(t_tile, da + 2 * k, db + 2 * k, idesc_s, k != 0);
            umma_commit(&bar_s[t]);
            need_pv[t] = 1;
          } else {
            if (!mbar_try_wait(&bar_p[t], i & 1)) continue;
            tc_fence_after();
            uint8_t* sV = st + 2 * kQBytes + kKBytesPad;
#pragma unroll
            for (int kk = 0; kk < 13; ++kk) {
              const uint64_t db = umma_smem_desc_kmajor<128>(smem_u32(sV + (kk >> 2) * kVBytes)) + 2 * (kk & 3);
              umma_bf16_ts(t_tile, t_tile + kTpP + 8 * kk, db, idesc_o, kk != 0);
            }
            umma_commit(&bar_o[t]);
            need_pv[t] = 0;
            item[t] = i + 1;
            const int done = item[0] < item[1] ? item[0] : item[1];
            while (released < done) {
              umma_commit(&bar_empty[released & 1]);
              ++released;
            }
          }
        }
        ps_poll_timeout(t0);
      }
    }
  } else {
    // ------------------------------------------------------------ softmax / output warpgroups
    const int t = warp >> 2;  // query tile
    const uint32_t t_row = tmem_base + t * kPsTile + (static_cast<uint32_t>((warp & 3) * 32) << 16);  // this thread's TMEM lane
    const int row = t * 128 + (warp & 3) * 32 + lane;  // token index of this thread's query row
    const bool warp_live = t * 128 + (warp & 3) * 32 < kS;  // warp 7 holds rows 224..255 only: nothing to compute
    constexpr float kLog2e = 1.4426950408889634f;
    for (int it = 0; it < n_items; ++it) {
      const int slab = blockIdx.x + it * gridDim.x;
      const int b = slab / MDE_VIT_HEADS, hd = slab - b * MDE_VIT_HEADS;
      mbar_wait(&bar_s[t], it & 1);
      tc_fence_after();
      if (it == 0) PS_T(3);
      if (it == 1) PS_T(8);
      float inv = 0.0f;
      if (warp_live) {
        // pass 1: row maximum over the 197 real keys; the TMEM load of chunk c+1 is in flight while chunk c is reduced
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        uint32_t ra[32], rb[32];
        tmem_ld_32x32b_x32(t_row, ra);
#pragma unroll
        for (int c = 0; c < 6; c += 2) {
          tmem_ld_wait();
          tmem_ld_32x32b_x32(t_row + (c + 1) * 32, rb);
#pragma unroll
          for (int j = 0; j < 32; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(ra[j]));
          tmem_ld_wait();
          if (c + 2 < 6) tmem_ld_32x32b_x32(t_row + (c + 2) * 32, ra);
#pragma unroll
          for (int j = 0; j < 32; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(rb[j]));
        }
        uint32_t tail[16];
        tmem_ld_32x32b_x16(t_row + 192, tail);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < kS - 192; ++j) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(tail[j]));
        const float mxs = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * kLog2e;
        if (it == 0) PS_T(4);

        // pass 2, last chunk first: exp, row sum, bf16 probabilities back into TMEM columns [104,208) of the tile
        float s4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        tmem_ld_32x32b_x32(t_row + 5 * 32, ra);
        {
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float e0 = (2 * j < kS - 192) ? fast_ex2(fmaf(__uint_as_float(tail[2 * j]), kLog2e, -mxs)) : 0.0f;
            const float e1 = (2 * j + 1 < kS - 192) ? fast_ex2(fmaf(__uint_as_float(tail[2 * j + 1]), kLog2e, -mxs)) : 0.0f;
            s4[j & 3] += e0 + e1;
            pk[j] = pack_bf16x2(e0, e1);
          }
          tmem_st_32x32b_x8(t_row + kTpP + 96, pk);
        }
#pragma unroll
        for (int c = 5; c >= 0; c -= 2) {
          uint32_t pk[16];
          tmem_ld_wait();
          tmem_ld_32x32b_x32(t_row + (c - 1) * 32, rb);  // scores of chunk c-1: P of chunk c only overwrites columns >= 104 + 16c >= 32c
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float e0 = fast_ex2(fmaf(__uint_as_float(ra[2 * j]), kLog2e, -mxs));
            const float e1 = fast_ex2(fmaf(__uint_as_float(ra[2 * j + 1]), kLog2e, -mxs));
            s4[j & 3] += e0 + e1;
            pk[j] = pack_bf16x2(e0, e1);
          }
          tmem_ld_wait();  // chunk c-1 is in registers before chunk c's probabilities may land on its columns
          tmem_st_32x32b_x16(t_row + kTpP + 16 * c, pk);
          if (c - 2 >= 0) tmem_ld_32x32b_x32(t_row + (c - 2) * 32, ra);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float e0 = fast_ex2(fmaf(__uint_as_float(rb[2 * j]), kLog2e, -mxs));
            const float e1 = fast_ex2(fmaf(__uint_as_float(rb[2 * j + 1]), kLog2e, -mxs));
            s4[j & 3] += e0 + e1;
            pk[j] = pack_bf16x2(e0, e1);
          }
          if (c - 2 >= 0) tmem_ld_wait();
          tmem_st_32x32b_x16(t_row + kTpP + 16 * (c - 1), pk);
        }
        tmem_st_wait();
        inv = 1.0f / ((s4[0] + s4[1]) + (s4[2] + s4[3]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[t]);
      if (it == 0) PS_T(5);

      mbar_wait(&bar_o[t], it & 1);
      tc_fence_after();
      if (it == 0) PS_T(6);
      uint32_t o0[32], o1[32];
      if (warp_live) {
        tmem_ld_32x32b_x32(t_row, o0);
        tmem_ld_32x32b_x32(t_row + 32, o1);
        tmem_ld_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_free[t]);  // the MMA thread may start the next pair's S on this tile
      if (warp_live && row < kS) {
        uint4* dst = reinterpret_cast<uint4*>(p.out + ((size_t)b * kS + row) * MDE_VIT_DIM + hd * 64);
#pragma unroll
        for (int g = 0; g < 4; ++g)
          dst[g] = make_uint4(pack_bf16x2(__uint_as_float(o0[8 * g]) * inv, __uint_as_float(o0[8 * g + 1]) * inv),
                              pack_bf16x2(__uint_as_float(o0[8 * g + 2]) * inv, __uint_as_float(o0[8 * g + 3]) * inv),
                              pack_bf16x2(__uint_as_float(o0[8 * g + 4]) * inv, __uint_as_float(o0[8 * g + 5]) * inv),
                              pack_bf16x2(__uint_as_float(o0[8 * g + 6]) * inv, __uint_as_float(o0[8 * g + 7]) * inv));
#pragma unroll
        for (int g = 0; g < 4; ++g)
          dst[4 + g] = make_uint4(pack_bf16x2(__uint_as_float(o1[8 * g]) * inv, __uint_as_float(o1[8 * g + 1]) * inv),
                                  pack_bf16x2(__uint_as_float(o1[8 * g + 2]) * inv, __uint_as_float(o1[8 * g + 3]) * inv),
                                  pack_bf16x2(__uint_as_float(o1[8 * g + 4]) * inv, __uint_as_float(o1[8 * g + 5]) * inv),
                                  pack_bf16x2(__uint_as_float(o1[8 * g + 6]) * inv, __uint_as_float(o1[8 * g + 7]) * inv));
      }
    }
  }
  PS_T(7);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 9) tmem_dealloc(tmem_base, 512);
#ifdef MDE_ATTN_PROF
  if (prof_on) {
    PS_T(9);
    printf("ps blk %d thr %d: setup %llu pdlwait %llu S0 %llu pass1 %llu pass2 %llu PV %llu rest %llu (S1 at %llu) dealloc %llu total %llu ns\n", blockIdx.x, threadIdx.x,
           tp[1] - tp[0], tp[2] - tp[1], tp[3] - tp[2], tp[4] - tp[3], tp[5] - tp[4], tp[6] - tp[5], tp[7] - tp[6], tp[8] ? tp[8] - tp[2] : 0ull, tp[9] - tp[7], tp[9] - tp[0]);
  }
#endif
}

// ------------------------------------------------------------------ spatial self-attention (enhanced_model.py:20-39)
constexpr int kSaQ = 16;       // queries per block (two per warp): every block streams all of V once, so fewer, fatter blocks halve the L2 traffic
constexpr int kSaMaxG = 224;   // grid positions supported (14x15 = 210, 14x14 = 196)
constexpr int kSaLd = 384;     // row pitch of the fp32 q|k|v projection buffer

__global__ void __launch_bounds__(256) spatial_attn_kernel(const bf16* __restrict__ x, const float* __restrict__ qkv, int gh, int gw,
                                                           const float* __restrict__ gamma_p, bf16* __restrict__ out) {
  pdl_launch_dependents();
  pdl_grid_wait();
  __shared__ float sK[kSaMaxG][33];
  __shared__ float sQ[kSaQ][32];
  __shared__ float sA[kSaQ][kSaMaxG];
  const int G = gh * gw;
  const int b = blockIdx.y, q0 = blockIdx.x * kSaQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* base = qkv + (size_t)b * G * kSaLd;
  for (int i = tid; i < G * 32; i += 256) sK[i >> 5][i & 31] = base[(size_t)(i >> 5) * kSaLd + 32 + (i & 31)];
  for (int t = tid; t < kSaQ * 32; t += 256) {
    const int i = t >> 5, d = t & 31;
    sQ[i][d] = (q0 + i < G) ? base[(size_t)(q0 + i) * kSaLd + d] : 0.0f;
  }
  __syncthreads();
  // energy[i][j] = q_i . k_j (no scaling, enhanced_model.py:28); warp w owns queries w and w + 8
  for (int i = warp; i < kSaQ; i += 8) {
    float mx = -INFINITY;
    for (int j = lane; j < G; j += 32) {
      float e = 0.0f;
#pragma unroll
      for (int d = 0; d < 32; ++d) e = fmaf(sQ[i][d], sK[j][d], e);
      sA[i][j] = e;
      mx = fmaxf(mx, e);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.0f;
    for (int j = lane; j < G; j += 32) {
      const float e = __expf(sA[i][j] - mx);
      sA[i][j] = e;
      sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.0f / sum;
    for (int j = lane; j < G; j += 32) sA[i][j] *= inv;
  }
  __syncthreads();
  // out[i][c] = sum_j att[i][j] * v[j][c]; thread = channel c (coalesced v reads)
  float acc[kSaQ];
#pragma unroll
  for (int i = 0; i < kSaQ; ++i) acc[i] = 0.0f;
  const float* vcol = base + 64 + tid;
#pragma unroll 4
  for (int j = 0; j < G; ++j) {
    const float v = __ldg(vcol + (size_t)j * kSaLd);
#pragma unroll
    for (int i = 0; i < kSaQ; ++i) acc[i] = fmaf(sA[i][j], v, acc[i]);
  }
  const float gamma = *gamma_p;
  const int H2 = gh + 2, W2 = gw + 2;
#pragma unroll  // fully unrolled: acc[] stays in registers (a rolled loop indexes it dynamically -> local memory)
  for (int i = 0; i < kSaQ; ++i) {
    const int g = q0 + i;
    if (g >= G) break;
    const float xin = __bfloat162float(x[((size_t)b * G + g) * 256 + tid]);
    const bf16 v = __float2bfloat16_rn(fmaf(gamma, acc[i], xin));
    const int y = g / gw, xx = g - y * gw;
    int ys[3], xs[3], ny = 0, nx = 0;
    ys[ny++] = y + 1;
    if (y == 1) ys[ny++] = 0;
    if (y == gh - 2) ys[ny++] = gh + 1;
    xs[nx++] = xx + 1;
    if (xx == 1) xs[nx++] = 0;
    if (xx == gw - 2) xs[nx++] = gw + 1;
    for (int iy = 0; iy < ny; ++iy)
      for (int ix = 0; ix < nx; ++ix) out[(((size_t)b * H2 + ys[iy]) * W2 + xs[ix]) * 256 + tid] = v;
  }
}

}  // namespace mde

using namespace mde;

extern "C" int mde_vit_attention(const mde_bf16* q, const mde_bf16* k, const mde_bf16* vt, int B, mde_bf16* out, mde_stream_t stream) {
  MDE_CHECK_ARG(q && k && vt && out && B > 0, "mde_vit_attention: bad arguments");
  MDE_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)vt | (uintptr_t)out) % 16 == 0, "mde_vit_attention: pointers must be 16-byte aligned");
  AttnParams p;
  memset(&p, 0, sizeof(p));
  const uint64_t slabs = (uint64_t)B * MDE_VIT_HEADS;
  {
    // token-major q / k: element (d, head, t, b) at ((b*197 + t)*768 + head*64 + d); rows t >= 197 of the box are zero-filled
    const uint64_t dims[4] = {64, MDE_VIT_HEADS, (uint64_t)kS, (uint64_t)B};
    const uint64_t strides[3] = {128, (uint64_t)MDE_VIT_DIM * 2, (uint64_t)kS * MDE_VIT_DIM * 2};
    const uint32_t boxq[4] = {64, 1, 128, 1};
    const uint32_t boxk[4] = {64, 1, (uint32_t)kKeys, 1};
    int rc = encode_tmap_bf16(&p.tmQ, q, 4, dims, strides, boxq, 128);
    if (rc != MDE_OK) return rc;
    rc = encode_tmap_bf16(&p.tmK, k, 4, dims, strides, boxk, 128);
    if (rc != MDE_OK) return rc;
    rc = encode_tmap_bf16(&p.tmO, out, 4, dims, strides, boxq, 128);  // the output has q's layout
    if (rc != MDE_OK) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)kS, 64, slabs};
    const uint64_t strides[2] = {(uint64_t)MDE_VT_PITCH * 2, (uint64_t)64 * MDE_VT_PITCH * 2};
    const uint32_t box[3] = {64, 64, 1};
    int rc = encode_tmap_bf16(&p.tmV, vt, 3, dims, strides, box, 128);
    if (rc != MDE_OK) return rc;
  }
  p.out = reinterpret_cast<bf16*>(out);
  p.B = B;
  static bool attr_set = false;
  if (!attr_set) {
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_tp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTpSmem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_tp_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_ps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPsSmem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_tp2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kT2Smem));
    MDE_CHECK_CUDA(cudaFuncSetAttribute(vit_attn_tp2_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  // MDE_ATTN_SMEM_P=1 selects the shared-memory-P kernel (A/B testing); default: probabilities in TMEM, 2 CTAs per SM
  static const bool smem_p = getenv("MDE_ATTN_SMEM_P") != nullptr && getenv("MDE_ATTN_SMEM_P")[0] == '1';
  // MDE_ATTN=ps selects the persistent kernel (K/V loaded once for both query tiles; A/B testing: 19.7 us against 18.1 us
  // at B=32, its 384 (image, head) pairs are 2.6 waves of 148 CTAs); default: one query tile per CTA, 2 CTAs per SM
  static const bool persistent = getenv("MDE_ATTN") != nullptr && strcmp(getenv("MDE_ATTN"), "ps") == 0;
  // MDE_ATTN_TP2=1 selects the two-threads-per-row kernel (A/B testing: 2.0990 ms per step against 2.0919 with one thread per row)
  static const bool two_per_row = getenv("MDE_ATTN_TP2") != nullptr && getenv("MDE_ATTN_TP2")[0] == '1';
  if (smem_p) launch_k(vit_attn_kernel, dim3((unsigned)slabs), dim3(256), kAttnSmem, stream, p);
  else if (!persistent && two_per_row) launch_k(vit_attn_tp2_kernel, dim3((unsigned)(2 * slabs)), dim3(256), kT2Smem, stream, p);
  else if (!persistent) launch_k(vit_attn_tp_kernel, dim3((unsigned)(2 * slabs)), dim3(128), kTpSmem, stream, p);
  else launch_k(vit_attn_ps_kernel, dim3((unsigned)(slabs < (uint64_t)sm_count() ? slabs : (uint64_t)sm_count())), dim3(320), kPsSmem, stream, p);
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}

extern "C" int mde_spatial_attention(const mde_bf16* x, const float* qkv, int B, int gh, int gw, const float* gamma, mde_bf16* out_padded,
                                     mde_stream_t stream) {
  MDE_CHECK_ARG(x && qkv && gamma && out_padded && B > 0 && gh >= 3 && gw >= 3, "mde_spatial_attention: bad arguments");
  MDE_CHECK_ARG(gh * gw <= kSaMaxG, "mde_spatial_attention: grid %dx%d exceeds %d positions", gh, gw, kSaMaxG);
  launch_k(spatial_attn_kernel, dim3(ceil_div(gh * gw, kSaQ), B), dim3(256), 0, stream, reinterpret_cast<const bf16*>(x), qkv, gh, gw, gamma,
                                                                            reinterpret_cast<bf16*>(out_padded));
  MDE_CHECK_LAUNCH();
  return MDE_OK;
}
