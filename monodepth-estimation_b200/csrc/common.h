// Internal host-side helpers shared by the .cu files of libmde.so (error reporting, TMA
// tensor-map encoding through the driver entry point, device properties).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/mde.h"

namespace mde {

// thread-local message behind mde_last_error()
void set_error(const char* fmt, ...);

#define MDE_CHECK_ARG(cond, ...)        \
  do {                                  \
    if (!(cond)) {                      \
      ::mde::set_error(__VA_ARGS__);    \
      return MDE_ERR_BAD_ARGUMENT;      \
    }                                   \
  } while (0)

#define MDE_CHECK_CUDA(expr)                                                                     \
  do {                                                                                           \
    cudaError_t err__ = (expr);                                                                  \
    if (err__ != cudaSuccess) {                                                                  \
      ::mde::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__), __FILE__, __LINE__); \
      return MDE_ERR_CUDA;                                                                       \
    }                                                                                            \
  } while (0)

// Launch-error check that does not synchronise (safe inside CUDA-graph capture).
#define MDE_CHECK_LAUNCH()                                                                       \
  do {                                                                                           \
    cudaError_t err__ = cudaGetLastError();                                                      \
    if (err__ != cudaSuccess) {                                                                  \
      ::mde::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(err__), __FILE__, __LINE__); \
      return MDE_ERR_CUDA;                                                                       \
    }                                                                                            \
  } while (0)

int sm_count();  // SMs of the current device (cached per device)

// Encode a bf16 tiled tensor map.  dims/box are innermost-first; strides_bytes has rank-1
// entries (strides of dims 1..rank-1).  swizzle_bytes in {32, 64, 128}.  Returns MDE_OK or error.
int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);
// Same for 4-byte (fp32) elements when elt_bytes == 4 (2 -> bf16).
int encode_tmap(CUtensorMap* out, int elt_bytes, const void* base, int rank, const uint64_t* dims,
                const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

// 2-CTA (cta_group::2) GEMM path of mde_gemm_bf16 (tc_gemm2.cu); same contract, needs N % 256 == 0.
// LayerNorm folded into the GEMMs on either side of it (mde_gemm_ln_bf16): a producer (residual / patch-embed epilogue)
// also writes a bf16 copy of its fp32 output and per-row partial sums; a consumer (QKV / fc1) reads A = that bf16 copy
// and applies (acc - mean*colsum) * rstd + bias in its epilogue.  All pointers optional (null = not used).
struct LnFuse {
  const float* stats_in;  // consumer: [K/32][M][2] (slot-major) partial (sum, sum of squares) per 32-column chunk of the normalised row
  const float* colsum;    // consumer: [N] sum over k of the (gamma-folded, bf16-rounded) weight row
  float eps;
  void* xb_out;           // producer: bf16 [M][N] copy of the fp32 output
  float* stats_out;       // producer: [N/32][M][2]
};
int launch_gemm_pair(const void* A, int lda, const void* W, int ldw, int M, int N, int K, int epilogue, const float* bias, void* out, int ldo,
                     const void* aux0, const void* aux1, const LnFuse* ln, cudaStream_t stream);

// Halo-resident 3x3 conv path of mde_conv_nhwc (conv_halo.cu).  MDE_OK = launched; MDE_ERR_UNSUPPORTED (no error text) =
// geometry not served there, the caller uses the tap-shifted kernel.
int launch_conv_halo(int f16, const void* in, int B, int h, int w, int cin, int in_cstride, const void* wpacked, const float* bias, int cout,
                     int padding, int activation, void* out_bf16, int out_cstride, int out_coff, float* out_f32, cudaStream_t stream);
// 3x3 conv on maps narrower than 48 px with cin % 64 == 0, cout % 32 == 0, bf16 output (conv_small.cu): halo tile + streamed weights.
// MDE_OK after launching, MDE_ERR_UNSUPPORTED (no error set) if the geometry is not served.
int launch_conv_small(int f16, const void* in, int B, int h, int w, int cin, int in_cstride, const void* wpacked, const float* bias, int cout,
                      int padding, int activation, void* out_bf16, int out_cstride, int out_coff, cudaStream_t stream);

// Kernel launch, with programmatic dependent launch unless MDE_PDL=0 (see pdl_enabled()).  Errors are picked up
// by MDE_CHECK_LAUNCH() at the call site.
bool pdl_enabled();
template <class... KArgs, class... Args>
static inline void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at;
  memset(&at, 0, sizeof(at));
  at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  (void)cudaLaunchKernelEx(&cfg, kernel, args...);
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace mde
