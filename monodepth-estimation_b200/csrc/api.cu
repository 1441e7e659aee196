// libmde.so: library-level entry points and host helpers (error text, device checks, TMA
// tensor-map encoding through the driver entry point so the library does not link libcuda).
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <cudaTypedefs.h>

#include <stdlib.h>

#include "common.h"

namespace mde {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (fn) return fn;
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess || ptr == nullptr)
    return nullptr;
  fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
  return fn;
}

bool pdl_enabled() {
  // Programmatic dependent launch on every kernel (MDE_PDL=0 turns it off).  Each kernel triggers its dependents at once and
  // waits for its predecessor only AFTER its own set-up (barrier init, TMEM allocation, descriptor prefetch), so launch
  // latency and prologue overlap the predecessor's tail.  Measured (r01x, B=32 1024 px): 2.388 ms/step with, 2.490 without;
  // with the wait at the top of the kernels (r01s) it was a loss.
  static const bool on = !(getenv("MDE_PDL") != nullptr && getenv("MDE_PDL")[0] == '0');
  return on;
}

int encode_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box, int swizzle_bytes) {
  return encode_tmap(out, 2, base, rank, dims, strides_bytes, box, swizzle_bytes);
}

int encode_tmap(CUtensorMap* out, int elt_bytes, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, int swizzle_bytes) {
  PFN_cuTensorMapEncodeTiled_v12000 fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point unavailable (driver too old?)");
    return MDE_ERR_CUDA;
  }
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                : CU_TENSOR_MAP_SWIZZLE_32B;
  cuuint64_t gdims[5];
  cuuint64_t gstrides[4];
  cuuint32_t gbox[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstrides[i] = strides_bytes[i];
  CUresult r = fn(out, elt_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdims, gstrides, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu %llu, box %u %u, swizzle %d)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0], rank > 1 ? box[1] : 0,
              swizzle_bytes);
    return MDE_ERR_CUDA;
  }
  return MDE_OK;
}

}  // namespace mde

extern "C" int mde_version(void) { return 100; }

extern "C" const char* mde_last_error(void) { return mde::g_err; }

extern "C" int mde_sm_count(void) { return mde::sm_count(); }

extern "C" int mde_device_check(void) {
  int dev = 0, major = 0;
  MDE_CHECK_CUDA(cudaGetDevice(&dev));
  MDE_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) {
    mde::set_error("libmde.so is built for sm_100a only; current device has compute capability %d.x", major);
    return MDE_ERR_DEVICE;
  }
  return MDE_OK;
}
