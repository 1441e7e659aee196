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

// Mirrors engine.Workspace buffer for buffer (tests/test_dropin_cpu.py compares the two on the meta device).
extern "C" long long mde_workspace_bytes(const mde_workspace_cfg* cfg) {
  if (!cfg || cfg->batch <= 0 || cfg->height <= 0 || cfg->width <= 0) {
    mde::set_error("mde_workspace_bytes: bad configuration");
    return MDE_ERR_BAD_ARGUMENT;
  }
  const long long B = cfg->batch;
  long long Hp = cfg->height, Wp = cfg->width;
  const bool padded = Hp % 16 || Wp % 16;
  if (padded) {
    Hp = (Hp / 16 + 1) * 16;
    Wp = (Wp / 16 + 1) * 16;
  }
  long long gh = Hp / 16, gw = Wp / 16;
  if (gh * gw == 197 || (Hp * Wp) % 4) {
    mde::set_error("mde_workspace_bytes: unsupported input size");
    return MDE_ERR_UNSUPPORTED;
  }
  if (gh * gw != 196) gh = 14, gw = 15;  // encoder.py:69-110 for 197 tokens
  const long long G = gh * gw, M = B * MDE_VIT_SEQ, P = (gh + 2) * (gw + 2);
  long long total = 0;
  auto add = [&](long long bytes) { total += (bytes + 255) / 256 * 256; };
  if (padded) add(B * 3 * Hp * Wp * 4);                 // xp
  add((long long)MDE_RANGE_SCRATCH_FLOATS * 4);          // scratch
  add(6 * 4);                                            // affine
  add(B * 196 * 768 * 2);                                // patches
  add(B * 3 * 224 * 224 * 4);                            // raw224
  for (int i = 0; i < 6; ++i) add(M * 768 * 4);          // xa, xb, taps[4]
  add(M * 768 * 2);                                      // h
  add(24 * M * 2 * 4);                                   // lnstats
  add(B * 12 * MDE_VIT_SEQ * 64 * 2);                    // q
  add(B * 12 * MDE_VIT_SEQ * 64 * 2);                    // k
  add(B * 12 * 64 * MDE_VT_PITCH * 2);                   // vt
  add(M * 768 * 2);                                      // attn
  add(M * 3072 * 2);                                     // mlp
  add(128 * 4);                                          // mlp_sync
  for (int i = 0; i < 4; ++i) add(B * G * 768 * 2);      // grid_a
  for (int i = 0; i < 4; ++i) add(B * G * 256 * 4);      // reduced
  add(B * P * (cfg->enhanced ? 320 : 256) * 2);          // feats[0]
  for (int i = 0; i < 3; ++i) add(B * P * 256 * 2);      // feats[1..3]
  add(B * P * 256 * 2);                                  // d_init
  static const int ch[5] = {256, 128, 64, 32, 16};
  long long h = gh, w = gw;
  for (int i = 0; i < 4; ++i) {
    const long long co = ch[i + 1];
    add(B * (h + 2) * (w + 2) * co * 2);                 // d_a
    if (i < 3) add(B * P * co * 2);                      // d_s
    h *= 2;
    w *= 2;
    add(B * (h + 2) * (w + 2) * co * (i < 3 ? 2 : 1) * 2);  // d_cat
    add(B * (h + 2) * (w + 2) * co * 2);                 // d_x
    add(B * h * w * 4);                                  // disp
  }
  add(B * (h + 2) * (w + 2) * 16 * 2);                   // d_x3b
  if (cfg->enhanced) {
    for (int i = 0; i < 2; ++i) {
      add(B * G * 256 * 2);                              // flat
      add(B * G * 384 * 4);                              // qkv
      add(B * P * 256 * 2);                              // enh
    }
    add(B * P * 128 * 2);                                // u1
    add(B * P * 128 * 2);                                // u2
    add(B * P * 256 * 2);                                // r1
    add(B * P * 128 * 2);                                // r2
    add(B * gh * gw * 4);                                // unc
    add(B * gh * gw * 4);                                // refn
    add(B * 4 * gh * gw * 4);                            // fused
  }
  return total;
}
