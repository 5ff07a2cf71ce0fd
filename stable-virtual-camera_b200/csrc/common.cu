// Host-side plumbing shared by all entry points: thread-local error string, device check,
// cuTensorMapEncodeTiled resolved through the runtime (no link-time libcuda dependency).
#include "common.cuh"

#include <cudaTypedefs.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "../../include/seva_b200.h"

namespace seva {

static thread_local char g_err[512] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
    set_error("cuTensorMapEncodeTiled unavailable: %s", cudaGetErrorString(e));
    return nullptr;
  }
  fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box) {
  return make_tmap_bf16_sw(out, base, rank, dims, strides_bytes, box, 128);
}

int make_tmap_bf16_sw(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
  auto enc = get_encode();
  if (!enc) return 1;
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                   gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                   : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                   : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d): rank=%d dims=[%llu,%llu,%llu] "
              "stride0=%llu box=[%u,%u,%u] base=%p",
              (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 2 ? dims[2] : 0),
              (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0], rank > 1 ? box[1] : 0,
              rank > 2 ? box[2] : 0, base);
    return 1;
  }
  return 0;
}

struct DevInfo {
  int valid = 0, major = 0, minor = 0, sms = 0;
};
static DevInfo& dev_info() {
  static thread_local DevInfo cache[16];
  static DevInfo bad;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return bad;
  DevInfo& d = cache[dev];
  if (!d.valid) {
    cudaDeviceGetAttribute(&d.major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&d.minor, cudaDevAttrComputeCapabilityMinor, dev);
    cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
    d.valid = 1;
  }
  return d;
}

int ensure_dyn_smem(const void* func, int bytes) {
  struct Entry { const void* f; int dev; int bytes; };
  static std::mutex mu;
  static std::vector<Entry> done;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    set_error("cudaGetDevice failed");
    return 2;
  }
  std::lock_guard<std::mutex> lock(mu);
  for (const Entry& e : done)
    if (e.f == func && e.dev == dev && e.bytes >= bytes) return 0;
  cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(MaxDynamicSharedMemorySize=%d) failed on device %d: %s", bytes, dev,
              cudaGetErrorString(e));
    return 2;
  }
  done.push_back({func, dev, bytes});
  return 0;
}

int sm_count() { return dev_info().sms > 0 ? dev_info().sms : 148; }

int check_device() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    set_error("seva_b200: no CUDA device visible; these kernels are sm_100a-only and have no CPU fallback");
    return 1;
  }
  DevInfo& d = dev_info();
  if (!d.valid || d.major != 10) {
    set_error("seva_b200: device is sm_%d%d; this library is built for sm_100a (B200) only", d.major,
              d.minor);
    return 1;
  }
  return 0;
}

}  // namespace seva

extern "C" const char* seva_last_error(void) { return seva::g_err; }
extern "C" int seva_check_device(void) { return seva::check_device(); }
extern "C" int seva_version(void) { return 1; }
