// Library-level entry points: ABI version, device check, status strings.
#include "common.cuh"
#include <cstdlib>

int ocv_require_sm100() {
  static int cached = 1;  // 1 = unknown
  if (cached <= 0) return cached;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) return (int)e;
  cached = (major == 10) ? OCV_OK : OCV_E_ARCH;
  return cached;
}

bool ocv_pdl_enabled() {
  static int cached = -1;
  if (cached < 0) {
    const char* e = getenv("OCV_NO_PDL");
    cached = (e && e[0] == '1') ? 0 : 1;
  }
  return cached == 1;
}

static const unsigned long long* g_dropout_salt = nullptr;
const unsigned long long* ocv_dropout_salt() { return g_dropout_salt; }
extern "C" int ocv_set_dropout_salt(const unsigned long long* device_salt) {
  g_dropout_salt = device_salt;
  return OCV_OK;
}
// Clears the process-wide salt only when `owner` still owns it (several trainers may be alive in one process).
extern "C" int ocv_release_dropout_salt(const unsigned long long* owner) {
  if (g_dropout_salt == owner) g_dropout_salt = nullptr;
  return OCV_OK;
}

extern "C" int ocv_abi_version(void) { return 1; }

extern "C" int ocv_device_check(int device) {
  int major = 0;
  cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  if (e != cudaSuccess) return (int)e;
  return major == 10 ? OCV_OK : OCV_E_ARCH;
}

extern "C" const char* ocv_status_string(int status) {
  switch (status) {
    case OCV_OK: return "ok";
    case OCV_E_ARG: return "invalid argument (null pointer or non-positive size)";
    case OCV_E_SHAPE: return "unsupported shape";
    case OCV_E_ALIGN: return "misaligned pointer or stride";
    case OCV_E_ARCH: return "device is not sm_100 (no fallback path exists)";
    case OCV_E_DRIVER: return "CUDA driver entry point unavailable";
    case OCV_E_CAPACITY: return "size exceeds the fixed capacity built into a no-allocation entry point";
    default: return status > 0 ? cudaGetErrorString((cudaError_t)status) : "unknown status";
  }
}
