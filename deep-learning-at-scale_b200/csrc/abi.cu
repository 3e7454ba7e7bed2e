// Library-level entry points: version, thread-local error text, device gate.
#include "common.cuh"
#include <string.h>

namespace dlas {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// No CPU fallback and no other architecture: the kernels are built for sm_100a only.
int device_ok() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("no CUDA device: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return DLAS_ERR_UNSUPPORTED_DEVICE;
  }
  static int cached[64];  // 0 unknown, 1 ok, 2 bad
  if (dev >= 0 && dev < 64 && cached[dev]) {
    if (cached[dev] == 1) return DLAS_OK;
    set_error("device %d is not compute capability 10.x (B200); libdlas_b200 has no fallback path", dev);
    return DLAS_ERR_UNSUPPORTED_DEVICE;
  }
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) {
    set_error("cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
    return DLAS_ERR_CUDA;
  }
  bool ok = (major == 10);
  if (dev >= 0 && dev < 64) cached[dev] = ok ? 1 : 2;
  if (!ok) {
    set_error("device %d has compute capability %d.x; libdlas_b200 needs 10.x (B200) and has no fallback path", dev, major);
    return DLAS_ERR_UNSUPPORTED_DEVICE;
  }
  return DLAS_OK;
}

int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  static int cached[64];
  if (dev >= 0 && dev < 64 && cached[dev]) return cached[dev];
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  if (dev >= 0 && dev < 64) cached[dev] = n;
  return n;
}

}  // namespace dlas

extern "C" {
int dlas_version(void) { return 100; }  // 0.1.0
const char* dlas_last_error(void) { return dlas::g_err; }
int dlas_device_check(void) { return dlas::device_ok(); }
int dlas_sm_count(void) { return dlas::sm_count(); }
}
