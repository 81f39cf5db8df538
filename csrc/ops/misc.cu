// Error reporting and device queries shared by every launcher in libzoo_cuda.so.
#include "../common/zoo_common.cuh"
#include <string.h>

static thread_local char g_last_error[512] = "";

void zoo_set_last_error(const char* msg, const char* file, int line) {
  snprintf(g_last_error, sizeof(g_last_error), "%s (%s:%d)", msg ? msg : "unknown", file ? file : "?", line);
}

ZOO_API const char* zoo_last_error() { return g_last_error; }

ZOO_API int zoo_abi_version() { return 1; }

// Synchronous device check used by the Python loader: returns the SM count, or a negative
// CUDA error code if no usable device is present.
ZOO_API int zoo_device_sm_count() {
  int dev = 0, sms = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return -(int)e;
  e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return -(int)e;
  return sms;
}
