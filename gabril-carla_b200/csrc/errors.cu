// Status / lifecycle entry points of libgabril_b200.
#include <cstdio>
#include <cstring>

#include "common.cuh"

namespace {
thread_local char g_last_error[256] = "";
}

int gabril_set_cuda_error(cudaError_t e) {
  snprintf(g_last_error, sizeof(g_last_error), "CUDA error %d: %s", (int)e, cudaGetErrorString(e));
  return GABRIL_ERR_CUDA;
}

void gabril_set_error(const char* msg) { snprintf(g_last_error, sizeof(g_last_error), "%s", msg); }

extern "C" {

int gabril_version(void) { return 100; }

const char* gabril_last_error(void) { return g_last_error; }

int gabril_init(void) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return gabril_set_cuda_error(e);
  int major = 0, minor = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  if (major != 10) {
    snprintf(g_last_error, sizeof(g_last_error),
             "libgabril_b200 is built for sm_100a only; device %d is sm_%d%d", dev, major, minor);
    return GABRIL_ERR_UNSUPPORTED_ARCH;
  }
  return gabril_tc_init();
}

}  // extern "C"
