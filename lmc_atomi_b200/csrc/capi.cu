// C-ABI plumbing shared by all entry points: error string, version, device query.
#include <stdarg.h>
#include <string.h>

#include "lmc_common.cuh"

static thread_local char g_err[512] = "";

void lmc_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* lmc_last_error(void) { return g_err; }

extern "C" int lmc_version(void) { return 100; }

extern "C" int lmc_device_sm_count(void) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    lmc_set_error("no CUDA device: %s", cudaGetErrorString(cudaGetLastError()));
    return -2;
  }
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    lmc_set_error("cudaDeviceGetAttribute failed: %s", cudaGetErrorString(cudaGetLastError()));
    return -2;
  }
  return sms;
}
