// libsmi: version, error state, device check.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void smi_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

SMI_API int smi_version(void) { return SMI_VERSION; }

SMI_API const char* smi_last_error(void) { return g_err; }

SMI_API int smi_device_check(void) {
  int dev = 0;
  SMI_CUDA_OK(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  SMI_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  SMI_CUDA_OK(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) SMI_FAIL(SMI_EARCH, "libsmi needs sm_100 (Blackwell B200); device is sm_%d%d", major, minor);
  return SMI_OK;
}

int smi_num_sms() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess) cached = n;
    cached_dev = dev;
  }
  return cached;
}
