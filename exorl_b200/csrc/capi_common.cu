// Error plumbing + device info for the C ABI (include/exorl_b200.h).
#include "common.cuh"

namespace exb {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}
}  // namespace exb

extern "C" int exb_abi_version(void) { return 2; }
extern "C" const char* exb_last_error(void) { return exb::g_err; }
extern "C" int exb_device_info(int* sms, int* cc) {
  int dev = 0, n = 0, maj = 0, min = 0;
  EXB_CUDA(cudaGetDevice(&dev));
  EXB_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  EXB_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev));
  EXB_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev));
  if (sms) *sms = n;
  if (cc) *cc = maj * 10 + min;
  return EXB_OK;
}
