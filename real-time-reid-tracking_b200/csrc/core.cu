// Error reporting, ABI version and device check.
#include "common.cuh"

namespace reidb200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace reidb200

extern "C" const char* reidb200_last_error(void) { return reidb200::g_err; }

extern "C" int reidb200_abi_version(void) { return 1; }

extern "C" int reidb200_check_device(void) {
  int dev = 0;
  RB_CUDA(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  RB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  RB_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) {
    reidb200::set_error("device %d is sm_%d%d; libreidb200 is built for sm_100a only", dev, major, minor);
    return REIDB200_ECUDA;
  }
  return REIDB200_OK;
}
