// ABI plumbing: version, thread-local error string, device check.
#include <stdarg.h>

#include "common.cuh"

namespace pygb {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace pygb

extern "C" {

int pygb_abi_version(void) { return PYGB_ABI_VERSION; }

const char* pygb_last_error(void) { return pygb::g_err; }

int pygb_check_device(int* sm_count, int* cc_major, int* cc_minor) {
  int dev = 0;
  PYGB_CUDA(cudaGetDevice(&dev));
  int major = 0, minor = 0, sms = 0;
  PYGB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  PYGB_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  PYGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  if (sm_count) *sm_count = sms;
  if (cc_major) *cc_major = major;
  if (cc_minor) *cc_minor = minor;
  if (major != 10) {
    pygb::set_error("libpygb is built for sm_100a only; device %d is sm_%d%d", dev, major, minor);
    return PYGB_E_UNSUPPORTED;
  }
  return PYGB_OK;
}

}  // extern "C"
