// Error reporting of the C-ABI: every entry point returns 0 on success, a positive
// cudaError_t value on a CUDA failure, or a negative kmls code; the text is kept per host
// thread and read through kmls_last_error().
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int fail_cuda(cudaError_t e, const char* what) {
  snprintf(error_buffer(), 512, "CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return (int)e;
}

}  // namespace kmls

extern "C" const char* kmls_last_error(void) { return kmls::error_buffer(); }
extern "C" int kmls_version(void) { return KMLS_B200_VERSION; }
