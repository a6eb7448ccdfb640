// Error reporting of the C-ABI: every entry point returns 0 on success, a positive
// cudaError_t value on a CUDA failure, or a negative kmls code; the text is kept per host
// thread and read through kmls_last_error().
#include <cstring>
#include <string>

#include "common.cuh"
#include "kmls_b200.h"
#include "profile.cuh"

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

namespace kmls {
Profiler& profiler() {
  static thread_local Profiler p;
  return p;
}
}  // namespace kmls

extern "C" int kmls_profile_begin(void) {
  kmls::Profiler& p = kmls::profiler();
  for (auto& r : p.records) {
    p.pool.push_back(r.e0);
    p.pool.push_back(r.e1);
  }
  p.records.clear();
  p.on = true;
  p.work = true;
  return 0;
}

extern "C" int kmls_profile_begin_events(void) {
  const int rc = kmls_profile_begin();
  kmls::profiler().work = false;
  return rc;
}

extern "C" int64_t kmls_profile_report(char* buf, size_t capacity) {
  kmls::Profiler& p = kmls::profiler();
  p.on = false;
  std::string out;
  char line[256];
  for (auto& r : p.records) {
    float ms = 0.f;
    cudaEventSynchronize(r.e1);
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) != cudaSuccess) {
      cudaGetLastError();
      ms = -1.f;
    }
    snprintf(line, sizeof(line), "%s,%d,%.6f,%.0f\n", r.name, r.level, (double)ms, r.work);
    out += line;
  }
  if (buf && capacity) {
    const size_t n = std::min(out.size(), capacity - 1);
    memcpy(buf, out.data(), n);
    buf[n] = 0;
  }
  return (int64_t)out.size() + 1;
}
