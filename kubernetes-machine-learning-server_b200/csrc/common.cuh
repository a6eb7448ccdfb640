// Shared helpers for the kmls_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

namespace kmls {

// Last error text of the calling host thread; read through kmls_last_error().
char* error_buffer();
int fail(int code, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* what);

constexpr int kArgError = -1;       // bad argument (sizes, alignment, null pointer)
constexpr int kCapacityError = -2;  // an output or scratch buffer is too small (never truncates)

}  // namespace kmls

#define KMLS_CUDA(expr)                                        \
  do {                                                         \
    cudaError_t e_ = (expr);                                   \
    if (e_ != cudaSuccess) return kmls::fail_cuda(e_, #expr);  \
  } while (0)

#define KMLS_LAUNCHED(name)                                    \
  do {                                                         \
    cudaError_t e_ = cudaGetLastError();                       \
    if (e_ != cudaSuccess) return kmls::fail_cuda(e_, name);   \
  } while (0)

#define KMLS_REQUIRE(cond, ...)                                        \
  do {                                                                 \
    if (!(cond)) return kmls::fail(kmls::kArgError, __VA_ARGS__);      \
  } while (0)

namespace kmls {

__device__ __forceinline__ int popc4(uint4 v) {
  return __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
}
__device__ __forceinline__ uint4 and4(uint4 a, uint4 b) {
  return make_uint4(a.x & b.x, a.y & b.y, a.z & b.z, a.w & b.w);
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Read-only 128-bit load that stays out of L1's way for streaming data.
__device__ __forceinline__ uint4 ldg4(const uint4* p) { return __ldg(p); }

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// Order-free fingerprint pieces: an itemset's key is finish(sum of item_term over members),
// so the key of "itemset minus one member" is one subtraction away.
__host__ __device__ __forceinline__ uint64_t item_term(int32_t item) {
  return mix64((uint64_t)(uint32_t)item + 0x51ED270B1ull);
}
__host__ __device__ __forceinline__ uint64_t finish_key(uint64_t sum, int k) {
  return mix64(sum ^ ((uint64_t)k << 56)) | 1ull;  // never 0 (0 marks an empty slot)
}

// SM count of the CURRENT device, cached per device ordinal (a process may use several GPUs).
inline int num_sms() {
  constexpr int kMaxDevices = 64;
  static std::atomic<int> cache[kMaxDevices];
  int dev = 0;
  cudaGetDevice(&dev);
  const bool cacheable = dev >= 0 && dev < kMaxDevices;
  int n = cacheable ? cache[dev].load(std::memory_order_relaxed) : 0;
  if (n == 0) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    if (cacheable) cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

}  // namespace kmls
