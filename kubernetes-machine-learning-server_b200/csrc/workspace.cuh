// Host-side helpers of the one-call entry points (kmls_mine / kmls_rules): a first-fit
// sub-allocator over the caller's device workspace and a pinned "mailbox" through which
// kernels hand a few sizes to the host between levels.
#pragma once
#include <cstdint>
#include <map>
#include <vector>

#include "common.cuh"

namespace kmls {

// Sub-allocator over [base, base + bytes): nothing here calls cudaMalloc. Results that
// outlive the call are taken from the low end, temporaries from the high end, so the long-
// lived tables do not fragment the space the per-level buffers come and go in.
class Workspace {
 public:
  Workspace(void* base, size_t bytes) : base_((char*)base) {
    const size_t lo = ((uintptr_t)base + kAlign - 1) / kAlign * kAlign - (uintptr_t)base;
    if (bytes > lo) free_[lo] = (bytes - lo) / kAlign * kAlign;
  }
  static constexpr size_t kAlign = 256;

  // nullptr when no free range is large enough (`failed_bytes` then holds the request)
  void* alloc(size_t bytes, bool high = false) {
    bytes = (std::max<size_t>(bytes, 1) + kAlign - 1) / kAlign * kAlign;
    size_t off = SIZE_MAX;
    if (!high) {
      for (auto& f : free_)
        if (f.second >= bytes) { off = f.first; break; }
    } else {
      for (auto it = free_.rbegin(); it != free_.rend(); ++it)
        if (it->second >= bytes) { off = it->first + it->second - bytes; break; }
    }
    if (off == SIZE_MAX) {
      failed_bytes = bytes;
      return nullptr;
    }
    // carve [off, off + bytes) out of the free range that contains it
    auto it = free_.upper_bound(off);
    --it;
    const size_t f_off = it->first, f_len = it->second;
    free_.erase(it);
    if (off > f_off) free_[f_off] = off - f_off;
    if (off + bytes < f_off + f_len) free_[off + bytes] = f_off + f_len - (off + bytes);
    used_[off] = bytes;
    in_use += bytes;
    peak = std::max(peak, in_use);
    return base_ + off;
  }

  void release(void* p) {
    if (!p) return;
    const size_t off = (char*)p - base_;
    auto u = used_.find(off);
    if (u == used_.end()) return;
    size_t len = u->second;
    used_.erase(u);
    in_use -= len;
    size_t o = off;
    auto nxt = free_.lower_bound(o);
    if (nxt != free_.end() && nxt->first == o + len) {   // merge with the range after
      len += nxt->second;
      nxt = free_.erase(nxt);
    }
    if (nxt != free_.begin()) {                          // merge with the range before
      auto prv = std::prev(nxt);
      if (prv->first + prv->second == o) {
        o = prv->first;
        len += prv->second;
        free_.erase(prv);
      }
    }
    free_[o] = len;
  }

  int64_t offset_of(const void* p) const { return p ? (const char*)p - base_ : -1; }
  size_t in_use = 0, peak = 0, failed_bytes = 0;

 private:
  char* base_;
  std::map<size_t, size_t> free_;   // offset -> length
  std::map<size_t, size_t> used_;
};

// 64 int64 words of pinned, device-visible host memory per host thread, allocated on the
// thread's first use (cudaHostAlloc - host memory, not device memory). A kernel writes the
// sizes the host needs next (k_post_*), the host synchronises the stream and reads them: no
// copy-engine transfer that could queue behind bulk device-to-host traffic.
volatile int64_t* mailbox();

}  // namespace kmls
