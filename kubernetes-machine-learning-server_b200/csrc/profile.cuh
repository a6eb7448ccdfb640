// Per-kernel device timing inside the one-call entry points, for bench.py's roofline objects.
// Off by default (one branch per launch). When switched on (kmls_profile_begin), every kernel
// launch that kmls_mine / kmls_rules make is bracketed by two CUDA events on the launching
// stream together with the algorithmic bytes (or operations) the call site states;
// kmls_profile_report synchronises and returns the records as text. Per host thread.
#pragma once
#include <string>
#include <vector>

#include "common.cuh"

namespace kmls {

struct ProfRecord {
  const char* name;
  cudaEvent_t e0, e1;
  double work;      // algorithmic bytes (or int8 operations for the tensor-core GEMM)
  int level;
};

struct Profiler {
  bool on = false;
  bool work = true;   // also account the data-dependent work figures (extra small kernels and host
                      // synchronisations per level): off for the timed steps of a benchmark
  std::vector<ProfRecord> records;
  std::vector<cudaEvent_t> pool;
  cudaEvent_t get() {
    if (!pool.empty()) {
      cudaEvent_t e = pool.back();
      pool.pop_back();
      return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
  }
};

Profiler& profiler();

// RAII bracket around one launch (or one library call that launches a known kernel)
struct ProfScope {
  Profiler& p;
  cudaStream_t s;
  int idx = -1;
  ProfScope(const char* name, cudaStream_t stream, double work, int level = 0) : p(profiler()), s(stream) {
    if (!p.on) return;
    ProfRecord r{name, p.get(), p.get(), work, level};
    cudaEventRecord(r.e0, s);
    idx = (int)p.records.size();
    p.records.push_back(r);
  }
  ~ProfScope() {
    if (idx >= 0) cudaEventRecord(p.records[idx].e1, s);
  }
};

}  // namespace kmls
