// Putting a class-partitioned mining result back together (multi-GPU jobs).
//
// Every rank mined the subtrees of its own frequent pairs (kmls_mine with n_parts > 1): its
// tables of sizes >= 3 hold only those subtrees, sorted, with parent indices local to the rank.
// The job's table of level k is the ranks' itemsets merged in the order of their ROOT pair
// (the frequent pair an itemset's first two members form) - which is exactly the sorted order
// a single GPU produces, because all itemsets under one pair live on one rank. These kernels
// compute the roots, count itemsets per root (the host sums those counts over the ranks and
// scans them) and write a rank's itemsets at their global positions into zero-initialised
// job-wide tables, which a sum-allreduce then completes on every rank.
// No counterpart in the reference (it has no parallelism).
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

__global__ void __launch_bounds__(256)
k_subtree_roots(const int32_t* __restrict__ par, int64_t n, const int32_t* __restrict__ root_prev,
                int32_t* __restrict__ root, int32_t* __restrict__ per_root) {
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < n; o += (int64_t)gridDim.x * blockDim.x) {
    const int32_t p = par[o];
    const int32_t r = root_prev ? __ldg(root_prev + p) : p;
    root[o] = r;
    // itemsets of one root are contiguous, so most lanes of a warp hold the same root: one
    // atomic per distinct root in the warp instead of one per lane
    const uint32_t active = __activemask();
    const uint32_t same = __match_any_sync(active, r);
    if ((int)(threadIdx.x & 31) == __ffs(same) - 1) atomicAdd(per_root + r, __popc(same));
  }
}

// g[o] = global_start[root[o]] + (o - local_start[root[o]]); tables written at g[o];
// parent: a pair index as is (gprev == nullptr, level 3), else the parent's global position
__global__ void __launch_bounds__(256)
k_place_level(const int32_t* __restrict__ items, const int32_t* __restrict__ counts,
              const int32_t* __restrict__ par, int64_t n, int k, const int32_t* __restrict__ root,
              const int64_t* __restrict__ global_start, const int64_t* __restrict__ local_start,
              const int32_t* __restrict__ gprev, int32_t* __restrict__ items_out,
              int32_t* __restrict__ counts_out, int32_t* __restrict__ par_out, int32_t* __restrict__ g_out) {
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < n; o += (int64_t)gridDim.x * blockDim.x) {
    const int32_t r = root[o];
    const int64_t g = __ldg(global_start + r) + (o - __ldg(local_start + r));
    g_out[o] = (int32_t)g;
    for (int t = 0; t < k; ++t) items_out[g * k + t] = items[o * k + t];
    counts_out[g] = counts[o];
    par_out[g] = gprev ? __ldg(gprev + par[o]) : par[o];
  }
}

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_subtree_roots(const int32_t* d_par, int64_t n, const int32_t* d_root_prev,
                                  int32_t* d_root, int32_t* d_per_root, kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0, "kmls_subtree_roots: negative size");
  if (n == 0) return 0;
  KMLS_REQUIRE(d_par && d_root && d_per_root, "kmls_subtree_roots: null pointer");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_subtree_roots<<<grid, 256, 0, (cudaStream_t)stream>>>(d_par, n, d_root_prev, d_root, d_per_root);
  KMLS_LAUNCHED("k_subtree_roots");
  return 0;
}

extern "C" int kmls_place_level(const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par,
                                int64_t n, int32_t k, const int32_t* d_root, const int64_t* d_global_start,
                                const int64_t* d_local_start, const int32_t* d_gprev, int32_t* d_items_out,
                                int32_t* d_counts_out, int32_t* d_par_out, int32_t* d_g_out,
                                kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && k >= 3, "kmls_place_level: bad arguments");
  if (n == 0) return 0;
  KMLS_REQUIRE(d_items && d_counts && d_par && d_root && d_global_start && d_local_start && d_items_out &&
                   d_counts_out && d_par_out && d_g_out, "kmls_place_level: null pointer");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_place_level<<<grid, 256, 0, (cudaStream_t)stream>>>(d_items, d_counts, d_par, n, k, d_root, d_global_start,
                                                       d_local_start, d_gprev, d_items_out, d_counts_out,
                                                       d_par_out, d_g_out);
  KMLS_LAUNCHED("k_place_level");
  return 0;
}

// ---------------------------------------------------------------------------------------
// Compact form of the merge: an itemset is its parent plus one member, so the ranks exchange
// only (last member, count, parent) per itemset - 12 bytes instead of 4k + 8 - and every rank
// rebuilds the member lists level by level from the merged parent links.
// ---------------------------------------------------------------------------------------
namespace kmls {

__global__ void __launch_bounds__(256)
k_place_level_compact(const int32_t* __restrict__ items, const int32_t* __restrict__ counts,
                      const int32_t* __restrict__ par, int64_t n, int k, const int32_t* __restrict__ root,
                      const int64_t* __restrict__ global_start, const int64_t* __restrict__ local_start,
                      const int32_t* __restrict__ gprev, int32_t* __restrict__ last_out,
                      int32_t* __restrict__ counts_out, int32_t* __restrict__ par_out, int32_t* __restrict__ g_out) {
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < n; o += (int64_t)gridDim.x * blockDim.x) {
    const int32_t r = root[o];
    const int64_t g = __ldg(global_start + r) + (o - __ldg(local_start + r));
    g_out[o] = (int32_t)g;
    last_out[g] = items[o * k + (k - 1)];
    counts_out[g] = counts[o];
    par_out[g] = gprev ? __ldg(gprev + par[o]) : par[o];
  }
}

// items[g] = items_prev[par[g]] ++ last[g]
__global__ void __launch_bounds__(256)
k_rebuild_items(const int32_t* __restrict__ items_prev, int k, const int32_t* __restrict__ par,
                const int32_t* __restrict__ last, int64_t n, int32_t* __restrict__ items) {
  const int64_t total = n * k;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t g = e / k;
    const int t = (int)(e - g * k);
    items[e] = t < k - 1 ? __ldg(items_prev + (int64_t)__ldg(par + g) * (k - 1) + t) : __ldg(last + g);
  }
}

}  // namespace kmls

extern "C" int kmls_place_level_compact(const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par,
                                        int64_t n, int32_t k, const int32_t* d_root,
                                        const int64_t* d_global_start, const int64_t* d_local_start,
                                        const int32_t* d_gprev, int32_t* d_last_out, int32_t* d_counts_out,
                                        int32_t* d_par_out, int32_t* d_g_out, kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && k >= 3, "kmls_place_level_compact: bad arguments");
  if (n == 0) return 0;
  KMLS_REQUIRE(d_items && d_counts && d_par && d_root && d_global_start && d_local_start && d_last_out &&
                   d_counts_out && d_par_out && d_g_out, "kmls_place_level_compact: null pointer");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_place_level_compact<<<grid, 256, 0, (cudaStream_t)stream>>>(d_items, d_counts, d_par, n, k, d_root,
                                                               d_global_start, d_local_start, d_gprev, d_last_out,
                                                               d_counts_out, d_par_out, d_g_out);
  KMLS_LAUNCHED("k_place_level_compact");
  return 0;
}

extern "C" int kmls_rebuild_items(const int32_t* d_items_prev, int32_t k, const int32_t* d_par,
                                  const int32_t* d_last, int64_t n, int32_t* d_items, kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && k >= 2, "kmls_rebuild_items: bad arguments");
  if (n == 0) return 0;
  KMLS_REQUIRE(d_items_prev && d_par && d_last && d_items, "kmls_rebuild_items: null pointer");
  int grid = (int)std::min<int64_t>((n * k + 255) / 256, (int64_t)num_sms() * 16);
  k_rebuild_items<<<grid, 256, 0, (cudaStream_t)stream>>>(d_items_prev, k, d_par, d_last, n, d_items);
  KMLS_LAUNCHED("k_rebuild_items");
  return 0;
}
