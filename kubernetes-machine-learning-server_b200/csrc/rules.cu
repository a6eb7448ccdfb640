// Rule generation: confidence and lift in fp64 from the exact integer counts.
//
// The reference's live path computes no rules; its dormant path
// (machine-learning/main.py:224-260) is the only place a min_confidence knob exists. The
// semantics implemented are those of mlxtend.frequent_patterns.association_rules(df,
// metric="confidence", min_threshold=c): for every frequent X,
// |X| >= 2, and every non-empty proper subset A: C = X \ A,
//     support = cX / N, confidence = (cX / N) / (cA / N), lift = confidence / (cC / N),
// kept iff confidence >= c. The operation order above is mlxtend's (it divides the already
// rounded supports), so the doubles are bit-identical, not merely within 1e-12.
//
// All levels' tables live in one arena: items of level k start at item_off[k], its
// itemsets have global ids id_off[k] .. id_off[k+1]-1, counts are indexed by global id.
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

constexpr int kMaxLevels = 32;

struct Arena {
  const int32_t* items;
  const int32_t* counts;
  int64_t item_off[kMaxLevels + 2];
  int64_t id_off[kMaxLevels + 2];
  int max_k;
};

__device__ __forceinline__ const int32_t* itemset_ptr(const Arena& a, int k, int64_t gid) {
  return a.items + a.item_off[k] + (gid - a.id_off[k]) * k;
}

// Exact map: fingerprint -> global id. Distinct itemsets with the same fingerprint occupy
// different slots; a lookup verifies the members.
__global__ void __launch_bounds__(256) k_map_build(Arena a, unsigned long long* __restrict__ keys,
                                                   int32_t* __restrict__ vals, uint64_t mask) {
  const int64_t n = a.id_off[a.max_k + 1];
  for (int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gid < n;
       gid += (int64_t)gridDim.x * blockDim.x) {
    int k = 1;
    while (gid >= a.id_off[k + 1]) ++k;
    const int32_t* it = itemset_ptr(a, k, gid);
    uint64_t s = 0;
    for (int t = 0; t < k; ++t) s += item_term(__ldg(it + t));
    const uint64_t key = finish_key(s, k);
    uint64_t h = (key >> 20) & mask;
    while (atomicCAS(keys + h, 0ull, (unsigned long long)key) != 0ull) h = (h + 1) & mask;
    vals[h] = (int32_t)gid;
  }
}

// count of the subset of `x` (k members) selected by `sel`; -1 if it is not in the map
// (cannot happen for a downward-closed table; reported by the caller).
__device__ __forceinline__ int32_t subset_count(const Arena& a, const unsigned long long* __restrict__ keys,
                                                const int32_t* __restrict__ vals, uint64_t mask,
                                                const int32_t* __restrict__ x, int k, uint32_t sel,
                                                uint64_t sum) {
  const int r = __popc(sel);
  const uint64_t key = finish_key(sum, r);
  uint64_t h = (key >> 20) & mask;
  while (true) {
    const unsigned long long v = __ldg(keys + h);
    if (v == 0ull) return -1;
    if (v == key) {
      const int32_t gid = __ldg(vals + h);
      if (gid >= a.id_off[r] && gid < a.id_off[r + 1]) {
        const int32_t* it = itemset_ptr(a, r, gid);
        bool same = true;
        int j = 0;
        for (int t = 0; t < k && same; ++t)
          if (sel >> t & 1u) same = __ldg(it + j++) == __ldg(x + t);
        if (same) return __ldg(a.counts + gid);
      }
    }
    h = (h + 1) & mask;
  }
}

// K5. One warp per itemset X (global ids of levels >= 2). Lanes enumerate the selections
// `sel` with the top member excluded; each handles the partition {A = sel, C = complement}
// and therefore two rules (A -> C and C -> A) from two lookups. Pass 1 counts the kept rules
// per itemset; pass 2 writes them at rule_off[x] + rank (ballot + popc rank), ordered by
// sel, A -> C before C -> A. Only the itemsets x_begin .. x_begin + x_count - 1 (numbered
// from the first 2-itemset) are processed, so ranks can split the table; rule_cnt/rule_off
// are indexed from x_begin.
__global__ void __launch_bounds__(256)
k_rules(Arena a, const unsigned long long* __restrict__ keys, const int32_t* __restrict__ vals,
        uint64_t mask, double n_total, double min_conf, int64_t x_begin, int64_t x_count, int write,
        int32_t* __restrict__ rule_cnt,
        const int64_t* __restrict__ rule_off, int32_t* __restrict__ out_itemset,
        uint32_t* __restrict__ out_antecedent, double* __restrict__ out_support,
        double* __restrict__ out_confidence, double* __restrict__ out_lift,
        int32_t* __restrict__ missing) {
  const int lane = threadIdx.x & 31;
  const int64_t first = a.id_off[2] + x_begin;
  for (int64_t xi = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); xi < x_count;
       xi += (int64_t)gridDim.x * 8) {
    const int64_t gid = first + xi;
    int k = 2;
    while (gid >= a.id_off[k + 1]) ++k;
    const int32_t* x = itemset_ptr(a, k, gid);
    const double s_x = __ddiv_rn((double)__ldg(a.counts + gid), n_total);
    uint64_t total = 0;
    for (int t = 0; t < k; ++t) total += item_term(__ldg(x + t));
    const uint32_t full = (1u << k) - 1u, half = 1u << (k - 1);
    int kept = 0;
    int64_t pos = write ? rule_off[xi] : 0;
    for (uint32_t s0 = 1; s0 < half; s0 += 32) {
      const uint32_t sel = s0 + lane;
      bool k1 = false, k2 = false;
      double c1 = 0, c2 = 0, l1 = 0, l2 = 0;
      if (sel < half) {
        uint64_t sa = 0;
        for (int t = 0; t < k; ++t)
          if (sel >> t & 1u) sa += item_term(__ldg(x + t));
        const int32_t ca = subset_count(a, keys, vals, mask, x, k, sel, sa);
        const int32_t cc = subset_count(a, keys, vals, mask, x, k, full ^ sel, total - sa);
        if (ca < 0 || cc < 0) {
          atomicAdd(missing, 1);
        } else {
          const double s_a = __ddiv_rn((double)ca, n_total), s_c = __ddiv_rn((double)cc, n_total);
          c1 = __ddiv_rn(s_x, s_a);
          c2 = __ddiv_rn(s_x, s_c);
          k1 = c1 >= min_conf;
          k2 = c2 >= min_conf;
          l1 = __ddiv_rn(c1, s_c);
          l2 = __ddiv_rn(c2, s_a);
        }
      }
      const uint32_t m1 = __ballot_sync(0xffffffffu, k1), m2 = __ballot_sync(0xffffffffu, k2);
      if (write) {
        const uint32_t below = (1u << lane) - 1u;
        int64_t o = pos + __popc(m1 & below) + __popc(m2 & below);
        if (k1) {
          out_itemset[o] = (int32_t)gid; out_antecedent[o] = sel; out_support[o] = s_x;
          out_confidence[o] = c1; out_lift[o] = l1;
          ++o;
        }
        if (k2) {
          out_itemset[o] = (int32_t)gid; out_antecedent[o] = full ^ sel; out_support[o] = s_x;
          out_confidence[o] = c2; out_lift[o] = l2;
        }
        pos += __popc(m1) + __popc(m2);
      } else {
        kept += __popc(m1) + __popc(m2);
      }
    }
    if (!write && lane == 0) rule_cnt[xi] = kept;
  }
}

static int make_arena(Arena& a, const int32_t* d_items, const int32_t* d_counts,
                      const int64_t* level_sizes, int32_t max_k) {
  KMLS_REQUIRE(max_k >= 0 && max_k <= kMaxLevels - 8, "rules: more than 24 levels are not supported");
  a.items = d_items;
  a.counts = d_counts;
  a.max_k = max_k;
  a.item_off[0] = a.item_off[1] = 0;
  a.id_off[0] = a.id_off[1] = 0;
  for (int k = 1; k <= max_k; ++k) {
    KMLS_REQUIRE(level_sizes[k] >= 0, "rules: negative level size");
    a.item_off[k + 1] = a.item_off[k] + level_sizes[k] * k;
    a.id_off[k + 1] = a.id_off[k] + level_sizes[k];
  }
  for (int k = max_k + 1; k <= kMaxLevels; ++k) { a.item_off[k + 1] = a.item_off[k]; a.id_off[k + 1] = a.id_off[k]; }
  KMLS_REQUIRE(a.id_off[max_k + 1] <= 0x7fffffff, "rules: more than 2^31-1 itemsets");
  return 0;
}

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_itemset_map_build(const int32_t* d_items, const int32_t* d_counts,
                                      const int64_t* level_sizes, int32_t max_k, uint64_t* d_keys,
                                      int32_t* d_vals, int64_t capacity, kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, level_sizes, max_k);
  if (rc) return rc;
  const int64_t n = a.id_off[max_k + 1];
  KMLS_REQUIRE(capacity >= 2 * n && capacity > 0 && (capacity & (capacity - 1)) == 0,
               "kmls_itemset_map_build: capacity must be a power of two >= 2n");
  cudaStream_t s = (cudaStream_t)stream;
  KMLS_CUDA(cudaMemsetAsync(d_keys, 0, sizeof(uint64_t) * (size_t)capacity, s));
  if (n == 0) return 0;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_map_build<<<grid, 256, 0, s>>>(a, (unsigned long long*)d_keys, d_vals, (uint64_t)capacity - 1);
  KMLS_LAUNCHED("k_map_build");
  return 0;
}

static int launch_rules(const int32_t* d_items, const int32_t* d_counts, const int64_t* level_sizes,
                        int32_t max_k, const uint64_t* d_keys, const int32_t* d_vals,
                        int64_t capacity, int64_t n_total, double min_conf, int64_t x_begin,
                        int64_t x_count, int write, int32_t* d_rule_cnt, const int64_t* d_rule_off, int32_t* o_itemset,
                        uint32_t* o_ante, double* o_sup, double* o_conf, double* o_lift,
                        int32_t* d_missing, cudaStream_t s) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(n_total > 0, "rules: n_playlists_total must be positive");
  const int64_t n_x = a.id_off[max_k + 1] - a.id_off[2];
  KMLS_REQUIRE(x_begin >= 0 && x_count >= 0 && x_begin + x_count <= std::max<int64_t>(n_x, 0),
               "rules: itemset range out of bounds");
  const int64_t n = x_count;
  if (max_k < 2 || n <= 0) return 0;
  KMLS_REQUIRE(capacity > 0 && (capacity & (capacity - 1)) == 0 && d_keys && d_vals && d_missing,
               "rules: bad map");
  int grid = (int)std::min<int64_t>((n + 7) / 8, (int64_t)num_sms() * 32);
  k_rules<<<grid, 256, 0, s>>>(a, (const unsigned long long*)d_keys, d_vals, (uint64_t)capacity - 1,
                               (double)n_total, min_conf, x_begin, x_count, write, d_rule_cnt, d_rule_off, o_itemset,
                               o_ante, o_sup, o_conf, o_lift, d_missing);
  KMLS_LAUNCHED("k_rules");
  return 0;
}

extern "C" int kmls_rules_count(const int32_t* d_items, const int32_t* d_counts,
                                const int64_t* level_sizes, int32_t max_k, const uint64_t* d_keys,
                                const int32_t* d_vals, int64_t capacity, int64_t n_playlists_total,
                                double min_confidence, int64_t x_begin, int64_t x_count,
                                int32_t* d_rule_cnt, int32_t* d_missing, kmls_stream_t stream) {
  return launch_rules(d_items, d_counts, level_sizes, max_k, d_keys, d_vals, capacity,
                      n_playlists_total, min_confidence, x_begin, x_count, 0, d_rule_cnt, nullptr,
                      nullptr, nullptr, nullptr, nullptr, nullptr, d_missing, (cudaStream_t)stream);
}

extern "C" int kmls_rules_write(const int32_t* d_items, const int32_t* d_counts,
                                const int64_t* level_sizes, int32_t max_k, const uint64_t* d_keys,
                                const int32_t* d_vals, int64_t capacity, int64_t n_playlists_total,
                                double min_confidence, int64_t x_begin, int64_t x_count,
                                const int64_t* d_rule_off, int32_t* d_out_itemset,
                                uint32_t* d_out_antecedent, double* d_out_support,
                                double* d_out_confidence, double* d_out_lift, int32_t* d_missing,
                                kmls_stream_t stream) {
  return launch_rules(d_items, d_counts, level_sizes, max_k, d_keys, d_vals, capacity,
                      n_playlists_total, min_confidence, x_begin, x_count, 1, nullptr, d_rule_off,
                      d_out_itemset, d_out_antecedent, d_out_support, d_out_confidence, d_out_lift,
                      d_missing, (cudaStream_t)stream);
}
