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
// Subset counts come from an EXACT index, not from fingerprints: the level tables form a
// trie (an itemset's parent is the itemset without its last member), and the index maps the
// 64-bit key (parent global id, last member) to (global id, count) - one 16-byte slot, one
// 32-byte sector per probe, nothing to verify. A warp that owns a class (siblings sharing
// the prefix P) first resolves all 2^(k-1) subsets of P by dynamic programming over that
// index (subset = smaller subset + its largest member) into shared memory, then needs ONE
// probe per (sibling, antecedent): the consequent's count.
//
// All levels' tables live in one arena: items of level k start at item_off[k], its itemsets
// have global ids id_off[k] .. id_off[k+1]-1, counts are indexed by global id; `par` holds,
// for every itemset of size >= 2 (indexed by global id - id_off[2]), the index of its parent
// inside level k-1.
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

constexpr int kMaxLevels = 32;
constexpr int kMaxRuleK = 15;  // 2^(k-1) subset table per warp must fit in shared memory
constexpr int kRuleIlp = 2;    // independent index probes in flight per lane

struct Arena {
  const int32_t* items;
  const int32_t* counts;
  const int32_t* par;
  int64_t item_off[kMaxLevels + 2];
  int64_t id_off[kMaxLevels + 2];
  int max_k;
};

__device__ __forceinline__ const int32_t* itemset_ptr(const Arena& a, int k, int64_t gid) {
  return a.items + a.item_off[k] + (gid - a.id_off[k]) * k;
}

__device__ __forceinline__ uint64_t edge_key(int32_t parent_gid, int32_t item) {
  return ((uint64_t)(uint32_t)parent_gid << 32) | (uint32_t)item;
}
constexpr unsigned long long kEmptyEdge = ~0ull;
// slot of a key; `mask` = capacity - 1 (capacity a power of two). A full 64-bit mixer: the
// keys are highly structured (small item ids, dense parent ids) and linear probing needs a
// hash without clustering (a plain multiplicative hash made probe chains 7x longer).
__device__ __forceinline__ uint64_t edge_slot(uint64_t key, uint64_t mask) {
  return mix64(key) & mask;
}

// slot = {key lo, key hi, child global id, child count}
__global__ void __launch_bounds__(256) k_trie_build(Arena a, uint4* __restrict__ slots, uint64_t mask) {
  const int64_t first = a.id_off[2], n = a.id_off[a.max_k + 1];
  for (int64_t gid = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gid < n;
       gid += (int64_t)gridDim.x * blockDim.x) {
    int k = 2;
    while (gid >= a.id_off[k + 1]) ++k;
    const int32_t parent = (int32_t)(a.id_off[k - 1] + __ldg(a.par + (gid - first)));
    const int32_t last = __ldg(itemset_ptr(a, k, gid) + (k - 1));
    const uint64_t key = edge_key(parent, last);
    uint64_t h = edge_slot(key, mask);
    while (atomicCAS(reinterpret_cast<unsigned long long*>(slots + h), kEmptyEdge,
                     (unsigned long long)key) != kEmptyEdge)
      h = (h + 1) & mask;
    reinterpret_cast<int2*>(slots + h)[1] = make_int2((int32_t)gid, __ldg(a.counts + gid));
  }
}

// (child gid, child count) of the edge (parent_gid, item); gid = -1 if absent.
__device__ __forceinline__ int2 trie_child(const uint4* __restrict__ slots, uint64_t mask,
                                           int32_t parent_gid, int32_t item) {
  const uint64_t key = edge_key(parent_gid, item);
  uint64_t h = edge_slot(key, mask);
  while (true) {
    const uint4 s = __ldg(slots + h);
    const uint64_t sk = ((uint64_t)s.y << 32) | s.x;
    if (sk == key) return make_int2((int32_t)s.z, (int32_t)s.w);
    if (sk == kEmptyEdge) return make_int2(-1, -1);
    h = (h + 1) & mask;
  }
}

struct RuleVals {
  bool k1, k2;
  double c1, c2, l1, l2;
};
__device__ __forceinline__ RuleVals rule_values(double s_x, int32_t ca, int32_t cc, double n_total,
                                                double min_conf) {
  RuleVals r;
  const double s_a = __ddiv_rn((double)ca, n_total), s_c = __ddiv_rn((double)cc, n_total);
  r.c1 = __ddiv_rn(s_x, s_a);
  r.c2 = __ddiv_rn(s_x, s_c);
  r.k1 = r.c1 >= min_conf;
  r.k2 = r.c2 >= min_conf;
  r.l1 = __ddiv_rn(r.c1, s_c);
  r.l2 = __ddiv_rn(r.c2, s_a);
  return r;
}

// confidence >= min_conf for confidence = fl(fl(cx/N) / fl(ca/N)), decided from the float
// ratio cx/ca whenever that is clearly (1e-5 relative, far above both the float and the
// double rounding errors) on one side of the threshold; the exact doubles otherwise.
__device__ __forceinline__ bool conf_keep(int32_t cx, int32_t ca, double s_x, double n_total,
                                          double min_conf) {
  const double r = (double)__fdividef((float)cx, (float)ca);
  if (r > min_conf * 1.00001) return true;
  if (r < min_conf * 0.99999) return false;
  return __ddiv_rn(s_x, __ddiv_rn((double)ca, n_total)) >= min_conf;
}

// ---------------------------------------------------------------------------------------
// K5 pass 1, one level k >= 2 per launch. One warp per itemset; only class heads (and the
// first itemset of the range) work: they resolve the subsets of the class prefix P, then
// walk their siblings. For every (itemset x, antecedent mask a over P) the pair
// (count(A), count(C)) goes to `part` (so that pass 2 needs no lookups) and the number of
// kept rules of x to rule_cnt. A = members of P selected by a; C = the rest of P plus the
// last member.
// ---------------------------------------------------------------------------------------
constexpr int kRuleChunkMax = 64;   // consecutive itemsets handled by one warp (at most)

__global__ void __launch_bounds__(256, 4)
k_rules_count(Arena a, int k, int64_t lvl_begin, int64_t lvl_count, const uint4* __restrict__ slots,
              uint64_t mask, double n_total, double min_conf, int chunk, int2* __restrict__ part,
              int32_t* __restrict__ rule_cnt, int32_t* __restrict__ missing) {
  extern __shared__ int32_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int n_sub = 1 << (k - 1);          // subsets of P (including the empty one)
  int32_t* tab_gid = smem + (size_t)warp * 2 * n_sub;
  int32_t* tab_cnt = tab_gid + n_sub;
  const int64_t lvl_first = a.id_off[k];   // global id of the level's first itemset
  const int n_part = n_sub - 1;            // antecedent masks 1 .. n_sub-1
  const int64_t n_chunks = (lvl_count + chunk - 1) / chunk;
  // A warp walks `chunk` consecutive itemsets. Consecutive classes of the sorted table
  // usually differ only in the LAST member of their prefix, so the subsets of P that avoid
  // the changed members are still in the table: only masks >= 2^d are resolved again, d =
  // first prefix position that changed (d = 0 for the first class of the chunk).
  for (int64_t ch = (int64_t)blockIdx.x * nwarps + warp; ch < n_chunks;
       ch += (int64_t)gridDim.x * nwarps) {
    const int64_t j_end = min(lvl_count, (ch + 1) * chunk);
    const int32_t* head = nullptr;         // prefix owner whose subsets are in the table
    for (int64_t j = ch * chunk; j < j_end; ++j) {
      const int32_t* x = itemset_ptr(a, k, lvl_first + lvl_begin + j);
      int d = 0;
      if (head != nullptr) {
        // first prefix position that differs from the table owner's: one parallel compare
        const bool ne = lane < k - 1 && __ldg(x + lane) != __ldg(head + lane);
        const uint32_t nm = __ballot_sync(0xffffffffu, ne);
        d = nm ? __ffs(nm) - 1 : k - 1;
      }
      if (head == nullptr || d < k - 1) {
        for (int h = d; h < k - 1; ++h) {
          const int32_t ph = __ldg(x + h);
          for (int m = (1 << h) + lane; m < (2 << h); m += 32) {
            const int rest = m ^ (1 << h);
            int2 gc;
            if (rest == 0) gc = make_int2(ph, __ldg(a.counts + ph));   // level-1 id = row index
            else gc = trie_child(slots, mask, tab_gid[rest], ph);
            if (gc.x < 0) atomicAdd(missing, 1);
            tab_gid[m] = gc.x;
            tab_cnt[m] = gc.y;
          }
          __syncwarp();
        }
        head = x;
      }
      const int32_t last = __ldg(x + (k - 1));
      const int32_t c_x = __ldg(a.counts + lvl_first + lvl_begin + j);
      const double s_x = __ddiv_rn((double)c_x, n_total);
      int kept = 0;
      // kRuleIlp antecedent masks per lane and iteration: their first index probes are all
      // issued before any is consumed
      for (int a0 = 1; a0 < n_sub; a0 += 32 * kRuleIlp) {
        int am[kRuleIlp];
        int32_t cc[kRuleIlp];
        uint64_t key[kRuleIlp], h[kRuleIlp];
        uint4 slot[kRuleIlp];
        bool probe[kRuleIlp];
#pragma unroll
        for (int u = 0; u < kRuleIlp; ++u) {
          am[u] = a0 + u * 32 + lane;
          probe[u] = false;
          cc[u] = -1;
          if (am[u] < n_sub) {
            const int rest = (n_sub - 1) ^ am[u];
            if (rest == 0) {
              cc[u] = __ldg(a.counts + last);
            } else {
              key[u] = edge_key(tab_gid[rest], last);
              h[u] = edge_slot(key[u], mask);
              slot[u] = __ldg(slots + h[u]);
              probe[u] = true;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < kRuleIlp; ++u) {
          if (probe[u]) {
            while (true) {
              const uint64_t sk = ((uint64_t)slot[u].y << 32) | slot[u].x;
              if (sk == key[u]) { cc[u] = (int32_t)slot[u].w; break; }
              if (sk == kEmptyEdge) break;
              h[u] = (h[u] + 1) & mask;
              slot[u] = __ldg(slots + h[u]);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < kRuleIlp; ++u) {
          bool k1 = false, k2 = false;
          if (am[u] < n_sub) {
            const int32_t ca = tab_cnt[am[u]];
            if (cc[u] < 0 || ca < 0) {
              atomicAdd(missing, 1);
            } else {
              k1 = conf_keep(c_x, ca, s_x, n_total, min_conf);
              k2 = conf_keep(c_x, cc[u], s_x, n_total, min_conf);
            }
            part[j * n_part + (am[u] - 1)] = make_int2(ca, cc[u]);
          }
          kept += __popc(__ballot_sync(0xffffffffu, k1)) + __popc(__ballot_sync(0xffffffffu, k2));
        }
      }
      if (lane == 0) rule_cnt[j] = kept;
    }
    __syncwarp();
  }
}

// K5 pass 2: streaming. One warp per itemset; reads the (count(A), count(C)) pairs of pass 1,
// recomputes the doubles and writes the kept rules at rule_off[x] + rank (ballot + popc
// rank): ordered by antecedent mask, A -> C before C -> A.
__global__ void __launch_bounds__(256)
k_rules_write(Arena a, int k, int64_t lvl_begin, int64_t lvl_count, double n_total, double min_conf,
              const int2* __restrict__ part, const int64_t* __restrict__ rule_off,
              int32_t* __restrict__ out_itemset, uint32_t* __restrict__ out_antecedent,
              double* __restrict__ out_support, double* __restrict__ out_confidence,
              double* __restrict__ out_lift) {
  const int lane = threadIdx.x & 31;
  const int n_sub = 1 << (k - 1), n_part = n_sub - 1;
  const uint32_t full = (1u << k) - 1u;
  const int64_t lvl_first = a.id_off[k];
  for (int64_t j = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); j < lvl_count;
       j += (int64_t)gridDim.x * 8) {
    const int64_t gid = lvl_first + lvl_begin + j;
    const double s_x = __ddiv_rn((double)__ldg(a.counts + gid), n_total);
    int64_t pos = rule_off[j];
    for (int a0 = 1; a0 < n_sub; a0 += 32) {
      const int am = a0 + lane;
      RuleVals r;
      r.k1 = r.k2 = false;
      if (am < n_sub) {
        const int2 cc = __ldg(part + j * n_part + (am - 1));
        if (cc.x >= 0 && cc.y >= 0) r = rule_values(s_x, cc.x, cc.y, n_total, min_conf);
      }
      const uint32_t m1 = __ballot_sync(0xffffffffu, r.k1), m2 = __ballot_sync(0xffffffffu, r.k2);
      const uint32_t below = (1u << lane) - 1u;
      int64_t o = pos + __popc(m1 & below) + __popc(m2 & below);
      if (r.k1) {
        out_itemset[o] = (int32_t)gid; out_antecedent[o] = (uint32_t)am; out_support[o] = s_x;
        out_confidence[o] = r.c1; out_lift[o] = r.l1;
        ++o;
      }
      if (r.k2) {
        out_itemset[o] = (int32_t)gid; out_antecedent[o] = full ^ (uint32_t)am; out_support[o] = s_x;
        out_confidence[o] = r.c2; out_lift[o] = r.l2;
      }
      pos += __popc(m1) + __popc(m2);
    }
  }
}

static int make_arena(Arena& a, const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par,
                      const int64_t* level_sizes, int32_t max_k) {
  KMLS_REQUIRE(max_k >= 0 && max_k <= kMaxLevels - 2, "rules: too many levels");
  a.items = d_items;
  a.counts = d_counts;
  a.par = d_par;
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

extern "C" int kmls_trie_index_build(const int32_t* d_items, const int32_t* d_counts,
                                     const int32_t* d_par, const int64_t* level_sizes,
                                     int32_t max_k, void* d_slots, int64_t capacity,
                                     kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, d_par, level_sizes, max_k);
  if (rc) return rc;
  const int64_t n = a.id_off[max_k + 1] - a.id_off[2];
  KMLS_REQUIRE(capacity >= 2 * n && capacity > 0 && (capacity & (capacity - 1)) == 0,
               "kmls_trie_index_build: capacity must be a power of two >= 2n");
  KMLS_REQUIRE(d_slots && ((uintptr_t)d_slots & 15) == 0, "kmls_trie_index_build: d_slots must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  KMLS_CUDA(cudaMemsetAsync(d_slots, 0xFF, sizeof(uint4) * (size_t)capacity, s));
  if (max_k < 2 || n <= 0) return 0;
  KMLS_REQUIRE(d_par, "kmls_trie_index_build: d_par is null");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_trie_build<<<grid, 256, 0, s>>>(a, (uint4*)d_slots, (uint64_t)capacity - 1);
  KMLS_LAUNCHED("k_trie_build");
  return 0;
}

extern "C" int64_t kmls_rules_parts_per_itemset(int32_t k) {
  return k >= 2 && k <= kMaxRuleK ? ((int64_t)1 << (k - 1)) - 1 : -1;
}

extern "C" int kmls_rules_count(const int32_t* d_items, const int32_t* d_counts,
                                const int64_t* level_sizes, int32_t max_k, int32_t k,
                                int64_t lvl_begin, int64_t lvl_count, const void* d_slots,
                                int64_t capacity, int64_t n_playlists_total, double min_confidence,
                                void* d_part, int32_t* d_rule_cnt, int32_t* d_missing,
                                kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, nullptr, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(k >= 2 && k <= max_k, "kmls_rules_count: level out of range");
  KMLS_REQUIRE(k <= kMaxRuleK, "kmls_rules_count: itemsets of more than 15 members are not supported");
  KMLS_REQUIRE(n_playlists_total > 0, "kmls_rules_count: n_playlists_total must be positive");
  KMLS_REQUIRE(lvl_begin >= 0 && lvl_count >= 0 && lvl_begin + lvl_count <= level_sizes[k],
               "kmls_rules_count: itemset range out of bounds");
  if (lvl_count == 0) return 0;
  KMLS_REQUIRE(d_slots && capacity > 0 && (capacity & (capacity - 1)) == 0 && d_part && d_rule_cnt && d_missing,
               "kmls_rules_count: bad index or null pointer");
  const size_t per_warp = sizeof(int32_t) * 2 * ((size_t)1 << (k - 1));
  int warps = (int)std::min<size_t>(8, (200 * 1024) / per_warp);
  KMLS_REQUIRE(warps >= 1, "kmls_rules_count: subset table does not fit in shared memory");
  const size_t smem = per_warp * warps;
  KMLS_CUDA(cudaFuncSetAttribute(k_rules_count, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  // chunks long enough to reuse the subset table, short enough to keep every SM busy
  const int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(kRuleChunkMax, lvl_count / ((int64_t)num_sms() * 64)));
  const int64_t n_chunks = (lvl_count + chunk - 1) / chunk;
  int grid = (int)std::min<int64_t>((n_chunks + warps - 1) / warps, (int64_t)num_sms() * 32);
  k_rules_count<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(
      a, k, lvl_begin, lvl_count, (const uint4*)d_slots, (uint64_t)capacity - 1,
      (double)n_playlists_total, min_confidence, chunk, (int2*)d_part, d_rule_cnt, d_missing);
  KMLS_LAUNCHED("k_rules_count");
  return 0;
}

extern "C" int kmls_rules_write(const int32_t* d_counts, const int64_t* level_sizes, int32_t max_k,
                                int32_t k, int64_t lvl_begin, int64_t lvl_count,
                                int64_t n_playlists_total, double min_confidence,
                                const void* d_part, const int64_t* d_rule_off,
                                int32_t* d_out_itemset, uint32_t* d_out_antecedent,
                                double* d_out_support, double* d_out_confidence, double* d_out_lift,
                                kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, nullptr, d_counts, nullptr, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(k >= 2 && k <= max_k && k <= kMaxRuleK, "kmls_rules_write: level out of range");
  KMLS_REQUIRE(lvl_begin >= 0 && lvl_count >= 0 && lvl_begin + lvl_count <= level_sizes[k],
               "kmls_rules_write: itemset range out of bounds");
  if (lvl_count == 0) return 0;
  KMLS_REQUIRE(d_part && d_rule_off && d_out_itemset && d_out_antecedent && d_out_support &&
                   d_out_confidence && d_out_lift, "kmls_rules_write: null pointer");
  int grid = (int)std::min<int64_t>((lvl_count + 7) / 8, (int64_t)num_sms() * 32);
  k_rules_write<<<grid, 256, 0, (cudaStream_t)stream>>>(
      a, k, lvl_begin, lvl_count, (double)n_playlists_total, min_confidence, (const int2*)d_part,
      d_rule_off, d_out_itemset, d_out_antecedent, d_out_support, d_out_confidence, d_out_lift);
  KMLS_LAUNCHED("k_rules_write");
  return 0;
}
