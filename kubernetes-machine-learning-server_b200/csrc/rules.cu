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
// trie (an itemset's parent is the itemset without its last member) and, the tables being
// sorted, the children of a node are CONTIGUOUS in the next level. The index is one 16-byte
// node per itemset, {first child's global id, number of children, last member, count};
// "child of p with last member x" = read node[p], then search x among its few children.
// Unlike a hash table this keeps everything a class needs close together: siblings (P, x_1),
// (P, x_2), ... walk the SAME child lists of the subsets of P, so their reads hit L1/L2.
// A warp that owns a class first resolves all 2^(k-1) subsets of P by dynamic programming
// over that index (subset = smaller subset + its largest member) into shared memory, then
// needs ONE lookup per (sibling, antecedent): the consequent's count.
//
// All levels' tables live in one arena: items of level k start at item_off[k], its itemsets
// have global ids id_off[k] .. id_off[k+1]-1, counts are indexed by global id; `par` holds,
// for every itemset of size >= 2 (indexed by global id - id_off[2]), the index of its parent
// inside level k-1.
#include <cstdlib>

#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

constexpr int kMaxLevels = 32;
constexpr int kMaxRuleK = 15;  // 2^(k-1) subset table per warp must fit in shared memory
constexpr int kRuleIlp = 2;    // independent index probes in flight per lane
constexpr int kGatherIlp = 2;  // blocks of 32 slots in flight per warp in the gather route of pass 1
constexpr int kWriteIlp = 2;   // groups of 32 splits whose pass-1 pairs are loaded ahead in pass 2

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

// node = {first child gid, number of children, last member, count}. The array is zeroed first
// (a node without children keeps {0, 0}); every itemset then writes its own (last member,
// count) and, if it is the first of its parent's children (the parent column is sorted, so a
// parent's children are one run), the parent's (first child, number of children) - the run's
// end found by galloping, not by a search per PARENT over the whole column (23 dependent probes
// for each of 34 M itemsets at Z1M: 1.07 ms).
__global__ void __launch_bounds__(256) k_trie_build(Arena a, int4* __restrict__ node) {
  const int64_t n = a.id_off[a.max_k + 1];
  for (int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gid < n;
       gid += (int64_t)gridDim.x * blockDim.x) {
    int k = 1;
    while (gid >= a.id_off[k + 1]) ++k;
    const int32_t last = __ldg(itemset_ptr(a, k, gid) + (k - 1));
    reinterpret_cast<int2*>(node + gid)[1] = make_int2(last, __ldg(a.counts + gid));
    if (k < 2) continue;
    const int64_t idx = gid - a.id_off[k], m = a.id_off[k + 1] - a.id_off[k];
    const int32_t* par = a.par + (a.id_off[k] - a.id_off[2]);
    const int32_t me = __ldg(par + idx);
    if (idx > 0 && __ldg(par + idx - 1) == me) continue;      // not the first child
    int64_t lo = idx, step = 1;                               // par[lo] == me
    while (lo + step < m && __ldg(par + lo + step) == me) {
      lo += step;
      step <<= 1;
    }
    int64_t hi = min(lo + step, m);                           // par[hi] != me or hi == m
    while (hi - lo > 1) {
      const int64_t mid = (lo + hi) >> 1;
      if (__ldg(par + mid) == me) lo = mid; else hi = mid;
    }
    reinterpret_cast<int2*>(node + a.id_off[k - 1] + me)[0] = make_int2((int32_t)gid, (int32_t)(hi - idx));
  }
}

// The child of a node whose last member is `item`, searched among the node's children
// node[first .. first + n) (sorted by last member); n < 0 = the node itself is missing. Returns
// the child's node, .y = -1 if absent. Indices are below 2^31, so unsigned 32-bit offsets
// address the index with one instruction.
__device__ __forceinline__ void trie_bisect(const int4* __restrict__ node, int32_t first, int32_t n,
                                            int32_t item, int& lo, int& hi) {
  lo = 0;
  hi = n;
  while (hi - lo > 4) {                      // long child lists: bisect down to 4 candidates
    const int mid = (lo + hi) >> 1;
    if (__ldg(&node[(uint32_t)(first + mid)].z) <= item) lo = mid; else hi = mid;
  }
}

__device__ __forceinline__ int4 trie_find(const int4* __restrict__ node, int32_t first, int32_t n,
                                          int32_t item, int32_t* gid = nullptr) {
  int lo, hi;
  trie_bisect(node, first, n, item, lo, hi);
  int4 r = make_int4(-1, -1, -1, -1);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (lo + i < hi) {
      const int4 c = __ldg(node + (uint32_t)(first + lo + i));
      if (c.z == item) {
        r = c;
        if (gid) *gid = first + lo + i;
      }
    }
  }
  return r;
}

struct RuleVals {
  bool k1, k2;
  double c1, c2, l1, l2;
};

// RN(a / b) from y = RN(1 / b) without a division: q = RN(a y), then twice
//   r = a - b q (exact, one FMA), q = RN(q + r y).
// Markstein's theorem (Markstein 1990; Muller et al., Handbook of Floating-Point Arithmetic,
// "division with an FMA"): if y approximates 1/b with relative error < 2^-53 and q is a
// faithful rounding of a/b, one such correction returns the correctly rounded quotient.
// RN(1/b) has relative error <= u/(1+u) < 2^-53; RN(a y) is within 1.5 ulp of a/b, not always
// faithful, but the first correction rounds a value within 2^-52 ulp of a/b and therefore IS
// faithful, so the second one meets the theorem's conditions for every normal a, b (all
// quantities here are between 2^-31 and 2^31). __ddiv_rn computes its own reciprocal for
// every quotient (about 30 instructions); pass 2 divides six times per split by only three
// different divisors (N, cA/N, cC/N), which made it issue-bound, not HBM-bound.
__device__ __forceinline__ double div_by_rcp(double a, double b, double y) {
  double q = __dmul_rn(a, y);
  double r = __fma_rn(-b, q, a);
  q = __fma_rn(r, y, q);
  r = __fma_rn(-b, q, a);
  return __fma_rn(r, y, q);
}

__device__ __forceinline__ RuleVals rule_values(double s_x, int32_t ca, int32_t cc, double n_total,
                                                double rcp_n, double min_conf) {
  RuleVals r;
  const double s_a = div_by_rcp((double)ca, n_total, rcp_n), s_c = div_by_rcp((double)cc, n_total, rcp_n);
  const double y_a = __drcp_rn(s_a), y_c = __drcp_rn(s_c);
  r.c1 = div_by_rcp(s_x, s_a, y_a);
  r.c2 = div_by_rcp(s_x, s_c, y_c);
  r.k1 = r.c1 >= min_conf;
  r.k2 = r.c2 >= min_conf;
  r.l1 = div_by_rcp(r.c1, s_c, y_c);
  r.l2 = div_by_rcp(r.c2, s_a, y_a);
  return r;
}

// confidence >= min_conf for confidence = fl(fl(cx/N) / fl(ca/N)), decided in fp32 whenever
// cx is clearly (1e-5 relative, far above the four float roundings involved) on one side of
// ca * min_conf; the exact doubles otherwise. thr_hi / thr_lo = (float)(min_conf * (1 +- 1e-5)).
__device__ __forceinline__ bool conf_keep(float fx, int32_t cx, int32_t ca, float thr_hi, float thr_lo,
                                          double n_total, double min_conf) {
  const float fa = (float)ca;
  if (fx > fa * thr_hi) return true;
  if (fx < fa * thr_lo) return false;
  asm volatile("" : "+r"(cx));   // keeps cx / N of this rare path from being hoisted out of the split loops
  return __ddiv_rn(__ddiv_rn((double)cx, n_total), __ddiv_rn((double)ca, n_total)) >= min_conf;
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
k_rules_count(Arena a, int k, int64_t lvl_begin, int64_t lvl_count, const int4* __restrict__ node,
              double n_total, double min_conf, int chunk, int32_t* __restrict__ part,
              int32_t* __restrict__ rule_cnt, int32_t* __restrict__ missing) {
  extern __shared__ int32_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int n_sub = 1 << (k - 1);          // subsets of P (including the empty one)
  // per warp, for every subset S of the class prefix P: its children in the index (first
  // child's global id, how many; -1 = S is not in the index) and count(S)
  int32_t* tab_first = smem + (size_t)warp * 3 * n_sub;
  int32_t* tab_n = tab_first + n_sub;
  int32_t* tab_cnt = tab_n + n_sub;
  const int64_t lvl_first = a.id_off[k];   // global id of the level's first itemset
  int32_t* pa = part;                      // planes of count(A) / count(C), see "Layout" below
  int32_t* pc = part + lvl_count * n_sub;
  const int64_t n_chunks = (lvl_count + chunk - 1) / chunk;
  const float thr_hi = (float)(min_conf * 1.00001), thr_lo = (float)(min_conf * 0.99999);
  const int32_t* items_lvl = a.items + a.item_off[k] + lvl_begin * k;
  const int32_t* counts_lvl = a.counts + lvl_first + lvl_begin;
  // Subsets of P by dynamic programming over the index: subset m with highest member h =
  // child `member h` of subset m without h. Only masks >= 2^d are resolved, d = first prefix
  // position that differs from the prefix the table was built for.
  auto resolve = [&](int d, int32_t my_item) {
    for (int h = d; h < k - 1; ++h) {
      const int32_t ph = __shfl_sync(0xffffffffu, my_item, h);
      for (int m = (1 << h) + lane; m < (2 << h); m += 32) {
        const int rest = m ^ (1 << h);
        int4 c;
        if (rest == 0) c = __ldg(node + (uint32_t)ph);          // level-1 id = row index
        else c = trie_find(node, tab_first[rest], tab_n[rest], ph);
        if (c.y < 0) atomicAdd(missing, 1);
        tab_first[m] = c.x;
        tab_n[m] = c.y;
        tab_cnt[m] = c.w;
      }
      __syncwarp();
    }
  };
  // A warp walks `chunk` consecutive itemsets. Consecutive classes of the sorted table
  // usually differ only in the LAST member of their prefix, so the subsets of P that avoid
  // the changed members are still in the table (d = 0 for the first class of the chunk).
  for (int64_t ch = (int64_t)blockIdx.x * nwarps + warp; ch < n_chunks;
       ch += (int64_t)gridDim.x * nwarps) {
    const int64_t j_end = min(lvl_count, (ch + 1) * chunk);
    const int64_t j0 = ch * chunk;
    if (n_sub <= 16) {
      // Short itemsets (k <= 5): lane = (sibling in the group, antecedent mask). The group is a
      // run of up to 32 / n_sub siblings of ONE class, so that all of them use the same
      // subset table; a class boundary simply ends the run early.
      const int per_warp = 32 / n_sub, g = lane / n_sub, am = lane % n_sub;
      const uint32_t gmask = ((1u << n_sub) - 1u) << (g * n_sub);
      int32_t head_item = -1;
      bool have_head = false;
      int64_t j = j0;
      while (j < j_end) {
        const int32_t my_item = lane < k ? __ldg(items_lvl + j * k + lane) : 0;
        int d = 0;
        if (have_head) {
          const uint32_t nm = __ballot_sync(0xffffffffu, lane < k - 1 && my_item != head_item);
          d = nm ? __ffs(nm) - 1 : k - 1;
        }
        if (!have_head || d < k - 1) {
          resolve(d, my_item);
          head_item = my_item;
          have_head = true;
        }
        // which of the next per_warp itemsets are siblings of j (same first k-1 members)?
        const int64_t sib = j + g;
        bool ok = sib < j_end;
        if (ok && am < k - 1) ok = __ldg(items_lvl + sib * k + am) == __ldg(items_lvl + j * k + am);
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        int run = 0;
        while (run < per_warp && ((okm >> (run * n_sub)) & ((1u << n_sub) - 1u)) == ((1u << n_sub) - 1u)) ++run;
        // run >= 1: group 0 is itemset j itself
        const bool active = g < run && am >= 1;
        bool k1 = false, k2 = false;
        if (active) {
          const int32_t last = __ldg(items_lvl + sib * k + (k - 1));
          const int32_t c_x = __ldg(counts_lvl + sib);
          const int32_t ca = tab_cnt[am];
          const int rest = (n_sub - 1) ^ am;
          const int32_t cc = rest == 0 ? __ldg(a.counts + last)
                                       : trie_find(node, tab_first[rest], tab_n[rest], last).w;
          if (cc < 0 || ca < 0) {
            atomicAdd(missing, 1);
          } else {
            const float fx = (float)c_x;
            k1 = conf_keep(fx, c_x, ca, thr_hi, thr_lo, n_total, min_conf);
            k2 = conf_keep(fx, c_x, cc, thr_hi, thr_lo, n_total, min_conf);
          }
          const int sl = (int)(__brev((unsigned)am) >> (33 - k));
          pa[sib * n_sub + sl] = ca;
          pc[sib * n_sub + sl] = cc;
          if (am == 1) pa[sib * n_sub] = pc[sib * n_sub] = c_x;      // slot 0 = count(X)
        }
        const uint32_t b1 = __ballot_sync(0xffffffffu, k1), b2 = __ballot_sync(0xffffffffu, k2);
        if (active && am == 1) rule_cnt[sib] = __popc(b1 & gmask) + __popc(b2 & gmask);
        j += run;
      }
      __syncwarp();
      continue;
    }
    // Lane t holds member t of the current itemset; the next itemset's members and count are
    // fetched one iteration ahead so that only the index probes expose memory latency.
    int32_t nx_item = lane < k ? __ldg(items_lvl + j0 * k + lane) : 0;
    int32_t nx_cnt = __ldg(counts_lvl + j0);
    int32_t head_item = -1;                // member `lane` of the prefix whose subsets are in the table
    bool have_head = false;
    for (int64_t j = j0; j < j_end; ++j) {
      const int32_t my_item = nx_item;
      const int32_t c_x = nx_cnt;
      if (j + 1 < j_end) {
        nx_item = lane < k ? __ldg(items_lvl + (j + 1) * k + lane) : 0;
        nx_cnt = __ldg(counts_lvl + j + 1);
      }
      int d = 0;
      if (have_head) {
        // first prefix position that differs from the table owner's: one parallel compare
        const uint32_t nm = __ballot_sync(0xffffffffu, lane < k - 1 && my_item != head_item);
        d = nm ? __ffs(nm) - 1 : k - 1;
      }
      if (!have_head || d < k - 1) {
        resolve(d, my_item);
        head_item = my_item;
        have_head = true;
      }
      const int32_t last = __shfl_sync(0xffffffffu, my_item, k - 1);
      const float fx = (float)c_x;
      int32_t* pa_j = pa + j * n_sub;      // indexed by slot = bit-reversed antecedent mask
      int32_t* pc_j = pc + j * n_sub;
      if (lane == 0) pa_j[0] = pc_j[0] = c_x;
      int kept = 0;
      // kRuleIlp antecedent masks per lane and iteration: the candidate children of all of
      // them are loaded before any is compared
      for (int a0 = 1; a0 < n_sub; a0 += 32 * kRuleIlp) {
        int32_t cc[kRuleIlp], first[kRuleIlp];
        int lo[kRuleIlp], hi[kRuleIlp];
#pragma unroll
        for (int u = 0; u < kRuleIlp; ++u) {
          const int am = a0 + u * 32 + lane;
          cc[u] = -1;
          lo[u] = hi[u] = 0;
          first[u] = 0;
          if (am < n_sub) {
            const int rest = (n_sub - 1) ^ am;
            if (rest == 0) {
              cc[u] = __ldg(a.counts + last);
            } else {
              first[u] = tab_first[rest];
              trie_bisect(node, first[u], tab_n[rest], last, lo[u], hi[u]);
            }
          }
        }
        int2 cand[kRuleIlp][4];              // (last member, count) of up to 4 candidate children
#pragma unroll
        for (int u = 0; u < kRuleIlp; ++u)
#pragma unroll
          for (int i = 0; i < 4; ++i)
            cand[u][i] = lo[u] + i < hi[u]
                             ? __ldg(reinterpret_cast<const int2*>(node + (uint32_t)(first[u] + lo[u] + i)) + 1)
                             : make_int2(-1, -1);
#pragma unroll
        for (int u = 0; u < kRuleIlp; ++u) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (cand[u][i].x == last) cc[u] = cand[u][i].y;
          const int am = a0 + u * 32 + lane;
          bool k1 = false, k2 = false;
          if (am < n_sub) {
            const int32_t ca = tab_cnt[am];
            if (cc[u] < 0 || ca < 0) {
              atomicAdd(missing, 1);
            } else {
              k1 = conf_keep(fx, c_x, ca, thr_hi, thr_lo, n_total, min_conf);
              k2 = conf_keep(fx, c_x, cc[u], thr_hi, thr_lo, n_total, min_conf);
            }
            const int sl = (int)(__brev((unsigned)am) >> (33 - k));
            pa_j[sl] = ca;
            pc_j[sl] = cc[u];
          }
          kept += __popc(__ballot_sync(0xffffffffu, k1)) + __popc(__ballot_sync(0xffffffffu, k2));
        }
      }
      if (lane == 0) rule_cnt[j] = kept;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------
// Layout of pass 1's output ("part") for level k, n rows (itemsets), S = 2^(k-1) slots per row:
// two int32 planes, planeA = part, planeC = part + n * S. Slot s >= 1 of row X stands for the
// antecedent mask am = bitreverse_{k-1}(s) over X's prefix P: planeA = count(A), planeC =
// count(C), A = members of P selected by am, C = X \ A. Slot 0 of both planes = count(X).
// Bit-REVERSED slots make everything the next level needs contiguous (see k_rules_gather).
//
// K5 pass 1 without searching per split ("gather" route, used when one GPU holds the whole
// previous level's output). Every subset count a k-itemset X = P + {x} needs was already
// produced by level k-1 (row stride S' = S / 2 there):
//   count(A), A a subset of P: P is X's parent, a (k-1)-itemset with prefix P' and last member
//     q. A without q is an antecedent of P (planeA of P's row, slot s >> 1), A with q is a
//     consequent of P (planeC, slot S' - 1 - (s >> 1); slot 0 when A = P). s even / odd decides.
//   count(C), C = R + {x}, R a proper subset of P: with p_i the first member of P not in R, C
//     is a consequent of X_i = X \ {p_i}, which has the same last member x. In slot terms:
//     h = highest set bit of s, i = k - 2 - h, and the slot inside X_i's row is s - 2^h
//     (slot 0, count(X_i), when A = {p_i}).
// So a block of 32 consecutive slots reads two half-blocks of the parent's row and one block
// of ONE X_i row, all contiguous, and writes two contiguous blocks. Level 2 reads level 1's
// counts as its planeC (S' = 1).
// What remains to be searched are the k-1 rows X_i per itemset (k_rules_subsets, one thread
// per (itemset, i)); X_i = child x of P \ {p_i}, and P \ {p_i} is what the parent's own
// k_rules_subsets output names (or the parent's parent for i = k-2).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_rules_subsets(Arena a, int k, const int4* __restrict__ node, const int32_t* __restrict__ sub_prev,
                int32_t* __restrict__ sub, int32_t* __restrict__ missing) {
  const int64_t n_k = a.id_off[k + 1] - a.id_off[k];
  const int i = blockIdx.y;                  // which member is dropped
  const int32_t* items_k = a.items + a.item_off[k];
  const int32_t* par_k = a.par + (a.id_off[k] - a.id_off[2]);
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_k;
       j += (int64_t)gridDim.x * blockDim.x) {
    const int32_t x = __ldg(items_k + j * k + (k - 1));
    int32_t out;
    if (k == 2) {
      out = x;                               // {p0, x} \ {p0} = {x}: level-1 index = row
    } else {
      const int32_t row_p = __ldg(par_k + j);
      // P \ {p_i} as an index into level k-2
      int32_t pi;
      if (i == k - 2) pi = __ldg(a.par + (a.id_off[k - 1] - a.id_off[2]) + row_p);
      else pi = __ldg(sub_prev + (int64_t)row_p * (k - 2) + i);
      out = -1;
      if (pi >= 0) {
        const int4 nd = __ldg(node + (uint32_t)(a.id_off[k - 2] + pi));
        int32_t gid = -1;
        trie_find(node, nd.x, nd.y, x, &gid);
        if (gid >= 0) out = gid - (int32_t)a.id_off[k - 1];
      }
      if (out < 0) atomicAdd(missing, 1);
    }
    sub[j * (k - 1) + i] = out;
  }
}

// The same, one thread per ITEMSET (k >= 3): the last member and the parent are loaded once for
// all k-1 dropped members, and two dropped members are looked up at a time - their loads
// (P \ {p_i}, its node, the candidate children) are independent and issued together.
__global__ void __launch_bounds__(256)
k_rules_subsets_row(Arena a, int k, const int4* __restrict__ node, const int32_t* __restrict__ sub_prev,
                    int32_t* __restrict__ sub, int32_t* __restrict__ missing) {
  const int64_t n_k = a.id_off[k + 1] - a.id_off[k];
  const int32_t* items_k = a.items + a.item_off[k];
  const int32_t* par_k = a.par + (a.id_off[k] - a.id_off[2]);
  const int32_t* par_km1 = a.par + (a.id_off[k - 1] - a.id_off[2]);
  const uint32_t base2 = (uint32_t)a.id_off[k - 2];
  const int32_t base1 = (int32_t)a.id_off[k - 1];
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_k;
       j += (int64_t)gridDim.x * blockDim.x) {
    const int32_t x = __ldg(items_k + j * k + (k - 1));
    const int32_t row_p = __ldg(par_k + j);
    const int32_t* sp = sub_prev + (int64_t)row_p * (k - 2);
    int32_t* out = sub + j * (k - 1);
    int bad = 0;
    for (int i0 = 0; i0 < k - 1; i0 += 2) {
      int32_t pi[2];
      int4 nd[2];
      int lo[2], hi[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i = i0 + u;
        pi[u] = -1;
        if (i < k - 1) pi[u] = i == k - 2 ? __ldg(par_km1 + row_p) : __ldg(sp + i);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        nd[u] = make_int4(0, 0, 0, 0);
        if (pi[u] >= 0) nd[u] = __ldg(node + (base2 + (uint32_t)pi[u]));
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) trie_bisect(node, nd[u].x, nd[u].y, x, lo[u], hi[u]);
      int32_t key[2][4];
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int c = 0; c < 4; ++c)
          key[u][c] = lo[u] + c < hi[u] ? __ldg(&node[(uint32_t)(nd[u].x + lo[u] + c)].z) : -1;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int i = i0 + u;
        if (i >= k - 1) continue;
        int32_t res = -1;
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (key[u][c] == x) res = nd[u].x + lo[u] + c - base1;
        if (res < 0) ++bad;
        out[i] = res;
      }
    }
    if (bad) atomicAdd(missing, bad);
  }
}

// antecedent mask of slot s (and back): bit reversal over the k-1 prefix positions
__device__ __forceinline__ int slot_mask(int s, int k) { return (int)(__brev((unsigned)s) >> (33 - k)); }

__global__ void __launch_bounds__(256)
k_rules_gather(Arena a, int k, const int32_t* __restrict__ sub, const int32_t* __restrict__ part_prev,
               double n_total, double min_conf, int32_t* __restrict__ part,
               int32_t* __restrict__ rule_cnt, int32_t* __restrict__ missing) {
  const int lane = threadIdx.x & 31;
  const int n_sub = 1 << (k - 1), top = n_sub >> 1;
  const int64_t n_k = a.id_off[k + 1] - a.id_off[k], n_p = a.id_off[k] - a.id_off[k - 1];
  const float thr_hi = (float)(min_conf * 1.00001), thr_lo = (float)(min_conf * 0.99999);
  const int32_t* counts_k = a.counts + a.id_off[k];
  const int32_t* par_k = a.par + (a.id_off[k] - a.id_off[2]);
  const int32_t* pa_prev = part_prev;
  const int32_t* pc_prev = k == 2 ? a.counts : part_prev + n_p * top;
  int32_t* pa = part;
  int32_t* pc = part + n_k * n_sub;
  if (n_sub <= 16) {
    // short itemsets: 32 / n_sub itemsets per warp, lane = (itemset in the group, slot)
    const int per_warp = 32 / n_sub, g = lane / n_sub, s = lane % n_sub;
    const uint32_t gmask = ((1u << n_sub) - 1u) << (g * n_sub);
    const int h = 31 - __clz(s | 1), i = k - 2 - h;
    const int64_t n_groups = (n_k + per_warp - 1) / per_warp;
    for (int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); w < n_groups;
         w += (int64_t)gridDim.x * 8) {
      const int64_t j = w * per_warp + g;
      const bool active = j < n_k;
      bool k1 = false, k2 = false;
      if (active) {
        const int32_t c_x = __ldg(counts_k + j);
        int32_t ca = c_x, cc = c_x;
        if (s >= 1) {
          const int64_t row_p = __ldg(par_k + j);
          const int64_t row_i = __ldg(sub + j * (k - 1) + i);
          ca = (s & 1) ? __ldg(pc_prev + row_p * top + (top - 1 - (s >> 1))) : __ldg(pa_prev + row_p * top + (s >> 1));
          cc = row_i < 0 ? -1 : __ldg(pc_prev + row_i * top + (s - (1 << h)));
          if (ca < 0 || cc < 0) {
            atomicAdd(missing, 1);
          } else {
            const float fx = (float)c_x;
            k1 = conf_keep(fx, c_x, ca, thr_hi, thr_lo, n_total, min_conf);
            k2 = conf_keep(fx, c_x, cc, thr_hi, thr_lo, n_total, min_conf);
          }
        }
        pa[j * n_sub + s] = ca;
        pc[j * n_sub + s] = cc;
      }
      const uint32_t b1 = __ballot_sync(0xffffffffu, k1), b2 = __ballot_sync(0xffffffffu, k2);
      if (active && s == 0) rule_cnt[j] = __popc(b1 & gmask) + __popc(b2 & gmask);
    }
    return;
  }
  // long itemsets: one warp per itemset, lane = slot (blocks of 32)
  for (int64_t j = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); j < n_k; j += (int64_t)gridDim.x * 8) {
    const int32_t c_x = __ldg(counts_k + j);
    const int64_t row_p = __ldg(par_k + j);
    const int32_t my_sub = lane < k - 1 ? __ldg(sub + j * (k - 1) + lane) : -1;
    const float fx = (float)c_x;
    const int32_t* pa_p = pa_prev + row_p * top;
    const int32_t* pc_p = pc_prev + row_p * top;
    int32_t* pa_j = pa + j * n_sub;
    int32_t* pc_j = pc + j * n_sub;
    int kept = 0;
    for (int s0 = 0; s0 < n_sub; s0 += 32 * kGatherIlp) {
      int32_t ca[kGatherIlp], cc[kGatherIlp];
#pragma unroll
      for (int u = 0; u < kGatherIlp; ++u) {
        const int s = s0 + u * 32 + lane;
        const int h = 31 - __clz(s | 1);
        const int64_t row_i = __shfl_sync(0xffffffffu, my_sub, (k - 2 - h) & 31);
        ca[u] = cc[u] = c_x;                               // slot 0
        if (s >= 1 && s < n_sub) {
          ca[u] = (s & 1) ? __ldg(pc_p + (top - 1 - (s >> 1))) : __ldg(pa_p + (s >> 1));
          cc[u] = row_i < 0 ? -1 : __ldg(pc_prev + row_i * top + (s - (1 << h)));
        }
      }
#pragma unroll
      for (int u = 0; u < kGatherIlp; ++u) {
        const int s = s0 + u * 32 + lane;
        bool k1 = false, k2 = false;
        if (s < n_sub) {
          if (s >= 1) {
            if (ca[u] < 0 || cc[u] < 0) {
              atomicAdd(missing, 1);
            } else {
              k1 = conf_keep(fx, c_x, ca[u], thr_hi, thr_lo, n_total, min_conf);
              k2 = conf_keep(fx, c_x, cc[u], thr_hi, thr_lo, n_total, min_conf);
            }
          }
          pa_j[s] = ca[u];
          pc_j[s] = cc[u];
        }
        kept += __popc(__ballot_sync(0xffffffffu, k1)) + __popc(__ballot_sync(0xffffffffu, k2));
      }
    }
    if (lane == 0) rule_cnt[j] = kept;
  }
}

// The slow side of conf_keep, out of line: taken by a handful of splits per million.
__device__ __noinline__ bool conf_keep_exact(int32_t cx, int32_t ca, double n_total, double min_conf) {
  return __ddiv_rn(__ddiv_rn((double)cx, n_total), __ddiv_rn((double)ca, n_total)) >= min_conf;
}

// K5 pass 2: streaming. One warp per itemset; reads the (count(A), count(C)) planes of pass 1,
// recomputes the doubles and writes the kept rules at rule_off[x] + rank (ballot + popc
// rank). Order inside an itemset: per block of 32 slots (the whole itemset for k <= 6),
// first the kept rules A -> C in slot order, then the kept rules C -> A in slot order - so
// that every store instruction of the warp writes one contiguous run.
__global__ void __launch_bounds__(256)
k_rules_write(Arena a, int k, int64_t lvl_begin, int64_t lvl_count, double n_total, double min_conf,
              const int32_t* __restrict__ pa, const int32_t* __restrict__ pc,
              const int64_t* __restrict__ rule_off,
              int2* __restrict__ out_rule, double* __restrict__ out_confidence,
              double* __restrict__ out_lift, int32_t* __restrict__ bad) {
  // `bad` counts the itemsets whose kept rules here differ in number from what pass 1 sized
  // for them (rule_off[j + 1] - rule_off[j]): the two passes decide "confidence >= c" with
  // different instruction sequences, and a disagreement would otherwise write outside the
  // itemset's range unnoticed. The caller treats a non-zero count as an error.
  const int lane = threadIdx.x & 31;
  const int n_sub = 1 << (k - 1);
  const uint32_t full = (1u << k) - 1u;
  const int64_t lvl_first = a.id_off[k];
  const double rcp_n = __drcp_rn(n_total);
  if (n_sub <= 16) {
    // short itemsets: 32 / n_sub itemsets per warp, lane = (itemset in the group, slot), so
    // that the lanes are not left idle by the few splits a small itemset has
    const int per_warp = 32 / n_sub, g = lane / n_sub, s = lane % n_sub;
    const int am = slot_mask(s, k);
    const uint32_t gmask = ((1u << n_sub) - 1u) << (g * n_sub);
    const uint32_t below = gmask & ((1u << lane) - 1u);
    const int64_t n_groups = (lvl_count + per_warp - 1) / per_warp;
    for (int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); w < n_groups;
         w += (int64_t)gridDim.x * 8) {
      const int64_t j = w * per_warp + g;
      RuleVals r;
      r.k1 = r.k2 = false;
      int64_t gid = 0;
      if (j < lvl_count && s >= 1) {
        gid = lvl_first + lvl_begin + j;
        const int32_t ca = __ldg(pa + j * n_sub + s), cc = __ldg(pc + j * n_sub + s);
        const double s_x = div_by_rcp((double)__ldg(a.counts + gid), n_total, rcp_n);
        if (ca >= 0 && cc >= 0) r = rule_values(s_x, ca, cc, n_total, rcp_n, min_conf);
      }
      const uint32_t m1 = __ballot_sync(0xffffffffu, r.k1), m2 = __ballot_sync(0xffffffffu, r.k2);
      if (j < lvl_count && s == 0 &&
          rule_off[j + 1] - rule_off[j] != (int64_t)(__popc(m1 & gmask) + __popc(m2 & gmask)))
        atomicAdd(bad, 1);
      if (r.k1 || r.k2) {
        const int64_t base = rule_off[j];
        if (r.k1) {
          const int64_t o = base + __popc(m1 & below);
          out_rule[o] = make_int2((int32_t)gid, am);
          out_confidence[o] = r.c1;
          out_lift[o] = r.l1;
        }
        if (r.k2) {
          const int64_t o = base + __popc(m1 & gmask) + __popc(m2 & below);
          out_rule[o] = make_int2((int32_t)gid, (int32_t)(full ^ (uint32_t)am));
          out_confidence[o] = r.c2;
          out_lift[o] = r.l2;
        }
      }
    }
    return;
  }
  const uint32_t below = (1u << lane) - 1u;
  for (int64_t j = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); j < lvl_count;
       j += (int64_t)gridDim.x * 8) {
    const int64_t gid = lvl_first + lvl_begin + j;
    const int32_t* pa_j = pa + j * n_sub;
    const int32_t* pc_j = pc + j * n_sub;
    // the loads of kWriteIlp blocks of 32 slots are issued before the first is consumed
    int32_t ca[kWriteIlp], cc[kWriteIlp];
#pragma unroll
    for (int u = 0; u < kWriteIlp; ++u) {
      const int s = u * 32 + lane;
      ca[u] = cc[u] = -1;
      if (s >= 1 && s < n_sub) {
        ca[u] = __ldg(pa_j + s);
        cc[u] = __ldg(pc_j + s);
      }
    }
    const double s_x = div_by_rcp((double)__ldg(a.counts + gid), n_total, rcp_n);
    int64_t pos = rule_off[j];
    for (int s0 = 0; s0 < n_sub; s0 += 32 * kWriteIlp) {
      int32_t na[kWriteIlp], nc[kWriteIlp];
#pragma unroll
      for (int u = 0; u < kWriteIlp; ++u) {     // next iteration's loads
        const int s = s0 + (kWriteIlp + u) * 32 + lane;
        na[u] = nc[u] = -1;
        if (s < n_sub) {
          na[u] = __ldg(pa_j + s);
          nc[u] = __ldg(pc_j + s);
        }
      }
#pragma unroll
      for (int u = 0; u < kWriteIlp; ++u) {
        const int s = s0 + u * 32 + lane;
        if (s0 + u * 32 < n_sub) {              // warp-uniform
          const int am = slot_mask(s, k);
          RuleVals r;
          r.k1 = r.k2 = false;
          if (ca[u] >= 0 && cc[u] >= 0) r = rule_values(s_x, ca[u], cc[u], n_total, rcp_n, min_conf);
          const uint32_t m1 = __ballot_sync(0xffffffffu, r.k1), m2 = __ballot_sync(0xffffffffu, r.k2);
          if (r.k1) {
            const int64_t o = pos + __popc(m1 & below);
            out_rule[o] = make_int2((int32_t)gid, am);
            out_confidence[o] = r.c1;
            out_lift[o] = r.l1;
          }
          if (r.k2) {
            const int64_t o = pos + __popc(m1) + __popc(m2 & below);
            out_rule[o] = make_int2((int32_t)gid, (int32_t)(full ^ (uint32_t)am));
            out_confidence[o] = r.c2;
            out_lift[o] = r.l2;
          }
          pos += __popc(m1) + __popc(m2);
        }
      }
#pragma unroll
      for (int u = 0; u < kWriteIlp; ++u) {
        ca[u] = na[u];
        cc[u] = nc[u];
      }
    }
    if (lane == 0 && pos != rule_off[j + 1]) atomicAdd(bad, 1);
  }
}

// Kept rules of one split, 0..2 (the pass-1 decision of conf_keep for A -> C and C -> A).
__device__ __forceinline__ int keep_pair(float fx, int32_t c_x, int32_t va, int32_t vc, float thr_hi, float thr_lo,
                                         double n_total, double min_conf, int32_t* missing) {
  if ((va | vc) < 0) {
    atomicAdd(missing, 1);
    return 0;
  }
  const float fa = (float)va, fc = (float)vc;
  bool k1 = fx > fa * thr_hi, k2 = fx > fc * thr_hi;
  const bool u1 = !k1 && !(fx < fa * thr_lo), u2 = !k2 && !(fx < fc * thr_lo);
  if (u1 | u2) {                        // within 1e-5 of the threshold: the exact doubles
    if (u1) k1 = conf_keep_exact(c_x, va, n_total, min_conf);
    if (u2) k2 = conf_keep_exact(c_x, vc, n_total, min_conf);
  }
  return (int)k1 + (int)k2;
}

// Gather route, FOUR consecutive slots per lane (k >= 4): the kernel above is bound by
// instruction issue - about 130 instructions per 32 slots, most of them 64-bit address
// arithmetic - not by memory. Here a lane owns slots s .. s+3 (s a multiple of 4): count(A)
// comes as two 8-byte loads (the even slots from the parent's plane A at s >> 1, the odd ones
// from its plane C at top - 2 - (s >> 1)), count(C) as ONE 16-byte load from the X_i row (all
// four slots share the highest set bit h when s >= 4), the results leave as two 16-byte
// stores: a quarter of the memory instructions and address computations per slot. LPI lanes
// serve one itemset (2^(k-1) / 4 of them, at most 32), 32 / LPI itemsets share a warp. The
// lane with s = 0 handles slots 0..3 separately (slot 0 makes no rule; slots 1, 2, 3 have
// h = 0, 1, 1). Same planes and counts as k_rules_gather.
// resident CTAs per SM asked of ptxas (A/B builds with tools/build_variant.sh: ptxas' own choice of
// 56 registers gives 8.46 ms over all levels at Z1M, 5 CTAs = 48 registers 7.96 ms, 6 CTAs = 40 registers 10.95 ms)
#ifndef KMLS_GATHER_MINB
#define KMLS_GATHER_MINB 5
#endif
template <int LPI>
__global__ void __launch_bounds__(256, KMLS_GATHER_MINB)
k_rules_gather_v4(Arena a, int k, const int32_t* __restrict__ sub, const int32_t* __restrict__ part_prev,
                  double n_total, double min_conf, int32_t* __restrict__ part,
                  int32_t* __restrict__ rule_cnt, int32_t* __restrict__ missing) {
  constexpr int IPW = 32 / LPI;
  const int lane = threadIdx.x & 31;
  const int g = lane / LPI, l = lane % LPI;
  const int n_sub = 1 << (k - 1), top = n_sub >> 1;
  const int n_iter = n_sub > 128 ? n_sub >> 7 : 1;
  const int64_t n_k = a.id_off[k + 1] - a.id_off[k], n_p = a.id_off[k] - a.id_off[k - 1];
  const float thr_hi = (float)(min_conf * 1.00001), thr_lo = (float)(min_conf * 0.99999);
  const int32_t* __restrict__ counts_k = a.counts + a.id_off[k];
  const int32_t* __restrict__ par_k = a.par + (a.id_off[k] - a.id_off[2]);
  const int32_t* __restrict__ pa_prev = part_prev;
  const int32_t* __restrict__ pc_prev = part_prev + n_p * top;
  int32_t* __restrict__ pa = part;
  int32_t* __restrict__ pc = part + n_k * n_sub;
  const int64_t stride = (int64_t)gridDim.x * 8 * IPW;
  // The warp's work is a flat stream of units (IPW itemsets, 128 slots of each); the header
  // words of the NEXT unit (count, parent row, the X_i row its slots read) are loaded before the
  // current unit's planes, so that consecutive units' dependent loads (header -> planes) overlap.
  int64_t j0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * IPW;
  int it = 0;
  int32_t c_x = 0, row_p = 0, r_a = -1, r_b = -1;      // r_a: X_i row of the lane's slots (s == 0: of slot 1), r_b: s == 0 only, of slots 2, 3
  auto load_rows = [&](int64_t j, int s, int32_t& ra, int32_t& rb) {
    const int32_t* subj = sub + j * (k - 1);
    if (s == 0) {
      ra = __ldg(subj + (k - 2));
      rb = __ldg(subj + (k - 3));
    } else {
      ra = __ldg(subj + (k - 2 - (31 - __clz(s))));
    }
  };
  if (j0 + g < n_k) {
    c_x = __ldg(counts_k + j0 + g);
    row_p = __ldg(par_k + j0 + g);
    load_rows(j0 + g, 4 * l, r_a, r_b);
  }
  int kept = 0;
  while (j0 < n_k) {
    const int64_t j = j0 + g;
    const bool active = j < n_k;
    // next unit's header
    int64_t j1 = j0;
    int it1 = it + 1;
    if (it1 >= n_iter) { it1 = 0; j1 = j0 + stride; }
    int32_t n_cx = c_x, n_rp = row_p, n_ra = -1, n_rb = -1;
    if (j1 + g < n_k) {
      if (it1 == 0) {
        n_cx = __ldg(counts_k + j1 + g);
        n_rp = __ldg(par_k + j1 + g);
      }
      load_rows(j1 + g, (it1 << 7) + 4 * l, n_ra, n_rb);
    }
    if (active) {
      const float fx = (float)c_x;
      const int32_t* pa_p = pa_prev + (int64_t)row_p * top;
      const int32_t* pc_p = pc_prev + (int64_t)row_p * top;
      const int s = (it << 7) + 4 * l;
      const int2 e = __ldg(reinterpret_cast<const int2*>(pa_p + (s >> 1)));
      const int2 o = __ldg(reinterpret_cast<const int2*>(pc_p + (top - 2 - (s >> 1))));
      int4 va = make_int4(e.x, o.y, e.y, o.x), vc;
      if (s == 0) {
        va.x = c_x;
        vc.x = c_x;
        vc.y = r_a < 0 ? -1 : __ldg(pc_prev + (int64_t)r_a * top);
        vc.z = vc.w = -1;
        if (r_b >= 0) {
          const int2 t2 = __ldg(reinterpret_cast<const int2*>(pc_prev + (int64_t)r_b * top));
          vc.z = t2.x;
          vc.w = t2.y;
        }
      } else {
        const int h = 31 - __clz(s);
        vc = make_int4(-1, -1, -1, -1);
        if (r_a >= 0) vc = __ldg(reinterpret_cast<const int4*>(pc_prev + (int64_t)r_a * top + (s - (1 << h))));
        kept += keep_pair(fx, c_x, va.x, vc.x, thr_hi, thr_lo, n_total, min_conf, missing);
      }
      kept += keep_pair(fx, c_x, va.y, vc.y, thr_hi, thr_lo, n_total, min_conf, missing);
      kept += keep_pair(fx, c_x, va.z, vc.z, thr_hi, thr_lo, n_total, min_conf, missing);
      kept += keep_pair(fx, c_x, va.w, vc.w, thr_hi, thr_lo, n_total, min_conf, missing);
      *reinterpret_cast<int4*>(pa + j * n_sub + s) = va;
      *reinterpret_cast<int4*>(pc + j * n_sub + s) = vc;
    }
    if (it1 == 0) {                                   // the itemsets of this unit are complete
#pragma unroll
      for (int d = LPI >> 1; d > 0; d >>= 1) kept += __shfl_xor_sync(0xffffffffu, kept, d);
      if (active && l == 0) rule_cnt[j] = kept;
      kept = 0;
    }
    c_x = n_cx;
    row_p = n_rp;
    r_a = n_ra;
    r_b = n_rb;
    j0 = j1;
    it = it1;
  }
}

// Pass 2 for k >= 6, software-pipelined like k_rules_gather_v4: the warp's work is a flat stream of units
// (IPW itemsets x BPI blocks of 32 slots, IPW * BPI = 2) and the plane words AND header words
// (count, rule offsets) of the NEXT unit are loaded before the current one is computed - across
// itemset boundaries too, where k_rules_write's look-ahead stops and every itemset's first
// loads are exposed to the full (write-loaded) memory latency. Same output, same order.
#ifndef KMLS_WRITE_MINB
#define KMLS_WRITE_MINB 4
#endif
template <int IPW>
__global__ void __launch_bounds__(256, KMLS_WRITE_MINB)
k_rules_write_mlp(Arena a, int k, int64_t lvl_begin, int64_t lvl_count, double n_total, double min_conf,
                  const int32_t* __restrict__ pa, const int32_t* __restrict__ pc,
                  const int64_t* __restrict__ rule_off,
                  int2* __restrict__ out_rule, double* __restrict__ out_confidence,
                  double* __restrict__ out_lift, int32_t* __restrict__ bad) {
  constexpr int BPI = 2 / IPW;
  constexpr uint32_t kFull = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int n_sub = 1 << (k - 1), nblk = n_sub >> 5;
  const uint32_t full = (1u << k) - 1u;
  const uint32_t below = (1u << lane) - 1u;
  const int64_t lvl_first = a.id_off[k];
  const int32_t* __restrict__ counts_l = a.counts + lvl_first + lvl_begin;
  const double rcp_n = __drcp_rn(n_total);
  const int64_t stride = (int64_t)gridDim.x * 8 * IPW;
  int64_t j0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * IPW;
  int b0 = 0;
  if (j0 >= lvl_count) return;

  int32_t ca[2], cc[2], n_ca[2], n_cc[2];
  int32_t cnt[IPW], n_cnt[IPW];
  int64_t off[IPW + 1], n_off[IPW + 1];
  // unit (jj, bb): block u belongs to itemset jj + (IPW == 2 ? u : 0), block bb + (IPW == 2 ? 0 : u)
  auto load_unit = [&](int64_t jj, int bb, int32_t* va, int32_t* vc) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int64_t j = jj + (IPW == 2 ? u : 0);
      const int s = ((bb + (IPW == 2 ? 0 : u)) << 5) + lane;
      va[u] = vc[u] = -1;
      if (j < lvl_count && s >= 1 && s < n_sub) {
        va[u] = __ldg(pa + j * n_sub + s);
        vc[u] = __ldg(pc + j * n_sub + s);
      }
    }
  };
  auto load_header = [&](int64_t jj, int32_t* c, int64_t* o) {
#pragma unroll
    for (int t = 0; t < IPW; ++t) {
      c[t] = 0;
      if (jj + t < lvl_count) c[t] = __ldg(counts_l + jj + t);
    }
#pragma unroll
    for (int t = 0; t <= IPW; ++t) {
      o[t] = 0;
      if (jj + t <= lvl_count) o[t] = __ldg(rule_off + jj + t);
    }
  };
  load_unit(j0, 0, ca, cc);
  load_header(j0, cnt, off);
  double s_x[IPW];
  int64_t pos[IPW];
#pragma unroll
  for (int t = 0; t < IPW; ++t) { s_x[t] = 0.0; pos[t] = 0; }

  while (j0 < lvl_count) {
    int64_t j1 = j0;
    int b1 = b0 + BPI;
    if (b1 >= nblk) { b1 = 0; j1 = j0 + stride; }
    const bool next_is_new = b1 == 0;                  // warp-uniform
    load_unit(j1, b1, n_ca, n_cc);
    if (next_is_new) load_header(j1, n_cnt, n_off);
    if (b0 == 0) {
#pragma unroll
      for (int t = 0; t < IPW; ++t) {
        s_x[t] = div_by_rcp((double)cnt[t], n_total, rcp_n);
        pos[t] = off[t];
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int t = IPW == 2 ? u : 0;
      const int b = b0 + (IPW == 2 ? 0 : u);
      if (b >= nblk || j0 + t >= lvl_count) continue;  // warp-uniform
      const int s = (b << 5) + lane;
      const int am = slot_mask(s, k);
      const int32_t gid = (int32_t)(lvl_first + lvl_begin + j0 + t);
      RuleVals r;
      r.k1 = r.k2 = false;
      if (ca[u] >= 0 && cc[u] >= 0) r = rule_values(s_x[t], ca[u], cc[u], n_total, rcp_n, min_conf);
      const uint32_t m1 = __ballot_sync(kFull, r.k1), m2 = __ballot_sync(kFull, r.k2);
      const int n1 = __popc(m1);
      if (r.k1) {
        const int64_t o = pos[t] + __popc(m1 & below);
        out_rule[o] = make_int2(gid, am);
        out_confidence[o] = r.c1;
        out_lift[o] = r.l1;
      }
      if (r.k2) {
        const int64_t o = pos[t] + n1 + __popc(m2 & below);
        out_rule[o] = make_int2(gid, (int32_t)(full ^ (uint32_t)am));
        out_confidence[o] = r.c2;
        out_lift[o] = r.l2;
      }
      pos[t] += n1 + __popc(m2);
    }
    if (next_is_new) {
#pragma unroll
      for (int t = 0; t < IPW; ++t)
        if (lane == 0 && j0 + t < lvl_count && pos[t] != off[t + 1]) atomicAdd(bad, 1);
#pragma unroll
      for (int t = 0; t < IPW; ++t) cnt[t] = n_cnt[t];
#pragma unroll
      for (int t = 0; t <= IPW; ++t) off[t] = n_off[t];
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      ca[u] = n_ca[u];
      cc[u] = n_cc[u];
    }
    j0 = j1;
    b0 = b1;
  }
}

// support column of the rule table, on demand: support[r] = count(itemset of rule r) / N
// (kept out of pass 2 because it is 8 of every 32 bytes the pass would write)
__global__ void __launch_bounds__(256)
k_rule_supports(const int2* __restrict__ rule, int64_t n, const int32_t* __restrict__ counts,
                double n_total, double* __restrict__ out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n;
       r += (int64_t)gridDim.x * blockDim.x)
    out[r] = __ddiv_rn((double)__ldg(counts + rule[r].x), n_total);
}

// Order-independent checksum of a rule table, for comparing a whole table with another
// implementation's without moving it: sum over rules of
//   mix(hA ^ mix(hC + 0x1234) ^ mix(bits(confidence)) ^ mix(bits(lift) + 1) ^ mix(bits(support) + 2)),
// hA / hC = sum of mix(track id + 1) over the antecedent / consequent members.
__global__ void __launch_bounds__(256)
k_rule_checksum(Arena a, const int2* __restrict__ rule, const double* __restrict__ conf,
                const double* __restrict__ lift, int64_t n, const int32_t* __restrict__ row_items,
                double n_total, unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n;
       r += (int64_t)gridDim.x * blockDim.x) {
    const int2 ru = rule[r];
    const int64_t gid = ru.x;
    int k = 2;
    while (gid >= a.id_off[k + 1]) ++k;
    const int32_t* x = itemset_ptr(a, k, gid);
    uint64_t ha = 0, hc = 0;
    for (int t = 0; t < k; ++t) {
      const uint64_t h = mix64((uint64_t)__ldg(row_items + __ldg(x + t)) + 1);
      if ((uint32_t)ru.y >> t & 1u) ha += h; else hc += h;
    }
    const double sx = __ddiv_rn((double)__ldg(a.counts + gid), n_total);
    acc += mix64(ha ^ mix64(hc + 0x1234) ^ mix64((uint64_t)__double_as_longlong(conf[r])) ^
                 mix64((uint64_t)__double_as_longlong(lift[r]) + 1) ^
                 mix64((uint64_t)__double_as_longlong(sx) + 2));
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

// every quotient of a split through div_by_rcp and through __ddiv_rn; counts differing bits
__global__ void __launch_bounds__(256)
k_division_selftest(double n_total, int64_t n_samples, uint64_t seed, unsigned long long* __restrict__ out) {
  const double rcp_n = __drcp_rn(n_total);
  const uint64_t n = (uint64_t)n_total;
  unsigned long long bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_samples;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t h0 = mix64(seed + 3 * (uint64_t)i), h1 = mix64(h0 + 1), h2 = mix64(h1 + 2);
    // cX in [1, N], skewed to small counts every other sample; cA, cC in [cX, N]
    uint64_t cx = 1 + h0 % n;
    if (i & 1) cx = 1 + cx % (1 + n / 1000);
    const uint64_t ca = cx + h1 % (n - cx + 1), cc = cx + h2 % (n - cx + 1);
    const double s_x = div_by_rcp((double)cx, n_total, rcp_n);
    const RuleVals r = rule_values(s_x, (int32_t)ca, (int32_t)cc, n_total, rcp_n, 0.0);
    const double e_x = __ddiv_rn((double)cx, n_total), e_a = __ddiv_rn((double)ca, n_total),
                 e_c = __ddiv_rn((double)cc, n_total);
    const double e_c1 = __ddiv_rn(e_x, e_a), e_c2 = __ddiv_rn(e_x, e_c);
    bad += (s_x != e_x) + (r.c1 != e_c1) + (r.c2 != e_c2) + (r.l1 != __ddiv_rn(e_c1, e_c)) +
           (r.l2 != __ddiv_rn(e_c2, e_a));
  }
  for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out, bad);
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
                                     int32_t max_k, void* d_nodes, kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, d_par, level_sizes, max_k);
  if (rc) return rc;
  const int64_t n = a.id_off[max_k + 1];
  if (n <= 0) return 0;
  KMLS_REQUIRE(d_nodes && ((uintptr_t)d_nodes & 15) == 0, "kmls_trie_index_build: d_nodes must be 16-byte aligned");
  KMLS_REQUIRE(max_k < 2 || d_par, "kmls_trie_index_build: d_par is null");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  KMLS_CUDA(cudaMemsetAsync(d_nodes, 0, sizeof(int4) * (size_t)n, (cudaStream_t)stream));
  k_trie_build<<<grid, 256, 0, (cudaStream_t)stream>>>(a, (int4*)d_nodes);
  KMLS_LAUNCHED("k_trie_build");
  return 0;
}

extern "C" int64_t kmls_rules_parts_per_itemset(int32_t k) {
  return k >= 2 && k <= kMaxRuleK ? ((int64_t)1 << (k - 1)) : -1;
}

extern "C" int kmls_rules_count(const int32_t* d_items, const int32_t* d_counts,
                                const int64_t* level_sizes, int32_t max_k, int32_t k,
                                int64_t lvl_begin, int64_t lvl_count, const void* d_nodes,
                                int64_t n_playlists_total, double min_confidence,
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
  KMLS_REQUIRE(d_nodes && d_part && d_rule_cnt && d_missing, "kmls_rules_count: null pointer");
  const size_t per_warp = sizeof(int32_t) * 3 * ((size_t)1 << (k - 1));
  int warps = (int)std::min<size_t>(8, (200 * 1024) / per_warp);
  KMLS_REQUIRE(warps >= 1, "kmls_rules_count: subset table does not fit in shared memory");
  const size_t smem = per_warp * warps;
  KMLS_CUDA(cudaFuncSetAttribute(k_rules_count, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  // chunks long enough to reuse the subset table, short enough to keep every SM busy
  const int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(kRuleChunkMax, lvl_count / ((int64_t)num_sms() * 64)));
  const int64_t n_chunks = (lvl_count + chunk - 1) / chunk;
  int grid = (int)std::min<int64_t>((n_chunks + warps - 1) / warps, (int64_t)num_sms() * 32);
  k_rules_count<<<grid, warps * 32, smem, (cudaStream_t)stream>>>(
      a, k, lvl_begin, lvl_count, (const int4*)d_nodes, (double)n_playlists_total, min_confidence,
      chunk, (int32_t*)d_part, d_rule_cnt, d_missing);
  KMLS_LAUNCHED("k_rules_count");
  return 0;
}

extern "C" int kmls_rules_subsets(const int32_t* d_items, const int32_t* d_par,
                                  const int64_t* level_sizes, int32_t max_k, int32_t k,
                                  const void* d_nodes, const int32_t* d_sub_prev, int32_t* d_sub,
                                  int32_t* d_missing, kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, nullptr, d_par, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(k >= 2 && k <= max_k && k <= kMaxRuleK, "kmls_rules_subsets: level out of range");
  const int64_t n_k = level_sizes[k];
  if (n_k == 0) return 0;
  KMLS_REQUIRE(d_items && d_par && d_nodes && d_sub && d_missing && (k == 2 || d_sub_prev),
               "kmls_rules_subsets: null pointer");
  static const bool old_kernel = [] { const char* e = getenv("KMLS_RULES_SUBSETS"); return e && e[0] == 'o'; }();
  if (k >= 3 && !old_kernel) {
    const int g = (int)std::min<int64_t>((n_k + 255) / 256, (int64_t)num_sms() * 32);
    k_rules_subsets_row<<<g, 256, 0, (cudaStream_t)stream>>>(a, k, (const int4*)d_nodes, d_sub_prev, d_sub, d_missing);
    KMLS_LAUNCHED("k_rules_subsets_row");
    return 0;
  }
  dim3 grid((unsigned)std::min<int64_t>((n_k + 255) / 256, (int64_t)num_sms() * 8), (unsigned)(k - 1));
  k_rules_subsets<<<grid, 256, 0, (cudaStream_t)stream>>>(a, k, (const int4*)d_nodes, d_sub_prev, d_sub,
                                                          d_missing);
  KMLS_LAUNCHED("k_rules_subsets");
  return 0;
}

extern "C" int kmls_rules_count_gather(const int32_t* d_items, const int32_t* d_counts,
                                       const int32_t* d_par, const int64_t* level_sizes,
                                       int32_t max_k, int32_t k, const int32_t* d_sub,
                                       const void* d_part_prev, int64_t n_playlists_total,
                                       double min_confidence, void* d_part, int32_t* d_rule_cnt,
                                       int32_t* d_missing, kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, d_par, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(k >= 2 && k <= max_k && k <= kMaxRuleK, "kmls_rules_count_gather: level out of range");
  KMLS_REQUIRE(n_playlists_total > 0, "kmls_rules_count_gather: n_playlists_total must be positive");
  const int64_t n_k = level_sizes[k];
  if (n_k == 0) return 0;
  KMLS_REQUIRE(d_items && d_counts && d_par && d_sub && d_part && d_rule_cnt && d_missing &&
                   (k == 2 || d_part_prev), "kmls_rules_count_gather: null pointer");
  static const bool old_long = [] { const char* e = getenv("KMLS_RULES_GATHER"); return e && e[0] == 'o'; }();
  const bool aligned16 = ((uintptr_t)d_part & 15) == 0 && ((uintptr_t)d_part_prev & 15) == 0;
  if (k >= 4 && !old_long && aligned16) {
    const int lpi = k >= 8 ? 32 : 1 << (k - 3);          // lanes per itemset = slots / 4, at most a warp
    const int64_t warps = (n_k * lpi + 31) / 32;
    const int grid = (int)std::min<int64_t>((warps + 7) / 8, (int64_t)num_sms() * 32);
    auto kern = lpi == 32 ? k_rules_gather_v4<32> : lpi == 16 ? k_rules_gather_v4<16> : lpi == 8 ? k_rules_gather_v4<8>
              : lpi == 4 ? k_rules_gather_v4<4> : k_rules_gather_v4<2>;
    kern<<<grid, 256, 0, (cudaStream_t)stream>>>(a, k, d_sub, (const int32_t*)d_part_prev, (double)n_playlists_total,
                                                 min_confidence, (int32_t*)d_part, d_rule_cnt, d_missing);
    KMLS_LAUNCHED("k_rules_gather_v4");
    return 0;
  }
  const int per_warp = k <= 5 ? 32 >> (k - 1) : 1;
  const int64_t warps = (n_k + per_warp - 1) / per_warp;
  int grid = (int)std::min<int64_t>((warps + 7) / 8, (int64_t)num_sms() * 32);
  k_rules_gather<<<grid, 256, 0, (cudaStream_t)stream>>>(
      a, k, d_sub, (const int32_t*)d_part_prev, (double)n_playlists_total, min_confidence, (int32_t*)d_part,
      d_rule_cnt, d_missing);
  KMLS_LAUNCHED("k_rules_gather");
  return 0;
}

extern "C" int kmls_rules_write(const int32_t* d_counts, const int64_t* level_sizes, int32_t max_k,
                                int32_t k, int64_t lvl_begin, int64_t lvl_count,
                                int64_t n_playlists_total, double min_confidence,
                                const void* d_part, const int64_t* d_rule_off, int32_t* d_out_rule,
                                double* d_out_confidence, double* d_out_lift, int32_t* d_mismatch,
                                kmls_stream_t stream) {
  KMLS_REQUIRE(k >= 2 && k <= kMaxRuleK && lvl_count >= 0, "kmls_rules_write: level out of range");
  const int32_t* pa = (const int32_t*)d_part;
  return kmls_rules_write_planes(d_counts, level_sizes, max_k, k, lvl_begin, lvl_count, n_playlists_total,
                                 min_confidence, pa, pa ? pa + lvl_count * ((int64_t)1 << (k - 1)) : nullptr,
                                 d_rule_off, d_out_rule, d_out_confidence, d_out_lift, d_mismatch, stream);
}

extern "C" int kmls_rules_write_planes(const int32_t* d_counts, const int64_t* level_sizes, int32_t max_k,
                                       int32_t k, int64_t lvl_begin, int64_t lvl_count,
                                       int64_t n_playlists_total, double min_confidence,
                                       const int32_t* d_plane_a, const int32_t* d_plane_c,
                                       const int64_t* d_rule_off, int32_t* d_out_rule,
                                       double* d_out_confidence, double* d_out_lift, int32_t* d_mismatch,
                                       kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, nullptr, d_counts, nullptr, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(k >= 2 && k <= max_k && k <= kMaxRuleK, "kmls_rules_write: level out of range");
  KMLS_REQUIRE(lvl_begin >= 0 && lvl_count >= 0 && lvl_begin + lvl_count <= level_sizes[k],
               "kmls_rules_write: itemset range out of bounds");
  if (lvl_count == 0) return 0;
  KMLS_REQUIRE(d_plane_a && d_plane_c && d_rule_off && d_out_rule && d_out_confidence && d_out_lift && d_mismatch &&
                   ((uintptr_t)d_out_rule & 7) == 0, "kmls_rules_write: null or misaligned pointer");
  static const bool old_long = [] { const char* e = getenv("KMLS_RULES_WRITE"); return e && e[0] == 'o'; }();
  if (k >= 6 && !old_long) {
    const int ipw = k == 6 ? 2 : 1;
    const int64_t warps = (lvl_count + ipw - 1) / ipw;
    const int grid = (int)std::min<int64_t>((warps + 7) / 8, (int64_t)num_sms() * 32);
    auto kern = ipw == 2 ? k_rules_write_mlp<2> : k_rules_write_mlp<1>;
    kern<<<grid, 256, 0, (cudaStream_t)stream>>>(a, k, lvl_begin, lvl_count, (double)n_playlists_total, min_confidence,
                                                 d_plane_a, d_plane_c, d_rule_off, (int2*)d_out_rule, d_out_confidence,
                                                 d_out_lift, d_mismatch);
    KMLS_LAUNCHED("k_rules_write_mlp");
    return 0;
  }
  int grid = (int)std::min<int64_t>((lvl_count + 7) / 8, (int64_t)num_sms() * 32);
  k_rules_write<<<grid, 256, 0, (cudaStream_t)stream>>>(
      a, k, lvl_begin, lvl_count, (double)n_playlists_total, min_confidence, d_plane_a, d_plane_c,
      d_rule_off, (int2*)d_out_rule, d_out_confidence, d_out_lift, d_mismatch);
  KMLS_LAUNCHED("k_rules_write");
  return 0;
}

extern "C" int kmls_rule_supports(const int32_t* d_rule, int64_t n_rules, const int32_t* d_counts,
                                  int64_t n_playlists_total, double* d_out, kmls_stream_t stream) {
  KMLS_REQUIRE(n_rules >= 0 && n_playlists_total > 0, "kmls_rule_supports: bad size");
  if (n_rules == 0) return 0;
  KMLS_REQUIRE(d_rule && d_counts && d_out, "kmls_rule_supports: null pointer");
  int grid = (int)std::min<int64_t>((n_rules + 255) / 256, (int64_t)num_sms() * 16);
  k_rule_supports<<<grid, 256, 0, (cudaStream_t)stream>>>((const int2*)d_rule, n_rules, d_counts,
                                                          (double)n_playlists_total, d_out);
  KMLS_LAUNCHED("k_rule_supports");
  return 0;
}

extern "C" int kmls_rule_checksum(const int32_t* d_items, const int32_t* d_counts,
                                  const int64_t* level_sizes, int32_t max_k, const int32_t* d_rule,
                                  const double* d_confidence, const double* d_lift, int64_t n_rules,
                                  const int32_t* d_row_items, int64_t n_playlists_total,
                                  uint64_t* d_out, kmls_stream_t stream) {
  Arena a;
  int rc = make_arena(a, d_items, d_counts, nullptr, level_sizes, max_k);
  if (rc) return rc;
  KMLS_REQUIRE(n_rules >= 0 && n_playlists_total > 0 && d_out, "kmls_rule_checksum: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  KMLS_CUDA(cudaMemsetAsync(d_out, 0, sizeof(uint64_t), s));
  if (n_rules == 0) return 0;
  KMLS_REQUIRE(d_rule && d_confidence && d_lift && d_row_items, "kmls_rule_checksum: null pointer");
  int grid = (int)std::min<int64_t>((n_rules + 255) / 256, (int64_t)num_sms() * 16);
  k_rule_checksum<<<grid, 256, 0, s>>>(a, (const int2*)d_rule, d_confidence, d_lift, n_rules, d_row_items,
                                       (double)n_playlists_total, (unsigned long long*)d_out);
  KMLS_LAUNCHED("k_rule_checksum");
  return 0;
}

extern "C" int kmls_rule_division_selftest(int64_t n_playlists_total, int64_t n_samples, uint64_t seed,
                                           uint64_t* d_mismatches, kmls_stream_t stream) {
  KMLS_REQUIRE(n_playlists_total > 0 && n_playlists_total <= 0x7fffffff && n_samples >= 0 && d_mismatches,
               "kmls_rule_division_selftest: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  KMLS_CUDA(cudaMemsetAsync(d_mismatches, 0, sizeof(uint64_t), s));
  if (n_samples == 0) return 0;
  int grid = (int)std::min<int64_t>((n_samples + 255) / 256, (int64_t)num_sms() * 16);
  k_division_selftest<<<grid, 256, 0, s>>>((double)n_playlists_total, n_samples, seed,
                                           (unsigned long long*)d_mismatches);
  KMLS_LAUNCHED("k_division_selftest");
  return 0;
}
