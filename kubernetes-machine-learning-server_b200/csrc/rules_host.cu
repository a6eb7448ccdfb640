// kmls_rules: rule generation behind one C call (the host loop around the kernels of
// rules.cu). Semantics: mlxtend.frequent_patterns.association_rules(df, metric="confidence",
// min_threshold=c) over the itemset table kmls_mine produced (the
// reference's only min_confidence site is its dormant path, machine-learning/main.py:224-260).
//
//   concatenate the level tables ("arena") -> subset index (k_trie_build)
//   pass 1 per level: subset counts of every (itemset, antecedent) split + kept-rule counts
//          gather route (whole table on this GPU) or search route (any slice of the table)
//   scan of the kept counts -> pass 2 per level: the doubles, written compacted
//
// A multi-GPU job gives every rank the whole itemset table and a slice (part i of n, cut by
// work) of the rule table.
#include <cstring>

#include "common.cuh"
#include "kmls_b200.h"
#include "profile.cuh"
#include "workspace.cuh"

namespace kmls {

__global__ void k_post2(const int64_t* a, const int32_t* c, volatile int64_t* dst) {
  if (a) dst[0] = *a;
  if (c) dst[2] = *c;
  __threadfence_system();
}

struct RulesHost {
  Workspace ws;
  cudaStream_t s;
  volatile int64_t* box;
  size_t needed_hint = 0;
  RulesHost(void* base, size_t bytes, cudaStream_t st) : ws(base, bytes), s(st), box(mailbox()) {}
  template <class T>
  int take(T** out, int64_t n, bool high, const char* what) {
    *out = (T*)ws.alloc(sizeof(T) * (size_t)std::max<int64_t>(n, 1), high);
    if (*out == nullptr)
      return fail(kCapacityError,
                  "kmls_rules: workspace too small for %s: %zu more bytes needed while %zu are in use "
                  "(call again with a larger workspace; nothing was truncated)", what, ws.failed_bytes, ws.in_use);
    return 0;
  }
};

#define KMLS_TRY(expr)        \
  do {                        \
    int rc_ = (expr);         \
    if (rc_) return rc_;      \
  } while (0)

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_rules(const kmls_mine_result* mined, const void* d_tables_base, double min_confidence,
                          int32_t part, int32_t n_parts, int32_t pass1, int64_t max_rules,
                          void* d_workspace, size_t workspace_bytes, kmls_rules_result* R,
                          kmls_stream_t stream) {
  KMLS_REQUIRE(mined && R, "kmls_rules: null argument");
  memset(R, 0, sizeof(*R));
  R->rule_off = R->confidence_off = R->lift_off = R->arena_items_off = R->arena_counts_off = -1;
  KMLS_REQUIRE(n_parts >= 1 && part >= 0 && part < n_parts, "kmls_rules: bad part");
  KMLS_REQUIRE(pass1 >= 0 && pass1 <= 2, "kmls_rules: pass1 must be 0 (auto), 1 (gather) or 2 (search)");
  KMLS_REQUIRE(d_workspace && ((uintptr_t)d_workspace & 255) == 0, "kmls_rules: workspace must be 256-byte aligned device memory");
  if (max_rules <= 0) max_rules = 4000000000ll;
  const int max_k = mined->max_k;
  KMLS_REQUIRE(max_k >= 0 && max_k <= KMLS_MAX_LEVELS, "kmls_rules: bad level count");
  const int64_t n_total = mined->n_playlists_total;
  cudaStream_t s = (cudaStream_t)stream;
  RulesHost H(d_workspace, workspace_bytes, s);
  KMLS_REQUIRE(H.box != nullptr, "kmls_rules: cannot allocate the pinned mailbox");
  char* const ws_base = (char*)d_workspace;
  auto off_of = [&](const void* p) { return (int64_t)((const char*)p - ws_base); };
  const char* tb = (const char*)d_tables_base;

  int64_t sizes[KMLS_MAX_LEVELS + 2] = {0};
  int64_t n_all = 0, n_members = 0;
  for (int k = 1; k <= max_k; ++k) {
    sizes[k] = mined->level_size[k];
    KMLS_REQUIRE(sizes[k] >= 0, "kmls_rules: negative level size");
    n_all += sizes[k];
    n_members += sizes[k] * k;
  }
  R->n_itemsets = n_all;
  int64_t n_x_all = n_all - (max_k >= 1 ? sizes[1] : 0);
  if (max_k < 2 || n_x_all == 0) return 0;
  KMLS_REQUIRE(n_total > 0, "kmls_rules: n_playlists_total must be positive");

  // ---- arena: the level tables back to back
  int32_t *a_items = nullptr, *a_counts = nullptr, *a_par = nullptr;
  KMLS_TRY(H.take(&a_items, n_members, false, "the itemset arena"));
  KMLS_TRY(H.take(&a_counts, n_all, false, "the itemset arena"));
  KMLS_TRY(H.take(&a_par, n_x_all, true, "the parent column"));
  {
    int64_t io = 0, co = 0, po = 0;
    for (int k = 1; k <= max_k; ++k) {
      if (sizes[k] == 0) continue;
      KMLS_REQUIRE(mined->items_off[k] >= 0 && mined->counts_off[k] >= 0 && (k == 1 || mined->parent_off[k] >= 0),
                   "kmls_rules: level %d has no table", k);
      KMLS_CUDA(cudaMemcpyAsync(a_items + io, tb + mined->items_off[k], sizeof(int32_t) * (size_t)(sizes[k] * k),
                                cudaMemcpyDeviceToDevice, s));
      KMLS_CUDA(cudaMemcpyAsync(a_counts + co, tb + mined->counts_off[k], sizeof(int32_t) * (size_t)sizes[k],
                                cudaMemcpyDeviceToDevice, s));
      if (k >= 2) {
        KMLS_CUDA(cudaMemcpyAsync(a_par + po, tb + mined->parent_off[k], sizeof(int32_t) * (size_t)sizes[k],
                                  cudaMemcpyDeviceToDevice, s));
        po += sizes[k];
      }
      io += sizes[k] * k;
      co += sizes[k];
    }
  }
  R->arena_items_off = off_of(a_items);
  R->arena_counts_off = off_of(a_counts);

  // ---- subset index
  void* nodes = nullptr;
  KMLS_TRY(H.take((char**)&nodes, n_all * 16, true, "the subset index"));
  {
    ProfScope ps("k_trie_build", s, 28.0 * (double)n_all, 0);
    KMLS_TRY(kmls_trie_index_build(a_items, a_counts, a_par, sizes, max_k, nodes, s));
  }
  int64_t launches = 1;

  // ---- this part's slice of the table, cut by work (2^(k-1) splits per k-itemset)
  auto boundary = [&](int64_t num, int64_t den) -> int64_t {
    unsigned __int128 tot = 0;
    for (int k = 2; k <= max_k; ++k) tot += (unsigned __int128)sizes[k] << (k - 1);
    unsigned __int128 target = tot * (unsigned __int128)num / (unsigned __int128)den;
    int64_t x = 0;
    for (int k = 2; k <= max_k; ++k) {
      const unsigned __int128 w = (unsigned __int128)sizes[k] << (k - 1);
      if (target >= w) {
        target -= w;
        x += sizes[k];
      } else {
        return x + (int64_t)(target >> (k - 1));
      }
    }
    return x;
  };
  const int64_t x_begin = boundary(part, n_parts);
  const int64_t x_end = part + 1 < n_parts ? boundary(part + 1, n_parts) : n_x_all;
  const int64_t n_x = x_end - x_begin;
  if (n_x <= 0) return 0;

  // Pass 1 by gathering reads the WHOLE previous level's output, so with n_parts > 1 it runs
  // over whole levels on every rank (replicated work: worth it for 2 or 3 ranks, where it
  // still beats searching a half or a third of the table - Z1M: 11.5 ms for the whole table
  // against about 46 ms / n_parts) and only pass 2 is restricted to the slice.
  if (pass1 == 0) pass1 = n_parts <= 3 ? 1 : 2;
  R->pass1_used = pass1;
  const bool full_pass1 = pass1 == 1 && n_parts > 1;
  struct Lvl {
    int k;
    int64_t begin, count, out0;
    int32_t* part_buf;
  };
  Lvl plan[KMLS_MAX_LEVELS + 2];
  int n_plan = 0;
  int64_t splits = 0;
  {
    int64_t lvl_start = 0;
    for (int k = 2; k <= max_k; ++k) {
      const int64_t lo = std::max(x_begin, lvl_start), hi = std::min(x_end, lvl_start + sizes[k]);
      if (hi > lo) {
        const int64_t ppi = kmls_rules_parts_per_itemset(k);
        if (ppi < 0) return fail(kArgError, "kmls_rules: rule generation supports itemsets of at most 15 members, got %d", k);
        plan[n_plan++] = Lvl{k, lo - lvl_start, hi - lo, lo - x_begin, nullptr};
        splits += (hi - lo) * ppi;
      }
      lvl_start += sizes[k];
    }
  }
  R->n_splits = splits;
  // what the whole call will need at least: say so before pass 1 spends time
  Lvl fplan[KMLS_MAX_LEVELS + 2];     // whole levels (pass 1 of a replicated gather)
  int n_fplan = 0;
  int64_t fsplits = 0;
  if (full_pass1) {
    int64_t lvl_start = 0;
    for (int k = 2; k <= max_k; ++k) {
      if (sizes[k] > 0) {
        const int64_t ppi = kmls_rules_parts_per_itemset(k);
        if (ppi < 0) return fail(kArgError, "kmls_rules: rule generation supports itemsets of at most 15 members, got %d", k);
        fplan[n_fplan++] = Lvl{k, 0, sizes[k], lvl_start, nullptr};
        fsplits += sizes[k] * ppi;
      }
      lvl_start += sizes[k];
    }
  }
  const int64_t p1_splits = full_pass1 ? fsplits : splits, p1_x = full_pass1 ? n_x_all : n_x;
  R->workspace_needed_bytes = H.ws.in_use + (size_t)p1_splits * 8 + (size_t)p1_x * 12 + (size_t)n_plan * 512 + (1u << 20);
  if (R->workspace_needed_bytes > workspace_bytes)
    return fail(kCapacityError,
                "kmls_rules: workspace too small: %lld (itemset, antecedent) splits need %zu bytes before the rule "
                "table itself, %zu given (raise min_support, shard over more GPUs, or give more memory; nothing "
                "was truncated)", (long long)splits, R->workspace_needed_bytes, workspace_bytes);
  Lvl* const p1 = full_pass1 ? fplan : plan;       // the ranges pass 1 runs over
  const int n_p1 = full_pass1 ? n_fplan : n_plan;
  for (int i = 0; i < n_p1; ++i)
    KMLS_TRY(H.take(&p1[i].part_buf, p1[i].count * kmls_rules_parts_per_itemset(p1[i].k) * 2, true,
                    "the subset counts of pass 1"));
  int32_t *rule_cnt_p1 = nullptr, *missing = nullptr;
  KMLS_TRY(H.take(&rule_cnt_p1, p1_x, true, "the rule counts"));
  KMLS_TRY(H.take(&missing, 2, true, "flags"));
  KMLS_CUDA(cudaMemsetAsync(missing, 0, 2 * sizeof(int32_t), s));
  int32_t* const rule_cnt = full_pass1 ? rule_cnt_p1 + x_begin : rule_cnt_p1;   // this part's slice

  if (pass1 == 1) {
    int32_t* sub_prev = nullptr;
    const int32_t* part_prev = nullptr;
    for (int i = 0; i < n_p1; ++i) {
      const Lvl& L = p1[i];
      int32_t* sub = nullptr;
      KMLS_TRY(H.take(&sub, L.count * (L.k - 1), true, "the subset links of pass 1"));
      {
        // one index lookup (32 B) + 4 B written per (itemset, dropped member)
        ProfScope ps("k_rules_subsets", s, 36.0 * (double)L.count * (L.k - 1), L.k);
        KMLS_TRY(kmls_rules_subsets(a_items, a_par, sizes, max_k, L.k, nodes, sub_prev, sub, missing, s));
      }
      {
        // two 4-byte counts read + 8 bytes written per split
        ProfScope ps("k_rules_gather", s, 16.0 * (double)L.count * (double)((int64_t)1 << (L.k - 1)), L.k);
        KMLS_TRY(kmls_rules_count_gather(a_items, a_counts, a_par, sizes, max_k, L.k, sub, part_prev, n_total,
                                         min_confidence, L.part_buf, rule_cnt_p1 + L.out0, missing, s));
      }
      launches += 2;
      H.ws.release(sub_prev);          // stream order: the kernels that read it are already enqueued
      sub_prev = sub;
      part_prev = L.part_buf;
    }
    H.ws.release(sub_prev);
  } else {
    for (int i = 0; i < n_plan; ++i) {
      const Lvl& L = plan[i];
      // 8 bytes written per split + one 32-byte index lookup per (itemset, antecedent) probe
      ProfScope ps("k_rules_count", s, 40.0 * (double)L.count * (double)((int64_t)1 << (L.k - 1)), L.k);
      KMLS_TRY(kmls_rules_count(a_items, a_counts, sizes, max_k, L.k, L.begin, L.count, nodes, n_total,
                                min_confidence, L.part_buf, rule_cnt + L.out0, missing, s));
      ++launches;
    }
  }
  int64_t* rule_off = nullptr;
  const size_t sbytes = kmls_scan_scratch_bytes(n_x);
  char* scratch = nullptr;
  KMLS_TRY(H.take(&rule_off, n_x + 1, true, "the rule offsets"));
  KMLS_TRY(H.take(&scratch, (int64_t)sbytes, true, "scan scratch"));
  KMLS_TRY(kmls_exclusive_scan_i32(rule_cnt, n_x, rule_off, scratch, sbytes, s));
  k_post2<<<1, 1, 0, s>>>(rule_off + n_x, missing, H.box);
  KMLS_LAUNCHED("k_post2");
  launches += 4;
  KMLS_CUDA(cudaStreamSynchronize(s));
  const int64_t n_rules = H.box[0];
  if (H.box[2] != 0)
    return fail(kArgError, "kmls_rules: a subset of a frequent itemset is missing from the itemset table "
                           "(table not downward closed)");
  R->n_rules = n_rules;
  if (n_rules > max_rules)
    return fail(kCapacityError, "kmls_rules: %lld rules exceed max_rules=%lld (never truncated)", (long long)n_rules,
                (long long)max_rules);
  H.ws.release(scratch);
  H.ws.release(rule_cnt_p1);
  H.ws.release(nodes);
  R->workspace_needed_bytes = H.ws.in_use + (size_t)n_rules * 24 + 4096;
  int32_t* o_rule = nullptr;
  double *o_conf = nullptr, *o_lift = nullptr;
  if (H.take(&o_rule, n_rules * 2, false, "the rule table") || H.take(&o_conf, n_rules, false, "the rule table") ||
      H.take(&o_lift, n_rules, false, "the rule table"))
    return fail(kCapacityError,
                "kmls_rules: workspace too small for the rule table: %lld rules need %zu bytes in total, %zu given "
                "(nothing was truncated)", (long long)n_rules, R->workspace_needed_bytes, workspace_bytes);
  if (n_rules) {
    int64_t lvl_rules[KMLS_MAX_LEVELS + 3] = {0};
    if (profiler().on && profiler().work) {   // rules per level, for the work figure of pass 2
      for (int i = 0; i < n_plan; ++i)
        KMLS_CUDA(cudaMemcpyAsync(&lvl_rules[i], rule_off + plan[i].out0, 8, cudaMemcpyDeviceToHost, s));
      KMLS_CUDA(cudaStreamSynchronize(s));
      lvl_rules[n_plan] = n_rules;
    }
    for (int i = 0; i < n_plan; ++i) {
      const Lvl& L = plan[i];
      // 8 bytes read per split + 24 bytes written per kept rule
      ProfScope ps("k_rules_write", s, 8.0 * (double)L.count * (double)((int64_t)1 << (L.k - 1)) +
                                           24.0 * (double)(lvl_rules[i + 1] - lvl_rules[i]), L.k);
      const int64_t n_sub = (int64_t)1 << (L.k - 1);
      const int32_t *pa = L.part_buf, *pc = L.part_buf ? L.part_buf + L.count * n_sub : nullptr;
      if (full_pass1) {       // the slice's rows inside the whole level's planes
        const int32_t* whole = nullptr;
        for (int f = 0; f < n_fplan; ++f)
          if (fplan[f].k == L.k) whole = fplan[f].part_buf;
        pa = whole + L.begin * n_sub;
        pc = whole + sizes[L.k] * n_sub + L.begin * n_sub;
      }
      KMLS_TRY(kmls_rules_write_planes(a_counts, sizes, max_k, L.k, L.begin, L.count, n_total, min_confidence, pa, pc,
                                       rule_off + L.out0, o_rule, o_conf, o_lift, missing + 1, s));
      ++launches;
    }
    k_post2<<<1, 1, 0, s>>>(nullptr, missing + 1, H.box);
    KMLS_LAUNCHED("k_post2");
    ++launches;
    KMLS_CUDA(cudaStreamSynchronize(s));
    if (H.box[2] != 0)
      return fail(kArgError, "kmls_rules: pass 2 kept a different number of rules than pass 1 sized for %lld "
                             "itemsets (the rule table would be corrupt)", (long long)H.box[2]);
  }
  R->rule_off = off_of(o_rule);
  R->confidence_off = off_of(o_conf);
  R->lift_off = off_of(o_lift);
  R->kernel_launches = launches;
  R->workspace_peak_bytes = H.ws.peak;
  return 0;
}
