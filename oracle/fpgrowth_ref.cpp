// CPU oracle at scale — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// FP-Growth with the reference's semantics, in C++17 + std::thread, so that parity can be
// checked (and a CPU baseline timed) at sizes the Python restatement
// (oracle/mlxtend_restated.py) cannot reach. PARITY UNPINNED UPSTREAM: the algorithm
// restated here is mlxtend 0.23.3's (machine-learning/uv.lock:302-303; not vendored, not
// importable in this image); the anchors are the reference's call site
// machine-learning/main.py:267-272 and the vectors in tests/golden. Only tests/,
// __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this library.
//
// What follows what (mlxtend symbol names):
//   build_root()        fpcommon.setup_fptree   level-1 keep test = integer T1
//                                               (== count/float(N) >= min_support, fp64)
//   Tree::conditional() FPTree.conditional_tree keep test count >= Tk = ceil(min_support*N)
//   Tree::is_path()     FPTree.is_path
//   mine()              fpgrowth.fpg_step       incl. the single-path combination shortcut
// Enumeration order is not part of parity; results are compared as maps
// or through an order-independent checksum.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>
#include <atomic>
#include <chrono>
#include <thread>

namespace {

inline uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

struct Sink {
  bool store = false;
  uint64_t level[64] = {0};
  uint64_t checksum = 0;
  uint64_t n = 0;
  bool overflow = false;  // a single path too long to enumerate was met: result incomplete
  std::vector<int64_t> offsets{0};
  std::vector<int32_t> items;
  std::vector<int64_t> counts;
  void emit(const int32_t* it, int k, int64_t c) {
    level[k < 63 ? k : 63]++;
    uint64_t h = 0;
    for (int i = 0; i < k; ++i) h += mix64((uint64_t)it[i] + 1);
    checksum += mix64(h ^ mix64((uint64_t)c * 0x100000001B3ull + (uint64_t)k));
    n++;
    if (store) {
      size_t b = items.size();
      items.insert(items.end(), it, it + k);
      std::sort(items.begin() + b, items.end());
      offsets.push_back((int64_t)items.size());
      counts.push_back(c);
    }
  }
  void merge(const Sink& o) {
    for (int i = 0; i < 64; ++i) level[i] += o.level[i];
    checksum += o.checksum;
    n += o.n;
    overflow = overflow || o.overflow;
    if (store) {
      int64_t base = (int64_t)items.size();
      items.insert(items.end(), o.items.begin(), o.items.end());
      for (size_t i = 1; i < o.offsets.size(); ++i) offsets.push_back(base + o.offsets[i]);
      counts.insert(counts.end(), o.counts.begin(), o.counts.end());
    }
  }
};

// A weighted path = items in insertion order (most frequent first), as local ids.
struct Paths {
  std::vector<int64_t> off{0};
  std::vector<int32_t> it;
  std::vector<int64_t> w;
  size_t size() const { return w.size(); }
};

struct Tree {
  // node 0 is the root. `label` maps local id -> original item id.
  std::vector<int32_t> item, parent, nchild;
  std::vector<int64_t> count;
  std::vector<int32_t> label;
  std::vector<std::vector<int32_t>> nodes;  // local id -> node list

  void init(size_t n_local) {
    item.assign(1, -1); parent.assign(1, -1); nchild.assign(1, 0); count.assign(1, 0);
    nodes.assign(n_local, {});
  }

  // Insert all paths. Sorting them lexicographically first means a path can only share
  // a prefix with its predecessor, so no child lookup structure is needed.
  void build(const Paths& p) {
    std::vector<int32_t> order(p.size());
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
      return std::lexicographical_compare(p.it.begin() + p.off[a], p.it.begin() + p.off[a + 1],
                                          p.it.begin() + p.off[b], p.it.begin() + p.off[b + 1]);
    });
    std::vector<int32_t> stack;
    const int32_t* prev = nullptr; int64_t prev_len = 0;
    for (int32_t idx : order) {
      const int32_t* cur = p.it.data() + p.off[idx];
      int64_t len = p.off[idx + 1] - p.off[idx], wt = p.w[idx];
      count[0] += wt;
      int64_t lcp = 0;
      while (lcp < len && lcp < prev_len && cur[lcp] == prev[lcp]) ++lcp;
      stack.resize(lcp);
      for (int64_t d = 0; d < lcp; ++d) count[stack[d]] += wt;
      for (int64_t d = lcp; d < len; ++d) {
        int32_t par = d ? stack[d - 1] : 0;
        int32_t id = (int32_t)item.size();
        item.push_back(cur[d]); parent.push_back(par); nchild.push_back(0); count.push_back(wt);
        nchild[par]++;
        nodes[cur[d]].push_back(id);
        stack.push_back(id);
      }
      prev = cur; prev_len = len;
    }
  }

  bool is_path() const {
    if (nchild[0] > 1) return false;
    for (const auto& nl : nodes) {
      if (nl.empty()) continue;
      if (nl.size() > 1 || nchild[nl[0]] > 1) return false;
    }
    return true;
  }

  // Conditional tree of local item x: ancestors' counts accumulated per item, items with
  // count >= tk kept, re-ranked by conditional count (most frequent inserted first).
  void conditional(int32_t x, int64_t tk, Tree& out, std::vector<int64_t>& cnt,
                   std::vector<int32_t>& remap) const {
    cnt.assign(label.size(), 0);
    for (int32_t n : nodes[x])
      for (int32_t a = parent[n]; a > 0; a = parent[a]) cnt[item[a]] += count[n];
    std::vector<int32_t> keep;
    for (size_t i = 0; i < cnt.size(); ++i) if (cnt[i] >= tk && cnt[i] > 0) keep.push_back((int32_t)i);
    std::sort(keep.begin(), keep.end(), [&](int32_t a, int32_t b) {
      return cnt[a] != cnt[b] ? cnt[a] > cnt[b] : a < b; });
    remap.assign(label.size(), -1);
    out.label.resize(keep.size());
    for (size_t i = 0; i < keep.size(); ++i) { remap[keep[i]] = (int32_t)i; out.label[i] = label[keep[i]]; }
    out.init(keep.size());
    Paths p;
    for (int32_t n : nodes[x]) {
      size_t b = p.it.size();
      for (int32_t a = parent[n]; a > 0; a = parent[a])
        if (remap[item[a]] >= 0) p.it.push_back(remap[item[a]]);
      std::sort(p.it.begin() + b, p.it.end());
      p.off.push_back((int64_t)p.it.size());
      p.w.push_back(count[n]);
    }
    out.build(p);
  }
};

void mine(const Tree& t, std::vector<int32_t>& suffix, int64_t tk, int max_len, Sink& sink) {
  size_t depth = suffix.size();
  if (t.is_path()) {
    // fpg_step's shortcut: every non-empty combination of the path's items, support =
    // min node count. (At the root this uses no Tk test — mlxtend's own edge)
    std::vector<int32_t> its; std::vector<int64_t> cs;
    for (size_t i = 0; i < t.nodes.size(); ++i)
      if (!t.nodes[i].empty()) { its.push_back(t.label[i]); cs.push_back(t.count[t.nodes[i][0]]); }
    int m = (int)its.size();
    int lim = max_len > 0 ? max_len - (int)depth : m;
    if (m > 40) { sink.overflow = true; return; }  // 2^40 combinations: reported, never silent
    for (uint64_t mask = 1; mask < (1ull << m); ++mask) {
      if (__builtin_popcountll(mask) > lim) continue;
      int64_t c = INT64_MAX;
      for (int i = 0; i < m; ++i) if (mask >> i & 1) { suffix.push_back(its[i]); c = std::min(c, cs[i]); }
      sink.emit(suffix.data(), (int)suffix.size(), c);
      suffix.resize(depth);
    }
    return;
  }
  if (max_len > 0 && (int)depth >= max_len) return;
  std::vector<int64_t> cnt; std::vector<int32_t> remap;
  for (size_t x = 0; x < t.nodes.size(); ++x) {
    if (t.nodes[x].empty()) continue;
    int64_t c = 0;
    for (int32_t n : t.nodes[x]) c += t.count[n];
    suffix.push_back(t.label[x]);
    sink.emit(suffix.data(), (int)suffix.size(), c);
    Tree ct;
    t.conditional((int32_t)x, tk, ct, cnt, remap);
    mine(ct, suffix, tk, max_len, sink);
    suffix.pop_back();
  }
}

struct Result { Sink sink; };

}  // namespace

extern "C" {

// Mines CSR transactions (duplicates inside a transaction collapse, as
// TransactionEncoder.transform does). t1/tk are the integer thresholds
// (oracle/mlxtend_restated.py: thresholds()). Returns an opaque handle.
void* oracle_fpgrowth_run(const int64_t* offsets, const int32_t* items, int64_t n_tx,
                          int32_t n_items, int64_t t1, int64_t tk, int max_len, int store,
                          int n_threads) {
  auto* res = new Result();
  res->sink.store = store != 0;
  if (n_tx == 0 || n_items == 0) return res;
  std::vector<int64_t> c1(n_items, 0);
  {
    std::vector<int32_t> tmp;
    for (int64_t r = 0; r < n_tx; ++r) {
      tmp.assign(items + offsets[r], items + offsets[r + 1]);
      std::sort(tmp.begin(), tmp.end());
      tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
      for (int32_t i : tmp) c1[i]++;
    }
  }
  std::vector<int32_t> keep;
  for (int32_t i = 0; i < n_items; ++i) if (c1[i] >= t1 && c1[i] > 0) keep.push_back(i);
  std::sort(keep.begin(), keep.end(), [&](int32_t a, int32_t b) {
    return c1[a] != c1[b] ? c1[a] > c1[b] : a < b; });
  std::vector<int32_t> remap(n_items, -1);
  Tree root;
  root.label.resize(keep.size());
  for (size_t i = 0; i < keep.size(); ++i) { remap[keep[i]] = (int32_t)i; root.label[i] = keep[i]; }
  root.init(keep.size());
  {
    Paths p;
    for (int64_t r = 0; r < n_tx; ++r) {
      size_t b = p.it.size();
      for (int64_t j = offsets[r]; j < offsets[r + 1]; ++j)
        if (remap[items[j]] >= 0) p.it.push_back(remap[items[j]]);
      std::sort(p.it.begin() + b, p.it.end());
      p.it.erase(std::unique(p.it.begin() + b, p.it.end()), p.it.end());
      p.off.push_back((int64_t)p.it.size());
      p.w.push_back(1);
    }
    root.build(p);
  }
  if (root.is_path() || keep.size() < 2) {
    std::vector<int32_t> suffix;
    mine(root, suffix, tk, max_len, res->sink);
    return res;
  }
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  std::vector<Sink> sinks(nt);
  for (auto& s : sinks) s.store = res->sink.store;
  int64_t nx = (int64_t)keep.size();
  // One task per first-level item (its whole conditional subtree), handed out dynamically.
  std::atomic<int64_t> next{0};
  auto worker = [&](int tid) {
    Sink& sink = sinks[tid];
    for (int64_t x = next.fetch_add(1); x < nx; x = next.fetch_add(1)) {
      if (root.nodes[x].empty()) continue;
      int64_t c = 0;
      for (int32_t n : root.nodes[x]) c += root.count[n];
      std::vector<int32_t> suffix{root.label[x]};
      sink.emit(suffix.data(), 1, c);
      if (max_len == 1) continue;
      Tree ct; std::vector<int64_t> cnt; std::vector<int32_t> rm;
      root.conditional((int32_t)x, tk, ct, cnt, rm);
      mine(ct, suffix, tk, max_len, sink);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nt; ++t) pool.emplace_back(worker, t);
  worker(0);
  for (auto& th : pool) th.join();
  for (auto& s : sinks) res->sink.merge(s);
  return res;
}

uint64_t oracle_fpgrowth_num_itemsets(void* h) { return ((Result*)h)->sink.n; }
uint64_t oracle_fpgrowth_checksum(void* h) { return ((Result*)h)->sink.checksum; }
int oracle_fpgrowth_overflow(void* h) { return ((Result*)h)->sink.overflow ? 1 : 0; }
void oracle_fpgrowth_level_counts(void* h, uint64_t* out64) {
  std::memcpy(out64, ((Result*)h)->sink.level, sizeof(uint64_t) * 64);
}
int64_t oracle_fpgrowth_total_items(void* h) { return (int64_t)((Result*)h)->sink.items.size(); }
// offsets[n+1], items[total], counts[n]; items of an itemset ascending by id.
void oracle_fpgrowth_copy(void* h, int64_t* offsets, int32_t* items, int64_t* counts) {
  auto& s = ((Result*)h)->sink;
  std::memcpy(offsets, s.offsets.data(), s.offsets.size() * sizeof(int64_t));
  if (!s.items.empty()) std::memcpy(items, s.items.data(), s.items.size() * sizeof(int32_t));
  if (!s.counts.empty()) std::memcpy(counts, s.counts.data(), s.counts.size() * sizeof(int64_t));
}
// mlxtend.frequent_patterns.association_rules(metric="confidence", min_threshold=c) over
// the stored itemsets (run with store=1): for every itemset X, |X| >= 2, and every non-empty
// proper subset A: confidence = (cX/N)/(cA/N), lift = confidence/(cC/N), kept iff
// confidence >= c. Returns the number of rules and an order-independent checksum over
// (antecedent, consequent, support, confidence, lift) bit patterns; -1 if a subset is missing.
// `stride` > 1 generates the rules of every stride-th block of itemsets only (a bounded
// sample for timing: runs of 65,536 consecutive itemsets, so that the sample keeps the
// cache locality a full pass has and still averages over the whole table);
// stride 1 = all.
static std::atomic<int64_t> g_last_splits{0};
static double g_last_index_seconds = 0.0;
// seconds the last oracle_rules* call spent building its (serial) subset index
double oracle_rules_last_index_seconds() { return g_last_index_seconds; }
// number of (itemset, antecedent) splits the last oracle_rules* call evaluated
int64_t oracle_rules_last_splits() { return g_last_splits.load(); }

int64_t oracle_rules_strided(void* h, int64_t n_total, double min_conf, int n_threads, int stride,
                             uint64_t* checksum) {
  auto& s = ((Result*)h)->sink;
  const int64_t n = (int64_t)s.counts.size();
  uint64_t cap = 1024;
  while (cap < 2 * (uint64_t)n) cap <<= 1;
  const auto t_index0 = std::chrono::steady_clock::now();
  std::vector<uint64_t> keys(cap, 0);
  std::vector<int32_t> vals(cap, -1);
  auto key_of = [&](const int32_t* it, int k) {
    uint64_t hsum = 0;
    for (int i = 0; i < k; ++i) hsum += mix64((uint64_t)it[i] + 1);
    return mix64(hsum ^ ((uint64_t)k << 56)) | 1ull;
  };
  for (int64_t i = 0; i < n; ++i) {
    int k = (int)(s.offsets[i + 1] - s.offsets[i]);
    uint64_t key = key_of(s.items.data() + s.offsets[i], k), hh = key & (cap - 1);
    while (keys[hh]) hh = (hh + 1) & (cap - 1);
    keys[hh] = key; vals[hh] = (int32_t)i;
  }
  g_last_index_seconds =
      std::chrono::duration<double>(std::chrono::steady_clock::now() - t_index0).count();
  auto lookup = [&](const int32_t* it, int k) -> int64_t {
    uint64_t key = key_of(it, k), hh = key & (cap - 1);
    while (keys[hh]) {
      if (keys[hh] == key) {
        int64_t j = vals[hh];
        if (s.offsets[j + 1] - s.offsets[j] == k && std::equal(it, it + k, s.items.data() + s.offsets[j]))
          return s.counts[j];
      }
      hh = (hh + 1) & (cap - 1);
    }
    return -1;
  };
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  std::vector<uint64_t> sums(nt, 0);
  std::vector<int64_t> cnts(nt, 0), splits(nt, 0);
  std::atomic<int64_t> next{0};
  std::atomic<int> missing{0};
  auto bits = [](double d) { uint64_t u; std::memcpy(&u, &d, 8); return u; };
  auto worker = [&](int tid) {
    int32_t a[64], c[64];
    const int64_t step = stride > 1 ? 256 : 4096;
    for (int64_t b = next.fetch_add(step); b < n; b = next.fetch_add(step)) {
      if (stride > 1 && ((b / step) >> 8) % stride != 0) continue;  // runs of 256 blocks (65,536 itemsets)
      for (int64_t i = b; i < std::min(n, b + step); ++i) {
        const int k = (int)(s.offsets[i + 1] - s.offsets[i]);
        if (k < 2 || k > 30) continue;
        const int32_t* x = s.items.data() + s.offsets[i];
        const double sx = (double)s.counts[i] / (double)n_total;
        splits[tid] += ((int64_t)1 << k) - 2;
        for (uint32_t m = 1; m + 1 < (1u << k); ++m) {
          int na = 0, nc = 0;
          uint64_t ha = 0, hc = 0;
          for (int t = 0; t < k; ++t) {
            if (m >> t & 1) { a[na++] = x[t]; ha += mix64((uint64_t)x[t] + 1); }
            else { c[nc++] = x[t]; hc += mix64((uint64_t)x[t] + 1); }
          }
          const int64_t ca = lookup(a, na);
          if (ca < 0) { missing = 1; continue; }
          const double sa = (double)ca / (double)n_total;
          const double conf = sx / sa;
          if (!(conf >= min_conf)) continue;
          const int64_t cc = lookup(c, nc);
          if (cc < 0) { missing = 1; continue; }
          const double lift = conf / ((double)cc / (double)n_total);
          cnts[tid]++;
          sums[tid] += mix64(ha ^ mix64(hc + 0x1234) ^ mix64(bits(conf)) ^ mix64(bits(lift) + 1) ^
                             mix64(bits(sx) + 2));
        }
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nt; ++t) pool.emplace_back(worker, t);
  worker(0);
  for (auto& th : pool) th.join();
  if (missing) return -1;
  int64_t total = 0; uint64_t cs = 0;
  int64_t sp = 0;
  for (int t = 0; t < nt; ++t) { total += cnts[t]; cs += sums[t]; sp += splits[t]; }
  g_last_splits = sp;
  *checksum = cs;
  return total;
}

int64_t oracle_rules(void* h, int64_t n_total, double min_conf, int n_threads, uint64_t* checksum) {
  return oracle_rules_strided(h, n_total, min_conf, n_threads, 1, checksum);
}

void oracle_fpgrowth_free(void* h) { delete (Result*)h; }

}  // extern "C"
