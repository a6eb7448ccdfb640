// Levels k >= 3 of the hot path: candidate generation, Apriori pruning, support counting.
//
// Replaces the recursion of mlxtend's fpgrowth.fpg_step / FPTree.conditional_tree (reached
// from reference machine-learning/main.py:272). The result set is the order-free one of
// L_k = {X subset of L1, |X| = k : c(X) >= Tk}; it is produced
// level by level (equivalence-class join of L_k with itself, prune, count, filter).
//
// A level table is `items[n][k]` (int32 row indices of the vertical bitmap, ascending
// inside an itemset, rows sorted lexicographically) plus `counts[n]`.
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

// ---------------------------------------------------------------------------------------
// Scans / selections (CUB, scratch supplied by the caller).
// ---------------------------------------------------------------------------------------
struct GeThreshold {
  int32_t thr;
  __host__ __device__ int64_t operator()(int32_t c) const { return c >= thr ? 1 : 0; }
};
struct ToI64 {
  __host__ __device__ int64_t operator()(int32_t c) const { return (int64_t)c; }
};
struct NonZeroAt {
  const int32_t* v;
  __host__ __device__ bool operator()(int32_t i) const { return v[i] > 0; }
};

template <class Op>
__global__ void k_scan_total(const int32_t* in, int64_t n, int64_t* out, Op op) {
  out[n] = n > 0 ? out[n - 1] + op(in[n - 1]) : 0;
}

template <class Op>
int scan_i32(const int32_t* d_in, int64_t n, int64_t* d_out, void* d_scratch,
             size_t scratch_bytes, Op op, cudaStream_t s) {
  if (n > 0) {
    KMLS_REQUIRE(n <= 0x7fffffff, "scan: more than 2^31-1 elements");
    cub::TransformInputIterator<int64_t, Op, const int32_t*> it(d_in, op);
    size_t need = 0;
    KMLS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, it, d_out, (int)n, s));
    if (need > scratch_bytes)
      return fail(kCapacityError, "scan: scratch too small (%zu > %zu bytes)", need, scratch_bytes);
    KMLS_CUDA(cub::DeviceScan::ExclusiveSum(d_scratch, need, it, d_out, (int)n, s));
  }
  k_scan_total<<<1, 1, 0, s>>>(d_in, n, d_out, op);
  KMLS_LAUNCHED("k_scan_total");
  return 0;
}

// ---------------------------------------------------------------------------------------
// Fingerprint set of a level table (open addressing, 64-bit keys, 0 = empty). A false
// "present" can only keep a candidate that then gets counted anyway, so fingerprints alone
// are safe for pruning.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t table_key(const int32_t* it, int k) {
  uint64_t s = 0;
  for (int t = 0; t < k; ++t) s += item_term(it[t]);
  return finish_key(s, k);
}

__global__ void __launch_bounds__(256) k_set_build(const int32_t* __restrict__ items, int64_t n,
                                                   int k, unsigned long long* __restrict__ slots,
                                                   uint64_t mask) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t key = table_key(items + i * k, k);
    uint64_t h = (key >> 20) & mask;
    while (true) {
      unsigned long long old = atomicCAS(slots + h, 0ull, (unsigned long long)key);
      if (old == 0ull || old == key) break;
      h = (h + 1) & mask;
    }
  }
}

__device__ __forceinline__ bool set_contains(const unsigned long long* __restrict__ slots,
                                             uint64_t mask, uint64_t key) {
  uint64_t h = (key >> 20) & mask;
  while (true) {
    const unsigned long long v = __ldg(slots + h);
    if (v == key) return true;
    if (v == 0ull) return false;
    h = (h + 1) & mask;
  }
}

// ---------------------------------------------------------------------------------------
// K3: join + prune. One warp per parent itemset p; the partners q are the later members of
// p's equivalence class (same first k-1 items). Candidate = items(p) + last(q). It survives
// if every k-subset obtained by dropping one of items(p)[0..k-2] is frequent (the two
// remaining k-subsets are p and q themselves). For k == 2 that is one lookup in the pair
// count matrix; for k >= 3 it is k-1 lookups in the fingerprint set of level k.
// Pass 1 (write == 0) counts survivors per p and remembers the first 32 decisions;
// pass 2 (write == 1) writes (p, q) at kept_off[p] + rank, in ascending q.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_join(const int32_t* __restrict__ items, int64_t n, int k,
       const unsigned long long* __restrict__ slots, uint64_t mask,
       const int32_t* __restrict__ pairC, int32_t n_rows, int32_t thr,
       const int32_t* __restrict__ par, int write,
       int32_t* __restrict__ kept_cnt, uint32_t* __restrict__ kept_mask,
       const int64_t* __restrict__ kept_off, int32_t* __restrict__ cand_p,
       int32_t* __restrict__ cand_q) {
  const int lane = threadIdx.x & 31;
  for (int64_t p = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); p < n;
       p += (int64_t)gridDim.x * 8) {
    const int32_t* ip = items + p * k;
    uint64_t sum_p = 0;
    if (k >= 3 && slots != nullptr)
      for (int t = 0; t < k; ++t) sum_p += item_term(__ldg(ip + t));
    int total = 0;
    int64_t pos = write ? kept_off[p] : 0;
    bool first = true;
    for (int64_t q0 = p + 1; q0 < n; q0 += 32) {
      const int64_t q = q0 + lane;
      bool in_class = q < n;
      if (in_class) {
        if (par != nullptr) {             // same parent <=> same first k-1 members
          in_class = __ldg(par + q) == __ldg(par + p);
        } else {
          const int32_t* iq = items + q * k;
          for (int t = 0; t < k - 1; ++t) in_class = in_class && (__ldg(iq + t) == __ldg(ip + t));
        }
      }
      const uint32_t cm = __ballot_sync(0xffffffffu, in_class);
      bool keep = false;
      if (write && first) {
        keep = (kept_mask[p] >> lane) & 1u;
      } else if (in_class) {
        const int32_t y = __ldg(items + q * k + (k - 1));
        if (k == 2) {
          keep = pairC == nullptr || __ldg(pairC + (int64_t)__ldg(ip + 1) * n_rows + y) >= thr;
        } else if (slots == nullptr) {
          keep = true;  // pruning switched off: every join survives and gets counted
        } else {
          const uint64_t sum = sum_p + item_term(y);
          keep = true;
          for (int t = 0; t < k - 1 && keep; ++t)
            keep = set_contains(slots, mask, finish_key(sum - item_term(__ldg(ip + t)), k));
        }
      }
      const uint32_t km = __ballot_sync(0xffffffffu, keep);
      if (!write) {
        if (first && lane == 0) kept_mask[p] = km;
        total += __popc(km);
      } else {
        if (keep) {
          const int64_t o = pos + __popc(km & ((1u << lane) - 1));
          cand_p[o] = (int32_t)p;
          cand_q[o] = (int32_t)q;
        }
        pos += __popc(km);
      }
      first = false;
      if (cm != 0xffffffffu) break;  // class members are contiguous: a miss ends the class
    }
    if (!write && lane == 0) kept_cnt[p] = total;
  }
}

// ---------------------------------------------------------------------------------------
// K4: support counting of the candidates of one level. Work unit = (parent p, chunk of
// kCountChunkWords bitmap words). The CTA stages the parent's bitset chunk - the AND of
// the k rows of items(p) - in shared memory; then each warp takes one candidate of p at a
// time, ANDs the staged chunk with the extension row's chunk (128-bit loads), __popc,
// warp-shuffle reduce. Units are ordered chunk-major so that CTAs running together touch
// the same column range of the bitmap (L2/L1 reuse of the popular rows).
// Algorithmic bytes per level: (C_k + G_k * k) * row_bytes  (C_k candidates, G_k parents).
// ---------------------------------------------------------------------------------------
constexpr int kCountChunkWords = 2048;  // 8 KB of shared memory

__global__ void __launch_bounds__(256)
k_count_candidates(const uint32_t* __restrict__ bitmap, int64_t row_words,
                   const int32_t* __restrict__ items, int k, const int32_t* __restrict__ grp_p,
                   int64_t n_groups, int64_t n_chunks, const int64_t* __restrict__ kept_off,
                   const int32_t* __restrict__ cand_q, int use_atomics,
                   int32_t* __restrict__ cand_cnt) {
  __shared__ uint4 S[kCountChunkWords / 4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t units = n_groups * n_chunks;
  for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
    const int64_t ch = u / n_groups, g = u - ch * n_groups;
    const int32_t p = __ldg(grp_p + g);
    const int64_t w0 = ch * kCountChunkWords;
    const int nv = (int)(min((int64_t)kCountChunkWords, row_words - w0) >> 2);
    const int32_t* ip = items + (int64_t)p * k;
    for (int v = threadIdx.x; v < nv; v += 256) {
      uint4 acc = ldg4(reinterpret_cast<const uint4*>(bitmap + (int64_t)__ldg(ip) * row_words + w0) + v);
      for (int t = 1; t < k; ++t)
        acc = and4(acc, ldg4(reinterpret_cast<const uint4*>(bitmap + (int64_t)__ldg(ip + t) * row_words + w0) + v));
      S[v] = acc;
    }
    __syncthreads();
    const int64_t c0 = kept_off[p], c1 = kept_off[p + 1];
    for (int64_t c = c0 + warp; c < c1; c += 8) {
      const int32_t y = __ldg(items + (int64_t)__ldg(cand_q + c) * k + (k - 1));
      const uint4* row = reinterpret_cast<const uint4*>(bitmap + (int64_t)y * row_words + w0);
      int acc = 0;
      for (int v = lane; v < nv; v += 32) acc += popc4(and4(S[v], ldg4(row + v)));
      acc = warp_sum(acc);
      if (lane == 0) {
        if (use_atomics) {
          if (acc) atomicAdd(cand_cnt + c, acc);
        } else {
          cand_cnt[c] = acc;
        }
      }
    }
    __syncthreads();
  }
}

// Filter + emit: candidates with count >= thr become level k+1, order preserved.
__global__ void __launch_bounds__(256)
k_emit_level(const int32_t* __restrict__ items, int k, const int32_t* __restrict__ cand_p,
             const int32_t* __restrict__ cand_q, const int32_t* __restrict__ cand_cnt,
             int64_t n_cand, int32_t thr, const int64_t* __restrict__ out_off,
             int32_t* __restrict__ out_items, int32_t* __restrict__ out_counts) {
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_cand;
       c += (int64_t)gridDim.x * blockDim.x) {
    const int32_t cnt = cand_cnt[c];
    if (cnt < thr) continue;
    const int64_t o = out_off[c];
    const int32_t* ip = items + (int64_t)cand_p[c] * k;
    int32_t* op = out_items + o * (k + 1);
    for (int t = 0; t < k; ++t) op[t] = __ldg(ip + t);
    op[k] = __ldg(items + (int64_t)cand_q[c] * k + (k - 1));
    out_counts[o] = cnt;
  }
}

}  // namespace kmls

using namespace kmls;

extern "C" size_t kmls_scan_scratch_bytes(int64_t n) {
  size_t a = 0, b = 0;
  if (n < 1) n = 1;
  if (n > 0x7fffffff) n = 0x7fffffff;
  cub::TransformInputIterator<int64_t, ToI64, const int32_t*> it(nullptr, ToI64());
  cub::DeviceScan::ExclusiveSum(nullptr, a, it, (int64_t*)nullptr, (int)n);
  cub::CountingInputIterator<int32_t> idx(0);
  cub::DeviceSelect::If(nullptr, b, idx, (int32_t*)nullptr, (int64_t*)nullptr, (int)n,
                        NonZeroAt{nullptr});
  return std::max(a, b) + 256;
}

extern "C" int kmls_exclusive_scan_i32(const int32_t* d_in, int64_t n, int64_t* d_out,
                                       void* d_scratch, size_t scratch_bytes,
                                       kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && d_out, "kmls_exclusive_scan_i32: bad arguments");
  KMLS_REQUIRE(n == 0 || d_in, "kmls_exclusive_scan_i32: d_in is null");
  return scan_i32(d_in, n, d_out, d_scratch, scratch_bytes, ToI64(), (cudaStream_t)stream);
}

extern "C" int64_t kmls_set_capacity(int64_t n) {
  int64_t cap = 1024;
  while (cap < 2 * n) cap <<= 1;
  return cap;
}

extern "C" int kmls_itemset_set_build(const int32_t* d_items, int64_t n, int32_t k,
                                      uint64_t* d_slots, int64_t capacity, kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && k >= 1 && capacity >= 2 * n && (capacity & (capacity - 1)) == 0,
               "kmls_itemset_set_build: capacity must be a power of two >= 2n");
  cudaStream_t s = (cudaStream_t)stream;
  KMLS_CUDA(cudaMemsetAsync(d_slots, 0, sizeof(uint64_t) * (size_t)capacity, s));
  if (n == 0) return 0;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_set_build<<<grid, 256, 0, s>>>(d_items, n, k, (unsigned long long*)d_slots, (uint64_t)capacity - 1);
  KMLS_LAUNCHED("k_set_build");
  return 0;
}

// Join without pruning when the parent column is known: the candidates of p are simply its
// later siblings, so a thread per itemset is enough - pass 1 counts them (a forward scan of
// the parent column), pass 2 writes (p, p+1), (p, p+2), ...
__global__ void __launch_bounds__(256)
k_join_siblings(const int32_t* __restrict__ par, int64_t n, int write, int32_t* __restrict__ kept_cnt,
                const int64_t* __restrict__ kept_off, int32_t* __restrict__ cand_p,
                int32_t* __restrict__ cand_q) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n;
       p += (int64_t)gridDim.x * blockDim.x) {
    if (!write) {
      const int32_t mine = __ldg(par + p);
      int64_t e = p + 1;
      while (e < n && __ldg(par + e) == mine) ++e;
      kept_cnt[p] = (int32_t)(e - p - 1);
    } else {
      const int64_t o = kept_off[p];
      const int32_t m = kept_cnt[p];
      for (int32_t i = 0; i < m; ++i) {
        cand_p[o + i] = (int32_t)p;
        cand_q[o + i] = (int32_t)(p + 1 + i);
      }
    }
  }
}

static int launch_join(const int32_t* d_items, int64_t n, int32_t k, const uint64_t* d_slots,
                       int64_t capacity, const int32_t* d_pair, int32_t n_rows, int32_t thr,
                       const int32_t* d_par, int write, int32_t* d_kept_cnt, uint32_t* d_kept_mask,
                       const int64_t* d_kept_off, int32_t* d_cand_p, int32_t* d_cand_q,
                       cudaStream_t s) {
  KMLS_REQUIRE(n >= 0 && n <= 0x7fffffff && k >= 2 && k <= 30, "kmls_join: bad level size");
  // a null d_pair (k == 2) or null d_slots (k >= 3) switches Apriori pruning off
  if (k == 2 && d_pair) KMLS_REQUIRE(n_rows > 0, "kmls_join: bad pair-count matrix size");
  if (k >= 3 && d_slots) KMLS_REQUIRE(capacity > 0 && (capacity & (capacity - 1)) == 0,
                                      "kmls_join: fingerprint set capacity must be a power of two");
  if (!d_slots) capacity = 1;
  if (n == 0) return 0;
  if (d_par && ((k == 2 && !d_pair) || (k >= 3 && !d_slots))) {
    int g = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 32);
    k_join_siblings<<<g, 256, 0, s>>>(d_par, n, write, d_kept_cnt, d_kept_off, d_cand_p, d_cand_q);
    KMLS_LAUNCHED("k_join_siblings");
    return 0;
  }
  int grid = (int)std::min<int64_t>((n + 7) / 8, (int64_t)num_sms() * 32);
  k_join<<<grid, 256, 0, s>>>(d_items, n, k, (const unsigned long long*)d_slots,
                              (uint64_t)capacity - 1, d_pair, n_rows, thr, d_par, write, d_kept_cnt,
                              d_kept_mask, d_kept_off, d_cand_p, d_cand_q);
  KMLS_LAUNCHED("k_join");
  return 0;
}

extern "C" int kmls_join_count(const int32_t* d_items, int64_t n, int32_t k,
                               const uint64_t* d_slots, int64_t capacity, const int32_t* d_pair,
                               int32_t n_rows, int32_t thr, const int32_t* d_par,
                               int32_t* d_kept_cnt, uint32_t* d_kept_mask, int64_t* d_kept_off,
                               void* d_scratch, size_t scratch_bytes, kmls_stream_t stream) {
  cudaStream_t s = (cudaStream_t)stream;
  KMLS_REQUIRE(d_kept_off, "kmls_join_count: null pointer");
  int rc = launch_join(d_items, n, k, d_slots, capacity, d_pair, n_rows, thr, d_par, 0, d_kept_cnt,
                       d_kept_mask, nullptr, nullptr, nullptr, s);
  if (rc) return rc;
  return scan_i32(d_kept_cnt, n, d_kept_off, d_scratch, scratch_bytes, ToI64(), s);
}

extern "C" int kmls_join_write(const int32_t* d_items, int64_t n, int32_t k,
                               const uint64_t* d_slots, int64_t capacity, const int32_t* d_pair,
                               int32_t n_rows, int32_t thr, const int32_t* d_par,
                               const int32_t* d_kept_cnt,
                               const uint32_t* d_kept_mask, const int64_t* d_kept_off,
                               int32_t* d_cand_p, int32_t* d_cand_q, int32_t* d_grp_p,
                               int64_t* d_n_groups, void* d_scratch, size_t scratch_bytes,
                               kmls_stream_t stream) {
  cudaStream_t s = (cudaStream_t)stream;
  int rc = launch_join(d_items, n, k, d_slots, capacity, d_pair, n_rows, thr, d_par, 1,
                       const_cast<int32_t*>(d_kept_cnt), const_cast<uint32_t*>(d_kept_mask),
                       d_kept_off, d_cand_p, d_cand_q, s);
  if (rc) return rc;
  if (d_grp_p == nullptr) return 0;  // caller does not need the parent list
  if (n == 0) {
    KMLS_CUDA(cudaMemsetAsync(d_n_groups, 0, sizeof(int64_t), s));
    return 0;
  }
  cub::CountingInputIterator<int32_t> idx(0);
  size_t need = 0;
  KMLS_CUDA(cub::DeviceSelect::If(nullptr, need, idx, d_grp_p, d_n_groups, (int)n,
                                  NonZeroAt{d_kept_cnt}, s));
  if (need > scratch_bytes)
    return fail(kCapacityError, "kmls_join_write: scratch too small (%zu > %zu bytes)", need, scratch_bytes);
  KMLS_CUDA(cub::DeviceSelect::If(d_scratch, need, idx, d_grp_p, d_n_groups, (int)n,
                                  NonZeroAt{d_kept_cnt}, s));
  return 0;
}

extern "C" int kmls_count_candidates(const uint32_t* d_bitmap, int64_t row_words,
                                     const int32_t* d_items, int32_t k, const int32_t* d_grp_p,
                                     int64_t n_groups, const int64_t* d_kept_off,
                                     const int32_t* d_cand_q, int64_t n_cand, int32_t* d_cand_cnt,
                                     kmls_stream_t stream) {
  KMLS_REQUIRE(n_groups >= 0 && n_cand >= 0 && k >= 2 && (row_words & 3) == 0 && row_words > 0,
               "kmls_count_candidates: bad size");
  if (n_cand == 0 || n_groups == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t n_chunks = (row_words + kCountChunkWords - 1) / kCountChunkWords;
  if (n_chunks > 1) KMLS_CUDA(cudaMemsetAsync(d_cand_cnt, 0, sizeof(int32_t) * (size_t)n_cand, s));
  const int64_t units = n_groups * n_chunks;
  int grid = (int)std::min<int64_t>(units, (int64_t)num_sms() * 64);
  k_count_candidates<<<grid, 256, 0, s>>>(d_bitmap, row_words, d_items, k, d_grp_p, n_groups,
                                          n_chunks, d_kept_off, d_cand_q, n_chunks > 1 ? 1 : 0,
                                          d_cand_cnt);
  KMLS_LAUNCHED("k_count_candidates");
  return 0;
}

extern "C" int kmls_emit_count(const int32_t* d_cand_cnt, int64_t n_cand, int32_t thr,
                               int64_t* d_out_off, void* d_scratch, size_t scratch_bytes,
                               kmls_stream_t stream) {
  KMLS_REQUIRE(n_cand >= 0 && d_out_off, "kmls_emit_count: bad arguments");
  return scan_i32(d_cand_cnt, n_cand, d_out_off, d_scratch, scratch_bytes, GeThreshold{thr},
                  (cudaStream_t)stream);
}

extern "C" int kmls_emit_level(const int32_t* d_items, int32_t k, const int32_t* d_cand_p,
                               const int32_t* d_cand_q, const int32_t* d_cand_cnt, int64_t n_cand,
                               int32_t thr, const int64_t* d_out_off, int32_t* d_out_items,
                               int32_t* d_out_counts, kmls_stream_t stream) {
  KMLS_REQUIRE(n_cand >= 0 && k >= 2, "kmls_emit_level: bad arguments");
  if (n_cand == 0) return 0;
  KMLS_REQUIRE(d_out_off && d_out_items && d_out_counts, "kmls_emit_level: null pointer");
  int grid = (int)std::min<int64_t>((n_cand + 255) / 256, (int64_t)num_sms() * 16);
  k_emit_level<<<grid, 256, 0, (cudaStream_t)stream>>>(d_items, k, d_cand_p, d_cand_q, d_cand_cnt,
                                                     n_cand, thr, d_out_off, d_out_items,
                                                     d_out_counts);
  KMLS_LAUNCHED("k_emit_level");
  return 0;
}

// ---------------------------------------------------------------------------------------
// Order-independent checksum of a level table, for comparing a whole result with another
// implementation's without moving it (same arithmetic as the CPU oracle's Sink::emit and
// oracle/fpgrowth_c.itemset_checksum): sum over itemsets of
//   mix( sum_i mix(track id_i + 1)  ^  mix(count * 0x100000001B3 + k) )   (mod 2^64).
// ---------------------------------------------------------------------------------------
namespace kmls {
__global__ void __launch_bounds__(256)
k_itemset_checksum(const int32_t* __restrict__ items, const int32_t* __restrict__ counts, int64_t n,
                   int k, const int32_t* __restrict__ row_items, unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t h = 0;
    for (int t = 0; t < k; ++t) h += mix64((uint64_t)__ldg(row_items + __ldg(items + i * k + t)) + 1);
    acc += mix64(h ^ mix64((uint64_t)__ldg(counts + i) * 0x100000001B3ull + (uint64_t)k));
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}
}  // namespace kmls

extern "C" int kmls_itemset_checksum(const int32_t* d_items, const int32_t* d_counts, int64_t n,
                                     int32_t k, const int32_t* d_row_items, uint64_t* d_out,
                                     kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && k >= 1 && d_out, "kmls_itemset_checksum: bad arguments");
  if (n == 0) return 0;
  KMLS_REQUIRE(d_items && d_counts && d_row_items, "kmls_itemset_checksum: null pointer");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  k_itemset_checksum<<<grid, 256, 0, (cudaStream_t)stream>>>(d_items, d_counts, n, k, d_row_items,
                                                            (unsigned long long*)d_out);
  KMLS_LAUNCHED("k_itemset_checksum");
  return 0;
}

// ---------------------------------------------------------------------------------------
// kmls_exclusive_scan_i32_noalloc: the 64-bit exclusive scan above for callers that have no
// way to hand over a scratch buffer.
//
// Who that is: code that only ever holds the input and output buffers somebody else gave
// it. Typical cases are a plugin running inside another program's CUDA context whose device
// memory was budgeted and carved up at start-up, a service loop that must not call
// cudaMalloc/cudaFree because those calls serialise the whole device and would stall its
// other streams, and a binding from a language that has a pointer to data but no device
// allocator behind it. kmls_exclusive_scan_i32 needs kmls_scan_scratch_bytes(n) bytes of
// scratch from such a caller; this variant brings its own.
//
// The scratch is the file-scope array below: it is part of the library image, so every GPU
// that loads the library sets aside kNoAllocScratchBytes (1 MiB) of device memory until the
// process exits, whether or not this entry point is ever called. The first call on a
// device looks up the array's device address (cudaGetSymbolAddress - the array's name in
// host code is not a device pointer) and caches it for that device; make that first call
// outside any stream capture. The cached address does not survive cudaDeviceReset(): a
// process that resets a device must not call this entry point on it afterwards. Nothing is
// allocated or freed, here or in CUB.
//
// Limits and sharing: n <= KMLS_NOALLOC_SCAN_MAX_N (4,194,304) - larger inputs are refused
// with an error, never truncated. All calls on a device use the same array. It keeps
// whatever the previous call left in it; that is harmless because the scan's own init
// kernel resets every tile state it reads before the scan kernel runs. Two calls that
// overlap on the GPU (different streams, or different host threads) would race on the
// array: serialise them on one stream.
// ---------------------------------------------------------------------------------------
namespace kmls {
constexpr size_t kNoAllocScratchBytes = 1u << 20;
static __device__ __align__(256) unsigned char g_noalloc_scan_scratch[kNoAllocScratchBytes];
}  // namespace kmls

extern "C" int kmls_exclusive_scan_i32_noalloc(const int32_t* d_in, int64_t n, int64_t* d_out,
                                               kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && d_out, "kmls_exclusive_scan_i32_noalloc: bad arguments");
  if (n > KMLS_NOALLOC_SCAN_MAX_N)
    return fail(kCapacityError,
                "kmls_exclusive_scan_i32_noalloc: n = %lld exceeds the built-in scratch limit of %d "
                "elements; use kmls_exclusive_scan_i32 with caller scratch",
                (long long)n, KMLS_NOALLOC_SCAN_MAX_N);
  KMLS_REQUIRE(n == 0 || d_in, "kmls_exclusive_scan_i32_noalloc: d_in is null");
  // device address of the array, looked up once per device with that device current
  constexpr int kMaxDevices = 64;
  // (relaxed atomics: two host threads making the first call together both look the address
  // up and store the same value)
  static std::atomic<void*> scratch_of_device[kMaxDevices];
  int dev = -1;
  KMLS_CUDA(cudaGetDevice(&dev));
  KMLS_REQUIRE(dev >= 0 && dev < kMaxDevices, "kmls_exclusive_scan_i32_noalloc: device index out of range");
  void* scratch = scratch_of_device[dev].load(std::memory_order_relaxed);
  if (scratch == nullptr) {
    KMLS_CUDA(cudaGetSymbolAddress(&scratch, g_noalloc_scan_scratch));
    KMLS_REQUIRE(scratch != nullptr, "kmls_exclusive_scan_i32_noalloc: no device address for the scratch array");
    scratch_of_device[dev].store(scratch, std::memory_order_relaxed);
  }
  return scan_i32(d_in, n, d_out, scratch, kNoAllocScratchBytes, ToI64(), (cudaStream_t)stream);
}
