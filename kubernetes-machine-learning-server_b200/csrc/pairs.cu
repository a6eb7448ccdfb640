// Level 2 of the hot path: pair co-occurrence counts and the frequent-pair table.
//
// Replaces the k=2 part of mlxtend's FP-tree mining (fpg_step / conditional_tree, reached
// from reference machine-learning/main.py:272) and supplies everything the reference's
// dict fold (main.py:282-296) needs: by anti-monotonicity the folded value of (a, b) is
// exactly support({a, b}).
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

// ---------------------------------------------------------------------------------------
// K2 (CUDA-core form): C[i][j] = popcount(row_i AND row_j) for a 64x64 tile of row pairs.
// 256 threads, each 4x4 pairs; both operand tiles staged in shared memory KC words at a
// time. Only tiles with tile_j >= tile_i are computed; both C[i][j] and C[j][i] are written
// so later lookups need no ordering. `ksplit` > 1 splits the playlist dimension over
// blockIdx.y and accumulates with atomicAdd into a zeroed C.
// The tensor-core form of the same contraction is in pair_gemm.cu.
// ---------------------------------------------------------------------------------------
constexpr int kPT = 64;   // tile edge (rows)
constexpr int kKC = 32;   // words per staged chunk

__global__ void __launch_bounds__(256) k_pair_counts(const uint32_t* __restrict__ bitmap,
                                                     int32_t n_rows, int64_t row_words,
                                                     int32_t n_tiles, int64_t words_per_split,
                                                     int use_atomics, int32_t* __restrict__ C) {
  // linear upper-triangular tile index -> (ti, tj), ti <= tj
  int t = blockIdx.x, ti = 0;
  {
    int rem = t, len = n_tiles;
    while (rem >= len) { rem -= len; --len; ++ti; }
    t = ti + rem;
  }
  const int tj = t;
  __shared__ uint32_t sa[kPT][kKC + 1];
  __shared__ uint32_t sb[kPT][kKC + 1];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // j index group, i index group
  int acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0;

  const int64_t w_begin = (int64_t)blockIdx.y * words_per_split;
  const int64_t w_end = min(row_words, w_begin + words_per_split);
  const int lr = threadIdx.x >> 5, lw = threadIdx.x & 31;  // loader: 8 rows x 32 words per pass
  for (int64_t w0 = w_begin; w0 < w_end; w0 += kKC) {
#pragma unroll
    for (int r = 0; r < kPT; r += 8) {
      const int ri = ti * kPT + r + lr, rj = tj * kPT + r + lr;
      const int64_t w = w0 + lw;
      sa[r + lr][lw] = (ri < n_rows && w < w_end) ? __ldg(bitmap + (int64_t)ri * row_words + w) : 0u;
      sb[r + lr][lw] = (rj < n_rows && w < w_end) ? __ldg(bitmap + (int64_t)rj * row_words + w) : 0u;
    }
    __syncthreads();
#pragma unroll 4
    for (int w = 0; w < kKC; ++w) {
      uint32_t a[4], b[4];
#pragma unroll
      for (int x = 0; x < 4; ++x) { a[x] = sa[ty * 4 + x][w]; b[x] = sb[tx + 16 * x][w]; }
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] += __popc(a[x] & b[y]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int x = 0; x < 4; ++x)
#pragma unroll
    for (int y = 0; y < 4; ++y) {
      const int i = ti * kPT + ty * 4 + x, j = tj * kPT + tx + 16 * y;
      if (i >= n_rows || j >= n_rows) continue;
      if (ti == tj && j < i) continue;  // the mirrored thread writes this one
      if (use_atomics) {
        atomicAdd(C + (int64_t)i * n_rows + j, acc[x][y]);
        if (i != j) atomicAdd(C + (int64_t)j * n_rows + i, acc[x][y]);
      } else {
        C[(int64_t)i * n_rows + j] = acc[x][y];
        C[(int64_t)j * n_rows + i] = acc[x][y];
      }
    }
}

// ---------------------------------------------------------------------------------------
// K6 part 1: frequent-pair table from C. Row a keeps the b > a with C[a][b] >= thr, in
// ascending b: per-row counts -> one-block scan -> ordered write (ballot + popc rank).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pair_row_counts(const int32_t* __restrict__ C,
                                                         int32_t n_rows, int32_t thr,
                                                         int64_t* __restrict__ row_cnt) {
  const int lane = threadIdx.x & 31;
  for (int a = blockIdx.x * 8 + (threadIdx.x >> 5); a < n_rows; a += gridDim.x * 8) {
    const int32_t* row = C + (int64_t)a * n_rows;
    int acc = 0;
    for (int b = a + 1 + lane; b < n_rows; b += 32) acc += __ldg(row + b) >= thr;
    acc = warp_sum(acc);
    if (lane == 0) row_cnt[a] = acc;
  }
}

// Exclusive scan of n int64 values by ONE block (n is a row count: at most ~1e5).
__global__ void __launch_bounds__(1024) k_scan_one_block(const int64_t* __restrict__ in, int32_t n,
                                                         int64_t* __restrict__ out) {
  __shared__ int64_t warp_tot[32];
  __shared__ int64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    int64_t v = i < n ? in[i] : 0, incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int64_t u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int64_t t = warp_tot[lane], ti = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        int64_t u = __shfl_up_sync(0xffffffffu, ti, o);
        if (lane >= o) ti += u;
      }
      warp_tot[lane] = ti - t;  // exclusive over warps
    }
    __syncthreads();
    const int64_t c = carry;
    if (i < n) out[i] = c + warp_tot[warp] + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = c + warp_tot[31] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) out[n] = carry;
}

__global__ void __launch_bounds__(256) k_pair_write(const int32_t* __restrict__ C, int32_t n_rows,
                                                    int32_t thr, const int64_t* __restrict__ row_off,
                                                    int64_t capacity, int32_t* __restrict__ items2,
                                                    int32_t* __restrict__ counts) {
  const int lane = threadIdx.x & 31;
  for (int a = blockIdx.x * 8 + (threadIdx.x >> 5); a < n_rows; a += gridDim.x * 8) {
    const int32_t* row = C + (int64_t)a * n_rows;
    int64_t pos = row_off[a];
    if (row_off[a + 1] > capacity) continue;  // host reports the overflow
    for (int b0 = a + 1; b0 < n_rows; b0 += 32) {
      const int b = b0 + lane;
      const int32_t c = b < n_rows ? __ldg(row + b) : -1;
      const bool keep = b < n_rows && c >= thr;
      const uint32_t m = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const int64_t o = pos + __popc(m & ((1u << lane) - 1));
        items2[2 * o] = a;
        items2[2 * o + 1] = b;
        counts[o] = c;
      }
      pos += __popc(m);
    }
  }
}

// support = count / N exactly as Python's int / int true division gives it for these
// magnitudes (both fit in 53 bits, so the quotient of the two doubles is correctly rounded).
__global__ void k_supports(const int32_t* __restrict__ counts, int64_t n, double n_total,
                           double* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    out[i] = __ddiv_rn((double)counts[i], n_total);
}


// ---------------------------------------------------------------------------------------
// K2, sparse form: pair counts straight from the playlists. A playlist with m kept tracks
// contributes to m(m-1)/2 pairs, so the whole level costs sum_p m_p^2 / 2 counter updates -
// for a sparse one-hot matrix (config 3: 1 % dense, m ~ 46, 1.0e9 updates) far less than the
// F^2 N / 2 multiply-adds of the dense contraction (3.6e13), most of which multiply zeros.
// One warp per playlist: track ids -> rows, duplicates dropped with a per-warp bit set over
// the rows (TransactionEncoder semantics: a repeated track counts once), the unique rows
// compacted into shared memory, then every pair (and the diagonal) is one RED.ADD into the
// pair matrix, which the hot (Zipf) pairs keep resident in L2. Only the upper triangle and
// the diagonal are written. Playlists with more than kCsrMaxLen distinct kept tracks raise
// *overflow instead (the caller then uses the dense kernels).
// ---------------------------------------------------------------------------------------
constexpr int kCsrMaxLen = 1024;     // distinct kept tracks of one playlist
constexpr int kCsrMaxRows = 16384;   // rows the per-warp bit set covers

constexpr int kCsrHot = 64;          // the kCsrHot most frequent rows' pairs are counted in shared memory

// MODE 0: every pair is one RED.ADD into the matrix. MODE 1: the pairs among the kCsrHot most
// frequent rows are counted in shared memory per CTA and flushed once. MODE 2: the pairs (and
// diagonal entries) among the last n_hot rows are LEFT OUT - the caller counts that block on
// the tensor cores (kmls_pair_counts_i8_block); a playlist's rows are kept as cold rows from
// the front and hot rows from the back of its list, so the left-out pairs cost no iterations.
template <int MODE>
__global__ void __launch_bounds__(256)
k_pair_counts_csr(const int64_t* __restrict__ offsets, const int32_t* __restrict__ items, int64_t n,
                  const int32_t* __restrict__ row_of_item, int32_t n_rows, int32_t n_hot,
                  int32_t* __restrict__ pair, int32_t* __restrict__ overflow) {
  extern __shared__ uint32_t csr_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int seen_words = (n_rows + 31) >> 5;
  uint32_t* seen = csr_smem + (size_t)warp * seen_words;
  int32_t* rows = reinterpret_cast<int32_t*>(csr_smem + (size_t)8 * seen_words) + (size_t)warp * kCsrMaxLen;
  for (int i = lane; i < seen_words; i += 32) seen[i] = 0u;
  __syncwarp();
  // Rows are in ascending-support order, so the last rows are the most frequent tracks; under
  // a Zipf law their pairs take a large share of all updates and would serialise on a few L2
  // atomic units.
  __shared__ int32_t hot[MODE == 1 ? kCsrHot * kCsrHot : 1];
  const int hot0 = n_rows - (MODE == 1 ? kCsrHot : MODE == 2 ? n_hot : 0);        // first hot row
  if (MODE == 1) {
    for (int i = threadIdx.x; i < kCsrHot * kCsrHot; i += blockDim.x) hot[i] = 0;
    __syncthreads();
  }
  const uint32_t below = (1u << lane) - 1u;
  for (int64_t p = (int64_t)blockIdx.x * 8 + warp; p < n; p += (int64_t)gridDim.x * 8) {
    const int64_t b = __ldg(offsets + p), e = __ldg(offsets + p + 1);
    int m = 0, mh = 0;          // rows kept from the front, (MODE 2) hot rows kept from the back
    bool over = false;
    for (int64_t q = b; q < e; q += 32) {
      int32_t r = -1;
      if (q + lane < e) r = __ldg(row_of_item + __ldg(items + q + lane));
      bool uniq = false;
      if (r >= 0) {
        const uint32_t bit = 1u << (r & 31);
        uniq = !(atomicOr(seen + (r >> 5), bit) & bit);     // first occurrence in this playlist
      }
      if (MODE == 2) {
        const bool is_hot = uniq && r >= hot0;
        const uint32_t mask_c = __ballot_sync(0xffffffffu, uniq && !is_hot);
        const uint32_t mask_h = __ballot_sync(0xffffffffu, is_hot);
        const int m1 = m + __popc(mask_c), mh1 = mh + __popc(mask_h);
        if (m1 + mh1 <= kCsrMaxLen) {
          if (uniq && !is_hot) rows[m + __popc(mask_c & below)] = r;
          if (is_hot) rows[kCsrMaxLen - 1 - (mh + __popc(mask_h & below))] = r;
        }
        m = m1;
        mh = mh1;
      } else {
        const uint32_t mask = __ballot_sync(0xffffffffu, uniq);
        const int pos = m + __popc(mask & below);
        if (uniq && pos < kCsrMaxLen) rows[pos] = r;
        m += __popc(mask);
      }
    }
    __syncwarp();
    if (m + mh > kCsrMaxLen) over = true;
    if (over) {
      // the dropped rows' bits must be cleared as well: rebuild from the items
      for (int64_t q = b + lane; q < e; q += 32) {
        const int32_t r = __ldg(row_of_item + __ldg(items + q));
        if (r >= 0) seen[r >> 5] = 0u;
      }
      if (lane == 0) atomicAdd(overflow, 1);
      __syncwarp();
      continue;
    }
    const int mt = m + mh;                  // list: cold rows [0, m), then hot rows from the array's end
    const int shift = kCsrMaxLen - mt;      // entry j >= m lives at rows[j + shift]
    for (int t = lane; t < mt; t += 32) seen[rows[t < m ? t : t + shift] >> 5] = 0u;     // leave the bit set clean
    __syncwarp();
    // diagonal, then all pairs i < j of the list with i < m: lane t takes pairs t, t+32, ...
    // of the row-major enumeration (i, c) with j = i + 1 + c
    for (int t = lane; t < m; t += 32) {
      const int64_t r = rows[t];
      atomicAdd(pair + r * n_rows + r, 1);
    }
    const int mi = min(m, mt - 1);          // first members that open pairs
    int i = 0, c = lane;
    while (i < mi && c >= mt - 1 - i) { c -= mt - 1 - i; ++i; }
    while (i < mi) {
      const int j = i + 1 + c;
      const int32_t ra = rows[i], rb = rows[j < m ? j : j + shift];
      const int64_t lo = min(ra, rb), hi = max(ra, rb);
      if (MODE == 1 && lo >= hot0) atomicAdd(hot + (lo - hot0) * kCsrHot + (hi - hot0), 1);
      else atomicAdd(pair + lo * n_rows + hi, 1);
      c += 32;
      while (i < mi && c >= mt - 1 - i) { c -= mt - 1 - i; ++i; }
    }
    __syncwarp();
  }
  if (MODE == 1) {
    __syncthreads();
    for (int i = threadIdx.x; i < kCsrHot * kCsrHot; i += blockDim.x) {
      const int v = hot[i];
      if (v) atomicAdd(pair + (int64_t)(hot0 + i / kCsrHot) * n_rows + (hot0 + i % kCsrHot), v);
    }
  }
}

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_pair_counts(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                                int32_t* d_pair, kmls_stream_t stream) {
  KMLS_REQUIRE(n_rows >= 0 && row_words >= 0, "kmls_pair_counts: negative size");
  if (n_rows == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_pair, "kmls_pair_counts: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  const int n_tiles = (n_rows + kPT - 1) / kPT;
  const int64_t tri = (int64_t)n_tiles * (n_tiles + 1) / 2;
  KMLS_REQUIRE(tri <= 0x7fffffff, "kmls_pair_counts: too many rows");
  const int64_t chunks = (row_words + kKC - 1) / kKC;
  int64_t ksplit = 1;
  const int64_t target = (int64_t)num_sms() * 4;
  if (tri < target) ksplit = std::min<int64_t>(std::max<int64_t>(chunks / 4, 1), (target + tri - 1) / tri);
  ksplit = std::min<int64_t>(ksplit, 65535);
  const int64_t words_per_split = ((chunks + ksplit - 1) / ksplit) * kKC;
  ksplit = (row_words + words_per_split - 1) / words_per_split;
  if (ksplit < 1) ksplit = 1;
  if (ksplit > 1) KMLS_CUDA(cudaMemsetAsync(d_pair, 0, sizeof(int32_t) * (size_t)n_rows * n_rows, s));
  dim3 grid((unsigned)tri, (unsigned)ksplit);
  k_pair_counts<<<grid, 256, 0, s>>>(d_bitmap, n_rows, row_words, n_tiles, words_per_split,
                                     ksplit > 1 ? 1 : 0, d_pair);
  KMLS_LAUNCHED("k_pair_counts");
  return 0;
}

static int pair_counts_csr_run(const char* who, const int64_t* d_offsets, const int32_t* d_items,
                               int64_t n_playlists, const int32_t* d_row_of_item, int32_t n_rows, int32_t n_hot,
                               int32_t* d_pair, int32_t* d_overflow, cudaStream_t s) {
  KMLS_REQUIRE(n_playlists >= 0 && n_rows >= 0, "kmls_pair_counts_csr: negative size");
  KMLS_REQUIRE(n_rows <= kCsrMaxRows, "kmls_pair_counts_csr: more than 16384 rows (use kmls_pair_counts_i8)");
  KMLS_REQUIRE(n_hot >= 0 && n_hot <= n_rows, "kmls_pair_counts_csr_cold: n_hot out of range");
  if (n_rows == 0) return 0;
  KMLS_REQUIRE(d_pair && d_overflow, "kmls_pair_counts_csr: null pointer");
  KMLS_CUDA(cudaMemsetAsync(d_pair, 0, sizeof(int32_t) * (size_t)n_rows * n_rows, s));
  KMLS_CUDA(cudaMemsetAsync(d_overflow, 0, sizeof(int32_t), s));
  if (n_playlists == 0) return 0;
  KMLS_REQUIRE(d_offsets && d_items && d_row_of_item, "kmls_pair_counts_csr: null pointer");
  (void)who;
  const size_t smem = sizeof(uint32_t) * 8 * (size_t)((n_rows + 31) / 32) + sizeof(int32_t) * 8 * kCsrMaxLen;
  static const bool no_hot = [] { const char* e = getenv("KMLS_PAIR_CSR_HOT"); return e && e[0] == '0'; }();
  const int mode = n_hot > 0 ? 2 : (!no_hot && n_rows >= kCsrHot) ? 1 : 0;
  auto kern = mode == 2 ? k_pair_counts_csr<2> : mode == 1 ? k_pair_counts_csr<1> : k_pair_counts_csr<0>;
  KMLS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  int grid = (int)std::min<int64_t>((n_playlists + 7) / 8, (int64_t)num_sms() * (mode == 1 ? 4 : 16));
  kern<<<grid, 256, smem, s>>>(d_offsets, d_items, n_playlists, d_row_of_item, n_rows, n_hot, d_pair, d_overflow);
  KMLS_LAUNCHED("k_pair_counts_csr");
  return 0;
}

extern "C" int kmls_pair_counts_csr(const int64_t* d_offsets, const int32_t* d_items,
                                    int64_t n_playlists, const int32_t* d_row_of_item,
                                    int32_t n_rows, int32_t* d_pair, int32_t* d_overflow,
                                    kmls_stream_t stream) {
  return pair_counts_csr_run("kmls_pair_counts_csr", d_offsets, d_items, n_playlists, d_row_of_item, n_rows, 0,
                             d_pair, d_overflow, (cudaStream_t)stream);
}

extern "C" int kmls_pair_counts_csr_cold(const int64_t* d_offsets, const int32_t* d_items,
                                         int64_t n_playlists, const int32_t* d_row_of_item,
                                         int32_t n_rows, int32_t n_hot, int32_t* d_pair,
                                         int32_t* d_overflow, kmls_stream_t stream) {
  return pair_counts_csr_run("kmls_pair_counts_csr_cold", d_offsets, d_items, n_playlists, d_row_of_item, n_rows,
                             n_hot, d_pair, d_overflow, (cudaStream_t)stream);
}

extern "C" int kmls_pair_select(const int32_t* d_pair, int32_t n_rows, int32_t thr,
                                int64_t capacity, int64_t* d_row_off, int32_t* d_items2,
                                int32_t* d_counts, kmls_stream_t stream) {
  KMLS_REQUIRE(n_rows >= 0 && capacity >= 0, "kmls_pair_select: negative size");
  KMLS_REQUIRE(d_row_off, "kmls_pair_select: d_row_off is null");
  cudaStream_t s = (cudaStream_t)stream;
  if (n_rows == 0) {
    KMLS_CUDA(cudaMemsetAsync(d_row_off, 0, sizeof(int64_t), s));
    return 0;
  }
  KMLS_REQUIRE(d_pair, "kmls_pair_select: null pointer");
  // d_row_off doubles as the count buffer: counts go to [1, n_rows], are scanned in place
  // from a shifted view. Keep it simple: counts into d_row_off[0..n), scan in place.
  int grid = std::min((n_rows + 7) / 8, num_sms() * 16);
  k_pair_row_counts<<<grid, 256, 0, s>>>(d_pair, n_rows, thr, d_row_off);
  KMLS_LAUNCHED("k_pair_row_counts");
  k_scan_one_block<<<1, 1024, 0, s>>>(d_row_off, n_rows, d_row_off);
  KMLS_LAUNCHED("k_scan_one_block");
  if (capacity > 0) {
    KMLS_REQUIRE(d_items2 && d_counts, "kmls_pair_select: null output");
    k_pair_write<<<grid, 256, 0, s>>>(d_pair, n_rows, thr, d_row_off, capacity, d_items2, d_counts);
    KMLS_LAUNCHED("k_pair_write");
  }
  return 0;
}

extern "C" int kmls_supports_f64(const int32_t* d_counts, int64_t n, int64_t n_playlists_total,
                                 double* d_out, kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && n_playlists_total > 0, "kmls_supports_f64: bad size");
  if (n == 0) return 0;
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 8);
  k_supports<<<grid, 256, 0, (cudaStream_t)stream>>>(d_counts, n, (double)n_playlists_total, d_out);
  KMLS_LAUNCHED("k_supports");
  return 0;
}

// ---------------------------------------------------------------------------------------
// K6: the fold of reference machine-learning/main.py:282-296 as a CSR built on the device.
// The reference walks every frequent itemset and keeps, for every ordered pair of its members,
// the largest support seen; by anti-monotonicity that is the support of the pair itself
// (SURVEY.md F4), so the map is fully described by the frequent-pair table:
//   track a -> [(track b, support({a, b})) for every frequent pair containing a]
// in both directions. Inner lists come out sorted by support DESCENDING, ties by track id:
// rest_api sorts the merged candidates with a stable sort (rest_api/app/main.py:250), so a
// deterministic inner order makes its top-K deterministic too (the reference's own order
// depends on Python's per-process string hashing).
// ---------------------------------------------------------------------------------------
#include <cub/device/device_segmented_sort.cuh>

#include "workspace.cuh"

namespace kmls {

__global__ void __launch_bounds__(256)
k_pair_degree(const int32_t* __restrict__ items2, int64_t n2, int32_t* __restrict__ degree) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < 2 * n2; i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(degree + items2[i], 1);
}

// key = (~count) << 32 | other track id: ascending keys = descending count, then ascending id
__global__ void __launch_bounds__(256)
k_pair_map_fill(const int32_t* __restrict__ items2, const int32_t* __restrict__ counts2, int64_t n2,
                const int32_t* __restrict__ row_items, const int64_t* __restrict__ indptr,
                int32_t* __restrict__ cursor, unsigned long long* __restrict__ keys) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t a = items2[2 * i], b = items2[2 * i + 1];
    const unsigned long long hi = (unsigned long long)(~(uint32_t)counts2[i]) << 32;
    keys[indptr[a] + atomicAdd(cursor + a, 1)] = hi | (uint32_t)__ldg(row_items + b);
    keys[indptr[b] + atomicAdd(cursor + b, 1)] = hi | (uint32_t)__ldg(row_items + a);
  }
}

__global__ void __launch_bounds__(256)
k_pair_map_decode(const unsigned long long* __restrict__ keys, int64_t n, double n_total,
                  int32_t* __restrict__ other, double* __restrict__ support) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long k = keys[i];
    other[i] = (int32_t)(uint32_t)k;
    support[i] = __ddiv_rn((double)(~(uint32_t)(k >> 32)), n_total);   // count / N as Python divides
  }
}

static size_t pair_map_sort_bytes(int64_t n_entries, int32_t n_rows) {
  size_t need = 0;
  cub::DeviceSegmentedSort::SortKeys((void*)nullptr, need, (const unsigned long long*)nullptr,
                                     (unsigned long long*)nullptr, n_entries, (int64_t)n_rows, (const int64_t*)nullptr,
                                     (const int64_t*)nullptr);
  return need;
}

}  // namespace kmls

extern "C" size_t kmls_pair_map_scratch_bytes(int64_t n_pairs, int32_t n_rows) {
  if (n_pairs < 0) n_pairs = 0;
  if (n_rows < 0) n_rows = 0;
  return 2 * (size_t)n_rows * 4 + 2 * (size_t)(2 * n_pairs) * 8 + kmls::pair_map_sort_bytes(2 * n_pairs, n_rows) +
         kmls_scan_scratch_bytes(n_rows) + 8 * 256;
}

extern "C" int kmls_pair_map(const int32_t* d_items2, const int32_t* d_counts2, int64_t n_pairs,
                             const int32_t* d_row_items, int32_t n_rows, int64_t n_playlists_total,
                             int64_t* d_indptr, int32_t* d_other, double* d_support, void* d_scratch,
                             size_t scratch_bytes, kmls_stream_t stream) {
  using namespace kmls;
  KMLS_REQUIRE(n_pairs >= 0 && n_rows >= 0 && d_indptr, "kmls_pair_map: bad arguments");
  KMLS_REQUIRE(2 * n_pairs <= 0x7fffffff, "kmls_pair_map: too many pairs");
  cudaStream_t s = (cudaStream_t)stream;
  if (n_rows == 0 || n_pairs == 0) {
    KMLS_CUDA(cudaMemsetAsync(d_indptr, 0, sizeof(int64_t) * (size_t)(n_rows + 1), s));
    return 0;
  }
  KMLS_REQUIRE(d_items2 && d_counts2 && d_row_items && d_other && d_support && n_playlists_total > 0,
               "kmls_pair_map: null pointer or bad playlist count");
  KMLS_REQUIRE(scratch_bytes >= kmls_pair_map_scratch_bytes(n_pairs, n_rows) && d_scratch,
               "kmls_pair_map: scratch smaller than kmls_pair_map_scratch_bytes()");
  Workspace ws(d_scratch, scratch_bytes);
  const int64_t n_e = 2 * n_pairs;
  int32_t* degree = (int32_t*)ws.alloc(sizeof(int32_t) * n_rows);
  int32_t* cursor = (int32_t*)ws.alloc(sizeof(int32_t) * n_rows);
  unsigned long long* k_in = (unsigned long long*)ws.alloc(8 * (size_t)n_e);
  unsigned long long* k_out = (unsigned long long*)ws.alloc(8 * (size_t)n_e);
  const size_t scan_bytes = kmls_scan_scratch_bytes(n_rows), sort_bytes = pair_map_sort_bytes(n_e, n_rows);
  void* scan_tmp = ws.alloc(scan_bytes);
  void* sort_tmp = ws.alloc(sort_bytes);
  KMLS_REQUIRE(degree && cursor && k_in && k_out && scan_tmp && sort_tmp, "kmls_pair_map: scratch too small");
  KMLS_CUDA(cudaMemsetAsync(degree, 0, sizeof(int32_t) * (size_t)n_rows, s));
  KMLS_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int32_t) * (size_t)n_rows, s));
  const int grid = (int)std::min<int64_t>((n_e + 255) / 256, (int64_t)num_sms() * 8);
  k_pair_degree<<<grid, 256, 0, s>>>(d_items2, n_pairs, degree);
  KMLS_LAUNCHED("k_pair_degree");
  int rc = kmls_exclusive_scan_i32(degree, n_rows, d_indptr, scan_tmp, scan_bytes, s);
  if (rc) return rc;
  k_pair_map_fill<<<grid, 256, 0, s>>>(d_items2, d_counts2, n_pairs, d_row_items, d_indptr, cursor, k_in);
  KMLS_LAUNCHED("k_pair_map_fill");
  size_t need = sort_bytes;
  KMLS_CUDA(cub::DeviceSegmentedSort::SortKeys(sort_tmp, need, (const unsigned long long*)k_in, k_out, n_e,
                                               (int64_t)n_rows, (const int64_t*)d_indptr,
                                               (const int64_t*)d_indptr + 1, s));
  k_pair_map_decode<<<grid, 256, 0, s>>>(k_out, n_e, (double)n_playlists_total, d_other, d_support);
  KMLS_LAUNCHED("k_pair_map_decode");
  return 0;
}

// ---------------------------------------------------------------------------------------
// Micro-benchmark behind the roofline of k_pair_counts_csr (no counterpart in the reference):
// how many 32-bit RED.ADD updates per second this GPU retires when warps scatter them over a
// counter matrix of the size the pair counts use. n_updates updates, addresses drawn by a
// counter-based hash over d_buf[n_words] (uniform: the worst case; the kernel's own addresses
// are skewed towards popular rows, which it absorbs in shared memory).
// ---------------------------------------------------------------------------------------
namespace kmls {
__global__ void __launch_bounds__(256)
k_redg_microbench(int32_t* __restrict__ buf, uint64_t n_words, int64_t n_updates, uint64_t seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_updates; i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(buf + mix64(seed + (uint64_t)i) % n_words, 1);
}
// The kernel's own pattern: the 32 updates of a warp instruction fall into ONE stretch of `span`
// consecutive counters (a pair (i, j) of a playlist keeps i - the matrix row - for many j) that
// starts at a random multiple of `span`; span = 32: one 128-byte line per warp instruction.
__global__ void __launch_bounds__(256)
k_redg_microbench_span(int32_t* __restrict__ buf, uint64_t n_words, int64_t n_updates, uint64_t span, uint64_t seed) {
  const uint64_t n_spans = n_words / span;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_updates; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t base = (mix64(seed + (uint64_t)(i >> 5)) % n_spans) * span;
    atomicAdd(buf + base + mix64(~seed + (uint64_t)i) % span, 1);
  }
}
}  // namespace kmls

extern "C" int kmls_redg_microbench(int32_t* d_buf, int64_t n_words, int64_t n_updates, uint64_t seed,
                                    kmls_stream_t stream) {
  KMLS_REQUIRE(d_buf && n_words > 0 && n_updates >= 0, "kmls_redg_microbench: bad arguments");
  if (n_updates == 0) return 0;
  const int grid = (int)std::min<int64_t>((n_updates + 255) / 256, (int64_t)kmls::num_sms() * 32);
  kmls::k_redg_microbench<<<grid, 256, 0, (cudaStream_t)stream>>>(d_buf, (uint64_t)n_words, n_updates, seed);
  KMLS_LAUNCHED("k_redg_microbench");
  return 0;
}

// Streaming micro-benchmark behind k_rules_write's roofline: every element is one 8-byte read
// (`reads` of the three inputs are used: 0..3) and `writes` 8-byte writes (0..3) into separate
// arrays, the traffic mix of pass 2 (8 bytes of plane words in, three 8-byte columns out).
namespace kmls {
__global__ void __launch_bounds__(256)
k_stream_mix(const int2* __restrict__ a, const int2* __restrict__ b, const int2* __restrict__ c, int64_t n,
             int reads, int writes, int2* __restrict__ x, int2* __restrict__ y, int2* __restrict__ z) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int2 v = make_int2((int)i, 1);
    if (reads > 0) { const int2 t = __ldg(a + i); v.x ^= t.x; v.y += t.y; }
    if (reads > 1) { const int2 t = __ldg(b + i); v.x ^= t.x; v.y += t.y; }
    if (reads > 2) { const int2 t = __ldg(c + i); v.x ^= t.x; v.y += t.y; }
    if (writes > 0) x[i] = v;
    if (writes > 1) y[i] = make_int2(v.y, v.x);
    if (writes > 2) z[i] = make_int2(v.x + 1, v.y);
    if (writes == 0 && v.x == 0x7fffffff && v.y == 0x7ffffffe) x[0] = v;   // keeps the loads alive
  }
}
}  // namespace kmls

extern "C" int kmls_stream_mix_microbench(const void* d_a, const void* d_b, const void* d_c, int64_t n,
                                          int32_t reads, int32_t writes, void* d_x, void* d_y, void* d_z,
                                          kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0 && reads >= 0 && reads <= 3 && writes >= 0 && writes <= 3 && d_a && d_b && d_c && d_x && d_y && d_z,
               "kmls_stream_mix_microbench: bad arguments");
  if (n == 0) return 0;
  const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)kmls::num_sms() * 32);
  kmls::k_stream_mix<<<grid, 256, 0, (cudaStream_t)stream>>>((const int2*)d_a, (const int2*)d_b, (const int2*)d_c, n,
                                                             reads, writes, (int2*)d_x, (int2*)d_y, (int2*)d_z);
  KMLS_LAUNCHED("k_stream_mix");
  return 0;
}

extern "C" int kmls_redg_microbench_span(int32_t* d_buf, int64_t n_words, int64_t n_updates, int64_t span,
                                         uint64_t seed, kmls_stream_t stream) {
  KMLS_REQUIRE(d_buf && span > 0 && n_words >= span && n_updates >= 0, "kmls_redg_microbench_span: bad arguments");
  if (n_updates == 0) return 0;
  const int grid = (int)std::min<int64_t>((n_updates + 255) / 256, (int64_t)kmls::num_sms() * 32);
  kmls::k_redg_microbench_span<<<grid, 256, 0, (cudaStream_t)stream>>>(d_buf, (uint64_t)n_words, n_updates,
                                                                       (uint64_t)span, seed);
  KMLS_LAUNCHED("k_redg_microbench_span");
  return 0;
}
