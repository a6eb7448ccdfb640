// Level 1 of the hot path: transaction encoding and single-item supports.
//
// Replaces mlxtend's TransactionEncoder.fit/transform and the column sum of
// fpcommon.setup_fptree (call sites: reference machine-learning/main.py:267-272).
// Data layout: playlists arrive as CSR by playlist (what `list(playlist_dict.values())`,
// main.py:265, is once names are ids); they leave as a VERTICAL bitmap: one row per kept
// track, one bit per playlist, rows padded to 256 playlists (32 B).
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

// ---------------------------------------------------------------------------------------
// K_hist: occurrences of every track id in the CSR item array. An upper bound of the
// track's support (a track repeated inside a playlist is counted twice here), used only to
// decide which rows are worth materialising; exact supports come from K1 below.
// ---------------------------------------------------------------------------------------
// Popular tracks make this a contended histogram (the most frequent track of a Zipfian
// catalogue is in most playlists: a million increments of one address), so every CTA first
// counts into a direct-mapped cache of kHistSlots (track id, count) pairs in shared memory and
// only ids that lose their slot to another id go straight to the global counters; the cache
// is flushed once at the end. Few, large CTAs keep the number of flushes small.
constexpr int kHistSlots = 4096;
constexpr int kHistThreads = 1024;

__device__ __forceinline__ void hist_add(int32_t v, int32_t n_items, int32_t* tag, int32_t* cnt,
                                         int32_t* __restrict__ occ) {
  if ((uint32_t)v >= (uint32_t)n_items) return;
  const int slot = v & (kHistSlots - 1);
  int32_t owner = tag[slot];                 // -1 -> an id exactly once, never changed afterwards
  if (owner == -1) {
    owner = atomicCAS(tag + slot, -1, v);
    if (owner == -1) owner = v;              // this thread claimed the slot
  }
  if (owner == v) atomicAdd(cnt + slot, 1);
  else atomicAdd(occ + v, 1);
}

__global__ void __launch_bounds__(kHistThreads) k_item_histogram(const int32_t* __restrict__ items,
                                                                 int64_t nnz, int32_t n_items,
                                                                 int32_t* __restrict__ occ) {
  __shared__ int32_t tag[kHistSlots], cnt[kHistSlots];
  for (int i = threadIdx.x; i < kHistSlots; i += blockDim.x) {
    tag[i] = -1;
    cnt[i] = 0;
  }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n4 = nnz >> 2;
  const int4* items4 = reinterpret_cast<const int4*>(items);
  for (int64_t i = tid; i < n4; i += stride) {
    const int4 v = __ldg(items4 + i);
    hist_add(v.x, n_items, tag, cnt, occ);
    hist_add(v.y, n_items, tag, cnt, occ);
    hist_add(v.z, n_items, tag, cnt, occ);
    hist_add(v.w, n_items, tag, cnt, occ);
  }
  for (int64_t i = (n4 << 2) + tid; i < nnz; i += stride) hist_add(__ldg(items + i), n_items, tag, cnt, occ);
  __syncthreads();
  for (int i = threadIdx.x; i < kHistSlots; i += blockDim.x)
    if (cnt[i]) atomicAdd(occ + tag[i], cnt[i]);
}

// ---------------------------------------------------------------------------------------
// K0: CSR-by-playlist -> vertical bitmap. One CTA owns a tile of 256 playlists x up to
// `rows_per_tile` rows in shared memory: it scatters the tile's (playlist, track) pairs
// into it with shared-memory atomicOr (duplicates collapse, as TransactionEncoder writes
// True once), then streams the tile out with 128-bit stores, two per row (one full 32-byte
// sector). Every word of the bitmap is written exactly once, so no memset is needed.
// ---------------------------------------------------------------------------------------
constexpr int kEncodeCols = 256;               // playlists per tile
constexpr int kEncodeWordsPerRow = kEncodeCols / 32;  // 8 words = 32 B
constexpr int kEncodeThreads = 512;
constexpr int kEncodeMaxRows = 7168;           // 7168 * 32 B = 224 KB of shared memory

__global__ void __launch_bounds__(kEncodeThreads, 1)
k_encode_bitmap(const int64_t* __restrict__ offsets, const int32_t* __restrict__ items,
                int64_t n_playlists, const int32_t* __restrict__ row_of_item, int32_t n_items,
                int32_t n_rows, int32_t rows_per_tile, uint32_t* __restrict__ bitmap,
                int64_t row_words) {
  extern __shared__ __align__(16) uint32_t tile[];
  const int32_t row0 = blockIdx.y * rows_per_tile;
  const int32_t rows = min(rows_per_tile, n_rows - row0);
  const int64_t p0 = (int64_t)blockIdx.x * kEncodeCols;
  const int64_t p1 = min(p0 + kEncodeCols, n_playlists);

  uint4* tile4 = reinterpret_cast<uint4*>(tile);
  for (int i = threadIdx.x; i < rows * 2; i += kEncodeThreads) tile4[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t p = p0 + warp; p < p1; p += kEncodeThreads / 32) {
    const int64_t e0 = __ldg(offsets + p), e1 = __ldg(offsets + p + 1);
    const int pl = (int)(p - p0);
    const uint32_t bit = 1u << (pl & 31);
    const int word = pl >> 5;
    for (int64_t e = e0 + lane; e < e1; e += 32) {
      const int32_t it = __ldg(items + e);
      if ((uint32_t)it >= (uint32_t)n_items) continue;
      const int32_t r = __ldg(row_of_item + it) - row0;
      if ((uint32_t)r < (uint32_t)rows) atomicOr(tile + r * kEncodeWordsPerRow + word, bit);
    }
  }
  __syncthreads();

  // two lanes per row: 2 x 16 B = one 32 B sector per row per tile
  uint4* out = reinterpret_cast<uint4*>(bitmap);
  const int64_t row_vec = row_words >> 2;
  for (int i = threadIdx.x; i < rows * 2; i += kEncodeThreads) {
    const int r = i >> 1, h = i & 1;
    out[(int64_t)(row0 + r) * row_vec + (int64_t)blockIdx.x * 2 + h] = tile4[i];
  }
}

// ---------------------------------------------------------------------------------------
// K1: exact support count of every bitmap row: 128-bit loads, __popc, warp-shuffle reduce.
// One CTA per row (long rows) or one warp per row (short rows).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_row_popcount_cta(const uint32_t* __restrict__ bitmap,
                                                          int64_t row_words,
                                                          int32_t* __restrict__ counts) {
  const uint4* row = reinterpret_cast<const uint4*>(bitmap + (int64_t)blockIdx.x * row_words);
  const int64_t nv = row_words >> 2;
  int acc = 0;
  for (int64_t v = threadIdx.x; v < nv; v += 256) acc += popc4(ldg4(row + v));
  acc = warp_sum(acc);
  __shared__ int part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = threadIdx.x < 8 ? part[threadIdx.x] : 0;
    v = warp_sum(v);
    if (threadIdx.x == 0) counts[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(256) k_row_popcount_warp(const uint32_t* __restrict__ bitmap,
                                                           int32_t n_rows, int64_t row_words,
                                                           int32_t* __restrict__ counts) {
  const int lane = threadIdx.x & 31;
  const int64_t nv = row_words >> 2;
  for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < n_rows;
       r += (int64_t)gridDim.x * 8) {
    const uint4* row = reinterpret_cast<const uint4*>(bitmap + r * row_words);
    int acc = 0;
    for (int64_t v = lane; v < nv; v += 32) acc += popc4(ldg4(row + v));
    acc = warp_sum(acc);
    if (lane == 0) counts[r] = acc;
  }
}

// Row gather: dst row r = src row src_row[r] (used to put rows into their final order when
// duplicate names made the provisional order differ from the exact one).
__global__ void __launch_bounds__(256) k_gather_rows(const uint32_t* __restrict__ src,
                                                     const int32_t* __restrict__ src_row,
                                                     int64_t row_words,
                                                     uint32_t* __restrict__ dst) {
  const uint4* s = reinterpret_cast<const uint4*>(src + (int64_t)__ldg(src_row + blockIdx.x) * row_words);
  uint4* d = reinterpret_cast<uint4*>(dst + (int64_t)blockIdx.x * row_words);
  const int64_t nv = row_words >> 2;
  for (int64_t v = threadIdx.x; v < nv; v += 256) d[v] = ldg4(s + v);
}

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_item_histogram(const int32_t* d_items, int64_t nnz, int32_t n_items,
                                   int32_t* d_occ, kmls_stream_t stream) {
  KMLS_REQUIRE(nnz >= 0 && n_items >= 0, "kmls_item_histogram: negative size");
  cudaStream_t s = (cudaStream_t)stream;
  if (n_items == 0) return 0;
  KMLS_REQUIRE(d_occ != nullptr, "kmls_item_histogram: d_occ is null");
  KMLS_CUDA(cudaMemsetAsync(d_occ, 0, sizeof(int32_t) * (size_t)n_items, s));
  if (nnz == 0) return 0;
  KMLS_REQUIRE(d_items != nullptr, "kmls_item_histogram: d_items is null");
  KMLS_REQUIRE(((uintptr_t)d_items & 15) == 0, "kmls_item_histogram: d_items must be 16-byte aligned");
  int64_t want = (nnz / 4 + kHistThreads - 1) / kHistThreads;
  int grid = (int)std::min<int64_t>(std::max<int64_t>(want, 1), (int64_t)num_sms() * 2);
  k_item_histogram<<<grid, kHistThreads, 0, s>>>(d_items, nnz, n_items, d_occ);
  KMLS_LAUNCHED("k_item_histogram");
  return 0;
}

extern "C" int64_t kmls_bitmap_row_words(int64_t n_playlists) {
  int64_t tiles = (n_playlists + kEncodeCols - 1) / kEncodeCols;
  if (tiles < 1) tiles = 1;
  return tiles * kEncodeWordsPerRow;
}

extern "C" int kmls_encode_bitmap(const int64_t* d_offsets, const int32_t* d_items,
                                  int64_t n_playlists, const int32_t* d_row_of_item,
                                  int32_t n_items, int32_t n_rows, uint32_t* d_bitmap,
                                  int64_t row_words, kmls_stream_t stream) {
  KMLS_REQUIRE(n_playlists >= 0 && n_rows >= 0 && n_items >= 0, "kmls_encode_bitmap: negative size");
  KMLS_REQUIRE(row_words == kmls_bitmap_row_words(n_playlists),
               "kmls_encode_bitmap: row_words must equal kmls_bitmap_row_words(n_playlists)");
  if (n_rows == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_offsets && d_row_of_item, "kmls_encode_bitmap: null pointer");
  KMLS_REQUIRE(((uintptr_t)d_bitmap & 15) == 0, "kmls_encode_bitmap: d_bitmap must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t col_tiles = row_words / kEncodeWordsPerRow;
  KMLS_REQUIRE(col_tiles <= 0x7fffffff, "kmls_encode_bitmap: too many playlists for one launch");
  int row_tiles = (n_rows + kEncodeMaxRows - 1) / kEncodeMaxRows;
  int rows_per_tile = (n_rows + row_tiles - 1) / row_tiles;
  KMLS_REQUIRE(row_tiles <= 65535, "kmls_encode_bitmap: too many rows");
  size_t smem = (size_t)rows_per_tile * kEncodeWordsPerRow * sizeof(uint32_t);
  // the opt-in is a per-device attribute: set it on every call (a process may mine on several GPUs)
  KMLS_CUDA(cudaFuncSetAttribute(k_encode_bitmap, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 kEncodeMaxRows * kEncodeWordsPerRow * (int)sizeof(uint32_t)));
  dim3 grid((unsigned)col_tiles, (unsigned)row_tiles);
  k_encode_bitmap<<<grid, kEncodeThreads, smem, s>>>(d_offsets, d_items, n_playlists,
                                                     d_row_of_item, n_items, n_rows,
                                                     rows_per_tile, d_bitmap, row_words);
  KMLS_LAUNCHED("k_encode_bitmap");
  return 0;
}

extern "C" int kmls_row_popcount(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                                 int32_t* d_counts, kmls_stream_t stream) {
  KMLS_REQUIRE(n_rows >= 0 && row_words >= 0 && (row_words & 3) == 0,
               "kmls_row_popcount: row_words must be a multiple of 4");
  if (n_rows == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_counts, "kmls_row_popcount: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  if (row_words >= 4096) {
    k_row_popcount_cta<<<n_rows, 256, 0, s>>>(d_bitmap, row_words, d_counts);
  } else {
    int grid = std::min((n_rows + 7) / 8, num_sms() * 32);
    k_row_popcount_warp<<<grid, 256, 0, s>>>(d_bitmap, n_rows, row_words, d_counts);
  }
  KMLS_LAUNCHED("k_row_popcount");
  return 0;
}

extern "C" int kmls_gather_rows(const uint32_t* d_src, const int32_t* d_src_row,
                                int32_t n_rows_out, int64_t row_words, uint32_t* d_dst,
                                kmls_stream_t stream) {
  KMLS_REQUIRE(n_rows_out >= 0 && (row_words & 3) == 0, "kmls_gather_rows: bad size");
  if (n_rows_out == 0) return 0;
  KMLS_REQUIRE(d_src && d_dst && d_src_row && d_src != d_dst, "kmls_gather_rows: null or aliased pointer");
  k_gather_rows<<<n_rows_out, 256, 0, (cudaStream_t)stream>>>(d_src, d_src_row, row_words, d_dst);
  KMLS_LAUNCHED("k_gather_rows");
  return 0;
}

// A few bytes from device memory to MAPPED (pinned, device-visible) host memory, written by
// a kernel instead of the copy engine: the sizes the host reads back between levels then do
// not queue behind bulk device-to-host transfers of earlier levels' tables.
namespace kmls {
__global__ void k_words_to_host(const uint32_t* __restrict__ src, int n_words, volatile uint32_t* dst) {
  for (int i = threadIdx.x; i < n_words; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();
}
}  // namespace kmls

extern "C" int kmls_words_to_host(const void* d_src, int32_t n_words, void* h_mapped_dst,
                                  kmls_stream_t stream) {
  KMLS_REQUIRE(n_words >= 0 && n_words <= 4096, "kmls_words_to_host: at most 4096 words");
  if (n_words == 0) return 0;
  KMLS_REQUIRE(d_src && h_mapped_dst && ((uintptr_t)d_src & 3) == 0 && ((uintptr_t)h_mapped_dst & 3) == 0,
               "kmls_words_to_host: null or misaligned pointer");
  kmls::k_words_to_host<<<1, 64, 0, (cudaStream_t)stream>>>((const uint32_t*)d_src, n_words,
                                                            (volatile uint32_t*)h_mapped_dst);
  KMLS_LAUNCHED("k_words_to_host");
  return 0;
}
