// Conditional (projected) bitmaps for levels k >= 3.
//
// mlxtend's FP-Growth gets its speed from conditional trees: once a prefix is fixed only
// the transactions containing it are looked at (FPTree.conditional_tree, reached from
// reference machine-learning/main.py:272). The bitmap analogue used here: every frequent
// k-itemset X = (P, x) owns a bit row over the playlists that contain its class prefix P
// (columns compacted, so the row is c(P) bits wide, not N). Then
//   * the support of a candidate (P, x_i, x_j) is popcount(row(P,x_i) AND row(P,x_j));
//   * the row of a new (k+1)-itemset (P, x_i, x_j) is row(P,x_j) with the columns where
//     row(P,x_i) is 0 squeezed out (a parallel bit extract), c(P,x_i) bits wide - or, once
//     rows are at most 128 words, simply row(P,x_i) AND row(P,x_j) in the same columns.
// Level 1 is the full bitmap (P = {}, all N columns). Work per candidate therefore shrinks
// with the support of its prefix instead of staying at N/8 bytes.
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

// ---------------------------------------------------------------------------------------
// K4 (projected form): one warp per candidate (p, q): popcount(row_p AND row_q) with 128-bit
// loads and a warp-shuffle reduce. Both rows live in the class's compacted column space and
// are roww[p] words wide. Algorithmic bytes: 2 * roww * 4 per candidate.
// ---------------------------------------------------------------------------------------
constexpr int kCountIlp = 4;   // candidates in flight per warp

__global__ void __launch_bounds__(256)
k_count_pairs(const uint32_t* __restrict__ arena, const int64_t* __restrict__ rowptr,
              const int32_t* __restrict__ roww, const int32_t* __restrict__ cand_p,
              const int32_t* __restrict__ cand_q, int64_t n_cand, int32_t* __restrict__ cand_cnt) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * 8;
  // Each candidate is a chain of dependent loads (ids -> row pointers -> row words), so a
  // warp keeps kCountIlp candidates in flight: all ids first, then all pointers, then the
  // first 128-bit word pair of every row, and only then the reductions.
  for (int64_t c0 = warp0 * kCountIlp; c0 < n_cand; c0 += stride * kCountIlp) {
    int32_t p[kCountIlp], q[kCountIlp];
#pragma unroll
    for (int u = 0; u < kCountIlp; ++u) {
      const int64_t c = min(c0 + u, n_cand - 1);
      p[u] = __ldg(cand_p + c);
      q[u] = __ldg(cand_q + c);
    }
    int nv[kCountIlp];
    const uint4 *rp[kCountIlp], *rq[kCountIlp];
#pragma unroll
    for (int u = 0; u < kCountIlp; ++u) {
      nv[u] = __ldg(roww + p[u]) >> 2;
      rp[u] = reinterpret_cast<const uint4*>(arena + __ldg(rowptr + p[u]));
      rq[u] = reinterpret_cast<const uint4*>(arena + __ldg(rowptr + q[u]));
    }
    uint4 a0[kCountIlp], b0[kCountIlp];
#pragma unroll
    for (int u = 0; u < kCountIlp; ++u) {
      a0[u] = lane < nv[u] ? ldg4(rp[u] + lane) : make_uint4(0, 0, 0, 0);
      b0[u] = lane < nv[u] ? ldg4(rq[u] + lane) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kCountIlp; ++u) {
      int acc = popc4(and4(a0[u], b0[u]));
      for (int v = lane + 32; v < nv[u]; v += 32) acc += popc4(and4(ldg4(rp[u] + v), ldg4(rq[u] + v)));
      acc = warp_sum(acc);
      if (lane == 0 && c0 + u < n_cand) cand_cnt[c0 + u] = acc;
    }
  }
}

// Filter + emit with parents: like k_emit_level (levelk.cu) and also records, for every new
// itemset, its two parents and its LOCAL count (this rank's playlists only), which sizes
// the rows of the next level.
__global__ void __launch_bounds__(256)
k_emit_level_px(const int32_t* __restrict__ items, int k, const int32_t* __restrict__ cand_p,
                const int32_t* __restrict__ cand_q, const int32_t* __restrict__ cand_cnt,
                const int32_t* __restrict__ cand_cnt_local, int64_t n_cand, int32_t thr,
                const int64_t* __restrict__ out_off, int32_t* __restrict__ out_items,
                int32_t* __restrict__ out_counts, int32_t* __restrict__ out_counts_local,
                int32_t* __restrict__ out_par_p, int32_t* __restrict__ out_par_q) {
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_cand;
       c += (int64_t)gridDim.x * blockDim.x) {
    const int32_t cnt = cand_cnt[c];
    if (cnt < thr) continue;
    const int64_t o = out_off[c];
    const int32_t p = cand_p[c], q = cand_q[c];
    const int32_t* ip = items + (int64_t)p * k;
    int32_t* op = out_items + o * (k + 1);
    for (int t = 0; t < k; ++t) op[t] = __ldg(ip + t);
    op[k] = __ldg(items + (int64_t)q * k + (k - 1));
    out_counts[o] = cnt;
    out_counts_local[o] = cand_cnt_local[c];
    out_par_p[o] = p;
    out_par_q[o] = q;
  }
}

// Columns are squeezed only while that shortens the work: a warp ANDs and popcounts up to
// kKeepWords words (one 128-bit load per lane) in a single step, so a row that is already
// that short is kept in its parent's column space (plain AND, no bit extract). The callers
// pass the limit (`keep_words`, default 256; 0 = always squeeze) to both kernels below.
constexpr int kKeepWordsDefault = 256;

// Row sizing: an itemset gets a row only if its class (same first parent) has at least two
// members - a lone member can never be joined again. Width: the parent's row width if that
// is already <= kKeepWords (the row stays in the parent's column space), else
// ceil(local count of the parent / 32) words, padded to 4 words (128-bit loads).
__global__ void __launch_bounds__(256)
k_row_alloc(const int32_t* __restrict__ par_p, int64_t n, const int32_t* __restrict__ parent_cnt_local,
            const int32_t* __restrict__ parent_roww, int keep_words, int32_t* __restrict__ roww) {
  for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < n;
       o += (int64_t)gridDim.x * blockDim.x) {
    const int32_t p = par_p[o];
    const bool shared = (o > 0 && par_p[o - 1] == p) || (o + 1 < n && par_p[o + 1] == p);
    const int32_t c = __ldg(parent_cnt_local + p);
    const int32_t w_old = parent_roww ? __ldg(parent_roww + p) : 0x7fffffff;
    const int32_t w = w_old <= keep_words ? w_old : ((((c + 31) >> 5) + 3) & ~3);
    roww[o] = shared ? w : 0;
  }
}

// Parallel bit extract of one word (Hacker's Delight 7-4 "compress"): the bits of x selected
// by m, packed towards bit 0.
__device__ __forceinline__ uint32_t pext32(uint32_t x, uint32_t m) {
  x &= m;
  uint32_t mk = ~m << 1;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    uint32_t mp = mk ^ (mk << 1);
    mp ^= mp << 2;
    mp ^= mp << 4;
    mp ^= mp << 8;
    mp ^= mp << 16;
    const uint32_t mv = mp & m;
    m = (m ^ mv) | (mv >> (1 << i));
    const uint32_t t = x & mv;
    x = (x ^ t) | (t >> (1 << i));
    mk &= ~mp;
  }
  return x;
}

// ---------------------------------------------------------------------------------------
// Projection: one warp per new itemset o. new_row(o) = row(par_q[o]) with the columns where
// row(par_p[o]) is 0 removed. 32 source words per iteration: per-lane pext, warp scan of the
// mask popcounts for the bit offsets, assembly of the output words in a small per-warp
// shared buffer (shared-memory atomicOr), coalesced store of the completed words.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_project_rows(const uint32_t* __restrict__ arena_old, const int64_t* __restrict__ rowptr_old,
               const int32_t* __restrict__ roww_old, const int32_t* __restrict__ par_p,
               const int32_t* __restrict__ par_q, int64_t n, uint32_t* __restrict__ arena_new,
               const int64_t* __restrict__ rowptr_new, const int32_t* __restrict__ roww_new,
               int keep_words) {
  __shared__ uint32_t sbuf[8][36];
  const int lane = threadIdx.x & 31;
  uint32_t* buf = sbuf[threadIdx.x >> 5];
  for (int64_t o = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); o < n;
       o += (int64_t)gridDim.x * 8) {
    const int w_new = __ldg(roww_new + o);
    if (w_new == 0) continue;  // warp-uniform
    const int32_t p = __ldg(par_p + o), q = __ldg(par_q + o);
    const int w_old = __ldg(roww_old + p);
    const uint32_t* msk = arena_old + __ldg(rowptr_old + p);
    const uint32_t* src = arena_old + __ldg(rowptr_old + q);
    uint32_t* dst = arena_new + __ldg(rowptr_new + o);
    if (w_old <= keep_words) {
      // short rows keep their column space: new row = row(p) AND row(q)   (w_new == w_old)
      const uint4* m4 = reinterpret_cast<const uint4*>(msk);
      const uint4* s4 = reinterpret_cast<const uint4*>(src);
      uint4* d4 = reinterpret_cast<uint4*>(dst);
      for (int v = lane; v < (w_old >> 2); v += 32) d4[v] = and4(ldg4(m4 + v), ldg4(s4 + v));
      continue;
    }
    for (int i = lane; i < 36; i += 32) buf[i] = 0;
    __syncwarp();
    int base = 0;  // bits emitted so far
    for (int blk = 0; blk < w_old; blk += 32) {
      const int w = blk + lane;
      const uint32_t m = w < w_old ? __ldg(msk + w) : 0u;
      const uint32_t x = w < w_old ? __ldg(src + w) : 0u;
      const int nb = __popc(m);
      int incl = nb;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += u;
      }
      const int tot = __shfl_sync(0xffffffffu, incl, 31);
      if (nb) {
        const uint32_t c = pext32(x, m);
        const int bitpos = (base & 31) + incl - nb;
        const int word = bitpos >> 5, sh = bitpos & 31;
        if (c) {
          atomicOr(buf + word, c << sh);
          if (sh + nb > 32) atomicOr(buf + word + 1, c >> (32 - sh));
        }
      }
      __syncwarp();
      const int nbits = (base & 31) + tot, full = nbits >> 5, w0 = base >> 5;
      for (int i = lane; i < full; i += 32)
        if (w0 + i < w_new) dst[w0 + i] = buf[i];
      const uint32_t partial = buf[full];
      __syncwarp();
      for (int i = lane; i < 36; i += 32) buf[i] = 0;
      __syncwarp();
      if (lane == 0) buf[0] = partial;
      __syncwarp();
      base += tot;
    }
    int done = base >> 5;
    if ((base & 31) && lane == 0 && done < w_new) dst[done] = buf[0];
    if (base & 31) ++done;
    for (int i = done + lane; i < w_new; i += 32) dst[i] = 0u;
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------
// First projection (level 1 -> level 2) by gathering. The masks there are the single-track
// rows, on average a few percent dense, so scanning all N/32 mask words per new row wastes
// almost all of the work. Instead each kept track's row is turned ONCE into the ascending
// list of the playlists that contain it (k_row_tidlists); a new row (a, b) is then
// row_b sampled at those positions: one gathered load + one ballot per 32 output bits.
// ---------------------------------------------------------------------------------------
// One CTA per row, one eighth of the row per warp: first the warp's number of set bits (128-bit
// loads, no block barrier inside the loop), one barrier to turn the eight totals into offsets,
// then the warp walks its part again (from L2) 128 words at a time - shuffle scan of the lanes'
// popcounts, every lane writes the positions of its four words' bits.
__global__ void __launch_bounds__(256)
k_row_tidlists(const uint32_t* __restrict__ bitmap, int64_t row_words,
               const int64_t* __restrict__ tid_off, int32_t* __restrict__ tids) {
  __shared__ int warp_tot[8];
  if (tid_off[blockIdx.x + 1] == tid_off[blockIdx.x]) return;   // no list wanted for this row (block-uniform)
  const uint32_t* row = bitmap + (int64_t)blockIdx.x * row_words;
  int32_t* out = tids + tid_off[blockIdx.x];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t part = (((row_words + 7) >> 3) + 127) & ~(int64_t)127;
  const int64_t w_lo = min(row_words, warp * part), w_hi = min(row_words, w_lo + part);
  int cnt = 0;
  for (int64_t w = w_lo + 4 * lane; w < w_hi; w += 128) cnt += popc4(ldg4(reinterpret_cast<const uint4*>(row + w)));
  cnt = warp_sum(cnt);
  if (lane == 0) warp_tot[warp] = cnt;
  __syncthreads();
  int pos = 0;
  for (int i = 0; i < warp; ++i) pos += warp_tot[i];
  for (int64_t w = w_lo; w < w_hi; w += 128) {
    const int64_t ww = w + 4 * lane;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (ww < w_hi) v = ldg4(reinterpret_cast<const uint4*>(row + ww));
    const int nb = popc4(v);
    int incl = nb;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += u;
    }
    int p = pos + incl - nb;
    const uint32_t m4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      uint32_t m = m4[t];
      const int32_t colbase = (int32_t)((ww + t) << 5);
      while (m) {
        out[p++] = colbase + (__ffs(m) - 1);
        m &= m - 1;
      }
    }
    pos += __shfl_sync(0xffffffffu, incl, 31);
  }
}

// Work unit = 256 consecutive words of the new arena (rows are 4-word multiples, so a group
// of 4 words never straddles two rows): every warp does the same amount of work no matter
// how long the rows are, and 4 independent gathers are in flight per lane.
// 32-word halves per work unit (A/B at Z1M: 2 / 4 / 8 / 16 halves = 2.12 / 1.88 / 1.78 / 1.71 ms - a unit starts
// with a 17-step binary search for its first row)
#ifndef KMLS_PG_HALVES
#define KMLS_PG_HALVES 8
#endif
constexpr int kPgWords = 32 * KMLS_PG_HALVES;   // words of the new arena per work unit
__global__ void __launch_bounds__(256)
k_project_gather(const uint32_t* __restrict__ bitmap, int64_t row_words,
                 const int32_t* __restrict__ par_p, const int32_t* __restrict__ par_q, int64_t n,
                 const int32_t* __restrict__ tids, const int64_t* __restrict__ tid_off,
                 uint32_t* __restrict__ arena_new, const int64_t* __restrict__ rowptr_new,
                 int64_t total_words) {
  const int lane = threadIdx.x & 31;
  const int64_t n_blocks = (total_words + kPgWords - 1) / kPgWords;
  for (int64_t blk = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); blk < n_blocks;
       blk += (int64_t)gridDim.x * 8) {
    const int64_t g0 = blk * kPgWords;
    // first row o with rowptr[o + 1] > g0 (rows without storage are skipped by construction)
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (__ldg(rowptr_new + mid + 1) > g0) hi = mid; else lo = mid + 1;
    }
    int64_t o = lo, rbeg = __ldg(rowptr_new + o), rend = __ldg(rowptr_new + o + 1);
    bool fresh = true;
    const int32_t* tl = nullptr;
    const uint32_t* src = nullptr;
    int c = 0;
#pragma unroll 1
    for (int half = 0; half < KMLS_PG_HALVES; ++half) {
      uint32_t mine = 0;
      for (int i = 0; i < 32; i += 4) {
        const int64_t g = g0 + half * 32 + i;
        if (g >= total_words) break;
        while (g >= rend) { ++o; rbeg = rend; rend = __ldg(rowptr_new + o + 1); fresh = true; }
        if (fresh) {
          const int32_t a = __ldg(par_p + o), b = __ldg(par_q + o);
          const int64_t t_begin = __ldg(tid_off + a);
          c = (int)(__ldg(tid_off + a + 1) - t_begin);
          tl = tids + t_begin;
          src = bitmap + (int64_t)b * row_words;
          fresh = false;
        }
        const int t0 = (int)(g - rbeg) * 32 + lane;
        int32_t pos[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) pos[j] = (t0 + 32 * j) < c ? __ldg(tl + t0 + 32 * j) : -1;
        uint32_t bit[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          bit[j] = pos[j] >= 0 ? (__ldg(src + (pos[j] >> 5)) >> (pos[j] & 31)) & 1u : 0u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t word = __ballot_sync(0xffffffffu, bit[j] != 0);
          if (lane == i + j) mine = word;
        }
      }
      const int64_t gw = g0 + half * 32 + lane;
      if (gw < total_words) arena_new[gw] = mine;
    }
  }
}

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_count_pairs(const uint32_t* d_arena, const int64_t* d_rowptr,
                                const int32_t* d_roww, const int32_t* d_cand_p,
                                const int32_t* d_cand_q, int64_t n_cand, int32_t* d_cand_cnt,
                                kmls_stream_t stream) {
  KMLS_REQUIRE(n_cand >= 0, "kmls_count_pairs: negative size");
  if (n_cand == 0) return 0;
  // d_arena may be NULL when no row of the level was materialised (all widths are 0)
  KMLS_REQUIRE(d_rowptr && d_roww && d_cand_p && d_cand_q && d_cand_cnt,
               "kmls_count_pairs: null pointer");
  int grid = (int)std::min<int64_t>((n_cand + 8 * kCountIlp - 1) / (8 * kCountIlp), (int64_t)num_sms() * 64);
  k_count_pairs<<<grid, 256, 0, (cudaStream_t)stream>>>(d_arena, d_rowptr, d_roww, d_cand_p,
                                                        d_cand_q, n_cand, d_cand_cnt);
  KMLS_LAUNCHED("k_count_pairs");
  return 0;
}

extern "C" int kmls_emit_level_px(const int32_t* d_items, int32_t k, const int32_t* d_cand_p,
                                  const int32_t* d_cand_q, const int32_t* d_cand_cnt,
                                  const int32_t* d_cand_cnt_local, int64_t n_cand, int32_t thr,
                                  const int64_t* d_out_off, int32_t* d_out_items,
                                  int32_t* d_out_counts, int32_t* d_out_counts_local,
                                  int32_t* d_out_par_p, int32_t* d_out_par_q,
                                  kmls_stream_t stream) {
  KMLS_REQUIRE(n_cand >= 0 && k >= 1, "kmls_emit_level_px: bad arguments");
  if (n_cand == 0) return 0;
  KMLS_REQUIRE(d_out_off && d_out_items && d_out_counts && d_out_counts_local && d_out_par_p &&
                   d_out_par_q && d_cand_cnt_local,
               "kmls_emit_level_px: null pointer");
  int grid = (int)std::min<int64_t>((n_cand + 255) / 256, (int64_t)num_sms() * 16);
  k_emit_level_px<<<grid, 256, 0, (cudaStream_t)stream>>>(
      d_items, k, d_cand_p, d_cand_q, d_cand_cnt, d_cand_cnt_local, n_cand, thr, d_out_off,
      d_out_items, d_out_counts, d_out_counts_local, d_out_par_p, d_out_par_q);
  KMLS_LAUNCHED("k_emit_level_px");
  return 0;
}

extern "C" int kmls_row_alloc(const int32_t* d_par_p, int64_t n, const int32_t* d_parent_cnt_local,
                              const int32_t* d_parent_roww, int32_t keep_words, int32_t* d_roww,
                              kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0, "kmls_row_alloc: negative size");
  if (n == 0) return 0;
  KMLS_REQUIRE(d_par_p && d_parent_cnt_local && d_roww, "kmls_row_alloc: null pointer");
  int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)num_sms() * 16);
  if (keep_words < 0) keep_words = kKeepWordsDefault;
  k_row_alloc<<<grid, 256, 0, (cudaStream_t)stream>>>(d_par_p, n, d_parent_cnt_local, d_parent_roww,
                                                      keep_words, d_roww);
  KMLS_LAUNCHED("k_row_alloc");
  return 0;
}

extern "C" int kmls_project_rows(const uint32_t* d_arena_old, const int64_t* d_rowptr_old,
                                 const int32_t* d_roww_old, const int32_t* d_par_p,
                                 const int32_t* d_par_q, int64_t n, uint32_t* d_arena_new,
                                 const int64_t* d_rowptr_new, const int32_t* d_roww_new,
                                 int32_t keep_words, kmls_stream_t stream) {
  KMLS_REQUIRE(n >= 0, "kmls_project_rows: negative size");
  if (n == 0) return 0;
  // an arena may be NULL when it holds no rows (all widths 0); rows of width 0 are skipped
  KMLS_REQUIRE(d_rowptr_old && d_roww_old && d_par_p && d_par_q && d_rowptr_new && d_roww_new,
               "kmls_project_rows: null pointer");
  KMLS_REQUIRE(d_arena_new == nullptr || (const void*)d_arena_old != (const void*)d_arena_new,
               "kmls_project_rows: source and destination arenas alias");
  if (d_arena_new == nullptr) return 0;
  int grid = (int)std::min<int64_t>((n + 7) / 8, (int64_t)num_sms() * 64);
  if (keep_words < 0) keep_words = kKeepWordsDefault;
  k_project_rows<<<grid, 256, 0, (cudaStream_t)stream>>>(d_arena_old, d_rowptr_old, d_roww_old,
                                                         d_par_p, d_par_q, n, d_arena_new,
                                                         d_rowptr_new, d_roww_new, keep_words);
  KMLS_LAUNCHED("k_project_rows");
  return 0;
}


extern "C" int kmls_row_tidlists(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                                 const int64_t* d_tid_off, int32_t* d_tids, kmls_stream_t stream) {
  using namespace kmls;
  KMLS_REQUIRE(n_rows >= 0 && row_words >= 0, "kmls_row_tidlists: negative size");
  if (n_rows == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_tid_off && d_tids, "kmls_row_tidlists: null pointer");
  KMLS_REQUIRE((row_words & 3) == 0 && ((uintptr_t)d_bitmap & 15) == 0,
               "kmls_row_tidlists: rows must be 16-byte aligned (row_words a multiple of 4, as kmls_bitmap_row_words gives)");
  k_row_tidlists<<<n_rows, 256, 0, (cudaStream_t)stream>>>(d_bitmap, row_words, d_tid_off, d_tids);
  KMLS_LAUNCHED("k_row_tidlists");
  return 0;
}

extern "C" int kmls_project_rows_gather(const uint32_t* d_bitmap, int64_t row_words,
                                        const int32_t* d_par_p, const int32_t* d_par_q, int64_t n,
                                        const int32_t* d_tids, const int64_t* d_tid_off,
                                        uint32_t* d_arena_new, const int64_t* d_rowptr_new,
                                        int64_t total_words, kmls_stream_t stream) {
  using namespace kmls;
  KMLS_REQUIRE(n >= 0 && total_words >= 0, "kmls_project_rows_gather: negative size");
  if (n == 0 || total_words == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_par_p && d_par_q && d_tids && d_tid_off && d_rowptr_new && d_arena_new,
               "kmls_project_rows_gather: null pointer");
  const int64_t n_blocks = (total_words + kPgWords - 1) / kPgWords;
  int grid = (int)std::min<int64_t>((n_blocks + 7) / 8, (int64_t)num_sms() * 64);
  k_project_gather<<<grid, 256, 0, (cudaStream_t)stream>>>(d_bitmap, row_words, d_par_p, d_par_q, n,
                                                           d_tids, d_tid_off, d_arena_new,
                                                           d_rowptr_new, total_words);
  KMLS_LAUNCHED("k_project_gather");
  return 0;
}
