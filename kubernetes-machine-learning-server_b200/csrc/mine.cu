// kmls_mine: the whole mining path behind one C call.
//
// Stands where the reference calls mlxtend (machine-learning/main.py:267-272:
// TransactionEncoder().fit().transform() + fpgrowth(df, min_support, use_colnames=True)) and
// produces the result set of SURVEY.md section 8c item 5, level by level:
//
//   level 1   item histogram -> select + order the frequent tracks -> bit-packed vertical
//             bitmap (k_encode_bitmap) -> exact supports (k_row_popcount)
//   level 2   pair counts (per playlist, CUDA-core or tcgen05 int8 GEMM) -> frequent pairs
//             -> conditional rows of the pairs (k_row_tidlists + k_project_gather)
//   level k   k_join_size -> k_count_project -> k_emit_rows          (until a level is empty)
//
// Levels >= 3 are where this file differs from the kernel-by-kernel entry points the Python
// loop of round 1 drove: the support of a candidate (P, x_i, x_j) is the popcount of
// row(P,x_i) AND row(P,x_j), and the conditional row of a SURVIVING candidate is that same
// AND (or its bit-extract). k_count_project forms it once: a warp owns a parent p = (P,x_i),
// keeps row(p) in registers, streams the rows of its later siblings through, and stores the
// AND of every candidate that reaches the threshold straight into the slot reserved for it
// in the next level's arena. Slots are sized before counting (all candidates of a parent have
// the same row width), so there is no second pass over the rows and no compaction of them:
// the new level table simply points at the slots of the survivors.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <cmath>
#include <cstring>

#include "common.cuh"
#include "kmls_b200.h"
#include "profile.cuh"
#include "workspace.cuh"

namespace kmls {

volatile int64_t* mailbox() {
  static thread_local int64_t* box = nullptr;
  if (box == nullptr) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, 64 * sizeof(int64_t), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    box = (int64_t*)p;
  }
  return box;
}

// up to four device words -> mailbox (null sources are skipped)
__global__ void k_post(const int64_t* a, const int64_t* b, const int32_t* c, const int32_t* d,
                       volatile int64_t* dst) {
  if (a) dst[0] = *a;
  if (b) dst[1] = *b;
  if (c) dst[2] = *c;
  if (d) dst[3] = *d;
  __threadfence_system();
}

// per parent: a = candidates, b = slot words, c = work units of the AND kernel (prefix-summed)
// (16 bytes: candidates and units of one level stay below 2^31 - k_join_size sums the
// candidates in 64 bits as well and the host refuses a level that would exceed that)
struct __align__(16) I2 {
  long long b;
  int a, c;
};
struct I2Sum {
  __host__ __device__ I2 operator()(const I2& x, const I2& y) const { return I2{x.b + y.b, x.a + y.a, x.c + y.c}; }
};

constexpr int kHotRows = 1024;      // level 2 hybrid: rows of the block counted on the tensor cores

// ---- level 1 helpers --------------------------------------------------------------------
struct OccAtLeast {
  const int32_t* occ;
  int32_t thr;
  __host__ __device__ bool operator()(int32_t i) const { return occ[i] >= thr; }
};

__global__ void __launch_bounds__(256)
k_gather_i32(const int32_t* __restrict__ src, const int32_t* __restrict__ idx, int64_t n,
             int32_t* __restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = src[idx[i]];
}

__global__ void __launch_bounds__(256) k_fill_i32(int32_t* __restrict__ dst, int64_t n, int32_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = v;
}

// row_of_item[cand[r]] = r; level-1 table: items1[r] = r
__global__ void __launch_bounds__(256)
k_number_rows(const int32_t* __restrict__ cand, int32_t n_rows, int32_t* __restrict__ row_of_item,
              int32_t* __restrict__ items1) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
    row_of_item[cand[r]] = r;
    if (items1) items1[r] = r;
  }
}

// flags[0] += 1 for every row whose exact support differs from the occurrence count it was
// ordered by; flags[1] (64-bit, at flags + 2) += the supports (mean kept tracks per playlist);
// 64-bit at flags + 8 += the supports of the rows from hot0 on (the most frequent tracks)
__global__ void __launch_bounds__(256)
k_check_counts(const int32_t* __restrict__ counts, const int32_t* __restrict__ keys, int32_t n_rows,
               int32_t hot0, int32_t* __restrict__ flags) {
  long long sum = 0, hot = 0;
  int bad = 0;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
    bad += counts[r] != keys[r];
    sum += counts[r];
    if (r >= hot0) hot += counts[r];
  }
  if (bad) atomicAdd(flags, bad);
  if (sum) atomicAdd((unsigned long long*)(flags + 2), (unsigned long long)sum);
  if (hot) atomicAdd((unsigned long long*)(flags + 8), (unsigned long long)hot);
}

// mlxtend's root single-path shortcut: the root FP-tree is one path iff
// the supports are nested, i.e. every pair count equals the rarer member's support (rows are
// in ascending-support order, so the rarer member of (i, j), i < j, is i). flag += violations.
__global__ void __launch_bounds__(256)
k_nested_check(const int32_t* __restrict__ pair, const int32_t* __restrict__ counts1, int32_t n_rows,
               int32_t* __restrict__ flag) {
  int bad = 0;
  const int64_t n2 = (int64_t)n_rows * n_rows;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n2; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(e / n_rows), j = (int)(e - (int64_t)i * n_rows);
    if (j > i && pair[e] != counts1[i]) ++bad;
  }
  if (bad) atomicAdd(flag, bad);
}

// the two members of every pair as separate columns (the first is the "parent" column of level 2)
__global__ void __launch_bounds__(256)
k_pair_parents(const int32_t* __restrict__ items2, int64_t n2, int32_t* __restrict__ par2,
               int32_t* __restrict__ second2) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) {
    par2[i] = items2[2 * i];
    second2[i] = items2[2 * i + 1];
  }
}

// Class partition: which conditional rows of the pairs does THIS part need? A pair's row is
// needed if the pair is one of the part's own (blocks part, part + n_parts, ... of `block` pairs)
// or a later sibling of an own pair - i.e. if some own pair lies between the start of its class
// and itself. Rows nobody here will read get width 0; need_row[a] marks the tracks whose
// playlist lists the remaining rows are sampled at.
__global__ void __launch_bounds__(256)
k_level2_needs(const int32_t* __restrict__ par, int64_t n2, int block, int n_parts, int part,
               int32_t* __restrict__ roww, int32_t* __restrict__ need_row) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x) {
    if (roww[i] == 0) continue;
    const int32_t a = __ldg(par + i);
    int64_t hi = i, step = 1;                      // par[hi] == a; gallop back to the class start
    while (hi - step >= 0 && __ldg(par + hi - step) == a) {
      hi -= step;
      step <<= 1;
    }
    int64_t lo = max(hi - step, (int64_t)-1);      // par[lo] != a or lo == -1
    while (hi - lo > 1) {
      const int64_t mid = (lo + hi) >> 1;
      if (__ldg(par + mid) == a) hi = mid; else lo = mid;
    }
    const int64_t cs = hi;                         // first pair of the class
    bool needed = false;
    for (int64_t b = cs / block; b <= i / block && !needed; ++b) needed = (int)(b % n_parts) == part;
    if (needed) need_row[a] = 1;
    else roww[i] = 0;
  }
}

__global__ void __launch_bounds__(256)
k_mask_counts(const int32_t* __restrict__ counts, const int32_t* __restrict__ need, int32_t n,
              int32_t* __restrict__ out) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x)
    out[r] = need[r] ? counts[r] : 0;
}

// Bit-extract route: a row wider than this is squeezed in SEGMENTS of this many words, one work
// unit each, so that one very wide row (a very popular prefix: hundreds of thousands of words at
// 10M playlists) is not a single warp's serial job. Multiple of 32.
constexpr int kSegWords = 8192;
constexpr int kUnitCands = 4;   // candidates of one parent per work unit of the AND kernel
__host__ __device__ __forceinline__ int seg_count(int w_old) { return (w_old + kSegWords - 1) / kSegWords; }

// ---- levels >= 3 --------------------------------------------------------------------------
// Join sizing, one thread per itemset p of level k. Its candidates are its later siblings
// (same parent; the table is sorted, so they follow it). ncand[p] = their number, wnew[p] =
// width in words of the conditional rows of p's children: p's own width when that is at
// most keep_words (the children stay in p's column space: AND), else ceil(count(p) / 32)
// padded to 4 words (columns where row(p) is 0 are squeezed out: bit extract).
// `n_parts > 1`: only itemsets of blocks part, part + n_parts, ... get candidates (level 2
// of a class-partitioned job; deeper levels hold only owned subtrees anyway).
__global__ void __launch_bounds__(256)
k_join_size(const int32_t* __restrict__ par, const int32_t* __restrict__ roww,
            const int32_t* __restrict__ counts, int64_t n, int keep_words, int squeeze_pct, int n_parts, int part,
            int part_block, int32_t* __restrict__ ncand, int32_t* __restrict__ wnew,
            I2* __restrict__ tuple, int32_t* __restrict__ n_wide) {
  int wide = 0;
  unsigned long long cands = 0;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const int32_t w_old = roww[p];
    int32_t m = 0, w = 0;
    const bool mine = n_parts <= 1 || (int)((p / part_block) % n_parts) == part;
    if (w_old > 0 && mine) {
      const int32_t cls = __ldg(par + p);
      // end of the class: gallop forward, then bisect the last stride
      int64_t lo = p, step = 1;              // par[lo] == cls
      while (lo + step < n && __ldg(par + lo + step) == cls) {
        lo += step;
        step <<= 1;
      }
      int64_t hi = min(lo + step, n);        // par[hi] != cls or hi == n
      while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(par + mid) == cls) lo = mid; else hi = mid;
      }
      m = (int32_t)(lo - p);
      // squeeze only rows that are wide AND shrink enough to pay for the bit extract: a dense
      // wide row (a very popular prefix) stays in its column space and takes the AND route
      const int32_t ws = (((__ldg(counts + p) + 31) >> 5) + 3) & ~3;
      const bool squeeze = w_old > keep_words && (int64_t)ws * 100 <= (int64_t)w_old * squeeze_pct;
      w = squeeze ? ws : w_old;
      if (m > 0 && squeeze) wide += seg_count(w_old);   // one work unit per row segment
    }
    ncand[p] = m;
    wnew[p] = w;
    // what the scan sums: slot words, candidates, work units of the AND route
    tuple[p] = I2{(long long)m * (long long)w, m, (m > 0 && w == w_old) ? (m + kUnitCands - 1) / kUnitCands : 0};
    cands += (unsigned long long)m;
  }
  // one atomic per warp, not per thread
  for (int o = 16; o > 0; o >>= 1) {
    wide += __shfl_xor_sync(0xffffffffu, wide, o);
    cands += __shfl_xor_sync(0xffffffffu, cands, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (wide) atomicAdd(n_wide, wide);   // work units (parent, chunk) of the bit-extract route
    if (cands) atomicAdd(reinterpret_cast<unsigned long long*>(n_wide + 1), cands);
  }
}


struct CandSlots {
  const int32_t* ncand;
  const int32_t* wnew;
  const int32_t* roww;
  __host__ __device__ I2 operator()(int64_t p) const {
    const int m = ncand[p];
    const int units = (m > 0 && roww[p] == wnew[p]) ? (m + kUnitCands - 1) / kUnitCands : 0;   // AND route
    return I2{(long long)m * (long long)wnew[p], m, units};
  }
};

__global__ void k_i2_total(CandSlots f, int64_t n, I2* off2, const int32_t* n_wide, volatile int64_t* box) {
  I2 t{0, 0, 0};
  if (n > 0) t = I2Sum()(off2[n - 1], f(n - 1));
  off2[n] = t;
  box[0] = t.a;
  box[1] = t.b;
  box[2] = t.c;
  box[3] = n_wide[0];
  box[4] = *reinterpret_cast<const long long*>(n_wide + 1);   // candidates of the level, summed in 64 bits
  __threadfence_system();
}

__device__ __forceinline__ uint4 or4z() { return make_uint4(0u, 0u, 0u, 0u); }

// ---------------------------------------------------------------------------------------
// AND route as a list of work units. A unit = up to kUnitCands consecutive candidates of ONE
// parent whose row is at most keep_words wide; k_expand_units writes, per unit, everything
// the counting kernel needs as one 64-byte record, so that its dependent-load chain is
// record -> rows (two latencies) instead of ids -> offsets -> pointers -> rows, and every
// unit it touches has work (parents without candidates produce no unit).
// ---------------------------------------------------------------------------------------
struct __align__(16) Unit {
  const uint32_t* rowp;             // row of the parent
  uint32_t* dst;                    // slot of the unit's first candidate in the new arena
  const uint32_t* rowq[kUnitCands]; // rows of the candidates (siblings p + 1 + i0 ...)
  int32_t p;                        // parent
  int32_t cand0;                    // number of the first candidate in the level's candidate list
  int32_t mc_nv;                    // candidates in this unit << 29 | row width in 128-bit words
  int32_t whole;                    // 1 = the unit holds ALL candidates of its parent
};
static_assert(sizeof(Unit) == 64, "Unit must be 64 bytes");

__global__ void __launch_bounds__(256)
k_expand_units(const uint32_t* __restrict__ arena_old, const int64_t* __restrict__ rowptr_old,
               const int32_t* __restrict__ roww_old, const int32_t* __restrict__ ncand,
               const int32_t* __restrict__ wnew, const I2* __restrict__ off2, int64_t n,
               uint32_t* __restrict__ arena_new,
               Unit* __restrict__ units, int2* __restrict__ wide_list, int32_t* __restrict__ wide_fill) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const int m = ncand[p];
    const int w = roww_old[p];
    if (m == 0) continue;
    if (wnew[p] != w) {         // bit-extract route: k_count_project walks this list (order is irrelevant)
      const int segs = seg_count(w);
      int at = atomicAdd(wide_fill, segs);
      for (int sg = 0; sg < segs; ++sg) wide_list[at++] = make_int2((int32_t)p, sg);
      continue;
    }
    const I2 o = off2[p];
    const uint32_t* rowp = arena_old + rowptr_old[p];
    for (int i0 = 0, u = 0; i0 < m; i0 += kUnitCands, ++u) {
      Unit un;
      un.rowp = rowp;
      un.dst = arena_new + o.b + (int64_t)i0 * w;
      const int mc = min(kUnitCands, m - i0);
#pragma unroll
      for (int t = 0; t < kUnitCands; ++t)
        un.rowq[t] = t < mc ? arena_old + __ldg(rowptr_old + p + 1 + i0 + t) : rowp;
      un.p = (int32_t)p;
      un.cand0 = (int32_t)(o.a + i0);
      un.mc_nv = (int32_t)(((uint32_t)mc << 29) | (uint32_t)(w >> 2));
      un.whole = (m <= kUnitCands) ? 1 : 0;
      uint4* d = reinterpret_cast<uint4*>(units + o.c + u);
      const uint4* sv = reinterpret_cast<const uint4*>(&un);
      d[0] = sv[0];
      d[1] = sv[1];
      d[2] = sv[2];
      d[3] = sv[3];
    }
  }
}

// One warp per unit: row(p) and the unit's candidate rows are all in flight at once; AND,
// popcount, warp reduction; the AND of a candidate that reaches the threshold is stored from
// the registers it is in (rows of more than 128 words: the second half is formed again from
// the L1-resident rows). When the unit holds all candidates of its parent and exactly one
// survives, nothing is stored: an only child is never joined again (k_emit_rows gives it
// width 0).
// (KMLS_AND_MINB / KMLS_CP_MINB: minimum resident CTAs per SM asked of ptxas, for A/B builds with
// tools/build_variant.sh; unset = ptxas' own choice, the measured best)
#ifdef KMLS_AND_MINB
__global__ void __launch_bounds__(256, KMLS_AND_MINB)
#else
__global__ void __launch_bounds__(256)
#endif
k_count_and(const Unit* __restrict__ units, int64_t n_units, int32_t thr, int32_t* __restrict__ cand_cnt,
            int32_t* __restrict__ cand_p) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * 8;
  for (int64_t ui = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); ui < n_units; ui += stride) {
    const uint4* u4 = reinterpret_cast<const uint4*>(units + ui);
    const uint4 d0 = __ldg(u4), d1 = __ldg(u4 + 1), d2 = __ldg(u4 + 2), d3 = __ldg(u4 + 3);
    const uint4* rowp = reinterpret_cast<const uint4*>(((uint64_t)d0.y << 32) | d0.x);
    uint4* dst = reinterpret_cast<uint4*>(((uint64_t)d0.w << 32) | d0.z);
    const uint4* rq[kUnitCands] = {reinterpret_cast<const uint4*>(((uint64_t)d1.y << 32) | d1.x),
                                   reinterpret_cast<const uint4*>(((uint64_t)d1.w << 32) | d1.z),
                                   reinterpret_cast<const uint4*>(((uint64_t)d2.y << 32) | d2.x),
                                   reinterpret_cast<const uint4*>(((uint64_t)d2.w << 32) | d2.z)};
    const int mc = (int)(d3.z >> 29), nv = (int)(d3.z & 0x1fffffffu);
    const bool h0 = lane < nv;
    const uint4 m0 = h0 ? ldg4(rowp + lane) : or4z();
    uint4 r0[kUnitCands];
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) r0[t] = (h0 && t < mc) ? ldg4(rq[t] + lane) : or4z();
    int cnt[kUnitCands];
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) {
      r0[t] = and4(r0[t], m0);
      cnt[t] = popc4(r0[t]);
    }
    for (int seg = 32; seg < nv; seg += 32) {    // rows wider than 128 words: warp-uniform trip count
      if (lane + seg < nv) {
        const uint4 m1 = ldg4(rowp + lane + seg);
#pragma unroll
        for (int t = 0; t < kUnitCands; ++t)
          if (t < mc) cnt[t] += popc4(and4(ldg4(rq[t] + lane + seg), m1));
      }
    }
    uint32_t surv = 0;
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) {
      cnt[t] = warp_sum(cnt[t]);
      if (t < mc && cnt[t] >= thr) surv |= 1u << t;
    }
    if (lane < mc) {
      int c = cnt[0];
#pragma unroll
      for (int t = 1; t < kUnitCands; ++t)
        if (lane == t) c = cnt[t];
      cand_cnt[d3.y + lane] = c;
      cand_p[d3.y + lane] = (int32_t)d3.x;
    }
    if (surv == 0 || (d3.w && (surv & (surv - 1)) == 0)) continue;   // nobody, or an only child
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) {
      if ((surv >> t) & 1u) {                    // warp-uniform
        uint4* d = dst + (size_t)t * nv;
        if (h0) d[lane] = r0[t];
        for (int seg = 32; seg < nv; seg += 32)   // formed again from the (L1/L2-resident) rows
          if (lane + seg < nv) d[lane + seg] = and4(ldg4(rq[t] + lane + seg), ldg4(rowp + lane + seg));
      }
    }
  }
}

// The same with HALF a warp per unit, for the deep levels where a unit holds one or two
// candidates: a warp per unit then has two or three rows in flight and levels 8-10 run at
// 3.3-4.8 TB/s where levels 4-6 reach 6-6.8. Two units per warp double the rows in flight per
// resident warp; rows wider than 64 words take several passes of 16 lanes (their tail is formed
// again at store time, as above).
__global__ void __launch_bounds__(256)
k_count_and_half(const Unit* __restrict__ units, int64_t n_units, int32_t thr, int32_t* __restrict__ cand_cnt,
                 int32_t* __restrict__ cand_p) {
  const int lane = threadIdx.x & 31, l = lane & 15;
  const int64_t stride = (int64_t)gridDim.x * 16;
  const int64_t n_iter = (n_units + stride - 1) / stride;
  int64_t ui = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + (lane >> 4);
  for (int64_t it = 0; it < n_iter; ++it, ui += stride) {     // warp-uniform trip count (full-mask shuffles inside)
    const bool live = ui < n_units;
    uint4 d0 = or4z(), d1 = or4z(), d2 = or4z(), d3 = or4z();
    if (live) {
      const uint4* u4 = reinterpret_cast<const uint4*>(units + ui);
      d0 = __ldg(u4);
      d1 = __ldg(u4 + 1);
      d2 = __ldg(u4 + 2);
      d3 = __ldg(u4 + 3);
    }
    const uint4* rowp = reinterpret_cast<const uint4*>(((uint64_t)d0.y << 32) | d0.x);
    uint4* dst = reinterpret_cast<uint4*>(((uint64_t)d0.w << 32) | d0.z);
    const uint4* rq[kUnitCands] = {reinterpret_cast<const uint4*>(((uint64_t)d1.y << 32) | d1.x),
                                   reinterpret_cast<const uint4*>(((uint64_t)d1.w << 32) | d1.z),
                                   reinterpret_cast<const uint4*>(((uint64_t)d2.y << 32) | d2.x),
                                   reinterpret_cast<const uint4*>(((uint64_t)d2.w << 32) | d2.z)};
    const int mc = (int)(d3.z >> 29), nv = (int)(d3.z & 0x1fffffffu);   // 0, 0 for a dead half
    const bool h0 = l < nv;
    const uint4 m0 = h0 ? ldg4(rowp + l) : or4z();
    uint4 r0[kUnitCands];
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) r0[t] = (h0 && t < mc) ? ldg4(rq[t] + l) : or4z();
    int cnt[kUnitCands];
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) {
      r0[t] = and4(r0[t], m0);
      cnt[t] = popc4(r0[t]);
    }
    for (int seg = 16; seg < nv; seg += 16) {          // rows wider than 64 words (trip count differs per half: no shuffles inside)
      if (l + seg < nv) {
        const uint4 m1 = ldg4(rowp + l + seg);
#pragma unroll
        for (int t = 0; t < kUnitCands; ++t)
          if (t < mc) cnt[t] += popc4(and4(ldg4(rq[t] + l + seg), m1));
      }
    }
    uint32_t surv = 0;
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) {
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) cnt[t] += __shfl_xor_sync(0xffffffffu, cnt[t], o);   // inside the half
      if (t < mc && cnt[t] >= thr) surv |= 1u << t;
    }
    if (l < mc) {
      int c = cnt[0];
#pragma unroll
      for (int t = 1; t < kUnitCands; ++t)
        if (l == t) c = cnt[t];
      cand_cnt[d3.y + l] = c;
      cand_p[d3.y + l] = (int32_t)d3.x;
    }
    if (surv == 0 || (d3.w && (surv & (surv - 1)) == 0)) continue;   // nobody, or an only child (no shuffles follow)
#pragma unroll
    for (int t = 0; t < kUnitCands; ++t) {
      if ((surv >> t) & 1u) {                          // uniform inside the half
        uint4* d = dst + (size_t)t * nv;
        if (h0) d[l] = r0[t];
        for (int seg = 16; seg < nv; seg += 16)
          if (l + seg < nv) d[l + seg] = and4(ldg4(rq[t] + l + seg), ldg4(rowp + l + seg));
      }
    }
  }
}

constexpr int kAndIlp = 4;   // candidates of a parent in flight per warp (AND route)
constexpr int kChunk = 8;    // candidates of a parent handled together (bit-extract route)
static_assert(kChunk == 8, "k_join_size and k_expand_units count chunks of 8");
constexpr int kKeepWordsMax = 1024;     // widest row the AND route takes
constexpr int kKeepWordsDefaultMine = 256;

// Count + project for parents whose rows are WIDER than keep_words, one warp per parent p (see
// the file header). Candidate i of p is its i-th later sibling q = p + 1 + i; its number in the
// level's candidate list is off2[p].a + i and its slot in the new arena starts at word
// off2[p].b + i * wnew[p]. Per chunk of up to 8 candidates, first the popcounts of row(p) AND
// row(q) (128-bit loads, all rows of the chunk per mask block), then, for the survivors only,
// new_row = row(q) with the columns where row(p) is 0 squeezed out: 32 mask words per step, the
// compress network's move masks (Hacker's Delight 7-4) derived ONCE per mask word for all
// survivors of the chunk, output words assembled in shared memory. (Rows of at most keep_words
// words take k_expand_units + k_count_and.)
#ifdef KMLS_CP_MINB
__global__ void __launch_bounds__(256, KMLS_CP_MINB)
#else
__global__ void __launch_bounds__(256)
#endif
k_count_project(const uint32_t* __restrict__ arena_old, const int64_t* __restrict__ rowptr_old,
                const int32_t* __restrict__ roww_old, const int32_t* __restrict__ ncand,
                const int32_t* __restrict__ wnew, const I2* __restrict__ off2,
                const int2* __restrict__ wide_list, int64_t n_wide, int32_t thr,
                int keep_words, uint32_t* __restrict__ arena_new, int32_t* __restrict__ cand_cnt,
                int32_t* __restrict__ cand_p) {
  __shared__ uint32_t sbuf[8][kChunk][36];
  __shared__ int64_t s_src[8][2 * kChunk], s_dst[8][2 * kChunk];   // queue of surviving candidates
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t li = (int64_t)blockIdx.x * 8 + warp; li < n_wide; li += (int64_t)gridDim.x * 8) {
    const int2 unit = __ldg(wide_list + li);   // (parent, row segment)
    const int64_t p = unit.x;
    const int m = __ldg(ncand + p);
    if (m == 0) continue;                        // warp-uniform
    const int w_old = __ldg(roww_old + p), w_new = __ldg(wnew + p);
    const I2 o = off2[p];
    const uint32_t* rowp = arena_old + __ldg(rowptr_old + p);
    if (w_old == w_new) continue;            // rows that keep their column space: k_count_and
    const int nv = w_old >> 2;
    const uint4* msk4 = reinterpret_cast<const uint4*>(rowp);
    const int seg = unit.y;
    const int seg_lo = seg * kSegWords, seg_hi = min(seg_lo + kSegWords, w_old);   // this unit's mask words
    const bool last_seg = seg_hi >= w_old;
    int own_lo = 0, own_hi = 0, bits_lo = 0;
    int queued = 0;                            // survivors waiting in s_src / s_dst (warp-uniform)

    // The squeeze of the first `ns` queued rows: 32 mask words per step; the compress network's
    // move masks are derived once per mask word for all `ns` rows.
    auto squeeze = [&](int ns) {
      for (int i = lane; i < ns * 36; i += 32) sbuf[warp][i / 36][i % 36] = 0;
      __syncwarp();
      int base = bits_lo;                      // bits emitted before this block (the same for every row)
      // past the segment's end only until the last owned word is complete
      for (int blk = seg_lo; blk < w_old && (blk < seg_hi || (base >> 5) < own_hi); blk += 32) {
        const int w = blk + lane;
        const uint32_t mw = w < w_old ? __ldg(rowp + w) : 0u;
        uint32_t xs[kChunk];
#pragma unroll
        for (int t = 0; t < kChunk; ++t)
          xs[t] = (t < ns && w < w_old) ? __ldg(arena_old + s_src[warp][t] + w) : 0u;
        const int nb = __popc(mw);
        int incl = nb;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int u = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d) incl += u;
        }
        const int tot = __shfl_sync(0xffffffffu, incl, 31);
        if (nb) {
          uint32_t mv[5];
          uint32_t mm = mw, mk = ~mw << 1;
#pragma unroll
          for (int i = 0; i < 5; ++i) {
            uint32_t mp = mk ^ (mk << 1);
            mp ^= mp << 2;
            mp ^= mp << 4;
            mp ^= mp << 8;
            mp ^= mp << 16;
            mv[i] = mp & mm;
            mm = (mm ^ mv[i]) | (mv[i] >> (1 << i));
            mk &= ~mp;
          }
          const int bitpos = (base & 31) + incl - nb;
          const int word = bitpos >> 5, sh = bitpos & 31;
#pragma unroll
          for (int t = 0; t < kChunk; ++t) {
            if (t < ns) {
              uint32_t x = xs[t] & mw;
#pragma unroll
              for (int i = 0; i < 5; ++i) {
                const uint32_t tt = x & mv[i];
                x = (x ^ tt) | (tt >> (1 << i));
              }
              if (x) {
                uint32_t* buf = sbuf[warp][t];
                atomicOr(buf + word, x << sh);
                if (sh + nb > 32) atomicOr(buf + word + 1, x >> (32 - sh));
              }
            }
          }
        }
        __syncwarp();
        const int nbits = (base & 31) + tot, full = nbits >> 5, w0 = base >> 5;
        for (int t = 0; t < ns; ++t) {
          uint32_t* buf = sbuf[warp][t];
          uint32_t* dst = arena_new + s_dst[warp][t];
          const uint32_t v0 = buf[lane], v1 = lane < 4 ? buf[32 + lane] : 0u;
          const uint32_t partial = buf[full];
          __syncwarp();
          if (lane < full && w0 + lane >= own_lo && w0 + lane < own_hi) dst[w0 + lane] = v0;
          if (lane < 4 && 32 + lane < full && w0 + 32 + lane >= own_lo && w0 + 32 + lane < own_hi) dst[w0 + 32 + lane] = v1;
          buf[lane] = lane == 0 ? partial : 0u;
          if (lane < 4) buf[32 + lane] = 0u;
        }
        __syncwarp();
        base += tot;
      }
      for (int t = 0; t < ns; ++t) {
        uint32_t* dst = arena_new + s_dst[warp][t];
        int done = base >> 5;
        if ((base & 31) && lane == 0 && done >= own_lo && done < own_hi) dst[done] = sbuf[warp][t][0];
        if (base & 31) ++done;
        if (last_seg)
          for (int i = max(done, own_lo) + lane; i < w_new; i += 32) dst[i] = 0u;
      }
      __syncwarp();
    };

    for (int i0 = 0; i0 < m; i0 += kChunk) {
      const int mc = min(kChunk, m - i0);
      const int64_t my_src = lane < mc ? __ldg(rowptr_old + p + 1 + i0 + lane) : 0;
      int acc[kChunk];
      const uint4* rows[kChunk];
#pragma unroll
      for (int t = 0; t < kChunk; ++t) {
        acc[t] = 0;
        rows[t] = reinterpret_cast<const uint4*>(arena_old + __shfl_sync(0xffffffffu, my_src, t));
      }
      // every segment's unit counts the WHOLE row (cheap next to the squeeze); the first chunk
      // also counts the mask bits before its segment: that is where its output starts
      int b_lo = 0, b_hi = 0;
      for (int v = lane; v < nv; v += 32) {
        const uint4 mw = ldg4(msk4 + v);
        if (i0 == 0) {
          const int pm = popc4(mw);
          if (4 * v < seg_lo) b_lo += pm;
          if (4 * v < seg_hi) b_hi += pm;
        }
#pragma unroll
        for (int t = 0; t < kChunk; ++t)
          if (t < mc) acc[t] += popc4(and4(ldg4(rows[t] + v), mw));
      }
      if (i0 == 0) {
        bits_lo = warp_sum(b_lo);
        const int bits_hi = warp_sum(b_hi);
        // output words this unit owns: those whose first bit lies in its segment's bit range
        own_lo = (bits_lo + 31) >> 5;
        own_hi = last_seg ? w_new : (bits_hi + 31) >> 5;
      }
      uint32_t surv = 0;
#pragma unroll
      for (int t = 0; t < kChunk; ++t) {
        if (t < mc) {
          const int cnt = warp_sum(acc[t]);
          if (lane == 0 && seg == 0) {
            cand_cnt[o.a + i0 + t] = cnt;
            cand_p[o.a + i0 + t] = (int32_t)p;
          }
          if (cnt >= thr) surv |= 1u << t;
        }
      }
      if (own_lo >= own_hi) {                  // this segment owns no output word: count only
        if (seg != 0) break;
        continue;
      }
      if (surv == 0) continue;
      __syncwarp();
      if (lane < mc && ((surv >> lane) & 1u)) {
        const int slot = queued + __popc(surv & ((1u << lane) - 1u));
        s_src[warp][slot] = my_src;
        s_dst[warp][slot] = o.b + (int64_t)(i0 + lane) * w_new;
      }
      queued += __popc(surv);
      __syncwarp();
      if (queued >= kChunk) {                  // a full batch: squeeze it, keep the rest queued
        squeeze(kChunk);
        const int rest = queued - kChunk;
        int64_t a_src = 0, a_dst = 0;
        if (lane < rest) {
          a_src = s_src[warp][kChunk + lane];
          a_dst = s_dst[warp][kChunk + lane];
        }
        __syncwarp();
        if (lane < rest) {
          s_src[warp][lane] = a_src;
          s_dst[warp][lane] = a_dst;
        }
        queued = rest;
        __syncwarp();
      }
    }
    if (queued > 0) squeeze(queued);
  }
}

// work figures for the profiler (launched only while it is on)
// out[0] / out[1]: words read on the AND / bit-extract route (row(p) once + one row per candidate)
__global__ void __launch_bounds__(256)
k_level_work(const int32_t* __restrict__ ncand, const int32_t* __restrict__ roww,
             const int32_t* __restrict__ wnew, int64_t n,
             unsigned long long* __restrict__ out) {
  unsigned long long acc = 0, acc_w = 0;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (int64_t)gridDim.x * blockDim.x) {
    const long long m = ncand[p];
    if (m) {
      const unsigned long long w = (unsigned long long)((m + 1) * (long long)roww[p]);
      if (roww[p] == wnew[p]) acc += w; else acc_w += w;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    acc_w += __shfl_xor_sync(0xffffffffu, acc_w, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (acc) atomicAdd(out, acc);
    if (acc_w) atomicAdd(out + 1, acc_w);
  }
}
// out[0] / out[1]: row words of the new level whose parent took the AND / bit-extract route
__global__ void __launch_bounds__(256)
k_written_work(const int32_t* __restrict__ roww_new, const int32_t* __restrict__ par_new,
               const int32_t* __restrict__ roww_old, const int32_t* __restrict__ wnew_old, int64_t n,
               unsigned long long* __restrict__ out) {
  unsigned long long acc = 0, acc_w = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long w = (unsigned long long)roww_new[i];
    if (roww_old[par_new[i]] == wnew_old[par_new[i]]) acc += w; else acc_w += w;
  }
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    acc_w += __shfl_xor_sync(0xffffffffu, acc_w, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (acc) atomicAdd(out, acc);
    if (acc_w) atomicAdd(out + 1, acc_w);
  }
}
__global__ void __launch_bounds__(256)
k_sum_i32(const int32_t* __restrict__ v, int64_t n, unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    acc += (unsigned long long)v[i];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}
// sum of the upper triangle (diagonal included) of the pair matrix = counter updates made
__global__ void __launch_bounds__(256)
k_sum_upper(const int32_t* __restrict__ pair, int32_t n_rows, unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  const int64_t n2 = (int64_t)n_rows * n_rows;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n2; e += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(e / n_rows), j = (int)(e - (int64_t)i * n_rows);
    if (j >= i) acc += (unsigned long long)pair[e];
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

struct GeThr {
  int32_t thr;
  __host__ __device__ int64_t operator()(int32_t c) const { return c >= thr ? 1 : 0; }
};

__global__ void k_scan_total_post(const int32_t* cnt, int64_t n, int32_t thr, int64_t* off,
                                  volatile int64_t* box) {
  const int64_t t = n > 0 ? off[n - 1] + (cnt[n - 1] >= thr ? 1 : 0) : 0;
  off[n] = t;
  box[0] = t;
  __threadfence_system();
}

// New level table from the survivors (order preserved, so the table stays sorted): members,
// count, parent, and the conditional row = the slot k_count_project filled. A survivor
// whose parent has no other survivor gets width 0: it has no sibling to be joined with.
__global__ void __launch_bounds__(256)
k_emit_rows(const int32_t* __restrict__ items, int k, const int32_t* __restrict__ cand_p,
            const int32_t* __restrict__ cand_cnt, int64_t n_cand, int32_t thr,
            const int64_t* __restrict__ out_off, const I2* __restrict__ off2,
            const int32_t* __restrict__ wnew, int32_t* __restrict__ out_items,
            int32_t* __restrict__ out_counts, int32_t* __restrict__ out_par,
            int64_t* __restrict__ out_rowptr, int32_t* __restrict__ out_roww) {
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_cand; c += (int64_t)gridDim.x * blockDim.x) {
    const int32_t cnt = cand_cnt[c];
    if (cnt < thr) continue;
    const int64_t o = out_off[c];
    const int32_t p = cand_p[c];
    const I2 b = off2[p], e = off2[p + 1];
    const int64_t i = c - b.a;
    const int64_t q = (int64_t)p + 1 + i;
    const int32_t* ip = items + (int64_t)p * k;
    int32_t* op = out_items + o * (k + 1);
    for (int t = 0; t < k; ++t) op[t] = __ldg(ip + t);
    op[k] = __ldg(items + q * k + (k - 1));
    out_counts[o] = cnt;
    out_par[o] = p;
    const int32_t w = __ldg(wnew + p);
    const int64_t survivors = out_off[e.a] - out_off[b.a];
    out_rowptr[o] = b.b + i * w;
    out_roww[o] = survivors >= 2 ? w : 0;
  }
}

// ---- host side -----------------------------------------------------------------------------
struct Miner {
  Workspace ws;
  cudaStream_t s;
  volatile int64_t* box;
  int64_t launches = 0;
  const char* stage = "start";
  Miner(void* base, size_t bytes, cudaStream_t stream) : ws(base, bytes), s(stream), box(mailbox()) {}

  template <class T>
  int take(T** out, int64_t n, bool high = true) {
    *out = (T*)ws.alloc(sizeof(T) * (size_t)std::max<int64_t>(n, 1), high);
    if (*out == nullptr)
      return fail(kCapacityError,
                  "kmls_mine: workspace too small at %s: %zu more bytes needed while %zu are in use "
                  "(call again with a larger workspace; nothing was truncated)",
                  stage, ws.failed_bytes, ws.in_use);
    return 0;
  }
  int sync() {
    KMLS_CUDA(cudaStreamSynchronize(s));
    return 0;
  }
};

#define KMLS_TRY(expr)        \
  do {                        \
    int rc_ = (expr);         \
    if (rc_) return rc_;      \
  } while (0)

static inline int grid_for(int64_t n, int per_block, int max_per_sm) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((n + per_block - 1) / per_block, (int64_t)num_sms() * max_per_sm));
}

// frequent tracks: ids with key[id] >= thr, ordered by (key, id). Returns their number.
static int select_and_order(Miner& M, const int32_t* d_key, int32_t n, int32_t thr, int32_t** d_ids_out,
                            int32_t** d_keys_out, int32_t* n_out) {
  int32_t *sel = nullptr, *keys = nullptr, *ids2 = nullptr, *keys2 = nullptr, *d_n = nullptr;
  KMLS_TRY(M.take(&sel, n));
  KMLS_TRY(M.take(&d_n, 2));
  cub::CountingInputIterator<int32_t> idx(0);
  size_t need = 0;
  KMLS_CUDA(cub::DeviceSelect::If(nullptr, need, idx, sel, d_n, n, OccAtLeast{d_key, thr}, M.s));
  void* tmp = nullptr;
  KMLS_TRY(M.take((char**)&tmp, (int64_t)need));
  KMLS_CUDA(cub::DeviceSelect::If(tmp, need, idx, sel, d_n, n, OccAtLeast{d_key, thr}, M.s));
  M.launches += 2;
  k_post<<<1, 1, 0, M.s>>>(nullptr, nullptr, d_n, nullptr, M.box);
  KMLS_LAUNCHED("k_post");
  KMLS_TRY(M.sync());
  M.ws.release(tmp);
  const int32_t m = (int32_t)M.box[2];
  *n_out = m;
  *d_ids_out = *d_keys_out = nullptr;
  if (m > 0) {
    KMLS_TRY(M.take(&keys, m));
    KMLS_TRY(M.take(&ids2, m, false));
    KMLS_TRY(M.take(&keys2, m));
    k_gather_i32<<<grid_for(m, 256, 8), 256, 0, M.s>>>(d_key, sel, m, keys);
    KMLS_LAUNCHED("k_gather_i32");
    need = 0;
    KMLS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, keys, keys2, sel, ids2, m, 0, 32, M.s));
    KMLS_TRY(M.take((char**)&tmp, (int64_t)need));
    KMLS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, need, keys, keys2, sel, ids2, m, 0, 32, M.s));   // stable
    M.launches += 8;
    M.ws.release(tmp);
    M.ws.release(keys);
    *d_ids_out = ids2;
    *d_keys_out = keys2;
  }
  M.ws.release(sel);
  M.ws.release(d_n);
  return 0;
}

}  // namespace kmls

using namespace kmls;

extern "C" int kmls_thresholds(double min_support, int64_t n, int64_t* t1, int64_t* tk) {
  KMLS_REQUIRE(t1 && tk && n >= 0, "kmls_thresholds: bad arguments");
  KMLS_REQUIRE(min_support > 0.0,
               "`min_support` must be a positive number within the interval `(0, 1]`. Got %g.", min_support);
  const double prod = min_support * (double)n;
  const int64_t k = (int64_t)std::ceil(prod);
  *tk = k;
  if (n == 0) {
    *t1 = 0;
    return 0;
  }
  int64_t c = std::max<int64_t>(0, std::min<int64_t>(n + 1, k));
  while (c > 0 && (double)(c - 1) / (double)n >= min_support) --c;
  while (c <= n && !((double)c / (double)n >= min_support)) ++c;
  *t1 = c;
  return 0;
}

extern "C" int kmls_mine(const int64_t* d_offsets, const int32_t* d_items, int64_t n_playlists,
                         int32_t n_items, int64_t n_total, int32_t t1, int32_t tk,
                         const kmls_mine_options* opt_in, void* d_workspace, size_t workspace_bytes,
                         kmls_mine_result* R, kmls_stream_t stream) {
  KMLS_REQUIRE(R != nullptr, "kmls_mine: result is null");
  kmls_mine_options opt{};
  if (opt_in) opt = *opt_in;
  memset(R, 0, sizeof(*R));
  for (int k = 0; k < KMLS_MAX_LEVELS + 2; ++k)
    R->items_off[k] = R->counts_off[k] = R->parent_off[k] = R->h_items_off[k] = R->h_counts_off[k] = -1;
  R->row_items_off = R->occurrences_off = R->pair_off = R->h_row_items_off = -1;
  R->n_playlists_total = n_total;
  R->t1 = t1;
  R->tk = R->thr_k = tk;
  KMLS_REQUIRE(n_playlists >= 0 && n_items >= 0 && n_total >= n_playlists && t1 >= 0 && tk >= 0,
               "kmls_mine: bad sizes or thresholds");
  const int keep_words = opt.keep_words > 0 ? opt.keep_words : kKeepWordsDefaultMine;
  KMLS_REQUIRE(keep_words <= kKeepWordsMax && (keep_words & 3) == 0, "kmls_mine: keep_words must be a multiple of 4, at most 1024");
  const int squeeze_pct = opt.squeeze_pct > 0 ? opt.squeeze_pct : 50;
  KMLS_REQUIRE(squeeze_pct <= 100, "kmls_mine: squeeze_pct is a percentage");
  KMLS_REQUIRE(kmls_bitmap_row_words(n_playlists) / 4 < (1 << 29), "kmls_mine: too many playlists for one row");
  KMLS_REQUIRE(opt.n_parts <= 1 || (opt.part >= 0 && opt.part < opt.n_parts && opt.part_block >= 1),
               "kmls_mine: bad class partition");
  const int64_t max_itemsets = opt.max_itemsets > 0 ? opt.max_itemsets : 200000000;
  const int max_len = opt.max_len > 0 ? opt.max_len : KMLS_MAX_LEVELS;
  if (n_total == 0 || n_items == 0 || n_playlists == 0) return 0;
  KMLS_REQUIRE(d_offsets && d_items, "kmls_mine: null input");
  KMLS_REQUIRE(d_workspace && ((uintptr_t)d_workspace & 255) == 0, "kmls_mine: workspace must be 256-byte aligned device memory");
  cudaStream_t s = (cudaStream_t)stream;
  Miner M(d_workspace, workspace_bytes, s);
  KMLS_REQUIRE(M.box != nullptr, "kmls_mine: cannot allocate the pinned mailbox");
  char* const ws_base = (char*)d_workspace;
  auto off_of = [&](const void* p) { return (int64_t)((const char*)p - ws_base); };

  // host-side copies of the tables (optional), on a side stream
  cudaStream_t side = nullptr;
  cudaEvent_t ev = nullptr;
  size_t h_used = 0;
  struct SideGuard {
    cudaStream_t& st;
    cudaEvent_t& e;
    ~SideGuard() {
      if (e) cudaEventDestroy(e);
      if (st) cudaStreamDestroy(st);
    }
  } side_guard{side, ev};
  if (opt.h_tables) {
    KMLS_CUDA(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
    KMLS_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  }
  auto to_host = [&](const void* d_src, size_t bytes, int64_t* h_off) -> int {
    if (!opt.h_tables) return 0;
    const size_t at = (h_used + 255) / 256 * 256;
    if (at + bytes > opt.h_tables_bytes)
      return fail(kCapacityError, "kmls_mine: h_tables too small (%zu bytes needed so far, %zu given)",
                  at + bytes, opt.h_tables_bytes);
    KMLS_CUDA(cudaEventRecord(ev, s));
    KMLS_CUDA(cudaStreamWaitEvent(side, ev, 0));
    // in pieces: a later small transfer is never stuck behind one long copy
    const size_t piece = (size_t)4 << 20;
    for (size_t a = 0; a < bytes; a += piece)
      KMLS_CUDA(cudaMemcpyAsync((char*)opt.h_tables + at + a, (const char*)d_src + a, std::min(piece, bytes - a),
                                cudaMemcpyDeviceToHost, side));
    *h_off = (int64_t)at;
    h_used = at + bytes;
    return 0;
  };

  // ---------------------------------------------------------------- level 1
  M.stage = "level 1";
  int32_t* occ = nullptr;
  KMLS_TRY(M.take(&occ, n_items, false));
  R->occurrences_off = off_of(occ);
  // nnz from the offsets' last entry
  k_post<<<1, 1, 0, s>>>(d_offsets + n_playlists, nullptr, nullptr, nullptr, M.box);
  KMLS_LAUNCHED("k_post");
  KMLS_TRY(M.sync());
  const int64_t nnz = M.box[0];
  KMLS_REQUIRE(nnz >= 0, "kmls_mine: offsets are not ascending");
  {
    // ranks of a class-partitioned job that can sum over the ranks each count a share of the
    // occurrences (cut at multiples of 4 ids: the kernel loads 128 bits at a time)
    const bool shared_hist = opt.allreduce_i32 != nullptr && opt.n_parts > 1;
    const int64_t i_lo = shared_hist ? (nnz * opt.part / opt.n_parts) & ~(int64_t)3 : 0;
    const int64_t i_hi = shared_hist && opt.part + 1 < opt.n_parts ? (nnz * (opt.part + 1) / opt.n_parts) & ~(int64_t)3 : nnz;
    {
      ProfScope ps("k_item_histogram", s, 4.0 * (double)(i_hi - i_lo), 1);
      KMLS_TRY(kmls_item_histogram(d_items + i_lo, i_hi - i_lo, n_items, occ, s));
    }
    if (shared_hist && opt.allreduce_i32(opt.allreduce_ctx, occ, n_items, s) != 0)
      return fail(kArgError, "kmls_mine: the allreduce callback failed");
  }
  ++M.launches;
  const int32_t thr1 = std::max(t1, 1);
  int32_t *cand = nullptr, *keys = nullptr, n_rows = 0;
  KMLS_TRY(select_and_order(M, occ, n_items, thr1, &cand, &keys, &n_rows));
  if (n_rows == 0) {
    R->workspace_peak_bytes = M.ws.peak;
    R->kernel_launches = M.launches;
    return 0;
  }
  const int64_t row_words = kmls_bitmap_row_words(n_playlists);
  int32_t* row_of_item = nullptr;
  uint32_t* bitmap = nullptr;
  int32_t *counts1 = nullptr, *flags = nullptr;
  KMLS_TRY(M.take(&row_of_item, n_items));
  KMLS_TRY(M.take(&bitmap, (int64_t)n_rows * row_words));
  KMLS_TRY(M.take(&counts1, n_rows, false));
  KMLS_TRY(M.take(&flags, 16));
  KMLS_CUDA(cudaMemsetAsync(flags, 0, 16 * sizeof(int32_t), s));
  k_fill_i32<<<grid_for(n_items, 256, 8), 256, 0, s>>>(row_of_item, n_items, -1);
  KMLS_LAUNCHED("k_fill_i32");
  k_number_rows<<<grid_for(n_rows, 256, 8), 256, 0, s>>>(cand, n_rows, row_of_item, nullptr);
  KMLS_LAUNCHED("k_number_rows");
  {
    ProfScope ps("k_encode_bitmap", s, 4.0 * (double)nnz + 8.0 * (double)n_playlists + 4.0 * (double)n_rows * (double)row_words, 1);
    KMLS_TRY(kmls_encode_bitmap(d_offsets, d_items, n_playlists, row_of_item, n_items, n_rows, bitmap, row_words, s));
  }
  {
    ProfScope ps("k_row_popcount", s, 4.0 * (double)n_rows * (double)row_words, 1);
    KMLS_TRY(kmls_row_popcount(bitmap, n_rows, row_words, counts1, s));
  }
  k_check_counts<<<grid_for(n_rows, 256, 4), 256, 0, s>>>(counts1, keys, n_rows, n_rows - kHotRows, flags);
  KMLS_LAUNCHED("k_check_counts");
  k_post<<<1, 1, 0, s>>>((const int64_t*)(flags + 2), (const int64_t*)(flags + 8), flags, nullptr, M.box);
  KMLS_LAUNCHED("k_post");
  M.launches += 6;
  KMLS_TRY(M.sync());
  int64_t sum_counts = M.box[0], sum_hot = M.box[1];
  if (M.box[2] != 0) {
    // some playlist repeats a track: exact supports are below the occurrence
    // counts, so drop the rows that fell under T1 and restore ascending-support order
    int32_t *src = nullptr, *keys2 = nullptr, n_keep = 0;
    KMLS_TRY(select_and_order(M, counts1, n_rows, thr1, &src, &keys2, &n_keep));
    if (n_keep == 0) {
      R->workspace_peak_bytes = M.ws.peak;
      R->kernel_launches = M.launches;
      return 0;
    }
    uint32_t* bitmap2 = nullptr;
    int32_t* cand2 = nullptr;
    KMLS_TRY(M.take(&bitmap2, (int64_t)n_keep * row_words));
    KMLS_TRY(M.take(&cand2, n_keep, false));
    KMLS_TRY(kmls_gather_rows(bitmap, src, n_keep, row_words, bitmap2, s));
    k_gather_i32<<<grid_for(n_keep, 256, 8), 256, 0, s>>>(cand, src, n_keep, cand2);
    KMLS_LAUNCHED("k_gather_i32");
    KMLS_CUDA(cudaMemcpyAsync(counts1, keys2, sizeof(int32_t) * (size_t)n_keep, cudaMemcpyDeviceToDevice, s));
    KMLS_CUDA(cudaMemsetAsync(flags, 0, 16 * sizeof(int32_t), s));
    k_check_counts<<<grid_for(n_keep, 256, 4), 256, 0, s>>>(counts1, keys2, n_keep, n_keep - kHotRows, flags);
    KMLS_LAUNCHED("k_check_counts");
    k_post<<<1, 1, 0, s>>>((const int64_t*)(flags + 2), (const int64_t*)(flags + 8), nullptr, nullptr, M.box);
    KMLS_LAUNCHED("k_post");
    KMLS_TRY(M.sync());
    sum_counts = M.box[0];
    sum_hot = M.box[1];
    M.launches += 5;
    M.ws.release(bitmap);
    M.ws.release(cand);
    M.ws.release(src);
    M.ws.release(keys2);
    bitmap = bitmap2;
    cand = cand2;
    n_rows = n_keep;
    k_fill_i32<<<grid_for(n_items, 256, 8), 256, 0, s>>>(row_of_item, n_items, -1);
    KMLS_LAUNCHED("k_fill_i32");
    k_number_rows<<<grid_for(n_rows, 256, 8), 256, 0, s>>>(cand, n_rows, row_of_item, nullptr);
    KMLS_LAUNCHED("k_number_rows");
    M.launches += 2;
  }
  M.ws.release(keys);
  int32_t* items1 = nullptr;
  KMLS_TRY(M.take(&items1, n_rows, false));
  k_number_rows<<<grid_for(n_rows, 256, 8), 256, 0, s>>>(cand, n_rows, row_of_item, items1);
  KMLS_LAUNCHED("k_number_rows");
  ++M.launches;
  R->n_rows = n_rows;
  R->row_items_off = off_of(cand);
  R->max_k = 1;
  R->level_size[1] = n_rows;
  R->level_candidates[1] = n_items;
  R->items_off[1] = off_of(items1);
  R->counts_off[1] = off_of(counts1);
  R->n_itemsets = n_rows;
  KMLS_TRY(to_host(cand, sizeof(int32_t) * (size_t)n_rows, &R->h_row_items_off));
  KMLS_TRY(to_host(items1, sizeof(int32_t) * (size_t)n_rows, &R->h_items_off[1]));
  KMLS_TRY(to_host(counts1, sizeof(int32_t) * (size_t)n_rows, &R->h_counts_off[1]));

  auto finish = [&]() -> int {
    if (opt.h_tables) {
      KMLS_CUDA(cudaStreamSynchronize(side));
      R->h_bytes = (int64_t)h_used;
    }
    KMLS_CUDA(cudaStreamSynchronize(s));
    R->workspace_peak_bytes = M.ws.peak;
    R->kernel_launches = M.launches;
    return 0;
  };
  if (n_rows < 2 || max_len <= 1) return finish();

  // ---------------------------------------------------------------- level 2
  M.stage = "level 2";
  int32_t* pair = nullptr;
  KMLS_TRY(M.take(&pair, (int64_t)n_rows * n_rows, false));
  R->pair_off = off_of(pair);
  int impl = opt.pair_impl;
  bool hybrid = false;
  const bool big = n_rows >= 256 && row_words >= 512;
  if (impl == 0) {
    // tensor cores once there are enough tiles and playlists to feed them, unless the
    // one-hot matrix is so sparse that counting each playlist's pairs is cheaper: about
    // 0.65 m^2 counter updates per playlist (m = mean kept tracks) at ~2e11 updates/s,
    // against F(F+1)N int8 operations at ~2.2e15/s
    impl = big ? 3 : 2;
    if (big && n_rows <= 16384) {
      const double m_mean = (double)sum_counts / (double)n_playlists;
      if (0.65 * m_mean * m_mean * (double)n_playlists / 2.0e11 <
          (double)n_rows * (n_rows + 1.0) * (double)n_playlists / 2.2e15)
        impl = 1;
    }
  }
  if (impl == 1) {
    int rec = -1;
    // a rank of a class-partitioned job counts the pairs of ITS share of the playlists only and
    // the ranks' matrices are summed (the job's host side supplies the allreduce); the shares
    // start at multiples of 128 playlists (one K block of the tensor-core kernel)
    const bool shared_pairs = opt.allreduce_i32 != nullptr && opt.n_parts > 1;
    const int64_t pl_lo = shared_pairs ? (n_playlists * opt.part / opt.n_parts) & ~(int64_t)127 : 0;
    const int64_t pl_hi = shared_pairs && opt.part + 1 < opt.n_parts
                              ? (n_playlists * (opt.part + 1) / opt.n_parts) & ~(int64_t)127 : n_playlists;
    // Hybrid: the block of the kHotRows most frequent rows on the tensor cores, the rest per
    // playlist. Under a Zipf law that block holds about (hot rows' share of all occurrences)^2
    // of the counter updates; it pays when those updates (~1.5e11/s) cost more than the block's
    // 20 tiles x 128 x 256 x playlists int8 multiply-adds (~2.2e15 ops/s, x 1.5 for the tail wave).
    int n_hot = 0;
    if ((row_words & 3) != 0 || pl_hi <= pl_lo) n_hot = 0;
    else if (opt.hot_rows > 0) n_hot = std::min<int>((opt.hot_rows + 127) / 128 * 128, n_rows);
    else if (opt.hot_rows == 0 && n_rows >= 2 * kHotRows && pl_hi - pl_lo >= 65536) {
      const double m_mean = (double)sum_counts / (double)n_playlists;
      const double share = ((double)sum_hot / (double)sum_counts) * ((double)sum_hot / (double)sum_counts);
      const double t_saved = share * 0.65 * m_mean * m_mean * (double)(pl_hi - pl_lo) / 1.5e11;
      const double t_block = 1.5 * 20.0 * 128.0 * 256.0 * 2.0 * (double)(pl_hi - pl_lo) / 2.2e15;
      if (t_saved > 1.5 * t_block) n_hot = kHotRows;
    }
    {
      ProfScope ps("k_pair_counts_csr", s, 0.0, 2);
      rec = ps.idx;
      KMLS_TRY(kmls_pair_counts_csr_cold(d_offsets + pl_lo, d_items, pl_hi - pl_lo, row_of_item, n_rows, n_hot, pair,
                                         flags + 4, s));
    }
    if (!profiler().work) rec = -1;
    if (rec >= 0) {   // work = counter updates made = sum of the upper triangle, 4 bytes each
      KMLS_CUDA(cudaMemsetAsync(flags + 6, 0, 8, s));
      k_sum_upper<<<grid_for((int64_t)n_rows * n_rows, 256, 8), 256, 0, s>>>(pair, n_rows, (unsigned long long*)(flags + 6));
      KMLS_LAUNCHED("k_sum_upper");
    }
    if (n_hot > 0 && pl_hi > pl_lo) {
      const int64_t w_lo = pl_lo >> 5, w_hi = pl_hi == n_playlists ? row_words : pl_hi >> 5;
      const size_t gbytes = kmls_pair_gemm_scratch_bytes(n_hot, w_hi - w_lo);
      char* gscratch = nullptr;
      KMLS_TRY(M.take(&gscratch, (int64_t)gbytes));
      const int64_t h0 = n_rows - n_hot;
      {
        ProfScope ps("k_pair_gemm_i8", s, (double)n_hot * (n_hot + 1.0) * (double)(pl_hi - pl_lo), 2);
        KMLS_TRY(kmls_pair_counts_i8_block(bitmap + h0 * row_words + w_lo, n_hot, w_hi - w_lo, row_words,
                                           pair + h0 * n_rows + h0, n_rows, gscratch, gbytes, s));
      }
      M.launches += 2;
      M.ws.release(gscratch);
    }
    if (shared_pairs) {
      ProfScope ps("allreduce_pair_counts", s, 4.0 * (double)n_rows * (double)n_rows, 2);
      if (opt.allreduce_i32(opt.allreduce_ctx, pair, (int64_t)n_rows * n_rows, s) != 0 ||
          opt.allreduce_i32(opt.allreduce_ctx, flags + 4, 1, s) != 0)
        return fail(kArgError, "kmls_mine: the allreduce callback failed");
    }
    k_post<<<1, 1, 0, s>>>(rec >= 0 ? (const int64_t*)(flags + 6) : nullptr, nullptr, flags + 4, nullptr, M.box);
    KMLS_LAUNCHED("k_post");
    M.launches += 2;
    KMLS_TRY(M.sync());
    if (rec >= 0) profiler().records[rec].work = 4.0 * (double)M.box[0];
    if (M.box[2] != 0) impl = big ? 3 : 2;   // a playlist with more than 1024 kept tracks
    else if (n_hot > 0) hybrid = true;
  }
  if (impl == 2) {
    ProfScope ps("k_pair_counts", s, 4.0 * (double)n_rows * (double)row_words, 2);
    KMLS_TRY(kmls_pair_counts(bitmap, n_rows, row_words, pair, s));
    ++M.launches;
  } else if (impl == 3) {
    const size_t gbytes = kmls_pair_gemm_scratch_bytes(n_rows, row_words);
    char* gscratch = nullptr;
    KMLS_TRY(M.take(&gscratch, (int64_t)gbytes));
    {
      ProfScope ps("k_pair_gemm_i8", s, (double)n_rows * (n_rows + 1.0) * (double)n_playlists, 2);
      KMLS_TRY(kmls_pair_counts_i8(bitmap, n_rows, row_words, pair, gscratch, gbytes, s));
    }
    M.launches += 2;
    M.ws.release(gscratch);
  } else if (impl != 1) {
    return fail(kArgError, "kmls_mine: unknown pair_impl %d", impl);
  }
  R->pair_impl_used = hybrid ? 4 : impl;
  M.ws.release(row_of_item);
  int32_t thr = tk;
  if (t1 < tk) {
    KMLS_CUDA(cudaMemsetAsync(flags, 0, sizeof(int32_t), s));
    k_nested_check<<<grid_for((int64_t)n_rows * n_rows, 256, 8), 256, 0, s>>>(pair, counts1, n_rows, flags);
    KMLS_LAUNCHED("k_nested_check");
    k_post<<<1, 1, 0, s>>>(nullptr, nullptr, flags, nullptr, M.box);
    KMLS_LAUNCHED("k_post");
    M.launches += 2;
    KMLS_TRY(M.sync());
    if (M.box[2] == 0) thr = t1;
  }
  R->thr_k = thr;
  int64_t* row_off = nullptr;
  KMLS_TRY(M.take(&row_off, n_rows + 1));
  {
    ProfScope ps("k_pair_select", s, 4.0 * (double)n_rows * (double)n_rows, 2);
    KMLS_TRY(kmls_pair_select(pair, n_rows, thr, 0, row_off, nullptr, nullptr, s));
  }
  k_post<<<1, 1, 0, s>>>(row_off + n_rows, nullptr, nullptr, nullptr, M.box);
  KMLS_LAUNCHED("k_post");
  M.launches += 3;
  KMLS_TRY(M.sync());
  const int64_t n2 = M.box[0];
  R->level_candidates[2] = (int64_t)n_rows * (n_rows - 1) / 2;
  if (n2 == 0) return finish();
  KMLS_REQUIRE(n2 <= 0x7fffffff, "kmls_mine: more than 2^31-1 frequent pairs");
  int32_t *items_k = nullptr, *counts_k = nullptr, *par_k = nullptr, *second2 = nullptr;
  KMLS_TRY(M.take(&items_k, n2 * 2, false));
  KMLS_TRY(M.take(&counts_k, n2, false));
  KMLS_TRY(M.take(&par_k, n2, false));
  KMLS_TRY(M.take(&second2, n2));
  {
    ProfScope ps("k_pair_select", s, 4.0 * (double)n_rows * (double)n_rows + 12.0 * (double)n2, 2);
    KMLS_TRY(kmls_pair_select(pair, n_rows, thr, n2, row_off, items_k, counts_k, s));
  }
  k_pair_parents<<<grid_for(n2, 256, 8), 256, 0, s>>>(items_k, n2, par_k, second2);
  KMLS_LAUNCHED("k_pair_parents");
  M.launches += 4;
  M.ws.release(row_off);
  R->max_k = 2;
  R->level_size[2] = n2;
  R->items_off[2] = off_of(items_k);
  R->counts_off[2] = off_of(counts_k);
  R->parent_off[2] = off_of(par_k);
  R->n_itemsets += n2;
  KMLS_TRY(to_host(items_k, sizeof(int32_t) * 2 * (size_t)n2, &R->h_items_off[2]));
  KMLS_TRY(to_host(counts_k, sizeof(int32_t) * (size_t)n2, &R->h_counts_off[2]));
  if (n2 < 2 || max_len <= 2) return finish();

  // conditional rows of the pairs: row(a, b) = row_b sampled at the playlists of a
  int64_t n_k = n2;
  uint32_t* arena_k = nullptr;
  int64_t* rowptr_k = nullptr;
  int32_t* roww_k = nullptr;
  {
    size_t sbytes = kmls_scan_scratch_bytes(std::max<int64_t>(n2, n_rows));
    char* scratch = nullptr;
    int64_t* tid_off = nullptr;
    int32_t* tids = nullptr;
    KMLS_TRY(M.take(&roww_k, n2));
    KMLS_TRY(M.take(&rowptr_k, n2 + 1));
    KMLS_TRY(M.take(&scratch, (int64_t)sbytes));
    KMLS_TRY(M.take(&tid_off, n_rows + 1));
    KMLS_TRY(kmls_row_alloc(par_k, n2, counts1, nullptr, keep_words, roww_k, s));
    const int32_t* tid_counts = counts1;
    int32_t *need_row = nullptr, *masked = nullptr;
    if (opt.n_parts > 1) {
      // a part of a class-partitioned job builds only the rows its own subtrees read
      KMLS_TRY(M.take(&need_row, n_rows));
      KMLS_TRY(M.take(&masked, n_rows));
      KMLS_CUDA(cudaMemsetAsync(need_row, 0, sizeof(int32_t) * (size_t)n_rows, s));
      k_level2_needs<<<grid_for(n2, 256, 8), 256, 0, s>>>(par_k, n2, std::max(opt.part_block, 1), opt.n_parts, opt.part,
                                                          roww_k, need_row);
      KMLS_LAUNCHED("k_level2_needs");
      k_mask_counts<<<grid_for(n_rows, 256, 4), 256, 0, s>>>(counts1, need_row, n_rows, masked);
      KMLS_LAUNCHED("k_mask_counts");
      tid_counts = masked;
      M.launches += 2;
    }
    KMLS_TRY(kmls_exclusive_scan_i32(roww_k, n2, rowptr_k, scratch, sbytes, s));
    KMLS_TRY(kmls_exclusive_scan_i32(tid_counts, n_rows, tid_off, scratch, sbytes, s));
    k_post<<<1, 1, 0, s>>>(rowptr_k + n2, tid_off + n_rows, nullptr, nullptr, M.box);
    KMLS_LAUNCHED("k_post");
    M.launches += 8;
    KMLS_TRY(M.sync());
    const int64_t words = M.box[0], n_tids = M.box[1];
    R->level_row_words[2] = words;
    KMLS_TRY(M.take(&arena_k, words));
    KMLS_TRY(M.take(&tids, n_tids));
    {
      ProfScope ps("k_row_tidlists", s, 4.0 * (double)n_rows * (double)row_words + 4.0 * (double)n_tids, 2);
      KMLS_TRY(kmls_row_tidlists(bitmap, n_rows, row_words, tid_off, tids, s));
    }
    ProfScope ps("k_project_gather", s, 4.0 * (double)words, 2);
    KMLS_TRY(kmls_project_rows_gather(bitmap, row_words, par_k, second2, n2, tids, tid_off, arena_k, rowptr_k,
                                      words, s));
    M.launches += 2;
    M.ws.release(tids);
    M.ws.release(tid_off);
    M.ws.release(scratch);
    M.ws.release(need_row);
    M.ws.release(masked);
  }
  M.ws.release(second2);
  M.ws.release(bitmap);
  bitmap = nullptr;

  // ---------------------------------------------------------------- levels k >= 3
  int k = 2;
  int64_t total = R->n_itemsets;
  while (k < max_len && n_k >= 2) {
    M.stage = "levels >= 3";
    KMLS_REQUIRE(k + 1 <= KMLS_MAX_LEVELS, "kmls_mine: itemsets longer than %d members", KMLS_MAX_LEVELS);
    KMLS_REQUIRE(n_k <= 0x7fffffff, "kmls_mine: more than 2^31-1 itemsets at level %d", k);
    int32_t *ncand = nullptr, *wnew = nullptr;
    I2* off2 = nullptr;
    KMLS_TRY(M.take(&ncand, n_k));
    KMLS_TRY(M.take(&wnew, n_k));
    KMLS_TRY(M.take(&off2, n_k + 1));
    const bool parted = k == 2 && opt.n_parts > 1;
    const bool prof = profiler().on && profiler().work;
    KMLS_CUDA(cudaMemsetAsync(flags + 5, 0, 3 * sizeof(int32_t), s));
    {
      ProfScope ps("k_join_size", s, 20.0 * (double)n_k, k + 1);
      k_join_size<<<grid_for(n_k, 256, 16), 256, 0, s>>>(par_k, roww_k, counts_k, n_k, keep_words, squeeze_pct,
                                                         parted ? opt.n_parts : 1, opt.part, std::max(opt.part_block, 1),
                                                         ncand, wnew, off2, flags + 5);
      KMLS_LAUNCHED("k_join_size");
    }
    if (prof) {
      KMLS_CUDA(cudaMemsetAsync(flags + 8, 0, 16, s));
      k_level_work<<<grid_for(n_k, 256, 8), 256, 0, s>>>(ncand, roww_k, wnew, n_k, (unsigned long long*)(flags + 8));
      KMLS_LAUNCHED("k_level_work");
    }
    {
      // in place over the tuples k_join_size wrote (one coalesced 16-byte stream)
      size_t need = 0;
      KMLS_CUDA(cub::DeviceScan::ExclusiveScan(nullptr, need, off2, off2, I2Sum(), I2{0, 0, 0}, (int)n_k, s));
      char* tmp = nullptr;
      KMLS_TRY(M.take(&tmp, (int64_t)need));
      ProfScope ps("scan_slots", s, 32.0 * (double)n_k, k + 1);
      KMLS_CUDA(cub::DeviceScan::ExclusiveScan(tmp, need, off2, off2, I2Sum(), I2{0, 0, 0}, (int)n_k, s));
      M.ws.release(tmp);
    }
    k_i2_total<<<1, 1, 0, s>>>(CandSlots{ncand, wnew, roww_k}, n_k, off2, flags + 5, M.box);
    KMLS_LAUNCHED("k_i2_total");
    M.launches += 4;
    KMLS_TRY(M.sync());
    KMLS_REQUIRE(M.box[4] <= 0x7fffffff, "kmls_mine: more than 2^31-1 candidates at level %d", k + 1);
    const int64_t n_cand = M.box[0], slot_words = M.box[1], n_units = M.box[2], n_wide = M.box[3];
    double read_and = 0.0, read_wide = 0.0;
    if (prof) {
      unsigned long long h[2] = {0, 0};
      KMLS_CUDA(cudaMemcpyAsync(h, flags + 8, 16, cudaMemcpyDeviceToHost, s));
      KMLS_TRY(M.sync());
      read_and = (double)h[0];
      read_wide = (double)h[1];
    }
    R->level_candidates[k + 1] = n_cand;
    if (n_cand == 0) break;
    if (total + n_cand > max_itemsets * 4)
      return fail(kCapacityError, "kmls_mine: level %d: %lld candidates exceed the max_itemsets guard", k + 1,
                  (long long)n_cand);
    KMLS_REQUIRE(n_cand <= 0x7fffffff, "kmls_mine: more than 2^31-1 candidates at level %d", k + 1);
    uint32_t* arena_n = nullptr;
    int32_t *cand_cnt = nullptr, *cand_p = nullptr;
    int64_t* out_off = nullptr;
    KMLS_TRY(M.take(&arena_n, slot_words));
    KMLS_TRY(M.take(&cand_cnt, n_cand));
    KMLS_TRY(M.take(&cand_p, n_cand));
    KMLS_TRY(M.take(&out_off, n_cand + 1));
    // algorithmic bytes of counting + projecting this level: row(p) once per parent + one row
    // per candidate read, 8 bytes per candidate written (count, parent), + the rows of the
    // survivors (added after the emit); split between the two kernels by their share of parents
    int rec_cp = -1, rec_and = -1;
    Unit* units = nullptr;
    int2* wide_list = nullptr;
    KMLS_TRY(M.take(&units, n_units));
    KMLS_TRY(M.take(&wide_list, n_wide));
    KMLS_CUDA(cudaMemsetAsync(flags + 5, 0, sizeof(int32_t), s));
    {
      ProfScope ps("k_expand_units", s, 64.0 * (double)n_units + 24.0 * (double)n_k, k + 1);
      k_expand_units<<<grid_for(n_k, 256, 16), 256, 0, s>>>(arena_k, rowptr_k, roww_k, ncand, wnew, off2, n_k, arena_n,
                                                            units, wide_list, flags + 5);
      KMLS_LAUNCHED("k_expand_units");
      ++M.launches;
    }
    if (n_units > 0) {
      ProfScope ps("k_count_and", s, 4.0 * read_and + 32.0 * (double)n_units, k + 1);
      rec_and = ps.idx;
      // half a warp per unit once the level's units are mostly single candidates (k_count_and_half)
      static const int and_half = [] { const char* e = getenv("KMLS_AND_HALF"); return e ? atoi(e) : -1; }();
      // (measured at Z1M: levels 5-10, where a unit holds 1-2 candidates, gain 5-8 %; levels 3-4 with their
      // full units of four lose as much)
      const bool half = and_half >= 0 ? and_half != 0 : (double)n_cand < 2.6 * (double)n_units;
      if (half) k_count_and_half<<<grid_for(n_units, 16, 48), 256, 0, s>>>(units, n_units, thr, cand_cnt, cand_p);
      else k_count_and<<<grid_for(n_units, 8, 48), 256, 0, s>>>(units, n_units, thr, cand_cnt, cand_p);
      KMLS_LAUNCHED("k_count_and");
      ++M.launches;
    }
    if (n_wide > 0) {
      ProfScope ps("k_count_project", s, 4.0 * read_wide + 8.0 * (double)n_wide, k + 1);
      rec_cp = ps.idx;
      k_count_project<<<grid_for(n_wide, 8, 64), 256, 0, s>>>(arena_k, rowptr_k, roww_k, ncand, wnew, off2, wide_list,
                                                              n_wide, thr, keep_words, arena_n, cand_cnt, cand_p);
      KMLS_LAUNCHED("k_count_project");
    }
    {
      cub::TransformInputIterator<int64_t, GeThr, const int32_t*> it(cand_cnt, GeThr{thr});
      size_t need = 0;
      KMLS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, it, out_off, (int)n_cand, s));
      char* tmp = nullptr;
      KMLS_TRY(M.take(&tmp, (int64_t)need));
      ProfScope ps("scan_survivors", s, 12.0 * (double)n_cand, k + 1);
      KMLS_CUDA(cub::DeviceScan::ExclusiveSum(tmp, need, it, out_off, (int)n_cand, s));
      M.ws.release(tmp);
    }
    k_scan_total_post<<<1, 1, 0, s>>>(cand_cnt, n_cand, thr, out_off, M.box);
    KMLS_LAUNCHED("k_scan_total_post");
    M.launches += 4;
    KMLS_TRY(M.sync());
    const int64_t n_next = M.box[0];
    if (n_next == 0) break;
    total += n_next;
    if (total > max_itemsets)
      return fail(kCapacityError,
                  "kmls_mine: more than max_itemsets=%lld frequent itemsets at level %d; raise the guard or "
                  "min_support (results are never truncated)", (long long)max_itemsets, k + 1);
    int32_t *items_n = nullptr, *counts_n = nullptr, *par_n = nullptr, *roww_n = nullptr;
    int64_t* rowptr_n = nullptr;
    KMLS_TRY(M.take(&items_n, n_next * (k + 1), false));
    KMLS_TRY(M.take(&counts_n, n_next, false));
    KMLS_TRY(M.take(&par_n, n_next, false));
    KMLS_TRY(M.take(&rowptr_n, n_next));
    KMLS_TRY(M.take(&roww_n, n_next));
    {
      ProfScope ps("k_emit_rows", s, 16.0 * (double)n_cand + (double)n_next * (4.0 * (2 * k + 1) + 20.0), k + 1);
      k_emit_rows<<<grid_for(n_cand, 256, 16), 256, 0, s>>>(items_k, k, cand_p, cand_cnt, n_cand, thr, out_off, off2,
                                                            wnew, items_n, counts_n, par_n, rowptr_n, roww_n);
      KMLS_LAUNCHED("k_emit_rows");
    }
    ++M.launches;
    if ((rec_cp >= 0 || rec_and >= 0) && profiler().work) {   // rows written = widths of the new itemsets that own a row
      unsigned long long h[2] = {0, 0};
      KMLS_CUDA(cudaMemsetAsync(flags + 8, 0, 16, s));
      k_written_work<<<grid_for(n_next, 256, 8), 256, 0, s>>>(roww_n, par_n, roww_k, wnew, n_next,
                                                             (unsigned long long*)(flags + 8));
      KMLS_LAUNCHED("k_written_work");
      KMLS_CUDA(cudaMemcpyAsync(h, flags + 8, 16, cudaMemcpyDeviceToHost, s));
      KMLS_TRY(M.sync());
      if (rec_and >= 0) profiler().records[rec_and].work += 4.0 * (double)h[0];
      if (rec_cp >= 0) profiler().records[rec_cp].work += 4.0 * (double)h[1];
    }
    // this level's scratch and the previous level's rows are done with (stream order)
    M.ws.release(units);
    M.ws.release(wide_list);
    M.ws.release(ncand);
    M.ws.release(wnew);
    M.ws.release(off2);
    M.ws.release(cand_cnt);
    M.ws.release(cand_p);
    M.ws.release(out_off);
    M.ws.release(arena_k);
    M.ws.release(rowptr_k);
    M.ws.release(roww_k);
    ncand = wnew = cand_cnt = cand_p = nullptr;
    arena_k = arena_n;
    rowptr_k = rowptr_n;
    roww_k = roww_n;
    items_k = items_n;
    counts_k = counts_n;
    par_k = par_n;
    n_k = n_next;
    ++k;
    R->max_k = k;
    R->level_size[k] = n_k;
    R->level_row_words[k] = slot_words;
    R->items_off[k] = off_of(items_k);
    R->counts_off[k] = off_of(counts_k);
    R->parent_off[k] = off_of(par_k);
    R->n_itemsets = total;
    KMLS_TRY(to_host(items_k, sizeof(int32_t) * (size_t)k * (size_t)n_k, &R->h_items_off[k]));
    KMLS_TRY(to_host(counts_k, sizeof(int32_t) * (size_t)n_k, &R->h_counts_off[k]));
  }
  return finish();
}
