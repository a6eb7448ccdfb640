// K2 on tensor cores: pair co-occurrence counts as an int8 GEMM, C = B * B^T, with
// B in {0,1}^(F x N) the vertical bitmap (row = track, column = playlist).
//
// Replaces the k = 2 part of mlxtend's FP-tree mining (fpg_step / conditional_tree, reached
// from reference machine-learning/main.py:272); same result as k_pair_counts (pairs.cu), on
// the 5th-generation tensor cores: tcgen05.mma kind::i8 with exact int32 accumulation in TMEM.
//
// The bitmap stays bit-packed in HBM and L2; it is expanded to int8 only inside the SM:
//   warp 0      TMA producer: cp.async.bulk of the 128-playlist bit slices of the CTA's
//               128 A rows and 256 B rows (2 KB + 4 KB per stage) into shared memory;
//   warps 2-9   expanders: 16 bits -> 16 bytes (0/1) with one multiply-and-mask per nibble,
//               written straight into the 128-byte-swizzled K-major operand tiles;
//   warp 1      one thread issues 4 x tcgen05.mma (M=128, N=256, K=32) per stage and commits
//               to the stage's "empty" mbarrier; accumulators: 256 TMEM columns;
//   warps 4-7   epilogue: tcgen05.ld 32 lanes x 32 columns at a time, int32 stores to C.
// Only tiles that touch the upper triangle (col >= row) are computed.
//
// Operand traffic: the bit slices come from a K-block-major copy of the bitmap made by
// k_bits_kblock_major (one pass, 2 * F * W bytes), so every stage is two contiguous bulk
// copies. Useful work: 2 * N * F(F+1)/2 int8 ops.
#include <cstdlib>

#include "bulk.cuh"
#include "common.cuh"
#include "kmls_b200.h"

namespace kmls {

constexpr int kGM = 128;          // A rows per tile (TMEM lanes)
constexpr int kGN = 256;          // B rows per tile (TMEM columns)
constexpr int kGKBytes = 128;     // int8 K per stage = one 128-byte swizzle span = 128 playlists
constexpr int kGStages = 4;
constexpr int kGExpanderWarps = 8;
constexpr int kGThreads = 32 * (2 + kGExpanderWarps);
constexpr int kGOpBytes = (kGM + kGN) * kGKBytes;      // 48 KB of int8 per stage
constexpr int kGBitBytes = (kGM + kGN) * 16;           // 6 KB of bits per stage
constexpr int kGSmemBytes = kGStages * (kGOpBytes + kGBitBytes) + 1024 /*align*/ + 256 /*barriers*/;

// ---- PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

// Shared-memory matrix descriptor: K-major, 128-byte swizzle, 8-row groups 1024 bytes apart.
__device__ __forceinline__ uint64_t smem_desc_k_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);        // start address
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused with swizzle)
  d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;                        // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
  return d;
}
// Instruction descriptor: D = s32, A = B = u8, both K-major, N = 256, M = 128.
constexpr uint32_t kIdescI8 = (2u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(kGN >> 3) << 17) |
                              ((uint32_t)(kGM >> 4) << 24);

// ---------------------------------------------------------------------------------------
// Re-layout: out[kb][row][4 words] = bitmap[row][4*kb .. 4*kb+3]; rows >= n_rows are zero.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_bits_kblock_major(const uint32_t* __restrict__ bitmap, int32_t n_rows, int32_t rows_pad,
                    int64_t row_words, int64_t row_stride, uint4* __restrict__ out) {
  // tile: 32 rows x 32 k-blocks through shared memory so that both sides are coalesced
  __shared__ uint4 tile[32][33];
  const int64_t kb0 = (int64_t)blockIdx.x * 32;
  const int r0 = blockIdx.y * 32;
  const int64_t n_kb = row_words >> 2;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 8 warps
  for (int rr = ty; rr < 32; rr += 8) {
    const int r = r0 + rr;
    const int64_t kb = kb0 + tx;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < n_rows && kb < n_kb) v = ldg4(reinterpret_cast<const uint4*>(bitmap + (int64_t)r * row_stride) + kb);
    tile[rr][tx] = v;
  }
  __syncthreads();
  for (int kk = ty; kk < 32; kk += 8) {
    const int64_t kb = kb0 + kk;
    const int r = r0 + tx;
    if (kb < n_kb && r < rows_pad) out[kb * rows_pad + r] = tile[tx][kk];
  }
}

// ---------------------------------------------------------------------------------------
// The GEMM. grid = (tiles of 256 B-rows, tiles of 128 A-rows); one tile per CTA, one CTA per SM.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kGThreads, 1)
k_pair_gemm_i8(const uint4* __restrict__ bits_km, int32_t n_rows, int32_t rows_pad, int64_t n_kb_total,
               int64_t kb_per_split, int32_t* __restrict__ C, int64_t ld) {
  const int tj = blockIdx.x, ti = blockIdx.y;
  if (tj * kGN + (kGN - 1) < ti * kGM) return;  // tile entirely below the diagonal
  // blockIdx.z selects a slice of the playlist dimension; with more than one slice the
  // epilogue accumulates into a zeroed C with atomic adds
  const int64_t kb0 = (int64_t)blockIdx.z * kb_per_split;
  const int64_t n_kb = min(kb_per_split, n_kb_total - kb0);
  if (n_kb <= 0) return;
  const bool split = gridDim.z > 1;
  bits_km += kb0 * rows_pad;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t op_base = base;                                  // stages of [A 16 KB | B 32 KB]
  const uint32_t bit_base = base + kGStages * kGOpBytes;          // stages of [A 2 KB | B 4 KB]
  const uint32_t bar_base = bit_base + kGStages * kGBitBytes;     // mbarriers
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));     // generic pointer to `base`
  auto bits_full = [&](int s) { return bar_base + 8u * s; };
  auto bits_empty = [&](int s) { return bar_base + 8u * (kGStages + s); };
  auto op_full = [&](int s) { return bar_base + 8u * (2 * kGStages + s); };
  auto op_empty = [&](int s) { return bar_base + 8u * (3 * kGStages + s); };
  const uint32_t tmem_full = bar_base + 8u * (4 * kGStages);
  const uint32_t tmem_slot = bar_base + 8u * (4 * kGStages + 1);
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kGStages; ++s) {
      mbar_init(bits_full(s), 1);
      mbar_init(bits_empty(s), kGExpanderWarps);
      mbar_init(op_full(s), kGExpanderWarps);
      mbar_init(op_empty(s), 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot_gen;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int64_t i = 0; i < n_kb; ++i) {
        const int s = (int)(i % kGStages);
        const uint32_t ph = (uint32_t)((i / kGStages) & 1);
        mbar_wait(bits_empty(s), ph ^ 1u);
        mbar_expect_tx(bits_full(s), kGBitBytes);
        const uint4* src = bits_km + i * rows_pad;
        const uint32_t dst = bit_base + s * kGBitBytes;
        tma_bulk_g2s(dst, src + (int64_t)ti * kGM, kGM * 16, bits_full(s));
        tma_bulk_g2s(dst + kGM * 16, src + (int64_t)tj * kGN, kGN * 16, bits_full(s));
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      for (int64_t i = 0; i < n_kb; ++i) {
        const int s = (int)(i % kGStages);
        const uint32_t ph = (uint32_t)((i / kGStages) & 1);
        mbar_wait(op_full(s), ph);
        tc_fence_after();
        const uint32_t a_addr = op_base + s * kGOpBytes;
        const uint64_t a_desc = smem_desc_k_sw128(a_addr);
        const uint64_t b_desc = smem_desc_k_sw128(a_addr + kGM * kGKBytes);
#pragma unroll
        for (int k = 0; k < kGKBytes / 32; ++k)   // advance 32 bytes (K = 32) inside the swizzle span
          tc_mma_i8(tmem_acc, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), kIdescI8,
                    (i > 0 || k > 0) ? 1u : 0u);
        tc_commit(op_empty(s));
      }
      tc_commit(tmem_full);
    }
  } else {
    // ===== expanders: bits -> int8 operand tiles =====
    const int et = threadIdx.x - 64;  // 0 .. 255
    for (int64_t i = 0; i < n_kb; ++i) {
      const int s = (int)(i % kGStages);
      const uint32_t ph = (uint32_t)((i / kGStages) & 1);
      mbar_wait(bits_full(s), ph);
      mbar_wait(op_empty(s), ph ^ 1u);
      const uint8_t* bits = gen_base + (bit_base - base) + s * kGBitBytes;
      uint8_t* ops = gen_base + s * kGOpBytes;
      // unit = (row, half): 64 bits in, four 16-byte chunks out. All loads first (so that no
      // store sits between them and they overlap), then pure ALU + 128-bit stores. Lanes 0-7
      // of a store hit 8 consecutive rows with the same logical chunk: the XOR swizzle puts
      // them in 8 different 16-byte bank groups (conflict-free); the 64-bit loads of 16
      // consecutive lanes cover 16 distinct bank pairs.
      constexpr int kUnits = (kGM + kGN) * 2 / (kGExpanderWarps * 32);   // 3 per thread
      uint2 in[kUnits];
#pragma unroll
      for (int it = 0; it < kUnits; ++it) {
        const int u = it * (kGExpanderWarps * 32) + et;
        const int r = ((u >> 4) << 3) + (u & 7), half = (u >> 3) & 1;
        in[it] = *reinterpret_cast<const uint2*>(bits + r * 16 + half * 8);
      }
#pragma unroll
      for (int it = 0; it < kUnits; ++it) {
        const int u = it * (kGExpanderWarps * 32) + et;
        const int r = ((u >> 4) << 3) + (u & 7), half = (u >> 3) & 1;
        uint8_t* orow = ops + r * kGKBytes;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const uint32_t h = ((jj & 2) ? in[it].y : in[it].x) >> ((jj & 1) * 16);
          uint4 v;
          v.x = ((h & 0xFu) * 0x00204081u) & 0x01010101u;
          v.y = (((h >> 4) & 0xFu) * 0x00204081u) & 0x01010101u;
          v.z = (((h >> 8) & 0xFu) * 0x00204081u) & 0x01010101u;
          v.w = (((h >> 12) & 0xFu) * 0x00204081u) & 0x01010101u;
          const int j = half * 4 + jj;
          *reinterpret_cast<uint4*>(orow + ((j ^ (r & 7)) << 4)) = v;
        }
      }
      fence_proxy_async_smem();   // generic-proxy stores -> visible to the tensor core's async proxy
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(op_full(s));
        mbar_arrive(bits_empty(s));
      }
    }
    // ===== epilogue (warps 4..7: TMEM lane quadrant = warp % 4) =====
    if (warp >= 4 && warp < 8) {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      const int q = warp & 3;
      const int row = ti * kGM + q * 32 + lane;
      for (int c0 = 0; c0 < kGN; c0 += 32) {
        uint32_t v[32];
        const uint32_t taddr = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (row < n_rows) {
          int32_t* crow = C + (int64_t)row * ld;
          const int col0 = tj * kGN + c0;
#pragma unroll
          for (int t = 0; t < 32; ++t)
            if (col0 + t < n_rows) {
              if (split) atomicAdd(crow + col0 + t, (int32_t)v[t]);
              else crow[col0 + t] = (int32_t)v[t];
            }
        }
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(256) : "memory");
  }
}


// ---------------------------------------------------------------------------------------
// Variant with the A operand in tensor memory ("TS"). In the kernel above every stage writes
// 48 KB of int8 into shared memory and the tensor core reads the same 48 KB back: 96 KB per
// 524-cycle stage against a 128 B/cycle shared-memory port, i.e. the port, not the tensor
// pipe, sets the pace. Here the 128 A rows are expanded into REGISTERS and stored to TMEM
// with tcgen05.st (lane = row, 32 columns of 4 int8 per stage); only the 256 B rows go through
// shared memory (32 KB written + 32 KB read per stage).
//   warp 0        TMA producer (bit slices)          warp 1      MMA issuer (A from TMEM)
//   warps 4-7     A expanders (quadrant = warp % 4) and, at the end, the epilogue
//   warps 2,3,8-13  B expanders (as above)
// TMEM: columns 0-255 accumulators, 256 + 32*stage: the stage's A tile.
// ---------------------------------------------------------------------------------------
constexpr int kTBWarps = 8;                                  // B expander warps
constexpr int kTThreads = 32 * 14;
constexpr int kTStages = 4;
constexpr int kTOpBytes = kGN * kGKBytes;                    // 32 KB of int8 (B only) per stage
constexpr int kTSmemBytes = kTStages * (kTOpBytes + kGBitBytes) + 1024 + 256;

__device__ __forceinline__ void tc_mma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

__global__ void __launch_bounds__(kTThreads, 1)
k_pair_gemm_i8_ts(const uint4* __restrict__ bits_km, int32_t n_rows, int32_t rows_pad, int64_t n_kb_total,
                  int64_t kb_per_split, int32_t* __restrict__ C, int64_t ld) {
  const int tj = blockIdx.x, ti = blockIdx.y;
  if (tj * kGN + (kGN - 1) < ti * kGM) return;
  const int64_t kb0 = (int64_t)blockIdx.z * kb_per_split;
  const int64_t n_kb = min(kb_per_split, n_kb_total - kb0);
  if (n_kb <= 0) return;
  const bool split = gridDim.z > 1;
  bits_km += kb0 * rows_pad;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t op_base = base;                                  // stages of B (32 KB)
  const uint32_t bit_base = base + kTStages * kTOpBytes;          // stages of [A 2 KB | B 4 KB] bits
  const uint32_t bar_base = bit_base + kTStages * kGBitBytes;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  auto bits_full = [&](int s) { return bar_base + 8u * s; };
  auto bits_empty = [&](int s) { return bar_base + 8u * (kTStages + s); };
  auto op_full = [&](int s) { return bar_base + 8u * (2 * kTStages + s); };
  auto op_empty = [&](int s) { return bar_base + 8u * (3 * kTStages + s); };
  const uint32_t tmem_full = bar_base + 8u * (4 * kTStages);
  const uint32_t tmem_slot = bar_base + 8u * (4 * kTStages + 1);
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < kTStages; ++s) {
      mbar_init(bits_full(s), 1);
      mbar_init(bits_empty(s), kTBWarps + 4);
      mbar_init(op_full(s), kTBWarps + 4);
      mbar_init(op_empty(s), 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot_gen;
  const uint32_t tmem_a0 = tmem_acc + 256;

  if (warp == 0) {
    if (lane == 0) {
      for (int64_t i = 0; i < n_kb; ++i) {
        const int s = (int)(i % kTStages);
        const uint32_t ph = (uint32_t)((i / kTStages) & 1);
        mbar_wait(bits_empty(s), ph ^ 1u);
        mbar_expect_tx(bits_full(s), kGBitBytes);
        const uint4* src = bits_km + i * rows_pad;
        const uint32_t dst = bit_base + s * kGBitBytes;
        tma_bulk_g2s(dst, src + (int64_t)ti * kGM, kGM * 16, bits_full(s));
        tma_bulk_g2s(dst + kGM * 16, src + (int64_t)tj * kGN, kGN * 16, bits_full(s));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      for (int64_t i = 0; i < n_kb; ++i) {
        const int s = (int)(i % kTStages);
        const uint32_t ph = (uint32_t)((i / kTStages) & 1);
        mbar_wait(op_full(s), ph);
        tc_fence_after();
        const uint64_t b_desc = smem_desc_k_sw128(op_base + s * kTOpBytes);
#pragma unroll
        for (int k = 0; k < kGKBytes / 32; ++k)   // 32 int8 of K = 8 TMEM columns of A = 32 bytes of B
          tc_mma_i8_ts(tmem_acc, tmem_a0 + (uint32_t)(s * 32 + k * 8), b_desc + (uint64_t)(k * 2), kIdescI8,
                       (i > 0 || k > 0) ? 1u : 0u);
        tc_commit(op_empty(s));
      }
      tc_commit(tmem_full);
    }
  } else if (warp >= 4 && warp < 8) {
    // ===== A expanders: row = 32 * quadrant + lane; 128 bits -> 32 columns of 4 int8 =====
    const int q = warp & 3;
    for (int64_t i = 0; i < n_kb; ++i) {
      const int s = (int)(i % kTStages);
      const uint32_t ph = (uint32_t)((i / kTStages) & 1);
      mbar_wait(bits_full(s), ph);
      mbar_wait(op_empty(s), ph ^ 1u);
      const uint4 b = *reinterpret_cast<const uint4*>(gen_base + (bit_base - base) + s * kGBitBytes +
                                                      (q * 32 + lane) * 16);
      uint32_t v[32];
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        const uint32_t w = c < 8 ? b.x : c < 16 ? b.y : c < 24 ? b.z : b.w;
        v[c] = (((w >> (4 * (c & 7))) & 0xFu) * 0x00204081u) & 0x01010101u;
      }
      const uint32_t taddr = tmem_a0 + ((uint32_t)(q * 32) << 16) + (uint32_t)(s * 32);
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
          "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
          "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
          ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
            "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
            "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
            "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
          : "memory");
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(op_full(s));
        mbar_arrive(bits_empty(s));
      }
    }
    // ===== epilogue =====
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row = ti * kGM + q * 32 + lane;
    for (int c0 = 0; c0 < kGN; c0 += 32) {
      uint32_t v[32];
      const uint32_t taddr = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
            "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
            "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
            "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(taddr) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (row < n_rows) {
        int32_t* crow = C + (int64_t)row * ld;
        const int col0 = tj * kGN + c0;
#pragma unroll
        for (int t = 0; t < 32; ++t)
          if (col0 + t < n_rows) {
            if (split) atomicAdd(crow + col0 + t, (int32_t)v[t]);
            else crow[col0 + t] = (int32_t)v[t];
          }
      }
    }
    tc_fence_before();
  } else {
    // ===== B expanders: bits -> int8 operand tile in shared memory =====
    const int bt = warp < 4 ? (warp - 2) * 32 + lane : 64 + (warp - 8) * 32 + lane;   // 0 .. 255
    for (int64_t i = 0; i < n_kb; ++i) {
      const int s = (int)(i % kTStages);
      const uint32_t ph = (uint32_t)((i / kTStages) & 1);
      mbar_wait(bits_full(s), ph);
      mbar_wait(op_empty(s), ph ^ 1u);
      const uint8_t* bits = gen_base + (bit_base - base) + s * kGBitBytes + kGM * 16;   // B rows' bits
      uint8_t* ops = gen_base + s * kTOpBytes;
      constexpr int kUnits = kGN * 2 / (kTBWarps * 32);   // 2 per thread
      uint2 in[kUnits];
#pragma unroll
      for (int it = 0; it < kUnits; ++it) {
        const int u = it * (kTBWarps * 32) + bt;
        const int r = ((u >> 4) << 3) + (u & 7), half = (u >> 3) & 1;
        in[it] = *reinterpret_cast<const uint2*>(bits + r * 16 + half * 8);
      }
#pragma unroll
      for (int it = 0; it < kUnits; ++it) {
        const int u = it * (kTBWarps * 32) + bt;
        const int r = ((u >> 4) << 3) + (u & 7), half = (u >> 3) & 1;
        uint8_t* orow = ops + r * kGKBytes;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const uint32_t h = ((jj & 2) ? in[it].y : in[it].x) >> ((jj & 1) * 16);
          uint4 v;
          v.x = ((h & 0xFu) * 0x00204081u) & 0x01010101u;
          v.y = (((h >> 4) & 0xFu) * 0x00204081u) & 0x01010101u;
          v.z = (((h >> 8) & 0xFu) * 0x00204081u) & 0x01010101u;
          v.w = (((h >> 12) & 0xFu) * 0x00204081u) & 0x01010101u;
          const int j = half * 4 + jj;
          *reinterpret_cast<uint4*>(orow + ((j ^ (r & 7)) << 4)) = v;
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(op_full(s));
        mbar_arrive(bits_empty(s));
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"(512) : "memory");
  }
}

}  // namespace kmls

using namespace kmls;

extern "C" size_t kmls_pair_gemm_scratch_bytes(int32_t n_rows, int64_t row_words) {
  if (n_rows <= 0 || row_words <= 0) return 0;
  const int64_t rows_pad = ((int64_t)n_rows + kGN - 1) / kGN * kGN;
  return (size_t)(row_words / 4) * (size_t)rows_pad * 16;
}

// Shared host side of the two entry points. `block`: C is a sub-block (leading dimension ld)
// of a larger, already zeroed matrix - no memset here, and the playlist dimension is split as
// finely as the SM count asks for (a block has few tiles).
static int pair_gemm_run(const char* who, const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                         int64_t row_stride, int32_t* d_pair, int64_t ld, bool block, void* d_scratch,
                         size_t scratch_bytes, cudaStream_t s) {
  const size_t need = kmls_pair_gemm_scratch_bytes(n_rows, row_words);
  if (scratch_bytes < need)
    return fail(kCapacityError, "%s: scratch too small (%zu > %zu bytes)", who, need, scratch_bytes);
  const int32_t rows_pad = (n_rows + kGN - 1) / kGN * kGN;
  const int64_t n_kb = row_words / 4;
  {
    dim3 grid((unsigned)((n_kb + 31) / 32), (unsigned)(rows_pad / 32));
    KMLS_REQUIRE(grid.y <= 65535, "kmls_pair_counts_i8: too many rows");
    k_bits_kblock_major<<<grid, 256, 0, s>>>(d_bitmap, n_rows, rows_pad, row_words, row_stride, (uint4*)d_scratch);
    KMLS_LAUNCHED("k_bits_kblock_major");
  }
  KMLS_CUDA(cudaFuncSetAttribute(k_pair_gemm_i8, cudaFuncAttributeMaxDynamicSharedMemorySize, kGSmemBytes));
  // One tile per CTA and one CTA per SM: split the playlist dimension when the last wave of
  // tiles would leave most SMs idle (or when there are fewer tiles than SMs).
  const int tx = rows_pad / kGN, ty = rows_pad / kGM;
  int64_t tiles = 0;
  for (int i = 0; i < ty; ++i)
    for (int j = 0; j < tx; ++j) tiles += (j * kGN + (kGN - 1) >= i * kGM);
  const int sms = num_sms();
  int ksplit = 1;
  if (!block) {
    for (int ks = 1; ks <= 16; ks *= 2) {
      if (n_kb / ks < 128 && ks > 1) break;
      ksplit = ks;
      const int64_t units = tiles * ks, waves = (units + sms - 1) / sms;
      if ((double)units / (double)(waves * sms) >= 0.93) break;
    }
  } else {
    // shortest makespan = waves / ks, the smallest such ks
    double best = 1e30;
    for (int ks = 1; ks <= 64; ++ks) {
      if (n_kb / ks < 64 && ks > 1) break;
      const int64_t units = tiles * ks, waves = (units + sms - 1) / sms;
      const double span = (double)waves / (double)ks;
      if (span < best * 0.98) best = span, ksplit = ks;
    }
  }
  const int64_t kb_per_split = (n_kb + ksplit - 1) / ksplit;
  if (ksplit > 1 && !block) KMLS_CUDA(cudaMemsetAsync(d_pair, 0, sizeof(int32_t) * (size_t)n_rows * n_rows, s));
  dim3 grid((unsigned)tx, (unsigned)ty, (unsigned)ksplit);
  // default: the variant with the A operand in tensor memory; KMLS_PAIR_GEMM=ss selects the
  // one that stages both operands in shared memory
  const char* variant = getenv("KMLS_PAIR_GEMM");
  if (!(variant && variant[0] == 's')) {
    KMLS_CUDA(cudaFuncSetAttribute(k_pair_gemm_i8_ts, cudaFuncAttributeMaxDynamicSharedMemorySize, kTSmemBytes));
    k_pair_gemm_i8_ts<<<grid, kTThreads, kTSmemBytes, s>>>((const uint4*)d_scratch, n_rows, rows_pad, n_kb,
                                                           kb_per_split, d_pair, ld);
    KMLS_LAUNCHED("k_pair_gemm_i8_ts");
    return 0;
  }
  k_pair_gemm_i8<<<grid, kGThreads, kGSmemBytes, s>>>((const uint4*)d_scratch, n_rows, rows_pad, n_kb,
                                                      kb_per_split, d_pair, ld);
  KMLS_LAUNCHED("k_pair_gemm_i8");
  return 0;
}

extern "C" int kmls_pair_counts_i8(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                                   int32_t* d_pair, void* d_scratch, size_t scratch_bytes,
                                   kmls_stream_t stream) {
  KMLS_REQUIRE(n_rows >= 0 && row_words >= 0 && (row_words & 3) == 0,
               "kmls_pair_counts_i8: row_words must be a multiple of 4");
  if (n_rows == 0 || row_words == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_pair && d_scratch, "kmls_pair_counts_i8: null pointer");
  KMLS_REQUIRE(((uintptr_t)d_scratch & 15) == 0 && ((uintptr_t)d_bitmap & 15) == 0,
               "kmls_pair_counts_i8: buffers must be 16-byte aligned");
  return pair_gemm_run("kmls_pair_counts_i8", d_bitmap, n_rows, row_words, row_words, d_pair, n_rows, false,
                       d_scratch, scratch_bytes, (cudaStream_t)stream);
}

extern "C" int kmls_pair_counts_i8_block(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                                         int64_t row_stride, int32_t* d_pair, int64_t ld, void* d_scratch,
                                         size_t scratch_bytes, kmls_stream_t stream) {
  KMLS_REQUIRE(n_rows >= 0 && row_words >= 0 && (row_words & 3) == 0 && (row_stride & 3) == 0 &&
                   row_stride >= row_words && ld >= n_rows,
               "kmls_pair_counts_i8_block: row_words and row_stride must be multiples of 4, ld >= n_rows");
  if (n_rows == 0 || row_words == 0) return 0;
  KMLS_REQUIRE(d_bitmap && d_pair && d_scratch, "kmls_pair_counts_i8_block: null pointer");
  KMLS_REQUIRE(((uintptr_t)d_scratch & 15) == 0 && ((uintptr_t)d_bitmap & 15) == 0,
               "kmls_pair_counts_i8_block: buffers must be 16-byte aligned");
  return pair_gemm_run("kmls_pair_counts_i8_block", d_bitmap, n_rows, row_words, row_stride, d_pair, ld, true,
                       d_scratch, scratch_bytes, (cudaStream_t)stream);
}
