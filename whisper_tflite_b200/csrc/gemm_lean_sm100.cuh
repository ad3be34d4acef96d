// Sixteen-epilogue-warp variant of the one-CTA tcgen05 GEMM (gemm_sm100.cuh) for the encoder's hot shapes: bf16 C through
// TMA stores, optional bias / GELU / residual through TMA loads — nothing else.
//
// Why: at K = 384 a 128 x 256 tile is 3072 tensor-clocks of MMA, but its epilogue is ~15 500 warp instructions of
// dependent chains (tcgen05.ld -> bias -> GELU -> pack -> st.shared -> bulk store). With two epilogue warps per scheduler
// (gemm_bf16_sm100_kernel: 8 warps, one group per accumulator stage) each warp issues at ~0.2 IPC and a tile takes ~9500
// clocks: the epilogue paces the kernel. Here FOUR warps per scheduler drain accumulators: per accumulator stage two groups
// of four warps, each group taking half of the tile's columns. The general epilogue (f32 / argmax / K-V scatter / RowMap
// addends) needs 168 registers and would spill under the 112-register cap of 576 threads; this one keeps only what the
// TMA path needs (the tile's 32 values, no staging registers) and stays below it.
#pragma once

#include "gemm_sm100.cuh"

#ifndef LEAN_PROBE
#define LEAN_PROBE 0   // timing probes (build.build_cuda_variant + WT_LIB): 1 = no epilogue work, 2 = no operand loads, 4 = no C
                       // stores (bits combine); results are garbage
#endif

namespace wt {

constexpr int kLeanEpiWarps = 16;
constexpr int kLeanThreads = 64 + kLeanEpiWarps * 32;   // TMA producer, MMA issuer, 16 epilogue warps

// Shared memory is carved at run time: `stages` operand stages, then per epilogue warp `boxes` 32 x 32 bf16 boxes of 2 KB
// (SWIZZLE_64B), then the barriers. The operand ring is what paces this kernel once the epilogue is out of the way (ncu
// r02j: 30 % of all warp samples are epilogue warps waiting for an accumulator, tensor pipe 46 % with three 48 KB stages
// at BN = 256), so a GEMM without a residual runs with ONE box per warp (a box is rewritten once its store has been read,
// three other warps per scheduler cover the wait) and the bias comes from global memory (same address for every lane, L1
// resident) instead of a staged copy: that buys the fourth stage at BN = 256 / 192 and six at BN = 128. A residual GEMM
// keeps three (two at BN = 256) boxes per warp so its addend arrives ahead of the chunk that needs it.
template <int BN>
struct GemmLeanSmem {
  static constexpr int kABytes = kGemmBM * kGemmBK * 2;
  static constexpr int kBBytes = BN * kGemmBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTileQ = 4;
  static constexpr int kBarBytes = 1024;
  static constexpr int kMaxStages = 6;
  // `add`: residual GEMM (EPI_TMA_ADD) or the K/V-stream scatter (EPI_SPLIT_KV: one 4 KB box of 32 keys x 128 bytes)
  __host__ __device__ static constexpr int boxes(bool add) { return add ? (BN <= 192 ? 3 : 2) : 1; }
  __host__ __device__ static constexpr int stages(bool add) {
    const int fit = (232448 - 1024 - kLeanEpiWarps * boxes(add) * 2048 - kBarBytes) / kStageBytes;
    return fit > kMaxStages ? kMaxStages : fit;
  }
  __host__ __device__ static constexpr int total(bool add) {
    return stages(add) * kStageBytes + kLeanEpiWarps * boxes(add) * 2048 + kBarBytes + 1024;
  }
  static constexpr int kTmemCols = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
  static_assert(stages(true) >= 3 && stages(false) >= 3, "operand ring too shallow");
};

// One warp's share of a tile: rows rowq .. rowq + 31 (its TMEM lane quarter), columns n0 + col0 .. + 32 * kChunks.
// t_row / bias_s already point at column col0. Boxes: chunk c of a tile with an addend sits in box c % kBoxes (the first
// kBoxes - 1 were requested by the caller before the accumulator was waited for); without one the rotation carries on
// across tiles in `rot`, so "all but the last kBoxes - 1 stores have been read" always frees the box about to be written.
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2, int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
               : "memory");
}

// The cross-attention K/V projection's epilogue (EPI_SPLIT_KV, streaming layout): the tile's 128 rows are one 128-key
// chunk of clip `batch`, a 64-column block is one head's K or V. A lane writes its key's 128-byte row into the warp's
// 4 KB box with the 16-byte pieces XOR-ed by (key & 7) — the layout the decoder's ldmatrix reads (gemm_sm100.cuh's
// scatter writes the same bytes with 16-byte global stores) — and the box goes out verbatim (tensor map without swizzle)
// to [pair][chunk][K | V][32 keys of this warp]. Keys past the clip's end are written as zeros.
template <int kChunks>
__device__ __forceinline__ void gemm_lean_epilogue_kv(const GemmBf16Params& p, const CUtensorMap* map_kv, uint32_t t_row,
                                                      uint8_t* scr, int lane, int lane_q, int batch, int row0, int ncol0) {
  static_assert(kChunks % 2 == 0, "whole 64-column blocks per warp");
  const bool row_ok = row0 + lane_q * 32 + lane < p.rows_per_batch;
  uint8_t* row = scr + lane * 128;
  const int sw = lane & 7;
#pragma unroll 1
  for (int c = 0; c < kChunks; ++c) {
    uint32_t v[32];
    tmem_ld32(t_row + c * 32, v);
    tmem_ld_wait();
    if (p.flags & EPI_BIAS) {
      const float4* bg = reinterpret_cast<const float4*>(p.bias + ncol0 + c * 32);
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 b4 = __ldg(bg + (j >> 2));
        v[j] = __float_as_uint(__uint_as_float(v[j]) + b4.x);
        v[j + 1] = __float_as_uint(__uint_as_float(v[j + 1]) + b4.y);
        v[j + 2] = __float_as_uint(__uint_as_float(v[j + 2]) + b4.z);
        v[j + 3] = __float_as_uint(__uint_as_float(v[j + 3]) + b4.w);
      }
    }
    const int half4 = (c & 1) << 2;
    if (half4 == 0) {
      if (lane == 0) bulk_wait_read0();   // the previous block's store has read the box
      __syncwarp();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 o = make_uint4(0u, 0u, 0u, 0u);
      if (row_ok)
        o = make_uint4(pack_bf16x2(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1])),
                       pack_bf16x2(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                       pack_bf16x2(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                       pack_bf16x2(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])));
      *reinterpret_cast<uint4*>(row + (((half4 + i) ^ sw) << 4)) = o;
    }
    if (half4 != 0) {
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        const int cb = (ncol0 + (c - 1) * 32) >> 6;
        const int which = cb >= p.kv_H ? 1 : 0;
        tma_store_5d(map_kv, scr, 0, lane_q * 32, which, row0 >> 7, batch * p.kv_H + (cb - which * p.kv_H));
        bulk_commit();
      }
    }
  }
}

template <int kChunks>
__device__ __forceinline__ void gemm_lean_epilogue(const GemmBf16Params& p, const CUtensorMap* map_c, const CUtensorMap* map_add,
                                                   uint64_t* bar_add, uint32_t& add_ph, int& rot, uint32_t t_row, int kBoxes,
                                                   uint8_t* scr, int lane, int rowq, int batch, int ncol0) {
  const bool tma_add = (p.flags & EPI_TMA_ADD) != 0;
  int b = tma_add ? 0 : rot;
#pragma unroll 1
  for (int c = 0; c < kChunks; ++c) {
    uint32_t v[32];
    tmem_ld32(t_row + c * 32, v);
    tmem_ld_wait();
    if (p.flags & EPI_BIAS) {
      const float4* bg = reinterpret_cast<const float4*>(p.bias + ncol0 + c * 32);   // N % BN == 0 on this path
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 b4 = __ldg(bg + (j >> 2));
        v[j] = __float_as_uint(__uint_as_float(v[j]) + b4.x);
        v[j + 1] = __float_as_uint(__uint_as_float(v[j + 1]) + b4.y);
        v[j + 2] = __float_as_uint(__uint_as_float(v[j + 2]) + b4.z);
        v[j + 3] = __float_as_uint(__uint_as_float(v[j + 3]) + b4.w);
      }
    }
    if (p.flags & EPI_GELU) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(gelu_tanh_fit(__uint_as_float(v[j])));
    }
    if (tma_add) {
      mbar_wait(bar_add + b, (add_ph >> b) & 1u);
      add_ph ^= 1u << b;
    } else {
      if (lane == 0) {   // the store that last used this box has read it
        if (kBoxes == 3) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
        else if (kBoxes == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      __syncwarp();
    }
    uint8_t* box = scr + b * 2048 + lane * 64;
    const int sw = (lane >> 1) & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4* pc = reinterpret_cast<uint4*>(box + ((i ^ sw) << 4));
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[i * 8 + j]);
      if (tma_add) {
        const uint4 val = *pc;
        const uint32_t w[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          f[2 * j] += __uint_as_float(w[j] << 16);
          f[2 * j + 1] += __uint_as_float(w[j] & 0xffff0000u);
        }
      }
      *pc = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
    }
    fence_proxy_async();
    __syncwarp();
    const int bp = b == 0 ? kBoxes - 1 : b - 1;   // the box of chunk c - 1 (== the box of chunk c + kBoxes - 1)
    if (lane == 0) {
#if !(LEAN_PROBE & 4)
      tma_store_3d(map_c, scr + b * 2048, ncol0 + c * 32, rowq, batch);
#endif
      bulk_commit();
      if (tma_add && c + kBoxes - 1 < kChunks) {   // kBoxes >= 2 with an addend
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // every store but chunk c's has read its box
        mbar_expect_tx(bar_add + bp, 2048);
        tma_load_3d(scr + bp * 2048, map_add, bar_add + bp, ncol0 + (c + kBoxes - 1) * 32, rowq, batch);
      }
    }
    __syncwarp();
    b = b == kBoxes - 1 ? 0 : b + 1;
  }
  if (!tma_add) rot = b;
}

template <int BN>
__global__ void __launch_bounds__(kLeanThreads, 1)
gemm_lean_sm100_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                       const __grid_constant__ CUtensorMap map_c, const __grid_constant__ CUtensorMap map_add,
                       GemmBf16Params p) {
  using S = GemmLeanSmem<BN>;
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  const bool has_add = (p.flags & EPI_TMA_ADD) != 0;
  const bool has_kv = (p.flags & EPI_SPLIT_KV) != 0;   // two 2 KB boxes = the one 4 KB box of the K/V epilogue (BN = 256 / 128)
  const int n_stages = S::stages(has_add || has_kv), n_boxes = has_kv ? 2 : S::boxes(has_add);
  uint8_t* s_stage = smem;
  uint8_t* s_scratch = s_stage + n_stages * S::kStageBytes;
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(s_scratch + kLeanEpiWarps * n_boxes * 2048);
  uint64_t* bar_empty = bar_full + 6;
  uint64_t* bar_tfull = bar_empty + 6;            // [2]
  uint64_t* bar_tempty = bar_tfull + 2;           // [2] 8 epilogue warps each
  uint64_t* bar_qfull = bar_tempty + 2;           // [kTileQ]
  uint64_t* bar_qempty = bar_qfull + S::kTileQ;   // [kTileQ] MMA issuer + the 8 epilogue warps of one accumulator stage
  uint64_t* bar_addend = bar_qempty + S::kTileQ;  // [16 epilogue warps][3] addend boxes landed (EPI_TMA_ADD)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_addend + kLeanEpiWarps * 3);
  volatile int* s_tileq = reinterpret_cast<volatile int*>(s_tmem + 4);
  static_assert((6 + 6 + 2 + 2 + 2 * S::kTileQ + kLeanEpiWarps * 3) * 8 + 16 + S::kTileQ * 4 <= S::kBarBytes, "barrier block");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tiles_m = p.tiles_per_batch * p.n_batches;
  const int n_tiles = n_tiles_m * p.n_tiles_n;
  const int n_kblocks = p.taps * p.kblocks_per_tap;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_c);
    if (p.flags & EPI_TMA_ADD) tma_prefetch_desc(&map_add);
    for (int i = 0; i < n_stages; ++i) {
      mbar_init(bar_full + i, 1);
      mbar_init(bar_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_tfull + i, 1);
      mbar_init(bar_tempty + i, 8);
    }
    for (int i = 0; i < S::kTileQ; ++i) {
      mbar_init(bar_qfull + i, 1);
      mbar_init(bar_qempty + i, 9);
    }
    for (int i = 0; i < kLeanEpiWarps * 3; ++i) mbar_init(bar_addend + i, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, S::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ================= TMA producer + tile scheduler (as in gemm_bf16_sm100_kernel) =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int n_end = 0;
      int next_tile = (int)blockIdx.x;
      for (int it = 0;; ++it) {
        int tile = (n_end == 0) ? next_tile : n_tiles;
        if (tile >= n_tiles) tile = -1;
        if (tile >= 0) next_tile = (int)gridDim.x + atomicAdd(p.sched, 1);
        const int qs = it % S::kTileQ;
        mbar_wait(bar_qempty + qs, ((uint32_t)(it / S::kTileQ) & 1u) ^ 1u);
        s_tileq[qs] = tile;
        mbar_arrive(bar_qfull + qs);
        if (tile < 0) {
          if (++n_end == 2) break;     // one end marker per accumulator stage
          continue;
        }
        const int tn = tile % p.n_tiles_n, tm = tile / p.n_tiles_n;
        const int batch = tm / p.tiles_per_batch, row0 = (tm % p.tiles_per_batch) * kGemmBM;
        for (int kb = 0; kb < n_kblocks; ++kb) {
          mbar_wait(bar_empty + stage, phase ^ 1);
          uint8_t* sa = s_stage + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
#if LEAN_PROBE & 2
          (void)sa; (void)sb; (void)tn; (void)batch; (void)row0;
          mbar_arrive(bar_full + stage);
#else
          mbar_expect_tx(bar_full + stage, S::kStageBytes);
          const int tap = kb / p.kblocks_per_tap, kc = kb - tap * p.kblocks_per_tap;
          tma_load_3d(sa, &map_a, bar_full + stage, p.tap_c0[tap] + kc * kGemmBK, row0 + p.tap_row[tap], batch);
          tma_load_2d(sb, &map_w, bar_full + stage, tap * p.tap_k + kc * kGemmBK, tn * BN);
#endif
          if (++stage == n_stages) { stage = 0; phase ^= 1; }
        }
      }
      __threadfence();
      if (atomicAdd(p.sched + 1, 1) == (int)gridDim.x - 1) {
        p.sched[0] = 0;
        p.sched[1] = 0;
        __threadfence();
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    constexpr uint32_t idesc = make_idesc_bf16(kGemmBM, BN);
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0;; ++it) {
      const int qs = it % S::kTileQ;
      mbar_wait(bar_qfull + qs, (uint32_t)(it / S::kTileQ) & 1u);
      const int tile = s_tileq[qs];
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_qempty + qs);
      if (tile < 0) break;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(bar_tempty + acc, acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = 0; kb < n_kblocks; ++kb) {
        mbar_wait(bar_full + stage, phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(s_stage + stage * S::kStageBytes);
          const uint64_t da = make_sw128_desc(sa);
          const uint64_t db = make_sw128_desc(sa + S::kABytes);
#pragma unroll
          for (int k = 0; k < kGemmBK / 16; ++k) umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(bar_empty + stage);
          if (kb == n_kblocks - 1) umma_commit(bar_tfull + acc);
        }
        __syncwarp();
        if (++stage == n_stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ================= epilogue: warps 2..17 = 4 sets of 4 (one per TMEM lane quarter); set s drains accumulator stage
    //                   s & 1, column half s >> 1 =================
    constexpr int kHalf = BN / 2, kChunks = kHalf / 32;
    const int n_pre = (n_boxes - 1 < kChunks) ? n_boxes - 1 : kChunks;   // addend boxes requested up front
    const int ew = warp - 2;
    const int set = ew >> 2;
    const int eg = set & 1, half = set >> 1;
    const int lane_q = warp & 3;
    uint8_t* scr = s_scratch + ew * n_boxes * 2048;
    uint64_t* bar_add = bar_addend + ew * 3;
    uint32_t add_ph = 0;
    int rot = 0;
    for (int it = eg;; it += 2) {
      const int acc = eg;
      const int qs = it % S::kTileQ;
      mbar_wait(bar_qfull + qs, (uint32_t)(it / S::kTileQ) & 1u);
      const int tile = s_tileq[qs];
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_qempty + qs);
      if (tile < 0) break;
      const uint32_t acc_phase = (it >> 1) & 1;
      const int tn = tile % p.n_tiles_n, tm = tile / p.n_tiles_n;
      const int batch = tm / p.tiles_per_batch, row0 = (tm % p.tiles_per_batch) * kGemmBM;
      const int n0 = tn * BN;
      const int rowq = row0 + lane_q * 32, ncol0 = n0 + half * kHalf;
      if (has_add && lane == 0) {
        bulk_wait_read0();   // the previous tile's stores have read their boxes
        for (int b = 0; b < n_pre; ++b) {
          mbar_expect_tx(bar_add + b, 2048);
          tma_load_3d(scr + b * 2048, &map_add, bar_add + b, ncol0 + b * 32, rowq, batch);
        }
      }
      __syncwarp();
      mbar_wait(bar_tfull + acc, acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(lane_q * 32) << 16) + acc * BN + half * kHalf;
#if LEAN_PROBE & 1
      (void)t_row; (void)rot; (void)add_ph;
#else
      if (has_kv) {
        if constexpr (kChunks % 2 == 0) gemm_lean_epilogue_kv<kChunks>(p, &map_c, t_row, scr, lane, lane_q, batch, row0, ncol0);
      } else {
        gemm_lean_epilogue<kChunks>(p, &map_c, &map_add, bar_add, add_ph, rot, t_row, n_boxes, scr, lane, rowq, batch, ncol0);
      }
#endif
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + acc);
    }
    if (lane == 0) bulk_wait0();   // the copy engine is done with this warp's boxes before the CTA's shared memory goes away
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, S::kTmemCols);
  }
}

}  // namespace wt
