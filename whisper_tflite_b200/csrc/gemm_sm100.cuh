// bf16 tensor-core GEMM for sm_100a: C = epi(A . W^T), f32 accumulate in TMEM.
//
// Every dense contraction of the Whisper graph that the reference runs inside
// tflite::Interpreter::Invoke (whisper.cpp:295, :375 — CONV_2D / FULLY_CONNECTED per
// export/generate_onnx.py:85-130) goes through this one kernel: conv1/conv2 (as GEMMs over overlapping
// rows of a zero-padded time-major activation), QKV, attention out-proj, MLP fc1/fc2, the cross-attention
// K/V projection, and the decoder's per-step projections and tied logits.
//
// Structure (one CTA per SM, persistent over output tiles, 6 warps):
//   warp 0      TMA producer: cp.async.bulk.tensor loads of A [128 x 64] and W [BN x 64] bf16 tiles,
//               128B-swizzled, into a kStages-deep ring guarded by full/empty mbarriers;
//   warp 1      TMEM allocator + MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16
//               (M = 128, N = BN, K = 16) x 4 per ring slot, tcgen05.commit frees the slot / publishes the
//               accumulator; two accumulator stages in TMEM so tile i+1's MMAs overlap tile i's epilogue;
//   warps 2-5   epilogue: tcgen05.ld the accumulator (lane = row), + bias, erf-GELU, + addend (residual or
//               positional rows), convert, transpose through a per-warp shared scratch so that global
//               stores are full 128-byte row segments; or, for the logits, a fused (max, last-index)
//               reduction per row (whisper.cpp:346-361's tie rule) instead of a store.
// Rows are addressed through RowMap (simt_kernels.cuh) on the C/addend side and through a 3-D tensor map
// {inner, rows, batches} on the A side, so M tiles never straddle a clip boundary and TMA's out-of-bounds
// zero fill provides the ragged last tile and the K tail. The k=3 convolutions are "3-tap" GEMMs: the same
// A buffer is read at three (column, row) offsets while K walks the [n][tap][c_in] weight matrix.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "simt_kernels.cuh"

namespace wt {

constexpr int kGemmBM = 128;
constexpr int kGemmBK = 64;
constexpr int kGemmEpiGroups = 2;   // epilogue warp groups (4 warps each): group e drains the accumulator stage of tiles it & 1 == e
constexpr int kGemmThreads = 64 + kGemmEpiGroups * 128;

enum : int { EPI_ADD = 4, EPI_OUT_F32 = 16, EPI_ARGMAX = 32, EPI_SPLIT_KV = 64 };  // EPI_BIAS = 1, EPI_GELU = 2 shared with SIMT
// set by the host when C (bf16) is expressible as a 3-D tensor map {N, rows_per_batch, batches}: the epilogue warps hand
// their 32 x 64 pieces of the tile to TMA (cp.async.bulk.tensor store) instead of issuing the global stores themselves
constexpr int EPI_TMA_C = 128;
constexpr int EPI_TMA_ADD = 256;   // with EPI_TMA_C and EPI_ADD: the addend has a tensor map too (see gemm_epilogue_tile)

struct GemmBf16Params {
  const float* bias;               // [N] f32 or null
  const __nv_bfloat16* addend;     // [M][N] via amap_add (EPI_ADD) — residual stream or positional rows
  RowMap add_map;
  void* C;                         // bf16 (or f32 with EPI_OUT_F32) via cmap; unused with EPI_ARGMAX alone
  RowMap cmap;
  unsigned long long* argmax_part; // [M][n_tiles_n] packed (value, index) (EPI_ARGMAX)
  // EPI_SPLIT_KV: N = 2*H*64 columns (k | v) of row m = clip*kv_S + s go to kv_k / kv_v [clip][H][kv_S][64] — the
  // per-head streams the decoder's cross-attention reads — instead of C
  // kv_chunk > 0 selects the streaming layout of the persistent decoder instead: one buffer (kv_k), per (clip, head)
  // pair kv_cpp chunks of kv_chunk keys, each chunk = [K rows | V rows] (kv_chunk * 128 bytes each), so a decode step
  // pulls a chunk with ONE contiguous 2*kv_chunk*128-byte bulk copy
  __nv_bfloat16 *kv_k, *kv_v;
  int kv_S, kv_H, kv_chunk, kv_cpp;
  int M, N, K;
  int rows_per_batch;              // A-side batching (== M for a plain matrix)
  int tiles_per_batch;             // ceil(rows_per_batch / 128)
  int n_batches;
  int n_tiles_n;
  int flags;
  // K is walked as `taps` groups of `kblocks_per_tap` 64-wide blocks. A plain GEMM has one tap. A k=3
  // convolution over a zero-padded time-major activation has three: tap t reads the A tile at
  // coordinates (tap_c0[t] + kc*64, row0 + tap_row[t]) and the W tile at column t*tap_k + kc*64.
  int taps, kblocks_per_tap, tap_k;
  int tap_c0[3], tap_row[3];
  // dynamic tile scheduler: sched[0] = next tile, sched[1] = CTAs done (the last one resets both for the next launch on
  // the stream). A CTA takes tiles as fast as it can drain them, so a launch that shares the GPU with another stream's
  // kernels (the other engine's decoder holds 64 SMs) needs no second wave of CTAs to cover the tile space.
  int* sched;
};

// ---------------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WT_WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
      "@P1 bra WT_WAIT_DONE;\n\t"
      "bra WT_WAIT_LOOP;\n\t"
      "WT_WAIT_DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity), "r"(0x989680)
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .b32 %%rx;\n\t"
      ".reg .pred %%px;\n\t"
      "elect.sync %%rx|%%px, %1;\n\t"
      "@%%px mov.s32 %0, 1;\n\t"
      "}"
      : "+r"(pred)
      : "r"(0xffffffffu));
  return pred != 0;
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// store of a {64 bf16, 32 rows, 1} box from shared memory; rows / columns outside the tensor map's extents are clipped
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, bf16 x bf16 -> f32
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128B-swizzled operand tile whose rows are 128 bytes (64 bf16): 8-row atoms of 1024 bytes.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);  // start address
  d |= (uint64_t)0 << 16;                      // leading byte offset: unused for swizzled K-major
  d |= (uint64_t)(1024 >> 4) << 32;            // stride byte offset between 8-row atoms
  d |= (uint64_t)1 << 46;                      // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                      // SWIZZLE_128B
  return d;
}
__host__ __device__ constexpr uint32_t make_idesc_bf16(int m, int n) {
  return (1u << 4)                    // D format f32
         | (1u << 7)                  // A format bf16
         | (1u << 10)                 // B format bf16
         | ((uint32_t)(n >> 3) << 17) // N / 8
         | ((uint32_t)(m >> 4) << 24);  // M / 16   (A, B both K-major: bits 15/16 = 0)
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

// Per-warp epilogue scratch on a 1024-byte boundary, used in one of two layouts. TMA store (EPI_TMA_C): 32 rows x 128
// bytes, 16-byte pieces XOR-swizzled by (row & 7) — what CU_TENSOR_MAP_SWIZZLE_128B expects of a {64 bf16, 32 rows} box,
// conflict-free for one-row-per-lane access. Otherwise: rows of 128 + 16 bytes (64 bf16 or 32 f32 + pad), which keeps both
// the row-per-lane and the coalesced four-pieces-per-row phases of a transpose conflict-free.
constexpr int kGemmScrBytes = 6144;   // three 2 KB boxes (EPI_TMA_ADD) / one 4 KB box (EPI_TMA_C) / 32 padded rows
constexpr int kGemmScrRow = 128 + 16;
__device__ __forceinline__ uint8_t* scr_piece(uint8_t* scr, int row, int chunk) {
  return scr + row * 128 + ((chunk ^ (row & 7)) << 4);
}

template <int BN>
struct GemmSmem {
  static constexpr int kStages = (BN <= 64) ? 6 : (BN <= 128) ? 5 : (BN <= 192) ? 4 : 3;
  static constexpr int kABytes = kGemmBM * kGemmBK * 2;  // 16 KB
  static constexpr int kBBytes = BN * kGemmBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kScratchBytes = kGemmEpiGroups * 4 * kGemmScrBytes;
  // bias staging: per epilogue group TWO buffers, alternating with the group's tiles. A warp that has finished tile i
  // writes tile i+2's bias (other n0) while a slower warp of the group may still be reading tile i's: with one buffer
  // that was a write-after-read race (the run-to-run token differences of round 1); with two, the writer of tile i+4's
  // bias has passed tile i+2's group barrier, which every warp reaches only after its last read of tile i's buffer.
  static constexpr int kBiasBytes = 4 * BN * 4;
  static constexpr int kTileQ = 4;      // tile-index queue producer -> MMA issuer + epilogue group (even: it & 1 keeps its group)
  static constexpr int kBarBytes = (2 * kStages + 4 + 2 * kTileQ + 24) * 8 + 16 + kTileQ * 4;
  static constexpr int kTotal = kStages * kStageBytes + kScratchBytes + kBiasBytes + kBarBytes + 1024;
  static constexpr int kTmemCols = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
};

// The epilogue of one 128 x BN accumulator tile for one warp (its 32 rows = its TMEM lane quarter): tcgen05.ld in
// 32-column chunks -> + bias -> GELU -> + addend -> bf16 (or f32) -> transpose through the warp's scratch -> 16-byte
// coalesced stores through the RowMap, or the K/V-stream scatter, or the fused (max, last index). Returns the row's
// packed winner (EPI_ARGMAX). Shared by the one-CTA and the CTA-pair kernels.
// EPI_TMA_ADD: the first two addend boxes of a tile, requested as soon as the tile is known (before its accumulator is)
__device__ __forceinline__ void gemm_epilogue_prefetch_add(const GemmBf16Params& p, const CUtensorMap* map_add, uint64_t* bar_add,
                                                           uint8_t* scr, int lane, int lane_q, int batch, int row0, int n0) {
  if ((p.flags & EPI_TMA_ADD) && lane == 0) {
    bulk_wait_read0();   // the previous tile's last two stores have read their boxes
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_expect_tx(bar_add + b, 2048);
      tma_load_3d(scr + b * 2048, map_add, bar_add + b, n0 + b * 32, row0 + lane_q * 32, batch);
    }
  }
  __syncwarp();
}

template <int BN>
__device__ __forceinline__ unsigned long long gemm_epilogue_tile(const GemmBf16Params& p, const CUtensorMap* map_c,
                                                                  const CUtensorMap* map_add, uint64_t* bar_add, uint32_t& add_ph,
                                                                  uint32_t t_row, const float* bias_s, uint8_t* scr, int lane,
                                                                  int lane_q, int batch, int row0, int n0) {
  const int r_in_tile = lane_q * 32 + lane;           // accumulator row owned by this thread
  const int r_batch = row0 + r_in_tile;               // row within the clip
  const bool row_ok = r_batch < p.rows_per_batch;
  const bool tma_c = (p.flags & EPI_TMA_C) != 0;
  unsigned long long best = 0ull;
  if (p.flags & EPI_TMA_ADD) {
    // Residual GEMMs: addend in and result out through 32 x 32 boxes (64-byte rows, 16-byte pieces XOR-swizzled by
    // (row >> 1) & 3 = CU_TENSOR_MAP_SWIZZLE_64B), three boxes per warp in rotation. Chunks 0 and 1 were requested by
    // gemm_epilogue_prefetch_add before the accumulator was waited for; chunk c + 2's addend is requested when chunk c's
    // result has been handed to the copy engine (into the box chunk c - 1 went out of: its store has had a whole chunk
    // to read it). A lane adds its row in place and the box goes back out with one bulk store.
    constexpr int kChunks = BN / 32;
    const int rowq = row0 + lane_q * 32;
    int b = 0;   // c % 3
#pragma unroll 1
    for (int c = 0; c < kChunks; ++c) {
      uint32_t v[32];
      tmem_ld32(t_row + c * 32, v);
      tmem_ld_wait();
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
      if (p.flags & EPI_BIAS) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b4 = *reinterpret_cast<const float4*>(bias_s + c * 32 + j);
          f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
        }
      }
      if (p.flags & EPI_GELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = gelu_tanh_fit(f[j]);
      }
      mbar_wait(bar_add + b, (add_ph >> b) & 1u);
      add_ph ^= 1u << b;
      uint8_t* box = scr + b * 2048 + lane * 64;
      const int sw = (lane >> 1) & 3;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4* pc = reinterpret_cast<uint4*>(box + ((i ^ sw) << 4));
        const uint4 val = *pc;
        const uint32_t w[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          f[i * 8 + 2 * j] += __uint_as_float(w[j] << 16);
          f[i * 8 + 2 * j + 1] += __uint_as_float(w[j] & 0xffff0000u);
        }
        *pc = make_uint4(pack_bf16x2(f[8 * i], f[8 * i + 1]), pack_bf16x2(f[8 * i + 2], f[8 * i + 3]),
                         pack_bf16x2(f[8 * i + 4], f[8 * i + 5]), pack_bf16x2(f[8 * i + 6], f[8 * i + 7]));
      }
      fence_proxy_async();
      __syncwarp();
      const int bn = b == 2 ? 0 : b + 1, bp = b == 0 ? 2 : b - 1;
      if (lane == 0) {
        tma_store_3d(map_c, scr + b * 2048, n0 + c * 32, rowq, batch);
        bulk_commit();
        if (c + 2 < kChunks) {
          asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // every store but chunk c's has read its box
          mbar_expect_tx(bar_add + bp, 2048);
          tma_load_3d(scr + bp * 2048, map_add, bar_add + bp, n0 + (c + 2) * 32, rowq, batch);
        }
      }
      __syncwarp();
      b = bn;
    }
    return 0ull;
  }
  // a tile never straddles a batch: one RowMap division per tile, rows are then row_stride apart
  const long long c_base = p.cmap.at(batch * p.rows_per_batch + row0);
  const long long a_base = (p.flags & EPI_ADD) ? p.add_map.at(batch * p.rows_per_batch + row0) : 0ll;

  for (int c0 = 0; c0 < BN; c0 += 32) {
    // the addend's 32 rows x 32 cols of this warp (coalesced 16-byte pieces) are requested first: their global
    // latency runs underneath the TMEM load and the bias / GELU math
    uint4 addv[4];
    if (p.flags & EPI_ADD) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int piece = i * 32 + lane;          // 128 pieces of 8 bf16: 4 per row
        const int rr = piece >> 2, pc = piece & 3;
        const int rb = row0 + lane_q * 32 + rr;
        const int col = n0 + c0 + pc * 8;
        addv[i] = make_uint4(0u, 0u, 0u, 0u);
        if (rb < p.rows_per_batch && col < p.N)
          addv[i] = *reinterpret_cast<const uint4*>(p.addend + a_base + (long long)(rb - row0) * p.add_map.row_stride + col);
      }
    }
    uint32_t v[32];
    tmem_ld32(t_row + c0, v);
    tmem_ld_wait();
    float f[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
    if (p.flags & EPI_BIAS) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bias_s + c0 + j);
        f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
      }
    }
    if (p.flags & EPI_GELU) {
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = gelu_tanh_fit(f[j]);
    }
    // TMA store: columns c0 .. c0+31 are the left (half 0) or right half of a 64-column store box; the box of the
    // previous pair must have been read out of the scratch before anything is written into it again
    const int half4 = tma_c ? ((c0 >> 3) & 4) : 0;
    if (tma_c && half4 == 0) {
      if (lane == 0) bulk_wait_read0();
      __syncwarp();
    }
    if (p.flags & EPI_ADD) {
      // coalesced fetch of this warp's 32 rows x 32 cols of the addend through the scratch
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int piece = i * 32 + lane;
        const int rr = piece >> 2, pc = piece & 3;
        *reinterpret_cast<uint4*>(tma_c ? scr_piece(scr, rr, half4 + pc) : scr + rr * kGemmScrRow + pc * 16) = addv[i];
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 val = *reinterpret_cast<const uint4*>(tma_c ? scr_piece(scr, lane, half4 + i) : scr + lane * kGemmScrRow + i * 16);
        const uint32_t w[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          f[i * 8 + 2 * j] += __uint_as_float(w[j] << 16);
          f[i * 8 + 2 * j + 1] += __uint_as_float(w[j] & 0xffff0000u);
        }
      }
      __syncwarp();
    }
    if (p.flags & EPI_ARGMAX) {
      if (row_ok) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int n = n0 + c0 + j;
          if (n < p.N) best = u64_max(best, pack_val_idx(f[j], (unsigned)n));
        }
      }
    }
    if (p.flags & EPI_OUT_F32) {
      // f32 output: 32 cols = 128 bytes per row
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<float4*>(scr + lane * kGemmScrRow + i * 16) = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int piece = i * 32 + lane;  // 256 pieces of 4 f32: 8 per row
        const int rr = piece >> 3, pc = piece & 7;
        const int rb = row0 + lane_q * 32 + rr;
        const int col = n0 + c0 + pc * 4;
        if (rb < p.rows_per_batch && col < p.N) {
          const float4 val = *reinterpret_cast<const float4*>(scr + rr * kGemmScrRow + pc * 16);
          float* dst = reinterpret_cast<float*>(p.C) + c_base + (long long)(rb - row0) * p.cmap.row_stride + col;
          if (col + 3 < p.N && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
            *reinterpret_cast<float4*>(dst) = val;
          } else {
            const float tmp[4] = {val.x, val.y, val.z, val.w};
            for (int q = 0; q < 4; ++q)
              if (col + q < p.N) dst[q] = tmp[q];
          }
        }
      }
      __syncwarp();
    } else if (tma_c) {
      // bf16 output through TMA: this lane's 64 bytes into its row of the swizzled box; every second chunk one lane
      // hands the 32 x 64 box to the copy engine (rows beyond the clip and columns beyond N are clipped by the map)
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(scr_piece(scr, lane, half4 + i)) =
            make_uint4(pack_bf16x2(f[8 * i], f[8 * i + 1]), pack_bf16x2(f[8 * i + 2], f[8 * i + 3]),
                       pack_bf16x2(f[8 * i + 4], f[8 * i + 5]), pack_bf16x2(f[8 * i + 6], f[8 * i + 7]));
      if (half4 != 0) {
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(map_c, scr, n0 + c0 - 32, row0 + lane_q * 32, batch);
          bulk_commit();
        }
      }
    } else if (p.C != nullptr || (p.flags & EPI_SPLIT_KV)) {
      // bf16 output: 32 cols = 64 bytes per row
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(scr + lane * kGemmScrRow + i * 16) =
            make_uint4(pack_bf16x2(f[8 * i], f[8 * i + 1]), pack_bf16x2(f[8 * i + 2], f[8 * i + 3]),
                       pack_bf16x2(f[8 * i + 4], f[8 * i + 5]), pack_bf16x2(f[8 * i + 6], f[8 * i + 7]));
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int piece = i * 32 + lane;
        const int rr = piece >> 2, pc = piece & 3;
        const int rb = row0 + lane_q * 32 + rr;
        const int col = n0 + c0 + pc * 8;
        if (rb < p.rows_per_batch && col < p.N) {
          const uint4 val = *reinterpret_cast<const uint4*>(scr + rr * kGemmScrRow + pc * 16);
          if (p.flags & EPI_SPLIT_KV) {
            int clip = batch, sq = rb;     // launched per clip (rows_per_batch == kv_S): no division
            if (p.rows_per_batch != p.kv_S) {
              const int mm = batch * p.rows_per_batch + rb;
              clip = mm / p.kv_S;
              sq = mm - clip * p.kv_S;
            }
            const int dm = p.kv_H * 64;
            const int which = col >= dm, cc = which ? col - dm : col;
            __nv_bfloat16* dst;
            if (p.kv_chunk > 0) {
              int ch = sq >> 7, kin = sq & 127;
              if (p.kv_chunk != 128) {
                ch = sq / p.kv_chunk;
                kin = sq - ch * p.kv_chunk;
              }
              // 16-byte segments of a key's 128-byte row are XOR-swizzled by key & 7: the decoder reads the chunk
              // with ldmatrix straight out of the bulk-copied stage (decoder_stream_sm100.cuh: ds_cross_attn)
              dst = p.kv_k + ((((size_t)(clip * p.kv_H + (cc >> 6)) * p.kv_cpp + ch) * 2 + which) * p.kv_chunk + kin) * 64 +
                    (((((cc & 63) >> 3) ^ (kin & 7)) << 3));
            } else {
              dst = (which ? p.kv_v : p.kv_k) + (((size_t)(clip * p.kv_H + (cc >> 6)) * p.kv_S + sq) * 64 + (cc & 63));
            }
            *reinterpret_cast<uint4*>(dst) = val;
          } else {
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.C) + c_base +
                                      (long long)(rb - row0) * p.cmap.row_stride + col) = val;
          }
        }
      }
      __syncwarp();
    }
  }
  return best;
}

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_sm100_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                       const __grid_constant__ CUtensorMap map_c, const __grid_constant__ CUtensorMap map_add,
    GemmBf16Params p) {
  using S = GemmSmem<BN>;
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  uint8_t* s_stage = smem;
  uint8_t* s_scratch = s_stage + S::kStages * S::kStageBytes;
  float* s_bias = reinterpret_cast<float*>(s_scratch + S::kScratchBytes);
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(s_bias) + S::kBiasBytes);
  uint64_t* bar_empty = bar_full + S::kStages;
  uint64_t* bar_tfull = bar_empty + S::kStages;   // [2]
  uint64_t* bar_tempty = bar_tfull + 2;           // [2]
  uint64_t* bar_qfull = bar_tempty + 2;           // [kTileQ] tile index published
  uint64_t* bar_qempty = bar_qfull + S::kTileQ;   // [kTileQ] read by the MMA issuer and the four warps of one epilogue group
  uint64_t* bar_addend = bar_qempty + S::kTileQ;  // [8 epilogue warps][3] addend boxes landed (EPI_TMA_ADD)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_addend + 24);
  volatile int* s_tileq = reinterpret_cast<volatile int*>(s_tmem + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tiles_m = p.tiles_per_batch * p.n_batches;
  const int n_tiles = n_tiles_m * p.n_tiles_n;
  const int n_kblocks = p.taps * p.kblocks_per_tap;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w);
    if (p.flags & EPI_TMA_C) tma_prefetch_desc(&map_c);
    if (p.flags & EPI_TMA_ADD) tma_prefetch_desc(&map_add);
    for (int i = 0; i < S::kStages; ++i) {
      mbar_init(bar_full + i, 1);
      mbar_init(bar_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_tfull + i, 1);
      mbar_init(bar_tempty + i, 4);  // one arrive per epilogue warp
    }
    for (int i = 0; i < S::kTileQ; ++i) {
      mbar_init(bar_qfull + i, 1);
      mbar_init(bar_qempty + i, 5);  // MMA issuer + 4 epilogue warps
    }
    for (int i = 0; i < 24; ++i) mbar_init(bar_addend + i, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, S::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int n_end = 0;
      // The first tile of a CTA is its block index; every further one is drawn from the counter ONE TILE AHEAD: the
      // atomic's round trip to L2 runs underneath the current tile's loads instead of in front of the next tile's.
      int next_tile = (int)blockIdx.x;
      for (int it = 0;; ++it) {
        int tile = (n_end == 0) ? next_tile : n_tiles;
        if (tile >= n_tiles) tile = -1;
        if (tile >= 0) next_tile = (int)gridDim.x + atomicAdd(p.sched, 1);
        // publish: slot it % kTileQ (free once the MMA issuer and the epilogue group of tile it - kTileQ have read it)
        const int qs = it % S::kTileQ;
        mbar_wait(bar_qempty + qs, ((uint32_t)(it / S::kTileQ) & 1u) ^ 1u);
        s_tileq[qs] = tile;
        mbar_arrive(bar_qfull + qs);   // release: the store above is visible to whoever observes the phase
        if (tile < 0) {
          if (++n_end == 2) break;     // one end marker per epilogue group
          continue;
        }
        const int tn = tile % p.n_tiles_n, tm = tile / p.n_tiles_n;
        const int batch = tm / p.tiles_per_batch, row0 = (tm % p.tiles_per_batch) * kGemmBM;
        for (int kb = 0; kb < n_kblocks; ++kb) {
          mbar_wait(bar_empty + stage, phase ^ 1);
          uint8_t* sa = s_stage + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          mbar_expect_tx(bar_full + stage, S::kStageBytes);
          const int tap = kb / p.kblocks_per_tap, kc = kb - tap * p.kblocks_per_tap;
          tma_load_3d(sa, &map_a, bar_full + stage, p.tap_c0[tap] + kc * kGemmBK, row0 + p.tap_row[tap], batch);
          tma_load_2d(sb, &map_w, bar_full + stage, tap * p.tap_k + kc * kGemmBK, tn * BN);
          if (++stage == S::kStages) { stage = 0; phase ^= 1; }
        }
      }
      // last CTA out re-arms the scheduler for the next launch on this stream
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
      mbar_wait(bar_tempty + acc, acc_phase ^ 1);  // epilogue has drained this accumulator
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
          for (int k = 0; k < kGemmBK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle atom: +2 in the (addr >> 4) field
            umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(bar_empty + stage);                       // slot reusable once these MMAs retire
          if (kb == n_kblocks - 1) umma_commit(bar_tfull + acc);  // accumulator complete
        }
        __syncwarp();
        if (++stage == S::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ================= epilogue (warps 2..9: two groups of four, one accumulator stage each) =================
    const int ew = warp - 2;            // 0..7 (scratch slot)
    const int eg = ew >> 2;             // epilogue group: drains tiles with (it & 1) == eg
    const int lane_q = warp & 3;        // TMEM lane quarter this warp may touch: lanes 32*(warp%4)..+31
    const int et = (threadIdx.x - 64) & 127;  // 0..127 within the group
    uint8_t* scr = s_scratch + ew * kGemmScrBytes;
    uint64_t* bar_add = bar_addend + ew * 3;   // this warp's three addend boxes (EPI_TMA_ADD)
    uint32_t add_ph = 0;
    for (int it = eg;; it += 2) {   // this group's tiles: it & 1 == eg
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
      float* bias_s = s_bias + (acc * 2 + ((it >> 1) & 1)) * BN;
      if (p.flags & EPI_BIAS)
        for (int i = et; i < BN; i += 128) bias_s[i] = (n0 + i < p.N) ? __ldg(p.bias + n0 + i) : 0.0f;
      gemm_epilogue_prefetch_add(p, &map_add, bar_add, scr, lane, lane_q, batch, row0, n0);
      asm volatile("bar.sync %0, 128;" ::"r"(1 + eg) : "memory");
      mbar_wait(bar_tfull + acc, acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(lane_q * 32) << 16) + acc * BN;
      const unsigned long long best = gemm_epilogue_tile<BN>(p, &map_c, &map_add, bar_add, add_ph, t_row, bias_s, scr, lane, lane_q, batch, row0, n0);
      // accumulator fully read: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + acc);
      if (p.flags & EPI_ARGMAX) {
        const int r_batch = row0 + lane_q * 32 + lane;
        if (r_batch < p.rows_per_batch) p.argmax_part[(long long)(batch * p.rows_per_batch + r_batch) * p.n_tiles_n + tn] = best;
      }
    }
    // the copy engine has read (and written) this warp's last boxes before the CTA's shared memory goes away
    if ((p.flags & EPI_TMA_C) && lane == 0) bulk_wait0();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, S::kTmemCols);
  }
}

// ---------------------------------------------------------------------------------------------------
// CTA-pair variant (tcgen05 cta_group::2): a cluster of two CTAs computes a 256 x BN tile. Each CTA loads ITS 128 rows
// of A and ITS half of the W tile (BN/2 rows); one tcgen05.mma.cta_group::2 (M = 256) issued by the leader drives both
// SMs' tensor cores, each reading its own A and only half of W from its own shared memory — the operand traffic per
// SM drops from 12 KB to 8 KB per MMA and the TMA writes from 48 KB to 32 KB per k-block, which is what the one-CTA
// kernel above is bound by at K = 384 (192 B/clk of operand traffic against a 128 B/clk shared-memory port). The
// accumulator of a CTA's 128 rows lives in its own tensor memory (2 stages), so the epilogue is unchanged.
// Plumbing that differs: TMA loads of both CTAs complete on the LEADER's full barrier (.cta_group::2, the barrier's
// address mapped into CTA 0); tcgen05.commit multicasts to both CTAs' empty / accumulator-full barriers; the peer's
// epilogue warps release an accumulator stage on the leader's barrier (remote arrive); the leader's producer draws the
// tile from the global counter and hands it to the peer's producer through a 4-entry inbox in the peer's shared memory.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {   // acquire at cluster scope: data a peer wrote
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WT_WAITC_LOOP:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra WT_WAITC_DONE;\n\t"
      "bra WT_WAITC_LOOP;\n\t"
      "WT_WAITC_DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma2_load_2d(void* smem_dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_3d(void* smem_dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs of the pair once all previously issued tcgen05.mma have completed
__device__ __forceinline__ void umma2_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}

template <int BN>
struct Gemm2Smem {
  static constexpr int kABytes = kGemmBM * kGemmBK * 2;      // this CTA's 128 rows of A: 16 KB
  static constexpr int kBBytes = (BN / 2) * kGemmBK * 2;     // this CTA's half of the W tile
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kScratchBytes = kGemmEpiGroups * 4 * kGemmScrBytes;
  static constexpr int kBiasBytes = 4 * BN * 4;
  static constexpr int kTileQ = 4;
  static constexpr int kBarBytes = 640;
  static constexpr int kFit = (232448 - 1024 - kScratchBytes - kBiasBytes - kBarBytes) / kStageBytes;
  static constexpr int kStages = kFit > 8 ? 8 : kFit;
  static constexpr int kTotal = kStages * kStageBytes + kScratchBytes + kBiasBytes + kBarBytes + 1024;
  static constexpr int kTmemCols = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
};

// p.tiles_per_batch = ceil(rows_per_batch / 256) here; map_w has a box of BN/2 rows. grid = 2 x pairs, cluster (2,1,1).
template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm2_bf16_sm100_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                        const __grid_constant__ CUtensorMap map_c, const __grid_constant__ CUtensorMap map_add,
    GemmBf16Params p) {
  using S = Gemm2Smem<BN>;
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  uint8_t* s_stage = smem;
  uint8_t* s_scratch = s_stage + S::kStages * S::kStageBytes;
  float* s_bias = reinterpret_cast<float*>(s_scratch + S::kScratchBytes);
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(s_bias) + S::kBiasBytes);   // [8] (used in the leader)
  uint64_t* bar_empty = bar_full + 8;              // [8] both CTAs (multicast commit)
  uint64_t* bar_tfull = bar_empty + 8;             // [2] both CTAs (multicast commit)
  uint64_t* bar_tempty = bar_tfull + 2;            // [2] leader: 4 local + 4 remote epilogue warps
  uint64_t* bar_qfull = bar_tempty + 2;            // [kTileQ] local tile queue
  uint64_t* bar_qempty = bar_qfull + S::kTileQ;    // [kTileQ]
  uint64_t* bar_ifull = bar_qempty + S::kTileQ;    // [kTileQ] peer: inbox entry written by the leader's producer
  uint64_t* bar_iempty = bar_ifull + S::kTileQ;    // [kTileQ] leader: inbox entry read by the peer's producer
  uint64_t* bar_addend = bar_iempty + S::kTileQ;   // [8 epilogue warps][3] addend boxes landed (EPI_TMA_ADD)
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_addend + 24);
  volatile int* s_tileq = reinterpret_cast<volatile int*>(s_tmem + 4);
  volatile int* s_inbox = s_tileq + S::kTileQ;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int n_tiles_m = p.tiles_per_batch * p.n_batches;   // 256-row tiles
  const int n_tiles = n_tiles_m * p.n_tiles_n;
  const int n_kblocks = p.taps * p.kblocks_per_tap;
  const int n_pairs = (int)gridDim.x >> 1, pair = (int)blockIdx.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w);
    if (p.flags & EPI_TMA_C) tma_prefetch_desc(&map_c);
    if (p.flags & EPI_TMA_ADD) tma_prefetch_desc(&map_add);
    for (int i = 0; i < 8; ++i) {
      mbar_init(bar_full + i, 1);
      mbar_init(bar_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_tfull + i, 1);
      mbar_init(bar_tempty + i, 8);
    }
    for (int i = 0; i < S::kTileQ; ++i) {
      mbar_init(bar_qfull + i, 1);
      mbar_init(bar_qempty + i, leader ? 5 : 4);   // (MMA issuer +) 4 epilogue warps
      mbar_init(bar_ifull + i, 1);
      mbar_init(bar_iempty + i, 1);
    }
    for (int i = 0; i < 24; ++i) mbar_init(bar_addend + i, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"((uint32_t)S::kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // both CTAs' barriers exist before any remote arrive / multicast commit / cross-CTA TMA signal
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ================= TMA producer (both CTAs; the leader also schedules) =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int n_end = 0;
      int next_tile = pair;
      const uint32_t peer_inbox = mapa_u32(smem_u32(const_cast<int*>(s_inbox)), 1);
      for (int it = 0;; ++it) {
        int tile;
        const int is = it % S::kTileQ;
        const uint32_t ipar = (uint32_t)(it / S::kTileQ) & 1u;
        if (leader) {
          tile = (n_end == 0) ? next_tile : n_tiles;
          if (tile >= n_tiles) tile = -1;
          if (tile >= 0) next_tile = n_pairs + atomicAdd(p.sched, 1);
          // hand the tile to the peer's producer (inbox slot free once the peer has read entry it - kTileQ)
          mbar_wait_cluster(bar_iempty + is, ipar ^ 1u);
          asm volatile("st.shared::cluster.b32 [%0], %1;" ::"r"(peer_inbox + is * 4), "r"(tile) : "memory");
          mbar_arrive_remote(mapa_u32(smem_u32(bar_ifull + is), 1));
        } else {
          mbar_wait_cluster(bar_ifull + is, ipar);
          tile = s_inbox[is];
          mbar_arrive_remote(mapa_u32(smem_u32(bar_iempty + is), 0));
        }
        const int qs = it % S::kTileQ;
        mbar_wait(bar_qempty + qs, ((uint32_t)(it / S::kTileQ) & 1u) ^ 1u);
        s_tileq[qs] = tile;
        mbar_arrive(bar_qfull + qs);
        if (tile < 0) {
          if (++n_end == 2) break;     // one end marker per epilogue group
          continue;
        }
        const int tn = tile % p.n_tiles_n, tm = tile / p.n_tiles_n;
        const int batch = tm / p.tiles_per_batch, row0 = (tm % p.tiles_per_batch) * (2 * kGemmBM) + (int)rank * kGemmBM;
        for (int kb = 0; kb < n_kblocks; ++kb) {
          mbar_wait(bar_empty + stage, phase ^ 1);
          uint8_t* sa = s_stage + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          const uint32_t lbar = mapa_u32(smem_u32(bar_full + stage), 0);
          if (leader) mbar_expect_tx(bar_full + stage, 2 * S::kStageBytes);   // both CTAs' bytes land on this barrier
          const int tap = kb / p.kblocks_per_tap, kc = kb - tap * p.kblocks_per_tap;
          tma2_load_3d(sa, &map_a, lbar, p.tap_c0[tap] + kc * kGemmBK, row0 + p.tap_row[tap], batch);
          tma2_load_2d(sb, &map_w, lbar, tap * p.tap_k + kc * kGemmBK, tn * BN + (int)rank * (BN / 2));
          if (++stage == S::kStages) { stage = 0; phase ^= 1; }
        }
      }
      if (leader) {   // last pair out re-arms the scheduler for the next launch on this stream
        __threadfence();
        if (atomicAdd(p.sched + 1, 1) == n_pairs - 1) {
          p.sched[0] = 0;
          p.sched[1] = 0;
          __threadfence();
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (leader only) =================
    if (leader) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * kGemmBM, BN);
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
        mbar_wait_cluster(bar_tempty + acc, acc_phase ^ 1);  // both CTAs' epilogues have drained this accumulator
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
            for (int k = 0; k < kGemmBK / 16; ++k) umma2_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            umma2_commit(bar_empty + stage);                       // both CTAs' slots reusable once these MMAs retire
            if (kb == n_kblocks - 1) umma2_commit(bar_tfull + acc);  // both CTAs' accumulator halves complete
          }
          __syncwarp();
          if (++stage == S::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ================= epilogue (both CTAs: their own 128 rows) =================
    const int ew = warp - 2;
    const int eg = ew >> 2;
    const int lane_q = warp & 3;
    const int et = (threadIdx.x - 64) & 127;
    uint8_t* scr = s_scratch + ew * kGemmScrBytes;
    uint64_t* bar_add = bar_addend + ew * 3;   // this warp's three addend boxes (EPI_TMA_ADD)
    uint32_t add_ph = 0;
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
      const int batch = tm / p.tiles_per_batch, row0 = (tm % p.tiles_per_batch) * (2 * kGemmBM) + (int)rank * kGemmBM;
      const int n0 = tn * BN;
      float* bias_s = s_bias + (acc * 2 + ((it >> 1) & 1)) * BN;
      if (p.flags & EPI_BIAS)
        for (int i = et; i < BN; i += 128) bias_s[i] = (n0 + i < p.N) ? __ldg(p.bias + n0 + i) : 0.0f;
      gemm_epilogue_prefetch_add(p, &map_add, bar_add, scr, lane, lane_q, batch, row0, n0);
      asm volatile("bar.sync %0, 128;" ::"r"(1 + eg) : "memory");
      mbar_wait(bar_tfull + acc, acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(lane_q * 32) << 16) + acc * BN;
      const unsigned long long best = gemm_epilogue_tile<BN>(p, &map_c, &map_add, bar_add, add_ph, t_row, bias_s, scr, lane, lane_q, batch, row0, n0);
      // accumulator fully read: hand it back to the leader's MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(mapa_u32(smem_u32(bar_tempty + acc), 0));
      if (p.flags & EPI_ARGMAX) {
        const int r_batch = row0 + lane_q * 32 + lane;
        if (r_batch < p.rows_per_batch) p.argmax_part[(long long)(batch * p.rows_per_batch + r_batch) * p.n_tiles_n + tn] = best;
      }
    }
    // the copy engine has read (and written) this warp's last boxes before the CTA's shared memory goes away
    if ((p.flags & EPI_TMA_C) && lane == 0) bulk_wait0();
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // no CTA leaves (or frees tensor memory) while its peer may still signal it
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)S::kTmemCols) : "memory");
  }
}

}  // namespace wt
