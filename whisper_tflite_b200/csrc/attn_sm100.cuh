// Encoder self-attention for sm_100a: softmax(Q K^T / sqrt(64)) V per (clip, head), head_dim 64,
// S = 1500 keys, non-causal — the BATCH_MATMUL / SOFTMAX / BATCH_MATMUL the reference runs inside
// Interpreter::Invoke (whisper.cpp:295; graph: export/generate_onnx.py:91-99 AudioEncoder).
//
// One CTA owns a 128-query tile of one (clip, head) and walks the keys in blocks of 128:
//   warp 0      TMA: Q tile once, then K_j / V_j tiles (128 x 64 bf16, 128B swizzle) through a 4-stage ring;
//   warp 1      tcgen05 issuer: S_j = Q K_j^T (M128 N128 K64) into one of two TMEM S buffers, and
//               O_j = P_j V_j (M128 N64 K128; V consumed MN-major straight from its [key][dim] tile) into one
//               of two TMEM O buffers; S_{j+2} is issued right behind PV_j so QK^T of the next blocks overlaps
//               the softmax of the current one;
//   warps 2-5   softmax, one thread per query row: tcgen05.ld the S row, running max / sum in the exp2 domain,
//               P_j = exp2(...) as bf16 written into a 128B-swizzled shared tile (the A operand of PV_j),
//               and the per-block partial O_j is folded into a register accumulator with the usual
//               exp2(m_old - m_new) correction — so the MMA pipe never waits for a rescale.
// Keys past the clip end (1500 = 11*128 + 92) are zero-filled by TMA (3-D map {3d, 1500, B}) and masked.
#pragma once

#include "gemm_sm100.cuh"

namespace wt {

constexpr int kAtThreads = 64 + 256;   // TMA warp, MMA warp, 2 softmax warpgroups (keys 0-63 / 64-127 of every block)
constexpr int kAtStages = 4;
constexpr int kAtTile = 128 * 64 * 2;  // one 128 x 64 bf16 tile: 16 KB
struct AttnSmem {
  static constexpr int kQ = 0;
  static constexpr int kP = kQ + kAtTile;                 // 2 buffers x (2 sub-tiles of 128 x 64) = 64 KB
  static constexpr int kKV = kP + 4 * kAtTile;            // kAtStages x (K tile + V tile)
  static constexpr int kBars = kKV + kAtStages * 2 * kAtTile;
  static constexpr int kXch = kBars + 256;                // row-max exchange between the two key halves: f32 [2 blocks][2][128]
  static constexpr int kTotal = kXch + 2 * 2 * 128 * 4 + 1024;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// MN-major (N contiguous), 128B-swizzled B operand: rows are K (keys), 128 bytes of N (64 dims) per row.
__device__ __forceinline__ uint64_t make_sw128_mn_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)(16384 >> 4) << 16;  // leading byte offset: next 64-wide N atom (unused, N = 64)
  d |= (uint64_t)(1024 >> 4) << 32;   // stride byte offset: next group of 8 K rows
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// grid = (ceil(S/128), H, B); block = 192
__global__ void __launch_bounds__(kAtThreads, 1)
attention_sm100_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, int S, int H) {
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnSmem::kBars);
  uint64_t* bar_q = bars;                       // 1
  uint64_t* bar_kv_full = bars + 1;             // kAtStages
  uint64_t* bar_kv_empty = bar_kv_full + kAtStages;
  uint64_t* bar_s_full = bar_kv_empty + kAtStages;  // 2
  uint64_t* bar_p_full = bar_s_full + 2;            // 2
  uint64_t* bar_o_full = bar_p_full + 2;            // 2
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_o_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int D = H * 64;
  const int q0 = qt * 128;
  const int nkb = (S + 127) / 128;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    mbar_init(bar_q, 1);
    for (int i = 0; i < kAtStages; ++i) {
      mbar_init(bar_kv_full + i, 1);
      mbar_init(bar_kv_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s_full + i, 1);
      mbar_init(bar_p_full + i, 8);
      mbar_init(bar_o_full + i, 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const uint32_t tmem_s = tmem_base;        // 2 x 128 columns
  const uint32_t tmem_o = tmem_base + 256;  // 2 x 64 columns

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar_q, kAtTile);
      tma_load_3d(smem + AttnSmem::kQ, &map_qkv, bar_q, h * 64, q0, b);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < nkb; ++j) {
        mbar_wait(bar_kv_empty + stage, phase ^ 1);
        uint8_t* sk = smem + AttnSmem::kKV + stage * 2 * kAtTile;
        mbar_expect_tx(bar_kv_full + stage, 2 * kAtTile);
        tma_load_3d(sk, &map_qkv, bar_kv_full + stage, D + h * 64, j * 128, b);
        tma_load_3d(sk + kAtTile, &map_qkv, bar_kv_full + stage, 2 * D + h * 64, j * 128, b);
        if (++stage == kAtStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = make_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = make_idesc_bf16(128, 64) | (1u << 16);  // B (= V) is MN-major
    const uint32_t q_addr = smem_u32(smem + AttnSmem::kQ);
    auto issue_s = [&](int j) {
      const int stage = j % kAtStages;
      mbar_wait(bar_kv_full + stage, (uint32_t)((j / kAtStages) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint64_t da = make_sw128_desc(q_addr);
        const uint64_t db = make_sw128_desc(smem_u32(smem + AttnSmem::kKV + stage * 2 * kAtTile));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tmem_s + (j & 1) * 128, da + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
        umma_commit(bar_s_full + (j & 1));
      }
      __syncwarp();
    };
    mbar_wait(bar_q, 0);
    issue_s(0);
    if (nkb > 1) issue_s(1);
    for (int j = 0; j < nkb; ++j) {
      const int stage = j % kAtStages;
      mbar_wait(bar_p_full + (j & 1), (uint32_t)((j >> 1) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t p_addr = smem_u32(smem + AttnSmem::kP + (j & 1) * 2 * kAtTile);
        const uint32_t v_addr = smem_u32(smem + AttnSmem::kKV + stage * 2 * kAtTile + kAtTile);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          // A = P: sub-tile kk/4 (64 keys each, 16 KB apart), 32 bytes per K step inside the swizzle atom
          const uint64_t da = make_sw128_desc(p_addr + (kk >> 2) * kAtTile) + 2 * (kk & 3);
          // B = V: 16 keys per K step = 16 rows of 128 bytes
          const uint64_t db = make_sw128_mn_desc(v_addr + kk * 2048);
          umma_bf16(tmem_o + (j & 1) * 64, da, db, idesc_o, kk != 0 ? 1u : 0u);
        }
        umma_commit(bar_o_full + (j & 1));
        umma_commit(bar_kv_empty + stage);
      }
      __syncwarp();
      if (j + 2 < nkb) issue_s(j + 2);
    }
  } else {
    // ---- softmax + accumulate: two threads per query row, one per 64-key half of every block (sg); each folds half
    //      of the O columns. The halves meet once per block to agree on the row maximum (shared memory + a 64-thread
    //      named barrier per lane quarter) and once at the end for the row sum. ----
    const int sg = (warp - 2) >> 2;       // key half of the block / dim half of O
    const int quarter = warp & 3;         // TMEM lane quarter this warp may touch
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
    const float c_scale = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) * log2(e)
    float* xch = reinterpret_cast<float*>(smem + AttnSmem::kXch);
    float o[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = 0.0f;
    float m_run = -INFINITY, l_run = 0.0f, alpha_prev = 0.0f;
    uint8_t* p_row_base = smem + AttnSmem::kP + sg * kAtTile + row * 128;
    const int sw = row & 7;
    for (int j = 0; j < nkb; ++j) {
      mbar_wait(bar_s_full + (j & 1), (uint32_t)((j >> 1) & 1));
      tc_fence_after();
      const uint32_t t_s = tmem_s + lane_sel + (j & 1) * 128 + sg * 64;
      const int valid = S - j * 128 - sg * 64;  // keys of this half that exist (>= 64 except in the clip's last block)
      uint32_t sv[64];
      tmem_ld32(t_s, sv);
      tmem_ld32(t_s + 32, sv + 32);
      tmem_ld_wait();
      float mx0 = -INFINITY, mx1 = -INFINITY;
      if (valid >= 64) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          mx0 = fmaxf(mx0, __uint_as_float(sv[i]));
          mx1 = fmaxf(mx1, __uint_as_float(sv[32 + i]));
        }
      } else {
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          if (i >= valid) sv[i] = 0xff800000u;  // -inf: exp2 -> 0, never the maximum
          mx0 = fmaxf(mx0, __uint_as_float(sv[i]));
        }
      }
      float mx = fmaxf(mx0, mx1);
      float* xr = xch + (j & 1) * 256;
      xr[sg * 128 + row] = mx;
      asm volatile("bar.sync %0, 64;" ::"r"(1 + quarter) : "memory");
      mx = fmaxf(mx, xr[(sg ^ 1) * 128 + row]);
      const float m_new = fmaxf(m_run, mx * c_scale);
      const float alpha = ex2_approx(m_run - m_new);
      float l0 = 0.0f, l1 = 0.0f, l2 = 0.0f, l3 = 0.0f;
      uint8_t* p_row = p_row_base + (j & 1) * 2 * kAtTile;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          pv[i] = ex2_approx(fmaf(__uint_as_float(sv[c * 32 + i]), c_scale, -m_new));
          pv[i + 1] = ex2_approx(fmaf(__uint_as_float(sv[c * 32 + i + 1]), c_scale, -m_new));
          pv[i + 2] = ex2_approx(fmaf(__uint_as_float(sv[c * 32 + i + 2]), c_scale, -m_new));
          pv[i + 3] = ex2_approx(fmaf(__uint_as_float(sv[c * 32 + i + 3]), c_scale, -m_new));
          l0 += pv[i]; l1 += pv[i + 1]; l2 += pv[i + 2]; l3 += pv[i + 3];
        }
        // this half's 64 keys = 8 chunks of 16 bytes in sub-tile sg; keys c*32.. are chunks c*4 + q
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int chunk = (c * 4 + q) ^ sw;
          *reinterpret_cast<uint4*>(p_row + chunk * 16) =
              make_uint4(pack_bf16x2(pv[8 * q], pv[8 * q + 1]), pack_bf16x2(pv[8 * q + 2], pv[8 * q + 3]),
                         pack_bf16x2(pv[8 * q + 4], pv[8 * q + 5]), pack_bf16x2(pv[8 * q + 6], pv[8 * q + 7]));
        }
      }
      l_run = l_run * alpha + ((l0 + l1) + (l2 + l3));
      m_run = m_new;
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full + (j & 1));
      // fold the previous block's partial O (this thread's 32 dims) into the register accumulator
      if (j >= 1) {
        const int jp = j - 1;
        mbar_wait(bar_o_full + (jp & 1), (uint32_t)((jp >> 1) & 1));
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tmem_o + lane_sel + (jp & 1) * 64 + sg * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = fmaf(o[i], alpha_prev, __uint_as_float(v[i]));
      }
      alpha_prev = alpha;
    }
    {
      const int jp = nkb - 1;
      mbar_wait(bar_o_full + (jp & 1), (uint32_t)((jp >> 1) & 1));
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_o + lane_sel + (jp & 1) * 64 + sg * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = fmaf(o[i], alpha_prev, __uint_as_float(v[i]));
    }
    // row sum = sum of the two halves' partial sums (both carry the same running maximum)
    float* xl = xch + (nkb & 1) * 256;
    xl[sg * 128 + row] = l_run;
    asm volatile("bar.sync %0, 64;" ::"r"(1 + quarter) : "memory");
    l_run += xl[(sg ^ 1) * 128 + row];
    const int q = q0 + row;
    if (q < S) {
      const float inv = 1.0f / l_run;
      uint4* dst = reinterpret_cast<uint4*>(out + ((long long)b * S + q) * D + h * 64 + sg * 32);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        dst[i] = make_uint4(pack_bf16x2(o[8 * i] * inv, o[8 * i + 1] * inv), pack_bf16x2(o[8 * i + 2] * inv, o[8 * i + 3] * inv),
                            pack_bf16x2(o[8 * i + 4] * inv, o[8 * i + 5] * inv), pack_bf16x2(o[8 * i + 6] * inv, o[8 * i + 7] * inv));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace wt
