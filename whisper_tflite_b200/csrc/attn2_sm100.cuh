// Encoder self-attention for sm_100a, two CTAs per SM (the default; attn_sm100.cuh keeps the one-CTA-per-SM kernel as
// WT_ATTN=1cta): softmax(Q K^T / sqrt(64)) V per (clip, head), head_dim 64, S = 1500 keys, non-causal
// (whisper.cpp:295; graph: export/generate_onnx.py:91-99 AudioEncoder).
//
// Same roles as attn_sm100.cuh (TMA warp, tcgen05 issuer, 4 softmax warps with one thread per query row), on half the
// resources — Q 16 KB + ONE P tile 32 KB + a 2-stage K/V ring 64 KB = 112 KB of shared memory, 256 TMEM columns (one S
// buffer, two O buffers), <= 168 registers — so that two CTAs share an SM. Inside a CTA the chain per key block is
// strictly serial (S_j -> softmax_j -> P_j -> PV_j | QK_{j+1} -> softmax_{j+1}); the overlap that keeps the tensor pipe
// and the 16 MUFU lanes busy comes from the OTHER resident CTA, which is at a different point of its own chain. At
// head_dim 64 a 128 x 128 block costs 512 clk of MMA and 1024 clk of ex2: two desynchronised CTAs are what one CTA
// with one softmax group cannot be.
#pragma once

#include "attn_sm100.cuh"

namespace wt {

// Softmax passes of one 128-key block for one query row, S read from TMEM in 32-column pieces. Both passes keep the NEXT
// piece's tcgen05.ld in flight while the current one is processed (tcgen05.wait::ld covers every outstanding load, so
// the order is wait -> issue next -> compute); the mask of a clip's last key block is a template flag behind a real
// branch — as a run-time condition inside the unrolled body ptxas if-converts it into three predicated-off
// instructions per element, which cost issue slots in every full block.
template <bool kMasked>
__device__ __forceinline__ float at2_piece_max(const uint32_t (&v)[32], int c, int valid) {
  if (!kMasked) {
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      m0 = fmaxf(m0, __uint_as_float(v[i]));
      m1 = fmaxf(m1, __uint_as_float(v[i + 1]));
    }
    return fmaxf(m0, m1);
  } else {
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, (c * 32 + i < valid) ? __uint_as_float(v[i]) : -INFINITY);
    return mx;
  }
}
// pass 1: row maximum of the block (raw scores)
template <bool kMasked, bool kPrefetch>
__device__ __forceinline__ float at2_pass1(uint32_t t_s, int valid) {
  float mx;
  if (kPrefetch) {
    uint32_t va[32], vb[32];
    tmem_ld32(t_s, va);
    tmem_ld_wait();
    tmem_ld32(t_s + 32, vb);
    mx = at2_piece_max<kMasked>(va, 0, valid);
    tmem_ld_wait();
    tmem_ld32(t_s + 64, va);
    mx = fmaxf(mx, at2_piece_max<kMasked>(vb, 1, valid));
    tmem_ld_wait();
    tmem_ld32(t_s + 96, vb);
    mx = fmaxf(mx, at2_piece_max<kMasked>(va, 2, valid));
    tmem_ld_wait();
    mx = fmaxf(mx, at2_piece_max<kMasked>(vb, 3, valid));
  } else {
    mx = -INFINITY;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t v[32];
      tmem_ld32(t_s + c * 32, v);
      tmem_ld_wait();
      mx = fmaxf(mx, at2_piece_max<kMasked>(v, c, valid));
    }
  }
  return mx;
}
// one piece of pass 2: P = exp2(s * c - m) -> bf16 into the 128B-swizzled P tile (32-bit shared addresses); returns the
// piece's row sum. 32 keys = 4 chunks of 16 bytes; key piece c lies in sub-tile c/2, chunk index (c&1)*4 + q.
template <bool kMasked>
__device__ __forceinline__ float at2_piece_exp(uint32_t (&v)[32], int c, uint32_t p_row_s, int sw, float c_scale, float m_new,
                                               int valid, float& mx) {
  float la = 0.0f, lb = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    float p0 = ex2_approx(fmaf(__uint_as_float(v[i]), c_scale, -m_new));
    float p1 = ex2_approx(fmaf(__uint_as_float(v[i + 1]), c_scale, -m_new));
    if (kMasked) {
      if (c * 32 + i >= valid) p0 = 0.0f; else mx = fmaxf(mx, __uint_as_float(v[i]));
      if (c * 32 + i + 1 >= valid) p1 = 0.0f; else mx = fmaxf(mx, __uint_as_float(v[i + 1]));
    } else {
      mx = fmaxf(fmaxf(mx, __uint_as_float(v[i])), __uint_as_float(v[i + 1]));   // one FMNMX3
    }
    la += p0;
    lb += p1;
    v[i >> 1] = pack_bf16x2(p0, p1);   // in place: word i/2 is dead by now (i/2 <= i)
  }
  const uint32_t sub = p_row_s + (uint32_t)((c >> 1) * kAtTile);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int chunk = ((c & 1) * 4 + q) ^ sw;
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sub + chunk * 16), "r"(v[4 * q]), "r"(v[4 * q + 1]),
                 "r"(v[4 * q + 2]), "r"(v[4 * q + 3])
                 : "memory");
  }
  return la + lb;
}
template <bool kMasked, bool kPrefetch>
__device__ __forceinline__ float at2_pass2(uint32_t t_s, uint32_t p_row_s, int sw, float c_scale, float m_new, int valid,
                                           float& mx) {
  float l_blk = 0.0f;
  mx = -INFINITY;
  if (kPrefetch) {
    uint32_t va[32], vb[32];
    tmem_ld32(t_s, va);
    tmem_ld_wait();
    tmem_ld32(t_s + 32, vb);
    l_blk = at2_piece_exp<kMasked>(va, 0, p_row_s, sw, c_scale, m_new, valid, mx);
    tmem_ld_wait();
    tmem_ld32(t_s + 64, va);
    l_blk += at2_piece_exp<kMasked>(vb, 1, p_row_s, sw, c_scale, m_new, valid, mx);
    tmem_ld_wait();
    tmem_ld32(t_s + 96, vb);
    l_blk += at2_piece_exp<kMasked>(va, 2, p_row_s, sw, c_scale, m_new, valid, mx);
    tmem_ld_wait();
    l_blk += at2_piece_exp<kMasked>(vb, 3, p_row_s, sw, c_scale, m_new, valid, mx);
  } else {
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t v[32];
      tmem_ld32(t_s + c * 32, v);
      tmem_ld_wait();
      l_blk += at2_piece_exp<kMasked>(v, c, p_row_s, sw, c_scale, m_new, valid, mx);
    }
  }
  return l_blk;
}

constexpr bool kAt2Prefetch1 = false;
constexpr bool kAt2Prefetch2 = false;   // pass-2 prefetch needs 64 + 64 (O row) + temporaries > 168 registers: spills

constexpr int kAt2Threads = 192;
constexpr int kAt2Stages = 2;
struct Attn2Smem {
  static constexpr int kQ = 0;
  static constexpr int kP = kQ + kAtTile;                 // one buffer: 2 sub-tiles of 128 x 64 = 32 KB
  static constexpr int kKV = kP + 2 * kAtTile;            // kAt2Stages x (K tile + V tile)
  static constexpr int kBars = kKV + kAt2Stages * 2 * kAtTile;
  // 112 KB + barriers; 512 B of alignment slack (the dynamic window of a kernel without static shared memory starts
  // 1024-aligned; the kernel traps if it ever does not) keeps two CTAs within the SM's 227 KB
  static constexpr int kTotal = kBars + 256 + 512;
};

// grid = (ceil(S/128), H, B); block = 192
__global__ void __launch_bounds__(kAt2Threads, 2)
attention2_sm100_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, int S, int H) {
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  if (smem - smem_dyn > 512) __trap();   // see Attn2Smem::kTotal
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Attn2Smem::kBars);
  uint64_t* bar_q = bars;                       // 1
  uint64_t* bar_kv_full = bars + 1;             // kAt2Stages
  uint64_t* bar_kv_empty = bar_kv_full + kAt2Stages;
  uint64_t* bar_s_full = bar_kv_empty + kAt2Stages;  // 1 (phase = j & 1)
  uint64_t* bar_p_full = bar_s_full + 2;            // 1 (phase = j & 1)
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
    for (int i = 0; i < kAt2Stages; ++i) {
      mbar_init(bar_kv_full + i, 1);
      mbar_init(bar_kv_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s_full + i, 1);
      mbar_init(bar_p_full + i, 4);
      mbar_init(bar_o_full + i, 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  const uint32_t tmem_s = tmem_base;        // 128 columns
  const uint32_t tmem_o = tmem_base + 128;  // 2 x 64 columns

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar_q, kAtTile);
      tma_load_3d(smem + Attn2Smem::kQ, &map_qkv, bar_q, h * 64, q0, b);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < nkb; ++j) {
        mbar_wait(bar_kv_empty + stage, phase ^ 1);
        uint8_t* sk = smem + Attn2Smem::kKV + stage * 2 * kAtTile;
        mbar_expect_tx(bar_kv_full + stage, 2 * kAtTile);
        tma_load_3d(sk, &map_qkv, bar_kv_full + stage, D + h * 64, j * 128, b);
        tma_load_3d(sk + kAtTile, &map_qkv, bar_kv_full + stage, 2 * D + h * 64, j * 128, b);
        if (++stage == kAt2Stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = make_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = make_idesc_bf16(128, 64) | (1u << 16);  // B (= V) is MN-major
    const uint32_t q_addr = smem_u32(smem + Attn2Smem::kQ);
    auto issue_s = [&](int j) {
      const int stage = j % kAt2Stages;
      mbar_wait(bar_kv_full + stage, (uint32_t)((j / kAt2Stages) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint64_t da = make_sw128_desc(q_addr);
        const uint64_t db = make_sw128_desc(smem_u32(smem + Attn2Smem::kKV + stage * 2 * kAtTile));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tmem_s, da + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
        umma_commit(bar_s_full);
      }
      __syncwarp();
    };
    mbar_wait(bar_q, 0);
    issue_s(0);
    for (int j = 0; j < nkb; ++j) {
      const int stage = j % kAt2Stages;
      mbar_wait(bar_p_full, (uint32_t)(j & 1));   // P_j written, so S_j has been read: the S buffer is free again
      tc_fence_after();
      // S_{j+1} goes to the tensor pipe ahead of PV_j: the softmax warps start on block j+1 one MMA earlier, and PV_j
      // runs underneath their pass 1 (they wait for its commit before they overwrite P)
      if (j + 1 < nkb) issue_s(j + 1);
      if (elect_one()) {
        const uint32_t p_addr = smem_u32(smem + Attn2Smem::kP);
        const uint32_t v_addr = smem_u32(smem + Attn2Smem::kKV + stage * 2 * kAtTile + kAtTile);
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
    }
  } else {
    // ---- softmax + accumulate: thread <-> query row ----
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
    const float c_scale = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) * log2(e)
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.0f;
    float m_run = -INFINITY, m_seen = -INFINITY, l_run = 0.0f, alpha_prev = 0.0f;
    const uint32_t p_row_s = smem_u32(smem + Attn2Smem::kP + row * 128);
    const int sw = row & 7;
    for (int j = 0; j < nkb; ++j) {
      mbar_wait(bar_s_full, (uint32_t)(j & 1));
      tc_fence_after();
      const uint32_t t_s = tmem_s + lane_sel;
      const int valid = S - j * 128;  // keys valid in this block (>= 128 except the last)
      // Reference exponent of the block. Block 0: its exact row maximum (pass 1). Later blocks: the maximum of the blocks
      // before it — P_j = exp2(s*c - m_ref) may then exceed 1, which is exact mathematics as long as nothing overflows;
      // the block's own maximum comes out of the same pass and becomes the next reference. A row whose scores jump by
      // more than 2^64 against everything seen before (never observed) redoes the block with its exact maximum.
      float m_ref = m_seen;
      if (j == 0) {
        const float mx0 = valid >= 128 ? at2_pass1<false, kAt2Prefetch1>(t_s, valid) : at2_pass1<true, false>(t_s, valid);
        m_ref = mx0 * c_scale;
      }
      // S_j was issued AHEAD of PV_{j-1} (see the MMA warp), so PV_{j-1} may still be reading the P tile: its commit is
      // awaited here, normally long since arrived
      if (j >= 1) mbar_wait(bar_o_full + ((j - 1) & 1), (uint32_t)(((j - 1) >> 1) & 1));
      // P = exp2(s*c - m_ref), row sum, block maximum, bf16 -> swizzled shared tile
      float mx_blk;
      float l_blk = valid >= 128 ? at2_pass2<false, kAt2Prefetch2>(t_s, p_row_s, sw, c_scale, m_ref, valid, mx_blk)
                                 : at2_pass2<true, false>(t_s, p_row_s, sw, c_scale, m_ref, valid, mx_blk);
      // The redo is decided per WARP: tcgen05.ld is .sync.aligned, so all 32 lanes must take the branch together (a lane
      // on its own — the round-1 code — is undefined behaviour and hung the kernel in tests/test_gpu_kernels.py's score
      // jump case). Lanes whose row did not jump repeat the block with their unchanged reference: identical results.
      const bool jump = mx_blk * c_scale > m_ref + 64.0f;
      if (__any_sync(0xffffffffu, jump)) {
        if (jump) m_ref = mx_blk * c_scale;
        l_blk = at2_pass2<true, false>(t_s, p_row_s, sw, c_scale, m_ref, valid, mx_blk);
      }
      const float alpha = ex2_approx(m_run - m_ref);   // previous reference -> this block's
      l_run = l_run * alpha + l_blk;
      m_run = m_ref;
      m_seen = fmaxf(m_ref, mx_blk * c_scale);
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full);
      // fold the previous block's partial O into the register accumulator
      if (j >= 1) {
        const int jp = j - 1;
        mbar_wait(bar_o_full + (jp & 1), (uint32_t)((jp >> 1) & 1));
        tc_fence_after();
        const uint32_t t_o = tmem_o + lane_sel + (jp & 1) * 64;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld32(t_o + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], alpha_prev, __uint_as_float(v[i]));
        }
      }
      alpha_prev = alpha;
    }
    {
      const int jp = nkb - 1;
      mbar_wait(bar_o_full + (jp & 1), (uint32_t)((jp >> 1) & 1));
      tc_fence_after();
      const uint32_t t_o = tmem_o + lane_sel + (jp & 1) * 64;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld32(t_o + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], alpha_prev, __uint_as_float(v[i]));
      }
    }
    const int q = q0 + row;
    if (q < S) {
      const float inv = 1.0f / l_run;
      uint4* dst = reinterpret_cast<uint4*>(out + ((long long)b * S + q) * D + h * 64);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        dst[i] = make_uint4(pack_bf16x2(o[8 * i] * inv, o[8 * i + 1] * inv), pack_bf16x2(o[8 * i + 2] * inv, o[8 * i + 3] * inv),
                            pack_bf16x2(o[8 * i + 4] * inv, o[8 * i + 5] * inv), pack_bf16x2(o[8 * i + 6] * inv, o[8 * i + 7] * inv));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace wt
