// Encoder self-attention for sm_100a, third design (the default): softmax(Q K^T / sqrt(64)) V per (clip, head),
// head_dim 64, S = 1500 keys, non-causal (whisper.cpp:295; graph: export/generate_onnx.py:91-99 AudioEncoder).
//
// What round 1's kernel (attn2_sm100.cuh: two CTAs per SM, one serial chain each) left on the table, per its ncu
// capture (XU 51 %, tensor 25 %, issue 44 %): a chain S_j -> softmax_j -> P_j (through shared memory + a proxy fence)
// -> PV_j with the partial O folded into 64 registers per row every block. Here ONE CTA per SM owns TWO 128-query
// tiles of a (clip, head) and runs them in ping-pong:
//   warp 0       TMA: Q0, Q1 once; K_j / V_j tiles (128 x 64 bf16, 128B swizzle) through a 4-stage ring — every K/V
//                tile is fetched once for 256 queries;
//   warp 1       tcgen05 issuer. Per key block j and query tile w: S_w,j+1 = Q_w K_j+1^T is issued as soon as the
//                softmax warps have read S_w,j (ahead of PV_w,j), then O_w += P_w,j V_j with the A operand (P) read
//                from TENSOR MEMORY and the accumulator staying in tensor memory across all key blocks;
//   warps 2-5    softmax of query tile 0, warps 6-9 of query tile 1 (thread = query row): tcgen05.ld the S row in
//                32-column pieces (the next piece's load in flight under the current piece's math — no O registers
//                to pay for it), P = exp2(s*c - m_ref) as packed bf16 straight back into tensor memory (tcgen05.st),
//                row sum and block maximum on the way. No shared-memory round trip, no fence.proxy.async on the chain.
// The reference exponent m_ref of a row is the maximum of the blocks BEFORE the current one (exact pass over block 0
// only), and it only moves when the seen maximum exceeds it by more than 2^8: P may exceed 1 (exact mathematics, f32
// accumulation), and the O accumulator in tensor memory is rescaled only on those rare moves — by the row's own
// softmax warp, between PV_j's completion and P_j+1's hand-over, decided per warp (tcgen05.ld/st are .sync.aligned).
// A row whose scores jump by more than 2^64 against everything seen before redoes the block with its exact maximum.
// Tensor memory (512 columns): S0 | S1 (128 each, f32 scores), P0 | P1 (64 each: 128 keys as bf16 pairs),
// O0 | O1 (64 each). Keys past the clip end (1500 = 11*128 + 92) are zero-filled by TMA and masked.
#pragma once

#include "attn2_sm100.cuh"   // softmax piece helpers' siblings, descriptors, ex2

namespace wt {

constexpr int kAt3Threads = 64 + 256;   // TMA warp, MMA warp, 2 softmax warpgroups (one per query tile)
constexpr int kAt3Stages = 4;
struct Attn3Smem {
  static constexpr int kQ = 0;                                   // 2 buffers x (Q0, Q1): 4 x 16 KB (this item's, the next item's)
  static constexpr int kKV = kQ + 4 * kAtTile;                   // kAt3Stages x (K tile + V tile)
  static constexpr int kBars = kKV + kAt3Stages * 2 * kAtTile;
  static constexpr int kTotal = kBars + 256 + 1024;              // + alignment slack
};

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] . B[smem]: A = 128 rows (lanes) x 16 bf16 (8 columns of packed pairs)
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

#ifndef AT3_PROBE
#define AT3_PROBE 0   // timing probes (results are WRONG): 1 = no MUFU (p = x), 2 = only the first S piece is loaded
#endif
#ifndef AT3_EXP
#define AT3_EXP 0     // 0 = ex2.approx.ftz.f32 per element; 1 = ex2.approx.ftz.bf16x2 per PAIR of elements
#endif
__device__ __forceinline__ uint32_t ex2_bf16x2(uint32_t x) {
  uint32_t y;
  asm("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}

// One 32-key piece of a row: P = exp2(s*c - m_ref) -> 16 packed bf16 pairs (in place in v[0..15]); row-sum and maximum.
template <bool kMasked>
__device__ __forceinline__ void at3_piece(uint32_t (&v)[32], int c, float c_scale, float m_ref, int valid, float& l_blk,
                                          float& mx) {
  float la = 0.0f, lb = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    const float s0 = __uint_as_float(v[i]), s1 = __uint_as_float(v[i + 1]);
#if AT3_EXP == 1
    // exponent arguments rounded to bf16 (|x| * 2^-9 absolute: at most 7e-4 of the row's largest weight), ONE MUFU op
    // per pair, the result is the packed P itself; the row sum adds the bf16 values PV multiplies by
    uint32_t pk = ex2_bf16x2(pack_bf16x2(fmaf(s0, c_scale, -m_ref), fmaf(s1, c_scale, -m_ref)));
    if (kMasked) {
      if (c * 32 + i >= valid) pk &= 0xffff0000u; else mx = fmaxf(mx, s0);
      if (c * 32 + i + 1 >= valid) pk &= 0x0000ffffu; else mx = fmaxf(mx, s1);
    } else {
      mx = fmaxf(fmaxf(mx, s0), s1);
    }
    la += __uint_as_float(pk << 16);
    lb += __uint_as_float(pk & 0xffff0000u);
    v[i >> 1] = pk;
#else
#if AT3_PROBE == 1
    float p0 = fmaf(s0, c_scale, -m_ref), p1 = fmaf(s1, c_scale, -m_ref);
#else
    float p0 = ex2_approx(fmaf(s0, c_scale, -m_ref));
    float p1 = ex2_approx(fmaf(s1, c_scale, -m_ref));
#endif
    if (kMasked) {
      if (c * 32 + i >= valid) p0 = 0.0f; else mx = fmaxf(mx, s0);
      if (c * 32 + i + 1 >= valid) p1 = 0.0f; else mx = fmaxf(mx, s1);
    } else {
      mx = fmaxf(fmaxf(mx, s0), s1);   // one FMNMX3
    }
    la += p0;
    lb += p1;
    v[i >> 1] = pack_bf16x2(p0, p1);   // word i/2 is dead by now
#endif
  }
  l_blk += la + lb;
}

// Pass 2 of a block: S (4 pieces of 32 columns, the next piece's load in flight under the current one's math) -> P in
// tensor memory. `pre` runs once, after the first piece has been computed and before the first P store: the wait for
// the previous block's PV (it reads the P buffer) and the rare rescale of O sit there, off the critical path.
template <bool kMasked, class Pre>
__device__ __forceinline__ void at3_pass2(uint32_t t_s, uint32_t t_p, float c_scale, float m_ref, int valid, float& l_blk,
                                          float& mx, Pre pre) {
  l_blk = 0.0f;
  mx = -INFINITY;
  uint32_t va[32], vb[32];
  tmem_ld32(t_s, va);
  tmem_ld_wait();
#if AT3_PROBE == 2
#pragma unroll
  for (int i = 0; i < 32; ++i) vb[i] = va[i] ^ 0x00010000u;
#else
  tmem_ld32(t_s + 32, vb);
#endif
  at3_piece<kMasked>(va, 0, c_scale, m_ref, valid, l_blk, mx);
  pre();
  tmem_st16(t_p, va);
  tmem_ld_wait();
#if AT3_PROBE == 2
#pragma unroll
  for (int i = 0; i < 32; ++i) va[i] = vb[i] ^ 0x00020000u;
#else
  tmem_ld32(t_s + 64, va);
#endif
  at3_piece<kMasked>(vb, 1, c_scale, m_ref, valid, l_blk, mx);
  tmem_st16(t_p + 16, vb);
  tmem_ld_wait();
#if AT3_PROBE == 2
#pragma unroll
  for (int i = 0; i < 32; ++i) vb[i] = va[i] ^ 0x00040000u;
#else
  tmem_ld32(t_s + 96, vb);
#endif
  at3_piece<kMasked>(va, 2, c_scale, m_ref, valid, l_blk, mx);
  tmem_st16(t_p + 32, va);
  tmem_ld_wait();
  at3_piece<kMasked>(vb, 3, c_scale, m_ref, valid, l_blk, mx);
  tmem_st16(t_p + 48, vb);
}

// O_w[row] *= alpha (this thread's row, 64 columns)
__device__ __forceinline__ void at3_rescale_o(uint32_t t_o, float alpha) {
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    uint32_t o[32];
    tmem_ld32(t_o + c * 32, o);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
    tmem_st32(t_o + c * 32, o);
  }
}

// Persistent: grid = min(items, SMs) CTAs, item = (clip, head, 256-query pair), drawn from a global counter one ahead
// (sched[0] next item, sched[1] CTAs done; the last CTA out re-arms both — the GEMM kernels' protocol and counters: the
// launches of a stream are serialised). A launch that shares the GPU with other engines' decoders needs no second wave of
// CTAs, the K/V ring and the next item's Q tiles keep loading across item boundaries, and tensor memory, barriers and
// the descriptor prefetch are set up once per SM instead of 15 times. block = 320.
__global__ void __launch_bounds__(kAt3Threads, 1)
attention3_sm100_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, int S, int H, int B,
                        int* __restrict__ sched) {
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Attn3Smem::kBars);
  uint64_t* bar_q_full = bars;                         // [2] Q buffer loaded + item index published (end marker: plain arrive)
  uint64_t* bar_q_empty = bars + 2;                    // [2] 9 arrivals: the item's last S product retired + 8 softmax warps
  uint64_t* bar_kv_full = bars + 4;                    // kAt3Stages
  uint64_t* bar_kv_empty = bar_kv_full + kAt3Stages;   // kAt3Stages
  uint64_t* bar_s_full = bar_kv_empty + kAt3Stages;    // [2] per query tile, phase = j & 1
  uint64_t* bar_p_full = bar_s_full + 2;               // [2] (4 arrivals: one per softmax warp of the tile)
  uint64_t* bar_o_done = bar_p_full + 2;               // [2] PV_w,j complete
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bar_o_done + 2);
  volatile int* s_item = reinterpret_cast<volatile int*>(s_tmem + 2);   // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int D = H * 64;
  const int nkb = (S + 127) / 128;
  const int n_qp = (S + 255) / 256;
  const int n_items = B * H * n_qp;   // item = (b * H + h) * n_qp + qp: the 6 items of a (clip, head) run on neighbouring CTAs

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_q_full + i, 1);
      mbar_init(bar_q_empty + i, 9);
    }
    for (int i = 0; i < kAt3Stages; ++i) {
      mbar_init(bar_kv_full + i, 1);
      mbar_init(bar_kv_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s_full + i, 1);
      mbar_init(bar_p_full + i, 4);
      mbar_init(bar_o_done + i, 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(s_tmem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *s_tmem;
  // columns: S0 0..127 | S1 128..255 | P0 256..319 | P1 320..383 | O0 384..447 | O1 448..511
  const uint32_t tmem_s = tmem_base, tmem_p = tmem_base + 256, tmem_o = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int next_item = (int)blockIdx.x;
      for (int it = 0;; ++it) {
        const int item = next_item < n_items ? next_item : -1;
        if (item >= 0) next_item = (int)gridDim.x + atomicAdd(sched, 1);
        const int qb = it & 1;
        mbar_wait(bar_q_empty + qb, (uint32_t)((it >> 1) & 1) ^ 1u);   // item it - 2 is done with this Q buffer and slot
        s_item[qb] = item;
        if (item < 0) {
          mbar_arrive(bar_q_full + qb);
          break;
        }
        const int qp = item % n_qp, bh = item / n_qp, h = bh % H, b = bh / H;
        uint8_t* sq = smem + Attn3Smem::kQ + qb * 2 * kAtTile;
        mbar_expect_tx(bar_q_full + qb, 2 * kAtTile);
        tma_load_3d(sq, &map_qkv, bar_q_full + qb, h * 64, qp * 256, b);
        tma_load_3d(sq + kAtTile, &map_qkv, bar_q_full + qb, h * 64, qp * 256 + 128, b);
        for (int j = 0; j < nkb; ++j) {
          mbar_wait(bar_kv_empty + stage, phase ^ 1);
          uint8_t* sk = smem + Attn3Smem::kKV + stage * 2 * kAtTile;
          mbar_expect_tx(bar_kv_full + stage, 2 * kAtTile);
          tma_load_3d(sk, &map_qkv, bar_kv_full + stage, D + h * 64, j * 128, b);
          tma_load_3d(sk + kAtTile, &map_qkv, bar_kv_full + stage, 2 * D + h * 64, j * 128, b);
          if (++stage == kAt3Stages) { stage = 0; phase ^= 1; }
        }
      }
      __threadfence();
      if (atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) {   // last CTA out re-arms the counters for the next launch
        sched[0] = 0;
        sched[1] = 0;
        __threadfence();
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = make_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = make_idesc_bf16(128, 64) | (1u << 16);  // B (= V) is MN-major
    uint32_t g = 0;   // key blocks processed by this CTA so far: ring stage g % kAt3Stages, barrier parities from g
    for (int it = 0;; ++it) {
      const int qb = it & 1;
      mbar_wait(bar_q_full + qb, (uint32_t)((it >> 1) & 1));
      if (s_item[qb] < 0) break;
      const uint32_t q_addr = smem_u32(smem + Attn3Smem::kQ + qb * 2 * kAtTile);
      auto issue_s = [&](int w, uint32_t gi, bool last) {   // S_w = Q_w K^T for ring block gi (the caller has waited for it)
        if (elect_one()) {
          const int stage = (int)(gi % kAt3Stages);
          const uint64_t da = make_sw128_desc(q_addr + w * kAtTile);
          const uint64_t db = make_sw128_desc(smem_u32(smem + Attn3Smem::kKV + stage * 2 * kAtTile));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_s + w * 128, da + 2 * k, db + 2 * k, idesc_s, k != 0 ? 1u : 0u);
          umma_commit(bar_s_full + w);
          if (last && w == 1) umma_commit(bar_q_empty + qb);   // the item's last products that read Q
        }
        __syncwarp();
      };
      mbar_wait(bar_kv_full + g % kAt3Stages, (g / kAt3Stages) & 1u);
      tc_fence_after();
      issue_s(0, g, nkb == 1);
      issue_s(1, g, nkb == 1);
      for (int j = 0; j < nkb; ++j, ++g) {
        const int stage = (int)(g % kAt3Stages);
        if (j + 1 < nkb) {   // K_j+1 (and V_j+1) landed: both query tiles' next S products need it
          mbar_wait(bar_kv_full + (g + 1) % kAt3Stages, ((g + 1) / kAt3Stages) & 1u);
        }
        for (int w = 0; w < 2; ++w) {
          mbar_wait(bar_p_full + w, g & 1u);   // P_w,j is in tensor memory, so S_w,j has been read
          tc_fence_after();
          if (j + 1 < nkb) issue_s(w, g + 1, j + 2 == nkb);   // ahead of PV_w,j: the softmax warps restart one MMA earlier
          if (elect_one()) {
            const uint32_t v_addr = smem_u32(smem + Attn3Smem::kKV + stage * 2 * kAtTile + kAtTile);
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
              // A = P_w: 16 keys per K step = 8 columns of packed bf16 pairs; B = V: 16 rows of 128 bytes per K step
              const uint64_t db = make_sw128_mn_desc(v_addr + kk * 2048);
              umma_bf16_ts(tmem_o + w * 64, tmem_p + w * 64 + kk * 8, db, idesc_o, (j | kk) != 0 ? 1u : 0u);
            }
            umma_commit(bar_o_done + w);
            if (w == 1) umma_commit(bar_kv_empty + stage);   // every MMA that reads K_j / V_j has been issued
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ---- softmax: thread <-> query row of tile w ----
    const int w = (warp - 2) >> 2;
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may touch
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
    const float c_scale = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) * log2(e)
    const uint32_t t_s = tmem_s + lane_sel + w * 128;
    const uint32_t t_p = tmem_p + lane_sel + w * 64;
    const uint32_t t_o = tmem_o + lane_sel + w * 64;
    constexpr float kTau = 8.0f;                  // the reference exponent moves only past 2^8 (P <= 256: exact in f32 sums)
    uint32_t g = 0;                               // key blocks processed by this CTA so far (barrier parities)
    for (int it = 0;; ++it) {
    const int qb = it & 1;
    mbar_wait(bar_q_full + qb, (uint32_t)((it >> 1) & 1));
    const int item = s_item[qb];
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_q_empty + qb);
    if (item < 0) break;
    float m_ref = -INFINITY, m_seen = -INFINITY, l_run = 0.0f;
    float alpha_pend = 1.0f;                      // O rescale owed from the previous block's reference move
    bool rescale_pend = false;                    // warp-uniform
    for (int j = 0; j < nkb; ++j, ++g) {
      mbar_wait(bar_s_full + w, g & 1u);
      tc_fence_after();
      const int valid = S - j * 128;  // keys valid in this block (>= 128 except the last)
      if (j == 0) {
        const float mx0 = valid >= 128 ? at2_pass1<false, false>(t_s, valid) : at2_pass1<true, false>(t_s, valid);
        m_ref = mx0 * c_scale;
        m_seen = m_ref;
      }
      // before the first P store: PV_w,j-1 (reads the P buffer, accumulates into O) must be complete; the O rescale owed
      // from the last reference move goes here too
      auto pre = [&]() {
        if (j >= 1) {
          mbar_wait(bar_o_done + w, (g - 1) & 1u);
          tc_fence_after();
          if (rescale_pend) at3_rescale_o(t_o, alpha_pend);
        }
      };
      float l_blk, mx_blk;
      if (valid >= 128) at3_pass2<false>(t_s, t_p, c_scale, m_ref, valid, l_blk, mx_blk, pre);
      else at3_pass2<true>(t_s, t_p, c_scale, m_ref, valid, l_blk, mx_blk, pre);
      // a jump of more than 2^64 above the reference would overflow P: redo the block with its exact maximum. Decided
      // per WARP (tcgen05.ld / st are .sync.aligned); lanes without a jump repeat it with their unchanged reference and
      // rescale by 1.
      const bool jump = mx_blk * c_scale > m_ref + 64.0f;
      if (__any_sync(0xffffffffu, jump)) {
        float a = 1.0f;
        if (jump) {
          const float m_new = mx_blk * c_scale;
          a = ex2_approx(m_ref - m_new);
          l_run *= a;
          m_ref = m_new;
        }
        if (j >= 1) at3_rescale_o(t_o, a);   // O is complete up to block j-1 (`pre`); PV_w,j accumulates after the hand-over
        at3_pass2<true>(t_s, t_p, c_scale, m_ref, valid, l_blk, mx_blk, [] {});
      }
      l_run += l_blk;
      rescale_pend = false;
      alpha_pend = 1.0f;
      m_seen = fmaxf(m_seen, mx_blk * c_scale);
      // lazy reference move (applies from block j+1 on): l now, O after PV_w,j completes (next block's `pre`)
      if (j + 1 < nkb) {
        const bool move = m_seen > m_ref + kTau;
        if (__any_sync(0xffffffffu, move)) {
          rescale_pend = true;
          if (move) {
            alpha_pend = ex2_approx(m_ref - m_seen);
            l_run *= alpha_pend;
            m_ref = m_seen;
          }
        }
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p_full + w);
    }
    // ---- O / l -> bf16 ----
    {
      mbar_wait(bar_o_done + w, (g - 1) & 1u);   // PV of the item's last key block
      tc_fence_after();
      const int qp = item % n_qp, bh = item / n_qp, h = bh % H, b = bh / H;   // decoded here: nothing of it is live in the loop
      const int q = qp * 256 + w * 128 + row;
      const float inv = 1.0f / l_run;
      uint4* dst = reinterpret_cast<uint4*>(out + ((long long)b * S + q) * D + h * 64);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        tmem_ld32(t_o + c * 32, o);
        tmem_ld_wait();
        if (q < S) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[c * 4 + i] = make_uint4(pack_bf16x2(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv),
                                        pack_bf16x2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv),
                                        pack_bf16x2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv),
                                        pack_bf16x2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv));
        }
      }
    }
    }   // items
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace wt
