// Row-maximum helpers of the encoder attention's softmax warps (attn3_sm100.cuh): one pass over a 128-key block of S in
// tensor memory, 32 columns at a time; the mask of a clip's last key block is a template flag behind a real branch (as
// a run-time condition inside the unrolled body ptxas if-converts it into predicated-off instructions per element,
// which cost issue slots in every full block).
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
}  // namespace wt
