// Helpers shared by the encoder attention kernel (attn2_sm100.cuh): tile size, ex2, the MN-major operand descriptor
// for V (consumed straight from its [key][dim] tile).
#pragma once

#include "gemm_sm100.cuh"

namespace wt {

constexpr int kAtTile = 128 * 64 * 2;  // one 128 x 64 bf16 tile: 16 KB

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

}  // namespace wt
