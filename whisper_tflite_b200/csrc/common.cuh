// Small helpers shared by every kernel header (nvcc and the CPU emulator both include this).
#pragma once
#ifndef WT_CUDA_EMU
#include <cuda_runtime.h>
#endif

namespace wt {
// Monotone float <-> int map so a float max can be an integer atomicMax.
__device__ __forceinline__ int float_to_ordered(float f) {
  int b = __float_as_int(f);
  return b >= 0 ? b : (b ^ 0x7fffffff);
}
__device__ __forceinline__ float ordered_to_float(int k) { return __int_as_float(k >= 0 ? k : (k ^ 0x7fffffff)); }
}  // namespace wt
