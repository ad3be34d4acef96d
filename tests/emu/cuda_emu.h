// TEST INFRASTRUCTURE — a minimal CPU emulation of the CUDA SIMT model, so that the repo's
// non-tensor-core kernels (csrc/*.cuh) can be compiled with g++ and logic-checked in the
// no-GPU container (`pytest -m "not gpu"`). One block runs at a time; its threads are real OS
// threads from a pool, `__syncthreads` is a barrier, warp shuffles go through a per-warp
// mailbox. It is NOT a product path: nothing in the package links or loads it.
#pragma once
#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#define WT_CUDA_EMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static

struct dim3 {
  unsigned x, y, z;
  dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct uint3_emu { unsigned x, y, z; };
extern thread_local uint3_emu threadIdx, blockIdx;
extern dim3 blockDim, gridDim;

struct float2 { float x, y; };
struct float4 { float x, y, z, w; };
struct int2 { int x, y; };
struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };
static inline float2 make_float2(float a, float b) { return {a, b}; }
static inline float4 make_float4(float a, float b, float c, float d) { return {a, b, c, d}; }
static inline int2 make_int2(int a, int b) { return {a, b}; }
static inline uint2 make_uint2(unsigned a, unsigned b) { return {a, b}; }
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { return {a, b, c, d}; }

// ---- bf16 -------------------------------------------------------------------------------
struct __nv_bfloat16 {
  uint16_t v;
  __nv_bfloat16() = default;
  __nv_bfloat16(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) { v = 0x7fc0; return; }
    u += 0x7fffu + ((u >> 16) & 1u);
    v = (uint16_t)(u >> 16);
  }
  operator float() const {
    uint32_t u = (uint32_t)v << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
  }
};
struct __nv_bfloat162 { __nv_bfloat16 x, y; };
static inline __nv_bfloat16 __float2bfloat16(float f) { return __nv_bfloat16(f); }
static inline __nv_bfloat16 __float2bfloat16_rn(float f) { return __nv_bfloat16(f); }
static inline float __bfloat162float(__nv_bfloat16 b) { return (float)b; }
static inline __nv_bfloat162 __floats2bfloat162_rn(float a, float b) { return {__nv_bfloat16(a), __nv_bfloat16(b)}; }
static inline float2 __bfloat1622float2(__nv_bfloat162 p) { return {(float)p.x, (float)p.y}; }

// ---- block / warp machinery -------------------------------------------------------------
namespace emu {
struct Barrier {
  std::mutex m;
  std::condition_variable cv;
  int count = 0, waiting = 0;
  unsigned gen = 0;
  void reset(int n) { count = n; waiting = 0; }
  void wait() {
    std::unique_lock<std::mutex> lk(m);
    unsigned g = gen;
    if (++waiting == count) {
      waiting = 0;
      ++gen;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return gen != g; });
    }
  }
};
struct WarpBox {
  Barrier bar;
  uint64_t slot[32];
};
extern Barrier g_block_bar;
extern std::vector<WarpBox>* g_warps;
extern thread_local int t_linear;  // linear thread id in block
extern unsigned char* g_dyn_smem;
inline WarpBox& my_warp() { return (*g_warps)[t_linear / 32]; }

void launch_impl(dim3 grid, dim3 block, size_t dyn_smem, const std::function<void()>& body);
template <class K, class... A>
void launch(K kernel, dim3 grid, dim3 block, size_t dyn_smem, A... args) {
  launch_impl(grid, block, dyn_smem, [=]() { kernel(args...); });
}
}  // namespace emu

static inline void __syncthreads() { emu::g_block_bar.wait(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::my_warp().bar.wait(); }

template <class T>
static inline T emu_shfl_from(T v, int src_lane) {
  static_assert(sizeof(T) <= 8, "shuffle payload");
  auto& w = emu::my_warp();
  int lane = emu::t_linear % 32;
  uint64_t raw = 0;
  std::memcpy(&raw, &v, sizeof(T));
  w.slot[lane] = raw;
  w.bar.wait();
  uint64_t got = w.slot[src_lane & 31];
  w.bar.wait();
  T out;
  std::memcpy(&out, &got, sizeof(T));
  return out;
}
template <class T>
static inline T __shfl_xor_sync(unsigned, T v, int mask, int width = 32) {
  int lane = emu::t_linear % 32;
  (void)width;
  return emu_shfl_from(v, lane ^ mask);
}
template <class T>
static inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
  int lane = emu::t_linear % 32;
  int base = lane / width * width;
  return emu_shfl_from(v, base + (src % width));
}
template <class T>
static inline T __shfl_down_sync(unsigned, T v, unsigned delta, int width = 32) {
  int lane = emu::t_linear % 32;
  int src = lane + (int)delta;
  if (src / width != lane / width) src = lane;
  return emu_shfl_from(v, src);
}

// ---- atomics (blocks run one at a time, threads of a block concurrently) -----------------
static inline int atomicMax(int* a, int v) {
  auto* p = reinterpret_cast<std::atomic<int>*>(a);
  int old = p->load();
  while (old < v && !p->compare_exchange_weak(old, v)) {}
  return old;
}
static inline int atomicAdd(int* a, int v) { return reinterpret_cast<std::atomic<int>*>(a)->fetch_add(v); }
static inline unsigned atomicAdd(unsigned* a, unsigned v) {
  return reinterpret_cast<std::atomic<unsigned>*>(a)->fetch_add(v);
}
static inline float atomicAdd(float* a, float v) {
  auto* p = reinterpret_cast<std::atomic<uint32_t>*>(a);
  uint32_t old = p->load(), nw;
  float f;
  do {
    std::memcpy(&f, &old, 4);
    f += v;
    std::memcpy(&nw, &f, 4);
  } while (!p->compare_exchange_weak(old, nw));
  std::memcpy(&f, &old, 4);
  return f;
}
static inline unsigned long long atomicAdd(unsigned long long* a, unsigned long long v) {
  return reinterpret_cast<std::atomic<unsigned long long>*>(a)->fetch_add(v);
}
static inline unsigned long long atomicMax(unsigned long long* a, unsigned long long v) {
  auto* p = reinterpret_cast<std::atomic<unsigned long long>*>(a);
  unsigned long long old = p->load();
  while (old < v && !p->compare_exchange_weak(old, v)) {}
  return old;
}

// ---- math / misc intrinsics ----------------------------------------------------------------
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline float emu_log2f(float x) { return std::log2(x); }
static inline float emu_expf(float x) { return std::exp(x); }
#define __log2f(x) emu_log2f(x)
#define __expf(x) emu_expf(x)
static inline float exp2f_emu(float x) { return std::exp2(x); }
static inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline float __frcp_rn(float a) { return 1.0f / a; }
static inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
static inline unsigned __float_as_uint(float f) { unsigned i; std::memcpy(&i, &f, 4); return i; }
static inline float __uint_as_float(unsigned i) { float f; std::memcpy(&f, &i, 4); return f; }
using std::fmaf;
using std::fmaxf;
using std::fminf;
