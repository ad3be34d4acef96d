// Small device helpers shared by the streaming decoder (decoder_stream_sm100.cuh): mma.sync wrapper, bf16
// pack/unpack, L2 cache-policy loads, the 1-D bulk (TMA) copy, cluster special registers.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "gemm_sm100.cuh"  // mbarrier / TMA wrappers
#include "simt_kernels.cuh"

namespace wt {

constexpr int kDecGroupMax = 16;   // clips owned by one cluster: rows of the one m16 MMA tile
constexpr int kDecMaxCluster = 16; // CTAs per cluster (non-portable size; 8 is the portable fallback)
constexpr int kDecMaxD = 768;      // d_model ceiling of the streaming decoder (tiny 384, base 512, small 768)
constexpr int kDecTokStride = 32;
constexpr int kXaChunk = 128;      // keys per K/V stream chunk (= one ring stage)

// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mma_bf16_16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// streaming read of data used once per step (cross K/V): no L1 allocation, first in line for L2 eviction so the
// decoder weights and the embedding matrix stay L2-resident between steps
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p, uint64_t policy) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p), "l"(policy));
  return v;
}
// 1-D bulk copy global -> this CTA's shared memory, completion counted on an mbarrier, L2 evict-first hint
__device__ __forceinline__ void bulk_load_stream(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float dec_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t dec_pack2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void bf16x8_to_f32(const uint4& v, float* f) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = __uint_as_float(w[j] << 16);
    f[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}

// All CTAs of one cluster (hardware barrier, release/acquire at cluster scope). A cluster owns its clips for the whole
// loop, so nothing in this kernel ever waits on another cluster: no cooperative launch, no software grid barrier.
// Inter-CTA buffers live in global memory (L2) and are read with loads that bypass L1 (__ldcg).
__device__ __forceinline__ unsigned long long dec_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned dec_cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned dec_cluster_nctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned dec_clusterid_x() {
  unsigned r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}

}  // namespace wt
