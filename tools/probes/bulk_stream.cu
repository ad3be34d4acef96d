// Micro-benchmark: how fast can N CTAs stream a large buffer from HBM into shared memory with cp.async.bulk
// (1-D TMA) rings? One thread per CTA issues and retires the copies; nothing consumes the data.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_stream bulk_stream.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\tW_LOOP:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t@P1 bra W_DONE;\n\tbra W_LOOP;\n\tW_DONE:\n\t}" ::"r"(
          smem_u32(bar)),
      "r"(parity), "r"(0x989680)
      : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy, int hint) {
  if (hint)
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                     smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
  else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// each CTA streams `n_streams` separate sequential regions of per_stream bytes (like K and V of a few heads)
__global__ void bulk_kernel(const uint8_t* src, size_t per_cta, int chunk, int stages, int hint, int split) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint8_t* ring = smem + 128;
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) mbar_init(bars + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    const uint8_t* base = src + (size_t)blockIdx.x * per_cta;
    const int n = (int)(per_cta / chunk);
    const int piece = chunk / split;
    for (int j = 0; j < n + stages; ++j) {
      if (j >= stages) mbar_wait(bars + (j % stages), ((j / stages) - 1) & 1);
      if (j < n) {
        mbar_expect_tx(bars + (j % stages), chunk);
        for (int s = 0; s < split; ++s)
          bulk_load(ring + (size_t)(j % stages) * chunk + s * piece, base + (size_t)j * chunk + s * piece, piece, bars + (j % stages), policy, hint);
      }
    }
  }
}

// reference: plain 16-byte loads, all threads
__global__ void ldg_kernel(const uint4* src, size_t per_cta_vec, uint4* sink) {
  const uint4* base = src + (size_t)blockIdx.x * per_cta_vec;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (size_t i = threadIdx.x; i < per_cta_vec; i += blockDim.x * 4) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = (i + u * blockDim.x < per_cta_vec) ? __ldg(base + i + u * blockDim.x) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int u = 0; u < 4; ++u) { acc.x ^= v[u].x; acc.y ^= v[u].y; acc.z ^= v[u].z; acc.w ^= v[u].w; }
  }
  if (acc.x == 0x12345678u) sink[0] = acc;
}

int main() {
  const size_t total = (size_t)1 << 30;  // 1 GiB source
  uint8_t* src;
  cudaMalloc(&src, total);
  cudaMemset(src, 1, total);
  uint4* sink;
  cudaMalloc(&sink, 64);
  cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  const int ctas_list[4] = {16, 48, 112, 148};
  for (int ci = 0; ci < 4; ++ci) {
    const int ctas = ctas_list[ci];
    const size_t per_cta = ((total / ctas) / (64 * 1024)) * (64 * 1024);
    for (int chunk = 8192; chunk <= 65536; chunk *= 2) {
      for (int stages = 2; stages <= 6; stages += 2) {
        if ((size_t)chunk * stages + 128 > 96 * 1024 && ctas > 148) continue;
        if ((size_t)chunk * stages + 128 > 200 * 1024) continue;
        for (int hint = 0; hint < 2; ++hint) {
          for (int split = 1; split <= 4; split *= 4) {
            float best = 1e9f;
            for (int rep = 0; rep < 3; ++rep) {
              cudaEventRecord(a);
              bulk_kernel<<<ctas, 32, chunk * stages + 128>>>(src, per_cta, chunk, stages, hint, split);
              cudaEventRecord(b);
              cudaEventSynchronize(b);
              float ms;
              cudaEventElapsedTime(&ms, a, b);
              if (ms < best) best = ms;
            }
            cudaError_t e = cudaGetLastError();
            printf("bulk ctas %3d chunk %5d stages %d hint %d split %d : %7.1f GB/s %s\n", ctas, chunk, stages, hint, split,
                   per_cta * ctas / best / 1e6, e == cudaSuccess ? "" : cudaGetErrorString(e));
          }
        }
      }
    }
    for (int thr = 256; thr <= 1024; thr *= 2) {
      float best = 1e9f;
      for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(a);
        ldg_kernel<<<ctas, thr>>>((const uint4*)src, per_cta / 16, sink);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      printf("ldg  ctas %3d threads %4d : %7.1f GB/s\n", ctas, thr, per_cta * ctas / best / 1e6);
    }
  }
  return 0;
}
