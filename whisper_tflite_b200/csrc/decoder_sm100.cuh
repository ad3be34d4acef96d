// Persistent greedy decoder for sm_100a: the whole loop of Decoder::forward (whisper.cpp:363-402 — up to 27
// Interpreter::Invoke calls in the reference, each a full re-forward of all T tokens) as ONE cooperative launch.
//
// Why one kernel: with a KV cache a decode step is a chain of ~55 tiny dependent ops on [B <= 64][d] activations
// (LN, QKV, self-attention over <= 31 cached keys, projections, MLP) wrapped around the one heavy part, cross-attention,
// which streams every clip's encoder K/V (B x 2 x 1500 x d bf16 per layer: 147 MB at tiny/B=64) from HBM. As separate
// launches the step is bound by launch latency (1680 launches per transcription, ~15 ms of a 26 ms pass in round-1
// profiles/r01a_launches.md); here the 148 CTAs stay resident, phases are separated by a grid barrier (one atomic
// + one acquire spin, ~0.5 us), weights are re-read from L2 and only the K/V stream touches HBM.
//
// Per step and layer, 8 phases:
//   1 LN1 -> QKV GEMM (k, v written straight into the self-attention cache)      5 cross-attention (HBM stream)
//   2 self-attention per (clip, head), one warp each                              6 out-proj + residual
//   3 out-proj + residual                                                         7 LN2 -> fc1 -> GELU
//   4 LNx -> cross query GEMM                                                     8 fc2 + residual
// then final LN -> tied logits GEMM with a fused (max, last index) reduction (whisper.cpp:346-361's tie rule), and a
// finalize phase that appends the token, tracks EOT per clip and embeds the next position.
//
// GEMMs here are [B <= 64] x [N] x [K]: M is tiny, so each CTA stages ALL rows of A (bf16, LayerNorm fused into the
// staging) in shared memory and owns a slice of N; a warp computes 64 x 8 outputs with mma.sync.m16n8k16 (bf16 in,
// f32 accumulate), streaming its 8 weight rows from L2 with 16-byte loads — tcgen05's 128-row tiles would idle on
// M = 64 and these phases are latency-, not FLOP-bound. K is permuted consistently between the A fragments and the
// weight rows (a lane owns 8 consecutive k of a 32-wide block), so every operand load is one 128-bit access.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "gemm_sm100.cuh"  // mbarrier / TMA wrappers
#include "simt_kernels.cuh"

namespace wt {

constexpr int kDecThreads = 256;   // 8 warps: 255 registers per thread, so nothing in the phases spills
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kDecGroupMax = 16;   // clips owned by one cluster: rows of one staged A tile (kDecMT m-tiles of 16)
constexpr int kDecMT = kDecGroupMax / 16;
constexpr int kDecMaxCluster = 16; // CTAs per cluster (non-portable size; 8 is the portable fallback)
constexpr int kDecMaxD = 768;      // d_model ceiling of this kernel (tiny 384, base 512, small 768)
constexpr int kDecTokStride = 32;
constexpr int kDecMaxLayers = 12;  // layer pointer tables kept in shared memory
constexpr int kXaChunk = 128;      // keys per ring stage
constexpr int kXaStages = 4;       // 4 x 32 KB in flight per SM
constexpr int kXaMaxSeg = 31;      // (clip, head) segments one CTA's key range may touch

struct __align__(256) DecLayerDev {
  const float *ln1_g, *ln1_b, *lnx_g, *lnx_b, *ln2_g, *ln2_b;
  const __nv_bfloat16* w_qkv; const float* b_qkv;    // [3d][d]
  const __nv_bfloat16* w_out; const float* b_out;    // [d][d]
  const __nv_bfloat16* w_xq; const float* b_xq;      // [d][d]
  const __nv_bfloat16* w_xout; const float* b_xout;  // [d][d]
  const __nv_bfloat16* w_fc1; const float* b_fc1;    // [4d][d]
  const __nv_bfloat16* w_fc2; const float* b_fc2;    // [d][4d]
  const __nv_bfloat16* kvx;                          // cross-attention K/V stream: [B][H][chunk][K rows | V rows]
  __nv_bfloat16 *kc, *vc;                            // [B][H][32][64] self-attention cache
};

struct DecParams {
  const DecLayerDev* layers;
  int n_layers, n_clips, clip0, d, H, S, V;
  int group_clips;                // clips owned by one cluster (<= kDecGroupMax)
  int rank, G;                    // filled in by the kernel: CTA rank in its cluster, cluster size
  int debug;                      // WT_DEC_DEBUG bits (timing experiments only): 1 = cross-attn skips its math, 2 = skips its loads
  const __nv_bfloat16* tok_emb;   // [V][d] (also the tied logits matrix)
  const __nv_bfloat16* pos_emb;   // [n_text_ctx][d]
  const float *lnf_g, *lnf_b;
  int* tokens;                    // [B][32]
  int* lens;
  int* done;
  int prompt_len, max_total, eot;
  float* x;                       // [64][d]   residual stream
  float* qs;                      // [64][d]   self-attention query
  float* att;                     // [64][d]   attention output (self, then cross)
  float* qx;                      // [64][d]   cross-attention query
  __nv_bfloat16* hmid;            // [64][4d]  MLP hidden
  float* xpart;                   // [pairs][xslots][66] cross-attention partials (acc[64], max, sum)
  unsigned* xtickets;             // [pairs]
  int xslots;
  unsigned long long* amax_part;  // [group][kDecGroupMax][cluster size]
  float* step_logits;             // optional [steps][max_batch][V]
  long long step_logits_stride;   // max_batch * V
  int* finished;                  // [n_groups] number of clips of the group that produced EOT (zeroed before launch)
  unsigned long long* trace;      // optional [32]: ns per phase kind seen by cluster 0 / CTA 0 ([k] total, [16+k] barrier wait)
};

struct DecSmem {
  static constexpr int kAStride = 2 * kDecMaxD + 64;            // bytes per staged row (== 64 mod 128: conflict-free LDS.128)
  static constexpr int kA = 0;
  static constexpr int kRedPerWarp = kDecMT * 16 * 8;           // floats
  static constexpr int kRed = kA + kDecGroupMax * kAStride;     // [16 warps][kDecMT*16][8] f32
  static constexpr int kMisc = kRed + kDecWarps * kRedPerWarp * 4;   // sq / cross-merge / argmax scratch
  static constexpr int kMiscBytes = kDecWarps * 66 * 4 + 64 * 8 + 64;
  // cross-attention K/V ring: kXaStages stages of kXaChunk keys, K rows then V rows (128 bytes per key each),
  // filled by cp.async.bulk (TMA) straight from the per-head streams and guarded by full/empty mbarriers
  static constexpr int kSeg = kMisc + ((kMiscBytes + 127) & ~127);       // segment table: int4[kXaMaxSeg] + header
  static constexpr int kBars = kSeg + 16 * (kXaMaxSeg + 2);           // 32-byte XaHdr, then int4 segments
  static constexpr int kRing = (kBars + 2 * kXaStages * 8 + 127) & ~127;
  static constexpr int kStageBytes = 2 * kXaChunk * 128;
  static constexpr int kXq = kRing + kXaStages * kStageBytes;            // staged, pre-scaled queries [kXaMaxSeg][64] f32
  static constexpr int kLnGB = kXq + kXaMaxSeg * 64 * 4;                 // LayerNorm gamma | beta of the running phase
  // this cluster's DecParams and every layer's pointer table live in shared memory: the cluster barrier's acquire
  // invalidates L1, so parameters left in local/global memory would cost an L2 round trip at the head of every phase
  static constexpr int kParams = kLnGB + 2 * kDecMaxD * 4;
  static constexpr int kLayers = kParams + 512;
  static constexpr int kTotal = kLayers + kDecMaxLayers * 256;
};

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
__device__ __forceinline__ void dec_cluster_sync(const DecParams& p, int kind, unsigned long long& tlast) {
  const bool tr = p.trace != nullptr && threadIdx.x == 0;
  const unsigned long long t_arr = tr ? dec_globaltimer() : 0ull;
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (tr) {
    const unsigned long long t = dec_globaltimer();
    atomicAdd(p.trace + kind, t - tlast);        // RED: fire-and-forget, the timed warp does not wait on L2
    atomicAdd(p.trace + 16 + kind, t - t_arr);
    tlast = t;
  }
}

// ---- A staging: all n_rows (padded to 16) x d into shared memory as bf16, row stride DecSmem::kAStride ----
// LayerNorm gamma/beta -> shared memory with cp.async (no registers held while the weights stream in). Call, then
// dec_prefetch_ln_wait() + __syncthreads() before dec_stage_ln.
__device__ __forceinline__ void dec_prefetch_ln(uint8_t* smem, const float* gamma, const float* beta, int d) {
  float* s_g = reinterpret_cast<float*>(smem + DecSmem::kLnGB);
  const int n4 = d >> 2;  // 16-byte pieces per vector
  for (int i = threadIdx.x; i < 2 * n4; i += kDecThreads) {
    const float* src = (i < n4) ? gamma + 4 * i : beta + 4 * (i - n4);
    float* dst = (i < n4) ? s_g + 4 * i : s_g + kDecMaxD + 4 * (i - n4);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void dec_prefetch_ln_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void dec_stage_ln(uint8_t* sA, const float* x, const float* gamma, const float* beta,
                                             int n_clips, int rows_pad, int d) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = d >> 6;  // float2 per lane
  for (int r = warp; r < rows_pad; r += kDecWarps) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(sA + (size_t)r * DecSmem::kAStride);
    if (r < n_clips) {
      float2 v[kDecMaxD / 64];
      float s = 0.0f;
#pragma unroll
      for (int i = 0; i < kDecMaxD / 64; ++i)
        if (i < per) {
          v[i] = __ldcg(reinterpret_cast<const float2*>(x + (size_t)r * d + i * 64 + 2 * lane));
          s += v[i].x + v[i].y;
        }
      const float mean = warp_sum(s) / (float)d;
      float q = 0.0f;
#pragma unroll
      for (int i = 0; i < kDecMaxD / 64; ++i)
        if (i < per) {
          const float a = v[i].x - mean, b = v[i].y - mean;
          q += a * a + b * b;
        }
      const float rstd = rsqrtf(warp_sum(q) / (float)d + 1e-5f);
#pragma unroll
      for (int i = 0; i < kDecMaxD / 64; ++i)
        if (i < per) {
          const int c = i * 64 + 2 * lane;
          const float2 g = *reinterpret_cast<const float2*>(gamma + c);   // shared memory (dec_prefetch_ln)
          const float2 bb = *reinterpret_cast<const float2*>(beta + c);
          dst[c >> 1] = dec_pack2((v[i].x - mean) * rstd * g.x + bb.x, (v[i].y - mean) * rstd * g.y + bb.y);
        }
    } else {
      for (int c = lane; c < (d >> 1); c += 32) dst[c] = 0u;
    }
  }
}
__device__ __forceinline__ void dec_stage_f32(uint8_t* sA, const float* src, int ld, int n_clips, int rows_pad, int d) {
  const int per_row = d >> 2;
  for (int i = threadIdx.x; i < rows_pad * per_row; i += kDecThreads) {
    const int r = i / per_row, c = (i - r * per_row) * 4;
    uint2 o = make_uint2(0u, 0u);
    if (r < n_clips) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(src + (size_t)r * ld + c));
      o = make_uint2(dec_pack2(v.x, v.y), dec_pack2(v.z, v.w));
    }
    *reinterpret_cast<uint2*>(sA + (size_t)r * DecSmem::kAStride + c * 2) = o;
  }
}
__device__ __forceinline__ void dec_stage_bf16(uint8_t* sA, const __nv_bfloat16* src, int ld, int n_clips, int rows_pad,
                                               int d) {
  const int per_row = d >> 3;
  for (int i = threadIdx.x; i < rows_pad * per_row; i += kDecThreads) {
    const int r = i / per_row, c = (i - r * per_row) * 8;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (r < n_clips) o = __ldcg(reinterpret_cast<const uint4*>(src + (size_t)r * ld + c));
    *reinterpret_cast<uint4*>(sA + (size_t)r * DecSmem::kAStride + c * 2) = o;
  }
}

// Cross-attention output = softmax-merge of the per-CTA partials (acc[64], max, sum) every CTA covering a (clip, head)
// pair left in its slot (dec_cross_attn), staged as the bf16 A operand of the cross out-projection.
__device__ __forceinline__ void dec_stage_xmerge(uint8_t* sA, const DecParams& p, int per, int cpp, int n_clips,
                                                 int rows_pad, int d) {
  const int half = d >> 1;
  for (int i = threadIdx.x; i < rows_pad * half; i += kDecThreads) {
    const int r = i / half, c = (i - r * half) * 2;
    uint32_t o = 0u;
    if (r < n_clips) {
      const int pr = r * p.H + (c >> 6), e = c & 63;
      const int first = (pr * cpp) / per, last = ((pr + 1) * cpp - 1) / per;
      const float* p0 = p.xpart + (size_t)pr * p.xslots * 66;
      float M = -1.0e30f;
      for (int k = 0; k <= last - first; ++k) M = fmaxf(M, __ldcg(p0 + k * 66 + 64));
      float Lw = 0.0f, o0 = 0.0f, o1 = 0.0f;
      for (int k = 0; k <= last - first; ++k) {
        const float s1 = dec_ex2(__ldcg(p0 + k * 66 + 64) - M);
        const float2 a2 = __ldcg(reinterpret_cast<const float2*>(p0 + k * 66 + e));
        Lw = fmaf(__ldcg(p0 + k * 66 + 65), s1, Lw);
        o0 = fmaf(a2.x, s1, o0);
        o1 = fmaf(a2.y, s1, o1);
      }
      const float inv = 1.0f / Lw;
      o = dec_pack2(o0 * inv, o1 * inv);
    }
    *reinterpret_cast<uint32_t*>(sA + (size_t)r * DecSmem::kAStride + c * 2) = o;
  }
}

// Weight rows are streamed from L2 in batches of up to kDecWB 32-wide k-blocks per lane (16 bytes each): all loads of
// a batch are in flight before the first mma, so a warp pays one L2 round trip per batch, not per 4 k-blocks.
constexpr int kDecWB = 12;   // projections: one batch covers K = 384 per warp
constexpr int kDecWBL = 6;   // logits: two batches in flight (double buffer)
template <int WB>
__device__ __forceinline__ void dec_load_w(uint4 (&w)[WB], const __nv_bfloat16* wrow, bool row_ok, int kb0, int kb1) {
  const int t = threadIdx.x & 3;
#pragma unroll
  for (int u = 0; u < WB; ++u) {
    w[u] = make_uint4(0u, 0u, 0u, 0u);
    if (row_ok && kb0 + u < kb1) w[u] = ldg_nc_v4(wrow + (kb0 + u) * 32 + t * 8);
  }
}
// acc[mt][0..3] += A[mt*16 + {g, g+8}][kb0 .. kb1) . w   (w = the batch loaded by dec_load_w for the same range)
template <int WB>
__device__ __forceinline__ void dec_mma_w(const uint4 (&w)[WB], const uint8_t* sA, int kb0, int kb1, int n_mt,
                                          float (&acc)[kDecMT][4]) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const uint8_t* a_base = sA + (size_t)g * DecSmem::kAStride + t * 16;
#pragma unroll
  for (int u = 0; u < WB; ++u) {
    if (kb0 + u < kb1) {
#pragma unroll
      for (int mt = 0; mt < kDecMT; ++mt) {
        if (mt < n_mt) {
          const uint8_t* ap = a_base + (size_t)(mt * 16) * DecSmem::kAStride + (kb0 + u) * 64;
          const uint4 lo = *reinterpret_cast<const uint4*>(ap);
          const uint4 hi = *reinterpret_cast<const uint4*>(ap + 8 * DecSmem::kAStride);
          mma_bf16_16816(acc[mt], lo.x, hi.x, lo.y, hi.y, w[u].x, w[u].y);
          mma_bf16_16816(acc[mt], lo.z, hi.z, lo.w, hi.w, w[u].z, w[u].w);
        }
      }
    }
  }
}

__device__ __forceinline__ int dec_pick_ksplit(int n_groups, int kblocks, int G) {
  const int gpc = (n_groups + G - 1) / G;
  const int maxk = kDecWarps / (gpc < kDecWarps ? gpc : kDecWarps);
  int best = 1;
  for (int k = 2; k <= maxk && k <= 8; ++k)
    if (kblocks % k == 0) best = k;
  return best;
}

// ---- cross-attention K/V stream: per-CTA segment table + ring bookkeeping (shared memory, thread 0 issues) ----
struct XaHdr {
  int nseg, nchunks;
  unsigned issued;   // chunks requested so far over the whole kernel (thread 0)
  int per;           // chunks per CTA of the group's (clip, head, chunk) space
  int cpp;           // chunks per (clip, head) pair
  int pad[3];
};
__device__ __forceinline__ XaHdr* xa_hdr(uint8_t* smem) { return reinterpret_cast<XaHdr*>(smem + DecSmem::kSeg); }
__device__ __forceinline__ int4* xa_segs(uint8_t* smem) { return reinterpret_cast<int4*>(smem + DecSmem::kSeg + 32); }
__device__ __forceinline__ uint64_t* xa_full(uint8_t* smem) { return reinterpret_cast<uint64_t*>(smem + DecSmem::kBars); }
__device__ __forceinline__ uint64_t* xa_empty(uint8_t* smem) { return xa_full(smem) + kXaStages; }

// out[r][n] = epi(sum_k A[r][k] * W[n][k]) for r < n_clips, n < N; K = n_chunks * d. ONE non-inlined body serves every
// projection of the layer (the staging and epilogue variants are switches, not template instantiations): the kernel's
// code stays small enough to live in the instruction caches across the ~1000 phases of a transcription.
enum : int { DG_STAGE_LN = 0, DG_STAGE_F32 = 1, DG_STAGE_BF16 = 2, DG_STAGE_XMERGE = 3 };
enum : int { DG_EPI_QKV = 0, DG_EPI_RESID = 1, DG_EPI_STORE = 2, DG_EPI_GELU = 3 };
struct DecGemmArgs {
  const __nv_bfloat16* W;     // [N][K]
  const float* bias;          // [N]
  int N, n_chunks;
  int stage_kind;
  const void* src;            // LN / F32: f32 [rows][d]; BF16: bf16 [rows][src_ld], chunk c at column c*d
  int src_ld;
  const float *ln_g, *ln_b;
  int xa_per, xa_cpp;         // XMERGE: chunks per CTA / per pair of the cross-attention split
  int epi_kind;
  float* out;                 // QKV: qs; RESID: x (read-modify-write); STORE: destination [rows][N]
  __nv_bfloat16* out_bf16;    // GELU: hmid [rows][N]
  __nv_bfloat16 *kc, *vc;     // QKV: self-attention cache
  int pos;
};

#define DEC_MARK(k)                                            \
  do {                                                         \
    if (p.trace != nullptr && threadIdx.x == 0) {              \
      const unsigned long long t_ = dec_globaltimer();         \
      atomicAdd(p.trace + (k), t_ - tmark);                    \
      tmark = t_;                                              \
    }                                                          \
  } while (0)
enum : int { DG_QKV = 0, DG_OUT = 1, DG_XQ = 2, DG_XOUT = 3, DG_FC1 = 4, DG_FC2 = 5 };
// `which` selects the projection of layer l; everything it needs comes from the shared-memory parameter block.
__device__ __noinline__ void dec_gemm(uint8_t* smem, const DecParams& p, int which, int l, int pos) {
  unsigned long long tmark = (p.trace != nullptr && threadIdx.x == 0) ? dec_globaltimer() : 0ull;
  const DecLayerDev& L = p.layers[l];
  const int d = p.d;
  DecGemmArgs a;
  a.n_chunks = 1; a.N = d; a.src = p.x; a.src_ld = d; a.ln_g = nullptr; a.ln_b = nullptr; a.out = p.x;
  a.out_bf16 = nullptr; a.kc = L.kc; a.vc = L.vc; a.pos = pos;
  a.xa_per = xa_hdr(smem)->per; a.xa_cpp = xa_hdr(smem)->cpp;
  switch (which) {
    case DG_QKV:
      a.W = L.w_qkv; a.bias = L.b_qkv; a.N = 3 * d; a.stage_kind = DG_STAGE_LN; a.ln_g = L.ln1_g; a.ln_b = L.ln1_b;
      a.epi_kind = DG_EPI_QKV; a.out = p.qs;
      break;
    case DG_OUT:
      a.W = L.w_out; a.bias = L.b_out; a.stage_kind = DG_STAGE_F32; a.src = p.att; a.epi_kind = DG_EPI_RESID;
      break;
    case DG_XQ:
      a.W = L.w_xq; a.bias = L.b_xq; a.stage_kind = DG_STAGE_LN; a.ln_g = L.lnx_g; a.ln_b = L.lnx_b;
      a.epi_kind = DG_EPI_STORE; a.out = p.qx;
      break;
    case DG_XOUT:
      a.W = L.w_xout; a.bias = L.b_xout; a.stage_kind = DG_STAGE_XMERGE; a.epi_kind = DG_EPI_RESID;
      break;
    case DG_FC1:
      a.W = L.w_fc1; a.bias = L.b_fc1; a.N = 4 * d; a.stage_kind = DG_STAGE_LN; a.ln_g = L.ln2_g; a.ln_b = L.ln2_b;
      a.epi_kind = DG_EPI_GELU; a.out_bf16 = p.hmid;
      break;
    default:
      a.W = L.w_fc2; a.bias = L.b_fc2; a.n_chunks = 4; a.stage_kind = DG_STAGE_BF16; a.src = p.hmid; a.src_ld = 4 * d;
      a.epi_kind = DG_EPI_RESID;
      break;
  }
  DEC_MARK(15);
  if (a.stage_kind == DG_STAGE_LN) dec_prefetch_ln(smem, a.ln_g, a.ln_b, d);
  uint8_t* sA = smem + DecSmem::kA;
  float* red = reinterpret_cast<float*>(smem + DecSmem::kRed);
  const float* s_g = reinterpret_cast<const float*>(smem + DecSmem::kLnGB);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2;
  const int t = lane & 3;
  const int N = a.N, n_chunks = a.n_chunks, K = n_chunks * d, kblocks = d >> 5;
  const int B = p.n_clips, rows_pad = ((B + 15) >> 4) << 4, n_mt = rows_pad >> 4;
  const int n_groups = (N + 7) >> 3;
  const int ksplit = dec_pick_ksplit(n_groups, kblocks, p.G);
  const int teams = kDecWarps / ksplit;
  const int team = warp / ksplit, kpart = warp - team * ksplit;
  const bool team_ok = team < teams;
  const int gpc = (n_groups + p.G - 1) / p.G;
  const int rounds = (gpc + teams - 1) / teams;
  const int kb_per = kblocks / ksplit;
  uint4 w_next[kDecWB];
  bool have_next = false;
  for (int round = 0; round < rounds; ++round) {
    const int grp = p.rank + p.G * (round * teams + team);
    const bool active = team_ok && grp < n_groups;
    const int n0 = grp * 8;
    float acc[kDecMT][4];
    float pre_b[4], pre_x[4];
#pragma unroll
    for (int i = 0; i < kDecMT; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (int chunk = 0; chunk < n_chunks; ++chunk) {
      const bool row_ok = active && n0 + g < N;
      const __nv_bfloat16* wrow = a.W + (size_t)(row_ok ? n0 + g : 0) * K + (size_t)chunk * d;
      const int kb_lo = kpart * kb_per, kb_hi = (kpart + 1) * kb_per;
      uint4 w[kDecWB];
      // the weights do not depend on the activations: their first batch is in flight while A is staged; with several
      // rounds or chunks the NEXT (round, chunk)'s first batch was requested before this one's mma (w_next)
      if (have_next) {
#pragma unroll
        for (int u = 0; u < kDecWB; ++u) w[u] = w_next[u];
      } else if (active) {
        dec_load_w(w, wrow, row_ok, kb_lo, kb_hi);
      }
      if (chunk == 0) {
        // this thread's epilogue elements (bias, and the residual it will update) ride the same L2 round trip
        const int tt = kpart * 32 + lane, nt = ksplit * 32;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int idx = tt + j * nt;
          const int r = idx >> 3, n = n0 + (idx & 7);
          pre_b[j] = 0.0f;
          pre_x[j] = 0.0f;
          if (active && idx < DecSmem::kRedPerWarp && r < B && n < N) {
            pre_b[j] = __ldg(a.bias + n);
            if (a.epi_kind == DG_EPI_RESID) pre_x[j] = __ldcg(a.out + (size_t)r * d + n);
          }
        }
      }
      DEC_MARK(11);
      if (n_chunks > 1 || round == 0) {
        if (a.stage_kind == DG_STAGE_LN) dec_prefetch_ln_wait();
        __syncthreads();
        if (a.stage_kind == DG_STAGE_LN)
          dec_stage_ln(sA, reinterpret_cast<const float*>(a.src), s_g, s_g + kDecMaxD, B, rows_pad, d);
        else if (a.stage_kind == DG_STAGE_F32)
          dec_stage_f32(sA, reinterpret_cast<const float*>(a.src), a.src_ld, B, rows_pad, d);
        else if (a.stage_kind == DG_STAGE_BF16)
          dec_stage_bf16(sA, reinterpret_cast<const __nv_bfloat16*>(a.src) + (size_t)chunk * d, a.src_ld, B, rows_pad, d);
        else
          dec_stage_xmerge(sA, p, a.xa_per, a.xa_cpp, B, rows_pad, d);
        __syncthreads();
      }
      DEC_MARK(12);
      {
        // request the first weight batch of the next (round, chunk) of this warp
        int nr = round, nc = chunk + 1;
        if (nc == n_chunks) { nc = 0; ++nr; }
        have_next = false;
        if (nr < rounds) {
          const int grp2 = p.rank + p.G * (nr * teams + team);
          const bool act2 = team_ok && grp2 < n_groups;
          const bool ok2 = act2 && grp2 * 8 + g < N;
          const __nv_bfloat16* wrow2 = a.W + (size_t)(ok2 ? grp2 * 8 + g : 0) * K + (size_t)nc * d;
          if (act2) dec_load_w(w_next, wrow2, ok2, kb_lo, kb_hi);
          have_next = act2;
        }
      }
      if (active) {
        dec_mma_w(w, sA, kb_lo, kb_hi, n_mt, acc);
        for (int kb = kb_lo + kDecWB; kb < kb_hi; kb += kDecWB) {
          dec_load_w(w, wrow, row_ok, kb, kb_hi);
          dec_mma_w(w, sA, kb, kb_hi, n_mt, acc);
        }
      }
    }
    DEC_MARK(13);
    // cross-warp (k-split) reduction through shared memory, then the epilogue by the whole team
    constexpr int kRW = DecSmem::kRedPerWarp;
    float* mine = red + warp * kRW;
#pragma unroll
    for (int mt = 0; mt < kDecMT; ++mt) {
      *reinterpret_cast<float2*>(mine + (mt * 16 + g) * 8 + 2 * t) = make_float2(acc[mt][0], acc[mt][1]);
      *reinterpret_cast<float2*>(mine + (mt * 16 + g + 8) * 8 + 2 * t) = make_float2(acc[mt][2], acc[mt][3]);
    }
    if (ksplit > 1) {
      if (team_ok) asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(ksplit * 32) : "memory");
    } else {
      __syncwarp();
    }
    if (active) {
      const float* base = red + team * ksplit * kRW;
      const int tt = kpart * 32 + lane, nt = ksplit * 32;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int idx = tt + j * nt;
        const int r = idx >> 3, n = n0 + (idx & 7);
        if (idx < kRW && r < B && n < N) {
          float v = pre_b[j];
          for (int k = 0; k < ksplit; ++k) v += base[k * kRW + idx];
          switch (a.epi_kind) {
            case DG_EPI_QKV:
              if (n < d) {
                a.out[(size_t)r * d + n] = v;
              } else {
                const int nn = n - d, which = nn >= d, c = which ? nn - d : nn;
                __nv_bfloat16* cache = which ? a.vc : a.kc;
                cache[(((size_t)(p.clip0 + r) * p.H + (c >> 6)) * 32 + a.pos) * 64 + (c & 63)] = __float2bfloat16(v);
              }
              break;
            case DG_EPI_RESID:
              a.out[(size_t)r * d + n] = pre_x[j] + v;
              break;
            case DG_EPI_STORE:
              a.out[(size_t)r * N + n] = v;
              break;
            default:
              a.out_bf16[(size_t)r * N + n] = __float2bfloat16(gelu_erf(v));
              break;
          }
        }
      }
    }
    if (ksplit > 1) {
      if (team_ok) asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(ksplit * 32) : "memory");
    } else {
      __syncwarp();
    }
    DEC_MARK(14);
  }
}

// ---------------------------------------------------------------------------------------------------
// Phases. Each is its own non-inlined function: registers are allocated per phase (no spills under the 128-register
// ceiling of a 512-thread CTA) and the persistent loop's code footprint stays small.
// ---------------------------------------------------------------------------------------------------

// ---- 2. self-attention over cache rows 0..pos: one warp per (clip, head), lane = key ----
__device__ __noinline__ void dec_self_attn(uint8_t* smem, const DecParams& p, const DecLayerDev& L, int pos) {
  float* s_misc = reinterpret_cast<float*>(smem + DecSmem::kMisc);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.d, H = p.H, B = p.n_clips;
  float* sq = s_misc + warp * 64;
  for (int pr = p.rank * kDecWarps + warp; pr < B * H; pr += p.G * kDecWarps) {
    const int b = pr / H, h = pr - b * H;
    const size_t base = ((size_t)(p.clip0 + b) * H + h) * 32 * 64;
    // every load of this pair is issued up front: K row `lane` (8 x 16 B) and the V column pair of all 32 rows
    uint4 kr[8];
    uint32_t vv[32];
    {
      // rows > pos of the cache hold stale (finite) data: loaded unconditionally, masked out of the softmax below
      const uint4* kp = reinterpret_cast<const uint4*>(L.kc + base + (size_t)lane * 64);
#pragma unroll
      for (int i = 0; i < 8; ++i) kr[i] = __ldcg(kp + i);
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) vv[j] = __ldcg(reinterpret_cast<const uint32_t*>(L.vc + base + (size_t)j * 64 + 2 * lane));
    const float2 q2 = __ldcg(reinterpret_cast<const float2*>(p.qs + (size_t)b * d + h * 64 + 2 * lane));
    sq[2 * lane] = q2.x * 0.125f;
    sq[2 * lane + 1] = q2.y * 0.125f;
    __syncwarp();
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float kf[8];
      bf16x8_to_f32(kr[i], kf);
#pragma unroll
      for (int j = 0; j < 8; ++j) s = fmaf(sq[i * 8 + j], kf[j], s);
    }
    if (lane > pos) s = -3.0e38f;
    const float m = warp_max(s);
    const float pw = (lane <= pos) ? __expf(s - m) : 0.0f;
    const float lsum = warp_sum(pw);
    float o0 = 0.0f, o1 = 0.0f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float pj = __shfl_sync(0xffffffffu, pw, j);  // 0 for j > pos
      o0 = fmaf(pj, __uint_as_float(vv[j] << 16), o0);
      o1 = fmaf(pj, __uint_as_float(vv[j] & 0xffff0000u), o1);
    }
    const float inv = 1.0f / lsum;
    *reinterpret_cast<float2*>(p.att + (size_t)b * d + h * 64 + 2 * lane) = make_float2(o0 * inv, o1 * inv);
    __syncwarp();
  }
}

// thread 0: request chunk j of layer l's K/V stream into the next ring stage (one contiguous 32 KB bulk copy)
__device__ __forceinline__ void xa_issue(uint8_t* smem, const DecParams& p, int l, int j, uint64_t policy) {
  XaHdr* hd = xa_hdr(smem);
  const int4* segs = xa_segs(smem);
  int si = 0;
  while (si + 1 < hd->nseg && segs[si + 1].w <= j) ++si;
  const int4 sg = segs[si];          // {pair, first chunk, end chunk, index of the first chunk in this CTA's sequence}
  const int c = sg.y + (j - sg.w);
  const unsigned n = hd->issued;
  const unsigned stg = n % kXaStages, use = n / kXaStages;
  mbar_wait(xa_empty(smem) + stg, (use & 1u) ^ 1u);
  uint8_t* st = smem + DecSmem::kRing + (size_t)stg * DecSmem::kStageBytes;
  const __nv_bfloat16* src =
      p.layers[l].kvx + (((size_t)p.clip0 * p.H + sg.x) * hd->cpp + c) * (size_t)(2 * kXaChunk * 64);
  mbar_expect_tx(xa_full(smem) + stg, (uint32_t)DecSmem::kStageBytes);
  bulk_load_stream(st, src, (uint32_t)DecSmem::kStageBytes, xa_full(smem) + stg, policy);
  hd->issued = n + 1;
}
// The first min(nchunks, kXaStages) chunks of layer l, requested before its query exists: they stream in underneath
// the projection phases that precede the cross-attention.
__device__ __noinline__ void dec_xa_prefetch(uint8_t* smem, const DecParams& p, int l) {
  if (threadIdx.x == 0 && !(p.debug & 2)) {
    const uint64_t policy = l2_evict_first_policy();
    const int nc = xa_hdr(smem)->nchunks;
    const int n = nc < kXaStages ? nc : kXaStages;
    for (int j = 0; j < n; ++j) xa_issue(smem, p, l, j, policy);
  }
}

// ---- 5. cross-attention: the group's (clip, head, key) space is cut into one contiguous range per CTA; K/V chunks
//         arrive through the TMA ring. `cons` = chunks consumed so far (uniform); returns the updated count. ----
__device__ __noinline__ unsigned dec_cross_attn(uint8_t* smem, const DecParams& p, int l, unsigned cons) {
  float* s_misc = reinterpret_cast<float*>(smem + DecSmem::kMisc);
  float* s_q = reinterpret_cast<float*>(smem + DecSmem::kXq);
  const uint8_t* ring = smem + DecSmem::kRing;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grp = lane >> 3, sub = lane & 7;
  const int d = p.d, H = p.H, S = p.S;
  const XaHdr* hd = xa_hdr(smem);
  const int4* segs = xa_segs(smem);
  const int nseg = hd->nseg, nchunks = hd->nchunks, per = hd->per, cpp = hd->cpp;
  uint64_t* full = xa_full(smem);
  uint64_t* empty = xa_empty(smem);
  float* sm_m = s_misc;               // [16]
  float* sm_l = s_misc + 16;          // [16]
  float* sm_acc = s_misc + 32;        // [16][64]
  const uint64_t policy = l2_evict_first_policy();
  // queries of every (clip, head) pair this CTA touches: one L2 round trip for the whole phase, pre-scaled
  {
    const float qscale = 0.125f * 1.4426950408889634f;
    for (int i = tid; i < nseg * 64; i += kDecThreads) {
      const int pr = segs[i >> 6].x, b = pr / H, h = pr - b * H;
      s_q[i] = __ldcg(p.qx + (size_t)b * d + h * 64 + (i & 63)) * qscale;
    }
  }
  __syncthreads();
  int next_issue = nchunks < kXaStages ? nchunks : kXaStages;  // already requested by dec_xa_prefetch
  for (int si = 0; si < nseg; ++si) {
    const int4 sg = segs[si];        // {pair, first chunk, end chunk, ...}
    const int pr = sg.x;
    float qv[8];
    {
      const float4 q0 = *reinterpret_cast<const float4*>(s_q + si * 64 + sub * 8);
      const float4 q1 = *reinterpret_cast<const float4*>(s_q + si * 64 + sub * 8 + 4);
      qv[0] = q0.x; qv[1] = q0.y; qv[2] = q0.z; qv[3] = q0.w;
      qv[4] = q1.x; qv[5] = q1.y; qv[6] = q1.z; qv[7] = q1.w;
    }
    float m = -1.0e30f, lsum = 0.0f, acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.0f;
    for (int c = sg.y; c < sg.z; ++c) {
      const int n_keys = (S - c * kXaChunk < kXaChunk) ? S - c * kXaChunk : kXaChunk;
      const unsigned stg = cons % kXaStages, use = cons / kXaStages;
      if (!(p.debug & 2)) mbar_wait(full + stg, use & 1u);
      const uint8_t* st = ring + (size_t)stg * DecSmem::kStageBytes;
      constexpr int kU = kXaChunk / (kDecWarps * 4);
      uint4 kq[kU], vq[kU];
      bool ok[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int kl = u * (kDecWarps * 4) + warp * 4 + grp;
        ok[u] = kl < n_keys;
        const int sw = (sub ^ (kl & 7)) * 16;   // the stream's 16-byte segments are swizzled by key & 7 (gemm_sm100.cuh)
        kq[u] = *reinterpret_cast<const uint4*>(st + (size_t)kl * 128 + sw);
        vq[u] = *reinterpret_cast<const uint4*>(st + (size_t)(kXaChunk + kl) * 128 + sw);
        if (!ok[u]) {  // rows past the clip's last key are padding (never written): keep them out of the math
          kq[u] = make_uint4(0u, 0u, 0u, 0u);
          vq[u] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      // values are in registers: hand the stage back; thread 0 then requests the chunk that will reuse it
      __syncwarp();
      if (!(p.debug & 2)) {
        if (lane == 0) mbar_arrive(empty + stg);
        ++cons;
        if (next_issue < nchunks) {
          if (tid == 0) xa_issue(smem, p, l, next_issue, policy);
          ++next_issue;
        }
      }
      if (p.debug & 1) continue;
      float sc[kU];
      float mx = m;
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        float s = 0.0f;
        float kf[8];
        bf16x8_to_f32(kq[u], kf);
#pragma unroll
        for (int i = 0; i < 8; ++i) s = fmaf(qv[i], kf[i], s);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        sc[u] = s;
        if (ok[u]) mx = fmaxf(mx, s);
      }
      const float scale = dec_ex2(m - mx);
      lsum *= scale;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] *= scale;
      m = mx;
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (ok[u]) {
          const float pw = dec_ex2(sc[u] - m);
          lsum += pw;
          float vf[8];
          bf16x8_to_f32(vq[u], vf);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = fmaf(pw, vf[i], acc[i]);
        }
      }
    }
    // merge the 4 key lanes of the warp (lanes with equal `sub` hold the same 8 dims)
#pragma unroll
    for (int o = 8; o <= 16; o <<= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float l2 = __shfl_xor_sync(0xffffffffu, lsum, o);
      const float mn = fmaxf(m, m2);
      const float s1 = dec_ex2(m - mn), s2 = dec_ex2(m2 - mn);
      lsum = lsum * s1 + l2 * s2;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float a2 = __shfl_xor_sync(0xffffffffu, acc[i], o);
        acc[i] = acc[i] * s1 + a2 * s2;
      }
      m = mn;
    }
    __syncthreads();  // s_misc free (previous segment's merge done)
    if (grp == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) sm_acc[warp * 64 + sub * 8 + i] = acc[i];
      if (sub == 0) {
        sm_m[warp] = m;
        sm_l[warp] = lsum;
      }
    }
    __syncthreads();
    // this CTA's partial (acc[64], max, sum) of pair `pr` goes to its slot; the out-projection's staging merges the
    // slots of all CTAs that cover the pair (DG_STAGE_XMERGE) after the cluster barrier — no tickets, no fences
    if (tid < 64) {
      const int first_cta = (pr * cpp) / per;
      float* pp = p.xpart + ((size_t)pr * p.xslots + (p.rank - first_cta)) * 66;
      float M = sm_m[0];
#pragma unroll
      for (int w = 1; w < kDecWarps; ++w) M = fmaxf(M, sm_m[w]);
      float Lw = 0.0f, O = 0.0f;
#pragma unroll
      for (int w = 0; w < kDecWarps; ++w) {
        const float s1 = dec_ex2(sm_m[w] - M);
        Lw = fmaf(sm_l[w], s1, Lw);
        O = fmaf(sm_acc[w * 64 + tid], s1, O);
      }
      pp[tid] = O;
      if (tid == 0) {
        pp[64] = M;
        pp[65] = Lw;
      }
    }
  }
  __syncthreads();  // every warp has consumed every chunk of this layer: the ring may be refilled for the next one
  return cons;
}

// ---- final LN -> tied logits with a fused (max, last index) per row (whisper.cpp:346-361's tie rule) ----
__device__ __noinline__ void dec_logits(uint8_t* smem, const DecParams& p, int pos) {
  uint8_t* sA = smem + DecSmem::kA;
  unsigned long long* s_best = reinterpret_cast<unsigned long long*>(smem + DecSmem::kRed);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int d = p.d, V = p.V, B = p.n_clips, G = p.G;
  const int rows_pad = ((B + 15) >> 4) << 4, n_mt = rows_pad >> 4;
  __syncthreads();
  for (int i = tid; i < kDecGroupMax; i += kDecThreads) s_best[i] = 0ull;
  dec_prefetch_ln(smem, p.lnf_g, p.lnf_b, d);
  dec_prefetch_ln_wait();
  __syncthreads();
  {
    const float* s_g = reinterpret_cast<const float*>(smem + DecSmem::kLnGB);
    dec_stage_ln(sA, p.x, s_g, s_g + kDecMaxD, B, rows_pad, d);
  }
  __syncthreads();
  const int n_groups = (V + 7) >> 3;
  const int step = pos - (p.prompt_len - 1);
  float* lg = p.step_logits ? p.step_logits + (size_t)step * p.step_logits_stride : nullptr;
  unsigned long long best[kDecMT][2];
#pragma unroll
  for (int i = 0; i < kDecMT; ++i) best[i][0] = best[i][1] = 0ull;
  // one vocabulary group of 8 rows per warp and iteration: all its weight loads (<= kDecWB 16-byte loads per lane)
  // are in flight together; the 16 warps of the CTA keep ~100 KB of the embedding stream in flight per SM
  const int kblocks = d >> 5;
  for (int grp = p.rank + G * warp; grp < n_groups; grp += G * kDecWarps) {
    const int n0 = grp * 8;
    const bool row_ok = n0 + g < V;
    const __nv_bfloat16* wrow = p.tok_emb + (size_t)(row_ok ? n0 + g : 0) * d;
    float acc[kDecMT][4];
#pragma unroll
    for (int i = 0; i < kDecMT; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (int kb = 0; kb < kblocks; kb += kDecWB) {
      uint4 w[kDecWB];
      dec_load_w(w, wrow, row_ok, kb, kblocks);
      dec_mma_w(w, sA, kb, kblocks, n_mt, acc);
    }
    const int c0 = n0 + 2 * t;
#pragma unroll
    for (int mt = 0; mt < kDecMT; ++mt) {
      if (mt < n_mt) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const int r = mt * 16 + g + hh * 8;
          if (r < B) {
            const float v0 = acc[mt][hh * 2], v1 = acc[mt][hh * 2 + 1];
            if (c0 < V) best[mt][hh] = u64_max(best[mt][hh], pack_val_idx(v0, (unsigned)c0));
            if (c0 + 1 < V) best[mt][hh] = u64_max(best[mt][hh], pack_val_idx(v1, (unsigned)(c0 + 1)));
            if (lg) {
              float* row = lg + (size_t)(p.clip0 + r) * V;
              if (c0 < V) row[c0] = v0;
              if (c0 + 1 < V) row[c0 + 1] = v1;
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int mt = 0; mt < kDecMT; ++mt)
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      unsigned long long v = best[mt][hh];
      v = u64_max(v, __shfl_xor_sync(0xffffffffu, v, 1));
      v = u64_max(v, __shfl_xor_sync(0xffffffffu, v, 2));
      if (t == 0 && v != 0ull) atomicMax(s_best + mt * 16 + g + hh * 8, v);
    }
  __syncthreads();
  for (int r = tid; r < B; r += kDecThreads) p.amax_part[(size_t)r * G + p.rank] = s_best[r];
}

// ---- finalize (CTA b): append the winner, EOT bookkeeping ----
__device__ __noinline__ void dec_finalize(uint8_t* smem, const DecParams& p, int pos) {
  unsigned long long* s_red = reinterpret_cast<unsigned long long*>(smem + DecSmem::kMisc);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = p.G;
  for (int b = p.rank; b < p.n_clips; b += G) {
    unsigned long long v = 0ull;
    for (int i = tid; i < G; i += kDecThreads) v = u64_max(v, __ldcg(p.amax_part + (size_t)b * G + i));
    v = warp_max_u64(v);
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < kDecWarps; ++w) v = u64_max(v, s_red[w]);
      const int tok = (int)(v & 0xffffffffull);
      const int cb = p.clip0 + b;
      if (!p.done[cb]) {
        p.tokens[(size_t)cb * kDecTokStride + pos + 1] = tok;
        p.lens[cb] = pos + 2;
        if (tok == p.eot) {
          p.done[cb] = 1;
          atomicAdd(p.finished, 1);
        }
      }
      __threadfence();
    }
    __syncthreads();
  }
}

// x[b] = tok_emb[token(b, pos)] + pos_emb[pos]  (CTA b; a finished clip keeps stepping on token 0, unused)
__device__ __noinline__ void dec_embed(const DecParams& p, int pos) {
  const int d = p.d;
  for (int b = p.rank; b < p.n_clips; b += p.G) {
    int tok = __ldcg(p.tokens + (size_t)(p.clip0 + b) * kDecTokStride + pos);
    if (tok < 0) tok = 0;
    const __nv_bfloat16* te = p.tok_emb + (size_t)tok * d;
    const __nv_bfloat16* pe = p.pos_emb + (size_t)pos * d;
    for (int c = threadIdx.x; c < d; c += kDecThreads)
      p.x[(size_t)b * d + c] = __bfloat162float(te[c]) + __bfloat162float(pe[c]);
  }
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kDecThreads, 1) decoder_persistent_kernel(const DecParams p_in) {
  extern __shared__ __align__(128) uint8_t dec_smem[];
  const int tid = threadIdx.x;
  // this cluster's view (its clip group, its slices of the scratch buffers) is built once in shared memory
  DecParams& p = *reinterpret_cast<DecParams*>(dec_smem + DecSmem::kParams);
  static_assert(sizeof(DecParams) <= 512, "DecParams outgrew its shared-memory slot");
  static_assert(sizeof(DecLayerDev) <= 256, "DecLayerDev outgrew its shared-memory slot");
  const int G = (int)dec_cluster_nctarank();
  const int rank = (int)dec_cluster_ctarank();
  const int group = (int)dec_clusterid_x();
  {
    const int gc = p_in.group_clips;
    const int first = group * gc;
    const int nb = (p_in.n_clips - first < gc) ? p_in.n_clips - first : gc;
    if (nb <= 0 || p_in.n_layers > kDecMaxLayers) return;  // whole cluster leaves together
    if (tid == 0) {
      p = p_in;
      p.G = G;
      p.rank = rank;
      p.clip0 = p_in.clip0 + first;
      p.n_clips = nb;
      const size_t row0 = (size_t)group * kDecGroupMax;
      p.x += row0 * p.d;
      p.qs += row0 * p.d;
      p.att += row0 * p.d;
      p.qx += row0 * p.d;
      p.hmid += row0 * 4 * p.d;
      p.xpart += row0 * p.H * (size_t)p.xslots * 66;
      p.xtickets += row0 * p.H;
      p.amax_part += row0 * kDecMaxCluster;
      p.finished += group;
      if (group != 0 || rank != 0) p.trace = nullptr;
      p.layers = reinterpret_cast<const DecLayerDev*>(dec_smem + DecSmem::kLayers);
    }
    // layer pointer tables: global -> shared (slots of 256 bytes)
    const int words = (int)(sizeof(DecLayerDev) / 8);
    for (int i = tid; i < p_in.n_layers * words; i += kDecThreads) {
      const int li = i / words, w = i - li * words;
      reinterpret_cast<unsigned long long*>(dec_smem + DecSmem::kLayers + li * 256)[w] =
          reinterpret_cast<const unsigned long long*>(p_in.layers + li)[w];
    }
    __syncthreads();
  }
  const int d = p.d, B = p.n_clips;
  unsigned long long tlast = (p.trace != nullptr && threadIdx.x == 0) ? dec_globaltimer() : 0ull;

  // ---- cross-attention stream set-up: this CTA's key range (the same for every layer and position) as a table of
  //      (pair, first key, end key, first chunk) segments; full/empty mbarriers of the K/V ring ----
  if (tid == 0) {
    XaHdr* hd = xa_hdr(dec_smem);
    int4* segs = xa_segs(dec_smem);
    const int cpp = (p.S + kXaChunk - 1) / kXaChunk;
    const int T = B * p.H * cpp;
    const int per = (T + G - 1) / G;
    int cur = rank * per;
    const int hi = (cur + per < T) ? cur + per : T;
    int nseg = 0, nchunks = 0;
    while (cur < hi && nseg < kXaMaxSeg) {
      const int pr = cur / cpp;
      const int c0 = cur - pr * cpp;
      const int c1 = (hi - cur < cpp - c0) ? c0 + (hi - cur) : cpp;
      segs[nseg] = make_int4(pr, c0, c1, nchunks);
      nchunks += c1 - c0;
      ++nseg;
      cur += c1 - c0;
    }
    hd->nseg = nseg;
    hd->nchunks = nchunks;
    hd->issued = 0u;
    hd->per = per;
    hd->cpp = cpp;
    for (int i = 0; i < kXaStages; ++i) {
      mbar_init(xa_full(dec_smem) + i, 1);
      mbar_init(xa_empty(dec_smem) + i, kDecWarps);
    }
    fence_barrier_init();
  }
  __syncthreads();
  unsigned xa_cons = 0;  // chunks consumed so far (uniform over the CTA)

  // ---- init: lens, done (CTA b), first embedding ----
  for (int b = rank; b < B; b += G) {
    if (tid == 0) {
      p.lens[p.clip0 + b] = p.prompt_len;
      p.done[p.clip0 + b] = 0;
    }
  }
  dec_embed(p, 0);
  dec_cluster_sync(p, 0, tlast);

  for (int pos = 0; pos <= p.max_total - 2; ++pos) {
    dec_xa_prefetch(dec_smem, p, 0);  // layer 0's first K/V chunks stream in underneath its self-attention phases
    for (int l = 0; l < p.n_layers; ++l) {
      const DecLayerDev& L = p.layers[l];
      dec_gemm(dec_smem, p, DG_QKV, l, pos);     // 1. LN1 -> QKV; q -> qs, k/v -> cache row `pos`
      dec_cluster_sync(p, 1, tlast);
      dec_self_attn(dec_smem, p, L, pos);        // 2. self-attention over cache rows 0..pos
      dec_cluster_sync(p, 2, tlast);
      dec_gemm(dec_smem, p, DG_OUT, l, pos);     // 3. out-proj + residual
      dec_cluster_sync(p, 3, tlast);
      dec_gemm(dec_smem, p, DG_XQ, l, pos);      // 4. LNx -> cross-attention query
      dec_cluster_sync(p, 4, tlast);
      if (p.debug & 4) {  // timing experiment: the same (idempotent) phase again, now with warm instruction caches
        dec_gemm(dec_smem, p, DG_XQ, l, pos);
        dec_cluster_sync(p, 15, tlast);
      }
      // 5. cross-attention (HBM stream); then the next layer's first chunks are requested
      xa_cons = dec_cross_attn(dec_smem, p, l, xa_cons);
      if (l + 1 < p.n_layers) dec_xa_prefetch(dec_smem, p, l + 1);
      dec_cluster_sync(p, 5, tlast);
      dec_gemm(dec_smem, p, DG_XOUT, l, pos);    // 6. merge of the cross-attention partials -> out-proj + residual
      dec_cluster_sync(p, 6, tlast);
      dec_gemm(dec_smem, p, DG_FC1, l, pos);     // 7. LN2 -> fc1 -> GELU
      dec_cluster_sync(p, 7, tlast);
      dec_gemm(dec_smem, p, DG_FC2, l, pos);     // 8. fc2 + residual (K = 4d in 4 staged chunks)
      dec_cluster_sync(p, 8, tlast);
    }
    if (pos >= p.prompt_len - 1) {
      dec_logits(dec_smem, p, pos);
      dec_cluster_sync(p, 9, tlast);
      dec_finalize(dec_smem, p, pos);
    }
    if (pos + 1 <= p.max_total - 2) dec_embed(p, pos + 1);
    dec_cluster_sync(p, 10, tlast);
    if (__ldcg(p.finished) >= B) break;  // every clip has produced EOT (whisper.cpp:397-399, per clip)
  }
}

}  // namespace wt
