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

#include "simt_kernels.cuh"

namespace wt {

constexpr int kDecThreads = 512;
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kDecGroupMax = 16;   // clips owned by one cluster: rows of one staged A tile (kDecMT m-tiles of 16)
constexpr int kDecMT = kDecGroupMax / 16;
constexpr int kDecMaxCluster = 16; // CTAs per cluster (non-portable size; 8 is the portable fallback)
constexpr int kDecMaxD = 768;      // d_model ceiling of this kernel (tiny 384, base 512, small 768)
constexpr int kDecTokStride = 32;

struct DecLayerDev {
  const float *ln1_g, *ln1_b, *lnx_g, *lnx_b, *ln2_g, *ln2_b;
  const __nv_bfloat16* w_qkv; const float* b_qkv;    // [3d][d]
  const __nv_bfloat16* w_out; const float* b_out;    // [d][d]
  const __nv_bfloat16* w_xq; const float* b_xq;      // [d][d]
  const __nv_bfloat16* w_xout; const float* b_xout;  // [d][d]
  const __nv_bfloat16* w_fc1; const float* b_fc1;    // [4d][d]
  const __nv_bfloat16* w_fc2; const float* b_fc2;    // [d][4d]
  const __nv_bfloat16 *kx, *vx;                      // [B][H][S][64] cross-attention K / V (encoder side)
  __nv_bfloat16 *kc, *vc;                            // [B][H][32][64] self-attention cache
};

struct DecParams {
  const DecLayerDev* layers;
  int n_layers, n_clips, clip0, d, H, S, V;
  int group_clips;                // clips owned by one cluster (<= kDecGroupMax)
  int rank, G;                    // filled in by the kernel: CTA rank in its cluster, cluster size
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
  static constexpr int kTotal = kMisc + ((kMiscBytes + 127) & ~127);
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
  __threadfence();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (tr) {
    const unsigned long long t = dec_globaltimer();
    p.trace[kind] += t - tlast;
    p.trace[16 + kind] += t - t_arr;
    tlast = t;
  }
}

// ---- A staging: all n_rows (padded to 16) x d into shared memory as bf16, row stride DecSmem::kAStride ----
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
          const float2 g = __ldg(reinterpret_cast<const float2*>(gamma + c));
          const float2 bb = __ldg(reinterpret_cast<const float2*>(beta + c));
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

// acc[mt][0..3] += A[mt*16 + {g, g+8}][kb-range] . W[n0 + g][kb-range]   (one warp: 64 x 8 outputs)
__device__ __forceinline__ void dec_warp_mma(const uint8_t* sA, const __nv_bfloat16* wrow, bool row_ok, int kb0, int kb1,
                                             int n_mt, float (&acc)[kDecMT][4]) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const uint8_t* a_base = sA + (size_t)g * DecSmem::kAStride + t * 16;
  for (int kb = kb0; kb < kb1; kb += 4) {
    uint4 w[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      w[u] = make_uint4(0u, 0u, 0u, 0u);
      if (row_ok && kb + u < kb1) w[u] = ldg_nc_v4(wrow + (kb + u) * 32 + t * 8);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (kb + u < kb1) {
#pragma unroll
        for (int mt = 0; mt < kDecMT; ++mt) {
          if (mt < n_mt) {
            const uint8_t* ap = a_base + (size_t)(mt * 16) * DecSmem::kAStride + (kb + u) * 64;
            const uint4 lo = *reinterpret_cast<const uint4*>(ap);
            const uint4 hi = *reinterpret_cast<const uint4*>(ap + 8 * DecSmem::kAStride);
            mma_bf16_16816(acc[mt], lo.x, hi.x, lo.y, hi.y, w[u].x, w[u].y);
            mma_bf16_16816(acc[mt], lo.z, hi.z, lo.w, hi.w, w[u].z, w[u].w);
          }
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

// out[r][n] = epi(sum_k A[r][k] * W[n][k]) for r < n_clips, n < N; K = n_chunks * d; stage(chunk) fills sA.
template <class StageFn, class EpiFn>
__device__ __forceinline__ void dec_gemm(uint8_t* smem, const DecParams& p, const __nv_bfloat16* W, int N, int n_chunks,
                                         StageFn stage, EpiFn epi) {
  uint8_t* sA = smem + DecSmem::kA;
  float* red = reinterpret_cast<float*>(smem + DecSmem::kRed);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int d = p.d, K = n_chunks * d, kblocks = d >> 5;
  const int n_mt = (p.n_clips + 15) >> 4;
  const int n_groups = (N + 7) >> 3;
  const int ksplit = dec_pick_ksplit(n_groups, kblocks, p.G);
  const int teams = kDecWarps / ksplit;
  const int team = warp / ksplit, kpart = warp - team * ksplit;
  const bool team_ok = team < teams;
  const int gpc = (n_groups + p.G - 1) / p.G;
  const int rounds = (gpc + teams - 1) / teams;
  const int kb_per = kblocks / ksplit;
  for (int round = 0; round < rounds; ++round) {
    const int grp = p.rank + p.G * (round * teams + team);
    const bool active = team_ok && grp < n_groups;
    const int n0 = grp * 8;
    float acc[kDecMT][4];
#pragma unroll
    for (int i = 0; i < kDecMT; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (int chunk = 0; chunk < n_chunks; ++chunk) {
      if (n_chunks > 1 || round == 0) {
        __syncthreads();
        stage(chunk);
        __syncthreads();
      }
      if (active) {
        const bool row_ok = n0 + g < N;
        const __nv_bfloat16* wrow = W + (size_t)(row_ok ? n0 + g : 0) * K + (size_t)chunk * d;
        dec_warp_mma(sA, wrow, row_ok, kpart * kb_per, (kpart + 1) * kb_per, n_mt, acc);
      }
    }
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
      for (int idx = tt; idx < kRW; idx += nt) {
        const int r = idx >> 3, c = idx & 7;
        if (r < p.n_clips && n0 + c < N) {
          float v = 0.0f;
          for (int k = 0; k < ksplit; ++k) v += base[k * kRW + idx];
          epi(r, n0 + c, v);
        }
      }
    }
    if (ksplit > 1) {
      if (team_ok) asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(ksplit * 32) : "memory");
    } else {
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kDecThreads, 1) decoder_persistent_kernel(const DecParams p_in) {
  extern __shared__ __align__(128) uint8_t dec_smem[];
  uint8_t* sA = dec_smem + DecSmem::kA;
  float* s_misc = reinterpret_cast<float*>(dec_smem + DecSmem::kMisc);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // this cluster's view: its clip group, its slices of the scratch buffers
  DecParams p = p_in;
  const int G = (int)dec_cluster_nctarank();
  const int rank = (int)dec_cluster_ctarank();
  const int group = (int)dec_clusterid_x();
  p.G = G;
  p.rank = rank;
  {
    const int gc = p_in.group_clips;
    const int first = group * gc;
    p.clip0 = p_in.clip0 + first;
    p.n_clips = (p_in.n_clips - first < gc) ? p_in.n_clips - first : gc;
    if (p.n_clips <= 0) return;  // whole cluster leaves together
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
  }
  const int d = p.d, H = p.H, S = p.S, V = p.V, B = p.n_clips;
  const int rows_pad = ((B + 15) >> 4) << 4;
  const int n_mt = rows_pad >> 4;
  unsigned long long tlast = (p.trace != nullptr && threadIdx.x == 0) ? dec_globaltimer() : 0ull;
  const uint64_t pol_stream = l2_evict_first_policy();

  // x[b] = tok_emb[token(b, pos)] + pos_emb[pos]  (CTA b; a finished clip keeps stepping on token 0, unused)
  auto embed = [&](int pos) {
    for (int b = rank; b < B; b += G) {
      int tok = __ldcg(p.tokens + (size_t)(p.clip0 + b) * kDecTokStride + pos);
      if (tok < 0) tok = 0;
      const __nv_bfloat16* te = p.tok_emb + (size_t)tok * d;
      const __nv_bfloat16* pe = p.pos_emb + (size_t)pos * d;
      for (int c = tid; c < d; c += kDecThreads) p.x[(size_t)b * d + c] = __bfloat162float(te[c]) + __bfloat162float(pe[c]);
    }
  };

  // ---- init: token rows, lens, done (CTA b), first embedding ----
  for (int b = rank; b < B; b += G) {
    if (tid == 0) {
      p.lens[p.clip0 + b] = p.prompt_len;
      p.done[p.clip0 + b] = 0;
    }
  }
  embed(0);
  dec_cluster_sync(p, 0, tlast);

  for (int pos = 0; pos <= p.max_total - 2; ++pos) {
    for (int l = 0; l < p.n_layers; ++l) {
      const DecLayerDev& L = p.layers[l];
      // ---- 1. LN1 -> QKV; q -> qs, k/v -> cache row `pos` ----
      dec_gemm(dec_smem, p, L.w_qkv, 3 * d, 1, [&](int) { dec_stage_ln(sA, p.x, L.ln1_g, L.ln1_b, B, rows_pad, d); },
               [&](int r, int n, float v) {
                 v += __ldg(L.b_qkv + n);
                 if (n < d) {
                   p.qs[(size_t)r * d + n] = v;
                 } else {
                   const int nn = n - d, which = nn >= d, c = which ? nn - d : nn;
                   __nv_bfloat16* cache = which ? L.vc : L.kc;
                   cache[(((size_t)(p.clip0 + r) * H + (c >> 6)) * 32 + pos) * 64 + (c & 63)] = __float2bfloat16(v);
                 }
               });
      dec_cluster_sync(p, 1, tlast);
      // ---- 2. self-attention over cache rows 0..pos: one warp per (clip, head), lane = key ----
      {
        float* sq = s_misc + warp * 64;
        for (int pr = rank * kDecWarps + warp; pr < B * H; pr += G * kDecWarps) {
          const int b = pr / H, h = pr - b * H;
          const float2 q2 = __ldcg(reinterpret_cast<const float2*>(p.qs + (size_t)b * d + h * 64 + 2 * lane));
          sq[2 * lane] = q2.x * 0.125f;
          sq[2 * lane + 1] = q2.y * 0.125f;
          __syncwarp();
          const size_t base = ((size_t)(p.clip0 + b) * H + h) * 32 * 64;
          float s = -3.0e38f;
          if (lane <= pos) {
            s = 0.0f;
            const uint4* kr = reinterpret_cast<const uint4*>(L.kc + base + (size_t)lane * 64);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float kf[8];
              bf16x8_to_f32(__ldcg(kr + i), kf);
#pragma unroll
              for (int j = 0; j < 8; ++j) s = fmaf(sq[i * 8 + j], kf[j], s);
            }
          }
          const float m = warp_max(s);
          const float pw = (lane <= pos) ? __expf(s - m) : 0.0f;
          const float lsum = warp_sum(pw);
          float o0 = 0.0f, o1 = 0.0f;
          for (int j = 0; j <= pos; ++j) {
            const float pj = __shfl_sync(0xffffffffu, pw, j);
            const uint32_t v2 = __ldcg(reinterpret_cast<const uint32_t*>(L.vc + base + (size_t)j * 64 + 2 * lane));
            o0 = fmaf(pj, __uint_as_float(v2 << 16), o0);
            o1 = fmaf(pj, __uint_as_float(v2 & 0xffff0000u), o1);
          }
          const float inv = 1.0f / lsum;
          *reinterpret_cast<float2*>(p.att + (size_t)b * d + h * 64 + 2 * lane) = make_float2(o0 * inv, o1 * inv);
          __syncwarp();
        }
      }
      dec_cluster_sync(p, 2, tlast);
      // ---- 3. out-proj + residual ----
      dec_gemm(dec_smem, p, L.w_out, d, 1, [&](int) { dec_stage_f32(sA, p.att, d, B, rows_pad, d); },
               [&](int r, int n, float v) {
                 float* xp = p.x + (size_t)r * d + n;
                 *xp = __ldcg(xp) + v + __ldg(L.b_out + n);
               });
      dec_cluster_sync(p, 3, tlast);
      // ---- 4. LNx -> cross-attention query ----
      dec_gemm(dec_smem, p, L.w_xq, d, 1, [&](int) { dec_stage_ln(sA, p.x, L.lnx_g, L.lnx_b, B, rows_pad, d); },
               [&](int r, int n, float v) { p.qx[(size_t)r * d + n] = v + __ldg(L.b_xq + n); });
      dec_cluster_sync(p, 4, tlast);
      // ---- 5. cross-attention: the group's (clip, head, key) space is cut into one contiguous range per CTA ----
      {
        const long long T = (long long)B * H * S;
        const long long per = (T + G - 1) / G;
        long long cur = (long long)rank * per;
        const long long hi = (cur + per < T) ? cur + per : T;
        const int grp = lane >> 3, sub = lane & 7;
        float* sm_m = s_misc;               // [16]
        float* sm_l = s_misc + 16;          // [16]
        float* sm_acc = s_misc + 32;        // [16][64]
        unsigned* s_flag = reinterpret_cast<unsigned*>(s_misc + 32 + kDecWarps * 64);
        const float qscale = 0.125f * 1.4426950408889634f;
        while (cur < hi) {
          const int pr = (int)(cur / S);
          const int a = (int)(cur - (long long)pr * S);
          const int bnd = (int)((hi - cur < (long long)(S - a)) ? a + (hi - cur) : S);
          const int b = pr / H, h = pr - b * H;
          float qv[8];
          {
            const float4 q0 = __ldcg(reinterpret_cast<const float4*>(p.qx + (size_t)b * d + h * 64 + sub * 8));
            const float4 q1 = __ldcg(reinterpret_cast<const float4*>(p.qx + (size_t)b * d + h * 64 + sub * 8 + 4));
            qv[0] = q0.x * qscale; qv[1] = q0.y * qscale; qv[2] = q0.z * qscale; qv[3] = q0.w * qscale;
            qv[4] = q1.x * qscale; qv[5] = q1.y * qscale; qv[6] = q1.z * qscale; qv[7] = q1.w * qscale;
          }
          const __nv_bfloat16* kb_ = L.kx + ((size_t)(p.clip0 + b) * H + h) * (size_t)S * 64 + sub * 8;
          const __nv_bfloat16* vb_ = L.vx + ((size_t)(p.clip0 + b) * H + h) * (size_t)S * 64 + sub * 8;
          float m = -1.0e30f, lsum = 0.0f, acc[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = 0.0f;
          for (int k0 = a; k0 < bnd; k0 += 64 * 4) {
            uint4 kq[4], vq[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int kk = k0 + u * 64 + warp * 4 + grp;
              ok[u] = kk < bnd;
              if (ok[u]) {
                kq[u] = ldg_stream_v4(kb_ + (size_t)kk * 64, pol_stream);
                vq[u] = ldg_stream_v4(vb_ + (size_t)kk * 64, pol_stream);
              }
            }
            float sc[4];
            float mx = m;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              float s = 0.0f;
              if (ok[u]) {
                float kf[8];
                bf16x8_to_f32(kq[u], kf);
#pragma unroll
                for (int i = 0; i < 8; ++i) s = fmaf(qv[i], kf[i], s);
              }
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
            for (int u = 0; u < 4; ++u) {
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
          // slot of this CTA among the CTAs that cover pair `pr`
          const int first_cta = (int)(((long long)pr * S) / per);
          const int last_cta = (int)((((long long)pr + 1) * S - 1) / per);
          const int slot = rank - first_cta;
          float* pp = p.xpart + ((size_t)pr * p.xslots + slot) * 66;
          if (tid < 64) {
            float M = sm_m[0];
            for (int w = 1; w < kDecWarps; ++w) M = fmaxf(M, sm_m[w]);
            float Lw = 0.0f, O = 0.0f;
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
          __threadfence();
          __syncthreads();
          if (tid == 0) {
            const unsigned n_slots = (unsigned)(last_cta - first_cta + 1);
            const unsigned tk = atomicAdd(p.xtickets + pr, 1u);
            const unsigned last = (tk == n_slots - 1) ? 1u : 0u;
            if (last) p.xtickets[pr] = 0u;  // re-arm
            *s_flag = last;
          }
          __syncthreads();
          if (*s_flag) {
            __threadfence();
            if (tid < 64) {
              const int n_slots = last_cta - first_cta + 1;
              const float* p0 = p.xpart + (size_t)pr * p.xslots * 66;
              float M = -1.0e30f;
              for (int i = 0; i < n_slots; ++i) M = fmaxf(M, __ldcg(p0 + i * 66 + 64));
              float Lw = 0.0f, O = 0.0f;
              for (int i = 0; i < n_slots; ++i) {
                const float s1 = dec_ex2(__ldcg(p0 + i * 66 + 64) - M);
                Lw = fmaf(__ldcg(p0 + i * 66 + 65), s1, Lw);
                O = fmaf(__ldcg(p0 + i * 66 + tid), s1, O);
              }
              p.att[(size_t)b * d + h * 64 + tid] = O / Lw;
            }
          }
          cur += bnd - a;
        }
      }
      dec_cluster_sync(p, 5, tlast);
      // ---- 6. cross out-proj + residual ----
      dec_gemm(dec_smem, p, L.w_xout, d, 1, [&](int) { dec_stage_f32(sA, p.att, d, B, rows_pad, d); },
               [&](int r, int n, float v) {
                 float* xp = p.x + (size_t)r * d + n;
                 *xp = __ldcg(xp) + v + __ldg(L.b_xout + n);
               });
      dec_cluster_sync(p, 6, tlast);
      // ---- 7. LN2 -> fc1 -> GELU ----
      dec_gemm(dec_smem, p, L.w_fc1, 4 * d, 1, [&](int) { dec_stage_ln(sA, p.x, L.ln2_g, L.ln2_b, B, rows_pad, d); },
               [&](int r, int n, float v) { p.hmid[(size_t)r * 4 * d + n] = __float2bfloat16(gelu_erf(v + __ldg(L.b_fc1 + n))); });
      dec_cluster_sync(p, 7, tlast);
      // ---- 8. fc2 + residual (K = 4d in 4 staged chunks) ----
      dec_gemm(dec_smem, p, L.w_fc2, d, 4, [&](int chunk) { dec_stage_bf16(sA, p.hmid + (size_t)chunk * d, 4 * d, B, rows_pad, d); },
               [&](int r, int n, float v) {
                 float* xp = p.x + (size_t)r * d + n;
                 *xp = __ldcg(xp) + v + __ldg(L.b_fc2 + n);
               });
      dec_cluster_sync(p, 8, tlast);
    }

    const bool want_logits = pos >= p.prompt_len - 1;
    if (want_logits) {
      // ---- final LN -> tied logits with fused (max, last index) per row ----
      unsigned long long* s_best = reinterpret_cast<unsigned long long*>(dec_smem + DecSmem::kRed);
      __syncthreads();
      for (int i = tid; i < kDecGroupMax; i += kDecThreads) s_best[i] = 0ull;
      dec_stage_ln(sA, p.x, p.lnf_g, p.lnf_b, B, rows_pad, d);
      __syncthreads();
      const int g = lane >> 2, t = lane & 3;
      const int n_groups = (V + 7) >> 3;
      const int step = pos - (p.prompt_len - 1);
      float* lg = p.step_logits ? p.step_logits + (size_t)step * p.step_logits_stride : nullptr;
      unsigned long long best[kDecMT][2];
#pragma unroll
      for (int i = 0; i < kDecMT; ++i) best[i][0] = best[i][1] = 0ull;
      for (int grp = rank + G * warp; grp < n_groups; grp += G * kDecWarps) {
        const int n0 = grp * 8;
        float acc[kDecMT][4];
#pragma unroll
        for (int i = 0; i < kDecMT; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
        const bool row_ok = n0 + g < V;
        dec_warp_mma(sA, p.tok_emb + (size_t)(row_ok ? n0 + g : 0) * d, row_ok, 0, d >> 5, n_mt, acc);
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
      for (int r = tid; r < B; r += kDecThreads) p.amax_part[(size_t)r * G + rank] = s_best[r];
      dec_cluster_sync(p, 9, tlast);
    }
    // ---- finalize (CTA b): append the winner, EOT bookkeeping; then embed the next position ----
    for (int b = rank; b < B; b += G) {
      if (want_logits) {
        unsigned long long* s_red = reinterpret_cast<unsigned long long*>(s_misc);
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
    if (pos + 1 <= p.max_total - 2) embed(pos + 1);
    dec_cluster_sync(p, 10, tlast);
    if (__ldcg(p.finished) >= B) break;  // every clip has produced EOT (whisper.cpp:397-399, per clip)
  }
}

}  // namespace wt
