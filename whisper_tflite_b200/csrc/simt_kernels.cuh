// SIMT kernels of the Whisper hot path: the fp32 "verification mode" GEMM, and the element type
// templated (float | bf16) LayerNorm, encoder attention, KV-cached decode attention, embedding and
// greedy argmax used by both precisions. They replace what the reference runs inside
// tflite::Interpreter::Invoke (whisper.cpp:295, :375) — graph per export/generate_onnx.py:85-130.
// Compiles under nvcc and under tests/emu (CPU SIMT emulator) for logic checks without a GPU.
#pragma once

#ifndef WT_CUDA_EMU
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#endif
#include <stdint.h>
#include "common.cuh"

namespace wt {

template <class T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <class T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
// Same function for the bf16 epilogues: erfc by Abramowitz & Stegun 7.1.28, erfc(z) = (1 + a1 z + ... + a6 z^6)^-16
// (|error| <= 3e-7; measured 9.4e-7 on x Phi(x) in f32, far below one bf16 ulp of the result): a 6-term Horner, four
// squarings and ONE MUFU reciprocal per element — the epilogue of a 128 x 256 fc1 tile is MUFU-bound with two.
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  float p = fmaf(0.0000430638f, z, 0.0002765672f);
  p = fmaf(p, z, 0.0001520143f);
  p = fmaf(p, z, 0.0092705272f);
  p = fmaf(p, z, 0.0422820123f);
  p = fmaf(p, z, 0.0705230784f);
  p = fmaf(p, z, 1.0f);
  p *= p; p *= p; p *= p; p *= p;                   // ^16 (overflows to inf for |x| > ~25: erfc -> 0, as it should)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p));
  const float c = 0.5f * r;                         // 0.5 erfc(|x| / sqrt 2) = Phi(-|x|)
  return x * (x < 0.0f ? c : 1.0f - c);             // x Phi(x), no cancellation in the negative tail
}

// The GEMM epilogues' GELU: x Phi(x) with Phi(x) = 0.5 (1 + tanh(x (a + b x^2 + c x^4))), (a, b, c) a minimax fit of
// the erf form itself (NOT the 0.044715 "tanh GELU": that one is 4.7e-4 off; this is 2.5e-5 off in exact arithmetic and
// 3.0e-5 on the device with tanh.approx, tools/probes/gelu_probe.cu — a small fraction of a bf16 ulp of any output
// that matters). x^2 is clamped at 49 because the quartic term turns the inner polynomial over beyond |x| = 7.25;
// tanh.approx saturates to exactly +-1, so the tails are x and -0. 8 instructions, one MUFU (A&S 7.1.28 above: 17).
__device__ __forceinline__ float gelu_tanh_fit(float x) {
  const float x2 = fminf(x * x, 49.0f);
  float p = fmaf(x2, -3.51516783e-4f, 0.037005646f);
  p = fmaf(x2, p, 0.797507884f);
  const float u = x * p;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  const float h = 0.5f * x;
  return fmaf(h, t, h);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// --------------------------------------------------------------------------------------------
// Row addressing shared by every GEMM in the engine. Row m of a logical [M][*] operand lives at
//   base + (m / rows_per_batch) * batch_stride + (m % rows_per_batch) * row_stride      (elements)
// which expresses (i) plain matrices (rows_per_batch = M), (ii) the k=3 convolutions as GEMMs over
// OVERLAPPING rows of a zero-padded time-major activation (row_stride < K), (iii) writes into the
// interior of such a padded buffer.
// --------------------------------------------------------------------------------------------
struct RowMap {
  long long batch_stride;
  int rows_per_batch;
  int row_stride;
  __host__ __device__ long long at(int m) const {
    return (long long)(m / rows_per_batch) * batch_stride + (long long)(m % rows_per_batch) * row_stride;
  }
};

enum : int { EPI_BIAS = 1, EPI_GELU = 2, EPI_RESIDUAL = 4, EPI_POS = 8 };

struct GemmF32Params {
  const float* A; RowMap amap;      // [M][K] via amap
  const float* W;                   // [N][K] row-major (torch Linear layout)
  const float* bias;                // [N] or null
  const float* residual; RowMap rmap;  // [M][N] via rmap (EPI_RESIDUAL)
  const float* pos; int pos_rows;   // [pos_rows][N], row = m % pos_rows (EPI_POS)
  float* C; RowMap cmap;            // [M][N] via cmap
  int M, N, K, flags;
};

// C = epi(A . W^T): 64x64 tile, BK = 16, 256 threads, 4x4 register micro-tile, f32 FMA.
constexpr int kSgBM = 64, kSgBN = 64, kSgBK = 16;
__global__ void __launch_bounds__(256) sgemm_f32_kernel(GemmF32Params p) {
  __shared__ float sA[kSgBK][kSgBM + 4];
  __shared__ float sB[kSgBK][kSgBN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * kSgBM, n0 = blockIdx.x * kSgBN;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, each 4 (m) x 4 (n)
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
  // loader mapping: 64 rows x 16 k = 1024 elements, 4 per thread
  const int lrow = tid >> 2;         // 0..63
  const int lk = (tid & 3) * 4;      // 0,4,8,12
  const int am = m0 + lrow;
  const float* arow = (am < p.M) ? p.A + p.amap.at(am) : nullptr;
  const int bn = n0 + lrow;
  const float* brow = (bn < p.N) ? p.W + (long long)bn * p.K : nullptr;
  for (int k0 = 0; k0 < p.K; k0 += kSgBK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = k0 + lk + i;
      sA[lk + i][lrow] = (arow && k < p.K) ? arow[k] : 0.0f;
      sB[lk + i][lrow] = (brow && k < p.K) ? brow[k] : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSgBK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sB[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
    float* crow = p.C + p.cmap.at(m);
    const float* rrow = (p.flags & EPI_RESIDUAL) ? p.residual + p.rmap.at(m) : nullptr;
    const float* prow = (p.flags & EPI_POS) ? p.pos + (long long)(m % p.pos_rows) * p.N : nullptr;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float v = acc[i][j];
      if (p.flags & EPI_BIAS) v += p.bias[n];
      if (p.flags & EPI_GELU) v = gelu_erf(v);
      if (prow) v += prow[n];
      if (rrow) v += rrow[n];
      crow[n] = v;
    }
  }
}

// --------------------------------------------------------------------------------------------
// LayerNorm over the last dim (eps 1e-5), one warp per row, two-pass statistics in f32.
// D <= 1024, D % 32 == 0.
// --------------------------------------------------------------------------------------------
template <class TI, class TO>
__global__ void __launch_bounds__(256) layernorm_kernel(const TI* x, const float* gamma, const float* beta, TO* y,
                                                         int rows, int D) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const TI* xr = x + (long long)warp * D;
  float v[32];
  const int per = D / 32;
  float s = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; ++i)
    if (i < per) {
      v[i] = to_f32<TI>(xr[i * 32 + lane]);
      s += v[i];
    }
  const float mean = warp_sum(s) / (float)D;
  float q = 0.0f;
#pragma unroll
  for (int i = 0; i < 32; ++i)
    if (i < per) {
      const float d = v[i] - mean;
      q += d * d;
    }
  const float rstd = rsqrtf(warp_sum(q) / (float)D + 1e-5f);
  TO* yr = y + (long long)warp * D;
#pragma unroll
  for (int i = 0; i < 32; ++i)
    if (i < per) {
      const int c = i * 32 + lane;
      yr[c] = from_f32<TO>((v[i] - mean) * rstd * gamma[c] + beta[c]);
    }
}

// --------------------------------------------------------------------------------------------
// Encoder self-attention, SIMT form (verification mode; also the bf16 path's stand-in until the
// tcgen05 kernel takes over): one thread owns one query row of one (clip, head), K/V tiles are
// staged in shared memory and read as warp-wide broadcasts, online softmax in f32.
// qkv: [B*S][3*D] with q | k | v blocks of D = H*64 each; q is pre-scaled by head_dim^-0.5 on read.
// out: [B*S][D].  grid = (ceil(S/128), H, B), block = 128.
// --------------------------------------------------------------------------------------------
constexpr int kAttKT = 32;  // keys per shared tile
template <class T>
__global__ void __launch_bounds__(128) attention_simt_kernel(const T* qkv, T* out, int S, int H) {
  __shared__ float sK[kAttKT][64];
  __shared__ float sV[kAttKT][64];
  const int D = H * 64;
  const int b = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * 128 + threadIdx.x;
  const bool active = qi < S;
  const T* base = qkv + (long long)b * S * 3 * D;
  float q[64], o[64];
  {
    const T* qr = base + (long long)(active ? qi : 0) * 3 * D + h * 64;
#pragma unroll
    for (int i = 0; i < 64; ++i) {
      q[i] = to_f32<T>(qr[i]) * 0.125f;
      o[i] = 0.0f;
    }
  }
  float mrun = -3.0e38f, lrun = 0.0f;
  for (int k0 = 0; k0 < S; k0 += kAttKT) {
    __syncthreads();
    for (int i = threadIdx.x; i < kAttKT * 64; i += 128) {
      const int kr = i >> 6, c = i & 63;
      const int key = k0 + kr;
      float kv = 0.0f, vv = 0.0f;
      if (key < S) {
        const T* row = base + (long long)key * 3 * D + h * 64 + c;
        kv = to_f32<T>(row[D]);
        vv = to_f32<T>(row[2 * D]);
      }
      sK[kr][c] = kv;
      sV[kr][c] = vv;
    }
    __syncthreads();
    const int kmax = (S - k0 < kAttKT) ? (S - k0) : kAttKT;
    for (int kr = 0; kr < kmax; ++kr) {
      float s = 0.0f;
#pragma unroll
      for (int i = 0; i < 64; ++i) s = fmaf(q[i], sK[kr][i], s);
      if (s > mrun) {
        const float sc = __expf(mrun - s);
        lrun *= sc;
#pragma unroll
        for (int i = 0; i < 64; ++i) o[i] *= sc;
        mrun = s;
      }
      const float pw = __expf(s - mrun);
      lrun += pw;
#pragma unroll
      for (int i = 0; i < 64; ++i) o[i] = fmaf(pw, sV[kr][i], o[i]);
    }
  }
  if (active) {
    const float inv = 1.0f / lrun;
    T* orow = out + ((long long)b * S + qi) * D + h * 64;
#pragma unroll
    for (int i = 0; i < 64; ++i) orow[i] = from_f32<T>(o[i] * inv);
  }
}

// --------------------------------------------------------------------------------------------
// Decoder step pieces (one new position per clip per step, KV cache instead of the reference's
// full re-forward of all T tokens — whisper.cpp:368-375 — same values, 1/23 of the work).
// --------------------------------------------------------------------------------------------

// x[b][:] = tok_emb[token[b]][:] + pos_emb[pos][:]   grid = B, block = 128
template <class T>
__global__ void embed_kernel(const int* tokens, int tok_stride, int pos, const T* tok_emb, const T* pos_emb, T* x,
                             int D) {
  const int b = blockIdx.x;
  int tok = tokens[b * tok_stride + pos];
  if (tok < 0) tok = 0;  // clip already finished (EOT seen): row is computed but never consumed
  for (int c = threadIdx.x; c < D; c += blockDim.x)
    x[(long long)b * D + c] = from_f32<T>(to_f32<T>(tok_emb[(long long)tok * D + c]) + to_f32<T>(pos_emb[(long long)pos * D + c]));
}

// Appends this step's k,v (from qkv[b][D..3D)) to the self-attention cache and attends q over
// positions 0..pos. cache layout [B][H][max_t][64] for k and v. grid = (H, B), block = 64.
template <class T>
__global__ void __launch_bounds__(64) decode_self_attn_kernel(const T* qkv, T* kcache, T* vcache, T* out, int pos,
                                                               int max_t, int H) {
  __shared__ float sq[64];
  __shared__ float sp[64];
  const int h = blockIdx.x, b = blockIdx.y, D = H * 64;
  const int c = threadIdx.x;
  const T* row = qkv + (long long)b * 3 * D + h * 64;
  T* kc = kcache + ((long long)(b * H + h) * max_t) * 64;
  T* vc = vcache + ((long long)(b * H + h) * max_t) * 64;
  sq[c] = to_f32<T>(row[c]) * 0.125f;
  kc[(long long)pos * 64 + c] = row[D + c];
  vc[(long long)pos * 64 + c] = row[2 * D + c];
  __syncthreads();
  // thread c scores key c (pos < 64 always: the loop stops at 30 positions, whisper.cpp:364)
  float s = -3.0e38f;
  if (c <= pos) {
    s = 0.0f;
    for (int i = 0; i < 64; ++i) s = fmaf(sq[i], to_f32<T>(kc[(long long)c * 64 + i]), s);
  }
  sp[c] = s;
  __syncthreads();
  float m = -3.0e38f;
  for (int i = 0; i <= pos; ++i) m = fmaxf(m, sp[i]);
  float l = 0.0f, o = 0.0f;
  for (int i = 0; i <= pos; ++i) {
    const float pw = __expf(sp[i] - m);
    l += pw;
    o = fmaf(pw, to_f32<T>(vc[(long long)i * 64 + c]), o);
  }
  out[(long long)b * D + h * 64 + c] = from_f32<T>(o / l);
}

// Cross-attention for one decode step: q [B][D] against the per-clip encoder K/V
// ([B][H][S][64], written once per clip by the cross-K/V projection). This is the HBM-bound part of a
// decode step (every step re-reads all K/V: 2*S*D elements per clip per layer), so the kernel is a pure
// stream: the S keys are split over `nsplit` blocks per (clip, head); 8 lanes share a key (one 16-byte
// load each, 4 keys = 512 contiguous bytes per warp instruction, 4 such loads in flight per lane);
// pass 1 parks the scores in shared memory, a block max/sum gives the softmax weights with no per-key
// rescaling chain, pass 2 streams V. Each block writes a partial (acc[64], max, sum); the last block to
// finish (atomic ticket) merges the partials. grid = (nsplit, H, B), block = 128 (4 warps).
constexpr int kXaMaxSplit = 16;
constexpr int kXaMaxKeys = 1536;

template <class T> __device__ __forceinline__ void load8(const T* p, float* f);
template <> __device__ __forceinline__ void load8<float>(const float* p, float* f) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float* f) {
  const uint4 v = *reinterpret_cast<const uint4*>(p);
  const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = __uint_as_float(w[j] << 16);
    f[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}

template <class T>
__global__ void __launch_bounds__(128) decode_cross_attn_kernel(const T* q, const T* kx, const T* vx, T* out,
                                                                 float* part, unsigned* tickets, int S, int H) {
  __shared__ float s_sc[kXaMaxKeys];
  __shared__ float s_red[4];
  __shared__ float s_o[4][64];
  __shared__ unsigned s_last;
  const int sp = blockIdx.x, nsplit = gridDim.x, h = blockIdx.y, b = blockIdx.z, D = H * 64;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grp = lane >> 3, sub = lane & 7;  // 4 keys per warp instruction, 8 lanes (x 8 dims) per key
  int per = (S + nsplit - 1) / nsplit;
  per = (per + 15) & ~15;
  const int k_begin = sp * per;
  const int k_end = (k_begin + per < S) ? (k_begin + per) : S;
  const int nk = k_end > k_begin ? k_end - k_begin : 0;
  const T* kb = kx + ((long long)(b * H + h) * S) * 64;
  const T* vb = vx + ((long long)(b * H + h) * S) * 64;
  float qv[8];
  load8<T>(q + (long long)b * D + h * 64 + sub * 8, qv);
#pragma unroll
  for (int i = 0; i < 8; ++i) qv[i] *= 0.125f;

  // ---- pass 1: scores -> shared ----
  for (int kw = warp * 4; kw < nk; kw += 64) {  // warp-uniform trip count (shuffles below)
    const int k0 = kw + grp;                      // 4 keys in flight per lane: k0, k0+16, k0+32, k0+48
    float kf[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int kk = k0 + 16 * u;
      if (kk < nk) load8<T>(kb + (long long)(k_begin + kk) * 64 + sub * 8, kf[u]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int kk = k0 + 16 * u;
      float s = 0.0f;
      if (kk < nk) {
#pragma unroll
        for (int i = 0; i < 8; ++i) s = fmaf(qv[i], kf[u][i], s);
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      if (sub == 0 && kk < nk) s_sc[kk] = s;
    }
  }
  __syncthreads();
  // ---- block max, weights, sum ----
  float m = -3.0e38f;
  for (int i = tid; i < nk; i += 128) m = fmaxf(m, s_sc[i]);
  m = warp_max(m);
  if (lane == 0) s_red[warp] = m;
  __syncthreads();
  m = fmaxf(fmaxf(s_red[0], s_red[1]), fmaxf(s_red[2], s_red[3]));
  __syncthreads();
  float l = 0.0f;
  for (int i = tid; i < nk; i += 128) {
    const float pw = __expf(s_sc[i] - m);
    s_sc[i] = pw;
    l += pw;
  }
  l = warp_sum(l);
  if (lane == 0) s_red[warp] = l;
  __syncthreads();
  l = s_red[0] + s_red[1] + s_red[2] + s_red[3];
  // ---- pass 2: acc[dims] += p * V ----
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.0f;
  for (int kw = warp * 4; kw < nk; kw += 64) {
    const int k0 = kw + grp;
    float vf[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int kk = k0 + 16 * u;
      if (kk < nk) load8<T>(vb + (long long)(k_begin + kk) * 64 + sub * 8, vf[u]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int kk = k0 + 16 * u;
      if (kk < nk) {
        const float pw = s_sc[kk];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fmaf(pw, vf[u][i], acc[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 8);
    acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], 16);
  }
  if (grp == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) s_o[warp][sub * 8 + i] = acc[i];
  }
  __syncthreads();
  float* pp = part + (((long long)(b * H + h) * kXaMaxSplit) + sp) * 66;
  if (tid < 64) {
    pp[tid] = s_o[0][tid] + s_o[1][tid] + s_o[2][tid] + s_o[3][tid];
    if (tid == 0) { pp[64] = m; pp[65] = l; }
  }
#ifndef WT_CUDA_EMU
  __threadfence();
#endif
  __syncthreads();
  if (tid == 0) {
    const unsigned t = atomicAdd(tickets + (b * H + h), 1u);
    s_last = (t == (unsigned)(nsplit - 1)) ? 1u : 0u;
    if (s_last) tickets[b * H + h] = 0u;  // re-arm for the next launch
  }
  __syncthreads();
  if (s_last) {
#ifndef WT_CUDA_EMU
    __threadfence();
#endif
    if (tid < 64) {
      const float* p0 = part + ((long long)(b * H + h) * kXaMaxSplit) * 66;
      float M = -3.0e38f;
      for (int i = 0; i < nsplit; ++i) M = fmaxf(M, p0[i * 66 + 64]);
      float L = 0.0f, O = 0.0f;
      for (int i = 0; i < nsplit; ++i) {
        const float sc = __expf(p0[i * 66 + 64] - M);
        L += p0[i * 66 + 65] * sc;
        O += p0[i * 66 + tid] * sc;
      }
      out[(long long)b * D + h * 64 + tid] = from_f32<T>(O / L);
    }
  }
}

// Re-lays one cross-attention projection out of the GEMM's [B*S][2D] (k | v) row-major result into
// the per-head streams the decode kernel reads: kx, vx [B][H][S][64].
template <class T>
__global__ void __launch_bounds__(256) split_heads_kv_kernel(const T* kv, T* kx, T* vx, int B, int S, int H) {
  const int D = H * 64;
  const long long total = (long long)B * S * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % D);
    const long long row = i / D;
    const int s = (int)(row % S), b = (int)(row / S);
    const int h = c >> 6, e = c & 63;
    const long long dst = (((long long)(b * H + h) * S) + s) * 64 + e;
    kx[dst] = kv[row * 2 * D + c];
    vx[dst] = kv[row * 2 * D + D + c];
  }
}

// --------------------------------------------------------------------------------------------
// Greedy argmax with the reference's tie rule (whisper.cpp:346-361: `>=` scan -> LAST maximal
// index wins; a NaN never wins unless it is element 0). Packs (ordered value, index) into one u64
// so "max value, then max index" is a single integer max.
// --------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long pack_val_idx(float v, unsigned idx) {
  unsigned u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // monotone map float -> unsigned
  // NaN never satisfies `*p >= max_value` (whisper.cpp:353), so it loses everywhere — except at index 0, where it IS the
  // initial max_value (:348) and no later element can replace it
  if (v != v) u = idx == 0u ? 0xffffffffu : 0u;
  return ((unsigned long long)u << 32) | idx;
}
__device__ __forceinline__ unsigned long long u64_max(unsigned long long a, unsigned long long b) { return a > b ? a : b; }
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v = u64_max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// logits [B][V] f32 -> appends the winner to tokens[b][pos+1] unless the clip is finished.
// state per clip: lens[b] (tokens so far), done[b]. grid = B, block = 256.
__global__ void __launch_bounds__(256) argmax_append_kernel(const float* logits, int V, int* tokens, int tok_stride,
                                                             int* lens, int* done, int pos, int eot) {
  __shared__ unsigned long long s_best[8];
  const int b = blockIdx.x;
  const float* row = logits + (long long)b * V;
  unsigned long long best = 0ull;
  for (int i = threadIdx.x; i < V; i += 256) best = u64_max(best, pack_val_idx(row[i], (unsigned)i));
  best = warp_max_u64(best);
  if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) best = u64_max(best, s_best[w]);
    const int tok = (int)(best & 0xffffffffull);
    if (!done[b]) {
      tokens[b * tok_stride + pos + 1] = tok;
      lens[b] = pos + 2;
      if (tok == eot) done[b] = 1;
    }
  }
}

// second stage for the fused-epilogue path: partial (value,index) maxima [B][n_part] u64
__global__ void __launch_bounds__(128) argmax_partials_kernel(const unsigned long long* part, int n_part, int* tokens,
                                                               int tok_stride, int* lens, int* done, int pos, int eot) {
  __shared__ unsigned long long s_best[4];
  const int b = blockIdx.x;
  unsigned long long best = 0ull;
  for (int i = threadIdx.x; i < n_part; i += 128) best = u64_max(best, part[(long long)b * n_part + i]);
  best = warp_max_u64(best);
  if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 4; ++w) best = u64_max(best, s_best[w]);
    const int tok = (int)(best & 0xffffffffull);
    if (!done[b]) {
      tokens[b * tok_stride + pos + 1] = tok;
      lens[b] = pos + 2;
      if (tok == eot) done[b] = 1;
    }
  }
}

// tokens[b][0..prompt_len) = prompt, lens = prompt_len, done = 0.  grid = B, block = 32
__global__ void init_decode_state_kernel(const int* prompt, int prompt_len, int* tokens, int tok_stride, int* lens,
                                         int* done) {
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < tok_stride; i += blockDim.x) tokens[b * tok_stride + i] = (i < prompt_len) ? prompt[i] : -1;
  if (threadIdx.x == 0) { lens[b] = prompt_len; done[b] = 0; }
}

// f32 [B][n_mel][n_len] (reference layout, whisper.cpp:184) -> zero-padded time-major
// [B][n_len+2][n_mel] so conv1 (k=3, pad 1) becomes a GEMM over overlapping rows.
// If raw != null, applies the per-clip clamp + affine (whisper.cpp:205-213) on the fly.
template <class T>
__global__ void __launch_bounds__(256) mel_to_time_major_kernel(const float* mel, const int* clip_max, T* out, int n_mel,
                                                                 int n_len) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * 32, j0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const float floor_v = clip_max ? (ordered_to_float(clip_max[b]) - 8.0f) : 0.0f;
  for (int r = ty; r < 32; r += 8) {
    const int j = j0 + r, t = t0 + tx;
    float v = 0.0f;
    if (j < n_mel && t < n_len) {
      v = mel[((long long)b * n_mel + j) * n_len + t];
      if (clip_max) v = (fmaxf(v, floor_v) + 4.0f) * 0.25f;
    }
    tile[r][tx] = v;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int t = t0 + r, j = j0 + tx;
    if (t < n_len && j < n_mel) out[((long long)b * (n_len + 2) + t + 1) * n_mel + j] = from_f32<T>(tile[tx][r]);
  }
}

// bf16 rows with D % 128 == 0 (384 / 512 / 768 / 1024): 8-byte loads and stores (256 contiguous bytes per warp access),
// two rows per warp in flight. HBM-bound: D * 4 bytes per row.
template <int PER>  // uint2 (4 bf16) per lane and row = D / 128
__global__ void __launch_bounds__(256) layernorm_bf16_vec_kernel(const __nv_bfloat16* __restrict__ x,
                                                                  const float* __restrict__ gamma,
                                                                  const float* __restrict__ beta,
                                                                  __nv_bfloat16* __restrict__ y, int rows) {
  constexpr int D = PER * 128;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int r0 = warp * 2;
  if (r0 >= rows) return;
  const bool two = r0 + 1 < rows;
  const uint2* xa = reinterpret_cast<const uint2*>(x + (long long)r0 * D);
  const uint2* xb = reinterpret_cast<const uint2*>(x + (long long)(two ? r0 + 1 : r0) * D);
  uint2 ra[PER], rb[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    ra[i] = __ldg(xa + i * 32 + lane);
    rb[i] = __ldg(xb + i * 32 + lane);
  }
  float va[PER][4], vb[PER][4];
  float sa = 0.0f, sb = 0.0f;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    va[i][0] = __uint_as_float(ra[i].x << 16); va[i][1] = __uint_as_float(ra[i].x & 0xffff0000u);
    va[i][2] = __uint_as_float(ra[i].y << 16); va[i][3] = __uint_as_float(ra[i].y & 0xffff0000u);
    vb[i][0] = __uint_as_float(rb[i].x << 16); vb[i][1] = __uint_as_float(rb[i].x & 0xffff0000u);
    vb[i][2] = __uint_as_float(rb[i].y << 16); vb[i][3] = __uint_as_float(rb[i].y & 0xffff0000u);
    sa += (va[i][0] + va[i][1]) + (va[i][2] + va[i][3]);
    sb += (vb[i][0] + vb[i][1]) + (vb[i][2] + vb[i][3]);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    sa += __shfl_xor_sync(0xffffffffu, sa, o);
    sb += __shfl_xor_sync(0xffffffffu, sb, o);
  }
  const float ma = sa * (1.0f / D), mb = sb * (1.0f / D);
  float qa = 0.0f, qb = 0.0f;
#pragma unroll
  for (int i = 0; i < PER; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      va[i][j] -= ma;
      vb[i][j] -= mb;
      qa = fmaf(va[i][j], va[i][j], qa);
      qb = fmaf(vb[i][j], vb[i][j], qb);
    }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    qa += __shfl_xor_sync(0xffffffffu, qa, o);
    qb += __shfl_xor_sync(0xffffffffu, qb, o);
  }
  const float ia = rsqrtf(qa * (1.0f / D) + 1e-5f), ib = rsqrtf(qb * (1.0f / D) + 1e-5f);
  uint2* ya = reinterpret_cast<uint2*>(y + (long long)r0 * D);
  uint2* yb = reinterpret_cast<uint2*>(y + (long long)(r0 + 1) * D);
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i * 32 + lane);
    const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + i * 32 + lane);
    __nv_bfloat162 o0 = __floats2bfloat162_rn(va[i][0] * ia * g.x + b.x, va[i][1] * ia * g.y + b.y);
    __nv_bfloat162 o1 = __floats2bfloat162_rn(va[i][2] * ia * g.z + b.z, va[i][3] * ia * g.w + b.w);
    ya[i * 32 + lane] = make_uint2(*reinterpret_cast<uint32_t*>(&o0), *reinterpret_cast<uint32_t*>(&o1));
    if (two) {
      o0 = __floats2bfloat162_rn(vb[i][0] * ib * g.x + b.x, vb[i][1] * ib * g.y + b.y);
      o1 = __floats2bfloat162_rn(vb[i][2] * ib * g.z + b.z, vb[i][3] * ib * g.w + b.w);
      yb[i * 32 + lane] = make_uint2(*reinterpret_cast<uint32_t*>(&o0), *reinterpret_cast<uint32_t*>(&o1));
    }
  }
}
template <class TI, class TO>
__global__ void __launch_bounds__(256) convert_kernel(const TI* in, TO* out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = from_f32<TO>(to_f32<TI>(in[i]));
}

// Debug: order-independent 64-bit digest of a buffer of 32-bit words: sum over words of mix(index, word). Two runs that
// left the same bytes in the buffer give the same digest whatever the launch geometry (wt_debug_hash).
__device__ __forceinline__ unsigned long long hash_mix64(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__global__ void __launch_bounds__(256) hash_words_kernel(const uint32_t* __restrict__ w, long long n, unsigned long long* out) {
  unsigned long long acc = 0ull;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    acc += hash_mix64(((unsigned long long)i << 32) ^ (unsigned long long)w[i]);
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

}  // namespace wt
