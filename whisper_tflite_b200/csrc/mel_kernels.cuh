// Batched log-mel frontend for sm_100a — replaces whisper::log_mel_spectrogram
// (/root/reference/whisper.tflite/whisper.cpp:109-216) for B clips per launch.
//
// Same definition as the reference (periodic Hann(400), hop 160, no centre padding, zero fill past
// the end, |X|^2 with the mirrored-bin fold = x2 on bins 1..199, 80x201 mel, log10(max(.,1e-10)),
// per-clip max-8 clamp, (x+4)/4) — but NOT its algorithm: the reference runs a recursive radix-2
// FFT with O(N^2) DFT leaves per frame on CPU threads; here one half-warp owns a frame and runs a
// real-input mixed-radix FFT 400 = 16 x 25 entirely in registers:
//   A. lane r (0..15) takes the residue class x[16m+r]*hann, m<25, and does a REAL 25-point DFT
//      (5x5, Hermitian half only: 13 complex outputs),
//   B. multiplies by W_400^{r k'} and parks 13 complex values in shared memory,
//   C. lane k' (0..12) runs a complex 16-point FFT over r -> bins k'+25q and, by conjugate
//      symmetry, (25-k')+25(15-q): every one of the 201 bins exactly once,
//   D. lanes then share the 80 mel rows (sparse rows: 391 non-zeros for the slaney bank, the fold's
//      factor 2 pre-multiplied into the weights), log10 via MUFU lg2, and stage a [mel][32-frame]
//      tile so global stores are full 128-byte lines of the reference's [n_mel][n_len] layout.
// PCM is staged once per 32-frame tile with 128-bit loads (frames overlap 2.5x) into a skewed
// shared layout that keeps the stride-16 residue reads of two frames per warp conflict-free.
// HBM traffic = algorithmic: 1.92 MB in + 0.96 MB out per clip (+ one float per clip for the max).
#pragma once

#ifndef WT_CUDA_EMU
#include <cuda_runtime.h>
#define WT_DYN_SMEM(name) extern __shared__ __align__(1024) unsigned char name[]
#else
#define WT_DYN_SMEM(name) unsigned char* name = emu::g_dyn_smem
#endif
#include "common.cuh"

namespace wt {

constexpr int kFFT = 400;          // whisper.h:35
constexpr int kHop = 160;          // whisper.h:37
constexpr int kBins = 201;         // whisper.cpp:129
constexpr int kTileFrames = 32;    // frames per tile (one 128-byte line per mel row)
constexpr int kMelThreads = 256;   // 16 frames in flight x 16 lanes
constexpr int kSpan = (kTileFrames - 1) * kHop + kFFT;  // 5360 samples feed one tile
constexpr int kSpanPhys = kSpan + 16 * (kSpan / kHop + 1);
constexpr int kExStride = 16 * 13;  // float2 per frame in the exchange buffer
constexpr int kPsStride = 208;      // floats per frame in the power buffer (201 used)
constexpr int kOutStride = 34;      // floats per mel row in the output stage (32 used)
constexpr int kMaxMel = 128;

struct MelParams {
  const float* pcm;        // [B][n_samples] f32, or null when pcm_i16 is given
  const short* pcm_i16;    // [B][n_samples] int16: converted like wav_read_legacy / wav_read do on the host,
                           // static_cast<float>(s) / static_cast<float>(INT16_MAX) (wav_util.cpp:70-73, :120-122)
  float* raw;              // [B][n_mel][n_len]  log10 mel before the per-clip clamp
  int* clip_max;           // [B] ordered-int encoded max of raw (init INT_MIN)
  const float* hann;       // [400] f32 of double 0.5*(1-cos(2*pi*i/400))  (whisper.cpp:117-120)
  const float2* tw400;     // [16][13]  (cos, -sin)(2*pi*r*k/400)
  const float* fw;         // packed non-zero filter weights, fold factor applied
  const int2* frow;        // [n_mel] {offset into fw, first bin}
  const int* flen;         // [n_mel] number of consecutive bins
  int n_samples, n_len, n_mel, n_tiles;
};

struct cpx { float re, im; };
__device__ __forceinline__ cpx cadd(cpx a, cpx b) { return {a.re + b.re, a.im + b.im}; }
__device__ __forceinline__ cpx csub(cpx a, cpx b) { return {a.re - b.re, a.im - b.im}; }
__device__ __forceinline__ cpx cmul(cpx a, float c, float s) {  // a * (c + i s)
  return {a.re * c - a.im * s, a.re * s + a.im * c};
}

#define WT_C1 0.30901699437494745f   /* cos(2pi/5) */
#define WT_C2 (-0.8090169943749473f) /* cos(4pi/5) */
#define WT_S1 0.9510565162951535f    /* sin(2pi/5) */
#define WT_S2 0.5877852522924731f    /* sin(4pi/5) */

// real 5-point DFT: X0 real, X1, X2 complex (X3 = conj X2, X4 = conj X1)
__device__ __forceinline__ void rdft5(float a0, float a1, float a2, float a3, float a4, float& x0, cpx& x1, cpx& x2) {
  float t1 = a1 + a4, t2 = a2 + a3, t3 = a1 - a4, t4 = a2 - a3;
  x0 = a0 + t1 + t2;
  x1.re = a0 + WT_C1 * t1 + WT_C2 * t2;
  x1.im = -(WT_S1 * t3 + WT_S2 * t4);
  x2.re = a0 + WT_C2 * t1 + WT_C1 * t2;
  x2.im = -(WT_S2 * t3 - WT_S1 * t4);
}

// complex 5-point DFT, forward (e^{-i...})
__device__ __forceinline__ void cdft5(cpx z0, cpx z1, cpx z2, cpx z3, cpx z4, cpx& Z0, cpx& Z1, cpx& Z2, cpx& Z3,
                                      cpx& Z4) {
  cpx t1 = cadd(z1, z4), t2 = cadd(z2, z3), t3 = csub(z1, z4), t4 = csub(z2, z3);
  Z0 = {z0.re + t1.re + t2.re, z0.im + t1.im + t2.im};
  cpx a1 = {z0.re + WT_C1 * t1.re + WT_C2 * t2.re, z0.im + WT_C1 * t1.im + WT_C2 * t2.im};
  cpx a2 = {z0.re + WT_C2 * t1.re + WT_C1 * t2.re, z0.im + WT_C2 * t1.im + WT_C1 * t2.im};
  cpx b1 = {WT_S1 * t3.re + WT_S2 * t4.re, WT_S1 * t3.im + WT_S2 * t4.im};
  cpx b2 = {WT_S2 * t3.re - WT_S1 * t4.re, WT_S2 * t3.im - WT_S1 * t4.im};
  Z1 = {a1.re + b1.im, a1.im - b1.re};  // a1 - i b1
  Z4 = {a1.re - b1.im, a1.im + b1.re};  // a1 + i b1
  Z2 = {a2.re + b2.im, a2.im - b2.re};
  Z3 = {a2.re - b2.im, a2.im + b2.re};
}

// (cos, -sin)(2*pi*j/25) for the inner 5x5 twiddles
__device__ __forceinline__ void tw25(int j, float& c, float& s) {
  switch (j) {
    case 1: c = 0.9685831611286311f; s = -0.2486898871648548f; break;
    case 2: c = 0.8763066800438636f; s = -0.4817536741017153f; break;
    case 3: c = 0.7289686274214116f; s = -0.6845471059286887f; break;
    case 4: c = 0.5358267949789965f; s = -0.8443279255020151f; break;
    case 6: c = 0.0627905195293133f; s = -0.9980267284282716f; break;
    case 8: c = -0.4257792915650727f; s = -0.9048270524660195f; break;
    default: c = 1.0f; s = 0.0f; break;
  }
}

// Real 25-point DFT of a[0..24] -> S[0..12] (Hermitian half).  m = 5*m1 + m2, k = k1 + 5*k2.
__device__ __forceinline__ void rdft25(const float* a, cpx* S) {
  float u0[5];
  cpx u1[5], u2[5];
#pragma unroll
  for (int m2 = 0; m2 < 5; ++m2) rdft5(a[m2], a[5 + m2], a[10 + m2], a[15 + m2], a[20 + m2], u0[m2], u1[m2], u2[m2]);
#pragma unroll
  for (int m2 = 1; m2 < 5; ++m2) {
    float c, s;
    tw25(m2, c, s);
    u1[m2] = cmul(u1[m2], c, s);
    tw25(2 * m2, c, s);
    u2[m2] = cmul(u2[m2], c, s);
  }
  cpx x1, x2;
  float x0;
  rdft5(u0[0], u0[1], u0[2], u0[3], u0[4], x0, x1, x2);  // k1 = 0 column is real
  S[0] = {x0, 0.0f};
  S[5] = x1;
  S[10] = x2;
  cpx Z0, Z1, Z2, Z3, Z4;
  cdft5(u1[0], u1[1], u1[2], u1[3], u1[4], Z0, Z1, Z2, Z3, Z4);  // k = 1, 6, 11, 16, 21
  S[1] = Z0;
  S[6] = Z1;
  S[11] = Z2;
  S[9] = {Z3.re, -Z3.im};  // S[16] = conj S[9]
  S[4] = {Z4.re, -Z4.im};  // S[21] = conj S[4]
  cdft5(u2[0], u2[1], u2[2], u2[3], u2[4], Z0, Z1, Z2, Z3, Z4);  // k = 2, 7, 12, 17, 22
  S[2] = Z0;
  S[7] = Z1;
  S[12] = Z2;
  S[8] = {Z3.re, -Z3.im};
  S[3] = {Z4.re, -Z4.im};
}

// (cos, -sin)(2*pi*j/16), j = 0..7
__device__ __forceinline__ void tw16(int j, float& c, float& s) {
  switch (j) {
    case 1: c = 0.9238795325112867f; s = -0.3826834323650898f; break;
    case 2: c = 0.7071067811865476f; s = -0.7071067811865476f; break;
    case 3: c = 0.3826834323650898f; s = -0.9238795325112867f; break;
    case 4: c = 0.0f; s = -1.0f; break;
    case 5: c = -0.3826834323650898f; s = -0.9238795325112867f; break;
    case 6: c = -0.7071067811865476f; s = -0.7071067811865476f; break;
    case 7: c = -0.9238795325112867f; s = -0.3826834323650898f; break;
    default: c = 1.0f; s = 0.0f; break;
  }
}

// In-place radix-2 decimation-in-frequency FFT-16; v[i] ends up holding X[bitrev4(i)].
__device__ __forceinline__ void fft16_dif(cpx* v) {
#pragma unroll
  for (int s = 8; s >= 1; s >>= 1) {
#pragma unroll
    for (int g = 0; g < 16; g += 2 * s) {
#pragma unroll
      for (int j = 0; j < s; ++j) {
        cpx a = v[g + j], b = v[g + j + s];
        v[g + j] = cadd(a, b);
        cpx d = csub(a, b);
        const int e = j * (8 / s);  // exponent of W_16
        if (e == 0) {
          v[g + j + s] = d;
        } else if (e == 4) {
          v[g + j + s] = {d.im, -d.re};  // * (-i)
        } else {
          float c, sn;
          tw16(e, c, sn);
          v[g + j + s] = cmul(d, c, sn);
        }
      }
    }
  }
}
__device__ __forceinline__ int bitrev4(int i) { return ((i & 1) << 3) | ((i & 2) << 1) | ((i & 4) >> 1) | ((i & 8) >> 3); }

__device__ __forceinline__ int pcm_phys(int i) { return i + 16 * (i / kHop); }

// Shared memory: PCM span | exchange (one slot per frame in flight; the frame's power spectrum is written back over its
// own slot once the FFT-16 has read it) | output stage | Hann window | W_400 twiddles [k][r] | reduction scratch.
// 71 KB: three blocks per SM (the tables live here instead of 51 registers per thread: 112 -> <= 80 registers).
constexpr int kMelSmemBytes =
    (kSpanPhys + 2 * 16 * kExStride + kMaxMel * kOutStride + kFFT + 2 * 16 * 13 + 8) * (int)sizeof(float);

// grid = (n_tiles, B); block = 256
__global__ void __launch_bounds__(kMelThreads, 3) mel_frames_kernel(MelParams p) {
  WT_DYN_SMEM(smem_raw);
  float* s_pcm = reinterpret_cast<float*>(smem_raw);
  float2* s_ex = reinterpret_cast<float2*>(s_pcm + kSpanPhys);
  float* s_out = reinterpret_cast<float*>(s_ex + 16 * kExStride);
  float* s_hann = s_out + kMaxMel * kOutStride;                   // [400]
  float2* s_tw = reinterpret_cast<float2*>(s_hann + kFFT);        // [13][16]: (cos, -sin)(2 pi r k / 400) at [k][r]
  float* s_red = reinterpret_cast<float*>(s_tw + 16 * 13);        // 8 floats

  const int tid = threadIdx.x;
  const int tile = blockIdx.x;
  const int clip = blockIdx.y;
  const int lane16 = tid & 15;   // residue r in phases A/B, k' in phase C, mel-row phase in D
  const int fslot = tid >> 4;    // frame slot 0..15 within a round
  const int base = tile * kTileFrames * kHop;

  // ---- stage PCM span (zero fill past the clip end: whisper.cpp:149-153) ----
  if (p.pcm_i16 != nullptr) {
    // int16 ingest: 8 samples per 128-bit load; the division is the host's, so the mel is bit-equal to the f32 path
    // on the converted samples
    const short* pcm16 = p.pcm_i16 + (size_t)clip * p.n_samples;
    const bool vec16 = ((p.n_samples & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.pcm_i16) & 15) == 0);
    if (vec16) {
      for (int i8 = tid; i8 < kSpan / 8; i8 += kMelThreads) {
        const int i = i8 * 8;
        alignas(16) short v[8];
        if (base + i + 7 < p.n_samples) {
          *reinterpret_cast<uint4*>(v) = *reinterpret_cast<const uint4*>(pcm16 + base + i);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = (base + i + j < p.n_samples) ? pcm16[base + i + j] : (short)0;
        }
        // 8 consecutive samples never straddle a hop boundary (160 = 8 * 20): one skew for all of them
        float* dst = s_pcm + pcm_phys(i);
        *reinterpret_cast<float4*>(dst) = make_float4((float)v[0] / 32767.0f, (float)v[1] / 32767.0f, (float)v[2] / 32767.0f, (float)v[3] / 32767.0f);
        *reinterpret_cast<float4*>(dst + 4) = make_float4((float)v[4] / 32767.0f, (float)v[5] / 32767.0f, (float)v[6] / 32767.0f, (float)v[7] / 32767.0f);
      }
    } else {
      for (int i = tid; i < kSpan; i += kMelThreads) s_pcm[pcm_phys(i)] = (base + i < p.n_samples) ? (float)pcm16[base + i] / 32767.0f : 0.0f;
    }
  } else {
  const float* pcm = p.pcm + (size_t)clip * p.n_samples;
  const bool vec_ok = ((p.n_samples & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.pcm) & 15) == 0);
  if (vec_ok) {
    for (int i4 = tid; i4 < kSpan / 4; i4 += kMelThreads) {
      const int i = i4 * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (base + i + 3 < p.n_samples) {
        v = *reinterpret_cast<const float4*>(pcm + base + i);
      } else {
        if (base + i + 0 < p.n_samples) v.x = pcm[base + i + 0];
        if (base + i + 1 < p.n_samples) v.y = pcm[base + i + 1];
        if (base + i + 2 < p.n_samples) v.z = pcm[base + i + 2];
      }
      *reinterpret_cast<float4*>(s_pcm + pcm_phys(i)) = v;
    }
  } else {
    for (int i = tid; i < kSpan; i += kMelThreads) s_pcm[pcm_phys(i)] = (base + i < p.n_samples) ? pcm[base + i] : 0.0f;
  }
  }

  // ---- tables: Hann window, W_400^{r k'} transposed to [k'][r] (lanes of a half-warp read consecutive words) ----
  for (int i = tid; i < kFFT; i += kMelThreads) s_hann[i] = __ldg(p.hann + i);
  if (tid < 16 * 13) s_tw[(tid % 13) * 16 + tid / 13] = __ldg(p.tw400 + tid);
  __syncthreads();

  float vmax = -3.0e38f;
  for (int round = 0; round < kTileFrames / 16; ++round) {
    const int fl = round * 16 + fslot;  // frame within tile
    const int frame = tile * kTileFrames + fl;
    // ---- A + B: real DFT-25 of this lane's residue class, twiddle, park ----
    {
      float a[25];
#pragma unroll
      for (int m = 0; m < 25; ++m) a[m] = s_hann[16 * m + lane16] * s_pcm[pcm_phys(fl * kHop + 16 * m + lane16)];
      cpx S[13];
      rdft25(a, S);
      float2* ex = s_ex + fslot * kExStride + lane16 * 13;
      ex[0] = make_float2(S[0].re, S[0].im);
#pragma unroll
      for (int k = 1; k < 13; ++k) {
        const float2 w = s_tw[k * 16 + lane16];
        cpx t = cmul(S[k], w.x, w.y);
        ex[k] = make_float2(t.re, t.im);
      }
    }
    __syncwarp();
    // ---- C: complex FFT-16 over r for k' = lane16 < 13, power, scatter to bins ----
    cpx v[16];
    if (lane16 < 13) {
      const float2* ex = s_ex + fslot * kExStride + lane16;
#pragma unroll
      for (int r = 0; r < 16; ++r) {
        float2 t = ex[r * 13];
        v[r] = {t.x, t.y};
      }
      fft16_dif(v);
    }
    // the power spectrum goes over the frame's own exchange slot: every lane of the half-warp has read its 16 values
    __syncwarp();
    if (lane16 < 13) {
      float* ps = reinterpret_cast<float*>(s_ex + fslot * kExStride);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int q = bitrev4(i);
        const float pw = v[i].re * v[i].re + v[i].im * v[i].im;
        if (q < 8) {
          ps[lane16 + 25 * q] = pw;
        } else if (lane16 > 0) {
          ps[(25 - lane16) + 25 * (15 - q)] = pw;  // |X[400-k]| = |X[k]|
        } else if (q == 8) {
          ps[200] = pw;
        }
      }
    }
    __syncwarp();
    // ---- D: mel rows lane16, lane16+16, ... ; log10(max(.,1e-10)) (whisper.cpp:169-184) ----
    {
      const float* ps = reinterpret_cast<const float*>(s_ex + fslot * kExStride);
      for (int j = lane16; j < p.n_mel; j += 16) {
        const int2 row = __ldg(p.frow + j);
        const int len = __ldg(p.flen + j);
        const float* w = p.fw + row.x;
        const float* pp = ps + row.y;
        float acc = 0.0f;
        for (int i = 0; i < len; ++i) acc = fmaf(pp[i], __ldg(w + i), acc);
        acc = fmaxf(acc, 1e-10f);
        const float v = __log2f(acc) * 0.30102999566398120f;
        s_out[j * kOutStride + fl] = v;
        if (frame < p.n_len) vmax = fmaxf(vmax, v);
      }
    }
    __syncwarp();
  }
  __syncthreads();

  // ---- coalesced store of the [n_mel][32] tile: one 128-byte line per warp-row ----
  {
    const int f = tid & 31;
    const int frame = tile * kTileFrames + f;
    if (frame < p.n_len) {
      float* dst = p.raw + (size_t)clip * p.n_mel * p.n_len + frame;
      for (int j = tid >> 5; j < p.n_mel; j += kMelThreads / 32) dst[(size_t)j * p.n_len] = s_out[j * kOutStride + f];
    }
  }
  // ---- block max -> per-clip atomic max (whisper.cpp:198-203) ----
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
  if ((tid & 31) == 0) s_red[tid >> 5] = vmax;
  __syncthreads();
  if (tid == 0) {
    float m = s_red[0];
    for (int w = 1; w < kMelThreads / 32; ++w) m = fmaxf(m, s_red[w]);
    atomicMax(p.clip_max + clip, float_to_ordered(m));
  }
}

// Clamp + affine (whisper.cpp:205-213): out = (max(raw, max-8) + 4) / 4, f32 [B][n_mel][n_len].
// grid-stride over float4 when aligned.
__global__ void __launch_bounds__(256) mel_finalize_f32_kernel(const float* raw, const int* clip_max, float* out,
                                                                int per_clip, int n_clips) {
  const size_t total = (size_t)per_clip * n_clips;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int clip = (int)(i / per_clip);
    const float floor_v = ordered_to_float(clip_max[clip]) - 8.0f;
    out[i] = (fmaxf(raw[i], floor_v) + 4.0f) * 0.25f;
  }
}

__global__ void fill_int_kernel(int* p, int v, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace wt
