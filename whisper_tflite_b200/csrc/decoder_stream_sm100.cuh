// Streaming persistent greedy decoder for sm_100a (second design of the loop of Decoder::forward,
// whisper.cpp:363-402): one launch, one thread-block cluster per group of <= 16 clips, and ONE producer warp per CTA
// that feeds everything the CTA will ever read in bulk — its slices of the packed projection weights, its range of
// the cross-attention K/V stream, its slice of the tied logits matrix — through a single shared-memory ring of
// cp.async.bulk (TMA) copies, in program order, as far ahead as the ring allows.
//
// What the ncu source page of the first design (round 1's decoder_sm100.cuh, since removed) showed: no hot spot, 3.1 G warp instructions,
// ~2200 per warp and projection phase, almost all of it address arithmetic, predicates and staging around 24 MMAs,
// executed by 2 warps per scheduler with nothing to hide fixed latencies behind (stall_wait 23 %, barrier 21 %,
// long scoreboard 20 %). The phases were instruction- and latency-bound, 10-17 us each. Here:
//   * weights are re-laid at engine creation in the exact order a warp consumes them (pack kernels below), with the
//     LayerNorm gain folded into the weight columns and the LayerNorm bias folded into the projection bias, so a
//     consumer warp's inner loop is LDS.128 (weights, ring) + 2 x LDS.128 (activations) + 2 x mma.sync, rolled;
//   * a warp owns whole 8-column output groups: accumulators never leave registers, no cross-warp reduction, the
//     epilogue (bias, residual, GELU, KV-cache scatter, or the fused (max, last index) of whisper.cpp:346-361) is
//     applied to the mma fragment directly;
//   * the producer is decoupled from the consumers' cluster barrier (an mbarrier in every CTA's shared memory that
//     the 16 CTAs arrive on remotely through DSMEM), so the next phases' weights and the next layer's K/V are in
//     flight while the current phase waits on its peers.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef DS_TRACE_FINE
#define DS_TRACE_FINE 0
#endif
#include "dec_common.cuh"  // shared small helpers (mma wrapper, bf16 unpack, ex2, globaltimer, cluster ids)

namespace wt {

constexpr int kDsConsWarps = 8;
constexpr int kDsConsThreads = kDsConsWarps * 32;
constexpr int kDsThreads = kDsConsThreads + 32;  // + the producer warp
constexpr int kDsRows = 16;                      // clips per cluster = rows of the one m16 MMA tile
constexpr int kDsStageBytes = 32768;             // ring stage: one K/V chunk (128 keys) or <= 8 warps x upw weight units
constexpr int kDsMaxStages = 6;
constexpr int kDsMaxSeg = 24;                     // (clip, head) pairs one CTA may own
constexpr int kDsMaxLayers = 12;
constexpr int kDsUnitBytes = 512;                // one (8-column group, 32-wide k block) weight unit: 32 lanes x 16 B

// weight units per warp and ring stage of a round with `nact` active warps: the largest divisor of the group's k blocks
// that keeps the stage (nact * upw units of 512 B) within kDsStageBytes — as few, as large bulk copies as possible
// (measured: <= 4 bulk copies in flight per SM at ~1.3 us each, whatever their size; tools/probes/bulk_stream.cu)
__host__ __device__ inline int ds_upw(int kbK, int nact) {
  const int cap = (kDsStageBytes / kDsUnitBytes) / nact;
  for (int u = (cap < kbK ? cap : kbK); u > 1; --u)
    if (kbK % u == 0) return u;
  return 1;
}

enum : int { DS_QKV = 0, DS_OUT = 1, DS_XQ = 2, DS_XOUT = 3, DS_FC1 = 4, DS_FC2 = 5, DS_LOGITS = 6, DS_NKIND = 7 };

struct __align__(128) DsLayerDev {
  const uint4* w[6];          // packed weight streams, DS_QKV .. DS_FC2 (whole matrix, CTA-slice-major)
  const float* b[6];          // biases with the LayerNorm bias folded in
  const __nv_bfloat16* kvx;   // cross-attention K/V stream: [B][H][chunk][K rows | V rows]
  __nv_bfloat16 *kc, *vc;     // [B][H][32][64] self-attention cache
};

struct DsParams {
  const DsLayerDev* layers;
  int n_layers, n_clips, clip0, d, H, S, V;
  int group_clips;
  int rank, G;                    // filled in by the kernel
  int upw;                        // weight units per warp and ring stage (6 for d = 384, 8 for d = 512 / 768)
  int n_stages;                   // ring depth (host: what fits next to the activation tiles)
  int debug;
  int kv_ahead;                   // K/V chunks the L2 prefetch runs ahead of the shared-memory copies (0 = off)
  const uint4* w_logits;          // packed tied embedding (final LayerNorm gain folded in)
  const float* b_logits;          // [V] tok_emb . lnf_beta
  const __nv_bfloat16* tok_emb;   // [V][d] (embedding lookups)
  const __nv_bfloat16* pos_emb;
  int* tokens;                    // [B][32]
  int* lens;
  int* done;
  int prompt_len, max_total, eot;
  float* qs;                      // [rows][d] self-attention query (global: read together with the KV cache)
  float* step_logits;
  long long step_logits_stride;
  unsigned long long* trace;
};

// per-CTA description of its stream (shared memory, built once by thread 0)
struct DsStream {
  long long slice_unit0[DS_NKIND];  // first weight unit of this CTA's slice, per kind
  int ng[DS_NKIND];                 // 8-column groups owned by this CTA
  int kbk[DS_NKIND];                // 32-wide k blocks per group (K / 32)
  int upw_full[DS_NKIND];           // ds_upw of a full round (8 active warps) and of the slice's last, partial round
  int upw_last[DS_NKIND];
  int pair0, npairs;                // cross-attention: whole (clip, head) pairs [pair0, pair0 + npairs) of the group
  int pair_base, pair_rem;          // balanced split of the group's pairs over the cluster
  int cpp;                          // K/V chunks per pair
  volatile int stop;                // consumers -> producer: leave after draining
  volatile unsigned consumed;       // ring stages consumed when `stop` was raised
  int lg_tiles, lg_row0, lg_row1;   // logits: 128-row tiles of this CTA's vocabulary slice [lg_row0, lg_row1)
  unsigned tmem_addr;               // 32 tensor-memory columns: two 128 x 16 f32 logits accumulators
  unsigned long long lg_bar[4];     // mbarriers: accumulator full [2] (tcgen05.commit), accumulator drained [2] (4 epilogue warps)
};

// Shared memory. Activations never go through global memory between phases: every CTA holds a full copy of the
// group's residual stream (sX, bf16, LayerNorm input), of the current K = d MMA operand (sA: LayerNorm output, or the
// attention outputs its peers pushed) and of the MLP hidden state (sH), and producers of a slice push it into all
// peers' copies with st.shared::cluster before the cluster barrier.
struct DsSmem {
  static constexpr int kParams = 0;                                   // DsParams (<= 256 B)
  static constexpr int kStream = 256;                                 // DsStream (<= 256 B)
  static constexpr int kBars = 512;                                   // full[6] empty[6] cbar
  static constexpr int kTok = 640;                                    // int[16] tokens of the step, int[16] done flags
  static constexpr int kLayers = 768;                                 // DsLayerDev[12]
  static constexpr int kMisc = kLayers + kDsMaxLayers * 128;          // merge scratch: m[8] l[8] acc[8][64] f32
  static constexpr int kAmax = kMisc + (16 + 8 * 64) * 4;             // u64[16 ranks][16 rows] partial winners
  static constexpr int kXq = kAmax + kDecMaxCluster * kDsRows * 8;    // queries of this CTA's pairs [kDsMaxSeg][64] f32
  static constexpr int kX = kXq + kDsMaxSeg * 64 * 4;                 // sX
  __host__ __device__ static constexpr int x_stride(int d) { return 2 * d; }
  __host__ __device__ static constexpr int a_stride(int d) { return 2 * d + 64; }   // == 64 mod 128
  __host__ __device__ static constexpr int h_stride(int d) { return 8 * d + 64; }   // == 64 mod 128
  __host__ __device__ static constexpr int a_off(int d) { return (kX + kDsRows * x_stride(d) + 127) & ~127; }
  __host__ __device__ static constexpr int h_off(int d) { return a_off(d) + kDsRows * a_stride(d); }
  // logits B operand (ds_logits): [d / 64 blocks][16 clips][128 B], inside the hidden-state tile; the 128-byte swizzle is a
  // function of ADDRESS bits 7..9, so the operand starts on a 1024-byte boundary
  __host__ __device__ static constexpr int lb_off(int d) { return (h_off(d) + 1023) & ~1023; }
  __host__ __device__ static constexpr int ring(int d) { return (h_off(d) + kDsRows * h_stride(d) + 1023) & ~1023; }
  __host__ __device__ static constexpr int total(int d, int n_stages) { return ring(d) + n_stages * kDsStageBytes; }
};

// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, P1;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(20000)   // suspend-time hint (ns): sleep in hardware, not in the issue slots
      : "memory");
  return ok != 0;
}
#ifdef DS_PROBE
constexpr bool kDsProbe = true;   // -DDS_PROBE: WT_DEC_DEBUG=4 / 8 skip the cross-attention / projection math (stream-rate probes)
#else
constexpr bool kDsProbe = false;
#endif
__device__ __forceinline__ void ds_cons_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kDsConsThreads) : "memory"); }
__device__ __forceinline__ void l2_prefetch(const void* gsrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_last_policy() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// address of the same shared-memory location in CTA `cta` of the cluster
__device__ __forceinline__ uint32_t ds_mapa(uint32_t addr, int cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
  return r;
}
__device__ __forceinline__ void ds_st_remote(uint32_t raddr, uint32_t v) {
  asm volatile("st.shared::cluster.b32 [%0], %1;" ::"r"(raddr), "r"(v) : "memory");
}
__device__ __forceinline__ void ds_st_remote(uint32_t raddr, float2 v) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(raddr), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ void ds_st_remote(uint32_t raddr, unsigned long long v) {
  asm volatile("st.shared::cluster.b64 [%0], %1;" ::"r"(raddr), "l"(v) : "memory");
}
// the same value into the same location of every CTA of the cluster (own copy included)
template <class T>
__device__ __forceinline__ void ds_push_all(uint32_t local_addr, T v, int G) {
  // fully unrolled and predicated: with a run-time trip count every mapa -> st pair issues behind the previous one's
  // latency (16 x ~30 clk per value; the projection epilogues were 1 us of that)
#pragma unroll
  for (int c = 0; c < kDecMaxCluster; ++c)
    if (c < G) ds_st_remote(ds_mapa(local_addr, c), v);
}

__device__ __forceinline__ uint64_t* ds_full(uint8_t* smem) { return reinterpret_cast<uint64_t*>(smem + DsSmem::kBars); }
__device__ __forceinline__ uint64_t* ds_empty(uint8_t* smem) { return ds_full(smem) + kDsMaxStages; }
__device__ __forceinline__ uint64_t* ds_cbar(uint8_t* smem) { return ds_full(smem) + 2 * kDsMaxStages; }
__device__ __forceinline__ DsStream* ds_stream(uint8_t* smem) { return reinterpret_cast<DsStream*>(smem + DsSmem::kStream); }
__device__ __forceinline__ int* ds_tok(uint8_t* smem) { return reinterpret_cast<int*>(smem + DsSmem::kTok); }
__device__ __forceinline__ int* ds_done(uint8_t* smem) { return ds_tok(smem) + kDsRows; }
// owner CTA of (clip, head) pair `pr` under the balanced split, and the pair's index inside the owner's range
__device__ __forceinline__ void ds_pair_owner(const DsStream* st, int pr, int& cta, int& idx) {
  const int base = st->pair_base, rem = st->pair_rem, cut = rem * (base + 1);
  if (pr < cut) {
    cta = pr / (base + 1);
    idx = pr - cta * (base + 1);
  } else {
    const int q = (pr - cut) / base;   // base >= 1 whenever pr >= cut exists
    cta = rem + q;
    idx = (pr - cut) - q * base;
  }
}

// Cluster-wide barrier of the CONSUMER warps. Every consumer thread first fences its own pushes / global stores at
// cluster scope; then a CTA-level named barrier, one remote arrive per peer (lane i of warp 0 -> CTA i's mbarrier,
// release at cluster scope), and every consumer thread waits for its own CTA's mbarrier phase (acquire at cluster
// scope). Data peers left in global memory (KV cache, qs) is read afterwards with L1-bypassing loads (__ldcg).
__device__ __forceinline__ void ds_cluster_sync(uint8_t* smem, const DsParams& p, unsigned& cphase, int kind,
                                                unsigned long long& tlast, int G, bool tr) {
  const unsigned long long t_arr = tr ? dec_globaltimer() : 0ull;
#ifndef DS_NO_FENCE
  asm volatile("fence.acq_rel.cluster;" ::: "memory");
#endif
#if DS_TRACE_FINE
  const unsigned long long t_f = tr ? dec_globaltimer() : 0ull;
#endif
  ds_cons_bar();
  const uint32_t local = smem_u32(ds_cbar(smem));
  if (threadIdx.x < (unsigned)G) {
    const uint32_t remote = ds_mapa(local, (int)threadIdx.x);
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
  }
#if DS_TRACE_FINE
  if (tr) {
    const unsigned long long t_b = dec_globaltimer();
    atomicAdd(p.trace + 14, t_f - t_arr);
    atomicAdd(p.trace + 15, t_b - t_f);
  }
#endif
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "DS_CW_LOOP:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DS_CW_DONE;\n\t"
      "bra DS_CW_LOOP;\n\t"
      "DS_CW_DONE:\n\t}" ::"r"(local),
      "r"(cphase)
      : "memory");
  cphase ^= 1u;
  if (tr) {
    const unsigned long long t = dec_globaltimer();
    atomicAdd(p.trace + kind, t - tlast);
    atomicAdd(p.trace + 16 + kind, t - t_arr);
    tlast = t;
  }
}

// ---------------------------------------------------------------------------------------------------
// Weight packing (engine creation). A "unit" is the 512 bytes one warp feeds to two m16n8k16 MMAs: for 8-column
// group g and 32-wide k block kb, lane (gq = lane / 4, t = lane % 4) holds W[g*8 + gq][kb*32 + t*8 .. +8) — the K
// permutation inside a k block is the one the activation fragments use (mma_bf16_16816 in dec_common.cuh). Units are
// ordered the way the stream delivers them: CTA slice (gpc groups) -> round of 8 groups (one per consumer warp) ->
// k phase (upw k blocks) -> warp -> unit, so a ring stage is one contiguous copy and warp w's share of it sits at
// w * upw * 512.
// ---------------------------------------------------------------------------------------------------
// gpw = column groups per warp and round: 1 for the projections; 2 for the logits matrix, whose K = d gives 12 or 16 k
// blocks — with one group per warp a stage is 8 x 6 units = 24 KB, with two it is 16 x 4 = the full 32 KB (the ring
// keeps <= 4 copies in flight whatever their size, so stage bytes are stream rate). Groups i and i + 8 of a round
// belong to warp i & 7; a warp's units of a stage are contiguous.
__host__ __device__ inline int ds_round_prefix(int w, int na) {   // groups of the warps before w in a round of na groups
  const int a = w < na ? w : na, b = na - 8 < w ? na - 8 : w;
  return a + (b > 0 ? b : 0);
}
__global__ void ds_pack_weights_kernel(const __nv_bfloat16* __restrict__ W, const float* __restrict__ gamma,
                                       uint4* __restrict__ out, int N, int K, int G, int gpw) {
  const int n_groups = (N + 7) >> 3, kbK = K >> 5, gpc = (n_groups + G - 1) / G;
  const long long total = (long long)n_groups * kbK * 32;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int lane = (int)(idx & 31);
    const long long gk = idx >> 5;
    const int kb = (int)(gk % kbK), g = (int)(gk / kbK);
    const int r = g / gpc, gl = g - r * gpc;
    const int ng_r = (n_groups - r * gpc < gpc) ? n_groups - r * gpc : gpc;
    const int R = 8 * gpw, rho = gl / R, i = gl - rho * R, w = i & 7, j = i >> 3;
    const int nact = (ng_r - R * rho < R) ? ng_r - R * rho : R;
    const int upw = ds_upw(kbK, nact);
    const int h = kb / upw, u = kb - h * upw;
    const long long unit = (long long)r * gpc * kbK + (long long)rho * R * kbK + (long long)h * nact * upw +
                           (ds_round_prefix(w, nact) + j) * upw + u;
    const int n = g * 8 + (lane >> 2), k0 = kb * 32 + (lane & 3) * 8;
    uint32_t o[4] = {0u, 0u, 0u, 0u};
    if (n < N) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        f[j] = __bfloat162float(W[(size_t)n * K + k0 + j]);
        if (gamma) f[j] *= gamma[k0 + j];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] = dec_pack2(f[2 * j], f[2 * j + 1]);
    }
    out[unit * 32 + lane] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
// b'[n] = b[n] + sum_k W[n][k] * beta[k]   (one warp per output; b and/or beta may be null)
__global__ void ds_fold_bias_kernel(const __nv_bfloat16* __restrict__ W, const float* __restrict__ b,
                                    const float* __restrict__ beta, float* __restrict__ out, int N, int K) {
  const int warp = (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (warp >= N) return;
  float s = 0.0f;
  if (beta)
    for (int k = lane; k < K; k += 32) s = fmaf(__bfloat162float(W[(size_t)warp * K + k]), beta[k], s);
  s = warp_sum(s);
  if (lane == 0) out[warp] = s + (b ? b[warp] : 0.0f);
}

// The tied logits matrix in the layout tcgen05.mma reads from shared memory (ds_logits): per CTA slice of the
// vocabulary (the same 8-row-group split as above), tiles of 128 vocabulary rows; a tile = K / 64 blocks of
// [128 rows][64 k] bf16, rows of 128 bytes, the eight 16-byte chunks of a row XOR-swizzled by (row & 7) — exactly what a
// 128B-swizzled TMA tile load would leave, written once here so that plain bulk copies (two blocks = one 32 KB ring
// stage) deliver MMA-ready operands. LayerNorm gain folded into the columns; rows past the slice / vocabulary are zero.
__host__ __device__ inline int ds_logit_rows_per_cta(int V, int G) { return (((V + 7) / 8 + G - 1) / G) * 8; }
__host__ __device__ inline int ds_logit_tiles_per_cta(int V, int G) { return (ds_logit_rows_per_cta(V, G) + 127) / 128; }
__global__ void ds_pack_logits_kernel(const __nv_bfloat16* __restrict__ W, const float* __restrict__ gamma,
                                      uint4* __restrict__ out, int V, int K, int G) {
  const int rpc = ds_logit_rows_per_cta(V, G), T = ds_logit_tiles_per_cta(V, G), nkb = K >> 6;
  const long long total = (long long)G * T * nkb * 128 * 8;   // 16-byte chunks
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx & 7), row = (int)((idx >> 3) & 127);
    long long rest = idx >> 10;
    const int kb = (int)(rest % nkb);
    rest /= nkb;
    const int t = (int)(rest % T), r = (int)(rest / T);
    const int n = r * rpc + t * 128 + row, k0 = kb * 64 + c * 8;
    uint32_t o[4] = {0u, 0u, 0u, 0u};
    if (n < V && t * 128 + row < rpc) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        f[j] = __bfloat162float(W[(size_t)n * K + k0 + j]);
        if (gamma) f[j] *= gamma[k0 + j];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) o[j] = dec_pack2(f[2 * j], f[2 * j + 1]);
    }
    out[(((long long)(r * T + t) * nkb + kb) * 128 + row) * 8 + (c ^ (row & 7))] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}


// ---------------------------------------------------------------------------------------------------
// LayerNorm (gain/bias folded into the packed weights) of the residual copy sX -> MMA operand tile sA, shared memory to
// shared memory. Half a warp per row, so a warp's two rows reduce side by side (two passes of 4 shuffle steps instead
// of 2 x 2 x 5 in sequence: the phase starts behind this chain); the row is re-read from shared memory in each pass
// instead of being held in registers. d is a multiple of 128 (16 lanes x 8 bf16).
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ float ds_half_sum(float v) {
#pragma unroll
  for (int o = 8; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// kUmmaB: write the operand as the B side of tcgen05.mma instead ([16 clips][64 k] blocks of 2 KB, rows of 128 bytes,
// 16-byte chunks XOR-swizzled by (row & 7)) into the hidden-state tile, which is idle outside fc1 / fc2.
template <bool kUmmaB = false>
__device__ __forceinline__ void ds_stage_norm(uint8_t* smem, int n_clips, int d) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = warp * 2 + (lane >> 4), l16 = lane & 15;    // kDsRows == 2 * kDsConsWarps rows, one pass
  static_assert(kDsRows == 2 * kDsConsWarps, "two rows per consumer warp");
  const bool on = r < n_clips;
  const uint4* src = reinterpret_cast<const uint4*>(smem + DsSmem::kX + (size_t)r * DsSmem::x_stride(d)) + l16;
  uint4* dst = reinterpret_cast<uint4*>(smem + DsSmem::a_off(d) + (size_t)r * DsSmem::a_stride(d)) + l16;
  const int nv = d >> 7;   // 16-byte pieces per lane
  float s = 0.0f;
  if (on)
    for (int i = 0; i < nv; ++i) {
      float f[8];
      bf16x8_to_f32(src[i * 16], f);
      s += ((f[0] + f[1]) + (f[2] + f[3])) + ((f[4] + f[5]) + (f[6] + f[7]));
    }
  const float mean = ds_half_sum(s) / (float)d;
  float q = 0.0f;
  if (on)
    for (int i = 0; i < nv; ++i) {
      float f[8];
      bf16x8_to_f32(src[i * 16], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float c = f[j] - mean;
        q = fmaf(c, c, q);
      }
    }
  const float rstd = rsqrtf(ds_half_sum(q) / (float)d + 1e-5f);
  if (on)
    for (int i = 0; i < nv; ++i) {
      float f[8];
      bf16x8_to_f32(src[i * 16], f);
      uint4* o = dst + i * 16;
      if (kUmmaB) {
        const int pi = i * 16 + l16;   // 16-byte piece of the row: k = 8 pi .. 8 pi + 7 -> block pi / 8, chunk pi % 8
        o = reinterpret_cast<uint4*>(smem + DsSmem::lb_off(d) + (pi >> 3) * 2048 + r * 128 + (((pi & 7) ^ (r & 7)) << 4));
      }
      *o = make_uint4(dec_pack2((f[0] - mean) * rstd, (f[1] - mean) * rstd), dec_pack2((f[2] - mean) * rstd, (f[3] - mean) * rstd),
                               dec_pack2((f[4] - mean) * rstd, (f[5] - mean) * rstd), dec_pack2((f[6] - mean) * rstd, (f[7] - mean) * rstd));
    }
}

// ring position of the consumers (uniform over the consumer threads)
struct DsRing {
  unsigned cons;  // stages consumed so far
  int slot;
  unsigned par;
  int n_stages;   // ring depth (a register copy: DsParams lives in shared memory and every barrier asm clobbers memory)
  unsigned lt;    // logits tiles processed so far in this launch (phase of the accumulator barriers)
  __device__ __forceinline__ void advance() {
    ++cons;
    if (++slot == n_stages) { slot = 0; par ^= 1u; }
  }
};

// One 8-column group per warp and round, accumulated over kph ring stages of upw units (2 MMAs each, two independent
// accumulator chains). Returns with the warp's 16 x 8 tile in acc (rows g / g+8, columns 2t, 2t+1 of the group).
__device__ __forceinline__ void ds_mma_round(uint8_t* smem, const DsParams& p, DsRing& rg, const uint8_t* ring,
                                             const uint8_t* a_lane, int as, int upw, int kph, int nact,
                                             float (&acc)[4]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool active = warp < nact;
  uint64_t* full = ds_full(smem);
  uint64_t* empty = ds_empty(smem);
  float a0[4] = {0.0f, 0.0f, 0.0f, 0.0f}, a1[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  for (int h = 0; h < kph; ++h) {
    mbar_wait(full + rg.slot, rg.par);
    if (active && !(kDsProbe && (p.debug & 8))) {   // WT_DEC_DEBUG=8 (timing probe build): consume the weight stream without the MMAs
      const uint8_t* wp = ring + (size_t)rg.slot * kDsStageBytes + (size_t)(warp * upw) * kDsUnitBytes + lane * 16;
      const uint8_t* ap = a_lane + (size_t)(h * upw) * 64;
#pragma unroll 2
      for (int u = 0; u < upw; ++u) {
        const uint4 w = *reinterpret_cast<const uint4*>(wp + u * kDsUnitBytes);
        const uint4 lo = *reinterpret_cast<const uint4*>(ap + u * 64);
        const uint4 hi = *reinterpret_cast<const uint4*>(ap + u * 64 + 8 * as);
        mma_bf16_16816(a0, lo.x, hi.x, lo.y, hi.y, w.x, w.y);
        mma_bf16_16816(a1, lo.z, hi.z, lo.w, hi.w, w.z, w.w);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + rg.slot);
    rg.advance();
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) acc[j] = a0[j] + a1[j];
}

// ---- projections: (LayerNorm into the operand tile,) rounds of MMA over the ring, epilogue on the fragment.
//      xr = this lane's elements of the residual stream (rows g / g+8 of its columns), kept in registers for the whole
//      launch: the residual projections (OUT, XOUT, FC2) always land on the same columns of the same warps. ----
__device__ __forceinline__ void ds_gemm(uint8_t* smem, const DsParams& p, DsRing& rg, int kind, int l, int pos,
                                        float2 (&xr)[2][2]) {
  const DsLayerDev& L = p.layers[l];
  const DsStream* st = ds_stream(smem);
  const int d = p.d, B = p.n_clips, G = p.G;
  const bool from_h = kind == DS_FC2;
#if DS_TRACE_FINE
  const bool tr = p.trace != nullptr && threadIdx.x == 0;
#else
  constexpr bool tr = false;
#endif
  unsigned long long tm = tr ? dec_globaltimer() : 0ull;
  if (kind == DS_QKV || kind == DS_XQ || kind == DS_FC1) {
    ds_stage_norm(smem, B, d);
    ds_cons_bar();
  }
  if (tr) { const unsigned long long t_ = dec_globaltimer(); atomicAdd(p.trace + 11, t_ - tm); tm = t_; }
  const int as = from_h ? DsSmem::h_stride(d) : DsSmem::a_stride(d);
  const uint8_t* sA = smem + (from_h ? DsSmem::h_off(d) : DsSmem::a_off(d));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int ng = st->ng[kind], kbK = st->kbk[kind];
  const int g0 = (int)(st->slice_unit0[kind] / kbK);  // first group of this CTA's slice
  const int upw_f = st->upw_full[kind], upw_l = st->upw_last[kind], kph_f = kbK / upw_f, kph_l = kbK / upw_l;
  const float* bias = L.b[kind];
  const uint8_t* ring = smem + DsSmem::ring(d);
  const uint8_t* a_lane = sA + (size_t)g * as + t * 16;
  const bool r0 = g < B, r1 = g + 8 < B;
  for (int rho = 0; rho * 8 < ng; ++rho) {
    const bool active = rho * 8 + warp < ng;
    const int n = (g0 + rho * 8 + warp) * 8 + 2 * t;  // this lane's first column
    float2 bz = make_float2(0.0f, 0.0f);
    if (active) bz = __ldg(reinterpret_cast<const float2*>(bias + n));
    float acc[4];
    {
      const int nact = (ng - rho * 8 < 8) ? ng - rho * 8 : 8;
      const int upw = nact == 8 ? upw_f : upw_l;
      ds_mma_round(smem, p, rg, ring, a_lane, as, upw, nact == 8 ? kph_f : kph_l, nact, acc);
    }
    if (tr) { const unsigned long long t_ = dec_globaltimer(); atomicAdd(p.trace + 12, t_ - tm); tm = t_; }
    if (!active) continue;
    const float v00 = acc[0] + bz.x, v01 = acc[1] + bz.y, v10 = acc[2] + bz.x, v11 = acc[3] + bz.y;
    switch (kind) {
      case DS_QKV:
        if (n < d) {
          if (r0) *reinterpret_cast<float2*>(p.qs + (size_t)g * d + n) = make_float2(v00, v01);
          if (r1) *reinterpret_cast<float2*>(p.qs + (size_t)(g + 8) * d + n) = make_float2(v10, v11);
        } else {
          const int nn = n - d, isv = nn >= d, c = isv ? nn - d : nn;
          __nv_bfloat16* cache = isv ? L.vc : L.kc;
          const size_t col = (size_t)pos * 64 + (c & 63), hh = c >> 6;
          if (r0) *reinterpret_cast<uint32_t*>(cache + ((size_t)(p.clip0 + g) * p.H + hh) * 32 * 64 + col) = dec_pack2(v00, v01);
          if (r1) *reinterpret_cast<uint32_t*>(cache + ((size_t)(p.clip0 + g + 8) * p.H + hh) * 32 * 64 + col) = dec_pack2(v10, v11);
        }
        break;
      case DS_XQ: {
        // the query goes (pre-scaled, f32) straight to the CTA that owns the (clip, head) pair's K/V range
        const float qscale = 0.125f * 1.4426950408889634f;
        const int hh = n >> 6, e = n & 63;
        const uint32_t xq = smem_u32(smem + DsSmem::kXq);
        int cta, idx;
        if (r0) {
          ds_pair_owner(st, g * p.H + hh, cta, idx);
          ds_st_remote(ds_mapa(xq + (idx * 64 + e) * 4, cta), make_float2(v00 * qscale, v01 * qscale));
        }
        if (r1) {
          ds_pair_owner(st, (g + 8) * p.H + hh, cta, idx);
          ds_st_remote(ds_mapa(xq + (idx * 64 + e) * 4, cta), make_float2(v10 * qscale, v11 * qscale));
        }
        break;
      }
      case DS_FC1: {
        const uint32_t sh = smem_u32(smem + DsSmem::h_off(d)) + n * 2;
        const int hs = DsSmem::h_stride(d);
        if (r0) ds_push_all(sh + g * hs, dec_pack2(gelu_erf(v00), gelu_erf(v01)), G);
        if (r1) ds_push_all(sh + (g + 8) * hs, dec_pack2(gelu_erf(v10), gelu_erf(v11)), G);
        break;
      }
      default: {  // residual update: registers, and a bf16 copy into every CTA's sX for the next LayerNorm
        const uint32_t sx = smem_u32(smem + DsSmem::kX) + n * 2;
        const int xs = DsSmem::x_stride(d);
        auto upd = [&](float2(&x)[2]) {   // constant register indices on both paths
          x[0].x += v00; x[0].y += v01; x[1].x += v10; x[1].y += v11;
          if (r0) ds_push_all(sx + g * xs, dec_pack2(x[0].x, x[0].y), G);
          if (r1) ds_push_all(sx + (g + 8) * xs, dec_pack2(x[1].x, x[1].y), G);
        };
        if (rho & 1) upd(xr[1]); else upd(xr[0]);
        break;
      }
    }
    if (tr) { const unsigned long long t_ = dec_globaltimer(); atomicAdd(p.trace + 13, t_ - tm); tm = t_; }
  }
}

// x[b] = tok_emb[token(b, pos)] + pos_emb[pos], computed by the owners of the residual columns (the DS_OUT slice)
__device__ __forceinline__ void ds_embed(uint8_t* smem, const DsParams& p, int pos, float2 (&xr)[2][2]) {
  const DsStream* st = ds_stream(smem);
  const int d = p.d, B = p.n_clips, G = p.G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int ng = st->ng[DS_OUT];
  const int g0 = (int)(st->slice_unit0[DS_OUT] / st->kbk[DS_OUT]);
  const int* s_tok = ds_tok(smem);
  const uint32_t sxb = smem_u32(smem + DsSmem::kX);
  const int xs = DsSmem::x_stride(d);
  for (int rho = 0; rho * 8 < ng; ++rho) {
    if (rho * 8 + warp >= ng) continue;
    const int n = (g0 + rho * 8 + warp) * 8 + 2 * t;
    const __nv_bfloat162 e = *reinterpret_cast<const __nv_bfloat162*>(p.pos_emb + (size_t)pos * d + n);
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int r = g + 8 * hh;
      if (r < B) {
        int tok = (pos < p.prompt_len) ? __ldg(p.tokens + (size_t)(p.clip0 + r) * kDecTokStride + pos) : s_tok[r];
        if (tok < 0) tok = 0;
        const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(p.tok_emb + (size_t)tok * d + n);
        const float2 x = make_float2(__bfloat162float(a.x) + __bfloat162float(e.x), __bfloat162float(a.y) + __bfloat162float(e.y));
        if (rho & 1) xr[1][hh] = x; else xr[0][hh] = x;
        ds_push_all(sxb + r * xs + n * 2, dec_pack2(x.x, x.y), G);
      }
    }
  }
}

// ---- final LayerNorm -> tied logits on tcgen05 with a fused (max, last index) per clip (whisper.cpp:346-361's tie
//      rule); this CTA's winners go to every peer's s_amax[rank][row].
//      The one decoder contraction with real width (N = 51 865): D[128 vocabulary rows][16 clips] = W_tile . X^T with the
//      weights as the A operand straight out of the ring (packed MMA-ready by ds_pack_logits_kernel, two 64-wide k blocks
//      per 32 KB stage) and the LayerNorm output as a 16-row B operand: no weight ever passes through a register. Warp 0
//      issues (one elected lane), every consumer warp paces the ring (its arrival frees the stage; warp 0's is the
//      tcgen05.commit), warps 4..7 drain the two accumulators (lane = vocabulary row: 16 clips' logits per thread) and
//      keep the running winners per clip in registers — the phase runs at the rate the stream delivers the matrix. ----
__device__ __noinline__ DsRing ds_logits(uint8_t* smem, const DsParams& p, DsRing rg, int pos) {
  DsStream* st = ds_stream(smem);
  const int d = p.d, B = p.n_clips, V = p.V;
  unsigned long long* s_best = reinterpret_cast<unsigned long long*>(smem + DsSmem::kMisc);  // [16], merge scratch is idle
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid < kDsRows) s_best[tid] = 0ull;
  ds_stage_norm<true>(smem, B, d);
  fence_proxy_async();   // the operand tile was written with ordinary stores; tcgen05.mma reads it through the async proxy
  ds_cons_bar();
  const int T = st->lg_tiles, spt = d >> 7;   // ring stages per 128-row tile (two 64-wide k blocks each)
  uint64_t* full = ds_full(smem);
  uint64_t* empty = ds_empty(smem);
  uint64_t* acc_full = reinterpret_cast<uint64_t*>(st->lg_bar);
  uint64_t* acc_empty = acc_full + 2;
  const uint32_t tmem = st->tmem_addr;
  const uint32_t ring = smem_u32(smem + DsSmem::ring(d)), sb = smem_u32(smem + DsSmem::lb_off(d));
  constexpr uint32_t idesc = make_idesc_bf16(128, 16);
  const int step = pos - (p.prompt_len - 1);
  float* lg = p.step_logits ? p.step_logits + (size_t)step * p.step_logits_stride : nullptr;
  unsigned long long best[kDsRows];
#pragma unroll
  for (int c = 0; c < kDsRows; ++c) best[c] = 0ull;
  // Roles (decoupled, so that nothing but the MMAs' own completion holds a ring stage): warp 0 issues, warp 1 paces the
  // ring for the seven warps that never touch the stage (its arrival counts 7; warp 0's is the tcgen05.commit), warps
  // 4..7 only drain accumulators. Every thread advances its copy of the ring position by the same T * spt stages.
  const bool probe = (p.debug & 16) != 0;   // WT_DEC_DEBUG=16: stream-rate probe (no MMAs; results are wrong)
  if (warp == 0) {
    DsRing r = rg;
    for (int t = 0; t < T; ++t, ++r.lt) {
      const unsigned acc = r.lt & 1u, use = r.lt >> 1;
      if (!probe) {   // the accumulator's previous tile has been drained (a fresh barrier passes the first two uses)
        mbar_wait(acc_empty + acc, (use & 1u) ^ 1u);
        tc_fence_after();
      }
      for (int s_ = 0; s_ < spt; ++s_) {
        mbar_wait(full + r.slot, r.par);
        tc_fence_after();
        if (probe) {
          if (lane == 0) mbar_arrive(empty + r.slot);
        } else if (elect_one()) {
          const uint32_t sa = ring + (uint32_t)r.slot * kDsStageBytes;
#pragma unroll
          for (int kb2 = 0; kb2 < 2; ++kb2) {
            const uint64_t da = make_sw128_desc(sa + kb2 * 16384), db = make_sw128_desc(sb + (s_ * 2 + kb2) * 2048);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem + acc * 16, da + 2 * k, db + 2 * k, idesc, (s_ | kb2 | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty + r.slot);                       // this warp's arrival: the stage is free once the MMAs retire
          if (s_ == spt - 1) umma_commit(acc_full + acc);
        }
        __syncwarp();
        r.advance();
      }
    }
  } else if (warp == 1) {
    DsRing r = rg;
    for (int i = 0; i < T * spt; ++i) {
      mbar_wait(full + r.slot, r.par);   // the stage's current use has begun: the arrival cannot land in an earlier phase
      if (lane == 0)
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(empty + r.slot)), "r"(kDsConsWarps - 1) : "memory");
      r.advance();
    }
  } else if (warp >= 4 && !probe) {
    unsigned lt = rg.lt;
    for (int t = 0; t < T; ++t, ++lt) {
      const unsigned acc = lt & 1u, use = lt >> 1;
      mbar_wait(acc_full + acc, use & 1u);
      tc_fence_after();
      uint32_t v[16];
      tmem_ld16(tmem + ((uint32_t)((warp & 3) * 32) << 16) + acc * 16, v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + acc);
      const int n = st->lg_row0 + t * 128 + (warp & 3) * 32 + lane;
      if (n < st->lg_row1 && n < V) {
        const float bz = __ldg(p.b_logits + n);
#pragma unroll
        for (int c = 0; c < kDsRows; ++c) {
          if (c < B) {
            const float val = __uint_as_float(v[c]) + bz;
            best[c] = u64_max(best[c], pack_val_idx(val, (unsigned)n));
            if (lg) lg[(size_t)(p.clip0 + c) * V + n] = val;
          }
        }
      }
    }
  }
  for (int i = 0; i < T * spt; ++i) rg.advance();
  rg.lt += (unsigned)T;
  if (warp >= 4) {
#pragma unroll
    for (int c = 0; c < kDsRows; ++c) {
      const unsigned long long m = warp_max_u64(best[c]);
      if (lane == 0 && m != 0ull) atomicMax(s_best + c, m);
    }
  }
  ds_cons_bar();
  if (tid < B) ds_push_all(smem_u32(smem + DsSmem::kAmax) + (p.rank * kDsRows + tid) * 8, s_best[tid], p.G);
  return rg;
}

// ---- every CTA reduces the cluster's partial winners itself (identical results everywhere): token of the step, EOT
//      bookkeeping; CTA 0 publishes tokens / lens / done. Returns the number of clips that have produced EOT. ----
__device__ __forceinline__ int ds_finalize(uint8_t* smem, const DsParams& p, int pos) {
  const int tid = threadIdx.x;
  const unsigned long long* s_amax = reinterpret_cast<const unsigned long long*>(smem + DsSmem::kAmax);
  int* s_tok = ds_tok(smem);
  int* s_done = ds_done(smem);
  if (tid < p.n_clips) {
    unsigned long long v = 0ull;
    for (int c = 0; c < p.G; ++c) v = u64_max(v, s_amax[c * kDsRows + tid]);
    const int tok = (int)(v & 0xffffffffull);
    if (!s_done[tid]) {
      s_tok[tid] = tok;
      if (tok == p.eot) s_done[tid] = 1;
      if (p.rank == 0) {
        const int cb = p.clip0 + tid;
        p.tokens[(size_t)cb * kDecTokStride + pos + 1] = tok;
        p.lens[cb] = pos + 2;
        if (tok == p.eot) p.done[cb] = 1;
      }
    }
  }
  ds_cons_bar();
  int fin = 0;
  for (int r = 0; r < p.n_clips; ++r) fin += s_done[r];
  return fin;
}

// ---- self-attention over cache rows 0..pos: one warp per (clip, head), lane = key; the output row segment goes, as
//      bf16, into every CTA's operand tile (it is the A operand of the out-projection) ----
__device__ __noinline__ void ds_self_attn(uint8_t* smem, const DsParams& p, const DsLayerDev& L, int pos) {
  float* s_misc = reinterpret_cast<float*>(smem + DsSmem::kMisc);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.d, H = p.H, B = p.n_clips;
  float* sq = s_misc + 16 + warp * 64;
  const uint32_t sa = smem_u32(smem + DsSmem::a_off(d));
  const int as = DsSmem::a_stride(d);
  for (int pr = p.rank * kDsConsWarps + warp; pr < B * H; pr += p.G * kDsConsWarps) {
    const int b = pr / H, h = pr - b * H;
    const size_t base = ((size_t)(p.clip0 + b) * H + h) * 32 * 64;
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
      const float pj = __shfl_sync(0xffffffffu, pw, j);
      o0 = fmaf(pj, __uint_as_float(vv[j] << 16), o0);
      o1 = fmaf(pj, __uint_as_float(vv[j] & 0xffff0000u), o1);
    }
    const float inv = 1.0f / lsum;
    ds_push_all(sa + b * as + (h * 64 + 2 * lane) * 2, dec_pack2(o0 * inv, o1 * inv), p.G);
    __syncwarp();
  }
}

// ---- cross-attention: this CTA owns whole (clip, head) pairs; their K/V chunks arrive through the ring in order,
//      their queries were pushed into s_q by the query projection's epilogue. 8 lanes share a key (8 dims each), 4 keys
//      per warp and load. The finished 64-wide output goes, as bf16, into every CTA's operand tile. ----
// four 8 x 8 bf16 matrices from shared memory: lanes 8i .. 8i+7 give the row addresses of matrix i
__device__ __forceinline__ void ds_ldsm4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ds_ldsm4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
// m16n8k16 with only row 0 of A alive (one query per (clip, head) pair): rows 8..15 of A are zero and their
// accumulators are dead, so the instruction carries two live accumulator registers instead of four
__device__ __forceinline__ void ds_mma_row0(float& c0, float& c1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  float d2, d3;   // dead accumulators of the zero rows
  (void)d2; (void)d3;
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%5}, {%7,%8}, {%0,%1,%9,%9};"
      : "+f"(c0), "+f"(c1), "=f"(d2), "=f"(d3)
      : "r"(a0), "r"(0u), "r"(a2), "r"(b0), "r"(b1), "f"(0.0f));
}

// One softmax step of a warp over its 16-key slabs of NCH consecutive chunks (ring stages) of a pair: with NCH = 2 the
// two score chains, one max reduction and one rescale cover 32 keys, which halves the dependent latency per chunk —
// the phase is bound by that chain, not by issue slots or the stream (WT_DEC_DEBUG=4 measures the stream alone).
// (NCH = 2 is only a win without spills: 5.17 vs 5.09 ms per launch's cross-attention while the kernel sat at the
// 168-register cap, 3.58 vs 3.70 ms once the LayerNorm staging stopped holding rows in registers.)
constexpr int kDsXaChunksPerStep = 2;
template <int NCH>
__device__ __forceinline__ void ds_xa_step(const DsParams& p, DsRing& rg, uint32_t ring, uint64_t* full, uint64_t* empty,
                                           uint32_t k_off, uint32_t v_off, const uint32_t (&qa)[4][2], int n_keys_last,
                                           float& m, float& lsum, float (&o)[8][2]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  uint32_t vf[NCH][4][4];
  float sc[NCH][2][2];
#pragma unroll
  for (int n = 0; n < NCH; ++n) {
    mbar_wait(full + rg.slot, rg.par);
    const uint32_t stg = ring + (uint32_t)rg.slot * kDsStageBytes;
    if (!kDsProbe || !(p.debug & 4)) {  // WT_DEC_DEBUG=4 (timing probe build): consume the stream without the math
      // K: load h of key tile j = matrices (keys 8j..8j+7) x segments 4h..4h+3 -> b0/b1 of k-steps 2h, 2h+1
      uint32_t kf[2][2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int h = 0; h < 2; ++h) ds_ldsm4(stg + ((k_off + j * 1024u) ^ (uint32_t)(h * 64)), kf[j][h]);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        sc[n][j][0] = sc[n][j][1] = 0.0f;
#pragma unroll
        for (int s = 0; s < 4; ++s)
          ds_mma_row0(sc[n][j][0], sc[n][j][1], qa[s][0], qa[s][1], kf[j][s >> 1][(s & 1) * 2], kf[j][s >> 1][(s & 1) * 2 + 1]);
      }
      // V (transposed on load): load q = matrices (keys lo | hi) x segments 2q, 2q+1 -> b0/b1 of dim tiles 2q, 2q+1.
      // After the score MMAs: the K fragments are dead by then (register pressure), the softmax hides the latency.
#pragma unroll
      for (int q = 0; q < 4; ++q) ds_ldsm4_t(stg + (v_off ^ (uint32_t)(q * 32)), vf[n][q]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + rg.slot);  // values are in registers: the producer may refill the stage
    rg.advance();
  }
  if (kDsProbe && (p.debug & 4)) return;
  const int key0 = warp * 16 + 2 * t;   // sc[.][j][e] belongs to key key0 + 8j + e of its chunk
  if (n_keys_last < kXaChunk) {         // only the last chunk of a pair can be short
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e)
        if (key0 + 8 * j + e >= n_keys_last) sc[NCH - 1][j][e] = -1.0e30f;
  }
  float mx = fmaxf(fmaxf(sc[0][0][0], sc[0][0][1]), fmaxf(sc[0][1][0], sc[0][1][1]));
  if (NCH == 2) mx = fmaxf(mx, fmaxf(fmaxf(sc[NCH - 1][0][0], sc[NCH - 1][0][1]), fmaxf(sc[NCH - 1][1][0], sc[NCH - 1][1][1])));
  // lanes 0..3 hold the query's scores (the other rows of the fragment are zero rows): their quad maximum is the row's
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
  mx = __shfl_sync(0xffffffffu, mx, 0);
  if (mx > m) {  // warp-uniform; rare after the first chunks
    const float scale = dec_ex2(m - mx);
    lsum *= scale;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      o[i][0] *= scale;
      o[i][1] *= scale;
    }
    m = mx;
  }
#pragma unroll
  for (int n = 0; n < NCH; ++n) {
    float pw[2][2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) pw[j][e] = dec_ex2(sc[n][j][e] - m);
    if (n == NCH - 1 && n_keys_last < kXaChunk) {  // a slab that is all padding has m == its masked score: force 0
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e)
          if (key0 + 8 * j + e >= n_keys_last) pw[j][e] = 0.0f;
    }
    lsum += (pw[0][0] + pw[0][1]) + (pw[1][0] + pw[1][1]);
    const uint32_t pa0 = g == 0 ? dec_pack2(pw[0][0], pw[0][1]) : 0u;
    const uint32_t pa2 = g == 0 ? dec_pack2(pw[1][0], pw[1][1]) : 0u;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
      ds_mma_row0(o[nt][0], o[nt][1], pa0, pa2, vf[n][nt >> 1][(nt & 1) * 2], vf[n][nt >> 1][(nt & 1) * 2 + 1]);
  }
}

// Cross-attention over the (clip, head) pairs this CTA owns. A ring stage = one chunk of 128 keys (K rows then V rows,
// 128 B each, 16-byte segments XOR-swizzled by key & 7 in the stream itself — gemm_sm100.cuh's K/V scatter — so that
// ldmatrix reads them without bank conflicts). Warp w takes keys 16w .. 16w+15 of every chunk: scores = 8 mma (2 key
// tiles x 4 k-steps over the 64 dims, A row 0 = the bf16 query), online softmax on the accumulator fragment (lanes
// 0..3 hold the 16 scores), P goes straight back in as the A fragment of 8 more mma against V (ldmatrix.trans). The
// 8 warps' partial (max, sum, O) are merged through shared memory once per pair.
__device__ __noinline__ DsRing ds_cross_attn(uint8_t* smem, const DsParams& p, DsRing rg) {
  static_assert(kXaChunk == kDsConsWarps * 16, "one 16-key slab per consumer warp and chunk");
  float* s_misc = reinterpret_cast<float*>(smem + DsSmem::kMisc);
  const float* s_q = reinterpret_cast<const float*>(smem + DsSmem::kXq);
  const uint32_t ring = smem_u32(smem + DsSmem::ring(p.d));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3, r8 = lane & 7, mi = lane >> 3;
  const int d = p.d, H = p.H, S = p.S;
  const DsStream* st = ds_stream(smem);
  const int cpp = st->cpp;
  uint64_t* full = ds_full(smem);
  uint64_t* empty = ds_empty(smem);
  float* sm_m = s_misc;             // [8]
  float* sm_l = s_misc + 8;         // [8]
  float* sm_acc = s_misc + 16;      // [8][64]
  const uint32_t sa = smem_u32(smem + DsSmem::a_off(d));
  const int as = DsSmem::a_stride(d);
  // this lane's ldmatrix row addresses inside a stage (see the fragment maps above the loads)
  const uint32_t k_off = (uint32_t)(warp * 16 + r8) * 128u + (uint32_t)((mi ^ r8) << 4);
  const uint32_t v_off = (uint32_t)(kXaChunk + warp * 16 + 8 * (mi & 1) + r8) * 128u + (uint32_t)(((mi >> 1) ^ r8) << 4);
  for (int si = 0; si < st->npairs; ++si) {
    const int pr = st->pair0 + si;
    uint32_t qa[4][2];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float2 x = *reinterpret_cast<const float2*>(s_q + si * 64 + 16 * s + 2 * t);
      const float2 y = *reinterpret_cast<const float2*>(s_q + si * 64 + 16 * s + 8 + 2 * t);
      qa[s][0] = g == 0 ? dec_pack2(x.x, x.y) : 0u;
      qa[s][1] = g == 0 ? dec_pack2(y.x, y.y) : 0u;
    }
    float m = -1.0e30f, lsum = 0.0f, o[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = 0.0f;
    const int n_last = S - (cpp - 1) * kXaChunk;   // keys in the pair's last chunk
    int c = 0;
    if (kDsXaChunksPerStep == 2)
      for (; c + 2 <= cpp; c += 2)
        ds_xa_step<2>(p, rg, ring, full, empty, k_off, v_off, qa, c + 2 == cpp ? n_last : kXaChunk, m, lsum, o);
    for (; c < cpp; ++c)
      ds_xa_step<1>(p, rg, ring, full, empty, k_off, v_off, qa, c + 1 == cpp ? n_last : kXaChunk, m, lsum, o);
    lsum += __shfl_xor_sync(0xffffffffu, lsum, 1);
    lsum += __shfl_xor_sync(0xffffffffu, lsum, 2);
    ds_cons_bar();  // s_misc free (previous pair's merge done)
    if (g == 0) {
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
        *reinterpret_cast<float2*>(sm_acc + warp * 64 + 8 * nt + 2 * t) = make_float2(o[nt][0], o[nt][1]);
      if (t == 0) {
        sm_m[warp] = m;
        sm_l[warp] = lsum;
      }
    }
    ds_cons_bar();
    if (tid < 32) {
      float M = sm_m[0];
#pragma unroll
      for (int w = 1; w < kDsConsWarps; ++w) M = fmaxf(M, sm_m[w]);
      float Lw = 0.0f, o0 = 0.0f, o1 = 0.0f;
#pragma unroll
      for (int w = 0; w < kDsConsWarps; ++w) {
        const float s1 = dec_ex2(sm_m[w] - M);
        const float2 a2 = *reinterpret_cast<const float2*>(sm_acc + w * 64 + 2 * tid);
        Lw = fmaf(sm_l[w], s1, Lw);
        o0 = fmaf(a2.x, s1, o0);
        o1 = fmaf(a2.y, s1, o1);
      }
      const float inv = 1.0f / Lw;
      const int b = pr / H, h = pr - b * H;
      ds_push_all(sa + b * as + (h * 64 + 2 * tid) * 2, dec_pack2(o0 * inv, o1 * inv), p.G);
    }
  }
  return rg;
}

// ---------------------------------------------------------------------------------------------------
// Producer (one thread): walks the CTA's whole stream in program order — every source is constant for the launch
// (packed weights, the K/V stream written by the encoder side), so it never depends on the consumers' data, only on
// free ring stages. Returns the number of stages issued.
// ---------------------------------------------------------------------------------------------------
struct DsProd {
  uint8_t* ring;
  uint64_t *full, *empty;
  DsStream* st;
  unsigned n;
  int slot, n_stages;
  unsigned par;
  __device__ __forceinline__ bool issue(const void* src, uint32_t bytes, uint64_t policy) {
    while (!mbar_try_wait(empty + slot, par ^ 1u)) {
      if (st->stop) return false;
    }
    mbar_expect_tx(full + slot, bytes);
    bulk_load_stream(ring + (size_t)slot * kDsStageBytes, src, bytes, full + slot, policy);
    ++n;
    if (++slot == n_stages) { slot = 0; par ^= 1u; }
    return true;
  }
  __device__ __forceinline__ bool weights(const uint4* base, int kind, uint64_t policy, int R = kDsConsWarps) {
    const int ng = st->ng[kind], kbK = st->kbk[kind];
    const uint4* src = base + st->slice_unit0[kind] * 32;
    for (int rho = 0; rho * R < ng; ++rho) {
      const int nact = (ng - rho * R < R) ? ng - rho * R : R;
      const int upw = nact == R ? st->upw_full[kind] : st->upw_last[kind], kph = kbK / upw;
      const uint32_t bytes = (uint32_t)(nact * upw * kDsUnitBytes);
      for (int h = 0; h < kph; ++h) {
        if (!issue(src, bytes, policy)) return false;
        src += bytes >> 4;
      }
    }
    return true;
  }
};

__device__ __noinline__ unsigned ds_producer(uint8_t* smem, const DsParams& p) {
  DsProd pr;
  pr.ring = smem + DsSmem::ring(p.d);
  pr.full = ds_full(smem);
  pr.empty = ds_empty(smem);
  pr.st = ds_stream(smem);
  pr.n = 0; pr.slot = 0; pr.par = 0; pr.n_stages = p.n_stages;
  const uint64_t pol_kv = l2_evict_first_policy(), pol_w = l2_evict_last_policy();
  const int pair0 = pr.st->pair0, npairs = pr.st->npairs, cpp = pr.st->cpp;
  const size_t chunk_elems = (size_t)2 * kXaChunk * 64;
  const int ahead = p.kv_ahead;
  for (int pos = 0; pos <= p.max_total - 2; ++pos) {
    for (int l = 0; l < p.n_layers; ++l) {
      const DsLayerDev& L = p.layers[l];
      if (!pr.weights(L.w[DS_QKV], DS_QKV, pol_w)) return pr.n;
      if (!pr.weights(L.w[DS_OUT], DS_OUT, pol_w)) return pr.n;
      if (!pr.weights(L.w[DS_XQ], DS_XQ, pol_w)) return pr.n;
      {
        // The K/V chunks come from HBM (each is read once per position): an L2 prefetch runs kDsKvAhead chunks ahead of
        // the copies into shared memory, so those find their lines in L2 — the <= 4 bulk copies a CTA keeps in flight
        // then turn over at L2 latency instead of DRAM latency (the stream rate of a CTA is bytes in flight / latency).
        const __nv_bfloat16* src = L.kvx + ((size_t)p.clip0 * p.H + pair0) * cpp * chunk_elems;
        const int n_ch = npairs * cpp;
        if (ahead > 0)
          for (int c = 0; c < ahead && c < n_ch; ++c) l2_prefetch(src + (size_t)c * chunk_elems, (uint32_t)kDsStageBytes);
        for (int c = 0; c < n_ch; ++c, src += chunk_elems) {
          if (ahead > 0 && c + ahead < n_ch) l2_prefetch(src + (size_t)ahead * chunk_elems, (uint32_t)kDsStageBytes);
          if (!pr.issue(src, (uint32_t)kDsStageBytes, pol_kv)) return pr.n;
        }
      }
      if (!pr.weights(L.w[DS_XOUT], DS_XOUT, pol_w)) return pr.n;
      if (!pr.weights(L.w[DS_FC1], DS_FC1, pol_w)) return pr.n;
      if (!pr.weights(L.w[DS_FC2], DS_FC2, pol_w)) return pr.n;
    }
    if (pos >= p.prompt_len - 1) {   // the tied logits matrix: this CTA's tiles, 32 KB (two k blocks) per stage
      const int spt = p.d >> 7;
      const uint8_t* src = reinterpret_cast<const uint8_t*>(p.w_logits) +
                           (size_t)p.rank * ds_logit_tiles_per_cta(p.V, p.G) * spt * kDsStageBytes;
      for (int i = 0; i < pr.st->lg_tiles * spt; ++i)
        if (!pr.issue(src + (size_t)i * kDsStageBytes, (uint32_t)kDsStageBytes, pol_w)) return pr.n;
    }
  }
  return pr.n;
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kDsThreads, 1) decoder_stream_kernel(const DsParams p_in) {
  extern __shared__ __align__(1024) uint8_t ds_smem[];
  const int tid = threadIdx.x, warp = tid >> 5;
  DsParams& p = *reinterpret_cast<DsParams*>(ds_smem + DsSmem::kParams);
  static_assert(sizeof(DsParams) <= 256, "DsParams outgrew its shared-memory slot");
  static_assert(sizeof(DsStream) <= 256, "DsStream outgrew its shared-memory slot");
  static_assert(sizeof(DsLayerDev) == 128, "DsLayerDev slot size");
  const int G = (int)dec_cluster_nctarank();
  const int rank = (int)dec_cluster_ctarank();
  const int group = (int)dec_clusterid_x();
  const int gc = p_in.group_clips;
  const int first = group * gc;
  const int nb = (p_in.n_clips - first < gc) ? p_in.n_clips - first : gc;
  if (nb <= 0 || p_in.n_layers > kDsMaxLayers) return;  // whole cluster leaves together
  if (tid == 0) {
    p = p_in;
    p.G = G;
    p.rank = rank;
    p.clip0 = p_in.clip0 + first;
    p.n_clips = nb;
    p.qs += (size_t)group * kDsRows * p.d;
    if (group != 0 || rank != 0) p.trace = nullptr;
    p.layers = reinterpret_cast<const DsLayerDev*>(ds_smem + DsSmem::kLayers);
    // ---- this CTA's stream: weight slices per kind, cross-attention pairs ----
    DsStream* st = ds_stream(ds_smem);
    const int d = p.d;
    for (int k = 0; k < DS_NKIND; ++k) {
      const int N = (k == DS_QKV) ? 3 * d : (k == DS_FC1) ? 4 * d : (k == DS_LOGITS) ? p.V : d;
      const int kbK = ((k == DS_FC2) ? 4 * d : d) >> 5;
      const int n_groups = (N + 7) >> 3, gpc = (n_groups + G - 1) / G;
      int ng = n_groups - rank * gpc;
      ng = ng < 0 ? 0 : (ng > gpc ? gpc : ng);
      st->ng[k] = ng;
      st->kbk[k] = kbK;
      const int R = k == DS_LOGITS ? 2 * kDsConsWarps : kDsConsWarps;   // groups per round (ds_pack_weights_kernel's gpw)
      st->upw_full[k] = ds_upw(kbK, R);
      st->upw_last[k] = (ng % R) ? ds_upw(kbK, ng % R) : ds_upw(kbK, R);
      st->slice_unit0[k] = (long long)rank * gpc * kbK;
    }
    const int P = nb * p.H, base = P / G, rem = P - base * G;
    st->pair_base = base;
    st->pair_rem = rem;
    st->npairs = base + (rank < rem ? 1 : 0);
    st->pair0 = rank * base + (rank < rem ? rank : rem);
    st->cpp = (p.S + kXaChunk - 1) / kXaChunk;
    {
      const int rpc = ds_logit_rows_per_cta(p.V, G);
      st->lg_row0 = rank * rpc;
      st->lg_row1 = (rank + 1) * rpc < p.V ? (rank + 1) * rpc : p.V;
      const int rows = st->lg_row1 - st->lg_row0;
      st->lg_tiles = rows > 0 ? (rows + 127) / 128 : 0;
      for (int i = 0; i < 4; ++i) mbar_init(reinterpret_cast<uint64_t*>(st->lg_bar) + i, i < 2 ? 1u : 4u);
    }
    st->stop = 0;
    st->consumed = 0u;
    for (int i = 0; i < kDsMaxStages; ++i) {
      mbar_init(ds_full(ds_smem) + i, 1);
      mbar_init(ds_empty(ds_smem) + i, kDsConsWarps);
    }
    mbar_init(ds_cbar(ds_smem), (uint32_t)G);
    fence_barrier_init();
  }
  {
    const int words = (int)(sizeof(DsLayerDev) / 8);
    for (int i = tid; i < p_in.n_layers * words; i += kDsThreads)
      reinterpret_cast<unsigned long long*>(ds_smem + DsSmem::kLayers)[i] =
          reinterpret_cast<const unsigned long long*>(p_in.layers)[i];
    // operand tiles start out zero: rows >= n_clips are never written, and the MMAs read all 16
    const int z0 = DsSmem::kTok, z1 = DsSmem::ring(p_in.d);
    for (int i = z0 + tid * 16; i < z1; i += kDsThreads * 16) {
      if (i >= DsSmem::kLayers && i < DsSmem::kMisc) continue;
      *reinterpret_cast<uint4*>(ds_smem + i) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  if (warp == 0) tmem_alloc(&ds_stream(ds_smem)->tmem_addr, 32);   // the logits accumulators (ds_logits)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // every CTA's barriers and tiles exist before any peer arrives on / pushes into them
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");

  unsigned issued = 0;
  if (warp == kDsConsWarps) {
    if (tid == kDsConsThreads) issued = ds_producer(ds_smem, p);
  } else {
    const int B = p.n_clips;
    const bool tracing = p.trace != nullptr && tid == 0;   // registers: DsParams sits in shared memory (see DsRing)
    unsigned long long tlast = tracing ? dec_globaltimer() : 0ull;
    unsigned cphase = 0;
    DsRing rg;
    rg.cons = 0; rg.slot = 0; rg.par = 0; rg.n_stages = p.n_stages; rg.lt = 0;
    float2 xr[2][2];
#pragma unroll
    for (int i = 0; i < 2; ++i) xr[i][0] = xr[i][1] = make_float2(0.0f, 0.0f);
    if (rank == 0 && tid < B) {
      p.lens[p.clip0 + tid] = p.prompt_len;
      p.done[p.clip0 + tid] = 0;
    }
    for (int pos = 0; pos <= p.max_total - 2; ++pos) {
      ds_embed(ds_smem, p, pos, xr);
      ds_cluster_sync(ds_smem, p, cphase, 10, tlast, G, tracing);
      for (int l = 0; l < p.n_layers; ++l) {
        // phases 1..8: QKV, self-attention, out-proj, cross query, cross-attention, cross out-proj, fc1, fc2
#pragma unroll 1
        for (int ph = 1; ph <= 8; ++ph) {
          if (ph == 2) {
            ds_self_attn(ds_smem, p, p.layers[l], pos);
          } else if (ph == 5) {
            rg = ds_cross_attn(ds_smem, p, rg);
          } else {
            const int kind = ph == 1 ? DS_QKV : ph == 3 ? DS_OUT : ph == 4 ? DS_XQ : ph == 6 ? DS_XOUT : ph == 7 ? DS_FC1 : DS_FC2;
            ds_gemm(ds_smem, p, rg, kind, l, pos, xr);
          }
          ds_cluster_sync(ds_smem, p, cphase, ph, tlast, G, tracing);
        }
      }
      if (pos >= p.prompt_len - 1) {
        rg = ds_logits(ds_smem, p, rg, pos);
        ds_cluster_sync(ds_smem, p, cphase, 9, tlast, G, tracing);
        if (ds_finalize(ds_smem, p, pos) >= B) break;  // every clip has produced EOT (whisper.cpp:397-399, per clip)
      }
    }
    if (tid == 0) {
      DsStream* st = ds_stream(ds_smem);
      st->consumed = rg.cons;
      __threadfence_block();
      st->stop = 1;
    }
  }
  __syncthreads();
  // stages requested but never consumed (early EOT) must land before this CTA's shared memory goes away
  if (tid == kDsConsThreads) {
    const unsigned consumed = ds_stream(ds_smem)->consumed;
    for (unsigned j = consumed; j < issued; ++j) mbar_wait(ds_full(ds_smem) + (j % p.n_stages), (j / p.n_stages) & 1u);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(ds_stream(ds_smem)->tmem_addr, 32);
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

}  // namespace wt
