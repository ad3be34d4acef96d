// Engine behind the C ABI in include/whisper_b200.h: device-resident weights + execution plan for the
// Whisper hot path (log-mel -> encoder -> KV-cached greedy decoder) on one B200.
//
// It takes the place of the reference's `Atom` (FlatBufferModel + InterpreterBuilder + Interpreter,
// whisper.cpp:261-271) and of the three Interpreter::Invoke call sites (whisper.cpp:295, :375, :719).
// There is no CPU fallback: every entry point fails with WT_ERR_CUDA when no sm_100 device is usable.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/whisper_b200.h"
#include "attn_sm100.cuh"
#include "attn2_sm100.cuh"
#include "attn3_sm100.cuh"
#include "decoder_stream_sm100.cuh"
#include "gemm_sm100.cuh"
#include "gemm_lean_sm100.cuh"
#include "mel_host.h"
#include "mel_kernels.cuh"
#include "simt_kernels.cuh"

namespace {

thread_local char g_err[512] = "";
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define WT_CUDA_OK(expr)                                                                              \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess) return fail(WT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                       __FILE__, __LINE__);                                           \
  } while (0)
#define WT_TRY(expr)             \
  do {                           \
    int _s = (expr);             \
    if (_s != WT_OK) return _s;  \
  } while (0)

using bf16 = __nv_bfloat16;
using wt::RowMap;

uint16_t f32_to_bf16_bits(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

// ------------------------------------------------------------------------------------------------
// weight file (<prefix>.b200w, written by whisper_tflite_b200/synth.py::write_weight_file)
// ------------------------------------------------------------------------------------------------
struct HostTensor {
  const float* data;
  std::vector<int64_t> shape;
  size_t numel() const {
    size_t n = 1;
    for (auto d : shape) n *= (size_t)d;
    return n;
  }
};
struct WeightFile {
  void* base = nullptr;
  size_t size = 0;
  wt_dims dims{};
  std::map<std::string, HostTensor> t;
  ~WeightFile() {
    if (base) munmap(base, size);
  }
  int open(const char* path) {
    int fd = ::open(path, O_RDONLY);
    if (fd < 0) return fail(WT_ERR_IO, "cannot open weight file %s", path);
    struct stat st;
    if (fstat(fd, &st) != 0) {
      ::close(fd);
      return fail(WT_ERR_IO, "cannot stat %s", path);
    }
    size = (size_t)st.st_size;
    base = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    ::close(fd);
    if (base == MAP_FAILED) {
      base = nullptr;
      return fail(WT_ERR_IO, "cannot mmap %s", path);
    }
    const uint8_t* p = (const uint8_t*)base;
    if (size < 48 || memcmp(p, "WTB200W\x01", 8) != 0) return fail(WT_ERR_FORMAT, "%s: bad magic", path);
    uint32_t version, n;
    memcpy(&version, p + 8, 4);
    memcpy(&n, p + 12, 4);
    if (version != 1) return fail(WT_ERR_FORMAT, "%s: unsupported version %u", path, version);
    int32_t d[8];
    memcpy(d, p + 16, 32);
    dims = {d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7]};
    size_t off = 48;
    for (uint32_t i = 0; i < n; ++i) {
      if (off + 2 > size) return fail(WT_ERR_FORMAT, "%s: truncated index", path);
      uint16_t ln;
      memcpy(&ln, p + off, 2);
      off += 2;
      if (off + ln + 2 > size) return fail(WT_ERR_FORMAT, "%s: truncated index", path);
      std::string name((const char*)p + off, ln);
      off += ln;
      uint8_t dt = p[off], nd = p[off + 1];
      off += 2;
      if (dt != 0 || nd > 4) return fail(WT_ERR_FORMAT, "%s: tensor %s has unsupported dtype/rank", path, name.c_str());
      HostTensor ht;
      ht.shape.resize(nd);
      if (off + 8 * nd + 16 > size) return fail(WT_ERR_FORMAT, "%s: truncated index", path);
      memcpy(ht.shape.data(), p + off, 8 * nd);
      off += 8 * nd;
      uint64_t o, nb;
      memcpy(&o, p + off, 8);
      memcpy(&nb, p + off + 8, 8);
      off += 16;
      if (o + nb > size || nb != ht.numel() * 4) return fail(WT_ERR_FORMAT, "%s: tensor %s out of bounds", path, name.c_str());
      ht.data = (const float*)(p + o);
      t[name] = ht;
    }
    return WT_OK;
  }
  const HostTensor* find(const std::string& name, std::initializer_list<int64_t> shape) const {
    auto it = t.find(name);
    if (it == t.end()) {
      fail(WT_ERR_FORMAT, "weight file lacks tensor %s", name.c_str());
      return nullptr;
    }
    if (it->second.shape != std::vector<int64_t>(shape)) {
      fail(WT_ERR_FORMAT, "tensor %s has the wrong shape", name.c_str());
      return nullptr;
    }
    return &it->second;
  }
};

// ------------------------------------------------------------------------------------------------
// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda link dependency)
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// bf16 tensor; inner box 64 elements with the 128B swizzle, or 32 elements with the 64B swizzle. dims/strides innermost
// first; strides in bytes for dims 1..rank-1.
int make_map_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box, bool no_swizzle = false) {
  const CUtensorMapSwizzle swz = no_swizzle ? CU_TENSOR_MAP_SWIZZLE_NONE
                                            : (box[0] == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) return fail(WT_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gd[5];
  cuuint64_t gs[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) gs[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(WT_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) rank=%d dims=[%llu,%llu,%llu] stride0=%llu", (int)r,
                rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 1 ? strides_bytes[0] : 0));
  return WT_OK;
}

struct Lin {            // y = x . W^T + b
  void* w = nullptr;    // [N][K] f32 or bf16 depending on precision
  float* b = nullptr;   // [N] f32 (may be null)
  int N = 0, K = 0;
};
struct LNp { float* g = nullptr; float* b = nullptr; };
struct EncLayer { LNp ln1, ln2; Lin qkv, out, fc1, fc2; };
struct DecLayer { LNp ln1, lnx, ln2; Lin qkv, out, xq, xkv, xout, fc1, fc2; void *kx = nullptr, *vx = nullptr, *kc = nullptr, *vc = nullptr, *kvx = nullptr; };

enum Stage {
  ST_MEL = 0, ST_MEL_LAYOUT, ST_CONV, ST_ENC_LN, ST_ENC_GEMM, ST_ENC_ATTN, ST_XKV, ST_DEC_EMBED, ST_DEC_LN, ST_DEC_GEMM,
  ST_DEC_SELF_ATTN, ST_DEC_CROSS_ATTN, ST_LOGITS, ST_ARGMAX, ST_DEC_PERSISTENT, ST_COUNT
};
const char* kStageNames[ST_COUNT] = {"mel_frames", "mel_layout", "conv_stem", "enc_layernorm", "enc_gemm", "enc_attention",
                                     "cross_kv", "dec_embed", "dec_layernorm", "dec_gemm", "dec_self_attn",
                                     "dec_cross_attn", "logits", "argmax", "dec_persistent"};

}  // namespace

// ================================================================================================
struct wt_engine {
  // ---- per-stage CUDA-event profiler (off by default; bench.py's roofline pass turns it on) ----
  bool profiling = false;
  bool decoding = false;       // which phase generic ops (linear / layernorm) are attributed to
  int gemm_stage = -1;         // override for the next linear() (conv stem, cross K/V, logits)
  void *next_kv_k = nullptr, *next_kv_v = nullptr;  // EPI_SPLIT_KV destinations of the next gemm_bf16()
  bool next_kv_chunked = false;
  int stage = ST_MEL;
  struct ProfRec { int stage; cudaEvent_t a, b; };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;
  cudaEvent_t get_event() {
    if (!ev_pool.empty()) { cudaEvent_t e = ev_pool.back(); ev_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
  // Host wait for the engine's stream. cudaStreamSynchronize spins by default: a pool of 8 engines per GPU on 8 GPUs is
  // 64 host threads polling on 32 cores, and the threads that should be enqueuing the next batch are the ones that lose
  // (measured at N = 8 with 8 engines per GPU: e2e 1.59 M audio-s/s polling, 1.70 M sleeping; one GPU: no difference). A
  // blocking event puts the waiting thread to sleep instead.
  // WT_SPIN_SYNC=1 restores the polling wait (lowest wake-up latency for a single call).
  bool defer_pcm_wait = false;   // log_mel: do not wait for the host PCM copy (the caller waits at the end of its pass)
  cudaEvent_t sync_ev = nullptr;
  cudaError_t wait_stream() {
    static const char* spin = getenv("WT_SPIN_SYNC");
    if (spin && spin[0] == '1') return cudaStreamSynchronize(stream);
    if (!sync_ev) {
      cudaError_t e = cudaEventCreateWithFlags(&sync_ev, cudaEventBlockingSync | cudaEventDisableTiming);
      if (e != cudaSuccess) return e;
    }
    cudaError_t e = cudaEventRecord(sync_ev, stream);
    if (e != cudaSuccess) return e;
    return cudaEventSynchronize(sync_ev);
  }
  cudaEvent_t pending_a = nullptr;
  void pre_launch() {
    if (!profiling) return;
    pending_a = get_event();
    cudaEventRecord(pending_a, stream);
  }

  int device = 0;
  int precision = WT_PRECISION_BF16;
  int max_batch = 1;
  int keep_logits = 0;
  bool has_model = false;
  wt_dims dims{};
  int num_sms = 148;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  int64_t launches = 0;
  std::vector<void*> allocs;

  // mel frontend
  bool filters_set = false;
  int n_mel = 80;
  float *d_hann = nullptr, *d_tw400 = nullptr, *d_fw = nullptr;
  int *d_frow = nullptr, *d_flen = nullptr;
  float *d_pcm = nullptr, *d_mel_raw = nullptr, *d_mel = nullptr;
  int* d_clip_max = nullptr;
  int mel_clips = 0, mel_len = 0;  // what d_mel_raw currently holds

  // model
  size_t esz = 2;  // activation/weight element size
  Lin conv1, conv2;
  void* enc_pos = nullptr;      // [1500][d] T
  float* enc_pos_f32 = nullptr; // verify mode addend
  std::vector<EncLayer> enc;
  LNp ln_post, dec_ln;
  void *tok_emb = nullptr, *dec_pos = nullptr;  // T
  std::vector<DecLayer> dec;
  // activations
  void *mel_tm = nullptr, *x1 = nullptr, *x = nullptr, *xn = nullptr, *att = nullptr, *qkv = nullptr, *hbuf = nullptr,
       *enc_out = nullptr;
  void *dx = nullptr, *dxn = nullptr, *datt = nullptr, *dqkv = nullptr, *dq = nullptr, *dh = nullptr;
  float *logits = nullptr, *step_logits = nullptr, *io_f32 = nullptr, *xa_part = nullptr;
  unsigned* xa_tickets = nullptr;
  unsigned long long* argmax_part = nullptr;
  int *tokens = nullptr, *lens = nullptr, *done = nullptr, *d_prompt = nullptr;
  // persistent decoder (bf16, d_model <= 768): one launch, one thread-block cluster per group of clips
  bool dec_persistent = false;
  int dec_cluster = 0, dec_group = 0, dec_groups_cap = 0, dec_active_clusters = 0;
  int dec_group_user = 0;  // wt_set_decoder_group (0 = automatic)
  int dec_cluster_cfg = 0; // wt_config.decoder_cluster (0 = default)
  unsigned long long* d_hash = nullptr;   // wt_debug_hash accumulator
  int* gemm_sched = nullptr;  // {next tile, CTAs done} of the GEMM kernel's dynamic tile scheduler (self-resetting)
  float* pd_qs = nullptr;
  int* pd_finished = nullptr;
  unsigned long long* pd_trace = nullptr;  // WT_DEC_TRACE=1: per-phase ns of the persistent decoder (CTA 0)
  // decoder_stream_sm100.cuh: packed weight streams + one producer warp per CTA
  wt::DsLayerDev* d_ds_layers = nullptr;
  uint4* ds_w_logits = nullptr;
  float* ds_b_logits = nullptr;
  int ds_upw = 0, ds_stages = 0, ds_smem = 0;
  int ds_kv_ahead = 0;   // L2 prefetch distance of the K/V stream in chunks (WT_DEC_KV_AHEAD, read at engine creation)
  int init_stream_weights();
  int enc_clips = 0;  // clips whose encoder states / cross K,V are resident
  std::map<std::tuple<const void*, uint64_t, uint64_t, uint64_t, uint64_t, uint64_t, uint32_t>, CUtensorMap> map_cache;

  ~wt_engine() {
    cudaSetDevice(device);
    if (own_stream) cudaStreamSynchronize(own_stream);
    if (sync_ev) cudaEventDestroy(sync_ev);
    for (void* p : allocs) cudaFree(p);
    if (own_stream) cudaStreamDestroy(own_stream);
  }

  template <class T>
  int alloc(T** out, size_t count, bool zero = false) {
    void* p = nullptr;
    size_t bytes = count * sizeof(T);
    if (bytes == 0) bytes = 16;
    WT_CUDA_OK(cudaMalloc(&p, bytes));
    allocs.push_back(p);
    if (zero) WT_CUDA_OK(cudaMemset(p, 0, bytes));
    *out = (T*)p;
    return WT_OK;
  }
  int alloc_bytes(void** out, size_t bytes, bool zero = false) {
    uint8_t* p;
    WT_TRY(alloc<uint8_t>(&p, bytes, zero));
    *out = p;
    return WT_OK;
  }
  int upload_f32(float** out, const float* host, size_t n) {
    WT_TRY(alloc<float>(out, n));
    WT_CUDA_OK(cudaMemcpy(*out, host, n * 4, cudaMemcpyHostToDevice));
    return WT_OK;
  }
  // uploads in the engine's operand precision
  int upload_T(void** out, const float* host, size_t n) {
    if (precision == WT_PRECISION_FP32_VERIFY) {
      float* p;
      WT_TRY(upload_f32(&p, host, n));
      *out = p;
      return WT_OK;
    }
    std::vector<uint16_t> tmp(n);
    for (size_t i = 0; i < n; ++i) tmp[i] = f32_to_bf16_bits(host[i]);
    uint16_t* p;
    WT_TRY(alloc<uint16_t>(&p, n));
    WT_CUDA_OK(cudaMemcpy(p, tmp.data(), n * 2, cudaMemcpyHostToDevice));
    *out = p;
    return WT_OK;
  }
  int check_launch(const char* what) {
    ++launches;
    if (profiling && pending_a) {
      cudaEvent_t b = get_event();
      cudaEventRecord(b, stream);
      prof.push_back({stage, pending_a, b});
      pending_a = nullptr;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(WT_ERR_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
    return WT_OK;
  }
  bool f32mode() const { return precision == WT_PRECISION_FP32_VERIFY; }

  int init_frontend();
  int init_kernels();
  int load_model(const char* path);
  int set_filters(const float* f, int n_mel_, int n_freq);
  int log_mel(const void* pcm, int n_clips, int n_samples, int pcm_mem, float* out, int out_mem, bool finalize, bool i16 = false);
  int encode(const float* mel, int n_clips, int mel_mem, float* states_out, int out_mem, bool from_raw);
  int decode(int n_clips, const int64_t* prompt, int prompt_len, int max_total, int64_t eot, int32_t* ids, int32_t* lens_o,
             int out_mem);
  int copy_decode_results(int n_clips, int max_total, int32_t* ids, int32_t* lens_o, int out_mem);
  bool persistent_decoder_wanted() const;
  int init_persistent_decoder();
  int decode_persistent(int n_clips, int prompt_len, int max_total, int eot);

  // ---- op wrappers ----
  int gemm_bf16(const void* A, const uint64_t* a_dims, const uint64_t* a_strides, int rows_per_batch, int n_batches,
                const Lin& L, int K_total, int taps, int tap_k, const int* tap_c0, const int* tap_row, void* C,
                RowMap cmap, int flags, const void* addend, RowMap add_map, unsigned long long* amax, int force_bn = 0);
  int linear(const void* A, int M, const Lin& L, void* C, int flags, const void* addend = nullptr,
             unsigned long long* amax = nullptr);
  int layernorm(const void* in, const LNp& ln, void* out, int rows);
  int attention_bf16(const void* qkv_, void* out, int n_clips, int S, int H);
  int get_map(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
              const uint32_t* box);
};

int wt_engine::get_map(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides,
                       const uint32_t* box) {
  auto key = std::make_tuple(base, dims[0], rank > 1 ? dims[1] : 0, rank > 2 ? dims[2] : 0, rank > 1 ? strides[0] : 0,
                             rank > 2 ? strides[1] : 0, (uint32_t)((box[0] << 16) + box[1] * 4 + rank));
  auto it = map_cache.find(key);
  if (it != map_cache.end()) {
    *out = it->second;
    return WT_OK;
  }
  WT_TRY(make_map_bf16(out, base, rank, dims, strides, box));
  map_cache[key] = *out;
  return WT_OK;
}

template <int BN>
static int launch_gemm_bn(wt_engine* e, const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mc, const CUtensorMap& mr, wt::GemmBf16Params& p) {
  using S = wt::GemmSmem<BN>;
  p.n_tiles_n = (p.N + BN - 1) / BN;
  const int n_tiles = p.tiles_per_batch * p.n_batches * p.n_tiles_n;
  const int grid = n_tiles < e->num_sms ? n_tiles : e->num_sms;
  if (!e->gemm_sched) return fail(WT_ERR_STATE, "GEMM tile scheduler not allocated (engine not initialised)");
  p.sched = e->gemm_sched;
  e->pre_launch();
  wt::gemm_bf16_sm100_kernel<BN><<<grid, wt::kGemmThreads, S::kTotal, e->stream>>>(ma, mw, mc, mr, p);
  return e->check_launch("gemm_bf16_sm100_kernel");
}

// sixteen-epilogue-warp kernel (gemm_lean_sm100.cuh): plain bf16 C through TMA, bias / GELU / TMA residual only
template <int BN>
static int launch_gemm_lean_bn(wt_engine* e, const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mc, const CUtensorMap& mr, wt::GemmBf16Params& p) {
  using S = wt::GemmLeanSmem<BN>;
  p.n_tiles_n = (p.N + BN - 1) / BN;
  const int n_tiles = p.tiles_per_batch * p.n_batches * p.n_tiles_n;
  const int grid = n_tiles < e->num_sms ? n_tiles : e->num_sms;
  if (!e->gemm_sched) return fail(WT_ERR_STATE, "GEMM tile scheduler not allocated (engine not initialised)");
  p.sched = e->gemm_sched;
  e->pre_launch();
  wt::gemm_lean_sm100_kernel<BN><<<grid, wt::kLeanThreads, S::total((p.flags & (wt::EPI_TMA_ADD | wt::EPI_SPLIT_KV)) != 0), e->stream>>>(ma, mw, mc, mr, p);
  return e->check_launch("gemm_lean_sm100_kernel");
}

// CTA-pair kernel (gemm2_bf16_sm100_kernel): 256-row tiles on a cluster of two CTAs
template <int BN>
static int launch_gemm2_bn(wt_engine* e, const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mc, const CUtensorMap& mr, wt::GemmBf16Params& p) {
  using S = wt::Gemm2Smem<BN>;
  p.n_tiles_n = (p.N + BN - 1) / BN;
  const int n_tiles = p.tiles_per_batch * p.n_batches * p.n_tiles_n;
  const int max_pairs = e->num_sms / 2;
  const int pairs = n_tiles < max_pairs ? n_tiles : max_pairs;
  if (!e->gemm_sched) return fail(WT_ERR_STATE, "GEMM tile scheduler not allocated (engine not initialised)");
  p.sched = e->gemm_sched;
  e->pre_launch();
  wt::gemm2_bf16_sm100_kernel<BN><<<2 * pairs, wt::kGemmThreads, S::kTotal, e->stream>>>(ma, mw, mc, mr, p);
  return e->check_launch("gemm2_bf16_sm100_kernel");
}
// Where the pair wins (tools/gemm_bench.py, M = 96 000): K >= 768 — fc2 at K = 1536 / 2048 / 3072: 119 -> 111, 193 -> 152,
// 391 -> 341 us (1.33 PF/s = 0.97 of the sustained cuBLAS figure), QKV at K = 768: 298 -> 285. At K <= 512 a tile is
// <= 8 k-blocks and the epilogue, not the operand traffic, paces the kernel: the one-CTA kernel is 3-8 % faster there.
static bool gemm_pair_wanted(int rows_per_batch, int n_batches, int K_total) {
  static const char* ov = getenv("WT_GEMM_PAIR");   // "0" = never, "1" = whenever the shape allows (A/B measurements)
  if (ov && ov[0] == '0') return false;
  const bool fits = (long long)rows_per_batch * n_batches >= 4096 && rows_per_batch >= 256;
  return fits && ((ov && ov[0] == '1') || K_total >= 768);
}

static int pick_bn(int M, int N) {
  if (M <= 128) return N >= 16384 ? 128 : 64;
  const int cands[3] = {256, 192, 128};
  for (int bn : cands)
    if (N % bn == 0) return bn;
  return 128;
}

int wt_engine::gemm_bf16(const void* A, const uint64_t* a_dims, const uint64_t* a_strides, int rows_per_batch,
                         int n_batches, const Lin& L, int K_total, int taps, int tap_k, const int* tap_c0,
                         const int* tap_row, void* C, RowMap cmap, int flags, const void* addend, RowMap add_map,
                         unsigned long long* amax, int force_bn) {
  const int M = rows_per_batch * n_batches;
  const int bn = force_bn ? force_bn : pick_bn(M, L.N);
  const bool pair = !force_bn && bn >= 128 && gemm_pair_wanted(rows_per_batch, n_batches, K_total);
  CUtensorMap ma, mw;
  const uint32_t abox[3] = {64, 128, 1};
  WT_TRY(get_map(&ma, A, 3, a_dims, a_strides, abox));
  const uint64_t wdims[2] = {(uint64_t)K_total, (uint64_t)L.N};
  const uint64_t wstr[1] = {(uint64_t)K_total * 2};
  const uint32_t wbox[2] = {64, (uint32_t)(pair ? bn / 2 : bn)};   // a CTA of a pair loads its half of the W tile
  WT_TRY(get_map(&mw, L.w, 2, wdims, wstr, wbox));
  wt::GemmBf16Params p{};
  p.bias = L.b;
  p.addend = (const bf16*)addend;
  p.add_map = add_map;
  p.C = C;
  p.cmap = cmap;
  p.argmax_part = amax;
  if (flags & wt::EPI_SPLIT_KV) {
    p.kv_k = (bf16*)next_kv_k;
    p.kv_v = (bf16*)next_kv_v;
    p.kv_S = dims.n_audio_ctx;
    p.kv_H = dims.n_heads;
    p.kv_chunk = next_kv_chunked ? wt::kXaChunk : 0;
    p.kv_cpp = (dims.n_audio_ctx + wt::kXaChunk - 1) / wt::kXaChunk;
  }
  p.M = M;
  p.N = L.N;
  p.K = K_total;
  p.rows_per_batch = rows_per_batch;
  p.tiles_per_batch = (rows_per_batch + 127) / 128;
  p.n_batches = n_batches;
  p.flags = flags | (L.b ? wt::EPI_BIAS : 0);
  if (!L.b) p.flags &= ~wt::EPI_BIAS;
  p.taps = taps;
  p.tap_k = tap_k;
  const int per_tap_k = taps == 1 ? K_total : tap_k;
  p.kblocks_per_tap = (per_tap_k + 63) / 64;
  for (int i = 0; i < 3; ++i) {
    p.tap_c0[i] = (i < taps && tap_c0) ? tap_c0[i] : 0;
    p.tap_row[i] = (i < taps && tap_row) ? tap_row[i] : 0;
  }
  // C through TMA (EPI_TMA_C): a bf16 C whose RowMap is a 3-D tensor {N, rows_per_batch, batches} with 16-byte aligned
  // base and strides. Rows past a clip's end and columns past N are clipped by the map's extents. With an addend that
  // has such a map as well (EPI_TMA_ADD: the residual GEMMs) both directions go through 32 x 32 boxes, the addend
  // prefetched by the copy engine a chunk ahead and updated in place in shared memory.
  CUtensorMap mc = ma, mr = ma;
  bool lean = false;
  {
    // The decoder's K/V stream through TMA (sixteen-epilogue-warp kernel): [pair][chunk][K | V][128 keys][64] as a 5-D tensor
    // WITHOUT swizzle — the warps write their box with the 16-byte pieces XOR-ed by (key & 7), which is exactly the
    // stream's layout (decoder_stream_sm100.cuh reads it with ldmatrix) once the copy engine moves the box verbatim
    static const char* ovk = getenv("WT_GEMM_LEAN");
    if ((flags & wt::EPI_SPLIT_KV) && next_kv_chunked && rows_per_batch == dims.n_audio_ctx && !pair && !force_bn &&
        (bn == 256 || bn == 128) && L.N == 2 * dims.n_heads * 64 && L.N % bn == 0 && !(ovk && ovk[0] == '0')) {
      const uint64_t cpp = (uint64_t)p.kv_cpp;
      const uint64_t kd[5] = {64, 128, 2, cpp, (uint64_t)n_batches * dims.n_heads};
      const uint64_t ks[4] = {128, 128 * 128, 2 * 128 * 128, cpp * 2 * 128 * 128};
      const uint32_t kbox[5] = {64, 32, 1, 1, 1};
      WT_TRY(make_map_bf16(&mc, next_kv_k, 5, kd, ks, kbox, true));
      lean = true;
    }
  }
  if (!lean) {
    static const char* ov = getenv("WT_GEMM_TMA_C");   // "0": the epilogue warps load / store themselves; "1": no TMA addend
    auto as_tensor = [&](const void* base, const RowMap& m, long long& bstride) {
      bstride = -1;
      if (m.rows_per_batch == rows_per_batch) bstride = n_batches > 1 ? m.batch_stride : (long long)rows_per_batch * m.row_stride;
      else if (m.rows_per_batch >= M) bstride = (long long)rows_per_batch * m.row_stride;
      return base != nullptr && (reinterpret_cast<uintptr_t>(base) & 15) == 0 && m.row_stride > 0 && (m.row_stride * 2) % 16 == 0 &&
             bstride > 0 && (bstride * 2) % 16 == 0;
    };
    const bool plain_out = !(flags & (wt::EPI_OUT_F32 | wt::EPI_ARGMAX | wt::EPI_SPLIT_KV));
    long long cb = 0, rb = 0;
    if (plain_out && as_tensor(C, cmap, cb) && L.N >= 64 && bn >= 64 && !(ov && ov[0] == '0')) {
      const bool tma_add = (flags & wt::EPI_ADD) && as_tensor(addend, add_map, rb) && !(ov && ov[0] == '1');
      // the sixteen-epilogue-warp kernel: everything it needs goes through TMA (32-column boxes in both directions)
      static const char* ovl = getenv("WT_GEMM_LEAN");   // "0": never (A/B measurements)
      lean = !pair && !force_bn && bn >= 128 && L.N % bn == 0 && M >= 1024 && (tma_add || !(flags & wt::EPI_ADD)) &&
             !(ovl && ovl[0] == '0');
      const uint64_t cdims[3] = {(uint64_t)L.N, (uint64_t)rows_per_batch, (uint64_t)n_batches};
      const uint64_t cstr[2] = {(uint64_t)cmap.row_stride * 2, (uint64_t)cb * 2};
      const uint32_t cbox[3] = {(tma_add || lean) ? 32u : 64u, 32, 1};
      WT_TRY(get_map(&mc, C, 3, cdims, cstr, cbox));
      p.flags |= wt::EPI_TMA_C;
      if (tma_add) {
        const uint64_t rstr[2] = {(uint64_t)add_map.row_stride * 2, (uint64_t)rb * 2};
        WT_TRY(get_map(&mr, addend, 3, cdims, rstr, cbox));
        p.flags |= wt::EPI_TMA_ADD;
      }
    }
  }
  if (lean) {
    switch (bn) {
      case 128: return launch_gemm_lean_bn<128>(this, ma, mw, mc, mr, p);
      case 192: return launch_gemm_lean_bn<192>(this, ma, mw, mc, mr, p);
      case 256: return launch_gemm_lean_bn<256>(this, ma, mw, mc, mr, p);
    }
  }
  if (pair) {
    p.tiles_per_batch = (rows_per_batch + 255) / 256;
    switch (bn) {
      case 128: return launch_gemm2_bn<128>(this, ma, mw, mc, mr, p);
      case 192: return launch_gemm2_bn<192>(this, ma, mw, mc, mr, p);
      case 256: return launch_gemm2_bn<256>(this, ma, mw, mc, mr, p);
    }
  }
  switch (bn) {
    case 64: return launch_gemm_bn<64>(this, ma, mw, mc, mr, p);
    case 128: return launch_gemm_bn<128>(this, ma, mw, mc, mr, p);
    case 192: return launch_gemm_bn<192>(this, ma, mw, mc, mr, p);
    case 256: return launch_gemm_bn<256>(this, ma, mw, mc, mr, p);
  }
  return fail(WT_ERR_INVALID, "bad BN %d", bn);
}

// y[M][N] = epi(x[M][K] . W^T) on plain row-major operands, either precision. addend = residual [M][N] (may alias C).
int wt_engine::linear(const void* A, int M, const Lin& L, void* C, int flags, const void* addend,
                      unsigned long long* amax) {
  RowMap plain{0, M, L.N};
  stage = gemm_stage >= 0 ? gemm_stage : (decoding ? ST_DEC_GEMM : ST_ENC_GEMM);
  if (f32mode()) {
    wt::GemmF32Params p{};
    p.A = (const float*)A;
    p.amap = RowMap{0, M, L.K};
    p.W = (const float*)L.w;
    p.bias = L.b;
    p.residual = (const float*)addend;
    p.rmap = plain;
    p.pos = nullptr;
    p.pos_rows = 1;
    p.C = (float*)C;
    p.cmap = plain;
    p.M = M;
    p.N = L.N;
    p.K = L.K;
    p.flags = (flags & (wt::EPI_GELU)) | (L.b ? wt::EPI_BIAS : 0) | (addend ? wt::EPI_RESIDUAL : 0);
    dim3 grid((L.N + 63) / 64, (M + 63) / 64);
    pre_launch();
    wt::sgemm_f32_kernel<<<grid, 256, 0, stream>>>(p);
    return check_launch("sgemm_f32_kernel");
  }
  const uint64_t ad[3] = {(uint64_t)L.K, (uint64_t)M, 1};
  const uint64_t as[2] = {(uint64_t)L.K * 2, (uint64_t)L.K * 2 * (uint64_t)M};
  return gemm_bf16(A, ad, as, M, 1, L, L.K, 1, 0, nullptr, nullptr, C, plain, flags | (addend ? wt::EPI_ADD : 0), addend,
                   plain, amax);
}

// softmax(Q K^T / 8) V per (clip, head) on qkv [n_clips][S][3*H*64] bf16 -> out [n_clips][S][H*64] bf16
int wt_engine::attention_bf16(const void* qkv_, void* out, int n_clips, int S, int H) {
  const int d = H * 64;
  CUtensorMap mq;
  const uint64_t qd[3] = {(uint64_t)3 * d, (uint64_t)S, (uint64_t)n_clips};
  const uint64_t qs[2] = {(uint64_t)3 * d * 2, (uint64_t)S * 3 * d * 2};
  const uint32_t qbox[3] = {64, 128, 1};
  WT_TRY(get_map(&mq, qkv_, 3, qd, qs, qbox));
  stage = ST_ENC_ATTN;
  pre_launch();
  if (!gemm_sched) return fail(WT_ERR_STATE, "tile scheduler counters not allocated (engine not initialised)");
  const int n_items = n_clips * H * ((S + 255) / 256);
  wt::attention3_sm100_kernel<<<n_items < num_sms ? n_items : num_sms, wt::kAt3Threads, wt::Attn3Smem::kTotal, stream>>>(
      mq, (bf16*)out, S, H, n_clips, gemm_sched);
  return check_launch("attention3_sm100_kernel");
}

int wt_engine::layernorm(const void* in, const LNp& ln, void* out, int rows) {
  const int D = dims.d_model;
  stage = decoding ? ST_DEC_LN : ST_ENC_LN;
  dim3 grid((rows + 7) / 8);
  pre_launch();
  if (f32mode())
    wt::layernorm_kernel<float, float><<<grid, 256, 0, stream>>>((const float*)in, ln.g, ln.b, (float*)out, rows, D);
  else if (D % 128 == 0 && D <= 1024) {
    // one warp per two rows, 8 warps per block
    dim3 g2((rows + 15) / 16);
    const bf16* xi = (const bf16*)in;
    bf16* yo = (bf16*)out;
    switch (D / 128) {
      case 3: wt::layernorm_bf16_vec_kernel<3><<<g2, 256, 0, stream>>>(xi, ln.g, ln.b, yo, rows); break;
      case 4: wt::layernorm_bf16_vec_kernel<4><<<g2, 256, 0, stream>>>(xi, ln.g, ln.b, yo, rows); break;
      case 6: wt::layernorm_bf16_vec_kernel<6><<<g2, 256, 0, stream>>>(xi, ln.g, ln.b, yo, rows); break;
      case 8: wt::layernorm_bf16_vec_kernel<8><<<g2, 256, 0, stream>>>(xi, ln.g, ln.b, yo, rows); break;
      default: wt::layernorm_kernel<bf16, bf16><<<grid, 256, 0, stream>>>(xi, ln.g, ln.b, yo, rows, D); break;
    }
  } else
    wt::layernorm_kernel<bf16, bf16><<<grid, 256, 0, stream>>>((const bf16*)in, ln.g, ln.b, (bf16*)out, rows, D);
  return check_launch("layernorm_kernel");
}

int wt_engine::init_frontend() {
  const size_t B = (size_t)max_batch;
  WT_TRY(alloc(&d_pcm, B * WT_N_SAMPLES));
  WT_TRY(alloc(&d_mel_raw, B * wt::kMaxMel * WT_N_FRAMES));
  WT_TRY(alloc(&d_mel, B * wt::kMaxMel * WT_N_FRAMES));
  WT_TRY(alloc(&d_clip_max, B));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::mel_frames_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::kMelSmemBytes));
  return init_kernels();
}

// Per-device kernel attributes and the GEMM tile scheduler's counters: set at engine creation (every engine sets them
// again: the call is idempotent and thread-safe, unlike a process-wide "already set" flag), never on the hot path.
int wt_engine::init_kernels() {
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<64>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<128>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<192>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<256>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_lean_sm100_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (wt::GemmLeanSmem<128>::total(false) > wt::GemmLeanSmem<128>::total(true) ? wt::GemmLeanSmem<128>::total(false) : wt::GemmLeanSmem<128>::total(true))));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_lean_sm100_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, (wt::GemmLeanSmem<192>::total(false) > wt::GemmLeanSmem<192>::total(true) ? wt::GemmLeanSmem<192>::total(false) : wt::GemmLeanSmem<192>::total(true))));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_lean_sm100_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (wt::GemmLeanSmem<256>::total(false) > wt::GemmLeanSmem<256>::total(true) ? wt::GemmLeanSmem<256>::total(false) : wt::GemmLeanSmem<256>::total(true))));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm2_bf16_sm100_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Gemm2Smem<128>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm2_bf16_sm100_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Gemm2Smem<192>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm2_bf16_sm100_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Gemm2Smem<256>::kTotal));
  WT_CUDA_OK(cudaFuncSetAttribute(wt::attention3_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::Attn3Smem::kTotal));
  WT_TRY(alloc(&gemm_sched, 2, true));   // cudaMemset here; wt_engine_create synchronises the device before returning
  return WT_OK;
}

int wt_engine::set_filters(const float* f, int n_mel_, int n_freq) {
  if (!f || n_freq != WT_N_FREQ || n_mel_ < 1 || n_mel_ > wt::kMaxMel)
    return fail(WT_ERR_INVALID, "filters must be [n_mel<=%d][201]", wt::kMaxMel);
  if (has_model && n_mel_ != dims.n_mels) return fail(WT_ERR_INVALID, "model expects %d mel bins", dims.n_mels);
  wt::MelHostTables t = wt::build_mel_tables(f, n_mel_, n_freq);
  n_mel = n_mel_;
  WT_TRY(upload_f32(&d_hann, t.hann.data(), t.hann.size()));
  WT_TRY(upload_f32(&d_tw400, t.tw400.data(), t.tw400.size()));
  if (t.fw.empty()) t.fw.push_back(0.0f);
  WT_TRY(upload_f32(&d_fw, t.fw.data(), t.fw.size()));
  WT_TRY(alloc(&d_frow, t.frow.size()));
  WT_CUDA_OK(cudaMemcpy(d_frow, t.frow.data(), t.frow.size() * 4, cudaMemcpyHostToDevice));
  WT_TRY(alloc(&d_flen, t.flen.size()));
  WT_CUDA_OK(cudaMemcpy(d_flen, t.flen.data(), t.flen.size() * 4, cudaMemcpyHostToDevice));
  filters_set = true;
  return WT_OK;
}

// pcm -> raw log10 mel + per-clip max (resident); optionally the reference-layout normalised f32 mel.
int wt_engine::log_mel(const void* pcm, int n_clips, int n_samples, int pcm_mem, float* out, int out_mem,
                       bool finalize, bool i16) {
  if (!filters_set) return fail(WT_ERR_STATE, "wt_set_mel_filters has not been called");
  if (n_clips < 1 || n_clips > max_batch) return fail(WT_ERR_INVALID, "n_clips %d outside [1, max_batch=%d]", n_clips, max_batch);
  if (n_samples < wt::kHop || n_samples > WT_N_SAMPLES) return fail(WT_ERR_INVALID, "n_samples %d outside [160, 480000]", n_samples);
  if (!pcm) return fail(WT_ERR_INVALID, "pcm is null");
  const void* src = pcm;
  if (pcm_mem == WT_HOST) {   // int16 samples take half the staging buffer (and half the PCIe traffic)
    WT_CUDA_OK(cudaMemcpyAsync(d_pcm, pcm, (size_t)n_clips * n_samples * (i16 ? 2 : 4), cudaMemcpyHostToDevice, stream));
    src = d_pcm;
  }
  const int n_len = n_samples / wt::kHop;
  pre_launch();
  stage = ST_MEL;
  wt::fill_int_kernel<<<(n_clips + 255) / 256, 256, 0, stream>>>(d_clip_max, INT_MIN, n_clips);
  WT_TRY(check_launch("fill_int_kernel"));
  wt::MelParams p{};
  p.pcm = i16 ? nullptr : (const float*)src;
  p.pcm_i16 = i16 ? (const short*)src : nullptr;
  p.raw = d_mel_raw;
  p.clip_max = d_clip_max;
  p.hann = d_hann;
  p.tw400 = (const float2*)d_tw400;
  p.fw = d_fw;
  p.frow = (const int2*)d_frow;
  p.flen = d_flen;
  p.n_samples = n_samples;
  p.n_len = n_len;
  p.n_mel = n_mel;
  p.n_tiles = (n_len + wt::kTileFrames - 1) / wt::kTileFrames;
  pre_launch();
  wt::mel_frames_kernel<<<dim3(p.n_tiles, n_clips), wt::kMelThreads, wt::kMelSmemBytes, stream>>>(p);
  WT_TRY(check_launch("mel_frames_kernel"));
  mel_clips = n_clips;
  mel_len = n_len;
  if (finalize) {
    pre_launch();
    wt::mel_finalize_f32_kernel<<<num_sms * 4, 256, 0, stream>>>(d_mel_raw, d_clip_max, d_mel, n_mel * n_len, n_clips);
    WT_TRY(check_launch("mel_finalize_f32_kernel"));
    if (out)
      WT_CUDA_OK(cudaMemcpyAsync(out, d_mel, (size_t)n_clips * n_mel * n_len * 4,
                                 out_mem == WT_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, stream));
  }
  // the caller's host buffers are free again when the call returns — unless the caller (wt_transcribe_tokens) ends
  // with a wait of its own and wants the whole pass enqueued first
  if ((pcm_mem == WT_HOST && !defer_pcm_wait) || (out && out_mem == WT_HOST)) WT_CUDA_OK(wait_stream());
  return WT_OK;
}

int wt_engine::load_model(const char* path) {
  WeightFile wf;
  WT_TRY(wf.open(path));
  dims = wf.dims;
  const int d = dims.d_model, H = dims.n_heads, V = dims.n_vocab, S = dims.n_audio_ctx;
  if (d % 64 != 0 || H * 64 != d || d > 1024) return fail(WT_ERR_FORMAT, "unsupported dims: d_model=%d heads=%d (head_dim must be 64)", d, H);
  if (dims.n_mels != 80 || S != 1500) return fail(WT_ERR_FORMAT, "unsupported n_mels=%d n_audio_ctx=%d", dims.n_mels, S);
  if (dims.n_text_ctx < 32) return fail(WT_ERR_FORMAT, "n_text_ctx too small");
  esz = f32mode() ? 4 : 2;
  const size_t B = (size_t)max_batch;

  auto up_lin = [&](Lin& L, const std::vector<float>& w, const std::vector<float>* b, int N, int K) -> int {
    L.N = N;
    L.K = K;
    WT_TRY(upload_T(&L.w, w.data(), w.size()));
    if (b) WT_TRY(upload_f32(&L.b, b->data(), b->size()));
    return WT_OK;
  };
  auto get = [&](const std::string& n, std::initializer_list<int64_t> shp) { return wf.find(n, shp); };
  auto vec = [](const HostTensor* t) { return std::vector<float>(t->data, t->data + t->numel()); };
  auto up_ln = [&](LNp& ln, const std::string& pfx) -> int {
    const HostTensor *g = get(pfx + ".weight", {d}), *b = get(pfx + ".bias", {d});
    if (!g || !b) return WT_ERR_FORMAT;
    WT_TRY(upload_f32(&ln.g, g->data, d));
    WT_TRY(upload_f32(&ln.b, b->data, d));
    return WT_OK;
  };
  // conv weights [n][c][3] -> [n][tap][c]
  auto up_conv = [&](Lin& L, const std::string& pfx, int cin) -> int {
    const HostTensor *w = get(pfx + ".weight", {d, cin, 3}), *b = get(pfx + ".bias", {d});
    if (!w || !b) return WT_ERR_FORMAT;
    std::vector<float> p((size_t)d * 3 * cin);
    for (int n = 0; n < d; ++n)
      for (int c = 0; c < cin; ++c)
        for (int k = 0; k < 3; ++k) p[((size_t)n * 3 + k) * cin + c] = w->data[((size_t)n * cin + c) * 3 + k];
    std::vector<float> bb = vec(b);
    return up_lin(L, p, &bb, d, 3 * cin);
  };
  auto up_plain = [&](Lin& L, const std::string& pfx, int N, int K, bool bias) -> int {
    const HostTensor* w = get(pfx + ".weight", {N, K});
    if (!w) return WT_ERR_FORMAT;
    std::vector<float> ww = vec(w);
    if (bias) {
      const HostTensor* b = get(pfx + ".bias", {N});
      if (!b) return WT_ERR_FORMAT;
      std::vector<float> bb = vec(b);
      return up_lin(L, ww, &bb, N, K);
    }
    return up_lin(L, ww, nullptr, N, K);
  };
  // fused q|k|v (k has no bias -> zeros) or k|v
  auto up_fused = [&](Lin& L, const std::string& pfx, bool with_q) -> int {
    const HostTensor *qw = get(pfx + ".query.weight", {d, d}), *qb = get(pfx + ".query.bias", {d});
    const HostTensor *kw = get(pfx + ".key.weight", {d, d});
    const HostTensor *vw = get(pfx + ".value.weight", {d, d}), *vb = get(pfx + ".value.bias", {d});
    if (!qw || !qb || !kw || !vw || !vb) return WT_ERR_FORMAT;
    std::vector<float> w, b;
    if (with_q) {
      w.insert(w.end(), qw->data, qw->data + (size_t)d * d);
      b.insert(b.end(), qb->data, qb->data + d);
    }
    w.insert(w.end(), kw->data, kw->data + (size_t)d * d);
    b.insert(b.end(), (size_t)d, 0.0f);
    w.insert(w.end(), vw->data, vw->data + (size_t)d * d);
    b.insert(b.end(), vb->data, vb->data + d);
    return up_lin(L, w, &b, with_q ? 3 * d : 2 * d, d);
  };

  WT_TRY(up_conv(conv1, "encoder.conv1", dims.n_mels));
  WT_TRY(up_conv(conv2, "encoder.conv2", d));
  {
    const HostTensor* pe = get("encoder.positional_embedding", {S, d});
    if (!pe) return WT_ERR_FORMAT;
    WT_TRY(upload_T(&enc_pos, pe->data, (size_t)S * d));
    if (f32mode()) enc_pos_f32 = (float*)enc_pos;
  }
  enc.resize(dims.n_enc_layers);
  for (int i = 0; i < dims.n_enc_layers; ++i) {
    const std::string p = "encoder.blocks." + std::to_string(i);
    WT_TRY(up_ln(enc[i].ln1, p + ".attn_ln"));
    WT_TRY(up_fused(enc[i].qkv, p + ".attn", true));
    WT_TRY(up_plain(enc[i].out, p + ".attn.out", d, d, true));
    WT_TRY(up_ln(enc[i].ln2, p + ".mlp_ln"));
    WT_TRY(up_plain(enc[i].fc1, p + ".mlp.0", 4 * d, d, true));
    WT_TRY(up_plain(enc[i].fc2, p + ".mlp.2", d, 4 * d, true));
  }
  WT_TRY(up_ln(ln_post, "encoder.ln_post"));
  {
    const HostTensor *te = get("decoder.token_embedding.weight", {V, d}), *pe = get("decoder.positional_embedding", {dims.n_text_ctx, d});
    if (!te || !pe) return WT_ERR_FORMAT;
    WT_TRY(upload_T(&tok_emb, te->data, (size_t)V * d));
    WT_TRY(upload_T(&dec_pos, pe->data, (size_t)dims.n_text_ctx * d));
  }
  dec.resize(dims.n_dec_layers);
  const size_t kv_elems = B * (size_t)H * S * 64;
  const bool want_persistent = persistent_decoder_wanted();
  const size_t cache_elems = B * (size_t)H * 32 * 64;
  for (int i = 0; i < dims.n_dec_layers; ++i) {
    const std::string p = "decoder.blocks." + std::to_string(i);
    DecLayer& L = dec[i];
    WT_TRY(up_ln(L.ln1, p + ".attn_ln"));
    WT_TRY(up_fused(L.qkv, p + ".attn", true));
    WT_TRY(up_plain(L.out, p + ".attn.out", d, d, true));
    WT_TRY(up_ln(L.lnx, p + ".cross_attn_ln"));
    WT_TRY(up_plain(L.xq, p + ".cross_attn.query", d, d, true));
    WT_TRY(up_fused(L.xkv, p + ".cross_attn", false));
    WT_TRY(up_plain(L.xout, p + ".cross_attn.out", d, d, true));
    WT_TRY(up_ln(L.ln2, p + ".mlp_ln"));
    WT_TRY(up_plain(L.fc1, p + ".mlp.0", 4 * d, d, true));
    WT_TRY(up_plain(L.fc2, p + ".mlp.2", d, 4 * d, true));
    if (want_persistent) {
      // streaming layout of the persistent decoder: [clip][head][chunk][K rows | V rows] (decoder_stream_sm100.cuh)
      const size_t cpp = (size_t)(S + wt::kXaChunk - 1) / wt::kXaChunk;
      WT_TRY(alloc_bytes(&L.kvx, B * (size_t)H * cpp * 2 * wt::kXaChunk * 64 * esz, true));
    } else {
      WT_TRY(alloc_bytes(&L.kx, kv_elems * esz));
      WT_TRY(alloc_bytes(&L.vx, kv_elems * esz));
    }
    WT_TRY(alloc_bytes(&L.kc, cache_elems * esz, true));
    WT_TRY(alloc_bytes(&L.vc, cache_elems * esz, true));
  }
  WT_TRY(up_ln(dec_ln, "decoder.ln"));

  const size_t rows = B * S;
  WT_TRY(alloc_bytes(&mel_tm, B * (WT_N_FRAMES + 2) * dims.n_mels * esz, true));
  WT_TRY(alloc_bytes(&x1, B * (WT_N_FRAMES + 2) * d * esz, true));
  WT_TRY(alloc_bytes(&x, rows * d * esz));
  WT_TRY(alloc_bytes(&xn, rows * d * esz));
  WT_TRY(alloc_bytes(&att, rows * d * esz));
  WT_TRY(alloc_bytes(&qkv, rows * 3 * d * esz));
  WT_TRY(alloc_bytes(&hbuf, rows * 4 * d * esz));
  WT_TRY(alloc_bytes(&enc_out, rows * d * esz));
  WT_TRY(alloc_bytes(&dx, B * d * esz));
  WT_TRY(alloc_bytes(&dxn, B * d * esz));
  WT_TRY(alloc_bytes(&datt, B * d * esz));
  WT_TRY(alloc_bytes(&dqkv, B * 3 * d * esz));
  WT_TRY(alloc_bytes(&dq, B * d * esz));
  WT_TRY(alloc_bytes(&dh, B * 4 * d * esz));
  WT_TRY(alloc(&logits, B * V));
  if (keep_logits) WT_TRY(alloc(&step_logits, (size_t)(WT_MAX_TOKENS - 1) * B * V));
  WT_TRY(alloc(&xa_part, B * H * wt::kXaMaxSplit * 66));
  WT_TRY(alloc(&xa_tickets, B * H, true));
  WT_TRY(alloc(&argmax_part, B * ((V + 63) / 64)));
  WT_TRY(alloc(&tokens, B * 32));
  WT_TRY(alloc(&lens, B));
  WT_TRY(alloc(&done, B));
  WT_TRY(alloc(&d_prompt, 32));
  has_model = true;
  return init_persistent_decoder();
}

// The persistent decoder: one thread-block cluster per group of clips, every cluster independent of the others
// (hardware cluster barriers only), so a plain launch with a cluster dimension — any number of groups queue up.
bool wt_engine::persistent_decoder_wanted() const {
  const char* ov = getenv("WT_DECODER");
  return !(f32mode() || dims.d_model > wt::kDecMaxD || (ov && strcmp(ov, "legacy") == 0));
}

int wt_engine::init_persistent_decoder() {
  if (!persistent_decoder_wanted()) return WT_OK;
  const void* kfn = (const void*)wt::decoder_stream_kernel;
  {
    const int kbK = dims.d_model / 32;
    if (dims.d_model % 64 != 0) return fail(WT_ERR_INVALID, "streaming decoder needs d_model %% 64 == 0");
    const int cand[6] = {8, 6, 4, 3, 2, 1};
    for (int i = 0; i < 6 && !ds_upw; ++i)
      if (kbK % cand[i] == 0) ds_upw = cand[i];
    int max_optin = 0;
    WT_CUDA_OK(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    ds_stages = (max_optin - wt::DsSmem::ring(dims.d_model)) / wt::kDsStageBytes;
    if (ds_stages > wt::kDsMaxStages) ds_stages = wt::kDsMaxStages;
    if (const char* os = getenv("WT_DEC_STAGES")) ds_stages = atoi(os) < ds_stages ? atoi(os) : ds_stages;
    if (ds_stages < 2) return fail(WT_ERR_INVALID, "streaming decoder: shared memory holds only %d ring stages at d_model %d", ds_stages, dims.d_model);
    ds_smem = wt::DsSmem::total(dims.d_model, ds_stages);
  }
  WT_CUDA_OK(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, ds_smem));
  if (const char* ka = getenv("WT_DEC_KV_AHEAD")) ds_kv_ahead = atoi(ka) < 0 ? 0 : atoi(ka);
  // cluster shape: 16 CTAs when the device can co-schedule a 16-CTA cluster per GPC, else the portable 8.
  // WT_DEC_CLUSTER / WT_DEC_GROUP override (probing).
  dec_cluster = 0;
  const int want[2] = {16, 8};
  const char* ocl = getenv("WT_DEC_CLUSTER");
  char ocl_buf[16];
  if (!ocl && dec_cluster_cfg > 0) {   // wt_config.decoder_cluster
    snprintf(ocl_buf, sizeof(ocl_buf), "%d", dec_cluster_cfg);
    ocl = ocl_buf;
  }
  for (int i = 0; i < 2 && !dec_cluster; ++i) {
    const int cs = ocl ? atoi(ocl) : want[i];
    if (cs < 1 || cs > wt::kDecMaxCluster) return fail(WT_ERR_INVALID, "WT_DEC_CLUSTER must be in [1, %d]", wt::kDecMaxCluster);
    if (cs > 8) {
      if (cudaFuncSetAttribute(kfn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
        cudaGetLastError();
        if (ocl) return fail(WT_ERR_CUDA, "cluster size %d is not allowed on this device", cs);
        continue;
      }
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(cs);
    cfg.blockDim = dim3(wt::kDsThreads);
    cfg.dynamicSmemBytes = ds_smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n_active = 0;
    if (cudaOccupancyMaxActiveClusters(&n_active, kfn, &cfg) != cudaSuccess) { cudaGetLastError(); n_active = 0; }
    if (n_active >= 1) { dec_cluster = cs; dec_active_clusters = n_active; }
    if (ocl) break;
  }
  if (!dec_cluster) return fail(WT_ERR_CUDA, "decoder_stream_kernel: no schedulable cluster shape");
  dec_group = dec_cluster >= 16 ? 8 : 4;
  if (const char* og = getenv("WT_DEC_GROUP")) dec_group = atoi(og);
  if (dec_group < 1 || dec_group > wt::kDecGroupMax) return fail(WT_ERR_INVALID, "WT_DEC_GROUP must be in [1, %d]", wt::kDecGroupMax);
  const int d = dims.d_model;
  dec_groups_cap = (max_batch + dec_group - 1) / dec_group;
  const size_t R = (size_t)dec_groups_cap * wt::kDecGroupMax;  // scratch rows: kDecGroupMax per group
  WT_TRY(alloc(&pd_qs, R * d));
  WT_TRY(alloc(&pd_finished, (size_t)dec_groups_cap, true));
  if (const char* tr = getenv("WT_DEC_TRACE")) {
    if (tr[0] == '1') {
      WT_TRY(alloc(&pd_trace, 32, true));
      fprintf(stderr, "[whisper_b200] streaming decoder: cluster %d CTAs x %d clips, %d clusters co-resident, smem %d B, ring %d stages, upw %d\n",
              dec_cluster, dec_group, dec_active_clusters, ds_smem, ds_stages, ds_upw);
    }
  }
  WT_TRY(init_stream_weights());
  dec_persistent = true;
  return WT_OK;
}

// Packed weight streams of the streaming decoder: every projection re-laid in consumption order for this cluster size,
// LayerNorm gains folded into the weight columns, LayerNorm biases into the projection biases.
int wt_engine::init_stream_weights() {
  const int d = dims.d_model, V = dims.n_vocab, G = dec_cluster;
  std::vector<wt::DsLayerDev> hl(dims.n_dec_layers);
  auto pack = [&](const Lin& lin, const LNp* ln, const uint4** w_out, const float** b_out, int gpw = 1) -> int {
    const int n_groups = (lin.N + 7) / 8, kbK = lin.K / 32;
    uint4* pw = nullptr;
    float* pb = nullptr;
    WT_TRY(alloc(&pw, (size_t)n_groups * kbK * 32));
    WT_TRY(alloc(&pb, (size_t)n_groups * 8, true));
    const long long total = (long long)n_groups * kbK * 32;
    const int blocks = (int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
    wt::ds_pack_weights_kernel<<<blocks, 256, 0, stream>>>((const bf16*)lin.w, ln ? ln->g : nullptr, pw, lin.N, lin.K, G, gpw);
    wt::ds_fold_bias_kernel<<<(lin.N * 32 + 255) / 256, 256, 0, stream>>>((const bf16*)lin.w, lin.b, ln ? ln->b : nullptr, pb, lin.N, lin.K);
    WT_CUDA_OK(cudaGetLastError());
    launches += 2;
    *w_out = pw;
    *b_out = pb;
    return WT_OK;
  };
  for (int i = 0; i < dims.n_dec_layers; ++i) {
    DecLayer& L = dec[i];
    wt::DsLayerDev& o = hl[i];
    WT_TRY(pack(L.qkv, &L.ln1, &o.w[wt::DS_QKV], &o.b[wt::DS_QKV]));
    WT_TRY(pack(L.out, nullptr, &o.w[wt::DS_OUT], &o.b[wt::DS_OUT]));
    WT_TRY(pack(L.xq, &L.lnx, &o.w[wt::DS_XQ], &o.b[wt::DS_XQ]));
    WT_TRY(pack(L.xout, nullptr, &o.w[wt::DS_XOUT], &o.b[wt::DS_XOUT]));
    WT_TRY(pack(L.fc1, &L.ln2, &o.w[wt::DS_FC1], &o.b[wt::DS_FC1]));
    WT_TRY(pack(L.fc2, nullptr, &o.w[wt::DS_FC2], &o.b[wt::DS_FC2]));
    o.kvx = (const bf16*)L.kvx;
    o.kc = (bf16*)L.kc; o.vc = (bf16*)L.vc;
  }
  {
    // tied logits matrix: MMA-ready tiles per CTA slice (ds_pack_logits_kernel), final LayerNorm folded in
    const int T = wt::ds_logit_tiles_per_cta(V, G), nkb = d / 64;
    const size_t chunks = (size_t)G * T * nkb * 128 * 8;   // 16-byte chunks
    WT_TRY(alloc(&ds_w_logits, chunks));
    WT_TRY(alloc(&ds_b_logits, (size_t)((V + 7) / 8) * 8, true));
    const int blocks = (int)((chunks + 255) / 256 < 8192 ? (chunks + 255) / 256 : 8192);
    wt::ds_pack_logits_kernel<<<blocks, 256, 0, stream>>>((const bf16*)tok_emb, dec_ln.g, ds_w_logits, V, d, G);
    wt::ds_fold_bias_kernel<<<(V * 32 + 255) / 256, 256, 0, stream>>>((const bf16*)tok_emb, nullptr, dec_ln.b, ds_b_logits, V, d);
    WT_CUDA_OK(cudaGetLastError());
    launches += 2;
  }
  WT_TRY(alloc(&d_ds_layers, hl.size()));
  WT_CUDA_OK(cudaMemcpyAsync(d_ds_layers, hl.data(), hl.size() * sizeof(wt::DsLayerDev), cudaMemcpyHostToDevice, stream));
  WT_CUDA_OK(wait_stream());
  return WT_OK;
}

int wt_engine::decode_persistent(int n_clips, int prompt_len, int max_total, int eot) {
  const int d = dims.d_model, H = dims.n_heads, S = dims.n_audio_ctx;
  // one wave: grow the group (up to kDecGroupMax clips) until every cluster of the launch is co-resident
  int gsz = dec_group;
  if (dec_group_user > 0) {
    gsz = dec_group_user;
  } else if (!getenv("WT_DEC_GROUP")) {
    const int need = (n_clips + dec_active_clusters - 1) / dec_active_clusters;
    if (need > gsz) gsz = need < wt::kDecGroupMax ? need : wt::kDecGroupMax;
  }
  const int n_groups = (n_clips + gsz - 1) / gsz;
  if (n_groups > dec_groups_cap)
    return fail(WT_ERR_INVALID, "decoder: %d clip groups of %d exceed the %d the scratch buffers were sized for", n_groups, gsz, dec_groups_cap);
  WT_CUDA_OK(cudaMemsetAsync(pd_finished, 0, (size_t)dec_groups_cap * sizeof(int), stream));
  {
    wt::DsParams q{};
    q.layers = d_ds_layers;
    q.n_layers = dims.n_dec_layers;
    q.n_clips = n_clips;
    q.clip0 = 0;
    q.group_clips = gsz;
    q.upw = ds_upw;
    q.n_stages = ds_stages;
    if (const char* dbg = getenv("WT_DEC_DEBUG")) q.debug = atoi(dbg);
    q.kv_ahead = ds_kv_ahead;
    q.d = d; q.H = H; q.S = S; q.V = dims.n_vocab;
    q.w_logits = ds_w_logits; q.b_logits = ds_b_logits;
    q.tok_emb = (const bf16*)tok_emb;
    q.pos_emb = (const bf16*)dec_pos;
    q.tokens = tokens; q.lens = lens; q.done = done;
    q.prompt_len = prompt_len; q.max_total = max_total; q.eot = eot;
    q.qs = pd_qs;
    {
      // whole (clip, head) pairs per CTA and the residual columns per warp must fit their fixed tables
      const int pairs = ((gsz < n_clips ? gsz : n_clips) * H + dec_cluster - 1) / dec_cluster;
      const int res_groups = (d / 8 + dec_cluster - 1) / dec_cluster;
      if (pairs > wt::kDsMaxSeg || res_groups > 16)
        return fail(WT_ERR_STATE, "streaming decoder: %d pairs / %d residual groups per CTA exceed the kernel's tables (cluster %d)", pairs, res_groups, dec_cluster);
    }
    q.step_logits = keep_logits ? step_logits : nullptr;
    q.step_logits_stride = (long long)max_batch * dims.n_vocab;
    q.trace = pd_trace;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(n_groups * dec_cluster);
    cfg.blockDim = dim3(wt::kDsThreads);
    cfg.dynamicSmemBytes = ds_smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = dec_cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    stage = ST_DEC_PERSISTENT;
    pre_launch();
    WT_CUDA_OK(cudaLaunchKernelEx(&cfg, wt::decoder_stream_kernel, q));
    return check_launch("decoder_stream_kernel");
  }
}

// mel -> encoder states (+ cross-attention K/V of every decoder layer), all resident.
int wt_engine::encode(const float* mel, int n_clips, int mel_mem, float* states_out, int out_mem, bool from_raw) {
  if (!has_model) return fail(WT_ERR_STATE, "engine was created without weights");
  if (n_clips < 1 || n_clips > max_batch) return fail(WT_ERR_INVALID, "n_clips %d outside [1, max_batch=%d]", n_clips, max_batch);
  const int d = dims.d_model, H = dims.n_heads, S = dims.n_audio_ctx, T = WT_N_FRAMES, nm = dims.n_mels;
  const float* src = nullptr;
  const int* cmax = nullptr;
  if (mel) {
    src = mel;
    if (mel_mem == WT_HOST) {
      WT_CUDA_OK(cudaMemcpyAsync(d_mel, mel, (size_t)n_clips * nm * T * 4, cudaMemcpyHostToDevice, stream));
      src = d_mel;
    }
  } else {
    if (mel_clips < n_clips || mel_len != T)
      return fail(WT_ERR_STATE, "no resident mel for %d clips of 3000 frames (call wt_log_mel on 480000-sample clips first)", n_clips);
    src = from_raw ? d_mel_raw : d_mel;
    cmax = from_raw ? d_clip_max : nullptr;
  }
  // [B][80][3000] f32 -> zero-padded time-major [B][3002][80] (clamp + affine fused when reading raw)
  {
    dim3 grid((T + 31) / 32, (nm + 31) / 32, n_clips);
    stage = ST_MEL_LAYOUT;
    decoding = false;
    pre_launch();
    if (f32mode())
      wt::mel_to_time_major_kernel<float><<<grid, 256, 0, stream>>>(src, cmax, (float*)mel_tm, nm, T);
    else
      wt::mel_to_time_major_kernel<bf16><<<grid, 256, 0, stream>>>(src, cmax, (bf16*)mel_tm, nm, T);
    WT_TRY(check_launch("mel_to_time_major_kernel"));
  }
  const int M1 = n_clips * T, M2 = n_clips * S;
  stage = ST_CONV;
  if (f32mode()) {
    wt::GemmF32Params p{};
    // conv1 + GELU -> interior rows of x1
    p.A = (const float*)mel_tm;
    p.amap = RowMap{(long long)(T + 2) * nm, T, nm};
    p.W = (const float*)conv1.w;
    p.bias = conv1.b;
    p.C = (float*)x1 + d;
    p.cmap = RowMap{(long long)(T + 2) * d, T, d};
    p.rmap = p.cmap;
    p.pos_rows = 1;
    p.M = M1;
    p.N = d;
    p.K = 3 * nm;
    p.flags = wt::EPI_BIAS | wt::EPI_GELU;
    pre_launch();
    wt::sgemm_f32_kernel<<<dim3((d + 63) / 64, (M1 + 63) / 64), 256, 0, stream>>>(p);
    WT_TRY(check_launch("sgemm_f32_kernel(conv1)"));
    // conv2 (stride 2) + GELU + positional rows -> x
    wt::GemmF32Params q{};
    q.A = (const float*)x1;
    q.amap = RowMap{(long long)(T + 2) * d, S, 2 * d};
    q.W = (const float*)conv2.w;
    q.bias = conv2.b;
    q.pos = enc_pos_f32;
    q.pos_rows = S;
    q.C = (float*)x;
    q.cmap = RowMap{0, M2, d};
    q.rmap = q.cmap;
    q.M = M2;
    q.N = d;
    q.K = 3 * d;
    q.flags = wt::EPI_BIAS | wt::EPI_GELU | wt::EPI_POS;
    pre_launch();
    wt::sgemm_f32_kernel<<<dim3((d + 63) / 64, (M2 + 63) / 64), 256, 0, stream>>>(q);
    WT_TRY(check_launch("sgemm_f32_kernel(conv2)"));
  } else {
    {
      const uint64_t ad[3] = {(uint64_t)nm, (uint64_t)(T + 2), (uint64_t)n_clips};
      const uint64_t as[2] = {(uint64_t)nm * 2, (uint64_t)(T + 2) * nm * 2};
      const int c0[3] = {0, 0, 0}, r0[3] = {0, 1, 2};
      WT_TRY(gemm_bf16(mel_tm, ad, as, T, n_clips, conv1, 3 * nm, 3, nm, c0, r0, (bf16*)x1 + d,
                       RowMap{(long long)(T + 2) * d, T, d}, wt::EPI_GELU, nullptr, RowMap{0, 1, 0}, nullptr));
    }
    {
      // rows of the map are PAIRS of padded time steps: pair p holds steps 2p, 2p+1 (2d elements)
      const uint64_t ad[3] = {(uint64_t)2 * d, (uint64_t)(T + 2) / 2, (uint64_t)n_clips};
      const uint64_t as[2] = {(uint64_t)2 * d * 2, (uint64_t)(T + 2) * d * 2};
      const int c0[3] = {0, d, 0}, r0[3] = {0, 0, 1};
      WT_TRY(gemm_bf16(x1, ad, as, S, n_clips, conv2, 3 * d, 3, d, c0, r0, x, RowMap{0, M2, d},
                       wt::EPI_GELU | wt::EPI_ADD, enc_pos, RowMap{0, S, d}, nullptr));
    }
  }
  gemm_stage = -1;
  for (int l = 0; l < dims.n_enc_layers; ++l) {
    EncLayer& L = enc[l];
    WT_TRY(layernorm(x, L.ln1, xn, M2));
    WT_TRY(linear(xn, M2, L.qkv, qkv, 0));
    stage = ST_ENC_ATTN;
    if (f32mode()) {
      pre_launch();
      wt::attention_simt_kernel<float><<<dim3((S + 127) / 128, H, n_clips), 128, 0, stream>>>((const float*)qkv, (float*)att, S, H);
      WT_TRY(check_launch("attention_simt_kernel"));
    } else {
      WT_TRY(attention_bf16(qkv, att, n_clips, S, H));
    }
    WT_TRY(linear(att, M2, L.out, x, 0, x));
    WT_TRY(layernorm(x, L.ln2, xn, M2));
    WT_TRY(linear(xn, M2, L.fc1, hbuf, wt::EPI_GELU));
    WT_TRY(linear(hbuf, M2, L.fc2, x, 0, x));
  }
  WT_TRY(layernorm(x, ln_post, enc_out, M2));
  // cross-attention K/V of every decoder layer, once per clip (the reference recomputes them inside every
  // decoder Invoke: whisper.cpp:375)
  for (int l = 0; l < dims.n_dec_layers; ++l) {
    DecLayer& L = dec[l];
    gemm_stage = ST_XKV;
    if (!f32mode()) {
      // the GEMM epilogue writes K and V straight into the per-head layout [clip][H][S][64]
      next_kv_k = dec_persistent ? L.kvx : L.kx;
      next_kv_v = L.vx;
      next_kv_chunked = dec_persistent;
      {
        // batched per clip (S rows each; the last of a clip's 12 row tiles is padded by TMA): clip = batch, and with
        // 128-key chunks the chunk index is the row tile — the scatter epilogue needs no division
        const uint64_t ad[3] = {(uint64_t)L.xkv.K, (uint64_t)S, (uint64_t)n_clips};
        const uint64_t as[2] = {(uint64_t)L.xkv.K * 2, (uint64_t)L.xkv.K * 2 * (uint64_t)S};
        RowMap plain{0, M2, L.xkv.N};
        stage = ST_XKV;
        WT_TRY(gemm_bf16(enc_out, ad, as, S, n_clips, L.xkv, L.xkv.K, 1, 0, nullptr, nullptr, nullptr, plain,
                         wt::EPI_SPLIT_KV, nullptr, plain, nullptr));
      }
      gemm_stage = -1;
      continue;
    }
    WT_TRY(linear(enc_out, M2, L.xkv, qkv, 0));
    gemm_stage = -1;
    stage = ST_XKV;
    pre_launch();
    wt::split_heads_kv_kernel<float><<<num_sms * 4, 256, 0, stream>>>((const float*)qkv, (float*)L.kx, (float*)L.vx, n_clips, S, H);
    WT_TRY(check_launch("split_heads_kv_kernel"));
  }
  enc_clips = n_clips;
  if (states_out) {
    const size_t n = (size_t)M2 * d;
    float* dst = states_out;
    if (f32mode()) {
      WT_CUDA_OK(cudaMemcpyAsync(dst, enc_out, n * 4, out_mem == WT_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, stream));
    } else {
      float* tmp = states_out;
      if (out_mem == WT_HOST) {
        if (!io_f32) WT_TRY(alloc(&io_f32, (size_t)max_batch * S * d));
        tmp = io_f32;
      }
      pre_launch();
      wt::convert_kernel<bf16, float><<<num_sms * 4, 256, 0, stream>>>((const bf16*)enc_out, tmp, (long long)n);
      WT_TRY(check_launch("convert_kernel"));
      if (out_mem == WT_HOST) WT_CUDA_OK(cudaMemcpyAsync(dst, tmp, n * 4, cudaMemcpyDeviceToHost, stream));
    }
  }
  if ((mel && mel_mem == WT_HOST) || (states_out && out_mem == WT_HOST)) WT_CUDA_OK(wait_stream());
  return WT_OK;
}

template <class T>
static int decode_step_attn(wt_engine* e, DecLayer& L, int n_clips, int pos, bool cross) {
  const int H = e->dims.n_heads, S = e->dims.n_audio_ctx;
  if (!cross) {
    e->stage = ST_DEC_SELF_ATTN;
    e->pre_launch();
    wt::decode_self_attn_kernel<T><<<dim3(H, n_clips), 64, 0, e->stream>>>((const T*)e->dqkv, (T*)L.kc, (T*)L.vc, (T*)e->datt, pos, 32, H);
    return e->check_launch("decode_self_attn_kernel");
  }
  // enough splits to keep >= 8 streaming CTAs on every SM
  int nsplit = (8 * e->num_sms + n_clips * H - 1) / (n_clips * H);
  if (nsplit < 1) nsplit = 1;
  if (nsplit > wt::kXaMaxSplit) nsplit = wt::kXaMaxSplit;
  e->stage = ST_DEC_CROSS_ATTN;
  e->pre_launch();
  wt::decode_cross_attn_kernel<T><<<dim3(nsplit, H, n_clips), 128, 0, e->stream>>>((const T*)e->dq, (const T*)L.kx, (const T*)L.vx, (T*)e->datt,
                                                                                   e->xa_part, e->xa_tickets, S, H);
  return e->check_launch("decode_cross_attn_kernel");
}

// Greedy loop of Decoder::forward (whisper.cpp:363-402) over the resident encoder states, one new position per
// step with a KV cache; per-clip `done` flags replace the reference's early break.
int wt_engine::decode(int n_clips, const int64_t* prompt, int prompt_len, int max_total, int64_t eot, int32_t* ids,
                      int32_t* lens_o, int out_mem) {
  if (!has_model) return fail(WT_ERR_STATE, "engine was created without weights");
  if (n_clips < 1 || n_clips > enc_clips) return fail(WT_ERR_STATE, "encoder states resident for %d clips, decode asked for %d", enc_clips, n_clips);
  if (!prompt || prompt_len < 1 || prompt_len > max_total || max_total > 32 || max_total < 2)
    return fail(WT_ERR_INVALID, "need 1 <= prompt_len <= max_total_tokens <= 32");
  if (!ids || !lens_o) return fail(WT_ERR_INVALID, "ids_out / lens_out are null");
  const int d = dims.d_model, V = dims.n_vocab;
  int hp[32];
  for (int i = 0; i < prompt_len; ++i) {
    if (prompt[i] < 0 || prompt[i] >= V) return fail(WT_ERR_INVALID, "prompt token %lld outside the vocabulary", (long long)prompt[i]);
    hp[i] = (int)prompt[i];
  }
  WT_CUDA_OK(cudaMemcpyAsync(d_prompt, hp, prompt_len * 4, cudaMemcpyHostToDevice, stream));
  decoding = true;
  gemm_stage = -1;
  stage = ST_DEC_EMBED;
  pre_launch();
  wt::init_decode_state_kernel<<<n_clips, 32, 0, stream>>>(d_prompt, prompt_len, tokens, 32, lens, done);
  WT_TRY(check_launch("init_decode_state_kernel"));
  if (dec_persistent) {
    WT_TRY(decode_persistent(n_clips, prompt_len, max_total, (int)eot));
    decoding = false;
    return copy_decode_results(n_clips, max_total, ids, lens_o, out_mem);
  }
  Lin logit_lin;
  logit_lin.w = tok_emb;
  logit_lin.b = nullptr;
  logit_lin.N = V;
  logit_lin.K = d;
  // position `pos` = index of the last token fed; the reference feeds i+1 tokens for i = prompt_len-1 .. 29 and
  // appends one (whisper.cpp:367-395), so positions 0 .. max_total-2 are consumed and logits start at prompt_len-1.
  for (int pos = 0; pos <= max_total - 2; ++pos) {
    stage = ST_DEC_EMBED;
    pre_launch();
    if (f32mode())
      wt::embed_kernel<float><<<n_clips, 128, 0, stream>>>(tokens, 32, pos, (const float*)tok_emb, (const float*)dec_pos, (float*)dx, d);
    else
      wt::embed_kernel<bf16><<<n_clips, 128, 0, stream>>>(tokens, 32, pos, (const bf16*)tok_emb, (const bf16*)dec_pos, (bf16*)dx, d);
    WT_TRY(check_launch("embed_kernel"));
    for (int l = 0; l < dims.n_dec_layers; ++l) {
      DecLayer& L = dec[l];
      WT_TRY(layernorm(dx, L.ln1, dxn, n_clips));
      WT_TRY(linear(dxn, n_clips, L.qkv, dqkv, 0));
      WT_TRY(f32mode() ? decode_step_attn<float>(this, L, n_clips, pos, false) : decode_step_attn<bf16>(this, L, n_clips, pos, false));
      WT_TRY(linear(datt, n_clips, L.out, dx, 0, dx));
      WT_TRY(layernorm(dx, L.lnx, dxn, n_clips));
      WT_TRY(linear(dxn, n_clips, L.xq, dq, 0));
      WT_TRY(f32mode() ? decode_step_attn<float>(this, L, n_clips, pos, true) : decode_step_attn<bf16>(this, L, n_clips, pos, true));
      WT_TRY(linear(datt, n_clips, L.xout, dx, 0, dx));
      WT_TRY(layernorm(dx, L.ln2, dxn, n_clips));
      WT_TRY(linear(dxn, n_clips, L.fc1, dh, wt::EPI_GELU));
      WT_TRY(linear(dh, n_clips, L.fc2, dx, 0, dx));
    }
    if (pos < prompt_len - 1) continue;
    WT_TRY(layernorm(dx, dec_ln, dxn, n_clips));
    const int step = pos - (prompt_len - 1);
    gemm_stage = ST_LOGITS;
    stage = ST_LOGITS;
    if (f32mode() || keep_logits) {
      float* lg = keep_logits ? step_logits + (size_t)step * max_batch * V : logits;
      if (f32mode()) {
        WT_TRY(linear(dxn, n_clips, logit_lin, lg, 0));
      } else {
        const uint64_t ad[3] = {(uint64_t)d, (uint64_t)n_clips, 1};
        const uint64_t as[2] = {(uint64_t)d * 2, (uint64_t)d * 2 * n_clips};
        WT_TRY(gemm_bf16(dxn, ad, as, n_clips, 1, logit_lin, d, 1, 0, nullptr, nullptr, lg, RowMap{0, n_clips, V},
                         wt::EPI_OUT_F32, nullptr, RowMap{0, 1, 0}, nullptr));
      }
      gemm_stage = -1;
      stage = ST_ARGMAX;
      pre_launch();
      wt::argmax_append_kernel<<<n_clips, 256, 0, stream>>>(lg, V, tokens, 32, lens, done, pos, (int)eot);
      WT_TRY(check_launch("argmax_append_kernel"));
    } else {
      // fused: per-(row, N-tile) (max, last index) in the GEMM epilogue, logits never reach HBM
      const int bn = 128, n_part = (V + bn - 1) / bn;
      const uint64_t ad[3] = {(uint64_t)d, (uint64_t)n_clips, 1};
      const uint64_t as[2] = {(uint64_t)d * 2, (uint64_t)d * 2 * n_clips};
      WT_TRY(gemm_bf16(dxn, ad, as, n_clips, 1, logit_lin, d, 1, 0, nullptr, nullptr, nullptr, RowMap{0, n_clips, V},
                       wt::EPI_ARGMAX, nullptr, RowMap{0, 1, 0}, argmax_part, bn));
      gemm_stage = -1;
      stage = ST_ARGMAX;
      pre_launch();
      wt::argmax_partials_kernel<<<n_clips, 128, 0, stream>>>(argmax_part, n_part, tokens, 32, lens, done, pos, (int)eot);
      WT_TRY(check_launch("argmax_partials_kernel"));
    }
  }
  decoding = false;
  return copy_decode_results(n_clips, max_total, ids, lens_o, out_mem);
}

// results: [n_clips][max_total] ids (tokens has stride 32), lens
int wt_engine::copy_decode_results(int n_clips, int max_total, int32_t* ids, int32_t* lens_o, int out_mem) {
  if (out_mem == WT_HOST) {
    std::vector<int> ht((size_t)n_clips * 32), hl(n_clips);
    WT_CUDA_OK(cudaMemcpyAsync(ht.data(), tokens, ht.size() * 4, cudaMemcpyDeviceToHost, stream));
    WT_CUDA_OK(cudaMemcpyAsync(hl.data(), lens, hl.size() * 4, cudaMemcpyDeviceToHost, stream));
    WT_CUDA_OK(wait_stream());
    for (int b = 0; b < n_clips; ++b) {
      for (int i = 0; i < max_total; ++i) ids[(size_t)b * max_total + i] = ht[(size_t)b * 32 + i];
      lens_o[b] = hl[b];
    }
  } else {
    WT_CUDA_OK(cudaMemcpy2DAsync(ids, (size_t)max_total * 4, tokens, 32 * 4, (size_t)max_total * 4, n_clips, cudaMemcpyDeviceToDevice, stream));
    WT_CUDA_OK(cudaMemcpyAsync(lens_o, lens, (size_t)n_clips * 4, cudaMemcpyDeviceToDevice, stream));
  }
  return WT_OK;
}

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

const char* wt_last_error(void) { return g_err; }
const char* wt_version(void) { return "whisper_b200 0.1 (sm_100a)"; }

int wt_engine_create(const wt_config* cfg, const char* weights_path, int device, wt_engine** out) {
  if (!out) return fail(WT_ERR_INVALID, "out is null");
  *out = nullptr;
  if (!cfg || cfg->struct_size != (int32_t)sizeof(wt_config)) return fail(WT_ERR_INVALID, "wt_config.struct_size mismatch");
  if (cfg->max_batch < 1 || cfg->max_batch > 4096) return fail(WT_ERR_INVALID, "max_batch outside [1, 4096]");
  if (cfg->precision != WT_PRECISION_BF16 && cfg->precision != WT_PRECISION_FP32_VERIFY) return fail(WT_ERR_INVALID, "unknown precision");
  if (cfg->decoder_cluster < 0 || cfg->decoder_cluster > 16) return fail(WT_ERR_INVALID, "decoder_cluster outside [0, 16]");
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1) return fail(WT_ERR_CUDA, "no CUDA device visible: this library has no CPU path");
  if (device < 0 || device >= n_dev) return fail(WT_ERR_INVALID, "device %d outside [0, %d)", device, n_dev);
  WT_CUDA_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  WT_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(WT_ERR_CUDA, "device %d is sm_%d%d; this build is sm_100a only", device, prop.major, prop.minor);
  wt_engine* e = new wt_engine();
  e->device = device;
  e->precision = cfg->precision;
  e->max_batch = cfg->max_batch;
  e->keep_logits = cfg->keep_logits;
  e->dec_cluster_cfg = cfg->decoder_cluster;
  e->num_sms = prop.multiProcessorCount;
  int st = WT_OK;
  if (cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking) != cudaSuccess) st = fail(WT_ERR_CUDA, "cudaStreamCreate failed");
  e->stream = e->own_stream;
  if (st == WT_OK) st = e->init_frontend();
  if (st == WT_OK && weights_path) st = e->load_model(weights_path);
  if (st == WT_OK && cudaDeviceSynchronize() != cudaSuccess) st = fail(WT_ERR_CUDA, "device sync after load failed");
  if (st != WT_OK) {
    delete e;
    return st;
  }
  *out = e;
  return WT_OK;
}

void wt_engine_destroy(wt_engine* e) { delete e; }

int wt_engine_dims(const wt_engine* e, wt_dims* out) {
  if (!e || !out) return fail(WT_ERR_INVALID, "null argument");
  if (!e->has_model) return fail(WT_ERR_STATE, "engine was created without weights");
  *out = e->dims;
  return WT_OK;
}

int wt_set_mel_filters(wt_engine* e, const float* filters, int n_mel, int n_freq) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  return e->set_filters(filters, n_mel, n_freq);
}

int wt_set_stream(wt_engine* e, void* s) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaStream_t ns = s ? (cudaStream_t)s : e->own_stream;
  if (ns != e->stream) {
    // the engine's buffers and its tile-scheduler counters belong to ONE stream at a time: work still in flight on the
    // old stream finishes before anything is enqueued on the new one
    cudaSetDevice(e->device);
    WT_CUDA_OK(e->wait_stream());
    e->stream = ns;
  }
  return WT_OK;
}

int wt_synchronize(wt_engine* e) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  WT_CUDA_OK(e->wait_stream());
  return WT_OK;
}

int wt_log_mel(wt_engine* e, const float* pcm, int n_clips, int n_samples, int pcm_mem, float* mel_out, int out_mem) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  return e->log_mel(pcm, n_clips, n_samples, pcm_mem, mel_out, out_mem, true);
}

int wt_encode(wt_engine* e, const float* mel, int n_clips, int mel_mem, float* states_out, int out_mem) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  return e->encode(mel, n_clips, mel_mem, states_out, out_mem, false);
}

int wt_decode_greedy(wt_engine* e, int n_clips, const int64_t* prompt, int prompt_len, int max_total_tokens, int64_t eot,
                     int32_t* ids_out, int32_t* lens_out, int out_mem) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  return e->decode(n_clips, prompt, prompt_len, max_total_tokens, eot, ids_out, lens_out, out_mem);
}

int wt_transcribe_tokens(wt_engine* e, const float* pcm, int n_clips, int n_samples, int pcm_mem, const int64_t* prompt,
                         int prompt_len, int max_total_tokens, int64_t eot, int32_t* ids_out, int32_t* lens_out,
                         int out_mem) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  if (n_samples != WT_N_SAMPLES) return fail(WT_ERR_INVALID, "wt_transcribe_tokens needs 480000-sample windows (pad/trim like whisper.cpp:753)");
  // one host wait per pass: mel, encoder and decoder are enqueued back to back, the wait that frees the caller's PCM
  // buffer is the one that delivers the ids (or, with device-resident outputs, a single wait at the end)
  e->defer_pcm_wait = true;
  int st = e->log_mel(pcm, n_clips, n_samples, pcm_mem == WT_HOST ? WT_HOST : WT_DEVICE, nullptr, WT_DEVICE, false);
  e->defer_pcm_wait = false;
  if (st != WT_OK) {
    if (pcm_mem == WT_HOST) e->wait_stream();
    return st;
  }
  st = e->encode(nullptr, n_clips, WT_DEVICE, nullptr, WT_DEVICE, true);
  if (st == WT_OK) st = e->decode(n_clips, prompt, prompt_len, max_total_tokens, eot, ids_out, lens_out, out_mem);
  if (pcm_mem == WT_HOST && (st != WT_OK || out_mem != WT_HOST)) {
    cudaError_t ce = e->wait_stream();
    if (st == WT_OK && ce != cudaSuccess) st = fail(WT_ERR_CUDA, "stream wait failed: %s", cudaGetErrorString(ce));
  }
  return st;
}

int wt_log_mel_i16(wt_engine* e, const int16_t* pcm, int n_clips, int n_samples, int pcm_mem, float* mel_out, int out_mem) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  return e->log_mel(pcm, n_clips, n_samples, pcm_mem, mel_out, out_mem, true, true);
}

int wt_transcribe_tokens_i16(wt_engine* e, const int16_t* pcm, int n_clips, int n_samples, int pcm_mem, const int64_t* prompt,
                             int prompt_len, int max_total_tokens, int64_t eot, int32_t* ids_out, int32_t* lens_out,
                             int out_mem) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  if (n_samples != WT_N_SAMPLES) return fail(WT_ERR_INVALID, "wt_transcribe_tokens_i16 needs 480000-sample windows (pad/trim like whisper.cpp:753)");
  // one host wait per pass: mel, encoder and decoder are enqueued back to back, the wait that frees the caller's PCM
  // buffer is the one that delivers the ids (or, with device-resident outputs, a single wait at the end)
  e->defer_pcm_wait = true;
  int st = e->log_mel(pcm, n_clips, n_samples, pcm_mem == WT_HOST ? WT_HOST : WT_DEVICE, nullptr, WT_DEVICE, false, true);
  e->defer_pcm_wait = false;
  if (st != WT_OK) {
    if (pcm_mem == WT_HOST) e->wait_stream();
    return st;
  }
  st = e->encode(nullptr, n_clips, WT_DEVICE, nullptr, WT_DEVICE, true);
  if (st == WT_OK) st = e->decode(n_clips, prompt, prompt_len, max_total_tokens, eot, ids_out, lens_out, out_mem);
  if (pcm_mem == WT_HOST && (st != WT_OK || out_mem != WT_HOST)) {
    cudaError_t ce = e->wait_stream();
    if (st == WT_OK && ce != cudaSuccess) st = fail(WT_ERR_CUDA, "stream wait failed: %s", cudaGetErrorString(ce));
  }
  return st;
}

int wt_get_step_logits(wt_engine* e, int step, int n_clips, float* out, int out_mem) {
  if (!e || !out) return fail(WT_ERR_INVALID, "null argument");
  if (!e->keep_logits || !e->step_logits) return fail(WT_ERR_STATE, "engine was not created with keep_logits");
  if (step < 0 || step >= WT_MAX_TOKENS - 1 || n_clips < 1 || n_clips > e->max_batch) return fail(WT_ERR_INVALID, "bad step / n_clips");
  cudaSetDevice(e->device);
  const size_t V = e->dims.n_vocab;
  WT_CUDA_OK(cudaMemcpyAsync(out, e->step_logits + (size_t)step * e->max_batch * V, (size_t)n_clips * V * 4,
                             out_mem == WT_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, e->stream));
  WT_CUDA_OK(e->wait_stream());
  return WT_OK;
}

int64_t wt_launch_count(const wt_engine* e) { return e ? e->launches : 0; }
int wt_set_decoder_group(wt_engine* e, int clips_per_cluster) {
  if (!e) return fail(WT_ERR_INVALID, "null engine");
  if (clips_per_cluster < 0 || clips_per_cluster > wt::kDecGroupMax)
    return fail(WT_ERR_INVALID, "clips_per_cluster must be in [0, %d]", wt::kDecGroupMax);
  e->dec_group_user = clips_per_cluster;
  return WT_OK;
}

int wt_debug_decoder_trace(wt_engine* e, uint64_t* out32, int reset) {
  if (!e || !out32) return fail(WT_ERR_INVALID, "null argument");
  if (!e->pd_trace) return fail(WT_ERR_STATE, "engine was not created under WT_DEC_TRACE=1 (or the persistent decoder is off)");
  cudaSetDevice(e->device);
  WT_CUDA_OK(e->wait_stream());
  WT_CUDA_OK(cudaMemcpy(out32, e->pd_trace, 32 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  if (reset) WT_CUDA_OK(cudaMemset(e->pd_trace, 0, 32 * sizeof(uint64_t)));
  return WT_OK;
}

// Debug: digest of a resident buffer as the last call left it (the bitwise-reproducibility tests and the stage bisect).
int wt_debug_hash(wt_engine* e, const char* what, int n_clips, uint64_t* out) {
  if (!e || !what || !out) return fail(WT_ERR_INVALID, "null argument");
  if (n_clips < 1 || n_clips > e->max_batch) return fail(WT_ERR_INVALID, "n_clips %d outside [1, max_batch=%d]", n_clips, e->max_batch);
  cudaSetDevice(e->device);
  const std::string w(what);
  const size_t B = (size_t)n_clips;
  const void* buf = nullptr;
  size_t bytes = 0;
  if (w == "mel_raw") { buf = e->d_mel_raw; bytes = B * e->n_mel * WT_N_FRAMES * 4; }
  else if (!e->has_model) return fail(WT_ERR_STATE, "engine was created without weights");
  else {
    const size_t d = e->dims.d_model, S = e->dims.n_audio_ctx, H = e->dims.n_heads, esz = e->esz;
    if (w == "mel_tm") { buf = e->mel_tm; bytes = B * (WT_N_FRAMES + 2) * e->dims.n_mels * esz; }
    else if (w == "conv1") { buf = e->x1; bytes = B * (WT_N_FRAMES + 2) * d * esz; }
    else if (w == "x") { buf = e->x; bytes = B * S * d * esz; }
    else if (w == "qkv") { buf = e->qkv; bytes = B * S * 3 * d * esz; }
    else if (w == "att") { buf = e->att; bytes = B * S * d * esz; }
    else if (w == "hidden") { buf = e->hbuf; bytes = B * S * 4 * d * esz; }
    else if (w == "enc_out") { buf = e->enc_out; bytes = B * S * d * esz; }
    else if (w == "tokens") { buf = e->tokens; bytes = B * 32 * 4; }
    else if (w.rfind("cross_kv", 0) == 0) {
      const int l = atoi(w.c_str() + 8);
      if (l < 0 || l >= (int)e->dec.size()) return fail(WT_ERR_INVALID, "no decoder layer %d", l);
      if (e->dec[l].kvx) {
        const size_t cpp = (S + wt::kXaChunk - 1) / wt::kXaChunk;
        buf = e->dec[l].kvx; bytes = B * H * cpp * 2 * wt::kXaChunk * 64 * esz;
      } else { buf = e->dec[l].kx; bytes = B * H * S * 64 * esz; }
    } else return fail(WT_ERR_INVALID, "unknown buffer '%s'", what);
  }
  if (!buf) return fail(WT_ERR_STATE, "buffer '%s' is not allocated in this engine", what);
  if (!e->d_hash) WT_TRY(e->alloc(&e->d_hash, 1));
  WT_CUDA_OK(cudaMemsetAsync(e->d_hash, 0, 8, e->stream));
  wt::hash_words_kernel<<<e->num_sms * 8, 256, 0, e->stream>>>((const uint32_t*)buf, (long long)(bytes / 4), e->d_hash);
  WT_CUDA_OK(cudaGetLastError());
  unsigned long long h = 0;
  WT_CUDA_OK(cudaMemcpyAsync(&h, e->d_hash, 8, cudaMemcpyDeviceToHost, e->stream));
  WT_CUDA_OK(e->wait_stream());
  *out = h;
  return WT_OK;
}

int wt_profile_enable(wt_engine* e, int enable) {
  if (!e) return fail(WT_ERR_INVALID, "engine is null");
  cudaSetDevice(e->device);
  WT_CUDA_OK(e->wait_stream());
  for (auto& r : e->prof) { e->ev_pool.push_back(r.a); e->ev_pool.push_back(r.b); }
  e->prof.clear();
  e->profiling = enable != 0;
  return WT_OK;
}

int wt_profile_read(wt_engine* e, float* ms_out, int32_t* launches_out, int n_stages) {
  if (!e || !ms_out || !launches_out || n_stages < ST_COUNT) return fail(WT_ERR_INVALID, "need arrays of >= %d stages", (int)ST_COUNT);
  cudaSetDevice(e->device);
  WT_CUDA_OK(e->wait_stream());
  for (int i = 0; i < n_stages; ++i) { ms_out[i] = 0.0f; launches_out[i] = 0; }
  for (auto& r : e->prof) {
    float ms = 0.0f;
    WT_CUDA_OK(cudaEventElapsedTime(&ms, r.a, r.b));
    ms_out[r.stage] += ms;
    launches_out[r.stage] += 1;
  }
  return WT_OK;
}

int wt_stage_count(void) { return ST_COUNT; }
const char* wt_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }

static int debug_scratch_engine(void* cuda_stream, wt_engine** out) {
  static thread_local wt_engine* scratch = nullptr;
  int dev = 0;
  WT_CUDA_OK(cudaGetDevice(&dev));
  if (!scratch || scratch->device != dev) {
    cudaDeviceProp prop;
    WT_CUDA_OK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) return fail(WT_ERR_CUDA, "sm_100a only");
    scratch = new wt_engine();
    scratch->device = dev;
    scratch->num_sms = prop.multiProcessorCount;
    WT_TRY(scratch->init_kernels());
    WT_CUDA_OK(cudaDeviceSynchronize());   // the scheduler counters' memset precedes the first launch on any stream
  }
  scratch->stream = (cudaStream_t)cuda_stream;
  scratch->precision = WT_PRECISION_BF16;
  *out = scratch;
  return WT_OK;
}

int wt_debug_gemm_bf16(void* cuda_stream, const void* A, const void* W, const float* bias, void* C, int M, int N, int K,
                       int flags) {
  if (!A || !W || !C || M < 1 || N < 8 || K < 8 || (K % 8) || (N % 8)) return fail(WT_ERR_INVALID, "bad GEMM arguments");
  wt_engine* scratch = nullptr;
  WT_TRY(debug_scratch_engine(cuda_stream, &scratch));
  Lin L;
  L.w = const_cast<void*>(W);
  L.b = const_cast<float*>(bias);
  L.N = N;
  L.K = K;
  return scratch->linear(A, M, L, C, flags & wt::EPI_GELU);
}

int wt_debug_gemm_ex(void* cuda_stream, const wt_debug_gemm_args* a) {
  if (!a || a->struct_size != (int32_t)sizeof(wt_debug_gemm_args)) return fail(WT_ERR_INVALID, "wt_debug_gemm_args.struct_size mismatch");
  if (!a->A || !a->W || a->N < 8 || (a->N % 8) || a->K_total < 8 || a->taps < 1 || a->taps > 3 || a->rows_per_batch < 1 || a->n_batches < 1)
    return fail(WT_ERR_INVALID, "bad GEMM arguments");
  wt_engine* scratch = nullptr;
  WT_TRY(debug_scratch_engine(cuda_stream, &scratch));
  Lin L;
  L.w = const_cast<void*>(a->W);
  L.b = (a->flags & wt::EPI_BIAS) ? const_cast<float*>(a->bias) : nullptr;
  L.N = a->N;
  L.K = a->K_total;
  scratch->next_kv_k = a->kv_out;
  scratch->next_kv_v = a->kv_out_v;
  scratch->next_kv_chunked = a->kv_chunked != 0;
  scratch->dims.n_audio_ctx = a->kv_S;
  scratch->dims.n_heads = a->kv_H;
  scratch->stage = ST_ENC_GEMM;
  const RowMap cmap{a->c_batch_stride, a->c_rows > 0 ? a->c_rows : 1, a->c_row_stride};
  const RowMap amap{a->add_batch_stride, a->add_rows > 0 ? a->add_rows : 1, a->add_row_stride};
  return scratch->gemm_bf16(a->A, a->a_dims, a->a_strides, a->rows_per_batch, a->n_batches, L, a->K_total, a->taps, a->tap_k,
                            a->tap_c0, a->tap_row, a->C, cmap, a->flags & ~wt::EPI_BIAS, a->addend, amap,
                            (unsigned long long*)a->argmax_part, a->force_bn);
}

int wt_debug_attention_bf16(void* cuda_stream, const void* qkv, void* out, int n_clips, int S, int H) {
  if (!qkv || !out || n_clips < 1 || S < 1 || H < 1) return fail(WT_ERR_INVALID, "bad attention arguments");
  wt_engine* scratch = nullptr;
  WT_TRY(debug_scratch_engine(cuda_stream, &scratch));
  return scratch->attention_bf16(qkv, out, n_clips, S, H);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// wt_gemm_argmax_noalloc: the fused logits + argmax GEMM (EPI_ARGMAX, whisper.cpp:346-361's rule) for callers that may
// not allocate — CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc. No counterpart in the
// reference. Everything the kernels need besides the caller's operands —
// the tile scheduler's two counters and the per-column-tile partial winners [M][ceil(N / bn)] — lives in ONE file-scope
// device array of fixed size (every GPU that loads the library holds these kNoAllocScratchBytes until exit). In host
// code the array's name is only a stand-in: its device address is taken once per device, with that device current,
// through cudaGetSymbolAddress, and it is that pointer the kernels get.
// All calls share the array and it keeps the last call's contents: the scheduler words are cleared on the caller's
// stream before every launch, the partials in use are fully overwritten; calls that overlap on the GPU (two streams,
// two host threads) race on it — serialise them.
// ------------------------------------------------------------------------------------------------
namespace {
constexpr int kNoAllocHeaderBytes = 256;                    // scheduler counters (sched[0] next tile, sched[1] CTAs done)
constexpr int kNoAllocMaxPartials = 131072;                 // u64 partial winners: M * ceil(N / bn) <= this (1 MiB)
constexpr size_t kNoAllocScratchBytes = kNoAllocHeaderBytes + (size_t)kNoAllocMaxPartials * 8;
constexpr int kNoAllocMaxDevices = 64;
struct NoAllocDevice {
  unsigned char* ptr = nullptr;   // device address of g_noalloc_scratch on this device
  int num_sms = 0;
};
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
std::mutex g_noalloc_mu;
}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace wt {
// winner of each row out of its per-column-tile partials (packed (ordered value, index): one integer max)
__global__ void __launch_bounds__(128) noalloc_argmax_rows_kernel(const unsigned long long* __restrict__ part, int n_part,
                                                                   int* __restrict__ ids_out) {
  __shared__ unsigned long long s_best[4];
  const int m = blockIdx.x;
  unsigned long long best = 0ull;
  for (int i = threadIdx.x; i < n_part; i += 128) best = u64_max(best, part[(long long)m * n_part + i]);
  best = warp_max_u64(best);
  if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 4; ++w) best = u64_max(best, s_best[w]);
    ids_out[m] = (int)(best & 0xffffffffull);
  }
}
}  // namespace wt

// once per device, with that device current: shared-memory attributes of the GEMM variants, SM count, scratch address
static int noalloc_device_setup(int dev, NoAllocDevice** out) {
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(WT_ERR_INVALID, "device %d outside [0, %d)", dev, kNoAllocMaxDevices);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  NoAllocDevice& nd = g_noalloc_dev[dev];
  if (!nd.ptr) {
    int major = 0, sms = 0;
    WT_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    WT_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (major != 10) return fail(WT_ERR_CUDA, "sm_100a only");
    WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<64>::kTotal));
    WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<128>::kTotal));
    WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<192>::kTotal));
    WT_CUDA_OK(cudaFuncSetAttribute(wt::gemm_bf16_sm100_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, wt::GemmSmem<256>::kTotal));
    void* sym = nullptr;
    WT_CUDA_OK(cudaGetSymbolAddress(&sym, g_noalloc_scratch));   // the array's address ON THE CURRENT DEVICE
    if (!sym) return fail(WT_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    nd.num_sms = sms;
    nd.ptr = (unsigned char*)sym;
  }
  *out = &nd;
  return WT_OK;
}

template <int BN>
static int noalloc_launch_gemm(const NoAllocDevice& nd, cudaStream_t st, const CUtensorMap& ma, const CUtensorMap& mw,
                               wt::GemmBf16Params& p) {
  using S = wt::GemmSmem<BN>;
  const int n_tiles = p.tiles_per_batch * p.n_batches * p.n_tiles_n;
  const int grid = n_tiles < nd.num_sms ? n_tiles : nd.num_sms;
  wt::gemm_bf16_sm100_kernel<BN><<<grid, wt::kGemmThreads, S::kTotal, st>>>(ma, mw, ma, ma, p);   // no C / addend maps
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(WT_ERR_CUDA, "launch of gemm_bf16_sm100_kernel failed: %s", cudaGetErrorString(e));
  return WT_OK;
}

extern "C" int wt_gemm_argmax_noalloc(void* cuda_stream, const void* A, const void* W, const float* bias, int M, int N,
                                      int K, int32_t* ids_out) {
  if (!A || !W || !ids_out || M < 1 || N < 8 || K < 8 || (K % 8) || (N % 8))
    return fail(WT_ERR_INVALID, "bad GEMM arguments (M >= 1; N, K multiples of 8; A, W, ids_out device pointers)");
  const int bn = pick_bn(M, N);
  const int n_tiles_n = (N + bn - 1) / bn;
  if ((long long)M * n_tiles_n > kNoAllocMaxPartials)
    return fail(WT_ERR_INVALID,
                "wt_gemm_argmax_noalloc: M * ceil(N / %d) = %lld partial winners exceed the static scratch of %d (%zu bytes); "
                "use the engine's allocating path for this size",
                bn, (long long)M * n_tiles_n, kNoAllocMaxPartials, kNoAllocScratchBytes);
  int dev = 0;
  WT_CUDA_OK(cudaGetDevice(&dev));
  NoAllocDevice* nd = nullptr;
  WT_TRY(noalloc_device_setup(dev, &nd));
  cudaStream_t st = (cudaStream_t)cuda_stream;
  CUtensorMap ma, mw;
  {
    const uint64_t ad[3] = {(uint64_t)K, (uint64_t)M, 1}, as[2] = {(uint64_t)K * 2, (uint64_t)K * 2 * (uint64_t)M};
    const uint32_t abox[3] = {64, 128, 1};
    WT_TRY(make_map_bf16(&ma, A, 3, ad, as, abox));
    const uint64_t wd[2] = {(uint64_t)K, (uint64_t)N}, ws[1] = {(uint64_t)K * 2};
    const uint32_t wbox[2] = {64, (uint32_t)bn};
    WT_TRY(make_map_bf16(&mw, W, 2, wd, ws, wbox));
  }
  wt::GemmBf16Params p{};
  p.bias = bias;
  p.argmax_part = reinterpret_cast<unsigned long long*>(nd->ptr + kNoAllocHeaderBytes);
  p.sched = reinterpret_cast<int*>(nd->ptr);
  p.cmap = RowMap{0, M, N};
  p.add_map = RowMap{0, 1, 0};
  p.M = M;
  p.N = N;
  p.K = K;
  p.rows_per_batch = M;
  p.tiles_per_batch = (M + 127) / 128;
  p.n_batches = 1;
  p.n_tiles_n = n_tiles_n;
  p.flags = wt::EPI_ARGMAX | (bias ? wt::EPI_BIAS : 0);
  p.taps = 1;
  p.kblocks_per_tap = (K + 63) / 64;
  // the scheduler words must start at zero whatever an earlier (possibly failed) call left there
  WT_CUDA_OK(cudaMemsetAsync(nd->ptr, 0, kNoAllocHeaderBytes, st));
  switch (bn) {
    case 64: WT_TRY(noalloc_launch_gemm<64>(*nd, st, ma, mw, p)); break;
    case 128: WT_TRY(noalloc_launch_gemm<128>(*nd, st, ma, mw, p)); break;
    case 192: WT_TRY(noalloc_launch_gemm<192>(*nd, st, ma, mw, p)); break;
    case 256: WT_TRY(noalloc_launch_gemm<256>(*nd, st, ma, mw, p)); break;
    default: return fail(WT_ERR_INVALID, "bad BN %d", bn);
  }
  wt::noalloc_argmax_rows_kernel<<<M, 128, 0, st>>>(p.argmax_part, n_tiles_n, ids_out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(WT_ERR_CUDA, "launch of noalloc_argmax_rows_kernel failed: %s", cudaGetErrorString(e));
  return WT_OK;
}
