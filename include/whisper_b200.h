/* whisper_b200.h — C ABI of the B200-native Whisper inference path.
 *
 * This is the drop-in boundary: the host-side C++ mirror of the reference's whisper.h
 * (whisper_tflite_b200/csrc/host/whisper.h: Engine / EncDec / Monolith / create_engine /
 * log_mel_spectrogram) calls ONLY these entry points, in the places where the reference calls
 * into tflite::Interpreter. Plain pointers and sizes; no C++ or torch types; every function
 * returns a wt_status, never throws, never exits. One engine per device; calls on one engine
 * must be serialised by the caller (same contract as the reference, whose Java layer holds a lock
 * around the engine: io/github/jerinphilip/whisper/Whisper.java:109-113).
 *
 * Reference interface each entry replaces (paths relative to /root/reference):
 *   wt_engine_create / _destroy   Atom::Atom — FlatBufferModel::BuildFromFile + InterpreterBuilder +
 *                                 AllocateTensors (whisper.tflite/whisper.cpp:261-271); Encoder/Decoder
 *                                 ctors (:284, :311-312); `delete engine` in JNI destroy
 *                                 (bindings/java/whisper.tflite.cpp:36-42)
 *   wt_set_mel_filters            Filters filled by Reader::read_filters (whisper.cpp:519-545) and passed to
 *                                 log_mel_spectrogram (whisper.h:121-123)
 *   wt_log_mel                    whisper::log_mel_spectrogram (whisper.cpp:109-216)
 *   wt_encode                     Encoder::forward: memcpy mel -> typed_input_tensor<float>(0), Invoke,
 *                                 output_tensor(0) (whisper.cpp:285-309)
 *   wt_decode_greedy              Decoder::forward: encoder states -> input 0, prompt, ResizeInputTensor /
 *                                 AllocateTensors / Invoke / argmax loop (whisper.cpp:314-403)
 *   wt_transcribe_tokens          Monolith::transcribe's mel -> Invoke -> typed_output_tensor<int>(0)
 *                                 (whisper.cpp:692-732) and EncDec::transcribe's mel -> encoder -> decoder
 *                                 (whisper.cpp:752-765), batched
 */
#ifndef WHISPER_B200_H_
#define WHISPER_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WT_API __attribute__((visibility("default")))

typedef struct wt_engine wt_engine;

typedef enum wt_status {
  WT_OK = 0,
  WT_ERR_INVALID = 1, /* bad argument */
  WT_ERR_CUDA = 2,    /* CUDA runtime/driver error (no device, launch failure, OOM) */
  WT_ERR_IO = 3,      /* weight file missing/unreadable */
  WT_ERR_FORMAT = 4,  /* weight file malformed / unsupported dims */
  WT_ERR_STATE = 5    /* call order (e.g. decode before encode, filters not set) */
} wt_status;

typedef enum wt_precision {
  WT_PRECISION_BF16 = 0,       /* bf16 operands, f32 accumulate: tcgen05/TMEM GEMMs (production) */
  WT_PRECISION_FP32_VERIFY = 1 /* f32 operands and accumulators end to end: token-identity checks */
} wt_precision;

typedef enum wt_mem { WT_HOST = 0, WT_DEVICE = 1 } wt_mem;

typedef struct wt_config {
  int32_t struct_size; /* = sizeof(wt_config) */
  int32_t precision;   /* wt_precision */
  int32_t max_batch;   /* clips per call the engine is sized for (>= 1) */
  int32_t keep_logits; /* != 0: keep the f32 logits of every decode step (debug/parity; costs B*27*V*4 bytes) */
  int32_t decoder_cluster; /* CTAs per decoder cluster: 0 = default (16: lowest latency of a single call), 8 or 4 = fewer SMs
                              per batch (a pool of engines keeps more batches in flight: highest throughput per GPU) */
} wt_config;

typedef struct wt_dims {
  int32_t n_mels, n_audio_ctx, d_model, n_heads, n_enc_layers, n_dec_layers, n_vocab, n_text_ctx;
} wt_dims;

#define WT_N_SAMPLES 480000 /* kSampleRate * kChunkSize, whisper.h:34,38 */
#define WT_N_FRAMES 3000    /* kMelLen, whisper.h:39 */
#define WT_N_MEL 80         /* kNMEL, whisper.h:36 */
#define WT_N_FREQ 201       /* 1 + kNFFT/2, whisper.cpp:129 */
#define WT_MAX_TOKENS 31    /* 4 prompt + 27 generated, whisper.cpp:364-367 */

/* Loads `weights_path` (<prefix>.b200w, see whisper_tflite_b200/synth.py) onto `device`, builds the execution
 * plan and returns a handle. weights_path == NULL creates a frontend-only engine (wt_log_mel works,
 * encode/decode return WT_ERR_STATE). */
WT_API int wt_engine_create(const wt_config* cfg, const char* weights_path, int device, wt_engine** out);
WT_API void wt_engine_destroy(wt_engine* e);
WT_API int wt_engine_dims(const wt_engine* e, wt_dims* out);

/* filters: [n_mel=80][n_freq=201] f32 row-major, host memory (Filters.data, whisper.h:96-101). */
WT_API int wt_set_mel_filters(wt_engine* e, const float* filters, int n_mel, int n_freq);

/* Work is enqueued on `cuda_stream` (a cudaStream_t; NULL = the engine's own stream). Calls taking only
 * WT_DEVICE buffers return after enqueueing; calls touching a WT_HOST buffer synchronise before returning. Changing the
 * stream waits for the work still in flight on the previous one (an engine's buffers belong to one stream at a time). */
WT_API int wt_set_stream(wt_engine* e, void* cuda_stream);

/* Clips per thread-block cluster of the persistent greedy decoder (1..16; 0 = automatic: spread the batch over every
 * cluster that is co-resident on an otherwise idle GPU — lowest latency of a single call). A serving loop that keeps
 * several engines busy on one GPU (batches alternating between engines on separate streams, so one batch's decoder
 * overlaps the next batch's mel + encoder) sets 16: the decoder then occupies the fewest SMs per batch.
 * New in this library (the reference decodes one clip per call: whisper.cpp:363-402). */
WT_API int wt_set_decoder_group(wt_engine* e, int clips_per_cluster);
WT_API int wt_synchronize(wt_engine* e);

/* pcm [n_clips][n_samples] f32 -> mel_out [n_clips][80][n_samples/160] f32 (reference layout j*n_len+i).
 * n_samples <= 480000 per clip. mel_out may be NULL (result stays resident for wt_encode(mel=NULL)). */
WT_API int wt_log_mel(wt_engine* e, const float* pcm, int n_clips, int n_samples, int pcm_mem, float* mel_out,
                      int out_mem);

/* mel [n_clips][80][3000] f32 (NULL = use the engine-resident mel of the last wt_log_mel) ->
 * states_out [n_clips][1500][d_model] f32 (may be NULL). Encoder states and the cross-attention K/V stay
 * resident for wt_decode_greedy. */
WT_API int wt_encode(wt_engine* e, const float* mel, int n_clips, int mel_mem, float* states_out, int out_mem);

/* Greedy loop of Decoder::forward over the resident encoder states. prompt: host int64[prompt_len].
 * ids_out [n_clips][max_total_tokens] int32 (prompt included, -1 padded), lens_out [n_clips] int32. */
WT_API int wt_decode_greedy(wt_engine* e, int n_clips, const int64_t* prompt, int prompt_len, int max_total_tokens,
                            int64_t eot, int32_t* ids_out, int32_t* lens_out, int out_mem);

/* pcm -> token ids in one call (mel + encoder + decoder; nothing but ids/lens leaves the device). */
WT_API int wt_transcribe_tokens(wt_engine* e, const float* pcm, int n_clips, int n_samples, int pcm_mem,
                                const int64_t* prompt, int prompt_len, int max_total_tokens, int64_t eot,
                                int32_t* ids_out, int32_t* lens_out, int out_mem);

/* int16 PCM ingest: the same two calls on 16-bit samples as a WAV file holds them. The
 * conversion the reference does on the host — static_cast<float>(s) / static_cast<float>(INT16_MAX), wav_util.cpp:70-73
 * (wav_read_legacy) and :120-122 (wav_read, mono) — happens in the mel kernel's staging load, so the mel is bit-equal to
 * wt_log_mel on the converted samples and a host batch costs 0.96 MB per clip on PCIe instead of 1.92 MB. */
WT_API int wt_log_mel_i16(wt_engine* e, const int16_t* pcm, int n_clips, int n_samples, int pcm_mem, float* mel_out,
                          int out_mem);
WT_API int wt_transcribe_tokens_i16(wt_engine* e, const int16_t* pcm, int n_clips, int n_samples, int pcm_mem,
                                    const int64_t* prompt, int prompt_len, int max_total_tokens, int64_t eot,
                                    int32_t* ids_out, int32_t* lens_out, int out_mem);

/* keep_logits engines only: f32 logits of decode step `step` (0 = first generated token), [n_clips][n_vocab]. */
WT_API int wt_get_step_logits(wt_engine* e, int step, int n_clips, float* out, int out_mem);

/* Number of kernel launches this engine has enqueued since creation (bench.py's gpu_launches). */
WT_API int64_t wt_launch_count(const wt_engine* e);

/* Per-stage CUDA-event profiler (events on the launching stream around every kernel launch). Enable, run
 * calls, then read: ms_out / launches_out are arrays of wt_stage_count() entries accumulated since enable. */
WT_API int wt_profile_enable(wt_engine* e, int enable);
WT_API int wt_profile_read(wt_engine* e, float* ms_out, int32_t* launches_out, int n_stages);
WT_API int wt_stage_count(void);
WT_API const char* wt_stage_name(int stage);

/* Thread-local description of the last failure on this thread. */
WT_API const char* wt_last_error(void);
WT_API const char* wt_version(void);

/* Stand-alone tcgen05 GEMM entry used by the kernel-level parity tests and the micro-benchmark:
 * C[M][N] (bf16) = epi(A[M][K] . W[N][K]^T), all device pointers, bf16 row-major, f32 bias.
 * flags: 1 bias, 2 GELU. */
WT_API int wt_debug_gemm_bf16(void* cuda_stream, const void* A, const void* W, const float* bias, void* C, int M, int N,
                              int K, int flags);

/* Kernel-level test hooks (tests/test_gpu_kernels.py compares them with fp32 torch references).
 * wt_debug_gemm_ex: the tcgen05 GEMM with every addressing / epilogue feature the engine uses. A is a bf16 tensor
 * {a_dims[0] inner, a_dims[1] rows, a_dims[2] batches} with byte strides a_strides[0..1]; K is walked as `taps` groups
 * (tap t reads A at column tap_c0[t], row offset tap_row[t]; W column t*tap_k) — taps = 1 is a plain GEMM, 3 the k=3
 * convolutions. Output row m lives at C + (m / c_rows) * c_batch_stride + (m % c_rows) * c_row_stride (elements);
 * the addend (flags & 4) likewise. flags: 1 bias, 2 GELU, 4 add, 16 f32 output, 32 fused (max, last index) into
 * argmax_part [M][ceil(N/bn)] u64, 64 K/V scatter into kv_out (kv_chunked != 0: the decoder's chunk stream with
 * XOR-swizzled 16-byte segments; else [clip][H][kv_S][64] K into kv_out, V into kv_out_v). All pointers device. */
typedef struct wt_debug_gemm_args {
  int32_t struct_size;
  const void* A; uint64_t a_dims[3]; uint64_t a_strides[2];
  int32_t rows_per_batch, n_batches;
  const void* W; const float* bias; int32_t N, K_total;
  int32_t taps, tap_k, tap_c0[3], tap_row[3];
  void* C; int64_t c_batch_stride; int32_t c_rows, c_row_stride;
  const void* addend; int64_t add_batch_stride; int32_t add_rows, add_row_stride;
  uint64_t* argmax_part;
  void *kv_out, *kv_out_v; int32_t kv_S, kv_H, kv_chunked;
  int32_t flags, force_bn;
} wt_debug_gemm_args;
WT_API int wt_debug_gemm_ex(void* cuda_stream, const wt_debug_gemm_args* a);

/* The encoder attention kernel alone: qkv [n_clips][S][3*H*64] bf16 (q | k | v) -> out [n_clips][S][H*64] bf16. */
WT_API int wt_debug_attention_bf16(void* cuda_stream, const void* qkv, void* out, int n_clips, int S, int H);

/* No-allocation entry point. It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * ids_out[m] (int32, device) = index of the maximum of row m of A[M][K] . W[N][K]^T (+ bias[N], may be NULL), with the
 * argmax rule of whisper.cpp:346-361 (last index among equal maxima; NaN only wins at index 0): the decoder's fused
 * tied-logits + argmax product. A, W bf16 row-major device pointers, N and K multiples of 8. The call allocates
 * nothing (no cudaMalloc / cudaFree, no allocating library routine): the GEMM's tile-scheduler words and its
 * per-column-tile partial winners live in one static __device__ array of 1 MiB + 256 B that every GPU loading the
 * library holds until exit. Limit: M * ceil(N / bn) <= 131072 partial winners (bn = 64 or 128 for M <= 128, else the
 * largest of 256 / 192 / 128 dividing N) — e.g. 64 rows x 51 865 columns uses 26 048; beyond it the call fails with
 * WT_ERR_INVALID and a message naming the limit. Asynchronous on cuda_stream. All calls share the scratch, which keeps
 * the previous call's contents (the words that must start at zero are cleared on the stream first): calls that overlap
 * on the GPU — different streams or host threads — race on it and must be serialised by the caller. */
WT_API int wt_gemm_argmax_noalloc(void* cuda_stream, const void* A, const void* W, const float* bias, int M, int N,
                                  int K, int32_t* ids_out);

/* Debug / test hook: 64-bit order-independent digest of a device-resident buffer as the last call on this engine left
 * it, over the first n_clips clips. what: "mel_raw", "mel_tm", "conv1", "x", "qkv", "att", "hidden", "enc_out",
 * "cross_kv<layer>", "tokens". Equal bytes give equal digests: the bitwise-reproducibility tests compare these across
 * repeats and engines (a window must map to ONE id list: whisper.cpp:367-400). Synchronises the engine's stream. */
WT_API int wt_debug_hash(wt_engine* e, const char* what, int n_clips, uint64_t* out);

/* Engines created with WT_DEC_TRACE=1 in the environment: nanoseconds per phase kind of the persistent decoder as
 * seen by CTA 0, accumulated over launches. out32[k] = phase k total (0 init, 1..8 the layer phases, 9 logits,
 * 10 finalize+embed), out32[16+k] = the part of it spent waiting in the grid barrier. */
WT_API int wt_debug_decoder_trace(wt_engine* e, uint64_t* out32, int reset);

#ifdef __cplusplus
}
#endif
#endif /* WHISPER_B200_H_ */
