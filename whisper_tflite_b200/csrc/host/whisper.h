// Host-side mirror of the reference's public C++ API (/root/reference/whisper.tflite/whisper.h) over the
// B200 C ABI (include/whisper_b200.h). Same namespace, type names, constants, constructor arguments,
// return types and error behaviour as the reference, so app/encdec.cpp, app/minimal.cpp and the JNI bridge
// (bindings/java/whisper.tflite.cpp) build against it unchanged. What differs is only what the reference
// could not keep once tflite::Interpreter is gone:
//   * Atom holds a wt_engine handle (device-resident weights + plan) instead of FlatBufferModel +
//     InterpreterBuilder + Interpreter (whisper.h:128-138);
//   * Encoder::forward returns EncoderOut (a count of clips whose states are resident on the device)
//     instead of std::tuple<TfLiteTensor*, float*> (whisper.h:143) — TfLite types do not exist here;
//   * model files: `<prefix>.b200w` in place of `<prefix>.tflite` / `.encoder.tflite` / `.decoder.tflite`
//     (whisper.cpp:669, :743-744); the vocab+filters blob is read exactly as before;
//   * ADDED (the reference API is batch-1, whisper.cpp:368): Engine::transcribe_batch, which is what the
//     batch-64 metric drives; Engine::transcribe_batch_i16 (16-bit samples as a WAV holds them, converted on the
//     device); Engine::set_language / set_task / set_omit_special_tokens for what the reference hard-codes
//     (language "de": whisper.cpp:327; omit_special_tokens = false: :733, :766). Defaults = the reference's.
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "mmap_file.h"
#include "wav_util.h"

struct wt_engine;  // include/whisper_b200.h

// Fatal model/device errors keep the reference's behaviour (whisper.h:13-17): message + exit(1).
#define WHISPER_B200_CHECK(x)                                \
  if (!(x)) {                                                \
    fprintf(stderr, "Error at %s:%d\n", __FILE__, __LINE__); \
    exit(1);                                                 \
  }

namespace whisper {

// whisper.h:27-32 — greedy IDs of whisper-tiny.en on LibriSpeech-dummy[0]; needs the real weights to reproduce.
static constexpr int kNumGoldenGeneratedIDs = 21;
static constexpr int kGoldenGeneratedIDs[kNumGoldenGeneratedIDs] = {50257, 50362, 1770, 13,  2264, 346,  353,
                                                                    318,   262,   46329, 286, 262,  3504, 6097,
                                                                    11,    290,   356,  389, 9675, 284,  7062};

static constexpr int kSampleRate = 16000;  // whisper.h:34-39
static constexpr int kNFFT = 400;
static constexpr int kNMEL = 80;
static constexpr int kHopLength = 160;
static constexpr int kChunkSize = 30;
static constexpr int kMelLen = 3000;
static constexpr int kVocabEnSize = 51864;  // whisper.h:41-42
static constexpr int kVocabMultilingualSize = 51865;

// Special-token ids of the English vocabulary; transform_vocab_multilingual shifts them (whisper.h:44-94).
struct Vocab {
  std::map<int, std::string> id_to_token;
  int n_vocab = 51864;
  int token_eot = 50256;
  int token_sot = 50257;
  int token_translate = 50358;
  int token_transcribe = 50359;
  int token_prev = 50360;
  int token_solm = 50361;
  int token_not = 50362;
  int token_beg = 50363;
};

struct Filters {  // whisper.h:96-101 — data[j * n_fft + k]
  int n_mel;
  int n_fft;
  std::vector<float> data;
};

struct Mel {  // whisper.h:103-107 — data[j * n_len + i]
  int n_len;
  int n_mel;
  std::vector<float> data;
};

// whisper.h:121-123. Same arguments and result layout; the frames are computed by the batched sm_100a
// frontend (wt_log_mel) on the current device. `sample_rate` and `n_threads` are accepted and unused (the
// reference ignores sample_rate too: whisper.cpp:110). Returns false where the reference cannot fail: no
// usable device, fft_size/fft_step other than 400/160, n_samples outside (0, 480000].
bool log_mel_spectrogram(const float* samples, int n_samples, int sample_rate, int fft_size, int fft_step, int n_mel,
                         int n_threads, Filters& filters, Mel& mel);

void transform_vocab_multilingual(Vocab& vocab);

// Device-resident model: what Atom's interpreter was to the reference (whisper.h:128-138).
struct Atom {
 public:
  // `path` is the weight file; precision is wt_precision (0 bf16 tcgen05 path, 1 fp32 verification).
  explicit Atom(const std::string& path, int max_batch = 1, int precision = -1, int device = -1);
  ~Atom();
  Atom(const Atom&) = delete;
  Atom& operator=(const Atom&) = delete;
  wt_engine* handle() const { return engine_; }
  int max_batch() const { return max_batch_; }
  void set_filters(const Filters& filters);

 private:
  wt_engine* engine_ = nullptr;
  int max_batch_ = 1;
};

struct EncoderOut {
  int n_clips = 0;  // encoder states + cross-attention K/V of this many clips are resident in the engine
};

struct Encoder {
 public:
  explicit Encoder(std::shared_ptr<Atom> atom) : atom_(std::move(atom)) {}
  // whisper.cpp:285-309: mel -> input 0, Invoke. Here: H2D of the mel, conv stem + blocks on the device.
  EncoderOut forward(const whisper::Mel& mel);
  // the engine-resident mel of the last wt_log_mel call (skips the mel round trip through the host)
  EncoderOut forward_resident(int n_clips);
  // host copy of the states [n_clips][1500][d_model] f32 (parity / debugging)
  std::vector<float> states(const EncoderOut& out);

 private:
  std::shared_ptr<Atom> atom_;
};

// What the reference fixes at compile time, as run-time options with the reference's values as defaults.
struct DecodeOptions {
  std::string language = "de";       // whisper.cpp:327 (language tag token = 50259 + language_id(language))
  bool translate = false;            // task token: vocab.token_transcribe (:336) or vocab.token_translate
  bool omit_special_tokens = false;  // decode(..., omit_special_tokens) at whisper.cpp:733, :766
};

struct Decoder {
 public:
  Decoder(std::shared_ptr<Atom> atom, const whisper::Vocab& vocab) : atom_(std::move(atom)), vocab_(vocab) {}
  // whisper.cpp:314-403 for clip 0; prompt + generated ids, <= 31 entries.
  std::vector<int64_t> forward(EncoderOut encoder_out);
  std::vector<std::vector<int64_t>> forward_batch(EncoderOut encoder_out);
  // [sot, 50259 + language_id(language), transcribe | translate, notimestamps] (whisper.cpp:327-339; "de" and
  // transcribe unless set_options changed them)
  std::vector<int64_t> prompt() const;
  void set_options(const DecodeOptions& o) { options_ = o; }

 private:
  std::shared_ptr<Atom> atom_;
  const whisper::Vocab& vocab_;
  DecodeOptions options_;
};

struct Engine {
  virtual std::string transcribe(std::vector<float>& samples) = 0;
  virtual std::string transcribe(const char* waveFile) = 0;
  // ADDED: one call for many windows; every vector is resized to 480000 like the single-clip call does.
  virtual std::vector<std::string> transcribe_batch(std::vector<std::vector<float>>& clips) = 0;
  // ADDED: the same on 16-bit PCM (what wav_read_legacy / wav_read read before their `/ INT16_MAX`, wav_util.cpp:63-76):
  // clips are resized to 480000 samples, the conversion runs in the mel kernel's staging load.
  virtual std::vector<std::string> transcribe_batch_i16(std::vector<std::vector<int16_t>>& clips) = 0;
  virtual ~Engine() = default;

  // ADDED: language / task / detokeniser options (defaults = the reference's hard-coded choices).
  // set_language returns false (and changes nothing) for a code that is not in language_meta.
  bool set_language(const std::string& code);
  void set_task(bool translate) { options_.translate = translate; on_options(); }
  void set_omit_special_tokens(bool omit) { options_.omit_special_tokens = omit; on_options(); }
  const DecodeOptions& options() const { return options_; }

 protected:
  virtual void on_options() {}
  DecodeOptions options_;
};

// Upper bound on clips per device call for engines built through the reference constructors; override with the
// environment variable WHISPER_B200_MAX_BATCH (activations are sized for it at construction).
int default_max_batch();

struct Monolith : public Engine {
 public:
  Monolith(const std::string& model_prefix, const std::string& vocab_path, bool multilingual);
  std::string transcribe(std::vector<float>& samples) final;
  std::string transcribe(const char* waveFile) final;
  std::vector<std::string> transcribe_batch(std::vector<std::vector<float>>& clips) final;
  std::vector<std::string> transcribe_batch_i16(std::vector<std::vector<int16_t>>& clips) final;
  // ADDED: token ids of the fused PCM -> ids call (the monolithic graph's int32 output, whisper.cpp:727-732)
  std::vector<std::vector<int32_t>> transcribe_tokens(std::vector<std::vector<float>>& clips);
  std::vector<std::vector<int32_t>> transcribe_tokens_i16(std::vector<std::vector<int16_t>>& clips);

 private:
  std::vector<int64_t> prompt() const;
  std::vector<int16_t> staging16_;
  MmapFile vocab_file_;
  Vocab vocab_;
  Filters filters_;
  std::shared_ptr<Atom> whisper_;
  bool multilingual_;
  std::vector<float> staging_;
};

struct EncDec : public Engine {
 public:
  EncDec(const std::string& model_prefix, const std::string& vocab_path, bool multilingual);
  std::string transcribe(std::vector<float>& samples) final;
  std::string transcribe(const char* waveFile) final;
  std::vector<std::string> transcribe_batch(std::vector<std::vector<float>>& clips) final;
  std::vector<std::string> transcribe_batch_i16(std::vector<std::vector<int16_t>>& clips) final;

 protected:
  void on_options() final { decoder_.set_options(options_); }

 private:
  std::vector<int16_t> staging16_;
  MmapFile vocab_file_;
  Vocab vocab_;
  Filters filters_;
  std::shared_ptr<Atom> atom_;
  Encoder encoder_;
  Decoder decoder_;
  std::vector<float> staging_;
};

enum class EngineType {
  Monolith = 0,
  EncDec = 1,
};

using LangKey = std::pair<std::string, std::string>;
extern std::vector<LangKey> language_meta;
int language_id(const std::string& code);  // index into language_meta; language_meta.size() when unknown
const std::string& lang_code(size_t id);

// Parser of the vocab+filters blob (layout: whisper.h:214-235; `head` points just past the u64 size prefix,
// whisper.cpp:676-681): u32 magic (unchecked) | i32 n_mel | i32 n_fft | f32[n_mel*n_fft] | i32 n_vocab |
// {u32 len, bytes}*n_vocab; ids in [n_vocab, 51864 + multilingual) get synthesised names.
struct Reader {
 public:
  explicit Reader(const char* head, bool multilingual) : head_(head), multilingual_(multilingual) {}
  void read(Filters& filters, Vocab& vocab);

 private:
  static const char* read_filters(Filters& filters, const char* head);
  static const char* read_vocab(Vocab& vocab, bool multilingual, const char* head);
  const char* head_;
  bool multilingual_;
};

std::string remove_extra_spaces(const std::string& input);

template <class Int>
std::string decode(const Vocab& vocab, const Int* begin, const Int* end, bool omit_special_tokens);
std::string decode(const Vocab& vocab, const std::vector<int64_t>& generated, bool omit_special_tokens);

// Raw owning pointer, nullptr for an unknown type (whisper.cpp:778-790); the caller deletes (JNI destroy).
Engine* create_engine(EngineType type, const char* model_prefix, const char* vocab_path, bool multilingual);

}  // namespace whisper
