// Host side of the drop-in: the reference's whisper.cpp (/root/reference/whisper.tflite/whisper.cpp) minus every
// tflite::Interpreter call, which become calls into the B200 C ABI (include/whisper_b200.h). No numerics live
// here: mel, encoder, decoder and the greedy argmax all run on the device; this file parses the vocab blob,
// builds the prompt, moves PCM in and token ids out, and detokenises.
#include "whisper.h"

#include <algorithm>
#include <cassert>
#include <cstring>
#include <iostream>
#include <mutex>

#include "whisper_b200.h"

namespace whisper {

namespace {

int env_int(const char* name, int fallback) {
  const char* v = std::getenv(name);
  if (!v || !*v) return fallback;
  char* end = nullptr;
  const long x = std::strtol(v, &end, 10);
  return (end && *end == 0) ? static_cast<int>(x) : fallback;
}

int env_precision() {
  const char* v = std::getenv("WHISPER_B200_PRECISION");
  if (v && (std::strcmp(v, "fp32") == 0 || std::strcmp(v, "f32") == 0 || std::strcmp(v, "verify") == 0))
    return WT_PRECISION_FP32_VERIFY;
  return WT_PRECISION_BF16;
}

bool debug_enabled() { return std::getenv("DEBUG") != nullptr; }  // whisper.cpp:274, :387

[[noreturn]] void die(const char* what) {
  fprintf(stderr, "whisper_b200: %s: %s\n", what, wt_last_error());
  exit(1);
}

// Frontend-only engine behind the free function log_mel_spectrogram (which has no engine argument in the
// reference API). One per process, created on first use, re-armed when the caller's filterbank changes.
struct Frontend {
  std::mutex mu;
  wt_engine* engine = nullptr;
  std::vector<float> filters;
  int n_mel = 0;
  ~Frontend() {
    if (engine) wt_engine_destroy(engine);
  }
};
Frontend& frontend() {
  static Frontend f;
  return f;
}

const int64_t* prompt_ptr(const std::vector<int64_t>& p) { return p.data(); }

}  // namespace

int default_max_batch() { return std::max(1, env_int("WHISPER_B200_MAX_BATCH", 64)); }

// ------------------------------------------------------------------------------------------------
// log-mel (whisper.cpp:109-216) — device computed
// ------------------------------------------------------------------------------------------------
bool log_mel_spectrogram(const float* samples, int n_samples, int /*sample_rate*/, int fft_size, int fft_step, int n_mel,
                         int /*n_threads*/, Filters& filters, Mel& mel) {
  if (fft_size != kNFFT || fft_step != kHopLength) return false;
  if (!samples || n_samples < fft_step || n_samples > kSampleRate * kChunkSize) return false;
  if (filters.n_mel != n_mel || filters.n_fft != 1 + fft_size / 2 ||
      filters.data.size() != static_cast<size_t>(n_mel) * filters.n_fft)
    return false;
  Frontend& f = frontend();
  std::lock_guard<std::mutex> lock(f.mu);
  if (!f.engine) {
    wt_config cfg{};
    cfg.struct_size = sizeof(cfg);
    cfg.precision = WT_PRECISION_BF16;
    cfg.max_batch = 1;
    if (wt_engine_create(&cfg, nullptr, std::max(0, env_int("WHISPER_B200_DEVICE", 0)), &f.engine) != WT_OK) {
      fprintf(stderr, "log_mel_spectrogram: %s\n", wt_last_error());
      return false;
    }
  }
  if (f.n_mel != n_mel || f.filters != filters.data) {
    if (wt_set_mel_filters(f.engine, filters.data.data(), n_mel, filters.n_fft) != WT_OK) {
      fprintf(stderr, "log_mel_spectrogram: %s\n", wt_last_error());
      return false;
    }
    f.filters = filters.data;
    f.n_mel = n_mel;
  }
  mel.n_mel = n_mel;
  mel.n_len = n_samples / fft_step;  // whisper.cpp:123
  mel.data.resize(static_cast<size_t>(mel.n_mel) * mel.n_len);
  if (wt_log_mel(f.engine, samples, 1, n_samples, WT_HOST, mel.data.data(), WT_HOST) != WT_OK) {
    fprintf(stderr, "log_mel_spectrogram: %s\n", wt_last_error());
    return false;
  }
  return true;
}

void transform_vocab_multilingual(Vocab& vocab) {  // whisper.cpp:218-226: translate/transcribe stay put
  vocab.n_vocab = kVocabMultilingualSize;
  for (int* id : {&vocab.token_eot, &vocab.token_sot, &vocab.token_prev, &vocab.token_solm, &vocab.token_not,
                  &vocab.token_beg})
    ++*id;
}

// ------------------------------------------------------------------------------------------------
// Atom / Encoder / Decoder
// ------------------------------------------------------------------------------------------------
Atom::Atom(const std::string& path, int max_batch, int precision, int device) : max_batch_(std::max(1, max_batch)) {
  wt_config cfg{};
  cfg.struct_size = sizeof(cfg);
  cfg.precision = precision < 0 ? env_precision() : precision;
  cfg.max_batch = max_batch_;
  cfg.keep_logits = 0;
  const int dev = device < 0 ? std::max(0, env_int("WHISPER_B200_DEVICE", 0)) : device;
  if (wt_engine_create(&cfg, path.c_str(), dev, &engine_) != WT_OK) die(path.c_str());  // whisper.cpp:268-270
}

Atom::~Atom() {
  if (engine_) wt_engine_destroy(engine_);
}

void Atom::set_filters(const Filters& filters) {
  if (wt_set_mel_filters(engine_, filters.data.data(), filters.n_mel, filters.n_fft) != WT_OK) die("mel filters");
}

EncoderOut Encoder::forward(const whisper::Mel& mel) {
  WHISPER_B200_CHECK(mel.n_mel == kNMEL && mel.n_len == kMelLen &&
                     mel.data.size() == static_cast<size_t>(kNMEL) * kMelLen);
  if (wt_encode(atom_->handle(), mel.data.data(), 1, WT_HOST, nullptr, WT_HOST) != WT_OK) die("Encoder::forward");
  return EncoderOut{1};
}

EncoderOut Encoder::forward_resident(int n_clips) {
  if (wt_encode(atom_->handle(), nullptr, n_clips, WT_DEVICE, nullptr, WT_DEVICE) != WT_OK) die("Encoder::forward");
  return EncoderOut{n_clips};
}

std::vector<float> Encoder::states(const EncoderOut& out) {
  wt_dims d{};
  if (wt_engine_dims(atom_->handle(), &d) != WT_OK) die("wt_engine_dims");
  std::vector<float> s(static_cast<size_t>(out.n_clips) * d.n_audio_ctx * d.d_model);
  // re-runs the encoder on the resident mel; states only leave the device when asked for
  if (wt_encode(atom_->handle(), nullptr, out.n_clips, WT_DEVICE, s.data(), WT_HOST) != WT_OK) die("Encoder::states");
  return s;
}

std::vector<int64_t> Decoder::prompt() const {
  // whisper.cpp:327: 50259 + language index; the reference hard-codes "de" and transcribe, which stay the defaults
  const int64_t target_lang_id = 50259 + language_id(options_.language);
  return {vocab_.token_sot, target_lang_id, options_.translate ? vocab_.token_translate : vocab_.token_transcribe,
          vocab_.token_not};
}

bool Engine::set_language(const std::string& code) {
  if (language_id(code) >= static_cast<int>(language_meta.size())) return false;  // unknown code (whisper.cpp:510-515)
  options_.language = code;
  on_options();
  return true;
}

std::vector<std::vector<int64_t>> Decoder::forward_batch(EncoderOut enc) {
  constexpr int kMaxTotal = WT_MAX_TOKENS;  // 4 prompt + 27 generated: whisper.cpp:364-367
  const std::vector<int64_t> p = prompt();
  std::vector<int32_t> ids(static_cast<size_t>(enc.n_clips) * kMaxTotal), lens(enc.n_clips);
  if (wt_decode_greedy(atom_->handle(), enc.n_clips, prompt_ptr(p), static_cast<int>(p.size()), kMaxTotal,
                       vocab_.token_eot, ids.data(), lens.data(), WT_HOST) != WT_OK)
    die("Decoder::forward");
  std::vector<std::vector<int64_t>> out(enc.n_clips);
  for (int b = 0; b < enc.n_clips; ++b) {
    const int32_t* row = ids.data() + static_cast<size_t>(b) * kMaxTotal;
    out[b].assign(row, row + lens[b]);
    if (debug_enabled())
      for (int i = static_cast<int>(p.size()); i < lens[b]; ++i) fprintf(stderr, "decode[%d]@clip%d = %d\n", i, b, row[i]);
  }
  return out;
}

std::vector<int64_t> Decoder::forward(EncoderOut encoder_out) {
  encoder_out.n_clips = 1;
  return forward_batch(encoder_out)[0];
}

// ------------------------------------------------------------------------------------------------
// language table (whisper.cpp:405-517): order defines the language-tag token ids 50259 + index
// ------------------------------------------------------------------------------------------------
namespace {
const char* const kLanguages[][2] = {
    {"en", "english"},    {"zh", "chinese"},   {"de", "german"},      {"es", "spanish"},    {"ru", "russian"},
    {"ko", "korean"},     {"fr", "french"},    {"ja", "japanese"},    {"pt", "portuguese"}, {"tr", "turkish"},
    {"pl", "polish"},     {"ca", "catalan"},   {"nl", "dutch"},       {"ar", "arabic"},     {"sv", "swedish"},
    {"it", "italian"},    {"id", "indonesian"}, {"hi", "hindi"},      {"fi", "finnish"},    {"vi", "vietnamese"},
    {"he", "hebrew"},     {"uk", "ukrainian"}, {"el", "greek"},       {"ms", "malay"},      {"cs", "czech"},
    {"ro", "romanian"},   {"da", "danish"},    {"hu", "hungarian"},   {"ta", "tamil"},      {"no", "norwegian"},
    {"th", "thai"},       {"ur", "urdu"},      {"hr", "croatian"},    {"bg", "bulgarian"},  {"lt", "lithuanian"},
    {"la", "latin"},      {"mi", "maori"},     {"ml", "malayalam"},   {"cy", "welsh"},      {"sk", "slovak"},
    {"te", "telugu"},     {"fa", "persian"},   {"lv", "latvian"},     {"bn", "bengali"},    {"sr", "serbian"},
    {"az", "azerbaijani"}, {"sl", "slovenian"}, {"kn", "kannada"},    {"et", "estonian"},   {"mk", "macedonian"},
    {"br", "breton"},     {"eu", "basque"},    {"is", "icelandic"},   {"hy", "armenian"},   {"ne", "nepali"},
    {"mn", "mongolian"},  {"bs", "bosnian"},   {"kk", "kazakh"},      {"sq", "albanian"},   {"sw", "swahili"},
    {"gl", "galician"},   {"mr", "marathi"},   {"pa", "punjabi"},     {"si", "sinhala"},    {"km", "khmer"},
    {"sn", "shona"},      {"yo", "yoruba"},    {"so", "somali"},      {"af", "afrikaans"},  {"oc", "occitan"},
    {"ka", "georgian"},   {"be", "belarusian"}, {"tg", "tajik"},      {"sd", "sindhi"},     {"gu", "gujarati"},
    {"am", "amharic"},    {"yi", "yiddish"},   {"lo", "lao"},         {"uz", "uzbek"},      {"fo", "faroese"},
    {"ht", "haitian creole"}, {"ps", "pashto"}, {"tk", "turkmen"},    {"nn", "nynorsk"},    {"mt", "maltese"},
    {"sa", "sanskrit"},   {"lb", "luxembourgish"}, {"my", "myanmar"}, {"bo", "tibetan"},    {"tl", "tagalog"},
    {"mg", "malagasy"},   {"as", "assamese"},  {"tt", "tatar"},       {"haw", "hawaiian"},  {"ln", "lingala"},
    {"ha", "hausa"},      {"ba", "bashkir"},   {"jw", "javanese"},    {"su", "sundanese"},  {"yue", "cantonese"},
};
std::vector<LangKey> make_language_meta() {
  std::vector<LangKey> v;
  for (const auto& row : kLanguages) v.emplace_back(row[0], row[1]);
  return v;
}
}  // namespace

std::vector<LangKey> language_meta = make_language_meta();

int language_id(const std::string& code) {
  for (size_t i = 0; i < language_meta.size(); ++i)
    if (language_meta[i].first == code) return static_cast<int>(i);
  return static_cast<int>(language_meta.size());
}

const std::string& lang_code(size_t id) { return language_meta[id].first; }

// ------------------------------------------------------------------------------------------------
// vocab + filters blob (whisper.cpp:519-611)
// ------------------------------------------------------------------------------------------------
namespace {
template <class T>
T take(const char*& head) {
  T v;
  std::memcpy(&v, head, sizeof(T));
  head += sizeof(T);
  return v;
}

std::string special_name(const Vocab& v, int id) {  // whisper.cpp:577-604
  if (id > v.token_beg) return "<|TT" + std::to_string(id - v.token_beg) + "|>";
  if (id == v.token_eot) return "<|endoftranscript|>";
  if (id == v.token_sot) return "<|startoftranscript_|>";
  if (id == v.token_prev) return "<|PREV|>";
  if (id == v.token_not) return "<|notimestamps|>";
  if (id == v.token_beg) return "<|timestampbegin|>";
  if (id == v.token_translate) return "<|translate|>";
  if (id == v.token_transcribe) return "<|transcribe|>";
  if (id > v.token_sot && id < v.token_translate) return "<|lang-" + lang_code(id - (v.token_sot + 1)) + "|>";
  return "<|e" + std::to_string(id) + "|>";
}
}  // namespace

const char* Reader::read_filters(Filters& filters, const char* head) {
  (void)take<uint32_t>(head);  // magic: read, never checked (whisper.cpp:521-529)
  filters.n_mel = take<int32_t>(head);
  filters.n_fft = take<int32_t>(head);
  const size_t n = static_cast<size_t>(filters.n_mel) * filters.n_fft;
  filters.data.resize(n);
  std::memcpy(filters.data.data(), head, n * sizeof(float));
  return head + n * sizeof(float);
}

const char* Reader::read_vocab(Vocab& vocab, bool multilingual, const char* head) {
  const int32_t n_vocab = take<int32_t>(head);
  vocab.n_vocab = n_vocab;
  printf("\nn_vocab:%d\n", static_cast<int>(n_vocab));  // whisper.cpp:555
  if (multilingual) transform_vocab_multilingual(vocab);
  for (int i = 0; i < n_vocab; ++i) {
    const uint32_t len = take<uint32_t>(head);
    vocab.id_to_token[i] = std::string(head, len).c_str();  // reference copies through a C string: stops at NUL
    head += len;
  }
  const int expected = kVocabEnSize + static_cast<int>(multilingual);
  for (int i = n_vocab; i < expected; ++i) vocab.id_to_token[i] = special_name(vocab, i);
  return head;
}

void Reader::read(Filters& filters, Vocab& vocab) {
  head_ = read_filters(filters, head_);
  head_ = read_vocab(vocab, multilingual_, head_);
}

// ------------------------------------------------------------------------------------------------
// detokenisation (whisper.cpp:613-665)
// ------------------------------------------------------------------------------------------------
std::string remove_extra_spaces(const std::string& input) {
  std::string out;
  out.reserve(input.size());
  char prev = 0;
  for (char c : input) {
    if (c != ' ' || prev != ' ') out.push_back(c);
    prev = c;
  }
  return out;
}

template <class Int>
std::string decode(const Vocab& vocab, const Int* begin, const Int* end, bool omit_special_tokens) {
  std::string surface;
  for (const Int* p = begin; p != end; ++p) {
    const int id = static_cast<int>(*p);
    if (!omit_special_tokens || id < vocab.token_eot) {
      const auto it = vocab.id_to_token.find(id);
      assert(it != vocab.id_to_token.end());
      if (it != vocab.id_to_token.end()) surface += it->second;
    }
    if (id == vocab.token_eot) break;
  }
  return surface;
}
template std::string decode(const Vocab&, const int*, const int*, bool);
template std::string decode(const Vocab&, const int64_t*, const int64_t*, bool);

std::string decode(const Vocab& vocab, const std::vector<int64_t>& generated, bool omit_special_tokens) {
  return decode(vocab, generated.data(), generated.data() + generated.size(), omit_special_tokens);
}

// ------------------------------------------------------------------------------------------------
// engines
// ------------------------------------------------------------------------------------------------
namespace {
constexpr int kWindow = kSampleRate * kChunkSize;  // 480000

const char* blob_head(const MmapFile& f) {
  // the first 8 bytes hold the payload size and are skipped (whisper.cpp:676-681, :746-747)
  return static_cast<const char*>(f.data()) + sizeof(uint64_t);
}

// copies `n` clips starting at `first` into one contiguous [n][480000] buffer, resizing each clip like
// samples.resize(kSampleRate * kChunkSize, 0) does (whisper.cpp:698, :753 — the caller's vectors are mutated)
const float* stage(std::vector<std::vector<float>>& clips, size_t first, size_t n, std::vector<float>& staging) {
  for (size_t i = 0; i < n; ++i) clips[first + i].resize(kWindow, 0.0f);
  if (n == 1) return clips[first].data();
  staging.resize(n * kWindow);
  for (size_t i = 0; i < n; ++i) std::memcpy(staging.data() + i * kWindow, clips[first + i].data(), kWindow * sizeof(float));
  return staging.data();
}
const int16_t* stage16(std::vector<std::vector<int16_t>>& clips, size_t first, size_t n, std::vector<int16_t>& staging) {
  for (size_t i = 0; i < n; ++i) clips[first + i].resize(kWindow, 0);
  if (n == 1) return clips[first].data();
  staging.resize(n * kWindow);
  for (size_t i = 0; i < n; ++i) std::memcpy(staging.data() + i * kWindow, clips[first + i].data(), kWindow * sizeof(int16_t));
  return staging.data();
}
}  // namespace

Monolith::Monolith(const std::string& model_prefix, const std::string& vocab_path, bool multilingual)
    : vocab_file_(vocab_path), multilingual_(multilingual) {
  Reader reader(blob_head(vocab_file_), multilingual);
  reader.read(filters_, vocab_);
  whisper_ = std::make_shared<Atom>(model_prefix + ".b200w", default_max_batch());
  whisper_->set_filters(filters_);
}

std::vector<int64_t> Monolith::prompt() const {
  // What HF generate() was forced to start with inside the monolithic graph (export/generate.py:24-30):
  // English-only models start [sot, notimestamps] (cf. kGoldenGeneratedIDs, whisper.h:27-32); multilingual ones
  // carry the language and task tags, for which the reference's only statement is Decoder::forward's prompt.
  if (!multilingual_) return {vocab_.token_sot, vocab_.token_not};
  return {vocab_.token_sot, 50259 + language_id(options_.language),
          options_.translate ? vocab_.token_translate : vocab_.token_transcribe, vocab_.token_not};
}

namespace {
// one device call per `cap` clips; `call` = wt_transcribe_tokens or its int16 twin on the staged batch
template <class Sample, class Stage, class Call>
std::vector<std::vector<int32_t>> monolith_tokens(std::vector<std::vector<Sample>>& clips, size_t cap, Stage stage_fn, Call call) {
  constexpr int kMaxTotal = WT_MAX_TOKENS;
  std::vector<std::vector<int32_t>> out(clips.size());
  std::vector<int32_t> ids(cap * kMaxTotal), lens(cap);
  for (size_t first = 0; first < clips.size(); first += cap) {
    const size_t n = std::min(cap, clips.size() - first);
    const Sample* pcm = stage_fn(clips, first, n);
    if (call(pcm, static_cast<int>(n), ids.data(), lens.data()) != WT_OK) {
      fprintf(stderr, "Monolith::transcribe: %s\n", wt_last_error());
      return {};  // soft failure -> empty result, like `return ""` at whisper.cpp:719-721
    }
    for (size_t i = 0; i < n; ++i) out[first + i].assign(ids.data() + i * kMaxTotal, ids.data() + i * kMaxTotal + lens[i]);
  }
  return out;
}
}  // namespace

std::vector<std::vector<int32_t>> Monolith::transcribe_tokens(std::vector<std::vector<float>>& clips) {
  const std::vector<int64_t> p = prompt();
  return monolith_tokens(
      clips, static_cast<size_t>(whisper_->max_batch()),
      [&](std::vector<std::vector<float>>& c, size_t first, size_t n) { return stage(c, first, n, staging_); },
      [&](const float* pcm, int n, int32_t* ids, int32_t* lens) {
        return wt_transcribe_tokens(whisper_->handle(), pcm, n, kWindow, WT_HOST, p.data(), static_cast<int>(p.size()),
                                    WT_MAX_TOKENS, vocab_.token_eot, ids, lens, WT_HOST);
      });
}

std::vector<std::vector<int32_t>> Monolith::transcribe_tokens_i16(std::vector<std::vector<int16_t>>& clips) {
  const std::vector<int64_t> p = prompt();
  return monolith_tokens(
      clips, static_cast<size_t>(whisper_->max_batch()),
      [&](std::vector<std::vector<int16_t>>& c, size_t first, size_t n) { return stage16(c, first, n, staging16_); },
      [&](const int16_t* pcm, int n, int32_t* ids, int32_t* lens) {
        return wt_transcribe_tokens_i16(whisper_->handle(), pcm, n, kWindow, WT_HOST, p.data(), static_cast<int>(p.size()),
                                        WT_MAX_TOKENS, vocab_.token_eot, ids, lens, WT_HOST);
      });
}

std::vector<std::string> Monolith::transcribe_batch(std::vector<std::vector<float>>& clips) {
  const auto tokens = transcribe_tokens(clips);
  std::vector<std::string> text(clips.size());
  for (size_t i = 0; i < tokens.size(); ++i)
    text[i] = decode(vocab_, tokens[i].data(), tokens[i].data() + tokens[i].size(), options_.omit_special_tokens);
  return text;
}

std::vector<std::string> Monolith::transcribe_batch_i16(std::vector<std::vector<int16_t>>& clips) {
  const auto tokens = transcribe_tokens_i16(clips);
  std::vector<std::string> text(clips.size());
  for (size_t i = 0; i < tokens.size(); ++i)
    text[i] = decode(vocab_, tokens[i].data(), tokens[i].data() + tokens[i].size(), options_.omit_special_tokens);
  return text;
}

std::string Monolith::transcribe(std::vector<float>& samples) {
  std::vector<std::vector<float>> one(1);
  one[0].swap(samples);
  std::vector<std::string> text = transcribe_batch(one);
  samples.swap(one[0]);  // the caller's vector comes back resized to 480000 (whisper.cpp:698)
  return text.empty() ? std::string() : text[0];
}

std::string Monolith::transcribe(const char* waveFile) {
  std::vector<float> pcmf32 = wav_read_legacy(waveFile);
  pcmf32.resize(kWindow, 0);
  return transcribe(pcmf32);
}

EncDec::EncDec(const std::string& model_prefix, const std::string& vocab_path, bool multilingual)
    : vocab_file_(vocab_path),
      atom_(std::make_shared<Atom>(model_prefix + ".b200w", default_max_batch())),
      encoder_(atom_),
      decoder_(atom_, vocab_) {
  Reader reader(blob_head(vocab_file_), multilingual);
  reader.read(filters_, vocab_);
  atom_->set_filters(filters_);
}

namespace {
// log_mel_spectrogram -> Encoder::forward -> Decoder::forward (whisper.cpp:756-765) per batch of `cap` clips; the mel
// and the encoder states stay on the device between the three calls. `mel` = wt_log_mel or its int16 twin.
template <class Sample, class Stage, class MelCall>
std::vector<std::string> encdec_batch(std::vector<std::vector<Sample>>& clips, size_t cap, Stage stage_fn, MelCall mel,
                                      Encoder& encoder, Decoder& decoder, const Vocab& vocab, bool omit_special) {
  std::vector<std::string> text(clips.size());
  for (size_t first = 0; first < clips.size(); first += cap) {
    const size_t n = std::min(cap, clips.size() - first);
    const Sample* pcm = stage_fn(clips, first, n);
    if (mel(pcm, static_cast<int>(n)) != WT_OK) {
      std::cerr << "Failed to compute mel_ spectrogram" << '\n';  // whisper.cpp:759-761
      return std::vector<std::string>(clips.size());
    }
    const EncoderOut enc = encoder.forward_resident(static_cast<int>(n));
    const auto generated = decoder.forward_batch(enc);
    for (size_t i = 0; i < n; ++i) text[first + i] = decode(vocab, generated[i], omit_special);
  }
  return text;
}
}  // namespace

std::vector<std::string> EncDec::transcribe_batch(std::vector<std::vector<float>>& clips) {
  return encdec_batch(
      clips, static_cast<size_t>(atom_->max_batch()),
      [&](std::vector<std::vector<float>>& c, size_t first, size_t n) { return stage(c, first, n, staging_); },
      [&](const float* pcm, int n) { return wt_log_mel(atom_->handle(), pcm, n, kWindow, WT_HOST, nullptr, WT_DEVICE); },
      encoder_, decoder_, vocab_, options_.omit_special_tokens);
}

std::vector<std::string> EncDec::transcribe_batch_i16(std::vector<std::vector<int16_t>>& clips) {
  return encdec_batch(
      clips, static_cast<size_t>(atom_->max_batch()),
      [&](std::vector<std::vector<int16_t>>& c, size_t first, size_t n) { return stage16(c, first, n, staging16_); },
      [&](const int16_t* pcm, int n) { return wt_log_mel_i16(atom_->handle(), pcm, n, kWindow, WT_HOST, nullptr, WT_DEVICE); },
      encoder_, decoder_, vocab_, options_.omit_special_tokens);
}

std::string EncDec::transcribe(std::vector<float>& samples) {
  std::vector<std::vector<float>> one(1);
  one[0].swap(samples);
  std::vector<std::string> text = transcribe_batch(one);
  samples.swap(one[0]);
  return text.empty() ? std::string() : text[0];
}

std::string EncDec::transcribe(const char* waveFile) {
  std::vector<float> pcmf32 = wav_read_legacy(waveFile);
  pcmf32.resize(kWindow, 0);
  return transcribe(pcmf32);
}

Engine* create_engine(EngineType type, const char* model_prefix, const char* vocab_path, bool multilingual) {
  switch (type) {
    case EngineType::Monolith:
      return new Monolith(model_prefix, vocab_path, multilingual);
    case EngineType::EncDec:
      return new EncDec(model_prefix, vocab_path, multilingual);
    default:
      fprintf(stderr, "Unknown engine-type\n");
      break;
  }
  return nullptr;
}

}  // namespace whisper
