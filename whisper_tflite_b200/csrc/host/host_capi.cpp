// Flat C entry points over the whisper.h mirror so Python (ctypes) tests and tools can drive the C++ host
// layer exactly as app/encdec / the JNI bridge would. Not part of the reference surface.
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "whisper.h"

#define WTH_API extern "C" __attribute__((visibility("default")))

namespace {
struct Blob {
  whisper::Vocab vocab;
  whisper::Filters filters;
};
int put(const std::string& s, char* buf, int cap) {
  if (static_cast<int>(s.size()) >= cap) return -2;
  std::memcpy(buf, s.data(), s.size());
  buf[s.size()] = 0;
  return static_cast<int>(s.size());
}
thread_local std::string g_error;
}  // namespace

WTH_API const char* wth_last_error() { return g_error.c_str(); }

// ---- Reader / decode / language table ----
WTH_API void* wth_blob_read(const char* head, int multilingual) {
  auto* b = new Blob();
  whisper::Reader(head, multilingual != 0).read(b->filters, b->vocab);
  return b;
}
WTH_API void wth_blob_free(void* b) { delete static_cast<Blob*>(b); }
WTH_API int wth_blob_info(void* bp, int* n_vocab, int* n_mel, int* n_fft, int* specials8) {
  auto* b = static_cast<Blob*>(bp);
  *n_vocab = b->vocab.n_vocab;
  *n_mel = b->filters.n_mel;
  *n_fft = b->filters.n_fft;
  const whisper::Vocab& v = b->vocab;
  const int sp[8] = {v.token_eot, v.token_sot, v.token_translate, v.token_transcribe, v.token_prev, v.token_solm, v.token_not, v.token_beg};
  std::memcpy(specials8, sp, sizeof(sp));
  return static_cast<int>(b->vocab.id_to_token.size());
}
WTH_API void wth_blob_filters(void* bp, float* out) {
  auto* b = static_cast<Blob*>(bp);
  std::memcpy(out, b->filters.data.data(), b->filters.data.size() * sizeof(float));
}
WTH_API int wth_blob_token(void* bp, int id, char* buf, int cap) {
  auto* b = static_cast<Blob*>(bp);
  auto it = b->vocab.id_to_token.find(id);
  return it == b->vocab.id_to_token.end() ? -1 : put(it->second, buf, cap);
}
WTH_API int wth_blob_decode(void* bp, const int64_t* ids, int n, int omit_special, char* buf, int cap) {
  auto* b = static_cast<Blob*>(bp);
  return put(whisper::decode(b->vocab, ids, ids + n, omit_special != 0), buf, cap);
}
WTH_API int wth_blob_decode_i32(void* bp, const int32_t* ids, int n, int omit_special, char* buf, int cap) {
  auto* b = static_cast<Blob*>(bp);
  return put(whisper::decode(b->vocab, ids, ids + n, omit_special != 0), buf, cap);
}
WTH_API int wth_language_id(const char* code) { return whisper::language_id(code); }
WTH_API int wth_lang_code(int id, char* buf, int cap) {
  if (id < 0 || id >= static_cast<int>(whisper::language_meta.size())) return -1;
  return put(whisper::lang_code(static_cast<size_t>(id)), buf, cap);
}
WTH_API int wth_remove_extra_spaces(const char* in, char* buf, int cap) { return put(whisper::remove_extra_spaces(in), buf, cap); }

// ---- WAV / mmap ----
WTH_API int wth_wav_read_legacy(const char* path, float* out, int cap) {
  const std::vector<float> v = whisper::wav_read_legacy(path);
  const int n = static_cast<int>(v.size());
  if (out) std::memcpy(out, v.data(), sizeof(float) * static_cast<size_t>(n < cap ? n : cap));
  return n;
}
WTH_API int wth_wav_read(const char* path, float* out, int cap) {
  const std::vector<float> v = whisper::wav_read(path);
  const int n = static_cast<int>(v.size());
  if (out) std::memcpy(out, v.data(), sizeof(float) * static_cast<size_t>(n < cap ? n : cap));
  return n;
}
WTH_API long long wth_mmap_probe(const char* path, unsigned char* first16) {
  try {
    whisper::MmapFile f(path);
    whisper::MmapFile g(std::move(f));  // exercise the move path
    if (f.data() != nullptr || f.size() != 0) return -3;
    const size_t n = g.size() < 16 ? g.size() : 16;
    std::memcpy(first16, g.data(), n);
    return static_cast<long long>(g.size());
  } catch (const std::runtime_error& e) {
    g_error = e.what();
    return -1;
  }
}

// ---- log_mel_spectrogram (device) ----
WTH_API int wth_log_mel(const float* pcm, int n_samples, const float* filters, int n_mel, int n_fft, float* out, int* n_len) {
  whisper::Filters f;
  f.n_mel = n_mel;
  f.n_fft = n_fft;
  f.data.assign(filters, filters + static_cast<size_t>(n_mel) * n_fft);
  whisper::Mel mel;
  const unsigned threads = 1;
  if (!whisper::log_mel_spectrogram(pcm, n_samples, whisper::kSampleRate, whisper::kNFFT, whisper::kHopLength, n_mel,
                                    static_cast<int>(threads), f, mel))
    return 1;
  *n_len = mel.n_len;
  std::memcpy(out, mel.data.data(), mel.data.size() * sizeof(float));
  return 0;
}

// ---- engines ----
WTH_API void* wth_engine_create(int type, const char* model_prefix, const char* vocab_path, int multilingual) {
  try {
    return whisper::create_engine(static_cast<whisper::EngineType>(type), model_prefix, vocab_path, multilingual != 0);
  } catch (const std::runtime_error& e) {  // MmapFile (mmap_file.cpp:15-30)
    g_error = e.what();
    return nullptr;
  }
}
WTH_API void wth_engine_destroy(void* e) { delete static_cast<whisper::Engine*>(e); }
WTH_API int wth_engine_transcribe(void* e, const float* pcm, int n, char* buf, int cap, int* size_after) {
  std::vector<float> v(pcm, pcm + n);
  const std::string s = static_cast<whisper::Engine*>(e)->transcribe(v);
  if (size_after) *size_after = static_cast<int>(v.size());
  return put(s, buf, cap);
}
WTH_API int wth_engine_transcribe_file(void* e, const char* path, char* buf, int cap) {
  return put(static_cast<whisper::Engine*>(e)->transcribe(path), buf, cap);
}
// clips: [n_clips][n_samples] contiguous; texts are written NUL-separated into buf
WTH_API int wth_engine_transcribe_batch(void* e, const float* pcm, int n_clips, int n_samples, char* buf, int cap) {
  std::vector<std::vector<float>> clips(n_clips);
  for (int i = 0; i < n_clips; ++i) clips[i].assign(pcm + static_cast<size_t>(i) * n_samples, pcm + static_cast<size_t>(i + 1) * n_samples);
  const std::vector<std::string> text = static_cast<whisper::Engine*>(e)->transcribe_batch(clips);
  int off = 0;
  for (const std::string& s : text) {
    if (off + static_cast<int>(s.size()) + 1 > cap) return -2;
    std::memcpy(buf + off, s.data(), s.size());
    off += static_cast<int>(s.size());
    buf[off++] = 0;
  }
  return off;
}
// ---- ADDED options / int16 ingest (whisper.h mirror: Engine::set_language / set_task / set_omit_special_tokens,
//      transcribe_batch_i16) ----
WTH_API int wth_engine_set_language(void* e, const char* code) { return static_cast<whisper::Engine*>(e)->set_language(code) ? 1 : 0; }
WTH_API void wth_engine_set_task(void* e, int translate) { static_cast<whisper::Engine*>(e)->set_task(translate != 0); }
WTH_API void wth_engine_set_omit_special_tokens(void* e, int omit) { static_cast<whisper::Engine*>(e)->set_omit_special_tokens(omit != 0); }
WTH_API int wth_engine_transcribe_batch_i16(void* e, const int16_t* pcm, int n_clips, int n_samples, char* buf, int cap) {
  std::vector<std::vector<int16_t>> clips(n_clips);
  for (int i = 0; i < n_clips; ++i) clips[i].assign(pcm + static_cast<size_t>(i) * n_samples, pcm + static_cast<size_t>(i + 1) * n_samples);
  const std::vector<std::string> text = static_cast<whisper::Engine*>(e)->transcribe_batch_i16(clips);
  int off = 0;
  for (const std::string& s : text) {
    if (off + static_cast<int>(s.size()) + 1 > cap) return -2;
    std::memcpy(buf + off, s.data(), s.size());
    off += static_cast<int>(s.size());
    buf[off++] = 0;
  }
  return off;
}
