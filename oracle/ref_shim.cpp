// TEST INFRASTRUCTURE — not product code.
// C shim over the UNMODIFIED reference sources compiled where they lie under
// /root/reference (see oracle/Makefile). Gives pytest/bench a ctypes handle on the
// reference's own log_mel_spectrogram / fft / Reader / decode / wav_read_legacy
// (whisper.tflite/whisper.cpp:37-216, :519-665; wav_util.cpp:18-87).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load the resulting oracle/_ref/libwhisper_ref.so.
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "wav_util.h"
#include "whisper.h"

namespace {
whisper::Vocab g_vocab;
whisper::Filters g_filters;
}  // namespace

extern "C" {

// whisper.cpp:109-216, called exactly as EncDec::transcribe does (whisper.cpp:754-758).
int ref_log_mel(const float* pcm, int n_samples, int n_threads, const float* filters_80x201,
                float* mel_out_80x3000) {
  whisper::Filters f;
  f.n_mel = whisper::kNMEL;
  f.n_fft = 1 + whisper::kNFFT / 2;
  f.data.assign(filters_80x201, filters_80x201 + f.n_mel * f.n_fft);
  whisper::Mel mel;
  if (n_threads <= 0) n_threads = static_cast<int>(std::thread::hardware_concurrency());
  bool ok = whisper::log_mel_spectrogram(pcm, n_samples, whisper::kSampleRate, whisper::kNFFT,
                                         whisper::kHopLength, whisper::kNMEL, n_threads, f, mel);
  if (!ok) return 1;
  std::memcpy(mel_out_80x3000, mel.data.data(), mel.data.size() * sizeof(float));
  return 0;
}

// whisper.cpp:58-106
void ref_fft(const float* in, int n, float* out_2n) {
  std::vector<float> vin(in, in + n), vout;
  whisper::fft(vin, vout);
  std::memcpy(out_2n, vout.data(), sizeof(float) * 2 * n);
}

// whisper.cpp:519-611 — `head` points just past the u64 size prefix (whisper.cpp:676-681).
// Fills the process-global vocab used by ref_token / ref_decode; returns n_vocab after the
// multilingual transform; specials_out[8] = eot,sot,translate,transcribe,prev,solm,not,beg.
int ref_read_blob(const char* head, int multilingual, float* filters_out, int* n_mel, int* n_fft,
                  int* specials_out) {
  g_vocab = whisper::Vocab();
  g_filters = whisper::Filters();
  whisper::Reader reader(head, multilingual != 0);
  reader.read(g_filters, g_vocab);
  *n_mel = g_filters.n_mel;
  *n_fft = g_filters.n_fft;
  if (filters_out) std::memcpy(filters_out, g_filters.data.data(), g_filters.data.size() * sizeof(float));
  const int sp[8] = {g_vocab.token_eot, g_vocab.token_sot, g_vocab.token_translate, g_vocab.token_transcribe,
                     g_vocab.token_prev, g_vocab.token_solm, g_vocab.token_not, g_vocab.token_beg};
  std::memcpy(specials_out, sp, sizeof(sp));
  return g_vocab.n_vocab;
}

int ref_token(int id, char* buf, int cap) {
  auto it = g_vocab.id_to_token.find(id);
  if (it == g_vocab.id_to_token.end()) return -1;
  int n = static_cast<int>(it->second.size());
  if (n >= cap) return -2;
  std::memcpy(buf, it->second.data(), n);
  buf[n] = 0;
  return n;
}

// whisper.cpp:634-665
int ref_decode(const int64_t* ids, int n, int omit_special, char* buf, int cap) {
  std::string s = whisper::decode(g_vocab, ids, ids + n, omit_special != 0);
  if (static_cast<int>(s.size()) >= cap) return -2;
  std::memcpy(buf, s.data(), s.size());
  buf[s.size()] = 0;
  return static_cast<int>(s.size());
}

int ref_language_id(const char* code) { return whisper::language_id(code); }

int ref_remove_extra_spaces(const char* in, char* buf, int cap) {
  std::string s = whisper::remove_extra_spaces(in);
  if (static_cast<int>(s.size()) >= cap) return -2;
  std::memcpy(buf, s.data(), s.size());
  buf[s.size()] = 0;
  return static_cast<int>(s.size());
}

// wav_util.cpp:18-87
int ref_wav_read_legacy(const char* path, float* out, int cap) {
  std::vector<float> v = whisper::wav_read_legacy(path);
  int n = static_cast<int>(v.size());
  if (out) std::memcpy(out, v.data(), sizeof(float) * (n < cap ? n : cap));
  return n;
}

// wav_util.cpp:89-129 (dr_wav based)
int ref_wav_read(const char* path, float* out, int cap) {
  std::vector<float> v = whisper::wav_read(path);
  int n = static_cast<int>(v.size());
  if (out) std::memcpy(out, v.data(), sizeof(float) * (n < cap ? n : cap));
  return n;
}

const char* ref_lang_code(int id) { return whisper::lang_code(static_cast<size_t>(id)).c_str(); }
int ref_language_count() { return static_cast<int>(whisper::language_meta.size()); }

int ref_hardware_concurrency() { return static_cast<int>(std::thread::hardware_concurrency()); }
}
