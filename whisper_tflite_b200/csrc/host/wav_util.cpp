#include "wav_util.h"

#include <climits>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <memory>

namespace whisper {
namespace {
struct FileCloser {
  void operator()(FILE* f) const {
    if (f) std::fclose(f);
  }
};
using File = std::unique_ptr<FILE, FileCloser>;
}  // namespace

std::vector<float> wav_read_legacy(const char* filename) {
  File f(std::fopen(filename, "rb"));
  if (!f) {
    std::cerr << "Failed to open file: " << filename << '\n';
    return {};
  }
  WAVHeader h{};
  const size_t got = std::fread(&h, 1, sizeof(h), f.get());
  if (got != sizeof(h) || std::memcmp(h.riff_header, "RIFF", 4) != 0 || std::memcmp(h.wave_header, "WAVE", 4) != 0 ||
      std::memcmp(h.fmt_header, "fmt ", 4) != 0) {
    std::cerr << "Not a valid WAV file: " << filename << '\n';
    return {};
  }
  const char* fmt = h.audio_format == 1 ? "PCM" : (h.audio_format == 3 ? "IEEE Float" : "Unknown");
  std::cout << "Audio Format: " << fmt << '\n'
            << "Num Channels: " << h.num_channels << '\n'
            << "Sample Rate: " << h.sample_rate << '\n'
            << "Bits Per Sample: " << h.bits_per_sample << '\n';
  if (h.block_align == 0) return {};
  // The reference sizes everything off the RIFF size and reads straight after the fmt chunk (wav_util.cpp:61-80):
  // samples = wav_size / block_align, then `wav_size` BYTES are requested; what the file cannot supply stays 0.
  const uint32_t n = h.wav_size / h.block_align;
  std::vector<float> out(n, 0.0f);
  if (h.audio_format == 1) {
    std::vector<int16_t> pcm(n, 0);
    const size_t want = std::min<size_t>(h.wav_size, pcm.size() * sizeof(int16_t));
    (void)!std::fread(pcm.data(), 1, want, f.get());
    for (uint32_t i = 0; i < n; ++i) out[i] = static_cast<float>(pcm[i]) / static_cast<float>(INT16_MAX);
  } else {
    const size_t want = std::min<size_t>(h.wav_size, out.size() * sizeof(float));
    (void)!std::fread(out.data(), 1, want, f.get());
  }
  return out;
}

std::vector<float> wav_read(const char* filename) {
  File f(std::fopen(filename, "rb"));
  if (!f) {
    std::fprintf(stderr, "failed to open WAV file '%s' - check your input\n", filename);
    return {};
  }
  char riff[12];
  if (std::fread(riff, 1, 12, f.get()) != 12 || std::memcmp(riff, "RIFF", 4) != 0 || std::memcmp(riff + 8, "WAVE", 4) != 0) {
    std::fprintf(stderr, "failed to open WAV file '%s' - check your input\n", filename);
    return {};
  }
  uint16_t channels = 0, bits = 0, format = 0;
  uint32_t rate = 0;
  std::vector<uint8_t> payload;
  for (;;) {
    char id[4];
    uint32_t sz = 0;
    if (std::fread(id, 1, 4, f.get()) != 4 || std::fread(&sz, 4, 1, f.get()) != 1) break;
    if (std::memcmp(id, "fmt ", 4) == 0) {
      uint8_t buf[16] = {};
      const size_t take = sz < 16 ? sz : 16;
      if (std::fread(buf, 1, take, f.get()) != take) break;
      std::memcpy(&format, buf, 2);
      std::memcpy(&channels, buf + 2, 2);
      std::memcpy(&rate, buf + 4, 4);
      std::memcpy(&bits, buf + 14, 2);
      std::fseek(f.get(), static_cast<long>(sz - take + (sz & 1)), SEEK_CUR);
    } else if (std::memcmp(id, "data", 4) == 0) {
      payload.resize(sz);
      payload.resize(std::fread(payload.data(), 1, sz, f.get()));
      break;
    } else {
      std::fseek(f.get(), static_cast<long>(sz + (sz & 1)), SEEK_CUR);
    }
  }
  if (channels != 1 && channels != 2) std::fprintf(stderr, "WAV file '%s' must be mono or stereo\n", filename);
  if (rate != 16000) std::fprintf(stderr, "WAV file '%s' must be 16 kHz\n", filename);
  if (bits != 16) std::fprintf(stderr, "WAV file '%s' must be 16-bit\n", filename);
  if (format != 1 || bits != 16 || (channels != 1 && channels != 2)) return {};
  const size_t frames = payload.size() / (2u * channels);
  const int16_t* s = reinterpret_cast<const int16_t*>(payload.data());
  std::vector<float> out(frames);
  if (channels == 1) {
    for (size_t i = 0; i < frames; ++i) out[i] = static_cast<float>(s[i]) / static_cast<float>(INT16_MAX);
  } else {
    // wav_util.cpp:121-124: the stereo sum is divided by INT32_MAX (sic) — kept
    for (size_t i = 0; i < frames; ++i)
      out[i] = static_cast<float>(static_cast<int>(s[2 * i]) + static_cast<int>(s[2 * i + 1])) / static_cast<float>(INT32_MAX);
  }
  return out;
}

}  // namespace whisper
