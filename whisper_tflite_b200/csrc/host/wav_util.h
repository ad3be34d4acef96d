// WAV ingest, host side. Mirrors /root/reference/whisper.tflite/wav_util.h:9-26.
//   wav_read_legacy — what the engines call (whisper.cpp:686, :772): fixed 36-byte header, NO chunk walk, so
//                     the 8-byte "data" sub-chunk header is decoded as 4 (int16) / 2 (f32) leading samples and
//                     the RIFF size is taken as the payload size (wav_util.cpp:29, :61-69). Kept bit-for-bit,
//                     quirks included, because file transcription results depend on it.
//   wav_read        — chunk-walking reader with the reference's dr_wav semantics (wav_util.cpp:89-129):
//                     int16, mono or stereo -> mono float.
#pragma once
#include <cstdint>
#include <vector>

namespace whisper {
#pragma pack(push, 1)
struct WAVHeader {
  char riff_header[4];
  uint32_t wav_size;
  char wave_header[4];
  char fmt_header[4];
  uint32_t fmt_chunk_size;
  uint16_t audio_format;
  uint16_t num_channels;
  uint32_t sample_rate;
  uint32_t byte_rate;
  uint16_t block_align;
  uint16_t bits_per_sample;
};
#pragma pack(pop)
static_assert(sizeof(WAVHeader) == 36, "packed RIFF/fmt header");

std::vector<float> wav_read_legacy(const char* filename);
std::vector<float> wav_read(const char* filename);
}  // namespace whisper
