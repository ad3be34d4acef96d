// TEST INFRASTRUCTURE — runs csrc/mel_kernels.cuh under the CPU SIMT emulator.
#include <climits>
#include "cuda_emu.h"
#include "../../whisper_tflite_b200/csrc/mel_host.h"
#include "../../whisper_tflite_b200/csrc/mel_kernels.cuh"

extern "C" int emu_log_mel(const float* pcm, int n_clips, int n_samples, const float* filters, int n_mel,
                           float* raw_out, float* mel_out) {
  using namespace wt;
  MelHostTables t = build_mel_tables(filters, n_mel);
  MelParams p{};
  const int n_len = n_samples / kHop;
  std::vector<int> cmax(n_clips, INT_MIN);
  p.pcm = pcm;
  p.raw = raw_out;
  p.clip_max = cmax.data();
  p.hann = t.hann.data();
  p.tw400 = reinterpret_cast<const float2*>(t.tw400.data());
  p.fw = t.fw.data();
  p.frow = reinterpret_cast<const int2*>(t.frow.data());
  p.flen = t.flen.data();
  p.n_samples = n_samples;
  p.n_len = n_len;
  p.n_mel = n_mel;
  p.n_tiles = (n_len + kTileFrames - 1) / kTileFrames;
  emu::launch(mel_frames_kernel, dim3(p.n_tiles, n_clips), dim3(kMelThreads), kMelSmemBytes, p);
  emu::launch(mel_finalize_f32_kernel, dim3(8), dim3(256), 0, (const float*)raw_out, (const int*)cmax.data(), mel_out,
              n_mel * n_len, n_clips);
  return 0;
}
