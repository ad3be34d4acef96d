// Host-side table construction for the mel frontend (plain C++, shared by the engine and the
// CPU-emulation tests).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace wt {

struct MelHostTables {
  std::vector<float> hann;      // [400]
  std::vector<float> tw400;     // [16][13][2]  (cos, -sin)(2*pi*r*k/400)
  std::vector<float> fw;        // packed weights
  std::vector<int32_t> frow;    // [n_mel][2] {offset, first bin}
  std::vector<int32_t> flen;    // [n_mel]
  int n_mel = 0;
};

// filters: [n_mel][201] row-major (whisper.h:96-101). Each row is packed as the contiguous bin range
// [first non-zero, last non-zero]; a dense bank simply packs whole rows. The reference adds the
// mirrored bin P[400-k] onto P[k] for k = 1..199 (whisper.cpp:164-166); |X[400-k]| == |X[k]| for
// real input, so that fold is a factor 2 folded into the packed weights here.
inline MelHostTables build_mel_tables(const float* filters, int n_mel, int n_bins = 201) {
  const double kPi = 3.14159265358979323846;
  MelHostTables t;
  t.n_mel = n_mel;
  t.hann.resize(400);
  for (int i = 0; i < 400; ++i) t.hann[i] = (float)(0.5 * (1.0 - std::cos((2.0 * kPi * i) / 400)));
  t.tw400.resize(16 * 13 * 2);
  for (int r = 0; r < 16; ++r)
    for (int k = 0; k < 13; ++k) {
      const double ang = 2.0 * kPi * (double)(r * k) / 400.0;
      t.tw400[(r * 13 + k) * 2 + 0] = (float)std::cos(ang);
      t.tw400[(r * 13 + k) * 2 + 1] = (float)(-std::sin(ang));
    }
  t.frow.resize(2 * n_mel);
  t.flen.resize(n_mel);
  for (int j = 0; j < n_mel; ++j) {
    int lo = n_bins, hi = -1;
    for (int k = 0; k < n_bins; ++k)
      if (filters[j * n_bins + k] != 0.0f) {
        if (k < lo) lo = k;
        hi = k;
      }
    t.frow[2 * j + 0] = (int32_t)t.fw.size();
    if (hi < 0) {
      t.frow[2 * j + 1] = 0;
      t.flen[j] = 0;
      continue;
    }
    t.frow[2 * j + 1] = lo;
    t.flen[j] = hi - lo + 1;
    for (int k = lo; k <= hi; ++k) {
      const float fold = (k >= 1 && k <= 199) ? 2.0f : 1.0f;
      t.fw.push_back(filters[j * n_bins + k] * fold);
    }
  }
  return t;
}

}  // namespace wt
