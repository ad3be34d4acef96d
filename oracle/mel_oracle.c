/* TEST INFRASTRUCTURE — not product code.
 *
 * Plain-C restatement of the reference's log-mel frontend, written to follow the
 * reference's arithmetic operation by operation (f32 butterflies, cosf/sinf twiddles per
 * butterfly, O(N^2) DFT leaves at odd sizes, f64 mel accumulation) so that its output is
 * expected to be bit-identical to the reference compiled with the same compiler.
 *
 *   reference: /root/reference/whisper.tflite/whisper.cpp
 *     dft                  :37-54
 *     fft                  :58-106
 *     log_mel_spectrogram  :109-216
 *
 * Pinned by tests/test_oracle_mel.py against oracle/_ref/libwhisper_ref.so (the reference's
 * own sources compiled in place) and against tests/golden/mel_*.npz generated from it.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call this.
 */
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* whisper.cpp:37-54 — naive DFT; the angle is formed in double then rounded to float, and
 * cos/sin are the float overloads (std::cos(float) == cosf). */
static void oracle_dft(const float* in, int n, float* out) {
  for (int k = 0; k < n; k++) {
    float re = 0.0f, im = 0.0f;
    for (int j = 0; j < n; j++) {
      float angle = (float)(2 * M_PI * k * j / n);
      re += in[j] * cosf(angle);
      im -= in[j] * sinf(angle);
    }
    out[2 * k + 0] = re;
    out[2 * k + 1] = im;
  }
}

/* whisper.cpp:58-106 — recursive radix-2 decimation in time; odd sizes fall to the DFT.
 * `scratch` must hold 6*n floats for the top-level call (even, odd, even_fft, odd_fft of
 * each level are carved from it: 2*(n/2) inputs + 2*(n) outputs = 3n, halving per level). */
static void oracle_fft_rec(const float* in, int n, float* out, float* scratch) {
  if (n == 1) {
    out[0] = in[0];
    out[1] = 0.0f;
    return;
  }
  if (n % 2 == 1) {
    oracle_dft(in, n, out);
    return;
  }
  const int h = n / 2;
  float* even = scratch;
  float* odd = even + h;
  float* even_fft = odd + h;
  float* odd_fft = even_fft + 2 * h;
  float* next = odd_fft + 2 * h;
  for (int i = 0; i < h; i++) {
    even[i] = in[2 * i];
    odd[i] = in[2 * i + 1];
  }
  oracle_fft_rec(even, h, even_fft, next);
  oracle_fft_rec(odd, h, odd_fft, next);
  for (int k = 0; k < h; k++) {
    float theta = (float)(2 * M_PI * k / n);
    float re = cosf(theta);
    float im = -sinf(theta);
    float re_odd = odd_fft[2 * k + 0];
    float im_odd = odd_fft[2 * k + 1];
    out[2 * k + 0] = even_fft[2 * k + 0] + re * re_odd - im * im_odd;
    out[2 * k + 1] = even_fft[2 * k + 1] + re * im_odd + im * re_odd;
    out[2 * (k + h) + 0] = even_fft[2 * k + 0] - re * re_odd + im * im_odd;
    out[2 * (k + h) + 1] = even_fft[2 * k + 1] - re * im_odd - im * re_odd;
  }
}

void oracle_fft(const float* in, int n, float* out_2n) {
  float* scratch = (float*)malloc(sizeof(float) * 6 * (size_t)n + 64);
  oracle_fft_rec(in, n, out_2n, scratch);
  free(scratch);
}

typedef struct {
  const float* samples;
  int n_samples, fft_size, fft_step, n_mel, n_len, n_fft;
  const float* hann;
  const float* filters;
  float* mel;
  int ith, n_threads;
} mel_job;

/* whisper.cpp:134-187 — one worker striding over frames. */
static void* mel_worker(void* arg) {
  mel_job* J = (mel_job*)arg;
  const int N = J->fft_size;
  float* fft_in = (float*)calloc((size_t)N, sizeof(float));
  float* fft_out = (float*)calloc((size_t)2 * N, sizeof(float));
  float* scratch = (float*)malloc(sizeof(float) * 6 * (size_t)N + 64);
  for (int i = J->ith; i < J->n_len; i += J->n_threads) {
    const int offset = i * J->fft_step;
    for (int j = 0; j < N; j++) { /* :148-154 Hann window, zero past the end */
      fft_in[j] = (offset + j < J->n_samples) ? J->hann[j] * J->samples[offset + j] : 0.0f;
    }
    oracle_fft_rec(fft_in, N, fft_out, scratch); /* :157 */
    for (int j = 0; j < N; j++) {                /* :159-162 power, in place */
      fft_out[j] = fft_out[2 * j + 0] * fft_out[2 * j + 0] + fft_out[2 * j + 1] * fft_out[2 * j + 1];
    }
    for (int j = 1; j < N / 2; j++) fft_out[j] += fft_out[N - j]; /* :164-166 mirrored fold */
    for (int j = 0; j < J->n_mel; j++) {                          /* :169-185 */
      double sum = 0.0;
      for (int k = 0; k < J->n_fft; k++) sum += fft_out[k] * J->filters[j * J->n_fft + k]; /* f32 product, f64 add */
      if (sum < 1e-10f) sum = 1e-10f; /* kEPS is a float constant in the reference (:176) */
      sum = log10(sum);
      J->mel[j * J->n_len + i] = (float)sum;
    }
  }
  free(fft_in);
  free(fft_out);
  free(scratch);
  return NULL;
}

/* whisper.cpp:109-216. mel_out is [n_mel][n_samples/fft_step] row-major f32.
 * If raw_out != NULL the un-normalised log10 values (before :198-213) are copied there. */
int oracle_log_mel(const float* samples, int n_samples, int fft_size, int fft_step, int n_mel, int n_threads,
                   const float* filters, float* mel_out, float* raw_out) {
  float* hann = (float*)malloc(sizeof(float) * (size_t)fft_size);
  for (int i = 0; i < fft_size; i++) hann[i] = (float)(0.5 * (1.0 - cos((2.0 * M_PI * i) / fft_size))); /* :117-120 */
  const int n_len = n_samples / fft_step; /* :123 */
  const int n_fft = 1 + fft_size / 2;     /* :129 */
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  pthread_t th[64];
  mel_job jobs[64];
  for (int t = 0; t < n_threads; t++) {
    mel_job j = {samples, n_samples, fft_size, fft_step, n_mel, n_len, n_fft, hann, filters, mel_out, t, n_threads};
    jobs[t] = j;
    pthread_create(&th[t], NULL, mel_worker, &jobs[t]);
  }
  for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  const int total = n_mel * n_len;
  if (raw_out) memcpy(raw_out, mel_out, sizeof(float) * (size_t)total);
  double mmax = -1e20; /* :198-205 */
  for (int i = 0; i < total; i++)
    if (mel_out[i] > mmax) mmax = mel_out[i];
  mmax -= 8.0;
  for (int i = 0; i < total; i++) { /* :207-213 — compare/assign go through the f32 store */
    if (mel_out[i] < mmax) mel_out[i] = (float)mmax;
    mel_out[i] = (float)((mel_out[i] + 4.0) / 4.0);
  }
  free(hann);
  return 0;
}

/* whisper.cpp:346-361 — greedy argmax, `>=` so the LAST maximal index wins. */
long long oracle_argmax_last(const float* begin, long long n, float* max_out) {
  float max_value = begin[0];
  long long max_index = 0;
  for (long long i = 1; i < n; i++) {
    if (begin[i] >= max_value) {
      max_value = begin[i];
      max_index = i;
    }
  }
  if (max_out) *max_out = max_value;
  return max_index;
}
