// TEST INFRASTRUCTURE — runs csrc/simt_kernels.cuh under the CPU SIMT emulator.
#include "cuda_emu.h"
#include "../../whisper_tflite_b200/csrc/simt_kernels.cuh"

using namespace wt;

extern "C" {

// C[M][N] = epi(A.W^T); A rows via (rows_per_batch, batch_stride, row_stride) likewise C.
void emu_sgemm(const float* A, long long a_bs, int a_rpb, int a_rs, const float* W, const float* bias,
               const float* residual, const float* pos, int pos_rows, float* C, long long c_bs, int c_rpb, int c_rs,
               int M, int N, int K, int flags) {
  GemmF32Params p;
  p.A = A; p.amap = {a_bs, a_rpb, a_rs};
  p.W = W; p.bias = bias;
  p.residual = residual; p.rmap = {c_bs, c_rpb, c_rs};
  p.pos = pos; p.pos_rows = pos_rows;
  p.C = C; p.cmap = {c_bs, c_rpb, c_rs};
  p.M = M; p.N = N; p.K = K; p.flags = flags;
  emu::launch(sgemm_f32_kernel, dim3((N + 63) / 64, (M + 63) / 64), dim3(256), 0, p);
}

void emu_layernorm(const float* x, const float* g, const float* b, float* y, int rows, int D) {
  emu::launch(layernorm_kernel<float, float>, dim3((rows + 7) / 8), dim3(256), 0, x, g, b, y, rows, D);
}

void emu_attention(const float* qkv, float* out, int B, int S, int H) {
  emu::launch(attention_simt_kernel<float>, dim3((S + 127) / 128, H, B), dim3(128), 0, qkv, out, S, H);
}

void emu_decode_self(const float* qkv, float* kc, float* vc, float* out, int pos, int max_t, int H, int B) {
  emu::launch(decode_self_attn_kernel<float>, dim3(H, B), dim3(64), 0, qkv, kc, vc, out, pos, max_t, H);
}

void emu_decode_cross(const float* q, const float* kx, const float* vx, float* out, float* part, unsigned* tickets,
                      int S, int H, int B, int nsplit) {
  emu::launch(decode_cross_attn_kernel<float>, dim3(nsplit, H, B), dim3(128), 0, q, kx, vx, out, part, tickets, S, H);
}

void emu_split_heads(const float* kv, float* kx, float* vx, int B, int S, int H) {
  emu::launch(split_heads_kv_kernel<float>, dim3(16), dim3(256), 0, kv, kx, vx, B, S, H);
}

void emu_argmax_append(const float* logits, int V, int* tokens, int stride, int* lens, int* done, int pos, int eot,
                       int B) {
  emu::launch(argmax_append_kernel, dim3(B), dim3(256), 0, logits, V, tokens, stride, lens, done, pos, eot);
}

void emu_embed(const int* tokens, int stride, int pos, const float* te, const float* pe, float* x, int D, int B) {
  emu::launch(embed_kernel<float>, dim3(B), dim3(128), 0, tokens, stride, pos, te, pe, x, D);
}

void emu_mel_to_tm(const float* mel, const int* cmax, float* out, int n_mel, int n_len, int B) {
  emu::launch(mel_to_time_major_kernel<float>, dim3((n_len + 31) / 32, (n_mel + 31) / 32, B), dim3(256), 0, mel, cmax,
              out, n_mel, n_len);
}
}
