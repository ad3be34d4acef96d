// Accuracy of the tanh.approx-based erf-GELU of the GEMM epilogue against the exact function, on the device.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probes/gelu_probe tools/probes/gelu_probe.cu
#include <cstdio>
#include <cmath>
#include <vector>
__device__ __forceinline__ float gelu_tanh_fast(float x) {
  const float x2 = fminf(x * x, 49.0f);
  float p = fmaf(x2, -3.51516783e-4f, 0.037005646f);
  p = fmaf(x2, p, 0.797507884f);
  const float u = x * p;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  const float h = 0.5f * x;
  return fmaf(h, t, h);
}
__global__ void k(const float* x, float* y, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = gelu_tanh_fast(x[i]);
}
int main() {
  const int n = 1 << 22;
  std::vector<float> hx(n), hy(n);
  for (int i = 0; i < n; ++i) hx[i] = -40.0f + 80.0f * (float)i / (float)(n - 1);
  float *dx, *dy;
  cudaMalloc(&dx, n * 4); cudaMalloc(&dy, n * 4);
  cudaMemcpy(dx, hx.data(), n * 4, cudaMemcpyHostToDevice);
  k<<<n / 256, 256>>>(dx, dy, n);
  cudaMemcpy(hy.data(), dy, n * 4, cudaMemcpyDeviceToHost);
  double worst = 0, wx = 0, worst_neg = 0, wnx = 0;
  for (int i = 0; i < n; ++i) {
    const double x = hx[i], g = 0.5 * x * (1.0 + erf(x / sqrt(2.0))), e = fabs(hy[i] - g);
    if (e > worst) { worst = e; wx = x; }
    if (x < -3.0 && e > worst_neg) { worst_neg = e; wnx = x; }
  }
  printf("max abs err %.3e at x=%.4f ; for x<-3: %.3e at x=%.4f ; y(-40)=%g y(-8)=%g y(40)=%g\n", worst, wx, worst_neg, wnx, hy[0],
         hy[(int)((32.0 / 80.0) * (n - 1))], hy[n - 1]);
  return 0;
}
