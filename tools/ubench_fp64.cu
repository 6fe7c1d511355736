// Microbenchmark: FP64 DFMA peak of the vector pipe (the FP64 roofline denominator is not in
// MEASURED_PEAKS.json and has to be measured on the box).  8 independent FMA chains per thread, 1024 threads/SM x 2.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
int main() {
  const int blocks = 148 * 8, threads = 256, iters = 20000;
  double* o; cudaMalloc(&o, (size_t)blocks * threads * 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0); k_dfma<<<blocks, threads>>>(o, iters, 0.999999, 1e-9); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double n = (double)blocks * threads * iters * 8;
    printf("DFMA: %.3f ms  %.2f T DFMA/s = %.2f TFLOP/s  (%.1f DFMA/clk/SM at 1.965 GHz)\n", ms, n / ms * 1e-9, 2 * n / ms * 1e-9, n / (ms * 1e-3) / 148 / 1.965e9);
  }
  return 0;
}
