// Store / load bandwidth of the per-element state layout of the update kernel.
//   soa   : component-major streams base[c*ld + e]  (what the C-ABI exposes; 130 written streams per tet)
//   aosoa : tiles of 32 elements, base[(e/32)*K*32 + c*32 + e%32]
// at the update kernel's occupancy (3 CTAs of 128 threads per SM, forced with dynamic shared memory) and at full
// occupancy, write-only and with R read streams beside the K written ones.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_streams ubench_streams.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

template <int K, int R, bool AOSOA>
__global__ void __launch_bounds__(128) k_streams(double* __restrict__ out, const double* __restrict__ in, int64_t n, int64_t ld) {
  extern __shared__ double sm[];
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const int lane = threadIdx.x & 31;
  double acc = (double)e;
#pragma unroll
  for (int c = 0; c < R; ++c) acc += AOSOA ? in[(e >> 5) * (R * 32) + c * 32 + lane] : in[c * ld + e];
  if (acc == 1.2345e300) sm[threadIdx.x] = acc;     // keep the shared-memory allocation alive
#pragma unroll
  for (int c = 0; c < K; ++c) {
    const double v = acc + c;
    if (AOSOA) out[(e >> 5) * (K * 32) + c * 32 + lane] = v; else out[c * ld + e] = v;
  }
}

template <int K, int R, bool AOSOA>
static void run(const char* name, double* out, const double* in, int64_t n, int smem) {
  const int64_t ld = (n + 31) & ~31ll;
  auto kern = k_streams<K, R, AOSOA>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  const int grid = (int)((n + 127) / 128);
  for (int i = 0; i < 2; ++i) kern<<<grid, 128, smem>>>(out, in, n, ld);
  CK(cudaEventRecord(a));
  const int reps = 5;
  for (int i = 0; i < reps; ++i) kern<<<grid, 128, smem>>>(out, in, n, ld);
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms; CK(cudaEventElapsedTime(&ms, a, b)); ms /= reps;
  const double bytes = (double)n * 8.0 * (K + R);
  printf("%-58s smem %6d  %7.3f ms  %7.1f GB/s\n", name, smem, ms, bytes / ms * 1e-6);
}

int main(int argc, char** argv) {
  const int64_t n = argc > 1 ? atoll(argv[1]) : 10110954;
  const int64_t ld = (n + 31) & ~31ll;
  double *out, *in;
  CK(cudaMalloc(&out, (size_t)ld * 130 * 8));
  CK(cudaMalloc(&in, (size_t)ld * 48 * 8));
  CK(cudaMemset(in, 0, (size_t)ld * 48 * 8));
  {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    CK(cudaMemset(out, 0, (size_t)ld * 130 * 8));
    CK(cudaEventRecord(a));
    for (int i = 0; i < 3; ++i) CK(cudaMemsetAsync(out, 0, (size_t)ld * 130 * 8));
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); ms /= 3;
    printf("%-58s              %7.3f ms  %7.1f GB/s\n", "cudaMemsetAsync 130 * ld doubles", ms, (double)ld * 130 * 8 / ms * 1e-6);
  }
  for (int smem : {0, 65536}) {
    run<130, 0, false>("write 130 streams, SoA", out, in, n, smem);
    run<130, 0, true>("write 130 streams, AoSoA tiles of 32", out, in, n, smem);
    run<54, 0, false>("write 54 streams, SoA", out, in, n, smem);
    run<16, 0, false>("write 16 streams, SoA", out, in, n, smem);
    run<130, 48, false>("read 48 + write 130 streams, SoA", out, in, n, smem);
    run<130, 48, true>("read 48 + write 130 streams, AoSoA", out, in, n, smem);
    run<48, 48, false>("read 48 + write 48 streams, SoA", out, in, n, smem);
    run<1, 48, false>("read 48 + write 1 stream, SoA", out, in, n, smem);
  }
  return 0;
}
