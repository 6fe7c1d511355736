// Microbenchmark (design input, not product): FP64 reduction throughput into a large global array for the access
// patterns the assembly kernel could use.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a ubench_red.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// each thread issues `per` REDs; run = number of consecutive doubles a LANE GROUP of `run` lanes covers contiguously
// mode 0: every lane random address.  mode 1: groups of `run` adjacent lanes hit `run` contiguous doubles at a random base.
// mode 2: one lane does `run` contiguous doubles with successive instructions (what the kernel does today: run = 3).
template <int MODE>
__global__ void k_red(double* A, uint32_t nslots, int per, int run, uint32_t seed) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31;
  for (int i = 0; i < per; ++i) {
    uint32_t idx;
    if (MODE == 0) idx = hash32(tid * 977u + i * 131071u + seed) % nslots;
    else if (MODE == 1) {
      const uint32_t grp = (tid - lane) + lane / run;       // group id within the warp
      const uint32_t base = hash32(grp * 977u + i * 131071u + seed) % (nslots - 64);
      idx = (base / run) * run + lane % run;
    } else {
      const uint32_t base = hash32(tid * 977u + (i / run) * 131071u + seed) % (nslots - 64);
      idx = (base / run) * run + (i % run);
    }
    atomicAdd(&A[idx], 1.0);
  }
}

// bulk reduce: each CTA stages `bytes` of doubles in shared memory and issues `per` cp.reduce.async.bulk ops of `bytes`
__global__ void k_bulk(double* A, uint32_t nslots, int per, int bytes, uint32_t seed) {
  extern __shared__ __align__(128) double sm[];
  for (int i = threadIdx.x; i < bytes / 8 * 4; i += blockDim.x) sm[i] = 1.0;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nd = bytes / 8;
  for (int i = 0; i < per; ++i) {
    uint32_t base = hash32(tid * 977u + i * 131071u + seed) % (nslots - 4096u);
    base = base / nd * nd;
    base &= ~1u;
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(sm);
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;" :: "l"(A + base), "r"(s), "r"(bytes) : "memory");
    if ((i & 7) == 7) { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory"); }
  }
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main(int argc, char** argv) {
  const uint32_t nslots = argc > 1 ? (uint32_t)atoll(argv[1]) : 230000000u;   // default 1.84 GB, like the C4 matrix values
  printf("window = %.1f MB\n", nslots * 8e-6);
  double* A; CK(cudaMalloc(&A, (size_t)nslots * 8)); CK(cudaMemset(A, 0, (size_t)nslots * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = 148 * 16, threads = 128, per = 64;
  const double total = (double)blocks * threads * per;
  auto report = [&](const char* name, double n, float ms) { printf("%-44s %8.3f ms  %8.1f G values/s  %6.3f cyc/lane/SM@1.965GHz\n", name, ms, n / ms * 1e-6, ms * 1e-3 * 1.965e9 * 148 / n); };
  float ms;
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0); k_red<0><<<blocks, threads>>>(A, nslots, per, 1, 1u + rep); cudaEventRecord(e1); CK(cudaDeviceSynchronize()); cudaEventElapsedTime(&ms, e0, e1);
    if (rep) report("RED.F64 random 8 B per lane", total, ms);
    for (int run : {2, 4, 8, 16, 32}) {
      cudaEventRecord(e0); k_red<1><<<blocks, threads>>>(A, nslots, per, run, 7u + rep); cudaEventRecord(e1); CK(cudaDeviceSynchronize()); cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "RED.F64 lane groups of %d contiguous doubles", run); if (rep) report(nm, total, ms);
    }
    for (int run : {3, 9}) {
      cudaEventRecord(e0); k_red<2><<<blocks, threads>>>(A, nslots, per - per % run, run, 9u + rep); cudaEventRecord(e1); CK(cudaDeviceSynchronize()); cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "RED.F64 per-lane runs of %d successive doubles", run); if (rep) report(nm, (double)blocks * threads * (per - per % run), ms);
    }
    for (int bytes : {32, 80, 144, 1024}) {
      const int bper = 16;
      cudaEventRecord(e0); k_bulk<<<blocks, threads, 4 * bytes + 128>>>(A, nslots, bper, bytes, 3u + rep); cudaEventRecord(e1);
      cudaError_t er = cudaDeviceSynchronize(); if (er != cudaSuccess) { printf("bulk %d B: %s\n", bytes, cudaGetErrorString(er)); return 1; }
      cudaEventElapsedTime(&ms, e0, e1);
      char nm[96]; snprintf(nm, 96, "cp.reduce.async.bulk add.f64 %d B per op", bytes);
      if (rep) { report(nm, (double)blocks * threads * bper * (bytes / 8), ms); printf("    = %.2f G ops/s, %.2f cyc/op/SM\n", (double)blocks * threads * bper / ms * 1e-6, ms * 1e-3 * 1.965e9 * 148 / ((double)blocks * threads * bper)); }
    }
  }
  return 0;
}
