// SpMV variants on the reference-layout scalar CSR of a Kuhn-grid tet mesh (every block-row: 3 scalar rows of equal
// length, 3x3 blocks; interior vertices have 15 blocks).  Design input for k_pcg_spmv_dot (mfb_pcg.cu).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o ubench_spmv ubench_spmv.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

// V0: one warp per block-row, lanes over the scalar row positions
__global__ void __launch_bounds__(256) v0(const int* __restrict__ bp, const int* __restrict__ bc, int nF, const double* __restrict__ A,
                                          const double* __restrict__ x, double* __restrict__ y) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= nF) return;
  const int b0 = bp[warp], nbr = bp[warp + 1] - b0, rowlen = 3 * nbr;
  const int64_t base = 9ll * b0;
  double acc[3] = {0, 0, 0};
  for (int q = lane; q < rowlen; q += 32) {
    const double xv = __ldg(&x[3 * bc[b0 + q / 3] + q % 3]);
#pragma unroll
    for (int r = 0; r < 3; ++r) acc[r] += A[base + (int64_t)r * rowlen + q] * xv;
  }
#pragma unroll
  for (int r = 0; r < 3; ++r)
    for (int o = 16; o > 0; o >>= 1) acc[r] += __shfl_down_sync(0xffffffffu, acc[r], o);
  if (lane == 0) for (int r = 0; r < 3; ++r) y[3 * warp + r] = acc[r];
}

// V1: half-warp per block-row, lane per block; PERSIST: grid-stride with row-pointer prefetch
template <bool PERSIST>
__global__ void __launch_bounds__(256) v1(const int* __restrict__ bp, const int* __restrict__ bc, int nF, const double* __restrict__ A,
                                          const double* __restrict__ x, double* __restrict__ y) {
  const int lane = threadIdx.x & 31, hl = lane & 15;
  const int stride = PERSIST ? gridDim.x * 16 : (1 << 30);
  int row = (blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + (lane >> 4);
  int b0 = 0, b1 = 0;
  if (row < nF) { b0 = bp[row]; b1 = bp[row + 1]; }
  for (int row0 = row - (lane >> 4); row0 < nF; row0 += stride, row += stride) {
    const int nxt = row + stride;
    int n0 = 0, n1 = 0;
    if (PERSIST && nxt < nF) { n0 = bp[nxt]; n1 = bp[nxt + 1]; }
    const int nbr = (row < nF) ? b1 - b0 : 0, rowlen = 3 * nbr;
    const double* Ar = A + 9ll * b0;
    double acc[3] = {0, 0, 0};
    for (int k = hl; k < nbr; k += 16) {
      const int col = 3 * bc[b0 + k];
      double xv[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) xv[c] = __ldg(&x[col + c]);
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[r] = fma(Ar[(int64_t)r * rowlen + 3 * k + c], xv[c], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) acc[r] += __shfl_down_sync(0xffffffffu, acc[r], o, 16);
    if (hl == 0 && row < nF) for (int r = 0; r < 3; ++r) y[3 * row + r] = acc[r];
    b0 = n0; b1 = n1;
    if (!PERSIST) break;
  }
}

// V2: a warp takes RW consecutive block-rows; the values of those rows are ONE contiguous run which the warp copies to
// shared memory with fully coalesced loads (all lanes, 8 B each, many in flight); then lane per block from shared memory.
template <int RW>
__global__ void __launch_bounds__(256) v2(const int* __restrict__ bp, const int* __restrict__ bc, int nF, const double* __restrict__ A,
                                          const double* __restrict__ x, double* __restrict__ y, int cap /* doubles per warp */) {
  extern __shared__ double sm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double* sA = sm + (size_t)w * cap;
  const int nwarps = gridDim.x * 8;
  for (int g = blockIdx.x * 8 + w; g * RW < nF; g += nwarps) {
    const int r0 = g * RW, r1 = min(r0 + RW, nF);
    const int B0 = bp[r0], B1 = bp[r1];
    const int nval = 9 * (B1 - B0);
    const double* Ag = A + 9ll * B0;
    __syncwarp();
    if (nval <= cap) {
      for (int i = lane; i < nval; i += 32) sA[i] = Ag[i];
    }
    __syncwarp();
    // lane per block over the group's blocks
    for (int j0 = 0; j0 < B1 - B0; j0 += 32) {
      const int j = j0 + lane;
      double acc[3] = {0, 0, 0};
      int row = -1;
      if (j < B1 - B0) {
        // find the row of block B0 + j (RW <= 8 pointers)
        row = r0;
#pragma unroll
        for (int t = 1; t < RW; ++t) if (r0 + t < r1 && bp[r0 + t] <= B0 + j) row = r0 + t;
        const int b0 = bp[row], nbr = bp[row + 1] - b0, rowlen = 3 * nbr, k = B0 + j - b0;
        const int col = 3 * bc[B0 + j];
        double xv[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) xv[c] = __ldg(&x[col + c]);
        const double* Ar = (nval <= cap) ? sA + 9 * (b0 - B0) : Ag + 9 * (b0 - B0);
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int c = 0; c < 3; ++c) acc[r] = fma(Ar[r * rowlen + 3 * k + c], xv[c], acc[r]);
      }
      // segmented reduction by row (rows are contiguous in lane order)
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int rr = __shfl_down_sync(0xffffffffu, row, o);
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const double v = __shfl_down_sync(0xffffffffu, acc[r], o);
          if (lane + o < 32 && rr == row) acc[r] += v;
        }
      }
      const int prev = __shfl_up_sync(0xffffffffu, row, 1);
      if (row >= 0 && (lane == 0 || prev != row)) {
        // a row that straddles two 32-block passes needs accumulation: passes are sequential in this warp, so += is safe
        for (int r = 0; r < 3; ++r) { if (j0 > 0 && bp[row] < B0 + j0) y[3 * row + r] += acc[r]; else y[3 * row + r] = acc[r]; }
      }
    }
  }
}

// plain streaming read of the value array: the bandwidth ceiling for this size in one launch
__global__ void __launch_bounds__(256) vread(const double* __restrict__ A, int64_t n, double* out) {
  double s = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += A[i];
  if (s == 1.2345e300) *out = s;
}

int main(int argc, char** argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 55;
  const int nv = n + 1;
  const int nF = nv * nv * nv;
  const int off[15][3] = {{0,0,0},{1,0,0},{-1,0,0},{0,1,0},{0,-1,0},{0,0,1},{0,0,-1},{1,1,0},{-1,-1,0},{0,1,1},{0,-1,-1},{1,0,1},{-1,0,-1},{1,1,1},{-1,-1,-1}};
  std::vector<int> bp(nF + 1, 0), bc;
  for (int k = 0; k < nv; ++k) for (int j = 0; j < nv; ++j) for (int i = 0; i < nv; ++i) {
    const int v = i + nv * (j + nv * k);
    std::vector<int> cols;
    for (auto& o : off) {
      const int a = i + o[0], b = j + o[1], c = k + o[2];
      if (a < 0 || b < 0 || c < 0 || a >= nv || b >= nv || c >= nv) continue;
      cols.push_back(a + nv * (b + nv * c));
    }
    std::sort(cols.begin(), cols.end());
    for (int c : cols) bc.push_back(c);
    bp[v + 1] = (int)bc.size();
  }
  const int64_t nb = bc.size(), nnz = 9 * nb;
  std::vector<double> A(nnz), x(3 * nF), yref(3 * nF, 0.0);
  srand(1);
  for (auto& a : A) a = rand() / (double)RAND_MAX - 0.5;
  for (auto& a : x) a = rand() / (double)RAND_MAX - 0.5;
  for (int v = 0; v < nF; ++v) {
    const int b0 = bp[v], nbr = bp[v + 1] - b0, rowlen = 3 * nbr;
    for (int r = 0; r < 3; ++r) { double s = 0; for (int q = 0; q < rowlen; ++q) s += A[9ll * b0 + r * rowlen + q] * x[3 * bc[b0 + q / 3] + q % 3]; yref[3 * v + r] = s; }
  }
  int *dbp, *dbc; double *dA, *dx, *dy;
  CK(cudaMalloc(&dbp, (nF + 1) * 4)); CK(cudaMalloc(&dbc, nb * 4)); CK(cudaMalloc(&dA, nnz * 8)); CK(cudaMalloc(&dx, 3 * nF * 8)); CK(cudaMalloc(&dy, 3 * nF * 8));
  CK(cudaMemcpy(dbp, bp.data(), (nF + 1) * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dbc, bc.data(), nb * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dA, A.data(), nnz * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dx, x.data(), 3 * nF * 8, cudaMemcpyHostToDevice));
  // L2 flush buffer
  double* flush; const size_t fl = 256ull << 20; CK(cudaMalloc(&flush, fl));
  printf("nF %d blocks %lld values %.1f MB\n", nF, (long long)nb, nnz * 8 / 1e6);
  std::vector<double> y(3 * nF);
  auto bench = [&](const char* name, auto launch) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float tot = 0; const int reps = 10;
    for (int i = 0; i < reps + 2; ++i) {
      CK(cudaMemsetAsync(flush, i, fl));
      CK(cudaMemsetAsync(dy, 0, 3 * nF * 8));
      CK(cudaEventRecord(a)); launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
      float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (i >= 2) tot += ms;
    }
    CK(cudaGetLastError());
    CK(cudaMemcpy(y.data(), dy, 3 * nF * 8, cudaMemcpyDeviceToHost));
    double err = 0; for (int i = 0; i < 3 * nF; ++i) err = fmax(err, fabs(y[i] - yref[i]));
    // back-to-back (L2 as the CG loop leaves it)
    CK(cudaEventRecord(a)); for (int i = 0; i < 20; ++i) launch(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms2; CK(cudaEventElapsedTime(&ms2, a, b));
    printf("%-44s cold %7.1f us  %6.0f GB/s | back-to-back %7.1f us   maxerr %.2e\n", name, tot / reps * 1e3, nnz * 8 / (tot / reps) * 1e-6, ms2 / 20 * 1e3, err);
  };
  bench("read values only (592 blocks)", [&] { vread<<<592, 256>>>(dA, nnz, dy); });
  bench("read values only (2368 blocks)", [&] { vread<<<2368, 256>>>(dA, nnz, dy); });
  bench("V0 warp per row", [&] { v0<<<(nF * 32 + 255) / 256, 256>>>(dbp, dbc, nF, dA, dx, dy); });
  bench("V1 half-warp per row, one shot", [&] { v1<false><<<(nF + 15) / 16, 256>>>(dbp, dbc, nF, dA, dx, dy); });
  for (int g : {592, 740, 888, 1184}) {
    char nm[64]; snprintf(nm, 64, "V1 half-warp per row, persistent %d", g);
    bench(nm, [&] { v1<true><<<g, 256>>>(dbp, dbc, nF, dA, dx, dy); });
  }
  {
    const int cap = 9 * 16 * 8 + 64;   // 8 rows x 16 blocks
    auto k = v2<8>; CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * cap * 8));
    for (int g : {444, 592, 888}) {
      char nm[64]; snprintf(nm, 64, "V2 8 rows per warp via smem, grid %d", g);
      bench(nm, [&] { k<<<g, 256, 8 * cap * 8>>>(dbp, dbc, nF, dA, dx, dy, cap); });
    }
  }
  {
    const int cap = 9 * 16 * 4 + 64;
    auto k = v2<4>; CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * cap * 8));
    for (int g : {592, 888, 1184}) {
      char nm[64]; snprintf(nm, 64, "V2 4 rows per warp via smem, grid %d", g);
      bench(nm, [&] { k<<<g, 256, 8 * cap * 8>>>(dbp, dbc, nF, dA, dx, dy, cap); });
    }
  }
  return 0;
}
