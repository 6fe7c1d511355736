// mfb_create / mfb_destroy: mesh-derived constants (volumes, dphi/dX), the Assembler's CSR pattern and the
// element->nonzero scatter map, all built once and kept on the device.
//
//   volumes ........ src/mesh/tet_mesh.cpp:31-34 (|igl::volume|), src/mesh/tri2d_mesh.cpp:32-34 (signed igl::doublearea)
//   dphi/dX ........ deps/Bartels/src/linear_tet_dphi_dX.cpp:5-17, linear_tri2d_dphi_dX.cpp:5-17
//   pattern + map .. src/utils/sparse_utils.cpp:19-117 (ctor) and :140-163 (value index of each block entry)
// The reference sorts the (row,col) node pairs on the host; here the same sort runs on the GPU as a radix sort of
// 64-bit keys row*nF+col (cub via thrust), which yields the identical pattern because the pattern only depends on the
// SET of pairs and on ascending (row, col) order.
#include "mfb_ctx.cuh"
#include <algorithm>
#include <thrust/device_ptr.h>
#include <thrust/sort.h>
#include <thrust/scan.h>
#include <thrust/binary_search.h>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/execution_policy.h>
#include <cmath>
#include <climits>

namespace mfb {
thread_local std::string g_last_error;

static const long long KEY_INVALID = LLONG_MAX;

__global__ void k_pair_keys(const int* __restrict__ fidx, int npe, int nE, int64_t ld, long long nF, long long* keys, int* vals) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int nn = npe * npe;
  if (p >= (int64_t)nE * nn) return;
  const int e = (int)(p / nn), jk = (int)(p % nn), j = jk / npe, k = jk % npe;
  const int gi = fidx[j * ld + e], gj = fidx[k * ld + e];
  keys[p] = (gi < 0 || gj < 0) ? KEY_INVALID : (long long)gi * nF + gj;
  vals[p] = (int)p;
}
__global__ void k_heads(const long long* __restrict__ keys, int64_t n, int* head) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}
__global__ void k_block_info(const long long* __restrict__ keys, const int* __restrict__ head, const int* __restrict__ bid,
                             int64_t n, long long nF, int* brow, int* bcol, int* diagblk) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !head[i]) return;
  const int b = bid[i] - 1;
  const int gi = (int)(keys[i] / nF), gj = (int)(keys[i] % nF);
  brow[b] = gi; bcol[b] = gj;
  if (gi == gj) diagblk[gi] = b;
}
__global__ void k_fill_maps(const int* __restrict__ vals, const int* __restrict__ bid, const int* __restrict__ brow,
                            const int* __restrict__ blockptr, int64_t n, int npe, int dim, int64_t ld, int* smap, int* sstride) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = bid[i] - 1;
  const int gi = brow[b];
  const int rank = b - blockptr[gi], nbr = blockptr[gi + 1] - blockptr[gi];
  const int p = vals[i];
  const int nn = npe * npe;
  const int e = p / nn, jk = p % nn, j = jk / npe;
  smap[(int64_t)jk * ld + e] = dim * dim * blockptr[gi] + dim * rank;
  sstride[(int64_t)j * ld + e] = dim * nbr;
}
__global__ void k_outer(const int* __restrict__ blockptr, int nF, int dim, int* outer) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > dim * nF) return;
  if (i == dim * nF) { outer[i] = dim * dim * blockptr[nF]; return; }
  const int gi = i / dim, r = i % dim;
  const int nbr = blockptr[gi + 1] - blockptr[gi];
  outer[i] = dim * dim * blockptr[gi] + r * dim * nbr;
}
__global__ void k_inner(const int* __restrict__ blockptr, const int* __restrict__ brow, const int* __restrict__ bcol, int nb, int dim, int* inner) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int dd = dim * dim;
  if (i >= (int64_t)nb * dd) return;
  const int b = (int)(i / dd), rc = (int)(i % dd), r = rc / dim, c = rc % dim;
  const int gi = brow[b];
  const int rank = b - blockptr[gi], nbr = blockptr[gi + 1] - blockptr[gi];
  inner[(int64_t)dd * blockptr[gi] + (int64_t)r * dim * nbr + dim * rank + c] = dim * bcol[b] + c;
}
// bT[b] = index of the transposed block (gj,gi) of block b = (gi,gj): binary search in block-row gj (columns ascending)
__global__ void k_block_transpose(const int* __restrict__ blockptr, const int* __restrict__ brow, const int* __restrict__ bcol, int nb, int* bT) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  const int gi = brow[b], gj = bcol[b];
  int lo = blockptr[gj], hi = blockptr[gj + 1] - 1;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (bcol[mid] < gi) lo = mid + 1; else hi = mid; }
  bT[b] = lo;
}
// mirror list: one int4 (dst0, dst row stride, src0, src row stride) per block of the lower triangle incl. diagonal
// blocks, so that k_mirror_lower needs a single index load per copied entry
__global__ void k_mirror_flags(const int* __restrict__ brow, const int* __restrict__ bcol, int nb, int* flag) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < nb) flag[b] = (brow[b] >= bcol[b]) ? 1 : 0;
}
__global__ void k_mirror_fill(const int* __restrict__ blockptr, const int* __restrict__ brow, const int* __restrict__ bcol,
                              const int* __restrict__ bT, const int* __restrict__ flag, const int* __restrict__ pos, int nb, int dim, int4* list) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb || !flag[b]) return;
  const int gi = brow[b], gj = bcol[b], dd = dim * dim;
  const int p0 = blockptr[gi], n0 = blockptr[gi + 1] - p0;
  const int p1 = blockptr[gj], n1 = blockptr[gj + 1] - p1;
  const int bt = (gi == gj) ? b : bT[b];
  int4 v;
  v.x = dd * p0 + dim * (b - p0); v.y = dim * n0;
  v.z = dd * p1 + dim * (bt - p1); v.w = dim * n1;
  list[pos[b]] = v;
}
// ---- v4 assembly structures (tets): block-CSR copy of the upper triangle
__global__ void k_upper_flags(const int* __restrict__ brow, const int* __restrict__ bcol, int nb, int* flag) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < nb) flag[b] = (brow[b] <= bcol[b]) ? 1 : 0;
}
// fcode[b] = 2 * (upper block holding the values of block b) + transposed ; ub = exclusive scan of the upper flags
__global__ void k_fcode_fill(const int* __restrict__ brow, const int* __restrict__ bcol, const int* __restrict__ bT, const int* __restrict__ ub, int nb, int* fcode) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  fcode[b] = (brow[b] <= bcol[b]) ? 2 * ub[b] : 2 * ub[bT[b]] + 1;
}
// local pair (j,k), j < k -> 0..5
__host__ __device__ inline int pair6_index(int j, int k) { return j * 3 - j * (j - 1) / 2 + (k - j - 1); }
// bmap[pair6(j,k)][e] for pairs with BOTH vertices free (sorted pair list: vals = e*16+j*4+k): 2 * upper block + transposed
__global__ void k_fill_bmap(const int* __restrict__ vals, const int* __restrict__ bid, const int* __restrict__ fcode, int64_t n, int64_t ld, int* bmap) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int p = vals[i];
  const int e = p / 16, j = (p % 16) / 4, k = p % 4;
  if (j >= k) return;
  bmap[(int64_t)pair6_index(j, k) * ld + e] = fcode[bid[i] - 1];
}
// pairs with exactly ONE pinned vertex: the block K_e(free, pinned) is accumulated into the DIAGONAL slot of the free
// vertex; k_finalize_rows turns (row sum of the off-diagonal blocks + that slot) into the diagonal block, using
// sum_k K_e(j,k) = 0 (Jloc = dphi^T (x) I and the four dphi rows sum to zero).
__global__ void k_fill_bmap_pinned(const int* __restrict__ fidx, const int* __restrict__ diagblk, const int* __restrict__ ub, int nE, int64_t ld, int* bmap) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nE) return;
  int f[4];
  for (int k = 0; k < 4; ++k) f[k] = fidx[k * ld + e];
  for (int j = 0; j < 4; ++j)
    for (int k = j + 1; k < 4; ++k) {
      if (f[j] >= 0 && f[k] < 0) bmap[(int64_t)pair6_index(j, k) * ld + e] = 2 * ub[diagblk[f[j]]];
      else if (f[j] < 0 && f[k] >= 0) bmap[(int64_t)pair6_index(j, k) * ld + e] = 2 * ub[diagblk[f[k]]] + 1;
    }
}
__global__ void k_fill_int(int* p, int64_t n, int v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

static int build_pattern(mfb_ctx* c) {
  const int npe = c->npe, dim = c->dim, nE = c->nE, nF = c->nF;
  const int64_t np = (int64_t)nE * npe * npe;
  cudaStream_t st = c->stream;
  DBuf<long long> keys; DBuf<int> vals, head, bid;
  DBuf<int>& brow = c->brow;
  int rc;
  if ((rc = keys.alloc(np)) || (rc = vals.alloc(np)) || (rc = head.alloc(np)) || (rc = bid.alloc(np))) return rc;
  k_pair_keys<<<nblk(np), 256, 0, st>>>(c->fidx.p, npe, nE, c->ld, nF, keys.p, vals.p);
  auto pol = thrust::cuda::par.on(st);
  thrust::sort_by_key(pol, thrust::device_ptr<long long>(keys.p), thrust::device_ptr<long long>(keys.p + np), thrust::device_ptr<int>(vals.p));
  const int64_t nvalid = thrust::lower_bound(pol, thrust::device_ptr<long long>(keys.p), thrust::device_ptr<long long>(keys.p + np), KEY_INVALID) -
                         thrust::device_ptr<long long>(keys.p);
  int nb = 0;
  if (nvalid > 0) {
    k_heads<<<nblk(nvalid), 256, 0, st>>>(keys.p, nvalid, head.p);
    thrust::inclusive_scan(pol, thrust::device_ptr<int>(head.p), thrust::device_ptr<int>(head.p + nvalid), thrust::device_ptr<int>(bid.p));
    MFB_CUDA(cudaMemcpyAsync(&nb, bid.p + (nvalid - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
    MFB_CUDA(cudaStreamSynchronize(st));
  }
  const int64_t nnz = (int64_t)nb * dim * dim;
  if (nnz > (int64_t)INT_MAX) { g_last_error = "nnz exceeds int32"; return MFB_ERR_TOO_LARGE; }
  c->nb = nb; c->nnz = nnz;
  if ((rc = brow.alloc(nb > 0 ? nb : 1)) || (rc = c->bT.alloc(nb > 0 ? nb : 1)) || (rc = c->bcol.alloc(nb > 0 ? nb : 1)) || (rc = c->diagblk.alloc(nF > 0 ? nF : 1)) ||
      (rc = c->blockptr.alloc((size_t)nF + 1)) || (rc = c->outer.alloc((size_t)dim * nF + 1)) || (rc = c->inner.alloc(nnz > 0 ? nnz : 1)) ||
      (rc = c->smap.alloc((size_t)npe * npe * c->ld)) || (rc = c->sstride.alloc((size_t)npe * c->ld)))
    return rc;
  k_fill_int<<<nblk((int64_t)npe * npe * c->ld), 256, 0, st>>>(c->smap.p, (int64_t)npe * npe * c->ld, -1);
  k_fill_int<<<nblk((int64_t)npe * c->ld), 256, 0, st>>>(c->sstride.p, (int64_t)npe * c->ld, 0);
  k_fill_int<<<nblk(nF > 0 ? nF : 1), 256, 0, st>>>(c->diagblk.p, nF > 0 ? nF : 1, -1);
  if (nvalid > 0) {
    k_block_info<<<nblk(nvalid), 256, 0, st>>>(keys.p, head.p, bid.p, nvalid, nF, brow.p, c->bcol.p, c->diagblk.p);
    // blockptr[gi] = first block whose row >= gi
    thrust::lower_bound(pol, thrust::device_ptr<int>(brow.p), thrust::device_ptr<int>(brow.p + nb), thrust::counting_iterator<int>(0),
                        thrust::counting_iterator<int>(nF + 1), thrust::device_ptr<int>(c->blockptr.p));
    k_fill_maps<<<nblk(nvalid), 256, 0, st>>>(vals.p, bid.p, brow.p, c->blockptr.p, nvalid, npe, dim, c->ld, c->smap.p, c->sstride.p);
    k_inner<<<nblk((int64_t)nb * dim * dim), 256, 0, st>>>(c->blockptr.p, brow.p, c->bcol.p, nb, dim, c->inner.p);
    k_block_transpose<<<nblk(nb), 256, 0, st>>>(c->blockptr.p, brow.p, c->bcol.p, nb, c->bT.p);
    {
      // head / bid are free again: reuse them as flag / position of the mirror list
      k_mirror_flags<<<nblk(nb), 256, 0, st>>>(brow.p, c->bcol.p, nb, head.p);
      thrust::exclusive_scan(pol, thrust::device_ptr<int>(head.p), thrust::device_ptr<int>(head.p + nb), thrust::device_ptr<int>(bid.p));
      int last_pos = 0, last_flag = 0;
      MFB_CUDA(cudaMemcpyAsync(&last_pos, bid.p + (nb - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
      MFB_CUDA(cudaMemcpyAsync(&last_flag, head.p + (nb - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
      MFB_CUDA(cudaStreamSynchronize(st));
      c->n_mirror = last_pos + last_flag;
      if ((rc = c->mirror.alloc(c->n_mirror > 0 ? c->n_mirror : 1))) return rc;
      k_mirror_fill<<<nblk(nb), 256, 0, st>>>(c->blockptr.p, brow.p, c->bcol.p, c->bT.p, head.p, bid.p, nb, dim, c->mirror.p);
    }
    if (c->kind == MFB_TET) {
      // v4 assembly: number the upper-triangle blocks, element->block map, finalize list
      DBuf<int> uflag, ub;
      if ((rc = uflag.alloc(nb)) || (rc = ub.alloc(nb)) || (rc = c->bmap.alloc((size_t)6 * c->ld)) || (rc = c->fcode.alloc(nb))) return rc;
      k_upper_flags<<<nblk(nb), 256, 0, st>>>(brow.p, c->bcol.p, nb, uflag.p);
      thrust::exclusive_scan(pol, thrust::device_ptr<int>(uflag.p), thrust::device_ptr<int>(uflag.p + nb), thrust::device_ptr<int>(ub.p));
      int last_pos = 0, last_flag = 0;
      MFB_CUDA(cudaMemcpyAsync(&last_pos, ub.p + (nb - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
      MFB_CUDA(cudaMemcpyAsync(&last_flag, uflag.p + (nb - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
      MFB_CUDA(cudaStreamSynchronize(st));
      c->nbu = last_pos + last_flag;
      k_fill_int<<<nblk((int64_t)6 * c->ld), 256, 0, st>>>(c->bmap.p, (int64_t)6 * c->ld, -1);
      // vals / bid still hold the sorted pair list (head was reused above, bid was overwritten by the mirror scan: redo it)
      k_heads<<<nblk(nvalid), 256, 0, st>>>(keys.p, nvalid, head.p);
      thrust::inclusive_scan(pol, thrust::device_ptr<int>(head.p), thrust::device_ptr<int>(head.p + nvalid), thrust::device_ptr<int>(bid.p));
      k_fcode_fill<<<nblk(nb), 256, 0, st>>>(brow.p, c->bcol.p, c->bT.p, ub.p, nb, c->fcode.p);
      k_fill_bmap<<<nblk(nvalid), 256, 0, st>>>(vals.p, bid.p, c->fcode.p, nvalid, c->ld, c->bmap.p);
      k_fill_bmap_pinned<<<nblk(nE), 256, 0, st>>>(c->fidx.p, c->diagblk.p, ub.p, nE, c->ld, c->bmap.p);
      MFB_CUDA(cudaStreamSynchronize(st));
      if ((int64_t)c->nbu * 9 > (int64_t)INT_MAX) { g_last_error = "block-CSR exceeds int32"; return MFB_ERR_TOO_LARGE; }
    }
  } else {
    MFB_CUDA(cudaMemsetAsync(c->blockptr.p, 0, ((size_t)nF + 1) * sizeof(int), st));
  }
  k_outer<<<nblk((int64_t)dim * nF + 1), 256, 0, st>>>(c->blockptr.p, nF, dim, c->outer.p);
  MFB_CUDA(cudaStreamSynchronize(st));
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

// ---- deterministic assembly lists (asm_mode 2) ----
__global__ void k_vert_keys(const int* __restrict__ fidx, int nE, int64_t ld, int* keys, int* vals) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (int64_t)nE * 4) return;
  const int e = (int)(q >> 2), k = (int)(q & 3);
  const int f = fidx[k * ld + e];
  keys[q] = f < 0 ? INT_MAX : f;
  vals[q] = (int)q;
}
__global__ void k_head_positions(const int* __restrict__ head, const int* __restrict__ bid, int64_t n, int nb, int* bptr) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (head[i]) bptr[bid[i] - 1] = (int)i;
  if (i == n - 1) bptr[nb] = (int)n;
}

int build_det_lists(mfb_ctx* c) {
  if (c->det_ready) return MFB_OK;
  if (c->kind != MFB_TET) { g_last_error = "asm_mode 2 is implemented for tetrahedra"; return MFB_ERR_UNSUPPORTED; }
  const int nE = c->nE, nF = c->nF, nb = c->nb;
  const int64_t np = (int64_t)nE * 16;
  if (np > (int64_t)INT_MAX) { g_last_error = "asm_mode 2: contribution list exceeds int32"; return MFB_ERR_TOO_LARGE; }
  cudaStream_t st = c->stream;
  auto pol = thrust::cuda::par.on(st);
  int rc;
  {   // matrix: all (e, j, k) with both vertices free, grouped by CSR block, ascending p = 16 e + 4 j + k inside a block
    DBuf<long long> keys; DBuf<int> head, bid;
    if ((rc = keys.alloc(np)) || (rc = c->det_bent.alloc(np)) || (rc = head.alloc(np)) || (rc = bid.alloc(np)) || (rc = c->det_bptr.alloc((size_t)nb + 1))) return rc;
    k_pair_keys<<<nblk(np), 256, 0, st>>>(c->fidx.p, 4, nE, c->ld, nF, keys.p, c->det_bent.p);
    thrust::stable_sort_by_key(pol, thrust::device_ptr<long long>(keys.p), thrust::device_ptr<long long>(keys.p + np), thrust::device_ptr<int>(c->det_bent.p));
    const int64_t nvalid = thrust::lower_bound(pol, thrust::device_ptr<long long>(keys.p), thrust::device_ptr<long long>(keys.p + np), KEY_INVALID) -
                           thrust::device_ptr<long long>(keys.p);
    if (nvalid > 0) {
      k_heads<<<nblk(nvalid), 256, 0, st>>>(keys.p, nvalid, head.p);
      thrust::inclusive_scan(pol, thrust::device_ptr<int>(head.p), thrust::device_ptr<int>(head.p + nvalid), thrust::device_ptr<int>(bid.p));
      k_head_positions<<<nblk(nvalid), 256, 0, st>>>(head.p, bid.p, nvalid, nb, c->det_bptr.p);
    } else {
      MFB_CUDA(cudaMemsetAsync(c->det_bptr.p, 0, ((size_t)nb + 1) * 4, st));
    }
    MFB_CUDA(cudaStreamSynchronize(st));
  }
  {   // vectors: all (e, k) with a free vertex, grouped by free vertex, ascending q = 4 e + k
    const int64_t nq = (int64_t)nE * 4;
    DBuf<int> keys;
    if ((rc = keys.alloc(nq)) || (rc = c->det_vent.alloc(nq)) || (rc = c->det_vptr.alloc((size_t)nF + 1))) return rc;
    k_vert_keys<<<nblk(nq), 256, 0, st>>>(c->fidx.p, nE, c->ld, keys.p, c->det_vent.p);
    thrust::stable_sort_by_key(pol, thrust::device_ptr<int>(keys.p), thrust::device_ptr<int>(keys.p + nq), thrust::device_ptr<int>(c->det_vent.p));
    thrust::lower_bound(pol, thrust::device_ptr<int>(keys.p), thrust::device_ptr<int>(keys.p + nq), thrust::counting_iterator<int>(0),
                        thrust::counting_iterator<int>(nF + 1), thrust::device_ptr<int>(c->det_vptr.p));
    MFB_CUDA(cudaStreamSynchronize(st));
  }
  if ((rc = c->Kelem.alloc((size_t)nE * 54 + 1)) || (rc = c->Velem.alloc((size_t)nE * 24 + 1))) return rc;
  MFB_CUDA(cudaGetLastError());
  c->det_ready = true;
  return MFB_OK;
}

template <typename T>
static int upload(DBuf<T>& d, const std::vector<T>& h, cudaStream_t st) {
  int rc = d.alloc(h.size());
  if (rc) return rc;
  MFB_CUDA(cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, st));
  return MFB_OK;
}

}  // namespace mfb

using namespace mfb;

// Everything about a descriptor that can be checked WITHOUT a device, so that a bad mesh is an error message and not an
// out-of-bounds read in a kernel or a NaN three calls later (the reference checks none of this: igl / Eigen index with the
// values as given).  Runs before the library touches CUDA: tests/test_abi_and_device_math.py calls it on the CPU box.
static int validate_desc(const mfb_mesh_desc* d) {
  const int dim = d->elem_kind == MFB_TRI2D ? 2 : 3, npe = d->elem_kind == MFB_TET ? 4 : 3;
  for (int64_t i = 0; i < (int64_t)dim * d->nV; ++i)
    if (!std::isfinite(d->Vref[i])) { g_last_error = "mfb_create: vertex " + std::to_string(i / dim) + " has a non-finite coordinate"; return MFB_ERR_INVALID; }
  // free_map: -1 = pinned, the other values are the ids 0 .. nF-1, each exactly once (mesh.cpp:25-50 builds it that way)
  int nF = 0;
  for (int v = 0; v < d->nV; ++v) {
    if (d->free_map[v] < -1) { g_last_error = "mfb_create: free_map[" + std::to_string(v) + "] < -1"; return MFB_ERR_INVALID; }
    nF = std::max(nF, d->free_map[v] + 1);
  }
  {
    std::vector<char> seen((size_t)nF, 0);
    int cnt = 0;
    for (int v = 0; v < d->nV; ++v) {
      const int f = d->free_map[v];
      if (f < 0) continue;
      if (seen[f]) { g_last_error = "mfb_create: free id " + std::to_string(f) + " appears twice in free_map"; return MFB_ERR_INVALID; }
      seen[f] = 1; ++cnt;
    }
    if (cnt != nF) { g_last_error = "mfb_create: free_map skips ids (" + std::to_string(cnt) + " free vertices, largest id " + std::to_string(nF - 1) + ")"; return MFB_ERR_INVALID; }
  }
  for (int e = 0; e < d->nE; ++e) {
    const int32_t* t = &d->T[(size_t)e * npe];
    for (int k = 0; k < npe; ++k)
      if (t[k] < 0 || t[k] >= d->nV) { g_last_error = "element references a vertex out of range"; return MFB_ERR_INVALID; }
    if (d->mat_id && (d->mat_id[e] < 0 || d->mat_id[e] >= d->n_mat)) { g_last_error = "mat_id out of range"; return MFB_ERR_INVALID; }
    // rest shape: the inverse of Dm (tets, 2D) / of T^T T (shells) must exist
    double det;
    if (d->elem_kind == MFB_TET) {
      double A[3][3];
      for (int k = 0; k < 3; ++k) for (int l = 0; l < 3; ++l) A[l][k] = d->Vref[3 * (size_t)t[k + 1] + l] - d->Vref[3 * (size_t)t[0] + l];
      det = A[0][0] * (A[1][1] * A[2][2] - A[2][1] * A[1][2]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
            A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
    } else if (d->elem_kind == MFB_TRI2D) {
      const double* a = &d->Vref[2 * (size_t)t[0]]; const double* b = &d->Vref[2 * (size_t)t[1]]; const double* c = &d->Vref[2 * (size_t)t[2]];
      det = (b[0] - a[0]) * (c[1] - a[1]) - (c[0] - a[0]) * (b[1] - a[1]);
    } else {
      const double* a = &d->Vref[3 * (size_t)t[0]]; const double* b = &d->Vref[3 * (size_t)t[1]]; const double* c = &d->Vref[3 * (size_t)t[2]];
      double t0[3], t1[3];
      for (int l = 0; l < 3; ++l) { t0[l] = b[l] - a[l]; t1[l] = c[l] - a[l]; }
      const double g00 = t0[0] * t0[0] + t0[1] * t0[1] + t0[2] * t0[2], g01 = t0[0] * t1[0] + t0[1] * t1[1] + t0[2] * t1[2],
                   g11 = t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2];
      det = g00 * g11 - g01 * g01;
    }
    if (!(std::fabs(det) > 0.0) || !std::isfinite(1.0 / det)) {
      g_last_error = "mfb_create: element " + std::to_string(e) + " is degenerate (zero rest volume / area)";
      return MFB_ERR_INVALID;
    }
  }
  return MFB_OK;
}

extern "C" int mfb_create(mfb_ctx** out, const mfb_mesh_desc* d) {
  if (!out || !d || !d->T || !d->Vref || !d->free_map || d->nV <= 0 || d->nE <= 0 || d->n_mat <= 0 || !d->materials) {
    g_last_error = "mfb_create: invalid descriptor";
    return MFB_ERR_INVALID;
  }
  if (d->elem_kind != MFB_TET && d->elem_kind != MFB_TRI2D && d->elem_kind != MFB_SHELL) { g_last_error = "unsupported elem_kind"; return MFB_ERR_UNSUPPORTED; }
  if (d->elem_kind == MFB_SHELL && (!d->normals || !(d->thickness > 0))) { g_last_error = "MFB_SHELL needs per-face normals and a positive thickness"; return MFB_ERR_INVALID; }
  for (int i = 0; i < d->n_mat; ++i) {
    const int mdl = d->materials[i].model;
    if (mdl < 0 || mdl > 5) { g_last_error = "unknown material model"; return MFB_ERR_INVALID; }
    if (d->elem_kind == MFB_TRI2D && (mdl == MFB_NH || mdl == MFB_FUNG)) {
      g_last_error = "Neohookean/Fung have no 3-vector (2D) form in the reference (src/energies/*.cpp)";
      return MFB_ERR_UNSUPPORTED;
    }
  }
  { const int rv = validate_desc(d); if (rv) return rv; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    g_last_error = "no CUDA device visible: this library has no CPU fallback";
    return MFB_ERR_NO_DEVICE;
  }
  if (d->device < 0 || d->device >= ndev) { g_last_error = "bad device ordinal"; return MFB_ERR_INVALID; }
  const int dim = d->elem_kind == MFB_TRI2D ? 2 : 3;
  MFB_CUDA(cudaSetDevice(d->device));
  mfb_ctx* c = new mfb_ctx();
  c->device = d->device;
  c->kind = d->elem_kind;
  c->dim = dim; c->npe = d->elem_kind == MFB_TET ? 4 : 3; c->N = dim == 3 ? 6 : 3; c->M = dim * dim; c->NP = c->N * (c->N + 1) / 2;
  c->nV = d->nV; c->nE = d->nE;
  c->ld = ((int64_t)d->nE + 31) / 32 * 32;
  const int npe = c->npe, nE = c->nE;
  const int64_t ld = c->ld;
  int nF = 0;
  for (int v = 0; v < d->nV; ++v) nF = d->free_map[v] + 1 > nF ? d->free_map[v] + 1 : nF;
  c->nF = nF;

  // host precompute, component-major
  std::vector<int> hT((size_t)npe * ld, 0), hf((size_t)npe * ld, -1), hmat;
  std::vector<double> hdphi((size_t)dim * dim * ld, 0.0), hvol((size_t)ld, 1.0), hnrm;
  if (d->elem_kind == MFB_SHELL) hnrm.assign((size_t)3 * ld, 0.0);
  for (int e = 0; e < nE; ++e) {
    const int32_t* t = &d->T[(size_t)e * npe];
    for (int k = 0; k < npe; ++k) {
      if (t[k] < 0 || t[k] >= d->nV) { delete c; g_last_error = "element references a vertex out of range"; return MFB_ERR_INVALID; }
      hT[(size_t)k * ld + e] = t[k];
      hf[(size_t)k * ld + e] = d->free_map[t[k]];
    }
    if (d->elem_kind == MFB_SHELL) {
      // TriMesh: vol = |t0 x t1| thickness / 2 (tri_mesh.cpp:38-41); rows 1..2 of dphi = (T^T T)^-1 T^T, T = [t0 t1]
      // (linear_tri3d_dphi_dX.cpp:5-17); components (k-1)*3 + c, k = 1, 2
      const double* a = &d->Vref[3 * (size_t)t[0]]; const double* b = &d->Vref[3 * (size_t)t[1]]; const double* cc = &d->Vref[3 * (size_t)t[2]];
      double t0[3], t1[3];
      for (int l = 0; l < 3; ++l) { t0[l] = b[l] - a[l]; t1[l] = cc[l] - a[l]; }
      const double cr[3] = {t0[1] * t1[2] - t0[2] * t1[1], t0[2] * t1[0] - t0[0] * t1[2], t0[0] * t1[1] - t0[1] * t1[0]};
      hvol[e] = std::sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]) * (d->thickness / 2);
      const double g00 = t0[0] * t0[0] + t0[1] * t0[1] + t0[2] * t0[2], g01 = t0[0] * t1[0] + t0[1] * t1[1] + t0[2] * t1[2],
                   g11 = t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2];
      const double id = 1.0 / (g00 * g11 - g01 * g01);
      for (int l = 0; l < 3; ++l) {
        hdphi[(size_t)l * ld + e] = (g11 * id) * t0[l] + (-g01 * id) * t1[l];
        hdphi[(size_t)(3 + l) * ld + e] = (-g01 * id) * t0[l] + (g00 * id) * t1[l];
        hnrm[(size_t)l * ld + e] = d->normals[3 * (size_t)e + l];
      }
    } else if (dim == 3) {
      const double* a = &d->Vref[3 * (size_t)t[0]]; const double* b = &d->Vref[3 * (size_t)t[1]];
      const double* cc = &d->Vref[3 * (size_t)t[2]]; const double* dd = &d->Vref[3 * (size_t)t[3]];
      double ad[3], bd[3], cd[3];
      for (int l = 0; l < 3; ++l) { ad[l] = a[l] - dd[l]; bd[l] = b[l] - dd[l]; cd[l] = cc[l] - dd[l]; }
      const double cr[3] = {bd[1] * cd[2] - bd[2] * cd[1], bd[2] * cd[0] - bd[0] * cd[2], bd[0] * cd[1] - bd[1] * cd[0]};
      hvol[e] = std::fabs(-(ad[0] * cr[0] + ad[1] * cr[1] + ad[2] * cr[2]) / 6.0);
      double A[3][3];   // Dm, columns = edges from vertex 0
      for (int k = 0; k < 3; ++k) for (int l = 0; l < 3; ++l) A[l][k] = d->Vref[3 * (size_t)t[k + 1] + l] - a[l];
      const double det = A[0][0] * (A[1][1] * A[2][2] - A[2][1] * A[1][2]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                         A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
      const double id = 1.0 / det;
      double inv[3][3];
      inv[0][0] = (A[1][1] * A[2][2] - A[2][1] * A[1][2]) * id;  inv[0][1] = -(A[0][1] * A[2][2] - A[0][2] * A[2][1]) * id;
      inv[0][2] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) * id;  inv[1][0] = -(A[1][0] * A[2][2] - A[1][2] * A[2][0]) * id;
      inv[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) * id;  inv[1][2] = -(A[0][0] * A[1][2] - A[1][0] * A[0][2]) * id;
      inv[2][0] = (A[1][0] * A[2][1] - A[2][0] * A[1][1]) * id;  inv[2][1] = -(A[0][0] * A[2][1] - A[2][0] * A[0][1]) * id;
      inv[2][2] = (A[0][0] * A[1][1] - A[1][0] * A[0][1]) * id;
      for (int k = 0; k < 3; ++k) for (int cix = 0; cix < 3; ++cix) hdphi[(size_t)(3 * k + cix) * ld + e] = inv[k][cix];
    } else {
      const double* a = &d->Vref[2 * (size_t)t[0]]; const double* b = &d->Vref[2 * (size_t)t[1]]; const double* cc = &d->Vref[2 * (size_t)t[2]];
      hvol[e] = (a[0] - cc[0]) * (b[1] - cc[1]) - (a[1] - cc[1]) * (b[0] - cc[0]);
      double A[2][2];
      for (int k = 0; k < 2; ++k) for (int l = 0; l < 2; ++l) A[l][k] = d->Vref[2 * (size_t)t[k + 1] + l] - a[l];
      const double id = 1.0 / (A[0][0] * A[1][1] - A[0][1] * A[1][0]);
      const double inv[2][2] = {{A[1][1] * id, -A[0][1] * id}, {-A[1][0] * id, A[0][0] * id}};
      for (int k = 0; k < 2; ++k) for (int cix = 0; cix < 2; ++cix) hdphi[(size_t)(2 * k + cix) * ld + e] = inv[k][cix];
    }
  }
  int rc = MFB_OK;
  cudaStream_t st = c->stream;
#define TRY(x) do { rc = (x); if (rc) { delete c; return rc; } } while (0)
  { std::vector<double> hv(d->Vref, d->Vref + (size_t)dim * d->nV); TRY(upload(c->vref, hv, st)); MFB_CUDA(cudaStreamSynchronize(st)); }
  {   // element chunks (multiples of the 128-thread CTA) and the vertices each one needs
    const int ctas = (nE + kV4Threads - 1) / kV4Threads;
    int vmax = -1;
    // one chunk per 1.25M elements, 1 .. 8: a rank of an 8-way split (1.26M tets, 0.6 ms of kernels) gains nothing from
    // pipelining a 5 MB upload and pays for the 8 x (copy, event, wait, launch) on the host
    c->nchunks = std::max(1, std::min(8, nE / 1250000));
    if (getenv("MFB_CHUNKS")) c->nchunks = std::max(1, std::min((int)mfb_ctx::kChunks, atoi(getenv("MFB_CHUNKS"))));   // experiments
    for (int k = 0; k < c->nchunks; ++k) {
      c->chunk_e0[k] = (int)((int64_t)ctas * k / c->nchunks) * kV4Threads;
      const int e1 = k + 1 == c->nchunks ? nE : (int)((int64_t)ctas * (k + 1) / c->nchunks) * kV4Threads;
      for (int e = c->chunk_e0[k]; e < e1 && e < nE; ++e)
        for (int j = 0; j < npe; ++j) vmax = std::max(vmax, (int)d->T[(size_t)e * npe + j]);
      c->chunk_vhi[k] = vmax + 1;
    }
    for (int k = c->nchunks; k <= mfb_ctx::kChunks; ++k) c->chunk_e0[k] = nE;
  }
  std::vector<int> hfin;
  if (d->elem_kind == MFB_TET && nF > 0) {   // staged finalize: block-rows grouped by the chunk of the last element that holds them
    std::vector<int> lastc((size_t)nF, 0);
    int k = 0;
    for (int e = 0; e < nE; ++e) {
      while (e >= c->chunk_e0[k + 1]) ++k;
      for (int j = 0; j < npe; ++j) { const int f = hf[(size_t)j * ld + e]; if (f >= 0) lastc[f] = k; }     // e ascends: the last write wins
    }
    for (int q = 0; q <= mfb_ctx::kChunks; ++q) c->fin_ptr[q] = 0;
    for (int f = 0; f < nF; ++f) c->fin_ptr[lastc[f] + 1]++;
    for (int q = 0; q < mfb_ctx::kChunks; ++q) c->fin_ptr[q + 1] += c->fin_ptr[q];
    hfin.resize((size_t)nF);
    std::vector<int> pos(c->fin_ptr, c->fin_ptr + mfb_ctx::kChunks);
    for (int f = 0; f < nF; ++f) hfin[(size_t)pos[lastc[f]]++] = f;
  }
  TRY(upload(c->T, hT, st)); TRY(upload(c->fidx, hf, st)); TRY(upload(c->dphi, hdphi, st)); TRY(upload(c->vol, hvol, st));
  if (d->elem_kind == MFB_SHELL) { TRY(upload(c->nrest, hnrm, st)); TRY(c->xupd.alloc((size_t)3 * d->nV)); MFB_CUDA(cudaMemcpyAsync(c->xupd.p, c->vref.p, (size_t)3 * d->nV * 8, cudaMemcpyDeviceToDevice, st)); }
  if (d->mat_id) {
    hmat.assign((size_t)ld, 0);
    for (int e = 0; e < nE; ++e) {
      if (d->mat_id[e] < 0 || d->mat_id[e] >= d->n_mat) { delete c; g_last_error = "mat_id out of range"; return MFB_ERR_INVALID; }
      hmat[e] = d->mat_id[e];
    }
    TRY(upload(c->matid, hmat, st));
    c->h_matid.assign(d->mat_id, d->mat_id + nE);
  }
  c->h_mat_model.resize(d->n_mat);
  for (int i = 0; i < d->n_mat; ++i) c->h_mat_model[i] = d->materials[i].model;
  c->h_mat_density.resize(d->n_mat);
  for (int i = 0; i < d->n_mat; ++i) c->h_mat_density[i] = d->materials[i].density > 0 ? d->materials[i].density : 1000.0;
  std::vector<MaterialDev> hm(d->n_mat);
  for (int i = 0; i < d->n_mat; ++i) { hm[i].model = d->materials[i].model; hm[i].mu = d->materials[i].mu; hm[i].la = d->materials[i].la; }
  TRY(upload(c->mats, hm, st));
  c->mat0_host = hm[0];
  TRY(build_pattern(c));
  const size_t N = c->N, M = c->M, NP = c->NP;
  TRY(c->R.alloc(M * ld)); TRY(c->s.alloc(N * ld)); TRY(c->la.alloc(N * ld)); TRY(c->ds.alloc(N * ld)); TRY(c->S.alloc(N * ld));
  TRY(c->dSdF.alloc(M * N * ld)); TRY(c->g.alloc(N * ld)); TRY(c->H.alloc(NP * ld)); TRY(c->Hinv.alloc(NP * ld));
  TRY(c->evals.alloc(ld)); TRY(c->gl.alloc(N * ld)); TRY(c->grad_la.alloc(N * ld)); TRY(c->iters.alloc(ld));
  TRY(c->grad_x.alloc((size_t)dim * nF + 1)); TRY(c->rhs.alloc((size_t)dim * nF + 1)); TRY(c->Avals.alloc(c->nnz > 0 ? c->nnz : 1));
  if (c->kind == MFB_TET) { TRY(c->Ablk.alloc((size_t)c->nbu * 9 + 32 * kV4SinkSlots)); TRY(c->gxrh.alloc((size_t)6 * nF + 32 * kV4SinkSlots));   /* + the reduction sinks */ TRY(c->slow_list.alloc((size_t)nE + 1)); }
  if (c->kind == MFB_TET) {
    TRY(c->Ablk2.alloc((size_t)c->nbu * 9 + 32 * kV4SinkSlots)); TRY(c->chunk_cnt.alloc(2 * mfb_ctx::kChunks));
    if (!hfin.empty()) TRY(upload(c->fin_rows, hfin, st));
  }
  TRY(c->counters.alloc(8)); TRY(c->xdev.alloc((size_t)dim * c->nV)); TRY(c->vecF.alloc((size_t)dim * nF + 1));
  TRY(c->partials.alloc((size_t)c->grid() + 1024)); TRY(c->scal.alloc(16));
  TRY(c->soa_tmp.alloc(N * ld)); TRY(c->soa_tmp2.alloc(N * ld));
#undef TRY
  cudaError_t e1 = cudaMallocHost((void**)&c->h_scal, 16 * sizeof(double));
  if (e1 != cudaSuccess) { delete c; g_last_error = "cudaMallocHost failed"; return MFB_ERR_CUDA; }
  // zero all state so that padding lanes hold finite values
  cudaMemsetAsync(c->R.p, 0, M * ld * 8, st); cudaMemsetAsync(c->s.p, 0, N * ld * 8, st); cudaMemsetAsync(c->la.p, 0, N * ld * 8, st);
  cudaMemsetAsync(c->ds.p, 0, N * ld * 8, st); cudaMemsetAsync(c->S.p, 0, N * ld * 8, st); cudaMemsetAsync(c->dSdF.p, 0, M * N * ld * 8, st);
  cudaMemsetAsync(c->g.p, 0, N * ld * 8, st); cudaMemsetAsync(c->H.p, 0, NP * ld * 8, st); cudaMemsetAsync(c->Hinv.p, 0, NP * ld * 8, st);
  cudaMemsetAsync(c->evals.p, 0, ld * 8, st); cudaMemsetAsync(c->gl.p, 0, N * ld * 8, st); cudaMemsetAsync(c->grad_la.p, 0, N * ld * 8, st);
  cudaMemsetAsync(c->iters.p, 0, ld * 4, st); cudaMemsetAsync(c->grad_x.p, 0, ((size_t)dim * nF + 1) * 8, st);
  cudaMemsetAsync(c->rhs.p, 0, ((size_t)dim * nF + 1) * 8, st); cudaMemsetAsync(c->Avals.p, 0, (c->nnz > 0 ? c->nnz : 1) * 8, st);
  cudaMemsetAsync(c->counters.p, 0, 64, st);
  MFB_CUDA(cudaStreamSynchronize(st));
  *out = c;
  return MFB_OK;
}

extern "C" void mfb_destroy(mfb_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  mfb::saddle_free(c);
  mfb::stretch_free(c);
  mfb_newton_free(c);
  mfb::comm_free(c);
  if (c->copy_stream) {
    cudaStreamDestroy(c->copy_stream); cudaEventDestroy(c->ev_in); cudaEventDestroy(c->ev_vec);
    for (int k = 0; k < mfb_ctx::kChunks; ++k) if (c->ev_chunk[k]) cudaEventDestroy(c->ev_chunk[k]);
  }
  for (int k = 0; k < mfb_ctx::kChunks; ++k) if (c->ev_stage[k]) cudaEventDestroy(c->ev_stage[k]);
  if (c->ev_fin) cudaEventDestroy(c->ev_fin);
  if (c->ev_upd) cudaEventDestroy(c->ev_upd);
  for (int k = 0; k < 2; ++k) if (c->ev_zero[k]) cudaEventDestroy(c->ev_zero[k]);
  if (c->fin_stream) cudaStreamDestroy(c->fin_stream);
  if (c->zero_stream) cudaStreamDestroy(c->zero_stream);
  if (c->h_scal) cudaFreeHost(c->h_scal);
  delete c;
}
