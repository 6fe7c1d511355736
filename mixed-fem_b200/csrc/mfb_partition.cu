// Element partitioning for the multi-GPU path, HOST code behind the C ABI so that a C++ caller
// can shard a mesh without the Python package.  The reference is a single-address-space OpenMP code: no counterpart.
//
//   method 0  slabs: elements sorted (stably) by centroid along the longest axis of the mesh, cut into parts of equal
//             element count - what the synthetic box grids use;
//   method 1  recursive coordinate bisection on the element centroids (longest axis of the current subset, element
//             counts split in proportion to the number of parts on each side) - for general meshes.
//
// Every rank computes the whole assignment from the global mesh (deterministic, no communication) and keeps its own part:
// locally renumbered vertices (ascending in the global numbering, so local free ids keep the relative order of the global
// free_map_), its elements with the INTERFACE ELEMENTS FIRST (those touching a free vertex shared with another part: the
// update launches them first and exchanges their vertex sums while the interior elements run), the description of
// the shared vertices in both forms the library can use (packed buffer over all shared vertices; per-neighbour lists
// for the point-to-point exchange), and the owner weights of the distributed dot products.
#include "mfb_ctx.cuh"
#include <algorithm>
#include <cstring>
#include <numeric>

using namespace mfb;

struct mfb_partition {
  int dim = 3, npe = 4, nparts = 1, rank = 0, method = 0;
  int32_t n_if_elems = 0, n_shared = 0;
  std::vector<int32_t> elems, vmap, T, free_map, if_local, if_pos, nbr_rank, nbr_count, nbr_local, part_of_elem;
  std::vector<int64_t> free_l2g;
  std::vector<double> V, dot_weight;
};

namespace {

// element -> part, slabs along the longest axis (same rule as numpy's array_split of the stably sorted centroids)
void assign_slabs(int dim, int npe, int nV, int nE, const int32_t* T, const double* V, int nparts, std::vector<int32_t>& part) {
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int v = 0; v < nV; ++v)
    for (int l = 0; l < dim; ++l) { lo[l] = std::min(lo[l], V[(size_t)v * dim + l]); hi[l] = std::max(hi[l], V[(size_t)v * dim + l]); }
  int axis = 0;
  for (int l = 1; l < dim; ++l) if (hi[l] - lo[l] > hi[axis] - lo[axis]) axis = l;
  std::vector<double> cen((size_t)nE);
  for (int e = 0; e < nE; ++e) {
    double s = 0.0;
    for (int k = 0; k < npe; ++k) s += V[(size_t)T[(size_t)e * npe + k] * dim + axis];
    cen[e] = s / npe;
  }
  std::vector<int32_t> order((size_t)nE);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return cen[a] < cen[b]; });
  const int base = nE / nparts, extra = nE % nparts;
  size_t pos = 0;
  for (int q = 0; q < nparts; ++q) {
    const int cnt = base + (q < extra ? 1 : 0);
    for (int i = 0; i < cnt; ++i) part[order[pos++]] = q;
  }
}

void rcb(std::vector<int32_t>& idx, size_t b, size_t e, int q0, int nq, int dim, const std::vector<double>& cen, std::vector<int32_t>& part) {
  if (nq == 1) { for (size_t i = b; i < e; ++i) part[idx[i]] = q0; return; }
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (size_t i = b; i < e; ++i)
    for (int l = 0; l < dim; ++l) { const double c = cen[(size_t)idx[i] * dim + l]; lo[l] = std::min(lo[l], c); hi[l] = std::max(hi[l], c); }
  int axis = 0;
  for (int l = 1; l < dim; ++l) if (hi[l] - lo[l] > hi[axis] - lo[axis]) axis = l;
  const int nl = nq / 2;
  const size_t mid = b + (size_t)((double)(e - b) * nl / nq + 0.5);
  std::nth_element(idx.begin() + b, idx.begin() + mid, idx.begin() + e, [&](int32_t x, int32_t y) {
    const double cx = cen[(size_t)x * dim + axis], cy = cen[(size_t)y * dim + axis];
    return cx < cy || (cx == cy && x < y);
  });
  rcb(idx, b, mid, q0, nl, dim, cen, part);
  rcb(idx, mid, e, q0 + nl, nq - nl, dim, cen, part);
}

void assign_rcb(int dim, int npe, int nE, const int32_t* T, const double* V, int nparts, std::vector<int32_t>& part) {
  std::vector<double> cen((size_t)nE * dim);
  for (int e = 0; e < nE; ++e)
    for (int l = 0; l < dim; ++l) {
      double s = 0.0;
      for (int k = 0; k < npe; ++k) s += V[(size_t)T[(size_t)e * npe + k] * dim + l];
      cen[(size_t)e * dim + l] = s / npe;
    }
  std::vector<int32_t> idx((size_t)nE);
  std::iota(idx.begin(), idx.end(), 0);
  rcb(idx, 0, (size_t)nE, 0, nparts, dim, cen, part);
}

}  // namespace

extern "C" int mfb_partition_create(mfb_partition** out, int32_t dim, int32_t npe, int32_t nV, int32_t nE, const int32_t* T, const double* V,
                                    const int32_t* free_map, int32_t nparts, int32_t rank, int32_t method) {
  if (!out || !T || !V || !free_map || nV <= 0 || nE <= 0 || nparts < 1 || nparts > 64 || rank < 0 || rank >= nparts || dim < 2 || dim > 3 ||
      npe < 3 || npe > 4 || (method != 0 && method != 1)) {
    g_last_error = "mfb_partition_create: bad arguments (1 <= nparts <= 64)";
    return MFB_ERR_INVALID;
  }
  for (int64_t i = 0; i < (int64_t)nE * npe; ++i)
    if (T[i] < 0 || T[i] >= nV) { g_last_error = "mfb_partition_create: vertex id out of range"; return MFB_ERR_INVALID; }
  if (nparts > nE) { g_last_error = "mfb_partition_create: more parts (" + std::to_string(nparts) + ") than elements (" + std::to_string(nE) + ")"; return MFB_ERR_INVALID; }
  {   // free_map: -1 = pinned, otherwise the ids 0 .. nF-1 each exactly once (the same rule as mfb_create)
    int nFg = 0;
    for (int v = 0; v < nV; ++v) {
      if (free_map[v] < -1) { g_last_error = "mfb_partition_create: free_map[" + std::to_string(v) + "] < -1"; return MFB_ERR_INVALID; }
      nFg = std::max(nFg, free_map[v] + 1);
    }
    std::vector<char> seen((size_t)nFg, 0);
    int cnt = 0;
    for (int v = 0; v < nV; ++v) {
      const int f = free_map[v];
      if (f < 0) continue;
      if (seen[f]) { g_last_error = "mfb_partition_create: free id " + std::to_string(f) + " appears twice in free_map"; return MFB_ERR_INVALID; }
      seen[f] = 1; ++cnt;
    }
    if (cnt != nFg) { g_last_error = "mfb_partition_create: free_map skips ids"; return MFB_ERR_INVALID; }
  }
  mfb_partition* p = new mfb_partition();
  p->dim = dim; p->npe = npe; p->nparts = nparts; p->rank = rank; p->method = method;
  p->part_of_elem.assign((size_t)nE, 0);
  if (nparts > 1) {
    if (method == 0) assign_slabs(dim, npe, nV, nE, T, V, nparts, p->part_of_elem);
    else assign_rcb(dim, npe, nE, T, V, nparts, p->part_of_elem);
  }
  // which parts touch each vertex
  std::vector<uint64_t> mask((size_t)nV, 0);
  for (int e = 0; e < nE; ++e)
    for (int k = 0; k < npe; ++k) mask[T[(size_t)e * npe + k]] |= (1ull << p->part_of_elem[e]);
  const uint64_t me = 1ull << rank;
  // shared free vertices, ascending (free ids ascend with the vertex id: mesh.cpp:95-101), their slot and owner
  std::vector<int32_t> slot((size_t)nV, -1);
  int32_t ns = 0;
  for (int v = 0; v < nV; ++v)
    if (free_map[v] >= 0 && (mask[v] & (mask[v] - 1))) slot[v] = ns++;
  p->n_shared = ns;
  // local vertices / elements
  std::vector<int32_t> g2l((size_t)nV, -1);
  for (int v = 0; v < nV; ++v)
    if (mask[v] & me) { g2l[v] = (int32_t)p->vmap.size(); p->vmap.push_back(v); }
  std::vector<int32_t> interior;
  for (int e = 0; e < nE; ++e) {
    if (p->part_of_elem[e] != rank) continue;
    bool iface = false;
    for (int k = 0; k < npe; ++k) iface = iface || slot[T[(size_t)e * npe + k]] >= 0;
    (iface ? p->elems : interior).push_back(e);
  }
  p->n_if_elems = (int32_t)p->elems.size();
  p->elems.insert(p->elems.end(), interior.begin(), interior.end());
  const size_t nEl = p->elems.size(), nVl = p->vmap.size();
  p->T.resize(nEl * npe);
  for (size_t i = 0; i < nEl; ++i)
    for (int k = 0; k < npe; ++k) p->T[i * npe + k] = g2l[T[(size_t)p->elems[i] * npe + k]];
  p->V.resize(nVl * dim);
  p->free_map.assign(nVl, -1);
  int32_t nf = 0;
  for (size_t i = 0; i < nVl; ++i) {
    const int32_t v = p->vmap[i];
    for (int l = 0; l < dim; ++l) p->V[i * dim + l] = V[(size_t)v * dim + l];
    if (free_map[v] >= 0) { p->free_map[i] = nf++; p->free_l2g.push_back(free_map[v]); }
  }
  p->dot_weight.assign((size_t)nf, 1.0);
  // interface description
  struct Ent { int32_t q; int32_t gfree; int32_t lfree; };
  std::vector<Ent> ents;
  for (size_t i = 0; i < nVl; ++i) {
    const int32_t v = p->vmap[i], lf = p->free_map[i];
    if (lf < 0 || slot[v] < 0) continue;
    p->if_local.push_back(lf);
    p->if_pos.push_back(slot[v]);
    const uint64_t m = mask[v];
    int owner = 0;
    while (!((m >> owner) & 1ull)) ++owner;                 // lowest rank touching the vertex
    if (owner != rank) p->dot_weight[lf] = 0.0;
    for (int q = 0; q < nparts; ++q)
      if (q != rank && ((m >> q) & 1ull)) ents.push_back({q, free_map[v], lf});
  }
  std::sort(ents.begin(), ents.end(), [](const Ent& a, const Ent& b) { return a.q < b.q || (a.q == b.q && a.gfree < b.gfree); });
  for (size_t i = 0; i < ents.size(); ++i) {
    if (p->nbr_rank.empty() || p->nbr_rank.back() != ents[i].q) { p->nbr_rank.push_back(ents[i].q); p->nbr_count.push_back(0); }
    p->nbr_count.back() += 1;
    p->nbr_local.push_back(ents[i].lfree);
  }
  *out = p;
  return MFB_OK;
}

extern "C" void mfb_partition_destroy(mfb_partition* p) { delete p; }

extern "C" int mfb_partition_sizes(const mfb_partition* p, int32_t* sizes) {
  if (!p || !sizes) return MFB_ERR_INVALID;
  sizes[0] = (int32_t)p->vmap.size(); sizes[1] = (int32_t)p->elems.size(); sizes[2] = (int32_t)p->free_l2g.size();
  sizes[3] = p->n_if_elems; sizes[4] = p->n_shared; sizes[5] = (int32_t)p->if_local.size();
  sizes[6] = (int32_t)p->nbr_rank.size(); sizes[7] = (int32_t)p->nbr_local.size();
  return MFB_OK;
}

extern "C" int mfb_partition_get(const mfb_partition* p, int what, const void** ptr, int64_t* count) {
  if (!p || !ptr || !count) return MFB_ERR_INVALID;
#define RET(vec) do { *ptr = (vec).data(); *count = (int64_t)(vec).size(); return MFB_OK; } while (0)
  switch (what) {
    case MFB_PART_ELEMS: RET(p->elems);
    case MFB_PART_VMAP: RET(p->vmap);
    case MFB_PART_T: RET(p->T);
    case MFB_PART_V: RET(p->V);
    case MFB_PART_FREE_MAP: RET(p->free_map);
    case MFB_PART_FREE_L2G: RET(p->free_l2g);
    case MFB_PART_IF_LOCAL: RET(p->if_local);
    case MFB_PART_IF_POS: RET(p->if_pos);
    case MFB_PART_DOT_WEIGHT: RET(p->dot_weight);
    case MFB_PART_NBR_RANK: RET(p->nbr_rank);
    case MFB_PART_NBR_COUNT: RET(p->nbr_count);
    case MFB_PART_NBR_LOCAL: RET(p->nbr_local);
    case MFB_PART_ELEM_PART: RET(p->part_of_elem);
  }
#undef RET
  return MFB_ERR_INVALID;
}

int mfb::partition_view(const mfb_partition* p, PartitionView* v) {
  if (!p || !v) return MFB_ERR_INVALID;
  v->nparts = p->nparts; v->rank = p->rank; v->nF = (int)p->free_l2g.size(); v->nE = (int)p->elems.size();
  v->n_if_elems = p->n_if_elems; v->n_shared = p->n_shared;
  v->if_local = &p->if_local; v->if_pos = &p->if_pos; v->dot_weight = &p->dot_weight;
  v->nbr_rank = &p->nbr_rank; v->nbr_count = &p->nbr_count; v->nbr_local = &p->nbr_local;
  return MFB_OK;
}
