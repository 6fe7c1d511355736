// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or executed from the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// CPU (OpenMP, FP64, Eigen-free) restatement of the mixed-fem per-element hot path, phase by phase in the
// reference's own structure (separate passes, materialised Aloc, gather assembly).  PARITY UNPINNED for the
// Eigen-dependent pieces (no golden vectors ship with the reference and it cannot be compiled here: no Eigen,
// SURVEY.md F4/§8c); materials ARE pinned against the reference's generated expressions (tests/golden).
//
// Follows, in order:
//   deps/Bartels/src/linear_tet_dphi_dX.cpp:5-17, deps/Bartels/include/EigenTypes.h:105-131 (inverse33/22)
//   deps/Bartels/src/linear_tri2d_dphi_dX.cpp:5-17
//   deps/libigl/include/igl/volume.cpp:15-32, igl::doublearea
//   src/mesh/tet_mesh.cpp:17-80, src/mesh/tri2d_mesh.cpp:16-77, src/mesh/mesh.cpp:44-102
//   src/utils/sparse_utils.cpp:19-171        (Assembler)
//   src/variables/mixed_stretch.cpp:93-468   (MixedStretch<DIM>)
//   src/utils/psd_fix.h:10-34, src/svd/newton_procrustes.h:87-189
#include "orc_linalg.hpp"
#include "orc_materials.hpp"
#include "orc_polar.hpp"

#include <vector>
#include <map>
#include <string>
#include <chrono>
#include <numeric>
#include <cstdint>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

struct Timer {
  std::map<std::string, double> total_ms;
  std::map<std::string, std::chrono::steady_clock::time_point> t0;
  void start(const std::string& k) { t0[k] = std::chrono::steady_clock::now(); }
  void stop(const std::string& k) {
    auto t1 = std::chrono::steady_clock::now();
    total_ms[k] += std::chrono::duration<double, std::milli>(t1 - t0[k]).count();
  }
};

// Assembler<double,DIM,-1>   sparse_utils.cpp:19-171
struct Assembler {
  int dim = 3, npe = 4;
  std::vector<int> element_ids, multiplicity, offsets, row_offsets;
  std::vector<std::pair<int, int>> global_pairs, local_pairs;
  std::vector<int> outer, inner;      // CSR pattern (RowMajor), size dim*m+1 / nnz
  std::vector<double> values;
  int nrows = 0;

  void init(const int* E, int nE, int npe_, int dim_, const std::vector<int>& free_map) {
    dim = dim_; npe = npe_;
    std::vector<int> ids;
    int cur_id = 0;
    for (int i = 0; i < nE; ++i)
      for (int j = 0; j < npe; ++j) {
        int vj = E[i * npe + j];
        if (vj == -1 || free_map[vj] == -1) continue;
        int id1 = free_map[vj];
        for (int k = 0; k < npe; ++k) {
          int vk = E[i * npe + k];
          if (vk == -1 || free_map[vk] == -1) continue;
          int id2 = free_map[vk];
          element_ids.push_back(i);
          global_pairs.push_back(std::make_pair(id1, id2));
          local_pairs.push_back(std::make_pair(j, k));
          ids.push_back(cur_id++);
        }
      }
    std::stable_sort(ids.begin(), ids.end(), [&](int a, int b) { return global_pairs[a] < global_pairs[b]; });
    {
      std::vector<int> e2(ids.size()); std::vector<std::pair<int, int>> g2(ids.size()), l2(ids.size());
      for (size_t i = 0; i < ids.size(); ++i) { e2[i] = element_ids[ids[i]]; g2[i] = global_pairs[ids[i]]; l2[i] = local_pairs[ids[i]]; }
      element_ids.swap(e2); global_pairs.swap(g2); local_pairs.swap(l2);
    }
    multiplicity.assign(element_ids.size(), 1);
    size_t i = 0; int curr_row = -1;
    while (i < multiplicity.size()) {
      std::pair<int, int> curr_pair = global_pairs[i];
      if (curr_pair.first != curr_row) { row_offsets.push_back((int)i); curr_row = curr_pair.first; }
      while ((i + multiplicity[i]) < multiplicity.size() && curr_pair == global_pairs[i + multiplicity[i]]) ++multiplicity[i];
      offsets.push_back((int)i);
      i += multiplicity[i];
    }
    row_offsets.push_back((int)multiplicity.size());
    // pattern == setFromTriplets of every dim x dim entry of each unique pair (RowMajor, inner ascending)
    int m = -1;
    for (int v : free_map) m = std::max(m, v);
    m += 1;
    nrows = dim * m;
    std::vector<int> blocks_in_row(m, 0);
    for (size_t b = 0; b < offsets.size(); ++b) blocks_in_row[global_pairs[offsets[b]].first]++;
    outer.assign(nrows + 1, 0);
    for (int r = 0; r < m; ++r) for (int a = 0; a < dim; ++a) outer[dim * r + a + 1] = dim * blocks_in_row[r];
    for (int r = 0; r < nrows; ++r) outer[r + 1] += outer[r];
    inner.assign(outer[nrows], 0);
    std::vector<int> seen(m, 0);
    for (size_t b = 0; b < offsets.size(); ++b) {
      const std::pair<int, int>& p = global_pairs[offsets[b]];
      const int bc = seen[p.first]++;
      for (int a = 0; a < dim; ++a) for (int c = 0; c < dim; ++c) inner[outer[dim * p.first + a] + dim * bc + c] = dim * p.second + c;
    }
    values.assign(inner.size(), 1.0);
  }

  // update_matrix   sparse_utils.cpp:120-171 ; blocks[e] is (dim*npe)^2 column-major
  void update_matrix(const std::vector<double>& blocks) {
    const int ld = dim * npe;
    #pragma omp parallel for
    for (long ii = 0; ii < (long)row_offsets.size() - 1; ++ii) {
      int row_beg = row_offsets[ii], row_end = row_offsets[ii + 1];
      int block_col = 0;
      while (row_beg < row_end) {
        const int n = multiplicity[row_beg];
        double local_block[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < n; ++i) {
          const int e = element_ids[row_beg + i];
          const std::pair<int, int>& l = local_pairs[row_beg + i];
          const double* B = &blocks[(size_t)e * ld * ld];
          for (int a = 0; a < dim; ++a) for (int c = 0; c < dim; ++c)
            local_block[a * dim + c] += B[(size_t)(dim * l.second + c) * ld + dim * l.first + a];
        }
        const std::pair<int, int>& g = global_pairs[row_beg];
        for (int a = 0; a < dim; ++a) {
          const int start = outer[dim * g.first + a] + dim * block_col;
          for (int c = 0; c < dim; ++c) values[start + c] = local_block[a * dim + c];
        }
        row_beg += n; ++block_col;
      }
    }
  }
};

// Mesh data the variable pulls (src/mesh/mesh.h:82-105,151-171) for linear tets (dim 3) and 2D triangles (dim 2)
struct Mesh {
  int dim = 3, npe = 4, nV = 0, nE = 0;
  std::vector<int> T;            // nE x npe
  std::vector<double> Vref;      // nV x dim row-major
  std::vector<int> free_map;     // nV
  std::vector<double> vols;      // nE
  std::vector<double> dphi;      // nE x npe x dim : dphi[e][k][c]
  std::vector<int> mat_id;       // nE
  std::vector<Material> mats;
  int nfree = 0;
  // shells (TriMesh, src/mesh/tri_mesh.cpp): triangles in 3D, dim = 3, npe = 3
  bool shell = false;
  double thickness = 1e-3;
  std::vector<double> Nrest;     // nE x 3 per-face rest normals (TriMesh::N_)
  std::vector<double> Jloc0, Jloc;   // nE x (9 x 9), column-major: entry (row j, column 3k+l) at [ (3k+l)*9 + j ]

  void init() {
    nfree = 0;
    for (int v : free_map) nfree = std::max(nfree, v + 1);
    vols.resize(nE); dphi.resize((size_t)nE * npe * dim);
    if (shell) { init_shell(); return; }
    for (int e = 0; e < nE; ++e) {
      const int* t = &T[(size_t)e * npe];
      if (dim == 3) {
        const double* a = &Vref[3 * t[0]]; const double* b = &Vref[3 * t[1]];
        const double* c = &Vref[3 * t[2]]; const double* d = &Vref[3 * t[3]];
        // igl::volume : -(a-d).dot((b-d).cross(c-d))/6 ; tet_mesh.cpp:31-34 takes |.|
        double ad[3], bd[3], cd[3];
        for (int l = 0; l < 3; ++l) { ad[l] = a[l] - d[l]; bd[l] = b[l] - d[l]; cd[l] = c[l] - d[l]; }
        double cr[3] = {bd[1] * cd[2] - bd[2] * cd[1], bd[2] * cd[0] - bd[0] * cd[2], bd[0] * cd[1] - bd[1] * cd[0]};
        vols[e] = std::fabs(-(ad[0] * cr[0] + ad[1] * cr[1] + ad[2] * cr[2]) / 6.0);
        // linear_tet_dphi_dX : Dm columns = edges from vertex 0, rows 1..3 of dphi = Dm^-1, row 0 = -column sums
        M3 Dm;
        for (int k = 0; k < 3; ++k) for (int l = 0; l < 3; ++l) Dm(l, k) = Vref[3 * t[k + 1] + l] - a[l];
        M3 inv = inverse33(Dm);
        double* P = &dphi[(size_t)e * 12];
        for (int c2 = 0; c2 < 3; ++c2) {
          double sum = 0;
          for (int k = 0; k < 3; ++k) { P[(k + 1) * 3 + c2] = inv(k, c2); sum += inv(k, c2); }
          P[c2] = -sum;
        }
      } else {
        const double* a = &Vref[2 * t[0]]; const double* b = &Vref[2 * t[1]]; const double* c = &Vref[2 * t[2]];
        // igl::doublearea for 2D triangles: SIGNED rx*sy - ry*sx with r=a-c, s=b-c (doublearea.cpp:33-42,60-67);
        // tri2d_mesh.cpp:32-34 takes it as is: the reference's 2D "volume" is the signed DOUBLE area
        const double r0x = a[0] - c[0], r0y = a[1] - c[1], r1x = b[0] - c[0], r1y = b[1] - c[1];
        vols[e] = r0x * r1y - r0y * r1x;
        M2 Dm;
        for (int k = 0; k < 2; ++k) for (int l = 0; l < 2; ++l) Dm(l, k) = Vref[2 * t[k + 1] + l] - a[l];
        const double det = Dm(0, 0) * Dm(1, 1) - Dm(0, 1) * Dm(1, 0);
        const double id = 1.0 / det;
        M2 inv; inv(0, 0) = Dm(1, 1) * id; inv(1, 0) = -Dm(1, 0) * id; inv(0, 1) = -Dm(0, 1) * id; inv(1, 1) = Dm(0, 0) * id;
        double* P = &dphi[(size_t)e * 6];
        for (int c2 = 0; c2 < 2; ++c2) {
          double sum = 0;
          for (int k = 0; k < 2; ++k) { P[(k + 1) * 2 + c2] = inv(k, c2); sum += inv(k, c2); }
          P[c2] = -sum;
        }
      }
    }
  }
  // TriMesh ctor / volumes / init_jacobian  (tri_mesh.cpp:31-41,71-102 ; linear_tri3d_dphi_dX.cpp:5-17)
  void init_shell() {
    Jloc0.assign((size_t)nE * 81, 0.0);
    for (int e = 0; e < nE; ++e) {
      const int* t = &T[(size_t)e * 3];
      const double* a = &Vref[3 * t[0]]; const double* b = &Vref[3 * t[1]]; const double* c = &Vref[3 * t[2]];
      double t0[3], t1[3];
      for (int l = 0; l < 3; ++l) { t0[l] = b[l] - a[l]; t1[l] = c[l] - a[l]; }
      // igl::doublearea (3D) = |t0 x t1| ; volumes = doublearea * thickness / 2
      const double cr[3] = {t0[1] * t1[2] - t0[2] * t1[1], t0[2] * t1[0] - t0[0] * t1[2], t0[0] * t1[1] - t0[1] * t1[0]};
      vols[e] = std::sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]) * (thickness / 2);
      // rows 1..2 of dphi = (T^T T)^-1 T^T with T = [t0 t1] (3 x 2), row 0 = -column sums
      const double g00 = t0[0] * t0[0] + t0[1] * t0[1] + t0[2] * t0[2], g01 = t0[0] * t1[0] + t0[1] * t1[1] + t0[2] * t1[2],
                   g11 = t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2];
      const double det = g00 * g11 - g01 * g01, id = 1.0 / det;
      double* P = &dphi[(size_t)e * 9];
      for (int l = 0; l < 3; ++l) {
        P[3 + l] = (g11 * id) * t0[l] + (-g01 * id) * t1[l];
        P[6 + l] = (-g01 * id) * t0[l] + (g00 * id) * t1[l];
        P[l] = -(P[3 + l] + P[6 + l]);
      }
      double* J = &Jloc0[(size_t)e * 81];
      for (int k = 0; k < 3; ++k) for (int cc = 0; cc < 3; ++cc) for (int r = 0; r < 3; ++r) J[(3 * k + r) * 9 + 3 * cc + r] = P[k * 3 + cc];
    }
    Jloc = Jloc0;
  }
  // TriMesh::update_jacobian (tri_mesh.cpp:128-190): Jloc = Jloc0 + N (I - n n^T) dn_dq / |v1 x v2|
  void update_jacobian(const double* x) {
    if (!shell) return;
    #pragma omp parallel for
    for (int e = 0; e < nE; ++e) {
      const int* t = &T[(size_t)e * 3];
      double v1[3], v2[3];
      for (int l = 0; l < 3; ++l) { v1[l] = x[3 * t[1] + l] - x[3 * t[0] + l]; v2[l] = x[3 * t[2] + l] - x[3 * t[0] + l]; }
      double n[3] = {v1[1] * v2[2] - v1[2] * v2[1], v1[2] * v2[0] - v1[0] * v2[2], v1[0] * v2[1] - v1[1] * v2[0]};
      const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
      for (int l = 0; l < 3; ++l) n[l] /= len;
      auto crossmat = [](const double* v, double* m) {   // row-major 3x3
        m[0] = 0; m[1] = -v[2]; m[2] = v[1]; m[3] = v[2]; m[4] = 0; m[5] = -v[0]; m[6] = -v[1]; m[7] = v[0]; m[8] = 0; };
      double dx1[9], dx2[9], dn[27];   // dn: 3 x 9 row-major
      crossmat(v1, dx1); crossmat(v2, dx2);
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) {
        dn[r * 9 + c] = dx2[r * 3 + c] - dx1[r * 3 + c];
        dn[r * 9 + 3 + c] = -dx2[r * 3 + c];
        dn[r * 9 + 6 + c] = dx1[r * 3 + c];
      }
      double Pn[9];
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Pn[r * 3 + c] = (r == c ? 1.0 : 0.0) - n[r] * n[c];
      double PD[27];   // (I - n n^T) dn_dq / l
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 9; ++c) { double a = 0; for (int k = 0; k < 3; ++k) a += Pn[r * 3 + k] * dn[k * 9 + c]; PD[r * 9 + c] = a / len; }
      const double* Nr = &Nrest[(size_t)e * 3];
      const double* J0 = &Jloc0[(size_t)e * 81];
      double* J = &Jloc[(size_t)e * 81];
      // N (9 x 3): row 3c + r has N_c in column r
      for (int col = 0; col < 9; ++col) for (int cc = 0; cc < 3; ++cc) for (int r = 0; r < 3; ++r)
        J[col * 9 + 3 * cc + r] = J0[col * 9 + 3 * cc + r] + Nr[cc] * PD[r * 9 + col];
    }
  }
  static M3 inverse33(const M3& A) {  // EigenTypes.h:105-120
    const double det = A(0, 0) * (A(1, 1) * A(2, 2) - A(2, 1) * A(1, 2)) - A(0, 1) * (A(1, 0) * A(2, 2) - A(1, 2) * A(2, 0)) +
                       A(0, 2) * (A(1, 0) * A(2, 1) - A(1, 1) * A(2, 0));
    const double id = 1 / det;
    M3 r;
    r(0, 0) = (A(1, 1) * A(2, 2) - A(2, 1) * A(1, 2)) * id;
    r(0, 1) = -(A(0, 1) * A(2, 2) - A(0, 2) * A(2, 1)) * id;
    r(0, 2) = (A(0, 1) * A(1, 2) - A(0, 2) * A(1, 1)) * id;
    r(1, 0) = -(A(1, 0) * A(2, 2) - A(1, 2) * A(2, 0)) * id;
    r(1, 1) = (A(0, 0) * A(2, 2) - A(0, 2) * A(2, 0)) * id;
    r(1, 2) = -(A(0, 0) * A(1, 2) - A(1, 0) * A(0, 2)) * id;
    r(2, 0) = (A(1, 0) * A(2, 1) - A(2, 0) * A(1, 1)) * id;
    r(2, 1) = -(A(0, 0) * A(2, 1) - A(2, 0) * A(0, 1)) * id;
    r(2, 2) = (A(0, 0) * A(1, 1) - A(1, 0) * A(0, 1)) * id;
    return r;
  }

  // F = J_ * x : row (M e + d c + r) sums x[d v + r] * dphi(k,c) over the element's vertices in ascending
  // global column order (Eigen RowMajor SpMV), tet_mesh.cpp:17-27,58-80
  void deformation_gradient(const double* x, std::vector<double>& F) const {
    const int M = dim * dim;
    F.resize((size_t)M * nE);
    #pragma omp parallel for
    for (int e = 0; e < nE; ++e) {
      const int* t = &T[(size_t)e * npe];
      int order[4] = {0, 1, 2, 3};
      std::sort(order, order + npe, [&](int a, int b) { return t[a] < t[b]; });
      const double* P = &dphi[(size_t)e * npe * dim];
      for (int c = 0; c < dim; ++c)
        for (int r = 0; r < dim; ++r) {
          double s = 0;
          for (int kk = 0; kk < npe; ++kk) { const int k = order[kk]; s += P[k * dim + c] * x[dim * t[k] + r]; }
          F[(size_t)M * e + dim * c + r] = s;
        }
      if (shell) {    // F += n(x) (x) N  (tri_mesh.cpp:104-126)
        double v1[3], v2[3];
        for (int l = 0; l < 3; ++l) { v1[l] = x[3 * t[1] + l] - x[3 * t[0] + l]; v2[l] = x[3 * t[2] + l] - x[3 * t[0] + l]; }
        double n[3] = {v1[1] * v2[2] - v1[2] * v2[1], v1[2] * v2[0] - v1[0] * v2[2], v1[0] * v2[1] - v1[1] * v2[0]};
        const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) F[(size_t)9 * e + 3 * c + r] += Nrest[(size_t)3 * e + c] * (n[r] / len);
      }
    }
  }
  // out (dim*nfree) = PJW_ * tmp   (PJW_ = P J^T W, mesh.cpp:63) ; tmp is M*nE
  void jacobian_mult(const std::vector<double>& tmp, std::vector<double>& out, double sign) const {
    const int M = dim * dim;
    out.assign((size_t)dim * nfree, 0.0);
    for (int e = 0; e < nE; ++e) {     // serial scatter: deterministic, element order
      const int* t = &T[(size_t)e * npe];
      const double* P = &dphi[(size_t)e * npe * dim];
      for (int k = 0; k < npe; ++k) {
        const int f = free_map[t[k]];
        if (f < 0) continue;
        for (int r = 0; r < dim; ++r) {
          double s = 0;
          if (shell) { const double* J = &Jloc[(size_t)e * 81 + (3 * k + r) * 9]; for (int j = 0; j < 9; ++j) s += J[j] * tmp[(size_t)9 * e + j]; }
          else for (int c = 0; c < dim; ++c) s += P[k * dim + c] * tmp[(size_t)M * e + dim * c + r];
          out[(size_t)dim * f + r] += sign * vols[e] * s;
        }
      }
    }
  }
  // out (M*nE) = PJW_^T * dx
  void jacobian_transpose_mult(const double* dx, std::vector<double>& out) const {
    const int M = dim * dim;
    out.assign((size_t)M * nE, 0.0);
    #pragma omp parallel for
    for (int e = 0; e < nE; ++e) {
      const int* t = &T[(size_t)e * npe];
      const double* P = &dphi[(size_t)e * npe * dim];
      if (shell) {
        for (int j = 0; j < 9; ++j) {
          double s = 0;
          for (int k = 0; k < 3; ++k) { const int f = free_map[t[k]]; if (f >= 0) for (int l = 0; l < 3; ++l) s += Jloc[(size_t)e * 81 + (3 * k + l) * 9 + j] * dx[(size_t)3 * f + l]; }
          out[(size_t)9 * e + j] = vols[e] * s;
        }
        continue;
      }
      for (int c = 0; c < dim; ++c)
        for (int r = 0; r < dim; ++r) {
          double s = 0;
          for (int k = 0; k < npe; ++k) { const int f = free_map[t[k]]; if (f >= 0) s += P[k * dim + c] * dx[(size_t)dim * f + r]; }
          out[(size_t)M * e + dim * c + r] = vols[e] * s;
        }
    }
  }
};

// MixedStretch<DIM>   src/variables/mixed_stretch.{h,cpp}
template <int DIM>
struct MixedStretch {
  static constexpr int N = DIM == 3 ? 6 : 3;
  static constexpr int M = DIM * DIM;
  static constexpr int NPEMAX = DIM + 1;
  static constexpr int LDMAX = DIM * NPEMAX;
  Mesh* mesh = nullptr;
  int nelem = 0;
  std::vector<double> s, ds, la, rhs_, grad, grad_x, grad_la, gl, Jdx, evals;
  std::vector<double> R, S, g, H, Hloc, Hinv, dSdF, Aloc;   // per element, column-major blocks
  std::vector<int> iters;   // procrustes iterations of the last update_rotations
  int status = 0;
  Assembler assembler;
  Timer timer;

  static double sym(int i) { return i < DIM ? 1.0 : 2.0; }
  static double syminv(int i) { return i < DIM ? 1.0 : 0.5; }

  // returns procrustes iteration count
  int polar(double* Rp, double* sp, const double* Fp, bool grads, double* dsdF_NxM, int* st) const {
    if constexpr (DIM == 3) {
      M3 Rm, Fm; std::memcpy(Rm.a, Rp, 72); std::memcpy(Fm.a, Fp, 72);
      Mat<6, 9> Js;
      int it = polar_svd3(Rm, sp, Fm, grads, Js, st);
      std::memcpy(Rp, Rm.a, 72);
      if (grads) std::memcpy(dsdF_NxM, Js.a, sizeof(Js.a));
      return it;
    } else {
      M2 Rm, Fm; std::memcpy(Rm.a, Rp, 32); std::memcpy(Fm.a, Fp, 32);
      Mat<3, 4> Js;
      polar_svd2(Rm, sp, Fm, Js);
      std::memcpy(Rp, Rm.a, 32);
      std::memcpy(dsdF_NxM, Js.a, sizeof(Js.a));
      *st = 0;
      return 0;
    }
  }
  double mat_energy(int e, const double* si) const {
    const Material& m = mesh->mats[mesh->mat_id[e]];
    if constexpr (DIM == 3) return energy6(m, si); else return energy3(m, si);
  }
  void mat_gradient(int e, const double* si, double* gi) const {
    const Material& m = mesh->mats[mesh->mat_id[e]];
    if constexpr (DIM == 3) gradient6(m, si, gi); else gradient3(m, si, gi);
  }
  void mat_hessian(int e, const double* si, double* Hi) const {
    const Material& m = mesh->mats[mesh->mat_id[e]];
    if constexpr (DIM == 3) hessian6(m, si, Hi); else hessian3(m, si, Hi);
  }

  // set dSdF_[i] = Js^T * Sym  (M x N col-major) from Js (N x M col-major)
  void store_dsdf(int i, const double* Js) {
    double* D = &dSdF[(size_t)i * M * N];
    for (int n = 0; n < N; ++n) for (int m = 0; m < M; ++m) D[n * M + m] = Js[m * N + n] * sym(n);
  }

  // reset()   mixed_stretch.cpp:294-338
  void reset() {
    nelem = mesh->nE;
    s.assign((size_t)N * nelem, 0); ds.assign((size_t)N * nelem, 0); la.assign((size_t)N * nelem, 0);
    R.assign((size_t)M * nelem, 0); S.assign((size_t)N * nelem, 0); H.assign((size_t)N * N * nelem, 0);
    g.assign((size_t)N * nelem, 0); dSdF.assign((size_t)M * N * nelem, 0); Hinv.assign((size_t)N * N * nelem, 0);
    Hloc.assign((size_t)N * N * nelem, 0); Aloc.assign((size_t)(DIM * mesh->npe) * (DIM * mesh->npe) * nelem, 0); evals.assign(nelem, 0);
    gl.assign((size_t)N * nelem, 0); grad_la.assign((size_t)N * nelem, 0); grad.assign((size_t)N * nelem, 0);
    iters.assign(nelem, 0);
    assembler = Assembler();
    assembler.init(mesh->T.data(), mesh->nE, mesh->npe, DIM, mesh->free_map);
    for (int i = 0; i < nelem; ++i) for (int k = 0; k < N; ++k) { H[(size_t)i * N * N + k * N + k] = 1.0; Hinv[(size_t)i * N * N + k * N + k] = 1.0; }
    std::vector<double> def_grad;
    mesh->deformation_gradient(mesh->Vref.data(), def_grad);   // x = vec(V^T): interleaved == row-major V
    int st_all = 0;
    #pragma omp parallel for reduction(| : st_all)
    for (int i = 0; i < nelem; ++i) {
      double Js[N * M]; int st = 0;
      double* Ri = &R[(size_t)M * i];
      for (int k = 0; k < M; ++k) Ri[k] = 0; for (int k = 0; k < DIM; ++k) Ri[k * DIM + k] = 1.0;
      iters[i] = polar(Ri, &S[(size_t)N * i], &def_grad[(size_t)M * i], true, Js, &st);
      store_dsdf(i, Js);
      for (int k = 0; k < N; ++k) s[(size_t)N * i + k] = S[(size_t)N * i + k];
      st_all |= st;
    }
    status = st_all;
  }

  // energy()   mixed_stretch.cpp:93-107
  double energy(const double* /*x*/, const double* sv) {
    timer.start("energy");
    double e = 0;
    #pragma omp parallel for reduction(+ : e)
    for (int i = 0; i < nelem; ++i) e += mat_energy(i, &sv[(size_t)N * i]) * mesh->vols[i];
    timer.stop("energy");
    return e;
  }
  // constraint_value()   mixed_stretch.cpp:109-136
  double constraint_value(const double* x, const double* sv) {
    timer.start("constraint");
    std::vector<double> def_grad;
    mesh->deformation_gradient(x, def_grad);
    double e = 0;
    #pragma omp parallel for reduction(+ : e)
    for (int i = 0; i < nelem; ++i) {
      double Rl[M], stmp[N], tmp[N * M]; int st;
      std::memcpy(Rl, &R[(size_t)M * i], sizeof(Rl));
      polar(Rl, stmp, &def_grad[(size_t)M * i], false, tmp, &st);
      double el = 0;
      for (int k = 0; k < N; ++k) el += la[(size_t)N * i + k] * sym(k) * (stmp[k] - sv[(size_t)N * i + k]);
      e += el * mesh->vols[i];
    }
    timer.stop("constraint");
    return e;
  }
  // evaluate_constraint()   mixed_stretch.cpp:345-366
  void evaluate_constraint(const double* x, std::vector<double>& c) {
    c.resize((size_t)N * nelem);
    std::vector<double> def_grad;
    mesh->deformation_gradient(x, def_grad);
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) {
      double Rl[M], stmp[N], tmp[N * M]; int st;
      std::memcpy(Rl, &R[(size_t)M * i], sizeof(Rl));
      polar(Rl, stmp, &def_grad[(size_t)M * i], false, tmp, &st);
      for (int k = 0; k < N; ++k) c[(size_t)N * i + k] = sym(k) * (stmp[k] - s[(size_t)N * i + k]) * mesh->vols[i];
    }
  }
  // update_rotations()   mixed_stretch.cpp:166-185
  void update_rotations(const double* x) {
    timer.start("rotations");
    std::vector<double> def_grad;
    mesh->deformation_gradient(x, def_grad);
    int st_all = 0;
    #pragma omp parallel for reduction(| : st_all)
    for (int i = 0; i < nelem; ++i) {
      double Js[N * M]; int st = 0;
      iters[i] = polar(&R[(size_t)M * i], &S[(size_t)N * i], &def_grad[(size_t)M * i], true, Js, &st);
      store_dsdf(i, Js);
      st_all |= st;
    }
    status = st_all;
    timer.stop("rotations");
  }
  // update_derivatives()   mixed_stretch.cpp:187-243
  void update_derivatives(double dt) {
    const double h2 = dt * dt;
    timer.start("derivatives");
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) {
      const double vol = mesh->vols[i];
      Mat<N, N> Hm, Hi; double ev[N];
      mat_hessian(i, &s[(size_t)N * i], Hm.a);
      for (int k = 0; k < N * N; ++k) Hm.a[k] *= h2;
      psd_fix_invert<N>(Hm, Hi, ev, 1e-8);
      std::memcpy(&Hinv[(size_t)N * N * i], Hi.a, sizeof(Hi.a));
      evals[i] = h2 * ev[N - 1];                 // sic: h2 applied twice (mixed_stretch.cpp:197,200)
      double gi[N]; mat_gradient(i, &s[(size_t)N * i], gi);
      for (int k = 0; k < N; ++k) g[(size_t)N * i + k] = h2 * gi[k];
      for (int c = 0; c < N; ++c) for (int r = 0; r < N; ++r) {
        H[(size_t)N * N * i + c * N + r] = (1.0 / vol) * (syminv(r) * Hm(r, c) * syminv(c));
        Hloc[(size_t)N * N * i + c * N + r] = vol * Hm(r, c);
      }
    }
    timer.stop("derivatives");

    timer.start("Aloc");
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) {
      const double vol = mesh->vols[i];
      const double* D = &dSdF[(size_t)M * N * i];   // M x N
      const double* Hh = &H[(size_t)N * N * i];
      // K = dSdF * H_ * dSdF^T  (M x M)
      double DH[M * N];
      for (int n = 0; n < N; ++n) for (int m = 0; m < M; ++m) { double a = 0; for (int k = 0; k < N; ++k) a += D[k * M + m] * Hh[n * N + k]; DH[n * M + m] = a; }
      double K[M * M];
      for (int c = 0; c < M; ++c) for (int r = 0; r < M; ++r) { double a = 0; for (int k = 0; k < N; ++k) a += DH[k * M + r] * D[k * M + c]; K[c * M + r] = a; }
      // Jloc (M x LD): entry (DIM*c + r, DIM*k + l) = delta_rl * dphi(k,c)   (tet_mesh.cpp:17-27)
      const int NPE = mesh->npe, LD = DIM * NPE;
      const double* P = &mesh->dphi[(size_t)i * NPE * DIM];
      double Jl[M * LDMAX];
      for (int q = 0; q < M * LD; ++q) Jl[q] = 0;
      if (mesh->shell) { for (int q = 0; q < 81; ++q) Jl[q] = mesh->Jloc[(size_t)i * 81 + q]; }   // TriMesh::local_jacobians()
      else for (int k = 0; k < NPE; ++k) for (int c = 0; c < DIM; ++c) for (int r = 0; r < DIM; ++r) Jl[(DIM * k + r) * M + DIM * c + r] = P[k * DIM + c];
      double KJ[M * LDMAX];
      for (int c = 0; c < LD; ++c) for (int r = 0; r < M; ++r) { double a = 0; for (int k = 0; k < M; ++k) a += K[k * M + r] * Jl[c * M + k]; KJ[c * M + r] = a; }
      double* A = &Aloc[(size_t)LD * LD * i];
      for (int c = 0; c < LD; ++c) for (int r = 0; r < LD; ++r) { double a = 0; for (int k = 0; k < M; ++k) a += Jl[r * M + k] * KJ[c * M + k]; A[c * LD + r] = a * (vol * vol); }
    }
    timer.stop("Aloc");

    timer.start("assemble");
    assembler.update_matrix(Aloc);
    timer.stop("assemble");

    std::vector<double> tmp((size_t)M * nelem);
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) {
      const double* D = &dSdF[(size_t)M * N * i];
      for (int m = 0; m < M; ++m) { double a = 0; for (int k = 0; k < N; ++k) a += D[k * M + m] * la[(size_t)N * i + k]; tmp[(size_t)M * i + m] = a; }
    }
    mesh->jacobian_mult(tmp, grad_x, 1.0);
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) for (int k = 0; k < N; ++k) grad[(size_t)N * i + k] = mesh->vols[i] * g[(size_t)N * i + k];
  }
  // update()   mixed_stretch.cpp:158-164
  void update(const double* x, double dt) {
    mesh->update_jacobian(x);      // shells only: NewtonOptimizer::update_system does this before var->update (newton_optimizer.cpp:119-121)
    update_rotations(x);
    update_derivatives(dt);
    evaluate_constraint(x, grad_la);
  }
  // rhs()   mixed_stretch.cpp:245-275
  void rhs() {
    timer.start("rhs");
    std::vector<double> tmp((size_t)M * nelem);
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) {
      const double vol = mesh->vols[i];
      const double* Hh = &H[(size_t)N * N * i];
      const double* D = &dSdF[(size_t)M * N * i];
      double diff[N];
      for (int k = 0; k < N; ++k) diff[k] = sym(k) * (S[(size_t)N * i + k] - s[(size_t)N * i + k]);
      for (int r = 0; r < N; ++r) {
        double a = 0; for (int k = 0; k < N; ++k) a += Hh[k * N + r] * diff[k];
        gl[(size_t)N * i + r] = vol * a + syminv(r) * g[(size_t)N * i + r];
      }
      for (int m = 0; m < M; ++m) { double a = 0; for (int k = 0; k < N; ++k) a += D[k * M + m] * gl[(size_t)N * i + k]; tmp[(size_t)M * i + m] = a; }
    }
    mesh->jacobian_mult(tmp, rhs_, -1.0);
    timer.stop("rhs");
  }
  // solve()   mixed_stretch.cpp:277-292
  void solve(const double* dx) {
    timer.start("solve");
    mesh->jacobian_transpose_mult(dx, Jdx);
    la = gl;
    ds.resize((size_t)N * nelem);
    #pragma omp parallel for
    for (int i = 0; i < nelem; ++i) {
      const double* Hh = &H[(size_t)N * N * i];
      const double* Hi = &Hinv[(size_t)N * N * i];
      const double* D = &dSdF[(size_t)M * N * i];
      double t[N];
      for (int n = 0; n < N; ++n) { double a = 0; for (int m = 0; m < M; ++m) a += D[n * M + m] * Jdx[(size_t)M * i + m]; t[n] = a; }
      for (int r = 0; r < N; ++r) { double a = 0; for (int k = 0; k < N; ++k) a += Hh[k * N + r] * t[k]; la[(size_t)N * i + r] += a; }
      double v[N];
      for (int k = 0; k < N; ++k) v[k] = g[(size_t)N * i + k] - sym(k) * la[(size_t)N * i + k];
      for (int r = 0; r < N; ++r) { double a = 0; for (int k = 0; k < N; ++k) a += Hi[k * N + r] * v[k]; ds[(size_t)N * i + r] = -a; }
    }
    timer.stop("solve");
  }
  // product_* operators   mixed_stretch.cpp:406-468
  void product_hessian(const double* x, double* out) const {
    for (int i = 0; i < nelem; ++i) for (int r = 0; r < N; ++r) { double a = 0; for (int k = 0; k < N; ++k) a += Hloc[(size_t)N * N * i + k * N + r] * x[(size_t)N * i + k]; out[(size_t)N * i + r] += a; }
  }
  void product_hessian_inv(const double* x, double* out) const {
    for (int i = 0; i < nelem; ++i) for (int r = 0; r < N; ++r) { double a = 0; for (int k = 0; k < N; ++k) a += Hinv[(size_t)N * N * i + k * N + r] * x[(size_t)N * i + k]; out[(size_t)N * i + r] += mesh->vols[i] * a; }
  }
  void product_jacobian_mixed(const double* x, double* out) const {
    for (int i = 0; i < nelem; ++i) for (int k = 0; k < N; ++k) out[(size_t)N * i + k] += mesh->vols[i] * sym(k) * x[(size_t)N * i + k];
  }
  void product_jacobian_x(const double* x, double* out, bool transposed) const {
    if (transposed) {   // x is N*nelem, out is dim*nfree
      std::vector<double> tmp((size_t)M * nelem), o;
      for (int i = 0; i < nelem; ++i) { const double* D = &dSdF[(size_t)M * N * i]; for (int m = 0; m < M; ++m) { double a = 0; for (int k = 0; k < N; ++k) a += D[k * M + m] * x[(size_t)N * i + k]; tmp[(size_t)M * i + m] = a; } }
      mesh->jacobian_mult(tmp, o, 1.0);
      for (size_t k = 0; k < o.size(); ++k) out[k] -= o[k];
    } else {           // x is dim*nfree, out is N*nelem
      std::vector<double> tmp;
      mesh->jacobian_transpose_mult(x, tmp);
      for (int i = 0; i < nelem; ++i) { const double* D = &dSdF[(size_t)M * N * i]; for (int n = 0; n < N; ++n) { double a = 0; for (int m = 0; m < M; ++m) a += D[n * M + m] * tmp[(size_t)M * i + m]; out[(size_t)N * i + n] -= a; } }
    }
  }
};

struct Ctx {
  Mesh mesh;
  MixedStretch<3> v3;
  MixedStretch<2> v2;
  int dim = 3;
};

}  // namespace orc

// =========================================================================================================
// C interface for ctypes (tests / bench only)
// =========================================================================================================
using namespace orc;

#define DISPATCH(ctx, call) do { if ((ctx)->dim == 3) { (ctx)->v3.call; } else { (ctx)->v2.call; } } while (0)

extern "C" {

struct orc_material { int32_t model; double mu, la; };

void* orc_create(int dim, int nV, int nE, const int32_t* T, const double* Vref, const int32_t* free_map,
                 const int32_t* mat_id, int n_mat, const orc_material* mats) {
  Ctx* c = new Ctx();
  c->dim = dim;
  Mesh& m = c->mesh;
  m.dim = dim; m.npe = dim + 1; m.nV = nV; m.nE = nE;
  m.T.assign(T, T + (size_t)nE * (dim + 1));
  m.Vref.assign(Vref, Vref + (size_t)nV * dim);
  m.free_map.assign(free_map, free_map + nV);
  m.mat_id.assign(nE, 0);
  if (mat_id) m.mat_id.assign(mat_id, mat_id + nE);
  for (int i = 0; i < n_mat; ++i) m.mats.push_back(Material{mats[i].model, mats[i].mu, mats[i].la});
  m.init();
  c->v3.mesh = &c->mesh; c->v2.mesh = &c->mesh;
  return c;
}
// TriMesh + MixedStretch<3> (tri_mesh.cpp): triangles in 3D with per-face rest normals and a thickness
void* orc_create_shell(int nV, int nE, const int32_t* T, const double* Vref, const int32_t* free_map, const double* normals, double thickness,
                       const int32_t* mat_id, int n_mat, const orc_material* mats) {
  Ctx* c = new Ctx();
  c->dim = 3;
  Mesh& m = c->mesh;
  m.dim = 3; m.npe = 3; m.nV = nV; m.nE = nE; m.shell = true; m.thickness = thickness;
  m.T.assign(T, T + (size_t)nE * 3);
  m.Vref.assign(Vref, Vref + (size_t)nV * 3);
  m.Nrest.assign(normals, normals + (size_t)nE * 3);
  m.free_map.assign(free_map, free_map + nV);
  m.mat_id.assign(nE, 0);
  if (mat_id) m.mat_id.assign(mat_id, mat_id + nE);
  for (int i = 0; i < n_mat; ++i) m.mats.push_back(Material{mats[i].model, mats[i].mu, mats[i].la});
  m.init();
  c->v3.mesh = &c->mesh; c->v2.mesh = &c->mesh;
  return c;
}
void orc_destroy(void* p) { delete (Ctx*)p; }

void orc_reset(void* p) { Ctx* c = (Ctx*)p; DISPATCH(c, reset()); }
void orc_update(void* p, const double* x, double dt) { Ctx* c = (Ctx*)p; DISPATCH(c, update(x, dt)); }
void orc_update_rotations(void* p, const double* x) { Ctx* c = (Ctx*)p; DISPATCH(c, update_rotations(x)); }
void orc_rhs(void* p, double* out) {
  Ctx* c = (Ctx*)p; DISPATCH(c, rhs());
  const std::vector<double>& r = c->dim == 3 ? c->v3.rhs_ : c->v2.rhs_;
  std::memcpy(out, r.data(), r.size() * 8);
}
void orc_solve(void* p, const double* dx) { Ctx* c = (Ctx*)p; DISPATCH(c, solve(dx)); }
double orc_energy(void* p, const double* x, const double* s) { Ctx* c = (Ctx*)p; return c->dim == 3 ? c->v3.energy(x, s) : c->v2.energy(x, s); }
double orc_constraint_value(void* p, const double* x, const double* s) { Ctx* c = (Ctx*)p; return c->dim == 3 ? c->v3.constraint_value(x, s) : c->v2.constraint_value(x, s); }
void orc_evaluate_constraint(void* p, const double* x, double* out) {
  Ctx* c = (Ctx*)p; std::vector<double> v;
  DISPATCH(c, evaluate_constraint(x, v));
  std::memcpy(out, v.data(), v.size() * 8);
}
void orc_product(void* p, int which, const double* x, double* out, int transposed) {
  Ctx* c = (Ctx*)p;
  switch (which) {
    case 0: DISPATCH(c, product_hessian(x, out)); break;
    case 1: DISPATCH(c, product_hessian_inv(x, out)); break;
    case 2: DISPATCH(c, product_jacobian_mixed(x, out)); break;
    case 3: DISPATCH(c, product_jacobian_x(x, out, transposed != 0)); break;
  }
}

int orc_nfree(void* p) { return ((Ctx*)p)->mesh.nfree; }
int orc_status(void* p) { Ctx* c = (Ctx*)p; return c->dim == 3 ? c->v3.status : c->v2.status; }
int64_t orc_nnz(void* p) { Ctx* c = (Ctx*)p; return (int64_t)(c->dim == 3 ? c->v3.assembler.inner.size() : c->v2.assembler.inner.size()); }
void orc_csr(void* p, int32_t* outer, int32_t* inner, double* values) {
  Ctx* c = (Ctx*)p;
  const Assembler& a = c->dim == 3 ? c->v3.assembler : c->v2.assembler;
  if (outer) std::memcpy(outer, a.outer.data(), a.outer.size() * 4);
  if (inner) std::memcpy(inner, a.inner.data(), a.inner.size() * 4);
  if (values) std::memcpy(values, a.values.data(), a.values.size() * 8);
}
// element -> nnz map: out[e*npe*npe + j*npe + k] = value index of entry (0,0) of block (j,k), -1 if pinned;
// stride[e*npe + j] = distance between consecutive rows of that block (= row length of block-row gi)
void orc_scatter_map(void* p, int32_t* out, int32_t* stride) {
  Ctx* c = (Ctx*)p;
  const Assembler& a = c->dim == 3 ? c->v3.assembler : c->v2.assembler;
  const int npe = a.npe, dim = a.dim;
  const int nE = c->mesh.nE;
  for (size_t i = 0; i < (size_t)nE * npe * npe; ++i) out[i] = -1;
  for (size_t i = 0; i < (size_t)nE * npe; ++i) stride[i] = 0;
  for (size_t ii = 0; ii + 1 < a.row_offsets.size(); ++ii) {
    int row_beg = a.row_offsets[ii], row_end = a.row_offsets[ii + 1], block_col = 0;
    while (row_beg < row_end) {
      const int n = a.multiplicity[row_beg];
      const int gi = a.global_pairs[row_beg].first;
      const int start = a.outer[dim * gi] + dim * block_col;
      const int len = a.outer[dim * gi + 1] - a.outer[dim * gi];
      for (int i = 0; i < n; ++i) {
        const int e = a.element_ids[row_beg + i];
        out[(size_t)e * npe * npe + a.local_pairs[row_beg + i].first * npe + a.local_pairs[row_beg + i].second] = start;
        stride[(size_t)e * npe + a.local_pairs[row_beg + i].first] = len;
      }
      row_beg += n; ++block_col;
    }
  }
}

// field access: names follow the reference members.  Returns element count, copies if out != NULL.
int64_t orc_get(void* p, const char* name, double* out) {
  Ctx* c = (Ctx*)p;
  const std::string n(name);
  const std::vector<double>* v = nullptr;
#define FIELD(nm, member) if (n == nm) v = (c->dim == 3) ? &c->v3.member : &c->v2.member;
  FIELD("s", s) FIELD("ds", ds) FIELD("la", la) FIELD("rhs", rhs_) FIELD("grad", grad) FIELD("grad_x", grad_x)
  FIELD("grad_la", grad_la) FIELD("gl", gl) FIELD("Jdx", Jdx) FIELD("evals", evals) FIELD("R", R) FIELD("S", S)
  FIELD("g", g) FIELD("H", H) FIELD("Hloc", Hloc) FIELD("Hinv", Hinv) FIELD("dSdF", dSdF) FIELD("Aloc", Aloc)
#undef FIELD
  if (n == "vols") v = &c->mesh.vols;
  if (n == "dphi") v = &c->mesh.dphi;
  if (!v) return -1;
  if (out) std::memcpy(out, v->data(), v->size() * 8);
  return (int64_t)v->size();
}
int orc_set(void* p, const char* name, const double* in) {
  Ctx* c = (Ctx*)p;
  const std::string n(name);
  std::vector<double>* v = nullptr;
#define FIELD(nm, member) if (n == nm) v = (c->dim == 3) ? &c->v3.member : &c->v2.member;
  FIELD("s", s) FIELD("la", la) FIELD("R", R) FIELD("ds", ds)
#undef FIELD
  if (!v) return -1;
  std::memcpy(v->data(), in, v->size() * 8);
  return 0;
}
void orc_get_iters(void* p, int32_t* out) {
  Ctx* c = (Ctx*)p;
  const std::vector<int>& it = c->dim == 3 ? c->v3.iters : c->v2.iters;
  std::memcpy(out, it.data(), it.size() * 4);
}
// timers: fills ms for the reference's Timer keys in order
// rotations, derivatives, Aloc, assemble, rhs, solve, energy, constraint   
void orc_timers(void* p, double* ms8, int reset) {
  Ctx* c = (Ctx*)p;
  Timer& t = c->dim == 3 ? c->v3.timer : c->v2.timer;
  const char* keys[8] = {"rotations", "derivatives", "Aloc", "assemble", "rhs", "solve", "energy", "constraint"};
  for (int i = 0; i < 8; ++i) ms8[i] = t.total_ms.count(keys[i]) ? t.total_ms[keys[i]] : 0.0;
  if (reset) t.total_ms.clear();
}
int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// ---- low-level probes used by the unit tests ----
int orc_newton_procrustes(double* R9, const double* F9, int compute_gradients, double* dRdF81, double tol, int max_iter, int* status) {
  M3 R, F, I; std::memcpy(R.a, R9, 72); std::memcpy(F.a, F9, 72); I.identity();
  M9 d; d.zero();
  int it = newton_procrustes(R, I, F, compute_gradients != 0, d, tol, max_iter, status);
  std::memcpy(R9, R.a, 72);
  if (compute_gradients && dRdF81) std::memcpy(dRdF81, d.a, sizeof(d.a));
  return it;
}
int orc_polar_svd3(double* R9, double* s6, const double* F9, int compute_gradients, double* dsdF_6x9, int* status) {
  M3 R, F; std::memcpy(R.a, R9, 72); std::memcpy(F.a, F9, 72);
  Mat<6, 9> Js; Js.zero();
  int it = polar_svd3(R, s6, F, compute_gradients != 0, Js, status);
  std::memcpy(R9, R.a, 72);
  if (compute_gradients && dsdF_6x9) std::memcpy(dsdF_6x9, Js.a, sizeof(Js.a));
  return it;
}
void orc_polar_svd2(double* R4, double* s3, const double* F4, double* dsdF_3x4) {
  M2 R, F; std::memcpy(R.a, R4, 32); std::memcpy(F.a, F4, 32);
  Mat<3, 4> Js;
  polar_svd2(R, s3, F, Js);
  std::memcpy(R4, R.a, 32); std::memcpy(dsdF_3x4, Js.a, sizeof(Js.a));
}
void orc_rodrigues(double* R9, const double* omega) { M3 R; rodrigues(R, omega); std::memcpy(R9, R.a, 72); }
void orc_colpiv_solve3(const double* A9, const double* b3, double* x3) {
  M3 A; std::memcpy(A.a, A9, 72); Mat<3, 1> b; std::memcpy(b.a, b3, 24);
  Mat<3, 1> x = colpiv_qr_solve<3, 1>(A, b); std::memcpy(x3, x.a, 24);
}
void orc_sym_eig(int n, const double* A, double* evals, double* V) {
  if (n == 6) { Mat<6, 6> a, v; std::memcpy(a.a, A, sizeof(a.a)); sym_eig<6>(a, evals, v); std::memcpy(V, v.a, sizeof(v.a)); }
  else if (n == 3) { Mat<3, 3> a, v; std::memcpy(a.a, A, sizeof(a.a)); sym_eig<3>(a, evals, v); std::memcpy(V, v.a, sizeof(v.a)); }
}
void orc_psd_fix_invert(int n, double* A, double* Ainv, double* evals, double tol) {
  if (n == 6) { Mat<6, 6> a, ai; std::memcpy(a.a, A, sizeof(a.a)); psd_fix_invert<6>(a, ai, evals, tol); std::memcpy(A, a.a, sizeof(a.a)); std::memcpy(Ainv, ai.a, sizeof(ai.a)); }
  else if (n == 3) { Mat<3, 3> a, ai; std::memcpy(a.a, A, sizeof(a.a)); psd_fix_invert<3>(a, ai, evals, tol); std::memcpy(A, a.a, sizeof(a.a)); std::memcpy(Ainv, ai.a, sizeof(ai.a)); }
}
void orc_svd2(const double* F4, double* sig2, double* U4, double* V4, int flipfix) {
  M2 F, U, V; std::memcpy(F.a, F4, 32);
  if (flipfix) svd2_flipfix(F, sig2, U, V); else svd2(F, sig2, U, V);
  std::memcpy(U4, U.a, 32); std::memcpy(V4, V.a, 32);
}
double orc_material_energy(int model, double mu, double la, int n, const double* s) {
  Material m{model, mu, la};
  return n == 6 ? energy6(m, s) : energy3(m, s);
}
void orc_material_gradient(int model, double mu, double la, int n, const double* s, double* g) {
  Material m{model, mu, la};
  if (n == 6) gradient6(m, s, g); else gradient3(m, s, g);
}
void orc_material_hessian(int model, double mu, double la, int n, const double* s, double* H) {
  Material m{model, mu, la};
  if (n == 6) hessian6(m, s, H); else hessian3(m, s, H);
}

}  // extern "C"
