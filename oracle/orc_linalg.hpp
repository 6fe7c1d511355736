// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or executed from the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// Small fixed-size dense linear algebra used by the CPU restatement of the mixed-fem hot path.
// The reference uses Eigen 3.4 (absent here); the three Eigen algorithms on the path are
// restated from their published algorithms:
//   * ColPivHouseholderQR::solve   (src/svd/newton_procrustes.h:141,186)
//   * SelfAdjointEigenSolver       (src/utils/psd_fix.h:16-33) : Householder tridiagonalisation + implicit QL
//   * JacobiSVD 2x2                (src/svd/svd_eigen.h:12-16)
// PARITY UNPINNED for these three: the reference ships no golden vectors; they are pinned
// against numpy/scipy (solve / eigh / svd) in tests/test_oracle_*.py instead.
#pragma once
#include <cmath>
#include <cstring>
#include <algorithm>
#include <limits>

namespace orc {

// Column-major fixed-size matrix, same storage order as Eigen's default.
template <int R, int C>
struct Mat {
  double a[R * C];
  double& operator()(int i, int j) { return a[j * R + i]; }
  double operator()(int i, int j) const { return a[j * R + i]; }
  double& operator[](int i) { return a[i]; }
  double operator[](int i) const { return a[i]; }
  void zero() { for (int i = 0; i < R * C; ++i) a[i] = 0.0; }
  void identity() { zero(); for (int i = 0; i < (R < C ? R : C); ++i) (*this)(i, i) = 1.0; }
  double norm() const { double s = 0; for (int i = 0; i < R * C; ++i) s += a[i] * a[i]; return std::sqrt(s); }
  double trace() const { double s = 0; for (int i = 0; i < (R < C ? R : C); ++i) s += (*this)(i, i); return s; }
};

template <int R, int K, int C>
inline Mat<R, C> mul(const Mat<R, K>& A, const Mat<K, C>& B) {
  Mat<R, C> out;
  for (int j = 0; j < C; ++j)
    for (int i = 0; i < R; ++i) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += A(i, k) * B(k, j);
      out(i, j) = s;
    }
  return out;
}
template <int R, int C>
inline Mat<C, R> transpose(const Mat<R, C>& A) {
  Mat<C, R> out;
  for (int j = 0; j < C; ++j) for (int i = 0; i < R; ++i) out(j, i) = A(i, j);
  return out;
}
template <int R, int C>
inline Mat<R, C> scaled(const Mat<R, C>& A, double s) {
  Mat<R, C> out; for (int i = 0; i < R * C; ++i) out.a[i] = A.a[i] * s; return out;
}
template <int R, int C>
inline Mat<R, C> add(const Mat<R, C>& A, const Mat<R, C>& B) {
  Mat<R, C> out; for (int i = 0; i < R * C; ++i) out.a[i] = A.a[i] + B.a[i]; return out;
}
template <int R, int C>
inline Mat<R, C> sub(const Mat<R, C>& A, const Mat<R, C>& B) {
  Mat<R, C> out; for (int i = 0; i < R * C; ++i) out.a[i] = A.a[i] - B.a[i]; return out;
}

// ---------------------------------------------------------------------------------------------------------
// Column-pivoted Householder QR solve, A (n x n) X = B (n x m).  Restates Eigen::ColPivHouseholderQR:
// at step k the remaining column of largest squared norm is swapped in, a Householder reflector annihilates
// the sub-column, rank = #pivots with |R_kk| > eps*n*|max pivot|; solve = apply Q^T, back-substitute the
// leading rank x rank triangle, zero the rest, undo the column permutation.
// ---------------------------------------------------------------------------------------------------------
template <int N, int M>
inline Mat<N, M> colpiv_qr_solve(const Mat<N, N>& Ain, const Mat<N, M>& Bin) {
  Mat<N, N> A = Ain;
  Mat<N, M> B = Bin;
  int perm[N];
  for (int i = 0; i < N; ++i) perm[i] = i;
  double maxpivot = 0.0;
  double diag[N];
  for (int k = 0; k < N; ++k) {
    // pick pivot column
    int best = k; double bestn = -1.0;
    for (int j = k; j < N; ++j) {
      double s = 0; for (int i = k; i < N; ++i) s += A(i, j) * A(i, j);
      if (s > bestn) { bestn = s; best = j; }
    }
    if (best != k) {
      for (int i = 0; i < N; ++i) std::swap(A(i, k), A(i, best));
      std::swap(perm[k], perm[best]);
    }
    // Householder: v = x + sign(x0)*|x| e0
    double normx = std::sqrt(bestn);
    double x0 = A(k, k);
    double beta = (x0 >= 0.0) ? -normx : normx;   // resulting R_kk
    if (normx == 0.0) { diag[k] = 0.0; continue; }
    double v[N];
    for (int i = 0; i < N; ++i) v[i] = 0.0;
    v[k] = x0 - beta;
    for (int i = k + 1; i < N; ++i) v[i] = A(i, k);
    double vtv = 0; for (int i = k; i < N; ++i) vtv += v[i] * v[i];
    if (vtv > 0.0) {
      for (int j = k; j < N; ++j) {
        double d = 0; for (int i = k; i < N; ++i) d += v[i] * A(i, j);
        d = 2.0 * d / vtv;
        for (int i = k; i < N; ++i) A(i, j) -= d * v[i];
      }
      for (int j = 0; j < M; ++j) {
        double d = 0; for (int i = k; i < N; ++i) d += v[i] * B(i, j);
        d = 2.0 * d / vtv;
        for (int i = k; i < N; ++i) B(i, j) -= d * v[i];
      }
    }
    A(k, k) = beta;
    for (int i = k + 1; i < N; ++i) A(i, k) = 0.0;
    diag[k] = beta;
    if (std::fabs(beta) > maxpivot) maxpivot = std::fabs(beta);
  }
  const double thresh = maxpivot * std::numeric_limits<double>::epsilon() * N;
  int rank = 0;
  for (int k = 0; k < N; ++k) if (std::fabs(diag[k]) > thresh) ++rank;
  Mat<N, M> X; X.zero();
  for (int j = 0; j < M; ++j) {
    double y[N];
    for (int i = 0; i < N; ++i) y[i] = 0.0;
    for (int i = rank - 1; i >= 0; --i) {
      double s = B(i, j);
      for (int k = i + 1; k < rank; ++k) s -= A(i, k) * y[k];
      y[i] = s / A(i, i);
    }
    for (int i = 0; i < N; ++i) X(perm[i], j) = y[i];
  }
  return X;
}

// ---------------------------------------------------------------------------------------------------------
// Symmetric eigen-decomposition (lower triangle is read, like Eigen::SelfAdjointEigenSolver):
// Householder reduction to tridiagonal form followed by the implicit-shift QL iteration (EISPACK tred2 /
// tql2 formulation), eigenvalues returned ascending, eigenvectors in the columns of V.
// ---------------------------------------------------------------------------------------------------------
template <int N>
inline void sym_eig(const Mat<N, N>& Ain, double* evals, Mat<N, N>& V) {
  double d[N], e[N];
  // symmetrise from the lower triangle
  for (int i = 0; i < N; ++i) for (int j = 0; j <= i; ++j) { V(i, j) = Ain(i, j); V(j, i) = Ain(i, j); }
  // --- tred2 ---
  for (int j = 0; j < N; ++j) d[j] = V(N - 1, j);
  for (int i = N - 1; i > 0; --i) {
    double scale = 0.0, h = 0.0;
    for (int k = 0; k < i; ++k) scale += std::fabs(d[k]);
    if (scale == 0.0) {
      e[i] = d[i - 1];
      for (int j = 0; j < i; ++j) { d[j] = V(i - 1, j); V(i, j) = 0.0; V(j, i) = 0.0; }
    } else {
      for (int k = 0; k < i; ++k) { d[k] /= scale; h += d[k] * d[k]; }
      double f = d[i - 1];
      double g = std::sqrt(h);
      if (f > 0) g = -g;
      e[i] = scale * g;
      h -= f * g;
      d[i - 1] = f - g;
      for (int j = 0; j < i; ++j) e[j] = 0.0;
      for (int j = 0; j < i; ++j) {
        f = d[j];
        V(j, i) = f;
        g = e[j] + V(j, j) * f;
        for (int k = j + 1; k <= i - 1; ++k) { g += V(k, j) * d[k]; e[k] += V(k, j) * f; }
        e[j] = g;
      }
      f = 0.0;
      for (int j = 0; j < i; ++j) { e[j] /= h; f += e[j] * d[j]; }
      double hh = f / (h + h);
      for (int j = 0; j < i; ++j) e[j] -= hh * d[j];
      for (int j = 0; j < i; ++j) {
        f = d[j]; g = e[j];
        for (int k = j; k <= i - 1; ++k) V(k, j) -= (f * e[k] + g * d[k]);
        d[j] = V(i - 1, j);
        V(i, j) = 0.0;
      }
    }
    d[i] = h;
  }
  for (int i = 0; i < N - 1; ++i) {
    V(N - 1, i) = V(i, i);
    V(i, i) = 1.0;
    double h = d[i + 1];
    if (h != 0.0) {
      for (int k = 0; k <= i; ++k) d[k] = V(k, i + 1) / h;
      for (int j = 0; j <= i; ++j) {
        double g = 0.0;
        for (int k = 0; k <= i; ++k) g += V(k, i + 1) * V(k, j);
        for (int k = 0; k <= i; ++k) V(k, j) -= g * d[k];
      }
    }
    for (int k = 0; k <= i; ++k) V(k, i + 1) = 0.0;
  }
  for (int j = 0; j < N; ++j) { d[j] = V(N - 1, j); V(N - 1, j) = 0.0; }
  V(N - 1, N - 1) = 1.0;
  e[0] = 0.0;
  // --- tql2 ---
  for (int i = 1; i < N; ++i) e[i - 1] = e[i];
  e[N - 1] = 0.0;
  double f = 0.0, tst1 = 0.0;
  const double eps = std::numeric_limits<double>::epsilon();
  for (int l = 0; l < N; ++l) {
    tst1 = std::max(tst1, std::fabs(d[l]) + std::fabs(e[l]));
    int m = l;
    while (m < N) { if (std::fabs(e[m]) <= eps * tst1) break; ++m; }
    if (m > l) {
      int iter = 0;
      do {
        ++iter;
        double g = d[l];
        double p = (d[l + 1] - g) / (2.0 * e[l]);
        double r = std::hypot(p, 1.0);
        if (p < 0) r = -r;
        d[l] = e[l] / (p + r);
        d[l + 1] = e[l] * (p + r);
        double dl1 = d[l + 1];
        double h = g - d[l];
        for (int i = l + 2; i < N; ++i) d[i] -= h;
        f += h;
        p = d[m];
        double c = 1.0, c2 = c, c3 = c;
        double el1 = e[l + 1];
        double s = 0.0, s2 = 0.0;
        for (int i = m - 1; i >= l; --i) {
          c3 = c2; c2 = c; s2 = s;
          g = c * e[i];
          h = c * p;
          r = std::hypot(p, e[i]);
          e[i + 1] = s * r;
          s = e[i] / r;
          c = p / r;
          p = c * d[i] - s * g;
          d[i + 1] = h + s * (c * g + s * d[i]);
          for (int k = 0; k < N; ++k) {
            h = V(k, i + 1);
            V(k, i + 1) = s * V(k, i) + c * h;
            V(k, i) = c * V(k, i) - s * h;
          }
        }
        p = -s * s2 * c3 * el1 * e[l] / dl1;
        e[l] = s * p;
        d[l] = c * p;
      } while (std::fabs(e[l]) > eps * tst1 && iter < 200);
    }
    d[l] = d[l] + f;
    e[l] = 0.0;
  }
  // sort ascending
  for (int i = 0; i < N - 1; ++i) {
    int k = i; double p = d[i];
    for (int j = i + 1; j < N; ++j) if (d[j] < p) { k = j; p = d[j]; }
    if (k != i) {
      d[k] = d[i]; d[i] = p;
      for (int j = 0; j < N; ++j) std::swap(V(j, i), V(j, k));
    }
  }
  for (int i = 0; i < N; ++i) evals[i] = d[i];
}

// ---------------------------------------------------------------------------------------------------------
// 2x2 SVD  F = U diag(sig) V^T with sig0 >= sig1 >= 0 (Eigen::JacobiSVD convention: non-negative, sorted
// descending; U, V orthogonal, possibly improper).  Closed form through the symmetric/antisymmetric split.
// ---------------------------------------------------------------------------------------------------------
inline void svd2(const Mat<2, 2>& F, double* sig, Mat<2, 2>& U, Mat<2, 2>& V) {
  const double a = F(0, 0), b = F(0, 1), c = F(1, 0), d = F(1, 1);
  const double E = 0.5 * (a + d), Fm = 0.5 * (a - d), G = 0.5 * (c + b), Hh = 0.5 * (c - b);
  const double Q = std::hypot(E, Hh), Rr = std::hypot(Fm, G);
  double sx = Q + Rr, sy = Q - Rr;
  const double a1 = std::atan2(G, Fm), a2 = std::atan2(Hh, E);
  const double theta = 0.5 * (a2 - a1), phi = 0.5 * (a2 + a1);
  // F = Rot(phi) diag(sx,sy) Rot(theta)
  U(0, 0) = std::cos(phi); U(0, 1) = -std::sin(phi); U(1, 0) = std::sin(phi); U(1, 1) = std::cos(phi);
  // V^T = Rot(theta)  =>  V = Rot(theta)^T
  V(0, 0) = std::cos(theta); V(1, 0) = -std::sin(theta); V(0, 1) = std::sin(theta); V(1, 1) = std::cos(theta);
  if (sy < 0.0) {  // JacobiSVD returns non-negative singular values: push the sign into U's last column
    sy = -sy; U(0, 1) = -U(0, 1); U(1, 1) = -U(1, 1);
  }
  sig[0] = sx; sig[1] = sy;
}

}  // namespace orc
