// projective_fit.cpp -- host-side least-squares fit of the 14-coefficient projective
// ground->image seed (once per line-scan model; its output is an INPUT of the per-point
// kernels).
//
// Replaces usgscsm::computeBestFitProjectiveTransform (src/EigenUtilities.cpp:14-66), which
// solves the 2N x 14 system with Eigen::colPivHouseholderQr().solve(b).  The same published
// algorithm is used here: Householder QR with column pivoting on the largest remaining
// column norm, numerical rank from |R_kk| > eps * min(m,n) * max|R_kk|, basic solution
// with the rank-deficient tail set to zero.
#include <cmath>
#include <limits>
#include <stdexcept>
#include <utility>
#include <vector>

#include "model.h"

namespace ub200 {

namespace {

struct ColMajor {
  int m, n;
  std::vector<double> a;
  ColMajor(int m_, int n_) : m(m_), n(n_), a((size_t)m_ * n_, 0.0) {}
  double &operator()(int i, int j) { return a[(size_t)j * m + i]; }
  double *col(int j) { return &a[(size_t)j * m]; }
};

double tailNorm2(const double *c, int from, int m) {
  double s = 0.0;
  for (int i = from; i < m; i++) s += c[i] * c[i];
  return s;
}

void leastSquaresColPivQr(ColMajor &A, std::vector<double> &b, std::vector<double> &u) {
  const int m = A.m, n = A.n;
  const int steps = m < n ? m : n;
  std::vector<int> perm(n);
  for (int j = 0; j < n; j++) perm[j] = j;
  std::vector<double> rdiag(steps, 0.0);
  double maxPivot = 0.0;

  for (int k = 0; k < steps; k++) {
    int piv = k;
    double pivNorm = -1.0;
    for (int j = k; j < n; j++) {
      double s = tailNorm2(A.col(j), k, m);
      if (s > pivNorm) {
        pivNorm = s;
        piv = j;
      }
    }
    if (piv != k) {
      for (int i = 0; i < m; i++) std::swap(A(i, k), A(i, piv));
      std::swap(perm[k], perm[piv]);
    }
    double *v = A.col(k);
    const double alpha = v[k];
    const double tail2 = tailNorm2(v, k + 1, m);
    double beta = alpha, tau = 0.0;
    if (tail2 != 0.0) {
      beta = std::sqrt(alpha * alpha + tail2);
      if (alpha >= 0.0) beta = -beta;
      const double scale = alpha - beta;
      for (int i = k + 1; i < m; i++) v[i] /= scale;
      tau = (beta - alpha) / beta;
    }
    v[k] = beta;
    rdiag[k] = beta;
    if (std::fabs(beta) > maxPivot) maxPivot = std::fabs(beta);
    if (tau == 0.0) continue;
    // H = I - tau w w^T with w = (1, v[k+1..]) applied to the trailing columns and to b
    auto reflect = [&](double *c) {
      double s = c[k];
      for (int i = k + 1; i < m; i++) s += v[i] * c[i];
      s *= tau;
      c[k] -= s;
      for (int i = k + 1; i < m; i++) c[i] -= s * v[i];
    };
    for (int j = k + 1; j < n; j++) reflect(A.col(j));
    reflect(b.data());
  }

  const double threshold = std::numeric_limits<double>::epsilon() * (double)steps * maxPivot;
  int rank = 0;
  for (int k = 0; k < steps; k++)
    if (std::fabs(rdiag[k]) > threshold) rank++;

  std::vector<double> y(n, 0.0);
  for (int k = rank - 1; k >= 0; k--) {
    double s = b[k];
    for (int j = k + 1; j < rank; j++) s -= A(k, j) * y[j];
    y[k] = s / A(k, k);
  }
  u.assign(n, 0.0);
  for (int k = 0; k < n; k++) u[perm[k]] = y[k];
}

}  // namespace

void fitProjectiveTransform(int numPts, const double *line, const double *samp, const double *x,
                            const double *y, const double *z, double coeffs[14]) {
  if (numPts < 8)
    throw std::invalid_argument(
        "At least 8 points are needed to fit a 3D-to-2D projective transform.");
  // (u0 + u1 x + u2 y + u3 z)  / (1 + u4 x  + u5 y  + u6 z)  = line
  // (u7 + u8 x + u9 y + u10 z) / (1 + u11 x + u12 y + u13 z) = samp
  // multiplied through by the denominators: linear in u.
  ColMajor M(2 * numPts, 14);
  std::vector<double> b(2 * numPts, 0.0);
  for (int p = 0; p < numPts; p++) {
    const double r = line[p], c = samp[p];
    const int r0 = 2 * p, r1 = 2 * p + 1;
    M(r0, 0) = 1;
    M(r0, 1) = x[p];
    M(r0, 2) = y[p];
    M(r0, 3) = z[p];
    M(r0, 4) = -x[p] * r;
    M(r0, 5) = -y[p] * r;
    M(r0, 6) = -z[p] * r;
    M(r1, 7) = 1;
    M(r1, 8) = x[p];
    M(r1, 9) = y[p];
    M(r1, 10) = z[p];
    M(r1, 11) = -x[p] * c;
    M(r1, 12) = -y[p] * c;
    M(r1, 13) = -z[p] * c;
    b[r0] = r;
    b[r1] = c;
  }
  std::vector<double> u;
  leastSquaresColPivQr(M, b, u);
  for (int i = 0; i < 14; i++) coeffs[i] = u[i];
}

}  // namespace ub200
