// Eigen-free restatement of usgscsm::computeBestFitProjectiveTransform for the oracle
// build (TEST INFRASTRUCTURE -- not part of the product).
//
// The reference solves an over-determined 2N x 14 linear system with
// Eigen::MatrixXd::colPivHouseholderQr().solve(b)
// (/root/reference/src/EigenUtilities.cpp:14-66, call at :58).  Eigen normally comes
// from the `ale/eigen` submodule, which is absent, so the published algorithm
// (Householder QR with column pivoting on the largest remaining column norm, rank
// decided by |R_kk| > eps * size * max|R_kk|, minimum-norm-free basic solution with the
// rank-deficient tail set to zero) is restated here.  Pinned by the known-answer test
// /root/reference/tests/EigenUtilitiesTests.cpp:8-45 (tests/test_oracle_kats.py replays it).
#include <cmath>
#include <limits>
#include <vector>

#include <Error.h>
#include <csm.h>

#include "EigenUtilities.h"

namespace {

// Least-squares solve of A(m x n, row-major) u = b by column-pivoted Householder QR.
void colPivHouseholderSolve(std::vector<double> A, std::vector<double> b, int m, int n,
                            std::vector<double> &u) {
  std::vector<int> perm(n);
  std::vector<double> colNorm2(n);
  for (int j = 0; j < n; j++) {
    perm[j] = j;
    double s = 0.0;
    for (int i = 0; i < m; i++) s += A[i * n + j] * A[i * n + j];
    colNorm2[j] = s;
  }
  const int steps = m < n ? m : n;
  std::vector<double> diag(steps, 0.0);
  double maxPivot = 0.0;
  int rank = steps;
  for (int k = 0; k < steps; k++) {
    // Recompute the trailing column norms exactly (cheap at this size and more robust
    // than down-dating).
    int best = k;
    double bestNorm = -1.0;
    for (int j = k; j < n; j++) {
      double s = 0.0;
      for (int i = k; i < m; i++) s += A[i * n + j] * A[i * n + j];
      colNorm2[j] = s;
      if (s > bestNorm) {
        bestNorm = s;
        best = j;
      }
    }
    if (best != k) {
      for (int i = 0; i < m; i++) std::swap(A[i * n + k], A[i * n + best]);
      std::swap(perm[k], perm[best]);
      std::swap(colNorm2[k], colNorm2[best]);
    }
    // Householder vector for column k, rows k..m-1.
    double alpha = A[k * n + k];
    double tail2 = 0.0;
    for (int i = k + 1; i < m; i++) tail2 += A[i * n + k] * A[i * n + k];
    double beta;
    double tau;
    if (tail2 == 0.0) {
      beta = alpha;
      tau = 0.0;
    } else {
      beta = std::sqrt(alpha * alpha + tail2);
      if (alpha >= 0.0) beta = -beta;
      for (int i = k + 1; i < m; i++) A[i * n + k] /= (alpha - beta);
      tau = (beta - alpha) / beta;
    }
    A[k * n + k] = beta;
    diag[k] = beta;
    if (std::fabs(beta) > maxPivot) maxPivot = std::fabs(beta);
    // Apply H = I - tau v v^T (v_k = 1) to the trailing columns and to b.
    if (tau != 0.0) {
      for (int j = k + 1; j < n; j++) {
        double s = A[k * n + j];
        for (int i = k + 1; i < m; i++) s += A[i * n + k] * A[i * n + j];
        s *= tau;
        A[k * n + j] -= s;
        for (int i = k + 1; i < m; i++) A[i * n + j] -= s * A[i * n + k];
      }
      double s = b[k];
      for (int i = k + 1; i < m; i++) s += A[i * n + k] * b[i];
      s *= tau;
      b[k] -= s;
      for (int i = k + 1; i < m; i++) b[i] -= s * A[i * n + k];
    }
  }
  const double threshold =
      std::numeric_limits<double>::epsilon() * static_cast<double>(steps) * maxPivot;
  rank = 0;
  for (int k = 0; k < steps; k++)
    if (std::fabs(diag[k]) > threshold) rank++;
  // Back substitution on the leading rank x rank triangle.
  std::vector<double> y(n, 0.0);
  for (int k = rank - 1; k >= 0; k--) {
    double s = b[k];
    for (int j = k + 1; j < rank; j++) s -= A[k * n + j] * y[j];
    y[k] = s / A[k * n + k];
  }
  u.assign(n, 0.0);
  for (int k = 0; k < n; k++) u[perm[k]] = y[k];
}

}  // namespace

void usgscsm::computeBestFitProjectiveTransform(std::vector<csm::ImageCoord> const &imagePts,
                                                std::vector<csm::EcefCoord> const &groundPts,
                                                std::vector<double> &transformCoeffs) {
  if (imagePts.size() != groundPts.size())
    throw csm::Error(csm::Error::INVALID_USE, "The number of inputs and outputs must agree.",
                     "computeBestFitProjectiveTransform");
  const int numPts = static_cast<int>(imagePts.size());
  if (numPts < 8)
    throw csm::Error(csm::Error::INVALID_USE,
                     "At least 8 points are needed to fit a 3D-to-2D projective transform.",
                     "computeBestFitProjectiveTransform");
  const int m = 2 * numPts, n = 14;
  std::vector<double> M(static_cast<size_t>(m) * n, 0.0), b(m, 0.0);
  for (int it = 0; it < numPts; it++) {
    const double x = groundPts[it].x, y = groundPts[it].y, z = groundPts[it].z;
    const double r = imagePts[it].line, c = imagePts[it].samp;
    double *r0 = &M[static_cast<size_t>(2 * it) * n];
    double *r1 = r0 + n;
    r0[0] = 1; r0[1] = x; r0[2] = y; r0[3] = z; r0[4] = -x * r; r0[5] = -y * r; r0[6] = -z * r;
    r1[7] = 1; r1[8] = x; r1[9] = y; r1[10] = z; r1[11] = -x * c; r1[12] = -y * c; r1[13] = -z * c;
    b[2 * it] = r;
    b[2 * it + 1] = c;
  }
  colPivHouseholderSolve(M, b, m, n, transformCoeffs);
}
