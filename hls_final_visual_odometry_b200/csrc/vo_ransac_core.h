// Shared by the host stage (vo_egomotion.cpp) and the device scoring kernel (vo_ransac.cu): the arithmetic of one RANSAC
// hypothesis of VisualOdometryMono::estimateMotion -- 8-point fundamental matrix (src/viso_mono.cpp:242-275) and its
// Sampson inlier count (:277-325) -- written once, on fixed-size arrays, so that both sides execute the same double
// operations in the same order (the library is compiled without FMA contraction on either side: -fmad=false for nvcc,
// -ffp-contract=off for the host compiler; sqrt and the divisions are IEEE on both).  A hypothesis scored on the GPU
// therefore gets exactly the inlier count the host would give it, and the winner of the reference's sequential
// "first strictly larger inlier set" rule is the same.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define VO_HD __host__ __device__ inline
#else
#define VO_HD inline
#endif

namespace vo_ransac {

constexpr int kMaxN = 9;

struct Pt { float u1p, v1p, u1c, v1c; };

// ---------------------------------------------------------------------------------------------
// One-sided Jacobi SVD of a row-major m x n matrix held in G (n <= kMaxN columns), in place.
// Output: w[n] singular values in descending order, V (n x n, row-major, orthogonal: a product of plane
// rotations), and (want_g, m * n <= 81 only) G = A*V with its columns in that order, whose column j has norm w[j].
// ---------------------------------------------------------------------------------------------
VO_HD void jacobi_svd_core(double* G, int m, int n, double* w, double* V, bool want_g) {
  double fro = 0;
  for (int i = 0; i < m * n; i++) fro += G[i] * G[i];
  // a column whose norm has dropped to rounding level of the whole matrix is a null direction already;
  // rotating it further only chases noise (this is what ends the sweeps for the 8 x 9 systems of RANSAC)
  const double tiny = fro * 1e-30;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) V[i * n + j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 60; sweep++) {
    bool rotated = false;
    for (int p = 0; p < n - 1; p++) {
      for (int q = p + 1; q < n; q++) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int k = 0; k < m; k++) {
          const double gp = G[k * n + p], gq = G[k * n + q];
          alpha += gp * gp; beta += gq * gq; gamma += gp * gq;
        }
        if (alpha <= tiny || beta <= tiny || gamma * gamma <= 1e-30 * alpha * beta) continue;
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int k = 0; k < m; k++) {
          const double gp = G[k * n + p], gq = G[k * n + q];
          G[k * n + p] = c * gp - s * gq;
          G[k * n + q] = s * gp + c * gq;
        }
        for (int k = 0; k < n; k++) {
          const double vp = V[k * n + p], vq = V[k * n + q];
          V[k * n + p] = c * vp - s * vq;
          V[k * n + q] = s * vp + c * vq;
        }
      }
    }
    if (!rotated) break;
  }
  double nrm[kMaxN];
  int order[kMaxN];
  for (int j = 0; j < n; j++) {
    double s = 0;
    for (int k = 0; k < m; k++) s += G[k * n + j] * G[k * n + j];
    nrm[j] = sqrt(s);
    order[j] = j;
  }
  // stable, descending (src/matrix.cpp:759): insertion sort == std::stable_sort's result
  for (int j = 1; j < n; j++) {
    const int key = order[j];
    int i = j - 1;
    while (i >= 0 && nrm[order[i]] < nrm[key]) { order[i + 1] = order[i]; i--; }
    order[i + 1] = key;
  }
  double V2[kMaxN * kMaxN];
  for (int j = 0; j < n; j++) {
    w[j] = nrm[order[j]];
    for (int k = 0; k < n; k++) V2[k * n + j] = V[k * n + order[j]];
  }
  for (int i = 0; i < n * n; i++) V[i] = V2[i];
  if (want_g) {  // only the 3 x 3 callers need the left vectors
    double G2[kMaxN * kMaxN];
    for (int j = 0; j < n; j++)
      for (int k = 0; k < m; k++) G2[k * n + j] = G[k * n + order[j]];
    for (int i = 0; i < m * n; i++) G[i] = G2[i];
  }
}

struct Svd3 { double U[9], w[3], V[9]; };

// 3x3 SVD with an orthogonal U even when the matrix has rank 2: u2 = u0 x u1.  (The reference's U comes
// from Householder/Givens products and is orthogonal too; its sign of u2 is arbitrary, and so is ours.)
VO_HD Svd3 svd3(const double* A) {
  Svd3 r;
  double G[9];
  for (int i = 0; i < 9; i++) G[i] = A[i];
  jacobi_svd_core(G, 3, 3, r.w, r.V, true);
  for (int j = 0; j < 2; j++) {
    const double inv = r.w[j] > 0 ? 1.0 / r.w[j] : 0.0;
    for (int k = 0; k < 3; k++) r.U[k * 3 + j] = G[k * 3 + j] * inv;
  }
  if (r.w[2] > 1e-9 * r.w[0]) {
    for (int k = 0; k < 3; k++) r.U[k * 3 + 2] = G[k * 3 + 2] / r.w[2];
  } else {
    r.U[0 * 3 + 2] = r.U[1 * 3 + 0] * r.U[2 * 3 + 1] - r.U[2 * 3 + 0] * r.U[1 * 3 + 1];
    r.U[1 * 3 + 2] = r.U[2 * 3 + 0] * r.U[0 * 3 + 1] - r.U[0 * 3 + 0] * r.U[2 * 3 + 1];
    r.U[2 * 3 + 2] = r.U[0 * 3 + 0] * r.U[1 * 3 + 1] - r.U[1 * 3 + 0] * r.U[0 * 3 + 1];
  }
  return r;
}

// U * diag(w0, w1, 0) * V^T : the rank-2 projection of src/viso_mono.cpp:271-274 and :93-97
VO_HD void rank2(const double* M, double* out) {
  const Svd3 s = svd3(M);
  double T[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      T[i * 3 + j] = s.U[i * 3 + 0] * s.w[0] * s.V[j * 3 + 0] + s.U[i * 3 + 1] * s.w[1] * s.V[j * 3 + 1];
  for (int i = 0; i < 9; i++) out[i] = T[i];
}

// one row of the 8-point constraint matrix (products rounded to float, as `m.u1c*m.u1p` on float fields is)
VO_HD void constraint_row(const Pt& m, double* a) {
  a[0] = m.u1c * m.u1p; a[1] = m.u1c * m.v1p; a[2] = m.u1c;
  a[3] = m.v1c * m.u1p; a[4] = m.v1c * m.v1p; a[5] = m.v1c;
  a[6] = m.u1p;         a[7] = m.v1p;         a[8] = 1;
}

// src/viso_mono.cpp:242-275 for the eight points of a RANSAC sample: null vector of the 8 x 9 system, then rank 2
VO_HD void fundamental8(const Pt* p, const int32_t* sample, double F[9]) {
  double G[8 * 9], w[9], V[81];
  for (int i = 0; i < 8; i++) constraint_row(p[sample[i]], &G[i * 9]);
  jacobi_svd_core(G, 8, 9, w, V, false);
  double F0[9];
  for (int k = 0; k < 9; k++) F0[k] = V[k * 9 + 8];
  rank2(F0, F);
}

// src/viso_mono.cpp:277-325: is the Sampson distance of match m below the threshold?
VO_HD bool sampson_inlier(const Pt& m, const double F[9], double thr) {
  const double u1 = m.u1p, v1 = m.v1p, u2 = m.u1c, v2 = m.v1c;
  const double Fx1u = F[0] * u1 + F[1] * v1 + F[2];
  const double Fx1v = F[3] * u1 + F[4] * v1 + F[5];
  const double Fx1w = F[6] * u1 + F[7] * v1 + F[8];
  const double Ftx2u = F[0] * u2 + F[3] * v2 + F[6];
  const double Ftx2v = F[1] * u2 + F[4] * v2 + F[7];
  const double x2tFx1 = u2 * Fx1u + v2 * Fx1v + Fx1w;
  const double d = x2tFx1 * x2tFx1 / (Fx1u * Fx1u + Fx1v * Fx1v + Ftx2u * Ftx2u + Ftx2v * Ftx2v);
  return fabs(d) < thr;
}

// inlier count of the hypothesis drawn from one sample set
VO_HD int32_t score_hypothesis(const Pt* p, int32_t n, const int32_t* sample, double thr) {
  double F[9];
  fundamental8(p, sample, F);
  int32_t c = 0;
  for (int32_t i = 0; i < n; i++) c += sampson_inlier(p[i], F, thr) ? 1 : 0;
  return c;
}

}  // namespace vo_ransac
