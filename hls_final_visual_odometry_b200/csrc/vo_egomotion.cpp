// Host stage F2 of SURVEY.md section 8: monocular egomotion from a bucketed match list.
//
// Replaces, within a stated floating-point tolerance (tests/test_egomotion.py: 1e-6 on the six motion
// parameters against the compiled reference), the reference's
//   VisualOdometryMono::estimateMotion        src/viso_mono.cpp:41-160
//   ::normalizeFeaturePoints                  src/viso_mono.cpp:187-240
//   ::fundamentalMatrix                       src/viso_mono.cpp:242-275
//   ::getInlier                               src/viso_mono.cpp:277-325
//   ::EtoRt / ::triangulateChieral            src/viso_mono.cpp:327-401
//   VisualOdometry::getRandomSample           src/viso.cpp:86-106   (glibc rand() stream after srand(0), :34)
//   VisualOdometry::transformationVectorToMatrix  src/viso.cpp:59-84
// This is tiny dense double-precision math on <= a few hundred points: a host stage by design (the
// reference's FPGA build keeps it on the host too, hls_opencl/host/main2.cpp:398-402).  What is new
// here: no heap Matrix class (fixed-size stack arrays), a one-sided Jacobi SVD instead of the
// reference's Golub-Reinsch routine (src/matrix.cpp:579-760; every quantity used downstream is
// invariant to the sign/rotation freedom of singular vectors, see the notes at each call), a private
// re-entrant copy of glibc's TYPE_3 rand() so engines do not share the process-wide generator, and
// RANSAC hypotheses evaluated by a pool of host threads with the reference's sequential
// "first strictly larger inlier set wins" rule applied afterwards.
//
// Mixed precision is mirrored where the reference has it: match coordinates live in float fields, so the
// centroid shift, the scaling and the products of the constraint matrix round to float exactly where
// src/viso_mono.cpp does.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/vo_b200.h"
#include "vo_ransac_core.h"

namespace {

// ---------------------------------------------------------------------------------------------
// glibc rand(): TYPE_3 additive feedback generator r[i] = r[i-3] + r[i-31], seeded by the Lehmer
// recurrence 16807 * x mod (2^31 - 1), first 310 outputs discarded, result >> 1.
// ---------------------------------------------------------------------------------------------
struct GlibcRand {
  int f, b;  // positions of r[i-3] and r[i-31] in the 31-word ring
  uint32_t ring[31];
  void seed(uint32_t s) {
    if (s == 0) s = 1;
    int32_t word = (int32_t)s;
    ring[0] = s;
    for (int i = 1; i < 31; i++) {
      const int32_t hi = word / 127773, lo = word % 127773;
      word = 16807 * lo - 2836 * hi;
      if (word < 0) word += 2147483647;
      ring[i] = (uint32_t)word;
    }
    f = 3; b = 0;
    for (int i = 0; i < 310; i++) next();
  }
  uint32_t next() {
    ring[f] += ring[b];
    const uint32_t out = ring[f] >> 1;
    f = (f + 1 == 31) ? 0 : f + 1;
    b = (b + 1 == 31) ? 0 : b + 1;
    return out;
  }
};

// One-sided Jacobi SVD, 3 x 3 SVD, rank-2 projection, the 8-point system and the Sampson test live in vo_ransac_core.h (shared
// with the device kernel that scores RANSAC hypotheses).  The general m x n entry keeps its std::vector workspace.
using vo_ransac::kMaxN;
using vo_ransac::Svd3;
using vo_ransac::svd3;

void jacobi_svd(const double* A, int m, int n, std::vector<double>& G, double* w, double* V) {
  G.assign(A, A + (size_t)m * n);
  vo_ransac::jacobi_svd_core(G.data(), m, n, w, V, false);
}

inline void mul3(const double* A, const double* B, double* C) {  // C = A*B, all 3x3 row-major
  double T[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += A[i * 3 + k] * B[k * 3 + j];
      T[i * 3 + j] = s;
    }
  std::memcpy(C, T, sizeof(T));
}
inline void transpose3(const double* A, double* T) {
  double R[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) R[j * 3 + i] = A[i * 3 + j];
  std::memcpy(T, R, sizeof(R));
}
inline double det3(const double* A) {
  return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
}

using vo_ransac::rank2;
using vo_ransac::Pt;

}  // namespace

struct vo_egomotion {
  GlibcRand rng;
  int threads;
  std::vector<int32_t> inliers;
};

namespace {

// src/viso_mono.cpp:187-240.  Float fields, double accumulators: rounding points as in the reference.
bool normalize_points(std::vector<Pt>& p, double Tp[9], double Tc[9]) {
  const double n = (double)p.size();
  double cpu = 0, cpv = 0, ccu = 0, ccv = 0;
  for (const Pt& m : p) { cpu += m.u1p; cpv += m.v1p; ccu += m.u1c; ccv += m.v1c; }
  cpu /= n; cpv /= n; ccu /= n; ccv /= n;
  for (Pt& m : p) {
    m.u1p = (float)(m.u1p - cpu); m.v1p = (float)(m.v1p - cpv);
    m.u1c = (float)(m.u1c - ccu); m.v1c = (float)(m.v1c - ccv);
  }
  double sp = 0, sc = 0;
  for (const Pt& m : p) {
    sp += std::sqrt(m.u1p * m.u1p + m.v1p * m.v1p);  // float product, float sum, float sqrt
    sc += std::sqrt(m.u1c * m.u1c + m.v1c * m.v1c);
  }
  if (std::fabs(sp) < 1e-10 || std::fabs(sc) < 1e-10) return false;
  sp = std::sqrt(2.0) * n / sp;
  sc = std::sqrt(2.0) * n / sc;
  for (Pt& m : p) {
    m.u1p = (float)(m.u1p * sp); m.v1p = (float)(m.v1p * sp);
    m.u1c = (float)(m.u1c * sc); m.v1c = (float)(m.v1c * sc);
  }
  const double tp[9] = {sp, 0, -sp * cpu, 0, sp, -sp * cpv, 0, 0, 1};
  const double tc[9] = {sc, 0, -sc * ccu, 0, sc, -sc * ccv, 0, 0, 1};
  std::memcpy(Tp, tp, sizeof(tp));
  std::memcpy(Tc, tc, sizeof(tc));
  return true;
}

// src/viso_mono.cpp:242-275: 8-point constraint matrix (products rounded to float, as `m.u1c*m.u1p` on float
// fields is), null vector = right singular vector of the smallest singular value, then rank 2.
// The null vector's sign is free; everything downstream (Sampson distance, U*diag*V^T of F) is even in it.
void fundamental(const std::vector<Pt>& p, const int32_t* active, int n_active, double F[9], std::vector<double>& A,
                 std::vector<double>& G) {
  if (n_active == 8) {  // a RANSAC sample: the fixed-size path the device kernel runs too
    vo_ransac::fundamental8(p.data(), active, F);
    return;
  }
  A.resize((size_t)n_active * 9);
  for (int i = 0; i < n_active; i++) vo_ransac::constraint_row(p[active[i]], &A[(size_t)i * 9]);
  double w[9], V[81];
  jacobi_svd(A.data(), n_active, 9, G, w, V);
  double F0[9];
  for (int k = 0; k < 9; k++) F0[k] = V[k * 9 + 8];
  rank2(F0, F);
}

// src/viso_mono.cpp:277-325: Sampson distance below the threshold
void sampson_inliers(const std::vector<Pt>& p, const double F[9], double thr, std::vector<int32_t>& out) {
  out.clear();
  for (int32_t i = 0; i < (int32_t)p.size(); i++)
    if (vo_ransac::sampson_inlier(p[i], F, thr)) out.push_back(i);
}

// src/viso_mono.cpp:363-401: linear triangulation (null vector of the 4x4 system) and chirality count.
// `mirror` != nullptr: the homogeneous points of the SAME rotation with the opposite translation, already triangulated.
// Negating t negates exactly the fourth column of P2 and of every point's 4 x 4 system (the fourth column of P1 is zero), i.e.
// J' = J D with D = diag(1, 1, 1, -1).  One-sided Jacobi is sign-symmetric under that: every rotation angle of a pair that
// involves column 4 flips sign, nothing else changes, V' = D V D -- so the null vector is the mirrored one with its fourth
// component negated (up to the overall sign, which no user of X can see: the chirality products and the x/w ratios are
// invariant to it), and half of the 4 N small SVDs per frame are not run (r2v: they were 70 % of the egomotion tail).  The
// only asymmetry in the arithmetic is the tie zeta == 0 (two columns of exactly equal norm), where either result is a valid
// decomposition; -DVO_EGO_CHECK_MIRROR runs both ways and aborts on any other difference.
int triangulate(const std::vector<Pt>& p, const double K[9], const double R[9], const double t[3], std::vector<double>& X,
                const std::vector<double>* mirror = nullptr) {
  const int n = (int)p.size();
  X.assign((size_t)4 * n, 0.0);
  double P1[12], P2[12], Rt[12];
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) { P1[i * 4 + j] = K[i * 3 + j]; Rt[i * 4 + j] = R[i * 3 + j]; }
    P1[i * 4 + 3] = 0; Rt[i * 4 + 3] = t[i];
  }
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 4; j++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += K[i * 3 + k] * Rt[k * 4 + j];
      P2[i * 4 + j] = s;
    }
  std::vector<double> G;
  double J[16], w[4], V[16];
  if (mirror != nullptr) {
    for (int i = 0; i < n; i++) {
      for (int k = 0; k < 3; k++) X[(size_t)k * n + i] = (*mirror)[(size_t)k * n + i];
      X[(size_t)3 * n + i] = -(*mirror)[(size_t)3 * n + i];
    }
  }
#ifndef VO_EGO_CHECK_MIRROR
  if (mirror == nullptr)
#endif
  for (int i = 0; i < n; i++) {
    for (int j = 0; j < 4; j++) {
      J[0 * 4 + j] = P1[2 * 4 + j] * p[i].u1p - P1[0 * 4 + j];
      J[1 * 4 + j] = P1[2 * 4 + j] * p[i].v1p - P1[1 * 4 + j];
      J[2 * 4 + j] = P2[2 * 4 + j] * p[i].u1c - P2[0 * 4 + j];
      J[3 * 4 + j] = P2[2 * 4 + j] * p[i].v1c - P2[1 * 4 + j];
    }
    jacobi_svd(J, 4, 4, G, w, V);
#ifdef VO_EGO_CHECK_MIRROR
    if (mirror != nullptr) {  // the mirrored vector must be +- the computed one, bit for bit
      bool same = true, neg = true;
      for (int k = 0; k < 4; k++) {
        same = same && X[(size_t)k * n + i] == V[k * 4 + 3];
        neg = neg && X[(size_t)k * n + i] == -V[k * 4 + 3];
      }
      if (!same && !neg) std::abort();
      continue;
    }
#endif
    for (int k = 0; k < 4; k++) X[(size_t)k * n + i] = V[k * 4 + 3];
  }
  int num = 0;
  for (int i = 0; i < n; i++) {
    double a = 0, b = 0;
    for (int k = 0; k < 4; k++) { a += P1[2 * 4 + k] * X[(size_t)k * n + i]; b += P2[2 * 4 + k] * X[(size_t)k * n + i]; }
    const double x3 = X[(size_t)3 * n + i];
    if (a * x3 > 0 && b * x3 > 0) num++;
  }
  return num;
}

// src/viso_mono.cpp:327-361.  T = U Z U^T and {U W V^T, U W^T V^T} do not depend on the rotation freedom of
// the two equal singular vectors; flipping u2 swaps the two rotations after the determinant fix, so only the
// ORDER in which the four (R, t) candidates are tried can differ from the reference -- and that matters
// only for an exact tie of the chirality counts.
void e_to_rt(const double E[9], const double K[9], const std::vector<Pt>& p, std::vector<double>& X, double R[9], double t[3]) {
  const Svd3 s = svd3(E);
  const double Wm[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1}, Zm[9] = {0, 1, 0, -1, 0, 0, 0, 0, 0};
  double Ut[9], Vt[9], Wt[9], T[9], Ra[9], Rb[9], tmp[9];
  transpose3(s.U, Ut); transpose3(s.V, Vt); transpose3(Wm, Wt);
  mul3(s.U, Zm, tmp); mul3(tmp, Ut, T);
  mul3(s.U, Wm, tmp); mul3(tmp, Vt, Ra);
  mul3(s.U, Wt, tmp); mul3(tmp, Vt, Rb);
  double t0[3] = {T[2 * 3 + 1], T[0 * 3 + 2], T[1 * 3 + 0]};
  if (det3(Ra) < 0) for (double& x : Ra) x = -x;
  if (det3(Rb) < 0) for (double& x : Rb) x = -x;
  int best = 0;
  std::vector<double> Xc, Xprev;
  double tb[3] = {t0[0], t0[1], t0[2]};  // the reference leaves t = +t0 and R, X empty when no candidate has an inlier
  std::memset(R, 0, sizeof(double) * 9);
  X.clear();
  for (int i = 0; i < 4; i++) {
    const double* Ri = (i < 2) ? Ra : Rb;
    const double sg = (i & 1) ? -1.0 : 1.0;
    const double ti[3] = {sg * t0[0], sg * t0[1], sg * t0[2]};
    const int num = (i & 1) ? triangulate(p, K, Ri, ti, Xc, &Xprev) : triangulate(p, K, Ri, ti, Xc);
    if (!(i & 1)) Xprev = Xc;  // candidates come as (R, +t), (R, -t): the second is the mirror image of the first
    if (num > best) {
      best = num;
      X = Xc;
      std::memcpy(R, Ri, sizeof(double) * 9);
      std::memcpy(tb, ti, sizeof(tb));
    }
  }
  std::memcpy(t, tb, sizeof(tb));
}

}  // namespace

namespace {

// VisualOdometry::getRandomSample(N, 8): src/viso.cpp:86-106 (draw without replacement from a shrinking set)
// The reference erases the drawn element from a vector of all N indices; element j of what is left is the j-th smallest index
// not drawn yet, which the (at most seven) earlier draws determine without touching N elements.
void draw_sample(GlibcRand& rng, int32_t N, std::vector<int32_t>& /*pool*/, int32_t out[8]) {
  int32_t taken[8];  // ascending
  for (int i = 0; i < 8; i++) {
    int32_t v = (int32_t)(rng.next() % (uint32_t)(N - i));
    int pos = 0;
    while (pos < i && taken[pos] <= v) { v++; pos++; }
    for (int k = i; k > pos; k--) taken[k] = taken[k - 1];
    taken[pos] = v;
    out[i] = v;
  }
}

// Does a match list reach the RANSAC loop of estimateMotion (N >= 10 and a non-degenerate normalisation,
// src/viso_mono.cpp:45-56)?  Only then does the reference consume rand() values: 8 per iteration.
bool prepare_points(const vo_p_match* matches, int32_t N, std::vector<Pt>& raw, std::vector<Pt>& p, double Tp[9], double Tc[9]) {
  if (N < 10) return false;
  raw.resize(N);
  for (int32_t i = 0; i < N; i++) raw[i] = Pt{matches[i].u1p, matches[i].v1p, matches[i].u1c, matches[i].v1c};
  p = raw;
  return normalize_points(p, Tp, Tc);
}

// src/viso_mono.cpp:41-160 with the RANSAC sample sets given (8 indices per iteration, in the order the
// reference draws them).  Stateless.  Returns false where the reference returns an empty vector.
constexpr int kScoreHere = -2;  // best_k_in: score the hypotheses on the host (any value >= -1: the winner is already known)

bool estimate_motion(const vo_p_match* matches, int32_t N, const vo_mono_params& prm, const int32_t* samples_in, int iters,
                     int threads, std::vector<int32_t>& inliers, double tr[6], int best_k_in = kScoreHere) {
  inliers.clear();
  std::vector<Pt> raw, p;
  const double K[9] = {prm.f, 0, prm.cu, 0, prm.f, prm.cv, 0, 0, 1};
  double Tp[9], Tc[9];
  if (!prepare_points(matches, N, raw, p, Tp, Tc)) return false;
  struct EgView { std::vector<int32_t>& inliers; int threads; } view{inliers, threads};
  EgView* eg = &view;
  const int32_t* samples = samples_in;

  // ---- RANSAC: the hypotheses are independent, so they are scored by a pool of threads; the winner is
  //      the FIRST hypothesis with the largest inlier count == the sequential `if (curr.size() > best.size())`.
  std::vector<int32_t> counts(best_k_in == kScoreHere ? iters : 0, 0);
  auto score = [&](int k0, int k1) {
    std::vector<double> A, G;
    std::vector<int32_t> inl;
    double F[9];
    for (int k = k0; k < k1; k++) {
      fundamental(p, &samples[(size_t)k * 8], 8, F, A, G);
      sampson_inliers(p, F, prm.inlier_threshold, inl);
      counts[k] = (int32_t)inl.size();
    }
  };
  const int nt = std::max(1, std::min(eg->threads, iters / 64 + 1));
  if (best_k_in != kScoreHere) {
    // scored elsewhere (vo_ransac.cu runs the same arithmetic, one GPU thread per hypothesis)
  } else if (nt == 1) {
    score(0, iters);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) th.emplace_back(score, (int)((int64_t)iters * t / nt), (int)((int64_t)iters * (t + 1) / nt));
    for (auto& t : th) t.join();
  }
  int best_k = -1, best_n = 0;
  if (best_k_in != kScoreHere) best_k = (best_k_in < iters) ? best_k_in : -1;
  else
    for (int k = 0; k < iters; k++)
      if (counts[k] > best_n) { best_n = counts[k]; best_k = k; }
  std::vector<double> A, G;
  double F[9];
  if (best_k >= 0) {
    fundamental(p, &samples[(size_t)best_k * 8], 8, F, A, G);
    sampson_inliers(p, F, prm.inlier_threshold, eg->inliers);
  }
  if (eg->inliers.size() < 10) return false;

  // refine on all inliers, denormalise, essential matrix, rank 2 (:85-97)
  fundamental(p, eg->inliers.data(), (int)eg->inliers.size(), F, A, G);
  double Tct[9], Kt[9], E[9], tmp[9];
  transpose3(Tc, Tct); transpose3(K, Kt);
  mul3(Tct, F, tmp); mul3(tmp, Tp, F);
  mul3(Kt, F, tmp); mul3(tmp, K, E);
  rank2(E, E);

  std::vector<double> X;
  double R[9], t[3];
  e_to_rt(E, K, raw, X, R, t);
  if (X.empty()) return false;

  // 3-D points with positive depth (:103-113)
  std::vector<double> d, dist;  // d = n^T (y, z), dist = |x|+|y|+|z|
  const double n0 = std::cos(-prm.pitch), n1 = std::sin(-prm.pitch);
  for (int32_t i = 0; i < N; i++) {
    const double w = X[(size_t)3 * N + i];
    const double x = X[i] / w, y = X[(size_t)N + i] / w, z = X[(size_t)2 * N + i] / w;
    if (z > 0) {
      dist.push_back(std::fabs(x) + std::fabs(y) + std::fabs(z));
      d.push_back(n0 * y + n1 * z);
    }
  }
  const int np = (int)d.size();
  if (np < 10) return false;
  std::vector<double> sorted = dist;
  std::nth_element(sorted.begin(), sorted.begin() + np / 2, sorted.end());
  const double median = sorted[np / 2];
  if (median > prm.motion_threshold) return false;

  // ground-plane distance by kernel density over the candidate heights (:127-151)
  const double sigma = median / 50.0, weight = 1.0 / (2.0 * sigma * sigma);
  double best_sum = 0;
  int best_idx = 0;
  for (int i = 0; i < np; i++) {
    if (d[i] > median / prm.motion_threshold) {
      double sum = 0;
      for (int j = 0; j < np; j++) {
        const double dd = d[j] - d[i];
        sum += std::exp(-dd * dd * weight);
      }
      if (sum > best_sum) { best_sum = sum; best_idx = i; }
    }
  }
  const double scale = prm.height / d[best_idx];
  const double ry = std::asin(R[2]);
  const double rx = std::asin(-R[5] / std::cos(ry));
  const double rz = std::asin(-R[1] / std::cos(ry));
  tr[0] = rx; tr[1] = ry; tr[2] = rz;
  tr[3] = t[0] * scale; tr[4] = t[1] * scale; tr[5] = t[2] * scale;
  return true;
}

}  // namespace

extern "C" {

vo_status vo_egomotion_create(uint32_t seed, int32_t threads, vo_egomotion** out) {
  if (!out) return VO_ERR_INVALID;
  vo_egomotion* eg = new vo_egomotion();
  eg->rng.seed(seed);  // VisualOdometry::VisualOdometry calls srand(0): src/viso.cpp:34
  int hw = (int)std::thread::hardware_concurrency();
  if (hw <= 0) hw = 1;
  eg->threads = threads > 0 ? threads : std::min(hw, 16);
  *out = eg;
  return VO_OK;
}

void vo_egomotion_destroy(vo_egomotion* eg) { delete eg; }

// Draws the sample sets the reference would draw for this match list (advancing the handle's stream) --
// nothing if the list never reaches RANSAC.  samples: int32[ransac_iters * 8]; *n_iters = iterations drawn.
vo_status vo_egomotion_draw_samples(vo_egomotion* eg, const vo_p_match* matches, int32_t n, const vo_mono_params* prm,
                                    int32_t* samples, int32_t* n_iters) {
  if (!eg || !prm || !samples || !n_iters || n < 0 || (n > 0 && !matches)) return VO_ERR_INVALID;
  std::vector<Pt> raw, p;
  double Tp[9], Tc[9];
  *n_iters = 0;
  if (!prepare_points(matches, n, raw, p, Tp, Tc)) return VO_OK;
  const int iters = std::max(0, prm->ransac_iters);
  std::vector<int32_t> pool;
  for (int k = 0; k < iters; k++) draw_sample(eg->rng, n, pool, samples + (size_t)k * 8);
  *n_iters = iters;
  return VO_OK;
}

// The same for the n_lists consecutive frame pairs of a batch.  Only the generator itself is sequential (one add per value);
// whether a list reaches RANSAC and what its sample sets look like are per-list work, done by `threads` host threads:
//   1. per list, in parallel: does it reach RANSAC?  (then it consumes ransac_iters * 8 values, else none)
//   2. serially: that many raw values of the stream, in frame order
//   3. per list, in parallel: the sample sets from its slice of raw values
// samples + i * samples_stride * 8 receives drawn[i] sets.  Results equal n_lists calls of vo_egomotion_draw_samples.
vo_status vo_egomotion_draw_samples_batch(vo_egomotion* eg, int32_t n_lists, const vo_p_match* lists, int64_t stride,
                                          const int32_t* counts, const vo_mono_params* prm, int32_t* samples,
                                          int64_t samples_stride, int32_t* drawn, int32_t threads) {
  if (!eg || !lists || !counts || !prm || !samples || !drawn || n_lists < 0) return VO_ERR_INVALID;
  const int iters = std::max(0, prm->ransac_iters);
  if (samples_stride < iters) return VO_ERR_INVALID;
  auto parallel = [&](auto fn) {
    const int nt = std::max(1, std::min(threads, (int)n_lists));
    if (nt == 1) { for (int i = 0; i < n_lists; i++) fn(i); return; }
    std::atomic<int> next(0);
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++)
      th.emplace_back([&] { for (int i; (i = next.fetch_add(1)) < n_lists;) fn(i); });
    for (auto& t : th) t.join();
  };
  parallel([&](int i) {
    std::vector<Pt> raw, p;
    double Tp[9], Tc[9];
    drawn[i] = (counts[i] >= 0 && prepare_points(lists + (int64_t)i * stride, counts[i], raw, p, Tp, Tc)) ? iters : 0;
  });
  std::vector<size_t> first(n_lists + 1, 0);
  for (int i = 0; i < n_lists; i++) first[i + 1] = first[i] + (size_t)drawn[i] * 8;
  std::vector<uint32_t> raw_values(first[n_lists]);
  for (uint32_t& v : raw_values) v = eg->rng.next();
  parallel([&](int i) {
    const uint32_t* rv = raw_values.data() + first[i];
    int32_t* out = samples + (size_t)i * samples_stride * 8;
    const int32_t N = counts[i];
    for (int k = 0; k < drawn[i]; k++, rv += 8, out += 8) {
      int32_t taken[8];  // ascending; see draw_sample
      for (int j = 0; j < 8; j++) {
        int32_t v = (int32_t)(rv[j] % (uint32_t)(N - j));
        int pos = 0;
        while (pos < j && taken[pos] <= v) { v++; pos++; }
        for (int q = j; q > pos; q--) taken[q] = taken[q - 1];
        taken[pos] = v;
        out[j] = v;
      }
    }
  });
  return VO_OK;
}

vo_status vo_estimate_motion_mono_samples(const vo_p_match* matches, int32_t n, const vo_mono_params* prm,
                                          const int32_t* samples, int32_t n_iters, int32_t threads, double tr_delta[6],
                                          int32_t* ok, int32_t* inliers, int32_t inliers_cap, int32_t* n_inliers) {
  if (!prm || !tr_delta || !ok || n < 0 || (n > 0 && !matches) || n_iters < 0 || (n_iters > 0 && !samples)) return VO_ERR_INVALID;
  std::vector<int32_t> inl;
  *ok = estimate_motion(matches, n, *prm, samples, n_iters, std::max(1, threads), inl, tr_delta) ? 1 : 0;
  if (n_inliers) *n_inliers = (int32_t)inl.size();
  if (inliers) {
    if (inliers_cap < (int32_t)inl.size()) return VO_ERR_CAPACITY;
    std::memcpy(inliers, inl.data(), sizeof(int32_t) * inl.size());
  }
  return VO_OK;
}

// As vo_estimate_motion_mono_samples with the RANSAC winner already known (vo_ransac_gpu_score): best_k = index of the first
// hypothesis with the largest inlier count, -1 if none.
vo_status vo_estimate_motion_mono_best(const vo_p_match* matches, int32_t n, const vo_mono_params* prm, const int32_t* samples,
                                       int32_t n_iters, int32_t best_k, double tr_delta[6], int32_t* ok, int32_t* inliers,
                                       int32_t inliers_cap, int32_t* n_inliers) {
  if (!prm || !tr_delta || !ok || n < 0 || (n > 0 && !matches) || n_iters < 0 || (n_iters > 0 && !samples) || best_k < -1) return VO_ERR_INVALID;
  std::vector<int32_t> inl;
  *ok = estimate_motion(matches, n, *prm, samples, n_iters, 1, inl, tr_delta, best_k) ? 1 : 0;
  if (n_inliers) *n_inliers = (int32_t)inl.size();
  if (inliers) {
    if (inliers_cap < (int32_t)inl.size()) return VO_ERR_CAPACITY;
    std::memcpy(inliers, inl.data(), sizeof(int32_t) * inl.size());
  }
  return VO_OK;
}

// normalised match coordinates as estimateMotion uses them (src/viso_mono.cpp:187-240): out4n = n x {u1p, v1p, u1c, v1c};
// returns 0 if the list never reaches RANSAC.  For vo_ransac.cu.
int32_t vo_egomotion_prepare_points(const vo_p_match* matches, int32_t n, float* out4n) {
  std::vector<Pt> raw, p;
  double Tp[9], Tc[9];
  if (!matches || !out4n || !prepare_points(matches, n, raw, p, Tp, Tc)) return 0;
  std::memcpy(out4n, p.data(), sizeof(Pt) * (size_t)n);
  return 1;
}

vo_status vo_estimate_motion_mono(vo_egomotion* eg, const vo_p_match* matches, int32_t n, const vo_mono_params* prm,
                                  double tr_delta[6], int32_t* ok) {
  if (!eg || !prm || !tr_delta || !ok || n < 0 || (n > 0 && !matches)) return VO_ERR_INVALID;
  std::vector<int32_t> samples((size_t)std::max(0, prm->ransac_iters) * 8);
  int32_t iters = 0;
  vo_status st = vo_egomotion_draw_samples(eg, matches, n, prm, samples.data(), &iters);
  if (st != VO_OK) return st;
  *ok = estimate_motion(matches, n, *prm, samples.data(), iters, eg->threads, eg->inliers, tr_delta) ? 1 : 0;
  return VO_OK;
}

vo_status vo_egomotion_inliers(const vo_egomotion* eg, int32_t* out, int32_t cap, int32_t* n) {
  if (!eg || !n) return VO_ERR_INVALID;
  *n = (int32_t)eg->inliers.size();
  if (out) {
    if (cap < *n) return VO_ERR_CAPACITY;
    std::memcpy(out, eg->inliers.data(), sizeof(int32_t) * eg->inliers.size());
  }
  return VO_OK;
}

uint32_t vo_egomotion_rand(vo_egomotion* eg) { return eg ? eg->rng.next() : 0u; }

// VisualOdometry::transformationVectorToMatrix: src/viso.cpp:59-84 (row-major 4x4)
void vo_transformation_vector_to_matrix(const double tr[6], double Tr[16]) {
  const double sx = std::sin(tr[0]), cx = std::cos(tr[0]), sy = std::sin(tr[1]), cy = std::cos(tr[1]);
  const double sz = std::sin(tr[2]), cz = std::cos(tr[2]);
  const double M[16] = {+cy * cz,                -cy * sz,                +sy,      tr[3],
                        +sx * sy * cz + cx * sz, -sx * sy * sz + cx * cz, -sx * cy, tr[4],
                        -cx * sy * cz + sx * sz, +cx * sy * sz + sx * cz, +cx * cy, tr[5],
                        0, 0, 0, 1};
  std::memcpy(Tr, M, sizeof(M));
}

}  // extern "C"

// =================================================================================================
// Stereo egomotion: VisualOdometryStereo::estimateMotion, src/viso_stereo.cpp:54-157 -- 3-point RANSAC over a
// Gauss-Newton minimisation of the reprojection error of the previous frame's triangulated points into the
// current left/right images, then a refit on the inliers.  A host stage like the monocular one (sequential,
// 6x6 systems).  Operation order follows the reference statement by statement (sums, Gauss-Jordan pivoting), and
// the library is compiled without FMA contraction, so results agree with the compiled reference to rounding.
// =================================================================================================
namespace {

enum GnResult { GN_UPDATED, GN_FAILED, GN_CONVERGED };  // src/viso_stereo.h:69

struct StereoWork {
  const vo_p_match* m;
  int32_t N;
  vo_stereo_params prm;
  std::vector<double> X, Y, Z, J, predict, observe, residual;
};

// Matrix::solve (Gauss-Jordan, full pivoting, eps = 1e-20): src/matrix.cpp:417-503, for a 6x6 system with one RHS
bool solve6(double A[6][6], double B[6]) {
  int indxc[6], indxr[6], ipiv[6] = {0, 0, 0, 0, 0, 0};
  for (int i = 0; i < 6; i++) {
    double big = 0.0;
    int irow = 0, icol = 0;
    for (int j = 0; j < 6; j++)
      if (ipiv[j] != 1)
        for (int k = 0; k < 6; k++)
          if (ipiv[k] == 0 && std::fabs(A[j][k]) >= big) { big = std::fabs(A[j][k]); irow = j; icol = k; }
    ++ipiv[icol];
    if (irow != icol) {
      for (int l = 0; l < 6; l++) std::swap(A[irow][l], A[icol][l]);
      std::swap(B[irow], B[icol]);
    }
    indxr[i] = irow; indxc[i] = icol;
    if (std::fabs(A[icol][icol]) < 1e-20) return false;
    const double pivinv = 1.0 / A[icol][icol];
    A[icol][icol] = 1.0;
    for (int l = 0; l < 6; l++) A[icol][l] *= pivinv;
    B[icol] *= pivinv;
    for (int ll = 0; ll < 6; ll++)
      if (ll != icol) {
        const double dum = A[ll][icol];
        A[ll][icol] = 0.0;
        for (int l = 0; l < 6; l++) A[ll][l] -= A[icol][l] * dum;
        B[ll] -= B[icol] * dum;
      }
  }
  (void)indxr; (void)indxc;  // the column unscrambling only touches A, which is not read afterwards
  return true;
}

// computeObservations + computeResidualsAndJacobian: src/viso_stereo.cpp:222-327
void residuals_and_jacobian(StereoWork& w, const double tr[6], const int32_t* active, int32_t na) {
  for (int32_t i = 0; i < na; i++) {
    const vo_p_match& q = w.m[active[i]];
    w.observe[4 * i + 0] = q.u1c; w.observe[4 * i + 1] = q.v1c; w.observe[4 * i + 2] = q.u2c; w.observe[4 * i + 3] = q.v2c;
  }
  const double rx = tr[0], ry = tr[1], rz = tr[2], tx = tr[3], ty = tr[4], tz = tr[5];
  const double sx = std::sin(rx), cx = std::cos(rx), sy = std::sin(ry), cy = std::cos(ry), sz = std::sin(rz), cz = std::cos(rz);
  const double r00 = +cy * cz, r01 = -cy * sz, r02 = +sy;
  const double r10 = +sx * sy * cz + cx * sz, r11 = -sx * sy * sz + cx * cz, r12 = -sx * cy;
  const double r20 = -cx * sy * cz + sx * sz, r21 = +cx * sy * sz + sx * cz, r22 = +cx * cy;
  const double rdrx10 = +cx * sy * cz - sx * sz, rdrx11 = -cx * sy * sz - sx * cz, rdrx12 = -cx * cy;
  const double rdrx20 = +sx * sy * cz + cx * sz, rdrx21 = -sx * sy * sz + cx * cz, rdrx22 = -sx * cy;
  const double rdry00 = -sy * cz, rdry01 = +sy * sz, rdry02 = +cy;
  const double rdry10 = +sx * cy * cz, rdry11 = -sx * cy * sz, rdry12 = +sx * sy;
  const double rdry20 = -cx * cy * cz, rdry21 = +cx * cy * sz, rdry22 = -cx * sy;
  const double rdrz00 = -cy * sz, rdrz01 = -cy * cz;
  const double rdrz10 = -sx * sy * sz + cx * cz, rdrz11 = -sx * sy * cz - cx * sz;
  const double rdrz20 = +cx * sy * sz + sx * cz, rdrz21 = +cx * sy * cz - sx * sz;
  const double f = w.prm.f, cu = w.prm.cu, cv = w.prm.cv, base = w.prm.base;
  for (int32_t i = 0; i < na; i++) {
    const double X1p = w.X[active[i]], Y1p = w.Y[active[i]], Z1p = w.Z[active[i]];
    const double X1c = r00 * X1p + r01 * Y1p + r02 * Z1p + tx;
    const double Y1c = r10 * X1p + r11 * Y1p + r12 * Z1p + ty;
    const double Z1c = r20 * X1p + r21 * Y1p + r22 * Z1p + tz;
    double weight = 1.0;
    if (w.prm.reweighting) weight = 1.0 / (std::fabs(w.observe[4 * i + 0] - cu) / std::fabs(cu) + 0.05);
    const double X2c = X1c - base;
    for (int j = 0; j < 6; j++) {
      double X1cd = 0, Y1cd = 0, Z1cd = 0;
      switch (j) {
        case 0: X1cd = 0; Y1cd = rdrx10 * X1p + rdrx11 * Y1p + rdrx12 * Z1p; Z1cd = rdrx20 * X1p + rdrx21 * Y1p + rdrx22 * Z1p; break;
        case 1: X1cd = rdry00 * X1p + rdry01 * Y1p + rdry02 * Z1p; Y1cd = rdry10 * X1p + rdry11 * Y1p + rdry12 * Z1p;
                Z1cd = rdry20 * X1p + rdry21 * Y1p + rdry22 * Z1p; break;
        case 2: X1cd = rdrz00 * X1p + rdrz01 * Y1p; Y1cd = rdrz10 * X1p + rdrz11 * Y1p; Z1cd = rdrz20 * X1p + rdrz21 * Y1p; break;
        case 3: X1cd = 1; Y1cd = 0; Z1cd = 0; break;
        case 4: X1cd = 0; Y1cd = 1; Z1cd = 0; break;
        case 5: X1cd = 0; Y1cd = 0; Z1cd = 1; break;
      }
      w.J[(4 * i + 0) * 6 + j] = weight * f * (X1cd * Z1c - X1c * Z1cd) / (Z1c * Z1c);
      w.J[(4 * i + 1) * 6 + j] = weight * f * (Y1cd * Z1c - Y1c * Z1cd) / (Z1c * Z1c);
      w.J[(4 * i + 2) * 6 + j] = weight * f * (X1cd * Z1c - X2c * Z1cd) / (Z1c * Z1c);
      w.J[(4 * i + 3) * 6 + j] = weight * f * (Y1cd * Z1c - Y1c * Z1cd) / (Z1c * Z1c);
    }
    w.predict[4 * i + 0] = f * X1c / Z1c + cu;
    w.predict[4 * i + 1] = f * Y1c / Z1c + cv;
    w.predict[4 * i + 2] = f * X2c / Z1c + cu;
    w.predict[4 * i + 3] = f * Y1c / Z1c + cv;
    for (int k = 0; k < 4; k++) w.residual[4 * i + k] = weight * (w.observe[4 * i + k] - w.predict[4 * i + k]);
  }
}

// updateParameters: src/viso_stereo.cpp:178-220
GnResult update_parameters(StereoWork& w, const int32_t* active, int32_t na, double tr[6], double step_size, double eps) {
  if (na < 3) return GN_FAILED;
  residuals_and_jacobian(w, tr, active, na);
  double A[6][6], B[6];
  for (int m = 0; m < 6; m++) {
    for (int n = 0; n < 6; n++) {
      double a = 0;
      for (int32_t i = 0; i < 4 * na; i++) a += w.J[i * 6 + m] * w.J[i * 6 + n];
      A[m][n] = a;
    }
    double b = 0;
    for (int32_t i = 0; i < 4 * na; i++) b += w.J[i * 6 + m] * w.residual[i];
    B[m] = b;
  }
  if (!solve6(A, B)) return GN_FAILED;
  bool converged = true;
  for (int m = 0; m < 6; m++) {
    tr[m] += step_size * B[m];
    if (std::fabs(B[m]) > eps) converged = false;
  }
  return converged ? GN_CONVERGED : GN_UPDATED;
}

// getInlier: src/viso_stereo.cpp:159-176
void stereo_inliers(StereoWork& w, const double tr[6], const std::vector<int32_t>& all, std::vector<int32_t>& out) {
  residuals_and_jacobian(w, tr, all.data(), w.N);
  out.clear();
  const double thr2 = w.prm.inlier_threshold * w.prm.inlier_threshold;
  for (int32_t i = 0; i < w.N; i++) {
    const double d0 = w.observe[4 * i + 0] - w.predict[4 * i + 0], d1 = w.observe[4 * i + 1] - w.predict[4 * i + 1];
    const double d2 = w.observe[4 * i + 2] - w.predict[4 * i + 2], d3 = w.observe[4 * i + 3] - w.predict[4 * i + 3];
    if (d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3 < thr2) out.push_back(i);
  }
}

}  // namespace

extern "C" vo_status vo_estimate_motion_stereo(vo_egomotion* eg, const vo_p_match* matches, int32_t n,
                                               const vo_stereo_params* prm, double tr_delta[6], int32_t* ok) {
  if (!eg || !prm || !tr_delta || !ok || n < 0 || (n > 0 && !matches)) return VO_ERR_INVALID;
  *ok = 0;
  eg->inliers.clear();
  for (int k = 0; k < 6; k++) tr_delta[k] = 0.0;
  if (n < 6) return VO_OK;  // src/viso_stereo.cpp:69-70 (before any rand() is consumed)
  StereoWork w;
  w.m = matches; w.N = n; w.prm = *prm;
  w.X.resize(n); w.Y.resize(n); w.Z.resize(n);
  w.J.resize((size_t)4 * n * 6); w.predict.resize((size_t)4 * n); w.observe.resize((size_t)4 * n); w.residual.resize((size_t)4 * n);
  for (int32_t i = 0; i < n; i++) {  // project matches of the previous image into 3d: src/viso_stereo.cpp:82-87
    const double d = std::max(matches[i].u1p - matches[i].u2p, 0.0001f);
    w.X[i] = (matches[i].u1p - prm->cu) * prm->base / d;
    w.Y[i] = (matches[i].v1p - prm->cv) * prm->base / d;
    w.Z[i] = prm->f * prm->base / d;
  }
  std::vector<int32_t> all(n), pool, inl_curr, best_inl;
  for (int32_t i = 0; i < n; i++) all[i] = i;
  double best_tr[6] = {0, 0, 0, 0, 0, 0};
  bool have_tr = false;
  for (int32_t k = 0; k < prm->ransac_iters; k++) {
    int32_t active[3];  // getRandomSample(N, 3): src/viso.cpp:86-106
    pool = all;
    for (int i = 0; i < 3; i++) {
      const int32_t j = (int32_t)(eg->rng.next() % (uint32_t)pool.size());
      active[i] = pool[j];
      pool.erase(pool.begin() + j);
    }
    double tr[6] = {0, 0, 0, 0, 0, 0};
    GnResult result = GN_UPDATED;
    int iter = 0;
    while (result == GN_UPDATED) {
      result = update_parameters(w, active, 3, tr, 1, 1e-6);
      if (iter++ > 20 || result == GN_CONVERGED) break;
    }
    if (result != GN_FAILED) {
      stereo_inliers(w, tr, all, inl_curr);
      if (inl_curr.size() > best_inl.size()) {
        best_inl = inl_curr;
        for (int q = 0; q < 6; q++) best_tr[q] = tr[q];
        have_tr = true;
      }
    }
  }
  bool success = true;
  if (best_inl.size() >= 6 && have_tr) {  // final optimisation on the inliers: src/viso_stereo.cpp:126-139
    int iter = 0;
    GnResult result = GN_UPDATED;
    while (result == GN_UPDATED) {
      result = update_parameters(w, best_inl.data(), (int32_t)best_inl.size(), best_tr, 1, 1e-8);
      if (iter++ > 100 || result == GN_CONVERGED) break;
    }
    if (result != GN_CONVERGED) success = false;
  } else {
    success = false;
  }
  eg->inliers = best_inl;
  if (success) {
    for (int q = 0; q < 6; q++) tr_delta[q] = best_tr[q];
    *ok = 1;
  }
  return VO_OK;
}
