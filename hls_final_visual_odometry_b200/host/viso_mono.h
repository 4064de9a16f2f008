// Drop-in surface of VisualOdometryMono (src/viso_mono.h:27-84, src/viso.h:31-113):
//   process(I, dims, replace) = pushBack -> matchFeatures -> bucketFeatures -> getMatches -> updateMotion
// (src/viso_mono.cpp:33-39).  The first four are the GPU hot path (+ the host Delaunay filter inside
// matchFeatures); updateMotion is the host egomotion stage vo_estimate_motion_mono (row F2,
// csrc/vo_egomotion.cpp) with the reference's sample stream (srand(0) in the constructor, src/viso.cpp:34).
// Header-only on top of the C ABI; no algorithm lives here.
#ifndef VO_B200_HOST_VISO_MONO_H
#define VO_B200_HOST_VISO_MONO_H

#include <cmath>
#include <cstdlib>
#include <functional>
#include <ostream>
#include <utility>
#include <vector>

#include "matcher.h"
#include "vo_b200.h"

// The slice of the reference's Matrix (src/matrix.h:37-133) that callers of getMotion() touch
// (src/demo.cpp:225-232: `pose = pose * Matrix::inv(viso.getMotion())`, `pose.val[i][j]`, operator<<).
class Matrix {
 public:
  int32_t m, n;
  double val[4][4];
  Matrix() : m(4), n(4) { for (auto& r : val) for (double& x : r) x = 0; }
  static Matrix eye(int32_t k = 4) { Matrix M; M.m = M.n = k; for (int i = 0; i < k; i++) M.val[i][i] = 1; return M; }
  Matrix operator*(const Matrix& B) const {
    Matrix C; C.m = m; C.n = B.n;
    for (int i = 0; i < m; i++)
      for (int j = 0; j < B.n; j++) { double s = 0; for (int k = 0; k < n; k++) s += val[i][k] * B.val[k][j]; C.val[i][j] = s; }
    return C;
  }
  Matrix operator~() const { Matrix T; T.m = n; T.n = m; for (int i = 0; i < m; i++) for (int j = 0; j < n; j++) T.val[j][i] = val[i][j]; return T; }
  // inverse by Gauss-Jordan elimination with partial pivoting (src/matrix.cpp:372-391 uses the same method)
  static Matrix inv(const Matrix& A) {
    const int k = A.m;
    Matrix L = A, R = eye(k);
    for (int c = 0; c < k; c++) {
      int piv = c;
      for (int r = c + 1; r < k; r++) if (std::fabs(L.val[r][c]) > std::fabs(L.val[piv][c])) piv = r;
      if (piv != c) for (int j = 0; j < k; j++) { std::swap(L.val[c][j], L.val[piv][j]); std::swap(R.val[c][j], R.val[piv][j]); }
      const double d = L.val[c][c];
      if (d == 0) return R;
      for (int j = 0; j < k; j++) { L.val[c][j] /= d; R.val[c][j] /= d; }
      for (int r = 0; r < k; r++) {
        if (r == c) continue;
        const double f = L.val[r][c];
        for (int j = 0; j < k; j++) { L.val[r][j] -= f * L.val[c][j]; R.val[r][j] -= f * R.val[c][j]; }
      }
    }
    return R;
  }
  friend std::ostream& operator<<(std::ostream& os, const Matrix& M) {
    for (int i = 0; i < M.m; i++) { for (int j = 0; j < M.n; j++) os << M.val[i][j] << (j + 1 < M.n ? " " : ""); if (i + 1 < M.m) os << "\n"; }
    return os;
  }
};

class VisualOdometryMono {
 public:
  struct bucketing {  // src/viso.h:43-53
    int32_t max_features;
    double bucket_width, bucket_height;
    bucketing() : max_features(2), bucket_width(50), bucket_height(50) {}
  };
  struct calibration {  // src/viso.h:56-64
    double f, cu, cv;
    calibration() : f(1), cu(0), cv(0) {}
  };
  struct parameters {  // src/viso.h:67-71 + src/viso_mono.h:32-46
    Matcher::parameters match;
    bucketing bucket;
    calibration calib;
    double height, pitch;
    int32_t ransac_iters;
    double inlier_threshold, motion_threshold;
    parameters() : height(1.0), pitch(0.0), ransac_iters(2000), inlier_threshold(0.00001), motion_threshold(100.0) {}
  };
  // optional replacement of estimateMotion: returns (rx,ry,rz,tx,ty,tz) or an empty vector on failure
  typedef std::function<std::vector<double>(const std::vector<Matcher::p_match>&)> MotionEstimator;

  explicit VisualOdometryMono(parameters param, int device = 0)
      : param(param), matcher(new Matcher(param.match, device)), ego(nullptr), Tr_delta(Matrix::eye(4)), Tr_valid(false) {
    vo_egomotion_create(0u, 0, &ego);  // srand(0): src/viso.cpp:34
  }
  ~VisualOdometryMono() { delete matcher; vo_egomotion_destroy(ego); }
  VisualOdometryMono(const VisualOdometryMono&) = delete;
  VisualOdometryMono& operator=(const VisualOdometryMono&) = delete;

  void setMotionEstimator(MotionEstimator f) { estimator = f; }

  // src/viso_mono.cpp:33-39
  bool process(uint8_t* I, int32_t* dims, bool replace = false) {
    matcher->pushBack(I, dims, replace);
    matcher->matchFeatures(0);
    matcher->bucketFeatures(param.bucket.max_features, (float)param.bucket.bucket_width, (float)param.bucket.bucket_height);
    p_matched = matcher->getMatches();
    return updateMotion();
  }
  // src/viso.h:75-78: egomotion from caller-supplied matches
  bool process(std::vector<Matcher::p_match> p_matched_) {
    p_matched = p_matched_;
    return updateMotion();
  }

  Matrix getMotion() { return Tr_delta; }                                       // src/viso.h:87
  std::vector<Matcher::p_match> getMatches() { return matcher->getMatches(); }  // src/viso.h:90
  int32_t getNumberOfMatches() { return (int32_t)p_matched.size(); }            // src/viso.h:93
  int32_t getNumberOfInliers() { return (int32_t)inliers.size(); }              // src/viso.h:96
  std::vector<int32_t> getInlierIndices() { return inliers; }                   // src/viso.h:99
  std::vector<double> getMotionVector() { return tr_delta; }
  Matcher* getMatcher() { return matcher; }

  // src/viso.cpp:59-84
  static Matrix transformationVectorToMatrix(const std::vector<double>& tr) {
    double T[16];
    vo_transformation_vector_to_matrix(tr.data(), T);
    Matrix M;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) M.val[i][j] = T[i * 4 + j];
    return M;
  }

 private:
  // src/viso.cpp:42-57 + estimateMotion (src/viso_mono.cpp:41-160) through the C ABI
  bool updateMotion() {
    if (estimator) {
      tr_delta = estimator(p_matched);
      inliers.clear();
    } else {
      vo_mono_params prm;
      prm.f = param.calib.f; prm.cu = param.calib.cu; prm.cv = param.calib.cv;
      prm.height = param.height; prm.pitch = param.pitch; prm.ransac_iters = param.ransac_iters; prm.reserved = 0;
      prm.inlier_threshold = param.inlier_threshold; prm.motion_threshold = param.motion_threshold;
      double tr[6];
      int32_t ok = 0, n = 0;
      static_assert(sizeof(Matcher::p_match) == sizeof(vo_p_match), "p_match layout");
      vo_estimate_motion_mono(ego, reinterpret_cast<const vo_p_match*>(p_matched.data()), (int32_t)p_matched.size(), &prm, tr, &ok);
      tr_delta = ok ? std::vector<double>(tr, tr + 6) : std::vector<double>();
      vo_egomotion_inliers(ego, nullptr, 0, &n);
      inliers.resize(n);
      if (n) vo_egomotion_inliers(ego, inliers.data(), n, &n);
    }
    if (tr_delta.size() != 6) return false;
    Tr_delta = transformationVectorToMatrix(tr_delta);
    Tr_valid = true;
    return true;
  }
  parameters param;
  Matcher* matcher;
  vo_egomotion* ego;
  MotionEstimator estimator;
  std::vector<Matcher::p_match> p_matched;
  std::vector<int32_t> inliers;
  std::vector<double> tr_delta;
  Matrix Tr_delta;
  bool Tr_valid;
};

#endif
