// Drop-in surface of VisualOdometryStereo (src/viso_stereo.h:27-96, src/viso.h:31-113):
//   process(I1, I2, dims, replace) = pushBack(I1,I2) -> [bootstrap while no valid motion: matchFeatures(2) ->
//   bucketFeatures -> updateMotion] -> matchFeatures(2) -> bucketFeatures -> getMatches -> updateMotion
// (src/viso_stereo.cpp:33-52), updateMotion = the host stage vo_estimate_motion_stereo (csrc/vo_egomotion.cpp,
// bit-identical to the reference's estimateMotion, src/viso_stereo.cpp:54-157).
//
// In the reference this class cannot work: its Matcher is flow-only, so the u2p/u2c fields estimateMotion reads are
// always -1.  Here the Matcher is constructed in stereo mode (EXTENSION, parity unpinned for the
// matching half): matchFeatures(2) is the quad matcher and the class does what its interface promises.
// Header-only on top of the C ABI; no algorithm lives here.
#ifndef VO_B200_HOST_VISO_STEREO_H
#define VO_B200_HOST_VISO_STEREO_H

#include "viso_mono.h"

class VisualOdometryStereo {
 public:
  typedef VisualOdometryMono::bucketing bucketing;
  typedef VisualOdometryMono::calibration calibration;
  struct parameters {  // src/viso.h:67-71 + src/viso_stereo.h:32-43
    Matcher::parameters match;
    bucketing bucket;
    calibration calib;
    double base;
    int32_t ransac_iters;
    double inlier_threshold;
    bool reweighting;
    parameters() : base(1.0), ransac_iters(200), inlier_threshold(2.0), reweighting(true) {}
  };

  explicit VisualOdometryStereo(parameters param, int device = 0)
      : param(param), matcher(new Matcher(param.match, device, true)), ego(nullptr), Tr_delta(Matrix::eye(4)), Tr_valid(false) {
    matcher->setIntrinsics(param.calib.f, param.calib.cu, param.calib.cv, param.base);  // src/viso_stereo.cpp:26-28
    vo_egomotion_create(0u, 0, &ego);                                                   // srand(0): src/viso.cpp:34
  }
  ~VisualOdometryStereo() { delete matcher; vo_egomotion_destroy(ego); }
  VisualOdometryStereo(const VisualOdometryStereo&) = delete;
  VisualOdometryStereo& operator=(const VisualOdometryStereo&) = delete;

  // src/viso_stereo.cpp:33-52
  bool process(uint8_t* I1, uint8_t* I2, int32_t* dims, bool replace = false) {
    matcher->pushBack(I1, I2, dims, replace);
    if (!Tr_valid) {  // bootstrap motion estimate if invalid
      matcher->matchFeatures(2);
      matcher->bucketFeatures(param.bucket.max_features, (float)param.bucket.bucket_width, (float)param.bucket.bucket_height);
      p_matched = matcher->getMatches();
      updateMotion();
    }
    matcher->matchFeatures(2, Tr_valid ? &Tr_delta : 0);
    matcher->bucketFeatures(param.bucket.max_features, (float)param.bucket.bucket_width, (float)param.bucket.bucket_height);
    p_matched = matcher->getMatches();
    return updateMotion();
  }
  bool process(std::vector<Matcher::p_match> p_matched_) {  // src/viso.h:75-78
    p_matched = p_matched_;
    return updateMotion();
  }

  Matrix getMotion() { return Tr_delta; }
  std::vector<Matcher::p_match> getMatches() { return matcher->getMatches(); }
  int32_t getNumberOfMatches() { return (int32_t)p_matched.size(); }
  int32_t getNumberOfInliers() { return (int32_t)inliers.size(); }
  std::vector<int32_t> getInlierIndices() { return inliers; }
  std::vector<double> getMotionVector() { return tr_delta; }
  Matcher* getMatcher() { return matcher; }

 private:
  bool updateMotion() {  // src/viso.cpp:42-57
    vo_stereo_params prm;
    prm.f = param.calib.f; prm.cu = param.calib.cu; prm.cv = param.calib.cv; prm.base = param.base;
    prm.ransac_iters = param.ransac_iters; prm.reweighting = param.reweighting ? 1 : 0;
    prm.inlier_threshold = param.inlier_threshold;
    double tr[6];
    int32_t ok = 0, n = 0;
    vo_estimate_motion_stereo(ego, reinterpret_cast<const vo_p_match*>(p_matched.data()), (int32_t)p_matched.size(), &prm, tr, &ok);
    tr_delta = ok ? std::vector<double>(tr, tr + 6) : std::vector<double>();
    vo_egomotion_inliers(ego, nullptr, 0, &n);
    inliers.resize(n);
    if (n) vo_egomotion_inliers(ego, inliers.data(), n, &n);
    if (tr_delta.size() != 6) return false;
    Tr_delta = VisualOdometryMono::transformationVectorToMatrix(tr_delta);
    Tr_valid = true;
    return true;
  }
  parameters param;
  Matcher* matcher;
  vo_egomotion* ego;
  std::vector<Matcher::p_match> p_matched;
  std::vector<int32_t> inliers;
  std::vector<double> tr_delta;
  Matrix Tr_delta;
  bool Tr_valid;
};

#endif
