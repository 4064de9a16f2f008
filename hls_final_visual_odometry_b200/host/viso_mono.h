// Drop-in surface of VisualOdometryMono::process (src/viso_mono.h:27-84, src/viso_mono.cpp:33-39):
// pushBack -> matchFeatures -> bucketFeatures -> getMatches -> updateMotion.  The first four are the
// GPU hot path; egomotion (8-point RANSAC + SVD, src/viso_mono.cpp:41-160) is host math outside this
// round's scope and is supplied as a plug-in: without one, process()
// returns false exactly as the reference does when estimateMotion fails (src/viso.cpp:47-49).
#ifndef VO_B200_HOST_VISO_MONO_H
#define VO_B200_HOST_VISO_MONO_H

#include <functional>
#include <vector>

#include "matcher.h"

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
  // returns the 6-vector (rx,ry,rz,tx,ty,tz) or an empty vector on failure, like estimateMotion
  typedef std::function<std::vector<double>(const std::vector<Matcher::p_match>&)> MotionEstimator;

  explicit VisualOdometryMono(parameters param, int device = 0) : param(param), matcher(new Matcher(param.match, device)) {}
  ~VisualOdometryMono() { delete matcher; }

  void setMotionEstimator(MotionEstimator f) { estimator = f; }

  // src/viso_mono.cpp:33-39
  bool process(uint8_t* I, int32_t* dims, bool replace = false) {
    matcher->pushBack(I, dims, replace);
    matcher->matchFeatures(0);
    matcher->bucketFeatures(param.bucket.max_features, (float)param.bucket.bucket_width, (float)param.bucket.bucket_height);
    p_matched = matcher->getMatches();
    return updateMotion();
  }

  std::vector<Matcher::p_match> getMatches() { return p_matched; }            // src/viso.h:90
  int32_t getNumberOfMatches() { return (int32_t)p_matched.size(); }          // src/viso.h:93
  std::vector<double> getMotionVector() { return tr_delta; }
  Matcher* getMatcher() { return matcher; }

 private:
  bool updateMotion() {  // src/viso.cpp:42-57
    tr_delta = estimator ? estimator(p_matched) : std::vector<double>();
    return tr_delta.size() == 6;
  }
  parameters param;
  Matcher* matcher;
  MotionEstimator estimator;
  std::vector<Matcher::p_match> p_matched;
  std::vector<double> tr_delta;
};

#endif
