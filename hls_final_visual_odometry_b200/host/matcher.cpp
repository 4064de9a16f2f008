#include "matcher.h"

#include <iostream>

#include "vo_b200.h"

static_assert(sizeof(Matcher::p_match) == sizeof(vo_p_match), "p_match must stay the 48-byte wire format");

Matcher::Matcher(parameters param, int device, bool stereo)
    : param(param), device(device), stereo(stereo), eng(nullptr), last_status(0) {}

Matcher::~Matcher() {
  if (eng) vo_engine_destroy(eng);
}

const char* Matcher::lastError() const { return vo_last_error(); }

bool Matcher::ensureEngine(int32_t width, int32_t height) {
  if (eng) return true;
  vo_config cfg = {};
  cfg.width = width;
  cfg.height = height;
  // the reference's executed path hard-codes tau = 50 whatever parameters::nms_tau says (src/myComputeFeature.cpp:157);
  // a drop-in does the same unless the caller opts in to the (unpinned) live parameter
  cfg.nms_tau = live_nms_tau ? param.nms_tau : 0;
  cfg.bin_w = live_bin_w;
  cfg.match_radius_u = live_rad_u;
  cfg.match_radius_v = live_rad_v;
  cfg.bin_h = live_bin_h;
  cfg.stereo = stereo ? 1 : 0;
  cfg.match_disp_tolerance = param.match_disp_tolerance;
  last_status = vo_engine_create(&cfg, device, nullptr, &eng);
  if (last_status != VO_OK) {
    std::cerr << "ERROR: " << vo_last_error() << std::endl;
    eng = nullptr;
    return false;
  }
  return true;
}

void Matcher::pushBack(uint8_t* I1, uint8_t* I2, int32_t* dims, const bool replace) {
  const int32_t width = dims[0], height = dims[1], bpl = dims[2];
  if (width <= 0 || height <= 0 || bpl < width || I1 == 0) {  // src/matcher.cpp:59-62
    std::cerr << "ERROR: Image dimension mismatch!" << std::endl;
    return;
  }
  if (!ensureEngine(width, height)) return;
  last_status = vo_push_back(eng, I1, I2, dims, replace ? 1 : 0);
  if (last_status != VO_OK) std::cerr << "ERROR: " << vo_last_error() << std::endl;
}

void Matcher::matchFeatures(int32_t method, Matrix* /*Tr_delta*/) {
  if (!eng) return;
  last_status = vo_match_features(eng, method);
  if (last_status != VO_OK) {
    std::cerr << "ERROR: " << vo_last_error() << std::endl;
    return;
  }
  if (outlier_filter) {  // caller-supplied replacement of the host stage
    std::vector<p_match> m = getMatches();
    outlier_filter(m);
    last_status = vo_set_matches(eng, reinterpret_cast<const vo_p_match*>(m.data()), (int32_t)m.size());
  } else {
    last_status = vo_remove_outliers(eng);  // removeOutliers(p_matched_2, p_matched_2_cnt), src/matcher.cpp:108
  }
  if (last_status == VO_OK && param.refinement == 2) last_status = vo_refine_matches(eng);  // extension (matcher.h)
  if (last_status != VO_OK) std::cerr << "ERROR: " << vo_last_error() << std::endl;
}

void Matcher::bucketFeatures(int32_t max_features, float bucket_width, float bucket_height) {
  if (!eng) return;
  last_status = vo_bucket_features(eng, max_features, bucket_width, bucket_height);
  if (last_status != VO_OK) std::cerr << "ERROR: " << vo_last_error() << std::endl;
}

std::vector<Matcher::p_match> Matcher::getMatches() {
  std::vector<p_match> out;
  if (!eng) return out;
  int32_t n = 0;
  if ((last_status = vo_num_matches(eng, &n)) != VO_OK || n == 0) return out;
  out.resize(n);
  last_status = vo_get_matches(eng, reinterpret_cast<vo_p_match*>(out.data()), n, &n);
  if (last_status != VO_OK) out.clear();
  return out;
}
