// Drop-in C++ host shim: the reference's Matcher class surface (src/matcher.h:40-143) on top of the
// C ABI of libvo_b200.so.  Same class name, nested types, method names, argument meaning and error
// behaviour, so a caller of the reference (src/viso_mono.cpp:33-39, src/demo.cpp:117) recompiles
// against this header unchanged.  No algorithm lives here: every method is one or two C-ABI calls.
//
// Differences that are visible on purpose:
//  * the image size is taken from `dims` at the first pushBack (the reference hard-wires
//    1024x284 in src/size.h and ignores dims);
//  * matchFeatures() runs the GPU circular matcher and then, as the reference does inside
//    matchFeatures (src/matcher.cpp:108), the host-side Delaunay support filter
//    (vo_remove_outliers, csrc/vo_outliers.cpp); setOutlierFilter() can replace that stage;
//  * frame 0 yields zero matches instead of matching against uninitialised counts;
//  * EXTENSION (absent from the reference): a Matcher constructed with stereo = true keeps the
//    right images pushBack receives and runs the quad matcher for matchFeatures(2); parameters::refinement == 2
//    (dead in the reference, default 0) adds the sub-pixel stage after removeOutliers.  With the defaults the
//    behaviour is the reference's: I2, `method` and `refinement` are ignored.
#ifndef VO_B200_HOST_MATCHER_H
#define VO_B200_HOST_MATCHER_H

#include <stdint.h>

#include <functional>
#include <vector>

struct vo_engine;
class Matrix;  // only ever passed through as an ignored pointer, as in the reference

class Matcher {
 public:
  // parameter settings (src/matcher.h:45-72) -- dead on the reference's executed path, kept for API parity
  struct parameters {
    int32_t nms_n;
    int32_t nms_tau;
    int32_t match_binsize;
    int32_t match_radius;
    int32_t match_disp_tolerance;
    int32_t outlier_disp_tolerance;
    int32_t outlier_flow_tolerance;
    int32_t multi_stage;
    int32_t half_resolution;
    int32_t refinement;
    double f, cu, cv, base;
    parameters() {
      nms_n = 2; nms_tau = 50; match_binsize = 50; match_radius = 200; match_disp_tolerance = 2;
      outlier_disp_tolerance = 5; outlier_flow_tolerance = 5; multi_stage = 0; half_resolution = 0;
      refinement = 0; f = cu = cv = base = 0;
    }
  };

  // structure for storing matches (src/matcher.h:89-104): 48 bytes, identical to vo_p_match
  class p_match {
   public:
    float u1p, v1p; int32_t i1p;
    float u2p, v2p; int32_t i2p;
    float u1c, v1c; int32_t i1c;
    float u2c, v2c; int32_t i2c;
    p_match() {}
    p_match(float u1p, float v1p, int32_t i1p, float u2p, float v2p, int32_t i2p, float u1c, float v1c,
            int32_t i1c, float u2c, float v2c, int32_t i2c)
        : u1p(u1p), v1p(v1p), i1p(i1p), u2p(u2p), v2p(v2p), i2p(i2p), u1c(u1c), v1c(v1c), i1c(i1c),
          u2c(u2c), v2c(v2c), i2c(i2c) {}
  };

  explicit Matcher(parameters param, int device = 0, bool stereo = false);
  ~Matcher();
  Matcher(const Matcher&) = delete;
  Matcher& operator=(const Matcher&) = delete;

  void setIntrinsics(double f, double cu, double cv, double base) {
    param.f = f; param.cu = cu; param.cv = cv; param.base = base;
  }

  // src/matcher.cpp:51-91.  dims = {width, height, bytes per line}; I2 is ignored as in the reference (unless stereo).
  void pushBack(uint8_t* I1, uint8_t* I2, int32_t* dims, const bool replace);
  void pushBack(uint8_t* I1, int32_t* dims, const bool replace) { pushBack(I1, 0, dims, replace); }

  // src/matcher.cpp:93-111.  method and Tr_delta are accepted and ignored, as in the reference (unless stereo: method 2).
  void matchFeatures(int32_t method, Matrix* Tr_delta = 0);

  // src/matcher.cpp:140-187
  void bucketFeatures(int32_t max_features, float bucket_width, float bucket_height);

  // src/matcher.h:138-143: returns a copy
  std::vector<Matcher::p_match> getMatches();

  // optional replacement of the host stage between matching and bucketing (default: the
  // reference's removeOutliers, src/remove_outliers.cpp:4); in/out list
  void setOutlierFilter(std::function<void(std::vector<Matcher::p_match>&)> f) { outlier_filter = f; }

  // EXTENSION (unpinned): parameters::nms_tau is dead on the reference's executed path, which always
  // thresholds at 50 (src/myComputeFeature.cpp:157), and this class ignores it the same way.  Call before the first
  // pushBack to make it live instead.  nms_n / match_binsize / match_radius stay ignored like in the reference.
  void useLiveNmsTau(bool on) { live_nms_tau = on; }

  // EXTENSION: the feature-list bin size, BIN_W x BIN_H of src/size.h:10-11 (60 x 45), as a run-time choice.
  // Supported: 60x45, 45x45, 60x30, 90x30 (bit-exact against the reference compiled with those numbers,
  // tests/golden/golden.json "*_b<BW>x<BH>"); anything else makes the first pushBack fail with VO_ERR_UNSUPPORTED.  The
  // search window stays 3 bins either way (SEARCH_RAD_U/V, src/size.h:14-15).  Call before the first pushBack.
  void useBinSize(int32_t bin_w, int32_t bin_h) { live_bin_w = bin_w; live_bin_h = bin_h; }

  // Opt-in: a live search radius in pixels (`match_radius`, src/matcher.h:50: dead on the reference's executed path, where the
  // |du| / |dv| window of find_match is SEARCH_RAD_U x SEARCH_RAD_V = 3 bins, src/size.h:14-15).  Bit-exact against the
  // reference compiled with those numbers (tests/golden/golden.json "*_r<RU>x<RV>"); the bin range of the search stays
  // 3 bins (src/myMatch.cpp:90-99).  0 = the reference's value.  Call before the first pushBack.
  void useMatchRadius(int32_t rad_u, int32_t rad_v) { live_rad_u = rad_u; live_rad_v = rad_v; }

  // last C-ABI status (0 = ok) and message, for callers that want more than the reference's cerr line
  int lastStatus() const { return last_status; }
  const char* lastError() const;
  vo_engine* engine() { return eng; }

 private:
  bool ensureEngine(int32_t width, int32_t height);
  parameters param;
  int device;
  bool stereo;
  vo_engine* eng;
  int last_status;
  bool live_nms_tau = false;
  int32_t live_bin_w = 0, live_bin_h = 0;  // 0 = the reference's 60 x 45
  int32_t live_rad_u = 0, live_rad_v = 0;  // 0 = the reference's three bins
  std::function<void(std::vector<Matcher::p_match>&)> outlier_filter;
};

#endif
