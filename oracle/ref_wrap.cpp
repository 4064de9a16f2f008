// TEST INFRASTRUCTURE ONLY -- thin extern "C" wrapper around the UNMODIFIED reference sources.
//
// This translation unit is compiled together with the reference's own .cpp files, taken where
// they lie under /root/reference/src (see oracle/build_ref.sh); nothing of the reference is
// copied into this repository.  The resulting shared object lands in oracle/_ref/ (git-ignored)
// and is used (a) to pin oracle/vo_oracle.c, (b) to generate tests/golden/*, and (c) as the timed
// CPU baseline of bench.py (`--impl reference`, cpu_baseline.kind == "reference").
// Only tests/, __graft_entry__.smoke() and bench.py's cpu legs may load it.
//
// The wrapper adds no algorithmic code: every export forwards to a reference symbol
//   myComputeFeatures   /root/reference/src/myComputeFeature.cpp:352
//   myMatching          /root/reference/src/myMatch.cpp:122
//   removeOutliers      /root/reference/src/remove_outliers.cpp:4
//   Matcher::*          /root/reference/src/matcher.cpp:32-187
//   VisualOdometryMono::estimateMotion   /root/reference/src/viso_mono.cpp:41  (row F2, host egomotion)
//   VisualOdometryStereo::estimateMotion /root/reference/src/viso_stereo.cpp:54 (row F3, stereo egomotion)
// The reference keeps multi-megabyte arrays on the stack, so every call is executed on a
// helper thread with a 1 GiB stack; std::cout chatter of the reference is muted.

#define private public      // reach Matcher::p_matched_2 / max2c for the parity dumps, VisualOdometryMono::estimateMotion
#define protected public
#include "matcher.h"
#include "viso_mono.h"
#include "viso_stereo.h"
#undef private
#undef protected

#include <pthread.h>
#include <functional>
#include <sstream>
#include <cstring>

namespace {

struct Thunk { std::function<void()> fn; };

void* trampoline(void* p) {
  static_cast<Thunk*>(p)->fn();
  return nullptr;
}

void on_big_stack(std::function<void()> fn) {
  static std::ostringstream sink;
  static bool muted = false;
  if (!muted) { std::cout.rdbuf(sink.rdbuf()); muted = true; }
  sink.str("");
  Thunk t{fn};
  pthread_attr_t attr;
  pthread_attr_init(&attr);
  pthread_attr_setstacksize(&attr, size_t(1) << 30);
  pthread_t th;
  pthread_create(&th, &attr, trampoline, &t);
  pthread_join(th, nullptr);
  pthread_attr_destroy(&attr);
}

}  // namespace

extern "C" {

// compile-time geometry of this build: {IMG_W, IMG_H, BIN_W, BIN_H, U_BIN_NUM, V_BIN_NUM,
// MAX_FP_IN_BIN, COL_BIN_FEATURE_MAX, POINT_L, sizeof(p_match)}
void ref_info(int32_t out[10]) {
  out[0] = IMG_W; out[1] = IMG_H; out[2] = BIN_W; out[3] = BIN_H;
  out[4] = U_BIN_NUM; out[5] = V_BIN_NUM; out[6] = MAX_FP_IN_BIN;
  out[7] = COL_BIN_FEATURE_MAX; out[8] = POINT_L; out[9] = (int32_t)sizeof(Matcher::p_match);
}

void ref_sobel5x5(const uint8_t* in, uint8_t* du, uint8_t* dv) {
  on_big_stack([&] { mySobel5x5(in, du, dv, IMG_W, IMG_H); });
}
void ref_blob5x5(const uint8_t* in, int16_t* out) {
  on_big_stack([&] { myBlob5x5(in, out, IMG_W, IMG_H); });
}
void ref_checkerboard5x5(const uint8_t* in, int16_t* out) {
  on_big_stack([&] { myCheckerboard5x5(in, out, IMG_W, IMG_H); });
}

// max2: int32[BIN_NUM*MAX_FP_IN_BIN*12], max2_num: int32[BIN_NUM]
void ref_compute_features(uint8_t* I, int32_t* max2, int32_t* max2_num) {
  on_big_stack([&] { myComputeFeatures(I, max2, max2_num); });
}

void ref_matching(int32_t* m1p, int32_t* m1c, int32_t* n1p, int32_t* n1c,
                  void* p_matched, int32_t* p_matched_num) {
  on_big_stack([&] {
    int32_t n = 0;
    myMatching(m1p, m1c, n1p, n1c, (Matcher::p_match*)p_matched, n);
    *p_matched_num = n;
  });
}

// in/out list, at most POINT_L entries
void ref_remove_outliers(void* p_matched, int32_t* cnt) {
  on_big_stack([&] {
    int32_t n = *cnt;
    removeOutliers((Matcher::p_match*)p_matched, n);
    *cnt = n;
  });
}

uint32_t ref_rand_number(uint32_t x) {
  Matcher::parameters p;
  static Matcher* m = new Matcher(p);
  return m->rand_number(x);
}

// Matcher::bucketFeatures on an arbitrary list (<= POINT_L entries, <= 126 buckets of <= 256)
void ref_bucket_features(void* p_matched, int32_t* cnt, int32_t max_features, float bw, float bh) {
  on_big_stack([&] {
    Matcher::parameters p;
    Matcher* m = new Matcher(p);
    m->p_matched_2_cnt = *cnt;
    std::memcpy(m->p_matched_2, p_matched, sizeof(Matcher::p_match) * size_t(*cnt));
    m->bucketFeatures(max_features, bw, bh);
    *cnt = m->p_matched_2_cnt;
    std::memcpy(p_matched, m->p_matched_2, sizeof(Matcher::p_match) * size_t(*cnt));
    delete m;
  });
}

// ---- the Matcher facade itself (pushBack / matchFeatures / bucketFeatures / getMatches) ----
void* ref_matcher_new() {
  Matcher::parameters p;
  Matcher* m = new Matcher(p);
  // UB fence: the reference never initialises the per-bin counts, so
  // frame 0 would be matched against heap garbage.  Zero them, as src_hls/demo.cpp:127 and
  // hls_opencl/kernel/myMatch.cpp:221-224 do in spirit.
  std::memset(m->max2c_num, 0, sizeof(m->max2c_num));
  std::memset(m->max2p_num, 0, sizeof(m->max2p_num));
  std::memset(m->max2c, 0, 4 * size_t(MAX_FEATURE_ARRAY_SIZE));
  std::memset(m->max2p, 0, 4 * size_t(MAX_FEATURE_ARRAY_SIZE));
  m->p_matched_2_cnt = 0;
  return m;
}
void ref_matcher_delete(void* h) { delete (Matcher*)h; }
void ref_matcher_push_back(void* h, uint8_t* I, int32_t* dims, int replace) {
  on_big_stack([&] { ((Matcher*)h)->pushBack(I, dims, replace != 0); });
}
void ref_matcher_match_features(void* h, int method) {
  on_big_stack([&] { ((Matcher*)h)->matchFeatures(method); });
}
void ref_matcher_bucket_features(void* h, int32_t max_features, float bw, float bh) {
  on_big_stack([&] { ((Matcher*)h)->bucketFeatures(max_features, bw, bh); });
}
int32_t ref_matcher_num_matches(void* h) { return ((Matcher*)h)->p_matched_2_cnt; }
void ref_matcher_get_matches(void* h, void* out) {
  Matcher* m = (Matcher*)h;
  std::memcpy(out, m->p_matched_2, sizeof(Matcher::p_match) * size_t(m->p_matched_2_cnt));
}
// parity dumps of the ring buffer: which = 0 prev, 1 curr
void ref_matcher_dump_features(void* h, int which, int32_t* max2, int32_t* max2_num) {
  Matcher* m = (Matcher*)h;
  std::memcpy(max2, which ? m->max2c : m->max2p, 4 * size_t(MAX_FEATURE_ARRAY_SIZE));
  std::memcpy(max2_num, which ? m->max2c_num : m->max2p_num, 4 * size_t(BIN_NUM));
}

// ---- row F2: monocular egomotion on injected match lists -----------------------------------------
// One VisualOdometryMono object (its constructor calls srand(0), src/viso.cpp:34) processes n_lists match
// lists in order, so the glibc rand() stream carries over from list to list exactly as in a sequence.
// prm = {f, cu, cv, height, pitch, ransac_iters, inlier_threshold, motion_threshold};
// tr_out[6*i..] = estimateMotion's vector, ok[i] = 1 if it had 6 entries, n_inliers[i] = inliers.size().
void ref_mono_motion_seq(const void* lists, const int32_t* counts, int32_t n_lists, const double* prm,
                         double* tr_out, int32_t* ok, int32_t* n_inliers) {
  on_big_stack([&] {
    VisualOdometryMono::parameters p;
    p.calib.f = prm[0]; p.calib.cu = prm[1]; p.calib.cv = prm[2];
    p.height = prm[3]; p.pitch = prm[4];
    p.ransac_iters = (int32_t)prm[5]; p.inlier_threshold = prm[6]; p.motion_threshold = prm[7];
    VisualOdometryMono vo(p);
    const Matcher::p_match* cur = (const Matcher::p_match*)lists;
    for (int32_t i = 0; i < n_lists; i++) {
      std::vector<Matcher::p_match> v(cur, cur + counts[i]);
      cur += counts[i];
      std::vector<double> tr = vo.estimateMotion(v);
      ok[i] = tr.size() == 6 ? 1 : 0;
      for (int k = 0; k < 6; k++) tr_out[6 * i + k] = ok[i] ? tr[k] : 0.0;
      n_inliers[i] = (int32_t)vo.inliers.size();
    }
  });
}

// ---- stereo egomotion (the egomotion half IS in the reference, src/viso_stereo.cpp:54-157) ----
// One VisualOdometryStereo object (srand(0) in its base constructor) processes the lists in order.
// prm = {f, cu, cv, base, ransac_iters, inlier_threshold, reweighting}
void ref_stereo_motion_seq(const void* lists, const int32_t* counts, int32_t n_lists, const double* prm,
                           double* tr_out, int32_t* ok, int32_t* n_inliers) {
  on_big_stack([&] {
    VisualOdometryStereo::parameters p;
    p.calib.f = prm[0]; p.calib.cu = prm[1]; p.calib.cv = prm[2];
    p.base = prm[3]; p.ransac_iters = (int32_t)prm[4]; p.inlier_threshold = prm[5]; p.reweighting = prm[6] != 0.0;
    VisualOdometryStereo vo(p);
    const Matcher::p_match* cur = (const Matcher::p_match*)lists;
    for (int32_t i = 0; i < n_lists; i++) {
      std::vector<Matcher::p_match> v(cur, cur + counts[i]);
      cur += counts[i];
      std::vector<double> tr = vo.estimateMotion(v);
      ok[i] = tr.size() == 6 ? 1 : 0;
      for (int k = 0; k < 6; k++) tr_out[6 * i + k] = ok[i] ? tr[k] : 0.0;
      n_inliers[i] = (int32_t)vo.inliers.size();
    }
  });
}

// the C library generator the reference samples from (src/viso.cpp:98), for pinning our private copy
void ref_srand(uint32_t seed) { srand(seed); }
int32_t ref_rand() { return rand(); }

}  // extern "C"
