/* vo_b200.h -- C ABI of the B200-native visual-odometry front end (libvo_b200.so).
 *
 * Drop-in boundary for ONE path of Chang-Tun-Yu/HLS-final-Visual-Odometry: feature extraction ->
 * circular (prev<->curr) descriptor matching -> match bucketing.  Plain pointers and sizes only, no
 * C++/torch types; every entry point returns a vo_status and never throws.  All device work is
 * hand-written CUDA for sm_100a; there is NO CPU fallback: if no CUDA device is usable,
 * vo_engine_create fails with VO_ERR_CUDA and nothing else can be called.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference
 * checkout).  Three levels, lowest first:
 *
 *  1. operator level  == the reference's three `extern "C"` HLS kernels
 *       vo_compute_features   <- myComputeFeatures   hls_opencl/kernel/myComputeFeature.hpp:19-21
 *                                                    src/myComputeFeature.cpp:352
 *       vo_matching           <- myMatching          hls_opencl/kernel/myMatch.hpp:51-57
 *                                                    src/myMatch.cpp:122
 *       vo_bucket_matches     <- bucketing half of removeOutliers
 *                                                    hls_opencl/kernel/remove_outliers.cpp:104-154
 *                                                    src/matcher.cpp:140-187
 *  2. Matcher level   == the stateful ring buffer behind Matcher::pushBack / matchFeatures /
 *       bucketFeatures / getMatches (src/matcher.h:116-143, src/matcher.cpp:51-187)
 *  3. batch level     == many frames / frame pairs per launch (what the reference's FPGA host
 *       pipelines by hand, hls_opencl/host/main2.cpp:395-519); this is the throughput path.
 *  4. host stage F2   == VisualOdometryMono::estimateMotion (src/viso_mono.cpp:41-160), pure host math.
 *
 * Threading: an engine is not re-entrant; different engines are independent (one CUDA stream
 * each, or a caller-supplied stream).
 */
#ifndef VO_B200_H
#define VO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum vo_status {
  VO_OK = 0,
  VO_ERR_INVALID = 1,      /* bad argument (the reference prints "Image dimension mismatch", src/matcher.cpp:59-62) */
  VO_ERR_UNSUPPORTED = 2,  /* geometry/parameter outside what the reference path defines */
  VO_ERR_CUDA = 3,         /* CUDA runtime/driver failure; vo_last_error() has the text */
  VO_ERR_CAPACITY = 4,     /* an output did not fit the caller's buffer or an engine capacity */
  VO_ERR_STATE = 5         /* call sequence error (e.g. matching with no previous frame) */
} vo_status;

/* == Matcher::p_match, src/matcher.h:89-104 (48 bytes, no padding).  The reference's flow matcher
 * fills u1p,v1p,u1c,v1c with integer-valued floats and every other field with -1
 * (src/myMatch.cpp:189). */
typedef struct vo_p_match {
  float u1p, v1p; int32_t i1p;
  float u2p, v2p; int32_t i2p;
  float u1c, v1c; int32_t i1c;
  float u2c, v2c; int32_t i2c;
} vo_p_match;

/* Geometry.  The reference hard-wires 1024x284 and these constants in src/size.h:4-22 and
 * src/myComputeFeature.cpp:157-158; here they are runtime values validated at create time.
 * Fields left 0 take the reference value. */
typedef struct vo_config {
  int32_t width, height;   /* image size in pixels; >= 64 x 64 */
  int32_t bin_w, bin_h;    /* 0 = 60 x 45 (BIN_W, BIN_H of src/size.h:10-11).  Also built: 45x45, 60x30, 90x30 (live bin size,
                              SURVEY.md row F4: bit-exact against the reference compiled with those numbers); anything else
                              -> VO_ERR_UNSUPPORTED */
  int32_t nms_tau;         /* 50 */
  int32_t max_frames;      /* frame slots held on the device (>= 2; default 2) */
  int32_t max_pairs;       /* frame pairs matched per batch call (>= 1; default 1) */
  int32_t max_matches;     /* capacity of one pair's raw match list (default: generous bound) */
  int32_t match_rows_per_cta; /* tuning: origin v-bin rows per matching CTA (0 = auto: whole column) */
  int32_t match_exhaustive;   /* 0: exact lower-bound pruning in the SAD search (default); 1: evaluate every SAD.
                                 Results are bit-identical either way. */
  int32_t match_threads;      /* tuning: threads per matching CTA (256, 384 or 512; 0 = library default) */
  int32_t stereo;             /* EXTENSION (absent from the reference): 1 = keep a right-image ring
                                 (vo_push_back extracts I2), vo_match_features(2) and vo_batch_match_quad run the
                                 quad matcher.  0 (default) = the reference's behaviour: I2 and `method` are ignored. */
  int32_t match_disp_tolerance; /* |dv| window of left<->right edges, src/matcher.h:51; 0 = the default 2 */
  int32_t match_seed_good;    /* tuning: an origin whose starting bound (SAD of its seed or hint) is above this is matched by
                                 evaluating every candidate exactly instead of filtering (0 = library default 96).
                                 Results do not depend on it. */
  /* live search radius in pixels (Matcher::parameters::match_radius, src/matcher.h:50): SEARCH_RAD_U /
   * SEARCH_RAD_V of src/size.h:14-15, the |du| / |dv| window of find_match (src/myMatch.cpp:84-87,101).  0 = the
   * reference's 3*bin_w / 3*bin_h; else 1..4095.  The bin range of the search stays 3 bins (src/myMatch.cpp:90-99), as in
   * the reference compiled with other numbers, against which the results are bit-exact (tests/golden/golden.json
   * "*_r<RU>x<RV>"). */
  int32_t match_radius_u;
  int32_t match_radius_v;
} vo_config;

/* Derived geometry (what src/size.h spells as macros). */
typedef struct vo_geometry {
  int32_t width, height, pitch;     /* pitch = device row pitch of the frame slots (multiple of 16) */
  int32_t bin_w, bin_h;
  int32_t u_bins, v_bins;           /* U_BIN_NUM, V_BIN_NUM */
  int32_t bin_cap;                  /* MAX_FP_IN_BIN: records per (v_bin,u_bin,class) list = (bin_w/3)*(bin_h/3) */
  int32_t col_cap;                  /* COL_BIN_FEATURE_MAX = bin_cap * v_bins */
  int32_t bin_num;                  /* BIN_NUM = u_bins*v_bins*4 */
  int32_t max_frames, max_pairs, max_matches;
  int32_t reserved[3];
} vo_geometry;

typedef struct vo_engine vo_engine;

/* ---------------------------------------------------------------------------------------- */
/* lifetime                                                                                  */
/* ---------------------------------------------------------------------------------------- */

/* `stream` is a cudaStream_t passed as void* (NULL: the engine creates its own stream).
 * Replaces Matcher::Matcher / ~Matcher (src/matcher.cpp:32-49): the engine owns all feature buffers. */
vo_status vo_engine_create(const vo_config* cfg, int32_t device, void* stream, vo_engine** out);
void      vo_engine_destroy(vo_engine* e);
vo_status vo_engine_geometry(const vo_engine* e, vo_geometry* out);
vo_status vo_engine_device(const vo_engine* e, int32_t* device);   /* the CUDA device ordinal the engine was created on */
vo_status vo_engine_synchronize(vo_engine* e);
/* Fences for the batch-level software pipeline (the double-buffered event graph of the reference's
 * FPGA host, hls_opencl/host/main2.cpp:449-518).  vo_fence_record(idx) marks "everything enqueued
 * so far, downloads included"; vo_fence_wait(idx) blocks the host until that point.  idx in [0,8). */
vo_status vo_fence_record(vo_engine* e, int32_t idx);
vo_status vo_fence_wait(vo_engine* e, int32_t idx);
const char* vo_last_error(void);            /* thread-local text of the last failure */
const char* vo_version(void);
/* Diagnostic, host only: the 8-byte lower-bound signature K1 stores beside every 32-byte descriptor
 * (src/myComputeFeature.cpp:11-55 defines the bytes) and K2 prunes with: sig[0] + sig[1] compared byte-wise (sum of the
 * eight absolute differences) never exceeds the SAD of the two descriptors -- tests/test_host_logic.py checks exactly that. */
void      vo_descriptor_signature(const uint8_t desc[32], uint32_t sig[2]);

/* ---------------------------------------------------------------------------------------- */
/* 1. operator level (stateless from the caller's point of view; host buffers in, host out)   */
/* ---------------------------------------------------------------------------------------- */

/* myComputeFeatures(I, max2, max2_num): src/myComputeFeature.cpp:352, layout src/size.h:17-20.
 *   image     : u8, `pitch_bytes` per row (>= width)
 *   max2      : int32[bin_num * cap_stride * 12], record {u,v,0,class,8 descriptor words}
 *   max2_num  : int32[bin_num], bin index (v_bin*u_bins + u_bin)*4 + class
 *   cap_stride: records per list in the caller's array (MAX_FP_IN_BIN; 160 in the reference).
 * Returns VO_ERR_CAPACITY (after filling what fits) if a list is longer than cap_stride -- the
 * reference silently overflows. */
vo_status vo_compute_features(vo_engine* e, const uint8_t* image, int32_t pitch_bytes,
                              int32_t* max2, int32_t* max2_num, int32_t cap_stride);

/* myMatching(m1p, m1c, n1p, n1c, p_matched, p_matched_num): src/myMatch.cpp:122.
 * Feature arrays in the layout above (same cap_stride for both). */
vo_status vo_matching(vo_engine* e, const int32_t* m1p, const int32_t* n1p, const int32_t* m1c,
                      const int32_t* n1c, int32_t cap_stride, vo_p_match* p_matched,
                      int32_t p_matched_cap, int32_t* p_matched_num);

/* removeOutliers(p_matched, p_matched_cnt): src/remove_outliers.cpp:4-94 (Delaunay triangulation of
 * the current-frame coordinates, flow-consistency votes per triangle, keep support >= 4).  A HOST
 * stage by design (sequential sweep-hull); single precision and visiting order
 * follow src/delaunator.cpp so the kept set is the reference's.  In place; no engine needed. */
vo_status vo_remove_outliers_host(vo_p_match* matches, int32_t* n);

/* Matcher::bucketFeatures on an explicit list (src/matcher.cpp:140-187; FPGA: second half of
 * removeOutliers, hls_opencl/kernel/remove_outliers.cpp:104-154).  in and out may alias. */
vo_status vo_bucket_matches(vo_engine* e, const vo_p_match* in, int32_t n_in, int32_t max_features,
                            float bucket_w, float bucket_h, vo_p_match* out, int32_t out_cap,
                            int32_t* n_out);

/* NO COUNTERPART IN THE REFERENCE.  vo_bucket_matches for callers that may not
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the same bucketing kernel,
 * with its scratch in a static device array of the library (624 KB, held by every GPU that loads the library until exit)
 * instead of engine-owned allocations.  The call allocates nothing, copies nothing and does not synchronise: it enqueues two
 * kernels on the engine's stream; read the results after vo_engine_synchronize (or your own stream ordering).
 *   d_in      DEVICE pointer, n_in records of 48 bytes, 16-byte aligned; LIMIT: n_in <= 32768, more -> VO_ERR_CAPACITY
 *   d_out     DEVICE pointer, room for out_cap records, 16-byte aligned, must not overlap d_in
 *   d_result  DEVICE pointer to two int32: {number of bucketed matches, status}; status 0 = ok, 2 = more than 4096 buckets,
 *             3 = more than min(out_cap, 16384) matches kept (then the count is 0)
 * Same results as vo_bucket_matches on the same list.  Every call on a device shares the one array: calls that OVERLAP on the
 * GPU (two engines / streams on one device) race -- serialise them.  Nothing has to be cleared between calls. */
vo_status vo_bucket_matches_noalloc(vo_engine* e, const vo_p_match* d_in, int32_t n_in, int32_t max_features,
                                    float bucket_w, float bucket_h, vo_p_match* d_out, int32_t out_cap,
                                    int32_t* d_result);

/* The reference's FPGA operator signatures, 1:1, for a host that drives the three kernels directly
 * (hls_opencl/host/main2.cpp:429-444,477-504):
 *   myMatching(m1p, m1c, n1p, n1c, p_matched, p_matched_num, iter)     hls_opencl/kernel/myMatch.hpp:51-57
 *     iter == 0 (first frame, no previous features) -> *p_matched_num = 0, nothing else touched (.cpp:221-224)
 *   removeOutliers(p_matched, p_matched_cnt, p_matched_res, p_matched_res_cnt)   hls_opencl/kernel/remove_outliers.hpp:9-11
 *     = Delaunay support filter, THEN bucketing with max_features = 2, 50 x 50 px (.cpp:104-154); with
 *     *in_cnt <= 3 it returns with out / *out_cnt untouched, exactly as the reference (.cpp:20-21). */
vo_status vo_matching_iter(vo_engine* e, const int32_t* m1p, const int32_t* n1p, const int32_t* m1c,
                           const int32_t* n1c, int32_t cap_stride, vo_p_match* p_matched,
                           int32_t p_matched_cap, int32_t* p_matched_num, int32_t iter);
vo_status vo_remove_outliers_bucket(vo_engine* e, const vo_p_match* in, const int32_t* in_cnt,
                                    vo_p_match* out, int32_t out_cap, int32_t* out_cnt);

/* ---------------------------------------------------------------------------------------- */
/* 2. Matcher level: a prev/curr ring on frame slots 0/1 of the engine                         */
/* ---------------------------------------------------------------------------------------- */

/* Matcher::pushBack(I1, I2, dims, replace): src/matcher.cpp:51-91.  dims = {width,height,bytes per
 * line}; I2 is accepted and ignored exactly as in the reference.  width/height must equal the
 * engine geometry (the reference ignores dims and uses its compile-time size). */
vo_status vo_push_back(vo_engine* e, const uint8_t* I1, const uint8_t* I2, const int32_t dims[3],
                       int32_t replace);
/* Matcher::matchFeatures(method): src/matcher.cpp:93-111, WITHOUT the host-side Delaunay filter,
 * which is a separate call (vo_set_matches lets a host removeOutliers stage sit in between).
 * `method` is accepted and ignored as in the reference.  With fewer than two frames pushed the
 * result is zero matches (frame-0 contract of hls_opencl/kernel/myMatch.cpp:221-224). */
vo_status vo_match_features(vo_engine* e, int32_t method);
/* the removeOutliers call inside Matcher::matchFeatures (src/matcher.cpp:108) on the engine's current
 * raw list: download, vo_remove_outliers_host, upload.  vo_match_features + vo_remove_outliers ==
 * the reference's matchFeatures. */
vo_status vo_remove_outliers(vo_engine* e);
/* Matcher::bucketFeatures(max_features, bucket_width, bucket_height): src/matcher.cpp:140-187 */
vo_status vo_bucket_features(vo_engine* e, int32_t max_features, float bucket_w, float bucket_h);
/* Matcher::getMatches(): src/matcher.h:138-143 */
vo_status vo_get_matches(vo_engine* e, vo_p_match* out, int32_t cap, int32_t* n);
vo_status vo_num_matches(vo_engine* e, int32_t* n);
/* replace the current match list (e.g. after a host-side removeOutliers, src/remove_outliers.cpp:4) */
vo_status vo_set_matches(vo_engine* e, const vo_p_match* in, int32_t n);
/* parity dump of the ring buffer in the reference layout; which: 0 = previous, 1 = current */
vo_status vo_dump_features(vo_engine* e, int32_t which, int32_t* max2, int32_t* max2_num,
                           int32_t cap_stride);

/* ---------------------------------------------------------------------------------------- */
/* 3. batch level                                                                            */
/* ---------------------------------------------------------------------------------------- */

/* Extract features of `n` frames into frame slots first_slot .. first_slot+n-1 (one launch).
 * Host variant: frames at images + i*frame_stride_bytes, rows pitch_bytes apart; the H2D copy is
 * enqueued on the engine's upload stream (it overlaps the kernels of the previous batch; the host
 * buffer must stay valid until a later fence / synchronize).  Device variant: same addressing in device memory. */
vo_status vo_batch_extract_host(vo_engine* e, int32_t first_slot, int32_t n, const uint8_t* images,
                                int64_t frame_stride_bytes, int32_t pitch_bytes);
vo_status vo_batch_extract_device(vo_engine* e, int32_t first_slot, int32_t n,
                                  const uint8_t* d_images, int64_t frame_stride_bytes,
                                  int32_t pitch_bytes);
/* Match n_pairs pairs (prev_slot[i] -> curr_slot[i]) and, if max_features > 0, bucket each raw
 * list (bucketing directly follows matching, as in the FPGA operator chain without Delaunay). */
vo_status vo_batch_match(vo_engine* e, int32_t n_pairs, const int32_t* prev_slot,
                         const int32_t* curr_slot, int32_t max_features, float bucket_w,
                         float bucket_h);
/* ---- EXTENSION, parity unpinned: named by the reference's API (Matcher::pushBack(I1,I2,..), matchFeatures(2),
 * parameters::match_disp_tolerance / refinement, the u2p..i2c fields of p_match: src/matcher.h:45-56,89-104;
 * VisualOdometryStereo::process, src/viso_stereo.cpp:33-52) but not implemented by it -- its matcher is flow-only.
 * Defined in DESIGN.md section 9; oracle: oracle/vo_oracle.c (EXTENSION section). --------------------------------- */
/* Quad matching of n_items stereo frame pairs: the cycle 1c -> 1p -> 2p -> 2c -> 1c over four searches of the
 * reference's find_match (flow window on prev<->curr edges, |dv| <= disp_tolerance on left<->right edges), both
 * disparities >= 0, then the reference's emission rules.  Matches carry all eight coordinates.  refinement: 0, or 2 =
 * sub-pixel (vo_batch_refine).  Needs an engine with vo_config.stereo = 1. */
vo_status vo_batch_match_quad(vo_engine* e, int32_t n_items, const int32_t* slot_1p, const int32_t* slot_2p,
                              const int32_t* slot_1c, const int32_t* slot_2c, int32_t disp_tolerance, int32_t refinement,
                              int32_t max_features, float bucket_w, float bucket_h);
/* Sub-pixel refinement (parameters::refinement == 2, src/matcher.h:56) of the raw lists of the last vo_batch_match /
 * vo_batch_match_quad, in place: libviso2's parabolic fit on the small descriptor (computeSmallDescriptor,
 * src/matcher.cpp:516-543) with the current left feature as anchor; rejected matches are dropped.  The frames must
 * still be resident in their slots (VO_ERR_STATE after an in-place vo_batch_extract_device). */
vo_status vo_batch_refine(vo_engine* e, int32_t n_items);
/* the same on the Matcher-level list (after vo_match_features / vo_remove_outliers, before vo_bucket_features) */
vo_status vo_refine_matches(vo_engine* e);

/* Replace the raw lists of pairs 0..n_pairs-1 by host lists (lists + i*stride_matches, counts[i] entries; e.g. the
 * raw lists after the host-side removeOutliers, src/matcher.cpp:108) and bucket them (src/matcher.cpp:140-187) in one
 * launch.  The host buffers must stay valid until a later fence / synchronize.  Results: vo_batch_counts /
 * vo_batch_get_matches / vo_batch_download_async with which = 1. */
vo_status vo_batch_bucket_host(vo_engine* e, int32_t n_pairs, const vo_p_match* lists, int64_t stride_matches,
                               const int32_t* counts, int32_t max_features, float bucket_w, float bucket_h);
/* Counts of the last vo_batch_match: raw[i], bucketed[i] (either may be NULL). Synchronises. */
vo_status vo_batch_counts(vo_engine* e, int32_t n_pairs, int32_t* raw, int32_t* bucketed);
/* Copy pair i's raw (which=0) or bucketed (which=1) list to the host. Synchronises. */
vo_status vo_batch_get_matches(vo_engine* e, int32_t pair, int32_t which, vo_p_match* out,
                               int32_t cap, int32_t* n);
/* Enqueue D2H (on the engine's download stream, after the kernels enqueued so far) of all n_pairs
 * count words + lists into caller (ideally pinned) memory without synchronising: counts int32[2*n_pairs] = {raw, bucketed} per pair; lists laid out
 * out + i*stride_matches.  which selects raw (0) / bucketed (1) lists. */
vo_status vo_batch_download_async(vo_engine* e, int32_t n_pairs, int32_t which, int32_t* counts,
                                  vo_p_match* out, int64_t stride_matches, int32_t max_per_pair);
/* parity dump of any frame slot in the reference layout */
vo_status vo_slot_dump_features(vo_engine* e, int32_t slot, int32_t* max2, int32_t* max2_num,
                                int32_t cap_stride);
/* number of features in a slot (sum of max2_num) -- for the algorithmic-bytes formula */
vo_status vo_slot_feature_count(vo_engine* e, int32_t slot, int32_t* n);

/* ---------------------------------------------------------------------------------------- */
/* 4. host stage F2: monocular egomotion from a bucketed match list (no engine, no GPU)        */
/* ---------------------------------------------------------------------------------------- */

/* VisualOdometryMono::parameters, src/viso_mono.h:32-46 + VisualOdometry::calibration, src/viso.h:56-64 */
typedef struct vo_mono_params {
  double f, cu, cv;          /* calibration: focal length, principal point (pixels) */
  double height, pitch;      /* camera height above ground (m) and pitch (rad); reference defaults 1.0 / 0.0 */
  int32_t ransac_iters;      /* 2000 */
  int32_t reserved;
  double inlier_threshold;   /* 1e-5 (Sampson distance in normalised coordinates) */
  double motion_threshold;   /* 100.0 */
} vo_mono_params;

typedef struct vo_egomotion vo_egomotion;

/* Owns the RANSAC sample stream: bit-identical to glibc's rand() after srand(seed), which is what the
 * reference uses (VisualOdometry::VisualOdometry calls srand(0), src/viso.cpp:34; getRandomSample,
 * src/viso.cpp:86-106), but private to the handle.  threads: host threads scoring RANSAC hypotheses
 * (0 = min(hardware threads, 16)); the result does not depend on it. */
vo_status vo_egomotion_create(uint32_t seed, int32_t threads, vo_egomotion** out);
void      vo_egomotion_destroy(vo_egomotion* eg);
/* VisualOdometryMono::estimateMotion(p_matched): src/viso_mono.cpp:41-160.  tr_delta = {rx,ry,rz,tx,ty,tz};
 * *ok = 0 where the reference returns an empty vector (process() == false): fewer than 10 matches or
 * inliers, degenerate normalisation, too little motion. */
vo_status vo_estimate_motion_mono(vo_egomotion* eg, const vo_p_match* matches, int32_t n,
                                  const vo_mono_params* prm, double tr_delta[6], int32_t* ok);
/* The same in two steps, for callers that estimate many frame pairs in parallel and still want the reference's
 * results: the reference's rand() stream runs through the frames of a sequence in order, so the sample sets are
 * DRAWN in frame order (cheap) and the estimation itself is stateless and may run on any thread.
 * samples: int32[ransac_iters * 8]; *n_iters = 0 if the list never reaches RANSAC (the reference then consumes
 * no random numbers either). */
vo_status vo_egomotion_draw_samples(vo_egomotion* eg, const vo_p_match* matches, int32_t n,
                                    const vo_mono_params* prm, int32_t* samples, int32_t* n_iters);
/* ... for the n_lists consecutive pairs of a batch at once (lists + i*stride, counts[i] matches; samples + i*samples_stride*8
 * receives drawn[i] sets): only the generator runs serially, the per-list work runs on `threads` host threads. */
vo_status vo_egomotion_draw_samples_batch(vo_egomotion* eg, int32_t n_lists, const vo_p_match* lists, int64_t stride,
                                          const int32_t* counts, const vo_mono_params* prm, int32_t* samples,
                                          int64_t samples_stride, int32_t* drawn, int32_t threads);
vo_status vo_estimate_motion_mono_samples(const vo_p_match* matches, int32_t n, const vo_mono_params* prm,
                                          const int32_t* samples, int32_t n_iters, int32_t threads,
                                          double tr_delta[6], int32_t* ok, int32_t* inliers,
                                          int32_t inliers_cap, int32_t* n_inliers);
/* ... with the RANSAC winner known already (vo_ransac_gpu_score): best_k = index of the first hypothesis with the largest
 * inlier count, -1 if no hypothesis has an inlier. */
vo_status vo_estimate_motion_mono_best(const vo_p_match* matches, int32_t n, const vo_mono_params* prm,
                                       const int32_t* samples, int32_t n_iters, int32_t best_k, double tr_delta[6],
                                       int32_t* ok, int32_t* inliers, int32_t inliers_cap, int32_t* n_inliers);
/* RANSAC hypothesis scoring on the GPU (csrc/vo_ransac.cu): the 2000 hypotheses of estimateMotion's loop
 * (src/viso_mono.cpp:63-77: fundamentalMatrix of an 8-point sample :242-275, getInlier :277-325) are independent; one
 * thread each runs the host stage's own arithmetic (csrc/vo_ransac_core.h, no FMA contraction on either side), so the inlier
 * counts and the winner -- the FIRST hypothesis with the largest count, as in the reference's sequential loop -- are the
 * host's.  Sample sets still come from vo_egomotion_draw_samples, in frame order.
 * lists + i*stride: counts[i] matches; samples + i*samples_stride*8: drawn[i] sample sets.  best_k[i] = -1 where no
 * hypothesis has an inlier or the list never reaches RANSAC (drawn[i] == 0); best_count (optional) = the winner's count. */
typedef struct vo_ransac_gpu vo_ransac_gpu;
vo_status vo_ransac_gpu_create(int32_t device, int32_t max_lists, int32_t max_points, int32_t max_iters, vo_ransac_gpu** out);
void      vo_ransac_gpu_destroy(vo_ransac_gpu* r);
vo_status vo_ransac_gpu_score(vo_ransac_gpu* r, int32_t n_lists, const vo_p_match* lists, int64_t stride, const int32_t* counts,
                              const vo_mono_params* prm, const int32_t* samples, int64_t samples_stride, const int32_t* drawn,
                              int32_t* best_k, int32_t* best_count);
/* VisualOdometryStereo::parameters, src/viso_stereo.h:31-43 + VisualOdometry::calibration, src/viso.h:56-64 */
typedef struct vo_stereo_params {
  double f, cu, cv;          /* calibration */
  double base;               /* baseline (m); reference default 1.0 */
  int32_t ransac_iters;      /* 200 */
  int32_t reweighting;       /* 1: lower weights towards the image border */
  double inlier_threshold;   /* 2.0 (pixels) */
} vo_stereo_params;
/* VisualOdometryStereo::estimateMotion(p_matched): src/viso_stereo.cpp:54-157 -- 3-point RANSAC over Gauss-Newton on
 * the reprojection error into the current left/right images, refit on the inliers.  Needs quad matches (u2p, u2c,
 * v2c filled).  Draws its samples from the handle's stream (3 per iteration, none for n < 6).  *ok = 0 where the
 * reference returns an empty vector. */
vo_status vo_estimate_motion_stereo(vo_egomotion* eg, const vo_p_match* matches, int32_t n,
                                    const vo_stereo_params* prm, double tr_delta[6], int32_t* ok);
/* VisualOdometry::getInlierIndices / getNumberOfInliers: src/viso.h:96-99 (out may be NULL) */
vo_status vo_egomotion_inliers(const vo_egomotion* eg, int32_t* out, int32_t cap, int32_t* n);
/* next value of the handle's sample stream (== glibc rand()); test hook */
uint32_t  vo_egomotion_rand(vo_egomotion* eg);
/* VisualOdometry::transformationVectorToMatrix: src/viso.cpp:59-84; Tr is a row-major 4x4 */
void      vo_transformation_vector_to_matrix(const double tr[6], double Tr[16]);

/* ---------------------------------------------------------------------------------------- */
/* 5. sequence level: VisualOdometryMono::process for a batch of consecutive frames           */
/* ---------------------------------------------------------------------------------------- */

/* process() (src/viso_mono.cpp:33-39) of up to max_pairs consecutive frames of ONE sequence per call, scheduled as
 * a pipeline (what the reference's FPGA host does by hand, hls_opencl/host/main2.cpp:395-519): GPU extraction +
 * circular matching of the whole batch, host removeOutliers of all pairs on `host_threads` threads, GPU bucketing
 * of all pairs in one launch, RANSAC sample sets drawn in frame order from the sequence's rand() stream, host
 * egomotion of all pairs on the threads.  Per-frame results equal those of process() called frame by frame.
 * The engine needs max_frames >= max_pairs + 1 and must not be used for anything else meanwhile. */
typedef struct vo_pipeline vo_pipeline;
vo_status vo_pipeline_create(vo_engine* e, const vo_mono_params* prm, int32_t max_features, float bucket_w,
                             float bucket_h, int32_t host_threads /* 0 = all */, vo_pipeline** out);
void      vo_pipeline_destroy(vo_pipeline* p);
/* frames at images + i*frame_stride_bytes (host), rows pitch_bytes apart.  Outputs per frame i: ok[i] = process()
 * result, tr_delta[6*i..6*i+5] valid when ok[i], n_matches[i] (after bucketing), n_inliers[i]; matches_out
 * (may be NULL): the bucketed list of frame i at matches_out + i*matches_stride. */
vo_status vo_pipeline_process(vo_pipeline* p, int32_t n, const uint8_t* images, int64_t frame_stride_bytes,
                              int32_t pitch_bytes, int32_t* ok, double* tr_delta, int32_t* n_matches,
                              int32_t* n_inliers, vo_p_match* matches_out, int64_t matches_stride);
/* Wall seconds vo_pipeline_process has spent per stage since creation: {extract + match (GPU), raw lists to the host,
   removeOutliers (host pool), host -> bucketing -> host (GPU), RANSAC sample sets (host, serial), RANSAC scoring (GPU),
   egomotion after the winner (host pool)}. */
vo_status vo_pipeline_stage_seconds(const vo_pipeline* p, double* out7);

/* ---------------------------------------------------------------------------------------- */
/* instrumentation                                                                           */
/* ---------------------------------------------------------------------------------------- */

/* Kernels launched by this engine since creation (every launch of one of OUR kernels counts 1). */
vo_status vo_launch_count(const vo_engine* e, int64_t* n);
/* Accumulated device time per kernel family, measured with CUDA events on the engine stream while
 * profiling is enabled: ms[0]=extract ms[1]=match ms[2]=cycle+emit ms[3]=bucket; calls[] likewise. */
vo_status vo_profile_enable(vo_engine* e, int32_t on);
vo_status vo_profile_read(vo_engine* e, double ms[4], int64_t calls[4], int32_t reset);
/* eight families: the four above, [4] = refine, the rest reserved */
vo_status vo_profile_read_ext(vo_engine* e, double ms[8], int64_t calls[8], int32_t reset);
/* pinned host memory helpers for callers without a CUDA runtime of their own */
vo_status vo_host_alloc(void** p, int64_t bytes);
vo_status vo_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* VO_B200_H */
