// Sequence level of the C ABI: VisualOdometryMono::process (src/viso_mono.cpp:33-39) for a whole batch of
// consecutive frames of ONE sequence, pipelined the way the reference's FPGA host does it by hand
// (hls_opencl/host/main2.cpp:395-519: kernels for frame f in flight while the host finishes frame f-2):
//
//   GPU   : extract n frames -> circular matching of the n pairs            (vo_batch_extract_host / vo_batch_match)
//   host  : removeOutliers (Delaunay support filter) of every pair           on a pool of threads, pairs in parallel
//   GPU   : bucketing of the n filtered lists, one launch                    (vo_batch_bucket_host)
//   host  : RANSAC sample sets drawn IN FRAME ORDER from the sequence's rand() stream (cheap), then
//           estimateMotion of every pair                                     on the pool, pairs in parallel
//
// Results per frame are exactly those of calling process() frame by frame (same stages, same order, same random
// samples); only the schedule differs.  Built on the public entry points of vo_b200.h alone.
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/vo_b200.h"

struct vo_pipeline {
  vo_engine* eng;
  vo_egomotion* ego;
  vo_ransac_gpu* ransac;    // RANSAC hypotheses of the whole batch are scored on the GPU (null: on the host pool)
  vo_mono_params prm;
  vo_geometry geo;
  int32_t max_features;
  float bw, bh;
  int threads;
  int64_t frames_seen;      // frames pushed so far (frame 0 has no previous frame: zero matches, process() == false)
  int32_t last_slot;        // slot holding the newest frame
  vo_p_match* raw;          // pinned: pairs x raw_stride, grown on demand
  int64_t raw_capacity;     // entries
  vo_p_match* bucketed;     // pinned: max_pairs x bucket_cap
  int32_t* counts;          // pinned: 2 x max_pairs
  int32_t bucket_cap;
  std::vector<int32_t> n_raw, n_filtered;
  double stage_s[7];        // wall seconds spent so far: extract+match (GPU), raw lists to host, removeOutliers (host pool),
                            // host->bucket->host (GPU), RANSAC sample sets (host, serial), RANSAC scoring (GPU), egomotion (host pool)
};

namespace {

inline double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <class F>
void parallel_for(int n, int threads, F fn) {
  if (n <= 0) return;
  const int nt = threads < n ? threads : n;
  if (nt <= 1) {
    for (int i = 0; i < n; i++) fn(i);
    return;
  }
  std::atomic<int> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; t++)
    th.emplace_back([&] {
      for (;;) {
        const int i = next.fetch_add(1);
        if (i >= n) break;
        fn(i);
      }
    });
  for (auto& t : th) t.join();
}

}  // namespace

extern "C" {

vo_status vo_pipeline_create(vo_engine* e, const vo_mono_params* prm, int32_t max_features, float bucket_w, float bucket_h,
                             int32_t host_threads, vo_pipeline** out) {
  if (!e || !prm || !out || max_features < 1 || !(bucket_w > 0.f) || !(bucket_h > 0.f)) return VO_ERR_INVALID;
  vo_pipeline* p = new vo_pipeline();
  p->eng = e;
  p->prm = *prm;
  p->max_features = max_features;
  p->bw = bucket_w;
  p->bh = bucket_h;
  vo_status st = vo_engine_geometry(e, &p->geo);
  if (st != VO_OK || p->geo.max_frames < p->geo.max_pairs + 1) {  // a batch of n pairs needs n new frames + the previous one
    delete p;
    return st != VO_OK ? st : VO_ERR_INVALID;
  }
  int hw = (int)std::thread::hardware_concurrency();
  if (hw <= 0) hw = 1;
  p->threads = host_threads > 0 ? host_threads : hw;
  p->frames_seen = 0;
  p->last_slot = -1;
  for (double& x : p->stage_s) x = 0.0;
  // buckets: (W / bw + 1) x (H / bh + 1), max_features each
  p->bucket_cap = ((int32_t)(p->geo.width / bucket_w) + 2) * ((int32_t)(p->geo.height / bucket_h) + 2) * max_features;
  if (p->bucket_cap > p->geo.max_matches) p->bucket_cap = p->geo.max_matches;
  const int64_t B = p->geo.max_pairs;
  void *a = nullptr, *b = nullptr, *c = nullptr;
  p->raw_capacity = B * 4096;
  if (vo_host_alloc(&a, (int64_t)sizeof(vo_p_match) * p->raw_capacity) != VO_OK ||
      vo_host_alloc(&b, (int64_t)sizeof(vo_p_match) * B * p->bucket_cap) != VO_OK ||
      vo_host_alloc(&c, (int64_t)sizeof(int32_t) * 2 * B) != VO_OK || vo_egomotion_create(0u, 1, &p->ego) != VO_OK) {
    if (a) vo_host_free(a);
    if (b) vo_host_free(b);
    if (c) vo_host_free(c);
    delete p;
    return VO_ERR_CUDA;
  }
  p->raw = (vo_p_match*)a;
  p->bucketed = (vo_p_match*)b;
  p->counts = (int32_t*)c;
  p->ransac = nullptr;
  int32_t dev = 0;
  if (prm->ransac_iters > 0 && vo_engine_device(e, &dev) == VO_OK &&
      vo_ransac_gpu_create(dev, (int32_t)B, p->bucket_cap, prm->ransac_iters, &p->ransac) != VO_OK)
    p->ransac = nullptr;  // (out of memory: the host pool scores instead, same results)
  *out = p;
  return VO_OK;
}

void vo_pipeline_destroy(vo_pipeline* p) {
  if (!p) return;
  vo_host_free(p->raw);
  vo_host_free(p->bucketed);
  vo_host_free(p->counts);
  vo_egomotion_destroy(p->ego);
  vo_ransac_gpu_destroy(p->ransac);
  delete p;
}

// n <= max_pairs consecutive frames (images + i*frame_stride_bytes, rows pitch_bytes apart).  Per frame i:
// ok[i] = process() result, tr_delta[6*i..] = (rx,ry,rz,tx,ty,tz) when ok, n_matches[i] = bucketed matches,
// n_inliers[i].  matches_out (optional): bucketed lists at matches_out + i*matches_stride.
vo_status vo_pipeline_process(vo_pipeline* p, int32_t n, const uint8_t* images, int64_t frame_stride_bytes, int32_t pitch_bytes,
                              int32_t* ok, double* tr_delta, int32_t* n_matches, int32_t* n_inliers, vo_p_match* matches_out,
                              int64_t matches_stride) {
  if (!p || !images || !ok || !tr_delta || !n_matches || !n_inliers || n < 1 || n > p->geo.max_pairs) return VO_ERR_INVALID;
  if (matches_out && matches_stride < p->bucket_cap) return VO_ERR_INVALID;
  const int S = p->geo.max_frames;
  int64_t M = p->geo.max_matches;
  vo_status st;
  double t_mark = now_s();
  auto lap = [&](int stage) { const double t = now_s(); p->stage_s[stage] += t - t_mark; t_mark = t; };
  // ---- frames into consecutive slots of the ring (split at the wrap) ------------------------------------------
  const int first = (p->last_slot + 1) % S;
  for (int done = 0; done < n;) {
    const int slot = (first + done) % S;
    const int cnt = (n - done < S - slot) ? n - done : S - slot;
    st = vo_batch_extract_host(p->eng, slot, cnt, images + (int64_t)done * frame_stride_bytes, frame_stride_bytes, pitch_bytes);
    if (st != VO_OK) return st;
    done += cnt;
  }
  // ---- pairs (frame i-1, frame i); the very first frame of the sequence has no partner ---------------------------
  const int skip = (p->frames_seen == 0) ? 1 : 0;
  const int np = n - skip;
  for (int i = 0; i < n; i++) { ok[i] = 0; n_matches[i] = 0; n_inliers[i] = 0; }
  if (np > 0) {
    std::vector<int32_t> prev(np), curr(np);
    for (int i = 0; i < np; i++) {
      curr[i] = (first + skip + i) % S;
      prev[i] = (curr[i] + S - 1) % S;
    }
    st = vo_batch_match(p->eng, np, prev.data(), curr.data(), 0, p->bw, p->bh);
    if (st != VO_OK) return st;
    p->n_raw.assign(np, 0);
    st = vo_batch_counts(p->eng, np, p->n_raw.data(), nullptr);
    if (st != VO_OK) return st;
    lap(0);
    int32_t widest = 1;
    for (int i = 0; i < np; i++) widest = p->n_raw[i] > widest ? p->n_raw[i] : widest;
    M = widest;  // host stride of the raw lists of this batch
    if ((int64_t)np * M > p->raw_capacity) {
      void* a = nullptr;
      if (vo_host_alloc(&a, (int64_t)sizeof(vo_p_match) * np * M * 2) != VO_OK) return VO_ERR_CUDA;
      vo_host_free(p->raw);
      p->raw = (vo_p_match*)a;
      p->raw_capacity = (int64_t)np * M * 2;
    }
    st = vo_batch_download_async(p->eng, np, 0, p->counts, p->raw, M, (int32_t)M);
    if (st != VO_OK) return st;
    st = vo_engine_synchronize(p->eng);
    if (st != VO_OK) return st;
    lap(1);
    // ---- host: Delaunay support filter, pairs in parallel (src/matcher.cpp:108, src/remove_outliers.cpp:4) --------
    p->n_filtered = p->n_raw;
    std::atomic<int> bad(0);
    parallel_for(np, p->threads, [&](int i) {
      if (vo_remove_outliers_host(p->raw + (int64_t)i * M, &p->n_filtered[i]) != VO_OK) bad = 1;
    });
    if (bad) return VO_ERR_INVALID;
    lap(2);
    // ---- GPU: bucketing of the filtered lists (src/matcher.cpp:140-187) -------------------------------------------
    st = vo_batch_bucket_host(p->eng, np, p->raw, M, p->n_filtered.data(), p->max_features, p->bw, p->bh);
    if (st != VO_OK) return st;
    st = vo_batch_download_async(p->eng, np, 1, p->counts, p->bucketed, p->bucket_cap, p->bucket_cap);
    if (st != VO_OK) return st;
    st = vo_engine_synchronize(p->eng);
    if (st != VO_OK) return st;
    lap(3);
    // ---- host: sample sets in frame order, then egomotion of every pair in parallel -------------------------------
    const int iters = p->prm.ransac_iters > 0 ? p->prm.ransac_iters : 0;
    std::vector<int32_t> samples((size_t)np * iters * 8), drawn(np, 0);
    for (int i = 0; i < np; i++)
      if (p->counts[np + i] > p->bucket_cap) return VO_ERR_CAPACITY;
    st = vo_egomotion_draw_samples_batch(p->ego, np, p->bucketed, p->bucket_cap, p->counts + np, &p->prm, samples.data(), iters,
                                         drawn.data(), p->threads);
    if (st != VO_OK) return st;
    lap(4);
    // ---- GPU: every hypothesis of every pair scored in one launch; the host keeps what follows the winner ----------
    std::vector<int32_t> best_k;
    if (p->ransac && iters > 0) {
      best_k.assign(np, -1);
      st = vo_ransac_gpu_score(p->ransac, np, p->bucketed, p->bucket_cap, p->counts + np, &p->prm, samples.data(), iters, drawn.data(),
                               best_k.data(), nullptr);
      if (st != VO_OK) return st;
    }
    lap(5);
    parallel_for(np, p->threads, [&](int i) {
      const int f = i + skip;
      const int32_t nb = p->counts[np + i];
      n_matches[f] = nb;
      const vo_p_match* lst = p->bucketed + (int64_t)i * p->bucket_cap;
      const int32_t* smp = samples.data() + (size_t)i * iters * 8;
      const vo_status s2 = best_k.empty()
          ? vo_estimate_motion_mono_samples(lst, nb, &p->prm, smp, drawn[i], 1, tr_delta + 6 * (size_t)f, &ok[f], nullptr, 0, &n_inliers[f])
          : vo_estimate_motion_mono_best(lst, nb, &p->prm, smp, drawn[i], best_k[i], tr_delta + 6 * (size_t)f, &ok[f], nullptr, 0, &n_inliers[f]);
      if (s2 != VO_OK)
        bad = 1;
      if (matches_out && nb > 0)
        std::memcpy(matches_out + (int64_t)f * matches_stride, p->bucketed + (int64_t)i * p->bucket_cap, sizeof(vo_p_match) * (size_t)nb);
    });
    if (bad) return VO_ERR_INVALID;
    lap(6);
  }
  p->frames_seen += n;
  p->last_slot = (first + n - 1) % S;
  return VO_OK;
}

// wall seconds per stage accumulated by vo_pipeline_process since creation (see vo_pipeline::stage_s)
vo_status vo_pipeline_stage_seconds(const vo_pipeline* p, double* out7) {
  if (!p || !out7) return VO_ERR_INVALID;
  for (int i = 0; i < 7; i++) out7[i] = p->stage_s[i];
  return VO_OK;
}

}  // extern "C"
