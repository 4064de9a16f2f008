// libvo_b200: engine + C ABI (include/vo_b200.h).  Host side only -- all arithmetic of the hot path
// lives in vo_extract.cu / vo_match.cu / vo_bucket.cu.  There is no CPU implementation of any
// operator in this library: without a usable CUDA device vo_engine_create fails.
#include "../../include/vo_b200.h"
#include "vo_common.cuh"

#include <cuda.h>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

namespace vo {
bool k1_supports(int bin_w, int bin_h);
int k2_cand_cap();
void k1_box(int bin_w, int bin_h, int* tw, int* th);
cudaError_t init_k1_device();
cudaError_t init_k2_device();
cudaError_t init_k4_device();
cudaError_t init_k4_noalloc_device();
int k4_noalloc_max_matches();
cudaError_t launch_k4_noalloc(int device, const void* in, int32_t n_in, int32_t max_features, float bw, float bh,
                              const uint32_t* lfsr, int32_t lfsr_len, void* out, int32_t out_cap, int32_t* result,
                              cudaStream_t st);
cudaError_t launch_k1(const CUtensorMap& tmap, const Geom& g, const FrameArrays& fa, int first_slot, int z_base, int n,
                      cudaStream_t st);
cudaError_t launch_dump(const Geom& g, const FrameArrays& fa, int slot, int32_t* max2, int32_t* max2_num,
                        int cap_stride, int32_t* overflow, cudaStream_t st);
cudaError_t launch_load(const Geom& g, const FrameArrays& fa, int slot, const int32_t* max2, const int32_t* max2_num,
                        int cap_stride, cudaStream_t st);
cudaError_t launch_k2(const Geom& g, const FrameArrays& fa, const int4* edges, int n_edges, uint32_t* best,
                      int rows_per_group, int prune, int threads, int seed_good, int split, int lean, cudaStream_t st);
cudaError_t launch_k3_quad(const Geom& g, const FrameArrays& fa, const int4* quads, const int2* pairs, int n_items,
                           const uint32_t* best, uint8_t* flags, int32_t* list_cnt, uint32_t* qxy, void* out,
                           int64_t out_stride_matches, int32_t max_matches, int32_t* n_raw, int32_t* status,
                           uint32_t* side, cudaStream_t st);
cudaError_t launch_k5(const Geom& g, const FrameArrays& fa, const int4* quads, int n_items, void* lists,
                      int64_t stride_matches, int32_t max_n, int32_t* n_list, uint8_t* keep, cudaStream_t st);
cudaError_t launch_k3(const Geom& g, const FrameArrays& fa, const int2* pairs, int n_pairs, const uint32_t* best,
                      uint8_t* flags, int32_t* list_cnt, void* out, int64_t out_stride_matches, int32_t max_matches,
                      int32_t* n_raw, int32_t* status, uint32_t* side, cudaStream_t st);
int64_t k4_scratch_ints_host(int nb_max, int q_cap);
cudaError_t launch_k4(const void* in, int64_t in_stride_matches, const int32_t* n_in, int n_pairs, int32_t max_features,
                      float bw, float bh, const uint32_t* lfsr, int32_t lfsr_len, int32_t* scratch, int64_t scratch_stride,
                      int32_t nb_max, int32_t q_cap, void* out, int64_t out_stride_matches, int32_t out_cap,
                      int32_t* n_out, int32_t* status, const uint32_t* side, cudaStream_t st);
}  // namespace vo

namespace {

thread_local std::string g_err;

vo_status fail(vo_status s, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return s;
}

#define CU_TRY(expr)                                                                              \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) return fail(VO_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e));     \
  } while (0)

constexpr int kLfsrLen = 1 << 18;
constexpr int kNbMax = 4096;   // buckets per list supported by K4
constexpr int kQCap = 16384;   // kept matches per list supported by K4

enum Family { F_EXTRACT = 0, F_MATCH = 1, F_EMIT = 2, F_BUCKET = 3, F_REFINE = 4, F_FAMILIES = 8 };
constexpr int kFlowReach = 1 << 10;  // row reach of a flow edge: no restriction beyond the reference's vb +- 3

}  // namespace

struct vo_engine {
  vo_config cfg{};
  vo::Geom g{};
  int device = 0;
  cudaStream_t stream = nullptr;       // compute (and every synchronous entry point)
  bool own_stream = false;
  // batch-level software pipeline: uploads and downloads run on their own streams so that the
  // H2D copy of batch s+1 and the D2H copy of batch s-1 overlap the kernels of batch s
  cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
  cudaEvent_t ev_h2d_done = nullptr, ev_k1_done = nullptr, ev_compute_done = nullptr, ev_d2h_done = nullptr;
  bool d2h_pending = false;
  cudaEvent_t fences[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int n_slots = 0, n_pairbuf = 0;       // incl. the operator-level scratch slots / pair
  int op_slot_a = 0, op_slot_b = 0, op_pair = 0;
  int rows_per_group = 0, prune = 1, match_threads = 0;
  int lean_match = 0;                   // K2 staging layout: 0 = descriptors in shared memory, 1 = in global memory (tall frames)
  int split_match = 1;  // mono pairs: prev->curr in full, then curr->prev from hints (VO_MATCH_SPLIT=0: one joint launch; developer A/B switch)
  vo::FrameArrays fa{};
  CUtensorMap tmap_slots;
  // pair buffers
  int2* d_pairs = nullptr;
  int2* h_pairs = nullptr;              // pinned
  int epi = 2;                          // K2 edges (and `best` arrays) per pair buffer: 2, or 4 on a stereo engine
  int4* d_edges = nullptr;              // [n_pairbuf * epi] {origin slot, target slot, rad_v, row reach}
  int4* h_edges = nullptr;              // pinned
  int4* d_quads = nullptr;              // [n_pairbuf] {1p, 2p, 1c, 2c} slots of the last match call (2p = 2c = -1: flow)
  int4* h_quads = nullptr;              // pinned
  // The pair/edge/quad tables (pinned host copy + device copy) rotate through kStageGens generations.  Host side: they
  // are read by ASYNCHRONOUS copies, and a pipelined caller (bench.py's e2e loop) fills the next call's tables while the
  // previous copy has not run yet.  Device side: the copies travel on the UPLOAD stream, behind the frame upload of the
  // same step -- on the compute stream they would queue on the H2D copy engine behind the next step's 60 MB upload and
  // stall the kernels for a millisecond (measured: e2e 64k -> 47k pairs/s) -- so a generation's device table may only
  // be rewritten once the kernels that read it are done (stage_begin / stage_end).
  static constexpr int kStageGens = 4;
  int2* h_pairs_base = nullptr; int4* h_edges_base = nullptr; int4* h_quads_base = nullptr;
  int2* d_pairs_base = nullptr; int4* d_edges_base = nullptr; int4* d_quads_base = nullptr;
  cudaEvent_t stage_done[kStageGens] = {nullptr, nullptr, nullptr, nullptr};   // copies of the generation done (upload stream)
  cudaEvent_t stage_idle[kStageGens] = {nullptr, nullptr, nullptr, nullptr};   // kernels reading the generation enqueued up to here
  bool stage_used[kStageGens] = {false, false, false, false};
  int stage_gen = 0;
  uint32_t* d_qxy = nullptr;            // stereo engines: partner coordinates per current-left feature (k3q_flags)
  uint8_t* d_keep = nullptr;            // refinement: keep flag per match
  std::vector<uint8_t> slot_has_image;  // frame bytes resident in the slot (not when extraction read caller memory in place)
  // Kernels that read slot IMAGES (K1 always, K5 for the frames of its items): an upload on the copy stream into a slot
  // waits for exactly the recent readers of that slot, so uploads into other slots overlap them.  A record pushed out of
  // the ring makes the copy stream wait for it unconditionally (the copy stream is FIFO, so that holds for good).
  struct ImgReader { cudaEvent_t ev = nullptr; std::vector<uint8_t> slots; bool used = false; };
  static constexpr int kReaders = 8;
  ImgReader readers[kReaders];
  int reader_next = 0;
  int ring_r_prev = 2, ring_r_curr = 3; // stereo engines: right-image ring
  int right_pushed = 0;
  bool last_was_quad = false;
  uint32_t* d_best = nullptr;
  uint8_t* d_flags = nullptr;
  int32_t* d_list_cnt = nullptr;
  vo_p_match* d_raw = nullptr;          // [n_pairbuf][max_matches]
  uint32_t* d_side = nullptr;           // [n_pairbuf][max_matches] (u1c | v1c << 16) of the records K3 emitted last (for K4)
  vo_p_match* d_bucketed = nullptr;     // [n_pairbuf][max_matches]
  int32_t* d_n_raw = nullptr;           // [n_pairbuf]
  int32_t* d_n_bucketed = nullptr;      // [n_pairbuf]
  int32_t* d_status = nullptr;          // sticky device-side status word
  int32_t* h_counts = nullptr;          // pinned [2*n_pairbuf + 4]
  int32_t* d_scratch = nullptr;
  int64_t scratch_stride = 0;
  uint32_t* d_lfsr = nullptr;
  // operator-level staging
  int32_t* d_max2 = nullptr; size_t d_max2_bytes = 0;
  int32_t* d_max2b = nullptr; size_t d_max2b_bytes = 0;
  int32_t* d_num = nullptr;             // [2][bin_num]
  // Matcher-level state
  int ring_prev = 0, ring_curr = 1;
  int frames_pushed = 0;
  int cur_which = 0;                    // 0: raw list is current, 1: bucketed list
  int last_batch_pairs = 0;
  // instrumentation
  int64_t launches = 0;
  bool profiling = false;
  struct Ev { cudaEvent_t a, b; int fam; };
  std::vector<Ev> evs;
  size_t ev_used = 0;
  double prof_ms[F_FAMILIES] = {0, 0, 0, 0, 0, 0, 0, 0};
  int64_t prof_calls[F_FAMILIES] = {0, 0, 0, 0, 0, 0, 0, 0};
};

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 3-D u8 tensor (x, y, frame) with the tile kernel's box (96x59x1 for 60x45 bins); out-of-bounds reads return zeros.
vo_status make_tmap(CUtensorMap* m, const void* base, int W, int H, int n, int64_t pitch, int64_t frame_stride, int bin_w, int bin_h) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail(VO_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n};
  cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frame_stride};
  int tw = 0, th = 0;
  vo::k1_box(bin_w, bin_h, &tw, &th);
  if (tw <= 0) return fail(VO_ERR_UNSUPPORTED, "bin size %dx%d", bin_w, bin_h);
  cuuint32_t box[3] = {(cuuint32_t)tw, (cuuint32_t)th, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(VO_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return VO_OK;
}

// bit model of Matcher::rand_number (src/matcher.cpp:113-124) on x86; see oracle/vo_oracle.c
uint32_t lfsr_next(uint32_t x) {
  uint32_t t0 = (x >= 0x80000000u) ? 0u : (x & 1u);
  uint32_t b = (t0 ^ (x >> 10) ^ (x >> 30) ^ (x >> 31)) & 1u;
  return (x >> 1) | (b << 31);
}

struct Scope {  // brackets one kernel family launch with CUDA events when profiling is on
  vo_engine* e; int fam; size_t idx = (size_t)-1;
  Scope(vo_engine* e_, int fam_) : e(e_), fam(fam_) {
    if (!e->profiling) return;
    if (e->ev_used == e->evs.size()) {
      vo_engine::Ev ev{};
      cudaEventCreate(&ev.a); cudaEventCreate(&ev.b);
      e->evs.push_back(ev);
    }
    idx = e->ev_used++;
    e->evs[idx].fam = fam;
    cudaEventRecord(e->evs[idx].a, e->stream);
  }
  ~Scope() {
    if (idx != (size_t)-1) cudaEventRecord(e->evs[idx].b, e->stream);
  }
};

void drain_events(vo_engine* e) {
  for (size_t i = 0; i < e->ev_used; i++) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e->evs[i].a, e->evs[i].b) == cudaSuccess) {
      e->prof_ms[e->evs[i].fam] += ms;
      e->prof_calls[e->evs[i].fam] += 1;
    }
  }
  e->ev_used = 0;
}

vo_status check_device_status(vo_engine* e) {
  int32_t st = 0;
  CU_TRY(cudaMemcpyAsync(&e->h_counts[2 * e->n_pairbuf], e->d_status, 4, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  st = e->h_counts[2 * e->n_pairbuf];
  if (st != 0) {
    cudaMemsetAsync(e->d_status, 0, 4, e->stream);
    return fail(VO_ERR_CAPACITY, "device capacity flag %d (1: raw match list, 2: bucket grid/LFSR table, 3: bucketed list)", st);
  }
  return VO_OK;
}

// registers a kernel just launched on the compute stream as a reader of the images of the slots `mark` flags
template <class Mark>
vo_status note_image_reader(vo_engine* e, Mark mark) {
  vo_engine::ImgReader& r = e->readers[e->reader_next];
  e->reader_next = (e->reader_next + 1) % vo_engine::kReaders;
  if (r.used) CU_TRY(cudaStreamWaitEvent(e->h2d_stream, r.ev, 0));
  r.slots.assign((size_t)e->n_slots, 0);
  mark(r.slots);
  CU_TRY(cudaEventRecord(r.ev, e->stream));
  r.used = true;
  return VO_OK;
}

vo_status do_extract(vo_engine* e, const CUtensorMap& tm, int first_slot, int z_base, int n) {
  const vo::Geom& g = e->g;
  // v-bins the reference never flushes read as empty: clear all counts of the target slots
  CU_TRY(cudaMemsetAsync(e->fa.cnt + (int64_t)first_slot * e->fa.cnt_stride, 0, sizeof(int32_t) * g.bin_num * (size_t)n,
                         e->stream));
  {
    Scope sc(e, F_EXTRACT);
    CU_TRY(vo::launch_k1(tm, g, e->fa, first_slot, z_base, n, e->stream));
    e->launches += 1;
  }
  if (&tm == &e->tmap_slots)  // (an extraction through a tensor map of caller memory reads no slot image)
    return note_image_reader(e, [&](std::vector<uint8_t>& sl) { for (int i = 0; i < n; i++) sl[first_slot + i] = 1; });
  return VO_OK;
}

vo_status upload_images(vo_engine* e, int first_slot, int n, const uint8_t* images, int64_t frame_stride, int pitch,
                        cudaMemcpyKind kind, cudaStream_t st = nullptr) {
  if (!st) st = e->stream;
  const vo::Geom& g = e->g;
  for (int i = 0; i < n; i++) e->slot_has_image[first_slot + i] = 1;
  uint8_t* dst = e->fa.img + (int64_t)first_slot * e->fa.img_stride;
  if (pitch == g.pitch && frame_stride == e->fa.img_stride) {
    // caller's layout == slot layout (16-byte row pitch): one linear copy at full link speed; the
    // pad bytes beyond W travel too but TMA never reads past the tensor's W
    CU_TRY(cudaMemcpyAsync(dst, images, (size_t)e->fa.img_stride * n, kind, st));
  } else if (frame_stride == (int64_t)pitch * g.H) {
    CU_TRY(cudaMemcpy2DAsync(dst, g.pitch, images, pitch, g.W, (size_t)g.H * n, kind, st));
  } else {
    for (int i = 0; i < n; i++)
      CU_TRY(cudaMemcpy2DAsync(dst + (int64_t)i * e->fa.img_stride, g.pitch, images + (int64_t)i * frame_stride, pitch,
                               g.W, g.H, kind, st));
  }
  return VO_OK;
}

vo_status do_match(vo_engine* e, int pair0, int n_pairs, int max_features, float bw, float bh) {
  const vo::Geom& g = e->g;
  if (e->d2h_pending) {  // the pair buffers are still being downloaded by the previous batch
    CU_TRY(cudaStreamWaitEvent(e->stream, e->ev_d2h_done, 0));
    e->d2h_pending = false;
  }
  {
    Scope sc(e, F_MATCH);
    CU_TRY(vo::launch_k2(g, e->fa, e->d_edges + (int64_t)pair0 * 2, 2 * n_pairs,
                         e->d_best + (int64_t)pair0 * 2 * g.U * 4 * g.col_cap, e->rows_per_group, e->prune,
                         e->match_threads, e->cfg.match_seed_good, e->split_match, e->lean_match, e->stream));
    e->launches += (e->prune && e->split_match) ? 3 : 1;
  }
  {
    Scope sc(e, F_EMIT);
    CU_TRY(vo::launch_k3(g, e->fa, e->d_pairs + pair0, n_pairs, e->d_best + (int64_t)pair0 * 2 * g.U * 4 * g.col_cap,
                         e->d_flags + (int64_t)pair0 * g.U * 4 * g.col_cap, e->d_list_cnt + (int64_t)pair0 * g.U * 4,
                         e->d_raw + (int64_t)pair0 * e->cfg.max_matches, e->cfg.max_matches, e->cfg.max_matches,
                         e->d_n_raw + pair0, e->d_status, e->d_side + (int64_t)pair0 * e->cfg.max_matches, e->stream));
    e->launches += 2;
  }
  if (max_features > 0) {
    Scope sc(e, F_BUCKET);
    CU_TRY(vo::launch_k4(e->d_raw + (int64_t)pair0 * e->cfg.max_matches, e->cfg.max_matches, e->d_n_raw + pair0, n_pairs,
                         max_features, bw, bh, e->d_lfsr, kLfsrLen, e->d_scratch + (int64_t)pair0 * e->scratch_stride,
                         e->scratch_stride, kNbMax, kQCap, e->d_bucketed + (int64_t)pair0 * e->cfg.max_matches,
                         e->cfg.max_matches, e->cfg.max_matches, e->d_n_bucketed + pair0, e->d_status,
                         e->d_side + (int64_t)pair0 * e->cfg.max_matches, e->stream));
    e->launches += 1;
  }
  return VO_OK;
}

// next generation of the pinned pair/edge/quad tables; blocks only if the host runs kStageGens calls ahead of the GPU
vo_status stage_begin(vo_engine* e) {
  CU_TRY(cudaEventRecord(e->stage_idle[e->stage_gen], e->stream));  // every kernel that reads the current tables is enqueued
  e->stage_gen = (e->stage_gen + 1) % vo_engine::kStageGens;
  const int gi = e->stage_gen;
  if (e->stage_used[gi]) {
    CU_TRY(cudaEventSynchronize(e->stage_done[gi]));                 // host table: its last copy has run
    CU_TRY(cudaStreamWaitEvent(e->h2d_stream, e->stage_idle[gi], 0));  // device table: its last readers are done
  }
  e->h_pairs = e->h_pairs_base + (int64_t)gi * e->n_pairbuf;
  e->h_edges = e->h_edges_base + (int64_t)gi * e->n_pairbuf * e->epi;
  e->h_quads = e->h_quads_base + (int64_t)gi * e->n_pairbuf;
  e->d_pairs = e->d_pairs_base + (int64_t)gi * e->n_pairbuf;
  e->d_edges = e->d_edges_base + (int64_t)gi * e->n_pairbuf * e->epi;
  e->d_quads = e->d_quads_base + (int64_t)gi * e->n_pairbuf;
  return VO_OK;
}
// after the table copies have been enqueued on the upload stream
vo_status stage_end(vo_engine* e) {
  CU_TRY(cudaEventRecord(e->stage_done[e->stage_gen], e->h2d_stream));
  CU_TRY(cudaStreamWaitEvent(e->stream, e->stage_done[e->stage_gen], 0));
  e->stage_used[e->stage_gen] = true;
  return VO_OK;
}

vo_status set_pairs(vo_engine* e, int pair0, int n, const int32_t* prev, const int32_t* curr) {
  // validate everything BEFORE stage_begin advances the table generation: an error must leave the current tables intact
  if (pair0 < 0 || n < 0 || pair0 + n > e->n_pairbuf) return fail(VO_ERR_INVALID, "pairs [%d, %d) outside the %d pair buffers", pair0, pair0 + n, e->n_pairbuf);
  for (int i = 0; i < n; i++)
    if (prev[i] < 0 || prev[i] >= e->n_slots || curr[i] < 0 || curr[i] >= e->n_slots)
      return fail(VO_ERR_INVALID, "pair %d: slot out of range", i);
  vo_status sb = stage_begin(e);
  if (sb != VO_OK) return sb;
  for (int i = 0; i < n; i++) {
    e->h_pairs[pair0 + i] = make_int2(prev[i], curr[i]);
    // the two searches of myMatching (src/myMatch.cpp:148-176): curr -> prev, prev -> curr, both with the flow window
    e->h_edges[2 * (pair0 + i)] = make_int4(curr[i], prev[i], e->g.rad_v, kFlowReach);
    e->h_edges[2 * (pair0 + i) + 1] = make_int4(prev[i], curr[i], e->g.rad_v, kFlowReach);
    e->h_quads[pair0 + i] = make_int4(prev[i], -1, curr[i], -1);
  }
  CU_TRY(cudaMemcpyAsync(e->d_pairs + pair0, e->h_pairs + pair0, sizeof(int2) * n, cudaMemcpyHostToDevice, e->h2d_stream));
  CU_TRY(cudaMemcpyAsync(e->d_edges + 2 * pair0, e->h_edges + 2 * pair0, sizeof(int4) * 2 * n, cudaMemcpyHostToDevice, e->h2d_stream));
  CU_TRY(cudaMemcpyAsync(e->d_quads + pair0, e->h_quads + pair0, sizeof(int4) * n, cudaMemcpyHostToDevice, e->h2d_stream));
  e->last_was_quad = false;
  return stage_end(e);
}

// EXTENSION (row F3): sub-pixel refinement of the raw lists of items [item0, item0 + n) in place
vo_status do_refine(vo_engine* e, int item0, int n_items, int32_t max_n) {
  for (int i = 0; i < n_items; i++) {
    const int4 q = e->h_quads[item0 + i];
    for (int s : {q.x, q.y, q.z, q.w})
      if (s >= 0 && !e->slot_has_image[s])
        return fail(VO_ERR_STATE, "refinement needs the frame of slot %d resident in the engine (the slot was never filled, or "
                    "it was extracted in place from caller memory by vo_batch_extract_device)", s);
  }
  if (!e->d_keep) {
    cudaError_t ce = cudaMalloc((void**)&e->d_keep, (size_t)e->n_pairbuf * e->cfg.max_matches);
    if (ce != cudaSuccess) return fail(VO_ERR_CUDA, "cudaMalloc(keep flags): %s", cudaGetErrorString(ce));
  }
  Scope sc(e, F_REFINE);
  CU_TRY(vo::launch_k5(e->g, e->fa, e->d_quads + item0, n_items, e->d_raw + (int64_t)item0 * e->cfg.max_matches,
                       e->cfg.max_matches, max_n, e->d_n_raw + item0, e->d_keep + (int64_t)item0 * e->cfg.max_matches,
                       e->stream));
  e->launches += 2;
  return note_image_reader(e, [&](std::vector<uint8_t>& sl) {
    for (int i = 0; i < n_items; i++) {
      const int4 q = e->h_quads[item0 + i];
      for (int s : {q.x, q.y, q.z, q.w}) if (s >= 0) sl[s] = 1;
    }
  });
}

// EXTENSION (row F3): quad matching of items [item0, item0 + n): four K2 edges, cycle + emit, optional refinement
// and bucketing.  h_quads/h_pairs/h_edges of the items must be set (set_quads).
vo_status do_match_quad(vo_engine* e, int item0, int n_items, int refinement, int max_features, float bw, float bh) {
  const vo::Geom& g = e->g;
  if (e->d2h_pending) {
    CU_TRY(cudaStreamWaitEvent(e->stream, e->ev_d2h_done, 0));
    e->d2h_pending = false;
  }
  const int64_t lists = (int64_t)g.U * 4 * g.col_cap;
  {
    Scope sc(e, F_MATCH);
    CU_TRY(vo::launch_k2(g, e->fa, e->d_edges + (int64_t)item0 * 4, 4 * n_items, e->d_best + (int64_t)item0 * 4 * lists,
                         e->rows_per_group, e->prune, e->match_threads, e->cfg.match_seed_good, 0, e->lean_match, e->stream));
    e->launches += 1;
  }
  {
    Scope sc(e, F_EMIT);
    CU_TRY(vo::launch_k3_quad(g, e->fa, e->d_quads + item0, e->d_pairs + item0, n_items, e->d_best + (int64_t)item0 * 4 * lists,
                              e->d_flags + (int64_t)item0 * lists, e->d_list_cnt + (int64_t)item0 * g.U * 4,
                              e->d_qxy + (int64_t)item0 * lists * 3, e->d_raw + (int64_t)item0 * e->cfg.max_matches,
                              e->cfg.max_matches, e->cfg.max_matches, e->d_n_raw + item0, e->d_status,
                              e->d_side + (int64_t)item0 * e->cfg.max_matches, e->stream));
    e->launches += 2;
  }
  if (refinement == 2) {
    vo_status s = do_refine(e, item0, n_items, e->cfg.max_matches);
    if (s != VO_OK) return s;
  }
  if (max_features > 0) {
    Scope sc(e, F_BUCKET);
    CU_TRY(vo::launch_k4(e->d_raw + (int64_t)item0 * e->cfg.max_matches, e->cfg.max_matches, e->d_n_raw + item0, n_items,
                         max_features, bw, bh, e->d_lfsr, kLfsrLen, e->d_scratch + (int64_t)item0 * e->scratch_stride,
                         e->scratch_stride, kNbMax, kQCap, e->d_bucketed + (int64_t)item0 * e->cfg.max_matches,
                         e->cfg.max_matches, e->cfg.max_matches, e->d_n_bucketed + item0, e->d_status,
                         // the side array holds the integer coordinates K3 emitted: stale once K5 refined or compacted the list
                         refinement == 2 ? nullptr : e->d_side + (int64_t)item0 * e->cfg.max_matches, e->stream));
    e->launches += 1;
  }
  return VO_OK;
}

vo_status set_quads(vo_engine* e, int item0, int n, const int32_t* s1p, const int32_t* s2p, const int32_t* s1c,
                    const int32_t* s2c, int disp_tolerance) {
  if (e->epi != 4) return fail(VO_ERR_STATE, "quad matching needs an engine created with vo_config.stereo = 1");
  if (disp_tolerance < 0 || disp_tolerance > 16) return fail(VO_ERR_INVALID, "match_disp_tolerance %d outside [0,16]", disp_tolerance);
  if (item0 < 0 || n < 0 || item0 + n > e->n_pairbuf) return fail(VO_ERR_INVALID, "items [%d, %d) outside the %d pair buffers", item0, item0 + n, e->n_pairbuf);
  for (int i = 0; i < n; i++)  // before stage_begin: an error must leave the current tables intact
    for (int s : {s1p[i], s2p[i], s1c[i], s2c[i]})
      if (s < 0 || s >= e->n_slots) return fail(VO_ERR_INVALID, "item %d: slot out of range", i);
  vo_status sb = stage_begin(e);
  if (sb != VO_OK) return sb;
  for (int i = 0; i < n; i++) {
    const int k = item0 + i;
    e->h_quads[k] = make_int4(s1p[i], s2p[i], s1c[i], s2c[i]);
    e->h_pairs[k] = make_int2(s1p[i], s1c[i]);
    e->h_edges[4 * k + 0] = make_int4(s1c[i], s1p[i], e->g.rad_v, kFlowReach);  // flow   1c -> 1p
    e->h_edges[4 * k + 1] = make_int4(s1p[i], s2p[i], disp_tolerance, 1);       // stereo 1p -> 2p
    e->h_edges[4 * k + 2] = make_int4(s2p[i], s2c[i], e->g.rad_v, kFlowReach);  // flow   2p -> 2c
    e->h_edges[4 * k + 3] = make_int4(s2c[i], s1c[i], disp_tolerance, 1);       // stereo 2c -> 1c
  }
  CU_TRY(cudaMemcpyAsync(e->d_pairs + item0, e->h_pairs + item0, sizeof(int2) * n, cudaMemcpyHostToDevice, e->h2d_stream));
  CU_TRY(cudaMemcpyAsync(e->d_edges + 4 * item0, e->h_edges + 4 * item0, sizeof(int4) * 4 * n, cudaMemcpyHostToDevice, e->h2d_stream));
  CU_TRY(cudaMemcpyAsync(e->d_quads + item0, e->h_quads + item0, sizeof(int4) * n, cudaMemcpyHostToDevice, e->h2d_stream));
  e->last_was_quad = true;
  return stage_end(e);
}

vo_status ensure(int32_t** p, size_t* have, size_t need) {
  if (*have >= need) return VO_OK;
  if (*p) cudaFree(*p);
  *p = nullptr; *have = 0;
  CU_TRY(cudaMalloc(p, need));
  *have = need;
  return VO_OK;
}

}  // namespace

extern "C" {

const char* vo_last_error(void) { return g_err.c_str(); }
const char* vo_version(void) { return "vo_b200 0.1 (sm_100a)"; }
void vo_descriptor_signature(const uint8_t desc[32], uint32_t sig[2]) {
  uint4 a, b;
  memcpy(&a, desc, 16);
  memcpy(&b, desc + 16, 16);
  const uint2 q = vo::desc_q(a, b);
  sig[0] = q.x;
  sig[1] = q.y;
}

vo_status vo_engine_create(const vo_config* cfg_in, int32_t device, void* stream, vo_engine** out) {
  if (!cfg_in || !out) return fail(VO_ERR_INVALID, "null argument");
  *out = nullptr;
  vo_config cfg = *cfg_in;
  if (cfg.bin_w == 0) cfg.bin_w = 60;
  if (cfg.bin_h == 0) cfg.bin_h = 45;
  if (cfg.nms_tau == 0) cfg.nms_tau = 50;
  if (cfg.match_disp_tolerance == 0) cfg.match_disp_tolerance = 2;  // src/matcher.h:64
  if (cfg.max_frames == 0) cfg.max_frames = cfg.stereo ? 4 : 2;
  if (cfg.stereo != 0 && cfg.stereo != 1) return fail(VO_ERR_INVALID, "vo_config.stereo must be 0 or 1");
  if (cfg.stereo && cfg.max_frames < 4) return fail(VO_ERR_INVALID, "a stereo engine needs max_frames >= 4 (left and right prev/curr)");
  if (cfg.max_pairs == 0) cfg.max_pairs = 1;
  if (cfg.width < 128 || cfg.height < 64 || cfg.width > 16384 || cfg.height > 16384)
    return fail(VO_ERR_INVALID, "image size %dx%d outside [128..16384]x[64..16384]", cfg.width, cfg.height);
  if (!vo::k1_supports(cfg.bin_w, cfg.bin_h))
    return fail(VO_ERR_UNSUPPORTED, "bin size %dx%d: the sm_100a tile kernel is built for 60x45 (the reference's, "
                "src/size.h:10-11), 45x45, 60x30 and 90x30", cfg.bin_w, cfg.bin_h);
  if (cfg.nms_tau < 1 || cfg.nms_tau > 32767) return fail(VO_ERR_INVALID, "nms_tau out of range");
  if (cfg.match_radius_u < 0 || cfg.match_radius_u > 4095 || cfg.match_radius_v < 0 || cfg.match_radius_v > 4095)
    return fail(VO_ERR_INVALID, "match_radius %d x %d outside [1,4095] (0 = three bins)", cfg.match_radius_u, cfg.match_radius_v);
  if (cfg.max_frames < 2 || cfg.max_pairs < 1) return fail(VO_ERR_INVALID, "max_frames >= 2 and max_pairs >= 1 required");

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(VO_ERR_CUDA, "no CUDA device: libvo_b200 has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(VO_ERR_INVALID, "device %d of %d", device, ndev);
  CU_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(VO_ERR_CUDA, "device is sm_%d%d; this library is built for sm_100a only", prop.major, prop.minor);
  // function attributes are per device: every engine sets them on its own GPU (a process may own engines on several)
  CU_TRY(vo::init_k1_device());
  CU_TRY(vo::init_k2_device());
  CU_TRY(vo::init_k4_device());
  CU_TRY(vo::init_k4_noalloc_device());  // address of the static scratch of vo_bucket_matches_noalloc on this device

  vo_engine* e = new (std::nothrow) vo_engine();
  if (!e) return fail(VO_ERR_INVALID, "out of host memory");
  e->device = device;
  vo::Geom& g = e->g;
  g.W = cfg.width; g.H = cfg.height; g.pitch = (cfg.width + 15) & ~15;
  g.bin_w = cfg.bin_w; g.bin_h = cfg.bin_h;
  const int last_i = 9 + 3 * ((g.W - 10 - 9) / 3);
  g.U = last_i / g.bin_w + 1;
  g.V = g.H / g.bin_h + 1;
  g.cap = (g.bin_w / 3) * (g.bin_h / 3);
  g.col_cap = g.cap * g.V;
  g.bin_num = g.U * g.V * 4;
  g.tau = cfg.nms_tau;
  g.j_last = 9 + 3 * ((g.H - 10 - 9) / 3);
  g.v_rem = g.H / g.bin_h;
  g.rad_u = cfg.match_radius_u > 0 ? cfg.match_radius_u : 3 * g.bin_w;  // SEARCH_RAD_U / SEARCH_RAD_V, src/size.h:14-15
  g.rad_v = cfg.match_radius_v > 0 ? cfg.match_radius_v : 3 * g.bin_h;
  if (g.V > 64 || g.col_cap > 65535 || g.U < 2) {
    const int gu = g.U, gv = g.V;
    delete e;
    return fail(VO_ERR_UNSUPPORTED, "geometry U=%d V=%d exceeds kernel tables (V<=64, cap*V<=65535)", gu, gv);
  }
  if (6 * ((g.cap + 3) & ~3) > vo::k2_cand_cap()) {
    const int cap = g.cap;
    delete e;
    return fail(VO_ERR_UNSUPPORTED, "bin capacity %d: six bin lists of a window row exceed the matcher's staging buffer", cap);
  }
  if (cfg.max_matches == 0) cfg.max_matches = g.U * g.V * g.cap;
  e->cfg = cfg;
  e->n_slots = cfg.max_frames + 2; e->op_slot_a = cfg.max_frames; e->op_slot_b = cfg.max_frames + 1;
  e->n_pairbuf = cfg.max_pairs + 1; e->op_pair = cfg.max_pairs;
  // origin v-bin rows per matching CTA: the whole column for KITTI-sized frames, groups of 8 rows for tall frames
  // (2048x1024: V = 23; measured 8.7k -> 9.7k pairs/s), so that a CTA's candidate window fits one staging chunk
  // tall frames use the lean staging layout (descriptors fetched from global memory by the filter's survivors), whose 4096-candidate
  // buffer holds the window of 12 origin rows at once.  VO_MATCH_LEAN = 0 / 1 overrides the choice (developer A/B switch).
  e->lean_match = g.V > 12 ? 1 : 0;
  if (const char* lm = getenv("VO_MATCH_LEAN")) e->lean_match = atoi(lm) != 0;
  e->rows_per_group = cfg.match_rows_per_cta > 0 ? cfg.match_rows_per_cta : (g.V <= 12 ? g.V : (e->lean_match ? 12 : 8));
  e->prune = cfg.match_exhaustive ? 0 : 1;
  if (const char* sp = getenv("VO_MATCH_SPLIT")) e->split_match = atoi(sp) != 0;
  e->epi = cfg.stereo ? 4 : 2;
  e->slot_has_image.assign((size_t)e->n_slots, 0);
  e->match_threads = cfg.match_threads;

  if (stream) { e->stream = (cudaStream_t)stream; e->own_stream = false; }
  else {
    if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) { delete e; return fail(VO_ERR_CUDA, "stream create"); }
    e->own_stream = true;
  }

  {
    bool ok = cudaStreamCreateWithFlags(&e->h2d_stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&e->d2h_stream, cudaStreamNonBlocking) == cudaSuccess;
    cudaEvent_t* evs[4] = {&e->ev_h2d_done, &e->ev_k1_done, &e->ev_compute_done, &e->ev_d2h_done};
    for (auto ev : evs) ok = ok && cudaEventCreateWithFlags(ev, cudaEventDisableTiming) == cudaSuccess;
    for (auto& r : e->readers) ok = ok && cudaEventCreateWithFlags(&r.ev, cudaEventDisableTiming) == cudaSuccess;
    for (auto& f : e->fences) ok = ok && cudaEventCreateWithFlags(&f, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { vo_engine_destroy(e); return fail(VO_ERR_CUDA, "pipeline stream/event create"); }
  }

#define ALLOC(ptr, bytes)                                                                        \
  do {                                                                                           \
    cudaError_t _e = cudaMalloc((void**)&(ptr), (bytes));                                        \
    if (_e != cudaSuccess) { vo_engine_destroy(e); return fail(VO_ERR_CUDA, "cudaMalloc(%zu): %s", (size_t)(bytes), cudaGetErrorString(_e)); } \
  } while (0)

  const int64_t S = e->n_slots, P = e->n_pairbuf;
  e->fa.img_stride = (int64_t)g.pitch * g.H;
  e->fa.cnt_stride = g.bin_num;
  e->fa.rec_stride = (int64_t)g.bin_num * g.cap;
  ALLOC(e->fa.img, (size_t)(S * e->fa.img_stride + 256));
  ALLOC(e->fa.cnt, sizeof(int32_t) * S * e->fa.cnt_stride);
  ALLOC(e->fa.xy, sizeof(uint32_t) * S * e->fa.rec_stride);
  ALLOC(e->fa.meta, sizeof(uint32_t) * S * e->fa.rec_stride);
  ALLOC(e->fa.desc, sizeof(uint4) * 2 * S * e->fa.rec_stride);
  ALLOC(e->fa.q, sizeof(uint2) * S * e->fa.rec_stride);
  if (cfg.stereo) {  // Sobel response planes K1 -> K5 (sub-pixel refinement of quad lists), 2 bytes per pixel and slot
    e->fa.duv_stride = (int64_t)g.pitch * g.H;
    ALLOC(e->fa.duv, sizeof(uint16_t) * S * e->fa.duv_stride);
    cudaMemsetAsync(e->fa.duv, 0, sizeof(uint16_t) * S * e->fa.duv_stride, e->stream);
  }
  cudaMemsetAsync(e->fa.cnt, 0, sizeof(int32_t) * S * e->fa.cnt_stride, e->stream);
  cudaMemsetAsync(e->fa.img, 0, (size_t)(S * e->fa.img_stride + 256), e->stream);
  const int64_t lists = (int64_t)g.U * 4 * g.col_cap;
  ALLOC(e->d_pairs_base, sizeof(int2) * P * vo_engine::kStageGens);
  ALLOC(e->d_best, sizeof(uint32_t) * P * e->epi * lists);
  ALLOC(e->d_edges_base, sizeof(int4) * P * e->epi * vo_engine::kStageGens);
  ALLOC(e->d_quads_base, sizeof(int4) * P * vo_engine::kStageGens);
  e->d_pairs = e->d_pairs_base; e->d_edges = e->d_edges_base; e->d_quads = e->d_quads_base;
  if (cfg.stereo) ALLOC(e->d_qxy, sizeof(uint32_t) * 3 * P * lists);
  ALLOC(e->d_flags, (size_t)(P * lists));
  ALLOC(e->d_list_cnt, sizeof(int32_t) * P * g.U * 4);
  ALLOC(e->d_raw, sizeof(vo_p_match) * P * cfg.max_matches);
  ALLOC(e->d_side, sizeof(uint32_t) * P * cfg.max_matches);
  ALLOC(e->d_bucketed, sizeof(vo_p_match) * P * cfg.max_matches);
  ALLOC(e->d_n_raw, sizeof(int32_t) * P);
  ALLOC(e->d_n_bucketed, sizeof(int32_t) * P);
  ALLOC(e->d_status, sizeof(int32_t));
  ALLOC(e->d_num, sizeof(int32_t) * 2 * g.bin_num);
  cudaMemsetAsync(e->d_n_raw, 0, sizeof(int32_t) * P, e->stream);
  cudaMemsetAsync(e->d_n_bucketed, 0, sizeof(int32_t) * P, e->stream);
  cudaMemsetAsync(e->d_status, 0, sizeof(int32_t), e->stream);
  cudaMemsetAsync(e->d_best, 0xFF, sizeof(uint32_t) * P * e->epi * lists, e->stream);
  e->scratch_stride = vo::k4_scratch_ints_host(kNbMax, kQCap);
  ALLOC(e->d_scratch, sizeof(int32_t) * P * e->scratch_stride);
  ALLOC(e->d_lfsr, sizeof(uint32_t) * kLfsrLen);
  {
    std::vector<uint32_t> tab(kLfsrLen);
    uint32_t x = 5;  // src/matcher.cpp:170
    for (int i = 0; i < kLfsrLen; i++) { tab[i] = x; x = lfsr_next(x); }
    if (cudaMemcpy(e->d_lfsr, tab.data(), sizeof(uint32_t) * kLfsrLen, cudaMemcpyHostToDevice) != cudaSuccess) {
      vo_engine_destroy(e);
      return fail(VO_ERR_CUDA, "lfsr upload");
    }
  }
  {
    bool ok = true;
    for (auto& ev : e->stage_done) ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
    for (auto& ev : e->stage_idle) ok = ok && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { vo_engine_destroy(e); return fail(VO_ERR_CUDA, "staging event create"); }
  }
  if (cudaMallocHost((void**)&e->h_edges_base, sizeof(int4) * P * e->epi * vo_engine::kStageGens) != cudaSuccess ||
      cudaMallocHost((void**)&e->h_quads_base, sizeof(int4) * P * vo_engine::kStageGens) != cudaSuccess ||
      cudaMallocHost((void**)&e->h_pairs_base, sizeof(int2) * P * vo_engine::kStageGens) != cudaSuccess ||
      cudaMallocHost((void**)&e->h_counts, sizeof(int32_t) * (2 * P + 4)) != cudaSuccess) {
    vo_engine_destroy(e);
    return fail(VO_ERR_CUDA, "pinned alloc");
  }
  e->h_pairs = e->h_pairs_base; e->h_edges = e->h_edges_base; e->h_quads = e->h_quads_base;
  for (int64_t i = 0; i < P * vo_engine::kStageGens; i++) e->h_quads_base[i] = make_int4(-1, -1, -1, -1);
  vo_status s = make_tmap(&e->tmap_slots, e->fa.img, g.W, g.H, (int)S, g.pitch, e->fa.img_stride, g.bin_w, g.bin_h);
  if (s != VO_OK) { vo_engine_destroy(e); return s; }
  if (cudaStreamSynchronize(e->stream) != cudaSuccess) { vo_engine_destroy(e); return fail(VO_ERR_CUDA, "init sync"); }
  *out = e;
  return VO_OK;
#undef ALLOC
}

void vo_engine_destroy(vo_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  if (e->h2d_stream) { cudaStreamSynchronize(e->h2d_stream); cudaStreamDestroy(e->h2d_stream); }
  if (e->d2h_stream) { cudaStreamSynchronize(e->d2h_stream); cudaStreamDestroy(e->d2h_stream); }
  for (cudaEvent_t ev : {e->ev_h2d_done, e->ev_k1_done, e->ev_compute_done, e->ev_d2h_done}) if (ev) cudaEventDestroy(ev);
  for (auto& f : e->fences) if (f) cudaEventDestroy(f);
  for (auto& r : e->readers) if (r.ev) cudaEventDestroy(r.ev);
  cudaFree(e->fa.img); cudaFree(e->fa.cnt); cudaFree(e->fa.xy); cudaFree(e->fa.meta); cudaFree(e->fa.desc); cudaFree(e->fa.q); cudaFree(e->fa.duv);
  cudaFree(e->d_pairs_base); cudaFree(e->d_best); cudaFree(e->d_flags); cudaFree(e->d_list_cnt); cudaFree(e->d_raw); cudaFree(e->d_side);
  cudaFree(e->d_bucketed); cudaFree(e->d_n_raw); cudaFree(e->d_n_bucketed); cudaFree(e->d_status);
  cudaFree(e->d_scratch); cudaFree(e->d_lfsr); cudaFree(e->d_max2); cudaFree(e->d_max2b); cudaFree(e->d_num);
  cudaFree(e->d_edges_base); cudaFree(e->d_quads_base); cudaFree(e->d_qxy); cudaFree(e->d_keep);
  if (e->h_edges_base) cudaFreeHost(e->h_edges_base);
  if (e->h_quads_base) cudaFreeHost(e->h_quads_base);
  if (e->h_pairs_base) cudaFreeHost(e->h_pairs_base);
  for (auto& ev : e->stage_done) if (ev) cudaEventDestroy(ev);
  for (auto& ev : e->stage_idle) if (ev) cudaEventDestroy(ev);
  if (e->h_counts) cudaFreeHost(e->h_counts);
  for (auto& ev : e->evs) { cudaEventDestroy(ev.a); cudaEventDestroy(ev.b); }
  if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

vo_status vo_engine_geometry(const vo_engine* e, vo_geometry* out) {
  if (!e || !out) return fail(VO_ERR_INVALID, "null argument");
  std::memset(out, 0, sizeof(*out));
  out->width = e->g.W; out->height = e->g.H; out->pitch = e->g.pitch;
  out->bin_w = e->g.bin_w; out->bin_h = e->g.bin_h;
  out->u_bins = e->g.U; out->v_bins = e->g.V; out->bin_cap = e->g.cap; out->col_cap = e->g.col_cap;
  out->bin_num = e->g.bin_num;
  out->max_frames = e->cfg.max_frames; out->max_pairs = e->cfg.max_pairs; out->max_matches = e->cfg.max_matches;
  return VO_OK;
}

vo_status vo_engine_synchronize(vo_engine* e) {
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  CU_TRY(cudaStreamSynchronize(e->h2d_stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  CU_TRY(cudaStreamSynchronize(e->d2h_stream));
  return VO_OK;
}

vo_status vo_fence_record(vo_engine* e, int32_t idx) {
  if (!e || idx < 0 || idx >= 8) return fail(VO_ERR_INVALID, "fence index outside [0,8)");
  CU_TRY(cudaSetDevice(e->device));
  // everything enqueued so far: the compute stream's tail, then the download stream's tail
  CU_TRY(cudaEventRecord(e->ev_compute_done, e->stream));
  CU_TRY(cudaStreamWaitEvent(e->d2h_stream, e->ev_compute_done, 0));
  CU_TRY(cudaEventRecord(e->fences[idx], e->d2h_stream));
  return VO_OK;
}

vo_status vo_fence_wait(vo_engine* e, int32_t idx) {
  if (!e || idx < 0 || idx >= 8) return fail(VO_ERR_INVALID, "fence index outside [0,8)");
  CU_TRY(cudaEventSynchronize(e->fences[idx]));
  return VO_OK;
}

// ---------------------------------------------------------------------------------------------
// operator level
// ---------------------------------------------------------------------------------------------

static vo_status dump_slot(vo_engine* e, int slot, int32_t* max2, int32_t* max2_num, int32_t cap_stride) {
  const vo::Geom& g = e->g;
  if (!max2 || !max2_num || cap_stride < 1) return fail(VO_ERR_INVALID, "bad dump arguments");
  const size_t bytes = sizeof(int32_t) * 12 * (size_t)g.bin_num * cap_stride;
  vo_status s = ensure(&e->d_max2, &e->d_max2_bytes, bytes);
  if (s != VO_OK) return s;
  int32_t* d_ovf = e->d_num + g.bin_num;  // reuse second half's first word as the overflow flag
  CU_TRY(cudaMemsetAsync(d_ovf, 0, 4, e->stream));
  CU_TRY(cudaMemsetAsync(e->d_max2, 0, bytes, e->stream));
  CU_TRY(vo::launch_dump(g, e->fa, slot, e->d_max2, e->d_num, cap_stride, d_ovf, e->stream));
  e->launches += 1;
  CU_TRY(cudaMemcpyAsync(max2, e->d_max2, bytes, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaMemcpyAsync(max2_num, e->d_num, sizeof(int32_t) * g.bin_num, cudaMemcpyDeviceToHost, e->stream));
  int32_t ovf = 0;
  CU_TRY(cudaMemcpyAsync(&ovf, d_ovf, 4, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  if (ovf) return fail(VO_ERR_CAPACITY, "a (bin,class) list is longer than cap_stride=%d (the reference would overflow)", cap_stride);
  return VO_OK;
}

vo_status vo_compute_features(vo_engine* e, const uint8_t* image, int32_t pitch_bytes, int32_t* max2,
                              int32_t* max2_num, int32_t cap_stride) {
  if (!e || !image) return fail(VO_ERR_INVALID, "null argument");
  if (pitch_bytes < e->g.W) return fail(VO_ERR_INVALID, "ERROR: Image dimension mismatch!");
  CU_TRY(cudaSetDevice(e->device));
  vo_status s = upload_images(e, e->op_slot_a, 1, image, (int64_t)pitch_bytes * e->g.H, pitch_bytes, cudaMemcpyHostToDevice);
  if (s != VO_OK) return s;
  s = do_extract(e, e->tmap_slots, e->op_slot_a, e->op_slot_a, 1);
  if (s != VO_OK) return s;
  return dump_slot(e, e->op_slot_a, max2, max2_num, cap_stride);
}

vo_status vo_matching(vo_engine* e, const int32_t* m1p, const int32_t* n1p, const int32_t* m1c, const int32_t* n1c,
                      int32_t cap_stride, vo_p_match* p_matched, int32_t p_matched_cap, int32_t* p_matched_num) {
  if (!e || !m1p || !n1p || !m1c || !n1c || !p_matched_num || cap_stride < 1) return fail(VO_ERR_INVALID, "null argument");
  CU_TRY(cudaSetDevice(e->device));
  const vo::Geom& g = e->g;
  const size_t bytes = sizeof(int32_t) * 12 * (size_t)g.bin_num * cap_stride;
  vo_status s = ensure(&e->d_max2, &e->d_max2_bytes, bytes);
  if (s != VO_OK) return s;
  s = ensure(&e->d_max2b, &e->d_max2b_bytes, bytes);
  if (s != VO_OK) return s;
  CU_TRY(cudaMemcpyAsync(e->d_max2, m1p, bytes, cudaMemcpyHostToDevice, e->stream));
  CU_TRY(cudaMemcpyAsync(e->d_max2b, m1c, bytes, cudaMemcpyHostToDevice, e->stream));
  CU_TRY(cudaMemcpyAsync(e->d_num, n1p, sizeof(int32_t) * g.bin_num, cudaMemcpyHostToDevice, e->stream));
  CU_TRY(cudaMemcpyAsync(e->d_num + g.bin_num, n1c, sizeof(int32_t) * g.bin_num, cudaMemcpyHostToDevice, e->stream));
  CU_TRY(vo::launch_load(g, e->fa, e->op_slot_a, e->d_max2, e->d_num, cap_stride, e->stream));
  CU_TRY(vo::launch_load(g, e->fa, e->op_slot_b, e->d_max2b, e->d_num + g.bin_num, cap_stride, e->stream));
  e->launches += 2;
  int32_t pv = e->op_slot_a, cu = e->op_slot_b;
  s = set_pairs(e, e->op_pair, 1, &pv, &cu);
  if (s != VO_OK) return s;
  s = do_match(e, e->op_pair, 1, 0, 0.f, 0.f);
  if (s != VO_OK) return s;
  CU_TRY(cudaMemcpyAsync(&e->h_counts[0], e->d_n_raw + e->op_pair, 4, cudaMemcpyDeviceToHost, e->stream));
  s = check_device_status(e);
  if (s != VO_OK) return s;
  const int32_t n = e->h_counts[0];
  *p_matched_num = n;
  if (n > p_matched_cap) return fail(VO_ERR_CAPACITY, "%d matches > caller capacity %d", n, p_matched_cap);
  if (n > 0) {
    CU_TRY(cudaMemcpyAsync(p_matched, e->d_raw + (int64_t)e->op_pair * e->cfg.max_matches, sizeof(vo_p_match) * n,
                           cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(cudaStreamSynchronize(e->stream));
  }
  return VO_OK;
}

vo_status vo_bucket_matches(vo_engine* e, const vo_p_match* in, int32_t n_in, int32_t max_features, float bucket_w,
                            float bucket_h, vo_p_match* out, int32_t out_cap, int32_t* n_out) {
  if (!e || (!in && n_in > 0) || !n_out || n_in < 0) return fail(VO_ERR_INVALID, "null argument");
  if (!(bucket_w > 0.f) || !(bucket_h > 0.f)) return fail(VO_ERR_INVALID, "bucket size must be positive");
  if (n_in > e->cfg.max_matches) return fail(VO_ERR_CAPACITY, "%d matches > engine capacity %d", n_in, e->cfg.max_matches);
  CU_TRY(cudaSetDevice(e->device));
  const int pb = e->op_pair;
  if (n_in > 0)
    CU_TRY(cudaMemcpyAsync(e->d_raw + (int64_t)pb * e->cfg.max_matches, in, sizeof(vo_p_match) * n_in, cudaMemcpyHostToDevice, e->stream));
  e->h_counts[1] = n_in;
  CU_TRY(cudaMemcpyAsync(e->d_n_raw + pb, &e->h_counts[1], 4, cudaMemcpyHostToDevice, e->stream));
  CU_TRY(vo::launch_k4(e->d_raw + (int64_t)pb * e->cfg.max_matches, e->cfg.max_matches, e->d_n_raw + pb, 1, max_features,
                       bucket_w, bucket_h, e->d_lfsr, kLfsrLen, e->d_scratch + (int64_t)pb * e->scratch_stride,
                       e->scratch_stride, kNbMax, kQCap, e->d_bucketed + (int64_t)pb * e->cfg.max_matches,
                       e->cfg.max_matches, e->cfg.max_matches, e->d_n_bucketed + pb, e->d_status, nullptr, e->stream));
  e->launches += 1;
  CU_TRY(cudaMemcpyAsync(&e->h_counts[0], e->d_n_bucketed + pb, 4, cudaMemcpyDeviceToHost, e->stream));
  vo_status s = check_device_status(e);
  if (s != VO_OK) return s;
  const int32_t n = e->h_counts[0];
  *n_out = n;
  if (n > out_cap) return fail(VO_ERR_CAPACITY, "%d bucketed matches > caller capacity %d", n, out_cap);
  if (n > 0) {
    CU_TRY(cudaMemcpyAsync(out, e->d_bucketed + (int64_t)pb * e->cfg.max_matches, sizeof(vo_p_match) * n,
                           cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(cudaStreamSynchronize(e->stream));
  }
  return VO_OK;
}

// No counterpart in the reference: vo_bucket_matches for callers that may not
// allocate.  Device-resident list in, device-resident list out, K4's scratch in a static device array of the library
// (csrc/vo_bucket.cu); two kernel launches on the engine's stream, no cudaMalloc / cudaFree / synchronisation in here.
vo_status vo_bucket_matches_noalloc(vo_engine* e, const vo_p_match* d_in, int32_t n_in, int32_t max_features, float bucket_w,
                                    float bucket_h, vo_p_match* d_out, int32_t out_cap, int32_t* d_result) {
  if (!e || !d_out || !d_result || (!d_in && n_in > 0) || n_in < 0 || out_cap < 0) return fail(VO_ERR_INVALID, "null argument");
  if (!(bucket_w > 0.f) || !(bucket_h > 0.f)) return fail(VO_ERR_INVALID, "bucket size must be positive");
  const int limit = vo::k4_noalloc_max_matches() < kLfsrLen ? vo::k4_noalloc_max_matches() : kLfsrLen;
  if (n_in > limit)
    return fail(VO_ERR_CAPACITY, "%d matches > %d, the limit of the static scratch of vo_bucket_matches_noalloc", n_in, limit);
  if (((uintptr_t)d_in | (uintptr_t)d_out) & 15u) return fail(VO_ERR_INVALID, "d_in and d_out must be 16-byte aligned");
  if (((uintptr_t)d_result) & 3u) return fail(VO_ERR_INVALID, "d_result must be 4-byte aligned");
  CU_TRY(cudaSetDevice(e->device));
  CU_TRY(vo::launch_k4_noalloc(e->device, d_in, n_in, max_features, bucket_w, bucket_h, e->d_lfsr, kLfsrLen, d_out, out_cap,
                               d_result, e->stream));
  e->launches += 2;
  return VO_OK;
}

// the reference's FPGA operator signatures, 1:1 (hls_opencl/kernel/myMatch.hpp:51-57, remove_outliers.hpp:9-11)
vo_status vo_matching_iter(vo_engine* e, const int32_t* m1p, const int32_t* n1p, const int32_t* m1c, const int32_t* n1c,
                           int32_t cap_stride, vo_p_match* p_matched, int32_t p_matched_cap, int32_t* p_matched_num, int32_t iter) {
  if (iter == 0) {  // frame 0 has no previous features: hls_opencl/kernel/myMatch.cpp:221-224
    if (!p_matched_num) return fail(VO_ERR_INVALID, "null argument");
    *p_matched_num = 0;
    return VO_OK;
  }
  return vo_matching(e, m1p, n1p, m1c, n1c, cap_stride, p_matched, p_matched_cap, p_matched_num);
}

vo_status vo_remove_outliers_bucket(vo_engine* e, const vo_p_match* in, const int32_t* in_cnt, vo_p_match* out, int32_t out_cap,
                                    int32_t* out_cnt) {
  if (!e || !in_cnt || !out_cnt || (!in && *in_cnt > 0)) return fail(VO_ERR_INVALID, "null argument");
  int32_t n = *in_cnt;
  if (n <= 3) return VO_OK;  // hls_opencl/kernel/remove_outliers.cpp:20-21: returns with the outputs untouched
  if (n > e->cfg.max_matches) return fail(VO_ERR_CAPACITY, "%d matches > engine capacity %d", n, e->cfg.max_matches);
  std::vector<vo_p_match> tmp(in, in + n);
  vo_status s = vo_remove_outliers_host(tmp.data(), &n);  // Delaunay support filter (host stage), then ...
  if (s != VO_OK) return s;
  return vo_bucket_matches(e, tmp.data(), n, 2, 50.f, 50.f, out, out_cap, out_cnt);  // ... bucketing 2 per 50x50: .cpp:107-109
}

// ---------------------------------------------------------------------------------------------
// Matcher level
// ---------------------------------------------------------------------------------------------

vo_status vo_push_back(vo_engine* e, const uint8_t* I1, const uint8_t* I2, const int32_t dims[3], int32_t replace) {
  // I2: ignored, as in the reference (src/matcher.cpp:51-91 never reads I2), unless the engine was created with
  // vo_config.stereo = 1 (extension): then the right frame goes through the same extraction
  if (!e || !dims) return fail(VO_ERR_INVALID, "null argument");
  const int32_t width = dims[0], height = dims[1], bpl = dims[2];
  if (width <= 0 || height <= 0 || bpl < width || I1 == nullptr)
    return fail(VO_ERR_INVALID, "ERROR: Image dimension mismatch!");  // src/matcher.cpp:59-62
  if (width != e->g.W || height != e->g.H)
    return fail(VO_ERR_INVALID, "image %dx%d does not match engine geometry %dx%d", width, height, e->g.W, e->g.H);
  CU_TRY(cudaSetDevice(e->device));
  if (!replace) {  // prev <- curr is a slot swap, not a 3.66 MB copy (src/matcher.cpp:68-73)
    const int t = e->ring_prev; e->ring_prev = e->ring_curr; e->ring_curr = t;
    e->frames_pushed += 1;
  } else if (e->frames_pushed == 0) {
    e->frames_pushed = 1;
  }
  vo_status s = upload_images(e, e->ring_curr, 1, I1, (int64_t)bpl * height, bpl, cudaMemcpyHostToDevice);
  if (s != VO_OK) return s;
  s = do_extract(e, e->tmap_slots, e->ring_curr, e->ring_curr, 1);
  if (s != VO_OK) return s;
  if (e->cfg.stereo) {
    if (I2) {
      if (!replace) {
        const int t = e->ring_r_prev; e->ring_r_prev = e->ring_r_curr; e->ring_r_curr = t;
        e->right_pushed += 1;
      } else if (e->right_pushed == 0) {
        e->right_pushed = 1;
      }
      s = upload_images(e, e->ring_r_curr, 1, I2, (int64_t)bpl * height, bpl, cudaMemcpyHostToDevice);
      if (s != VO_OK) return s;
      s = do_extract(e, e->tmap_slots, e->ring_r_curr, e->ring_r_curr, 1);
      if (s != VO_OK) return s;
    } else if (!replace) {
      e->right_pushed = 0;  // a mono frame breaks the right-image history
    }
  }
  // the caller's image is borrowed only for the duration of the call
  CU_TRY(cudaStreamSynchronize(e->stream));
  return VO_OK;
}

vo_status vo_match_features(vo_engine* e, int32_t method) {
  // `method` is ignored, as in the reference (src/matcher.cpp:93-111: always the flow matcher), except on a stereo
  // engine, where method 2 runs the quad matcher once two stereo frames have been pushed (extension, row F3)
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  CU_TRY(cudaSetDevice(e->device));
  e->cur_which = 0;
  if (e->frames_pushed < 2) {  // frame-0 contract: zero matches
    CU_TRY(cudaMemsetAsync(e->d_n_raw + 0, 0, 4, e->stream));
    return VO_OK;
  }
  if (e->cfg.stereo && method == 2) {
    if (e->right_pushed < 2) {
      CU_TRY(cudaMemsetAsync(e->d_n_raw + 0, 0, 4, e->stream));
      return VO_OK;
    }
    const int32_t s1p = e->ring_prev, s2p = e->ring_r_prev, s1c = e->ring_curr, s2c = e->ring_r_curr;
    vo_status s = set_quads(e, 0, 1, &s1p, &s2p, &s1c, &s2c, e->cfg.match_disp_tolerance);
    if (s != VO_OK) return s;
    s = do_match_quad(e, 0, 1, 0, 0, 0.f, 0.f);
    if (s != VO_OK) return s;
    return check_device_status(e);
  }
  int32_t pv = e->ring_prev, cu = e->ring_curr;
  vo_status s = set_pairs(e, 0, 1, &pv, &cu);
  if (s != VO_OK) return s;
  s = do_match(e, 0, 1, 0, 0.f, 0.f);
  if (s != VO_OK) return s;
  return check_device_status(e);
}

vo_status vo_bucket_features(vo_engine* e, int32_t max_features, float bucket_w, float bucket_h) {
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  if (!(bucket_w > 0.f) || !(bucket_h > 0.f)) return fail(VO_ERR_INVALID, "bucket size must be positive");
  CU_TRY(cudaSetDevice(e->device));
  const int64_t M = e->cfg.max_matches;
  if (e->cur_which == 1) {  // bucketing an already bucketed list: feed it back as the input list
    CU_TRY(cudaMemcpyAsync(e->d_raw, e->d_bucketed, sizeof(vo_p_match) * M, cudaMemcpyDeviceToDevice, e->stream));
    CU_TRY(cudaMemcpyAsync(e->d_n_raw, e->d_n_bucketed, 4, cudaMemcpyDeviceToDevice, e->stream));
  }
  {
    Scope sc(e, F_BUCKET);
    CU_TRY(vo::launch_k4(e->d_raw, M, e->d_n_raw, 1, max_features, bucket_w, bucket_h, e->d_lfsr, kLfsrLen, e->d_scratch,
                         e->scratch_stride, kNbMax, kQCap, e->d_bucketed, M, (int32_t)M, e->d_n_bucketed, e->d_status, nullptr,
                         e->stream));
    e->launches += 1;
  }
  e->cur_which = 1;
  return check_device_status(e);
}

vo_status vo_num_matches(vo_engine* e, int32_t* n) {
  if (!e || !n) return fail(VO_ERR_INVALID, "null argument");
  CU_TRY(cudaSetDevice(e->device));
  CU_TRY(cudaMemcpyAsync(&e->h_counts[0], e->cur_which ? e->d_n_bucketed : e->d_n_raw, 4, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  *n = e->h_counts[0];
  return VO_OK;
}

vo_status vo_get_matches(vo_engine* e, vo_p_match* out, int32_t cap, int32_t* n) {
  if (!e || !n) return fail(VO_ERR_INVALID, "null argument");
  vo_status s = vo_num_matches(e, n);
  if (s != VO_OK) return s;
  if (*n > cap) return fail(VO_ERR_CAPACITY, "%d matches > caller capacity %d", *n, cap);
  if (*n > 0) {
    if (!out) return fail(VO_ERR_INVALID, "null output");
    CU_TRY(cudaMemcpyAsync(out, e->cur_which ? e->d_bucketed : e->d_raw, sizeof(vo_p_match) * (size_t)*n, cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(cudaStreamSynchronize(e->stream));
  }
  return VO_OK;
}

vo_status vo_set_matches(vo_engine* e, const vo_p_match* in, int32_t n) {
  if (!e || n < 0 || (!in && n > 0)) return fail(VO_ERR_INVALID, "bad argument");
  if (n > e->cfg.max_matches) return fail(VO_ERR_CAPACITY, "%d matches > engine capacity %d", n, e->cfg.max_matches);
  CU_TRY(cudaSetDevice(e->device));
  if (n > 0) CU_TRY(cudaMemcpyAsync(e->d_raw, in, sizeof(vo_p_match) * (size_t)n, cudaMemcpyHostToDevice, e->stream));
  e->h_counts[1] = n;
  CU_TRY(cudaMemcpyAsync(e->d_n_raw, &e->h_counts[1], 4, cudaMemcpyHostToDevice, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  e->cur_which = 0;
  return VO_OK;
}

vo_status vo_remove_outliers(vo_engine* e) {
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  if (e->cur_which != 0) return fail(VO_ERR_STATE, "removeOutliers runs on the raw match list, before bucketing");
  CU_TRY(cudaSetDevice(e->device));
  int32_t n = 0;
  vo_status s = vo_num_matches(e, &n);
  if (s != VO_OK) return s;
  if (n <= 3) return VO_OK;  // src/remove_outliers.cpp:6-7
  std::vector<vo_p_match> host((size_t)n);
  CU_TRY(cudaMemcpyAsync(host.data(), e->d_raw, sizeof(vo_p_match) * (size_t)n, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  s = vo_remove_outliers_host(host.data(), &n);
  if (s != VO_OK) return fail(s, "removeOutliers failed");
  return vo_set_matches(e, host.data(), n);
}

vo_status vo_dump_features(vo_engine* e, int32_t which, int32_t* max2, int32_t* max2_num, int32_t cap_stride) {
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  CU_TRY(cudaSetDevice(e->device));
  return dump_slot(e, which ? e->ring_curr : e->ring_prev, max2, max2_num, cap_stride);
}

// ---------------------------------------------------------------------------------------------
// batch level
// ---------------------------------------------------------------------------------------------

vo_status vo_batch_extract_host(vo_engine* e, int32_t first_slot, int32_t n, const uint8_t* images,
                                int64_t frame_stride_bytes, int32_t pitch_bytes) {
  if (!e || !images) return fail(VO_ERR_INVALID, "null argument");
  if (n < 1 || first_slot < 0 || first_slot + n > e->cfg.max_frames) return fail(VO_ERR_INVALID, "slot range [%d,%d) outside [0,%d)", first_slot, first_slot + n, e->cfg.max_frames);
  if (pitch_bytes < e->g.W || frame_stride_bytes < (int64_t)pitch_bytes * e->g.H) return fail(VO_ERR_INVALID, "ERROR: Image dimension mismatch!");
  CU_TRY(cudaSetDevice(e->device));
  // upload on the copy stream: it only has to wait for the last extraction that read the image
  // slots, so it overlaps the matching kernels of the previous batch
  for (const vo_engine::ImgReader& r : e->readers) {
    if (!r.used) continue;
    static const bool conservative = getenv("VO_CONSERVATIVE_UPLOAD") != nullptr;  // debugging aid: wait for every reader
    bool hit = conservative;
    for (int i = 0; i < n && !hit; i++) hit = r.slots[first_slot + i] != 0;
    if (hit) CU_TRY(cudaStreamWaitEvent(e->h2d_stream, r.ev, 0));
  }
  vo_status s = upload_images(e, first_slot, n, images, frame_stride_bytes, pitch_bytes, cudaMemcpyHostToDevice,
                              e->h2d_stream);
  if (s != VO_OK) return s;
  CU_TRY(cudaEventRecord(e->ev_h2d_done, e->h2d_stream));
  CU_TRY(cudaStreamWaitEvent(e->stream, e->ev_h2d_done, 0));
  return do_extract(e, e->tmap_slots, first_slot, first_slot, n);
}

vo_status vo_batch_extract_device(vo_engine* e, int32_t first_slot, int32_t n, const uint8_t* d_images,
                                  int64_t frame_stride_bytes, int32_t pitch_bytes) {
  if (!e || !d_images) return fail(VO_ERR_INVALID, "null argument");
  if (n < 1 || first_slot < 0 || first_slot + n > e->cfg.max_frames) return fail(VO_ERR_INVALID, "slot range [%d,%d) outside [0,%d)", first_slot, first_slot + n, e->cfg.max_frames);
  if (pitch_bytes < e->g.W || frame_stride_bytes < (int64_t)pitch_bytes * e->g.H) return fail(VO_ERR_INVALID, "ERROR: Image dimension mismatch!");
  CU_TRY(cudaSetDevice(e->device));
  const bool tma_ok = ((uintptr_t)d_images % 16 == 0) && (pitch_bytes % 16 == 0) && (frame_stride_bytes % 16 == 0);
  if (tma_ok) {  // read the caller's HBM-resident frames in place through a tensor map of their own
    CUtensorMap tm;
    vo_status s = make_tmap(&tm, d_images, e->g.W, e->g.H, n, pitch_bytes, frame_stride_bytes, e->g.bin_w, e->g.bin_h);
    if (s != VO_OK) return s;
    for (int i = 0; i < n; i++) e->slot_has_image[first_slot + i] = 0;  // the slots hold features only
    return do_extract(e, tm, first_slot, 0, n);
  }
  vo_status s = upload_images(e, first_slot, n, d_images, frame_stride_bytes, pitch_bytes, cudaMemcpyDeviceToDevice);
  if (s != VO_OK) return s;
  return do_extract(e, e->tmap_slots, first_slot, first_slot, n);
}

vo_status vo_batch_match(vo_engine* e, int32_t n_pairs, const int32_t* prev_slot, const int32_t* curr_slot,
                         int32_t max_features, float bucket_w, float bucket_h) {
  if (!e || !prev_slot || !curr_slot) return fail(VO_ERR_INVALID, "null argument");
  if (n_pairs < 1 || n_pairs > e->cfg.max_pairs) return fail(VO_ERR_INVALID, "n_pairs %d outside [1,%d]", n_pairs, e->cfg.max_pairs);
  if (max_features > 0 && (!(bucket_w > 0.f) || !(bucket_h > 0.f))) return fail(VO_ERR_INVALID, "bucket size must be positive");
  CU_TRY(cudaSetDevice(e->device));
  vo_status s = set_pairs(e, 0, n_pairs, prev_slot, curr_slot);
  if (s != VO_OK) return s;
  e->last_batch_pairs = n_pairs;
  e->cur_which = max_features > 0 ? 1 : 0;
  return do_match(e, 0, n_pairs, max_features, bucket_w, bucket_h);
}

vo_status vo_batch_match_quad(vo_engine* e, int32_t n_items, const int32_t* slot_1p, const int32_t* slot_2p,
                              const int32_t* slot_1c, const int32_t* slot_2c, int32_t disp_tolerance, int32_t refinement,
                              int32_t max_features, float bucket_w, float bucket_h) {
  if (!e || !slot_1p || !slot_2p || !slot_1c || !slot_2c) return fail(VO_ERR_INVALID, "null argument");
  if (n_items < 1 || n_items > e->cfg.max_pairs) return fail(VO_ERR_INVALID, "n_items %d outside [1,%d]", n_items, e->cfg.max_pairs);
  if (refinement != 0 && refinement != 2) return fail(VO_ERR_UNSUPPORTED, "refinement %d: 0 (none) or 2 (sub-pixel)", refinement);
  if (max_features > 0 && (!(bucket_w > 0.f) || !(bucket_h > 0.f))) return fail(VO_ERR_INVALID, "bucket size must be positive");
  CU_TRY(cudaSetDevice(e->device));
  vo_status s = set_quads(e, 0, n_items, slot_1p, slot_2p, slot_1c, slot_2c, disp_tolerance);
  if (s != VO_OK) return s;
  e->last_batch_pairs = n_items;
  e->cur_which = max_features > 0 ? 1 : 0;
  return do_match_quad(e, 0, n_items, refinement, max_features, bucket_w, bucket_h);
}

vo_status vo_batch_refine(vo_engine* e, int32_t n_items) {
  if (!e || n_items < 1 || n_items > e->cfg.max_pairs) return fail(VO_ERR_INVALID, "bad argument");
  CU_TRY(cudaSetDevice(e->device));
  return do_refine(e, 0, n_items, e->cfg.max_matches);
}

vo_status vo_refine_matches(vo_engine* e) {
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  if (e->cur_which != 0) return fail(VO_ERR_STATE, "refinement runs on the match list before bucketing");
  if (e->frames_pushed < 2) return VO_OK;
  CU_TRY(cudaSetDevice(e->device));
  int32_t n = 0;
  vo_status s = vo_num_matches(e, &n);
  if (s != VO_OK || n == 0) return s;
  s = do_refine(e, 0, 1, n);
  if (s != VO_OK) return s;
  return check_device_status(e);
}

vo_status vo_batch_bucket_host(vo_engine* e, int32_t n_pairs, const vo_p_match* lists, int64_t stride_matches,
                               const int32_t* counts, int32_t max_features, float bucket_w, float bucket_h) {
  if (!e || !lists || !counts) return fail(VO_ERR_INVALID, "null argument");
  if (n_pairs < 1 || n_pairs > e->cfg.max_pairs) return fail(VO_ERR_INVALID, "n_pairs %d outside [1,%d]", n_pairs, e->cfg.max_pairs);
  if (max_features < 1 || !(bucket_w > 0.f) || !(bucket_h > 0.f)) return fail(VO_ERR_INVALID, "bad bucketing parameters");
  CU_TRY(cudaSetDevice(e->device));
  if (e->d2h_pending) {  // the pair buffers may still be downloading
    CU_TRY(cudaStreamWaitEvent(e->stream, e->ev_d2h_done, 0));
    e->d2h_pending = false;
  }
  for (int i = 0; i < n_pairs; i++) {
    if (counts[i] < 0 || counts[i] > e->cfg.max_matches || counts[i] > stride_matches)
      return fail(VO_ERR_CAPACITY, "pair %d: %d matches > capacity", i, counts[i]);
    if (counts[i] > 0)
      CU_TRY(cudaMemcpyAsync(e->d_raw + (int64_t)i * e->cfg.max_matches, lists + (int64_t)i * stride_matches,
                             sizeof(vo_p_match) * (size_t)counts[i], cudaMemcpyHostToDevice, e->stream));
  }
  CU_TRY(cudaMemcpyAsync(e->d_n_raw, counts, 4 * (size_t)n_pairs, cudaMemcpyHostToDevice, e->stream));
  {
    Scope sc(e, F_BUCKET);
    CU_TRY(vo::launch_k4(e->d_raw, e->cfg.max_matches, e->d_n_raw, n_pairs, max_features, bucket_w, bucket_h, e->d_lfsr,
                         kLfsrLen, e->d_scratch, e->scratch_stride, kNbMax, kQCap, e->d_bucketed, e->cfg.max_matches,
                         e->cfg.max_matches, e->d_n_bucketed, e->d_status, nullptr, e->stream));
    e->launches += 1;
  }
  e->last_batch_pairs = n_pairs;
  e->cur_which = 1;
  return VO_OK;
}

vo_status vo_batch_counts(vo_engine* e, int32_t n_pairs, int32_t* raw, int32_t* bucketed) {
  if (!e || n_pairs < 1 || n_pairs > e->cfg.max_pairs) return fail(VO_ERR_INVALID, "bad argument");
  CU_TRY(cudaSetDevice(e->device));
  CU_TRY(cudaMemcpyAsync(e->h_counts, e->d_n_raw, 4 * n_pairs, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaMemcpyAsync(e->h_counts + e->n_pairbuf, e->d_n_bucketed, 4 * n_pairs, cudaMemcpyDeviceToHost, e->stream));
  vo_status s = check_device_status(e);
  if (s != VO_OK) return s;
  for (int i = 0; i < n_pairs; i++) {
    if (raw) raw[i] = e->h_counts[i];
    if (bucketed) bucketed[i] = e->h_counts[e->n_pairbuf + i];
  }
  return VO_OK;
}

vo_status vo_batch_get_matches(vo_engine* e, int32_t pair, int32_t which, vo_p_match* out, int32_t cap, int32_t* n) {
  if (!e || !n || pair < 0 || pair >= e->cfg.max_pairs) return fail(VO_ERR_INVALID, "bad argument");
  CU_TRY(cudaSetDevice(e->device));
  CU_TRY(cudaMemcpyAsync(&e->h_counts[0], (which ? e->d_n_bucketed : e->d_n_raw) + pair, 4, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  *n = e->h_counts[0];
  if (*n > cap) return fail(VO_ERR_CAPACITY, "%d matches > caller capacity %d", *n, cap);
  if (*n > 0) {
    if (!out) return fail(VO_ERR_INVALID, "null output");
    CU_TRY(cudaMemcpyAsync(out, (which ? e->d_bucketed : e->d_raw) + (int64_t)pair * e->cfg.max_matches,
                           sizeof(vo_p_match) * (size_t)*n, cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(cudaStreamSynchronize(e->stream));
  }
  return VO_OK;
}

vo_status vo_batch_download_async(vo_engine* e, int32_t n_pairs, int32_t which, int32_t* counts, vo_p_match* out,
                                  int64_t stride_matches, int32_t max_per_pair) {
  if (!e || !counts || n_pairs < 1 || n_pairs > e->cfg.max_pairs) return fail(VO_ERR_INVALID, "bad argument");
  if (out && (max_per_pair < 1 || max_per_pair > e->cfg.max_matches || stride_matches < max_per_pair))
    return fail(VO_ERR_INVALID, "bad list geometry");
  CU_TRY(cudaSetDevice(e->device));
  CU_TRY(cudaEventRecord(e->ev_compute_done, e->stream));
  CU_TRY(cudaStreamWaitEvent(e->d2h_stream, e->ev_compute_done, 0));
  CU_TRY(cudaMemcpyAsync(counts, e->d_n_raw, 4 * n_pairs, cudaMemcpyDeviceToHost, e->d2h_stream));
  CU_TRY(cudaMemcpyAsync(counts + n_pairs, e->d_n_bucketed, 4 * n_pairs, cudaMemcpyDeviceToHost, e->d2h_stream));
  if (out) {
    const vo_p_match* src = which ? e->d_bucketed : e->d_raw;
    CU_TRY(cudaMemcpy2DAsync(out, sizeof(vo_p_match) * stride_matches, src, sizeof(vo_p_match) * (size_t)e->cfg.max_matches,
                             sizeof(vo_p_match) * (size_t)max_per_pair, n_pairs, cudaMemcpyDeviceToHost, e->d2h_stream));
  }
  CU_TRY(cudaEventRecord(e->ev_d2h_done, e->d2h_stream));
  e->d2h_pending = true;
  return VO_OK;
}

vo_status vo_slot_dump_features(vo_engine* e, int32_t slot, int32_t* max2, int32_t* max2_num, int32_t cap_stride) {
  if (!e || slot < 0 || slot >= e->cfg.max_frames) return fail(VO_ERR_INVALID, "bad slot");
  CU_TRY(cudaSetDevice(e->device));
  return dump_slot(e, slot, max2, max2_num, cap_stride);
}

vo_status vo_slot_feature_count(vo_engine* e, int32_t slot, int32_t* n) {
  if (!e || !n || slot < 0 || slot >= e->cfg.max_frames) return fail(VO_ERR_INVALID, "bad slot");
  CU_TRY(cudaSetDevice(e->device));
  std::vector<int32_t> c(e->g.bin_num);
  CU_TRY(cudaMemcpyAsync(c.data(), e->fa.cnt + (int64_t)slot * e->fa.cnt_stride, 4 * (size_t)e->g.bin_num, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  int64_t s = 0;
  for (int32_t v : c) s += v;
  *n = (int32_t)s;
  return VO_OK;
}

// ---------------------------------------------------------------------------------------------
// instrumentation
// ---------------------------------------------------------------------------------------------

vo_status vo_engine_device(const vo_engine* e, int32_t* device) {
  if (!e || !device) return fail(VO_ERR_INVALID, "null argument");
  *device = e->device;
  return VO_OK;
}

vo_status vo_launch_count(const vo_engine* e, int64_t* n) {
  if (!e || !n) return fail(VO_ERR_INVALID, "null argument");
  *n = e->launches;
  return VO_OK;
}

vo_status vo_profile_enable(vo_engine* e, int32_t on) {
  if (!e) return fail(VO_ERR_INVALID, "null engine");
  CU_TRY(cudaStreamSynchronize(e->stream));
  drain_events(e);
  e->profiling = on != 0;
  return VO_OK;
}

vo_status vo_profile_read(vo_engine* e, double ms[4], int64_t calls[4], int32_t reset) {
  if (!e || !ms || !calls) return fail(VO_ERR_INVALID, "null argument");
  CU_TRY(cudaStreamSynchronize(e->stream));
  drain_events(e);
  for (int i = 0; i < 4; i++) { ms[i] = e->prof_ms[i]; calls[i] = e->prof_calls[i]; }
  if (reset) for (int i = 0; i < F_FAMILIES; i++) { e->prof_ms[i] = 0; e->prof_calls[i] = 0; }
  return VO_OK;
}

vo_status vo_profile_read_ext(vo_engine* e, double ms[8], int64_t calls[8], int32_t reset) {
  if (!e || !ms || !calls) return fail(VO_ERR_INVALID, "null argument");
  CU_TRY(cudaStreamSynchronize(e->stream));
  drain_events(e);
  for (int i = 0; i < F_FAMILIES; i++) { ms[i] = e->prof_ms[i]; calls[i] = e->prof_calls[i]; }
  if (reset) for (int i = 0; i < F_FAMILIES; i++) { e->prof_ms[i] = 0; e->prof_calls[i] = 0; }
  return VO_OK;
}

vo_status vo_host_alloc(void** p, int64_t bytes) {
  if (!p || bytes <= 0) return fail(VO_ERR_INVALID, "bad argument");
  CU_TRY(cudaMallocHost(p, (size_t)bytes));
  return VO_OK;
}

vo_status vo_host_free(void* p) {
  if (p) CU_TRY(cudaFreeHost(p));
  return VO_OK;
}

}  // extern "C"
