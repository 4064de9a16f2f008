// RANSAC hypothesis scoring of VisualOdometryMono::estimateMotion on the GPU (sm_100a): the 2000 hypotheses of a frame
// pair are independent -- an 8-point fundamental matrix (src/viso_mono.cpp:242-275) and a Sampson inlier count over the
// bucketed matches (:277-325) each -- and the reference's sequential loop keeps the FIRST hypothesis with the largest
// count (src/viso_mono.cpp:63-77).  One thread per hypothesis runs vo_ransac_core.h, the very code the host stage runs
// (same double operations in the same order, no FMA contraction on either side), a second kernel takes the arg-max with
// the first-wins rule.  The sample sets are still drawn on the host, in frame order, from the sequence's glibc-identical
// stream (vo_egomotion_draw_samples); everything after the winner is known (refit on the inliers, essential matrix,
// chirality, scale) stays in vo_egomotion.cpp.  SURVEY.md row F2, config 3 (the process() loop).
#include <cuda_runtime.h>

#include <cstring>
#include <string>
#include <vector>

#include "../../include/vo_b200.h"
#include "vo_ransac_core.h"

extern "C" int32_t vo_egomotion_prepare_points(const vo_p_match* matches, int32_t n, float* out4n);  // vo_egomotion.cpp

namespace {

constexpr int kScoreThreads = 128;

// grid (ceil(max_iters / kScoreThreads), lists); dynamic shared memory = max_points * sizeof(Pt)
__global__ void __launch_bounds__(kScoreThreads)
k_ransac_score(const vo_ransac::Pt* __restrict__ pts, int max_points, const uint16_t* __restrict__ samples, int max_iters,
               const int32_t* __restrict__ n_pts, const int32_t* __restrict__ n_iters, double thr, int32_t* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  vo_ransac::Pt* p = reinterpret_cast<vo_ransac::Pt*>(smem_raw);
  const int list = blockIdx.y, n = n_pts[list], iters = n_iters[list];
  if ((int)(blockIdx.x * kScoreThreads) >= iters) return;
  const float4* src = reinterpret_cast<const float4*>(pts + (int64_t)list * max_points);
  for (int i = threadIdx.x; i < n; i += kScoreThreads) reinterpret_cast<float4*>(p)[i] = src[i];
  __syncthreads();
  const int k = blockIdx.x * kScoreThreads + threadIdx.x;
  if (k >= iters) return;
  const uint4 s16 = *reinterpret_cast<const uint4*>(samples + ((int64_t)list * max_iters + k) * 8);
  const int32_t smp[8] = {(int32_t)(s16.x & 0xFFFFu), (int32_t)(s16.x >> 16), (int32_t)(s16.y & 0xFFFFu), (int32_t)(s16.y >> 16),
                          (int32_t)(s16.z & 0xFFFFu), (int32_t)(s16.z >> 16), (int32_t)(s16.w & 0xFFFFu), (int32_t)(s16.w >> 16)};
  counts[(int64_t)list * max_iters + k] = vo_ransac::score_hypothesis(p, n, smp, thr);
}

// one CTA per list: the first hypothesis with the largest (non-zero) inlier count, -1 if there is none
__global__ void __launch_bounds__(256)
k_ransac_best(const int32_t* __restrict__ counts, int max_iters, const int32_t* __restrict__ n_iters, int32_t* __restrict__ best) {
  __shared__ unsigned long long red[256 / 32];
  const int list = blockIdx.x, iters = n_iters[list];
  unsigned long long key = 0;  // count << 32 | ~k : the maximum is the largest count, then the smallest k
  for (int k = threadIdx.x; k < iters; k += 256) {
    const unsigned long long c = (unsigned long long)(uint32_t)counts[(int64_t)list * max_iters + k];
    const unsigned long long cand = (c << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)k);
    key = cand > key ? cand : key;
  }
  for (int d = 16; d > 0; d >>= 1) {
    const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, key, d);
    key = o > key ? o : key;
  }
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = key;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 256 / 32; w++) key = red[w] > key ? red[w] : key;
    const uint32_t c = (uint32_t)(key >> 32);
    best[2 * list] = (iters > 0 && c > 0) ? (int32_t)(0xFFFFFFFFu - (uint32_t)key) : -1;
    best[2 * list + 1] = (int32_t)c;
  }
}

}  // namespace

struct vo_ransac_gpu {
  int device = 0;
  cudaStream_t stream = nullptr;
  int max_lists = 0, max_points = 0, max_iters = 0;
  vo_ransac::Pt *h_pts = nullptr, *d_pts = nullptr;
  uint16_t *h_smp = nullptr, *d_smp = nullptr;
  int32_t *h_n = nullptr, *d_n = nullptr;       // [0, max_lists): points per list, [max_lists, 2 max_lists): iterations
  int32_t *h_best = nullptr, *d_best = nullptr;  // {best_k, count} per list
  int32_t* d_counts = nullptr;
  long long launches = 0;
};

extern "C" {

void vo_ransac_gpu_destroy(vo_ransac_gpu* r) {
  if (!r) return;
  cudaSetDevice(r->device);
  if (r->stream) { cudaStreamSynchronize(r->stream); cudaStreamDestroy(r->stream); }
  cudaFreeHost(r->h_pts); cudaFreeHost(r->h_smp); cudaFreeHost(r->h_n); cudaFreeHost(r->h_best);
  cudaFree(r->d_pts); cudaFree(r->d_smp); cudaFree(r->d_n); cudaFree(r->d_best); cudaFree(r->d_counts);
  delete r;
}

vo_status vo_ransac_gpu_create(int32_t device, int32_t max_lists, int32_t max_points, int32_t max_iters, vo_ransac_gpu** out) {
  if (!out || max_lists < 1 || max_points < 8 || max_points > 65535 || max_iters < 1) return VO_ERR_INVALID;
  *out = nullptr;
  if (cudaSetDevice(device) != cudaSuccess) return VO_ERR_CUDA;
  vo_ransac_gpu* r = new vo_ransac_gpu();
  r->device = device; r->max_lists = max_lists; r->max_points = max_points; r->max_iters = max_iters;
  const size_t pts = sizeof(vo_ransac::Pt) * (size_t)max_lists * max_points, smp = sizeof(uint16_t) * 8 * (size_t)max_lists * max_iters;
  bool ok = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaHostAlloc((void**)&r->h_pts, pts, cudaHostAllocDefault) == cudaSuccess && cudaMalloc((void**)&r->d_pts, pts) == cudaSuccess;
  ok = ok && cudaHostAlloc((void**)&r->h_smp, smp, cudaHostAllocDefault) == cudaSuccess && cudaMalloc((void**)&r->d_smp, smp) == cudaSuccess;
  ok = ok && cudaHostAlloc((void**)&r->h_n, sizeof(int32_t) * 2 * max_lists, cudaHostAllocDefault) == cudaSuccess &&
       cudaMalloc((void**)&r->d_n, sizeof(int32_t) * 2 * max_lists) == cudaSuccess;
  ok = ok && cudaHostAlloc((void**)&r->h_best, sizeof(int32_t) * 2 * max_lists, cudaHostAllocDefault) == cudaSuccess &&
       cudaMalloc((void**)&r->d_best, sizeof(int32_t) * 2 * max_lists) == cudaSuccess;
  ok = ok && cudaMalloc((void**)&r->d_counts, sizeof(int32_t) * (size_t)max_lists * max_iters) == cudaSuccess;
  if (ok && sizeof(vo_ransac::Pt) * (size_t)max_points > 48 * 1024)
    ok = cudaFuncSetAttribute(k_ransac_score, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(vo_ransac::Pt) * max_points)) == cudaSuccess;
  if (!ok) { vo_ransac_gpu_destroy(r); return VO_ERR_CUDA; }
  *out = r;
  return VO_OK;
}

// best_k[i] = index of the first hypothesis of list i with the largest non-zero inlier count, -1 if there is none or the
// list never reaches RANSAC (fewer than 10 matches, degenerate normalisation: drawn[i] == 0); best_count[i] (optional) = its
// inlier count.  lists + i * stride: counts[i] matches; samples + i * samples_stride * 8: drawn[i] sample sets.
vo_status vo_ransac_gpu_score(vo_ransac_gpu* r, int32_t n_lists, const vo_p_match* lists, int64_t stride, const int32_t* counts,
                              const vo_mono_params* prm, const int32_t* samples, int64_t samples_stride, const int32_t* drawn,
                              int32_t* best_k, int32_t* best_count) {
  if (!r || !lists || !counts || !prm || !samples || !drawn || !best_k || n_lists < 1 || n_lists > r->max_lists) return VO_ERR_INVALID;
  if (cudaSetDevice(r->device) != cudaSuccess) return VO_ERR_CUDA;
  int most_iters = 0, most_pts = 0;
  for (int i = 0; i < n_lists; i++) {
    const int32_t n = counts[i];
    int32_t iters = drawn[i];
    if (n > r->max_points || iters > r->max_iters || n < 0 || iters < 0) return VO_ERR_CAPACITY;
    if (iters > 0 && !vo_egomotion_prepare_points(lists + (int64_t)i * stride, n, reinterpret_cast<float*>(r->h_pts + (size_t)i * r->max_points)))
      iters = 0;  // (cannot happen: the samples were drawn only because the list reaches RANSAC)
    r->h_n[i] = n;
    r->h_n[r->max_lists + i] = iters;
    uint16_t* dst = r->h_smp + (size_t)i * r->max_iters * 8;
    const int32_t* src = samples + (size_t)i * samples_stride * 8;
    for (int k = 0; k < iters * 8; k++) dst[k] = (uint16_t)src[k];
    most_iters = iters > most_iters ? iters : most_iters;
    most_pts = (iters > 0 && n > most_pts) ? n : most_pts;
  }
  bool ok = true;
  if (most_iters > 0) {
    const size_t pts = sizeof(vo_ransac::Pt) * (size_t)n_lists * r->max_points, smp = sizeof(uint16_t) * 8 * (size_t)n_lists * r->max_iters;
    ok = ok && cudaMemcpyAsync(r->d_pts, r->h_pts, pts, cudaMemcpyHostToDevice, r->stream) == cudaSuccess;
    ok = ok && cudaMemcpyAsync(r->d_smp, r->h_smp, smp, cudaMemcpyHostToDevice, r->stream) == cudaSuccess;
    ok = ok && cudaMemcpyAsync(r->d_n, r->h_n, sizeof(int32_t) * 2 * r->max_lists, cudaMemcpyHostToDevice, r->stream) == cudaSuccess;
    if (ok) {
      dim3 grid((most_iters + kScoreThreads - 1) / kScoreThreads, n_lists);
      k_ransac_score<<<grid, kScoreThreads, sizeof(vo_ransac::Pt) * (size_t)r->max_points, r->stream>>>(
          r->d_pts, r->max_points, r->d_smp, r->max_iters, r->d_n, r->d_n + r->max_lists, prm->inlier_threshold, r->d_counts);
      k_ransac_best<<<n_lists, 256, 0, r->stream>>>(r->d_counts, r->max_iters, r->d_n + r->max_lists, r->d_best);
      r->launches += 2;
      ok = cudaGetLastError() == cudaSuccess;
    }
    ok = ok && cudaMemcpyAsync(r->h_best, r->d_best, sizeof(int32_t) * 2 * n_lists, cudaMemcpyDeviceToHost, r->stream) == cudaSuccess;
    ok = ok && cudaStreamSynchronize(r->stream) == cudaSuccess;
    if (!ok) return VO_ERR_CUDA;
  }
  for (int i = 0; i < n_lists; i++) {
    const bool ran = most_iters > 0 && r->h_n[r->max_lists + i] > 0;
    best_k[i] = ran ? r->h_best[2 * i] : -1;
    if (best_count) best_count[i] = ran ? r->h_best[2 * i + 1] : 0;
  }
  return VO_OK;
}

long long vo_ransac_gpu_launches(const vo_ransac_gpu* r) { return r ? r->launches : 0; }

}  // extern "C"
