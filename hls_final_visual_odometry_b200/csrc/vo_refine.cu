// K5: sub-pixel refinement of a match list (sm_100a).
//
// EXTENSION -- PARITY UNPINNED.  The reference names the stage (`refinement`, src/matcher.h:56:
// "0=none,1=pixel,2=subpixel") but never executes one; this is libviso2's published parabolicFitting on the 16-byte
// "small" descriptor the reference does define (computeSmallDescriptor, src/matcher.cpp:516-543; margin 7,
// src/matcher.cpp:38), over the Sobel responses of the reference's executed filter (mySobel5x5,
// src/myComputeFeature.cpp:58-109).  Oracle: oracle/vo_oracle.c parabolic_fit / vo_oracle_refine (same arithmetic:
// integer costs, the least-squares quadric through the 3x3 costs around the minimum in closed form, one IEEE
// double division per offset), so the two agree bit for bit.
//
// One WARP per match.  The frames are still resident in their slots, so nothing but 13x13-byte patches is read:
// per target (1p, and for quad lists 2c and 2p) the warp pulls the patch into shared memory, evaluates du/dv on
// the 9x9 pixels the 25 shifted descriptors touch, lanes 0..24 each take one shift's SAD, a warp min picks the first
// minimum in raster order, and lane 0 solves the fit.  k5_compact then drops rejected matches, order preserved.
#include "vo_common.cuh"

namespace vo {

constexpr int K5_WARPS = 8;
constexpr int kPatch = 13;   // 2*(2 shift + 2 sample + 2 sobel) + 1
constexpr int kResp = 9;     // 2*(2 shift + 2 sample) + 1

struct K5Warp {
  uint8_t patch[kPatch * 16];
  uint8_t du[kResp * kResp + 3];
  uint8_t dv[kResp * kResp + 3];
  uint8_t d1[16];
};

// computeSmallDescriptor sample offsets (src/matcher.cpp:527-542): 12 du then 4 dv
__constant__ int8_t c_small_dx[16] = {0, -2, 0, 2, -1, 0, 0, 1, -2, 0, 2, 0, 0, -1, 1, 0};
__constant__ int8_t c_small_dy[16] = {-2, -1, -1, -1, 0, 0, 0, 0, 1, 1, 1, 2, -1, 0, 0, 1};

__device__ __forceinline__ uint8_t sobel_u8(int acc) {  // src/myComputeFeature.cpp:84-106
  const int q = acc / 128 + 128;
  return (uint8_t)min(max(q, 0), 255);
}

// du/dv on the (2R+1)^2 pixels around (cx, cy) of one frame; R = 2 (anchor) or 4 (target)
__device__ __forceinline__ void responses(K5Warp& w, const uint8_t* __restrict__ img, int pitch, int cx, int cy, int R,
                                          int lane) {
  const int P = 2 * R + 5, S = 2 * R + 1;
  __syncwarp();
  for (int i = lane; i < P * P; i += 32) {
    const int r = i / P, c = i - r * P;
    w.patch[r * 16 + c] = img[(int64_t)(cy - R - 2 + r) * pitch + (cx - R - 2 + c)];
  }
  __syncwarp();
  for (int i = lane; i < S * S; i += 32) {
    const int r = i / S, c = i - r * S;
    int h[5], s[5];  // per patch row: derivative [1 2 0 -2 -1] and smoothing [1 4 6 4 1] along x
#pragma unroll
    for (int k = 0; k < 5; k++) {
      const uint8_t* p = &w.patch[(r + k) * 16 + c];
      const int a = p[0], b = p[1], m = p[2], d = p[3], e = p[4];
      h[k] = a + 2 * b - 2 * d - e;
      s[k] = a + 4 * b + 6 * m + 4 * d + e;
    }
    w.du[i] = sobel_u8(h[0] + 4 * h[1] + 6 * h[2] + 4 * h[3] + h[4]);   // v_weight, src/myComputeFeature.cpp:59-65
    w.dv[i] = sobel_u8(s[0] + 2 * s[1] - 2 * s[3] - s[4]);              // h_weight, src/myComputeFeature.cpp:66-72
  }
  __syncwarp();
}

// one parabolicFitting call: anchor descriptor in w.d1, target frame `img`, (u2, v2) updated by lane 0's result
__device__ __forceinline__ bool fit_target(K5Warp& w, const Geom& g, const uint8_t* __restrict__ img, float& u2, float& v2,
                                           int lane) {
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  const float margin = 7.f;  // src/matcher.cpp:38
  if (u2 - 2.f < margin || u2 + 2.f > (float)(g.W - 1) - margin || v2 - 2.f < margin || v2 + 2.f > (float)(g.H - 1) - margin)
    return false;
  responses(w, img, g.pitch, (int)u2, (int)v2, 4, lane);
  int cost = 0x7FFFFF;
  if (lane < 25) {
    const int a = lane % 5, b = lane / 5;
    cost = 0;
#pragma unroll
    for (int s = 0; s < 16; s++) {
      const int idx = (4 + b - 2 + c_small_dy[s]) * kResp + (4 + a - 2 + c_small_dx[s]);
      const int val = (s < 12) ? w.du[idx] : w.dv[idx];
      cost += abs((int)w.d1[s] - val);
    }
  }
  const int best = __reduce_min_sync(FULL, cost * 32 + lane);  // first minimum in raster order
  const int mi = best & 31, a0 = mi % 5, b0 = mi / 5;
  if (a0 == 0 || a0 == 4 || b0 == 0 || b0 == 4) return false;
  long long P = 0, Q = 0, R = 0, B2 = 0, B3 = 0, B4 = 0;
#pragma unroll
  for (int i = -1; i <= 1; i++)
#pragma unroll
    for (int j = -1; j <= 1; j++) {
      const long long c = __shfl_sync(FULL, cost, (b0 + i) * 5 + (a0 + j));
      P += j * j * c; Q += i * i * c; R += c; B2 += i * j * c; B3 += j * c; B4 += i * c;
    }
  const long long B0 = 3 * P - 2 * R, B1 = 3 * Q - 2 * R;
  const long long D = 9 * B2 * B2 - 16 * B0 * B1, NV = 8 * B0 * B4 - 6 * B2 * B3;
  if (D == 0 || B2 == 0) return false;
  const float ddv = (float)__ddiv_rn((double)NV, (double)D);
  const float ddu = (float)__ddiv_rn((double)(-2 * (B4 * D + 2 * B1 * NV)), (double)(3 * B2 * D));
  if (fabsf(ddu) >= 1.0f || fabsf(ddv) >= 1.0f) return false;
  u2 = (float)__dadd_rn((double)u2, __dadd_rn((double)(a0 - 2), (double)ddu));
  v2 = (float)__dadd_rn((double)v2, __dadd_rn((double)(b0 - 2), (double)ddv));
  return true;
}

// quads[item] = {1p, 2p, 1c, 2c} slots (2p/2c = -1 for flow lists)
__global__ void __launch_bounds__(K5_WARPS * 32)
k5_refine(Geom g, FrameArrays fa, const int4* __restrict__ quads, float* __restrict__ lists, int64_t stride_matches,
          const int32_t* __restrict__ n_list, uint8_t* __restrict__ keep) {
  __shared__ K5Warp sw[K5_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int item = blockIdx.y;
  K5Warp& w = sw[warp];
  const int4 qd = quads[item];
  const int n = n_list[item];
  for (int i = blockIdx.x * K5_WARPS + warp; i < n; i += gridDim.x * K5_WARPS) {
    float* rec = lists + ((int64_t)item * stride_matches + i) * 12;
    float u1p = rec[0], v1p = rec[1], u2p = rec[3], v2p = rec[4];
    const float u1c = rec[6], v1c = rec[7];
    float u2c = rec[9], v2c = rec[10];
    const int cx = (int)u1c, cy = (int)v1c;
    bool ok = cx >= 4 && cy >= 4 && cx <= g.W - 5 && cy <= g.H - 5;  // the anchor's descriptor must lie inside the image
    if (ok) {
      responses(w, fa.img + (int64_t)qd.z * fa.img_stride, g.pitch, cx, cy, 2, lane);
      if (lane < 16) {
        const int idx = (2 + c_small_dy[lane]) * 5 + (2 + c_small_dx[lane]);
        w.d1[lane] = (lane < 12) ? w.du[idx] : w.dv[idx];
      }
      __syncwarp();
      ok = fit_target(w, g, fa.img + (int64_t)qd.x * fa.img_stride, u1p, v1p, lane);
      if (ok && u2c >= 0.f && qd.y >= 0 && qd.w >= 0)
        ok = fit_target(w, g, fa.img + (int64_t)qd.w * fa.img_stride, u2c, v2c, lane) &&
             fit_target(w, g, fa.img + (int64_t)qd.y * fa.img_stride, u2p, v2p, lane);
    }
    __syncwarp();
    if (lane == 0) {
      keep[(int64_t)item * stride_matches + i] = ok ? 1 : 0;
      if (ok) { rec[0] = u1p; rec[1] = v1p; rec[3] = u2p; rec[4] = v2p; rec[9] = u2c; rec[10] = v2c; }
    }
  }
}

constexpr int K5C_THREADS = 1024;

// ordered in-place compaction of every list by its keep flags (one CTA per list; a chunk's records are read
// into registers before any of them is written, and a record never moves up)
__global__ void __launch_bounds__(K5C_THREADS)
k5_compact(uint4* __restrict__ lists, int64_t stride_matches, int32_t* __restrict__ n_list, const uint8_t* __restrict__ keep) {
  __shared__ int32_t warp_tot[K5C_THREADS / 32];
  const int item = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = n_list[item];
  uint4* L = lists + (int64_t)item * stride_matches * 3;
  const uint8_t* kp = keep + (int64_t)item * stride_matches;
  int running = 0;
  for (int base = 0; base < n; base += K5C_THREADS) {
    const int i = base + tid;
    const bool k = (i < n) && kp[i];
    uint4 r0, r1, r2;
    if (k) { r0 = L[(int64_t)i * 3]; r1 = L[(int64_t)i * 3 + 1]; r2 = L[(int64_t)i * 3 + 2]; }
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, k);
    if (lane == 0) warp_tot[warp] = __popc(b);
    __syncthreads();
    int pos = running + __popc(b & ((1u << lane) - 1u)), tot = 0;
#pragma unroll
    for (int x = 0; x < K5C_THREADS / 32; x++) {
      const int t = warp_tot[x];
      if (x < warp) pos += t;
      tot += t;
    }
    if (k) { L[(int64_t)pos * 3] = r0; L[(int64_t)pos * 3 + 1] = r1; L[(int64_t)pos * 3 + 2] = r2; }
    running += tot;
    __syncthreads();
  }
  if (tid == 0) n_list[item] = running;
}

cudaError_t launch_k5(const Geom& g, const FrameArrays& fa, const int4* quads, int n_items, void* lists,
                      int64_t stride_matches, int32_t max_n, int32_t* n_list, uint8_t* keep, cudaStream_t st) {
  if (max_n < 1) return cudaSuccess;
  k5_refine<<<dim3(min((max_n + K5_WARPS - 1) / K5_WARPS, 592), n_items), K5_WARPS * 32, 0, st>>>(g, fa, quads, (float*)lists,
                                                                                         stride_matches, n_list, keep);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  k5_compact<<<n_items, K5C_THREADS, 0, st>>>((uint4*)lists, stride_matches, n_list, keep);
  return cudaGetLastError();
}

}  // namespace vo
