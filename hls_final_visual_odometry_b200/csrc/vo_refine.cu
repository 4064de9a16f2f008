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
// per target (1p, and for quad lists 2c and 2p) the warp pulls the patch into shared memory as aligned words, 27 lanes
// evaluate du/dv on the 9x9 pixels the 25 shifted descriptors touch (column x 3-row segment each, separable, in
// registers), lanes 0..24 each take one shift's SAD (one VABSDIFF4 per byte against the anchor descriptor held in
// registers), a warp min picks the first minimum in raster order, and the fit is exact double arithmetic.  k5_compact then drops rejected matches, order preserved.
#include "vo_common.cuh"
#include <cstdlib>

namespace vo {

constexpr int K5_WARPS = 8;
constexpr int kPatch = 13;   // 2*(2 shift + 2 sample + 2 sobel) + 1
constexpr int kResp = 9;     // 2*(2 shift + 2 sample) + 1

struct K5Warp {
  __align__(16) uint8_t patch[kPatch * 16];  // row stride 16 bytes; a row starts at the 4-byte boundary below its first pixel
  uint8_t du[kResp * kResp + 3];
  uint8_t dv[kResp * kResp + 3];
};

// computeSmallDescriptor sample offsets (src/matcher.cpp:527-542): 12 du then 4 dv
__constant__ int8_t c_small_dx[16] = {0, -2, 0, 2, -1, 0, 0, 1, -2, 0, 2, 0, 0, -1, 1, 0};
__constant__ int8_t c_small_dy[16] = {-2, -1, -1, -1, 0, 0, 0, 0, 1, 1, 1, 2, -1, 0, 0, 1};

// acc/128 (C truncation toward zero) + 128; the reference's clamp to [0,255] cannot fire (|acc| <= 12240):
// src/myComputeFeature.cpp:84-106
__device__ __forceinline__ uint8_t sobel_u8(int acc) { return (uint8_t)(((acc + ((acc >> 31) & 127)) >> 7) + 128); }

// du/dv on the (2R+1)^2 pixels around (cx, cy) of one frame; R = 2 (anchor) or 4 (target).
// Patch rows arrive as aligned 32-bit words (<= 4 per row).  A lane owns one output column and a segment of three
// output rows: it walks the 7 patch rows under them once, keeping the horizontal 5-tap sums of the last five rows in
// registers (the separable form of the reference's dense 5x5 sums -- same integers).
template <int R>
__device__ __forceinline__ void responses(K5Warp& w, const uint8_t* __restrict__ img, int pitch, int cx, int cy, int lane) {
  constexpr int P = 2 * R + 5, S = 2 * R + 1, NSEG = (S + 2) / 3, WPR = (P + 3 + 3) / 4;  // words per row
  const int x0 = cx - R - 2, off = x0 & 3;
  const uint8_t* src = img + (int64_t)(cy - R - 2) * pitch + (x0 - off);
  __syncwarp();
  for (int i = lane; i < P * WPR; i += 32) {
    const int r = i / WPR, j = i - r * WPR;
    *reinterpret_cast<uint32_t*>(&w.patch[r * 16 + 4 * j]) = *reinterpret_cast<const uint32_t*>(src + (int64_t)r * pitch + 4 * j);
  }
  __syncwarp();
  if (lane < S * NSEG) {
    const int seg = lane / S, c = lane - seg * S, r0 = 3 * seg;
    int h[7], sm[7];  // per patch row: derivative [1 2 0 -2 -1] and smoothing [1 4 6 4 1] along x
#pragma unroll
    for (int k = 0; k < 7; k++) {
      const uint8_t* p = &w.patch[min(r0 + k, P - 1) * 16 + off + c];
      const int a = p[0], b = p[1], m = p[2], d = p[3], e = p[4];
      h[k] = a + 2 * b - 2 * d - e;
      sm[k] = a + 4 * b + 6 * m + 4 * d + e;
    }
#pragma unroll
    for (int k = 0; k < 3; k++)
      if (r0 + k < S) {
        w.du[(r0 + k) * S + c] = sobel_u8(h[k] + 4 * h[k + 1] + 6 * h[k + 2] + 4 * h[k + 3] + h[k + 4]);  // v_weight, src/myComputeFeature.cpp:59-65
        w.dv[(r0 + k) * S + c] = sobel_u8(sm[k] + 2 * sm[k + 1] - 2 * sm[k + 3] - sm[k + 4]);             // h_weight, src/myComputeFeature.cpp:66-72
      }
  }
  __syncwarp();
}

// The same responses read back from the du | dv plane K1 left in global memory (stereo engines, FrameArrays::duv): identical
// bytes (K1 computes the reference's Sobel on the whole frame), none of the patch loads and filter arithmetic above.
template <int R>
__device__ __forceinline__ void responses_from_plane(K5Warp& w, const uint16_t* __restrict__ plane, int pitch, int cx, int cy, int lane) {
  constexpr int S = 2 * R + 1;
  const uint16_t* src = plane + (int64_t)(cy - R) * pitch + (cx - R);
  __syncwarp();
#pragma unroll
  for (int i0 = 0; i0 < S * S; i0 += 32) {
    const int i = i0 + lane;
    if (i < S * S) {
      const int r = i / S, c = i - r * S;
      const uint32_t v = __ldg(&src[(int64_t)r * pitch + c]);
      w.du[i] = (uint8_t)(v & 255u);
      w.dv[i] = (uint8_t)(v >> 8);
    }
  }
  __syncwarp();
}

// |a - b| + acc on zero-extended bytes: one VABSDIFF4
__device__ __forceinline__ uint32_t absdiff_acc(uint32_t a, uint32_t b, uint32_t acc) {
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
  return d;
}

// one parabolicFitting call: anchor descriptor in d1[] (registers), target frame `img`, (u2, v2) updated
template <bool PLANES>
__device__ __forceinline__ bool fit_target(K5Warp& w, const Geom& g, const FrameArrays& fa, int slot, const uint32_t (&d1)[16],
                                           float& u2, float& v2, int lane) {
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  const float margin = 7.f;  // src/matcher.cpp:38
  if (u2 - 2.f < margin || u2 + 2.f > (float)(g.W - 1) - margin || v2 - 2.f < margin || v2 + 2.f > (float)(g.H - 1) - margin)
    return false;
  if constexpr (PLANES) responses_from_plane<4>(w, fa.duv + (int64_t)slot * fa.duv_stride, g.pitch, (int)u2, (int)v2, lane);
  else responses<4>(w, fa.img + (int64_t)slot * fa.img_stride, g.pitch, (int)u2, (int)v2, lane);
  int cost = 0x7FFFFF;
  if (lane < 25) {
    const int a = lane % 5, b = lane / 5;
    const uint8_t* pu = &w.du[(2 + b) * kResp + (2 + a)];
    const uint8_t* pv = &w.dv[(2 + b) * kResp + (2 + a)];
    uint32_t c = 0;
#pragma unroll
    for (int s = 0; s < 16; s++) {
      const int idx = c_small_dy[s] * kResp + c_small_dx[s];
      c = absdiff_acc(d1[s], (s < 12) ? pu[idx] : pv[idx], c);
    }
    cost = (int)c;
  }
  const int best = __reduce_min_sync(FULL, cost * 32 + lane);  // first minimum in raster order
  const int mi = best & 31, a0 = mi % 5, b0 = mi / 5;
  if (a0 == 0 || a0 == 4 || b0 == 0 || b0 == 4) return false;
  // every intermediate below is an integer of magnitude < 2^53, so double arithmetic is exact (see the oracle for the bounds)
  int P = 0, Q = 0, R = 0, B2 = 0, B3 = 0, B4 = 0;
#pragma unroll
  for (int i = -1; i <= 1; i++)
#pragma unroll
    for (int j = -1; j <= 1; j++) {
      const int c = __shfl_sync(FULL, cost, (b0 + i) * 5 + (a0 + j));
      P += j * j * c; Q += i * i * c; R += c; B2 += i * j * c; B3 += j * c; B4 += i * c;
    }
  if (B2 == 0) return false;
  const double b0d = (double)(3 * P - 2 * R), b1d = (double)(3 * Q - 2 * R), b2d = (double)B2, b3d = (double)B3, b4d = (double)B4;
  const double D = 9.0 * b2d * b2d - 16.0 * b0d * b1d, NV = 8.0 * b0d * b4d - 6.0 * b2d * b3d;
  if (D == 0.0) return false;
  const float ddv = (float)__ddiv_rn(NV, D);
  const float ddu = (float)__ddiv_rn(-2.0 * (b4d * D + 2.0 * b1d * NV), 3.0 * b2d * D);
  if (fabsf(ddu) >= 1.0f || fabsf(ddv) >= 1.0f) return false;
  u2 = (float)__dadd_rn((double)u2, __dadd_rn((double)(a0 - 2), (double)ddu));
  v2 = (float)__dadd_rn((double)v2, __dadd_rn((double)(b0 - 2), (double)ddv));
  return true;
}

// quads[item] = {1p, 2p, 1c, 2c} slots (2p/2c = -1 for flow lists)
#ifndef VO_K5_PLANES_CTAS
#define VO_K5_PLANES_CTAS 4
#endif
template <bool PLANES>
__global__ void __launch_bounds__(K5_WARPS * 32, PLANES ? VO_K5_PLANES_CTAS : 4)
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
      if constexpr (PLANES) responses_from_plane<2>(w, fa.duv + (int64_t)qd.z * fa.duv_stride, g.pitch, cx, cy, lane);
      else responses<2>(w, fa.img + (int64_t)qd.z * fa.img_stride, g.pitch, cx, cy, lane);
      uint32_t d1[16];  // the anchor's small descriptor, zero-extended bytes (every lane holds all 16)
#pragma unroll
      for (int s = 0; s < 16; s++) {
        const int idx = (2 + c_small_dy[s]) * 5 + (2 + c_small_dx[s]);
        d1[s] = (s < 12) ? w.du[idx] : w.dv[idx];
      }
      ok = fit_target<PLANES>(w, g, fa, qd.x, d1, u1p, v1p, lane);
      if (ok && u2c >= 0.f && qd.y >= 0 && qd.w >= 0)
        ok = fit_target<PLANES>(w, g, fa, qd.w, d1, u2c, v2c, lane) && fit_target<PLANES>(w, g, fa, qd.y, d1, u2p, v2p, lane);
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
  const dim3 grid(min((max_n + K5_WARPS - 1) / K5_WARPS, 592), n_items);
  if (fa.duv != nullptr && !getenv("VO_REFINE_RECOMPUTE"))  // (developer A/B switch: recompute the responses as a mono engine does)
    k5_refine<true><<<grid, K5_WARPS * 32, 0, st>>>(g, fa, quads, (float*)lists, stride_matches, n_list, keep);
  else
    k5_refine<false><<<grid, K5_WARPS * 32, 0, st>>>(g, fa, quads, (float*)lists, stride_matches, n_list, keep);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  k5_compact<<<n_items, K5C_THREADS, 0, st>>>((uint4*)lists, stride_matches, n_list, keep);
  return cudaGetLastError();
}

}  // namespace vo
