// K1: fused feature extraction for one bin tile (60x45 pixels in the reference) per CTA (sm_100a).
//
// Replaces, bit-exactly, the reference's
//   mySobel5x5 / myBlob5x5 / myCheckerboard5x5           src/myComputeFeature.cpp:58-151
//   myNonMaximumSuppression_and_ComputeDescriptors       src/myComputeFeature.cpp:153-350
//   myComputeDescriptor                                  src/myComputeFeature.cpp:11-55
// The reference makes four full-frame passes through four full-size planes; here a CTA pulls
// its 74x59 pixel tile (bin + 7 px halo, inside a 16-byte aligned 96x59 TMA box) into shared memory with one
// cp.async.bulk.tensor, derives du/dv/f1/f2 there (separably: identical integers, fewer MACs), runs
// the 3x3-block NMS + 5x5 validity test, and writes the surviving features straight into the
// per-(v_bin,u_bin,class) lists in block-raster order via warp ballots + an ordered CTA scan --
// never atomics, so list order equals the reference's sequential append order.
#include "vo_common.cuh"
#include <cuda.h>

namespace vo {

constexpr int K1_THREADS = 320;
constexpr int K1_WARPS = K1_THREADS / 32;
constexpr int kStripRows = 7;
constexpr uint32_t kHBias = 0x04000400u;                 // +1024 per half: keeps signed horizontal sums non-negative
constexpr uint32_t kFBias = 4096;                        // f1/f2 are stored as value + 4096 (range [16, 8176])

// Tile geometry for a bin of BW_ x BH_ pixels (60 x 45 is the reference's, src/size.h:10-11; the others are the live
// `match_binsize` of SURVEY.md row F4, checked against the reference compiled with those numbers).
template <int BW_, int BH_>
struct K1Tile {
  static constexpr int BW = BW_, BH = BH_;
  // TMA box: (BW+14) x (BH+14) needed; the box must START on a 16-byte boundary in x (u8 elements: measured, unaligned
  // inner coordinates fault), so up to 15 lead-in bytes are added and the width rounded to 16 (60 -> 74 + 15 -> 96)
  static constexpr int TW = (BW + 14 + 15 + 15) & ~15, TH = BH + 14;
  static constexpr int OW = (BW + 10 + 1) & ~1, OH = BH + 10;   // du/dv/f1/f2 region: tile cols 2.., rows 2..; even width
  static constexpr int NBX = BW / 3, NBY = BH / 3;              // NMS blocks per tile (20 x 15)
  // Filter phase work split: a thread owns two adjacent output columns (packed into the two 16-bit halves of a
  // register) and slides down a strip of kStripRows output rows.
  static constexpr int kColPairs = OW / 2;                        // 35
  static constexpr int kStrips = (OH + kStripRows - 1) / kStripRows;  // 8
  static_assert(BW % 3 == 0 && BH % 3 == 0, "bins are whole 3x3 NMS blocks");
  static_assert(kColPairs * kStrips <= K1_THREADS, "filter phase needs one thread per (column pair, strip)");
  static_assert(NBX * NBY <= K1_THREADS, "NMS phase needs one thread per 3x3 block");
  static_assert(TW <= 256 && TH <= 256, "TMA box dimensions");
  static_assert(OW + 19 <= TW, "the box holds the widest lead-in (15) + the last column pair's six bytes");
};

template <class T>
struct __align__(128) K1SmemT {
  uint8_t img[T::TH * T::TW];
  unsigned long long bar;
  uint16_t duv[T::OH * T::OW];  // du | dv<<8
  uint16_t f1[T::OH * T::OW];   // blob response + kFBias
  uint16_t f2[T::OH * T::OW];   // checkerboard response + kFBias
  __align__(16) int32_t warp_tot[K1_WARPS][4];
  uint32_t work[4][T::NBX * T::NBY];  // per class, in list order: ox | oy<<8 | twin rank<<16 | twin class<<30 | has twin<<31
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// Two Sobel sums at once.  x holds acc + 16384 in each 16-bit half (|acc| <= 12240, so every half is in
// [4144, 28624]).  The reference computes acc /= 128 with C truncation toward zero, then + 128 and a clamp
// that can never fire (src/myComputeFeature.cpp:84-106):  trunc(acc/128) + 128 = floor((acc + 16384 +
// (acc < 0 ? 127 : 0)) / 128), and acc < 0  <=>  bit 14 of the biased half is clear.
__device__ __forceinline__ uint32_t sobel_u8x2(uint32_t x) {
  const uint32_t neg = (~x >> 14) & 0x00010001u;
  return ((x + neg * 127u) >> 7) & 0x01FF01FFu;
}

#ifndef VO_K1_MIN_CTAS
#define VO_K1_MIN_CTAS 4
#endif
template <class T>
__global__ void __launch_bounds__(K1_THREADS, VO_K1_MIN_CTAS)
k1_extract(const __grid_constant__ CUtensorMap tmap, Geom g, FrameArrays fa, int first_slot, int z_base) {
  constexpr int BW = T::BW, BH = T::BH, TW = T::TW, TH = T::TH, OW = T::OW, OH = T::OH, NBX = T::NBX, NBY = T::NBY;
  constexpr int kColPairs = T::kColPairs, kStrips = T::kStrips;
  using K1Smem = K1SmemT<T>;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  K1Smem& s = *reinterpret_cast<K1Smem*>(smem_raw);
  const int tid = threadIdx.x;
  const int ub = blockIdx.x, r = blockIdx.y, slot = first_slot + blockIdx.z;
  const int x0 = ub * BW, y0 = r * BH;
  const int xs = (x0 - 7) & ~15;   // 16-byte aligned box origin (floor, also for the negative halo of ub == 0)
  const int xoff = x0 - 7 - xs;    // tile column of image x0-7

  // ---- TMA: one 96x59 box at (xs, y0-7, z); out-of-image bytes arrive as zeros -------------
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s.bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s.bar)),
                 "r"(TW * TH)
                 : "memory");
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(s.img)),
        "l"(&tmap), "r"(xs), "r"(y0 - 7), "r"(z_base + (int)blockIdx.z), "r"(smem_u32(&s.bar))
        : "memory");
  }
  {
    uint32_t done = 0, spins = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(smem_u32(&s.bar))
          : "memory");
      if (!done && ++spins > (1u << 24)) __trap();  // a TMA load that never completes fails the launch instead of hanging the GPU
    }
  }

  // ---- filters: du/dv (Sobel), f1 (blob), f2 (checkerboard) for tile columns 2..71, rows 2..56 ---------
  // Separable, sliding down the strip with the last five horizontal rows in registers; all sums are the
  // same integers as the reference's dense 5x5 correlations (src/myComputeFeature.cpp:58-151), two
  // columns per 32-bit operation.  Halves never borrow from each other because every packed value is
  // biased to be non-negative (kHBias on the horizontal sums, 16384 / kFBias on the vertical ones).
  if (tid < kColPairs * kStrips) {
    const int k = tid % kColPairs, strip = tid / kColPairs;
    const uint8_t* col = &s.img[xoff + 2 * k];  // image column of output column 2k, minus 2
    uint32_t HD[5], HG[5], HC[5], HS[5], HT[5], PC[5];
#pragma unroll
    for (int i = 0; i < kStripRows + 4; i++) {
      const int ty = min(strip * kStripRows + i, TH - 1);  // the clamped row only feeds outputs that are not stored
      const uint8_t* p = col + ty * TW;
      const uint32_t b0 = p[0], b1 = p[1], b2 = p[2], b3 = p[3], b4 = p[4], b5 = p[5];
      const uint32_t P0 = b0 | (b1 << 16), P1 = b1 | (b2 << 16), P2 = b2 | (b3 << 16), P3 = b3 | (b4 << 16), P4 = b4 | (b5 << 16);
      const int w = i % 5;
      const uint32_t S3 = P1 + P2 + P3, S5 = S3 + P0 + P4;
      HD[w] = P0 + 2u * P1 + kHBias - 2u * P3 - P4;        // [1 2 0 -2 -1] + 1024
      HG[w] = P0 + P4 + 4u * (P1 + P3) + 6u * P2;          // [1 4 6 4 1]
      HC[w] = P3 + P4 + kHBias - P0 - P1;                  // [-1 -1 0 1 1] + 1024
      HS[w] = S5;                                           // box 5
      HT[w] = 2u * S3 + kHBias - S5;                        // 2*box3 - box5 + 1024
      PC[w] = P2;
      if (i >= 4) {
        const int oy = strip * kStripRows + i - 4;
        // window rows: r0 = oldest (i-4) ... r4 = newest (i)
        const int r0 = (i + 1) % 5, r1 = (i + 2) % 5, r2 = (i + 3) % 5, r3 = (i + 4) % 5, r4 = w;
        const uint32_t du = sobel_u8x2(HD[r0] + HD[r4] + 4u * (HD[r1] + HD[r3]) + 6u * HD[r2]);  // 16 * 1024 = 16384 of bias
        const uint32_t dv = sobel_u8x2(HG[r0] + 2u * HG[r1] + 0x40004000u - 2u * HG[r3] - HG[r4]);
        const uint32_t f2v = HC[r0] + HC[r1] + (kFBias * 0x10001u) - HC[r3] - HC[r4];
        const uint32_t f1v = HT[r1] + HT[r2] + HT[r3] + 7u * PC[r2] + ((kFBias - 3072u) * 0x10001u) - HS[r0] - HS[r4];
        if (oy < OH) {
          const int idx = oy * OW + 2 * k;
          *reinterpret_cast<uint32_t*>(&s.duv[idx]) = du | (dv << 8);
          *reinterpret_cast<uint32_t*>(&s.f1[idx]) = f1v;
          *reinterpret_cast<uint32_t*>(&s.f2[idx]) = f2v;
        }
      }
    }
  }
  __syncthreads();

  // ---- optional: the du | dv plane of this tile's own pixels goes to global memory (stereo engines: K5 refines sub-pixel
  //      positions on exactly these responses and would otherwise recompute them per match) ------------------------------
  if (fa.duv != nullptr) {
    uint16_t* plane = fa.duv + (int64_t)slot * fa.duv_stride;
    for (int i = tid; i < BW * BH; i += K1_THREADS) {
      const int dy = i / BW, dx = i - dy * BW;  // compile-time divisor
      const int x = x0 + dx, y = y0 + dy;
      if (x < g.W && y < g.H) plane[(int64_t)y * g.pitch + x] = s.duv[(dy + 5) * OW + dx + 5];
    }
    if (ub >= g.U || r > g.j_last / BH) return;  // a tile launched only to cover the image's last rows / columns: no 3x3 block in it
  }

  // ---- NMS: one thread per 3x3 block, block-raster order == thread order -------------------
  const int bx = tid % NBX, by = tid / NBX;
  const int bi = x0 + 3 * bx, bj = y0 + 3 * by;  // block origin (i, j) of the reference loops
  const bool block_ok = (tid < NBX * NBY) && bi >= 9 && bi < g.W - 9 && bj >= 9 && bj < g.H - 9;
  bool flag[4] = {false, false, false, false};
  int pos[4] = {0, 0, 0, 0};  // ox | oy<<8 of the extremum, plane coordinates
  if (block_ok) {
    const int ox0 = 3 * bx + 5, oy0 = 3 * by + 5;
#pragma unroll
    for (int pl = 0; pl < 2; pl++) {
      const uint16_t* f = pl ? s.f2 : s.f1;  // value + kFBias
      // strict < / >: the first extremum in raster order wins (:199-224).  The raster index rides in the low four bits of a
      // key (value * 16 + idx for the minimum, value * 16 + 15 - idx for the maximum: ties resolve to the smallest idx), so
      // the scan is nine multiply-adds and eight min / max per extremum instead of compare-and-select pairs
      int kmn = 0x7FFFFFFF, kmx = -1;
#pragma unroll
      for (int dy = 0; dy < 3; dy++)
#pragma unroll
        for (int dx = 0; dx < 3; dx++) {
          const int v = f[(oy0 + dy) * OW + ox0 + dx];
          kmn = min(kmn, v * 16 + (dy * 3 + dx));
          kmx = max(kmx, v * 16 + (15 - (dy * 3 + dx)));
        }
      const int mn = kmn >> 4, mx = kmx >> 4;
      const int imn = kmn & 15, imx = 15 - (kmx & 15);
      const int ymn = (imn * 11) >> 5, ymx = (imx * 11) >> 5;  // idx / 3 for idx in 0..8
      const int pmn = (ox0 + imn - 3 * ymn) | ((oy0 + ymn) << 8), pmx = (ox0 + imx - 3 * ymx) | ((oy0 + ymx) << 8);
      bool vmn = mn <= (int)kFBias - g.tau, vmx = mx >= (int)kFBias + g.tau;
      // 5x5 validity (:226-260): no neighbour strictly below the minimum / above the maximum, i.e. the extremum of the 5x5
      // window around it (which contains it) equals it
      if (vmn) {
        const uint16_t* w = f + ((pmn >> 8) - 2) * OW + (pmn & 255) - 2;
        int m = 65535;
#pragma unroll
        for (int ky = 0; ky < 5; ky++)
#pragma unroll
          for (int kx = 0; kx < 5; kx++) m = min(m, (int)w[ky * OW + kx]);
        vmn = m >= mn;
      }
      if (vmx) {
        const uint16_t* w = f + ((pmx >> 8) - 2) * OW + (pmx & 255) - 2;
        int m = 0;
#pragma unroll
        for (int ky = 0; ky < 5; ky++)
#pragma unroll
          for (int kx = 0; kx < 5; kx++) m = max(m, (int)w[ky * OW + kx]);
        vmx = m <= mx;
      }
      flag[2 * pl] = vmn; pos[2 * pl] = pmn;
      flag[2 * pl + 1] = vmx; pos[2 * pl + 1] = pmx;
    }
  }

  // ---- ordered compaction: 4 ballots per warp + scan over warps in raster order -----------
  const int lane = tid & 31, warp = tid >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  int rank[4];
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, flag[c]);
    rank[c] = __popc(b & lt);
    if (lane == 0) s.warp_tot[warp][c] = __popc(b);
  }
  __syncthreads();
  // lane w holds warp w's four counts; two warp reductions per class give the list total and this warp's offset
  // (r2t: every thread used to walk all 40 counts, 7.5 % of the kernel's instructions)
  static_assert(K1_WARPS <= 32, "one lane per warp count");
  const int4 wt = (lane < K1_WARPS) ? *reinterpret_cast<const int4*>(s.warp_tot[lane]) : make_int4(0, 0, 0, 0);
  const bool before = lane < warp;
  int total[4];
  total[0] = __reduce_add_sync(0xFFFFFFFFu, wt.x); rank[0] += __reduce_add_sync(0xFFFFFFFFu, before ? wt.x : 0);
  total[1] = __reduce_add_sync(0xFFFFFFFFu, wt.y); rank[1] += __reduce_add_sync(0xFFFFFFFFu, before ? wt.y : 0);
  total[2] = __reduce_add_sync(0xFFFFFFFFu, wt.z); rank[2] += __reduce_add_sync(0xFFFFFFFFu, before ? wt.z : 0);
  total[3] = __reduce_add_sync(0xFFFFFFFFu, wt.w); rank[3] += __reduce_add_sync(0xFFFFFFFFu, before ? wt.w : 0);

  const int bj_first = max(y0, 9);
  if (bj_first >= g.H - 9) return;  // no block row in this tile (grid is sized to avoid it)
  // dest_v_bin(g, bj_first) with the tile's compile-time bin height (g.bin_h == BH: launch_k1 picks the instantiation by it)
  const int vrow = bj_first / BH;
  const int vd = (BH * (vrow + 1) - 3 <= g.j_last) ? vrow : g.v_rem;
  const int bin0 = (vd * g.U + ub) * 4;
  int32_t* cnt = fa.cnt + (int64_t)slot * fa.cnt_stride;
  if (tid < 4) cnt[bin0 + tid] = total[tid];

  // every survivor leaves a one-word note at its list position; the records are then built densely,
  // one thread per feature in list order (no divergence over empty blocks, coalesced record writes)
#pragma unroll
  for (int c = 0; c < 4; c++) {
    if (!flag[c]) continue;
    uint32_t note = (uint32_t)pos[c];
    if (c >= 2) {  // lower-class feature sharing this pixel (dedupe key of the reference's M[][])
      if (flag[0] && pos[0] == pos[c]) note |= 0x80000000u | ((uint32_t)rank[0] << 16);
      else if (flag[1] && pos[1] == pos[c]) note |= 0xC0000000u | ((uint32_t)rank[1] << 16);
    }
    s.work[c][rank[c]] = note;
  }
  __syncthreads();

  uint32_t* xy = fa.xy + (int64_t)slot * fa.rec_stride;
  uint32_t* meta = fa.meta + (int64_t)slot * fa.rec_stride;
  uint4* desc = fa.desc + (int64_t)slot * fa.rec_stride * 2;
  uint2* qsig = fa.q + (int64_t)slot * fa.rec_stride;
  constexpr int sdx[16] = {-3, -3, -1, -1, 3, 3, 1, 1, -1, -1, 1, 1, -5, -5, 5, 5};
  constexpr int sdy[16] = {-1, 1, -1, 1, -1, 1, -1, 1, -5, 5, -5, 5, -3, 3, -3, 3};
  const int e1 = total[0], e2 = e1 + total[1], e3 = e2 + total[2], e4 = e3 + total[3];
  for (int f = tid; f < e4; f += K1_THREADS) {
    const int c = (f >= e1) + (f >= e2) + (f >= e3);
    const int r = f - (c == 0 ? 0 : (c == 1 ? e1 : (c == 2 ? e2 : e3)));
    const uint32_t note = s.work[c][r];
    const int ox = (int)(note & 255u), oy = (int)((note >> 8) & 255u);
    const int64_t rec = (int64_t)(bin0 + c) * g.cap + r;
    xy[rec] = (uint32_t)(x0 - 5 + ox) | ((uint32_t)(y0 - 5 + oy) << 16);
    meta[rec] = (note & 0x80000000u) ? (((note >> 30) & 1u) << 16) | ((note >> 16) & 0x3FFFu) : kNone;
    uint32_t w[8];
    const uint16_t* base = &s.duv[oy * OW + ox];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint32_t lo = base[sdy[2 * k] * OW + sdx[2 * k]];
      const uint32_t hi = base[sdy[2 * k + 1] * OW + sdx[2 * k + 1]];
      w[k] = lo | (hi << 16);
    }
    const uint4 d0 = make_uint4(w[0], w[1], w[2], w[3]), d1 = make_uint4(w[4], w[5], w[6], w[7]);
    desc[rec * 2] = d0;
    desc[rec * 2 + 1] = d1;
    qsig[rec] = desc_q(d0, d1);
  }
}

// Convert a frame slot to the reference's AoS layout (parity dumps, operator-level ABI):
// max2 int32[bin_num][cap_stride][12] = {u, v, 0, class, 8 descriptor words}; src/size.h:17-20,
// src/myComputeFeature.cpp:262-270.  Lists longer than cap_stride are truncated and flagged.
__global__ void k_dump_features(Geom g, FrameArrays fa, int slot, int32_t* __restrict__ max2,
                                int32_t* __restrict__ max2_num, int cap_stride, int32_t* __restrict__ overflow) {
  const int bin = blockIdx.x;
  const int32_t n = fa.cnt[(int64_t)slot * fa.cnt_stride + bin];
  if (threadIdx.x == 0) {
    max2_num[bin] = n;
    if (n > cap_stride) atomicExch(overflow, 1);
  }
  const uint32_t* xy = fa.xy + (int64_t)slot * fa.rec_stride + (int64_t)bin * g.cap;
  const uint4* desc = fa.desc + ((int64_t)slot * fa.rec_stride + (int64_t)bin * g.cap) * 2;
  for (int k = threadIdx.x; k < min(n, cap_stride); k += blockDim.x) {
    int32_t* rec = max2 + ((int64_t)bin * cap_stride + k) * 12;
    const uint32_t p = xy[k];
    const uint4 a = desc[2 * k], b = desc[2 * k + 1];
    rec[0] = (int32_t)(p & 0xFFFFu); rec[1] = (int32_t)(p >> 16); rec[2] = 0; rec[3] = bin & 3;
    rec[4] = (int32_t)a.x; rec[5] = (int32_t)a.y; rec[6] = (int32_t)a.z; rec[7] = (int32_t)a.w;
    rec[8] = (int32_t)b.x; rec[9] = (int32_t)b.y; rec[10] = (int32_t)b.z; rec[11] = (int32_t)b.w;
  }
}

// Inverse of k_dump_features: load reference-layout feature arrays into a frame slot (operator
// level vo_matching).  The twin links are rebuilt from equal pixels inside a bin.
__global__ void k_load_features(Geom g, FrameArrays fa, int slot, const int32_t* __restrict__ max2,
                                const int32_t* __restrict__ max2_num, int cap_stride) {
  const int bin = blockIdx.x;
  const int32_t n = min(max2_num[bin], g.cap);
  if (threadIdx.x == 0) fa.cnt[(int64_t)slot * fa.cnt_stride + bin] = n;
  uint32_t* xy = fa.xy + (int64_t)slot * fa.rec_stride + (int64_t)bin * g.cap;
  uint32_t* meta = fa.meta + (int64_t)slot * fa.rec_stride + (int64_t)bin * g.cap;
  uint4* desc = fa.desc + ((int64_t)slot * fa.rec_stride + (int64_t)bin * g.cap) * 2;
  uint2* qsig = fa.q + (int64_t)slot * fa.rec_stride + (int64_t)bin * g.cap;
  const int c = bin & 3;
  for (int k = threadIdx.x; k < n; k += blockDim.x) {
    const int32_t* rec = max2 + ((int64_t)bin * cap_stride + k) * 12;
    xy[k] = (uint32_t)rec[0] | ((uint32_t)rec[1] << 16);
    const uint4 d0 = make_uint4(rec[4], rec[5], rec[6], rec[7]), d1 = make_uint4(rec[8], rec[9], rec[10], rec[11]);
    desc[2 * k] = d0;
    desc[2 * k + 1] = d1;
    qsig[k] = desc_q(d0, d1);
    uint32_t tw = kNone;
    if (c >= 2) {
      for (int c2 = 0; c2 < 2 && tw == kNone; c2++) {
        const int bin2 = (bin & ~3) | c2;
        const int32_t n2 = min(max2_num[bin2], g.cap);
        for (int k2 = 0; k2 < n2; k2++) {
          const int32_t* r2 = max2 + ((int64_t)bin2 * cap_stride + k2) * 12;
          if (r2[0] == rec[0] && r2[1] == rec[1]) { tw = ((uint32_t)c2 << 16) | (uint32_t)k2; break; }
        }
      }
    }
    meta[k] = tw;
  }
}

// ---- bin sizes this library is built for: the reference's 60 x 45 and the live `match_binsize` values of row F4 ----------
// (a list can hold (BW/3)*(BH/3) records; K2 stages six lists of a window row at once, so 6 * cap must fit its buffer)
#define VO_K1_TILES(X) X(60, 45) X(45, 45) X(60, 30) X(90, 30)

bool k1_supports(int bin_w, int bin_h) {
#define X(w, h) if (bin_w == w && bin_h == h) return true;
  VO_K1_TILES(X)
#undef X
  return false;
}

// TMA box of the tile kernel for a bin size (vo_engine builds the tensor maps)
void k1_box(int bin_w, int bin_h, int* tw, int* th) {
  *tw = 0; *th = 0;
#define X(w, h) if (bin_w == w && bin_h == h) { *tw = K1Tile<w, h>::TW; *th = K1Tile<w, h>::TH; }
  VO_K1_TILES(X)
#undef X
}

// The opt-in to more than 48 KB of dynamic shared memory is a per-DEVICE function attribute: vo_engine_create calls this
// (and its K2 / K4 counterparts) after cudaSetDevice, so a process that owns engines on several GPUs sets it on each.
cudaError_t init_k1_device() {
  cudaError_t e = cudaSuccess;
#define X(w, h) if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_extract<K1Tile<w, h>>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(K1SmemT<K1Tile<w, h>>));
  VO_K1_TILES(X)
#undef X
  return e;
}

cudaError_t launch_k1(const CUtensorMap& tmap, const Geom& g, const FrameArrays& fa, int first_slot, int z_base, int n,
                      cudaStream_t st) {
  int cols = g.U, rows = g.j_last / g.bin_h + 1;  // tile rows that contain at least one block row
  if (fa.duv != nullptr) {  // the response planes must cover every pixel K5 can touch: x <= W-6, y <= H-6
    cols = max(cols, (g.W - 5 + g.bin_w - 1) / g.bin_w);
    rows = max(rows, (g.H - 5 + g.bin_h - 1) / g.bin_h);
  }
  dim3 grid(cols, rows, n);
#define X(w, h)                                                                                                          \
  if (g.bin_w == w && g.bin_h == h) {                                                                                    \
    k1_extract<K1Tile<w, h>><<<grid, K1_THREADS, sizeof(K1SmemT<K1Tile<w, h>>), st>>>(tmap, g, fa, first_slot, z_base);  \
    return cudaGetLastError();                                                                                           \
  }
  VO_K1_TILES(X)
#undef X
  return cudaErrorInvalidValue;
}

cudaError_t launch_dump(const Geom& g, const FrameArrays& fa, int slot, int32_t* max2, int32_t* max2_num,
                        int cap_stride, int32_t* overflow, cudaStream_t st) {
  k_dump_features<<<g.bin_num, 128, 0, st>>>(g, fa, slot, max2, max2_num, cap_stride, overflow);
  return cudaGetLastError();
}

cudaError_t launch_load(const Geom& g, const FrameArrays& fa, int slot, const int32_t* max2, const int32_t* max2_num,
                        int cap_stride, cudaStream_t st) {
  k_load_features<<<g.bin_num, 128, 0, st>>>(g, fa, slot, max2, max2_num, cap_stride);
  return cudaGetLastError();
}

}  // namespace vo
