// Shared device-side definitions of libvo_b200 (sm_100a only).
//
// HBM layout of one frame slot (all arrays are engine-owned, slot-major):
//   img   u8  [H][pitch]                 pitch = roundup(W,16)  (TMA needs 16-byte row strides)
//   cnt   i32 [bin_num]                  bin = (v_bin*U + u_bin)*4 + class     == max2_num
//   xy    u32 [bin_num][cap]             u | v<<16
//   meta  u32 [bin_num][cap]             "twin": lower-class feature on the same pixel
//                                        (class<<16 | index in its list) or 0xFFFFFFFF
//   desc  uint4[bin_num][cap][2]         the 32 descriptor bytes in reference byte order
//   q     uint2[bin_num][cap]            2 x 4 coded group sums of the descriptor (SAD lower bound, K2)
//   duv   u16 [H][pitch]                 stereo engines only: Sobel responses du | dv << 8 per pixel (K1 -> K5)
// This is the reference's max2/max2_num (src/size.h:17-20) as padded SoA; vo_dump_features
// converts back to the 12 x int32 AoS records for parity.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace vo {

constexpr uint32_t kNone = 0xFFFFFFFFu;

struct Geom {
  int32_t W, H, pitch;
  int32_t bin_w, bin_h;
  int32_t U, V;
  int32_t cap;        // records per (bin,class) list
  int32_t col_cap;    // cap * V
  int32_t bin_num;    // U*V*4
  int32_t tau;
  int32_t j_last;     // last block-row origin (9 + 3k < H-9)
  int32_t v_rem;      // H / bin_h : v-bin of the trailing flush (src/myComputeFeature.cpp:337)
  int32_t rad_u, rad_v;  // 3*bin_w, 3*bin_h (src/size.h:14-15)
};

struct FrameArrays {   // base pointers of slot 0; slot s is at + s * stride
  uint8_t* img;   int64_t img_stride;    // bytes
  int32_t* cnt;   int64_t cnt_stride;    // elements
  uint32_t* xy;   int64_t rec_stride;    // elements (bin_num*cap) -- shared by xy/meta
  uint32_t* meta;
  uint4* desc;                            // 2 * rec_stride uint4 per slot
  uint2* q;                               // rec_stride uint2 per slot: lower-bound signature (desc_q)
  uint16_t* duv;  int64_t duv_stride;     // optional (stereo engines; else nullptr): du | dv << 8 of every pixel, u16[H][pitch] per
                                          // slot, written by K1 from the planes it holds anyway and read by K5 (sub-pixel refinement)
};

// destination v-bin of block-row origin j (see oracle/vo_oracle.c dest_v_bin;
// src/myComputeFeature.cpp:319-345)
__host__ __device__ inline int32_t dest_v_bin(const Geom& g, int32_t j) {
  int32_t r = j / g.bin_h;
  int32_t flush_row = g.bin_h * (r + 1) - 3;
  return (flush_row <= g.j_last) ? r : g.v_rem;
}

// Lower-bound signature of a descriptor: two 32-bit words of four coded group sums each.  The 32 bytes are (du,dv) pairs of
// 16 sample points (src/myComputeFeature.cpp:11-55; offsets (dx,dy) of sample k in vo_extract.cu sdx/sdy); eight groups of
// four bytes partition them, so by the triangle inequality  SAD(a,b) >= sum_g |S_g(a) - S_g(b)|  with S_g the group sums
// (<= 1020).  Each sum is coded into one byte by a monotone integer map f with f(S+1)-f(S) in {0,1} (unit steps inside
// [410, 610], where sums of four Sobel responses live; one step per 16 outside), so |f(x) - f(y)| <= |x - y| and
//   SAD(a,b) >= Q(a,b) = Q0 + Q1,   Q_w = sum over the 4 groups of word w of |f(S_g(a)) - f(S_g(b))| = ONE VABSDIFF4 each,
// with no slack term; Q0 alone is a (weaker) bound as well.  ANY partition gives a valid bound; how tight it is depends on
// how little the byte differences inside a group cancel, i.e. on how correlated the grouped samples are.  The groups
// (r2t, chosen by annealing over partitions on the oracle's features, tools/signature_study.py) follow the smoothing
// direction of each Sobel response -- du is smooth along y, dv along x:
//   du: left  {0,1,12,13} = (-3,-+1),(-5,-+3)   right {4,5,14,15} = (3,-+1),(5,-+3)
//       top   {2,6,8,10}  = (-+1,-1),(-+1,-5)   bottom {3,7,9,11} = (-+1,1),(-+1,5)
//   dv: row y=-1 {0,2,4,6}   row y=+1 {1,3,5,7}   upper ring {8,10,12,14} (y<0)   lower ring {9,11,13,15} (y>0)
// With the frames' own best SAD as the bound, 2.7-3.0 % of an unmatched origin's window passes Q <= bound (sensor noise
// +-2 and +-8); the r1e partition by sample index ({4h..4h+3} for du and for dv) let 6.8-7.1 % through.
// (r1c used four groups of 8 bytes in one word: 11 % on noisy frames.)
__host__ __device__ inline uint32_t sum_code(uint32_t s) {
  const uint32_t lo = (s < 410u ? s : 410u) >> 4;                          // 0..25
  const uint32_t mid = (s < 410u ? 410u : (s > 610u ? 610u : s)) - 410u;   // 0..200
  const uint32_t hi = ((s > 610u ? s : 610u) - 610u) >> 4;                 // 0..25
  return lo + mid + hi;                                                    // 0..250
}
#ifdef VO_SIG_R1E  // the r1e partition, kept for A/B builds (tools/ab_build.py)
// sums of the even (sh = 0: du) or odd (sh = 8: dv) bytes of two descriptor words
__host__ __device__ inline uint32_t group_q(uint32_t w0, uint32_t w1, int sh) {
  const uint32_t x = ((w0 >> sh) & 0x00FF00FFu) + ((w1 >> sh) & 0x00FF00FFu);
  return sum_code((x + (x >> 16)) & 0xFFFFu);
}
__host__ __device__ inline uint32_t half_q(const uint4& a) {
  return group_q(a.x, a.y, 0) | (group_q(a.x, a.y, 8) << 8) | (group_q(a.z, a.w, 0) << 16) | (group_q(a.z, a.w, 8) << 24);
}
__host__ __device__ inline uint2 desc_q(const uint4& a, const uint4& b) { return make_uint2(half_q(a), half_q(b)); }
#else
// descriptor word w holds samples 2w, 2w+1 as bytes {du(2w), dv(2w), du(2w+1), dv(2w+1)}
__host__ __device__ inline uint2 desc_q(const uint4& a, const uint4& b) {
  constexpr uint32_t M = 0x00FF00FFu;
  // du, both samples of a word: bytes 0 and 2
  const uint32_t l = (a.x & M) + (b.z & M);                       // words 0, 6: samples 0,1,12,13
  const uint32_t r = (a.z & M) + (b.w & M);                       // words 2, 7: samples 4,5,14,15
  const uint32_t tb = (a.y & M) + (a.w & M) + (b.x & M) + (b.y & M);  // words 1,3,4,5: even samples (top) | odd samples (bottom)
  // dv: byte 1 = even sample, byte 3 = odd sample
  const uint32_t ud = ((a.x >> 8) & M) + ((a.y >> 8) & M) + ((a.z >> 8) & M) + ((a.w >> 8) & M);  // words 0..3: y=-1 | y=+1
  const uint32_t ring = ((b.x >> 8) & M) + ((b.y >> 8) & M) + ((b.z >> 8) & M) + ((b.w >> 8) & M);  // words 4..7: y<0 | y>0
  const uint32_t du_l = sum_code((l + (l >> 16)) & 0xFFFFu), du_r = sum_code((r + (r >> 16)) & 0xFFFFu);
  const uint32_t du_t = sum_code(tb & 0xFFFFu), du_b = sum_code(tb >> 16);
  const uint32_t dv_u = sum_code(ud & 0xFFFFu), dv_d = sum_code(ud >> 16);
  const uint32_t dv_t = sum_code(ring & 0xFFFFu), dv_b = sum_code(ring >> 16);
  // word 0 (the one-word filter of exact-partner tasks): left, right, row y=-1, lower ring; word 1: the other four
  return make_uint2(du_l | (du_r << 8) | (dv_u << 16) | (dv_b << 24), du_t | (du_b << 8) | (dv_d << 16) | (dv_t << 24));
}
#endif

}  // namespace vo
