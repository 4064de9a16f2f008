// Shared device-side definitions of libvo_b200 (sm_100a only).
//
// HBM layout of one frame slot (all arrays are engine-owned, slot-major):
//   img   u8  [H][pitch]                 pitch = roundup(W,16)  (TMA needs 16-byte row strides)
//   cnt   i32 [bin_num]                  bin = (v_bin*U + u_bin)*4 + class     == max2_num
//   xy    u32 [bin_num][cap]             u | v<<16
//   meta  u32 [bin_num][cap]             "twin": lower-class feature on the same pixel
//                                        (class<<16 | index in its list) or 0xFFFFFFFF
//   desc  uint4[bin_num][cap][2]         the 32 descriptor bytes in reference byte order
//   q     uint2[bin_num][cap]            8 quantised group sums of the descriptor (SAD lower bound, K2)
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
};

// destination v-bin of block-row origin j (see oracle/vo_oracle.c dest_v_bin;
// src/myComputeFeature.cpp:319-345)
__host__ __device__ inline int32_t dest_v_bin(const Geom& g, int32_t j) {
  int32_t r = j / g.bin_h;
  int32_t flush_row = g.bin_h * (r + 1) - 3;
  return (flush_row <= g.j_last) ? r : g.v_rem;
}

// Lower-bound signature of a descriptor.  The 32 bytes are (du,dv) pairs of 16 sample points; the
// eight groups {du of samples 4j..4j+3}, {dv of samples 4j..4j+3} partition them, so by the
// triangle inequality  SAD(a,b) >= sum_g |S_g(a) - S_g(b)|  with S_g the group sums (<= 1020).
// With q_g = S_g >> 2 (one byte):  |S_g(a)-S_g(b)| >= 4|q_g(a)-q_g(b)| - 3, hence
//   SAD(a,b) >= 4 * Q(a,b) - 24,   Q = sum_g |q_g(a) - q_g(b)|  = two VABSDIFF4.
__host__ __device__ inline uint32_t group_q(uint32_t w0, uint32_t w1) {
  const uint32_t x = (w0 & 0x00FF00FFu) + (w1 & 0x00FF00FFu);
  return ((x + (x >> 16)) & 0xFFFFu) >> 2;
}
__host__ __device__ inline uint2 desc_q(const uint4& a, const uint4& b) {
  uint2 q;
  q.x = group_q(a.x, a.y) | (group_q(a.z, a.w) << 8) | (group_q(b.x, b.y) << 16) | (group_q(b.z, b.w) << 24);
  q.y = group_q(a.x >> 8, a.y >> 8) | (group_q(a.z >> 8, a.w >> 8) << 8) | (group_q(b.x >> 8, b.y >> 8) << 16) |
        (group_q(b.z >> 8, b.w >> 8) << 24);
  return q;
}

}  // namespace vo
