// K4: match bucketing (sm_100a).
//
// Replaces, bit-exactly, Matcher::bucketFeatures + random_shuffle + rand_number
// (src/matcher.cpp:113-187; FPGA: second half of removeOutliers,
// hls_opencl/kernel/remove_outliers.cpp:104-182).
//
// The reference distributes the matches into 2-D buckets in input order, Fisher-Yates-shuffles
// every bucket with one LFSR stream that starts at 5 and carries across buckets, and keeps the
// first max_features of each.  The LFSR stream is input independent, so it is a precomputed table
// (built with the bit model of the reference's x86 double->int behaviour);
// bucket b consumes entries [S_b, S_b + len_b - 1) with S_b = sum over earlier buckets of
// max(len-1, 0).  Nothing is physically shuffled: for each kept output slot the swap sequence is
// traced BACKWARDS to the bucket-local rank of the match that ends up there, and a stable
// segmented rank (32 ordered segments, warp match_any inside) tells every match its own rank.
#include "vo_common.cuh"

namespace vo {

constexpr int K4_THREADS = 1024;
constexpr int K4_WARPS = K4_THREADS / 32;
constexpr int K4_SMEM_BUCKETS = 1024;  // bucket grids up to this size keep the (bucket, segment) histogram in shared memory (128 KB)

// per-pair scratch layout (int32 units): hist[nb_max][32] | cnt[nb_max] | soff[nb_max+1] | ooff[nb_max+1] | q[q_cap]
__host__ __device__ inline int64_t k4_scratch_ints(int nb_max, int q_cap) {
  return (int64_t)nb_max * 32 + nb_max + 2 * (int64_t)(nb_max + 1) + q_cap;
}

__device__ __forceinline__ float block_max(float v, float* red) {
  for (int d = 16; d > 0; d >>= 1) v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, d));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = red[0];
  for (int w = 1; w < K4_WARPS; w++) r = fmaxf(r, red[w]);
  __syncthreads();
  return r;
}

// current-image coordinates of match i: from the 4-byte side array when there is one, else from the record
__device__ __forceinline__ uint2 k4_coord(const uint32_t* __restrict__ side, const uint4* __restrict__ src, int i) {
  if (side) { const uint32_t x = side[i]; return make_uint2(__float_as_uint((float)(x & 0xFFFFu)), __float_as_uint((float)(x >> 16))); }
  const uint4 w1 = src[(int64_t)i * 3 + 1];
  return make_uint2(w1.z, w1.w);
}

// in: lists of 48-byte p_match records (3 x uint4), u1c/v1c are words 6/7
__global__ void __launch_bounds__(K4_THREADS)
k4_bucket(const uint4* __restrict__ in, int64_t in_stride_matches, const int32_t* __restrict__ n_in,
          int32_t max_features, float bucket_w, float bucket_h, const uint32_t* __restrict__ lfsr, int32_t lfsr_len,
          int32_t* __restrict__ scratch, int64_t scratch_stride, int32_t nb_max, int32_t q_cap,
          uint4* __restrict__ out, int64_t out_stride_matches, int32_t out_cap, int32_t* __restrict__ n_out,
          int32_t* __restrict__ status, const uint32_t* __restrict__ side_all) {
  extern __shared__ int32_t sh_hist[];
  __shared__ float red[K4_WARPS];
  __shared__ int32_t s_carry[2];
  __shared__ int32_t s_scan[K4_WARPS][2];
  const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = n_in[pair];
  const uint4* src = in + (int64_t)pair * in_stride_matches * 3;
  uint4* dst = out + (int64_t)pair * out_stride_matches * 3;
  // optional 4-byte (u1c | v1c << 16) per record, written by K3's emit for lists of integer coordinates: the three
  // passes then stream 4 bytes per match instead of the 48-byte records (r1f: 208 MB of DRAM reads per 128 pairs)
  const uint32_t* side = side_all ? side_all + (int64_t)pair * in_stride_matches : nullptr;
  int32_t* ghist = scratch + (int64_t)pair * scratch_stride;
  int32_t* cnt = ghist + (int64_t)nb_max * 32;
  int32_t* soff = cnt + nb_max;
  int32_t* ooff = soff + nb_max + 1;
  int32_t* q = ooff + nb_max + 1;

  // ---- extent of the current-image coordinates (src/matcher.cpp:143-152) -------------------
  float um = 0.f, vm = 0.f;
  // (the passes are latency bound -- one CTA per pair, ~17 matches per thread: four independent loads in flight per thread)
  for (int i = tid; i < n; i += 4 * K4_THREADS) {
    uint2 x[4];
#pragma unroll
    for (int k = 0; k < 4; k++) x[k] = (i + k * K4_THREADS < n) ? k4_coord(side, src, i + k * K4_THREADS) : make_uint2(0u, 0u);
#pragma unroll
    for (int k = 0; k < 4; k++) { um = fmaxf(um, __uint_as_float(x[k].x)); vm = fmaxf(vm, __uint_as_float(x[k].y)); }
  }
  um = block_max(um, red);
  vm = block_max(vm, red);
  const int cols = (int)floorf(__fdiv_rn(um, bucket_w)) + 1;
  const int rows = (int)floorf(__fdiv_rn(vm, bucket_h)) + 1;
  const int nb = cols * rows;
  const int keep = max(max_features, 1);  // the refill loop always takes one (src/matcher.cpp:178-184)
  if (nb > nb_max || n > lfsr_len || nb <= 0) {
    if (tid == 0) { n_out[pair] = 0; atomicExch(status, 2); }
    return;
  }

  int32_t* hist = (nb <= K4_SMEM_BUCKETS) ? sh_hist : ghist;
  // ---- pass 1: histogram per (bucket, ordered segment) --------------------------------------
  for (int i = tid; i < nb * 32; i += K4_THREADS) hist[i] = 0;
  __syncthreads();
  const int seg_len = (n + K4_WARPS - 1) / K4_WARPS;
  const int seg_lo = min(warp * seg_len, n), seg_hi = min(seg_lo + seg_len, n);
  for (int i = seg_lo + lane; i < seg_hi; i += 4 * 32) {
    uint2 x[4];
#pragma unroll
    for (int k = 0; k < 4; k++) x[k] = (i + k * 32 < seg_hi) ? k4_coord(side, src, i + k * 32) : make_uint2(0u, 0u);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (i + k * 32 < seg_hi) {
        const int bu = (int)floorf(__fdiv_rn(__uint_as_float(x[k].x), bucket_w));
        const int bv = (int)floorf(__fdiv_rn(__uint_as_float(x[k].y), bucket_h));
        atomicAdd(&hist[(bv * cols + bu) * 32 + warp], 1);
      }
    }
  }
  __syncthreads();
  // exclusive scan over the 32 segments of every bucket (warp per bucket)
  for (int b = warp; b < nb; b += K4_WARPS) {
    const int v = hist[b * 32 + lane];
    int incl = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= d) incl += o;
    }
    hist[b * 32 + lane] = incl - v;
    if (lane == 31) cnt[b] = incl;
  }
  __syncthreads();
  // ---- exclusive scans over buckets: LFSR offsets and output offsets -----------------------
  if (tid == 0) { s_carry[0] = 0; s_carry[1] = 0; }
  __syncthreads();
  for (int base = 0; base < nb; base += K4_THREADS) {
    const int b = base + tid;
    const int len = (b < nb) ? cnt[b] : 0;
    const int a0 = max(len - 1, 0), a1 = min(len, keep);
    int i0 = a0, i1 = a1;
    for (int d = 1; d < 32; d <<= 1) {
      const int o0 = __shfl_up_sync(0xFFFFFFFFu, i0, d), o1 = __shfl_up_sync(0xFFFFFFFFu, i1, d);
      if (lane >= d) { i0 += o0; i1 += o1; }
    }
    if (lane == 31) { s_scan[warp][0] = i0; s_scan[warp][1] = i1; }
    __syncthreads();
    int p0 = s_carry[0], p1 = s_carry[1];
    for (int w = 0; w < warp; w++) { p0 += s_scan[w][0]; p1 += s_scan[w][1]; }
    if (b < nb) { soff[b] = p0 + i0 - a0; ooff[b] = p1 + i1 - a1; }
    __syncthreads();
    if (tid == K4_THREADS - 1) { s_carry[0] = p0 + i0; s_carry[1] = p1 + i1; }
    __syncthreads();
  }
  const int total_keep = s_carry[1];
  if (total_keep > q_cap || total_keep > out_cap) {
    if (tid == 0) { n_out[pair] = 0; atomicExch(status, 3); }
    return;
  }
  // ---- backward trace of the shuffle for every kept slot -----------------------------------
  for (int b = tid; b < nb; b += K4_THREADS) {
    const int len = cnt[b], kept = min(len, keep), s0 = soff[b], o0 = ooff[b];
    for (int t = 0; t < kept; t++) {
      int pos = t;
      for (int i = len - 1; i >= 1; i--) {  // undo swap(p[i], p[rand % (i+1)]) (src/matcher.cpp:128-136)
        const int j = (int)(lfsr[s0 + i - 1] % (uint32_t)(i + 1));
        pos = (pos == i) ? j : ((pos == j) ? i : pos);
      }
      q[o0 + t] = pos;
    }
  }
  __syncthreads();
  // ---- pass 2: stable rank of every match inside its bucket; copy the selected ones -------
  uint2 xa = (seg_lo + lane < seg_hi) ? k4_coord(side, src, seg_lo + lane) : make_uint2(0u, 0u);
  uint2 xb = (seg_lo + 32 + lane < seg_hi) ? k4_coord(side, src, seg_lo + 32 + lane) : make_uint2(0u, 0u);
  for (int i0 = seg_lo; i0 < seg_hi; i0 += 32) {
    const int i = i0 + lane;
    const bool act = i < seg_hi;
    const uint32_t amask = __ballot_sync(0xFFFFFFFFu, act);
    const uint2 xc = xa;  // coordinates of this round; the next two rounds' loads are already in flight
    xa = xb;
    xb = (i + 64 < seg_hi) ? k4_coord(side, src, i + 64) : make_uint2(0u, 0u);
    if (act) {
      const int bu = (int)floorf(__fdiv_rn(__uint_as_float(xc.x), bucket_w));
      const int bv = (int)floorf(__fdiv_rn(__uint_as_float(xc.y), bucket_h));
      const int b = bv * cols + bu;
      const uint32_t same = __match_any_sync(amask, b);
      const int rank = hist[b * 32 + warp] + __popc(same & ((1u << lane) - 1u));
      __syncwarp(amask);
      if ((same & ((1u << lane) - 1u)) == 0) hist[b * 32 + warp] += __popc(same);
      const int kept = min(cnt[b], keep), o0 = ooff[b];
      for (int t = 0; t < kept; t++)
        if (q[o0 + t] == rank) {  // the few selected records are the only ones read in full
          uint4* rec = dst + (int64_t)(o0 + t) * 3;
          rec[0] = src[(int64_t)i * 3]; rec[1] = src[(int64_t)i * 3 + 1]; rec[2] = src[(int64_t)i * 3 + 2];
        }
    }
    __syncwarp();
  }
  if (tid == 0) n_out[pair] = total_keep;
}

int64_t k4_scratch_ints_host(int nb_max, int q_cap) { return k4_scratch_ints(nb_max, q_cap); }

// per-device opt-in to K4's dynamic shared memory (see init_k1_device)
cudaError_t init_k4_device() {
  return cudaFuncSetAttribute(k4_bucket, cudaFuncAttributeMaxDynamicSharedMemorySize, K4_SMEM_BUCKETS * 32 * (int)sizeof(int32_t));
}

cudaError_t launch_k4(const void* in, int64_t in_stride_matches, const int32_t* n_in, int n_pairs, int32_t max_features,
                      float bw, float bh, const uint32_t* lfsr, int32_t lfsr_len, int32_t* scratch, int64_t scratch_stride,
                      int32_t nb_max, int32_t q_cap, void* out, int64_t out_stride_matches, int32_t out_cap,
                      int32_t* n_out, int32_t* status, const uint32_t* side, cudaStream_t st) {
  const int smem = K4_SMEM_BUCKETS * 32 * (int)sizeof(int32_t);
  k4_bucket<<<n_pairs, K4_THREADS, smem, st>>>((const uint4*)in, in_stride_matches, n_in, max_features, bw, bh, lfsr,
                                            lfsr_len, scratch, scratch_stride, nb_max, q_cap, (uint4*)out,
                                            out_stride_matches, out_cap, n_out, status, side);
  return cudaGetLastError();
}

// ---- allocation-free variant (vo_bucket_matches_noalloc; no counterpart in the reference)
// ----------------------------------------------------------------------------------------------------
// Same kernel, but its per-list scratch (histogram | counts | LFSR offsets | output offsets | traced ranks) lives in a
// file-scope device array instead of engine-owned cudaMalloc memory, so a caller that may not allocate (CUDA graph capture,
// a serving loop, a harness that forbids it) can bucket a device-resident list.  Every GPU that loads the library holds the
// array until exit (kNoAllocInts * 4 bytes = 624 KB).  ONE array per device: calls that overlap on the GPU race, and the
// array keeps the previous call's contents -- the kernel zeroes the histogram it uses and writes every count / offset /
// rank entry before reading it; the header word (list length) is rewritten by k4_noalloc_prepare on every call.
constexpr int kNoAllocMaxMatches = 32768;  // stated limit of the entry point (a 1241x376 pair has ~17 k raw matches)
constexpr int kNoAllocNbMax = 4096;        // buckets per list, as the engine's own scratch
constexpr int kNoAllocQCap = 16384;        // kept matches per list, as the engine's own scratch
constexpr int kNoAllocHdr = 64;            // header ints in front of the kernel's scratch (256 bytes: keeps its alignment)
constexpr int64_t kNoAllocInts = kNoAllocHdr + (int64_t)kNoAllocNbMax * 32 + kNoAllocNbMax + 2 * (int64_t)(kNoAllocNbMax + 1) + kNoAllocQCap;
static __device__ __align__(256) int32_t g_k4_noalloc[kNoAllocInts];

// In host code `g_k4_noalloc` is only a stand-in for the symbol, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, once per device with that device current (init_k4_device, called by vo_engine_create)
constexpr int kMaxDevices = 64;
static int32_t* g_k4_noalloc_dev[kMaxDevices];

// header word 0 = list length (k4_bucket reads it through its n_in pointer); result = {n_out, status} starts at zero
__global__ void k4_noalloc_prepare(int32_t* __restrict__ hdr, int32_t n_in, int32_t* __restrict__ result) {
  if (threadIdx.x == 0) { hdr[0] = n_in; result[0] = 0; result[1] = 0; }
}

cudaError_t init_k4_noalloc_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  int32_t* ptr = nullptr;
  e = cudaGetSymbolAddress((void**)&ptr, g_k4_noalloc);
  if (e != cudaSuccess) return e;
  g_k4_noalloc_dev[dev] = ptr;
  return cudaSuccess;
}

int k4_noalloc_max_matches() { return kNoAllocMaxMatches; }

// in / out / result are DEVICE pointers of the caller; two launches on `st`, nothing else (no allocation, no synchronisation)
cudaError_t launch_k4_noalloc(int device, const void* in, int32_t n_in, int32_t max_features, float bw, float bh,
                              const uint32_t* lfsr, int32_t lfsr_len, void* out, int32_t out_cap, int32_t* result,
                              cudaStream_t st) {
  if (device < 0 || device >= kMaxDevices || g_k4_noalloc_dev[device] == nullptr) return cudaErrorNotReady;
  int32_t* base = g_k4_noalloc_dev[device];
  k4_noalloc_prepare<<<1, 32, 0, st>>>(base, n_in, result);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int smem = K4_SMEM_BUCKETS * 32 * (int)sizeof(int32_t);
  k4_bucket<<<1, K4_THREADS, smem, st>>>((const uint4*)in, (int64_t)n_in, base, max_features, bw, bh, lfsr, lfsr_len,
                                         base + kNoAllocHdr, kNoAllocInts - kNoAllocHdr, kNoAllocNbMax, kNoAllocQCap,
                                         (uint4*)out, (int64_t)out_cap, out_cap, result, result + 1, nullptr);
  return cudaGetLastError();
}

}  // namespace vo
