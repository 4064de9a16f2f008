// K2 / K3: bucketed circular descriptor matching (sm_100a).
//
// Replaces, bit-exactly, the reference's
//   fetch_col_bin   src/myMatch.cpp:17-68     (column lists, col_idx)
//   find_match      src/myMatch.cpp:70-119    (windowed SAD argmin, first-found-wins)
//   myMatching      src/myMatch.cpp:122-204   (both directions, cycle check, pixel dedupe, emit)
// in the ring-buffer-free formulation of SURVEY.md appendix A.
//
// K2 (k2_match): one CTA per (origin u-bin column, group of origin v-bin rows, class, direction,
// frame pair).  The CTA stages the target window of the OTHER frame -- up to 6 u-bin columns x
// the v-bin rows its origins can see -- into shared memory ordered (v_bin, u_bin, slot), so that
// every origin's candidate set is ONE contiguous range.  A warp owns an origin; its lanes stride
// the range, each candidate costing 8 VABSDIFF4.U8.ACC (packed-byte SAD) + a packed window test,
// and keeps min(SAD<<19 | u_rel<<16 | col_idx), which is exactly the reference's "strict <,
// first candidate in (u_bin asc, col_idx asc) order wins".  One REDUX.MIN per origin.
// Optional exact pruning (PRUNE): a 2-instruction lower bound of the SAD (see desc_q) skips the
// 8-instruction exact SAD for candidates that provably cannot win or tie; results are identical.
// SAD is not a contraction: no tensor cores here by design.
//
// K3 (k3_flags, k3_emit): closed-cycle test, same-pixel dedupe (lowest class wins, via the twin
// links written by K1), ordered emit of p_match records in (u_bin, class, col_idx) order.
#include "vo_common.cuh"

namespace vo {

constexpr int K2_THREADS = 512;
constexpr int K2_WARPS = K2_THREADS / 32;
constexpr int kMaxV = 64;       // v-bin rows supported by the shared-memory tables
constexpr int kMaxCols = 6;     // target u-bin columns: [ub-3, ub+3) clipped
constexpr int kCandCap = 2176;  // staged candidates per chunk; >= 6*300 so one window row always fits

struct K2Tables {
  int32_t cntT[kMaxV][kMaxCols];      // target counts (row, rel. column)
  int32_t colT[kMaxCols][kMaxV + 1];  // col_idx prefix of each target column
  int32_t rowTot[kMaxV];              // candidates per staged row
  int32_t stageOff[kMaxV + 1];        // staged position of the first candidate of a row (chunk-relative)
  int32_t cntO[kMaxV];
  int32_t prefO[kMaxV + 1];
  int32_t lagEnd[kMaxV];              // origins o < lagEnd[v] of row v may carry a lagging v_bin (slow path)
  int32_t posLo[kMaxV], posHi[kMaxV]; // staged candidate range of a non-lagging origin of row v (this chunk)
};

template <bool PRUNE>
struct __align__(128) K2Smem {
  K2Tables t;
  __align__(16) uint4 lo[kCandCap];   // descriptor bytes 0..15
  uint4 hi[kCandCap];                 // descriptor bytes 16..31
  uint2 xr[kCandCap];                 // {xy + 0x80008000, u_rel<<16 | col_idx}
  uint2 q[PRUNE ? kCandCap : 1];      // 8 quantised group sums (lower-bound filter)
};

// acc + sum of |a_i - b_i| over the four packed bytes: one VABSDIFF4.U8.ACC
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t acc) {
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
  return d;
}

__device__ __forceinline__ uint32_t sad32(const uint4& a, const uint4& b, const uint4& x, const uint4& y) {
  uint32_t s = sad4(a.x, x.x, 0u);
  s = sad4(a.y, x.y, s);
  s = sad4(a.z, x.z, s);
  s = sad4(a.w, x.w, s);
  s = sad4(b.x, y.x, s);
  s = sad4(b.y, y.y, s);
  s = sad4(b.z, y.z, s);
  return sad4(b.w, y.w, s);
}

// v_buffer_idx of the reference (src/myMatch.cpp:153-158): starts at 0, bumped at most once per
// origin => vb(k) = min(k+1, v+1, min over non-empty rows b<v of b+2+k-prefix[b+1]).
__device__ __forceinline__ int origin_vb(const K2Tables& t, int col_idx, int v) {
  int vb = min(col_idx + 1, v + 1);
  for (int b = 0; b < v; b++)
    if (t.cntO[b] > 0) vb = min(vb, b + 2 + col_idx - t.prefO[b + 1]);
  return vb;
}

template <bool PRUNE>
__global__ void __launch_bounds__(K2_THREADS, 2)
k2_match(Geom g, FrameArrays fa, const int2* __restrict__ pairs, uint32_t* __restrict__ best, int rows_per_group,
         int ngroups) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  K2Smem<PRUNE>& S = *reinterpret_cast<K2Smem<PRUNE>*>(smem_raw);
  K2Tables& t = S.t;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ub = blockIdx.x / ngroups, grp = blockIdx.x - ub * ngroups;
  const int c = blockIdx.y >> 1, dir = blockIdx.y & 1;
  const int2 pr = pairs[blockIdx.z];  // x = prev slot, y = curr slot
  const int from_slot = dir ? pr.x : pr.y, to_slot = dir ? pr.y : pr.x;
  const int U = g.U, V = g.V;
  const int ulo = max(ub - 3, 0), uhi = min(ub + 3, U - 1);  // [ulo, uhi): src/myMatch.cpp:90-91,97
  const int ncol = uhi - ulo;

  const int32_t* cnt_to = fa.cnt + (int64_t)to_slot * fa.cnt_stride;
  const int32_t* cnt_from = fa.cnt + (int64_t)from_slot * fa.cnt_stride;
  for (int idx = tid; idx < V * kMaxCols; idx += K2_THREADS) {
    const int v = idx / kMaxCols, j = idx - v * kMaxCols;
    t.cntT[v][j] = (j < ncol) ? cnt_to[(v * U + ulo + j) * 4 + c] : 0;
  }
  for (int v = tid; v < V; v += K2_THREADS) t.cntO[v] = cnt_from[(v * U + ub) * 4 + c];
  __syncthreads();
  if (tid < kMaxCols) {
    int acc = 0;
    t.colT[tid][0] = 0;
    for (int v = 0; v < V; v++) { acc += t.cntT[v][tid]; t.colT[tid][v + 1] = acc; }
  } else if (tid == 32) {
    int acc = 0;
    t.prefO[0] = 0;
    for (int v = 0; v < V; v++) { acc += t.cntO[v]; t.prefO[v + 1] = acc; }
  } else if (tid >= 64 && tid < 64 + V) {
    const int v = tid - 64;
    int acc = 0;
#pragma unroll
    for (int j = 0; j < kMaxCols; j++) acc += t.cntT[v][j];
    t.rowTot[v] = acc;
  }
  __syncthreads();
  if (tid < V) {  // first origin index of row v whose v_bin is the nominal v+1 (see origin_vb)
    int le = tid;
    for (int b = 0; b < tid; b++)
      if (t.cntO[b] > 0) le = max(le, t.prefO[b + 1] + tid - 1 - b);
    t.lagEnd[tid] = le;
  }

  const int r0 = grp * rows_per_group, r1 = min(r0 + rows_per_group, V);
  const int o_begin = t.prefO[r0], o_end = t.prefO[r1];
  if (o_begin >= o_end) return;

  // rows any origin of this group can see: vb is non-decreasing along the column list
  int v_first = r0;
  while (t.prefO[v_first + 1] <= o_begin) v_first++;
  const int vb_first = origin_vb(t, o_begin, v_first);
  const int row_lo = max(vb_first - 3, 0);
  const int row_hi = min(r1 + 3, V - 1);  // exclusive: src/myMatch.cpp:93,99

  const uint32_t* xy_from = fa.xy + (int64_t)from_slot * fa.rec_stride;
  const uint4* desc_from = fa.desc + (int64_t)from_slot * fa.rec_stride * 2;
  const uint32_t* xy_to = fa.xy + (int64_t)to_slot * fa.rec_stride;
  const uint4* desc_to = fa.desc + (int64_t)to_slot * fa.rec_stride * 2;
  const uint2* q_from = fa.q + (int64_t)from_slot * fa.rec_stride;
  const uint2* q_to = fa.q + (int64_t)to_slot * fa.rec_stride;
  uint32_t* best_out = best + ((int64_t)(blockIdx.z * 2 + dir) * U * 4 + (ub * 4 + c)) * g.col_cap;
  const int rec_row_stride = U * 4 * g.cap;              // records between consecutive rows of one (u,c)
  const int rec_col0 = (ub * 4 + c) * g.cap;             // record index of row 0 of the origin column

  bool first_chunk = true;
  for (int ra = row_lo;;) {
    // ---- chunk = as many whole rows as fit the staging buffer ------------------------------
    int rb = ra, staged = 0;
    while (rb < row_hi && staged + t.rowTot[rb] <= kCandCap) { staged += t.rowTot[rb]; rb++; }
    __syncthreads();  // previous chunk fully consumed (and lagEnd visible)
    if (tid == 0) {
      int acc = 0;
      for (int v = ra; v < rb; v++) { t.stageOff[v] = acc; acc += t.rowTot[v]; }
      t.stageOff[rb] = acc;
    }
    __syncthreads();
    if (tid < V) {  // nominal window of row tid: v-bins [v-2, v+4) clipped (vb = v+1)
      const int vlo = max(max(tid - 2, 0), ra), vhi = min(min(tid + 4, V - 1), rb);
      t.posLo[tid] = (vlo < vhi) ? t.stageOff[vlo] : 0;
      t.posHi[tid] = (vlo < vhi) ? t.stageOff[vhi] : 0;
    }
    // ---- stage: warp per (row, column) bin list, coalesced 16-byte copies ------------------
    for (int b = warp; b < (rb - ra) * ncol; b += K2_WARPS) {
      const int v = ra + b / ncol, j = b % ncol;
      const int n = t.cntT[v][j];
      int dst = t.stageOff[v];
      for (int jj = 0; jj < j; jj++) dst += t.cntT[v][jj];
      const int64_t src = (int64_t)((v * U + ulo + j) * 4 + c) * g.cap;
      const uint32_t rank0 = ((uint32_t)j << 16) | (uint32_t)t.colT[j][v];
      for (int k = lane; k < n; k += 32) {
        const uint4 d0 = desc_to[(src + k) * 2], d1 = desc_to[(src + k) * 2 + 1];
        S.lo[dst + k] = d0;
        S.hi[dst + k] = d1;
        S.xr[dst + k] = make_uint2(xy_to[src + k] + 0x80008000u, rank0 + (uint32_t)k);
        if (PRUNE) S.q[dst + k] = q_to[src + k];
      }
    }
    __syncthreads();

    // ---- origins: one warp each, next origin's record prefetched while this one is matched ---
    int o = o_begin + warp, vl = v_first;
    uint32_t n_xy = 0;
    uint4 n_a = make_uint4(0, 0, 0, 0), n_b = n_a;
    uint2 n_q = make_uint2(0, 0);
    if (o < o_end) {
      while (t.prefO[vl + 1] <= o) vl++;
      const int rec = rec_col0 + vl * rec_row_stride + (o - t.prefO[vl]);
      n_xy = __ldg(&xy_from[rec]); n_a = __ldg(&desc_from[2 * rec]); n_b = __ldg(&desc_from[2 * rec + 1]);
      if (PRUNE) n_q = __ldg(&q_from[rec]);
    }
    while (o < o_end) {
      const uint32_t oxy = n_xy;
      const uint4 a = n_a, b = n_b;
      const uint2 qa = n_q;
      const int v = vl;  // row of origin o (found when it was prefetched)
      const int on = o + K2_WARPS;
      if (on < o_end) {
        while (t.prefO[vl + 1] <= on) vl++;
        const int rec = rec_col0 + vl * rec_row_stride + (on - t.prefO[vl]);
        n_xy = __ldg(&xy_from[rec]); n_a = __ldg(&desc_from[2 * rec]); n_b = __ldg(&desc_from[2 * rec + 1]);
        if (PRUNE) n_q = __ldg(&q_from[rec]);
      }
      int lo, hi;
      if (o >= t.lagEnd[v]) {
        lo = t.posLo[v]; hi = t.posHi[v];
      } else {  // rare: the first origins after empty bins carry a lagging v_bin
        const int vb = origin_vb(t, o, v);
        const int vlo = max(max(vb - 3, 0), ra), vhi = min(min(vb + 3, V - 1), rb);
        lo = (vlo < vhi) ? t.stageOff[vlo] : 0;
        hi = (vlo < vhi) ? t.stageOff[vhi] : 0;
      }
      const int ou = (int)(oxy & 0xFFFFu), ov = (int)(oxy >> 16);
      // packed window test: both 16-bit fields carry a guard bit (bit 15); DESIGN.md "window test"
      const uint32_t lo_p = (uint32_t)max(ou - g.rad_u, 0) | ((uint32_t)max(ov - g.rad_v, 0) << 16);
      const uint32_t hi_p = (uint32_t)(ou + g.rad_u + 1) | ((uint32_t)(ov + g.rad_v + 1) << 16);
      uint32_t bestk = kNone;
      if (!PRUNE) {
#pragma unroll 4
        for (int pos = lo + lane; pos < hi; pos += 32) {
          const uint2 xr = S.xr[pos];
          const uint32_t sad = sad32(a, b, S.lo[pos], S.hi[pos]);
          const uint32_t t1 = xr.x - lo_p, t2 = xr.x - hi_p;
          const uint32_t key = sad * 0x80000u + xr.y;
          if (((t1 & ~t2) & 0x80008000u) == 0x80008000u) bestk = min(bestk, key);
        }
      } else {
        // pass 1: the candidate with the smallest lower bound, per lane
        uint32_t k1 = kNone;
#pragma unroll 4
        for (int pos = lo + lane; pos < hi; pos += 32) {
          const uint2 q = S.q[pos];
          const uint32_t Q = sad4(qa.y, q.y, sad4(qa.x, q.x, 0u));
          k1 = min(k1, Q * 0x100000u + (uint32_t)(pos - lo));
        }
        uint32_t tq = 0;
        // exact SAD + window test of one staged candidate; tightens this lane's bound
        auto exact = [&](int pos) {
          const uint2 xr = S.xr[pos];
          const uint32_t sad = sad32(a, b, S.lo[pos], S.hi[pos]);
          const uint32_t t1 = xr.x - lo_p, t2 = xr.x - hi_p;
          const uint32_t key = sad * 0x80000u + xr.y;
          if (((t1 & ~t2) & 0x80008000u) == 0x80008000u && key < bestk) {
            bestk = key;
            tq = min(tq, (sad + 24u) >> 2);
          }
        };
        tq = kNone;
        if (k1 != kNone) exact(lo + (int)(k1 & 0xFFFFFu));
        // any candidate that can still win (or tie) has 4Q-24 <= best SAD so far
        uint32_t tsad = __reduce_min_sync(0xFFFFFFFFu, bestk) >> 19;
        if (!first_chunk) tsad = min(tsad, best_out[o] >> 19);
        tq = (tsad + 24u) >> 2;
        // pass 2: exact SAD only where the bound allows; four candidates share one branch
        int pos = lo + lane;
        for (; pos + 96 < hi; pos += 128) {
          const uint2 q0 = S.q[pos], q1 = S.q[pos + 32], q2 = S.q[pos + 64], q3 = S.q[pos + 96];
          const uint32_t Q0 = sad4(qa.y, q0.y, sad4(qa.x, q0.x, 0u));
          const uint32_t Q1 = sad4(qa.y, q1.y, sad4(qa.x, q1.x, 0u));
          const uint32_t Q2 = sad4(qa.y, q2.y, sad4(qa.x, q2.x, 0u));
          const uint32_t Q3 = sad4(qa.y, q3.y, sad4(qa.x, q3.x, 0u));
          if (min(min(Q0, Q1), min(Q2, Q3)) <= tq) {
            if (Q0 <= tq) exact(pos);
            if (Q1 <= tq) exact(pos + 32);
            if (Q2 <= tq) exact(pos + 64);
            if (Q3 <= tq) exact(pos + 96);
          }
        }
#pragma unroll 1
        for (; pos < hi; pos += 32) {
          const uint2 q = S.q[pos];
          const uint32_t Q = sad4(qa.y, q.y, sad4(qa.x, q.x, 0u));
          if (Q <= tq) exact(pos);
        }
      }
      bestk = __reduce_min_sync(0xFFFFFFFFu, bestk);
      if (lane == 0) {
        if (!first_chunk) bestk = min(bestk, best_out[o]);
        best_out[o] = bestk;
      }
      o = on;
    }
    first_chunk = false;
    ra = rb;
    if (ra >= row_hi) break;
  }
}

// decode a K2 key of an origin in column `ub`: target column and col_idx
__device__ __forceinline__ void decode_key(uint32_t key, int ub, int& tu, int& tk) {
  tu = max(ub - 3, 0) + (int)((key >> 16) & 7u);
  tk = (int)(key & 0xFFFFu);
}

__device__ __forceinline__ bool cycle_closed(const uint32_t* __restrict__ best_cp, const uint32_t* __restrict__ best_pc,
                                             int col_cap, int u, int c, int k, uint32_t* key_out) {
  const uint32_t key = best_cp[(int64_t)(u * 4 + c) * col_cap + k];
  if (key_out) *key_out = key;
  if (key == kNone) return false;
  int pu, pk;
  decode_key(key, u, pu, pk);
  const uint32_t key2 = best_pc[(int64_t)(pu * 4 + c) * col_cap + pk];
  if (key2 == kNone) return false;
  int cu, ck;
  decode_key(key2, pu, cu, ck);
  return cu == u && ck == k;  // src/myMatch.cpp:186
}

constexpr int K3_THREADS = 256;

// K3a: per current-frame column u: emit flag of every feature + per-(u,class) counts.
__global__ void __launch_bounds__(K3_THREADS)
k3_flags(Geom g, FrameArrays fa, const int2* __restrict__ pairs, const uint32_t* __restrict__ best,
         uint8_t* __restrict__ flags, int32_t* __restrict__ list_cnt) {
  __shared__ int32_t pref[4][kMaxV + 1];
  __shared__ int32_t warp_cnt[K3_THREADS / 32];
  const int u = blockIdx.x, pair = blockIdx.y, tid = threadIdx.x;
  const int U = g.U, V = g.V;
  const int curr = pairs[pair].y;
  const int32_t* cnt = fa.cnt + (int64_t)curr * fa.cnt_stride;
  if (tid < 4) {
    int acc = 0;
    pref[tid][0] = 0;
    for (int v = 0; v < V; v++) { acc += cnt[(v * U + u) * 4 + tid]; pref[tid][v + 1] = acc; }
  }
  __syncthreads();
  const uint32_t* best_cp = best + (int64_t)(pair * 2 + 0) * U * 4 * g.col_cap;
  const uint32_t* best_pc = best + (int64_t)(pair * 2 + 1) * U * 4 * g.col_cap;
  const uint32_t* meta = fa.meta + (int64_t)curr * fa.rec_stride;
  uint8_t* fl = flags + (int64_t)pair * U * 4 * g.col_cap;
  for (int c = 0; c < 4; c++) {
    const int n = pref[c][V];
    int mine = 0;
    if (u < U - 1) {  // the last column is never matched (src/myMatch.cpp:141-146)
      for (int k = tid; k < n; k += K3_THREADS) {
        bool emit = cycle_closed(best_cp, best_pc, g.col_cap, u, c, k, nullptr);
        if (emit && c >= 2) {
          int v = 0;
          while (pref[c][v + 1] <= k) v++;
          const uint32_t tw = meta[(int64_t)((v * U + u) * 4 + c) * g.cap + (k - pref[c][v])];
          if (tw != kNone) {
            const int tc = (int)(tw >> 16), tk = pref[tc][v] + (int)(tw & 0xFFFFu);
            // a lower class already claimed this current pixel (M[][] of src/myMatch.cpp:188-191)
            if (cycle_closed(best_cp, best_pc, g.col_cap, u, tc, tk, nullptr)) emit = false;
          }
        }
        fl[(int64_t)(u * 4 + c) * g.col_cap + k] = emit ? 1 : 0;
        mine += emit ? 1 : 0;
      }
    }
    for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((tid & 31) == 0) warp_cnt[tid >> 5] = mine;
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < K3_THREADS / 32; w++) tot += warp_cnt[w];
      list_cnt[(int64_t)pair * U * 4 + u * 4 + c] = tot;
    }
    __syncthreads();
  }
}

// K3b: ordered emit.  One CTA per (u,class) list; its output offset is the sum of the counts of
// all lists before it in (u_bin, class) order (src/myMatch.cpp:181-196).
__global__ void __launch_bounds__(K3_THREADS)
k3_emit(Geom g, FrameArrays fa, const int2* __restrict__ pairs, const uint32_t* __restrict__ best,
        const uint8_t* __restrict__ flags, const int32_t* __restrict__ list_cnt, uint4* __restrict__ out,
        int64_t out_stride_matches, int32_t max_matches, int32_t* __restrict__ n_raw, int32_t* __restrict__ status) {
  __shared__ int32_t pref[kMaxV + 1];
  __shared__ int32_t warp_tot[K3_THREADS / 32];
  __shared__ int32_t s_off, s_total;
  const int list = blockIdx.x, pair = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int U = g.U, V = g.V, u = list >> 2, c = list & 3;
  const int2 pr = pairs[pair];
  const int32_t* lc = list_cnt + (int64_t)pair * U * 4;
  if (warp == 0) {
    int before = 0, total = 0;
    for (int l = lane; l < U * 4; l += 32) {
      const int n = lc[l];
      total += n;
      if (l < list) before += n;
    }
    for (int d = 16; d > 0; d >>= 1) {
      before += __shfl_xor_sync(0xFFFFFFFFu, before, d);
      total += __shfl_xor_sync(0xFFFFFFFFu, total, d);
    }
    if (lane == 0) { s_off = before; s_total = total; }
  } else if (tid == 32) {
    const int32_t* cnt = fa.cnt + (int64_t)pr.y * fa.cnt_stride;
    int acc = 0;
    pref[0] = 0;
    for (int v = 0; v < V; v++) { acc += cnt[(v * U + u) * 4 + c]; pref[v + 1] = acc; }
  }
  __syncthreads();
  const int total = s_total;
  if (list == 0 && tid == 0) {
    n_raw[pair] = (total < 5) ? 0 : min(total, max_matches);  // "<5 -> 0": src/myMatch.cpp:198-203
    if (total > max_matches) atomicExch(status, 1);
  }
  if (lc[list] == 0) return;
  const int n = pref[V];
  const uint32_t* best_cp = best + (int64_t)(pair * 2 + 0) * U * 4 * g.col_cap;
  const uint8_t* fl = flags + ((int64_t)pair * U * 4 + list) * g.col_cap;
  const uint32_t* xy_c = fa.xy + (int64_t)pr.y * fa.rec_stride;
  const uint32_t* xy_p = fa.xy + (int64_t)pr.x * fa.rec_stride;
  const int32_t* cnt_p = fa.cnt + (int64_t)pr.x * fa.cnt_stride;
  uint4* dst = out + (int64_t)pair * out_stride_matches * 3;
  int running = s_off;
  for (int base = 0; base < n; base += K3_THREADS) {
    const int k = base + tid;
    const bool emit = (k < n) && fl[k];
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, emit);
    if (lane == 0) warp_tot[warp] = __popc(b);
    __syncthreads();
    int posn = running + __popc(b & ((1u << lane) - 1u));
    int chunk_total = 0;
#pragma unroll
    for (int w = 0; w < K3_THREADS / 32; w++) {
      const int tw = warp_tot[w];
      if (w < warp) posn += tw;
      chunk_total += tw;
    }
    if (emit && posn < max_matches) {
      int v = 0;
      while (pref[v + 1] <= k) v++;
      const uint32_t cxy = xy_c[(int64_t)((v * U + u) * 4 + c) * g.cap + (k - pref[v])];
      int pu, pk;
      decode_key(best_cp[(int64_t)list * g.col_cap + k], u, pu, pk);
      int pv = 0, rem = pk;
      for (;; pv++) {  // locate the previous feature's (row, slot) from its col_idx
        const int nn = cnt_p[(pv * U + pu) * 4 + c];
        if (rem < nn) break;
        rem -= nn;
      }
      const uint32_t pxy = xy_p[(int64_t)((pv * U + pu) * 4 + c) * g.cap + rem];
      // p_match(u1p,v1p,-1,-1,-1,-1,u1c,v1c,-1,-1,-1,-1): src/myMatch.cpp:189
      const uint32_t m1f = __float_as_uint(-1.0f), m1i = 0xFFFFFFFFu;
      const uint32_t u1p = __float_as_uint((float)(pxy & 0xFFFFu)), v1p = __float_as_uint((float)(pxy >> 16));
      const uint32_t u1c = __float_as_uint((float)(cxy & 0xFFFFu)), v1c = __float_as_uint((float)(cxy >> 16));
      uint4* rec = dst + (int64_t)posn * 3;
      rec[0] = make_uint4(u1p, v1p, m1i, m1f);
      rec[1] = make_uint4(m1f, m1i, u1c, v1c);
      rec[2] = make_uint4(m1i, m1f, m1f, m1i);
    }
    running += chunk_total;
    __syncthreads();
  }
}

template <bool PRUNE>
static cudaError_t launch_k2_t(const Geom& g, const FrameArrays& fa, const int2* pairs, int n_pairs, uint32_t* best,
                               int rows_per_group, cudaStream_t st) {
  static bool attr_set = false;
  const int bytes = (int)sizeof(K2Smem<PRUNE>);
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k2_match<PRUNE>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const int ngroups = (g.V + rows_per_group - 1) / rows_per_group;
  dim3 grid((g.U - 1) * ngroups, 8, n_pairs);
  k2_match<PRUNE><<<grid, K2_THREADS, bytes, st>>>(g, fa, pairs, best, rows_per_group, ngroups);
  return cudaGetLastError();
}

cudaError_t launch_k2(const Geom& g, const FrameArrays& fa, const int2* pairs, int n_pairs, uint32_t* best,
                      int rows_per_group, int prune, cudaStream_t st) {
  return prune ? launch_k2_t<true>(g, fa, pairs, n_pairs, best, rows_per_group, st)
               : launch_k2_t<false>(g, fa, pairs, n_pairs, best, rows_per_group, st);
}

cudaError_t launch_k3(const Geom& g, const FrameArrays& fa, const int2* pairs, int n_pairs, const uint32_t* best,
                      uint8_t* flags, int32_t* list_cnt, void* out, int64_t out_stride_matches, int32_t max_matches,
                      int32_t* n_raw, int32_t* status, cudaStream_t st) {
  k3_flags<<<dim3(g.U, n_pairs), K3_THREADS, 0, st>>>(g, fa, pairs, best, flags, list_cnt);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  k3_emit<<<dim3(g.U * 4, n_pairs), K3_THREADS, 0, st>>>(g, fa, pairs, best, flags, list_cnt, (uint4*)out,
                                                         out_stride_matches, max_matches, n_raw, status);
  return cudaGetLastError();
}

}  // namespace vo
