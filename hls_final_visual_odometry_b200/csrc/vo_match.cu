// K2 / K3: bucketed circular descriptor matching (sm_100a).
//
// Replaces, bit-exactly, the reference's
//   fetch_col_bin   src/myMatch.cpp:17-68     (column lists, col_idx)
//   find_match      src/myMatch.cpp:70-119    (windowed SAD argmin, first-found-wins)
//   myMatching      src/myMatch.cpp:122-204   (both directions, cycle check, pixel dedupe, emit)
// in the ring-buffer-free formulation of SURVEY.md appendix A.
//
// K2 (k2_match): one CTA per (pair of origin u-bin columns, group of origin v-bin rows, class,
// direction, frame pair).  The CTA stages the target window of the OTHER frame -- up to 7 u-bin
// columns x the v-bin rows its origins can see -- into shared memory ordered (v_bin, u_bin, slot),
// every bin list padded to a multiple of 4.  A LANE owns an origin: the 32 lanes of a warp hold 32
// consecutive origins of one column list and walk the same candidates, so every shared-memory
// read is a broadcast and no per-origin work is replicated across lanes.  Per candidate and lane:
// a 2-instruction lower bound Q of the SAD (see desc_q) against the lane's running threshold;
// only candidates that could still win or tie pay the exact SAD (8 VABSDIFF4.U8.ACC) + packed
// window test, and the key min(SAD<<19 | u_rel<<16 | col_idx) is exactly the reference's
// "strict <, first candidate in (u_bin asc, col_idx asc) order wins".  The threshold is seeded
// from the bins the origins themselves sit in (zero-flow prior).  Pruning is conservative, so the
// result is identical to evaluating every SAD (PRUNE=false does that, for the tests).
// The bound has two words (one per descriptor half, desc_q).  A warp whose origins found exact partners
// (bit-identical patches: rendered or noise-free frames) filters with the first word only; a warp whose
// seeds carry non-zero bounds (sensor noise) filters with both words -- twice the shared-memory reads, six
// times fewer false hits -- and keeps its poorly seeded origins on the lane path.  Queued groups are
// re-tested candidate by candidate with both words against the owner's CURRENT bound before an exact SAD
// is spent.  Origins without a usable seed go to a warp-per-origin pass whose survivors are compacted
// through the warp's queue so that exact SADs run 32 at a time (tools/noise_probe.py: 1.3-1.7x over the
// one-word filter on noisy frames, -4 % on noise-free ones).
// SAD is not a contraction: no tensor cores here by design.
//
// K3 (k3_flags, k3_emit): closed-cycle test, same-pixel dedupe (lowest class wins, via the twin
// links written by K1), ordered emit of p_match records in (u_bin, class, col_idx) order.
#include "vo_common.cuh"
#include <cstdio>

namespace vo {
// -DVO_DEBUG_BOUNDS: every shared-memory index K2 computes is checked and a violation traps the launch (the pool's
// compute-sanitizer is closed, profiles/r2_sanitizer_refused.txt; tools/bounds_check.sh runs the parity and stress suites
// against such a build)
#ifdef VO_DEBUG_BOUNDS
#define K2CHECK(cond) do { if (!(cond)) { printf("K2CHECK failed: %s (line %d, block %d,%d,%d thread %d)\n", #cond, __LINE__, blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x); __trap(); } } while (0)
#else
#define K2CHECK(cond)
#endif
#ifdef VO_K2_STATS
__device__ unsigned long long g_k2_stats[8];
#define K2STAT(i, n) atomicAdd(&g_k2_stats[i], (unsigned long long)(n))
#else
#define K2STAT(i, n)
#endif

constexpr int kMaxV = 64;       // v-bin rows supported by the shared-memory tables
constexpr int kOC = 2;          // origin u-bin columns per CTA (they share the staged window)
constexpr int kMaxCols = kOC + 5;  // staged target columns: [ub0-3, ub_last+3) clipped
constexpr int kSegW = kMaxCols + 1;
constexpr int kCandCap = 1920;  // staged candidates per chunk; >= 6*300 so a row of a one-column window always fits
constexpr int kMaxWarps = 16;
constexpr int kTwoWordVotes = 3;  // a warp filters with both signature words when more than 3/4 of its searching lanes hold a non-zero bound
constexpr int kQueue = 128;     // deferred exact evaluations per warp: < 32 left over + up to kAppend per lane and pass
#ifndef VO_K2_APPEND
#define VO_K2_APPEND 3
#endif
#ifndef VO_K2_FLOOD
#define VO_K2_FLOOD 20
#endif
constexpr int kAppend = VO_K2_APPEND;  // queue entries a lane contributes per compaction pass (1..3: the count travels in two ballots)
constexpr int kFlood = VO_K2_FLOOD;    // a lane with at least this many filter hits in a chunk of 32 groups is expanded by the whole warp
constexpr int kBig = 1 << 20;
static_assert(kCandCap <= 2048, "k2 hard pass packs the staged position into 11 bits");
static_assert(kAppend >= 1 && kAppend <= 3 && 31 + 32 * kAppend <= kQueue, "a compaction pass must fit the queue");
constexpr int kHardCap = 1024;  // origins per chunk handed to the warp-per-origin pass
constexpr uint32_t kSkip = 0xFFFFFFFEu;  // best[] between the chunks of a hinted search: nobody points at this origin (never a real key: SAD <= 8160)
constexpr int kSeedGood = 128;  // an origin whose seed SAD is above this has no partner worth a lane-per-origin walk: warp-per-origin pass
constexpr int kHitBudget = 4;   // default of vo_config.match_hit_budget: filter hits (groups of 4) a lane may queue before its origin is handed over

struct K2Tables {
  int32_t cntT[kMaxV][kSegW];         // target counts (row, staged column)
  int32_t segOff[kMaxV][kSegW];       // this chunk: staged position of bin (row, column); [ncol] = end of the row
  int32_t colT[kMaxCols][kMaxV + 1];  // col_idx prefix of each target column
  int32_t rowTot[kMaxV];              // staged (padded) candidates per row
  int32_t rowStart[kMaxV + 1];        // this chunk: staged position of the first candidate of a row
  int32_t cntO[kOC][kMaxV];
  int32_t prefO[kOC][kMaxV + 1];
  int32_t lagEnd[kOC][kMaxV];         // origins o < lagEnd[v] of row v may carry a lagging v_bin
};

struct __align__(128) K2Smem {
  K2Tables t;
  uint32_t sbest[kMaxWarps][32];      // running best key of the 32 origins of a warp's task
  uint32_t queue[kMaxWarps][kQueue];  // deferred exact evaluations: owner lane << 16 | candidate group
  int32_t hard_n;                     // origins handed to the warp-per-origin pass in this chunk
  uint16_t hard[kHardCap];            // origin column (bit 15) | col_idx
  __align__(16) uint4 lo[kCandCap];   // descriptor bytes 0..15
  uint4 hi[kCandCap];                 // descriptor bytes 16..31
  uint2 xr[kCandCap];                 // {xy + 0x80008000, u_rel<<16 | col_idx}
  __align__(16) uint32_t q[kCandCap];  // signature word 0: 4 coded group sums of bytes 0..15 (first-level filter); read four at a time
  __align__(16) uint32_t q1[kCandCap]; // signature word 1: bytes 16..31 (second level: Q0 + Q1)
};
static_assert(sizeof(K2Smem) <= 113 * 1024, "two K2 CTAs per SM need <= 113 KB each (227 KB minus the per-CTA reserve)");

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

__device__ __forceinline__ int sadq(uint32_t a, uint32_t x) { return (int)sad4(a, x, 0u); }

// key of one staged candidate for an origin (a, b, window lo_p/hi_p), kNone if outside the window.
// packed window test: both 16-bit fields carry a guard bit (bit 15); DESIGN.md "window test"
__device__ __forceinline__ uint32_t cand_key(const K2Smem& S, int p, const uint4& a, const uint4& b, uint32_t lo_p,
                                             uint32_t hi_p) {
  K2CHECK(p >= 0 && p < kCandCap);
  const uint2 xr = S.xr[p];
  const uint32_t sad = sad32(a, b, S.lo[p], S.hi[p]);
  const uint32_t t1 = xr.x - lo_p, t2 = xr.x - hi_p;
  const uint32_t key = sad * 0x80000u + xr.y;
  return (((t1 & ~t2) & 0x80008000u) == 0x80008000u) ? key : kNone;
}

// v_buffer_idx of the reference (src/myMatch.cpp:153-158): starts at 0, bumped at most once per
// origin => vb(k) = min(k+1, v+1, min over non-empty rows b<v of b+2+k-prefix[b+1]).
__device__ __forceinline__ int origin_vb(const K2Tables& t, int cc, int col_idx, int v) {
  int vb = min(col_idx + 1, v + 1);
  for (int b = 0; b < v; b++)
    if (t.cntO[cc][b] > 0) vb = min(vb, b + 2 + col_idx - t.prefO[cc][b + 1]);
  return vb;
}

__device__ __forceinline__ int pad4(int n) { return (n + 3) & ~3; }

// Matches the origins of columns [ub0, ub0+noc) x rows of group `grp`.  Returns false (having
// written nothing) when a row of the shared (noc+5)-column window does not fit the staging buffer;
// the caller then runs the columns one at a time (a 6-column row always fits).
template <bool PRUNE, int THREADS>
__device__ __forceinline__ bool k2_columns(K2Smem& S, const Geom& g, const FrameArrays& fa, int from_slot, int to_slot,
                                           uint32_t* __restrict__ best_dir, int c, int ub0, int noc, int grp,
                                           int rows_per_group, int rad_v, int reach, int hit_budget, bool hinted,
                                           uint32_t* __restrict__ clear_dir) {
  constexpr int NW = THREADS / 32;
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  K2Tables& t = S.t;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int U = g.U, V = g.V;
  const int ulo = max(ub0 - 3, 0), uhi = min(ub0 + noc - 1 + 3, U - 1);  // staged target columns [ulo, uhi)
  const int ncol = uhi - ulo;

  const int32_t* cnt_to = fa.cnt + (int64_t)to_slot * fa.cnt_stride;
  const int32_t* cnt_from = fa.cnt + (int64_t)from_slot * fa.cnt_stride;
  __syncthreads();  // tables may still be in use by a previous call
  for (int idx = tid; idx < V * kSegW; idx += THREADS) {
    const int v = idx / kSegW, j = idx - v * kSegW;
    t.cntT[v][j] = (j < ncol) ? cnt_to[(v * U + ulo + j) * 4 + c] : 0;
  }
  for (int idx = tid; idx < V * kOC; idx += THREADS) {
    const int cc = idx / V, v = idx - cc * V;
    t.cntO[cc][v] = (cc < noc) ? cnt_from[(v * U + ub0 + cc) * 4 + c] : 0;
  }
  __syncthreads();
  if (tid < kMaxCols) {
    int acc = 0;
    t.colT[tid][0] = 0;
    for (int v = 0; v < V; v++) { acc += t.cntT[v][tid]; t.colT[tid][v + 1] = acc; }
  } else if (tid >= 32 && tid < 32 + kOC) {
    const int cc = tid - 32;
    int acc = 0;
    t.prefO[cc][0] = 0;
    for (int v = 0; v < V; v++) { acc += t.cntO[cc][v]; t.prefO[cc][v + 1] = acc; }
  } else if (tid >= 64 && tid < 64 + V) {
    const int v = tid - 64;
    int acc = 0;
#pragma unroll
    for (int j = 0; j < kMaxCols; j++) acc += pad4(t.cntT[v][j]);
    t.rowTot[v] = acc;
  }
  __syncthreads();
  for (int idx = tid; idx < V * kOC; idx += THREADS) {  // first origin of row v whose v_bin is the nominal v+1
    const int cc = idx / V, v = idx - cc * V;
    int le = v;
    for (int b = 0; b < v; b++)
      if (t.cntO[cc][b] > 0) le = max(le, t.prefO[cc][b + 1] + v - 1 - b);
    t.lagEnd[cc][v] = le;
  }

  if (clear_dir != nullptr && grp == 0) {
    // mode 1 (prev -> curr in full): the lists of the TARGET frame in this CTA's own columns are the origins of the hinted
    // launch that follows; reset their best[] entries here so that k2_hint_scatter can atomicMin into them
    for (int cc = 0; cc < noc; cc++) {
      const int j = ub0 + cc - ulo;
      uint32_t* bc = clear_dir + (int64_t)((ub0 + cc) * 4 + c) * g.col_cap;
      for (int k = tid, n = t.colT[j][V]; k < n; k += THREADS) bc[k] = kNone;
    }
  }
  const int r0 = grp * rows_per_group, r1 = min(r0 + rows_per_group, V);
  int oB0 = 0, nT0 = 0, ntasks = 0, row_lo = V;
#pragma unroll
  for (int cc = 0; cc < kOC; cc++) {
    const int ob = t.prefO[cc][r0];
    const int n = t.prefO[cc][r1] - ob;
    if (cc == 0) { oB0 = ob; nT0 = (n + 31) >> 5; }
    ntasks += (n + 31) >> 5;
    if (n > 0) {  // rows any origin of this group can see: vb is non-decreasing along the column list
      int v_first = r0;
      while (t.prefO[cc][v_first + 1] <= ob) v_first++;
      row_lo = min(row_lo, max(max(origin_vb(t, cc, ob, v_first) - 3, v_first - reach), 0));
    }
  }
  if (ntasks == 0) return true;
  // exclusive: src/myMatch.cpp:93,99.  `reach` (>= V on flow edges: no restriction) narrows the rows of a
  // left<->right edge to those that can hold a candidate inside its +-rad_v window: a bin row spans 45 block
  // rows, so |dv| <= rad_v <= 16 keeps the candidate's row within +-1 of the origin's own row
  const int row_hi = min(min(r1 + 3, V - 1), r1 + reach);
  if (noc > 1)
    for (int v = row_lo; v < row_hi; v++)
      if (t.rowTot[v] > kCandCap) return false;

  const uint32_t* xy_from = fa.xy + (int64_t)from_slot * fa.rec_stride;
  const uint4* desc_from = fa.desc + (int64_t)from_slot * fa.rec_stride * 2;
  const uint2* q_from = fa.q + (int64_t)from_slot * fa.rec_stride;
  const uint32_t* xy_to = fa.xy + (int64_t)to_slot * fa.rec_stride;
  const uint4* desc_to = fa.desc + (int64_t)to_slot * fa.rec_stride * 2;
  const uint2* q_to = fa.q + (int64_t)to_slot * fa.rec_stride;
  const uint4* Q4 = reinterpret_cast<const uint4*>(S.q);
  const uint4* Q14 = reinterpret_cast<const uint4*>(S.q1);
  uint32_t* sbest = S.sbest[warp];
  uint32_t* queue = S.queue[warp];
  const uint32_t lt_mask = (1u << lane) - 1u, lane16 = (uint32_t)lane << 16;

  bool first_chunk = true;
  for (int ra = row_lo;;) {
    // ---- chunk = as many whole rows as fit the staging buffer ------------------------------
    int rb = ra, staged = 0;
    while (rb < row_hi && staged + t.rowTot[rb] <= kCandCap) { staged += t.rowTot[rb]; rb++; }
    // hints (see k2_hint_scatter) are validated against the candidate's record in global memory, so they work whether or
    // not the window is staged in one piece; an origin nobody points at is parked as kSkip until the last chunk
    const bool use_hint = hinted;
    const bool last_chunk = rb >= row_hi;
    __syncthreads();  // previous chunk fully consumed (and lagEnd visible)
    if (tid == THREADS - 1) S.hard_n = 0;
    if (tid <= rb - ra) {
      const int v = ra + tid;
      int acc = 0;
      for (int x = ra; x < v; x++) acc += t.rowTot[x];
      t.rowStart[v] = acc;
      if (v < rb) {
#pragma unroll
        for (int j = 0; j < kSegW; j++) { t.segOff[v][j] = acc; acc += pad4(t.cntT[v][j]); }
      }
    }
    __syncthreads();
    // ---- stage: warp per (row, column) bin list, coalesced 16-byte copies; every list is padded
    //      to a multiple of 4 with dummies that can never pass the window test -------------------
    for (int b = warp; b < (rb - ra) * ncol; b += NW) {
      const int v = ra + b / ncol, j = b % ncol;
      const int n = t.cntT[v][j], npad = pad4(n);
      const int dst = t.segOff[v][j];
      const int64_t src = (int64_t)((v * U + ulo + j) * 4 + c) * g.cap;
      const uint32_t rank0 = ((uint32_t)j << 16) | (uint32_t)t.colT[j][v];
      K2CHECK(dst >= 0 && dst + npad <= kCandCap && n <= g.cap);
      for (int k = lane; k < npad; k += 32) {
        if (k < n) {
          const uint4 d0 = desc_to[(src + k) * 2], d1 = desc_to[(src + k) * 2 + 1];
          S.lo[dst + k] = d0;
          S.hi[dst + k] = d1;
          S.xr[dst + k] = make_uint2(xy_to[src + k] + 0x80008000u, rank0 + (uint32_t)k);
          const uint2 qv = q_to[src + k];
          S.q[dst + k] = qv.x;
          S.q1[dst + k] = qv.y;
        } else {
          S.lo[dst + k] = make_uint4(0, 0, 0, 0);
          S.hi[dst + k] = make_uint4(0, 0, 0, 0);
          S.xr[dst + k] = make_uint2(kNone, kNone);
          S.q[dst + k] = kNone;
          S.q1[dst + k] = kNone;
        }
      }
    }
    __syncthreads();

    // ---- tasks: 32 consecutive origins of one column list, one per lane.  Every lane walks the
    //      same candidates (shared-memory broadcast); its own search rows are [vlo, vhi). ----------
    for (int task = warp; task < ntasks; task += NW) {
      const int cc = (kOC > 1 && task >= nT0) ? 1 : 0;
      const int tb = task - (cc ? nT0 : 0);
      const int ub = ub0 + cc;
      const int j0 = max(ub - 3, 0) - ulo, j1 = min(ub + 3, U - 1) - ulo, jown = ub - ulo;  // src/myMatch.cpp:90-91,97
      const int o = (cc ? t.prefO[cc][r0] : oB0) + tb * 32 + lane;
      const bool valid = o < t.prefO[cc][r1];
      uint32_t* bo = best_dir + (int64_t)(ub * 4 + c) * g.col_cap;

      int v = r0, vlo = kBig, vhi = 0;
      uint32_t oxy = 0;
      uint4 a = make_uint4(0, 0, 0, 0), b = a;
      uint32_t qa = 0, qb = 0;
      uint32_t bestk = kNone;
      uint32_t hint = kNone;
      bool full = valid;  // this lane runs an unseeded search (seed pass below)
      bool skip = false;  // hinted direction: no feature of the other frame points here, the cycle cannot close
      int lo_full = 0, hi_full = 0;  // the origin's rows before clipping to the staged chunk
      if (valid) {
        while (t.prefO[cc][v + 1] <= o) v++;
        const int64_t rec = (int64_t)((v * U + ub) * 4 + c) * g.cap + (o - t.prefO[cc][v]);
        oxy = __ldg(&xy_from[rec]);
        a = __ldg(&desc_from[2 * rec]);
        b = __ldg(&desc_from[2 * rec + 1]);
        const uint2 qo = __ldg(&q_from[rec]);
        qa = qo.x; qb = qo.y;
        // lagging v_bin (src/myMatch.cpp:153-158): only the first origins after empty bins
        const int vb = (o >= t.lagEnd[cc][v]) ? v + 1 : origin_vb(t, cc, o, v);
        const int lo_r = max(max(max(vb - 3, 0), ra), v - reach);  // src/myMatch.cpp:92-93,99
        const int hi_r = min(min(min(vb + 3, V - 1), rb), v + reach + 1);
        if (lo_r < hi_r) { vlo = lo_r; vhi = hi_r; }
        lo_full = max(max(vb - 3, 0), v - reach);
        hi_full = min(min(vb + 3, V - 1), v + reach + 1);
        if (use_hint && first_chunk) {
          hint = bo[o];
          full = false;
          skip = hint == kNone;
        } else if (!first_chunk) {
          bestk = bo[o];
          skip = bestk == kSkip;
          if (skip) bestk = kNone;
          if (bestk != kNone) bestk += (uint32_t)j0 << 16;
          if (use_hint) full = bestk == kNone;  // a validated hint (or better) is already the bound
        }
        if (skip) { vlo = kBig; vhi = 0; full = false; }
      }
      const int rlo = __reduce_min_sync(FULL, vlo), rhi = __reduce_max_sync(FULL, vhi);
      const int ou = (int)(oxy & 0xFFFFu), ov = (int)(oxy >> 16);
      const uint32_t lo_p = (uint32_t)max(ou - g.rad_u, 0) | ((uint32_t)max(ov - rad_v, 0) << 16);  // src/myMatch.cpp:84-87,101
      const uint32_t hi_p = (uint32_t)(ou + g.rad_u + 1) | ((uint32_t)(ov + rad_v + 1) << 16);

      if (PRUNE && use_hint && first_chunk && valid && !skip) {
        // the hinted candidate must be one this origin may legally take (bin ranges are not symmetric between the two
        // directions): locate its record in global memory and evaluate it; anything unexpected -> plain search
        const int jr = (int)((hint >> 16) & 7u), kb = (int)(hint & 0xFFFFu), j = j0 + jr;
        uint32_t k = kNone;
        if (jr < 6 && j < j1) {
          for (int r = lo_full; r < hi_full; r++) {
            if (kb >= t.colT[j][r] && kb < t.colT[j][r + 1]) {
              const int64_t rec = (int64_t)((r * U + ulo + j) * 4 + c) * g.cap + (kb - t.colT[j][r]);
              const uint32_t x = __ldg(&xy_to[rec]) + 0x80008000u;
              const uint32_t sad = sad32(a, b, __ldg(&desc_to[2 * rec]), __ldg(&desc_to[2 * rec + 1]));
              const uint32_t t1 = x - lo_p, t2 = x - hi_p;
              if (((t1 & ~t2) & 0x80008000u) == 0x80008000u) k = sad * 0x80000u + (((uint32_t)j << 16) | (uint32_t)kb);
              break;
            }
          }
        }
        if (k != kNone && k == hint + ((uint32_t)j0 << 16)) bestk = k;
        else full = true;
      }
      if (PRUNE && __any_sync(FULL, full && vlo < vhi)) {
        // seed: the candidate with the smallest lower bound among the bins the origins themselves
        // sit in (zero-flow prior); any legal candidate gives a valid bound
        const int smin = max(__reduce_min_sync(FULL, valid ? v : kBig), ra);
        const int smax = min(__reduce_max_sync(FULL, valid ? v : -1), rb - 1);
        uint32_t k1 = kNone;
        int sa = jown, sb = jown + 1;  // seed bins of the own rows
        if (reach >= kMaxV) {  // flow edge: the +-135-row window admits practically every candidate of the own bin
          for (int r = smin; r <= smax; r++) {
            const int p1 = t.segOff[r][jown + 1];
            for (int p = t.segOff[r][jown]; p < p1; p += 4) {  // bin lists are padded to 4
              const uint4 qq = Q4[p >> 2], q1 = Q14[p >> 2];
              const uint32_t ka = sad4(qb, q1.x, sad4(qa, qq.x, 0u)) * 0x100000u + (uint32_t)p;
              const uint32_t kb = sad4(qb, q1.y, sad4(qa, qq.y, 0u)) * 0x100000u + (uint32_t)(p + 1);
              const uint32_t kc = sad4(qb, q1.z, sad4(qa, qq.z, 0u)) * 0x100000u + (uint32_t)(p + 2);
              const uint32_t kd = sad4(qb, q1.w, sad4(qa, qq.w, 0u)) * 0x100000u + (uint32_t)(p + 3);
              k1 = min(k1, min(min(ka, kb), min(kc, kd)));
            }
          }
        } else {  // narrow (left<->right) edge: only candidates inside the +-rad_v rows can seed; a disparity may cross
                  // into the neighbouring bins, so the own row's bins jown-1 .. jown+1 are searched
          const int ja = max(jown - 1, j0), jb = min(jown + 2, j1);
          sa = ja; sb = jb;
          for (int r = smin; r <= smax; r++) {
            const int p1 = t.segOff[r][jb];
            for (int p = t.segOff[r][ja]; p < p1; p++) {
              const uint32_t x = S.xr[p].x;
              const uint32_t t1 = x - lo_p, t2 = x - hi_p;
              const uint32_t kq = sad4(qb, S.q1[p], sad4(qa, S.q[p], 0u)) * 0x100000u + (uint32_t)p;
              k1 = min(k1, (((t1 & ~t2) & 0x80008000u) == 0x80008000u) ? kq : kNone);
            }
          }
        }
        if (k1 != kNone && vlo < vhi && full) {
          const int p = (int)(k1 & 0xFFFFFu);
          if (p >= t.rowStart[vlo] && p < t.rowStart[vhi]) bestk = min(bestk, cand_key(S, p, a, b, lo_p, hi_p));
        }
#ifdef VO_K2_SEEDX
        if (reach >= kMaxV) {  // second look at the own bins: every candidate the first seed's SAD cannot exclude, exactly
          int tqs = (bestk == kNone) ? 0x7FFFFFFF : (int)(bestk >> 19);
          const bool me = full && vlo < vhi;
          const int pa = me ? t.rowStart[vlo] : 0, pb = me ? t.rowStart[vhi] : 0;
          for (int r = smin; r <= smax; r++) {
            const int p1 = t.segOff[r][jown + 1];
            for (int p = t.segOff[r][jown]; p < p1; p++) {
              const int q = (int)sad4(qb, S.q1[p], sad4(qa, S.q[p], 0u));
              if (p >= pa && p < pb && q <= tqs) {
                bestk = min(bestk, cand_key(S, p, a, b, lo_p, hi_p));
                tqs = (bestk == kNone) ? 0x7FFFFFFF : (int)(bestk >> 19);
              }
            }
          }
        }
#endif
      }
      sbest[lane] = bestk;
      __syncwarp();
      // tq: a candidate can still win (or tie) only if its lower bound Q <= best SAD so far
      int tq = (bestk == kNone) ? 0x7FFFFFFF : (int)(bestk >> 19);
      int qn = 0;  // queued groups (warp-uniform)
      int hits = 0;
      bool handed = false;
      // An origin the filter cannot prune (no near-exact match: every candidate may still win) would
      // flood the queue; it leaves the lane-per-origin walk and is matched by a whole warp after the
      // tasks of this chunk.  It takes the best key found so far along (through best[]) as its starting bound.
      auto hand_over = [&]() {
        const int hs = atomicAdd(&S.hard_n, 1);
        hits = -kBig;  // once per task (if the list is full the origin stays on the queue path)
        if (hs < kHardCap) {
          S.hard[hs] = (uint16_t)((cc << 15) | o);
          bo[o] = (bestk == kNone) ? kNone : bestk - ((uint32_t)j0 << 16);
          vlo = kBig; vhi = 0;
          handed = true;
          return true;
        }
        return false;
      };
      K2STAT(0, valid ? 1 : 0);
      K2STAT(2, (valid && vlo < vhi && bestk == kNone) ? 1 : 0);
      K2STAT(3, (valid && vlo < vhi && bestk != kNone && (bestk >> 19) == 0) ? 1 : 0);
      // Filter width, chosen per warp: on clean frames most origins find an exact partner (bound 0) and the first
      // signature word alone prunes everything; once sensor noise leaves the seeds with non-zero bounds the filter
      // tests both words (twice the shared-memory reads, ~6x fewer false hits) and can afford to keep poorly seeded
      // origins on the lane path.
      const uint32_t searching = __ballot_sync(FULL, vlo < vhi);
      const bool two = PRUNE && __popc(__ballot_sync(FULL, vlo < vhi && tq > 0)) * 4 > __popc(searching) * kTwoWordVotes;
      K2STAT(1, (valid && two) ? 1 : 0);
      if (PRUNE && vlo < vhi && (two ? bestk == kNone : tq > kSeedGood)) hand_over();  // no seed or a poor one: nearly every group would hit

      // exact SAD + window test of up to 32 queued (origin, group of 4 candidates) entries, one per
      // lane; the origin's record comes from its owner lane by shuffle, the result goes back
      // through a shared-memory atomicMin.  Every lane then refreshes its bound.
      auto flush = [&]() {
        const int n = min(qn, 32);
        qn -= n;
        const bool on = lane < n;
        K2CHECK(qn >= 0 && qn + n <= kQueue);
        const uint32_t e = on ? queue[qn + lane] : 0u;
        const int owner = (int)(e >> 16), p = (int)(e & 0xFFFFu) << 2;
        K2CHECK(!on || (owner < 32 && p + 3 < kCandCap));
        uint4 oa, ob;
        oa.x = __shfl_sync(FULL, a.x, owner); oa.y = __shfl_sync(FULL, a.y, owner);
        oa.z = __shfl_sync(FULL, a.z, owner); oa.w = __shfl_sync(FULL, a.w, owner);
        ob.x = __shfl_sync(FULL, b.x, owner); ob.y = __shfl_sync(FULL, b.y, owner);
        ob.z = __shfl_sync(FULL, b.z, owner); ob.w = __shfl_sync(FULL, b.w, owner);
        const uint32_t olo = __shfl_sync(FULL, lo_p, owner), ohi = __shfl_sync(FULL, hi_p, owner);
        const uint32_t oq = __shfl_sync(FULL, qa, owner), oq1 = __shfl_sync(FULL, qb, owner);
        const int otq = __shfl_sync(FULL, tq, owner);  // the owner's bound NOW (it may have dropped since the group was queued)
        if (on) {
          // a group was queued because ONE of its four candidates may win: re-test each against the bound first
          const uint4 qq = Q4[p >> 2], q1 = Q14[p >> 2];
          uint32_t k = kNone;
          if ((int)sad4(oq1, q1.x, sad4(oq, qq.x, 0u)) <= otq) k = cand_key(S, p, oa, ob, olo, ohi);
          if ((int)sad4(oq1, q1.y, sad4(oq, qq.y, 0u)) <= otq) k = min(k, cand_key(S, p + 1, oa, ob, olo, ohi));
          if ((int)sad4(oq1, q1.z, sad4(oq, qq.z, 0u)) <= otq) k = min(k, cand_key(S, p + 2, oa, ob, olo, ohi));
          if ((int)sad4(oq1, q1.w, sad4(oq, qq.w, 0u)) <= otq) k = min(k, cand_key(S, p + 3, oa, ob, olo, ohi));
          if (k != kNone) atomicMin(&sbest[owner], k);
        }
        __syncwarp();
        bestk = sbest[lane];
        tq = (bestk == kNone) ? 0x7FFFFFFF : (int)(bestk >> 19);
      };

      // The set bits of lane L's hit mask, evaluated by the whole warp at once: lane j takes the group of bit 31 - j
      // (group g0 + j), re-tests its four candidates against L's bound and evaluates the survivors exactly.
      auto expand = [&](int L, uint32_t mine, int g0) {
        const uint32_t mL = __shfl_sync(FULL, mine, L);
        uint4 oa, ob;
        oa.x = __shfl_sync(FULL, a.x, L); oa.y = __shfl_sync(FULL, a.y, L);
        oa.z = __shfl_sync(FULL, a.z, L); oa.w = __shfl_sync(FULL, a.w, L);
        ob.x = __shfl_sync(FULL, b.x, L); ob.y = __shfl_sync(FULL, b.y, L);
        ob.z = __shfl_sync(FULL, b.z, L); ob.w = __shfl_sync(FULL, b.w, L);
        const uint32_t olo = __shfl_sync(FULL, lo_p, L), ohi = __shfl_sync(FULL, hi_p, L);
        const uint32_t oq = __shfl_sync(FULL, qa, L), oq1 = __shfl_sync(FULL, qb, L);
        const int otq = __shfl_sync(FULL, tq, L);
        uint32_t k = kNone;
        if ((mL << lane) & 0x80000000u) {
          const int p = (g0 + lane) << 2;
          K2CHECK(p >= 0 && p + 3 < kCandCap);
          const uint4 qq = Q4[p >> 2], q1 = Q14[p >> 2];
          if ((int)sad4(oq1, q1.x, sad4(oq, qq.x, 0u)) <= otq) k = cand_key(S, p, oa, ob, olo, ohi);
          if ((int)sad4(oq1, q1.y, sad4(oq, qq.y, 0u)) <= otq) k = min(k, cand_key(S, p + 1, oa, ob, olo, ohi));
          if ((int)sad4(oq1, q1.z, sad4(oq, qq.z, 0u)) <= otq) k = min(k, cand_key(S, p + 2, oa, ob, olo, ohi));
          if ((int)sad4(oq1, q1.w, sad4(oq, qq.w, 0u)) <= otq) k = min(k, cand_key(S, p + 3, oa, ob, olo, ohi));
        }
        k = __reduce_min_sync(FULL, k);
        if (lane == L && k < bestk) {  // (queued entries of L still pending add to sbest[L] by atomicMin)
          atomicMin(&sbest[lane], k);
          bestk = min(bestk, k);
          tq = (int)(bestk >> 19);
        }
        __syncwarp();
      };

      // rows in the order own row, then the rest: the partner of a small flow sits in the own row, so bounds tighten first
      const int rown = min(max(__shfl_sync(FULL, v, 0), rlo), rhi - 1);
      for (int ri = rlo - 1; ri < rhi; ri++) {
        if (ri == rown) continue;
        const int r = (ri < rlo) ? rown : ri;
        bool act = (r >= vlo) && (r < vhi);
        int tqe = act ? tq : -1;  // tq while this lane searches the current row, else -1 (never passes)
        int pg = t.segOff[r][j0] >> 2;            // candidate groups of 4 (every bin list is padded to 4)
        const int pg1 = t.segOff[r][j1] >> 2;
        // Chunks of up to 32 groups: every lane first collects a private bit mask of the groups that may
        // still hold a winner (no votes, no branches in the filter loop), then the warp compacts the set
        // bits into the exact-evaluation queue, one entry per lane and round.
        while (pg < pg1) {
          const int nb = min(32, pg1 - pg);
          K2CHECK(pg >= 0 && (pg + nb) * 4 <= kCandCap);
          uint32_t m = 0;  // MISS bits while filtering, newest group in bit 0
          // bias + Q has its sign bit set exactly when Q exceeds the lane's bound (tqe = -1: always; no bound yet: never), so a
          // group of four candidates misses iff the AND of its four sums is negative: 2 LOP3 + 1 SHF per group on top of
          // the VABSDIFF4s, no minimum and no subtraction
          const uint32_t bias = 0x7FFFFFFFu - (uint32_t)tqe;
          if (PRUNE && two) {
            const uint4* qp = Q4 + pg;
            const uint4* rp = Q14 + pg;
            auto gmiss = [&](const uint4& x, const uint4& y) {
              return (sad4(qb, y.x, sad4(qa, x.x, bias)) & sad4(qb, y.y, sad4(qa, x.y, bias))) &
                     (sad4(qb, y.z, sad4(qa, x.z, bias)) & sad4(qb, y.w, sad4(qa, x.w, bias)));
            };
            int i = 0;
#pragma unroll 1
            for (; i + 3 < nb; i += 4) {
              const uint32_t ma = gmiss(qp[i], rp[i]), mb = gmiss(qp[i + 1], rp[i + 1]);
              const uint32_t mc = gmiss(qp[i + 2], rp[i + 2]), md = gmiss(qp[i + 3], rp[i + 3]);
              m = __funnelshift_l(ma, m, 1);
              m = __funnelshift_l(mb, m, 1);
              m = __funnelshift_l(mc, m, 1);
              m = __funnelshift_l(md, m, 1);
            }
#pragma unroll 1
            for (; i < nb; i++) m = __funnelshift_l(gmiss(qp[i], rp[i]), m, 1);
            m = ~m;
          } else if (PRUNE) {
            const uint4* qp = Q4 + pg;  // one uint4 = the signatures of one group of 4 candidates
            auto gmiss1 = [&](const uint4& x) {
              return (sad4(qa, x.x, bias) & sad4(qa, x.y, bias)) & (sad4(qa, x.z, bias) & sad4(qa, x.w, bias));
            };
            int i = 0;
#pragma unroll 1
            for (; i + 3 < nb; i += 4) {
              const uint32_t ma = gmiss1(qp[i]), mb = gmiss1(qp[i + 1]), mc = gmiss1(qp[i + 2]), md = gmiss1(qp[i + 3]);
              m = __funnelshift_l(ma, m, 1);  // shifts in the sign: 1 = lower bound above the threshold
              m = __funnelshift_l(mb, m, 1);
              m = __funnelshift_l(mc, m, 1);
              m = __funnelshift_l(md, m, 1);
            }
#pragma unroll 1
            for (; i < nb; i++) m = __funnelshift_l(gmiss1(qp[i]), m, 1);
            m = ~m;
          } else {
            m = act ? kNone : 0u;
          }
          m &= kNone >> (32 - nb);  // bit b <-> group pg + nb-1-b
          if (PRUNE) {
            hits += __popc(m);
            K2STAT(4, __popc(m));
            if (!two && hits > hit_budget && hand_over()) { m = 0u; act = false; tqe = -1; K2STAT(5, 1); }
          }
          const uint32_t gfirst = lane16 + (uint32_t)(pg + nb - 32);  // owner lane << 16 + group of bit 31 (mod 2^32)
          uint32_t anyb = __ballot_sync(FULL, m != 0u);
          if (anyb != 0u) {
            // (1) a lane the bound hardly prunes (no partner in sight yet, or none at all) would feed the queue for many
            //     rounds while 31 lanes wait: the whole warp takes its mask instead, lane j its group j
            int left = __popc(m);
            uint32_t flood = __ballot_sync(FULL, left >= kFlood);
            while (flood != 0u) {
              const int L = __ffs((int)flood) - 1;
              flood &= flood - 1u;
              expand(L, m, pg + nb - 32);
              if (lane == L) { m = 0u; left = 0; }
            }
            // (2) the other lanes append up to kAppend groups per pass at positions given by two ballots (the count is a
            //     2-bit number), then the queue is drained 32 entries at a time
            anyb = __ballot_sync(FULL, m != 0u);
            while (anyb != 0u) {
              const int cn = min(left, kAppend);
              const uint32_t b0 = __ballot_sync(FULL, cn & 1), b1 = __ballot_sync(FULL, cn & 2);
              const int pos = qn + __popc(b0 & lt_mask) + 2 * __popc(b1 & lt_mask);
#pragma unroll
              for (int k = 0; k < kAppend; k++) {
                if (k < cn) {
                  const int z = __clz((int)m);
                  K2CHECK(pos + k >= 0 && pos + k < kQueue && m != 0u);
                  queue[pos + k] = gfirst + (uint32_t)z;
                  m &= ~(0x80000000u >> z);
                }
              }
              left -= cn;
              qn += __popc(b0) + 2 * __popc(b1);
              __syncwarp();
              while (qn >= 32) flush();
              anyb = __ballot_sync(FULL, m != 0u);
            }
            tqe = act ? tq : -1;
          }
          pg += nb;
        }
      }
      __syncwarp();
      while (qn > 0) flush();
      if (valid && !handed) bo[o] = (skip && !last_chunk) ? kSkip : (bestk == kNone) ? kNone : bestk - ((uint32_t)j0 << 16);
      __syncwarp();  // sbest/queue are reused by this warp's next task
    }
    if (PRUNE) {
      // ---- handed-over origins: one warp per origin, lanes stride its candidates ----------------------
      __syncthreads();
      const int nh = min(S.hard_n, kHardCap);
      for (int h = warp; h < nh; h += NW) {
        const uint32_t e = S.hard[h];
        const int cc = (int)(e >> 15), o = (int)(e & 0x7FFFu);
        const int ub = ub0 + cc;
        const int j0 = max(ub - 3, 0) - ulo, j1 = min(ub + 3, U - 1) - ulo;
        uint32_t* bo = best_dir + (int64_t)(ub * 4 + c) * g.col_cap;
        int v = r0;
        while (t.prefO[cc][v + 1] <= o) v++;
        const int64_t rec = (int64_t)((v * U + ub) * 4 + c) * g.cap + (o - t.prefO[cc][v]);
        const uint32_t oxy = __ldg(&xy_from[rec]);
        const uint4 a = __ldg(&desc_from[2 * rec]), b = __ldg(&desc_from[2 * rec + 1]);
        const uint2 qo = __ldg(&q_from[rec]);
        const uint32_t qa = qo.x, qb = qo.y;
        const int vb = (o >= t.lagEnd[cc][v]) ? v + 1 : origin_vb(t, cc, o, v);
        const int lo_r = max(max(max(vb - 3, 0), ra), v - reach), hi_r = min(min(min(vb + 3, V - 1), rb), v + reach + 1);
        const int ou = (int)(oxy & 0xFFFFu), ov = (int)(oxy >> 16);
        const uint32_t lo_p = (uint32_t)max(ou - g.rad_u, 0) | ((uint32_t)max(ov - rad_v, 0) << 16);
        const uint32_t hi_p = (uint32_t)(ou + g.rad_u + 1) | ((uint32_t)(ov + rad_v + 1) << 16);
        uint32_t bk = kNone;
        {  // the best key the lane-per-origin walk (and the earlier chunks) left in best[]
          const uint32_t prev = bo[o];
          if (prev != kNone) bk = prev + ((uint32_t)j0 << 16);
        }
        int tl = (int)(bk >> 19);  // kNone >> 19 = 8191 > any SAD
        if (lane == 0) K2STAT(6, 1);
        // One pass, the origin's own row first (zero-flow prior).  Lanes test the two-word lower bound of 32 candidates at
        // a time against the best SAD known to the warp and COMPACT the survivors into the warp's queue; exact SADs are
        // then taken 32 at a time with every lane busy (evaluating in place would run the exact path whenever ANY lane
        // survives: with a few per cent survivors that is nearly every iteration).
        int hn = 0;
        auto drain = [&](int n) {  // exact evaluation of the last n (<= 32) queued candidates
          hn -= n;
          if (lane < n) {
            const uint32_t e2 = queue[hn + lane];
            K2STAT(7, 1);
            if ((int)(e2 >> 11) <= tl) bk = min(bk, cand_key(S, (int)(e2 & 0x7FFu), a, b, lo_p, hi_p));  // bound re-tested: tl may have dropped
          }
          tl = min(tl, (int)(__reduce_min_sync(FULL, bk) >> 19));
          __syncwarp();
        };
        const bool own = (v >= lo_r) && (v < hi_r);
        const bool dense = tl > kSeedGood;  // no partner in sight: the bound prunes little, most lanes would evaluate anyway
        for (int rr = own ? lo_r - 1 : lo_r; rr < hi_r; rr++) {
          const int r = (rr < lo_r) ? v : rr;
          if (own && rr == v) continue;
          const int p1 = t.segOff[r][j1];
          if (dense) {
            // evaluate in place; the bound is lane-private inside a row and shared warp-wide after it
#pragma unroll 2
            for (int p = t.segOff[r][j0] + lane; p < p1; p += 32) {
              if ((int)sad4(qb, S.q1[p], sad4(qa, S.q[p], 0u)) <= tl) {
                bk = min(bk, cand_key(S, p, a, b, lo_p, hi_p));
                tl = min(tl, (int)(bk >> 19));
              }
            }
            tl = __reduce_min_sync(FULL, tl);
            continue;
          }
          // 64 candidates per iteration, two per lane; loads are unconditional (positions past p1 are clamped into the
          // staging arrays and masked out of the ballots)
          for (int pb = t.segOff[r][j0]; pb < p1; pb += 64) {
            const int pA = pb + lane, pB = pA + 32;
            const int cA = min(pA, kCandCap - 1), cB = min(pB, kCandCap - 1);
            const int qA = (int)sad4(qb, S.q1[cA], sad4(qa, S.q[cA], 0u));
            const int qB = (int)sad4(qb, S.q1[cB], sad4(qa, S.q[cB], 0u));
            const bool hA = (pA < p1) && (qA <= tl), hB = (pB < p1) && (qB <= tl);
            const uint32_t balA = __ballot_sync(FULL, hA), balB = __ballot_sync(FULL, hB);
            const int nA = __popc(balA);
            K2CHECK(hn >= 0 && hn + nA + __popc(balB) <= kQueue && p1 <= kCandCap);
            if (hA) queue[hn + __popc(balA & lt_mask)] = ((uint32_t)qA << 11) | (uint32_t)pA;
            if (hB) queue[hn + nA + __popc(balB & lt_mask)] = ((uint32_t)qB << 11) | (uint32_t)pB;
            hn += nA + __popc(balB);
            while (hn >= 32) { __syncwarp(); drain(32); }
          }
        }
        __syncwarp();
        if (hn > 0) drain(hn);
        bk = __reduce_min_sync(FULL, bk);
        if (lane == 0) bo[o] = (bk == kNone) ? kNone : bk - ((uint32_t)j0 << 16);
      }
    }
    first_chunk = false;
    ra = rb;
    if (ra >= row_hi) break;
  }
  return true;
}

template <bool PRUNE, int THREADS>
__global__ void __launch_bounds__(THREADS, 2)
k2_match(Geom g, FrameArrays fa, const int4* __restrict__ edges, uint32_t* __restrict__ best, int rows_per_group,
         int ngroups, int hit_budget, int mode) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  K2Smem& S = *reinterpret_cast<K2Smem*>(smem_raw);
  const int cg = blockIdx.x / ngroups, grp = blockIdx.x - cg * ngroups;
  // mode 0: grid.y = class x {even, odd edge}, every search in full.  Modes 1 and 2 split a mono pair's two directions
  // over two launches (grid.y = class): 1 = the odd edges (prev -> curr) in full, 2 = the even edges (curr -> prev)
  // starting from the hints k2_hint_scatter derived from the odd edges' results.
  const int c = mode ? (int)blockIdx.y : (int)(blockIdx.y >> 1);
  const int edge = blockIdx.z * 2 + (mode ? (mode == 1 ? 1 : 0) : (int)(blockIdx.y & 1));
  const int4 ed = edges[edge];  // x = origin slot, y = target slot, z = rad_v, w = row reach (MatchEdge)
  const int from_slot = ed.x, to_slot = ed.y;
  const int ub0 = cg * kOC;
  const int noc = min(kOC, (g.U - 1) - ub0);  // origin columns 0..U-2 only (src/myMatch.cpp:141-146)
  uint32_t* best_dir = best + (int64_t)edge * g.U * 4 * g.col_cap;
  // joint pass over both columns; if a row of the shared window does not fit, one pass per column
  for (int it = -1;;) {
    const bool ok = k2_columns<PRUNE, THREADS>(S, g, fa, from_slot, to_slot, best_dir, c, it < 0 ? ub0 : ub0 + it,
                                               it < 0 ? noc : 1, grp, rows_per_group, ed.z, ed.w, hit_budget, mode == 2,
                                               mode == 1 ? best + (int64_t)(edge - 1) * g.U * 4 * g.col_cap : nullptr);
    if (it < 0 && ok) break;
    if (++it >= noc) break;
  }
}

// ---- hints for the second direction of a mono pair ---------------------------------------------------------------
// A match is emitted only when the cycle closes: A (curr) -> argmin B (prev) and B -> argmin A (src/myMatch.cpp:160-186).
// Once the prev -> curr direction is known, the curr -> prev search of A matters only if some B points at A, and it can
// only close on the pointer with the smallest key; that key is what A's own search would compute for B, so it is a
// ready-made bound.  The full launch resets best[curr->prev] of every origin on its way, k2_hint_scatter writes min over the pointers
// B of key(A, B) = SAD << 19 | (u_bin(B) - first column of A's window) << 16 | col_idx(B) -- or a marker that cannot be
// validated when B's column lies outside A's (half-open, clipped) column range.  K2 mode 2 validates the hint against
// the staged window (rows, pixel window, key equality) and falls back to a plain search if anything differs; origins
// nobody points at are skipped (their result cannot be observed).
constexpr int KH_THREADS = 256;
__device__ __forceinline__ int column_total(const Geom& g, const int32_t* __restrict__ cnt, int u, int c) {
  __shared__ int tot;
  if (threadIdx.x == 0) tot = 0;
  __syncthreads();
  int s = 0;
  for (int v = threadIdx.x; v < g.V; v += blockDim.x) s += cnt[(v * g.U + u) * 4 + c];
  if (s) atomicAdd(&tot, s);
  __syncthreads();
  return tot;
}

__global__ void __launch_bounds__(KH_THREADS)
k2_hint_scatter(Geom g, FrameArrays fa, const int4* __restrict__ edges, uint32_t* __restrict__ best) {
  const int ub = blockIdx.x >> 2, c = blockIdx.x & 3, pair = blockIdx.y;
  const int4 ed = edges[2 * pair + 1];  // prev -> curr: origins B are the features of ed.x
  const int n = column_total(g, fa.cnt + (int64_t)ed.x * fa.cnt_stride, ub, c);
  const uint32_t* bsrc = best + ((int64_t)(2 * pair + 1) * g.U * 4 + (ub * 4 + c)) * g.col_cap;
  uint32_t* bdst = best + (int64_t)(2 * pair) * g.U * 4 * g.col_cap;
  const int a0 = max(ub - 3, 0);
  for (int k = threadIdx.x; k < n; k += KH_THREADS) {
    const uint32_t key = bsrc[k];
    if (key == kNone) continue;
    const int ua = a0 + (int)((key >> 16) & 7u), ka = (int)(key & 0xFFFFu);
    const int w0 = max(ua - 3, 0), w1 = min(ua + 3, g.U - 1);  // A's target columns [w0, w1): src/myMatch.cpp:90-91,97
    const uint32_t low = (ub >= w0 && ub < w1) ? (((uint32_t)(ub - w0) << 16) | (uint32_t)k) : 0x7FFFFu;
    atomicMin(&bdst[(int64_t)(ua * 4 + c) * g.col_cap + ka], (key & ~0x7FFFFu) | low);
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

// ---- EXTENSION (parity unpinned: absent from the reference; oracle = oracle/vo_oracle.c
// vo_oracle_quad_matching) -- quad cycle 1c -> 1p -> 2p -> 2c -> 1c over four K2 edges ---------------------------
// coordinates of feature (u_bin, class, col_idx) of a slot: walk the column's v-bin counts
__device__ __forceinline__ uint32_t feature_xy(const Geom& g, const FrameArrays& fa, int slot, int u, int c, int k) {
  const int32_t* cnt = fa.cnt + (int64_t)slot * fa.cnt_stride;
  int v = 0, rem = k;
  for (; v < g.V - 1; v++) {
    const int nn = cnt[(v * g.U + u) * 4 + c];
    if (rem < nn) break;
    rem -= nn;
  }
  return fa.xy[(int64_t)slot * fa.rec_stride + (int64_t)((v * g.U + u) * 4 + c) * g.cap + rem];
}

// one hop of the cycle: best target of origin (ou, k) on an edge, false if none
__device__ __forceinline__ bool hop(const uint32_t* __restrict__ best_e, int col_cap, int c, int ou, int ok, int& tu, int& tk) {
  const uint32_t key = best_e[(int64_t)(ou * 4 + c) * col_cap + ok];
  if (key == kNone) return false;
  decode_key(key, ou, tu, tk);
  return true;
}

// K3q: per current-left column u: does the cycle close with non-negative disparities (libviso2's quad rule)?
// then the reference's same-pixel dedupe (lowest class wins) and the per-(u,class) counts, as k3_flags.
__global__ void __launch_bounds__(K3_THREADS)
k3q_flags(Geom g, FrameArrays fa, const int4* __restrict__ quads, const uint32_t* __restrict__ best,
          uint8_t* __restrict__ flags, int32_t* __restrict__ list_cnt, uint32_t* __restrict__ qxy) {
  __shared__ int32_t pref[4][kMaxV + 1];
  __shared__ int32_t warp_cnt[K3_THREADS / 32];
  const int u = blockIdx.x, item = blockIdx.y, tid = threadIdx.x;
  const int U = g.U, V = g.V;
  const int4 qd = quads[item];  // x = 1p, y = 2p, z = 1c, w = 2c
  const int32_t* cnt = fa.cnt + (int64_t)qd.z * fa.cnt_stride;
  if (tid < 4) {
    int acc = 0;
    pref[tid][0] = 0;
    for (int v = 0; v < V; v++) { acc += cnt[(v * U + u) * 4 + tid]; pref[tid][v + 1] = acc; }
  }
  __syncthreads();
  const int64_t L = (int64_t)U * 4 * g.col_cap;
  const uint32_t* e0 = best + ((int64_t)item * 4 + 0) * L;  // 1c -> 1p (flow)
  const uint32_t* e1 = best + ((int64_t)item * 4 + 1) * L;  // 1p -> 2p (stereo)
  const uint32_t* e2 = best + ((int64_t)item * 4 + 2) * L;  // 2p -> 2c (flow)
  const uint32_t* e3 = best + ((int64_t)item * 4 + 3) * L;  // 2c -> 1c (stereo)
  const uint32_t* xy_c = fa.xy + (int64_t)qd.z * fa.rec_stride;
  const uint32_t* meta = fa.meta + (int64_t)qd.z * fa.rec_stride;
  uint8_t* fl = flags + (int64_t)item * L;
  uint32_t* qx = qxy + (int64_t)item * L * 3;
  if (u < U - 1) {  // the last column is never an origin (src/myMatch.cpp:141-146)
    for (int c = 0; c < 4; c++) {
      const int n = pref[c][V];
      for (int k = tid; k < n; k += K3_THREADS) {
        int au, ak, bu, bk, du, dk, su, sk;
        bool ok = hop(e0, g.col_cap, c, u, k, au, ak) && hop(e1, g.col_cap, c, au, ak, bu, bk) &&
                  hop(e2, g.col_cap, c, bu, bk, du, dk) && hop(e3, g.col_cap, c, du, dk, su, sk) && su == u && sk == k;
        if (ok) {
          int v = 0;
          while (pref[c][v + 1] <= k) v++;
          const uint32_t x1c = xy_c[(int64_t)((v * U + u) * 4 + c) * g.cap + (k - pref[c][v])];
          const uint32_t x1p = feature_xy(g, fa, qd.x, au, c, ak), x2p = feature_xy(g, fa, qd.y, bu, c, bk);
          const uint32_t x2c = feature_xy(g, fa, qd.w, du, c, dk);
          ok = (x1p & 0xFFFFu) >= (x2p & 0xFFFFu) && (x1c & 0xFFFFu) >= (x2c & 0xFFFFu);  // u1p >= u2p, u1c >= u2c
          if (ok) {
            uint32_t* q3 = qx + ((int64_t)(u * 4 + c) * g.col_cap + k) * 3;
            q3[0] = x1p; q3[1] = x2p; q3[2] = x2c;
          }
        }
        fl[(int64_t)(u * 4 + c) * g.col_cap + k] = ok ? 1 : 0;
      }
    }
  }
  __syncthreads();
  if (u < U - 1) {  // a lower class already claimed this current pixel (M[][] of src/myMatch.cpp:188-191)
    for (int c = 2; c < 4; c++) {
      const int n = pref[c][V];
      for (int k = tid; k < n; k += K3_THREADS) {
        uint8_t* f = &fl[(int64_t)(u * 4 + c) * g.col_cap + k];
        if (!*f) continue;
        int v = 0;
        while (pref[c][v + 1] <= k) v++;
        const uint32_t tw = meta[(int64_t)((v * U + u) * 4 + c) * g.cap + (k - pref[c][v])];
        if (tw != kNone) {
          const int tc = (int)(tw >> 16), tk = pref[tc][v] + (int)(tw & 0xFFFFu);
          if (fl[(int64_t)(u * 4 + tc) * g.col_cap + tk]) *f = 0;
        }
      }
    }
  }
  __syncthreads();
  for (int c = 0; c < 4; c++) {
    const int n = pref[c][V];
    int mine = 0;
    if (u < U - 1)
      for (int k = tid; k < n; k += K3_THREADS) mine += fl[(int64_t)(u * 4 + c) * g.col_cap + k];
    for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((tid & 31) == 0) warp_cnt[tid >> 5] = mine;
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < K3_THREADS / 32; w++) tot += warp_cnt[w];
      list_cnt[(int64_t)item * U * 4 + u * 4 + c] = tot;
    }
    __syncthreads();
  }
}

// K3b: ordered emit.  One CTA per (u,class) list; its output offset is the sum of the counts of
// all lists before it in (u_bin, class) order (src/myMatch.cpp:181-196).
// QUAD: the partners' coordinates come from the qxy scratch k3q_flags filled ({1p, 2p, 2c} per current-left feature).
template <bool QUAD>
__global__ void __launch_bounds__(K3_THREADS)
k3_emit(Geom g, FrameArrays fa, const int2* __restrict__ pairs, const uint32_t* __restrict__ best,
        const uint8_t* __restrict__ flags, const int32_t* __restrict__ list_cnt, const uint32_t* __restrict__ qxy,
        uint4* __restrict__ out, int64_t out_stride_matches, int32_t max_matches, int32_t* __restrict__ n_raw,
        int32_t* __restrict__ status, uint32_t* __restrict__ side) {
  __shared__ int32_t pref[kMaxV + 1];
  __shared__ int32_t ppref[6][kMaxV + 1];
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
  } else if (!QUAD && tid >= 64 && tid < 70) {
    // col_idx prefixes of the six previous-frame columns a match of this list can point into
    const int j = tid - 64, pu = max(u - 3, 0) + j;
    const int32_t* cnt = fa.cnt + (int64_t)pr.x * fa.cnt_stride;
    int acc = 0;
    ppref[j][0] = 0;
    for (int v = 0; v < V; v++) { acc += (pu < U) ? cnt[(v * U + pu) * 4 + c] : 0; ppref[j][v + 1] = acc; }
  }
  __syncthreads();
  const int total = s_total;
  if (list == 0 && tid == 0) {
    n_raw[pair] = (total < 5) ? 0 : min(total, max_matches);  // "<5 -> 0": src/myMatch.cpp:198-203
    if (total > max_matches) atomicExch(status, 1);
  }
  if (lc[list] == 0) return;
  const int n = pref[V];
  const uint32_t* best_cp = QUAD ? nullptr : best + (int64_t)(pair * 2 + 0) * U * 4 * g.col_cap;
  const uint8_t* fl = flags + ((int64_t)pair * U * 4 + list) * g.col_cap;
  const uint32_t* xy_c = fa.xy + (int64_t)pr.y * fa.rec_stride;
  const uint32_t* xy_p = fa.xy + (int64_t)pr.x * fa.rec_stride;
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
      const uint32_t m1f = __float_as_uint(-1.0f), m1i = 0xFFFFFFFFu;
      const uint32_t u1c = __float_as_uint((float)(cxy & 0xFFFFu)), v1c = __float_as_uint((float)(cxy >> 16));
      uint4* rec = dst + (int64_t)posn * 3;
      if (side) side[(int64_t)pair * out_stride_matches + posn] = cxy;  // (u1c, v1c) of every record: all K4 needs to bucket
      if (QUAD) {
        const uint32_t* qq = qxy + (((int64_t)pair * U * 4 + list) * g.col_cap + k) * 3;
        const uint32_t x1p = qq[0], x2p = qq[1], x2c = qq[2];
        // p_match(u1p,v1p,-1,u2p,v2p,-1,u1c,v1c,-1,u2c,v2c,-1): the fields of src/matcher.h:89-104, indices as src/myMatch.cpp:189
        rec[0] = make_uint4(__float_as_uint((float)(x1p & 0xFFFFu)), __float_as_uint((float)(x1p >> 16)), m1i,
                            __float_as_uint((float)(x2p & 0xFFFFu)));
        rec[1] = make_uint4(__float_as_uint((float)(x2p >> 16)), m1i, u1c, v1c);
        rec[2] = make_uint4(m1i, __float_as_uint((float)(x2c & 0xFFFFu)), __float_as_uint((float)(x2c >> 16)), m1i);
      } else {
        int pu, pk;
        decode_key(best_cp[(int64_t)list * g.col_cap + k], u, pu, pk);
        const int32_t* pp = ppref[pu - max(u - 3, 0)];  // locate the previous feature's (row, slot) from its col_idx
        int pv = 0;
        while (pv < V - 1 && pp[pv + 1] <= pk) pv++;
        const int rem = pk - pp[pv];
        const uint32_t pxy = xy_p[(int64_t)((pv * U + pu) * 4 + c) * g.cap + rem];
        // p_match(u1p,v1p,-1,-1,-1,-1,u1c,v1c,-1,-1,-1,-1): src/myMatch.cpp:189
        const uint32_t u1p = __float_as_uint((float)(pxy & 0xFFFFu)), v1p = __float_as_uint((float)(pxy >> 16));
        rec[0] = make_uint4(u1p, v1p, m1i, m1f);
        rec[1] = make_uint4(m1f, m1i, u1c, v1c);
        rec[2] = make_uint4(m1i, m1f, m1f, m1i);
      }
    }
    running += chunk_total;
    __syncthreads();
  }
}

template <bool PRUNE, int THREADS>
static cudaError_t launch_k2_t(const Geom& g, const FrameArrays& fa, const int4* edges, int n_edges, uint32_t* best,
                               int rows_per_group, int hit_budget, bool split, cudaStream_t st) {
  const int bytes = (int)sizeof(K2Smem);
  const int ngroups = (g.V + rows_per_group - 1) / rows_per_group;
  const int cgroups = (g.U - 1 + kOC - 1) / kOC;
  const int budget = hit_budget > 0 ? hit_budget : kHitBudget;
  if (!split) {
    dim3 grid(cgroups * ngroups, 8, n_edges / 2);  // edges come in twos (mono: curr->prev, prev->curr; quad: four)
    k2_match<PRUNE, THREADS><<<grid, THREADS, bytes, st>>>(g, fa, edges, best, rows_per_group, ngroups, budget, 0);
    return cudaGetLastError();
  }
  // mono pairs: prev -> curr in full (its CTAs also reset best[curr -> prev]), hints, then curr -> prev from the hints
  dim3 grid(cgroups * ngroups, 4, n_edges / 2), hgrid((g.U - 1) * 4, n_edges / 2);
  k2_match<PRUNE, THREADS><<<grid, THREADS, bytes, st>>>(g, fa, edges, best, rows_per_group, ngroups, budget, 1);
  k2_hint_scatter<<<hgrid, KH_THREADS, 0, st>>>(g, fa, edges, best);
  k2_match<PRUNE, THREADS><<<grid, THREADS, bytes, st>>>(g, fa, edges, best, rows_per_group, ngroups, budget, 2);
  return cudaGetLastError();
}

// per-device opt-in to K2's dynamic shared memory, every instantiation launch_k2 can pick (see init_k1_device)
cudaError_t init_k2_device() {
  const int bytes = (int)sizeof(K2Smem);
  cudaError_t e = cudaFuncSetAttribute(k2_match<false, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 384>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  return e;
}

cudaError_t launch_k2(const Geom& g, const FrameArrays& fa, const int4* edges, int n_edges, uint32_t* best,
                      int rows_per_group, int prune, int threads, int hit_budget, int split, cudaStream_t st) {
  // split != 0: the edges are mono pairs (curr->prev, prev->curr) and may run as two dependent launches
  if (!prune) return launch_k2_t<false, 256>(g, fa, edges, n_edges, best, rows_per_group, hit_budget, false, st);
  if (threads == 256) return launch_k2_t<true, 256>(g, fa, edges, n_edges, best, rows_per_group, hit_budget, split, st);
  if (threads == 384) return launch_k2_t<true, 384>(g, fa, edges, n_edges, best, rows_per_group, hit_budget, split, st);
  return launch_k2_t<true, 512>(g, fa, edges, n_edges, best, rows_per_group, hit_budget, split, st);
}

cudaError_t launch_k3(const Geom& g, const FrameArrays& fa, const int2* pairs, int n_pairs, const uint32_t* best,
                      uint8_t* flags, int32_t* list_cnt, void* out, int64_t out_stride_matches, int32_t max_matches,
                      int32_t* n_raw, int32_t* status, uint32_t* side, cudaStream_t st) {
  k3_flags<<<dim3(g.U, n_pairs), K3_THREADS, 0, st>>>(g, fa, pairs, best, flags, list_cnt);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  k3_emit<false><<<dim3(g.U * 4, n_pairs), K3_THREADS, 0, st>>>(g, fa, pairs, best, flags, list_cnt, nullptr, (uint4*)out,
                                                                out_stride_matches, max_matches, n_raw, status, side);
  return cudaGetLastError();
}

// quad items: quads[i] = {1p, 2p, 1c, 2c} slots, pairs[i] = {1p, 1c} (for the shared emit), best = 4 edges per item
cudaError_t launch_k3_quad(const Geom& g, const FrameArrays& fa, const int4* quads, const int2* pairs, int n_items,
                           const uint32_t* best, uint8_t* flags, int32_t* list_cnt, uint32_t* qxy, void* out,
                           int64_t out_stride_matches, int32_t max_matches, int32_t* n_raw, int32_t* status,
                           uint32_t* side, cudaStream_t st) {
  k3q_flags<<<dim3(g.U, n_items), K3_THREADS, 0, st>>>(g, fa, quads, best, flags, list_cnt, qxy);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  k3_emit<true><<<dim3(g.U * 4, n_items), K3_THREADS, 0, st>>>(g, fa, pairs, best, flags, list_cnt, qxy, (uint4*)out,
                                                               out_stride_matches, max_matches, n_raw, status, side);
  return cudaGetLastError();
}

}  // namespace vo
#ifdef VO_K2_STATS
extern "C" int vo_debug_k2_stats(unsigned long long* out) {
  cudaDeviceSynchronize();
  cudaError_t e = cudaMemcpyFromSymbol(out, vo::g_k2_stats, sizeof(unsigned long long) * 8);
  unsigned long long z[8] = {0};
  cudaMemcpyToSymbol(vo::g_k2_stats, z, sizeof(z));
  return (int)e;
}
#endif
