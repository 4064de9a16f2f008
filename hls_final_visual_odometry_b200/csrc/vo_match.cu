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
// every bin list padded to a multiple of 4.  A LANE owns an origin: the 32 lanes of a warp walk the
// same candidates, so every shared-memory read is a broadcast and no per-origin work is replicated.
// The key min(SAD<<19 | u_rel<<16 | col_idx) is exactly the reference's "strict <, first candidate in
// (u_bin asc, col_idx asc) order wins".
//
// Two passes per CTA (r2f):
//   pass 1  every origin gets a bound -- the hint of the other direction (mono pairs, second launch) or the
//           candidate with the smallest signature distance in its own bin (zero-flow prior), evaluated exactly --
//           and is SORTED by what that bound is worth into three lists:
//             Z  bound 0   (an exact partner: rendered or noise-free frames)   -> one-word filter
//             N  0 < bound <= seed_good (a partner under sensor noise)         -> two-word filter
//             P  no usable bound (unmatched feature, partner outside the own bin, exhaustive mode)
//   pass 2  warps pull work items from a shared counter: Z/N tasks of 32 sorted origins run the lower-bound
//           filter (desc_q: SAD >= Q, one VABSDIFF4 per signature word) over their window and collect one bit
//           per CANDIDATE; each lane then evaluates its own few hits exactly, no queue, no shuffles.  P tasks
//           are split by window row into items that evaluate EVERY candidate exactly (8 VABSDIFF4.U8.ACC +
//           packed window test, ~13 instructions per lane and candidate): for an origin whose bound prunes
//           little that is cheaper than filtering, and it bounds the cost of any input at about twice a clean
//           frame's.  Pruning is conservative, so the result equals evaluating every SAD (PRUNE=false sends
//           every origin down the P path, for the tests).
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
__device__ unsigned long long g_k2_stats[16];
#define K2STAT(i, n) atomicAdd(&g_k2_stats[i], (unsigned long long)(n))
#else
#define K2STAT(i, n)
#endif

constexpr int kMaxV = 64;       // v-bin rows supported by the shared-memory tables
constexpr int kOC = 2;          // origin u-bin columns per CTA (they share the staged window)
constexpr int kMaxCols = kOC + 5;  // staged target columns: [ub0-3, ub_last+3) clipped
constexpr int kSegW = kMaxCols + 1;
constexpr int kCandCap = 1896;  // staged candidates per chunk; >= 6*300 so a row of a one-column window always fits
constexpr int kMaxWarps = 16;
constexpr int kQueue = 128;     // pending hits per warp (M tasks): < 32 left over + what one 32-candidate word adds in one go
static_assert(kMaxV <= 64, "list entries carry the row in 6 bits");
constexpr int kMaxTasks = 72;   // filter tasks of one segment: per column <= 16 Z/N (+1 for a split) + 16 M, two columns
constexpr int kListCap = 512;   // origins per column sorted in one go (a longer column list is processed in segments)
constexpr int kMergeBelow = 12; // a Z or N list shorter than this joins the other list's tasks (either filter is exact for both)
#ifndef VO_K2_FLOOD
#define VO_K2_FLOOD 12
#endif
constexpr int kFlood = VO_K2_FLOOD;  // M tasks: a lane with at least this many hits in a 32-candidate word is expanded by the whole warp
constexpr int kBig = 1 << 20;
constexpr uint32_t kSkip = 0xFFFFFFFEu;  // best[] between the chunks of a hinted search: nobody points at this origin (never a real key: SAD <= 8160)
#ifndef VO_K2_SEED_GOOD
#define VO_K2_SEED_GOOD 96
#endif
constexpr int kSeedGood = VO_K2_SEED_GOOD;  // default of vo_config.match_seed_good: an origin whose bound is above this is matched exhaustively

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

// Staged candidates, two layouts (template parameter LEAN of the kernel):
//   fat  (frames whose windows fit: V <= 12 bin rows, e.g. 1241x376) -- descriptor, window word, rank and signature in shared
//        memory, 48 bytes per candidate, 1896 candidates per chunk;
//   lean (tall frames, e.g. 2048x1024: V = 23) -- the 32 descriptor bytes stay in global memory (L2-resident: a frame's
//        records are 1-3 MB) and are fetched by the few candidates that pass the filter; shared memory keeps 20 bytes per
//        candidate (window word, rank, record index, signature), 4096 candidates per chunk, so a whole window of 16 origin rows
//        is staged at once instead of in chunks that every origin has to be set up for again.
template <bool LEAN> struct K2Cands;
template <> struct K2Cands<false> {
  static constexpr int kCap = kCandCap, kPosBits = 11;
  using QEntry = uint16_t;
  QEntry queue[kMaxWarps][kQueue];    // M tasks: filter hits awaiting exact evaluation, owner lane << kPosBits | staged position
  __align__(16) uint4 lo[kCap];       // descriptor bytes 0..15
  uint4 hi[kCap];                     // descriptor bytes 16..31
  uint2 xr[kCap];                     // {xy + 0x80008000, u_rel<<16 | col_idx}
  __align__(16) uint32_t q[kCap];     // signature word 0 (desc_q); read four at a time
  __align__(16) uint32_t q1[kCap];    // signature word 1
};
template <> struct K2Cands<true> {
  static constexpr int kCap = 4096, kPosBits = 12;
  using QEntry = uint32_t;
  QEntry queue[kMaxWarps][kQueue];
  uint2 xr[kCap];
  uint32_t rec[kCap];                 // record index of the candidate inside its frame slot (descriptor = desc[2 * rec], [2 * rec + 1])
  __align__(16) uint32_t q[kCap];
  __align__(16) uint32_t q1[kCap];
};

template <bool LEAN>
struct __align__(128) K2SmemT : K2Cands<LEAN> {
  K2Tables t;
  int32_t n_z[kOC], n_n[kOC], n_p[kOC];  // entries of the three lists of this segment
  int32_t next_item;                  // pass 2: next unclaimed task
  int32_t task_ctr[kMaxTasks];        // pass 2: next unclaimed window row of a task (in its own row order)
  int32_t task_nr[kMaxTasks];         // pass 2: window rows of a task, once some warp has looked (-1 before)
  uint16_t glist[kOC][kListCap];      // well-bounded origins (index within the segment): Z from the front, N from the back
  uint32_t gkey[kOC][kListCap];       // bound of every origin of the segment (key, staged-column form), by index
  uint16_t plist[kOC][kListCap];      // M: origins with a poor bound or none (exhaustive mode: every origin)
  uint8_t cls[kOC][kListCap];         // pass 1 -> ordered scan: 0 nothing to search, 1 Z, 2 N, 3 M
  uint32_t sbest[kMaxWarps][32];      // M tasks: best key of the warp's 32 origins (results of the cooperative evaluations)
};
static_assert(K2Cands<false>::kCap <= (1 << K2Cands<false>::kPosBits) && K2Cands<true>::kCap <= (1 << K2Cands<true>::kPosBits),
              "queue entries pack the staged position into kPosBits bits");
static_assert(sizeof(K2SmemT<false>) <= 113 * 1024 && sizeof(K2SmemT<true>) <= 113 * 1024,
              "two K2 CTAs per SM need <= 113 KB each (227 KB minus the per-CTA reserve)");

// acc + sum of |a_i - b_i| over the four packed bytes: one VABSDIFF4.U8.ACC
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t acc) {
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
  return d;
}

__device__ __forceinline__ uint32_t sad32(const uint4& a, const uint4& b, const uint4& x, const uint4& y) {
  uint32_t s = sad4(a.x, x.x, 0u), r = sad4(b.x, y.x, 0u);  // two chains: the accumulating form serialises
  s = sad4(a.y, x.y, s);
  r = sad4(b.y, y.y, r);
  s = sad4(a.z, x.z, s);
  r = sad4(b.z, y.z, r);
  s = sad4(a.w, x.w, s);
  r = sad4(b.w, y.w, r);
  return s + r;
}

// key of one staged candidate for an origin (a, b, window lo_p/hi_p), kNone if outside the window.
// packed window test: both 16-bit fields carry a guard bit (bit 15); DESIGN.md "window test"
template <bool LEAN>
__device__ __forceinline__ uint32_t cand_key(const K2SmemT<LEAN>& S, const uint4* __restrict__ desc_to, int p, const uint4& a,
                                             const uint4& b, uint32_t lo_p, uint32_t hi_p) {
  K2CHECK(p >= 0 && p < K2Cands<LEAN>::kCap);
  const uint2 xr = S.xr[p];
  uint32_t sad;
  if constexpr (LEAN) {
    const uint32_t rec = S.rec[p];
    sad = sad32(a, b, __ldg(&desc_to[2 * rec]), __ldg(&desc_to[2 * rec + 1]));
  } else {
    sad = sad32(a, b, S.lo[p], S.hi[p]);
  }
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

// one origin in a lane's registers
struct K2Origin {
  uint4 a, b;            // descriptor
  uint32_t qa, qb;       // signature words
  uint32_t lo_p, hi_p;   // packed pixel window (src/myMatch.cpp:84-87,101)
  int v;                 // own v-bin row
  int vlo, vhi;          // candidate rows clipped to the staged chunk (vlo >= vhi: nothing to search here)
  int lo_full, hi_full;  // candidate rows before clipping
};

// Matches the origins of columns [ub0, ub0+noc) x rows of group `grp`.  Returns false (having
// written nothing) when a row of the shared (noc+5)-column window does not fit the staging buffer;
// the caller then runs the columns one at a time (a 6-column row always fits).
template <bool PRUNE, int THREADS, bool LEAN>
__device__ __forceinline__ bool k2_columns(K2SmemT<LEAN>& S, const Geom& g, const FrameArrays& fa, int from_slot, int to_slot,
                                           uint32_t* __restrict__ best_dir, int c, int ub0, int noc, int grp,
                                           int rows_per_group, int rad_v, int reach, int seed_good, bool hinted,
                                           uint32_t* __restrict__ clear_dir) {
  constexpr int NW = THREADS / 32;
  constexpr uint32_t FULL = 0xFFFFFFFFu;
  constexpr int kCap = K2Cands<LEAN>::kCap, kPosBits = K2Cands<LEAN>::kPosBits;  // staged candidates per chunk
  using QEntry = typename K2Cands<LEAN>::QEntry;
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
  static_assert(kOC == 2, "the per-column scalars below are written out for two origin columns");
  int row_lo = V;
  const int oB0 = t.prefO[0][r0], oB1 = t.prefO[1][r0];
  const int nO0 = t.prefO[0][r1] - oB0, nO1 = t.prefO[1][r1] - oB1;
  const int n_max = max(nO0, nO1);
#pragma unroll
  for (int cc = 0; cc < kOC; cc++) {
    const int ob = cc ? oB1 : oB0;
    if ((cc ? nO1 : nO0) > 0) {  // rows any origin of this group can see: vb is non-decreasing along the column list
      int v_first = r0;
      while (t.prefO[cc][v_first + 1] <= ob) v_first++;
      row_lo = min(row_lo, max(max(origin_vb(t, cc, ob, v_first) - 3, v_first - reach), 0));
    }
  }
  if (n_max == 0) return true;
  // exclusive: src/myMatch.cpp:93,99.  `reach` (>= V on flow edges: no restriction) narrows the rows of a
  // left<->right edge to those that can hold a candidate inside its +-rad_v window: a bin row spans 45 block
  // rows, so |dv| <= rad_v <= 16 keeps the candidate's row within +-1 of the origin's own row
  const int row_hi = min(min(r1 + 3, V - 1), r1 + reach);
  if (noc > 1)
    for (int v = row_lo; v < row_hi; v++)
      if (t.rowTot[v] > kCap) return false;

  const uint32_t* xy_from = fa.xy + (int64_t)from_slot * fa.rec_stride;
  const uint4* desc_from = fa.desc + (int64_t)from_slot * fa.rec_stride * 2;
  const uint2* q_from = fa.q + (int64_t)from_slot * fa.rec_stride;
  const uint32_t* xy_to = fa.xy + (int64_t)to_slot * fa.rec_stride;
  const uint4* desc_to = fa.desc + (int64_t)to_slot * fa.rec_stride * 2;
  const uint2* q_to = fa.q + (int64_t)to_slot * fa.rec_stride;
  const uint4* Q4 = reinterpret_cast<const uint4*>(S.q);
  const uint4* Q14 = reinterpret_cast<const uint4*>(S.q1);
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int nseg = (n_max + kListCap - 1) / kListCap;

  bool first_chunk = true;
  for (int ra = row_lo;;) {
    // ---- chunk = as many whole rows as fit the staging buffer ------------------------------
    int rb = ra, staged = 0;
    while (rb < row_hi && staged + t.rowTot[rb] <= kCap) { staged += t.rowTot[rb]; rb++; }
    const bool last_chunk = rb >= row_hi;
    __syncthreads();  // previous chunk fully consumed (and lagEnd visible)
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
      K2CHECK(dst >= 0 && dst + npad <= kCap && n <= g.cap);
      for (int k = lane; k < npad; k += 32) {
        if (k < n) {
          if constexpr (LEAN) {
            S.rec[dst + k] = (uint32_t)(src + k);
          } else {
            const uint4 d0 = desc_to[(src + k) * 2], d1 = desc_to[(src + k) * 2 + 1];
            S.lo[dst + k] = d0;
            S.hi[dst + k] = d1;
          }
          S.xr[dst + k] = make_uint2(xy_to[src + k] + 0x80008000u, rank0 + (uint32_t)k);
          const uint2 qv = q_to[src + k];
          S.q[dst + k] = qv.x;
          S.q1[dst + k] = qv.y;
        } else {
          if constexpr (LEAN) {
            S.rec[dst + k] = 0u;  // (never fetched: a dummy cannot pass the filter, and its window word rejects it anyway)
          } else {
            S.lo[dst + k] = make_uint4(0, 0, 0, 0);
            S.hi[dst + k] = make_uint4(0, 0, 0, 0);
          }
          S.xr[dst + k] = make_uint2(kNone, kNone);
          S.q[dst + k] = kNone;
          S.q1[dst + k] = kNone;
        }
      }
    }

    // origin (column cc, col_idx o) into a lane's registers; its candidate rows clipped to this chunk
    auto load_origin = [&](int cc, int o, bool valid, int v_known, K2Origin& og) {  // v_known < 0: find the row
      og.v = r0; og.vlo = kBig; og.vhi = 0; og.lo_full = 0; og.hi_full = 0;
      og.a = make_uint4(0, 0, 0, 0); og.b = og.a; og.qa = 0; og.qb = 0;
      uint32_t oxy = 0;
      if (valid) {
        int v = r0;
        if (v_known >= 0) v = v_known;
        else while (t.prefO[cc][v + 1] <= o) v++;
        K2CHECK(v >= r0 && v < r1 && o >= t.prefO[cc][v] && o < t.prefO[cc][v + 1]);
        const int64_t rec = (int64_t)((v * U + ub0 + cc) * 4 + c) * g.cap + (o - t.prefO[cc][v]);
        oxy = __ldg(&xy_from[rec]);
        og.a = __ldg(&desc_from[2 * rec]);
        og.b = __ldg(&desc_from[2 * rec + 1]);
        const uint2 qo = __ldg(&q_from[rec]);
        og.qa = qo.x; og.qb = qo.y;
        // lagging v_bin (src/myMatch.cpp:153-158): only the first origins after empty bins
        const int vb = (o >= t.lagEnd[cc][v]) ? v + 1 : origin_vb(t, cc, o, v);
        og.lo_full = max(max(vb - 3, 0), v - reach);  // src/myMatch.cpp:92-93,99
        og.hi_full = min(min(vb + 3, V - 1), v + reach + 1);
        const int lo_r = max(og.lo_full, ra), hi_r = min(og.hi_full, rb);
        if (lo_r < hi_r) { og.vlo = lo_r; og.vhi = hi_r; }
        og.v = v;
      }
      const int ou = (int)(oxy & 0xFFFFu), ov = (int)(oxy >> 16);
      og.lo_p = (uint32_t)max(ou - g.rad_u, 0) | ((uint32_t)max(ov - rad_v, 0) << 16);  // src/myMatch.cpp:84-87,101
      og.hi_p = (uint32_t)(ou + g.rad_u + 1) | ((uint32_t)(ov + rad_v + 1) << 16);
    };

    for (int seg = 0; seg < nseg; seg++) {
      __syncthreads();  // staging complete / previous segment's lists consumed
      if (tid == 32) S.next_item = 0;
      if (tid >= 64 && tid < 64 + kMaxTasks) { S.task_ctr[tid - 64] = 0; S.task_nr[tid - 64] = -1; }
      __syncthreads();

      // ================= pass 1: a bound for every origin, origins sorted by what it is worth =================
      // this segment's origins of column cc: [sO, sO + sN)
      const int sO0 = oB0 + seg * kListCap, sO1 = oB1 + seg * kListCap;
      const int sN0 = max(min(nO0 - seg * kListCap, kListCap), 0), sN1 = max(min(nO1 - seg * kListCap, kListCap), 0);
      const int nT10 = (sN0 + 31) >> 5, nT11 = (sN1 + 31) >> 5;
      for (int task = warp; task < nT10 + nT11; task += NW) {
        const int cc = (task >= nT10) ? 1 : 0;
        const int tb = task - (cc ? nT10 : 0);
        const int ub = ub0 + cc;
        const int j0 = max(ub - 3, 0) - ulo, j1 = min(ub + 3, U - 1) - ulo, jown = ub - ulo;  // src/myMatch.cpp:90-91,97
        const int o = (cc ? sO1 : sO0) + tb * 32 + lane;
        const bool valid = tb * 32 + lane < (cc ? sN1 : sN0);
        uint32_t* bo = best_dir + (int64_t)(ub * 4 + c) * g.col_cap;
        K2Origin og;
        load_origin(cc, o, valid, -1, og);
        uint32_t bestk = kNone, hint = kNone;
        bool need_seed = valid;  // no bound yet
        bool skip = false;       // hinted direction: no feature of the other frame points here, the cycle cannot close
        if (valid) {
          if (hinted && first_chunk) {
            hint = bo[o];
            need_seed = false;
            skip = hint == kNone;
          } else if (!first_chunk) {
            bestk = bo[o];
            skip = bestk == kSkip;
            if (skip) bestk = kNone;
            if (bestk != kNone) bestk += (uint32_t)j0 << 16;
            need_seed = bestk == kNone && !skip;
          }
          if (skip) { og.vlo = kBig; og.vhi = 0; }
        }
        if (PRUNE && hinted && first_chunk && valid && !skip) {
          // the hinted candidate must be one this origin may legally take (bin ranges are not symmetric between the two
          // directions): locate its record in global memory and evaluate it; anything unexpected -> plain search
          const int jr = (int)((hint >> 16) & 7u), kb = (int)(hint & 0xFFFFu), j = j0 + jr;
          uint32_t k = kNone;
          if (jr < 6 && j < j1) {
            for (int r = og.lo_full; r < og.hi_full; r++) {
              if (kb >= t.colT[j][r] && kb < t.colT[j][r + 1]) {
                const int64_t rec = (int64_t)((r * U + ulo + j) * 4 + c) * g.cap + (kb - t.colT[j][r]);
                const uint32_t x = __ldg(&xy_to[rec]) + 0x80008000u;
                const uint32_t sad = sad32(og.a, og.b, __ldg(&desc_to[2 * rec]), __ldg(&desc_to[2 * rec + 1]));
                const uint32_t t1 = x - og.lo_p, t2 = x - og.hi_p;
                if (((t1 & ~t2) & 0x80008000u) == 0x80008000u) k = sad * 0x80000u + (((uint32_t)j << 16) | (uint32_t)kb);
                break;
              }
            }
          }
          if (k != kNone && k == hint + ((uint32_t)j0 << 16)) bestk = k;
          else need_seed = true;
        }
        const bool rows_here = og.vlo < og.vhi;
        if (PRUNE && __any_sync(FULL, need_seed && rows_here)) {
          // seed: the candidate with the smallest lower bound among the bins the origins themselves
          // sit in (zero-flow prior); any legal candidate gives a valid bound
          const int smin = max(__reduce_min_sync(FULL, valid ? og.v : kBig), ra);
          const int smax = min(__reduce_max_sync(FULL, valid ? og.v : -1), rb - 1);
          uint32_t k1 = kNone;
          if (reach >= kMaxV) {  // flow edge: the +-135-row window admits practically every candidate of the own bin
            for (int r = smin; r <= smax; r++) {
              const int p1 = t.segOff[r][jown + 1];
              for (int p = t.segOff[r][jown]; p < p1; p += 4) {  // bin lists are padded to 4
                const uint4 qq = Q4[p >> 2], q1 = Q14[p >> 2];
                const uint32_t ka = sad4(og.qb, q1.x, sad4(og.qa, qq.x, 0u)) * 0x100000u + (uint32_t)p;
                const uint32_t kb = sad4(og.qb, q1.y, sad4(og.qa, qq.y, 0u)) * 0x100000u + (uint32_t)(p + 1);
                const uint32_t kc = sad4(og.qb, q1.z, sad4(og.qa, qq.z, 0u)) * 0x100000u + (uint32_t)(p + 2);
                const uint32_t kd = sad4(og.qb, q1.w, sad4(og.qa, qq.w, 0u)) * 0x100000u + (uint32_t)(p + 3);
                k1 = min(k1, min(min(ka, kb), min(kc, kd)));
              }
            }
          } else {  // narrow (left<->right) edge: only candidates inside the +-rad_v rows can seed; a disparity may cross
                    // into the neighbouring bins, so the own row's bins jown-1 .. jown+1 are searched
            const int ja = max(jown - 1, j0), jb = min(jown + 2, j1);
            for (int r = smin; r <= smax; r++) {
              const int p1 = t.segOff[r][jb];
              for (int p = t.segOff[r][ja]; p < p1; p++) {
                const uint32_t x = S.xr[p].x;
                const uint32_t t1 = x - og.lo_p, t2 = x - og.hi_p;
                const uint32_t kq = sad4(og.qb, S.q1[p], sad4(og.qa, S.q[p], 0u)) * 0x100000u + (uint32_t)p;
                k1 = min(k1, (((t1 & ~t2) & 0x80008000u) == 0x80008000u) ? kq : kNone);
              }
            }
          }
          if (k1 != kNone && rows_here && need_seed) {
            const int p = (int)(k1 & 0xFFFFFu);
            if (p >= t.rowStart[og.vlo] && p < t.rowStart[og.vhi]) bestk = min(bestk, cand_key(S, desc_to, p, og.a, og.b, og.lo_p, og.hi_p));
          }
        }
        // ---- class: Z (bound 0), N (bound <= seed_good), M (everything else that has rows to search) ----
        const int tq = (bestk == kNone) ? 0x7FFFFFFF : (int)(bestk >> 19);
        const bool searching = valid && rows_here;
        const bool poor = searching && (!PRUNE || tq > seed_good);
        const uint32_t ext = (bestk == kNone) ? kNone : bestk - ((uint32_t)j0 << 16);
        K2STAT(0, valid ? 1 : 0); K2STAT(1, (searching && !poor && tq == 0) ? 1 : 0); K2STAT(2, (searching && !poor && tq) ? 1 : 0);
        K2STAT(3, poor ? 1 : 0);
        if (valid) {
          const int i = tb * 32 + lane;
          S.cls[cc][i] = (uint8_t)((!searching ? 0 : poor ? 3 : (tq == 0 ? 1 : 2)) | (og.v << 2));  // kind | row << 2
          S.gkey[cc][i] = bestk;
          if (poor) bo[o] = ext;  // exhaustive mode merges into it with atomicMin
          else if (!searching) bo[o] = (skip && !last_chunk) ? kSkip : ext;  // nothing to search in this chunk
        }
      }
      __syncthreads();
      // ordered lists (a task of pass 2 takes 32 CONSECUTIVE origins of one class, so its lanes share most of their rows):
      // one warp per column scans the class codes in origin order
      if (warp < kOC) {
        const int cc = warp, n = cc ? sN1 : sN0;
        int tz = 0, tn = 0, tm = 0;
        for (int i0 = 0; i0 < n; i0 += 128) {  // four chunks of 32 per trip: the loads and votes of a trip are independent
          int k[4];
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int i = i0 + u * 32 + lane;
            k[u] = (i < n) ? (int)S.cls[cc][i] : 0;
          }
#pragma unroll
          for (int u = 0; u < 4; u++) {
            const int i = i0 + u * 32 + lane;
            const int kind = k[u] & 3;
            const uint16_t ent = (uint16_t)(i | ((k[u] >> 2) << 9));  // list entry: index within the segment | row << 9
            const uint32_t bz = __ballot_sync(FULL, kind == 1), bn = __ballot_sync(FULL, kind == 2), bm = __ballot_sync(FULL, kind == 3);
            if (kind == 1) S.glist[cc][tz + __popc(bz & lt_mask)] = ent;
            if (kind == 2) S.glist[cc][kListCap - 1 - (tn + __popc(bn & lt_mask))] = ent;
            if (kind == 3) S.plist[cc][tm + __popc(bm & lt_mask)] = ent;
            tz += __popc(bz); tn += __popc(bn); tm += __popc(bm);
          }
        }
        if (lane == 0) { S.n_z[cc] = tz; S.n_n[cc] = tn; S.n_p[cc] = tm; }
      }
      __syncthreads();

      // ================= pass 2: work items from a shared counter, long ones first =================
      // filter tasks (32 sorted origins, whole window) of kind M (poor or no bound), N, Z per column; without pruning the M
      // lists become P tasks split by window row instead
      // few exact-partner origins ride the two-word tasks (zAsN), few noisy ones the one-word tasks (nAsZ): either filter is exact
      const int nz0 = S.n_z[0], nn0 = S.n_n[0], nz1 = S.n_z[1], nn1 = S.n_n[1], np0 = S.n_p[0], np1 = S.n_p[1];
      const bool zAsN0 = nz0 > 0 && nn0 > 0 && nz0 < kMergeBelow, nAsZ0 = !zAsN0 && nz0 > 0 && nn0 > 0 && nn0 < kMergeBelow;
      const bool zAsN1 = nz1 > 0 && nn1 > 0 && nz1 < kMergeBelow, nAsZ1 = !zAsN1 && nz1 > 0 && nn1 > 0 && nn1 < kMergeBelow;
      const int nTN0 = nAsZ0 ? 0 : (nn0 + (zAsN0 ? nz0 : 0) + 31) >> 5, nTZ0 = zAsN0 ? 0 : (nz0 + (nAsZ0 ? nn0 : 0) + 31) >> 5;
      const int nTN1 = nAsZ1 ? 0 : (nn1 + (zAsN1 ? nz1 : 0) + 31) >> 5, nTZ1 = zAsN1 ? 0 : (nz1 + (nAsZ1 ? nn1 : 0) + 31) >> 5;
      const int nTM0 = PRUNE ? (np0 + 31) >> 5 : 0, nTM1 = PRUNE ? (np1 + 31) >> 5 : 0;
      const int nrows = rb - ra;
      const int iM = nTM0 + nTM1;
      const int i1 = iM + nTN0, i2 = i1 + nTN1, i3 = i2 + nTZ0, i4 = i3 + nTZ1;
      // Scheduling: a warp claims a task (32 sorted origins; M first: the long ones), loads its origins once and then takes
      // the task's window rows one at a time from a per-task counter -- own row first, where a small flow puts the partner.
      // A warp that finds no unclaimed task JOINS unfinished ones and takes rows from the same counters, so the CTA's
      // warps finish within a row of each other; the warps working on a task share its bounds through gkey[].
      const int nFilterTasks = i4;
      K2CHECK(nFilterTasks <= kMaxTasks);
      const int i5 = (PRUNE ? 0 : ((np0 + 31) >> 5) * nrows), n_items = i5 + (PRUNE ? 0 : ((np1 + 31) >> 5) * nrows);
      if (PRUNE) {
        auto run_task = [&](const int item) {
          // ---------- filter task: lower-bound filter, one bit per candidate, every lane evaluates its own hits ----------
          const bool kindM = item < iM;
          const int cc = kindM ? (item < nTM0 ? 0 : 1) : (item < i1) ? 0 : (item < i2) ? 1 : (item < i3) ? 0 : 1;
          const bool two = item < i2;  // M and N kinds: both signature words
          const int tb = item - (kindM ? (cc ? nTM0 : 0) : (item < i1) ? iM : (item < i2) ? i1 : (item < i3) ? i2 : i3);
          const int ub = ub0 + cc;
          const int j0 = max(ub - 3, 0) - ulo, j1 = min(ub + 3, U - 1) - ulo;
          // logical list of this kind: its own entries first, then the merged-in entries of the other kind
          const int nz = cc ? nz1 : nz0, nn = cc ? nn1 : nn0;
          const bool merged = two ? (cc ? zAsN1 : zAsN0) : (cc ? nAsZ1 : nAsZ0);
          const int n_own = kindM ? (cc ? np1 : np0) : two ? nn : nz;
          const int n_all = n_own + ((!kindM && merged) ? (two ? nz : nn) : 0);
          const int idx = tb * 32 + lane;
          const bool valid = idx < n_all;
          int slot = 0;
          if (valid && !kindM) {
            const bool n_entry = two ? (idx < n_own) : (idx >= n_own);
            const int k = (idx < n_own) ? idx : idx - n_own;
            slot = n_entry ? kListCap - 1 - k : k;
          } else if (valid) {
            slot = idx;
          }
          K2CHECK(slot >= 0 && slot < kListCap);
          const int ent = valid ? (int)(kindM ? S.plist[cc][slot] : S.glist[cc][slot]) : 0;
          const int ei = ent & (kListCap - 1);
          const int o = (cc ? sO1 : sO0) + ei;
          uint32_t bestk = valid ? S.gkey[cc][ei] : kNone;
          K2Origin og;
          load_origin(cc, o, valid, ent >> 9, og);
          const uint4 a = og.a, b = og.b;
          const uint32_t qa = og.qa, qb = og.qb, lo_p = og.lo_p, hi_p = og.hi_p;
          const int vlo = og.vlo, vhi = og.vhi;
          const int rlo = __reduce_min_sync(FULL, vlo), rhi = __reduce_max_sync(FULL, vhi);
          int tq = (bestk == kNone) ? 0x7FFFFFFF : (int)(bestk >> 19);
          // M tasks evaluate their hits COOPERATIVELY: a lane of such a task still has ~1 hit per 32 candidates, but the
          // lanes' counts differ widely, so private loops would run as long as the unluckiest lane's.  Hits go to a per-warp
          // queue at positions from a warp scan; 32 queued (origin, candidate) pairs are then evaluated at once, one per
          // lane, the origin's record coming from its owner lane by shuffle and the result going back through shared memory.
          uint32_t* sbest = S.sbest[warp];
          QEntry* queue = S.queue[warp];
          int qn = 0;  // queued hits (warp-uniform)
          if (kindM) { sbest[lane] = bestk; __syncwarp(); }
          auto flush = [&]() {
            const int n = min(qn, 32);
            qn -= n;
            const bool on = lane < n;
            K2CHECK(qn >= 0 && qn + n <= kQueue);
            const uint32_t e = on ? (uint32_t)queue[qn + lane] : 0u;
            const int owner = (int)(e >> kPosBits), p = (int)(e & ((1u << kPosBits) - 1u));
            uint4 oa, ob;
            oa.x = __shfl_sync(FULL, a.x, owner); oa.y = __shfl_sync(FULL, a.y, owner);
            oa.z = __shfl_sync(FULL, a.z, owner); oa.w = __shfl_sync(FULL, a.w, owner);
            ob.x = __shfl_sync(FULL, b.x, owner); ob.y = __shfl_sync(FULL, b.y, owner);
            ob.z = __shfl_sync(FULL, b.z, owner); ob.w = __shfl_sync(FULL, b.w, owner);
            const uint32_t olo = __shfl_sync(FULL, lo_p, owner), ohi = __shfl_sync(FULL, hi_p, owner);
            if (on) {
              const uint32_t k = cand_key(S, desc_to, p, oa, ob, olo, ohi);
              if (k < sbest[owner]) atomicMin(&sbest[owner], k);
            }
            if (lane == 0) K2STAT(7, 1);
            __syncwarp();
            const uint32_t sb = sbest[lane];
            if (sb < bestk) { bestk = sb; tq = (int)(sb >> 19); }
          };
          // rows in the order own row (of the first lane), then the rest, each claimed from the task's counter
          const int rown = min(max(__shfl_sync(FULL, og.v, 0), rlo), rhi - 1);
          const int nr = rhi - rlo;
          if (lane == 0) *(volatile int32_t*)&S.task_nr[item] = nr;
          for (;;) {
            int sq = 0;
            if (lane == 0) sq = atomicAdd(&S.task_ctr[item], 1);
            sq = __shfl_sync(FULL, sq, 0);
            if (sq >= nr) break;
            const int r = (sq == 0) ? rown : (rlo + sq - 1 >= rown ? rlo + sq : rlo + sq - 1);
            if (valid) {  // bounds found meanwhile by the other warps working on this task
              const uint32_t gk = *(volatile uint32_t*)&S.gkey[cc][ei];
              if (gk < bestk) { bestk = gk; tq = (int)(gk >> 19); }
            }
            const bool act = (r >= vlo) && (r < vhi);
            int pg = t.segOff[r][j0] >> 2;            // candidate groups of 4 (every bin list is padded to 4)
            const int pg1 = t.segOff[r][j1] >> 2;
            K2CHECK(pg >= 0 && pg1 * 4 <= kCap);
            // 32 candidates (8 groups) at a time: bias + Q has its sign bit set exactly when Q exceeds the lane's bound (lane not
            // searching this row: always; no bound yet: never), and one funnel shift per candidate collects the signs
            for (; pg < pg1; pg += 8) {
              const int nbw = min(8, pg1 - pg);
              const uint32_t bias = 0x7FFFFFFFu - (uint32_t)(act ? tq : -1);
              const uint4* qp = Q4 + pg;
              const uint4* rp = Q14 + pg;
              uint32_t m = 0;  // MISS bits, candidate i of the word in bit 31 - i once the word is complete
              if (nbw == 8) {
                // two independent 16-candidate chains (the funnel shifts of one chain depend on each other)
                uint32_t mA = 0, mB = 0;
                if (two) {
#pragma unroll
                  for (int i = 0; i < 4; i++) {
                    const uint4 x = qp[i], y = rp[i], x2 = qp[i + 4], y2 = rp[i + 4];
                    mA = __funnelshift_l(sad4(qb, y.x, sad4(qa, x.x, bias)), mA, 1);
                    mB = __funnelshift_l(sad4(qb, y2.x, sad4(qa, x2.x, bias)), mB, 1);
                    mA = __funnelshift_l(sad4(qb, y.y, sad4(qa, x.y, bias)), mA, 1);
                    mB = __funnelshift_l(sad4(qb, y2.y, sad4(qa, x2.y, bias)), mB, 1);
                    mA = __funnelshift_l(sad4(qb, y.z, sad4(qa, x.z, bias)), mA, 1);
                    mB = __funnelshift_l(sad4(qb, y2.z, sad4(qa, x2.z, bias)), mB, 1);
                    mA = __funnelshift_l(sad4(qb, y.w, sad4(qa, x.w, bias)), mA, 1);
                    mB = __funnelshift_l(sad4(qb, y2.w, sad4(qa, x2.w, bias)), mB, 1);
                  }
                } else {
#pragma unroll
                  for (int i = 0; i < 4; i++) {
                    const uint4 x = qp[i], x2 = qp[i + 4];
                    mA = __funnelshift_l(sad4(qa, x.x, bias), mA, 1);
                    mB = __funnelshift_l(sad4(qa, x2.x, bias), mB, 1);
                    mA = __funnelshift_l(sad4(qa, x.y, bias), mA, 1);
                    mB = __funnelshift_l(sad4(qa, x2.y, bias), mB, 1);
                    mA = __funnelshift_l(sad4(qa, x.z, bias), mA, 1);
                    mB = __funnelshift_l(sad4(qa, x2.z, bias), mB, 1);
                    mA = __funnelshift_l(sad4(qa, x.w, bias), mA, 1);
                    mB = __funnelshift_l(sad4(qa, x2.w, bias), mB, 1);
                  }
                }
                m = mA * 0x10000u + mB;  // both hold 16 bits
              } else if (two) {
#pragma unroll 1
                for (int i = 0; i < nbw; i++) {
                  const uint4 x = qp[i], y = rp[i];
                  m = __funnelshift_l(sad4(qb, y.x, sad4(qa, x.x, bias)), m, 1);
                  m = __funnelshift_l(sad4(qb, y.y, sad4(qa, x.y, bias)), m, 1);
                  m = __funnelshift_l(sad4(qb, y.z, sad4(qa, x.z, bias)), m, 1);
                  m = __funnelshift_l(sad4(qb, y.w, sad4(qa, x.w, bias)), m, 1);
                }
              } else {
#pragma unroll 1
                for (int i = 0; i < nbw; i++) {
                  const uint4 x = qp[i];
                  m = __funnelshift_l(sad4(qa, x.x, bias), m, 1);
                  m = __funnelshift_l(sad4(qa, x.y, bias), m, 1);
                  m = __funnelshift_l(sad4(qa, x.z, bias), m, 1);
                  m = __funnelshift_l(sad4(qa, x.w, bias), m, 1);
                }
              }
              uint32_t hit = ~m;
              if (nbw < 8) hit <<= 32 - 4 * nbw;  // nbw in 1..7: candidate 0 to bit 31, absent candidates never hit
              bool own = true;  // every lane evaluates its own hits (Z / N tasks: a couple per origin and scan)
              if (kindM) {
                // (1) a lane whose bound hardly prunes (none yet, or a flat patch that is close to everything) would keep the
                //     queue writes below running for its many hits while 31 lanes wait: the whole warp takes its word
                //     instead, lane j its candidate j
                int n = __popc(hit);
#ifdef VO_K2_STATS
                { const int mx = __reduce_max_sync(FULL, n), tot = __reduce_add_sync(FULL, n); const int nl = __popc(__ballot_sync(FULL, n > 0));
                  if (lane == 0) { K2STAT(4, 1); K2STAT(5, tot); K2STAT(6, mx); K2STAT(7, nl); } }
#endif
                uint32_t flood = __ballot_sync(FULL, n >= kFlood);
                while (flood != 0u) {
                  const int L = __ffs((int)flood) - 1;
                  flood &= flood - 1u;
                  const uint32_t mL = __shfl_sync(FULL, hit, L);
                  uint4 oa, ob;
                  oa.x = __shfl_sync(FULL, a.x, L); oa.y = __shfl_sync(FULL, a.y, L);
                  oa.z = __shfl_sync(FULL, a.z, L); oa.w = __shfl_sync(FULL, a.w, L);
                  ob.x = __shfl_sync(FULL, b.x, L); ob.y = __shfl_sync(FULL, b.y, L);
                  ob.z = __shfl_sync(FULL, b.z, L); ob.w = __shfl_sync(FULL, b.w, L);
                  const uint32_t olo = __shfl_sync(FULL, lo_p, L), ohi = __shfl_sync(FULL, hi_p, L);
                  uint32_t k = kNone;
                  if ((mL << lane) & 0x80000000u) { k = cand_key(S, desc_to, (pg << 2) + lane, oa, ob, olo, ohi); }
                  k = __reduce_min_sync(FULL, k);
                  if (lane == L) {
                    hit = 0u; n = 0;
                    if (k < bestk) { bestk = k; tq = (int)(k >> 19); }
                  }
                }
                // (2) the other lanes' hits join the queue at positions from a warp scan
                int incl = n;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                  const int up = __shfl_up_sync(FULL, incl, d);
                  if (lane >= d) incl += up;
                }
                const int total = __shfl_sync(FULL, incl, 31);
                own = qn + total > kQueue;  // lanes still without a bound hit everything: too many for the queue at once
                if (!own && total > 0) {
                  QEntry* qw = queue + (qn + incl - n);
                  const uint32_t ebase = ((uint32_t)lane << kPosBits) | (uint32_t)(pg << 2);
                  K2CHECK(qn + incl <= kQueue);
                  while (hit != 0u) {
                    const int z = __clz((int)hit);
                    hit &= ~(0x80000000u >> z);
                    *qw++ = (QEntry)(ebase + (uint32_t)z);
                  }
                  qn += total;
                  __syncwarp();
                  while (qn >= 32) flush();
                }
              }
#ifdef VO_K2_STATS
              if (!kindM) { const int n = __popc(hit); const int mx = __reduce_max_sync(FULL, n), tot = __reduce_add_sync(FULL, n);
                if (lane == 0) { K2STAT(8, 1); K2STAT(9, tot); K2STAT(10, mx); K2STAT(11, mx > 0 ? 1 : 0); K2STAT(12, two ? 1 : 0); } }
#endif
              if (own) {
                while (hit != 0u) {
                  const int z = __clz((int)hit);
                  hit &= ~(0x80000000u >> z);
                  const uint32_t k = cand_key(S, desc_to, (pg << 2) + z, a, b, lo_p, hi_p);
                  if (k < bestk) { bestk = k; tq = (int)(k >> 19); }
                }
              }
            }
            if (valid && bestk < *(volatile uint32_t*)&S.gkey[cc][ei]) atomicMin(&S.gkey[cc][ei], bestk);  // publish
          }
          if (lane == 0) K2STAT(13, 1);
          if (kindM) {
            __syncwarp();
            while (qn > 0) flush();
          }
          if (valid && bestk < *(volatile uint32_t*)&S.gkey[cc][ei]) atomicMin(&S.gkey[cc][ei], bestk);
        };
        for (int help = -1;;) {
          int item = 0;
          if (help < 0) {  // fresh tasks first
            if (lane == 0) item = atomicAdd(&S.next_item, 1);
            item = __shfl_sync(FULL, item, 0);
            if (item >= nFilterTasks) { help = 0; continue; }
          } else {         // then help with what is still running, longest (M) first
            for (; help < nFilterTasks; help++) {
              const int nr = *(volatile int32_t*)&S.task_nr[help];
              if (nr < 0 || *(volatile int32_t*)&S.task_ctr[help] < nr) break;  // rows of it are still unclaimed
            }
            if (help >= nFilterTasks) break;
            item = help++;
          }
          run_task(item);
        }
      }
      for (; !PRUNE;) {
        int item_raw = 0;
        if (lane == 0) item_raw = atomicAdd(&S.next_item, 1);
        item_raw = __shfl_sync(FULL, item_raw, 0);
        if (item_raw >= n_items) break;
        {
          const int item = item_raw;
          // ---------- P item (exhaustive mode): 32 origins x one window row, every candidate exactly ----------
          const int cc = (item < i5) ? 0 : 1;
          const int pi = item - (cc ? i5 : 0);
          const int tb = pi / nrows, r = ra + (pi - tb * nrows);
          const int ub = ub0 + cc;
          const int j0 = max(ub - 3, 0) - ulo, j1 = min(ub + 3, U - 1) - ulo;
          uint32_t* bo = best_dir + (int64_t)(ub * 4 + c) * g.col_cap;
          const int idx = tb * 32 + lane;
          const bool valid = idx < S.n_p[cc];
          const int ent = valid ? (int)S.plist[cc][idx] : 0;
          const int o = (cc ? sO1 : sO0) + (ent & (kListCap - 1));
          K2Origin og;
          load_origin(cc, o, valid, ent >> 9, og);
          const bool act = valid && r >= og.vlo && r < og.vhi;
          if (!__any_sync(FULL, act)) continue;
          const uint4 a = og.a, b = og.b;
          const uint32_t lo_p = og.lo_p, hi_p = og.hi_p;
          const int p0 = t.segOff[r][j0], p1 = t.segOff[r][j1];
          K2CHECK(p0 >= 0 && p1 <= kCap);
          uint32_t k = kNone;
#pragma unroll 2
          for (int p = p0; p < p1; p++) k = min(k, cand_key(S, desc_to, p, a, b, lo_p, hi_p));
          if (act && k != kNone) atomicMin(&bo[o], k - ((uint32_t)j0 << 16));
        }
      }
      if (PRUNE) {
        // every searched origin's key, from the shared bounds to best[] (origin-column form)
        __syncthreads();
#pragma unroll
        for (int cc = 0; cc < kOC; cc++) {
          const int ub = ub0 + cc, j0 = max(ub - 3, 0) - ulo;
          uint32_t* bo = best_dir + (int64_t)(ub * 4 + c) * g.col_cap;
          for (int i = tid, n = cc ? sN1 : sN0; i < n; i += THREADS) {
            if ((S.cls[cc][i] & 3) == 0) continue;
            const uint32_t k = S.gkey[cc][i];
            bo[(cc ? sO1 : sO0) + i] = (k == kNone) ? kNone : k - ((uint32_t)j0 << 16);
          }
        }
      }
    }
    first_chunk = false;
    ra = rb;
    if (ra >= row_hi) break;
  }
  return true;
}

template <bool PRUNE, int THREADS, bool LEAN>
__global__ void __launch_bounds__(THREADS, 2)
k2_match(Geom g, FrameArrays fa, const int4* __restrict__ edges, uint32_t* __restrict__ best, int rows_per_group,
         int ngroups, int seed_good, int mode) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  K2SmemT<LEAN>& S = *reinterpret_cast<K2SmemT<LEAN>*>(smem_raw);
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
    const bool ok = k2_columns<PRUNE, THREADS, LEAN>(S, g, fa, from_slot, to_slot, best_dir, c, it < 0 ? ub0 : ub0 + it,
                                               it < 0 ? noc : 1, grp, rows_per_group, ed.z, ed.w, seed_good, mode == 2,
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

template <bool PRUNE, int THREADS, bool LEAN>
static cudaError_t launch_k2_t(const Geom& g, const FrameArrays& fa, const int4* edges, int n_edges, uint32_t* best,
                               int rows_per_group, int seed_good, bool split, cudaStream_t st) {
  const int bytes = (int)sizeof(K2SmemT<LEAN>);
  const int ngroups = (g.V + rows_per_group - 1) / rows_per_group;
  const int cgroups = (g.U - 1 + kOC - 1) / kOC;
  const int budget = seed_good > 0 ? seed_good : kSeedGood;
  if (!split) {
    dim3 grid(cgroups * ngroups, 8, n_edges / 2);  // edges come in twos (mono: curr->prev, prev->curr; quad: four)
    k2_match<PRUNE, THREADS, LEAN><<<grid, THREADS, bytes, st>>>(g, fa, edges, best, rows_per_group, ngroups, budget, 0);
    return cudaGetLastError();
  }
  // mono pairs: prev -> curr in full (its CTAs also reset best[curr -> prev]), hints, then curr -> prev from the hints
  dim3 grid(cgroups * ngroups, 4, n_edges / 2), hgrid((g.U - 1) * 4, n_edges / 2);
  k2_match<PRUNE, THREADS, LEAN><<<grid, THREADS, bytes, st>>>(g, fa, edges, best, rows_per_group, ngroups, budget, 1);
  k2_hint_scatter<<<hgrid, KH_THREADS, 0, st>>>(g, fa, edges, best);
  k2_match<PRUNE, THREADS, LEAN><<<grid, THREADS, bytes, st>>>(g, fa, edges, best, rows_per_group, ngroups, budget, 2);
  return cudaGetLastError();
}

// staged candidates per chunk: a window row of one origin column (six bin lists, each padded to 4) must fit either layout
int k2_cand_cap() { return kCandCap; }

// per-device opt-in to K2's dynamic shared memory, every instantiation launch_k2 can pick (see init_k1_device)
cudaError_t init_k2_device() {
  const int fat = (int)sizeof(K2SmemT<false>), lean = (int)sizeof(K2SmemT<true>);
  cudaError_t e = cudaFuncSetAttribute(k2_match<false, 256, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fat);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 256, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fat);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 384, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fat);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 512, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fat);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<false, 256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lean);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k2_match<true, 512, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lean);
  return e;
}

// lean != 0: the staging layout for tall frames (descriptors stay in global memory, K2Cands<true>); built for the default
// thread count only
cudaError_t launch_k2(const Geom& g, const FrameArrays& fa, const int4* edges, int n_edges, uint32_t* best,
                      int rows_per_group, int prune, int threads, int seed_good, int split, int lean, cudaStream_t st) {
  // split != 0: the edges are mono pairs (curr->prev, prev->curr) and may run as two dependent launches
  if (lean) {
    if (!prune) return launch_k2_t<false, 256, true>(g, fa, edges, n_edges, best, rows_per_group, seed_good, false, st);
    return launch_k2_t<true, 512, true>(g, fa, edges, n_edges, best, rows_per_group, seed_good, split, st);
  }
  if (!prune) return launch_k2_t<false, 256, false>(g, fa, edges, n_edges, best, rows_per_group, seed_good, false, st);
  if (threads == 256) return launch_k2_t<true, 256, false>(g, fa, edges, n_edges, best, rows_per_group, seed_good, split, st);
  if (threads == 384) return launch_k2_t<true, 384, false>(g, fa, edges, n_edges, best, rows_per_group, seed_good, split, st);
  return launch_k2_t<true, 512, false>(g, fa, edges, n_edges, best, rows_per_group, seed_good, split, st);
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
  cudaError_t e = cudaMemcpyFromSymbol(out, vo::g_k2_stats, sizeof(unsigned long long) * 16);
  unsigned long long z[16] = {0};
  cudaMemcpyToSymbol(vo::g_k2_stats, z, sizeof(z));
  return (int)e;
}
#endif
