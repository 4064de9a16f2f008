/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of the reference's hot path.
 *
 * This file is a plain-C, runtime-sized restatement of the algorithm the reference executes in
 *   /root/reference/src/myComputeFeature.cpp   (feature extraction)
 *   /root/reference/src/myMatch.cpp            (circular prev<->curr matching)
 *   /root/reference/src/matcher.cpp:113-187    (LFSR shuffle + bucketing)
 * It is NOT part of the product: only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
 * legs may build, load or call it.  The shipped library (hls_final_visual_odometry_b200/csrc)
 * never links or falls back to it.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py checks every function here against the
 * reference's own sources compiled in place (oracle/_ref, built by oracle/build_ref.sh) on seeded
 * synthetic frames at 1024x284 (the reference's native, unmodified geometry), 1241x376, 2048x1024
 * and 500x150, bit for bit; tests/golden/ holds digests of those reference outputs so the same
 * check runs where /root/reference is absent.  The reference ships no executable golden vectors of
 * its own.
 *
 * Defined behaviour where the reference has undefined behaviour:
 *   - an origin with no candidate in its window gets the sentinel "none" (the reference leaves
 *     `min_cand` uninitialised, src/myMatch.cpp:79); it can never close a cycle;
 *   - the last u-bin column is never an origin nor a target (the reference re-processes a stale
 *     ring-buffer slot for it, src/myMatch.cpp:141-146, which finds no candidate);
 *   - v-bins that the reference never flushes (possible only for image heights where the last
 *     row of blocks does not fall in v-bin height/BIN_H) have count 0.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef struct vo_oracle_geom {
  int32_t width, height, pitch;
  int32_t bin_w, bin_h;  /* 60, 45: src/size.h:10-11 */
  int32_t u_bins, v_bins;
  int32_t cap;           /* stride of one (bin,class) list in records (MAX_FP_IN_BIN) */
  int32_t tau;           /* 50: src/myComputeFeature.cpp:157 */
} vo_oracle_geom;

typedef struct vo_oracle_match { /* == Matcher::p_match, src/matcher.h:89-104 */
  float u1p, v1p; int32_t i1p;
  float u2p, v2p; int32_t i2p;
  float u1c, v1c; int32_t i1c;
  float u2c, v2c; int32_t i2c;
} vo_oracle_match;

/* Geometry for a WxH image with the reference's bin size.  U = bin of the last block origin + 1
 * (block origins are i = 9, 12, ... < W-9, src/myComputeFeature.cpp:175,259); V = H/BIN_H + 1
 * (the trailing flush targets v_bin = IMG_H/BIN_H, src/myComputeFeature.cpp:337). */
void vo_oracle_geom_init(vo_oracle_geom* g, int32_t width, int32_t height, int32_t pitch) {
  g->width = width; g->height = height; g->pitch = pitch;
  g->bin_w = 60; g->bin_h = 45;
  int32_t last_i = 9 + 3 * ((width - 10 - 9) / 3);
  g->u_bins = last_i / g->bin_w + 1;
  g->v_bins = height / g->bin_h + 1;
  g->cap = (g->bin_w / 3) * (g->bin_h / 3);
  g->tau = 50;
}

/* ---------------------------------------------------------------------------------------------
 * Filters.  One generic 5x5 integer correlation; the four kernels are written as weight tables.
 * Only 2 <= x < w-2, 2 <= y < h-2 is produced (src/myComputeFeature.cpp:73-74,119-120,140-141);
 * the border is left at 0 here (the reference leaves it uninitialised and never reads it).
 * ------------------------------------------------------------------------------------------ */
static int32_t corr5x5_at(const uint8_t* in, int32_t pitch, int32_t x, int32_t y, const int8_t wt[25]) {
  int32_t acc = 0;
  for (int32_t r = 0; r < 5; r++)
    for (int32_t c = 0; c < 5; c++)
      acc += (int32_t)in[(y + r - 2) * pitch + (x + c - 2)] * wt[r * 5 + c];
  return acc;
}

static void outer5(const int8_t col[5], const int8_t row[5], int8_t wt[25]) {
  for (int r = 0; r < 5; r++) for (int c = 0; c < 5; c++) wt[r * 5 + c] = (int8_t)(col[r] * row[c]);
}

static const int8_t k_smooth[5] = {1, 4, 6, 4, 1};
static const int8_t k_deriv[5]  = {1, 2, 0, -2, -1};
static const int8_t k_check[5]  = {-1, -1, 0, 1, 1};

static uint8_t sobel_to_u8(int32_t acc) {
  /* int16 accumulate (|acc| <= 12240, no wrap), C division by 128 truncating toward zero,
   * +128, clamp: src/myComputeFeature.cpp:84-106 */
  int32_t q = acc / 128 + 128;
  return (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
}

void vo_oracle_sobel5x5(const uint8_t* in, int32_t w, int32_t h, int32_t pitch, uint8_t* du, uint8_t* dv) {
  int8_t w_du[25], w_dv[25];
  outer5(k_smooth, k_deriv, w_du);  /* v_weight, src/myComputeFeature.cpp:59-65 */
  outer5(k_deriv, k_smooth, w_dv);  /* h_weight, src/myComputeFeature.cpp:66-72 */
  memset(du, 0, (size_t)w * h); memset(dv, 0, (size_t)w * h);
  for (int32_t y = 2; y < h - 2; y++)
    for (int32_t x = 2; x < w - 2; x++) {
      du[y * w + x] = sobel_to_u8(corr5x5_at(in, pitch, x, y, w_du));
      dv[y * w + x] = sobel_to_u8(corr5x5_at(in, pitch, x, y, w_dv));
    }
}

void vo_oracle_blob5x5(const uint8_t* in, int32_t w, int32_t h, int32_t pitch, int16_t* out) {
  int8_t wt[25];  /* -1 ring, +1 inner ring, +8 centre: src/myComputeFeature.cpp:133-139 */
  for (int r = 0; r < 5; r++) for (int c = 0; c < 5; c++) {
    int ring = (abs(r - 2) > abs(c - 2)) ? abs(r - 2) : abs(c - 2);
    wt[r * 5 + c] = (int8_t)(ring == 2 ? -1 : (ring == 1 ? 1 : 8));
  }
  memset(out, 0, sizeof(int16_t) * (size_t)w * h);
  for (int32_t y = 2; y < h - 2; y++)
    for (int32_t x = 2; x < w - 2; x++)
      out[y * w + x] = (int16_t)corr5x5_at(in, pitch, x, y, wt);
}

void vo_oracle_checkerboard5x5(const uint8_t* in, int32_t w, int32_t h, int32_t pitch, int16_t* out) {
  int8_t wt[25];  /* src/myComputeFeature.cpp:112-118 == [1 1 0 -1 -1]^T (x) [-1 -1 0 1 1] */
  const int8_t col[5] = {1, 1, 0, -1, -1};
  outer5(col, k_check, wt);
  memset(out, 0, sizeof(int16_t) * (size_t)w * h);
  for (int32_t y = 2; y < h - 2; y++)
    for (int32_t x = 2; x < w - 2; x++)
      out[y * w + x] = (int16_t)corr5x5_at(in, pitch, x, y, wt);
}

/* ---------------------------------------------------------------------------------------------
 * Descriptor: 16 sample points, (du,dv) byte pair each; src/myComputeFeature.cpp:11-55.
 * Offsets are (dx,dy) relative to the feature pixel, in emission order.
 * ------------------------------------------------------------------------------------------ */
static const int8_t k_desc_dx[16] = {-3, -3, -1, -1,  3,  3,  1,  1, -1, -1,  1,  1, -5, -5,  5,  5};
static const int8_t k_desc_dy[16] = {-1,  1, -1,  1, -1,  1, -1,  1, -5,  5, -5,  5, -3,  3, -3,  3};

static void descriptor32(const uint8_t* du, const uint8_t* dv, int32_t w, int32_t u, int32_t v, uint8_t d[32]) {
  for (int s = 0; s < 16; s++) {
    int32_t a = (v + k_desc_dy[s]) * w + (u + k_desc_dx[s]);
    d[2 * s] = du[a];
    d[2 * s + 1] = dv[a];
  }
}

/* destination v-bin of block row j: rows are staged per bin row and flushed when (j+3)%BIN_H==0
 * with v_bin=j/BIN_H; whatever is still staged after the last row goes to v_bin=H/BIN_H
 * (src/myComputeFeature.cpp:319-345). */
static int32_t dest_v_bin(const vo_oracle_geom* g, int32_t j, int32_t j_last) {
  int32_t r = j / g->bin_h;
  int32_t flush_row = g->bin_h * (r + 1) - 3;
  return (flush_row <= j_last) ? r : g->height / g->bin_h;
}

/* NMS over 3x3 blocks + 5x5 validity + descriptor + per-(v_bin,u_bin,class) lists.
 * src/myComputeFeature.cpp:153-350.  Returns the number of features dropped for lack of capacity
 * (the reference has no capacity check; a non-zero return means "reference would overflow"). */
static int32_t nms_and_descriptors(const int16_t* f1, const int16_t* f2, const uint8_t* du, const uint8_t* dv,
                                   const vo_oracle_geom* g, int32_t* max2, int32_t* max2_num) {
  const int32_t w = g->width, h = g->height, U = g->u_bins, V = g->v_bins;
  int32_t dropped = 0;
  memset(max2_num, 0, sizeof(int32_t) * (size_t)U * V * 4);
  if (h - 9 <= 9 || w - 9 <= 9) return 0;
  const int32_t j_last = 9 + 3 * ((h - 10 - 9) / 3);
  for (int32_t j = 9; j < h - 9; j += 3) {
    const int32_t v_bin = dest_v_bin(g, j, j_last);
    for (int32_t i = 9; i < w - 9; i += 3) {
      const int32_t u_bin = i / g->bin_w;  /* from the block origin: src/myComputeFeature.cpp:259 */
      const int16_t* plane[2] = {f1, f2};
      for (int32_t cls = 0; cls < 4; cls++) {
        const int16_t* f = plane[cls >> 1];
        const int is_max = cls & 1;
        /* 3x3 scan, raster order, strict comparison => first extremum wins ties (:196-226) */
        int32_t best = is_max ? -32767 : 32767, bu = 0, bv = 0;
        for (int32_t y = j; y < j + 3; y++)
          for (int32_t x = i; x < i + 3; x++) {
            int32_t val = f[y * w + x];
            if (is_max ? (val > best) : (val < best)) { best = val; bu = x; bv = y; }
          }
        /* 5x5 validity around the extremum, not clipped (:233-256) */
        int valid = 1;
        for (int32_t y = bv - 2; y <= bv + 2 && valid; y++)
          for (int32_t x = bu - 2; x <= bu + 2; x++) {
            int32_t val = f[y * w + x];
            if (is_max ? (val > best) : (val < best)) { valid = 0; break; }
          }
        if (!valid) continue;
        if (is_max ? (best < g->tau) : (best > -g->tau)) continue;  /* :261,276,290,304 */
        const int32_t bin = (v_bin * U + u_bin) * 4 + cls;
        if (max2_num[bin] >= g->cap) { dropped++; continue; }
        int32_t* rec = max2 + ((size_t)bin * g->cap + max2_num[bin]) * 12;
        rec[0] = bu; rec[1] = bv; rec[2] = 0; rec[3] = cls;
        descriptor32(du, dv, w, bu, bv, (uint8_t*)(rec + 4));
        max2_num[bin]++;
      }
    }
  }
  return dropped;
}

/* myComputeFeatures, src/myComputeFeature.cpp:352-379.  max2: int32[V*U*4*cap*12], max2_num: int32[V*U*4]. */
int32_t vo_oracle_compute_features(const uint8_t* img, const vo_oracle_geom* g, int32_t* max2, int32_t* max2_num) {
  const size_t n = (size_t)g->width * g->height;
  uint8_t* du = (uint8_t*)malloc(n); uint8_t* dv = (uint8_t*)malloc(n);
  int16_t* f1 = (int16_t*)malloc(2 * n); int16_t* f2 = (int16_t*)malloc(2 * n);
  vo_oracle_sobel5x5(img, g->width, g->height, g->pitch, du, dv);
  vo_oracle_blob5x5(img, g->width, g->height, g->pitch, f1);
  vo_oracle_checkerboard5x5(img, g->width, g->height, g->pitch, f2);
  int32_t dropped = nms_and_descriptors(f1, f2, du, dv, g, max2, max2_num);
  free(du); free(dv); free(f1); free(f2);
  return dropped;
}

/* ---------------------------------------------------------------------------------------------
 * Circular matching, src/myMatch.cpp:17-204, restated without the 7-slot column ring buffer
 * (SURVEY.md appendix A).  A feature is addressed by (u_bin, class, col_idx) where col_idx is its
 * position in the concatenation over v_bin = 0..V-1 of the (v_bin,u_bin,class) lists
 * (fetch_col_bin, src/myMatch.cpp:29-66).
 * ------------------------------------------------------------------------------------------ */
typedef struct col_list {
  int32_t n;
  int32_t* prefix;       /* V+1 prefix sums over v_bin (mc_buffer_num) */
  const int32_t** rec;   /* n pointers to 12-int records */
  int32_t* best_ubin;    /* best target per origin: u_bin, col_idx, or -1 */
  int32_t* best_idx;
} col_list;

static void build_columns(const vo_oracle_geom* g, const int32_t* m, const int32_t* num, col_list* cols) {
  const int32_t U = g->u_bins, V = g->v_bins;
  for (int32_t u = 0; u < U; u++)
    for (int32_t c = 0; c < 4; c++) {
      col_list* L = &cols[u * 4 + c];
      L->prefix = (int32_t*)malloc(sizeof(int32_t) * (V + 1));
      L->prefix[0] = 0;
      for (int32_t v = 0; v < V; v++) L->prefix[v + 1] = L->prefix[v] + num[(v * U + u) * 4 + c];
      L->n = L->prefix[V];
      L->rec = (const int32_t**)malloc(sizeof(int32_t*) * (L->n + 1));
      L->best_ubin = (int32_t*)malloc(sizeof(int32_t) * (L->n + 1));
      L->best_idx = (int32_t*)malloc(sizeof(int32_t) * (L->n + 1));
      int32_t k = 0;
      for (int32_t v = 0; v < V; v++) {
        int32_t bin = (v * U + u) * 4 + c;
        for (int32_t e = 0; e < num[bin]; e++) L->rec[k++] = m + ((size_t)bin * g->cap + e) * 12;
      }
    }
}

static void free_columns(const vo_oracle_geom* g, col_list* cols) {
  for (int32_t i = 0; i < g->u_bins * 4; i++) {
    free(cols[i].prefix); free((void*)cols[i].rec); free(cols[i].best_ubin); free(cols[i].best_idx);
  }
}

static int32_t sad32(const uint8_t* a, const uint8_t* b) {
  int32_t s = 0;
  for (int i = 0; i < 32; i++) s += abs((int32_t)a[i] - (int32_t)b[i]);
  return s;
}

/* One direction: every origin column u < U-1 of `from` searches `to` (find_match, :70-119, driven
 * by the per-column loops :148-176).  stats[0] += SAD evaluations, stats[1] += origins without any
 * candidate (outside the dead column), stats[2] += candidates range-tested. */
static void match_direction(const vo_oracle_geom* g, col_list* from, const col_list* to, int64_t* stats) {
  const int32_t U = g->u_bins, V = g->v_bins;
  const int32_t rad_u = g->bin_w * 3, rad_v = g->bin_h * 3;  /* src/size.h:14-15 */
  for (int32_t ub = 0; ub < U; ub++)
    for (int32_t c = 0; c < 4; c++) {
      col_list* L = &from[ub * 4 + c];
      int32_t vb = 0;  /* v_buffer_idx: bumped at most once per feature (:153-158) */
      for (int32_t k = 0; k < L->n; k++) {
        if (k >= L->prefix[vb]) vb++;
        L->best_ubin[k] = -1; L->best_idx[k] = -1;
        if (ub == U - 1) continue;  /* dead column */
        const int32_t* o = L->rec[k];
        const int32_t ub_lo = ub - 3 < 0 ? 0 : ub - 3;
        const int32_t ub_hi = ub + 3 > U - 1 ? U - 1 : ub + 3;  /* exclusive (:97) */
        const int32_t vb_lo = vb - 3 < 0 ? 0 : vb - 3;
        const int32_t vb_hi = vb + 3 > V - 1 ? V - 1 : vb + 3;  /* exclusive via prefix (:99) */
        int32_t best = 32766;
        for (int32_t tu = ub_lo; tu < ub_hi; tu++) {
          const col_list* T = &to[tu * 4 + c];
          for (int32_t t = T->prefix[vb_lo]; t < T->prefix[vb_hi]; t++) {
            const int32_t* q = T->rec[t];
            stats[2]++;
            if (q[0] < o[0] - rad_u || q[0] > o[0] + rad_u || q[1] < o[1] - rad_v || q[1] > o[1] + rad_v) continue;
            int32_t s = sad32((const uint8_t*)(o + 4), (const uint8_t*)(q + 4));
            stats[0]++;
            if (s < best) { best = s; L->best_ubin[k] = tu; L->best_idx[k] = t; }  /* strict: first wins */
          }
        }
        if (L->best_ubin[k] < 0) stats[1]++;
      }
    }
}

/* myMatching, src/myMatch.cpp:122-204.  Returns 0, or 1 if `out_cap` was too small.
 * stats: int64[3], see match_direction. */
int32_t vo_oracle_matching(const vo_oracle_geom* g, const int32_t* m1p, const int32_t* n1p, const int32_t* m1c,
                           const int32_t* n1c, vo_oracle_match* out, int32_t out_cap, int32_t* n_out, int64_t* stats) {
  const int32_t U = g->u_bins;
  int64_t local_stats[3] = {0, 0, 0};
  if (!stats) stats = local_stats;
  col_list* cur = (col_list*)calloc((size_t)U * 4, sizeof(col_list));
  col_list* prv = (col_list*)calloc((size_t)U * 4, sizeof(col_list));
  build_columns(g, m1c, n1c, cur);
  build_columns(g, m1p, n1p, prv);
  match_direction(g, cur, prv, stats);
  match_direction(g, prv, cur, stats);

  uint8_t* taken = (uint8_t*)calloc((size_t)g->width * g->height, 1);  /* M[][] (:133) */
  int32_t n = 0, overflow = 0;
  for (int32_t u = 0; u < U; u++)
    for (int32_t c = 0; c < 4; c++) {
      const col_list* L = &cur[u * 4 + c];
      for (int32_t k = 0; k < L->n; k++) {
        if (L->best_ubin[k] < 0) continue;
        const col_list* P = &prv[L->best_ubin[k] * 4 + c];
        const int32_t pk = L->best_idx[k];
        if (P->best_ubin[pk] != u || P->best_idx[pk] != k) continue;  /* cycle closes? (:186) */
        const int32_t* fc = L->rec[k];
        const int32_t* fp = P->rec[pk];
        uint8_t* flag = &taken[(size_t)fc[1] * g->width + fc[0]];
        if (*flag) continue;  /* current pixel already emitted by a lower class (:188) */
        *flag = 1;
        if (n < out_cap) {
          vo_oracle_match m;
          m.u1p = (float)fp[0]; m.v1p = (float)fp[1]; m.i1p = -1;
          m.u2p = -1.f; m.v2p = -1.f; m.i2p = -1;
          m.u1c = (float)fc[0]; m.v1c = (float)fc[1]; m.i1c = -1;
          m.u2c = -1.f; m.v2c = -1.f; m.i2c = -1;
          out[n] = m;
        } else {
          overflow = 1;
        }
        n++;
      }
    }
  *n_out = (n < 5) ? 0 : n;  /* :198-203 */
  free(taken);
  free_columns(g, cur); free_columns(g, prv);
  free(cur); free(prv);
  return overflow;
}

/* ---------------------------------------------------------------------------------------------
 * Bucketing, src/matcher.cpp:113-187.
 * rand_number is an LFSR written with double pow/floor; its tap-32 term is int(double(x)) which,
 * for x >= 2^31, is an out-of-range conversion and yields INT_MIN (bit 0 == 0) on x86
 * (cvttsd2si).  Bit model, validated against the compiled reference.
 * ------------------------------------------------------------------------------------------ */
uint32_t vo_oracle_rand_number(uint32_t x) {
  uint32_t t0 = (x >= 0x80000000u) ? 0u : (x & 1u);
  uint32_t b = (t0 ^ (x >> 10) ^ (x >> 30) ^ (x >> 31)) & 1u;
  return (x >> 1) | (b << 31);
}

/* in-place; list has *n entries on entry and the bucketed entries on exit */
void vo_oracle_bucket_features(vo_oracle_match* list, int32_t* n, int32_t max_features, float bucket_w, float bucket_h) {
  const int32_t cnt = *n;
  float u_max = 0, v_max = 0;
  for (int32_t i = 0; i < cnt; i++) {
    if (list[i].u1c > u_max) u_max = list[i].u1c;
    if (list[i].v1c > v_max) v_max = list[i].v1c;
  }
  const int32_t cols = (int32_t)floorf(u_max / bucket_w) + 1;
  const int32_t rows = (int32_t)floorf(v_max / bucket_h) + 1;
  const int32_t nb = cols * rows;
  int32_t* id = (int32_t*)malloc(sizeof(int32_t) * (cnt + 1));
  int32_t* start = (int32_t*)calloc((size_t)nb + 1, sizeof(int32_t));
  for (int32_t i = 0; i < cnt; i++) {
    int32_t bu = (int32_t)floorf(list[i].u1c / bucket_w);
    int32_t bv = (int32_t)floorf(list[i].v1c / bucket_h);
    id[i] = bv * cols + bu;
    start[id[i] + 1]++;
  }
  for (int32_t b = 0; b < nb; b++) start[b + 1] += start[b];
  /* stable distribution: order inside a bucket == input order (:161-166) */
  vo_oracle_match* sorted = (vo_oracle_match*)malloc(sizeof(vo_oracle_match) * (cnt + 1));
  int32_t* fill = (int32_t*)calloc((size_t)nb + 1, sizeof(int32_t));
  for (int32_t i = 0; i < cnt; i++) sorted[start[id[i]] + fill[id[i]]++] = list[i];
  uint32_t rnd = 5;  /* restarts at 5 on every call, carries across buckets (:170) */
  int32_t out = 0;
  for (int32_t b = 0; b < nb; b++) {
    vo_oracle_match* p = sorted + start[b];
    const int32_t len = start[b + 1] - start[b];
    for (int32_t i = 1; i < len; i++) {  /* random_shuffle (:126-138) */
      int32_t j = (int32_t)(rnd % (uint32_t)(i + 1));
      rnd = vo_oracle_rand_number(rnd);
      vo_oracle_match t = p[i]; p[i] = p[j]; p[j] = t;
    }
    /* the refill loop appends first and tests `k>=max_features` afterwards (:178-184), so a
     * non-empty bucket always contributes at least one match */
    const int32_t keep = max_features < 1 ? 1 : max_features;
    for (int32_t k = 0; k < len && k < keep; k++) list[out++] = p[k];
  }
  *n = out;
  free(id); free(start); free(sorted); free(fill);
}
