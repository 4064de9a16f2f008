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
  int32_t rad_u, rad_v;  /* SEARCH_RAD_U / SEARCH_RAD_V: src/size.h:14-15 (BIN_W*3, BIN_H*3) */
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
void vo_oracle_geom_init_bins(vo_oracle_geom* g, int32_t width, int32_t height, int32_t pitch, int32_t bin_w, int32_t bin_h) {
  g->width = width; g->height = height; g->pitch = pitch;
  g->bin_w = bin_w; g->bin_h = bin_h;  /* BIN_W, BIN_H: src/size.h:10-11 (multiples of 3: the flush rule of :319 needs it) */
  int32_t last_i = 9 + 3 * ((width - 10 - 9) / 3);
  g->u_bins = last_i / g->bin_w + 1;
  g->v_bins = height / g->bin_h + 1;
  g->cap = (g->bin_w / 3) * (g->bin_h / 3);
  g->rad_u = g->bin_w * 3; g->rad_v = g->bin_h * 3;  /* src/size.h:14-15 */
  g->tau = 50;
}

void vo_oracle_geom_init(vo_oracle_geom* g, int32_t width, int32_t height, int32_t pitch) {
  vo_oracle_geom_init_bins(g, width, height, pitch, 60, 45);
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
static void match_direction(const vo_oracle_geom* g, col_list* from, const col_list* to, int64_t* stats, int32_t rad_v) {
  const int32_t U = g->u_bins, V = g->v_bins;
  const int32_t rad_u = g->rad_u;  /* SEARCH_RAD_U, src/size.h:14; rad_v = SEARCH_RAD_V (src/size.h:15) for flow edges */
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
  match_direction(g, cur, prv, stats, g->rad_v);
  match_direction(g, prv, cur, stats, g->rad_v);

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

/* =============================================================================================
 * EXTENSION -- PARITY UNPINNED.
 *
 * Quad (left/right x prev/curr) circular matching and sub-pixel refinement are named by the
 * reference's API (Matcher::pushBack(I1,I2,..), matchFeatures(method=2), parameters
 * match_disp_tolerance / refinement, p_match fields u2p..i2c: src/matcher.h:45-56,89-104;
 * VisualOdometryStereo::process calls matchFeatures(2): src/viso_stereo.cpp:33-52) but NOT implemented by it:
 * pushBack ignores I2, matchFeatures ignores `method`, `refinement` is never read.  What follows is OUR
 * restatement, built from
 *   - the reference's own search (find_match, src/myMatch.cpp:70-119, as match_direction above) for every
 *     edge of the cycle, with the window |dv| <= match_disp_tolerance on left<->right edges (the meaning
 *     src/matcher.h:51 gives the field; libviso2's published rule), |dv| <= SEARCH_RAD_V on flow edges;
 *   - libviso2's published quad rule (Geiger et al., libviso2 matcher.cpp `matching`, method 2): the cycle
 *     1c -> 1p -> 2p -> 2c -> 1c must close and both disparities must be non-negative;
 *   - the reference's emission rules (order, same-pixel dedupe, "<5 -> 0": src/myMatch.cpp:181-203);
 *   - libviso2's published parabolicFitting on the 16-byte "small" descriptor, whose definition the
 *     reference does carry (computeSmallDescriptor, src/matcher.cpp:516-543; margin = 7, src/matcher.cpp:38),
 *     over the Sobel responses of the reference's executed filter (mySobel5x5).
 * There is no reference output to pin these against; tests compare the CUDA path with THIS code only and
 * report the results separately from the pinned mono path.
 * ============================================================================================= */

/* matchFeatures(method = 2).  Returns 0, or 1 if out_cap was too small. */
int32_t vo_oracle_quad_matching(const vo_oracle_geom* g, const int32_t* m1p, const int32_t* n1p, const int32_t* m2p,
                                const int32_t* n2p, const int32_t* m1c, const int32_t* n1c, const int32_t* m2c,
                                const int32_t* n2c, int32_t disp_tolerance, vo_oracle_match* out, int32_t out_cap,
                                int32_t* n_out) {
  const int32_t U = g->u_bins;
  int64_t stats[3] = {0, 0, 0};
  col_list* c1 = (col_list*)calloc((size_t)U * 4, sizeof(col_list));
  col_list* p1 = (col_list*)calloc((size_t)U * 4, sizeof(col_list));
  col_list* p2 = (col_list*)calloc((size_t)U * 4, sizeof(col_list));
  col_list* c2 = (col_list*)calloc((size_t)U * 4, sizeof(col_list));
  build_columns(g, m1c, n1c, c1); build_columns(g, m1p, n1p, p1);
  build_columns(g, m2p, n2p, p2); build_columns(g, m2c, n2c, c2);
  match_direction(g, c1, p1, stats, g->rad_v);   /* flow   1c -> 1p */
  match_direction(g, p1, p2, stats, disp_tolerance); /* stereo 1p -> 2p */
  match_direction(g, p2, c2, stats, g->rad_v);   /* flow   2p -> 2c */
  match_direction(g, c2, c1, stats, disp_tolerance); /* stereo 2c -> 1c */

  uint8_t* taken = (uint8_t*)calloc((size_t)g->width * g->height, 1);
  int32_t n = 0, overflow = 0;
  for (int32_t u = 0; u < U; u++)
    for (int32_t c = 0; c < 4; c++) {
      const col_list* L = &c1[u * 4 + c];
      for (int32_t k = 0; k < L->n; k++) {
        if (L->best_ubin[k] < 0) continue;
        const col_list* A = &p1[L->best_ubin[k] * 4 + c]; const int32_t ak = L->best_idx[k];
        if (A->best_ubin[ak] < 0) continue;
        const col_list* B = &p2[A->best_ubin[ak] * 4 + c]; const int32_t bk = A->best_idx[ak];
        if (B->best_ubin[bk] < 0) continue;
        const col_list* D = &c2[B->best_ubin[bk] * 4 + c]; const int32_t dk = B->best_idx[bk];
        if (D->best_ubin[dk] != u || D->best_idx[dk] != k) continue;   /* circle closes */
        const int32_t* f1c = L->rec[k]; const int32_t* f1p = A->rec[ak];
        const int32_t* f2p = B->rec[bk]; const int32_t* f2c = D->rec[dk];
        if (f1p[0] < f2p[0] || f1c[0] < f2c[0]) continue;               /* disparities >= 0 */
        uint8_t* flag = &taken[(size_t)f1c[1] * g->width + f1c[0]];
        if (*flag) continue;
        *flag = 1;
        if (n < out_cap) {
          vo_oracle_match m;
          m.u1p = (float)f1p[0]; m.v1p = (float)f1p[1]; m.i1p = -1;
          m.u2p = (float)f2p[0]; m.v2p = (float)f2p[1]; m.i2p = -1;
          m.u1c = (float)f1c[0]; m.v1c = (float)f1c[1]; m.i1c = -1;
          m.u2c = (float)f2c[0]; m.v2c = (float)f2c[1]; m.i2c = -1;
          out[n] = m;
        } else {
          overflow = 1;
        }
        n++;
      }
    }
  *n_out = (n < 5) ? 0 : n;
  free(taken);
  free_columns(g, c1); free_columns(g, p1); free_columns(g, p2); free_columns(g, c2);
  free(c1); free(p1); free(p2); free(c2);
  return overflow;
}

/* computeSmallDescriptor, src/matcher.cpp:516-543: 12 du + 4 dv samples (the centre du twice) */
static const int8_t k_small_dx[16] = {0, -2, 0, 2, -1, 0, 0, 1, -2, 0, 2, 0,  0, -1, 1, 0};
static const int8_t k_small_dy[16] = {-2, -1, -1, -1, 0, 0, 0, 0, 1, 1, 1, 2, -1, 0, 0, 1};

static void small_descriptor(const uint8_t* du, const uint8_t* dv, int32_t w, int32_t u, int32_t v, uint8_t d[16]) {
  for (int s = 0; s < 16; s++) {
    const int32_t a = (v + k_small_dy[s]) * w + (u + k_small_dx[s]);
    d[s] = (s < 12) ? du[a] : dv[a];
  }
}

/* parabolicFitting (libviso2): 5x5 SAD costs of the anchor's small descriptor against the target shifted by
 * (-2..2)^2, minimum on the border -> fail, least-squares quadric c ~ b0 du^2 + b1 dv^2 + b2 du dv + b3 du +
 * b4 dv + b5 through the 3x3 costs around the minimum, stationary point -> (ddu, ddv), |.| >= 1 -> fail.
 * The 9x6 design of libviso2's `refinement` has closed-form normal equations; with integer costs
 *   6 b0 = B0 = 3 P - 2 R,  6 b1 = B1 = 3 Q - 2 R,  4 b2 = B2,  6 b3 = B3,  6 b4 = B4
 * (P = sum du^2 c, Q = sum dv^2 c, R = sum c, B2 = sum du dv c, B3 = sum du c, B4 = sum dv c), so
 *   ddv = NV / D,  ddu = -2 (B4 D + 2 B1 NV) / (3 B2 D),   D = 9 B2^2 - 16 B0 B1,  NV = 8 B0 B4 - 6 B2 B3
 * are ratios of integers: one IEEE double division each, rounded to float -- the same bits on any machine.
 * libviso2's tests |divisor| < 1e-8, |b2| < 1e-8 become D == 0, B2 == 0. */
static int parabolic_fit(const uint8_t* du1, const uint8_t* dv1, const uint8_t* du2, const uint8_t* dv2, int32_t w,
                         int32_t h, float u1, float v1, float* u2, float* v2) {
  const int32_t margin = 7;  /* src/matcher.cpp:38 */
  if (*u2 - 2 < margin || *u2 + 2 > w - 1 - margin || *v2 - 2 < margin || *v2 + 2 > h - 1 - margin) return 0;
  uint8_t d1[16], d2[16];
  small_descriptor(du1, dv1, w, (int32_t)u1, (int32_t)v1, d1);
  int32_t cost[25], min_ind = 0;
  for (int32_t b = 0; b < 5; b++)
    for (int32_t a = 0; a < 5; a++) {
      small_descriptor(du2, dv2, w, (int32_t)*u2 + a - 2, (int32_t)*v2 + b - 2, d2);
      int32_t s = 0;
      for (int i = 0; i < 16; i++) s += abs((int32_t)d1[i] - (int32_t)d2[i]);
      cost[b * 5 + a] = s;
    }
  for (int32_t i = 1; i < 25; i++) if (cost[i] < cost[min_ind]) min_ind = i;
  const int32_t a0 = min_ind % 5, b0 = min_ind / 5;
  if (a0 == 0 || a0 == 4 || b0 == 0 || b0 == 4) return 0;
  int64_t P = 0, Q = 0, R = 0, B2 = 0, B3 = 0, B4 = 0;
  for (int32_t i = -1; i <= 1; i++)
    for (int32_t j = -1; j <= 1; j++) {
      const int64_t c = cost[(b0 + i) * 5 + (a0 + j)];
      P += j * j * c; Q += i * i * c; R += c; B2 += i * j * c; B3 += j * c; B4 += i * c;
    }
  const int64_t B0 = 3 * P - 2 * R, B1 = 3 * Q - 2 * R;
  const int64_t D = 9 * B2 * B2 - 16 * B0 * B1, NV = 8 * B0 * B4 - 6 * B2 * B3;
  if (D == 0 || B2 == 0) return 0;
  const float ddv = (float)((double)NV / (double)D);
  const float ddu = (float)((double)(-2 * (B4 * D + 2 * B1 * NV)) / (double)(3 * B2 * D));
  if (fabsf(ddu) >= 1.0f || fabsf(ddv) >= 1.0f) return 0;
  *u2 = (float)((double)*u2 + ((double)(a0 - 2) + (double)ddu));
  *v2 = (float)((double)*v2 + ((double)(b0 - 2) + (double)ddv));
  return 1;
}

/* refinement(p_matched, method) with param.refinement == 2 (libviso2): the current left feature is the anchor;
 * 1p (flow), then -- quad lists only, i.e. u2c >= 0 -- 2c and 2p are refined against it; a match any fit
 * rejects is dropped.  In place, order preserved.  img_2p / img_2c may be NULL for flow lists. */
void vo_oracle_refine(const vo_oracle_geom* g, const uint8_t* img_1p, const uint8_t* img_2p, const uint8_t* img_1c,
                      const uint8_t* img_2c, vo_oracle_match* list, int32_t* n) {
  const int32_t w = g->width, h = g->height;
  const uint8_t* imgs[4] = {img_1p, img_2p, img_1c, img_2c};
  uint8_t* du[4] = {0, 0, 0, 0};
  uint8_t* dv[4] = {0, 0, 0, 0};
  for (int i = 0; i < 4; i++)
    if (imgs[i]) {
      du[i] = (uint8_t*)malloc((size_t)w * h); dv[i] = (uint8_t*)malloc((size_t)w * h);
      vo_oracle_sobel5x5(imgs[i], w, h, g->pitch, du[i], dv[i]);
    }
  int32_t kept = 0;
  for (int32_t i = 0; i < *n; i++) {
    vo_oracle_match m = list[i];
    int ok = parabolic_fit(du[2], dv[2], du[0], dv[0], w, h, m.u1c, m.v1c, &m.u1p, &m.v1p);
    if (ok && m.u2c >= 0.f && du[3] && du[1]) {
      ok = parabolic_fit(du[2], dv[2], du[3], dv[3], w, h, m.u1c, m.v1c, &m.u2c, &m.v2c) &&
           parabolic_fit(du[2], dv[2], du[1], dv[1], w, h, m.u1c, m.v1c, &m.u2p, &m.v2p);
    }
    if (ok) list[kept++] = m;
  }
  *n = kept;
  for (int i = 0; i < 4; i++) { free(du[i]); free(dv[i]); }
}
