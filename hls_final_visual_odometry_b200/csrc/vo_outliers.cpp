// Host stage between matching and bucketing: the Delaunay "support" filter of the reference.
//
// Replaces, result-for-result,
//   removeOutliers            src/remove_outliers.cpp:4-94
//   delaunator::Delaunator    src/delaunator.cpp:178-612, src/delaunator.hpp:78-140
// the triangulation is an incremental sweep-hull (inherently sequential,
// a few ms per frame) and its result on integer-lattice input depends on every float rounding and
// on the visiting order, so it stays a host C++ stage and follows the reference's arithmetic to the
// letter: single precision (src/delaunator.hpp:15) except where the reference mixes in double
// literals, first-wins strict comparisons, the same hash probe / hull walk / flip order.  No FMA
// contraction (this file is compiled with -ffp-contract=off).  What differs is engineering only:
// growable arrays instead of POINT_L/TRIANGLE_L fixed ones, 32-bit hull links instead of the
// int16 links overlaid on the distance array (src/delaunator.hpp:16-22; identical below 32768
// points), a stable O(n log n) sort instead of the O(n^2) insertion sort (same order: both are
// stable on the same float keys), a growable flip stack instead of EDGE_L = 13 entries, and
// defined behaviour (no triangles) for degenerate inputs on which the reference indexes out of
// bounds.
#include "../../include/vo_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>

namespace vo_host {

typedef float real;  // src/delaunator.hpp:15
static const int kNil = 2147483647;
static const real kInf = std::numeric_limits<real>::infinity();  // (float)1.797e308, src/delaunator.cpp:14

struct Pt { real x, y; };

static inline Pt sub(const Pt& a, const Pt& b) { Pt r; r.x = b.x - a.x; r.y = b.y - a.y; return r; }  // vector(a,b)
static inline real mag2(const Pt& p) { return p.x * p.x + p.y * p.y; }
static inline real cross(const Pt& a, const Pt& b) { return a.x * b.y - a.y * b.x; }
static inline real dist2(const Pt& a, const Pt& b) { return mag2(sub(a, b)); }

// src/delaunator.hpp:60-67: "same point" test relative to the bounding-box diagonal
static inline bool same_point(const Pt& a, const Pt& b, real span) {
  const real d2 = dist2(a, b);
  // shortcuts that decide the quotient test without the division (every hull-walk step runs it twice): 0 / span is 0, and
  // d2 > span * 1e-19 puts the rounded quotient at 0.99e-19 or more (span * 1e-19f underflowing to 0 only widens the margin)
  if (d2 == 0.0f) return span == span && span != 0.0f ? true : (d2 / span) < 1e-20;
  if (d2 > span * 1e-19f) return false;
  const real d = d2 / span;
  return d < 1e-20;
}

// src/delaunator.cpp:23-39.  The 0.5 is a double literal: numerator*0.5/det is evaluated in double
// and rounded to float once per component.
static inline real circumradius2(const Pt& p1, const Pt& p2, const Pt& p3) {
  const Pt d = sub(p1, p2), e = sub(p1, p3);
  const real bl = mag2(d), cl = mag2(e), det = cross(d, e);
  Pt r;
  r.x = (real)((e.y * bl - d.y * cl) * 0.5 / det);
  r.y = (real)((d.x * cl - e.x * bl) * 0.5 / det);
  if ((bl > 0.0 || bl < 0.0) && (cl > 0.0 || cl < 0.0) && (det > 0.0 || det < 0.0)) return mag2(r);
  return kInf;
}

// src/delaunator.cpp:91-103 (and :105-113)
static inline bool ccw(const Pt& p0, const Pt& p1, const Pt& p2) {
  const Pt v0 = sub(p0, p1), v1 = sub(p0, p2);
  const real det = cross(v0, v1);
  const real dist = mag2(v0) + mag2(v1);
  if (det == 0) return false;
  // dist <= |det| * 1e13 keeps the rounded quotient below 1.0000002e13 < 1e14: no division needed (always the case on an
  // integer lattice, where |det| >= 1 and dist < 2^25)
  if (!(dist <= std::fabs(det) * 1e13f)) {
    const real rel = std::fabs(dist / det);
    if (rel > 1e14) return false;
  }
  return det > 0;
}

// src/delaunator.cpp:116-137
static inline Pt circumcenter(const Pt& a, const Pt& b, const Pt& c) {
  const real dx = b.x - a.x, dy = b.y - a.y, ex = c.x - a.x, ey = c.y - a.y;
  const real bl = dx * dx + dy * dy, cl = ex * ex + ey * ey;
  const real d = dx * ey - dy * ex;
  Pt r;
  r.x = (real)(a.x + (ey * bl - dy * cl) * 0.5 / d);
  r.y = (real)(a.y + (dx * cl - ex * bl) * 0.5 / d);
  return r;
}

// src/delaunator.cpp:139-162: is p inside the circumcircle of (a,b,c)?  All single precision.
static inline bool in_circle(const Pt& a, const Pt& b, const Pt& c, const Pt& p) {
  const real dx = a.x - p.x, dy = a.y - p.y, ex = b.x - p.x, ey = b.y - p.y, fx = c.x - p.x, fy = c.y - p.y;
  const real ap = dx * dx + dy * dy, bp = ex * ex + ey * ey, cp = fx * fx + fy * fy;
  return (dx * (ey * cp - bp * fy) - dy * (ex * cp - bp * fx) + ap * (ex * fy - ey * fx)) < 0.0;
}

// src/delaunator.cpp:166-169: monotone in the angle; the constants are double literals
static inline real pseudo_angle(real dx, real dy) {
  const real p = dx / (std::fabs(dx) + std::fabs(dy));
  return (real)((dy > 0.0 ? 3.0 - p : 1.0 + p) / 4.0);
}

class SweepHull {
 public:
  std::vector<int> triangles;  // 3 point indices per triangle

  explicit SweepHull(const std::vector<Pt>& pts) : p_(pts) {}

  void run() {
    const int n = (int)p_.size();
    if (n < 3) return;
    hash_size_ = (int)std::ceil(std::sqrt((double)n));  // src/delaunator.cpp:186
    hash_.assign(hash_size_, kNil);
    prev_.assign(n, 0); next_.assign(n, 0); hull_tri_.assign(n, 0);
    triangles.reserve(6 * (size_t)n);
    half_.reserve(6 * (size_t)n);

    real max_x = -kInf, max_y = -kInf, min_x = kInf, min_y = kInf;
    for (int i = 0; i < n; i++) {
      min_x = std::min(p_[i].x, min_x); min_y = std::min(p_[i].y, min_y);
      max_x = std::max(p_[i].x, max_x); max_y = std::max(p_[i].y, max_y);
    }
    const real w = max_x - min_x, h = max_y - min_y;
    span_ = w * w + h * h;
    Pt mid; mid.x = (min_x + max_x) / 2; mid.y = (min_y + max_y) / 2;

    // seed: closest to the bounding-box centre, first minimum wins (src/delaunator.cpp:218-230)
    std::vector<real> dist(n);
    int i0 = kNil, i1 = kNil, i2 = kNil;
    real best = kInf;
    for (int i = 0; i < n; i++) {
      dist[i] = dist2(mid, p_[i]);
      if (dist[i] < best) { i0 = i; best = dist[i]; }
    }
    // visiting order: ascending distance from the bounding-box centre; the reference's insertion
    // sort (src/delaunator.cpp:409-424) is stable, so a stable sort gives the same permutation
    std::vector<int> ids(n);
    std::iota(ids.begin(), ids.end(), 0);
    if (!radix_order(dist, ids))
      std::stable_sort(ids.begin(), ids.end(), [&](int a, int b) { return dist[a] < dist[b]; });

    if (i0 == kNil) return;
    best = kInf;
    for (int i = 0; i < n; i++) {  // closest distinct point (src/delaunator.cpp:238-247)
      if (i == i0) continue;
      const real d = dist2(p_[i0], p_[i]);
      if (d < best && d > 0.0) { i1 = i; best = d; }
    }
    if (i1 == kNil) return;  // all points coincide: the reference reads out of bounds here
    best = kInf;
    for (int i = 0; i < n; i++) {  // smallest circumcircle with the first two (:251-259)
      const real r = circumradius2(p_[i0], p_[i1], p_[i]);
      if (r < best && !(i == i0 || i == i1)) { i2 = i; best = r; }
    }
    if (i2 == kNil) return;  // all points collinear: ditto
    if (ccw(p_[i0], p_[i1], p_[i2])) std::swap(i1, i2);
    center_ = circumcenter(p_[i0], p_[i1], p_[i2]);

    hull_start_ = i0;
    next_[i0] = prev_[i2] = i1;
    next_[i1] = prev_[i0] = i2;
    next_[i2] = prev_[i1] = i0;
    hull_tri_[i0] = 0; hull_tri_[i1] = 1; hull_tri_[i2] = 2;
    hash_[key(p_[i0])] = i0; hash_[key(p_[i1])] = i1; hash_[key(p_[i2])] = i2;
    add_triangle(i0, i1, i2, kNil, kNil, kNil);

    for (int k = 0; k < n; k++) {  // src/delaunator.cpp:298-405
      const int i = ids[k];
      const Pt pi = p_[i];
      // a hull vertex near the point's angle, through the edge hash
      int start = 0;
      const int hk = key(pi);
      for (int j = 0; j < hash_size_; j++) {
        start = hash_[wrap(hk + j)];
        if (start != kNil && start != next_[start]) break;
      }
      if (start == kNil) continue;  // the reference would index out of bounds
      start = prev_[start];
      int e = start, q;
      for (;;) {  // advance to an edge the point sees from outside
        q = next_[e];
        if (same_point(pi, p_[e], span_) || same_point(pi, p_[q], span_)) { e = kNil; break; }
        if (ccw(pi, p_[e], p_[q])) break;
        e = q;
        if (e == start) { e = kNil; break; }
      }
      if (e == kNil) continue;  // on the hull already (the seed points) or inside it

      int t = add_triangle(e, i, next_[e], kNil, kNil, hull_tri_[e]);
      hull_tri_[i] = legalize(t + 2);
      hull_tri_[e] = t;

      int nxt = next_[e];
      for (;;) {  // walk forward along the hull
        q = next_[nxt];
        if (!ccw(pi, p_[nxt], p_[q])) break;
        t = add_triangle(nxt, i, q, hull_tri_[i], kNil, hull_tri_[nxt]);
        hull_tri_[i] = legalize(t + 2);
        next_[nxt] = nxt;  // removed from the hull
        nxt = q;
      }
      if (e == start) {
        for (;;) {  // and backward
          q = prev_[e];
          if (!ccw(pi, p_[q], p_[e])) break;
          t = add_triangle(q, i, e, kNil, hull_tri_[e], hull_tri_[q]);
          legalize(t + 2);
          hull_tri_[q] = t;
          next_[e] = e;
          e = q;
        }
      }
      prev_[i] = e; hull_start_ = e;
      prev_[nxt] = i; next_[e] = i; next_[i] = nxt;
      hash_[hk] = i;  // key(pi), computed above (center_ and hash_size_ do not change inside the loop)
      hash_[key(p_[e])] = e;
    }
  }

 private:
  const std::vector<Pt>& p_;
  std::vector<int> half_, prev_, next_, hull_tri_, hash_, stack_;
  Pt center_{0, 0};
  real span_ = 0;
  int hash_size_ = 0, hull_start_ = 0;

  // Stable ascending order of non-negative finite-or-infinite float keys: their bit patterns order like unsigned integers,
  // so three stable counting passes over 11 + 11 + 10 bits give exactly std::stable_sort's permutation (r2u: the comparison
  // sort was 11 % of the stage).  Refuses (caller falls back) if a key is negative or NaN.
  static bool radix_order(const std::vector<real>& key, std::vector<int>& ids) {
    const size_t n = key.size();
    std::vector<uint32_t> k(n);
    for (size_t i = 0; i < n; i++) {
      uint32_t u;
      static_assert(sizeof(real) == 4, "float keys");
      std::memcpy(&u, &key[i], 4);
      if (u > 0x7F800000u) return false;  // sign bit set (incl. -0.0: equal to +0.0 for `<`, different as bits) or NaN
      k[i] = u;
    }
    std::vector<int> tmp(n);
    int* src = ids.data();
    int* dst = tmp.data();
    const int shifts[3] = {0, 11, 22}, bits[3] = {11, 11, 10};
    for (int pass = 0; pass < 3; pass++) {
      const uint32_t mask = (1u << bits[pass]) - 1u;
      std::vector<uint32_t> cnt((size_t)mask + 2, 0);
      for (size_t i = 0; i < n; i++) cnt[((k[src[i]] >> shifts[pass]) & mask) + 1]++;
      for (size_t b = 1; b < cnt.size(); b++) cnt[b] += cnt[b - 1];
      for (size_t i = 0; i < n; i++) dst[cnt[(k[src[i]] >> shifts[pass]) & mask]++] = src[i];
      std::swap(src, dst);
    }
    if (src != ids.data()) std::memcpy(ids.data(), src, n * sizeof(int));  // (three passes: the result sits in tmp)
    return true;
  }

  int wrap(int i) const { return i >= hash_size_ ? i % hash_size_ : i; }  // fast_mod, src/delaunator.cpp:19-21

  // src/delaunator.cpp:568-574: float product, float floor, truncation to int
  int key(const Pt& p) const {
    const real a = pseudo_angle(p.x - center_.x, p.y - center_.y);
    const real f = a * hash_size_;  // a is in [0, 1]: floor() and the truncation of the cast agree (src: std::floor, then (int))
    if (!(f >= 0.0f)) return 0;  // NaN (point on the circumcentre): undefined in the reference
    return wrap((int)f);
  }

  void link(int a, int b) {  // src/delaunator.cpp:594-610
    const int s = (int)half_.size();
    if (a == s) half_.push_back(b);
    else if (a < s) half_[a] = b;
    if (b != kNil) {
      const int s2 = (int)half_.size();
      if (b == s2) half_.push_back(a);
      else if (b < s2) half_[b] = a;
    }
  }

  // add_triangle's three links: a new triangle's half-edges t, t+1, t+2 are appended in that order and their partners
  // (a, b, c) are older edges or kNil, so link()'s case analysis reduces to "append, then point the partner back"
  void link_new(int t, int partner) {
    half_.push_back(partner);
    if (partner != kNil) half_[partner] = t;
  }

  void relink(int a, int b) {
    half_[a] = b;
    if (b != kNil) half_[b] = a;
  }

  int add_triangle(int i0, int i1, int i2, int a, int b, int c) {  // src/delaunator.cpp:576-592
    const int t = (int)triangles.size();
    triangles.push_back(i0); triangles.push_back(i1); triangles.push_back(i2);
    if ((int)half_.size() == t && (a == kNil || a < t) && (b == kNil || b < t) && (c == kNil || c < t)) {
      link_new(t, a); link_new(t + 1, b); link_new(t + 2, c);
    } else {  // (never taken by run(); kept so that the function stays the reference's for any caller)
      link(t, a); link(t + 1, b); link(t + 2, c);
    }
    return t;
  }

  // src/delaunator.cpp:465-566: flip until locally Delaunay; explicit stack, same visiting order
  int legalize(int a) {
    size_t depth = 0;
    int ar = 0;
    stack_.clear();
    for (;;) {
      static const int kNext[3] = {1, 2, 0}, kPrev[3] = {2, 0, 1};  // (x + 1) % 3, (x + 2) % 3
      const int b = half_[a];
      const int a0 = 3 * (a / 3), ra = a - a0;
      ar = a0 + kPrev[ra];
      if (b == kNil) {
        if (depth == 0) break;
        a = stack_[--depth];
        continue;
      }
      const int b0 = 3 * (b / 3), rb = b - b0;
      const int al = a0 + kNext[ra];
      const int bl = b0 + kPrev[rb];
      const int p0 = triangles[ar], pr = triangles[a], pl = triangles[al], p1 = triangles[bl];
      if (in_circle(p_[p0], p_[pr], p_[pl], p_[p1])) {
        triangles[a] = p1;
        triangles[b] = p0;
        const int hbl = half_[bl];
        if (hbl == kNil) {  // the flipped edge was a hull edge: repoint its hull triangle
          int e = hull_start_;
          do {
            if (hull_tri_[e] == bl) { hull_tri_[e] = a; break; }
            e = prev_[e];
          } while (e != hull_start_);
        }
        relink(a, hbl);  // every half-edge touched here exists already: plain stores (link() also grows the array)
        relink(b, half_[ar]);
        relink(ar, bl);
        const int br = b0 + kNext[rb];
        if (depth < stack_.size()) stack_[depth] = br; else stack_.push_back(br);
        depth++;
      } else {
        if (depth == 0) break;
        a = stack_[--depth];
      }
    }
    return ar;
  }
};

// src/remove_outliers.cpp:4-94
static void remove_outliers(vo_p_match* m, int32_t* count) {
  const int n = *count;
  if (n <= 3) return;
  std::vector<Pt> pts(n);
  for (int i = 0; i < n; i++) { pts[i].x = m[i].u1c; pts[i].y = m[i].v1c; }
  SweepHull tri(pts);
  tri.run();

  const int tol = 5;  // outlier_flow_tolerance
  std::vector<int32_t> support(n, 0);
  std::vector<real> fu(n), fv(n);
  for (int i = 0; i < n; i++) { fu[i] = m[i].u1c - m[i].u1p; fv[i] = m[i].v1c - m[i].v1p; }
  const size_t nt = tri.triangles.size() / 3;
  for (size_t t = 0; t < nt; t++) {
    const int p1 = tri.triangles[3 * t], p2 = tri.triangles[3 * t + 1], p3 = tri.triangles[3 * t + 2];
    const bool f1 = std::fabs(fu[p1] - fu[p2]) + std::fabs(fv[p1] - fv[p2]) < tol;
    const bool f2 = std::fabs(fu[p2] - fu[p3]) + std::fabs(fv[p2] - fv[p3]) < tol;
    const bool f3 = std::fabs(fu[p1] - fu[p3]) + std::fabs(fv[p1] - fv[p3]) < tol;
    support[p1] += (f1 && f3) ? 2 : ((f1 || f3) ? 1 : 0);
    support[p2] += (f1 && f2) ? 2 : ((f1 || f2) ? 1 : 0);
    support[p3] += (f2 && f3) ? 2 : ((f2 || f3) ? 1 : 0);
  }
  int k = 0;
  for (int i = 0; i < n; i++)
    if (support[i] >= 4) m[k++] = m[i];
  *count = k;
}

}  // namespace vo_host

extern "C" vo_status vo_remove_outliers_host(vo_p_match* matches, int32_t* n) {
  if (!n || *n < 0 || (!matches && *n > 0)) return VO_ERR_INVALID;
  vo_host::remove_outliers(matches, n);
  return VO_OK;
}
