// cx_narrow.cuh — device narrowphase: supports, GJK, EPA, contact generation.
//
// One thread evaluates one shape pair.  Replaces (paths in the reference):
//   cotix/_convex_shapes.py  get_support :22-26,62-66,149-155; contains :28-29,105-106,168-175
//   cotix/_collisions.py     GJK :20-112 + :277-310, EPA :115-273 + :313-329
//   cotix/_contacts.py       the six contact functions :30-315
//
// Design notes (B200): polygon vertices stay in registers (every access is unrolled, the
// first-maximum support scan is 6 predicated FFMA-free dot products); the EPA edge list and
// its cached squared distances live in a per-thread local array (L1-resident) — the
// reference recomputes every edge distance each iteration (O(iters^2)), here only the two
// edges an expansion creates are evaluated and the arg-min is a compare scan over the
// cache, which yields identical results because an edge's distance depends on its
// endpoints only.  A miss never runs EPA or contact generation (the reference computes and
// discards them).
#pragma once
#include "cx_math.cuh"

#define CX_MAXV 6

struct Poly {
  float x[CX_MAXV], y[CX_MAXV];
  int n;
};
struct Box {
  V2 lo, hi;
};
struct Circ {
  V2 c;
  float r;
};

// ---- supports ---------------------------------------------------------------------
CX_DEV V2 support(const Circ& s, V2 d) {
  V2 nd = d / norm2(d);
  return nd * s.r + s.c;
}
CX_DEV V2 support(const Box& s, V2 d) {
  return mk2(d.x >= 0.0f ? s.hi.x : s.lo.x, d.y >= 0.0f ? s.hi.y : s.lo.y);
}
CX_DEV V2 support(const Poly& s, V2 d) {
  // first maximum of the dot products, the first NaN winning (jnp.argmax).  The sum of the products is
  // NaN exactly when one of them is (or +inf meets -inf): only then the NaN-aware scan is needed
  float dd[CX_MAXV];
  dd[0] = s.x[0] * d.x + s.y[0] * d.y;
  float acc = dd[0];
#pragma unroll
  for (int k = 1; k < CX_MAXV; ++k) {
    dd[k] = s.x[k] * d.x + s.y[k] * d.y;
    if (k < s.n) acc = acc + dd[k];
  }
  float best = dd[0], bx = s.x[0], by = s.y[0];
  if (acc == acc) {
#pragma unroll
    for (int k = 1; k < CX_MAXV; ++k) {
      if (k < s.n && dd[k] > best) {
        best = dd[k];
        bx = s.x[k];
        by = s.y[k];
      }
    }
    return mk2(bx, by);
  }
  if (isnan2(d)) return nan2();
  bool bnan = best != best;
#pragma unroll
  for (int k = 1; k < CX_MAXV; ++k) {
    if (k < s.n) {
      const bool dnan = dd[k] != dd[k];
      const bool take = !bnan && (dnan || dd[k] > best);
      if (take) {
        best = dd[k];
        bx = s.x[k];
        by = s.y[k];
        bnan = dnan;
      }
    }
  }
  return mk2(bx, by);
}
// A polygon whose vertex count is a compile-time constant (the LunarLander bodies: 4 or 6): no per-vertex
// `k < n` predicate, and — the support function is half of all instructions the narrowphase executes — a
// support that takes the plain first-maximum scan whenever no dot product is NaN.  The sum of the products
// is NaN exactly when one of them is (or when +inf and -inf meet); then, and only then, the NaN-aware scan
// of support(Poly) decides, so the result is the same vertex in every case.
template <int N>
struct PolyN {
  float x[N], y[N];
};
template <int N>
CX_DEV V2 support(const PolyN<N>& s, V2 d) {
  float dd[N];
#pragma unroll
  for (int k = 0; k < N; ++k) dd[k] = s.x[k] * d.x + s.y[k] * d.y;
  float acc = dd[0];
#pragma unroll
  for (int k = 1; k < N; ++k) acc = acc + dd[k];
  float best = dd[0], bx = s.x[0], by = s.y[0];
  if (acc == acc) {
#pragma unroll
    for (int k = 1; k < N; ++k) {
      if (dd[k] > best) {
        best = dd[k];
        bx = s.x[k];
        by = s.y[k];
      }
    }
    return mk2(bx, by);
  }
  if (isnan2(d)) return nan2();
  bool bnan = best != best;
#pragma unroll
  for (int k = 1; k < N; ++k) {
    const bool dnan = dd[k] != dd[k];
    const bool take = !bnan && (dnan || dd[k] > best);
    if (take) {
      best = dd[k];
      bx = s.x[k];
      by = s.y[k];
      bnan = dnan;
    }
  }
  return mk2(bx, by);
}
template <int N>
CX_DEV PolyN<N> poly_fixed(const Poly& p) {
  PolyN<N> q;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    q.x[k] = p.x[k];
    q.y[k] = p.y[k];
  }
  return q;
}
template <class SA, class SB>
CX_DEV V2 mdiff(const SA& a, const SB& b, V2 d) {  // minkowski_diff, _geometry_utils.py:49-57
  return support(a, d) - support(b, -d);
}

// ---- edge lists (local memory; cold path) ------------------------------------------
struct EdgeList {
  float ax[CX_MAXV], ay[CX_MAXV], bx[CX_MAXV], by[CX_MAXV];  // edge k = (a_k, b_k); vertex k = a_k
  int n;
};
CX_DEV void edges_of(const Poly& s, EdgeList& e) {  // (v_k, v_{k-1}), _convex_shapes.py:160-163
  e.n = s.n;
  float lx = s.x[0], ly = s.y[0];
#pragma unroll
  for (int k = 1; k < CX_MAXV; ++k)
    if (k < s.n) {
      lx = s.x[k];
      ly = s.y[k];
    }
#pragma unroll
  for (int k = 0; k < CX_MAXV; ++k) {
    e.ax[k] = s.x[k];
    e.ay[k] = s.y[k];
    e.bx[k] = (k == 0) ? lx : s.x[k > 0 ? k - 1 : 0];
    e.by[k] = (k == 0) ? ly : s.y[k > 0 ? k - 1 : 0];
  }
}
CX_DEV void edges_of(const Box& s, EdgeList& e) {  // AABB.get_edges, :82-93
  e.n = 4;
  const float vx[4] = {s.hi.x, s.hi.x, s.lo.x, s.lo.x};
  const float vy[4] = {s.hi.y, s.lo.y, s.lo.y, s.hi.y};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    e.ax[k] = vx[k];
    e.ay[k] = vy[k];
    e.bx[k] = vx[(k + 1) & 3];
    e.by[k] = vy[(k + 1) & 3];
  }
  e.ax[4] = e.ay[4] = e.bx[4] = e.by[4] = 0.0f;
  e.ax[5] = e.ay[5] = e.bx[5] = e.by[5] = 0.0f;
}

template <int N>
CX_DEV void edges_of(const PolyN<N>& s, EdgeList& e) {
  e.n = N;
#pragma unroll
  for (int k = 0; k < CX_MAXV; ++k) {
    if (k < N) {
      e.ax[k] = s.x[k];
      e.ay[k] = s.y[k];
      e.bx[k] = s.x[k == 0 ? N - 1 : k - 1];
      e.by[k] = s.y[k == 0 ? N - 1 : k - 1];
    } else {
      e.ax[k] = e.ay[k] = e.bx[k] = e.by[k] = 0.0f;
    }
  }
}

// ---- contains ------------------------------------------------------------------------
CX_DEV bool contains(const Circ& s, V2 p) {
  float re = s.r + 1e-6f;
  return sumsq2(p - s.c) <= re * re;
}
CX_DEV bool contains(const Box& s, V2 p) {
  const float eps = 1e-6f;
  return (p.x >= s.lo.x - eps) && (p.x <= s.hi.x + eps) && (p.y >= s.lo.y - eps) &&
         (p.y <= s.hi.y + eps);
}
// polygon: all edge-side signs equal to the first (a point ON an edge is outside), :168-175
CX_DEV bool poly_contains(const EdgeList& e, V2 p) {
  float s0 = 0.0f;
  bool ok = true;
  for (int k = 0; k < e.n; ++k) {
    V2 e0 = mk2(e.ax[k], e.ay[k]);
    V2 e1 = mk2(e.bx[k], e.by[k]);
    float sg = jsign(dot2(p - e0, perp(e0 - e1)));
    if (k == 0) s0 = sg;
    if (!(sg == s0)) ok = false;
  }
  return ok;
}
CX_DEV bool contains_el(const Poly&, const EdgeList& e, V2 p) { return poly_contains(e, p); }
template <int N>
CX_DEV bool contains_el(const PolyN<N>&, const EdgeList& e, V2 p) {
  return poly_contains(e, p);
}
CX_DEV bool contains_el(const Box& s, const EdgeList&, V2 p) { return contains(s, p); }

// ---- GJK — cotix/_collisions.py:20-112, 277-310 --------------------------------------
CX_DEV bool point_in_triangle0(V2 v1, V2 v2, V2 v3) {  // is_point_in_triangle(0, ...)
  const V2 pt = mk2(0.0f, 0.0f);
  float d1 = (pt.x - v2.x) * (v1.y - v2.y) - (v1.x - v2.x) * (pt.y - v2.y);
  float d2 = (pt.x - v3.x) * (v2.y - v3.y) - (v2.x - v3.x) * (pt.y - v3.y);
  float d3 = (pt.x - v1.x) * (v3.y - v1.y) - (v3.x - v1.x) * (pt.y - v1.y);
  bool has_neg = (d1 < 0) || (d2 < 0) || (d3 < 0);
  bool has_pos = (d1 > 0) || (d2 > 0) || (d3 > 0);
  return !(has_neg && has_pos);
}

// returns exists; simplex s0,s1,s2 valid iff exists.  *capped = loop hit max_steps.
template <class SA, class SB>
CX_DEV bool gjk(const SA& A, const SB& B, V2 start, int max_steps, V2& s0, V2& s1, V2& s2,
                bool& capped) {
  s0 = mdiff(A, B, start);
  s1 = mdiff(A, B, -s0);
  V2 dir = perp(s1 - s0);
  if (dot2(dir, -s1) > 0) {
    V2 t = s0;
    s0 = s1;
    s1 = t;
  } else {
    dir = -dir;
  }
  s2 = mdiff(A, B, dir);
  int steps = 0;
  for (;;) {
    bool c1 = dot2(s2, dir) <= 0;
    bool c2 = dot2(perp(s2 - s0), -s2) < 0;
    bool c3 = dot2(perp(s1 - s2), -s2) < 0;
    bool go = !(c1 || (c2 && c3));
    if (!go || steps >= max_steps) break;
    V2 acn = perp(s2 - s0);
    V2 cbn = perp(s1 - s2);
    if (dot2(acn, -s2) >= 0) {
      s1 = s2;
      dir = acn;
    } else {
      s0 = s2;
      dir = cbn;
    }
    s2 = mdiff(A, B, dir);
    ++steps;
  }
  capped = steps >= max_steps;
  if (!point_in_triangle0(s0, s1, s2)) return false;  // simplex := zeros -> all-zero -> no
  float area = cross2(s1 - s0, s2 - s0);
  bool all_zero = (s0.x == 0.0f) && (s0.y == 0.0f) && (s1.x == 0.0f) && (s1.y == 0.0f) &&
                  (s2.x == 0.0f) && (s2.y == 0.0f);
  bool any_nan = isnan2(s0) || isnan2(s1) || isnan2(s2);
  return !(all_zero || any_nan || (area == 0));
}

// ---- EPA — cotix/_collisions.py:115-273 ----------------------------------------------
CX_DEV V2 closest_to_origin(V2 a, V2 b) {  // get_closest_point_on_edge_to_point(a, b, 0)
  const V2 point = mk2(0.0f, 0.0f);
  V2 ab = a - b;
  float length = sumsq2(ab);
  if (length == 0.0f) return point - a;
  float t = dot2(point - b, ab) / length;
  t = jclip(t, 0.0f, 1.0f);
  V2 projection = b + t * ab;
  return point - projection;
}
CX_DEV float edge_dist2(V2 a, V2 b) {  // distance_to_origin :168-169 via :137-154
  bool both_zero = (a.x == 0.0f) && (a.y == 0.0f) && (b.x == 0.0f) && (b.y == 0.0f);
  if (both_zero) return CX_INF;
  const V2 point = mk2(0.0f, 0.0f);
  V2 ab = a - b;
  float length = sumsq2(ab);
  V2 disp;
  if (length == 0) {
    disp = -a;
  } else {
    float t = dot2(point - b, ab) / length;
    t = jclip(t, 0.0f, 1.0f);
    V2 projection = b + t * ab;
    disp = point - projection;
  }
  return sumsq2(disp);
}
// first minimum, NaN wins (jnp.argmin)
CX_DEV int argmin_first(const float* d, int n) {
  int best = 0;
  float bv = d[0];
  if (bv != bv) return 0;
  for (int k = 1; k < n; ++k) {
    float v = d[k];
    if (v != v) return k;
    if (v < bv) {
      bv = v;
      best = k;
    }
  }
  return best;
}
// the same over an EPA edge list.  No early exit: the loads of d[] (local memory) do not depend on
// the running minimum, so they overlap instead of forming a chain of load -> compare -> branch.
CX_DEV int argmin_scan(const float* d, int n) {
  int best = 0;
  float bv = d[0];
  bool bnan = bv != bv;
#pragma unroll 4
  for (int k = 1; k < n; ++k) {
    const float v = d[k];
    const bool vnan = v != v;
    const bool take = !bnan && (vnan || v < bv);
    if (take) {
      bv = v;
      best = k;
      bnan = vnan;
    }
  }
  return best;
}

// epa_edge leaves the closest edge (ba, bb) of the expanded polytope in oa / ob; the penetration
// vector is closest_to_origin(oa, ob).  The adjoint kernel (lunar_adjoint.cu) needs the edge.
template <int MAXE, class SA, class SB>
CX_DEV void epa_edge(const SA& A, const SB& B, V2 s0, V2 s1, V2 s2, int iters, bool& capped, V2& oa,
                     V2& ob) {
  float eax[MAXE], eay[MAXE], ebx[MAXE], eby[MAXE], dist[MAXE];
  eax[0] = s0.x; eay[0] = s0.y; ebx[0] = s1.x; eby[0] = s1.y;
  eax[1] = s1.x; eay[1] = s1.y; ebx[1] = s2.x; eby[1] = s2.y;
  eax[2] = s2.x; eay[2] = s2.y; ebx[2] = s0.x; eby[2] = s0.y;
  dist[0] = edge_dist2(s0, s1);
  dist[1] = edge_dist2(s1, s2);
  dist[2] = edge_dist2(s2, s0);
  int cnt = 3;  // entries >= cnt are the zero padding: distance inf, never the arg-min
                // unless every live distance is inf/NaN-free larger — impossible here
  // short lists (polygon pairs: <= 16 entries) are scanned with the early-exit form, long ones
  // (circle-polygon) with the overlapped one: measured, each is slower on the other's kernels
  int bi = MAXE > 16 ? argmin_scan(dist, cnt) : argmin_first(dist, cnt);
  V2 ba = mk2(eax[bi], eay[bi]), bb = mk2(ebx[bi], eby[bi]);
  V2 newp = s2;
  V2 pa = s0, pb = s1;  // prev_edge = edges[0]
  // Two values the reference recomputes every iteration are already known here, bit for bit:
  //  * the unit normal of prev_edge is the expansion direction n2 of the previous iteration
  //    (same expression on the same end points);
  //  * |closest_to_origin(prev_edge)| is the square root of that edge's cached squared distance
  //    (edge_dist2 and closest_to_origin evaluate the same displacement; the only difference,
  //    the all-zero edge that edge_dist2 maps to inf, falls back to the direct evaluation).
  V2 nprev = mk2(0.0f, 0.0f);
  float dsel = dist[0];    // cached squared distance of prev_edge when it was the best edge
  float dcur = dist[bi];   // cached squared distance of the current best edge
  int ran = 0;
  for (int it = 0; it < iters; ++it) {
    bool c1 = sumsq2(ba - bb) > 1e-9f;
    bool c2 = cross2(ba, bb) >= 0;
    V2 nrm;
    if (it == 0) {
      nrm = perp(pa - pb);
      nrm = nrm / norm2(nrm);
    } else {
      nrm = nprev;
    }
    float d = dot2(newp, nrm);
    float ed = (dsel <= 3.0e38f) ? sqrtf(dsel) : norm2(closest_to_origin(pa, pb));
    bool c4 = (d - ed > 1e-6f) || (d <= 0);
    bool bad = isnan2(ba) || isnan2(bb);
    if (!(c4 && !bad && c1 && c2)) break;  // the scan carries the state unchanged from here
    V2 n2 = perp(ba - bb);
    n2 = n2 / norm2(n2);
    V2 np = mdiff(A, B, n2);
    // edges[bi] = (best.a, np); edges[cnt] = (np, best.b)
    ebx[bi] = np.x; eby[bi] = np.y;
    dist[bi] = edge_dist2(ba, np);
    eax[cnt] = np.x; eay[cnt] = np.y; ebx[cnt] = bb.x; eby[cnt] = bb.y;
    dist[cnt] = edge_dist2(np, bb);
    ++cnt;
    pa = ba;
    pb = bb;
    nprev = n2;
    dsel = dcur;
    bi = MAXE > 16 ? argmin_scan(dist, cnt) : argmin_first(dist, cnt);
    dcur = dist[bi];
    ba = mk2(eax[bi], eay[bi]);
    bb = mk2(ebx[bi], eby[bi]);
    newp = np;
    ++ran;
  }
  capped = ran >= iters;
  oa = ba;
  ob = bb;
}
// The same expansion with the edge list as a RING in caller-provided (shared) memory: the polytope
// is a closed polygon, so the end point b of edge i is the start point a of another edge, next[i].
// An entry is (a.x, a.y, squared distance, next) = 13 bytes instead of 20, and lives in shared memory
// laid out [entry][thread] (bank = thread: conflict-free whatever entry each lane looks at) instead
// of in local memory, whose footprint over all resident threads exceeds the L2 (the sweep's
// circle-polygon pass moved 910 B of DRAM traffic per hit for 170 algorithmic).  Invariant
// b_i == a[next_i]: true for the initial triangle; an expansion of edge bi by np keeps a[bi], makes
// its end point np (= a[cnt], next[bi] = cnt) and gives the new edge cnt the old end point
// (next[cnt] = old next[bi]).  The values, the list order (arg-min ties!) and every operation are
// those of epa_edge: identical bits.  STRIDE = threads per block.
template <int STRIDE>
struct EpaRing {
  float* ax;      // [MAXE][STRIDE]
  float* ay;      // [MAXE][STRIDE]
  float* d;       // [MAXE][STRIDE]
  uint8_t* next;  // [MAXE][STRIDE]
  CX_DEV V2 a(int k) const { return mk2(ax[k * STRIDE], ay[k * STRIDE]); }
  CX_DEV void set(int k, V2 p, int nx, float dist) const {
    ax[k * STRIDE] = p.x;
    ay[k * STRIDE] = p.y;
    next[k * STRIDE] = (uint8_t)nx;
    d[k * STRIDE] = dist;
  }
  // first minimum, NaN wins (argmin_scan on the strided list)
  CX_DEV int argmin(int n) const {
    int best = 0;
    float bv = d[0];
    bool bnan = bv != bv;
#pragma unroll 4
    for (int k = 1; k < n; ++k) {
      const float v = d[k * STRIDE];
      const bool vnan = v != v;
      const bool take = !bnan && (vnan || v < bv);
      if (take) {
        bv = v;
        best = k;
        bnan = vnan;
      }
    }
    return best;
  }
};
template <int STRIDE, class SA, class SB>
CX_DEV void epa_edge_ring(const SA& A, const SB& B, V2 s0, V2 s1, V2 s2, int iters, bool& capped, V2& oa, V2& ob,
                          const EpaRing<STRIDE>& R) {
  R.set(0, s0, 1, edge_dist2(s0, s1));
  R.set(1, s1, 2, edge_dist2(s1, s2));
  R.set(2, s2, 0, edge_dist2(s2, s0));
  int cnt = 3;
  int bi = R.argmin(cnt);
  V2 ba = R.a(bi), bb = R.a(R.next[bi * STRIDE]);
  V2 newp = s2;
  V2 pa = s0, pb = s1;  // prev_edge = edges[0]
  V2 nprev = mk2(0.0f, 0.0f);
  float dsel = R.d[0];
  float dcur = R.d[bi * STRIDE];
  int ran = 0;
  for (int it = 0; it < iters; ++it) {
    bool c1 = sumsq2(ba - bb) > 1e-9f;
    bool c2 = cross2(ba, bb) >= 0;
    V2 nrm;
    if (it == 0) {
      nrm = perp(pa - pb);
      nrm = nrm / norm2(nrm);
    } else {
      nrm = nprev;
    }
    float d = dot2(newp, nrm);
    float ed = (dsel <= 3.0e38f) ? sqrtf(dsel) : norm2(closest_to_origin(pa, pb));
    bool c4 = (d - ed > 1e-6f) || (d <= 0);
    bool bad = isnan2(ba) || isnan2(bb);
    if (!(c4 && !bad && c1 && c2)) break;
    V2 n2 = perp(ba - bb);
    n2 = n2 / norm2(n2);
    V2 np = mdiff(A, B, n2);
    // edges[bi] = (best.a, np); edges[cnt] = (np, best.b)
    const int nb = R.next[bi * STRIDE];
    R.d[bi * STRIDE] = edge_dist2(ba, np);
    R.next[bi * STRIDE] = (uint8_t)cnt;
    R.set(cnt, np, nb, edge_dist2(np, bb));
    ++cnt;
    pa = ba;
    pb = bb;
    nprev = n2;
    dsel = dcur;
    bi = R.argmin(cnt);
    dcur = R.d[bi * STRIDE];
    ba = R.a(bi);
    bb = R.a(R.next[bi * STRIDE]);
    newp = np;
    ++ran;
  }
  capped = ran >= iters;
  oa = ba;
  ob = bb;
}

template <int MAXE, class SA, class SB>
CX_DEV V2 epa(const SA& A, const SB& B, V2 s0, V2 s1, V2 s2, int iters, bool& capped) {
  V2 ba, bb;
  epa_edge<MAXE>(A, B, s0, s1, s2, iters, capped, ba, bb);
  return closest_to_origin(ba, bb);
}

// ---- contacts — cotix/_contacts.py -----------------------------------------------------
struct Contact {
  V2 pen, cp;       // ContactInfo: NaN contact_point == "no contact" (:19-24)
  uint32_t status;  // COTIX_ST_* bits
};
#define CX_ST_GJK_CAP 0x1u
#define CX_ST_EPA_CAP 0x2u
#define CX_ST_CCP_ERR 0x10u
CX_DEV Contact contact_none() {
  Contact c;
  c.pen = mk2(0.0f, 0.0f);
  c.cp = nan2();
  c.status = 0u;
  return c;
}
CX_DEV bool contact_isnan(const Contact& c) { return isnan2(c.cp); }

CX_DEV Contact circle_vs_circle(const Circ& a, const Circ& b) {  // :30-58
  V2 delta = a.c - b.c;
  float distance = norm2(delta);
  Contact c = contact_none();
  if (!(distance <= a.r + b.r)) return c;  // nothing below is kept when the circles are apart
  V2 dir = (distance == 0.0f) ? mk2(1.0f, 0.0f) : delta / distance;
  V2 pv = dir * jmin(distance - (a.r + b.r), 0.0f);
  V2 cp = (b.c + dir * (b.r - a.r) + a.c) / 2.0f;
  if (!(dot2(a.c - cp, b.c - cp) <= 0)) cp = contains(a, b.c) ? b.c : a.c;
  c.pen = -pv;
  c.cp = cp;
  return c;
}

CX_DEV Contact aabb_vs_aabb(const Box& a, const Box& b) {  // :61-96
  const float eps = 1e-8f;
  bool below = a.hi.y <= b.lo.y, above = a.lo.y >= b.hi.y;
  bool left = a.hi.x <= b.lo.x, right = a.lo.x >= b.hi.x;
  Contact c = contact_none();
  if (!(below || left || above || right)) {
    float dep[4] = {jmax(a.hi.y - b.lo.y, -eps), jmax(b.hi.y - a.lo.y, -eps),
                    jmax(a.hi.x - b.lo.x, -eps), jmax(b.hi.x - a.lo.x, -eps)};
    int idx = argmin_first(dep, 4);
    float md = jmax(dep[idx], 0.0f);
    const float dx[4] = {0.0f, 0.0f, -1.0f, 1.0f};
    const float dy[4] = {-1.0f, 1.0f, 0.0f, 0.0f};
    c.pen = mk2(md * dx[idx], md * dy[idx]);
    V2 mu = mk2(jmin(a.hi.x, b.hi.x), jmin(a.hi.y, b.hi.y));
    V2 ml = mk2(jmax(a.lo.x, b.lo.x), jmax(a.lo.y, b.lo.y));
    c.cp = (mu + ml) / 2.0f;
  }
  return c;
}

CX_DEV Contact circle_vs_aabb(const Circ& a, const Box& b) {  // :99-154
  const float eps = 1e-6f;
  V2 bc = (b.lo + b.hi) / 2.0f;
  V2 disp = a.c - bc;
  V2 lo = b.lo - bc, hi = b.hi - bc;
  V2 cd = mk2(jclip(disp.x, lo.x, hi.x), jclip(disp.y, lo.y, hi.y));
  V2 ccp = bc + cd;
  Contact c = contact_none();
  if (!contains(b, ccp)) c.status |= CX_ST_CCP_ERR;
  if (contains(a, ccp)) {
    bool pv = (norm2(b.lo - ccp) < eps) || (norm2(mk2(b.lo.x, b.hi.y) - ccp) < eps) ||
              (norm2(b.hi - ccp) < eps) || (norm2(mk2(b.hi.x, b.lo.y) - ccp) < eps);
    if (pv) {
      V2 dir = ccp - a.c;
      V2 dn = dir / norm2(dir);
      c.pen = -(a.c + a.r * dn - ccp);
      c.cp = ccp;
    } else {
      float sh[4] = {a.c.y + a.r - b.lo.y, b.hi.y - (a.c.y - a.r), a.c.x + a.r - b.lo.x,
                     b.hi.x - (a.c.x - a.r)};
      int best = argmin_first(sh, 4);
      const float dx[4] = {0.0f, 0.0f, 1.0f, -1.0f};
      const float dy[4] = {1.0f, -1.0f, 0.0f, 0.0f};
      float ms = -sh[best];
      c.pen = mk2(ms * dx[best], ms * dy[best]);
      c.cp = ccp;
    }
  }
  return c;
}

// contact point of circle_vs_polygon (:166-197): the point of the polygon's boundary nearest to
// the circle centre, or the centre itself when that point is farther than the radius
CX_DEV V2 circle_poly_contact_point(const Circ& circle, const Poly& poly) {
  EdgeList e;
  edges_of(poly, e);
  float bestd = 0.0f;
  V2 bestdisp = mk2(0.0f, 0.0f);
  bool bnan = false;
  for (int k = 0; k < e.n; ++k) {
    V2 a = mk2(e.ax[k], e.ay[k]), b = mk2(e.bx[k], e.by[k]);
    V2 disp;
    bool both_zero = (a.x == 0.0f) && (a.y == 0.0f) && (b.x == 0.0f) && (b.y == 0.0f);
    if (both_zero) {
      disp = mk2(CX_INF, CX_INF);
    } else {
      V2 ab = a - b;
      float length = sumsq2(ab);
      float t = dot2(circle.c - b, ab) / length;
      t = jclip(t, 0.0f, 1.0f);
      V2 projection = b + t * ab;
      disp = circle.c - projection;
    }
    float d = disp.x * disp.x + disp.y * disp.y;
    bool dn = d != d;
    if (k == 0 || (!bnan && (dn || d < bestd))) {
      bestd = d;
      bestdisp = disp;
      bnan = dn;
    }
  }
  V2 cp = circle.c + bestdisp;
  if (bestd > circle.r * circle.r) cp = circle.c;
  return cp;
}

// second half of circle_vs_polygon (:162-202), given the GJK simplex of a hit
template <int MAXE>
CX_DEV void circle_poly_finish(const Circ& circle, const Poly& poly, V2 s0, V2 s1, V2 s2, int epa_iters,
                               Contact& c) {
  bool ecap;
  if (epa_iters > MAXE - 3) epa_iters = MAXE - 3;
  V2 pen = epa<MAXE>(circle, poly, s0, s1, s2, epa_iters, ecap);
  if (ecap) c.status |= CX_ST_EPA_CAP;
  c.pen = pen;
  c.cp = circle_poly_contact_point(circle, poly);
}

template <int MAXE>
CX_DEV Contact circle_vs_polygon(const Circ& circle, const Poly& poly, V2 start, int gjk_max,
                                 int epa_iters) {  // :157-202
  V2 s0, s1, s2;
  bool gcap;
  Contact c = contact_none();
  bool ex = gjk(circle, poly, start, gjk_max, s0, s1, s2, gcap);
  if (gcap) c.status |= CX_ST_GJK_CAP;
  if (!ex) return c;
  circle_poly_finish<MAXE>(circle, poly, s0, s1, s2, epa_iters, c);
  return c;
}

// _contact_from_edges :205-267 — summation order: A-vertices inside B, B-vertices inside A,
// then intersections for eb in edges_b: for ea in edges_a
template <class SA, class SB>
CX_DEV V2 contact_from_edges(const SA& A, const SB& B) {
  EdgeList ea, eb;
  edges_of(A, ea);
  edges_of(B, eb);
  float n = 0.0f;
  V2 avg = mk2(0.0f, 0.0f);
  for (int k = 0; k < ea.n; ++k) {
    V2 v = mk2(ea.ax[k], ea.ay[k]);
    if (contains_el(B, eb, v)) {
      avg = avg + v;
      n = n + 1.0f;
    }
  }
  for (int k = 0; k < eb.n; ++k) {
    V2 v = mk2(eb.ax[k], eb.ay[k]);
    if (contains_el(A, ea, v)) {
      avg = avg + v;
      n = n + 1.0f;
    }
  }
  for (int ib = 0; ib < eb.n; ++ib) {
    V2 q = mk2(eb.ax[ib], eb.ay[ib]);
    V2 s = mk2(eb.bx[ib], eb.by[ib]) - q;
    for (int ia = 0; ia < ea.n; ++ia) {
      V2 p = mk2(ea.ax[ia], ea.ay[ia]);
      V2 r = mk2(ea.bx[ia], ea.by[ia]) - p;
      float c = r.x * s.y - s.x * r.y;
      V2 qp = q - p;
      float t = (qp.x * s.y - s.x * qp.y) / c;
      float u = (qp.x * r.y - r.x * qp.y) / c;
      if ((c != 0.0f) && (t >= 0.0f) && (t <= 1.0f) && (u >= 0.0f) && (u <= 1.0f)) {
        V2 x = p + r * t;
        if (!isnan2(x)) {
          avg = avg + x;
          n = n + 1.0f;
        }
      }
    }
  }
  if (n > 0.0f) return avg / n;
  return nan2();
}

CX_DEV int nverts(const Poly& p) { return p.n; }
CX_DEV int nverts(const Box&) { return 4; }
template <int N>
CX_DEV int nverts(const PolyN<N>&) { return N; }

// aabb_vs_polygon :270-291 / polygon_vs_polygon :294-315
template <class SA, class SB>
CX_DEV Contact convex_vs_polygon(const SA& A, const SB& B, V2 start, int gjk_max) {
  int iters = nverts(A) + nverts(B) + 1;
  if (iters > 13) iters = 13;  // min(48, .) never binds for <= 6 vertices
  V2 s0, s1, s2;
  bool gcap, ecap;
  Contact c = contact_none();
  bool ex = gjk(A, B, start, gjk_max, s0, s1, s2, gcap);
  if (gcap) c.status |= CX_ST_GJK_CAP;
  if (!ex) return c;
  c.pen = epa<16>(A, B, s0, s1, s2, iters, ecap);
  if (ecap) c.status |= CX_ST_EPA_CAP;
  c.cp = contact_from_edges(A, B);
  return c;
}
// second half of aabb_vs_polygon / polygon_vs_polygon, given the GJK simplex of a hit
template <class SA>
CX_DEV void convex_poly_finish(const SA& A, const Poly& B, V2 s0, V2 s1, V2 s2, Contact& c) {
  int iters = nverts(A) + B.n + 1;
  if (iters > 13) iters = 13;
  bool ecap;
  c.pen = epa<16>(A, B, s0, s1, s2, iters, ecap);
  if (ecap) c.status |= CX_ST_EPA_CAP;
  c.cp = contact_from_edges(A, B);
}
