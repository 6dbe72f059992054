// cx_lunar_step.cuh — the LunarLander env step as per-environment device functions, shared by
// the two launch strategies:
//   lunar_kernel.cu    one fused kernel (block-level narrowphase queue in shared memory); used for
//                      K > 1 sub-steps per launch and for small batches
//   lunar_pipeline.cu  wavefront pipeline (stream-compacted queues in HBM between small dense
//                      kernels); used for large batches at K = 1
// Both run exactly these functions on exactly the same inputs, so their results are identical
// bit for bit (and identical to the CPU oracle in exact-trig mode).
//
// The step (examples/test_viz.py:25-44) is cut at the two points where the work becomes
// irregular:
//   integrate_and_cull   ExplicitEulerPhysics.step + gravity + engines + exact broadphase
//   [narrowphase of the candidate pairs: make_item_polys + gjk / epa / contact_from_edges]
//   collide              RandomizedCollider.resolve given the contacts (coins, choice, impulses)
//   finish_step          key split, LunarLander.step joints, judge
//   lunar_reset_env      LunarLander.__init__ (auto-reset)
#pragma once
#include "cx_lunar.cuh"

struct EnvRegs {
  BodyS b[3];
  float h[8];
  float kn[4];
  Key2 key;
  uint32_t steps, status;
  float ret, shaping;
};

// pair ids = bits of the candidate mask:
//  0 (rl,L)  1 (ll,L)  2..8 (g_k,L)  9 (ll,rl)  10 (ll,ll)  11..17 (g_k,rl)  18..24 (g_k,ll)
constexpr uint32_t CX_GROUND_BITS = 0x01FFF9FCu;
constexpr uint32_t CX_PAIR_BITS = 0x01FFFFFFu;
constexpr int CX_NPAIR = 25;

struct StepCtx {
  float cs[3], sn[3];
  float thr_main, thr_side;
  uint32_t cand;
  int base;  // first queue slot of this env's items, or -1 (evaluate inline)
};

struct SubInfo {
  float reward;
  int done;
  bool rc, lc;
};

struct Aabb {
  float lx, ly, hx, hy;
};

template <int N>
CX_DEV void world_poly(const float (*local)[2], const BodyS& b, float c, float s, Poly& out) {
  out.n = N;
  const float ms = -s;
#pragma unroll
  for (int k = 0; k < CX_MAXV; ++k) {
    if (k < N) {
      // forward_vector (cotix/_geometry_utils.py:134-138): matrix @ [x, y, 1] / w with
      // w = 0*x + 0*y + 1 = 1 exactly for finite local vertices, so the division is dropped
      float x = local[k][0], y = local[k][1];
      out.x[k] = c * x + ms * y + b.px * 1.0f;
      out.y[k] = s * x + c * y + b.py * 1.0f;
    } else {
      out.x[k] = 0.0f;
      out.y[k] = 0.0f;
    }
  }
}
template <int N>
CX_DEV Aabb poly_aabb(const Poly& p) {
  Aabb a;
  a.lx = a.hx = p.x[0];
  a.ly = a.hy = p.y[0];
#pragma unroll
  for (int k = 1; k < N; ++k) {
    a.lx = fminf(a.lx, p.x[k]);
    a.hx = fmaxf(a.hx, p.x[k]);
    a.ly = fminf(a.ly, p.y[k]);
    a.hy = fmaxf(a.hy, p.y[k]);
  }
  return a;
}
CX_DEV bool aabb_near(const Aabb& a, const Aabb& b) {
  const float m = CX_CULL_MARGIN;
  return !(a.hx + m < b.lx || b.hx + m < a.lx || a.hy + m < b.ly || b.hy + m < a.ly);
}
// Separating-axis test against the top edge TL=(x0,h0) -> TR=(x1,h1) of a ground trapezoid
// (the trapezoid lies in n.(v - TL) <= 0, n = perp(TR - TL)): true when every vertex of the
// body polygon is above that line by more than the cull margin, i.e. the pair cannot touch.
template <int N>
CX_DEV bool above_top_edge(const Poly& p, float x0, float h0, float x1, float h1) {
  const float nx = -(h1 - h0), ny = x1 - x0;
  const float thr = CX_CULL_MARGIN * (fabsf(nx) + fabsf(ny));
  bool all_above = true;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    float s = nx * (p.x[k] - x0) + ny * (p.y[k] - h0);
    all_above = all_above && (s > thr);
  }
  return all_above;
}
CX_DEV float terrain_x(const float kn[4], const CotixLunarParams& P, int i) {
  switch (i) {
    case 0: return -P.terrain_x_outer;
    case 1: return kn[0];
    case 2: return kn[1];
    case 3: return -P.pad_half_width;
    case 4: return P.pad_half_width;
    case 5: return kn[2];
    case 6: return kn[3];
    default: return P.terrain_x_outer;
  }
}

CX_DEV float lunar_shaping(const BodyS& L, const CotixLunarParams& P, bool rc, bool lc) {
  float dy = L.py - P.pad_y;
  float dist = sqrtf(L.px * L.px + dy * dy);
  float speed = sqrtf(L.vx * L.vx + L.vy * L.vy);
  float s = -10.0f * dist;
  s = s - 100.0f * speed;
  s = s - 100.0f * fabsf(L.ang);
  s = s + (rc ? 10.0f : 0.0f);
  s = s + (lc ? 10.0f : 0.0f);
  return s;
}

// all vertices of leg polygon `w`, expressed in the lander frame, fall outside the lander's
// local extents (by more than the cull margin) along one of its two axes
CX_DEV bool outside_lander_obb(const Poly& w, const BodyS& L, float c, float s, const LunarDerived& D) {
  float lox = CX_INF, loy = CX_INF, hix = -CX_INF, hiy = -CX_INF;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float dx = w.x[k] - L.px, dy = w.y[k] - L.py;
    const float lx = c * dx + s * dy, ly = c * dy - s * dx;
    lox = fminf(lox, lx);
    hix = fmaxf(hix, lx);
    loy = fminf(loy, ly);
    hiy = fmaxf(hiy, ly);
  }
  const float m = CX_CULL_MARGIN;
  return (lox > D.lext[2] + m) || (hix < D.lext[0] - m) || (loy > D.lext[3] + m) || (hiy < D.lext[1] - m);
}

// element k (dynamic) of an 8-vector held in registers
CX_DEV float sel8(const float v[8], int k) {
  float r = v[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) r = (k == i) ? v[i] : r;
  return r;
}

// part.transform() re-sorts the world vertices (cotix/_convex_shapes.py:181-187); body = 0
// lander (6 vertices), 1 right leg, 2 left leg (4 vertices)
CX_DEV void order_body_poly(int body, Poly& p, const LunarDerived& D) {
  if (body == 0) {
    if (!(D.rigid_ok[0] && poly_order_rigid<6>(p, D.orient[0]))) poly_sort_cw<6>(p);
  } else {
    const int b = body == 1 ? 1 : 2;
    if (!(D.rigid_ok[b] && poly_order_rigid<4>(p, D.orient[b]))) poly_sort_cw<4>(p);
  }
}

// ---- integrate + broadphase -----------------------------------------------------------------
// Updates E (poses, lander velocity, status), fills cx (sin/cos, throttles, candidate mask) and
// returns the world polygons: with ORDER, in the order part.transform() gives them when there
// are candidates; without, always in local vertex order (the consumer calls order_body_poly).
// the first half of integrate_and_cull on its own: Euler, gravity, trig of the three angles, engines
CX_DEV void integrate_only(EnvRegs& E, float a_main, float a_side, const CotixLunarParams& P, StepCtx& cx) {
  // ExplicitEulerPhysics.step (cotix/_physics_solvers.py:17-24); the ground stays at 0
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    E.b[i].px = E.b[i].px + E.b[i].vx * P.dt;
    E.b[i].py = E.b[i].py + E.b[i].vy * P.dt;
    E.b[i].ang = E.b[i].ang + E.b[i].w * P.dt;
  }
  // gravity on the lander only (examples/test_viz.py:27-31)
  E.b[0].vx = E.b[0].vx + P.gravity_dv[0];
  E.b[0].vy = E.b[0].vy + P.gravity_dv[1];
#pragma unroll
  for (int i = 0; i < 3; ++i) cx_sincosf(E.b[i].ang, &cx.sn[i], &cx.cs[i]);
  if (!(fabsf(E.b[0].ang) < 1.0e9f) || !(fabsf(E.b[1].ang) < 1.0e9f) || !(fabsf(E.b[2].ang) < 1.0e9f))
    E.status |= COTIX_ST_ANGLE_RANGE;
  // engines (new): a zero action leaves the reference step untouched
  cx.thr_main = jclip(a_main, 0.0f, 1.0f);
  cx.thr_side = jclip(a_side, -1.0f, 1.0f);
  if (cx.thr_main != 0.0f) {
    float dv = P.main_engine_dv * cx.thr_main;
    E.b[0].vx = E.b[0].vx + (-cx.sn[0]) * dv;
    E.b[0].vy = E.b[0].vy + cx.cs[0] * dv;
  }
  if (cx.thr_side != 0.0f) {
    float dv = P.side_engine_dv * cx.thr_side;
    E.b[0].vx = E.b[0].vx + cx.cs[0] * dv;
    E.b[0].vy = E.b[0].vy + cx.sn[0] * dv;
    E.b[0].w = E.b[0].w - P.side_engine_dw * cx.thr_side;
  }
}

template <bool ORDER>
CX_DEV void integrate_and_cull(EnvRegs& E, float a_main, float a_side, const CotixLunarParams& P,
                               const LunarDerived& D, StepCtx& cx, Poly& wL, Poly& wr, Poly& wl) {
  integrate_only(E, a_main, a_side, P, cx);
  // world polygons (local vertex order) and the broadphase — new, exact
  world_poly<6>(P.lander_local, E.b[0], cx.cs[0], cx.sn[0], wL);
  world_poly<4>(P.leg_local[0], E.b[1], cx.cs[1], cx.sn[1], wr);
  world_poly<4>(P.leg_local[1], E.b[2], cx.cs[2], cx.sn[2], wl);
  const Aabb bL = poly_aabb<6>(wL), br = poly_aabb<4>(wr), bl = poly_aabb<4>(wl);
  uint32_t cand = 0u;
  if (aabb_near(br, bL) && !outside_lander_obb(wr, E.b[0], cx.cs[0], cx.sn[0], D)) cand |= 1u << 0;
  if (aabb_near(bl, bL) && !outside_lander_obb(wl, E.b[0], cx.cs[0], cx.sn[0], D)) cand |= 1u << 1;
  if (aabb_near(bl, br)) cand |= 1u << 9;
  {
    // ground trapezoids: only those whose x-range comes near the x-range of the assembly are
    // looked at (the same test aabb_near makes first), in a loop whose trip count is the number
    // of such segments (1-3) rather than 7, so the lanes of a warp stay together
    const float m = CX_CULL_MARGIN;
    const float ulx = fminf(bL.lx, fminf(br.lx, bl.lx)), uhx = fmaxf(bL.hx, fmaxf(br.hx, bl.hx));
    const float uly = fminf(bL.ly, fminf(br.ly, bl.ly));
    float X[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) X[k] = terrain_x(E.kn, P, k);
    uint32_t segs = 0u;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const float glx = fminf(X[k], X[k + 1]), ghx = fmaxf(X[k], X[k + 1]);
      // the segment's box against the box of the whole assembly, in x AND in y (its top is the higher of its
      // two heights and the floor): what the per-polygon aabb_near below would reject anyway, decided here
      // with static indices so that an env in flight enters the loop with no segment at all -- the loop's
      // trip count is the maximum over the warp.  Negated form: NaN bounds keep the segment.
      const float ghy = fmaxf(fmaxf(E.h[k], E.h[k + 1]), P.ground_floor);
      if (!(uhx + m < glx || ghx + m < ulx || ghy + m < uly)) segs |= 1u << k;
    }
    while (segs) {
      const int k = __ffs(segs) - 1;
      segs &= segs - 1u;
      const float x0 = sel8(X, k), x1 = sel8(X, k + 1);
      const float h0 = sel8(E.h, k), h1 = sel8(E.h, k + 1);
      Aabb g;
      g.lx = fminf(x0, x1);
      g.hx = fmaxf(x0, x1);
      g.ly = fminf(fminf(h0, h1), P.ground_floor);
      g.hy = fmaxf(fmaxf(h0, h1), P.ground_floor);
      const bool regular = (x1 > x0) && (h0 > P.ground_floor) && (h1 > P.ground_floor);
      if (aabb_near(g, bL) && !(regular && above_top_edge<6>(wL, x0, h0, x1, h1))) cand |= 1u << (2 + k);
      if (aabb_near(g, br) && !(regular && above_top_edge<4>(wr, x0, h0, x1, h1))) cand |= 1u << (11 + k);
      if (aabb_near(g, bl) && !(regular && above_top_edge<4>(wl, x0, h0, x1, h1))) cand |= 1u << (18 + k);
    }
  }
  bool bad = (E.status & COTIX_ST_GROUND_NAN) != 0u;
#pragma unroll
  for (int i = 0; i < 3; ++i)
    bad = bad || !(fabsf(E.b[i].px) <= 3.0e38f) || !(fabsf(E.b[i].py) <= 3.0e38f) ||
          !(fabsf(E.b[i].ang) <= 3.0e38f);
  if (bad) cand = 0u;  // NaN vertices: every GJK of the reference reports "no" (any-NaN rule)
  // the self pair (left leg vs itself) can only matter when body 2 has another
  // live cell, i.e. when pair 1 or pair 9 is a candidate
  if (cand & ((1u << 1) | (1u << 9))) cand |= 1u << 10;
  cx.cand = cand;
  cx.base = -1;
  if (ORDER && cand != 0u) {
    order_body_poly(0, wL, D);
    order_body_poly(1, wr, D);
    order_body_poly(2, wl, D);
  }
}

// ---- the two polygons of candidate pair `pid` ---------------------------------------------------
// PS (polygon source) provides body(which, Poly&) for which = 0 lander, 1 right leg, 2 left leg
// and terrain(h[8], kn[4]).
CX_DEV void pair_bodies(int pid, int& wa, int& wb, int& gk) {
  gk = -1;
  if (pid == 0) { wa = 1; wb = 0; }
  else if (pid == 1) { wa = 2; wb = 0; }
  else if (pid <= 8) { gk = pid - 2; wa = -1; wb = 0; }
  else if (pid == 9) { wa = 2; wb = 1; }
  else if (pid == 10) { wa = 2; wb = 2; }
  else if (pid <= 17) { gk = pid - 11; wa = -1; wb = 1; }
  else { gk = pid - 18; wa = -1; wb = 2; }
}
// ground part k as the collider sees it: Polygon4.__init__ sort (cotix/_lunar_lander.py:132)
// followed by the identity transform + re-sort of part.transform (cotix/_colliders.py:92-94)
#ifndef CX_GROUND_FAST
#define CX_GROUND_FAST 1
#endif
// One order_clockwise pass over a quadrilateral given in COUNTER-CLOCKWISE cyclic order, without computing an
// angle: the ascending-atan2 order about the centroid is the same cycle started at the one vertex below the
// centroid whose predecessor is not (cx_atan2f's branch cut: dy < 0 gives a negative angle, dy = -0 gives +pi).
// The centroid and the differences are computed with exactly the operations of poly_sort_cw (running sums
// from 0 in vertex order, / 4), so the sign of every dy is the sign poly_sort_cw's atan2 argument has.
// check_gaps: also require every angular gap between neighbours to exceed 0.05 rad (atan2 errors are 1e-6),
// which makes the cyclic order the sorted order.  Returns false (c untouched) when the pattern is not found.
CX_DEV bool quad_order_pass(float cxv[4], float cyv[4], bool check_gaps) {
  float sx = 0.0f, sy = 0.0f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    sx = sx + cxv[k];
    sy = sy + cyv[k];
  }
  const float mx = sx / 4.0f, my = sy / 4.0f;
  float dx[4], dy[4];
  bool neg[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    dy[k] = cyv[k] - my;
    dx[k] = cxv[k] - mx;
    neg[k] = dy[k] < 0.0f;
  }
  bool ok = true;
  if (check_gaps) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int j = (k + 1) & 3;
      const float cr = dx[k] * dy[j] - dy[k] * dx[j];
      const float dt = dx[k] * dx[j] + dy[k] * dy[j];
      ok = ok && (cr > 0.0f) && (cr > 0.05f * dt);
    }
  }
  int start = 0, nstart = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (neg[k] && !neg[(k + 3) & 3]) {
      start = k;
      ++nstart;
    }
  }
  if (!ok || nstart != 1) return false;
  float ox[4], oy[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    ox[t] = cxv[0];
    oy[t] = cyv[0];
#pragma unroll
    for (int k = 1; k < 4; ++k)
      if (((start + t) & 3) == k) {
        ox[t] = cxv[k];
        oy[t] = cyv[k];
      }
  }
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    cxv[t] = ox[t];
    cyv[t] = oy[t];
  }
  return true;
}
CX_DEV void ground_poly(float x0, float h0, float x1, float h1, const CotixLunarParams& P, Poly& g) {
  g.n = 4;
  g.x[4] = g.y[4] = g.x[5] = g.y[5] = 0.0f;
  const float f = P.ground_floor;
#if CX_GROUND_FAST
  // A regular trapezoid (left of right, tops above the floor) of ordinary size: its two order_clockwise passes
  // (8 atan2 + 2 sorts, a tenth of the narrowphase's instructions) reduce to two rotations of the cycle
  // bottom-left -> bottom-right -> top-right -> top-left.  The bounds keep the rounding of the second pass's
  // centroid (a different summation order) three orders of magnitude below the gap that was checked.
  const float lim = 1024.0f;
  if ((x1 - x0 >= 0.5f) && (h0 > f) && (h1 > f) && (fabsf(x0) <= lim) && (fabsf(x1) <= lim) && (fabsf(h0) <= lim) &&
      (fabsf(h1) <= lim) && (fabsf(f) <= lim)) {
    // poly_sort_cw's centroid sums the vertices in the order they are stored: raw order first
    const float sx = ((0.0f + x0) + x0 + x1) + x1, sy = ((0.0f + h0) + f + h1) + f;
    const float mx = sx / 4.0f, my = sy / 4.0f;
    // counter-clockwise cycle of the raw vertices: (x0, f), (x1, f), (x1, h1), (x0, h0)
    float cxv[4] = {x0, x1, x1, x0}, cyv[4] = {f, f, h1, h0};
    float dx[4], dy[4];
    bool neg[4], ok = true;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      dy[k] = cyv[k] - my;
      dx[k] = cxv[k] - mx;
      neg[k] = dy[k] < 0.0f;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int j = (k + 1) & 3;
      const float cr = dx[k] * dy[j] - dy[k] * dx[j];
      const float dt = dx[k] * dx[j] + dy[k] * dy[j];
      ok = ok && (cr > 0.0f) && (cr > 0.05f * dt);
    }
    int start = 0, nstart = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (neg[k] && !neg[(k + 3) & 3]) {
        start = k;
        ++nstart;
      }
    }
    if (ok && nstart == 1) {
      float px[4], py[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        px[t] = cxv[0];
        py[t] = cyv[0];
#pragma unroll
        for (int k = 1; k < 4; ++k)
          if (((start + t) & 3) == k) {
            px[t] = cxv[k];
            py[t] = cyv[k];
          }
      }
      if (quad_order_pass(px, py, false)) {  // the second pass, on the stored order of the first
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          g.x[t] = px[t];
          g.y[t] = py[t];
        }
        return;
      }
    }
  }
#endif
  g.x[0] = x0; g.y[0] = h0;
  g.x[1] = x0; g.y[1] = f;
  g.x[2] = x1; g.y[2] = h1;
  g.x[3] = x1; g.y[3] = f;
  poly_sort_cw<4>(g);
  poly_sort_cw<4>(g);
}
template <class PS>
CX_DEV void make_item_polys(const PS& src, int pid, const CotixLunarParams& P, Poly& A, Poly& B) {
  int wa, wb, gk;
  pair_bodies(pid, wa, wb, gk);
  if (gk >= 0) {
    float x0, h0, x1, h1;
    src.ground_edge(gk, P, x0, h0, x1, h1);
    ground_poly(x0, h0, x1, h1, P, A);
  } else {
    src.body(wa, A);
  }
  src.body(wb, B);
}

// polygon_vs_polygon of one candidate pair.  The first polygon is a leg or a ground trapezoid — always 4
// vertices (pair_bodies) — and is compiled with a static vertex count; the second (a leg, 4, or the lander,
// 6) keeps its dynamic count so that the lanes of a warp run ONE instruction stream whatever their pair
// (two static instantiations were measured slower: narrow stage 56 -> 61 us).
static __device__ __noinline__ Contact poly_pair(const Poly& a, const Poly& b, V2 start, int gjk_max) {
  return convex_vs_polygon(poly_fixed<4>(a), b, start, gjk_max);
}

// ---- collider given the contacts — cotix/_colliders.py:137-351 ------------------------------------
// The 7 all_contacts cells that can be populated for the LunarLander body set,
// with the scatter-scan entries that feed each of them:
//   cell   (i, j)          group  N1  N2  ind2 range  ind1 = a0 + q + nq*r, q < nq, r < reps
//   0 v10  (rl, L)           0     9   9    0..8       a0 0, nq 1, reps 1
//   1 v20  (ll, L)           0     9   9    0..8       a0 1, nq 1, reps 1
//   2 w30  (ground, L)       0     9   9    0..8       a0 2, nq 7, reps 1
//   3 v21  (ll, rl)          1    15  15    0..7       a0 0, nq 1, reps 1
//   4 w31  (ground, rl)      1    15  15    0..7       a0 1, nq 7, reps 2
//   5 w32  (ground, ll)      1    15  15    8..14      a0 1, nq 7, reps 2
//   6 v22  (ll, ll) self     1    15  15    8..14      a0 0, nq 1, reps 1
struct CellDesc {
  int g, N1, N2, b_lo, b_hi, a0, nq, reps;
};
CX_DEV CellDesc cell_desc(int c) {
  CellDesc d;
  d.g = c >= 3 ? 1 : 0;
  d.N1 = d.N2 = c >= 3 ? 15 : 9;
  d.b_lo = c >= 5 ? 8 : 0;
  d.b_hi = c <= 2 ? 8 : (c <= 4 ? 7 : 14);
  d.a0 = (c == 0 || c == 3 || c == 6) ? 0 : (c == 2 ? 2 : 1);
  d.nq = (c == 2 || c == 4 || c == 5) ? 7 : 1;
  d.reps = (c == 4 || c == 5) ? 2 : 1;
  return d;
}

// live masks of the cells (cotix/_colliders.py:137-140) from `live` = the candidate pairs whose
// contact is non-NaN (bit = pair id): m[c] bit q = entry q of cell c is live.  m[6] (the self
// pair) is filled only when want_self.
CX_DEV bool cell_masks(uint32_t live, bool want_self, uint32_t m[7], bool& crash, bool& rc, bool& lc) {
  m[0] = live & 1u;
  m[1] = (live >> 1) & 1u;
  m[2] = (live >> 2) & 0x7Fu;
  m[3] = (live >> 9) & 1u;
  m[4] = (live >> 11) & 0x7Fu;
  m[5] = (live >> 18) & 0x7Fu;
  crash = m[2] != 0u;
  rc = m[4] != 0u;
  lc = m[5] != 0u;
  const bool any = (m[0] | m[1] | m[2] | m[3] | m[4] | m[5]) != 0u;
  m[6] = (any && want_self) ? ((live >> 10) & 1u) : 0u;
  return any;
}
// the same, asking the contact source pair by pair (the self pair, bit 10, only if want_self)
template <class CS>
CX_DEV bool collide_masks(uint32_t cand, const CS& cs, bool want_self, uint32_t m[7], bool& crash,
                          bool& rc, bool& lc) {
  uint32_t live = 0u;
  uint32_t todo = want_self ? cand : (cand & ~(1u << 10));
  while (todo) {
    const int bit = __ffs(todo) - 1;
    todo &= todo - 1u;
    if (cs.live(bit)) live |= 1u << bit;
  }
  return cell_masks(live, want_self, m, crash, rc, lc);
}

// key chain: skey = split(rkey)[0] (:142); per group skey = split(skey)[0] (:175)
CX_DEV void make_coin_ctx(const EnvRegs& E, const CotixLunarParams& P, CoinCtx& cc) {
  cc.mode = P.rng_mode;
  cc.p = P.coin_p;
  cc.steps = E.steps;
  cc.episode = E.status >> 8;
  cc.pkey.a = P.philox_seed[0] ^ E.key.a;
  cc.pkey.b = P.philox_seed[1] ^ E.key.b;
  cc.gkey[0] = cc.gkey[1] = E.key;
  if (P.rng_mode == COTIX_RNG_THREEFRY) {
    Key2 sk = tf_split(E.key, 2u, 0u);
    cc.gkey[0] = tf_split(sk, 2u, 0u);
    cc.gkey[1] = tf_split(cc.gkey[0], 2u, 0u);
  }
}

// choose_random_contact (:274-295) + the sequential resolution scan (:310-336), given the winner
// of every cell (win[c] = winning q or -1).  Updates E (velocities, status).
template <class CS, bool INLINE_RESOLVE = false>
CX_DEV void collide_apply(EnvRegs& E, const CotixLunarParams& P, const LunarDerived& D, const CS& cs,
                          const CoinCtx& cc, const int win[7]) {
#define CX_RESOLVE(...)                                        \
  do {                                                         \
    if (INLINE_RESOLVE) resolve_collision_inl<true>(__VA_ARGS__); \
    else resolve_collision(__VA_ARGS__);                       \
  } while (0)
  MassP mp[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) mp[i] = mk_massp(P.body_props[i], D, i);
  const bool poisoned = (E.status & COTIX_ST_GROUND_NAN) != 0u;
  BodyS G;  // the static ground body (mass = inertia = inf); NaN once poisoned
  G.px = G.py = G.vx = G.vy = G.ang = G.w = poisoned ? CX_NAN : 0.0f;
  G.px = G.px + G.vx * P.dt;
  G.py = G.py + G.vy * P.dt;
  G.ang = G.ang + G.w * P.dt;
  const bool v10 = win[0] >= 0, v20 = win[1] >= 0, v21 = win[3] >= 0, v22 = win[6] >= 0;
  const int w30 = win[2], w31 = win[4], w32 = win[5];
  // keys split(skey, n_bodies) with the last skey
  int j2 = -1, j3 = -1;
  {
    const int cnt = (v20 ? 1 : 0) + (v21 ? 1 : 0) + (v22 ? 1 : 0);
    if (cnt == 1) j2 = v20 ? 0 : (v21 ? 1 : 2);
    else if (cnt > 1) {
      float u;
      if (P.rng_mode == COTIX_RNG_THREEFRY) u = tf_uniform(tf_split(cc.gkey[1], 4u, 2u), 0.0f, 1.0f);
      else u = bits_to_uniform(philox_word(cc.pkey, cc.steps, cc.episode, 0u, 2u, CX_TAG_CHOICE), 0.0f, 1.0f);
      j2 = choose_index(v20, v21, v22, u);
    }
  }
  {
    const bool m0 = w30 >= 0, m1 = w31 >= 0, m2 = w32 >= 0;
    const int cnt = (m0 ? 1 : 0) + (m1 ? 1 : 0) + (m2 ? 1 : 0);
    if (cnt == 1) j3 = m0 ? 0 : (m1 ? 1 : 2);
    else if (cnt > 1) {
      float u;
      if (P.rng_mode == COTIX_RNG_THREEFRY) u = tf_uniform(tf_split(cc.gkey[1], 4u, 3u), 0.0f, 1.0f);
      else u = bits_to_uniform(philox_word(cc.pkey, cc.steps, cc.episode, 0u, 3u, CX_TAG_CHOICE), 0.0f, 1.0f);
      j3 = choose_index(m0, m1, m2, u);
    }
  }
  // sequential resolution scan i = 0..3 (:310-336); row 0 is always empty
  if (v10) {
    const Contact c = cs.get(0);
    CX_RESOLVE(E.b[1], mp[1], E.b[0], mp[0], c.pen, c.cp, P.baumgarte, P.baumgarte_dt);
  }
  if (j2 == 0 && v20) {
    const Contact c = cs.get(1);
    CX_RESOLVE(E.b[2], mp[2], E.b[0], mp[0], c.pen, c.cp, P.baumgarte, P.baumgarte_dt);
  }
  if (j2 == 1 && v21) {
    const Contact c = cs.get(9);
    CX_RESOLVE(E.b[2], mp[2], E.b[1], mp[1], c.pen, c.cp, P.baumgarte, P.baumgarte_dt);
  }
  if (j3 >= 0 && j3 <= 2) {
    const int w = j3 == 0 ? w30 : (j3 == 1 ? w31 : w32);
    if (w >= 0) {
      const Contact c = cs.get((j3 == 0 ? 2 : (j3 == 1 ? 11 : 18)) + w);
      if (j3 == 0) CX_RESOLVE(G, mp[3], E.b[0], mp[0], c.pen, c.cp, P.baumgarte, P.baumgarte_dt);
      else if (j3 == 1) CX_RESOLVE(G, mp[3], E.b[1], mp[1], c.pen, c.cp, P.baumgarte, P.baumgarte_dt);
      else CX_RESOLVE(G, mp[3], E.b[2], mp[2], c.pen, c.cp, P.baumgarte, P.baumgarte_dt);
    }
  }
  if ((G.vx != 0.0f) || (G.vy != 0.0f) || (G.w != 0.0f) || (G.px != 0.0f) || (G.py != 0.0f) || (G.ang != 0.0f))
    E.status |= COTIX_ST_GROUND_NAN;
#undef CX_RESOLVE
}

// one thread does everything for its env (the fused kernel)
template <class CS>
CX_DEV void collide(EnvRegs& E, const CotixLunarParams& P, const LunarDerived& D, uint32_t cand, const CS& cs,
                    bool& crash, bool& rc, bool& lc) {
  uint32_t m[7];
  if (!collide_masks(cand, cs, false, m, crash, rc, lc)) return;
  CoinCtx cc;
  make_coin_ctx(E, P, cc);
  int win[7];
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const CellDesc d = cell_desc(c);
    win[c] = cell_winner(cc, d.g, d.N1, d.N2, d.b_lo, d.b_hi, d.a0, d.nq, d.reps, m[c]);
  }
  win[6] = -1;
  if (win[1] >= 0 || win[3] >= 0) {  // the self pair only matters when body 2 has another live cell
    const CellDesc d = cell_desc(6);
    win[6] = cell_winner(cc, d.g, d.N1, d.N2, d.b_lo, d.b_hi, d.a0, d.nq, d.reps, cs.live(10) ? 1u : 0u);
  }
  collide_apply(E, P, D, cs, cc, win);
}

// ---- key split, joints, judge --------------------------------------------------------------------
// cs/sn: sin/cos of the three body angles of THIS step (after integration).
template <bool INLINE_JOINTS = false>
CX_DEV void finish_step(EnvRegs& E, const CotixLunarParams& P, const LunarDerived& D, const float cs[3], const float sn[3],
                        float thr_main, float thr_side, bool crash, bool rc, bool lc, SubInfo& info) {
  MassP mp[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) mp[i] = mk_massp(P.body_props[i], D, i);
  // key, next_key = jr.split(key) (examples/test_viz.py:39)
  if (P.rng_mode == COTIX_RNG_THREEFRY) E.key = tf_split(E.key, 2u, 0u);

  // LunarLander.step (cotix/_lunar_lander.py:145-218); angles are those of this step
  {
    V2 p0 = mk2(E.b[0].px, E.b[0].py);
    V2 lj0 = rot_cs(mk2(P.joint_local[0][0], P.joint_local[0][1]), cs[0], sn[0]) + p0;
    V2 lj1 = rot_cs(mk2(P.joint_local[1][0], P.joint_local[1][1]), cs[0], sn[0]) + p0;
    V2 lj2 = rot_cs(mk2(P.joint_local[2][0], P.joint_local[2][1]), cs[0], sn[0]) + p0;
    V2 lj3 = rot_cs(mk2(P.joint_local[3][0], P.joint_local[3][1]), cs[0], sn[0]) + p0;
    V2 j2l = mk2(P.leg_joint2_local[0], P.leg_joint2_local[1]);
    V2 ll1 = mk2(E.b[2].px, E.b[2].py);
    V2 ll2 = ll1 + rot_cs(j2l, cs[2], sn[2]);
    V2 rl1 = mk2(E.b[1].px, E.b[1].py);
    V2 rl2 = rl1 + rot_cs(j2l, cs[1], sn[1]);
    // SimpleConstraintSolver.solve (cotix/_constraint_solvers.py:13-17): `loops` sequential passes
    // over the constraint list; the anchors are those of the step (positions do not move here)
    const int loops = P.solver_loops < 1 ? 1 : P.solver_loops;
    for (int it = 0; it < loops; ++it) {
      if (INLINE_JOINTS) {
        joint_constraint_inl(E.b[0], mp[0], lj0, E.b[2], mp[2], ll1, P);
        joint_constraint_inl(E.b[0], mp[0], lj1, E.b[2], mp[2], ll2, P);
        joint_constraint_inl(E.b[0], mp[0], lj2, E.b[1], mp[1], rl1, P);
        joint_constraint_inl(E.b[0], mp[0], lj3, E.b[1], mp[1], rl2, P);
      } else {
        joint_constraint(E.b[0], mp[0], lj0, E.b[2], mp[2], ll1, P);
        joint_constraint(E.b[0], mp[0], lj1, E.b[2], mp[2], ll2, P);
        joint_constraint(E.b[0], mp[0], lj2, E.b[1], mp[1], rl1, P);
        joint_constraint(E.b[0], mp[0], lj3, E.b[1], mp[1], rl2, P);
      }
    }
    E.b[1].w = E.b[1].w * P.leg_damping;
    E.b[2].w = E.b[2].w * P.leg_damping;
  }

  bool any_nan = false;
#pragma unroll
  for (int i = 0; i < 3; ++i)
    any_nan = any_nan || !(fabsf(E.b[i].px) <= 3.0e38f) || !(fabsf(E.b[i].py) <= 3.0e38f) ||
              !(fabsf(E.b[i].vx) <= 3.0e38f) || !(fabsf(E.b[i].vy) <= 3.0e38f) ||
              !(fabsf(E.b[i].ang) <= 3.0e38f) || !(fabsf(E.b[i].w) <= 3.0e38f);
  if (any_nan) E.status |= COTIX_ST_NAN;
  E.steps += 1u;

  // judge (new; template cotix/_envs.py:9-28)
  float reward = 0.0f;
  int done = COTIX_DONE_NO;
  if (P.judge_enabled) {
    float sh = lunar_shaping(E.b[0], P, rc, lc);
    if (E.shaping == E.shaping) reward = sh - E.shaping;
    E.shaping = sh;
    reward = reward - 0.3f * thr_main;
    reward = reward - 0.03f * fabsf(thr_side);
    const BodyS& L = E.b[0];
    if (crash) {
      done = COTIX_DONE_CRASH;
      reward = -100.0f;
    } else if (!(fabsf(L.px) < P.x_limit) || !(L.py < P.y_limit)) {
      done = COTIX_DONE_OOB;
      reward = -100.0f;
    } else if (rc && lc && sqrtf(L.vx * L.vx + L.vy * L.vy) < P.land_speed && fabsf(L.w) < P.land_angvel) {
      done = COTIX_DONE_LANDED;
      reward = 100.0f;
    }
  }
  if (any_nan) {
    done = COTIX_DONE_NAN;
    reward = P.judge_enabled ? -100.0f : 0.0f;
  }
  if (done == COTIX_DONE_NO && P.max_episode_steps > 0 && E.steps >= (uint32_t)P.max_episode_steps)
    done = COTIX_DONE_TIMEOUT;
  E.ret = E.ret + reward;
  info.reward = reward;
  info.done = done;
  info.rc = rc;
  info.lc = lc;
}

CX_DEV void lunar_reset_finish(EnvRegs& E, const uint32_t pert[8], const CotixLunarParams& P, uint32_t episode);
// reset(key): LunarLander.__init__, cotix/_lunar_lander.py:29-143 (+ optional pose
// perturbation, new).  Leaves E.key untouched; the caller owns the key chain.
static __device__ __noinline__ void lunar_reset_env(EnvRegs& E, Key2 key, const CotixLunarParams& P,
                                                    uint32_t episode) {
  uint32_t pert[8];
  if (P.rng_mode == COTIX_RNG_THREEFRY) {
    Key2 k0 = tf_split(key, 5u, 0u);
#pragma unroll
    for (int i = 0; i < 8; ++i) E.h[i] = bits_to_uniform(tf_bits_word(k0, 8u, (uint32_t)i), -5.0f, 5.0f);
    E.kn[0] = tf_uniform(tf_split(key, 5u, 1u), -12.0f, -9.0f);
    E.kn[1] = tf_uniform(tf_split(key, 5u, 2u), -8.0f, -4.0f);
    E.kn[2] = tf_uniform(tf_split(key, 5u, 3u), 4.0f, 8.0f);
    E.kn[3] = tf_uniform(tf_split(key, 5u, 4u), 9.0f, 12.0f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint2 o = P.reset_perturb ? tf_call(key.a, key.b, 0x70657274u + (uint32_t)i, (uint32_t)i)
                                : make_uint2(0u, 0u);
      pert[2 * i] = o.x;
      pert[2 * i + 1] = o.y;
    }
  } else {
    uint32_t w[20];
#pragma unroll
    for (int q = 0; q < 5; ++q)
      philox4x32_10(P.philox_seed[0] ^ key.a, P.philox_seed[1] ^ key.b, 0u, episode, (uint32_t)q,
                    CX_TAG_RESET, w + 4 * q);
#pragma unroll
    for (int i = 0; i < 8; ++i) E.h[i] = bits_to_uniform(w[i], -5.0f, 5.0f);
    E.kn[0] = bits_to_uniform(w[8], -12.0f, -9.0f);
    E.kn[1] = bits_to_uniform(w[9], -8.0f, -4.0f);
    E.kn[2] = bits_to_uniform(w[10], 4.0f, 8.0f);
    E.kn[3] = bits_to_uniform(w[11], 9.0f, 12.0f);
#pragma unroll
    for (int i = 0; i < 8; ++i) pert[i] = w[12 + i];
  }
  lunar_reset_finish(E, pert, P, episode);
}
// the part of reset(key) after the random draws: E.h[0..7] hold the uniform(-5, 5) heights, E.kn the
// x-knots, pert the eight words of the pose perturbation
CX_DEV void lunar_reset_finish(EnvRegs& E, const uint32_t pert[8], const CotixLunarParams& P, uint32_t episode) {
  E.h[0] = E.h[0] * 10.0f;
  E.h[3] = P.pad_y;
  E.h[4] = P.pad_y;
  E.h[7] = E.h[7] * 10.0f;

  float lx = P.lander_start[0], ly = P.lander_start[1], la = P.lander_start[2];
  float lvx = 0.0f, lvy = 0.0f, lw = 0.0f;
  if (!P.reset_perturb) {
#pragma unroll
    for (int l = 0; l < 2; ++l) {
      E.b[1 + l].px = P.leg_offset[l][0] + lx;
      E.b[1 + l].py = P.leg_offset[l][1] + ly;
      E.b[1 + l].vx = E.b[1 + l].vy = E.b[1 + l].ang = E.b[1 + l].w = 0.0f;
    }
  } else {
    lx = bits_to_uniform(pert[0], P.reset_x[0], P.reset_x[1]);
    ly = bits_to_uniform(pert[1], P.reset_y[0], P.reset_y[1]);
    la = bits_to_uniform(pert[2], P.reset_angle[0], P.reset_angle[1]);
    lvx = bits_to_uniform(pert[3], -P.reset_v, P.reset_v);
    lvy = bits_to_uniform(pert[4], -P.reset_v, P.reset_v);
    lw = bits_to_uniform(pert[5], -P.reset_w, P.reset_w);
    float da = la - P.lander_start[2];
    float c, s;
    cx_sincosf(da, &s, &c);
#pragma unroll
    for (int l = 0; l < 2; ++l) {
      float ox = P.leg_offset[l][0], oy = P.leg_offset[l][1];
      float rx = c * ox - s * oy, ry = s * ox + c * oy;
      E.b[1 + l].px = rx + lx;
      E.b[1 + l].py = ry + ly;
      E.b[1 + l].ang = da;
      E.b[1 + l].vx = lvx + (-ry) * lw;
      E.b[1 + l].vy = lvy + rx * lw;
      E.b[1 + l].w = lw;
    }
  }
  E.b[0].px = lx;
  E.b[0].py = ly;
  E.b[0].vx = lvx;
  E.b[0].vy = lvy;
  E.b[0].ang = la;
  E.b[0].w = lw;
  E.steps = 0u;
  E.status = episode << 8;
  E.ret = 0.0f;
  E.shaping = CX_NAN;
}
// lunar_auto_reset for one env by EIGHT consecutive lanes (sub = lane & 7; all 32 lanes of the warp
// must call it, `valid` says whether this group has an env).  The ~30 dependent threefry calls of
// a reset (key chain, 8 heights, 4 knots, 4 perturbation pairs) are independent of each other, so
// each lane takes its share — at most 8 in a row — and the group's first lane collects the draws
// by shuffle and finishes.  Same calls, same arguments: same bits.  E is complete in lane sub == 0.
CX_DEV void lunar_auto_reset_coop(EnvRegs& E, const CotixLunarParams& P, bool valid) {
  const unsigned lane = threadIdx.x & 31u, sub = lane & 7u, g0 = lane & ~7u;
  const uint32_t episode = ((E.status >> 8) + 1u) & 0xFFFFFFu;
  Key2 rk = E.key;
  float hv = 0.0f, kv = 0.0f;
  uint32_t w0 = 0u, w1 = 0u, w2 = 0u, w3 = 0u;
  if (valid) {
    if (P.rng_mode == COTIX_RNG_THREEFRY) {
      rk = tf_split(E.key, 2u, 1u);
      if (sub == 0u) E.key = tf_split(E.key, 2u, 0u);
      const Key2 k0 = tf_split(rk, 5u, 0u);
      hv = bits_to_uniform(tf_bits_word(k0, 8u, sub), -5.0f, 5.0f);
      if (sub < 4u) {
        const float lo = sub == 0u ? -12.0f : (sub == 1u ? -8.0f : (sub == 2u ? 4.0f : 9.0f));
        const float hi = sub == 0u ? -9.0f : (sub == 1u ? -4.0f : (sub == 2u ? 8.0f : 12.0f));
        kv = tf_uniform(tf_split(rk, 5u, 1u + sub), lo, hi);
      } else if (P.reset_perturb) {
        const uint32_t i = sub - 4u;
        const uint2 o = tf_call(rk.a, rk.b, 0x70657274u + i, i);
        w0 = o.x;
        w1 = o.y;
      }
    } else if (sub < 5u) {
      uint32_t w[4];
      philox4x32_10(P.philox_seed[0] ^ rk.a, P.philox_seed[1] ^ rk.b, 0u, episode, sub, CX_TAG_RESET, w);
      w0 = w[0]; w1 = w[1]; w2 = w[2]; w3 = w[3];
    }
  }
  uint32_t pert[8];
  if (P.rng_mode == COTIX_RNG_THREEFRY) {
#pragma unroll
    for (int i = 0; i < 8; ++i) E.h[i] = __shfl_sync(0xFFFFFFFFu, hv, (int)g0 + i);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      E.kn[i] = __shfl_sync(0xFFFFFFFFu, kv, (int)g0 + i);
      pert[2 * i] = __shfl_sync(0xFFFFFFFFu, w0, (int)g0 + 4 + i);
      pert[2 * i + 1] = __shfl_sync(0xFFFFFFFFu, w1, (int)g0 + 4 + i);
    }
  } else {
    uint32_t w[20];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
      w[4 * q + 0] = __shfl_sync(0xFFFFFFFFu, w0, (int)g0 + q);
      w[4 * q + 1] = __shfl_sync(0xFFFFFFFFu, w1, (int)g0 + q);
      w[4 * q + 2] = __shfl_sync(0xFFFFFFFFu, w2, (int)g0 + q);
      w[4 * q + 3] = __shfl_sync(0xFFFFFFFFu, w3, (int)g0 + q);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) E.h[i] = bits_to_uniform(w[i], -5.0f, 5.0f);
    E.kn[0] = bits_to_uniform(w[8], -12.0f, -9.0f);
    E.kn[1] = bits_to_uniform(w[9], -8.0f, -4.0f);
    E.kn[2] = bits_to_uniform(w[10], 4.0f, 8.0f);
    E.kn[3] = bits_to_uniform(w[11], 9.0f, 12.0f);
#pragma unroll
    for (int i = 0; i < 8; ++i) pert[i] = w[12 + i];
  }
  if (valid && sub == 0u) lunar_reset_finish(E, pert, P, episode);
}
// auto-reset of a finished env (same launch): fresh key from the chain, next episode number
CX_DEV void lunar_auto_reset(EnvRegs& E, const CotixLunarParams& P) {
  Key2 rk = E.key;
  const uint32_t episode = ((E.status >> 8) + 1u) & 0xFFFFFFu;
  if (P.rng_mode == COTIX_RNG_THREEFRY) {
    rk = tf_split(E.key, 2u, 1u);
    E.key = tf_split(E.key, 2u, 0u);
  }
  lunar_reset_env(E, rk, P, episode);
}

// ---- state I/O: struct-of-arrays over the env axis, one coalesced transaction per field -------
CX_DEV void load_dyn(EnvRegs& E, const float* __restrict__ s, int64_t ld, int64_t e) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    E.b[i].px = s[(i * 6 + 0) * ld + e];
    E.b[i].py = s[(i * 6 + 1) * ld + e];
    E.b[i].vx = s[(i * 6 + 2) * ld + e];
    E.b[i].vy = s[(i * 6 + 3) * ld + e];
    E.b[i].ang = s[(i * 6 + 4) * ld + e];
    E.b[i].w = s[(i * 6 + 5) * ld + e];
  }
}
CX_DEV void load_terrain(EnvRegs& E, const float* __restrict__ s, int64_t ld, int64_t e) {
#pragma unroll
  for (int i = 0; i < 8; ++i) E.h[i] = s[(COTIX_F_HEIGHT0 + i) * ld + e];
#pragma unroll
  for (int i = 0; i < 4; ++i) E.kn[i] = s[(COTIX_F_KNOT0 + i) * ld + e];
}
CX_DEV void load_meta(EnvRegs& E, const float* __restrict__ s, int64_t ld, int64_t e) {
  E.key.a = __float_as_uint(s[COTIX_F_KEY0 * ld + e]);
  E.key.b = __float_as_uint(s[COTIX_F_KEY1 * ld + e]);
  E.steps = __float_as_uint(s[COTIX_F_STEPS * ld + e]);
  E.status = __float_as_uint(s[COTIX_F_STATUS * ld + e]);
  E.ret = s[COTIX_F_RETURN * ld + e];
  E.shaping = s[COTIX_F_SHAPING * ld + e];
}
CX_DEV void load_env(EnvRegs& E, const float* __restrict__ s, int64_t ld, int64_t e) {
  load_dyn(E, s, ld, e);
  load_terrain(E, s, ld, e);
  load_meta(E, s, ld, e);
}
CX_DEV void store_dyn(const EnvRegs& E, float* __restrict__ s, int64_t ld, int64_t e) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    s[(i * 6 + 0) * ld + e] = E.b[i].px;
    s[(i * 6 + 1) * ld + e] = E.b[i].py;
    s[(i * 6 + 2) * ld + e] = E.b[i].vx;
    s[(i * 6 + 3) * ld + e] = E.b[i].vy;
    s[(i * 6 + 4) * ld + e] = E.b[i].ang;
    s[(i * 6 + 5) * ld + e] = E.b[i].w;
  }
}
CX_DEV void store_terrain(const EnvRegs& E, float* __restrict__ s, int64_t ld, int64_t e) {
#pragma unroll
  for (int i = 0; i < 8; ++i) s[(COTIX_F_HEIGHT0 + i) * ld + e] = E.h[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) s[(COTIX_F_KNOT0 + i) * ld + e] = E.kn[i];
}
CX_DEV void store_meta(const EnvRegs& E, float* __restrict__ s, int64_t ld, int64_t e) {
  s[COTIX_F_KEY0 * ld + e] = __uint_as_float(E.key.a);
  s[COTIX_F_KEY1 * ld + e] = __uint_as_float(E.key.b);
  s[COTIX_F_STEPS * ld + e] = __uint_as_float(E.steps);
  s[COTIX_F_STATUS * ld + e] = __uint_as_float(E.status);
  s[COTIX_F_RETURN * ld + e] = E.ret;
  s[COTIX_F_SHAPING * ld + e] = E.shaping;
}
CX_DEV void store_env(const EnvRegs& E, float* __restrict__ s, int64_t ld, int64_t e, bool terrain) {
  store_dyn(E, s, ld, e);
  if (terrain) store_terrain(E, s, ld, e);
  store_meta(E, s, ld, e);
}
CX_DEV void store_obs(const EnvRegs& E, float* __restrict__ obs, int64_t ld, int64_t e, bool rc, bool lc) {
  obs[0 * ld + e] = E.b[0].px;
  obs[1 * ld + e] = E.b[0].py;
  obs[2 * ld + e] = E.b[0].vx;
  obs[3 * ld + e] = E.b[0].vy;
  obs[4 * ld + e] = E.b[0].ang;
  obs[5 * ld + e] = E.b[0].w;
  obs[6 * ld + e] = rc ? 1.0f : 0.0f;
  obs[7 * ld + e] = lc ? 1.0f : 0.0f;
}

// Host-side analysis of the constant local shapes (see LunarDerived); lunar_kernel.cu
extern "C" void cx_lunar_derive(const CotixLunarParams* p, LunarDerived* out);
