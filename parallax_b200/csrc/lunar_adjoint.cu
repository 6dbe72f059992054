// lunar_adjoint.cu — K5: hand-written adjoint (reverse mode) of the LunarLander step, for
// jax.grad / jax.custom_vjp of a differentiable rollout (BASELINE config 5).
//
// What is differentiated: the composed step f of examples/test_viz.py:25-44 with engines,
//   ExplicitEulerPhysics.step   cotix/_physics_solvers.py:16-33
//   gravity + engines           examples/test_viz.py:27-31 (+ the action model of this engine)
//   RandomizedCollider.resolve  cotix/_colliders.py:74-351  -> resolve_collision
//                               cotix/_collision_resolution.py:52-151 on ContactInfo from
//                               polygon_vs_polygon cotix/_contacts.py:294-315 (EPA penetration
//                               vector cotix/_collisions.py:313-329, contact point :205-267)
//   LunarLander.step (joints)   cotix/_lunar_lander.py:145-218
// with respect to the 18 dynamic leaves of the state (3 bodies x px, py, vx, vy, angle, angular
// velocity) and the 2 action components.  This is the derivative jax.grad takes of the
// reference: every discrete decision (broadphase, GJK, the EPA expansion sequence, argmax /
// argmin / argsort selections, coin flips, jr.choice, the `pen . v_rel < 0` early-out, clip
// saturation) is held fixed, everything else is differentiated exactly:
//   * the penetration vector is closest_to_origin(a, b) of the final EPA edge, whose end points
//     are differences of one world vertex of each polygon (supports are argmax selections);
//   * the contact point is the mean of the vertices of one polygon inside the other and of the
//     edge-edge intersections;
//   * world vertices move with their body: d vertex = d p + perp(vertex - p) d angle.
//
// Strategy: recompute-and-reverse.  The tape of a rollout is just the state before every step
// (the forward kernels write it anyway); this kernel reloads state_t, re-runs the step with the
// SAME device functions the forward kernels use (so every decision is reproduced bit for bit —
// checked against state_{t+1} when the caller passes it), keeps the few intermediates the
// adjoint needs in registers / local memory and walks the step backwards.  One env per thread.
#include <cstdlib>

#include "cx_launch.cuh"
#include "cx_lunar_step.cuh"

namespace {

#ifndef CX_ADJ_TPB
#define CX_ADJ_TPB 64
#endif
constexpr int TPB = CX_ADJ_TPB;

struct G6 {  // cotangent of a BodyS
  float px, py, vx, vy, ang, w;
};
CX_DEV G6 g6_zero() {
  G6 g;
  g.px = g.py = g.vx = g.vy = g.ang = g.w = 0.0f;
  return g;
}

// polygon source: this thread's world polygons and terrain, from registers
struct RegPolys {
  const Poly& wL;
  const Poly& wr;
  const Poly& wl;
  const float* h;
  const float* kn;
  CX_DEV void body(int which, Poly& p) const { p = which == 0 ? wL : (which == 1 ? wr : wl); }
  CX_DEV void ground_edge(int k, const CotixLunarParams& P, float& x0, float& h0, float& x1,
                          float& h1) const {
    x0 = terrain_x(kn, P, k);
    x1 = terrain_x(kn, P, k + 1);
    h0 = sel8(h, k);
    h1 = sel8(h, k + 1);
  }
};
struct LocalContacts {
  const Contact* res;
  uint32_t cand;
  CX_DEV Contact get(int bit) const { return res[bit]; }
  CX_DEV bool live(int bit) const { return ((cand >> bit) & 1u) && !contact_isnan(res[bit]); }
};

// The resolve_collision calls RandomizedCollider.resolve makes this step, in scan order
// (cotix/_colliders.py:310-336): pair id, first body, second body (3 = the static ground).
struct Plan {
  int n;
  int pid[4], b1[4], b2[4];
};
CX_DEV void plan_add(Plan& pl, int pid, int b1, int b2) {
  pl.pid[pl.n] = pid;
  pl.b1[pl.n] = b1;
  pl.b2[pl.n] = b2;
  ++pl.n;
}
// same decisions as collide() + collide_apply() of cx_lunar_step.cuh, recorded instead of applied
static __device__ __noinline__ void collide_plan(const EnvRegs& E, const CotixLunarParams& P,
                                                 uint32_t cand, const LocalContacts& cs, Plan& pl) {
  pl.n = 0;
  uint32_t m[7];
  bool crash, rc, lc;
  if (!collide_masks(cand, cs, false, m, crash, rc, lc)) return;
  CoinCtx cc;
  make_coin_ctx(E, P, cc);
  int win[7];
  for (int c = 0; c < 6; ++c) {
    const CellDesc d = cell_desc(c);
    win[c] = cell_winner(cc, d.g, d.N1, d.N2, d.b_lo, d.b_hi, d.a0, d.nq, d.reps, m[c]);
  }
  win[6] = -1;
  if (win[1] >= 0 || win[3] >= 0) {
    const CellDesc d = cell_desc(6);
    win[6] = cell_winner(cc, d.g, d.N1, d.N2, d.b_lo, d.b_hi, d.a0, d.nq, d.reps, cs.live(10) ? 1u : 0u);
  }
  const bool v10 = win[0] >= 0, v20 = win[1] >= 0, v21 = win[3] >= 0, v22 = win[6] >= 0;
  const int w30 = win[2], w31 = win[4], w32 = win[5];
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
  if (v10) plan_add(pl, 0, 1, 0);
  if (j2 == 0 && v20) plan_add(pl, 1, 2, 0);
  if (j2 == 1 && v21) plan_add(pl, 9, 2, 1);
  if (j3 >= 0 && j3 <= 2) {
    const int w = j3 == 0 ? w30 : (j3 == 1 ? w31 : w32);
    if (w >= 0) plan_add(pl, (j3 == 0 ? 2 : (j3 == 1 ? 11 : 18)) + w, 3, j3);
  }
}

// ---- adjoint of apply_impulse (cotix/_collision_resolution.py:68-73) --------------------------
// forward: arm = point - p; v += imp / m; w += cross(arm, imp) / I.   g is the cotangent of the
// body AFTER the call on entry and BEFORE it on exit (v, w pass through; p collects -g_arm).
CX_DEV void apply_impulse_bwd(const MassP& m, V2 arm, V2 imp, G6& g, V2& g_imp, V2& g_point) {
  const float gt = g.w / m.inertia;
  g_imp.x += g.vx / m.mass + (-gt * arm.y);
  g_imp.y += g.vy / m.mass + gt * arm.x;
  const V2 g_arm = mk2(gt * imp.y, -gt * imp.x);
  g_point = g_point + g_arm;
  g.px -= g_arm.x;
  g.py -= g_arm.y;
}
// adjoint of velocity_at (cotix/_bodies.py:50-55): v_at = v + perp(point - p) * w
CX_DEV void velocity_at_bwd(V2 arm, float w, V2 g_vat, G6& g, V2& g_point) {
  g.vx += g_vat.x;
  g.vy += g_vat.y;
  g.w += dot2(g_vat, perp(arm));
  const V2 g_arm = mk2(w * g_vat.y, -w * g_vat.x);
  g_point = g_point + g_arm;
  g.px -= g_arm.x;
  g.py -= g_arm.y;
}

// ---- adjoint of resolve_collision (cotix/_collision_resolution.py:52-151) -----------------------
// b1, b2: the bodies BEFORE the call.  g1, g2: cotangents after -> before.  g_pen / g_cp are
// written (zero when the call changed nothing).
static __device__ __noinline__ void resolve_collision_bwd(const BodyS& b1, const MassP& m1,
                                                          const BodyS& b2, const MassP& m2, V2 pen,
                                                          V2 cp, float baumgarte, float baumgarte_dt,
                                                          G6& g1, G6& g2, V2& g_pen, V2& g_cp) {
  g_pen = mk2(0.0f, 0.0f);
  g_cp = mk2(0.0f, 0.0f);
  if (isnan2(cp)) return;
  // forward, same expressions as resolve_collision
  const V2 p1 = mk2(b1.px, b1.py), p2 = mk2(b2.px, b2.py);
  const V2 r1 = cp - p1, r2 = cp - p2;
  const V2 v1 = mk2(b1.vx, b1.vy) + perp(r1) * b1.w;
  const V2 v2 = mk2(b2.vx, b2.vy) + perp(r2) * b2.w;
  const V2 vrel = v2 - v1;
  if (dot2(pen, vrel) < 0) return;
  const float pn = norm2(pen);
  const V2 nd = pen / pn;
  const float vn = dot2(vrel, nd);
  const float el = jmin(m1.e, m2.e);
  const float ang = sumsq2(r1) / m1.inertia + sumsq2(r2) / m2.inertia;
  const float nim = -(1.0f + el) * vn - baumgarte * pn / baumgarte_dt;
  const float kden = 1.0f / m1.mass + 1.0f / m2.mass + ang;
  const float nimp = nim / kden;
  const float fr = (m1.mu + m2.mu) / 2.0f;
  const V2 vdrag = vrel + vn * nd;
  const float vdn = norm2(vdrag);
  const V2 vdu = vdrag / vdn;
  const float idrag0 = -vdn / kden;
  const float hi = nimp * fr;
  const float mx = jmax(idrag0, 0.0f);
  const float idrag = jmin(mx, hi);
  const V2 ivec = nimp * nd + idrag * vdu;

  // reverse
  V2 g_ivec = mk2(0.0f, 0.0f), g_r1 = mk2(0.0f, 0.0f), g_r2 = mk2(0.0f, 0.0f);
  {
    // apply_impulse(b2, ivec, cp); apply_impulse(b1, -ivec, cp): the arms are r2 / r1, their
    // cotangents go to cp and to the body position (inside apply_impulse_bwd)
    apply_impulse_bwd(m2, r2, ivec, g2, g_ivec, g_cp);
    V2 g_nivec = mk2(0.0f, 0.0f);
    apply_impulse_bwd(m1, r1, -ivec, g1, g_nivec, g_cp);
    g_ivec = g_ivec - g_nivec;
  }
  float g_nimp = dot2(g_ivec, nd);
  V2 g_nd = g_ivec * nimp;
  float g_kden = 0.0f;
  V2 g_vrel = mk2(0.0f, 0.0f);
  float g_vn = 0.0f;
  if (idrag != 0.0f && vdn > 0.0f) {  // idrag == 0: the drag term is the constant 0 here
    const float g_idrag = dot2(g_ivec, vdu);
    const V2 g_vdu = g_ivec * idrag;
    float g_idrag0 = 0.0f;
    if (mx <= hi) {
      if (idrag0 > 0.0f) g_idrag0 = g_idrag;
    } else {
      g_nimp += g_idrag * fr;
    }
    float g_vdn = -g_idrag0 / kden;
    g_kden += g_idrag0 * vdn / (kden * kden);
    V2 g_vdrag = g_vdu / vdn;
    g_vdn += -dot2(g_vdu, vdrag) / (vdn * vdn);
    g_vdrag = g_vdrag + vdrag * (g_vdn / vdn);
    g_vrel = g_vrel + g_vdrag;
    g_vn += dot2(g_vdrag, nd);
    g_nd = g_nd + g_vdrag * vn;
  }
  const float g_nim = g_nimp / kden;
  g_kden += -g_nimp * nim / (kden * kden);
  g_vn += -(1.0f + el) * g_nim;
  float g_pn = -(baumgarte / baumgarte_dt) * g_nim;
  g_r1 = g_r1 + r1 * (2.0f * (g_kden / m1.inertia));
  g_r2 = g_r2 + r2 * (2.0f * (g_kden / m2.inertia));
  g_vrel = g_vrel + nd * g_vn;
  g_nd = g_nd + vrel * g_vn;
  g_pen = g_nd / pn;
  g_pn += -dot2(g_nd, pen) / (pn * pn);
  g_pen = g_pen + pen * (g_pn / pn);
  // vrel = v2 - v1, v_i = v_i_lin + perp(r_i) * w_i
  {
    V2 gp = mk2(0.0f, 0.0f);
    velocity_at_bwd(r2, b2.w, g_vrel, g2, gp);
    g_cp = g_cp + gp;
    gp = mk2(0.0f, 0.0f);
    velocity_at_bwd(r1, b1.w, -g_vrel, g1, gp);
    g_cp = g_cp + gp;
  }
  g_cp = g_cp + g_r1 + g_r2;
  g1.px -= g_r1.x;
  g1.py -= g_r1.y;
  g2.px -= g_r2.x;
  g2.py -= g_r2.y;
}

// ---- adjoint of joint_constraint (cotix/_lunar_lander.py:176-205) -------------------------------
// b1, b2: the bodies BEFORE the call (positions never change inside a step after integration).
static __device__ __noinline__ void joint_constraint_bwd(const BodyS& b1, const MassP& m1, V2 c1,
                                                         const BodyS& b2, const MassP& m2, V2 c2,
                                                         const CotixLunarParams& P, G6& g1, G6& g2,
                                                         V2& g_c1, V2& g_c2) {
  const V2 a1 = c1 - mk2(b1.px, b1.py), a2 = c2 - mk2(b2.px, b2.py);
  const V2 dp = c1 - c2;
  const V2 dv = (mk2(b1.vx, b1.vy) + perp(a1) * b1.w) - (mk2(b2.vx, b2.vy) + perp(a2) * b2.w);
  const float n = norm2(dv);
  const float s = (n + P.spring_kv0) * P.spring_kv;
  const V2 imp = dp * P.spring_kp + dv * (n + P.spring_kv0) * P.spring_kv;
  g_c1 = mk2(0.0f, 0.0f);
  g_c2 = mk2(0.0f, 0.0f);
  V2 g_imp = mk2(0.0f, 0.0f), g_nimp = mk2(0.0f, 0.0f);
  apply_impulse_bwd(m2, a2, imp, g2, g_imp, g_c2);
  apply_impulse_bwd(m1, a1, -imp, g1, g_nimp, g_c1);
  g_imp = g_imp - g_nimp;
  const V2 g_dp = g_imp * P.spring_kp;
  V2 g_dv = g_imp * s;
  const float g_n = dot2(g_imp, dv) * P.spring_kv;
  if (n > 0.0f) g_dv = g_dv + dv * (g_n / n);
  velocity_at_bwd(a1, b1.w, g_dv, g1, g_c1);
  velocity_at_bwd(a2, b2.w, -g_dv, g2, g_c2);
  g_c1 = g_c1 + g_dp;
  g_c2 = g_c2 - g_dp;
}

// a world point rigidly attached to a body: d point = d p + perp(point - p) d angle
CX_DEV void attached_point_bwd(V2 point, const BodyS& b, V2 g_point, G6& g) {
  g.px += g_point.x;
  g.py += g_point.y;
  g.ang += dot2(g_point, perp(point - mk2(b.px, b.py)));
}

// ---- adjoint of the narrowphase outputs of one polygon pair -------------------------------------
struct PolyGrad {
  float x[CX_MAXV], y[CX_MAXV];
};
// adjoint of closest_to_origin(a, b) (get_closest_point_on_edge_to_point, cotix/_collisions.py:137-154)
CX_DEV void closest_to_origin_bwd(V2 a, V2 b, V2 g_out, V2& g_a, V2& g_b) {
  const V2 ab = a - b;
  const float length = sumsq2(ab);
  if (length == 0.0f) {
    g_a = -g_out;
    g_b = mk2(0.0f, 0.0f);
    return;
  }
  const float num = dot2(mk2(0.0f, 0.0f) - b, ab);
  const float t0 = num / length;
  const float t = jclip(t0, 0.0f, 1.0f);
  const V2 g_proj = -g_out;
  g_b = g_proj;
  V2 g_ab = g_proj * t;
  if (t0 > 0.0f && t0 < 1.0f) {
    const float g_t0 = dot2(g_proj, ab);
    const float g_num = g_t0 / length;
    const float g_len = -g_t0 * num / (length * length);
    g_b = g_b - ab * g_num;
    g_ab = g_ab - b * g_num + ab * (2.0f * g_len);
  }
  g_a = g_ab;
  g_b = g_b - g_ab;
}
// m = sa - sb with sa a vertex of A and sb a vertex of B (minkowski_diff of two argmax supports):
// route g_m to the vertices it came from (first exact match)
CX_DEV void mdiff_point_bwd(const Poly& A, const Poly& B, V2 m, V2 g_m, PolyGrad& gA, PolyGrad& gB) {
  for (int i = 0; i < A.n; ++i)
    for (int k = 0; k < B.n; ++k)
      if (A.x[i] - B.x[k] == m.x && A.y[i] - B.y[k] == m.y) {
        gA.x[i] += g_m.x;
        gA.y[i] += g_m.y;
        gB.x[k] -= g_m.x;
        gB.y[k] -= g_m.y;
        return;
      }
}
// adjoint of _contact_from_edges (cotix/_contacts.py:205-267)
CX_DEV void contact_from_edges_bwd(const Poly& A, const Poly& B, V2 g_cp, PolyGrad& gA, PolyGrad& gB) {
  EdgeList ea, eb;
  edges_of(A, ea);
  edges_of(B, eb);
  // pass 1: the count, exactly as the forward does
  float n = 0.0f;
  for (int k = 0; k < ea.n; ++k)
    if (poly_contains(eb, mk2(ea.ax[k], ea.ay[k]))) n = n + 1.0f;
  for (int k = 0; k < eb.n; ++k)
    if (poly_contains(ea, mk2(eb.ax[k], eb.ay[k]))) n = n + 1.0f;
  for (int ib = 0; ib < eb.n; ++ib) {
    const V2 q = mk2(eb.ax[ib], eb.ay[ib]);
    const V2 s = mk2(eb.bx[ib], eb.by[ib]) - q;
    for (int ia = 0; ia < ea.n; ++ia) {
      const V2 p = mk2(ea.ax[ia], ea.ay[ia]);
      const V2 r = mk2(ea.bx[ia], ea.by[ia]) - p;
      const float c = r.x * s.y - s.x * r.y;
      const V2 qp = q - p;
      const float t = (qp.x * s.y - s.x * qp.y) / c;
      const float u = (qp.x * r.y - r.x * qp.y) / c;
      if ((c != 0.0f) && (t >= 0.0f) && (t <= 1.0f) && (u >= 0.0f) && (u <= 1.0f) && !isnan2(p + r * t))
        n = n + 1.0f;
    }
  }
  if (!(n > 0.0f)) return;
  const V2 g = g_cp / n;
  // pass 2: every term of the mean gets g
  for (int k = 0; k < ea.n; ++k)
    if (poly_contains(eb, mk2(ea.ax[k], ea.ay[k]))) {
      gA.x[k] += g.x;
      gA.y[k] += g.y;
    }
  for (int k = 0; k < eb.n; ++k)
    if (poly_contains(ea, mk2(eb.ax[k], eb.ay[k]))) {
      gB.x[k] += g.x;
      gB.y[k] += g.y;
    }
  for (int ib = 0; ib < eb.n; ++ib) {
    const int ibp = ib == 0 ? eb.n - 1 : ib - 1;  // edge ib = (v_ib, v_{ib-1})
    const V2 q = mk2(eb.ax[ib], eb.ay[ib]);
    const V2 s = mk2(eb.bx[ib], eb.by[ib]) - q;
    for (int ia = 0; ia < ea.n; ++ia) {
      const int iap = ia == 0 ? ea.n - 1 : ia - 1;
      const V2 p = mk2(ea.ax[ia], ea.ay[ia]);
      const V2 r = mk2(ea.bx[ia], ea.by[ia]) - p;
      const float c = r.x * s.y - s.x * r.y;
      const V2 qp = q - p;
      const float num = qp.x * s.y - s.x * qp.y;
      const float t = num / c;
      const float u = (qp.x * r.y - r.x * qp.y) / c;
      if (!((c != 0.0f) && (t >= 0.0f) && (t <= 1.0f) && (u >= 0.0f) && (u <= 1.0f) && !isnan2(p + r * t)))
        continue;
      // x = p + r * t
      V2 g_p = g;
      V2 g_r = g * t;
      const float g_t = dot2(g, r);
      const float g_num = g_t / c;
      const float g_c = -g_t * num / (c * c);
      const V2 g_qp = mk2(g_num * s.y, -g_num * s.x);
      V2 g_s = mk2(-g_num * qp.y, g_num * qp.x);
      g_r = g_r + mk2(g_c * s.y, -g_c * s.x);
      g_s = g_s + mk2(-g_c * r.y, g_c * r.x);
      // qp = q - p; r = p_prev - p; s = q_prev - q
      V2 g_q = g_qp - g_s;
      g_p = g_p - g_qp - g_r;
      gA.x[ia] += g_p.x;
      gA.y[ia] += g_p.y;
      gA.x[iap] += g_r.x;
      gA.y[iap] += g_r.y;
      gB.x[ib] += g_q.x;
      gB.y[ib] += g_q.y;
      gB.x[ibp] += g_s.x;
      gB.y[ibp] += g_s.y;
    }
  }
}

// g_pen, g_cp of candidate pair `pid` -> cotangents of the poses of the dynamic bodies involved.
// (ba, bb) is the final EPA edge of the pair (the penetration vector is closest_to_origin(ba, bb)).
static __device__ __noinline__ void pair_contact_bwd(const RegPolys& src, int pid, const EnvRegs& E,
                                                     const CotixLunarParams& P, V2 ba, V2 bb, V2 g_pen,
                                                     V2 g_cp, G6 g[4]) {
  Poly A, B;
  make_item_polys(src, pid, P, A, B);
  int wa, wb, gk;
  pair_bodies(pid, wa, wb, gk);
  PolyGrad gA, gB;
#pragma unroll
  for (int k = 0; k < CX_MAXV; ++k) gA.x[k] = gA.y[k] = gB.x[k] = gB.y[k] = 0.0f;
  V2 g_a, g_b;
  closest_to_origin_bwd(ba, bb, g_pen, g_a, g_b);
  mdiff_point_bwd(A, B, ba, g_a, gA, gB);
  mdiff_point_bwd(A, B, bb, g_b, gA, gB);
  contact_from_edges_bwd(A, B, g_cp, gA, gB);
  if (wa >= 0)
    for (int k = 0; k < A.n; ++k)
      attached_point_bwd(mk2(A.x[k], A.y[k]), E.b[wa], mk2(gA.x[k], gA.y[k]), g[wa]);
  for (int k = 0; k < B.n; ++k)
    attached_point_bwd(mk2(B.x[k], B.y[k]), E.b[wb], mk2(gB.x[k], gB.y[k]), g[wb]);
}

// polygon_vs_polygon (cotix/_contacts.py:294-315) of candidate pair `pid`, as poly_pair computes it,
// also returning the final EPA edge
static __device__ __noinline__ Contact eval_pair_edge(const RegPolys& src, int pid,
                                                      const CotixLunarParams& P, V2& ea, V2& eb) {
  Poly A, B;
  make_item_polys(src, pid, P, A, B);
  Contact c = contact_none();
  ea = eb = mk2(0.0f, 0.0f);
  V2 s0, s1, s2;
  bool gcap, ecap;
  if (!gjk(A, B, mk2(P.init_dir[0], P.init_dir[1]), P.gjk_max_iters, s0, s1, s2, gcap)) return c;
  int iters = A.n + B.n + 1;
  if (iters > 13) iters = 13;
  epa_edge<16>(A, B, s0, s1, s2, iters, ecap, ea, eb);
  c.pen = closest_to_origin(ea, eb);
  c.cp = contact_from_edges(A, B);
  return c;
}

// What the collider decided in one step of one env: the resolve_collision calls in scan order with
// their ContactInfo and final EPA edge.  132 bytes per env-step on the decision tape:
// word 0 = n | pid0 << 3 | pid1 << 8 | pid2 << 13 | pid3 << 18, then 8 floats per call.
struct Decisions {
  Plan pl;
  V2 pen[4], cp[4], ea[4], eb[4];
};
constexpr int DT_WORDS = COTIX_ADJ_TAPE_WORDS;

// replay of the irregular part of the step: narrowphase of the candidates + the collider's choices
static __device__ __noinline__ void decide(const EnvRegs& E, uint32_t cand, const RegPolys& src,
                                           const CotixLunarParams& P, Decisions& d) {
  Contact res[CX_NPAIR];
  V2 ea[CX_NPAIR], eb[CX_NPAIR];
  uint32_t todo = cand;
  while (todo) {
    const int bit = __ffs(todo) - 1;
    todo &= todo - 1u;
    res[bit] = eval_pair_edge(src, bit, P, ea[bit], eb[bit]);
  }
  const LocalContacts cs{res, cand};
  collide_plan(E, P, cand, cs, d.pl);
  for (int i = 0; i < d.pl.n; ++i) {
    const int pid = d.pl.pid[i];
    d.pen[i] = res[pid].pen;
    d.cp[i] = res[pid].cp;
    d.ea[i] = ea[pid];
    d.eb[i] = eb[pid];
  }
}
CX_DEV void store_decisions(const Decisions& d, float* __restrict__ dt, int64_t ld, int64_t e) {
  uint32_t w = (uint32_t)d.pl.n;
  for (int i = 0; i < d.pl.n; ++i) w |= (uint32_t)d.pl.pid[i] << (3 + 5 * i);
  dt[e] = __uint_as_float(w);
  for (int i = 0; i < d.pl.n; ++i) {
    float* q = dt + (1 + 8 * i) * ld + e;
    q[0 * ld] = d.pen[i].x; q[1 * ld] = d.pen[i].y; q[2 * ld] = d.cp[i].x; q[3 * ld] = d.cp[i].y;
    q[4 * ld] = d.ea[i].x;  q[5 * ld] = d.ea[i].y;  q[6 * ld] = d.eb[i].x; q[7 * ld] = d.eb[i].y;
  }
}
CX_DEV void load_decisions(Decisions& d, const float* __restrict__ dt, int64_t ld, int64_t e) {
  const uint32_t w = __float_as_uint(dt[e]);
  d.pl.n = (int)(w & 7u);
  if (d.pl.n > 4) d.pl.n = 4;
  for (int i = 0; i < d.pl.n; ++i) {
    const int pid = (int)((w >> (3 + 5 * i)) & 31u);
    int wa, wb, gk;
    pair_bodies(pid, wa, wb, gk);
    d.pl.pid[i] = pid;
    d.pl.b1[i] = wa >= 0 ? wa : 3;
    d.pl.b2[i] = wb;
    const float* q = dt + (1 + 8 * i) * ld + e;
    d.pen[i] = mk2(q[0 * ld], q[1 * ld]);
    d.cp[i] = mk2(q[2 * ld], q[3 * ld]);
    d.ea[i] = mk2(q[4 * ld], q[5 * ld]);
    d.eb[i] = mk2(q[6 * ld], q[7 * ld]);
  }
}

CX_DEV bool same_bits(float a, float b) { return (a == b) || (a != a && b != b); }

// ---- one step backwards, given the collider's decisions -------------------------------------------
// E, cx, src: the env after integrate_and_cull of the replay.  g: cotangent of the next state on
// entry, of this state on exit.  snext (or NULL): the next state the forward kernel wrote.
static __device__ __noinline__ void step_bwd_core(const EnvRegs& E, const StepCtx& cx, const RegPolys& src,
                                                  float am, float as, const Decisions& d,
                                                  const CotixLunarParams& P, const float* __restrict__ snext,
                                                  int64_t ld, int64_t e, G6 g[4], float& g_am, float& g_as,
                                                  uint32_t& flags) {
  const Plan& pl = d.pl;
  MassP mp[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    mp[i].mass = P.body_props[i][0];
    mp[i].inertia = P.body_props[i][1];
    mp[i].e = P.body_props[i][2];
    mp[i].mu = P.body_props[i][3];
    mp[i].inv_mass = 1.0f / P.body_props[i][0];
    mp[i].dmass = divc_plain(P.body_props[i][0]);
    mp[i].dinertia = divc_plain(P.body_props[i][1]);
  }
  const bool poisoned = (E.status & COTIX_ST_GROUND_NAN) != 0u;
  BodyS bd[4];
  bd[0] = E.b[0];
  bd[1] = E.b[1];
  bd[2] = E.b[2];
  bd[3].px = bd[3].py = bd[3].vx = bd[3].vy = bd[3].ang = bd[3].w = 0.0f;
  // forward: the resolve_collision calls in scan order, then the joints (finish_step)
  BodyS pre1[4], pre2[4];
  for (int i = 0; i < pl.n; ++i) {
    pre1[i] = bd[pl.b1[i]];
    pre2[i] = bd[pl.b2[i]];
    resolve_collision(bd[pl.b1[i]], mp[pl.b1[i]], bd[pl.b2[i]], mp[pl.b2[i]], d.pen[i], d.cp[i], P.baumgarte,
                      P.baumgarte_dt);
  }
  V2 c1[4], c2[4];
  {
    const V2 p0 = mk2(bd[0].px, bd[0].py);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      c1[k] = rot_cs(mk2(P.joint_local[k][0], P.joint_local[k][1]), cx.cs[0], cx.sn[0]) + p0;
    const V2 j2l = mk2(P.leg_joint2_local[0], P.leg_joint2_local[1]);
    c2[0] = mk2(bd[2].px, bd[2].py);
    c2[1] = c2[0] + rot_cs(j2l, cx.cs[2], cx.sn[2]);
    c2[2] = mk2(bd[1].px, bd[1].py);
    c2[3] = c2[2] + rot_cs(j2l, cx.cs[1], cx.sn[1]);
  }
  BodyS jpre1[4], jpre2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int leg = k < 2 ? 2 : 1;
    jpre1[k] = bd[0];
    jpre2[k] = bd[leg];
    joint_constraint(bd[0], mp[0], c1[k], bd[leg], mp[leg], c2[k], P);
  }
  bd[1].w = bd[1].w * P.leg_damping;
  bd[2].w = bd[2].w * P.leg_damping;
  if (snext) {  // the replay must land on the state the forward kernel wrote
    bool same = true;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      same = same && same_bits(bd[i].px, snext[(i * 6 + 0) * ld + e]) &&
             same_bits(bd[i].py, snext[(i * 6 + 1) * ld + e]) &&
             same_bits(bd[i].vx, snext[(i * 6 + 2) * ld + e]) &&
             same_bits(bd[i].vy, snext[(i * 6 + 3) * ld + e]) &&
             same_bits(bd[i].ang, snext[(i * 6 + 4) * ld + e]) &&
             same_bits(bd[i].w, snext[(i * 6 + 5) * ld + e]);
    }
    if (!same) flags |= COTIX_ADJ_REPLAY_MISMATCH;
  }
  flags |= (uint32_t)pl.n << COTIX_ADJ_CONTACT_SHIFT;
  bool finite = !poisoned;
#pragma unroll
  for (int i = 0; i < 3; ++i)
    finite = finite && (fabsf(bd[i].px) <= 3.0e38f) && (fabsf(bd[i].py) <= 3.0e38f) &&
             (fabsf(bd[i].vx) <= 3.0e38f) && (fabsf(bd[i].vy) <= 3.0e38f) &&
             (fabsf(bd[i].ang) <= 3.0e38f) && (fabsf(bd[i].w) <= 3.0e38f);
  g_am = g_as = 0.0f;
  if (!finite) {
    // a non-finite step has no derivative: the cotangent is dropped (and flagged)
    flags |= COTIX_ADJ_NONFINITE;
    g[0] = g[1] = g[2] = g6_zero();
    return;
  }
  // reverse
  g[1].w = g[1].w * P.leg_damping;
  g[2].w = g[2].w * P.leg_damping;
#pragma unroll
  for (int k = 3; k >= 0; --k) {
    const int leg = k < 2 ? 2 : 1;
    V2 gc1, gc2;
    joint_constraint_bwd(jpre1[k], mp[0], c1[k], jpre2[k], mp[leg], c2[k], P, g[0], g[leg], gc1, gc2);
    attached_point_bwd(c1[k], bd[0], gc1, g[0]);
    attached_point_bwd(c2[k], bd[leg], gc2, g[leg]);
  }
  for (int i = pl.n - 1; i >= 0; --i) {
    V2 g_pen, g_cp;
    g[3] = g6_zero();  // the ground is static: its cotangent is discarded
    resolve_collision_bwd(pre1[i], mp[pl.b1[i]], pre2[i], mp[pl.b2[i]], d.pen[i], d.cp[i], P.baumgarte,
                          P.baumgarte_dt, g[pl.b1[i]], g[pl.b2[i]], g_pen, g_cp);
    if (!isnan2(d.cp[i])) pair_contact_bwd(src, pl.pid[i], E, P, d.ea[i], d.eb[i], g_pen, g_cp, g);
  }
  // engines: v0 += (-sn, cs) * dvm + (cs, sn) * dvs; w0 -= side_dw * thr_side
  {
    const float cs0 = cx.cs[0], sn0 = cx.sn[0];
    float g_tm = 0.0f, g_ts = 0.0f;
    if (cx.thr_main != 0.0f) {
      const float dvm = P.main_engine_dv * cx.thr_main;
      g_tm = P.main_engine_dv * (-sn0 * g[0].vx + cs0 * g[0].vy);
      g[0].ang += (-cs0 * g[0].vx - sn0 * g[0].vy) * dvm;
    }
    if (cx.thr_side != 0.0f) {
      const float dvs = P.side_engine_dv * cx.thr_side;
      g_ts = P.side_engine_dv * (cs0 * g[0].vx + sn0 * g[0].vy) - P.side_engine_dw * g[0].w;
      g[0].ang += (-sn0 * g[0].vx + cs0 * g[0].vy) * dvs;
    }
    if (am > 0.0f && am < 1.0f) g_am = g_tm;
    if (as > -1.0f && as < 1.0f) g_as = g_ts;
  }
  // ExplicitEulerPhysics.step: p' = p + v dt, angle' = angle + w dt (gravity is a constant)
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    g[i].vx += g[i].px * P.dt;
    g[i].vy += g[i].py * P.dt;
    g[i].w += g[i].ang * P.dt;
  }
}

CX_DEV void load_cotangent(G6 g[4], const float* g_out, int64_t ld, int64_t e) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    g[i].px = g_out[(i * 6 + 0) * ld + e];
    g[i].py = g_out[(i * 6 + 1) * ld + e];
    g[i].vx = g_out[(i * 6 + 2) * ld + e];
    g[i].vy = g_out[(i * 6 + 3) * ld + e];
    g[i].ang = g_out[(i * 6 + 4) * ld + e];
    g[i].w = g_out[(i * 6 + 5) * ld + e];
  }
  g[3] = g6_zero();
}
CX_DEV void store_cotangent(const G6 g[4], float* g_in, int64_t ld, int64_t e) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    g_in[(i * 6 + 0) * ld + e] = g[i].px;
    g_in[(i * 6 + 1) * ld + e] = g[i].py;
    g_in[(i * 6 + 2) * ld + e] = g[i].vx;
    g_in[(i * 6 + 3) * ld + e] = g[i].vy;
    g_in[(i * 6 + 4) * ld + e] = g[i].ang;
    g_in[(i * 6 + 5) * ld + e] = g[i].w;
  }
}

// Transposed Jacobians of the env-steps with contacts (see lunar_jacobian_kernel)
constexpr int JAC_ROWS = COTIX_LUNAR_NDYN + 2, JAC_FLOATS = JAC_ROWS * COTIX_LUNAR_NDYN;  // 20 x 18
constexpr uint32_t DT_HAS_MATRIX = 1u << 23;  // bit of decision word 0; the slot is in word 32
struct JacWs {
  uint32_t* ctr;    // [1] env-steps with contacts in this group
  uint32_t* hlist;  // [cap] item index (step_local * B + env) of slot s
  float* M;         // [cap][18][20], column-major: (g_state; g_action) = M g_next, M[k][j] = d out_j / d g_next_k
  uint32_t cap;
};

// One step backwards for every env.  dtape == NULL: the collider's decisions are replayed here
// (narrowphase of all candidates + coins); else they are read from the decision tape of this step,
// and an env-step whose transposed Jacobian was formed in advance is a matrix-vector product.
__global__ void __launch_bounds__(TPB)
lunar_step_bwd_kernel(int64_t B, const float* __restrict__ sin_, const float* __restrict__ snext,
                      int64_t ld, const float* __restrict__ action, const float* g_out, float* g_in,
                      float* __restrict__ g_action, uint32_t* __restrict__ flags_out,
                      const float* __restrict__ dtape, const float* __restrict__ jacM,
                      const __grid_constant__ CotixLunarParams P, const LunarDerived D) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  const bool in = e < B;
  bool heavy = false;
  uint32_t slot = 0u;
  if (dtape && jacM && in && (__float_as_uint(dtape[e]) & DT_HAS_MATRIX)) {
    heavy = true;
    slot = __float_as_uint(dtape[32 * ld + e]);
    // the warp takes its matrices one after the other below: have them all on their way to L2 now
    const char* Mb = reinterpret_cast<const char*>(jacM + (size_t)slot * JAC_FLOATS);
#pragma unroll
    for (int l = 0; l < JAC_FLOATS * 4; l += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(Mb + l));
  }
  G6 g[4];
  g[0] = g[1] = g[2] = g[3] = g6_zero();
  if (in) load_cotangent(g, g_out, ld, e);
  if (dtape && jacM) {
    // The matrix-vector products of the warp's env-steps with a Jacobian, one after the other with
    // the whole warp on each: lane j < 20 owns output row j and reads column-major M[k][j] (18
    // independent, coalesced loads; the next matrix is loaded while this one is used) while the
    // cotangent comes round by shuffle.  A lane working through its own 360 entries keeps only a few
    // loads in flight: ~20 rounds of memory latency per launch, which was most of the sweep's time.
    const float gv[COTIX_LUNAR_NDYN] = {g[0].px, g[0].py, g[0].vx, g[0].vy, g[0].ang, g[0].w,
                                        g[1].px, g[1].py, g[1].vx, g[1].vy, g[1].ang, g[1].w,
                                        g[2].px, g[2].py, g[2].vx, g[2].vy, g[2].ang, g[2].w};
    const int j = (int)(threadIdx.x & 31u);
    uint32_t hm = __ballot_sync(0xFFFFFFFFu, heavy);
    float m[COTIX_LUNAR_NDYN];
    auto fetch = [&](int src_lane) {
      const uint32_t sl = __shfl_sync(0xFFFFFFFFu, slot, src_lane);
      const float* M = jacM + (size_t)sl * JAC_FLOATS;
#pragma unroll
      for (int k = 0; k < COTIX_LUNAR_NDYN; ++k) m[k] = (j < JAC_ROWS) ? M[k * JAC_ROWS + j] : 0.0f;
    };
    if (hm) fetch(__ffs(hm) - 1);
    while (hm) {
      const int src = __ffs(hm) - 1;
      hm &= hm - 1u;
      float mc[COTIX_LUNAR_NDYN];
#pragma unroll
      for (int k = 0; k < COTIX_LUNAR_NDYN; ++k) mc[k] = m[k];
      if (hm) fetch(__ffs(hm) - 1);
      const int64_t es = e - j + src;
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < COTIX_LUNAR_NDYN; ++k) acc += mc[k] * __shfl_sync(0xFFFFFFFFu, gv[k], src);
      if (j < COTIX_LUNAR_NDYN) g_in[j * ld + es] = acc;
      else if (j < JAC_ROWS && g_action) g_action[(j - COTIX_LUNAR_NDYN) * ld + es] = acc;
    }
  }
  if (!in || heavy) return;  // a heavy env-step's flags were set when its matrix was formed
  EnvRegs E;
  load_dyn(E, sin_, ld, e);
  E.status = __float_as_uint(sin_[COTIX_F_STATUS * ld + e]);
  float am = 0.0f, as = 0.0f;
  if (action) {
    am = action[e];
    as = action[ld + e];
  }
  StepCtx cx;
  Poly wL, wr, wl;
  Decisions d;
  // a taped env-step without a resolved contact (most of them) needs neither the terrain, the world
  // polygons nor the broadphase: Euler, trig and engines only
  const bool contact_free = dtape && (__float_as_uint(dtape[e]) & 7u) == 0u;
  if (contact_free) {
    integrate_only(E, am, as, P, cx);
    cx.cand = 0u;
    cx.base = -1;
    d.pl.n = 0;
  } else {
    load_terrain(E, sin_, ld, e);
    load_meta(E, sin_, ld, e);
    integrate_and_cull<true>(E, am, as, P, D, cx, wL, wr, wl);
  }
  const RegPolys src{wL, wr, wl, E.h, E.kn};
  if (!contact_free) {
    if (dtape) {
      load_decisions(d, dtape, ld, e);
      // integrate_and_cull orders the world polygons only when it found candidates; a taped contact
      // implies it did
    } else {
      decide(E, cx.cand, src, P, d);
    }
  }
  uint32_t flags = 0u;
  float g_am, g_as;
  step_bwd_core(E, cx, src, am, as, d, P, snext, ld, e, g, g_am, g_as, flags);
  store_cotangent(g, g_in, ld, e);
  if (g_action) {
    g_action[e] = g_am;
    g_action[ld + e] = g_as;
  }
  // atomic: the Jacobian kernel of an earlier run of steps may be setting flags concurrently
  if (flags_out && flags) atomicOr(&flags_out[e], flags);
}

// Steps s_hi-1 .. s_lo backwards in ONE launch for a small batch: the reverse sweep is sequential in time,
// and launched step by step a batch of a few thousand envs spends its time in launch gaps and in warps that
// work through the Jacobians of eight heavy env-steps one after the other.  Envs never interact, so a warp
// keeps the cotangents of its `epw` envs (lanes 0 .. epw-1; few, so that few matrices queue up per step) in
// registers for the whole run of steps.  Needs the decision tape and the Jacobians of the run (taped mode).
// The arithmetic is that of lunar_step_bwd_kernel, operation for operation: identical gradients.
#ifndef CX_SWEEP_MINB
#define CX_SWEEP_MINB 1
#endif
__global__ void __launch_bounds__(TPB, CX_SWEEP_MINB)
lunar_sweep_bwd_kernel(int64_t B, int s_lo, int s_hi, const float* __restrict__ tape, int64_t ld,
                       const float* __restrict__ actions, float* g_io, float* __restrict__ g_actions,
                       uint32_t* __restrict__ flags_out, const float* __restrict__ dtape0,
                       const float* __restrict__ jacM, const __grid_constant__ CotixLunarParams P, const LunarDerived D,
                       int epw) {
  __shared__ float s_row[TPB / 32][JAC_ROWS][32];  // matrix-vector results: [output row][env lane]
  const int j = (int)(threadIdx.x & 31u), wib = (int)(threadIdx.x >> 5);
  const int64_t e = ((int64_t)blockIdx.x * (TPB / 32) + wib) * epw + j;
  const bool in = j < epw && e < B;
  G6 g[4];
  g[0] = g[1] = g[2] = g[3] = g6_zero();
  if (in) load_cotangent(g, g_io, ld, e);
  uint32_t flags = 0u;
  const int64_t slot_f = (int64_t)COTIX_LUNAR_NF * ld, dslot = (int64_t)DT_WORDS * ld;
  for (int s = s_hi - 1; s >= s_lo; --s) {
    const float* sin_ = tape + (int64_t)s * slot_f;
    const float* snext = sin_ + slot_f;
    const float* dtape = dtape0 + (int64_t)s * dslot;
    const float* action = actions ? actions + (int64_t)s * 2 * ld : nullptr;
    float* g_action = g_actions ? g_actions + (int64_t)s * 2 * ld : nullptr;
    bool heavy = false;
    uint32_t slot = 0u;
    uint32_t w0 = 0u;
    if (in) {
      w0 = __float_as_uint(dtape[e]);
      if (w0 & DT_HAS_MATRIX) {
        heavy = true;
        slot = __float_as_uint(dtape[32 * ld + e]);
      }
      if (s > s_lo) {  // next iteration's matrix on its way to L2 while this step is worked on
        const float* dn = dtape - dslot;
        if (__float_as_uint(dn[e]) & DT_HAS_MATRIX) {
          const char* Mb = reinterpret_cast<const char*>(jacM + (size_t)__float_as_uint(dn[32 * ld + e]) * JAC_FLOATS);
#pragma unroll
          for (int l = 0; l < JAC_FLOATS * 4; l += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(Mb + l));
        }
      }
    }
    uint32_t hm = __ballot_sync(0xFFFFFFFFu, heavy);
    if (hm) {
      const float gv[COTIX_LUNAR_NDYN] = {g[0].px, g[0].py, g[0].vx, g[0].vy, g[0].ang, g[0].w,
                                          g[1].px, g[1].py, g[1].vx, g[1].vy, g[1].ang, g[1].w,
                                          g[2].px, g[2].py, g[2].vx, g[2].vy, g[2].ang, g[2].w};
      float m[COTIX_LUNAR_NDYN];
      auto fetch = [&](int src_lane) {
        const uint32_t sl = __shfl_sync(0xFFFFFFFFu, slot, src_lane);
        const float* M = jacM + (size_t)sl * JAC_FLOATS;
#pragma unroll
        for (int k = 0; k < COTIX_LUNAR_NDYN; ++k) m[k] = (j < JAC_ROWS) ? M[k * JAC_ROWS + j] : 0.0f;
      };
      fetch(__ffs(hm) - 1);
      while (hm) {
        const int src = __ffs(hm) - 1;
        hm &= hm - 1u;
        float mc[COTIX_LUNAR_NDYN];
#pragma unroll
        for (int k = 0; k < COTIX_LUNAR_NDYN; ++k) mc[k] = m[k];
        if (hm) fetch(__ffs(hm) - 1);
        float acc = 0.0f;
#pragma unroll
        for (int k = 0; k < COTIX_LUNAR_NDYN; ++k) acc += mc[k] * __shfl_sync(0xFFFFFFFFu, gv[k], src);
        if (j < JAC_ROWS) s_row[wib][j][src] = acc;
      }
      __syncwarp();
      if (heavy) {
        g[0].px = s_row[wib][0][j]; g[0].py = s_row[wib][1][j]; g[0].vx = s_row[wib][2][j];
        g[0].vy = s_row[wib][3][j]; g[0].ang = s_row[wib][4][j]; g[0].w = s_row[wib][5][j];
        g[1].px = s_row[wib][6][j]; g[1].py = s_row[wib][7][j]; g[1].vx = s_row[wib][8][j];
        g[1].vy = s_row[wib][9][j]; g[1].ang = s_row[wib][10][j]; g[1].w = s_row[wib][11][j];
        g[2].px = s_row[wib][12][j]; g[2].py = s_row[wib][13][j]; g[2].vx = s_row[wib][14][j];
        g[2].vy = s_row[wib][15][j]; g[2].ang = s_row[wib][16][j]; g[2].w = s_row[wib][17][j];
        if (g_action) {
          g_action[e] = s_row[wib][18][j];
          g_action[ld + e] = s_row[wib][19][j];
        }
      }
      __syncwarp();
    }
    if (in && !heavy) {
      EnvRegs E;
      load_dyn(E, sin_, ld, e);
      E.status = __float_as_uint(sin_[COTIX_F_STATUS * ld + e]);
      float am = 0.0f, as = 0.0f;
      if (action) {
        am = action[e];
        as = action[ld + e];
      }
      StepCtx cx;
      Poly wL, wr, wl;
      Decisions d;
      const bool contact_free = (w0 & 7u) == 0u;
      if (contact_free) {
        integrate_only(E, am, as, P, cx);
        cx.cand = 0u;
        cx.base = -1;
        d.pl.n = 0;
      } else {
        load_terrain(E, sin_, ld, e);
        load_meta(E, sin_, ld, e);
        integrate_and_cull<true>(E, am, as, P, D, cx, wL, wr, wl);
        load_decisions(d, dtape, ld, e);
      }
      const RegPolys src{wL, wr, wl, E.h, E.kn};
      float g_am, g_as;
      step_bwd_core(E, cx, src, am, as, d, P, snext, ld, e, g, g_am, g_as, flags);
      if (g_action) {
        g_action[e] = g_am;
        g_action[ld + e] = g_as;
      }
    }
  }
  if (in) {
    store_cotangent(g, g_io, ld, e);
    if (flags_out && flags) atomicOr(&flags_out[e], flags);
  }
}

// The VJP of an env-step is linear in the cotangent and its coefficients depend only on forward
// quantities (the state before the step, the action, the collider's decisions).  For the env-steps
// WITH resolved contacts — whose adjoint is the expensive, divergent one — the 20 x 18 matrix
// (g_state; g_action) = M g_next is formed here for all steps of a group at once: thread = (env-step
// with contacts, k) evaluates the VJP of unit cotangent k.  The sequential sweep then spends a
// matrix-vector product on such an env-step instead of the contact adjoint.
// 20 blocks of 64 threads per SM (51 registers; unconstrained: 96 registers, 10 blocks): the kernel is throughput
// work over (contact env-step, unit cotangent) and gains from occupancy what it loses to spills --
// backward pass of 4096 envs x 100 steps 3.72 -> 3.51 ms (12 / 16 / 20 / 24 blocks: 3.63 / 3.59 / 3.51 / 3.51)
#ifndef CX_JAC_MINB
#define CX_JAC_MINB 20
#endif
__global__ void __launch_bounds__(TPB, CX_JAC_MINB)
lunar_jacobian_kernel(int64_t B, const float* __restrict__ tape, int64_t ld,
                      const float* __restrict__ actions, const float* __restrict__ dtape,
                      uint32_t* __restrict__ flags_out, const JacWs jw,
                      const __grid_constant__ CotixLunarParams P, const LunarDerived D) {
  const uint32_t nh = *jw.ctr < jw.cap ? *jw.ctr : jw.cap;
  // grid-stride: the launcher may cap the grid so that a concurrent sweep launch finds room on every SM
  for (int64_t idx = (int64_t)blockIdx.x * TPB + threadIdx.x; idx < (int64_t)nh * COTIX_LUNAR_NDYN;
       idx += (int64_t)gridDim.x * TPB) {
  const uint32_t slot = (uint32_t)(idx / COTIX_LUNAR_NDYN);
  const int k = (int)(idx - (int64_t)slot * COTIX_LUNAR_NDYN);
  const int64_t i = jw.hlist[slot];
  const int64_t t = i / B, e = i - t * B;
  const float* sin_ = tape + t * (int64_t)COTIX_LUNAR_NF * ld;
  EnvRegs E;
  load_env(E, sin_, ld, e);
  float am = 0.0f, as = 0.0f;
  if (actions) {
    am = actions[t * 2 * ld + e];
    as = actions[t * 2 * ld + ld + e];
  }
  StepCtx cx;
  Poly wL, wr, wl;
  integrate_and_cull<true>(E, am, as, P, D, cx, wL, wr, wl);
  const RegPolys src{wL, wr, wl, E.h, E.kn};
  Decisions d;
  load_decisions(d, dtape + t * (int64_t)DT_WORDS * ld, ld, e);
  G6 g[4];
  g[0] = g[1] = g[2] = g[3] = g6_zero();
  {
    G6& gb = g[k / 6];
    const int l = k % 6;
    if (l == 0) gb.px = 1.0f;
    else if (l == 1) gb.py = 1.0f;
    else if (l == 2) gb.vx = 1.0f;
    else if (l == 3) gb.vy = 1.0f;
    else if (l == 4) gb.ang = 1.0f;
    else gb.w = 1.0f;
  }
  uint32_t flags = 0u;
  float g_am, g_as;
  step_bwd_core(E, cx, src, am, as, d, P, k == 0 ? sin_ + (int64_t)COTIX_LUNAR_NF * ld : nullptr, ld, e, g, g_am,
                g_as, flags);
  // An env that is about to blow up (velocities ~1e20 a step or two before the state goes NaN) has
  // Jacobian entries that overflow.  Its cotangent was dropped downstream, i.e. is zero, and the
  // direct sweep then returns zero; 0 x inf would poison the matrix form, so such entries are stored
  // as 0.
  float* M = jw.M + (size_t)slot * JAC_FLOATS;
  auto fin = [](float v) { return (fabsf(v) <= 3.0e38f) ? v : 0.0f; };
  // column-major: the 20 results of unit cotangent k are contiguous
  float* Mk = M + k * JAC_ROWS;
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    Mk[b * 6 + 0] = fin(g[b].px);
    Mk[b * 6 + 1] = fin(g[b].py);
    Mk[b * 6 + 2] = fin(g[b].vx);
    Mk[b * 6 + 3] = fin(g[b].vy);
    Mk[b * 6 + 4] = fin(g[b].ang);
    Mk[b * 6 + 5] = fin(g[b].w);
  }
  Mk[18] = fin(g_am);
  Mk[19] = fin(g_as);
  if (k == 0 && flags_out && flags) atomicOr(&flags_out[e], flags);
  }
}

// The collider's decisions of nt consecutive steps of a taped rollout in ONE launch: item = (step,
// env).  The replay of a step depends only on the state before it, which is on the tape, so unlike
// the reverse sweep it is not sequential in time — this is where the irregular work of the backward
// pass (narrowphase, coins) runs, at full occupancy instead of one small launch per step.
#ifndef CX_DECIDE_MUL
#define CX_DECIDE_MUL 8
#endif
constexpr int DECIDE_TILE = TPB * CX_DECIDE_MUL;
#ifndef CX_DEC_MINB
#define CX_DEC_MINB 1
#endif
__global__ void __launch_bounds__(TPB, CX_DEC_MINB)
lunar_decide_kernel(int64_t B, int nt, const float* __restrict__ tape, int64_t ld,
                    const float* __restrict__ actions, float* __restrict__ dtape, const JacWs jw,
                    const __grid_constant__ CotixLunarParams P, const LunarDerived D) {
  // Only ~10 % of the items have candidate pairs.  Pass 1 integrates every item of a tile and
  // collects the ones with candidates in shared memory; pass 2 runs the narrowphase and the
  // collider's choices on those with consecutive lanes, so the warps of the irregular part are full.
  __shared__ uint32_t s_list[DECIDE_TILE];
  __shared__ int s_n;
  const int64_t N = B * nt;
  for (int64_t t0 = (int64_t)blockIdx.x * DECIDE_TILE; t0 < N; t0 += (int64_t)gridDim.x * DECIDE_TILE) {
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    for (int j = 0; j < DECIDE_TILE / TPB; ++j) {
      const int64_t i = t0 + j * TPB + threadIdx.x;
      if (i >= N) break;
      const int64_t t = i / B, e = i - t * B;
      EnvRegs E;
      load_env(E, tape + t * (int64_t)COTIX_LUNAR_NF * ld, ld, e);
      float am = 0.0f, as = 0.0f;
      if (actions) {
        am = actions[t * 2 * ld + e];
        as = actions[t * 2 * ld + ld + e];
      }
      StepCtx cx;
      Poly wL, wr, wl;
      integrate_and_cull<false>(E, am, as, P, D, cx, wL, wr, wl);
      if (cx.cand == 0u) dtape[t * (int64_t)DT_WORDS * ld + e] = 0.0f;
      else s_list[atomicAdd(&s_n, 1)] = (uint32_t)(i - t0);
    }
    __syncthreads();
    const int n = s_n;
    for (int q = threadIdx.x; q < n; q += TPB) {
      const int64_t i = t0 + s_list[q];
      const int64_t t = i / B, e = i - t * B;
      EnvRegs E;
      load_env(E, tape + t * (int64_t)COTIX_LUNAR_NF * ld, ld, e);
      float am = 0.0f, as = 0.0f;
      if (actions) {
        am = actions[t * 2 * ld + e];
        as = actions[t * 2 * ld + ld + e];
      }
      StepCtx cx;
      Poly wL, wr, wl;
      integrate_and_cull<true>(E, am, as, P, D, cx, wL, wr, wl);
      const RegPolys src{wL, wr, wl, E.h, E.kn};
      Decisions d;
      decide(E, cx.cand, src, P, d);
      float* dt = dtape + t * (int64_t)DT_WORDS * ld;
      store_decisions(d, dt, ld, e);
      if (jw.M && d.pl.n > 0) {  // an env-step with contacts gets a slot for its transposed Jacobian
        const uint32_t slot = atomicAdd(jw.ctr, 1u);
        if (slot < jw.cap) {
          jw.hlist[slot] = (uint32_t)i;
          dt[e] = __uint_as_float(__float_as_uint(dt[e]) | DT_HAS_MATRIX);
          dt[32 * ld + e] = __uint_as_float(slot);
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace

extern "C" int cx_launch_lunar_step_bwd(cudaStream_t st, int64_t B, const float* state_in,
                                        const float* state_next, int64_t ld, const float* action,
                                        const float* g_out, float* g_in, float* g_action,
                                        uint32_t* flags, const float* dtape, const float* jacM,
                                        const CotixLunarParams* p) {
  if (B == 0) return 0;
  LunarDerived D;
  cx_lunar_derive(p, &D);
  lunar_step_bwd_kernel<<<(unsigned)((B + TPB - 1) / TPB), TPB, 0, st>>>(
      B, state_in, state_next, ld, action, g_out, g_in, g_action, flags, dtape, jacM, *p, D);
  cx_count_launch();
  return cx_check_launch("lunar_step_bwd_kernel");
}

// steps [s_lo, s_hi) backwards in one launch (small batches, taped mode with Jacobians; see the kernel)
extern "C" int cx_launch_lunar_sweep_bwd(cudaStream_t st, int64_t B, int s_lo, int s_hi, const float* tape, int64_t ld,
                                         const float* actions, float* g, float* g_actions, uint32_t* flags,
                                         const float* dtape0, const float* jacM, const CotixLunarParams* p) {
  if (B == 0 || s_hi <= s_lo) return 0;
  LunarDerived D;
  cx_lunar_derive(p, &D);
  int epw = 16;  // measured, 4096 envs x 100 steps, 2 / 4 / 8 / 16 / 32 envs per warp: 5.08 / 4.57 / 4.02 / 3.74 / 4.05 ms
  if (const char* env = getenv("COTIX_SWEEP_EPW")) {  // measurement aid
    const int v = atoi(env);
    if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) epw = v;
  }
  const int64_t warps = (B + epw - 1) / epw;
  lunar_sweep_bwd_kernel<<<(unsigned)((warps + TPB / 32 - 1) / (TPB / 32)), TPB, 0, st>>>(
      B, s_lo, s_hi, tape, ld, actions, g, g_actions, flags, dtape0, jacM, *p, D, epw);
  cx_count_launch();
  return cx_check_launch("lunar_sweep_bwd_kernel");
}

// decisions of steps [t0, t0 + nt) of a taped rollout -> dtape [nt][COTIX_ADJ_TAPE_WORDS][ld]; with a
// Jacobian area (jac_ctr, jac_list, jac_M, jac_cap; or all NULL / 0) also the transposed Jacobians of
// the env-steps with contacts
extern "C" int cx_launch_lunar_decide(cudaStream_t st, int64_t B, int nt, const float* tape_t0, int64_t ld,
                                      const float* actions_t0, float* dtape, uint32_t* flags,
                                      uint32_t* jac_ctr, uint32_t* jac_list, float* jac_M, uint32_t jac_cap,
                                      const CotixLunarParams* p, int blocks_per_sm) {
  // blocks_per_sm > 0: both grids are capped at that many blocks per SM (they loop), which leaves
  // registers on every SM for sweep launches running concurrently on another stream
  if (B == 0 || nt == 0) return 0;
  LunarDerived D;
  cx_lunar_derive(p, &D);
  const int64_t n = B * nt;
  JacWs jw;
  jw.ctr = jac_ctr;
  jw.hlist = jac_list;
  jw.M = jac_M;
  jw.cap = jac_M ? jac_cap : 0u;
  if (jw.M && cudaMemsetAsync(jw.ctr, 0, 4, st) != cudaSuccess) return cx_check_launch("adjoint memset");
  int64_t blocks = (n + DECIDE_TILE - 1) / DECIDE_TILE;
  const int64_t cap = (int64_t)cx_sm_count() * (blocks_per_sm > 0 ? blocks_per_sm : 16);
  if (blocks > cap) blocks = cap;
  lunar_decide_kernel<<<(unsigned)blocks, TPB, 0, st>>>(B, nt, tape_t0, ld, actions_t0, dtape, jw, *p, D);
  cx_count_launch();
  if (jw.M) {
    const int64_t threads = (int64_t)jw.cap * COTIX_LUNAR_NDYN;
    int64_t jblocks = (threads + TPB - 1) / TPB;
    if (blocks_per_sm > 0 && jblocks > (int64_t)cx_sm_count() * blocks_per_sm)
      jblocks = (int64_t)cx_sm_count() * blocks_per_sm;
    lunar_jacobian_kernel<<<(unsigned)jblocks, TPB, 0, st>>>(B, tape_t0, ld, actions_t0, dtape,
                                                                                flags, jw, *p, D);
    cx_count_launch();
  }
  return cx_check_launch("lunar_decide_kernel");
}
