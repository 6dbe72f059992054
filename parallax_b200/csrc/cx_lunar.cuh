// cx_lunar.cuh — device functions of the fused LunarLander step.
//
// Reference step (examples/test_viz.py:25-44):
//   ExplicitEulerPhysics.step   cotix/_physics_solvers.py:16-33
//   gravity on the lander       examples/test_viz.py:27-31
//   RandomizedCollider.resolve  cotix/_colliders.py:74-351
//   key = split(key)[0]         examples/test_viz.py:39
//   LunarLander.step (joints)   cotix/_lunar_lander.py:145-218
//
// How the collider is restructured (results are identical):
//  * the reference evaluates 81 + 225 narrowphase calls per step for 25 distinct part pairs
//    (24 real + the left leg against itself); here each distinct pair is evaluated at most
//    once, and only if its AABBs come within CULL_MARGIN of each other;
//  * every (ind2, ind1) scatter entry of the reference draws a coin from its own threefry
//    key; only the coins of entries whose contact is non-NaN can change the outcome, and for
//    a cell the LAST heads in scan order wins — so coins are evaluated lazily, from the last
//    entry backwards, stopping at the first heads;
//  * the self pair (left leg vs itself) only dilutes body 2's random choice, so
//    it is evaluated only when body 2 has another live contact.
#pragma once
#include "../../include/cotix_b200.h"
#include "cx_narrow.cuh"
#include "cx_rng.cuh"

#define CX_CULL_MARGIN 1e-3f

struct BodyS {
  float px, py, vx, vy, ang, w;
};
// Exact division by a launch constant c in three operations instead of the ~22 of an IEEE division:
//   q = x * r;  rem = fma(-q, c, x);  x / c == fma(rem, r, q)        with r = RN(1 / c)
// (the remainder is exact, the correction lands on the correctly rounded quotient).  Whether that holds
// for EVERY x is a property of c: the host checks it exhaustively over one binade of x (2^23 values; the
// result is scale-invariant while nothing under- or overflows) the first time a constant is seen
// (cx_divc_make, lunar_kernel.cu) and sets `ok`; x outside [1e-20, 1e20], x = 0 and non-finite x take
// the plain division, as does every constant that fails the check.  The results are the bits of x / c.
// ok == 2: c is +-inf (the static ground's mass and inertia), where x / c == x * (+-0) for EVERY x — a
// finite x gives the zero of the right sign, an infinite or NaN x gives NaN — so r = +-0 and one multiply
// replaces the division, whose special-operand path was 4 % of the contact stage's instructions.
struct DivC {
  float c, r;
  int ok;
};
// the IEEE division out of line: after the shortcuts below it runs for out-of-range operands only, and its
// ~70 inlined instructions per call site were a seventh of the contact stage's code
static __device__ __noinline__ float div_ieee(float x, float c) { return x / c; }
CX_DEV DivC divc_plain(float c) {
  DivC d;
  d.c = c;
  d.r = 0.0f;
  d.ok = 0;
  return d;
}
CX_DEV float div_c(float x, const DivC& d) {
  if (d.c == 1.0f) return x;  // x / 1 == x bit for bit (the legs' mass and inertia); uniform branch
  if (d.ok == 2) return x * d.r;
  if (d.ok) {
    const float ax = fabsf(x);
    if (ax >= 1e-20f && ax <= 1e20f) {  // false for NaN
      const float q = x * d.r;
      const float rem = __fmaf_rn(-q, d.c, x);
      return __fmaf_rn(rem, d.r, q);
    }
  }
  return div_ieee(x, d.c);
}
struct MassP {
  float mass, inertia, e, mu;
  float inv_mass;  // 1.0f / mass, IEEE: the same bits whether the host or a thread divides
  DivC dmass, dinertia;
};
CX_DEV MassP mk_massp(const float props[4], float inv_mass) {
  MassP m;
  m.mass = props[0];
  m.inertia = props[1];
  m.e = props[2];
  m.mu = props[3];
  m.inv_mass = inv_mass;
  m.dmass = divc_plain(props[0]);
  m.dinertia = divc_plain(props[1]);
  return m;
}
CX_DEV MassP mk_massp(const float props[4]) { return mk_massp(props, 1.0f / props[0]); }

// ---- order_clockwise on a register polygon of N vertices (cotix/_geometry_utils.py:60-67) ----
template <int N>
static __device__ __noinline__ void poly_sort_cw(Poly& p) {
  float sx = 0.0f, sy = 0.0f;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    sx = sx + p.x[k];
    sy = sy + p.y[k];
  }
  const float mx = sx / (float)N, my = sy / (float)N;
  float ang[N];
#pragma unroll
  for (int k = 0; k < N; ++k) ang[k] = cx_atan2f(p.y[k] - my, p.x[k] - mx);
  int rank[N];
#pragma unroll
  for (int q = 0; q < N; ++q) {
    int r = 0;
    const float aq = ang[q];
    const bool qn = aq != aq;
#pragma unroll
    for (int t = 0; t < N; ++t) {
      if (t != q) {
        const float at = ang[t];
        const bool tn = at != at;
        bool before;  // stable, NaN last (jnp.argsort)
        if (tn || qn) before = (!tn && qn) || (tn == qn && t < q);
        else before = (at < aq) || (at == aq && t < q);
        if (before) ++r;
      }
    }
    rank[q] = r;
  }
  float ox[N], oy[N];
#pragma unroll
  for (int t = 0; t < N; ++t) {
    ox[t] = p.x[0];
    oy[t] = p.y[0];
#pragma unroll
    for (int q = 0; q < N; ++q)
      if (rank[q] == t) {
        ox[t] = p.x[q];
        oy[t] = p.y[q];
      }
  }
#pragma unroll
  for (int t = 0; t < N; ++t) {
    p.x[t] = ox[t];
    p.y[t] = oy[t];
  }
}

// Facts about the (constant) local shapes of the three dynamic bodies, derived on the host
// once per launch (lunar_kernel.cu: derive()).
struct LunarDerived {
  int rigid_ok[3];  // the world-space order_clockwise of body i is a cyclic rotation of its
                    // local vertex order (convex, centroid strictly inside, angular gaps > 1e-3)
  int orient[3];    // +1: local order is counter-clockwise (ascending angle), -1: clockwise
  float lext[4];    // lander local extents: min x, min y, max x, max y
  float inv_mass[4];  // 1.0f / body_props[i][0] (launch constants: divided once on the host, not per thread)
  DivC dmass[4], dinertia[4];  // exact division by the bodies' mass / inertia (div_c)
};
CX_DEV MassP mk_massp(const float props[4], const LunarDerived& D, int i) {
  MassP m = mk_massp(props, D.inv_mass[i]);
  m.dmass = D.dmass[i];
  m.dinertia = D.dinertia[i];
  return m;
}

// order_clockwise of a RIGIDLY transformed convex polygon without computing any angle.
// The vertices keep their cyclic order under rotation + translation, so the ascending-atan2
// order about the centroid is the counter-clockwise cyclic order started at the first vertex
// below the centroid whose predecessor is not (the +-pi branch cut of atan2).  Angular gaps
// between neighbours are > 1e-3 rad (checked on the host), 1000x the error of any fp32
// atan2, so this is exactly the permutation the stable argsort produces.  Returns false (and
// leaves p untouched) when the pattern is not found (NaN, degenerate): caller sorts generically.
template <int N>
CX_DEV bool poly_order_rigid(Poly& p, int orient) {
  float sy = 0.0f;
#pragma unroll
  for (int k = 0; k < N; ++k) sy = sy + p.y[k];
  const float my = sy / (float)N;
  float cx_[N], cy_[N];
  bool neg[N];
#pragma unroll
  for (int k = 0; k < N; ++k) {  // counter-clockwise copy
    const int src = orient > 0 ? k : N - 1 - k;
    cx_[k] = p.x[src];
    cy_[k] = p.y[src];
    neg[k] = (cy_[k] - my) < 0.0f;
  }
  int start = -1, nstart = 0;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    const bool prev_neg = neg[(k + N - 1) % N];
    if (neg[k] && !prev_neg) {
      start = k;
      ++nstart;
    }
  }
  bool finite = true;
#pragma unroll
  for (int k = 0; k < N; ++k) finite = finite && (fabsf(cx_[k]) <= 3.0e38f) && (fabsf(cy_[k]) <= 3.0e38f);
  if (nstart != 1 || !finite) return false;
#pragma unroll
  for (int t = 0; t < N; ++t) {
    float ox = cx_[0], oy = cy_[0];
#pragma unroll
    for (int k = 0; k < N; ++k)
      if ((start + t) % N == k) {
        ox = cx_[k];
        oy = cy_[k];
      }
    p.x[t] = ox;
    p.y[t] = oy;
  }
  return true;
}

// ---- resolution — cotix/_collision_resolution.py:52-151 -------------------------------
// x / m; a divisor of exactly 1 (the legs' mass and inertia) is skipped — x / 1 == x bit for bit —
// and since m is a launch constant the branch is uniform
CX_DEV void apply_impulse(BodyS& b, const MassP& m, V2 imp, V2 point) {  // :68-73
  V2 arm = point - mk2(b.px, b.py);
  float torque = cross2(arm, imp);
  b.vx = b.vx + div_c(imp.x, m.dmass);
  b.vy = b.vy + div_c(imp.y, m.dmass);
  b.w = b.w + div_c(torque, m.dinertia);
}
// the same with plain IEEE divisions: resolve_collision is out of line and gets its MassP by reference,
// i.e. from local memory, where the three extra loads per division of div_c cost more than they save
// (contact stage 64.8 -> 70.2 us); the inlined joints of the finish stage are where div_c pays (64 -> 53.5)
CX_DEV float div_m(float x, float m) { return (m == 1.0f) ? x : x / m; }
CX_DEV void apply_impulse_plain(BodyS& b, const MassP& m, V2 imp, V2 point) {
  V2 arm = point - mk2(b.px, b.py);
  float torque = cross2(arm, imp);
  b.vx = b.vx + div_m(imp.x, m.mass);
  b.vy = b.vy + div_m(imp.y, m.mass);
  b.w = b.w + div_m(torque, m.inertia);
}
CX_DEV V2 velocity_at(const BodyS& b, V2 point) {  // cotix/_bodies.py:50-55
  return mk2(b.vx, b.vy) + perp(point - mk2(b.px, b.py)) * b.w;
}
template <bool DIVC>
CX_DEV void resolve_collision_inl(BodyS& b1, const MassP& m1, BodyS& b2, const MassP& m2, V2 pen, V2 cp,
                                  float baumgarte, float baumgarte_dt) {
  if (isnan2(cp)) return;
  V2 p1 = mk2(b1.px, b1.py), p2 = mk2(b2.px, b2.py);
  V2 v1 = mk2(b1.vx, b1.vy) + perp(cp - p1) * b1.w;
  V2 v2 = mk2(b2.vx, b2.vy) + perp(cp - p2) * b2.w;
  V2 vrel = v2 - v1;
  float pn = norm2(pen);
  V2 nd = pen / pn;
  float vn = dot2(vrel, nd);
  float el = jmin(m1.e, m2.e);
  V2 r1 = cp - p1, r2 = cp - p2;
  float lever1 = sumsq2(r1), lever2 = sumsq2(r2);
  float ang = DIVC ? div_c(lever1, m1.dinertia) + div_c(lever2, m2.dinertia)
                   : lever1 / m1.inertia + lever2 / m2.inertia;
  float nim = -(1.0f + el) * vn - baumgarte * norm2(pen) / baumgarte_dt;
  float kden = m1.inv_mass + m2.inv_mass + ang;  // 1.0f / m1.mass + 1.0f / m2.mass + ang
  float nimp = nim / kden;
  V2 ivec = nimp * nd;
  float fr = (m1.mu + m2.mu) / 2.0f;
  V2 vdrag = vrel + vn * nd;
  float vdn = norm2(vdrag);
  V2 vdu = vdrag / vdn;
  float idrag = -norm2(vdrag) / (m1.inv_mass + m2.inv_mass + ang);
  idrag = jclip(idrag, 0.0f, nimp * fr);
  V2 idv = idrag * vdu;
  ivec = ivec + idv;
  if (dot2(pen, vrel) < 0) return;
  if (DIVC) {
    apply_impulse(b1, m1, -ivec, cp);
    apply_impulse(b2, m2, ivec, cp);
  } else {
    apply_impulse_plain(b1, m1, -ivec, cp);
    apply_impulse_plain(b2, m2, ivec, cp);
  }
}
static __device__ __noinline__ void resolve_collision(BodyS& b1, const MassP& m1, BodyS& b2,
                                               const MassP& m2, V2 pen, V2 cp, float baumgarte,
                                               float baumgarte_dt) {
  resolve_collision_inl<false>(b1, m1, b2, m2, pen, cp, baumgarte, baumgarte_dt);
}

// ---- joints — cotix/_lunar_lander.py:145-218 -------------------------------------------
CX_DEV void joint_constraint_inl(BodyS& b1, const MassP& m1, V2 c1, BodyS& b2, const MassP& m2, V2 c2,
                                 const CotixLunarParams& P) {
  V2 dp = c1 - c2;
  V2 dv = velocity_at(b1, c1) - velocity_at(b2, c2);
  V2 imp = dp * P.spring_kp + dv * (norm2(dv) + P.spring_kv0) * P.spring_kv;
  apply_impulse(b1, m1, -imp, c1);
  apply_impulse(b2, m2, imp, c2);
}
// out of line (the fused kernel and the adjoint: code size); the pipeline's finish stage inlines it so that
// the masses and their DivC fold into constant-bank operands instead of local-memory loads
static __device__ __noinline__ void joint_constraint(BodyS& b1, const MassP& m1, V2 c1, BodyS& b2, const MassP& m2, V2 c2,
                                                     const CotixLunarParams& P) {
  joint_constraint_inl(b1, m1, c1, b2, m2, c2, P);
}
CX_DEV V2 rot_cs(V2 v, float c, float s) {  // rotate(): [[c,-s],[s,c]] @ v
  return mk2(c * v.x + (-s) * v.y, s * v.x + c * v.y);
}

// ---- coin flips of the scatter scans — cotix/_colliders.py:208-268 --------------------
struct CoinCtx {
  int mode;           // COTIX_RNG_*
  Key2 gkey[2];       // threefry: skey of group 0 / group 1 (after its split, :175)
  Key2 pkey;          // philox key
  uint32_t steps, episode;
  float p;
};
// heads of entry (group g, ind2, ind1); k2 = split(gkey[g], N2)[ind2] is passed in
CX_DEV bool coin_heads(const CoinCtx& cc, int g, int N1, Key2 k2, int ind2, int ind1) {
  if (cc.mode == COTIX_RNG_THREEFRY) {
    Key2 k1 = tf_split(k2, (uint32_t)N1, (uint32_t)ind1);
    Key2 kk = tf_split(k1, 2u, 0u);
    return tf_bernoulli(kk, cc.p);
  }
  uint32_t w = philox_word(cc.pkey, cc.steps, cc.episode, (uint32_t)g,
                           (uint32_t)(ind2 * N1 + ind1), CX_TAG_COIN);
  return bits_to_uniform(w, 0.0f, 1.0f) < cc.p;
}
// Winner of one all_contacts cell: the entry with the largest (ind2, ind1) whose contact is
// live and whose coin is heads.  Entries of the cell: ind2 in [b_lo, b_hi], ind1 in
// {a0 + q + 7*r : q live in `mask` (bit q), r < reps}.  Returns q or -1.
static __device__ __noinline__ int cell_winner(const CoinCtx& cc, int g, int N1, int N2, int b_lo,
                                        int b_hi, int a0, int nq, int reps, uint32_t mask) {
  if (mask == 0u) return -1;
  for (int ind2 = b_hi; ind2 >= b_lo; --ind2) {
    Key2 k2;
    k2.a = k2.b = 0u;
    if (cc.mode == COTIX_RNG_THREEFRY) k2 = tf_split(cc.gkey[g], (uint32_t)N2, (uint32_t)ind2);
    for (int r = reps - 1; r >= 0; --r)
      for (int q = nq - 1; q >= 0; --q) {
        if (!((mask >> q) & 1u)) continue;
        int ind1 = a0 + q + nq * r;
        if (coin_heads(cc, g, N1, k2, ind2, ind1)) return q;
      }
  }
  return -1;
}

// jr.choice over the live cells of one all_contacts row — cotix/_colliders.py:274-291.
// m[j] = cell (i, j) live; u = the uniform draw of body i's key.
CX_DEV int choose_index(bool m0, bool m1, bool m2, float u) {
  int cnt = (m0 ? 1 : 0) + (m1 ? 1 : 0) + (m2 ? 1 : 0);
  float fc = (float)cnt;
  float p0 = (m0 ? 1.0f : 0.0f) / fc, p1 = (m1 ? 1.0f : 0.0f) / fc, p2 = (m2 ? 1.0f : 0.0f) / fc;
  float p3 = 0.0f / fc;
  float c0 = p0, c1 = c0 + p1, c2 = c1 + p2, c3 = c2 + p3;
  float r = c3 * (1.0f - u);
  int ind = (c0 < r ? 1 : 0) + (c1 < r ? 1 : 0) + (c2 < r ? 1 : 0) + (c3 < r ? 1 : 0);
  return ind > 3 ? 3 : ind;
}
