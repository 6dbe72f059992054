/*
 * cotix_oracle.cpp — CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library, and only as the checker / the timed CPU baseline.  Nothing under
 * parallax_b200/ links, imports or calls it.
 *
 * What it restates (paths relative to the reference root, the `cotix` package):
 *   cotix/_geometry_utils.py   :12-27 is_point_in_triangle, :30-34 fast_normal,
 *                              :49-57 minkowski_diff, :60-67 order_clockwise,
 *                              :81-88 rotate, :102-138 HomogenuousTransformer
 *   cotix/_convex_shapes.py    :22-29 Circle, :62-117 AABB, :149-187 AbstractPolygon
 *   cotix/_collisions.py       :20-112 GJK, :115-273 EPA, :277-329 public wrappers
 *   cotix/_contacts.py         :30-315 all six contact functions
 *   cotix/_collision_resolution.py :52-151
 *   cotix/_colliders.py        :74-351 RandomizedCollider.resolve
 *   cotix/_physics_solvers.py  :16-33
 *   cotix/_lunar_lander.py     :11-218
 *   examples/test_viz.py       :25-44 (the composed step)
 * and, because they live in the un-vendored dependency jax (pyproject.toml:23, no pin;
 * the reference needs jax < 0.5 because it calls jax.tree_map, cotix/_utils.py:104, so the
 * default PRNG is non-partitionable threefry2x32), the published jax.random algorithms:
 * threefry2x32 (Random123), split, random bits, uniform, bernoulli, choice, normal.
 *
 * PARITY PINNING: the reference cannot be imported in the build container (no jax, no
 * equinox).  The restatement is pinned by (1) the reference tests' own inputs and
 * invariants (tests/test_oracle_reference_cases.py), (2) Random123 known-answer vectors
 * for threefry, (3) golden vectors produced by executing the reference's OWN source files
 * under a numpy float32 stand-in for the jax API (oracle/jaxlite, oracle/gen_golden.py,
 * tests/golden/): contacts, resolution, reset and 24-step contact-rich rollouts reproduce
 * bit for bit, including the RNG-dependent behaviour (coin flips, choice, terrain).  Agreement
 * with a real jax/XLA run cannot be checked here (no jax): for XLA's rounding choices the bar
 * is the 1e-5 relative tolerance of north_star.
 *
 * Style: literal.  Every duplicated narrowphase evaluation, every coin flip and every
 * fixed-size EPA buffer of the reference is reproduced as written (set `dedupe` to memoise
 * identical shape pairs — results are identical, see tests).  All arithmetic is float32,
 * built with -ffp-contract=off.
 */
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/cotix_b200.h"
#include "../parallax_b200/csrc/cx_trig.h"

namespace {

const float kNaN = std::numeric_limits<float>::quiet_NaN();
const float kInf = std::numeric_limits<float>::infinity();

// ------------------------------------------------------------------------------------
// trig: libm by default; mode 1 = the engine's own definition (cx_trig.h), which lets the
// parity tests ask for exact equality instead of a tolerance.
// ------------------------------------------------------------------------------------
int g_trig_mode = 0;
inline float o_sin(float x) {
  if (g_trig_mode == 0) return sinf(x);
  float s, c;
  cx_sincosf(x, &s, &c);
  return s;
}
inline float o_cos(float x) {
  if (g_trig_mode == 0) return cosf(x);
  float s, c;
  cx_sincosf(x, &s, &c);
  return c;
}
inline float o_atan2(float y, float x) {
  if (g_trig_mode == 0) return atan2f(y, x);
  return cx_atan2f(y, x);
}

// ------------------------------------------------------------------------------------
// small float32 vector helpers with jnp semantics
// ------------------------------------------------------------------------------------
struct V2 {
  float x, y;
};
inline V2 mk(float x, float y) {
  V2 r;
  r.x = x;
  r.y = y;
  return r;
}
inline V2 operator+(V2 a, V2 b) { return mk(a.x + b.x, a.y + b.y); }
inline V2 operator-(V2 a, V2 b) { return mk(a.x - b.x, a.y - b.y); }
inline V2 operator-(V2 a) { return mk(-a.x, -a.y); }
inline V2 operator*(V2 a, float s) { return mk(a.x * s, a.y * s); }
inline V2 operator*(float s, V2 a) { return mk(s * a.x, s * a.y); }
inline V2 operator/(V2 a, float s) { return mk(a.x / s, a.y / s); }
inline float dot(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }        // jnp.dot
inline float cross2(V2 a, V2 b) { return a.x * b.y - a.y * b.x; }     // jnp.cross (2-D)
inline float sumsq(V2 a) { return a.x * a.x + a.y * a.y; }            // jnp.sum(a**2)
inline float norm(V2 a) { return sqrtf(a.x * a.x + a.y * a.y); }      // jnp.linalg.norm
inline bool isnan2(V2 a) { return (a.x != a.x) || (a.y != a.y); }
inline V2 nan2() { return mk(kNaN, kNaN); }
// _geometry_utils.py:30-34 and :70-72
inline V2 fast_normal(V2 a) { return mk(-a.y, a.x); }
// jnp.minimum / jnp.maximum propagate NaN
inline float jmin(float a, float b) {
  if (a != a || b != b) return kNaN;
  return a < b ? a : b;
}
inline float jmax(float a, float b) {
  if (a != a || b != b) return kNaN;
  return a > b ? a : b;
}
inline float jclip(float t, float lo, float hi) { return jmin(jmax(t, lo), hi); }
// jnp.argmin / jnp.argmax: first extremum, NaN wins (numpy semantics)
int jargmin(const float* v, int n) {
  int best = 0;
  for (int i = 0; i < n; ++i) {
    if (v[i] != v[i]) return i;
    if (v[i] < v[best]) best = i;
  }
  return best;
}
int jargmax(const float* v, int n) {
  int best = 0;
  for (int i = 0; i < n; ++i) {
    if (v[i] != v[i]) return i;
    if (v[i] > v[best]) best = i;
  }
  return best;
}
inline float jsign(float v) {
  if (v != v) return kNaN;
  return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f);
}

// _geometry_utils.py:81-88 rotate(vector, angle) = [[c,-s],[s,c]] @ vector
inline V2 rotate(V2 v, float angle) {
  float c = o_cos(angle), s = o_sin(angle);
  return mk(c * v.x + (-s) * v.y, s * v.x + c * v.y);
}

// ------------------------------------------------------------------------------------
// shapes
// ------------------------------------------------------------------------------------
enum { T_CIRCLE = 0, T_AABB = 1, T_POLY = 2 };
struct Shape {
  int type;
  int n;          // polygon vertex count
  float radius;   // circle
  V2 v[COTIX_MAX_POLY];  // circle: v[0] = position; aabb: v[0] = lower, v[1] = upper
};

// _geometry_utils.py:60-67 order_clockwise: stable argsort of atan2 about the mean
void order_clockwise(V2* v, int n) {
  float sx = 0.0f, sy = 0.0f;
  for (int i = 0; i < n; ++i) {
    sx = sx + v[i].x;
    sy = sy + v[i].y;
  }
  float mx = sx / (float)n, my = sy / (float)n;
  float ang[COTIX_MAX_POLY];
  int idx[COTIX_MAX_POLY];
  for (int i = 0; i < n; ++i) {
    ang[i] = o_atan2(v[i].y - my, v[i].x - mx);
    idx[i] = i;
  }
  // stable insertion sort, NaN last (jnp.argsort)
  for (int i = 1; i < n; ++i) {
    int k = idx[i];
    float a = ang[k];
    int j = i - 1;
    while (j >= 0) {
      float b = ang[idx[j]];
      bool b_gt_a = (b != b) ? !(a != a) : (a == a && b > a);
      if (!b_gt_a) break;
      idx[j + 1] = idx[j];
      --j;
    }
    idx[j + 1] = k;
  }
  V2 out[COTIX_MAX_POLY];
  for (int i = 0; i < n; ++i) out[i] = v[idx[i]];
  for (int i = 0; i < n; ++i) v[i] = out[i];
}

// get_support: _convex_shapes.py:22-26 (Circle), :62-66 (AABB), :149-155 (polygon)
V2 support(const Shape& s, V2 d) {
  if (s.type == T_CIRCLE) {
    V2 nd = d / norm(d);
    return nd * s.radius + s.v[0];
  } else if (s.type == T_AABB) {
    return mk(d.x >= 0.0f ? s.v[1].x : s.v[0].x, d.y >= 0.0f ? s.v[1].y : s.v[0].y);
  } else {
    if (isnan2(d)) return nan2();
    float dots[COTIX_MAX_POLY];
    for (int i = 0; i < s.n; ++i) dots[i] = dot(s.v[i], d);
    return s.v[jargmax(dots, s.n)];
  }
}
// _geometry_utils.py:49-57
inline V2 minkowski_diff(const Shape& a, const Shape& b, V2 d) {
  return support(a, d) - support(b, -d);
}

// contains: Circle :28-29, AABB :105-106, polygon :168-175
bool contains(const Shape& s, V2 p) {
  if (s.type == T_CIRCLE) {
    const float eps = 1e-6f;
    float re = s.radius + eps;
    return sumsq(p - s.v[0]) <= re * re;
  } else if (s.type == T_AABB) {
    const float eps = 1e-6f;
    return (p.x >= s.v[0].x - eps) && (p.x <= s.v[1].x + eps) && (p.y >= s.v[0].y - eps) &&
           (p.y <= s.v[1].y + eps);
  } else {
    float sg[COTIX_MAX_POLY];
    for (int k = 0; k < s.n; ++k) {
      V2 e0 = s.v[k];
      V2 e1 = s.v[(k + s.n - 1) % s.n];
      sg[k] = jsign(dot(p - e0, fast_normal(e0 - e1)));
    }
    for (int k = 0; k < s.n; ++k)
      if (!(sg[k] == sg[0])) return false;
    return true;
  }
}

struct Edge {
  V2 a, b;
};
// get_edges / get_vertices: AABB :82-103, polygon :160-166
int get_vertices(const Shape& s, V2* out) {
  if (s.type == T_AABB) {
    out[0] = s.v[1];
    out[1] = mk(s.v[1].x, s.v[0].y);
    out[2] = s.v[0];
    out[3] = mk(s.v[0].x, s.v[1].y);
    return 4;
  }
  for (int k = 0; k < s.n; ++k) out[k] = s.v[k];
  return s.n;
}
int get_edges(const Shape& s, Edge* out) {
  if (s.type == T_AABB) {
    V2 vs[4];
    get_vertices(s, vs);
    for (int k = 0; k < 4; ++k) {
      out[k].a = vs[k];
      out[k].b = vs[(k + 1) % 4];
    }
    return 4;
  }
  for (int k = 0; k < s.n; ++k) {
    out[k].a = s.v[k];
    out[k].b = s.v[(k + s.n - 1) % s.n];  // jnp.roll(vertices, shift=1)
  }
  return s.n;
}

// ------------------------------------------------------------------------------------
// GJK — _collisions.py:20-112
// ------------------------------------------------------------------------------------
// _geometry_utils.py:12-27
bool is_point_in_triangle(V2 pt, V2 v1, V2 v2, V2 v3) {
  auto sign = [](V2 p1, V2 p2, V2 p3) {
    return (p1.x - p3.x) * (p2.y - p3.y) - (p2.x - p3.x) * (p1.y - p3.y);
  };
  float d1 = sign(pt, v1, v2), d2 = sign(pt, v2, v3), d3 = sign(pt, v3, v1);
  bool has_neg = (d1 < 0) || (d2 < 0) || (d3 < 0);
  bool has_pos = (d1 > 0) || (d2 > 0) || (d3 > 0);
  return !(has_neg && has_pos);
}

void get_collision_simplex(const Shape& A, const Shape& B, V2 initial_direction,
                           int max_steps, V2 simplex[3], int* iters_out) {
  simplex[0] = minkowski_diff(A, B, initial_direction);
  simplex[1] = minkowski_diff(A, B, -simplex[0]);
  simplex[2] = mk(0.0f, 0.0f);
  V2 direction = fast_normal(simplex[1] - simplex[0]);
  if (dot(direction, -simplex[1]) > 0) {  // reverse_simplex
    V2 tmp = simplex[0];
    simplex[0] = simplex[1];
    simplex[1] = tmp;
  } else {  // reverse_direction
    direction = -direction;
  }
  simplex[2] = minkowski_diff(A, B, direction);
  int steps = 0;
  for (;;) {
    // cond_fn :85-97
    bool c1 = dot(simplex[2], direction) <= 0;
    bool c2 = dot(fast_normal(simplex[2] - simplex[0]), -simplex[2]) < 0;
    bool c3 = dot(fast_normal(simplex[1] - simplex[2]), -simplex[2]) < 0;
    bool go = !(c1 || (c2 && c3));
    if (!go || steps >= max_steps) break;
    // body_fn :62-83
    V2 a = simplex[0], b = simplex[1], c = simplex[2];
    V2 ac_normal = fast_normal(c - a);
    V2 cb_normal = fast_normal(b - c);
    if (dot(ac_normal, -c) >= 0) {
      simplex[1] = c;
      direction = ac_normal;
    } else {
      simplex[0] = c;
      direction = cb_normal;
    }
    simplex[2] = minkowski_diff(A, B, direction);
    ++steps;
  }
  if (iters_out) *iters_out = steps;
  if (!is_point_in_triangle(mk(0.0f, 0.0f), simplex[0], simplex[1], simplex[2])) {
    simplex[0] = simplex[1] = simplex[2] = mk(0.0f, 0.0f);
  }
}

// check_for_collision_convex — _collisions.py:277-310.  `start` is the already-resolved
// start direction (random_direction(PRNGKey(1)) when the caller passes none, :287-298).
bool check_for_collision_convex(const Shape& A, const Shape& B, V2 start, int max_steps,
                                V2 simplex[3], int* iters_out) {
  get_collision_simplex(A, B, start, max_steps, simplex, iters_out);
  float area = cross2(simplex[1] - simplex[0], simplex[2] - simplex[0]);
  bool all_zero = true, any_nan = false;
  for (int k = 0; k < 3; ++k) {
    if (!(simplex[k].x == 0.0f && simplex[k].y == 0.0f)) all_zero = false;
    if (isnan2(simplex[k])) any_nan = true;
  }
  bool c = all_zero || any_nan || (area == 0);
  if (c) {
    simplex[0] = simplex[1] = simplex[2] = nan2();
    return false;
  }
  return true;
}

// ------------------------------------------------------------------------------------
// EPA — _collisions.py:115-273
// ------------------------------------------------------------------------------------
V2 displacement_to_origin(V2 a, V2 b) {  // :137-154
  bool both_zero = (a.x == 0.0f) && (a.y == 0.0f) && (b.x == 0.0f) && (b.y == 0.0f);
  if (both_zero) return mk(kInf, kInf);
  V2 point = mk(0.0f, 0.0f);
  V2 ab = a - b;
  float length = sumsq(ab);
  float t = dot(point - b, ab) / length;
  t = jclip(t, 0.0f, 1.0f);
  V2 projection = b + t * ab;
  V2 displacement = point - projection;
  if (length == 0) return -a;
  return displacement;
}
V2 closest_point_on_edge_to_point(V2 a, V2 b, V2 point) {  // :156-166
  V2 ab = a - b;
  float length = sumsq(ab);
  if (length == 0.0f) return point - a;
  float t = dot(point - b, ab) / length;
  t = jclip(t, 0.0f, 1.0f);
  V2 projection = b + t * ab;
  return point - projection;
}
inline float distance_to_origin(const Edge& e) { return sumsq(displacement_to_origin(e.a, e.b)); }

int closest_edge_to_origin(const std::vector<Edge>& edges) {  // :171-175
  std::vector<float> d(edges.size());
  for (size_t k = 0; k < edges.size(); ++k) d[k] = distance_to_origin(edges[k]);
  return jargmin(d.data(), (int)d.size());
}

V2 get_closest_minkowski_diff(const Shape& A, const Shape& B, const V2 simplex[3],
                              int solver_iterations, int* iters_out) {
  std::vector<Edge> edges((size_t)solver_iterations + 3);
  for (auto& e : edges) e.a = e.b = mk(0.0f, 0.0f);
  edges[0].a = simplex[0];
  edges[0].b = simplex[1];
  edges[1].a = simplex[1];
  edges[1].b = simplex[2];
  edges[2].a = simplex[2];
  edges[2].b = simplex[0];
  int best_idx = closest_edge_to_origin(edges);
  Edge best = edges[(size_t)best_idx];
  V2 new_point = simplex[2];
  int i = 0;
  Edge prev = edges[0];
  int ran = 0;
  for (int it = 0; it < solver_iterations; ++it) {
    // cond_fn :178-212
    bool c1 = sumsq(best.a - best.b) > 1e-9f;
    bool c2 = cross2(best.a, best.b) >= 0;
    V2 normal = fast_normal(prev.a - prev.b);
    normal = normal / norm(normal);
    float d = dot(new_point, normal);
    float edistance = norm(closest_point_on_edge_to_point(prev.a, prev.b, mk(0.0f, 0.0f)));
    bool c4 = (d - edistance > 1e-6f) || (d <= 0);
    bool any_nan = isnan2(best.a) || isnan2(best.b);
    bool go = c4 && !any_nan && c1 && c2;
    if (!go) continue;  // lax.cond(..., lambda: x): the state is carried unchanged
    // body_fn :215-236
    V2 n2 = fast_normal(best.a - best.b);
    n2 = n2 / norm(n2);
    V2 np = minkowski_diff(A, B, n2);
    Edge ea, eb;
    ea.a = best.a;
    ea.b = np;
    eb.a = np;
    eb.b = best.b;
    edges[(size_t)best_idx] = ea;
    edges[(size_t)i + 3] = eb;
    int nb = closest_edge_to_origin(edges);
    prev = best;
    best = edges[(size_t)nb];
    best_idx = nb;
    new_point = np;
    i = i + 1;
    ++ran;
  }
  if (iters_out) *iters_out = ran;
  int fb = closest_edge_to_origin(edges);
  return closest_point_on_edge_to_point(edges[(size_t)fb].a, edges[(size_t)fb].b, mk(0.0f, 0.0f));
}

// ------------------------------------------------------------------------------------
// contacts — _contacts.py
// ------------------------------------------------------------------------------------
struct Contact {
  V2 pen;
  V2 cp;
  unsigned status;
  int gjk_iters, epa_iters;
};
inline Contact contact_nan() {  // ContactInfo.nan() :19-21
  Contact c;
  c.pen = mk(0.0f, 0.0f);
  c.cp = nan2();
  c.status = 0;
  c.gjk_iters = c.epa_iters = 0;
  return c;
}
inline bool contact_isnan(const Contact& c) { return isnan2(c.cp); }  // :23-24

Contact circle_vs_circle(const Shape& a, const Shape& b) {  // :30-58
  V2 apos = a.v[0], bpos = b.v[0];
  V2 delta = apos - bpos;
  float distance = norm(delta);
  V2 dir = (distance == 0.0f) ? mk(1.0f, 0.0f) : delta / distance;
  V2 penetration_vector = dir * jmin(distance - (a.radius + b.radius), 0.0f);
  V2 contact_point = (bpos + dir * (b.radius - a.radius) + apos) / 2.0f;
  if (!(dot(apos - contact_point, bpos - contact_point) <= 0)) {
    contact_point = contains(a, bpos) ? bpos : apos;
  }
  if (distance <= a.radius + b.radius) {
    Contact c = contact_nan();
    c.pen = -penetration_vector;
    c.cp = contact_point;
    return c;
  }
  return contact_nan();
}

Contact aabb_vs_aabb(const Shape& a, const Shape& b) {  // :61-96
  const float eps = 1e-8f;
  V2 al = a.v[0], au = a.v[1], bl = b.v[0], bu = b.v[1];
  bool below = au.y <= bl.y, above = al.y >= bu.y, left = au.x <= bl.x, right = al.x >= bu.x;
  if (!(below || left || above || right)) {
    float depths[4] = {jmax(au.y - bl.y, -eps), jmax(bu.y - al.y, -eps),
                       jmax(au.x - bl.x, -eps), jmax(bu.x - al.x, -eps)};
    const float dirs[4][2] = {{0, -1}, {0, 1}, {-1, 0}, {1, 0}};
    int index = jargmin(depths, 4);
    float min_depth = jmax(depths[index], 0.0f);  // jnp.clip(a_min=0.0)
    Contact c = contact_nan();
    c.pen = mk(min_depth * dirs[index][0], min_depth * dirs[index][1]);
    V2 min_upper = mk(jmin(au.x, bu.x), jmin(au.y, bu.y));
    V2 max_lower = mk(jmax(al.x, bl.x), jmax(al.y, bl.y));
    c.cp = (min_upper + max_lower) / 2.0f;
    return c;
  }
  return contact_nan();
}

Contact circle_vs_aabb(const Shape& a, const Shape& b) {  // :99-154
  const float eps = 1e-6f;
  V2 apos = a.v[0];
  V2 lower = b.v[0], upper = b.v[1];
  V2 bc = (lower + upper) / 2.0f;
  V2 disp = apos - bc;
  V2 lo = lower - bc, hi = upper - bc;
  V2 clamp_disp = mk(jclip(disp.x, lo.x, hi.x), jclip(disp.y, lo.y, hi.y));
  V2 ccp = bc + clamp_disp;
  unsigned status = 0;
  if (!contains(b, ccp)) status |= COTIX_ST_CCP_ERR;  // eqx.error_if :105-107
  V2 vs[4] = {lower, mk(lower.x, upper.y), upper, mk(upper.x, lower.y)};
  bool perfect_vertex = false;
  for (int k = 0; k < 4; ++k)
    if (norm(vs[k] - ccp) < eps) perfect_vertex = true;
  Contact c = contact_nan();
  c.status = status;
  if (contains(a, ccp)) {
    if (perfect_vertex) {  // circle_dir_move
      V2 dir = ccp - apos;
      V2 dir_norm = dir / norm(dir);
      c.pen = -(apos + a.radius * dir_norm - ccp);
      c.cp = ccp;
    } else {  // aligned_move
      const float dirs[4][2] = {{0, 1}, {0, -1}, {1, 0}, {-1, 0}};
      float shifts[4] = {apos.y + a.radius - lower.y, upper.y - (apos.y - a.radius),
                         apos.x + a.radius - lower.x, upper.x - (apos.x - a.radius)};
      int best = jargmin(shifts, 4);
      float ms = -shifts[best];
      c.pen = mk(ms * dirs[best][0], ms * dirs[best][1]);
      c.cp = ccp;
    }
  }
  return c;
}

Contact circle_vs_polygon(const Shape& circle, const Shape& polygon, V2 start, int gjk_max,
                          int epa_iters) {  // :157-202
  V2 simplex[3];
  int gi = 0, ei = 0;
  bool exists = check_for_collision_convex(circle, polygon, start, gjk_max, simplex, &gi);
  if (!exists) {  // EPA on the NaN simplex is computed and discarded by the reference
    Contact c = contact_nan();
    c.gjk_iters = gi;
    if (gi >= gjk_max) c.status |= COTIX_ST_GJK_CAP;
    return c;
  }
  V2 pen = get_closest_minkowski_diff(circle, polygon, simplex, epa_iters, &ei);
  Edge edges[COTIX_MAX_POLY];
  int ne = get_edges(polygon, edges);
  V2 point = circle.v[0];
  V2 disps[COTIX_MAX_POLY];
  float dists[COTIX_MAX_POLY];
  for (int k = 0; k < ne; ++k) {
    V2 a = edges[k].a, b = edges[k].b;
    bool both_zero = (a.x == 0.0f) && (a.y == 0.0f) && (b.x == 0.0f) && (b.y == 0.0f);
    if (both_zero) {
      disps[k] = mk(kInf, kInf);
    } else {
      V2 ab = a - b;
      float length = sumsq(ab);
      float t = dot(point - b, ab) / length;
      t = jclip(t, 0.0f, 1.0f);
      V2 projection = b + t * ab;
      disps[k] = point - projection;
    }
    dists[k] = disps[k].x * disps[k].x + disps[k].y * disps[k].y;
  }
  int mi = jargmin(dists, ne);
  V2 contact_point = point + disps[mi];
  if (dists[mi] > circle.radius * circle.radius) contact_point = point;
  Contact c = contact_nan();
  c.pen = pen;
  c.cp = contact_point;
  c.gjk_iters = gi;
  c.epa_iters = ei;
  if (gi >= gjk_max) c.status |= COTIX_ST_GJK_CAP;
  if (ei >= epa_iters) c.status |= COTIX_ST_EPA_CAP;
  return c;
}

// _contact_from_edges :205-267
V2 contact_from_edges(const Shape& A, const Shape& B) {
  Edge ea[COTIX_MAX_POLY], eb[COTIX_MAX_POLY];
  V2 va[COTIX_MAX_POLY], vb[COTIX_MAX_POLY];
  int na = get_edges(A, ea), nb = get_edges(B, eb);
  get_vertices(A, va);
  get_vertices(B, vb);
  float n = 0.0f;
  V2 avg = mk(0.0f, 0.0f);
  for (int k = 0; k < na; ++k)
    if (contains(B, va[k])) {
      avg = avg + va[k];
      n = n + 1.0f;
    }
  for (int k = 0; k < nb; ++k)
    if (contains(A, vb[k])) {
      avg = avg + vb[k];
      n = n + 1.0f;
    }
  for (int ib = 0; ib < nb; ++ib)
    for (int ia = 0; ia < na; ++ia) {
      V2 p = ea[ia].a;
      V2 r = ea[ia].b - ea[ia].a;
      V2 q = eb[ib].a;
      V2 s = eb[ib].b - eb[ib].a;
      auto crs = [](V2 a, V2 b) { return a.x * b.y - b.x * a.y; };
      float c = crs(r, s);
      float t = crs(q - p, s) / c;
      float u = crs(q - p, r) / c;
      if ((c != 0.0f) && (t >= 0.0f) && (t <= 1.0f) && (u >= 0.0f) && (u <= 1.0f)) {
        V2 x = p + r * t;
        if (!isnan2(x)) {
          avg = avg + x;
          n = n + 1.0f;
        }
      }
    }
  if (n > 0.0f) return avg / n;
  return nan2();
}

// aabb_vs_polygon :270-291 and polygon_vs_polygon :294-315
Contact convex_vs_polygon(const Shape& a, const Shape& b, V2 start, int gjk_max) {
  int na = (a.type == T_AABB) ? 4 : a.n;
  int iters = na + b.n + 1;
  if (iters > 48) iters = 48;
  V2 simplex[3];
  int gi = 0, ei = 0;
  bool exists = check_for_collision_convex(a, b, start, gjk_max, simplex, &gi);
  if (!exists) {
    Contact c = contact_nan();
    c.gjk_iters = gi;
    if (gi >= gjk_max) c.status |= COTIX_ST_GJK_CAP;
    return c;
  }
  Contact c = contact_nan();
  c.pen = get_closest_minkowski_diff(a, b, simplex, iters, &ei);
  c.cp = contact_from_edges(a, b);
  c.gjk_iters = gi;
  c.epa_iters = ei;
  if (gi >= gjk_max) c.status |= COTIX_ST_GJK_CAP;
  if (ei >= iters) c.status |= COTIX_ST_EPA_CAP;
  if (contact_isnan(c)) c.pen = c.pen;  // ContactInfo keeps pen; isnan() looks at cp only
  return c;
}

// dispatch of _contact_funcs, _colliders.py:21-35.  ok=false when the pair is not a key.
Contact contact_dispatch(const Shape& a, const Shape& b, V2 start, int gjk_max,
                         int epa_circle_poly, bool* ok) {
  *ok = true;
  if (a.type == T_CIRCLE && b.type == T_CIRCLE) return circle_vs_circle(a, b);
  if (a.type == T_AABB && b.type == T_AABB) return aabb_vs_aabb(a, b);
  if (a.type == T_CIRCLE && b.type == T_AABB) return circle_vs_aabb(a, b);
  if (a.type == T_CIRCLE && b.type == T_POLY)
    return circle_vs_polygon(a, b, start, gjk_max, epa_circle_poly);
  if (a.type == T_AABB && b.type == T_POLY) return convex_vs_polygon(a, b, start, gjk_max);
  if (a.type == T_POLY && b.type == T_POLY) return convex_vs_polygon(a, b, start, gjk_max);
  *ok = false;
  return contact_nan();
}

Shape shape_from_rec(const float* r) {
  Shape s;
  s.type = (int)r[0];
  s.n = (int)r[1];
  s.radius = r[2];
  for (int k = 0; k < COTIX_MAX_POLY; ++k) s.v[k] = mk(r[4 + 2 * k], r[5 + 2 * k]);
  if (s.type != T_POLY) s.n = 0;
  if (s.n > COTIX_MAX_POLY) s.n = COTIX_MAX_POLY;
  return s;
}

// ------------------------------------------------------------------------------------
// jax.random (non-partitionable threefry2x32), restated from the published algorithm
// ------------------------------------------------------------------------------------
inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
void threefry2x32(uint32_t k0, uint32_t k1, uint32_t x0, uint32_t x1, uint32_t* o0,
                  uint32_t* o1) {
  const int R0[4] = {13, 15, 26, 6}, R1[4] = {17, 29, 16, 24};
  uint32_t ks[3] = {k0, k1, k0 ^ k1 ^ 0x1BD11BDAu};
  x0 += ks[0];
  x1 += ks[1];
  for (int g = 1; g <= 5; ++g) {
    const int* R = (g & 1) ? R0 : R1;
    for (int r = 0; r < 4; ++r) {
      x0 += x1;
      x1 = rotl32(x1, R[r]);
      x1 ^= x0;
    }
    x0 += ks[g % 3];
    x1 += ks[(g + 1) % 3] + (uint32_t)g;
  }
  *o0 = x0;
  *o1 = x1;
}
struct Key {
  uint32_t a, b;
};
// random bits for a flat shape of n words: counts iota(n), zero-padded to even, split in halves
void random_bits(Key key, int n, uint32_t* out) {
  int npad = n + (n & 1);
  int half = npad / 2;
  for (int p = 0; p < half; ++p) {
    uint32_t c0 = (uint32_t)p;
    uint32_t c1 = (uint32_t)(half + p);
    if ((int)c1 >= n) c1 = 0;  // the padding element
    uint32_t o0, o1;
    threefry2x32(key.a, key.b, c0, c1, &o0, &o1);
    out[p] = o0;
    if (half + p < n) out[half + p] = o1;
  }
}
void split(Key key, int num, Key* out) {
  std::vector<uint32_t> bits((size_t)num * 2);
  random_bits(key, num * 2, bits.data());
  for (int k = 0; k < num; ++k) {
    out[k].a = bits[(size_t)2 * k];
    out[k].b = bits[(size_t)2 * k + 1];
  }
}
inline float bits_to_uniform(uint32_t bits, float lo, float hi) {
  uint32_t fb = (bits >> 9) | 0x3F800000u;
  float f;
  memcpy(&f, &fb, 4);
  f = f - 1.0f;
  float v = f * (hi - lo) + lo;
  return jmax(lo, v);
}
float uniform_scalar(Key key, float lo, float hi) {
  uint32_t b;
  random_bits(key, 1, &b);
  return bits_to_uniform(b, lo, hi);
}
void uniform_vec(Key key, int n, float lo, float hi, float* out) {
  std::vector<uint32_t> b((size_t)n);
  random_bits(key, n, b.data());
  for (int i = 0; i < n; ++i) out[i] = bits_to_uniform(b[(size_t)i], lo, hi);
}
inline bool bernoulli(Key key, float p) { return uniform_scalar(key, 0.0f, 1.0f) < p; }
// XLA's float32 erf_inv (Giles' single-precision approximation)
float erf_inv32(float x) {
  float w = -log1pf(-x * x);
  float p;
  if (w < 5.0f) {
    w = w - 2.5f;
    p = 2.81022636e-08f;
    p = 3.43273939e-07f + p * w;
    p = -3.5233877e-06f + p * w;
    p = -4.39150654e-06f + p * w;
    p = 0.00021858087f + p * w;
    p = -0.00125372503f + p * w;
    p = -0.00417768164f + p * w;
    p = 0.246640727f + p * w;
    p = 1.50140941f + p * w;
  } else {
    w = sqrtf(w) - 3.0f;
    p = -0.000200214257f;
    p = 0.000100950558f + p * w;
    p = 0.00134934322f + p * w;
    p = -0.00367342844f + p * w;
    p = 0.00573950773f + p * w;
    p = -0.0076224613f + p * w;
    p = 0.00943887047f + p * w;
    p = 1.00167406f + p * w;
    p = 2.83297682f + p * w;
  }
  if (fabsf(x) == 1.0f) return x * kInf;
  return p * x;
}
void normal_vec(Key key, int n, float* out) {
  float lo = nextafterf(-1.0f, 0.0f);
  uniform_vec(key, n, lo, 1.0f, out);
  for (int i = 0; i < n; ++i) out[i] = 1.41421356237309504880f * erf_inv32(out[i]);
}
// jr.choice(key, arange(n), p=probs) for shape (), given its uniform draw u
int choice_from_u(float u, const float* probs, int n) {
  std::vector<float> cum((size_t)n);
  float acc = 0.0f;
  for (int i = 0; i < n; ++i) {
    acc = (i == 0) ? probs[0] : acc + probs[i];
    cum[(size_t)i] = acc;
  }
  float r = cum[(size_t)n - 1] * (1.0f - u);
  int ind = 0;  // searchsorted side='left': number of entries strictly below r
  for (int i = 0; i < n; ++i)
    if (cum[(size_t)i] < r) ++ind;
  return ind;
}
int choice_p(Key key, const float* probs, int n) {
  return choice_from_u(uniform_scalar(key, 0.0f, 1.0f), probs, n);
}

// Philox4x32-10 (Random123) — used by the engine's throughput mode; restated here so the
// oracle can mirror that mode too.
void philox4x32_10(uint32_t k0, uint32_t k1, const uint32_t c[4], uint32_t out[4]) {
  uint32_t c0 = c[0], c1 = c[1], c2 = c[2], c3 = c[3];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n1 = lo1;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    uint32_t n3 = lo0;
    c0 = n0;
    c1 = n1;
    c2 = n2;
    c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0;
  out[1] = c1;
  out[2] = c2;
  out[3] = c3;
}

// ------------------------------------------------------------------------------------
// bodies, resolution — _bodies.py, _collision_resolution.py
// ------------------------------------------------------------------------------------
struct Body {  // DynamicBody leaves + parts
  float mass, inertia;
  V2 pos, vel;
  float angle, angvel;
  float elasticity, friction;
  std::vector<Shape> parts;  // local frame
};

inline void apply_impulse(Body& b, V2 impulse, V2 point) {  // :68-73
  V2 arm = point - b.pos;
  float torque = cross2(arm, impulse);
  b.vel = b.vel + impulse / b.mass;
  b.angvel = b.angvel + torque / b.inertia;
}
inline V2 velocity_at(const Body& b, V2 point) {  // _bodies.py:50-55
  return b.vel + fast_normal(point - b.pos) * b.angvel;
}

void resolve_collision(Body& b1, Body& b2, const Contact& ci, float baumgarte,
                       float baumgarte_dt) {  // :52-151
  if (contact_isnan(ci)) return;
  V2 cp = ci.cp;
  V2 v1 = b1.vel + fast_normal(cp - b1.pos) * b1.angvel;
  V2 v2 = b2.vel + fast_normal(cp - b2.pos) * b2.angvel;
  V2 vrel = v2 - v1;
  float pn = norm(ci.pen);
  V2 nd = ci.pen / pn;
  float vn = dot(vrel, nd);
  float elasticity = jmin(b1.elasticity, b2.elasticity);
  V2 r1 = cp - b1.pos, r2 = cp - b2.pos;
  float lever1 = sumsq(r1), lever2 = sumsq(r2);
  float ang = lever1 / b1.inertia + lever2 / b2.inertia;
  float nim = -(1.0f + elasticity) * vn - baumgarte * norm(ci.pen) / baumgarte_dt;
  float kden = 1.0f / b1.mass + 1.0f / b2.mass + ang;
  float normal_impulse = nim / kden;
  V2 impulse_vec = normal_impulse * nd;
  float friction = (b1.friction + b2.friction) / 2.0f;
  V2 vel_drag = vrel + vn * nd;
  float vdn = norm(vel_drag);
  V2 vel_drag_unit = vel_drag / vdn;
  float impulse_drag = -norm(vel_drag) / (1.0f / b1.mass + 1.0f / b2.mass + ang);
  impulse_drag = jclip(impulse_drag, 0.0f, normal_impulse * friction);
  V2 impulse_d_vec = impulse_drag * vel_drag_unit;
  impulse_vec = impulse_vec + impulse_d_vec;
  bool cond = dot(ci.pen, vrel) < 0;
  if (cond) return;
  apply_impulse(b1, -impulse_vec, cp);
  apply_impulse(b2, impulse_vec, cp);
}

// part.transform(body.shape._transformer): Circle :37-41, AABB :113-117, polygon :181-187
Shape transform_part(const Shape& local, V2 pos, float angle) {
  Shape w = local;
  if (local.type == T_CIRCLE) {
    w.v[0] = local.v[0] + pos;
  } else if (local.type == T_AABB) {
    w.v[0] = local.v[0] + pos;
    w.v[1] = local.v[1] + pos;
  } else {
    float c = o_cos(angle), s = o_sin(angle);
    float ms = -s;
    for (int k = 0; k < local.n; ++k) {
      // forward_vector: matrix @ [x, y, 1], then / transformed[2] (= 1)
      float x = local.v[k].x, y = local.v[k].y;
      float tx = c * x + ms * y + pos.x * 1.0f;
      float ty = s * x + c * y + pos.y * 1.0f;
      float tw = 0.0f * x + 0.0f * y + 1.0f * 1.0f;
      w.v[k] = mk(tx / tw, ty / tw);
    }
    order_clockwise(w.v, w.n);
  }
  return w;
}

// type label used as dictionary key in RandomizedCollider.resolve (Polygon4 and Polygon6
// are distinct classes, _convex_shapes.py:197-229)
inline int type_label(const Shape& s) {
  if (s.type == T_CIRCLE) return 0;
  if (s.type == T_AABB) return 1;
  return 2 + s.n;  // Polygon<n>
}
// is (t1, t2) a key of _contact_funcs (_colliders.py:21-35)?  Generic `Polygon` is
// represented by any polygon label that is not 4 or 6 vertices.
bool in_contact_funcs(int t1, int t2) {
  bool p1 = t1 >= 2, p2 = t2 >= 2;
  if (t1 == 1 && t2 == 1) return true;
  if (t1 == 0 && t2 == 0) return true;
  if (t1 == 0 && t2 == 1) return true;
  if (t1 == 0 && p2) return true;   // (Circle, Polygon|Polygon4|Polygon6) (others: generic)
  if (t1 == 1 && p2) return true;   // (AABB, Polygon|Polygon4|Polygon6)
  if (p1 && p2) {
    int n1 = t1 - 2, n2 = t2 - 2;
    if (n1 == 4 && n2 == 4) return true;
    if (n1 == 4 && n2 == 6) return true;
    if (n1 == 6 && n2 == 6) return true;
    bool g1 = (n1 != 4 && n1 != 6), g2 = (n2 != 4 && n2 != 6);
    if (g1 && g2) return true;  // (Polygon, Polygon)
    return false;
  }
  return false;
}

struct ColliderOut {
  int error;               // 1: illegal shape pair (RuntimeError in the reference)
  unsigned status;         // OR of narrowphase status bits
  bool cell[COTIX_SCENE_MAX_BODIES][COTIX_SCENE_MAX_BODIES];       // all_contacts[i][j] non-NaN after the coin flips
  bool candidate[COTIX_SCENE_MAX_BODIES][COTIX_SCENE_MAX_BODIES];  // some part pair of (i, j) gave a non-NaN contact
  int chosen[COTIX_SCENE_MAX_BODIES];                              // contact_index per body
};

// RandomizedCollider.resolve — _colliders.py:74-351.  Canonical order = insertion order
// (list(set(..)) at :116-120 is treated as the identity).
// RNG of the collider.  mode THREEFRY: the reference's key chain, literally.  mode PHILOX
// (engine-defined, same distributions): word(g, ind2, ind1) and the choice uniforms come
// from Philox4x32-10 counters (steps, episode, (g << 16) | (w >> 2), tag).
struct RngCtx {
  int mode;
  Key key;            // threefry: rkey;  philox: the Philox key
  uint32_t steps, episode;
};
inline uint32_t philox_word(const RngCtx& r, uint32_t g, uint32_t w, uint32_t tag) {
  uint32_t ctr[4] = {r.steps, r.episode, (g << 16) | (w >> 2), tag};
  uint32_t o[4];
  philox4x32_10(r.key.a, r.key.b, ctr, o);
  return o[w & 3u];
}
#define COTIX_TAG_COIN 0x434F494Eu
#define COTIX_TAG_CHOICE 0x43484F49u
#define COTIX_TAG_RESET 0x52535400u

void collider_resolve(std::vector<Body>& bodies, const RngCtx& rng, V2 start, int gjk_max,
                      float coin_p, float baumgarte, float baumgarte_dt, bool dedupe,
                      ColliderOut* out) {
  const Key rkey = rng.key;
  const int nb = (int)bodies.size();
  memset(out, 0, sizeof(*out));
  struct Ent {
    int body;
    int part;
    Shape world;
  };
  typedef std::pair<int, int> TK;
  std::vector<TK> order;
  std::map<TK, std::pair<std::vector<Ent>, std::vector<Ent>>> groups;
  for (int i = 0; i < nb; ++i)
    for (int j = 0; j < nb; ++j) {
      if (i <= j) continue;
      for (size_t pa = 0; pa < bodies[(size_t)i].parts.size(); ++pa) {
        Shape a = transform_part(bodies[(size_t)i].parts[pa], bodies[(size_t)i].pos,
                                 bodies[(size_t)i].angle);
        for (size_t pb = 0; pb < bodies[(size_t)j].parts.size(); ++pb) {
          Shape b = transform_part(bodies[(size_t)j].parts[pb], bodies[(size_t)j].pos,
                                   bodies[(size_t)j].angle);
          int t1 = type_label(a), t2 = type_label(b);
          if (in_contact_funcs(t1, t2)) {
          } else if (in_contact_funcs(t2, t1)) {
            // the reference swaps the labels but still appends (i, a) to list 0, which
            // later fails when the lists are stacked; treat as illegal
            out->error = 1;
            return;
          } else {
            out->error = 1;
            return;
          }
          TK key(t1, t2);
          if (groups.find(key) == groups.end()) order.push_back(key);
          Ent e0 = {i, (int)pa, a}, e1 = {j, (int)pb, b};
          groups[key].first.push_back(e0);
          groups[key].second.push_back(e1);
        }
      }
    }

  std::vector<Contact> all((size_t)nb * nb, contact_nan());
  Key sk[2];
  split(rkey, 2, sk);
  Key skey = sk[0];
  for (size_t g = 0; g < order.size(); ++g) {
    auto& lists = groups[order[g]];
    const std::vector<Ent>& l0 = lists.first;
    const std::vector<Ent>& l1 = lists.second;
    const int N1 = (int)l0.size(), N2 = (int)l1.size();
    std::vector<Contact> cur((size_t)N1 * N2);
    std::map<std::pair<std::pair<int, int>, std::pair<int, int>>, Contact> memo;
    for (int a = 0; a < N1; ++a)
      for (int b = 0; b < N2; ++b) {
        int i = l0[(size_t)a].body, j = l1[(size_t)b].body;
        Contact c;
        if (i < j) {
          c = contact_nan();  // jax.lax.cond(i < j, nan, contacts) :163
        } else {
          bool have = false;
          auto mk_ = std::make_pair(std::make_pair(i, l0[(size_t)a].part),
                                    std::make_pair(j, l1[(size_t)b].part));
          if (dedupe) {
            auto it = memo.find(mk_);
            if (it != memo.end()) {
              c = it->second;
              have = true;
            }
          }
          if (!have) {
            bool ok;
            c = contact_dispatch(l0[(size_t)a].world, l1[(size_t)b].world, start, gjk_max, 128,
                                 &ok);
            if (dedupe) memo[mk_] = c;
          }
          out->status |= c.status;
          if (!contact_isnan(c)) out->candidate[i][j] = true;
        }
        cur[(size_t)a * N2 + b] = c;
      }
    split(skey, 2, sk);
    skey = sk[0];
    std::vector<Key> keys2((size_t)N2), keys1((size_t)N1);
    split(skey, N2, keys2.data());
    for (int ind2 = 0; ind2 < N2; ++ind2) {
      split(keys2[(size_t)ind2], N1, keys1.data());
      for (int ind1 = 0; ind1 < N1; ++ind1) {
        const Contact& upd = cur[(size_t)ind1 * N2 + ind2];
        bool cond1;
        if (rng.mode == COTIX_RNG_THREEFRY) {
          Key kk[2];
          split(keys1[(size_t)ind1], 2, kk);
          cond1 = bernoulli(kk[0], coin_p);
        } else {
          uint32_t w = philox_word(rng, (uint32_t)g, (uint32_t)(ind2 * N1 + ind1), COTIX_TAG_COIN);
          cond1 = bits_to_uniform(w, 0.0f, 1.0f) < coin_p;
        }
        if (cond1 && !contact_isnan(upd)) {
          int i = l0[(size_t)ind1].body, j = l1[(size_t)ind2].body;
          all[(size_t)i * nb + j] = upd;
        }
      }
    }
  }
  // choose_random_contact :274-295
  std::vector<Key> ckeys((size_t)nb);
  split(skey, nb, ckeys.data());
  std::vector<int> chosen((size_t)nb);
  for (int i = 0; i < nb; ++i) {
    int cnt = 0;
    std::vector<float> probs((size_t)nb);
    for (int j = 0; j < nb; ++j) {
      bool good = !contact_isnan(all[(size_t)i * nb + j]);
      out->cell[i][j] = good;
      if (good) ++cnt;
    }
    for (int j = 0; j < nb; ++j)
      probs[(size_t)j] = (contact_isnan(all[(size_t)i * nb + j]) ? 0.0f : 1.0f) / (float)cnt;
    int idx;
    if (cnt == 0) {
      idx = i;
    } else if (rng.mode == COTIX_RNG_THREEFRY) {
      idx = choice_p(ckeys[(size_t)i], probs.data(), nb);
    } else {
      uint32_t w = philox_word(rng, 0u, (uint32_t)i, COTIX_TAG_CHOICE);
      idx = choice_from_u(bits_to_uniform(w, 0.0f, 1.0f), probs.data(), nb);
    }
    if (idx >= nb) idx = nb - 1;  // jnp indexing clamps
    chosen[(size_t)i] = idx;
    out->chosen[i] = idx;
  }
  // sequential resolution scan :310-336
  for (int i = 0; i < nb; ++i) {
    int j = chosen[(size_t)i];
    if (i == j) continue;
    resolve_collision(bodies[(size_t)i], bodies[(size_t)j], all[(size_t)i * nb + j], baumgarte,
                      baumgarte_dt);
  }
}

// ExplicitEulerPhysics._single_body_step — _physics_solvers.py:17-24
inline void euler_step(Body& b, float dt) {
  b.pos = b.pos + b.vel * dt;
  b.angle = b.angle + b.angvel * dt;
}

// ------------------------------------------------------------------------------------
// LunarLander — _lunar_lander.py
// ------------------------------------------------------------------------------------
void lunar_shapes(float lander[6][2], float legs[2][4][2]) {
  // :11-18, :30  Polygon6(jnp.array(LANDER_POLY) * 0.05) -> order_clockwise
  const int LP[6][2] = {{-14, 17}, {-17, 0}, {-17, -10}, {17, -10}, {17, 0}, {14, 17}};
  V2 lv[6];
  for (int k = 0; k < 6; ++k) lv[k] = mk((float)LP[k][0] * 0.05f, (float)LP[k][1] * 0.05f);
  order_clockwise(lv, 6);
  for (int k = 0; k < 6; ++k) {
    lander[k][0] = lv[k].x;
    lander[k][1] = lv[k].y;
  }
  // :32-41 Polygon4 of the integer rectangle -> order_clockwise on the integers
  const float LEG_W = 2, LEG_H = 8, LEG_ANGLE = -0.3f;
  V2 rv[4] = {mk(-LEG_W, -LEG_H), mk(+LEG_W, -LEG_H), mk(+LEG_W, LEG_H), mk(-LEG_W, LEG_H)};
  order_clockwise(rv, 4);
  // :44-60 vertices @ [[c, -s], [s, c]] (no re-sort: eqx.tree_at bypasses __init__)
  float c = o_cos(LEG_ANGLE), s = o_sin(LEG_ANGLE);
  V2 left[4];
  for (int k = 0; k < 4; ++k) {
    float x = rv[k].x * c + rv[k].y * s;
    float y = rv[k].x * (-s) + rv[k].y * c;
    left[k] = mk(x * 0.05f, y * 0.05f);  // :63-67
  }
  for (int k = 0; k < 4; ++k) {
    // right_leg_shape = left * [-1, 1] (:69-73); body 1 = right_leg, body 2 = left_leg
    legs[0][k][0] = left[k].x * -1.0f;
    legs[0][k][1] = left[k].y * 1.0f;
    legs[1][k][0] = left[k].x;
    legs[1][k][1] = left[k].y;
  }
}

void lunar_default_params(CotixLunarParams* p) {
  memset(p, 0, sizeof(*p));
  p->dt = 1e-2f;
  p->gravity_dv[0] = 0.0f;
  p->gravity_dv[1] = -0.002f;
  // random_direction(PRNGKey(1)) :_collisions.py:287-298, _geometry_utils.py:37-46
  {
    Key k1 = {0u, 1u};
    float nrm[2];
    normal_vec(k1, 2, nrm);
    V2 d = mk(nrm[0], nrm[1]);
    d = d / norm(d);
    p->init_dir[0] = d.x;
    p->init_dir[1] = d.y;
  }
  p->baumgarte = 0.3f;
  p->baumgarte_dt = 0.01f;
  p->coin_p = 0.5f;
  p->spring_kp = 1.0f;
  p->spring_kv0 = 0.1f;
  p->spring_kv = 0.05f;
  p->leg_damping = 0.95f;
  const float props[4][4] = {{30.0f, 30.0f, 1.0f, 0.1f},   // lander :78-87
                             {1.0f, 1.0f, 1.0f, 0.1f},     // right_leg :89-96
                             {1.0f, 1.0f, 1.0f, 0.1f},     // left_leg :98-105
                             {kInf, kInf, 0.1f, 0.1f}};    // ground :134-141
  memcpy(p->body_props, props, sizeof(props));
  lunar_shapes(p->lander_local, p->leg_local);
  const float LEG_AWAY = 24, LEG_DOWN = 8;
  // :146-166 lander-side anchors, in the order the constraints are applied (:193-204)
  p->joint_local[0][0] = LEG_AWAY * 0.05f;
  p->joint_local[0][1] = -LEG_DOWN * 0.05f;
  p->joint_local[1][0] = LEG_AWAY * 0.05f;
  p->joint_local[1][1] = (-LEG_DOWN + 8) * 0.05f;
  p->joint_local[2][0] = -LEG_AWAY * 0.05f;
  p->joint_local[2][1] = -LEG_DOWN * 0.05f;
  p->joint_local[3][0] = -LEG_AWAY * 0.05f;
  p->joint_local[3][1] = (-LEG_DOWN + 8) * 0.05f;
  p->leg_joint2_local[0] = 0.0f;
  p->leg_joint2_local[1] = 0.4f;
  p->lander_start[0] = 0.0f;
  p->lander_start[1] = 5.0f;
  p->lander_start[2] = 0.01f;
  // :75-76  array([+-LEG_AWAY, -LEG_DOWN]) * 0.05 (+ lander_center, added at reset)
  p->leg_offset[0][0] = -LEG_AWAY * 0.05f;
  p->leg_offset[0][1] = -LEG_DOWN * 0.05f;
  p->leg_offset[1][0] = LEG_AWAY * 0.05f;
  p->leg_offset[1][1] = -LEG_DOWN * 0.05f;
  p->ground_floor = -10.0f;
  p->terrain_x_outer = 100.0f;
  p->pad_half_width = 2.0f;
  p->pad_y = -2.0f;
  p->gjk_max_iters = 32;
  p->rng_mode = COTIX_RNG_THREEFRY;
  p->substeps = 1;
  p->auto_reset = 0;
  p->judge_enabled = 0;
  p->max_episode_steps = 0;
  p->reset_perturb = 0;
  p->main_engine_dv = 0.004f;
  p->side_engine_dv = 0.0005f;
  p->side_engine_dw = 0.002f;
  p->x_limit = 12.0f;
  p->y_limit = 20.0f;
  p->land_speed = 0.05f;
  p->land_angvel = 0.05f;
  p->reset_x[0] = -6.0f;
  p->reset_x[1] = 6.0f;
  p->reset_y[0] = -1.5f;
  p->reset_y[1] = 8.0f;
  p->reset_angle[0] = -0.5f;
  p->reset_angle[1] = 0.5f;
  p->reset_v = 1.7320508f;
  p->reset_w = 0.8660254f;
  p->philox_seed[0] = 0x243F6A88u;
  p->philox_seed[1] = 0x85A308D3u;
  p->solver_loops = 1;
}

// ---- the engine's env-logic spec (new; the reference has none): RNG streams ----------
// A "stream" hands out 32-bit words for reset perturbation.  Threefry mode derives them
// from the reset key; Philox mode from (seed, env, episode).
struct Env {
  float f[COTIX_LUNAR_NF];
};
inline uint32_t f2u(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
}
inline float u2f(uint32_t u) {
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// Reset: LunarLander.__init__ — _lunar_lander.py:29-143 (terrain from 5 sub-keys).
// `words`: in Philox mode, 12 terrain words + 8 perturbation words from the counter stream.
void lunar_reset_env(Env& e, Key key, const CotixLunarParams& P, int64_t env_index,
                     uint32_t episode) {
  float heights[8];
  float knots[4];
  uint32_t pert[8];
  if (P.rng_mode == COTIX_RNG_THREEFRY) {
    Key k[5];
    split(key, 5, k);
    uniform_vec(k[0], 8, -5.0f, 5.0f, heights);
    knots[0] = uniform_scalar(k[1], -12.0f, -9.0f);
    knots[1] = uniform_scalar(k[2], -8.0f, -4.0f);
    knots[2] = uniform_scalar(k[3], 4.0f, 8.0f);
    knots[3] = uniform_scalar(k[4], 9.0f, 12.0f);
    // perturbation words: threefry2x32(key, (0x70657274 + i, i)) — engine-defined stream
    for (int i = 0; i < 4; ++i) {
      uint32_t o0, o1;
      threefry2x32(key.a, key.b, 0x70657274u + (uint32_t)i, (uint32_t)i, &o0, &o1);
      pert[2 * i] = o0;
      pert[2 * i + 1] = o1;
    }
  } else {
    (void)env_index;
    uint32_t w[20];
    for (int q = 0; q < 5; ++q) {
      uint32_t ctr[4] = {0u, episode, (uint32_t)q, COTIX_TAG_RESET};
      philox4x32_10(P.philox_seed[0] ^ key.a, P.philox_seed[1] ^ key.b, ctr, w + 4 * q);
    }
    for (int i = 0; i < 8; ++i) heights[i] = bits_to_uniform(w[i], -5.0f, 5.0f);
    knots[0] = bits_to_uniform(w[8], -12.0f, -9.0f);
    knots[1] = bits_to_uniform(w[9], -8.0f, -4.0f);
    knots[2] = bits_to_uniform(w[10], 4.0f, 8.0f);
    knots[3] = bits_to_uniform(w[11], 9.0f, 12.0f);
    for (int i = 0; i < 8; ++i) pert[i] = w[12 + i];
  }
  heights[0] = heights[0] * 10.0f;
  heights[3] = P.pad_y;
  heights[4] = P.pad_y;
  heights[7] = heights[7] * 10.0f;
  for (int i = 0; i < 8; ++i) e.f[COTIX_F_HEIGHT0 + i] = heights[i];
  for (int i = 0; i < 4; ++i) e.f[COTIX_F_KNOT0 + i] = knots[i];

  float lx = P.lander_start[0], ly = P.lander_start[1], la = P.lander_start[2];
  float lvx = 0.0f, lvy = 0.0f, lw = 0.0f;
  V2 legp[2];
  float lega[2] = {0.0f, 0.0f};
  V2 legv[2] = {mk(0.0f, 0.0f), mk(0.0f, 0.0f)};
  float legw[2] = {0.0f, 0.0f};
  if (!P.reset_perturb) {
    for (int l = 0; l < 2; ++l) legp[l] = mk(P.leg_offset[l][0] + lx, P.leg_offset[l][1] + ly);
  } else {
    lx = bits_to_uniform(pert[0], P.reset_x[0], P.reset_x[1]);
    ly = bits_to_uniform(pert[1], P.reset_y[0], P.reset_y[1]);
    la = bits_to_uniform(pert[2], P.reset_angle[0], P.reset_angle[1]);
    lvx = bits_to_uniform(pert[3], -P.reset_v, P.reset_v);
    lvy = bits_to_uniform(pert[4], -P.reset_v, P.reset_v);
    lw = bits_to_uniform(pert[5], -P.reset_w, P.reset_w);
    // legs placed rigidly: leg angle = lander angle - start angle, offset rotated with it
    float da = la - P.lander_start[2];
    float c, s;
    if (g_trig_mode == 0) {
      c = cosf(da);
      s = sinf(da);
    } else {
      cx_sincosf(da, &s, &c);
    }
    for (int l = 0; l < 2; ++l) {
      float ox = P.leg_offset[l][0], oy = P.leg_offset[l][1];
      float rx = c * ox - s * oy, ry = s * ox + c * oy;
      legp[l] = mk(rx + lx, ry + ly);
      lega[l] = da;
      legv[l] = mk(lvx + (-ry) * lw, lvy + rx * lw);
      legw[l] = lw;
    }
  }
  e.f[COTIX_F_BODY(0, COTIX_F_PX)] = lx;
  e.f[COTIX_F_BODY(0, COTIX_F_PY)] = ly;
  e.f[COTIX_F_BODY(0, COTIX_F_VX)] = lvx;
  e.f[COTIX_F_BODY(0, COTIX_F_VY)] = lvy;
  e.f[COTIX_F_BODY(0, COTIX_F_ANG)] = la;
  e.f[COTIX_F_BODY(0, COTIX_F_ANGVEL)] = lw;
  for (int l = 0; l < 2; ++l) {
    int b = 1 + l;
    e.f[COTIX_F_BODY(b, COTIX_F_PX)] = legp[l].x;
    e.f[COTIX_F_BODY(b, COTIX_F_PY)] = legp[l].y;
    e.f[COTIX_F_BODY(b, COTIX_F_VX)] = legv[l].x;
    e.f[COTIX_F_BODY(b, COTIX_F_VY)] = legv[l].y;
    e.f[COTIX_F_BODY(b, COTIX_F_ANG)] = lega[l];
    e.f[COTIX_F_BODY(b, COTIX_F_ANGVEL)] = legw[l];
  }
  e.f[COTIX_F_STEPS] = u2f(0u);
  e.f[COTIX_F_STATUS] = u2f(episode << 8);  // sticky flags cleared, episode index kept
  e.f[COTIX_F_RETURN] = 0.0f;
  e.f[COTIX_F_SHAPING] = kNaN;
}

void build_bodies(const Env& e, const CotixLunarParams& P, std::vector<Body>& bodies) {
  bodies.assign(4, Body());
  for (int b = 0; b < 4; ++b) {
    Body& B = bodies[(size_t)b];
    B.mass = P.body_props[b][0];
    B.inertia = P.body_props[b][1];
    B.elasticity = P.body_props[b][2];
    B.friction = P.body_props[b][3];
    if (b < 3) {
      B.pos = mk(e.f[COTIX_F_BODY(b, COTIX_F_PX)], e.f[COTIX_F_BODY(b, COTIX_F_PY)]);
      B.vel = mk(e.f[COTIX_F_BODY(b, COTIX_F_VX)], e.f[COTIX_F_BODY(b, COTIX_F_VY)]);
      B.angle = e.f[COTIX_F_BODY(b, COTIX_F_ANG)];
      B.angvel = e.f[COTIX_F_BODY(b, COTIX_F_ANGVEL)];
    } else {
      bool poisoned = (f2u(e.f[COTIX_F_STATUS]) & COTIX_ST_GROUND_NAN) != 0;
      float z = poisoned ? kNaN : 0.0f;
      B.pos = mk(z, z);
      B.vel = mk(z, z);
      B.angle = z;
      B.angvel = z;
    }
  }
  Shape ls;
  ls.type = T_POLY;
  ls.n = 6;
  ls.radius = 0.0f;
  for (int k = 0; k < 6; ++k) ls.v[k] = mk(P.lander_local[k][0], P.lander_local[k][1]);
  bodies[0].parts.push_back(ls);
  for (int l = 0; l < 2; ++l) {
    Shape s;
    s.type = T_POLY;
    s.n = 4;
    s.radius = 0.0f;
    for (int k = 0; k < 4; ++k) s.v[k] = mk(P.leg_local[l][k][0], P.leg_local[l][k][1]);
    for (int k = 4; k < 6; ++k) s.v[k] = mk(0.0f, 0.0f);
    bodies[(size_t)1 + l].parts.push_back(s);
  }
  // ground: 7 Polygon4 trapezoids :116-132
  float xs[8] = {-P.terrain_x_outer, e.f[COTIX_F_KNOT0 + 0], e.f[COTIX_F_KNOT0 + 1],
                 -P.pad_half_width,  P.pad_half_width,       e.f[COTIX_F_KNOT0 + 2],
                 e.f[COTIX_F_KNOT0 + 3], P.terrain_x_outer};
  for (int i = 0; i < 7; ++i) {
    Shape s;
    s.type = T_POLY;
    s.n = 4;
    s.radius = 0.0f;
    s.v[0] = mk(xs[i], e.f[COTIX_F_HEIGHT0 + i]);
    s.v[1] = mk(xs[i], P.ground_floor);
    s.v[2] = mk(xs[i + 1], e.f[COTIX_F_HEIGHT0 + i + 1]);
    s.v[3] = mk(xs[i + 1], P.ground_floor);
    s.v[4] = s.v[5] = mk(0.0f, 0.0f);
    order_clockwise(s.v, 4);  // Polygon4.__init__
    bodies[3].parts.push_back(s);
  }
}

// LunarLander.step — _lunar_lander.py:145-218
void lunar_joints(std::vector<Body>& bodies, const CotixLunarParams& P) {
  Body &lander = bodies[0], &right_leg = bodies[1], &left_leg = bodies[2];
  V2 lj[4];
  for (int k = 0; k < 4; ++k)
    lj[k] = rotate(mk(P.joint_local[k][0], P.joint_local[k][1]), lander.angle) + lander.pos;
  V2 left_leg_joint_1 = left_leg.pos;
  V2 left_leg_joint_2 =
      left_leg.pos + rotate(mk(P.leg_joint2_local[0], P.leg_joint2_local[1]), left_leg.angle);
  V2 right_leg_joint_1 = right_leg.pos;
  V2 right_leg_joint_2 =
      right_leg.pos + rotate(mk(P.leg_joint2_local[0], P.leg_joint2_local[1]), right_leg.angle);
  auto constraint = [&](Body& b1, V2 c1, Body& b2, V2 c2) {
    V2 delta_pos = c1 - c2;
    V2 delta_vel = velocity_at(b1, c1) - velocity_at(b2, c2);
    V2 impulse = delta_pos * P.spring_kp + delta_vel * (norm(delta_vel) + P.spring_kv0) * P.spring_kv;
    apply_impulse(b1, -impulse, c1);
    apply_impulse(b2, impulse, c2);
  };
  // SimpleConstraintSolver.solve (_constraint_solvers.py:13-17): `loops` passes over the list
  for (int it = 0; it < (P.solver_loops < 1 ? 1 : P.solver_loops); ++it) {
    constraint(lander, lj[0], left_leg, left_leg_joint_1);
    constraint(lander, lj[1], left_leg, left_leg_joint_2);
    constraint(lander, lj[2], right_leg, right_leg_joint_1);
    constraint(lander, lj[3], right_leg, right_leg_joint_2);
  }
  right_leg.angvel = right_leg.angvel * P.leg_damping;
  left_leg.angvel = left_leg.angvel * P.leg_damping;
}

// shaping potential of the judge (new env logic; see DESIGN.md "Judge")
float lunar_shaping(const Env& e, const CotixLunarParams& P, bool rc, bool lc) {
  float x = e.f[COTIX_F_BODY(0, COTIX_F_PX)], y = e.f[COTIX_F_BODY(0, COTIX_F_PY)];
  float vx = e.f[COTIX_F_BODY(0, COTIX_F_VX)], vy = e.f[COTIX_F_BODY(0, COTIX_F_VY)];
  float a = e.f[COTIX_F_BODY(0, COTIX_F_ANG)];
  float dy = y - P.pad_y;
  float dist = sqrtf(x * x + dy * dy);
  float speed = sqrtf(vx * vx + vy * vy);
  float s = -10.0f * dist;
  s = s - 100.0f * speed;
  s = s - 100.0f * fabsf(a);
  s = s + (rc ? 10.0f : 0.0f);
  s = s + (lc ? 10.0f : 0.0f);
  return s;
}

struct StepInfo {
  float reward;
  int done;
  bool contact_right, contact_left;
};

// One physics sub-step = f(env, key) of examples/test_viz.py:25-44, plus engines + judge.
void lunar_substep(Env& e, float a_main, float a_side, const CotixLunarParams& P, bool dedupe,
                   StepInfo* info) {
  std::vector<Body> bodies;
  build_bodies(e, P, bodies);
  // physics.step(env.bodies, dt) :26
  for (auto& b : bodies) euler_step(b, P.dt);
  for (int b = 0; b < 3; ++b)
    if (!(fabsf(bodies[(size_t)b].angle) < 1.0e9f)) e.f[COTIX_F_STATUS] = u2f(f2u(e.f[COTIX_F_STATUS]) | COTIX_ST_ANGLE_RANGE);
  // gravity on the lander only :27-31
  bodies[0].vel = bodies[0].vel + mk(P.gravity_dv[0], P.gravity_dv[1]);
  // engines (new): zero action leaves the step bit-identical to the reference
  float thr_main = jclip(a_main, 0.0f, 1.0f);
  float thr_side = jclip(a_side, -1.0f, 1.0f);
  if (thr_main != 0.0f || thr_side != 0.0f) {
    float c = o_cos(bodies[0].angle), s = o_sin(bodies[0].angle);
    if (thr_main != 0.0f) {
      float dv = P.main_engine_dv * thr_main;
      bodies[0].vel = bodies[0].vel + mk((-s) * dv, c * dv);
    }
    if (thr_side != 0.0f) {
      float dv = P.side_engine_dv * thr_side;
      bodies[0].vel = bodies[0].vel + mk(c * dv, s * dv);
      bodies[0].angvel = bodies[0].angvel - P.side_engine_dw * thr_side;
    }
  }
  // collider.resolve(new_bodies, key) :37
  Key key = {f2u(e.f[COTIX_F_KEY0]), f2u(e.f[COTIX_F_KEY1])};
  uint32_t steps = f2u(e.f[COTIX_F_STEPS]);
  uint32_t status = f2u(e.f[COTIX_F_STATUS]);
  RngCtx rng;
  rng.mode = P.rng_mode;
  rng.steps = steps;
  rng.episode = status >> 8;
  rng.key = key;
  if (P.rng_mode == COTIX_RNG_PHILOX) {
    rng.key.a = P.philox_seed[0] ^ key.a;
    rng.key.b = P.philox_seed[1] ^ key.b;
  }
  ColliderOut co;
  collider_resolve(bodies, rng, mk(P.init_dir[0], P.init_dir[1]), P.gjk_max_iters, P.coin_p,
                   P.baumgarte, P.baumgarte_dt, dedupe, &co);
  // key, next_key = jr.split(key) :39
  if (P.rng_mode == COTIX_RNG_THREEFRY) {
    Key ks[2];
    split(key, 2, ks);
    key = ks[0];
  }
  // env.step() :41
  lunar_joints(bodies, P);
  // write back
  for (int b = 0; b < 3; ++b) {
    const Body& B = bodies[(size_t)b];
    e.f[COTIX_F_BODY(b, COTIX_F_PX)] = B.pos.x;
    e.f[COTIX_F_BODY(b, COTIX_F_PY)] = B.pos.y;
    e.f[COTIX_F_BODY(b, COTIX_F_VX)] = B.vel.x;
    e.f[COTIX_F_BODY(b, COTIX_F_VY)] = B.vel.y;
    e.f[COTIX_F_BODY(b, COTIX_F_ANG)] = B.angle;
    e.f[COTIX_F_BODY(b, COTIX_F_ANGVEL)] = B.angvel;
  }
  {
    const Body& G = bodies[3];
    bool bad = (G.vel.x != 0.0f) || (G.vel.y != 0.0f) || (G.angvel != 0.0f) ||
               (G.pos.x != 0.0f) || (G.pos.y != 0.0f) || (G.angle != 0.0f);
    if (bad) status |= COTIX_ST_GROUND_NAN;
  }
  bool any_nan = false;
  for (int k = 0; k < 18; ++k)
    if (!(fabsf(e.f[k]) <= 3.0e38f)) any_nan = true;
  if (any_nan) status |= COTIX_ST_NAN;
  steps += 1u;
  e.f[COTIX_F_KEY0] = u2f(key.a);
  e.f[COTIX_F_KEY1] = u2f(key.b);
  e.f[COTIX_F_STEPS] = u2f(steps);
  e.f[COTIX_F_STATUS] = u2f(status);

  // judge (new).  Contacts are the narrowphase results BEFORE the coin flips.
  bool rc = co.candidate[3][1], lc = co.candidate[3][2], crash = co.candidate[3][0];
  info->contact_right = rc;
  info->contact_left = lc;
  float reward = 0.0f;
  int done = COTIX_DONE_NO;
  if (P.judge_enabled) {
    float sh = lunar_shaping(e, P, rc, lc);
    float prev = e.f[COTIX_F_SHAPING];
    if (prev == prev) reward = sh - prev;
    e.f[COTIX_F_SHAPING] = sh;
    reward = reward - 0.3f * thr_main;
    reward = reward - 0.03f * fabsf(thr_side);
    float x = e.f[COTIX_F_BODY(0, COTIX_F_PX)], y = e.f[COTIX_F_BODY(0, COTIX_F_PY)];
    float vx = e.f[COTIX_F_BODY(0, COTIX_F_VX)], vy = e.f[COTIX_F_BODY(0, COTIX_F_VY)];
    float w = e.f[COTIX_F_BODY(0, COTIX_F_ANGVEL)];
    if (crash) {
      done = COTIX_DONE_CRASH;
      reward = -100.0f;
    } else if (!(fabsf(x) < P.x_limit) || !(y < P.y_limit)) {
      done = COTIX_DONE_OOB;
      reward = -100.0f;
    } else if (rc && lc && sqrtf(vx * vx + vy * vy) < P.land_speed && fabsf(w) < P.land_angvel) {
      done = COTIX_DONE_LANDED;
      reward = 100.0f;
    }
  }
  if (any_nan) {
    done = COTIX_DONE_NAN;
    reward = P.judge_enabled ? -100.0f : 0.0f;
  }
  if (done == COTIX_DONE_NO && P.max_episode_steps > 0 && steps >= (uint32_t)P.max_episode_steps)
    done = COTIX_DONE_TIMEOUT;
  e.f[COTIX_F_RETURN] = e.f[COTIX_F_RETURN] + reward;
  info->reward = reward;
  info->done = done;
}

void lunar_step_env(Env& e, float a_main, float a_side, const CotixLunarParams& P, bool dedupe,
                    int64_t env_index, float* reward, uint8_t* done, float* obs, int64_t ld_obs) {
  float rsum = 0.0f;
  int d = COTIX_DONE_NO;
  bool rc = false, lc = false;
  int K = P.substeps < 1 ? 1 : P.substeps;
  for (int k = 0; k < K; ++k) {
    StepInfo si;
    lunar_substep(e, a_main, a_side, P, dedupe, &si);
    rsum = rsum + si.reward;
    rc = si.contact_right;
    lc = si.contact_left;
    if (si.done != COTIX_DONE_NO) {  // freeze on done (_envs.py:81-103)
      d = si.done;
      break;
    }
  }
  if (reward) *reward = rsum;
  if (done) *done = (uint8_t)d;
  if (d != COTIX_DONE_NO && P.auto_reset) {
    // auto-reset (new): threefry mode consumes the sibling key of the step chain,
    // key, reset_key = split(key); Philox mode keys the reset by the episode counter.
    Key key = {f2u(e.f[COTIX_F_KEY0]), f2u(e.f[COTIX_F_KEY1])};
    Key rk = key;
    uint32_t episode = ((f2u(e.f[COTIX_F_STATUS]) >> 8) + 1u) & 0xFFFFFFu;
    if (P.rng_mode == COTIX_RNG_THREEFRY) {
      Key ks[2];
      split(key, 2, ks);
      key = ks[0];
      rk = ks[1];
    }
    lunar_reset_env(e, rk, P, env_index, episode);
    e.f[COTIX_F_KEY0] = u2f(key.a);
    e.f[COTIX_F_KEY1] = u2f(key.b);
    rc = lc = false;
  }
  if (obs) {  // observation of the state the NEXT action applies to
    for (int k = 0; k < 6; ++k) obs[k * ld_obs] = e.f[COTIX_F_BODY(0, k)];
    obs[6 * ld_obs] = rc ? 1.0f : 0.0f;
    obs[7 * ld_obs] = lc ? 1.0f : 0.0f;
  }
}

thread_local char g_err[256] = "";

}  // namespace

// ======================================================================================
// C API (loaded through ctypes by tests/ and bench.py)
// ======================================================================================
extern "C" {

int oracle_set_trig_mode(int mode) {
  g_trig_mode = mode ? 1 : 0;
  return 0;
}

int oracle_narrowphase(int64_t n, const float* a, const float* b, uint8_t* exists, float* pen,
                       float* cp, uint8_t* status, int32_t* iters,
                       const CotixNarrowParams* params) {
  V2 start = mk(params->init_dir[0], params->init_dir[1]);
#pragma omp parallel for schedule(dynamic, 256)
  for (int64_t k = 0; k < n; ++k) {
    Shape A = shape_from_rec(a + k * COTIX_SHAPE_REC_FLOATS);
    Shape B = shape_from_rec(b + k * COTIX_SHAPE_REC_FLOATS);
    bool ok;
    Contact c = contact_dispatch(A, B, start, params->gjk_max_iters,
                                 params->epa_iters_circle_poly, &ok);
    bool ex = !contact_isnan(c);
    if (exists) exists[k] = ex ? 1 : 0;
    if (pen) {
      pen[2 * k] = c.pen.x;
      pen[2 * k + 1] = c.pen.y;
    }
    if (cp) {
      cp[2 * k] = c.cp.x;
      cp[2 * k + 1] = c.cp.y;
    }
    if (status) status[k] = (uint8_t)(c.status | (ok ? 0u : 0x80u));
    if (iters) {
      iters[2 * k] = c.gjk_iters;
      iters[2 * k + 1] = c.epa_iters;
    }
  }
  return 0;
}

int oracle_order_clockwise(int64_t n, float* recs) {
  for (int64_t k = 0; k < n; ++k) {
    float* r = recs + k * COTIX_SHAPE_REC_FLOATS;
    if ((int)r[0] != T_POLY) continue;
    int nv = (int)r[1];
    V2 v[COTIX_MAX_POLY];
    for (int q = 0; q < nv; ++q) v[q] = mk(r[4 + 2 * q], r[5 + 2 * q]);
    order_clockwise(v, nv);
    for (int q = 0; q < nv; ++q) {
      r[4 + 2 * q] = v[q].x;
      r[5 + 2 * q] = v[q].y;
    }
  }
  return 0;
}

int oracle_support(int64_t n, const float* recs, const float* dirs, float* out) {
  for (int64_t k = 0; k < n; ++k) {
    Shape s = shape_from_rec(recs + k * COTIX_SHAPE_REC_FLOATS);
    V2 r = support(s, mk(dirs[2 * k], dirs[2 * k + 1]));
    out[2 * k] = r.x;
    out[2 * k + 1] = r.y;
  }
  return 0;
}

int oracle_contains(int64_t n, const float* recs, const float* pts, uint8_t* out) {
  for (int64_t k = 0; k < n; ++k) {
    Shape s = shape_from_rec(recs + k * COTIX_SHAPE_REC_FLOATS);
    out[k] = contains(s, mk(pts[2 * k], pts[2 * k + 1])) ? 1 : 0;
  }
  return 0;
}

int oracle_gjk(int64_t n, const float* a, const float* b, const CotixNarrowParams* params,
               uint8_t* exists, float* simplex /*[n][6]*/) {
  V2 start = mk(params->init_dir[0], params->init_dir[1]);
  for (int64_t k = 0; k < n; ++k) {
    Shape A = shape_from_rec(a + k * COTIX_SHAPE_REC_FLOATS);
    Shape B = shape_from_rec(b + k * COTIX_SHAPE_REC_FLOATS);
    V2 s[3];
    bool ex = check_for_collision_convex(A, B, start, params->gjk_max_iters, s, nullptr);
    exists[k] = ex;
    for (int q = 0; q < 3; ++q) {
      simplex[6 * k + 2 * q] = s[q].x;
      simplex[6 * k + 2 * q + 1] = s[q].y;
    }
  }
  return 0;
}

int oracle_euler_step(int64_t n, float* px, float* py, const float* vx, const float* vy,
                      float* ang, const float* angvel, float dt) {
  for (int64_t k = 0; k < n; ++k) {
    px[k] = px[k] + vx[k] * dt;
    py[k] = py[k] + vy[k] * dt;
    ang[k] = ang[k] + angvel[k] * dt;
  }
  return 0;
}

int oracle_resolve_collision(int64_t n, float* body1, float* body2, const float* pen,
                             const float* cp, float baumgarte, float baumgarte_dt) {
  for (int64_t k = 0; k < n; ++k) {
    Body b[2];
    float* src[2] = {body1 + 10 * k, body2 + 10 * k};
    for (int q = 0; q < 2; ++q) {
      b[q].mass = src[q][0];
      b[q].inertia = src[q][1];
      b[q].pos = mk(src[q][2], src[q][3]);
      b[q].vel = mk(src[q][4], src[q][5]);
      b[q].angle = src[q][6];
      b[q].angvel = src[q][7];
      b[q].elasticity = src[q][8];
      b[q].friction = src[q][9];
    }
    Contact c = contact_nan();
    c.pen = mk(pen[2 * k], pen[2 * k + 1]);
    c.cp = mk(cp[2 * k], cp[2 * k + 1]);
    resolve_collision(b[0], b[1], c, baumgarte, baumgarte_dt);
    for (int q = 0; q < 2; ++q) {
      src[q][4] = b[q].vel.x;
      src[q][5] = b[q].vel.y;
      src[q][7] = b[q].angvel;
    }
  }
  return 0;
}

// ---- RNG ----
int oracle_threefry2x32(int64_t n, uint32_t k0, uint32_t k1, const uint32_t* x0,
                        const uint32_t* x1, uint32_t* out) {
  for (int64_t i = 0; i < n; ++i) threefry2x32(k0, k1, x0[i], x1[i], &out[2 * i], &out[2 * i + 1]);
  return 0;
}
int oracle_philox4x32(int64_t n, uint32_t k0, uint32_t k1, const uint32_t* ctr, uint32_t* out) {
  for (int64_t i = 0; i < n; ++i) philox4x32_10(k0, k1, ctr + 4 * i, out + 4 * i);
  return 0;
}
int oracle_split(uint32_t k0, uint32_t k1, int num, uint32_t* out) {
  std::vector<Key> ks((size_t)num);
  Key k = {k0, k1};
  split(k, num, ks.data());
  for (int i = 0; i < num; ++i) {
    out[2 * i] = ks[(size_t)i].a;
    out[2 * i + 1] = ks[(size_t)i].b;
  }
  return 0;
}
int oracle_random_bits(uint32_t k0, uint32_t k1, int n, uint32_t* out) {
  Key k = {k0, k1};
  random_bits(k, n, out);
  return 0;
}
int oracle_uniform(uint32_t k0, uint32_t k1, int n, float lo, float hi, float* out) {
  Key k = {k0, k1};
  uniform_vec(k, n, lo, hi, out);
  return 0;
}
int oracle_normal(uint32_t k0, uint32_t k1, int n, float* out) {
  Key k = {k0, k1};
  normal_vec(k, n, out);
  return 0;
}
int oracle_bernoulli(uint32_t k0, uint32_t k1, float p) {
  Key k = {k0, k1};
  return bernoulli(k, p) ? 1 : 0;
}
int oracle_choice(uint32_t k0, uint32_t k1, const float* probs, int n) {
  Key k = {k0, k1};
  return choice_p(k, probs, n);
}

// ---- LunarLander ----
int oracle_lunar_default_params(CotixLunarParams* p) {
  lunar_default_params(p);
  return 0;
}

int oracle_lunar_reset(int64_t B, const uint32_t* keys, float* state, int64_t ld,
                       const CotixLunarParams* p) {
#pragma omp parallel for
  for (int64_t e = 0; e < B; ++e) {
    Env env;
    Key key = {keys[2 * e], keys[2 * e + 1]};
    lunar_reset_env(env, key, *p, e, 0u);
    env.f[COTIX_F_KEY0] = u2f(key.a);
    env.f[COTIX_F_KEY1] = u2f(key.b);
    for (int f = 0; f < COTIX_LUNAR_NF; ++f) state[f * ld + e] = env.f[f];
  }
  return 0;
}

int oracle_lunar_step(int64_t B, const float* state_in, float* state_out, int64_t ld,
                      const float* action, float* reward, uint8_t* done, float* obs,
                      const CotixLunarParams* p, int dedupe) {
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t e = 0; e < B; ++e) {
    Env env;
    for (int f = 0; f < COTIX_LUNAR_NF; ++f) env.f[f] = state_in[f * ld + e];
    float am = action ? action[e] : 0.0f;
    float as = action ? action[ld + e] : 0.0f;
    lunar_step_env(env, am, as, *p, dedupe != 0, e, reward ? reward + e : nullptr,
                   done ? done + e : nullptr, obs ? obs + e : nullptr, ld);
    for (int f = 0; f < COTIX_LUNAR_NF; ++f) state_out[f * ld + e] = env.f[f];
  }
  return 0;
}

// The collider alone, on the LunarLander body set, with its bookkeeping exposed:
// cell[e][16] (row-major 4x4 all_contacts non-NaN), candidate[e][16], chosen[e][4].
int oracle_lunar_collider_debug(int64_t B, const float* state, int64_t ld,
                                const CotixLunarParams* p, int dedupe, uint8_t* cell,
                                uint8_t* candidate, int32_t* chosen) {
  for (int64_t e = 0; e < B; ++e) {
    Env env;
    for (int f = 0; f < COTIX_LUNAR_NF; ++f) env.f[f] = state[f * ld + e];
    std::vector<Body> bodies;
    build_bodies(env, *p, bodies);
    Key key = {f2u(env.f[COTIX_F_KEY0]), f2u(env.f[COTIX_F_KEY1])};
    RngCtx rng;
    rng.mode = p->rng_mode;
    rng.steps = f2u(env.f[COTIX_F_STEPS]);
    rng.episode = f2u(env.f[COTIX_F_STATUS]) >> 8;
    rng.key = key;
    if (p->rng_mode == COTIX_RNG_PHILOX) {
      rng.key.a = p->philox_seed[0] ^ key.a;
      rng.key.b = p->philox_seed[1] ^ key.b;
    }
    ColliderOut co;
    collider_resolve(bodies, rng, mk(p->init_dir[0], p->init_dir[1]), p->gjk_max_iters,
                     p->coin_p, p->baumgarte, p->baumgarte_dt, dedupe != 0, &co);
    for (int i = 0; i < 4; ++i) {
      for (int j = 0; j < 4; ++j) {
        cell[e * 16 + i * 4 + j] = co.cell[i][j];
        candidate[e * 16 + i * 4 + j] = co.candidate[i][j];
      }
      chosen[e * 4 + i] = co.chosen[i];
    }
  }
  return 0;
}

// Generic scene step for arbitrary bodies (used for the RoboCup scene and small hand-made
// scenes in tests): bodies [nb][10] DynamicBody leaves, parts given as records with a body
// index each.  Runs physics.step -> collider.resolve (-> nothing else).
int oracle_scene_step(int nb, float* bodies10, int nparts, const int32_t* part_body,
                      const float* part_recs, uint32_t k0, uint32_t k1, float dt,
                      const float* init_dir, int gjk_max, float coin_p, float baumgarte,
                      float baumgarte_dt, int dedupe, uint8_t* cell, int32_t* chosen,
                      uint32_t* status) {
  if (nb > COTIX_SCENE_MAX_BODIES) return -1;
  std::vector<Body> bodies((size_t)nb);
  for (int b = 0; b < nb; ++b) {
    float* s = bodies10 + 10 * b;
    Body& B = bodies[(size_t)b];
    B.mass = s[0];
    B.inertia = s[1];
    B.pos = mk(s[2], s[3]);
    B.vel = mk(s[4], s[5]);
    B.angle = s[6];
    B.angvel = s[7];
    B.elasticity = s[8];
    B.friction = s[9];
  }
  for (int q = 0; q < nparts; ++q)
    bodies[(size_t)part_body[q]].parts.push_back(shape_from_rec(part_recs + 16 * q));
  for (auto& b : bodies) euler_step(b, dt);
  RngCtx rng;
  rng.mode = COTIX_RNG_THREEFRY;
  rng.steps = rng.episode = 0;
  rng.key.a = k0;
  rng.key.b = k1;
  ColliderOut co;
  collider_resolve(bodies, rng, mk(init_dir[0], init_dir[1]), gjk_max, coin_p, baumgarte,
                   baumgarte_dt, dedupe != 0, &co);
  if (co.error) return -3;
  for (int b = 0; b < nb; ++b) {
    float* s = bodies10 + 10 * b;
    const Body& B = bodies[(size_t)b];
    s[2] = B.pos.x;
    s[3] = B.pos.y;
    s[4] = B.vel.x;
    s[5] = B.vel.y;
    s[6] = B.angle;
    s[7] = B.angvel;
  }
  if (cell)
    for (int i = 0; i < nb; ++i)
      for (int j = 0; j < nb; ++j) cell[i * nb + j] = co.cell[i][j];
  if (chosen)
    for (int i = 0; i < nb; ++i) chosen[i] = co.chosen[i];
  if (status) *status = co.status;
  return 0;
}

// sin/cos/atan2 under the current trig mode (tests/test_trig.py checks cx_trig.h against libm)
int oracle_trig(int64_t n, const float* x, const float* y, float* out_sin, float* out_cos,
                float* out_atan2) {
  for (int64_t k = 0; k < n; ++k) {
    out_sin[k] = o_sin(x[k]);
    out_cos[k] = o_cos(x[k]);
    out_atan2[k] = o_atan2(y[k], x[k]);
  }
  return 0;
}

// ---- headless rasteriser (the stand-in for cotix/_viz.py's pygame window) -------------------
// PARITY UNPINNED against pygame / SDL (not installed here): the pixel rules are the engine's own
// definition (parallax_b200/csrc/raster_kernel.cu); this is their sequential restatement — the
// painter's algorithm one primitive and one pixel at a time, lines by an incremental
// error-accumulating walk (the CUDA kernel evaluates every pixel in closed form).
//   prims int32 [n_img][max_prims][16], counts int32 [n_img] or NULL, image uint8 [n_img][600][800][3]
static void o_put(uint8_t* img, long long x, long long y, uint32_t rgb) {
  if (x < 0 || x >= COTIX_IMG_W || y < 0 || y >= COTIX_IMG_H) return;
  uint8_t* p = img + (y * COTIX_IMG_W + x) * 3;
  p[0] = (uint8_t)(rgb >> 16);
  p[1] = (uint8_t)(rgb >> 8);
  p[2] = (uint8_t)rgb;
}
int oracle_raster_draw(int64_t n_img, const int32_t* prims, const int32_t* counts, int32_t max_prims,
                       uint8_t* image) {
  const int64_t img_bytes = (int64_t)COTIX_IMG_W * COTIX_IMG_H * 3;
  for (int64_t im = 0; im < n_img; ++im) {
    uint8_t* img = image + im * img_bytes;
    memset(img, 0, (size_t)img_bytes);  // PyPainter.next: screen.fill((0, 0, 0))
    int cnt = counts ? counts[im] : max_prims;
    if (cnt > max_prims) cnt = max_prims;
    for (int k = 0; k < cnt; ++k) {
      const int32_t* q = prims + (im * max_prims + k) * COTIX_PRIM_WORDS;
      const uint32_t rgb = (uint32_t)q[1];
      if (q[0] == COTIX_PRIM_LINE) {
        long long x0 = q[4], y0 = q[5], x1 = q[6], y1 = q[7];
        if (x1 < x0 || (x1 == x0 && y1 < y0)) {
          std::swap(x0, x1);
          std::swap(y0, y1);
        }
        const long long dx = x1 - x0, dy = y1 >= y0 ? y1 - y0 : y0 - y1, sy = y1 >= y0 ? 1 : -1;
        if (dx >= dy) {
          long long y = y0, err = dx;  // err / (2 dx) = the fractional position + 1/2
          for (long long x = x0; x <= x1; ++x) {
            // skip the off-screen run quickly (terrain edges extend far beyond the screen)
            if (x >= 0 && x < COTIX_IMG_W) o_put(img, x, y, rgb);
            if (x == x1) break;
            err += 2 * dy;
            while (dx > 0 && err >= 2 * dx) {
              err -= 2 * dx;
              y += sy;
            }
          }
        } else {
          long long x = x0, err = dy;
          for (long long i = 0; i <= dy; ++i) {
            o_put(img, x, y0 + sy * i, rgb);
            err += 2 * dx;
            while (err >= 2 * dy) {
              err -= 2 * dy;
              x += 1;
            }
          }
        }
      } else if (q[0] == COTIX_PRIM_CIRCLE) {
        const long long cx = q[4], cy = q[5], r = q[2];
        if (r < 1) continue;
        for (long long y = cy - r; y <= cy + r; ++y)
          for (long long x = cx - r; x <= cx + r; ++x)
            if ((x - cx) * (x - cx) + (y - cy) * (y - cy) <= r * r) o_put(img, x, y, rgb);
      } else if (q[0] == COTIX_PRIM_POLYGON) {
        const int nv = q[2];
        if (nv < 3 || nv > 6) continue;
        for (long long y = 0; y < COTIX_IMG_H; ++y)
          for (long long x = 0; x < COTIX_IMG_W; ++x) {
            bool pos = true, neg = true;
            for (int a = 0; a < nv; ++a) {
              const int b = (a + 1) % nv;
              const long long ex = (long long)q[4 + 2 * b] - q[4 + 2 * a], ey = (long long)q[5 + 2 * b] - q[5 + 2 * a];
              const long long e = ex * (y - q[5 + 2 * a]) - ey * (x - q[4 + 2 * a]);
              pos = pos && e >= 0;
              neg = neg && e <= 0;
            }
            if (pos || neg) o_put(img, x, y, rgb);
          }
      }
    }
  }
  return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
// explicit thread count for the timed CPU arm (torchrun exports OMP_NUM_THREADS=1); n <= 0: all
// logical cores of the box
int oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n <= 0) n = omp_get_num_procs();
  omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}
int oracle_sizeof_lunar_params(void) { return (int)sizeof(CotixLunarParams); }
const char* oracle_last_error(void) { return g_err; }

}  // extern "C"
