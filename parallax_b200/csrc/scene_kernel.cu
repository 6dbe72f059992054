// scene_kernel.cu — K2: the composed step for an ARBITRARY static scene description (bodies with
// circle / AABB / polygon parts), one env per thread.  This is the path the RoboCup scene takes
// (cotix/_robocup.py:14-130: AABB areas and goals + a circle ball, optionally robots).
//
//   ExplicitEulerPhysics.step            cotix/_physics_solvers.py:16-33
//   part.transform(body transform)       cotix/_convex_shapes.py:37-41, 113-117, 181-187
//   RandomizedCollider.resolve           cotix/_colliders.py:74-351
//   resolve_collision                    cotix/_collision_resolution.py:52-151
//   key, _ = jr.split(key)               examples/test_viz.py:39 (optional)
//
// What is static is analysed once on the host (cotix_scene_create): the groups of part pairs per
// shape-type pair in first-appearance order (:86-131), the entry lists of each group, and the
// table of DISTINCT (part, part) evaluations with body_i >= body_j — the reference evaluates the
// full n_g x n_g grid of every group (:149-173), most of it duplicates or i < j (NaN).  Per env
// and step the kernel evaluates each distinct pair once, flips a coin only where the contact is
// live (a coin on a NaN contact cannot change all_contacts, :231-246; every key is a counter-
// mode function of (group, ind2, ind1), so skipping the others changes nothing), then runs the
// per-body choice (:274-295) and the sequential resolve scan (:310-336).
#include <cstring>
#include <algorithm>
#include <new>
#include <vector>

#include "../../include/cotix_b200.h"
#include "cx_launch.cuh"
#include "cx_lunar.cuh"
#include "cx_narrow.cuh"
#include "cx_rng.cuh"

#define CXS_REQUIRE(cond, msg)               \
  do {                                       \
    if (!(cond)) {                           \
      cx_set_error("%s: %s", __func__, msg); \
      return COTIX_EINVAL;                   \
    }                                        \
  } while (0)
#define CXS_CUDA(call)                                                            \
  do {                                                                            \
    cudaError_t e_ = (call);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      cx_set_error("%s: %s failed: %s", __func__, #call, cudaGetErrorString(e_)); \
      return COTIX_ECUDA;                                                         \
    }                                                                             \
  } while (0)

namespace {

constexpr int MAXB = COTIX_SCENE_MAX_BODIES;
constexpr int MAXP = COTIX_SCENE_MAX_PARTS;
constexpr int MAXPAIR = COTIX_SCENE_MAX_PAIRS;
constexpr int MAXENT = COTIX_SCENE_MAX_ENTRIES;
constexpr int MAXG = 9;  // at most 9 keys of _contact_funcs can appear

// the static analysis of a scene; lives in device memory, read uniformly by all threads
struct SceneDesc {
  int32_t nb, nparts, ngroups, npairs;
  float props[MAXB][4];       // mass, inertia, elasticity, friction_coefficient
  float rec[MAXP][16];        // local shape records (include/cotix_b200.h)
  uint8_t part_body[MAXP];
  uint8_t pair_a[MAXPAIR], pair_b[MAXPAIR];  // distinct evaluations: part ids
  uint8_t pair_id[MAXP][MAXP];               // (part of list 0, part of list 1) -> pair or 0xFF
  int32_t g_n[MAXG], g_off[MAXG];            // entries of group g: ent0/ent1[g_off .. g_off + g_n)
  uint8_t ent0[MAXENT], ent1[MAXENT];        // part id of the k-th entry of list 0 / list 1
  // all_contacts cells that have grid entries: cell c = (i, j) holds its grid entries in
  // DESCENDING scan order (group, ind2, ind1), each packed as
  // pair | ind1 << 8 | ind2 << 16 | group << 24 (device arrays of the scene).  The distinct pairs
  // are numbered in cell order, so the pairs of one cell are consecutive.
  int32_t ncells;
  uint16_t pair_cell[MAXPAIR];  // pair -> its cell
  const uint16_t* cell_ij;      // [ncells] i << 8 | j
  const int32_t* cell_ent_off;  // [ncells + 1] into cell_ents
  const uint32_t* cell_ents;
};

CX_DEV void sort_generic(Poly& p) {
  switch (p.n) {
    case 3: poly_sort_cw<3>(p); break;
    case 4: poly_sort_cw<4>(p); break;
    case 5: poly_sort_cw<5>(p); break;
    case 6: poly_sort_cw<6>(p); break;
    default: break;  // n <= 2: nothing to order for a point; a segment keeps argsort order below
  }
}

// part.transform(transformer of the body): circles and boxes translate only (the angle is
// ignored, cotix/_convex_shapes.py:37-41, 113-117); polygons rotate, translate and re-sort
CX_DEV void world_part(const float* r, float px, float py, float c, float s, float out[12]) {
  const int type = (int)r[0], n = (int)r[1];
  if (type == COTIX_SHAPE_CIRCLE) {
    out[0] = r[4] + px;
    out[1] = r[5] + py;
  } else if (type == COTIX_SHAPE_AABB) {
    out[0] = r[4] + px;
    out[1] = r[5] + py;
    out[2] = r[6] + px;
    out[3] = r[7] + py;
  } else {
    Poly p;
    p.n = n < 1 ? 1 : (n > CX_MAXV ? CX_MAXV : n);
    const float ms = -s;
#pragma unroll
    for (int k = 0; k < CX_MAXV; ++k) {
      if (k < p.n) {
        const float x = r[4 + 2 * k], y = r[5 + 2 * k];
        // forward_vector: matrix @ [x, y, 1] / w, w = 1 exactly (cotix/_geometry_utils.py:134-138)
        p.x[k] = c * x + ms * y + px * 1.0f;
        p.y[k] = s * x + c * y + py * 1.0f;
      } else {
        p.x[k] = 0.0f;
        p.y[k] = 0.0f;
      }
    }
    sort_generic(p);
#pragma unroll
    for (int k = 0; k < CX_MAXV; ++k) {
      out[2 * k] = p.x[k];
      out[2 * k + 1] = p.y[k];
    }
  }
}

// a part in world coordinates; type, vertex count and radius are the scene's (rec[0..2])
struct WPart {
  float v[12];
};
CX_DEV int rec_n(const float* r) {
  const int n = (int)r[1];
  return n < 1 ? 1 : (n > CX_MAXV ? CX_MAXV : n);
}
CX_DEV Poly wp_poly(const WPart& w, const float* r) {
  Poly p;
  p.n = rec_n(r);
#pragma unroll
  for (int k = 0; k < CX_MAXV; ++k) {
    p.x[k] = w.v[2 * k];
    p.y[k] = w.v[2 * k + 1];
  }
  return p;
}
CX_DEV Circ wp_circ(const WPart& w, const float* r) {
  Circ c;
  c.c = mk2(w.v[0], w.v[1]);
  c.r = r[2];
  return c;
}
CX_DEV Box wp_box(const WPart& w) {
  Box b;
  b.lo = mk2(w.v[0], w.v[1]);
  b.hi = mk2(w.v[2], w.v[3]);
  return b;
}

// _contact_funcs[(type(a), type(b))](a, b), cotix/_colliders.py:21-35 (the host has checked that
// every pair of the scene is a key in this order)
static __device__ __noinline__ Contact contact_of(const WPart& a, const float* ra, const WPart& b,
                                                  const float* rb, V2 start, int gjk_max) {
  const int ta = (int)ra[0], tb = (int)rb[0];
  if (ta == COTIX_SHAPE_POLYGON && tb == COTIX_SHAPE_POLYGON)
    return convex_vs_polygon(wp_poly(a, ra), wp_poly(b, rb), start, gjk_max);
  if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_POLYGON)
    return circle_vs_polygon<131>(wp_circ(a, ra), wp_poly(b, rb), start, gjk_max, 128);
  if (ta == COTIX_SHAPE_AABB && tb == COTIX_SHAPE_POLYGON)
    return convex_vs_polygon(wp_box(a), wp_poly(b, rb), start, gjk_max);
  if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_CIRCLE)
    return circle_vs_circle(wp_circ(a, ra), wp_circ(b, rb));
  if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_AABB) return circle_vs_aabb(wp_circ(a, ra), wp_box(b));
  return aabb_vs_aabb(wp_box(a), wp_box(b));
}

// jr.choice(key, nb, p=probs) for one body: cumsum / uniform / searchsorted(side='left').
// probs[j] is 1/cnt on the live cells (bit j of `live`) and 0/cnt = 0 elsewhere.
CX_DEV int choose_generic(uint32_t live, int nb, float u) {
  const float fc = (float)__popc(live);
  const float p1 = 1.0f / fc;
  float total = 0.0f;
  for (int j = 0; j < nb; ++j) {
    const float p = ((live >> j) & 1u) ? p1 : 0.0f;
    total = (j == 0) ? p : total + p;
  }
  const float r = total * (1.0f - u);
  int ind = 0;
  float acc = 0.0f;
  for (int j = 0; j < nb; ++j) {
    const float p = ((live >> j) & 1u) ? p1 : 0.0f;
    acc = (j == 0) ? p : acc + p;
    if (acc < r) ++ind;
  }
  return ind >= nb ? nb - 1 : ind;  // jnp indexing clamps
}

#ifndef CXS_MINB
#define CXS_MINB 6  // 4 / 5 / 6 / 8 / 10 / 12 blocks per SM: 0.728 / 0.690 / 0.614 / 0.627 / 0.662 / 0.705 ms per RoboCup step of 262144 envs
#endif
__global__ void __launch_bounds__(128, CXS_MINB)
scene_step_kernel(const SceneDesc* __restrict__ S, int64_t B, float* __restrict__ state, int64_t ld,
                  uint32_t* __restrict__ keys, const CotixSceneParams P, uint8_t* __restrict__ cell,
                  int32_t* __restrict__ chosen_out, uint32_t* __restrict__ status_out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= B) return;
  const int nb = S->nb, nparts = S->nparts, npairs = S->npairs;
  BodyS body[MAXB];
  for (int b = 0; b < nb; ++b) {
    body[b].px = state[(b * 6 + 0) * ld + e];
    body[b].py = state[(b * 6 + 1) * ld + e];
    body[b].vx = state[(b * 6 + 2) * ld + e];
    body[b].vy = state[(b * 6 + 3) * ld + e];
    body[b].ang = state[(b * 6 + 4) * ld + e];
    body[b].w = state[(b * 6 + 5) * ld + e];
    // ExplicitEulerPhysics._single_body_step
    body[b].px = body[b].px + body[b].vx * P.dt;
    body[b].py = body[b].py + body[b].vy * P.dt;
    body[b].ang = body[b].ang + body[b].w * P.dt;
  }
  uint32_t status = 0u;
  // world parts
  WPart W[MAXP];
  for (int q = 0; q < nparts; ++q) {
    const int b = S->part_body[q];
    float sn = 0.0f, cs = 1.0f;  // only polygons rotate
    if ((int)S->rec[q][0] == COTIX_SHAPE_POLYGON) cx_sincosf(body[b].ang, &sn, &cs);
    world_part(S->rec[q], body[b].px, body[b].py, cs, sn, W[q].v);
  }
  // distinct narrowphase evaluations.  A contact is live when its contact point is not NaN
  // (:255-262); the cells with a live contact are collected on the way (the pairs of a cell are
  // consecutive), so nothing later scans the whole cell table.
  const V2 start = mk2(P.init_dir[0], P.init_dir[1]);
  float cpen[MAXPAIR][2], ccp[MAXPAIR][2];
  constexpr int LIVE_CAP = 64;
  uint16_t lcell[LIVE_CAP];
  int nlive = 0, last_cell = -1;
  bool spill = false;  // more live cells than LIVE_CAP: fall back to a scan of the cell table
  uint32_t lm[(MAXPAIR + 31) / 32];  // bit q: contact q is live (only live contacts are stored)
  for (int w = 0; w < (MAXPAIR + 31) / 32; ++w) lm[w] = 0u;
  for (int q = 0; q < npairs; ++q) {
    const int pa = S->pair_a[q], pb = S->pair_b[q];
    const Contact c = contact_of(W[pa], S->rec[pa], W[pb], S->rec[pb], start, P.gjk_max_iters);
    status |= c.status;
    if ((c.cp.x != c.cp.x) || (c.cp.y != c.cp.y)) continue;
    lm[q >> 5] |= 1u << (q & 31);
    ccp[q][0] = c.cp.x; ccp[q][1] = c.cp.y;
    cpen[q][0] = c.pen.x; cpen[q][1] = c.pen.y;
    const int cc = S->pair_cell[q];
    if (cc == last_cell) continue;
    last_cell = cc;
    if (nlive < LIVE_CAP) lcell[nlive++] = (uint16_t)cc;
    else spill = true;
  }
  // coin-flip scatter into all_contacts (:208-268): the last heads in (ind2, ind1) order wins.
  // row[i] bit j = all_contacts[i][j] is live; allq[i][j] is its pair (written only when live).
  uint32_t row[MAXB];
  uint8_t allq[MAXB][MAXB];
  for (int i = 0; i < nb; ++i) row[i] = 0u;
  Key2 rkey;
  rkey.a = keys[e];
  rkey.b = keys[ld + e];
  // group keys: skey is split once more before every group (:175)
  Key2 skey = tf_split(rkey, 2u, 0u);
  Key2 gk[MAXG];
  for (int g = 0; g < S->ngroups; ++g) {
    skey = tf_split(skey, 2u, 0u);
    gk[g] = skey;
  }
  // The scans write cell (i, j) at every grid entry whose coin is heads and whose contact is
  // live, in (group, ind2, ind1) order, so the cell ends up with the LAST such entry.  Every
  // coin is a counter-mode function of its (group, ind2, ind1), hence: walk the entries of a
  // cell from the last backwards, skip dead contacts, stop at the first heads.
  //
  // A coin costs ~700 instructions and the number of coins per cell is geometric, so the walk is
  // ONE flat loop over (live cell, entry): every trip flips one coin in every lane that still has
  // one to flip, whichever of its cells that lane has reached.  (A loop per cell would run
  // max-over-lanes trips for every cell.)
  {
    int t = 0, k = 0, kend = 0, ij = 0, scan = 0;
    uint32_t c_g = 0xFFFFFFFFu, c_ind2 = 0u;  // the (group, ind2) that k2 below belongs to
    Key2 k2 = rkey;
    for (;;) {
      uint32_t ent = 0u;
      int q = 0;
      bool have = false;
      while (!have) {
        if (k >= kend) {  // next live cell
          int c;
          if (!spill) {
            if (t >= nlive) break;
            c = lcell[t++];
          } else {
            if (scan >= S->ncells) break;
            c = scan++;
          }
          k = S->cell_ent_off[c];
          kend = S->cell_ent_off[c + 1];
          ij = S->cell_ij[c];
          continue;
        }
        ent = S->cell_ents[k++];
        q = (int)(ent & 0xFFu);
        have = (lm[q >> 5] >> (q & 31)) & 1u;
      }
      if (!have) break;
      const uint32_t ind1 = (ent >> 8) & 0xFFu, ind2 = (ent >> 16) & 0xFFu, g = ent >> 24;
      const uint32_t n = (uint32_t)S->g_n[g];
      if (g != c_g || ind2 != c_ind2) {
        k2 = tf_split(gk[g], n, ind2);
        c_g = g;
        c_ind2 = ind2;
      }
      const Key2 k1 = tf_split(k2, n, ind1);
      const Key2 kk = tf_split(k1, 2u, 0u);
      if (tf_bernoulli(kk, P.coin_p)) {
        row[ij >> 8] |= 1u << (ij & 0xFF);
        allq[ij >> 8][ij & 0xFF] = (uint8_t)q;
        k = kend;
      }
    }
  }
  // choose_random_contact (:274-295)
  int chosen[MAXB];
  for (int i = 0; i < nb; ++i) {
    const uint32_t live = row[i];
    if (cell)
      for (int j = 0; j < nb; ++j) cell[(int64_t)(i * nb + j) * ld + e] = (uint8_t)((live >> j) & 1u);
    int idx = i;
    if (live != 0u) idx = choose_generic(live, nb, tf_uniform(tf_split(skey, (uint32_t)nb, (uint32_t)i), 0.0f, 1.0f));
    chosen[i] = idx;
    if (chosen_out) chosen_out[(int64_t)i * ld + e] = idx;
  }
  // sequential resolution scan (:310-336)
  for (int i = 0; i < nb; ++i) {
    const int j = chosen[i];
    if (i == j) continue;
    if (!((row[i] >> j) & 1u)) continue;  // NaN contact: resolve_collision returns the bodies unchanged
    const int q = allq[i][j];
    MassP mi, mj;
    mi.mass = S->props[i][0]; mi.inertia = S->props[i][1]; mi.e = S->props[i][2]; mi.mu = S->props[i][3];
    mi.inv_mass = 1.0f / mi.mass;  // (dmass / dinertia: not read by the out-of-line resolve_collision)
    mj.mass = S->props[j][0]; mj.inertia = S->props[j][1]; mj.e = S->props[j][2]; mj.mu = S->props[j][3];
    mj.inv_mass = 1.0f / mj.mass;
    resolve_collision(body[i], mi, body[j], mj, mk2(cpen[q][0], cpen[q][1]),
                      mk2(ccp[q][0], ccp[q][1]), P.baumgarte, P.baumgarte_dt);
  }
  for (int b = 0; b < nb; ++b) {
    state[(b * 6 + 0) * ld + e] = body[b].px;
    state[(b * 6 + 1) * ld + e] = body[b].py;
    state[(b * 6 + 2) * ld + e] = body[b].vx;
    state[(b * 6 + 3) * ld + e] = body[b].vy;
    state[(b * 6 + 4) * ld + e] = body[b].ang;
    state[(b * 6 + 5) * ld + e] = body[b].w;
  }
  if (P.advance_key) {
    const Key2 nk = tf_split(rkey, 2u, 0u);
    keys[e] = nk.a;
    keys[ld + e] = nk.b;
  }
  if (status_out) status_out[e] = status;
}

// ---- host-side analysis ----------------------------------------------------------------
// dictionary label of a part (Polygon4 and Polygon6 are distinct classes,
// cotix/_convex_shapes.py:197-229; any other vertex count is the generic Polygon)
int type_label(const float* r) {
  const int t = (int)r[0];
  if (t == COTIX_SHAPE_CIRCLE) return 0;
  if (t == COTIX_SHAPE_AABB) return 1;
  return 2 + (int)r[1];  // Polygon<n>
}
// is (t1, t2) a key of _contact_funcs (cotix/_colliders.py:21-35)?
bool legal_pair(int t1, int t2) {
  const bool p1 = t1 >= 2, p2 = t2 >= 2;
  if (t1 == 0) return true;                 // (Circle, Circle|AABB|Polygon*)
  if (t1 == 1) return t2 == 1 || p2;        // (AABB, AABB|Polygon*)
  if (p1 && p2) {
    const int n1 = t1 - 2, n2 = t2 - 2;
    if (n1 == 4) return n2 == 4 || n2 == 6;  // (Polygon4, Polygon4|Polygon6)
    if (n1 == 6) return n2 == 6;             // (Polygon6, Polygon6)
    return n2 != 4 && n2 != 6;               // (Polygon, Polygon): the generic class
  }
  return false;
}

}  // namespace

struct CotixScene {
  int device;
  SceneDesc host;
  SceneDesc* dev;   // one device block: the description followed by its cell tables
  int64_t n_evals;  // sum of n_g^2: what the reference's vmap^2 evaluates per step
};

extern "C" {

int cotix_scene_create(CotixScene** out, int device, int32_t nb, const float* body_props, int32_t nparts,
                       const int32_t* part_body, const float* part_recs) {
  CXS_REQUIRE(out && body_props && part_body && part_recs, "null argument");
  CXS_REQUIRE(nb >= 1 && nb <= MAXB, "number of bodies out of range");
  CXS_REQUIRE(nparts >= 1 && nparts <= MAXP, "number of parts out of range");
  CotixScene* sc = new (std::nothrow) CotixScene();
  CXS_REQUIRE(sc != nullptr, "out of host memory");
  SceneDesc& S = sc->host;
  memset(&S, 0, sizeof(S));
  memset(S.pair_id, 0xFF, sizeof(S.pair_id));
  S.nb = nb;
  S.nparts = nparts;
  memcpy(S.props, body_props, sizeof(float) * 4 * (size_t)nb);
  for (int q = 0; q < nparts; ++q) {
    if (part_body[q] < 0 || part_body[q] >= nb) {
      delete sc;
      CXS_REQUIRE(false, "part_body out of range");
    }
    const int t = (int)part_recs[16 * q];
    const int n = (int)part_recs[16 * q + 1];
    if (t < 0 || t > 2 || (t == COTIX_SHAPE_POLYGON && (n < 3 || n > COTIX_MAX_POLY))) {
      delete sc;
      CXS_REQUIRE(false, "bad shape record");
    }
    S.part_body[q] = (uint8_t)part_body[q];
    memcpy(S.rec[q], part_recs + 16 * q, sizeof(float) * 16);
  }
  // groups in first-appearance order over (i, j, part of i, part of j), i > j (:86-113)
  int gkey[MAXG][2];
  std::vector<uint8_t> l0[MAXG], l1[MAXG];
  int ng = 0;
  for (int i = 0; i < nb; ++i)
    for (int j = 0; j < i; ++j)
      for (int pa = 0; pa < nparts; ++pa) {
        if (S.part_body[pa] != i) continue;
        for (int pb = 0; pb < nparts; ++pb) {
          if (S.part_body[pb] != j) continue;
          const int t1 = type_label(S.rec[pa]), t2 = type_label(S.rec[pb]);
          if (!legal_pair(t1, t2)) {
            delete sc;
            cx_set_error("cotix_scene_create: parts %d (body %d) and %d (body %d): this order of shape "
                         "types is not a key of _contact_funcs (the reference raises)", pa, i, pb, j);
            return COTIX_EINVAL;
          }
          int g = 0;
          while (g < ng && !(gkey[g][0] == t1 && gkey[g][1] == t2)) ++g;
          if (g == ng) {
            gkey[g][0] = t1;
            gkey[g][1] = t2;
            ++ng;
          }
          l0[g].push_back((uint8_t)pa);
          l1[g].push_back((uint8_t)pb);
        }
      }
  S.ngroups = ng;
  int off = 0, npairs = 0;
  int64_t evals = 0;
  for (int g = 0; g < ng; ++g) {
    const int n = (int)l0[g].size();
    if (off + n > MAXENT) {
      delete sc;
      CXS_REQUIRE(false, "too many part pairs for COTIX_SCENE_MAX_ENTRIES");
    }
    S.g_n[g] = n;
    S.g_off[g] = off;
    for (int k = 0; k < n; ++k) {
      S.ent0[off + k] = l0[g][(size_t)k];
      S.ent1[off + k] = l1[g][(size_t)k];
    }
    off += n;
    evals += (int64_t)n * n;
    // the vmap^2 grid: entry a of list 0 against entry b of list 1, kept when body_i >= body_j
    for (int a = 0; a < n; ++a)
      for (int b = 0; b < n; ++b) {
        const int pa = l0[g][(size_t)a], pb = l1[g][(size_t)b];
        if (S.part_body[pa] < S.part_body[pb]) continue;
        if (S.pair_id[pa][pb] != 0xFF) continue;
        if (npairs >= MAXPAIR) {
          delete sc;
          CXS_REQUIRE(false, "too many distinct part pairs for COTIX_SCENE_MAX_PAIRS");
        }
        S.pair_id[pa][pb] = (uint8_t)npairs;
        S.pair_a[npairs] = (uint8_t)pa;
        S.pair_b[npairs] = (uint8_t)pb;
        ++npairs;
      }
  }
  S.npairs = npairs;
  {
    // renumber the pairs in cell order (body_a, body_b), stable inside a cell
    std::vector<int> order((size_t)npairs);
    for (int q = 0; q < npairs; ++q) order[(size_t)q] = q;
    auto cell_of = [&](int q) { return (int)S.part_body[S.pair_a[q]] * nb + (int)S.part_body[S.pair_b[q]]; };
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return cell_of(x) < cell_of(y); });
    uint8_t na[MAXPAIR], nbp[MAXPAIR];
    for (int q = 0; q < npairs; ++q) {
      na[q] = S.pair_a[order[(size_t)q]];
      nbp[q] = S.pair_b[order[(size_t)q]];
    }
    for (int q = 0; q < npairs; ++q) {
      S.pair_a[q] = na[q];
      S.pair_b[q] = nbp[q];
      S.pair_id[na[q]][nbp[q]] = (uint8_t)q;
    }
  }
  sc->n_evals = evals;
  sc->device = device;
  sc->dev = nullptr;
  // per-cell grid entries (ascending here, reversed below)
  std::vector<std::vector<uint32_t>> cents((size_t)nb * nb);
  for (int g = 0; g < ng; ++g) {
    const int n = S.g_n[g];
    for (int ind2 = 0; ind2 < n; ++ind2)
      for (int ind1 = 0; ind1 < n; ++ind1) {
        const int pa = l0[g][(size_t)ind1], pb = l1[g][(size_t)ind2];
        const int q = S.pair_id[pa][pb];
        if (q == 0xFF) continue;  // i < j: NaN (:163)
        const size_t c = (size_t)S.part_body[pa] * nb + S.part_body[pb];
        cents[c].push_back((uint32_t)q | ((uint32_t)ind1 << 8) | ((uint32_t)ind2 << 16) | ((uint32_t)g << 24));
      }
  }
  std::vector<uint16_t> h_ij;
  std::vector<int32_t> h_eoff(1, 0);
  std::vector<uint32_t> h_ents;
  for (int i = 0; i < nb; ++i)
    for (int j = 0; j < nb; ++j) {
      const size_t c = (size_t)i * nb + j;
      if (cents[c].empty()) continue;
      for (uint32_t ent : cents[c]) S.pair_cell[ent & 0xFFu] = (uint16_t)h_ij.size();
      h_ij.push_back((uint16_t)((i << 8) | j));
      h_ents.insert(h_ents.end(), cents[c].rbegin(), cents[c].rend());
      h_eoff.push_back((int32_t)h_ents.size());
    }
  S.ncells = (int32_t)h_ij.size();
  if (device >= 0) {
    // one device block: [SceneDesc | cell_ij | cell_ent_off | cell_ents]
    auto al = [](size_t x) { return (x + 15) / 16 * 16; };
    const size_t o_ij = al(sizeof(SceneDesc)), o_eoff = o_ij + al(2 * h_ij.size()),
                 o_ents = o_eoff + al(4 * h_eoff.size()), total = o_ents + al(4 * h_ents.size());
    std::vector<char> blob(total, 0);
    cudaError_t ce = cudaSetDevice(device);
    char* d = nullptr;
    if (ce == cudaSuccess) ce = cudaMalloc(&d, total);
    if (ce == cudaSuccess) {
      S.cell_ij = reinterpret_cast<const uint16_t*>(d + o_ij);
      S.cell_ent_off = reinterpret_cast<const int32_t*>(d + o_eoff);
      S.cell_ents = reinterpret_cast<const uint32_t*>(d + o_ents);
      memcpy(blob.data(), &S, sizeof(SceneDesc));
      memcpy(blob.data() + o_ij, h_ij.data(), 2 * h_ij.size());
      memcpy(blob.data() + o_eoff, h_eoff.data(), 4 * h_eoff.size());
      memcpy(blob.data() + o_ents, h_ents.data(), 4 * h_ents.size());
      sc->dev = reinterpret_cast<SceneDesc*>(d);
      ce = cudaMemcpy(d, blob.data(), total, cudaMemcpyHostToDevice);
    }
    if (ce != cudaSuccess) {
      cx_set_error("cotix_scene_create: %s", cudaGetErrorString(ce));
      cudaGetLastError();
      if (d) cudaFree(d);
      delete sc;
      return COTIX_ECUDA;
    }
  }
  *out = sc;
  return COTIX_OK;
}

int cotix_scene_destroy(CotixScene* sc) {
  if (!sc) return COTIX_OK;
  if (sc->dev) {
    cudaSetDevice(sc->device);
    cudaFree(sc->dev);
  }
  delete sc;
  return COTIX_OK;
}

int cotix_scene_info(const CotixScene* sc, int32_t* n_groups, int32_t* n_pairs, int64_t* n_reference_evals,
                     int32_t* group_sizes) {
  CXS_REQUIRE(sc != nullptr, "scene is null");
  if (n_groups) *n_groups = sc->host.ngroups;
  if (n_pairs) *n_pairs = sc->host.npairs;
  if (n_reference_evals) *n_reference_evals = sc->n_evals;
  if (group_sizes)
    for (int g = 0; g < sc->host.ngroups; ++g) group_sizes[g] = sc->host.g_n[g];
  return COTIX_OK;
}

void cotix_scene_default_params(CotixSceneParams* p) {
  if (!p) return;
  CotixLunarParams lp;
  cotix_lunar_default_params(&lp);
  memset(p, 0, sizeof(*p));
  p->dt = lp.dt;
  p->init_dir[0] = lp.init_dir[0];
  p->init_dir[1] = lp.init_dir[1];
  p->gjk_max_iters = lp.gjk_max_iters;
  p->coin_p = lp.coin_p;
  p->baumgarte = lp.baumgarte;
  p->baumgarte_dt = lp.baumgarte_dt;
  p->advance_key = 1;
}

int cotix_scene_step(void* stream, const CotixScene* sc, int64_t B, float* state, int64_t ld, uint32_t* keys,
                     const CotixSceneParams* p, uint8_t* cell, int32_t* chosen, uint32_t* status) {
  CXS_REQUIRE(sc != nullptr && p != nullptr, "null argument");
  CXS_REQUIRE(sc->dev != nullptr, "the scene was created without a device");
  CXS_REQUIRE(B >= 0 && ld >= B, "bad batch size / leading dimension");
  CXS_REQUIRE(B == 0 || (state && keys), "null buffer");
  CXS_REQUIRE(p->gjk_max_iters >= 0, "bad params");
  if (B == 0) return COTIX_OK;
  const int tpb = 128;
  const int64_t blocks = (B + tpb - 1) / tpb;
  CXS_REQUIRE(blocks <= 0x7FFFFFFF, "batch too large");
  scene_step_kernel<<<(unsigned)blocks, tpb, 0, (cudaStream_t)stream>>>(sc->dev, B, state, ld, keys, *p, cell,
                                                                          chosen, status);
  cx_count_launch();
  return cx_check_launch("scene_step_kernel");
}

}  // extern "C"
