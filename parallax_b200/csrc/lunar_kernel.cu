// lunar_kernel.cu — K1 (fused LunarLander step, one launch) and K4 (reset).
//
// Mapping: one environment per THREAD for the regular part of the
// step, and a block-level work queue for the irregular part:
//   phase A  (thread = env)   integrate, engines, world polygons, exact broadphase cull;
//                             envs with candidate pairs publish their sorted world polygons to
//                             shared memory and push (env, pair) items
//   phase B  (thread = item)  GJK + EPA + contact generation, one candidate pair per thread —
//                             the 128 threads of the block drain the queue together, so the
//                             divergent narrowphase runs at full lane occupancy instead of on
//                             the few lanes whose env happens to be in contact
//   phase C  (thread = env)   coin flips / random choice (lazy threefry or Philox), sequential
//                             impulse resolution, joints, judge, auto-reset, store
// State is struct-of-arrays over the env axis (include/cotix_b200.h): every field access of a
// warp is one coalesced 128-byte transaction.  This kernel keeps the state in registers across
// K sub-steps; for large batches at K = 1 the wavefront pipeline (lunar_pipeline.cu) is faster.
#include <cmath>
#include <cstring>
#include <map>
#include <mutex>

#include "cx_launch.cuh"
#include "cx_lunar_step.cuh"

namespace {

constexpr int TPB = 128;    // threads (= envs) per block
constexpr int QCAP = 1024;  // narrowphase items per block and sub-step; overflow falls back to
                            // the owning thread evaluating its own pairs (same results)
constexpr uint32_t ITEM_SKIP = 0xFFFFFFFFu;

struct SharedBlock {
  float poly[28][TPB];  // sorted world vertices: lander 0..11, right leg 12..19, left leg 20..27
  float terr[12][TPB];  // heights 0..7, knots 8..11
  uint32_t item[QCAP];  // (env_local << 8) | pair id
  float res[4][QCAP];   // pen.x, pen.y, cp.x, cp.y   (cp NaN = no contact)
  int qcount;
};

// polygon source: env `el` of this block, from shared memory
struct SmemPolys {
  const SharedBlock& S;
  int el;
  CX_DEV void body(int which, Poly& p) const {
    const int off = which == 0 ? 0 : (which == 1 ? 12 : 20);
    const int n = which == 0 ? 6 : 4;
    p.n = n;
#pragma unroll
    for (int k = 0; k < CX_MAXV; ++k) {
      if (k < n) {
        p.x[k] = S.poly[off + 2 * k][el];
        p.y[k] = S.poly[off + 2 * k + 1][el];
      } else {
        p.x[k] = 0.0f;
        p.y[k] = 0.0f;
      }
    }
  }
  CX_DEV void ground_edge(int k, const CotixLunarParams& P, float& x0, float& h0, float& x1, float& h1) const {
    float kn[4] = {S.terr[8][el], S.terr[9][el], S.terr[10][el], S.terr[11][el]};
    x0 = terrain_x(kn, P, k);
    x1 = terrain_x(kn, P, k + 1);
    h0 = S.terr[k][el];
    h1 = S.terr[k + 1][el];
  }
};

// narrowphase of one (env, pair) item
static __device__ __noinline__ Contact eval_item(const SharedBlock& S, int el, int pid,
                                                 const CotixLunarParams& P) {
  Poly A, B;
  SmemPolys src{S, el};
  make_item_polys(src, pid, P, A, B);
  return poly_pair(A, B, mk2(P.init_dir[0], P.init_dir[1]), P.gjk_max_iters);
}

// ---- phase A: regular part up to the broadphase; publishes candidates --------------------
__device__ void phase_a(EnvRegs& E, float a_main, float a_side, const CotixLunarParams& P,
                        const LunarDerived& D, SharedBlock& S, StepCtx& cx) {
  const int tid = threadIdx.x;
  Poly wL, wr, wl;
  integrate_and_cull<true>(E, a_main, a_side, P, D, cx, wL, wr, wl);
  const uint32_t cand = cx.cand;
  if (cand != 0u) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      S.poly[2 * k][tid] = wL.x[k];
      S.poly[2 * k + 1][tid] = wL.y[k];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      S.poly[12 + 2 * k][tid] = wr.x[k];
      S.poly[13 + 2 * k][tid] = wr.y[k];
      S.poly[20 + 2 * k][tid] = wl.x[k];
      S.poly[21 + 2 * k][tid] = wl.y[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) S.terr[k][tid] = E.h[k];
#pragma unroll
    for (int k = 0; k < 4; ++k) S.terr[8 + k][tid] = E.kn[k];
    const int cnt = __popc(cand);
    const int base = atomicAdd(&S.qcount, cnt);
    if (base + cnt <= QCAP) {
      cx.base = base;
      uint32_t m = cand;
      int slot = base;
      while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        S.item[slot++] = ((uint32_t)tid << 8) | (uint32_t)bit;
      }
    } else {
      for (int slot = base; slot < QCAP && slot < base + cnt; ++slot) S.item[slot] = ITEM_SKIP;
    }
  }
}

// ---- phase B: the block drains the narrowphase queue, one pair per thread ----------------
__device__ void phase_b(SharedBlock& S, const CotixLunarParams& P) {
  const int n = S.qcount < QCAP ? S.qcount : QCAP;
  for (int i = threadIdx.x; i < n; i += TPB) {
    const uint32_t it = S.item[i];
    if (it == ITEM_SKIP) continue;
    const Contact c = eval_item(S, (int)(it >> 8), (int)(it & 0xFFu), P);
    S.res[0][i] = c.pen.x;
    S.res[1][i] = c.pen.y;
    S.res[2][i] = c.cp.x;
    S.res[3][i] = c.cp.y;
  }
}

// contact source of this thread's env: the block queue's results (or inline on overflow)
struct SmemContacts {
  const SharedBlock& S;
  const StepCtx& cx;
  const CotixLunarParams& P;
  CX_DEV Contact get(int bit) const {
    if (cx.base >= 0) {
      const int i = cx.base + __popc(cx.cand & ((1u << bit) - 1u));
      Contact c;
      c.pen = mk2(S.res[0][i], S.res[1][i]);
      c.cp = mk2(S.res[2][i], S.res[3][i]);
      c.status = 0u;
      return c;
    }
    return eval_item(S, threadIdx.x, bit, P);  // queue overflow: evaluate inline (same result)
  }
  CX_DEV bool live(int bit) const {
    if (!((cx.cand >> bit) & 1u)) return false;
    if (cx.base >= 0) {
      const int i = cx.base + __popc(cx.cand & ((1u << bit) - 1u));
      const float x = S.res[2][i], y = S.res[3][i];
      return !((x != x) || (y != y));
    }
    return !contact_isnan(eval_item(S, threadIdx.x, bit, P));
  }
};

// ---- phase C: contact selection, resolution, joints, judge --------------------------------
__device__ void phase_c(EnvRegs& E, const CotixLunarParams& P, const LunarDerived& D, const SharedBlock& S,
                        const StepCtx& cx, SubInfo& info) {
  bool crash, rc, lc;
  SmemContacts cs{S, cx, P};
  collide(E, P, D, cx.cand, cs, crash, rc, lc);
  finish_step(E, P, D, cx.cs, cx.sn, cx.thr_main, cx.thr_side, crash, rc, lc, info);
}

__global__ void __launch_bounds__(TPB)
lunar_step_kernel(int64_t B, const float* __restrict__ sin_, float* __restrict__ sout, int64_t ld,
                  const float* __restrict__ action, float* __restrict__ reward,
                  uint8_t* __restrict__ done, float* __restrict__ obs,
                  const __grid_constant__ CotixLunarParams P, const LunarDerived D) {
  __shared__ SharedBlock S;
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  const bool valid = e < B;
  EnvRegs E;
  float am = 0.0f, as = 0.0f;
  if (valid) {
    load_env(E, sin_, ld, e);
    if (action) {
      am = action[e];
      as = action[ld + e];
    }
  }
  float rsum = 0.0f;
  int d = COTIX_DONE_NO;
  bool rc = false, lc = false;
  bool running = valid;
  const int K = P.substeps < 1 ? 1 : P.substeps;
  for (int k = 0; k < K; ++k) {
    if (threadIdx.x == 0) S.qcount = 0;
    __syncthreads();
    StepCtx cx;
    cx.cand = 0u;
    cx.base = -1;
    if (running) phase_a(E, am, as, P, D, S, cx);
    __syncthreads();
    phase_b(S, P);
    __syncthreads();
    if (running) {
      SubInfo si;
      phase_c(E, P, D, S, cx, si);
      rsum = rsum + si.reward;
      rc = si.rc;
      lc = si.lc;
      if (si.done != COTIX_DONE_NO) {  // freeze on done (cotix/_envs.py:81-103)
        d = si.done;
        running = false;
      }
    }
  }
  if (!valid) return;
  bool terrain = (sin_ != sout);
  if (d != COTIX_DONE_NO && P.auto_reset) {
    lunar_auto_reset(E, P);
    rc = lc = false;
    terrain = true;
  }
  store_env(E, sout, ld, e, terrain);
  if (reward) reward[e] = rsum;
  if (done) done[e] = (uint8_t)d;
  if (obs) store_obs(E, obs, ld, e, rc, lc);
}

__global__ void __launch_bounds__(TPB)
lunar_reset_kernel(int64_t B, const uint32_t* __restrict__ keys, float* __restrict__ state,
                   int64_t ld, const __grid_constant__ CotixLunarParams P) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= B) return;
  EnvRegs E;
  Key2 key;
  key.a = keys[2 * e];
  key.b = keys[2 * e + 1];
  lunar_reset_env(E, key, P, 0u);
  E.key = key;
  store_env(E, state, ld, e, true);
}

}  // namespace

// Host-side analysis of the constant local shapes (see LunarDerived).
static void derive_body(const float (*v)[2], int n, int* ok, int* orient) {
  double cx = 0, cy = 0, area2 = 0;
  for (int k = 0; k < n; ++k) {
    cx += v[k][0];
    cy += v[k][1];
    const int j = (k + 1) % n;
    area2 += (double)v[k][0] * v[j][1] - (double)v[j][0] * v[k][1];
  }
  cx /= n;
  cy /= n;
  *orient = area2 >= 0 ? 1 : -1;
  bool good = std::isfinite(area2) && fabs(area2) > 1e-6;
  for (int k = 0; k < n && good; ++k) {
    const int j = (k + 1) % n, l = (k + 2) % n;
    const double ax = v[k][0] - cx, ay = v[k][1] - cy, bx = v[j][0] - cx, by = v[j][1] - cy;
    // angular gap between neighbours about the centroid, same turning direction all around
    const double gap = atan2(ax * by - ay * bx, ax * bx + ay * by) * (*orient);
    if (!(gap > 1e-3 && gap < 3.1)) good = false;
    // convexity: consecutive edges turn the same way
    const double e1x = v[j][0] - v[k][0], e1y = v[j][1] - v[k][1];
    const double e2x = v[l][0] - v[j][0], e2y = v[l][1] - v[j][1];
    if (!((e1x * e2y - e1y * e2x) * (*orient) > 0)) good = false;
  }
  *ok = good ? 1 : 0;
}
// DivC for constant c (see cx_lunar.cuh): r = RN(1 / c) and the exhaustive check that
// fma(fma(-x r, c, x), r, x r) == x / c for all 2^23 floats x of one binade (the property is invariant
// under scaling x by powers of two while nothing under- or overflows, which div_c's range test excludes).
// One check per distinct constant and process (a few tens of milliseconds), cached.
static DivC cx_divc_make(float c) {
  DivC d;
  d.c = c;
  d.r = 0.0f;
  d.ok = 0;
  const float ac = fabsf(c);
  if (std::isinf(c)) {  // static bodies: x / +-inf == x * +-0 for every x (finite -> signed zero, inf / NaN -> NaN)
    d.r = c > 0 ? 0.0f : -0.0f;
    d.ok = 2;
    return d;
  }
  if (!(ac >= 1e-6f && ac <= 1e6f) || c == 1.0f) return d;  // NaN, extreme: plain division
  static std::mutex mu;
  static std::map<uint32_t, int> cache;
  uint32_t bits;
  memcpy(&bits, &c, 4);
  d.r = 1.0f / c;
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(bits);
  if (it == cache.end()) {
    int ok = 1;
    for (uint32_t m = 0; m < (1u << 23) && ok; ++m) {
      const uint32_t xb = 0x3F800000u | m;  // x in [1, 2)
      float x;
      memcpy(&x, &xb, 4);
      const float q = x * d.r;
      const float rem = fmaf(-q, c, x);
      if (fmaf(rem, d.r, q) != x / c) ok = 0;
    }
    it = cache.emplace(bits, ok).first;
  }
  d.ok = it->second;
  return d;
}

extern "C" void cx_lunar_derive(const CotixLunarParams* p, LunarDerived* out) {
  LunarDerived& D = *out;
  derive_body(p->lander_local, 6, &D.rigid_ok[0], &D.orient[0]);
  derive_body(p->leg_local[0], 4, &D.rigid_ok[1], &D.orient[1]);
  derive_body(p->leg_local[1], 4, &D.rigid_ok[2], &D.orient[2]);
  D.lext[0] = D.lext[1] = INFINITY;
  D.lext[2] = D.lext[3] = -INFINITY;
  for (int k = 0; k < 6; ++k) {
    D.lext[0] = fminf(D.lext[0], p->lander_local[k][0]);
    D.lext[1] = fminf(D.lext[1], p->lander_local[k][1]);
    D.lext[2] = fmaxf(D.lext[2], p->lander_local[k][0]);
    D.lext[3] = fmaxf(D.lext[3], p->lander_local[k][1]);
  }
  for (int i = 0; i < 4; ++i) {
    D.inv_mass[i] = 1.0f / p->body_props[i][0];
    D.dmass[i] = cx_divc_make(p->body_props[i][0]);
    D.dinertia[i] = cx_divc_make(p->body_props[i][1]);
  }
}

extern "C" int cx_launch_lunar_step(cudaStream_t st, int64_t B, const float* sin_, float* sout,
                                    int64_t ld, const float* action, float* reward, uint8_t* done,
                                    float* obs, const CotixLunarParams* p) {
  if (B == 0) return 0;
  LunarDerived D;
  cx_lunar_derive(p, &D);
  lunar_step_kernel<<<(unsigned)((B + TPB - 1) / TPB), TPB, 0, st>>>(B, sin_, sout, ld, action, reward,
                                                                     done, obs, *p, D);
  cx_count_launch();
  return cx_check_launch("lunar_step_kernel");
}
extern "C" int cx_launch_lunar_reset(cudaStream_t st, int64_t B, const uint32_t* keys, float* state,
                                     int64_t ld, const CotixLunarParams* p) {
  if (B == 0) return 0;
  lunar_reset_kernel<<<(unsigned)((B + TPB - 1) / TPB), TPB, 0, st>>>(B, keys, state, ld, *p);
  cx_count_launch();
  return cx_check_launch("lunar_reset_kernel");
}
