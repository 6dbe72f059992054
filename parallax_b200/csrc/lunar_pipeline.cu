// lunar_pipeline.cu — the LunarLander step as a WAVEFRONT pipeline of small dense kernels.
//
// Why (profiles/r1c_step_v3.txt): in the fused kernel the regular part of the step costs
// ~160 us per 2^20 envs, but a contact-rich batch takes ~1400 us — the block waits at a barrier
// while a handful of lanes run GJK/EPA, coin-flip chains or an auto-reset (11 of 32 lanes active
// on average, 36 % of the samples at the barrier, 17 % issue-slot use).  HBM is nowhere near
// the limit (3 % of peak), so the cure is the classic one for divergent GPU work: cut the step
// at the points where it becomes irregular, stream-compact the irregular work into queues in
// HBM, and run every stage as a dense kernel over its queue.
//
//   memset   counters
//   K_A  integrate  (thread = env)       Euler + gravity + engines + exact broadphase; envs with
//                                        candidate pairs write their world polygons + terrain to
//                                        a 192-byte record and append (env, pair) items
//   K_B1 gjk        (thread = item)      GJK; hits append to the hit queue, their envs to the
//                                        contact-env queue (first hit of an env wins the slot)
//   K_B2 epa        (thread = hit)       EPA + contact point
//   K_C1 collide    (thread = contact env)  coins, random choice, sequential impulses
//   K_C2 finish     (thread = env)       key split, joints, judge, reward/done; finished envs
//                                        append to the reset queue
//   K_R  reset      (thread = finished env)  LunarLander.__init__ (threefry or Philox)
//
// Queue counters live in device memory; the queue kernels are launched with a fixed grid (a
// multiple of the SM count) and loop over the count they read on the device, so the whole step
// is 1 memset + 6 launches with no host synchronisation and is CUDA-Graph capturable.  All
// stages call the same device functions as the fused kernel (cx_lunar_step.cuh): results are
// identical bit for bit.
#include "cx_launch.cuh"
#include "cx_lunar_step.cuh"

namespace {

constexpr int TPB = 128;
constexpr int REC_FLOATS = 48;  // 192 B per env: polys 0..27, heights 28..35, knots 36..39
constexpr uint32_t BASE_INLINE = 0xFFFFFFFFu;
constexpr uint32_t F_QUEUED = 0x80000000u;
constexpr int F_CRASH_SHIFT = 25, F_RC_SHIFT = 26, F_LC_SHIFT = 27;

enum { CTR_ITEMS = 0, CTR_HITS = 1, CTR_CENV = 2, CTR_RESET = 3, CTR_OVERFLOW = 4, CTR_WORDS = 16 };

struct PipeWs {
  uint32_t* ctr;   // [CTR_WORDS]
  uint32_t* cand;  // [B] candidate mask | flags
  uint32_t* base;  // [B] first item slot, or BASE_INLINE
  float* rec;      // [B][REC_FLOATS]
  uint32_t* item;  // [cap] (env << 5) | pair id
  float* res;      // [6][cap] simplex after K_B1; pen.xy, cp.xy after K_B2
  uint32_t* hit;   // [cap] item slots whose GJK said yes
  uint32_t* cenv;  // [B] envs with at least one GJK hit
  uint32_t* renv;  // [B] envs to reset
  uint32_t cap;
};

CX_DEV uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
// warp-aggregated queue append; must be reached by all 32 lanes.  Returns the slot of this
// lane (valid only where pred).
CX_DEV uint32_t queue_push(uint32_t* counter, bool pred) {
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, pred);
  if (m == 0u) return 0u;
  const int leader = __ffs(m) - 1;
  uint32_t b = 0u;
  if ((int)(threadIdx.x & 31u) == leader) b = atomicAdd(counter, (uint32_t)__popc(m));
  b = __shfl_sync(0xFFFFFFFFu, b, leader);
  return b + (uint32_t)__popc(m & lanemask_lt());
}

// polygon source: the record K_A wrote for env e (world polygons in LOCAL vertex order; the
// re-sort of part.transform() is applied here, per use, where the lanes are dense)
struct RecPolys {
  const float4* r;
  const LunarDerived& D;
  CX_DEV void body(int which, Poly& p) const {
    body_raw(which, p);
    order_body_poly(which, p, D);
  }
  CX_DEV void body_raw(int which, Poly& p) const {
    if (which == 0) {
      const float4 a = r[0], b = r[1], c = r[2];
      p.n = 6;
      p.x[0] = a.x; p.y[0] = a.y; p.x[1] = a.z; p.y[1] = a.w;
      p.x[2] = b.x; p.y[2] = b.y; p.x[3] = b.z; p.y[3] = b.w;
      p.x[4] = c.x; p.y[4] = c.y; p.x[5] = c.z; p.y[5] = c.w;
    } else {
      const float4 a = r[which == 1 ? 3 : 5], b = r[which == 1 ? 4 : 6];
      p.n = 4;
      p.x[0] = a.x; p.y[0] = a.y; p.x[1] = a.z; p.y[1] = a.w;
      p.x[2] = b.x; p.y[2] = b.y; p.x[3] = b.z; p.y[3] = b.w;
      p.x[4] = p.y[4] = p.x[5] = p.y[5] = 0.0f;
    }
  }
  CX_DEV void ground_edge(int k, const CotixLunarParams& P, float& x0, float& h0, float& x1, float& h1) const {
    const float* f = reinterpret_cast<const float*>(r);
    const float4 q = r[9];
    const float kn[4] = {q.x, q.y, q.z, q.w};
    x0 = terrain_x(kn, P, k);
    x1 = terrain_x(kn, P, k + 1);
    h0 = f[28 + k];
    h1 = f[29 + k];
  }
};

// ---- K_A -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_integrate_kernel(int64_t B, const float* __restrict__ sin_, float* __restrict__ sout, int64_t ld,
                      const float* __restrict__ action, const __grid_constant__ CotixLunarParams P,
                      const LunarDerived D, const PipeWs ws) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  const bool valid = e < B;
  const bool inplace = (sin_ == sout);
  uint32_t cand = 0u;
  int cnt = 0;
  if (valid) {
    EnvRegs E;
    load_dyn(E, sin_, ld, e);
    load_terrain(E, sin_, ld, e);
    E.status = __float_as_uint(sin_[COTIX_F_STATUS * ld + e]);
    const uint32_t status0 = E.status;
    float am = 0.0f, as = 0.0f;
    if (action) {
      am = action[e];
      as = action[ld + e];
    }
    StepCtx cx;
    Poly wL, wr, wl;
    integrate_and_cull<false>(E, am, as, P, D, cx, wL, wr, wl);
    cand = cx.cand;
    cnt = __popc(cand);
    if (inplace) {
      // only what changed: the three poses and the lander's velocity
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        sout[(i * 6 + 0) * ld + e] = E.b[i].px;
        sout[(i * 6 + 1) * ld + e] = E.b[i].py;
        sout[(i * 6 + 4) * ld + e] = E.b[i].ang;
      }
      sout[2 * ld + e] = E.b[0].vx;
      sout[3 * ld + e] = E.b[0].vy;
      sout[5 * ld + e] = E.b[0].w;
      if (E.status != status0) sout[COTIX_F_STATUS * ld + e] = __uint_as_float(E.status);
    } else {
      store_dyn(E, sout, ld, e);
      store_terrain(E, sout, ld, e);
      sout[COTIX_F_KEY0 * ld + e] = sin_[COTIX_F_KEY0 * ld + e];
      sout[COTIX_F_KEY1 * ld + e] = sin_[COTIX_F_KEY1 * ld + e];
      sout[COTIX_F_STEPS * ld + e] = sin_[COTIX_F_STEPS * ld + e];
      sout[COTIX_F_STATUS * ld + e] = __uint_as_float(E.status);
      sout[COTIX_F_RETURN * ld + e] = sin_[COTIX_F_RETURN * ld + e];
      sout[COTIX_F_SHAPING * ld + e] = sin_[COTIX_F_SHAPING * ld + e];
    }
    if (cand != 0u) {
      float4* r = reinterpret_cast<float4*>(ws.rec + e * REC_FLOATS);
      r[0] = make_float4(wL.x[0], wL.y[0], wL.x[1], wL.y[1]);
      r[1] = make_float4(wL.x[2], wL.y[2], wL.x[3], wL.y[3]);
      r[2] = make_float4(wL.x[4], wL.y[4], wL.x[5], wL.y[5]);
      r[3] = make_float4(wr.x[0], wr.y[0], wr.x[1], wr.y[1]);
      r[4] = make_float4(wr.x[2], wr.y[2], wr.x[3], wr.y[3]);
      r[5] = make_float4(wl.x[0], wl.y[0], wl.x[1], wl.y[1]);
      r[6] = make_float4(wl.x[2], wl.y[2], wl.x[3], wl.y[3]);
      r[7] = make_float4(E.h[0], E.h[1], E.h[2], E.h[3]);
      r[8] = make_float4(E.h[4], E.h[5], E.h[6], E.h[7]);
      r[9] = make_float4(E.kn[0], E.kn[1], E.kn[2], E.kn[3]);
    }
  }
  // warp-aggregated allocation of `cnt` consecutive item slots per env
  const unsigned lane = threadIdx.x & 31u;
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int n = __shfl_up_sync(0xFFFFFFFFu, incl, d);
    if ((int)lane >= d) incl += n;
  }
  const int total = __shfl_sync(0xFFFFFFFFu, incl, 31);
  if (total == 0) {
    if (valid) ws.cand[e] = 0u;
    return;
  }
  uint32_t wbase = 0u;
  if (lane == 31u) wbase = atomicAdd(&ws.ctr[CTR_ITEMS], (uint32_t)total);
  wbase = __shfl_sync(0xFFFFFFFFu, wbase, 31);
  if (!valid) return;
  if (cnt == 0) {
    ws.cand[e] = 0u;
    return;
  }
  const uint32_t first = wbase + (uint32_t)(incl - cnt);
  if (first + (uint32_t)cnt <= ws.cap) {
    ws.cand[e] = cand;
    ws.base[e] = first;
    uint32_t m = cand, slot = first;
    while (m) {
      const int bit = __ffs(m) - 1;
      m &= m - 1u;
      ws.item[slot++] = ((uint32_t)e << 5) | (uint32_t)bit;
    }
  } else {
    // item queue overflow: this env evaluates its own pairs inside K_C1 (same results);
    // slots it was given inside the queue are marked empty
    for (uint32_t slot = first; slot < ws.cap && slot < first + (uint32_t)cnt; ++slot)
      ws.item[slot] = 0xFFFFFFFFu;
    ws.cand[e] = cand | F_QUEUED;
    ws.base[e] = BASE_INLINE;
    ws.cenv[atomicAdd(&ws.ctr[CTR_CENV], 1u)] = (uint32_t)e;
    atomicAdd(&ws.ctr[CTR_OVERFLOW], 1u);
  }
}

// ---- K_B1 ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_gjk_kernel(const __grid_constant__ CotixLunarParams P, const LunarDerived D, const PipeWs ws) {
  const uint32_t n = min(ws.ctr[CTR_ITEMS], ws.cap);
  const uint32_t stride = gridDim.x * TPB;
  const V2 start = mk2(P.init_dir[0], P.init_dir[1]);
  for (uint32_t i0 = blockIdx.x * TPB + (threadIdx.x & ~31u); i0 < n; i0 += stride) {
    const uint32_t i = i0 + (threadIdx.x & 31u);
    bool hit = false, first_hit = false;
    uint32_t e = 0u;
    if (i < n) {
      const uint32_t it = ws.item[i];
      if (it != 0xFFFFFFFFu) {
        e = it >> 5;
        Poly A, Bp;
        RecPolys src{reinterpret_cast<const float4*>(ws.rec + (size_t)e * REC_FLOATS), D};
        make_item_polys(src, (int)(it & 31u), P, A, Bp);
        V2 s0, s1, s2;
        bool capped;
        hit = gjk(A, Bp, start, P.gjk_max_iters, s0, s1, s2, capped);
        if (hit) {
          ws.res[0 * (size_t)ws.cap + i] = s0.x;
          ws.res[1 * (size_t)ws.cap + i] = s0.y;
          ws.res[2 * (size_t)ws.cap + i] = s1.x;
          ws.res[3 * (size_t)ws.cap + i] = s1.y;
          ws.res[4 * (size_t)ws.cap + i] = s2.x;
          ws.res[5 * (size_t)ws.cap + i] = s2.y;
          first_hit = (atomicOr(&ws.cand[e], F_QUEUED) & F_QUEUED) == 0u;
        } else {
          ws.res[0 * (size_t)ws.cap + i] = 0.0f;
          ws.res[1 * (size_t)ws.cap + i] = 0.0f;
          ws.res[2 * (size_t)ws.cap + i] = CX_NAN;
          ws.res[3 * (size_t)ws.cap + i] = CX_NAN;
        }
      }
    }
    const uint32_t hs = queue_push(&ws.ctr[CTR_HITS], hit);
    if (hit) ws.hit[hs] = i;
    const uint32_t cs = queue_push(&ws.ctr[CTR_CENV], first_hit);
    if (first_hit) ws.cenv[cs] = e;
  }
}

// ---- K_B2 ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_epa_kernel(const __grid_constant__ CotixLunarParams P, const LunarDerived D, const PipeWs ws) {
  const uint32_t n = ws.ctr[CTR_HITS];
  const uint32_t stride = gridDim.x * TPB;
  for (uint32_t h = blockIdx.x * TPB + threadIdx.x; h < n; h += stride) {
    const uint32_t i = ws.hit[h];
    const uint32_t it = ws.item[i];
    const uint32_t e = it >> 5;
    Poly A, Bp;
    RecPolys src{reinterpret_cast<const float4*>(ws.rec + (size_t)e * REC_FLOATS), D};
    make_item_polys(src, (int)(it & 31u), P, A, Bp);
    const V2 s0 = mk2(ws.res[0 * (size_t)ws.cap + i], ws.res[1 * (size_t)ws.cap + i]);
    const V2 s1 = mk2(ws.res[2 * (size_t)ws.cap + i], ws.res[3 * (size_t)ws.cap + i]);
    const V2 s2 = mk2(ws.res[4 * (size_t)ws.cap + i], ws.res[5 * (size_t)ws.cap + i]);
    int iters = A.n + Bp.n + 1;  // polygon_vs_polygon: min(48, nA + nB + 1), cotix/_contacts.py:295
    if (iters > 13) iters = 13;
    bool ecap;
    const V2 pen = epa<16>(A, Bp, s0, s1, s2, iters, ecap);
    const V2 cp = contact_from_edges(A, Bp);
    ws.res[0 * (size_t)ws.cap + i] = pen.x;
    ws.res[1 * (size_t)ws.cap + i] = pen.y;
    ws.res[2 * (size_t)ws.cap + i] = cp.x;
    ws.res[3 * (size_t)ws.cap + i] = cp.y;
  }
}

// contact source of one env: the queue results in HBM (or inline after a queue overflow)
struct QueueContacts {
  const PipeWs& ws;
  const CotixLunarParams& P;
  const LunarDerived& D;
  uint32_t e, cand, base;
  CX_DEV Contact inline_eval(int bit) const {
    Poly A, Bp;
    RecPolys src{reinterpret_cast<const float4*>(ws.rec + (size_t)e * REC_FLOATS), D};
    make_item_polys(src, bit, P, A, Bp);
    return poly_pair(A, Bp, mk2(P.init_dir[0], P.init_dir[1]), P.gjk_max_iters);
  }
  CX_DEV Contact get(int bit) const {
    if (base != BASE_INLINE) {
      const size_t i = base + (uint32_t)__popc(cand & ((1u << bit) - 1u));
      Contact c;
      c.pen = mk2(ws.res[0 * (size_t)ws.cap + i], ws.res[1 * (size_t)ws.cap + i]);
      c.cp = mk2(ws.res[2 * (size_t)ws.cap + i], ws.res[3 * (size_t)ws.cap + i]);
      c.status = 0u;
      return c;
    }
    return inline_eval(bit);
  }
  CX_DEV bool live(int bit) const {
    if (!((cand >> bit) & 1u)) return false;
    if (base != BASE_INLINE) {
      const size_t i = base + (uint32_t)__popc(cand & ((1u << bit) - 1u));
      const float x = ws.res[2 * (size_t)ws.cap + i], y = ws.res[3 * (size_t)ws.cap + i];
      return !((x != x) || (y != y));
    }
    return !contact_isnan(inline_eval(bit));
  }
};

// ---- K_C1 ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_collide_kernel(float* __restrict__ s, int64_t ld, const __grid_constant__ CotixLunarParams P,
                    const LunarDerived D, const PipeWs ws) {
  const uint32_t n = ws.ctr[CTR_CENV];
  const uint32_t stride = gridDim.x * TPB;
  for (uint32_t c = blockIdx.x * TPB + threadIdx.x; c < n; c += stride) {
    const uint32_t e = ws.cenv[c];
    EnvRegs E;
    load_dyn(E, s, ld, e);
    E.key.a = __float_as_uint(s[COTIX_F_KEY0 * ld + e]);
    E.key.b = __float_as_uint(s[COTIX_F_KEY1 * ld + e]);
    E.steps = __float_as_uint(s[COTIX_F_STEPS * ld + e]);
    E.status = __float_as_uint(s[COTIX_F_STATUS * ld + e]);
    const uint32_t status0 = E.status;
    const uint32_t cand = ws.cand[e] & CX_PAIR_BITS;
    QueueContacts qc{ws, P, D, e, cand, ws.base[e]};
    bool crash, rc, lc;
    collide(E, P, cand, qc, crash, rc, lc);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      s[(i * 6 + 2) * ld + e] = E.b[i].vx;
      s[(i * 6 + 3) * ld + e] = E.b[i].vy;
      s[(i * 6 + 5) * ld + e] = E.b[i].w;
    }
    if (E.status != status0) s[COTIX_F_STATUS * ld + e] = __uint_as_float(E.status);
    ws.cand[e] = cand | F_QUEUED | ((crash ? 1u : 0u) << F_CRASH_SHIFT) | ((rc ? 1u : 0u) << F_RC_SHIFT) |
                 ((lc ? 1u : 0u) << F_LC_SHIFT);
  }
}

// ---- K_C2 ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_finish_kernel(int64_t B, float* __restrict__ s, int64_t ld, const float* __restrict__ action,
                   float* __restrict__ reward, uint8_t* __restrict__ done, float* __restrict__ obs,
                   const __grid_constant__ CotixLunarParams P, const PipeWs ws) {
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  const bool valid = e < B;
  bool want_reset = false;
  if (valid) {
    EnvRegs E;
    load_dyn(E, s, ld, e);
    load_meta(E, s, ld, e);
    float am = 0.0f, as = 0.0f;
    if (action) {
      am = action[e];
      as = action[ld + e];
    }
    const uint32_t cw = ws.cand[e];
    const bool crash = (cw >> F_CRASH_SHIFT) & 1u, rc = (cw >> F_RC_SHIFT) & 1u, lc = (cw >> F_LC_SHIFT) & 1u;
    float cs[3], sn[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) cx_sincosf(E.b[i].ang, &sn[i], &cs[i]);
    SubInfo si;
    finish_step(E, P, cs, sn, jclip(am, 0.0f, 1.0f), jclip(as, -1.0f, 1.0f), crash, rc, lc, si);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      s[(i * 6 + 2) * ld + e] = E.b[i].vx;
      s[(i * 6 + 3) * ld + e] = E.b[i].vy;
      s[(i * 6 + 5) * ld + e] = E.b[i].w;
    }
    store_meta(E, s, ld, e);
    if (reward) reward[e] = si.reward;
    if (done) done[e] = (uint8_t)si.done;
    want_reset = (si.done != COTIX_DONE_NO) && P.auto_reset;
    if (obs && !want_reset) store_obs(E, obs, ld, e, si.rc, si.lc);
  }
  const uint32_t slot = queue_push(&ws.ctr[CTR_RESET], want_reset);
  if (want_reset) ws.renv[slot] = (uint32_t)e;
}

// ---- K_R -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_reset_kernel(float* __restrict__ s, int64_t ld, float* __restrict__ obs,
                  const __grid_constant__ CotixLunarParams P, const PipeWs ws) {
  const uint32_t n = ws.ctr[CTR_RESET];
  const uint32_t stride = gridDim.x * TPB;
  for (uint32_t r = blockIdx.x * TPB + threadIdx.x; r < n; r += stride) {
    const uint32_t e = ws.renv[r];
    EnvRegs E;
    E.key.a = __float_as_uint(s[COTIX_F_KEY0 * ld + e]);
    E.key.b = __float_as_uint(s[COTIX_F_KEY1 * ld + e]);
    E.status = __float_as_uint(s[COTIX_F_STATUS * ld + e]);
    lunar_auto_reset(E, P);
    store_env(E, s, ld, e, true);
    if (obs) store_obs(E, obs, ld, e, false, false);
  }
}

// optional per-stage timing (bench.py's roofline of the dominant kernel): CUDA events recorded
// between the stages on the launching stream while profiling is enabled
constexpr int N_STAGES = 6;
bool g_profile = false;
cudaEvent_t g_ev[N_STAGES + 1];
bool g_ev_made = false;
double g_stage_ms[N_STAGES];
int64_t g_stage_steps = 0;
bool g_ev_pending = false;

void profile_collect() {
  if (!g_ev_pending) return;
  if (cudaEventSynchronize(g_ev[N_STAGES]) == cudaSuccess) {
    for (int k = 0; k < N_STAGES; ++k) {
      float ms = 0.0f;
      if (cudaEventElapsedTime(&ms, g_ev[k], g_ev[k + 1]) == cudaSuccess) g_stage_ms[k] += ms;
    }
    ++g_stage_steps;
  }
  g_ev_pending = false;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
uint32_t item_capacity(int64_t B) { return (uint32_t)(4 * B + 4096); }

// carve the workspace; returns the bytes needed
size_t carve(int64_t B, void* base, PipeWs* ws) {
  const size_t cap = item_capacity(B);
  size_t off = 0;
  char* p = static_cast<char*>(base);
  auto take = [&](size_t bytes) {
    char* q = p ? p + off : nullptr;
    off = align_up(off + bytes, 256);
    return q;
  };
  uint32_t* ctr = reinterpret_cast<uint32_t*>(take(CTR_WORDS * 4));
  uint32_t* cand = reinterpret_cast<uint32_t*>(take((size_t)B * 4));
  uint32_t* bs = reinterpret_cast<uint32_t*>(take((size_t)B * 4));
  float* rec = reinterpret_cast<float*>(take((size_t)B * REC_FLOATS * 4));
  uint32_t* item = reinterpret_cast<uint32_t*>(take(cap * 4));
  float* res = reinterpret_cast<float*>(take(cap * 6 * 4));
  uint32_t* hit = reinterpret_cast<uint32_t*>(take(cap * 4));
  uint32_t* cenv = reinterpret_cast<uint32_t*>(take((size_t)B * 4));
  uint32_t* renv = reinterpret_cast<uint32_t*>(take((size_t)B * 4));
  if (ws) {
    ws->ctr = ctr; ws->cand = cand; ws->base = bs; ws->rec = rec; ws->item = item; ws->res = res;
    ws->hit = hit; ws->cenv = cenv; ws->renv = renv; ws->cap = (uint32_t)cap;
  }
  return off;
}

}  // namespace

extern "C" int64_t cx_lunar_pipeline_workspace_bytes(int64_t B) {
  if (B <= 0) return 0;
  return (int64_t)carve(B, nullptr, nullptr);
}

extern "C" int cx_launch_lunar_pipeline(cudaStream_t st, int64_t B, const float* sin_, float* sout,
                                        int64_t ld, const float* action, float* reward, uint8_t* done,
                                        float* obs, const CotixLunarParams* p, void* workspace,
                                        int64_t workspace_bytes) {
  if (B == 0) return 0;
  if (B >= (1ll << 27)) {
    cx_set_error("lunar pipeline: B must be < 2^27");
    return COTIX_EINVAL;
  }
  PipeWs ws;
  const size_t need = carve(B, workspace, &ws);
  if (!workspace || (size_t)workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255u)) {
    cx_set_error("lunar pipeline: workspace must be 256-byte aligned and >= %zu bytes", need);
    return COTIX_EINVAL;
  }
  LunarDerived D;
  cx_lunar_derive(p, &D);
  const unsigned env_blocks = (unsigned)((B + TPB - 1) / TPB);
  const int sms = cx_sm_count();
  auto qgrid = [&](int64_t max_items) {
    int64_t b = (max_items + TPB - 1) / TPB;
    const int64_t cap = (int64_t)sms * 8;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
  };
  if (cudaMemsetAsync(ws.ctr, 0, CTR_WORDS * 4, st) != cudaSuccess) return cx_check_launch("pipeline memset");
  const bool prof = g_profile;
  if (prof) {
    profile_collect();
    if (!g_ev_made) {
      for (int k = 0; k <= N_STAGES; ++k) cudaEventCreate(&g_ev[k]);
      g_ev_made = true;
    }
  }
#define CX_STAGE_MARK(k) \
  if (prof) cudaEventRecord(g_ev[k], st)
  CX_STAGE_MARK(0);
  pipe_integrate_kernel<<<env_blocks, TPB, 0, st>>>(B, sin_, sout, ld, action, *p, D, ws);
  CX_STAGE_MARK(1);
  pipe_gjk_kernel<<<qgrid((int64_t)ws.cap), TPB, 0, st>>>(*p, D, ws);
  CX_STAGE_MARK(2);
  pipe_epa_kernel<<<qgrid((int64_t)ws.cap), TPB, 0, st>>>(*p, D, ws);
  CX_STAGE_MARK(3);
  pipe_collide_kernel<<<qgrid(B), TPB, 0, st>>>(sout, ld, *p, D, ws);
  CX_STAGE_MARK(4);
  pipe_finish_kernel<<<env_blocks, TPB, 0, st>>>(B, sout, ld, action, reward, done, obs, *p, ws);
  CX_STAGE_MARK(5);
  if (p->auto_reset) {
    pipe_reset_kernel<<<qgrid(B), TPB, 0, st>>>(sout, ld, obs, *p, ws);
    cx_count_launch_n(6);
  } else {
    cx_count_launch_n(5);
  }
  CX_STAGE_MARK(6);
#undef CX_STAGE_MARK
  if (prof) g_ev_pending = true;
  return cx_check_launch("lunar pipeline");
}

// ---- per-stage timing (single-threaded use: bench.py) ----------------------------------------
extern "C" int cx_lunar_pipeline_profile(int enable) {
  if (!enable) profile_collect();
  g_profile = enable != 0;
  if (enable) {
    for (int k = 0; k < N_STAGES; ++k) g_stage_ms[k] = 0.0;
    g_stage_steps = 0;
    g_ev_pending = false;
  }
  return COTIX_OK;
}
extern "C" int64_t cx_lunar_pipeline_stage_ms(double* out) {
  profile_collect();
  for (int k = 0; k < N_STAGES; ++k) out[k] = g_stage_steps ? g_stage_ms[k] / (double)g_stage_steps : 0.0;
  return g_stage_steps;
}
