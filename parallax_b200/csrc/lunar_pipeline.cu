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
//   memset   counters (once per enqueue; between steps the finish stage zeroes them)
//   K_A  integrate  (thread = env)        Euler + gravity + engines + exact broadphase; envs with
//                                         candidate pairs write their world polygons + terrain to
//                                         a 192-byte record, append (env, pair) items and join
//                                         the contact-env queue
//   K_B  narrow     (thread = item)       vertex re-sort, GJK, EPA + contact point
//   K_C  contact    (thread = contact env; coins pooled per warp)  coin flips, random choice,
//                                         sequential impulses
//   K_C2 finish     (thread = env)        key split, joints, judge, reward/done; finished envs
//                                         append to the reset queue
//   K_R  reset      (thread = finished env)  LunarLander.__init__ (threefry or Philox)
//
// Queue counters live in device memory; the queue kernels are launched with a fixed grid (a
// multiple of the SM count) and loop over the count they read on the device, so the whole step
// is 6 launches (+ 1 memset per enqueue) with no host synchronisation and is CUDA-Graph capturable.  All
// stages call the same device functions as the fused kernel (cx_lunar_step.cuh): results are
// identical bit for bit.
#include <mutex>

#include "cx_launch.cuh"
#include "cx_lunar_coll.cuh"
#include "cx_lunar_step.cuh"
#include "cx_pipeline.h"

namespace {

// threads per block of the env-indexed stages.  64, not 128: the integrate stage has two block barriers (queue
// allocation), and with two warps per block instead of four fewer warps wait for the block's slowest one
// (1 M envs: 196.4 -> 192.1 us per step; 32: the contact stage loses 12 us; 256: integrate 72 -> 112 us)
#ifndef CX_PIPE_TPB
#define CX_PIPE_TPB 64
#endif
constexpr int TPB = CX_PIPE_TPB;
// Lanes of a warp that take work in the two sparse stages = 32 / SPREAD (template parameter, chosen per
// launch by sparse_spread()).  At 1 M envs these stages are bound by warp-instruction throughput and
// SPREAD = 1 is best (2 / 4: 61 -> 117 -> 251 us); at a few thousand envs they are ONE thin wave whose time
// is the longest warp's chain, a warp serialises the divergent paths of its lanes, and the SMs are empty:
// fewer items per warp make the chain shorter.
// block size of the narrow stage: the items a block takes are binned by pair class, so a bigger block
// has fewer warps that straddle a class boundary (3 boundaries per block)
#ifndef CX_NARROW_TPB
#define CX_NARROW_TPB 128
#endif
constexpr int NTPB = CX_NARROW_TPB;
constexpr int REC_FLOATS = 48;  // 192 B per env: polys 0..27, heights 28..35, knots 36..39
constexpr uint32_t BASE_INLINE = 0xFFFFFFFFu;
constexpr int F_CRASH_SHIFT = 25, F_RC_SHIFT = 26, F_LC_SHIFT = 27;

// Queue counters.  They are zeroed by a memset once per enqueue; between the steps of a multi-step enqueue the
// FINISH stage zeroes them for the next step (after copying them to CTR_SNAP.. for cx_lunar_pipeline_counters),
// so that a step is kernels only and every stage can be a programmatic dependent launch of the one before.  The
// reset queue's counter is written by finish and read by the reset stage that follows it, so it alternates
// between two words with the parity of the step: finish(t) zeroes the one reset(t - 1) has long finished with.
enum { CTR_ITEMS = 0, CTR_HITS = 1, CTR_CENV = 2, CTR_RESET = 3, CTR_OVERFLOW = 4, CTR_RESET_ODD = 5, CTR_SNAP = 8,
       CTR_WORDS = 16 };

// minimum resident blocks per SM asked of ptxas for the two dense kernels (register caps)
#ifndef CX_INTEGRATE_MINB
// 640 threads per SM = 96 registers per thread: the kernel keeps its 28 polygon floats next to the state without
// squeezing (80 registers at 768 threads: 70.0 us; 96 at 576 - 640: 65.2 - 65.9; 119 at 512: 71.2)
#define CX_INTEGRATE_MINB (640 / CX_PIPE_TPB)
#endif
#ifndef CX_FINISH_MINB
#define CX_FINISH_MINB (1024 / CX_PIPE_TPB)  // 64 registers: 72.5 -> 64 us (profiles/r2h_variants.jsonl); integrate is unchanged by 5..8 blocks of 128
#endif

struct PipeWs {
  uint32_t* ctr;   // [CTR_WORDS]
  uint32_t* cand;  // [B] candidate mask | flags
  uint32_t* base;  // [B] first item slot, or BASE_INLINE
  float* rec;      // [B][REC_FLOATS]
  uint32_t* item;  // [cap] (env << 5) | pair id
  float* res;      // [6][cap] simplex after K_B1; pen.xy, cp.xy after K_B2
  uint32_t* hit;   // [cap >= B] envs whose items overflowed the item queue (list of CTR_OVERFLOW entries)
  uint32_t* cenv;  // [B] envs with at least one GJK hit
  uint32_t* renv;  // [B] envs to reset
  float* racc;     // [B] reward accumulated over the sub-steps of this call (K > 1)
  uint32_t* latch; // [B] done code | rc << 8 | lc << 9 of an env that finished in an earlier sub-step
  uint32_t cap;
  uint32_t nenv;   // envs of this range
};

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
__global__ void __launch_bounds__(TPB, CX_INTEGRATE_MINB)
pipe_integrate_kernel(int64_t B, const float* __restrict__ sin_, float* __restrict__ sout, int64_t ld,
                      const float* __restrict__ action, const __grid_constant__ CotixLunarParams P,
                      const LunarDerived D, const PipeWs ws, int k, int K) {
  cudaGridDependencySynchronize();  // after the first step of an enqueue this stage follows reset / finish as a PDL
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  // sub-step k of K: an env that finished in an earlier sub-step of this call is frozen
  // (cotix/_envs.py:81-103) — nothing of it is read, written or queued
  const bool valid = e < B && !(k > 0 && (ws.latch[e] & 0xFFu) != 0u);
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
  // block-aggregated allocation of `cnt` consecutive item slots per env and of one contact-env slot
  // per env with candidates: ONE atomic per counter and block (the counters are hot addresses —
  // same-address atomics retire at about one per nanosecond, 2 x 30 k of them per step otherwise)
  __shared__ uint32_t s_items[TPB / 32], s_envs[TPB / 32], s_base[2];
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int n = __shfl_up_sync(0xFFFFFFFFu, incl, d);
    if ((int)lane >= d) incl += n;
  }
  const uint32_t em = __ballot_sync(0xFFFFFFFFu, cnt != 0);
  if (lane == 31u) {
    s_items[warp] = (uint32_t)incl;
    s_envs[warp] = (uint32_t)__popc(em);
  }
  if (e < B && !valid) ws.cand[e] = 0u;  // frozen
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t ti = 0u, te = 0u;
#pragma unroll
    for (int j = 0; j < TPB / 32; ++j) {
      ti += s_items[j];
      te += s_envs[j];
    }
    s_base[0] = ti ? atomicAdd(&ws.ctr[CTR_ITEMS], ti) : 0u;
    s_base[1] = te ? atomicAdd(&ws.ctr[CTR_CENV], te) : 0u;
  }
  __syncthreads();
  uint32_t wbase = s_base[0], cslot = s_base[1];
  for (unsigned j = 0; j < warp; ++j) {
    wbase += s_items[j];
    cslot += s_envs[j];
  }
  cslot += (uint32_t)__popc(em & lanemask_lt());
  // every env with candidates goes to the contact-env queue (K_C finishes those envs itself)
  if (cnt != 0) ws.cenv[cslot] = (uint32_t)e;
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
    ws.cand[e] = cand;
    ws.base[e] = BASE_INLINE;
    ws.hit[atomicAdd(&ws.ctr[CTR_OVERFLOW], 1u)] = (uint32_t)e;  // pipe_contact_kernel<true> takes these
  }
}

// ---- K_B -----------------------------------------------------------------------------------
// GJK, and for hits (82 % of the candidate pairs on the bench state: resting leg/lander and
// leg/ground contacts persist) EPA + contact point, in one pass over the item queue.
template <int SPREAD>
__global__ void __launch_bounds__(NTPB)
pipe_narrow_kernel(const __grid_constant__ CotixLunarParams P, const LunarDerived D, const PipeWs ws) {
  cudaGridDependencySynchronize();  // programmatic dependent launch: the previous stage is complete past here
  const uint32_t n = min(ws.ctr[CTR_ITEMS], ws.cap);
  // The stage is ONE wave that fills a third of the thread slots, so its time is the longest warp's
  // instruction chain, and a warp serialises the divergent paths of its lanes: a block takes only
  // IPB = NTPB / SPREAD items and gives every warp 32 / SPREAD of them (the other
  // lanes idle) -- more, shorter warps on SM slots that would otherwise be empty.
  constexpr uint32_t LPW = 32u / SPREAD, IPB = NTPB / SPREAD;
  const uint32_t stride = gridDim.x * IPB;
  const V2 start = mk2(P.init_dir[0], P.init_dir[1]);
  uint32_t hits = 0u;
  // The NTPB items a block takes are binned by pair class (leg-lander, ground-lander, ground-leg,
  // leg-leg) through a counting sort in shared memory before they are handed to the threads, so
  // that a warp mostly runs one kind of pair: same polygon sizes, same EPA trip bound, ground
  // polygon built or not.  Results go to the item's own slot, so nothing downstream changes.
  __shared__ uint32_t s_slot[NTPB];
  __shared__ uint32_t s_cnt[4];
  for (uint32_t base = blockIdx.x * IPB; base < n; base += stride) {
    if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t i0 = base + threadIdx.x;
    uint32_t it0 = 0xFFFFFFFFu;
    if (threadIdx.x < IPB && i0 < n) it0 = ws.item[i0];
    int cls = -1;
    uint32_t rank = 0u;
    if (it0 != 0xFFFFFFFFu) {
      const uint32_t pid = it0 & 31u;
      cls = pid <= 1u ? 0 : (pid <= 8u ? 1 : (pid <= 10u ? 3 : 2));
      rank = atomicAdd(&s_cnt[cls], 1u);
    }
    __syncthreads();
    const uint32_t c0 = s_cnt[0], c1 = s_cnt[1], c2 = s_cnt[2], c3 = s_cnt[3];
    if (cls >= 0) {
      const uint32_t off = cls == 0 ? 0u : (cls == 1 ? c0 : (cls == 2 ? c0 + c1 : c0 + c1 + c2));
      s_slot[off + rank] = i0;
    }
    __syncthreads();
    const uint32_t lane_ = threadIdx.x & 31u, j = (threadIdx.x >> 5) * LPW + lane_;
    if (lane_ < LPW && j < c0 + c1 + c2 + c3) {
      const uint32_t i = s_slot[j];
      const uint32_t it = ws.item[i];
      const uint32_t e = it >> 5;
      Poly A, Bp;
      RecPolys src{reinterpret_cast<const float4*>(ws.rec + (size_t)e * REC_FLOATS), D};
      make_item_polys(src, (int)(it & 31u), P, A, Bp);
      const Contact c = poly_pair(A, Bp, start, P.gjk_max_iters);
      ws.res[0 * (size_t)ws.cap + i] = c.pen.x;
      ws.res[1 * (size_t)ws.cap + i] = c.pen.y;
      ws.res[2 * (size_t)ws.cap + i] = c.cp.x;
      ws.res[3 * (size_t)ws.cap + i] = c.cp.y;
      if (!contact_isnan(c)) ++hits;
    }
  }
  // statistics only (scripts/exp_pipeline.py)
  hits = __reduce_add_sync(0xFFFFFFFFu, hits);
  if ((threadIdx.x & 31u) == 0u && hits) atomicAdd(&ws.ctr[CTR_HITS], hits);
}

// contact source of one env: the queue results in HBM (INLINE = false), or, for an env whose items
// overflowed the queue, its pairs evaluated on the spot (INLINE = true).  Two instantiations of the
// contact kernel: the narrowphase code (90 KB of SASS) costs the common one 12 us per step when it is
// compiled in, although it never runs there (the stage is instruction-fetch bound).
template <bool INLINE>
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
    if (!INLINE) {
      const size_t i = base + (uint32_t)__popc(cand & ((1u << bit) - 1u));
      Contact c;
      c.pen = mk2(ws.res[0 * (size_t)ws.cap + i], ws.res[1 * (size_t)ws.cap + i]);
      c.cp = mk2(ws.res[2 * (size_t)ws.cap + i], ws.res[3 * (size_t)ws.cap + i]);
      c.status = 0u;
      return c;
    }
    return inline_eval(bit);
  }
  // pairs with a non-NaN contact; the env's items sit in consecutive queue slots
  CX_DEV uint32_t live_bits() const {
    uint32_t live = 0u;
    if (!INLINE) {
      const float* cpx = ws.res + 2 * (size_t)ws.cap + base;
      const float* cpy = ws.res + 3 * (size_t)ws.cap + base;
      uint32_t todo = cand;
      int k = 0;
      while (todo) {
        const int bit = __ffs(todo) - 1;
        todo &= todo - 1u;
        const float x = cpx[k], y = cpy[k];
        ++k;
        if (!((x != x) || (y != y))) live |= 1u << bit;
      }
    } else {
      uint32_t todo = cand;
      while (todo) {
        const int bit = __ffs(todo) - 1;
        todo &= todo - 1u;
        if (!contact_isnan(inline_eval(bit))) live |= 1u << bit;
      }
    }
    return live;
  }
  CX_DEV bool live(int bit) const {
    if (!((cand >> bit) & 1u)) return false;
    if (!INLINE) {
      const size_t i = base + (uint32_t)__popc(cand & ((1u << bit) - 1u));
      const float x = ws.res[2 * (size_t)ws.cap + i], y = ws.res[3 * (size_t)ws.cap + i];
      return !((x != x) || (y != y));
    }
    return !contact_isnan(inline_eval(bit));
  }
};

// key split, joints, judge and the stores of one env whose dynamic state and meta are in E
// (cotix_lunar_step's tail), for sub-step k of K: rewards add up over the sub-steps, an env that
// finishes is latched (done code and leg contacts) and frozen, and reward / done / obs leave with
// the last sub-step.  Returns true when the env must be reset (auto-reset on, last sub-step).
CX_DEV bool finish_and_store(EnvRegs& E, int64_t e, float* __restrict__ s, int64_t ld,
                             const float* __restrict__ action, float* __restrict__ reward,
                             uint8_t* __restrict__ done, float* __restrict__ obs,
                             const CotixLunarParams& P, const LunarDerived& D, bool crash, bool rc, bool lc,
                             const PipeWs& ws, int k, int K) {
  float am = 0.0f, as = 0.0f;
  if (action) {
    am = action[e];
    as = action[ld + e];
  }
  float cs[3], sn[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) cx_sincosf(E.b[i].ang, &sn[i], &cs[i]);
  SubInfo si;
  finish_step<true>(E, P, D, cs, sn, jclip(am, 0.0f, 1.0f), jclip(as, -1.0f, 1.0f), crash, rc, lc, si);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    s[(i * 6 + 2) * ld + e] = E.b[i].vx;
    s[(i * 6 + 3) * ld + e] = E.b[i].vy;
    s[(i * 6 + 5) * ld + e] = E.b[i].w;
  }
  store_meta(E, s, ld, e);
  float rsum = si.reward;
  if (K > 1) {
    rsum = (k > 0 ? ws.racc[e] : 0.0f) + si.reward;
    if (k + 1 < K) {
      ws.racc[e] = rsum;
      ws.latch[e] = (uint32_t)si.done | ((si.rc ? 1u : 0u) << 8) | ((si.lc ? 1u : 0u) << 9);
      return false;
    }
  }
  if (reward) reward[e] = rsum;
  if (done) done[e] = (uint8_t)si.done;
  const bool want_reset = (si.done != COTIX_DONE_NO) && P.auto_reset;
  if (obs && !want_reset) store_obs(E, obs, ld, e, si.rc, si.lc);
  return want_reset;
}

// ---- K_C1 (the warp-pooled coin flips: cx_lunar_coll.cuh) ------------------------------------
template <bool INLINE, int SPREAD>
__global__ void __launch_bounds__(TPB)
pipe_contact_kernel(float* __restrict__ s, int64_t ld, const __grid_constant__ CotixLunarParams P,
                    const LunarDerived D, const PipeWs ws) {
  cudaGridDependencySynchronize();  // programmatic dependent launch: the previous stage is complete past here
  __shared__ WarpColl Wall[TPB / 32];
  WarpColl& W = Wall[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  // INLINE: the (rare) envs of the overflow list; otherwise the contact-env queue minus those envs
  const uint32_t n = INLINE ? min(ws.ctr[CTR_OVERFLOW], (uint32_t)ws.nenv) : ws.ctr[CTR_CENV];
  if (INLINE && n == 0u) return;
  const uint32_t* __restrict__ list = INLINE ? ws.hit : ws.cenv;
  constexpr uint32_t CLPW = 32u / SPREAD;  // envs per warp (the other lanes only help flipping coins)
  const uint32_t stride = gridDim.x * (TPB / 32u) * CLPW;
  for (uint32_t c0 = (blockIdx.x * (TPB / 32u) + (threadIdx.x >> 5)) * CLPW; c0 < n; c0 += stride) {
    const uint32_t c = c0 + (uint32_t)lane;
    bool valid = (uint32_t)lane < CLPW && c < n;
    uint32_t e = 0u, cand = 0u, status0 = 0u;
    EnvRegs E;
    uint32_t m[7] = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
    bool crash = false, rc = false, lc = false, any = false;
    CoinCtx cc;
    uint32_t base = 0u;
    if (valid) {
      e = list[c];
      base = ws.base[e];
      if (!INLINE && base == BASE_INLINE) valid = false;  // left to the INLINE instantiation
    }
    if (valid) {
      cand = ws.cand[e] & CX_PAIR_BITS;
      E.key.a = __float_as_uint(s[COTIX_F_KEY0 * ld + e]);
      E.key.b = __float_as_uint(s[COTIX_F_KEY1 * ld + e]);
      E.steps = __float_as_uint(s[COTIX_F_STEPS * ld + e]);
      E.status = __float_as_uint(s[COTIX_F_STATUS * ld + e]);
      status0 = E.status;
    }
    QueueContacts<INLINE> qc{ws, P, D, e, cand, base};
    if (valid) any = cell_masks(qc.live_bits(), true, m, crash, rc, lc);
    if (any) {
      make_coin_ctx(E, P, cc);
      W.gk[lane][0] = cc.gkey[0].a; W.gk[lane][1] = cc.gkey[0].b;
      W.gk[lane][2] = cc.gkey[1].a; W.gk[lane][3] = cc.gkey[1].b;
      W.pk[lane][0] = cc.pkey.a; W.pk[lane][1] = cc.pkey.b;
      W.st[lane][0] = cc.steps; W.st[lane][1] = cc.episode;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) W.win[lane][k] = -1;
    int head = 0, tail = 0;
#pragma unroll
    for (int k = 0; k < 6; ++k) coll_push(W, tail, m[k] != 0u, (uint32_t)lane | ((uint32_t)k << 5), m[k], 0u);
    coll_rounds(W, head, tail, P);
    // the self pair only matters when body 2 has another live cell
    coll_push(W, tail, m[6] != 0u && (W.win[lane][1] >= 0 || W.win[lane][3] >= 0),
              (uint32_t)lane | (6u << 5), m[6], 0u);
    coll_rounds(W, head, tail, P);
    if (valid) {
      if (any) {
        int win[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) win[k] = W.win[lane][k];
        load_dyn(E, s, ld, e);  // only envs with a live contact need their bodies
        // the common instantiation inlines resolve_collision (constant-bank masses, exact div_c): 65.6 -> 63.5 us
        collide_apply<QueueContacts<INLINE>, !INLINE>(E, P, D, qc, cc, win);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          s[(i * 6 + 2) * ld + e] = E.b[i].vx;
          s[(i * 6 + 3) * ld + e] = E.b[i].vy;
          s[(i * 6 + 5) * ld + e] = E.b[i].w;
        }
        if (E.status != status0) s[COTIX_F_STATUS * ld + e] = __uint_as_float(E.status);
      }
      if (crash || rc || lc)
        ws.cand[e] = cand | ((crash ? 1u : 0u) << F_CRASH_SHIFT) | ((rc ? 1u : 0u) << F_RC_SHIFT) |
                     ((lc ? 1u : 0u) << F_LC_SHIFT);
    }
    __syncwarp();
  }
}

// ---- K_C2 ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB, CX_FINISH_MINB)
pipe_finish_kernel(int64_t B, float* __restrict__ s, int64_t ld, const float* __restrict__ action,
                   float* __restrict__ reward, uint8_t* __restrict__ done, float* __restrict__ obs,
                   const __grid_constant__ CotixLunarParams P, const __grid_constant__ LunarDerived D, const PipeWs ws,
                   int k, int K, int par) {
  cudaGridDependencySynchronize();  // programmatic dependent launch: the previous stage is complete past here
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // the narrow / contact stages are done with their counters: keep a copy for the host, zero them for the next
    // step's integrate stage (no memset node between the steps)
    ws.ctr[CTR_SNAP + 0] = ws.ctr[CTR_ITEMS];
    ws.ctr[CTR_SNAP + 1] = ws.ctr[CTR_HITS];
    ws.ctr[CTR_SNAP + 2] = ws.ctr[CTR_CENV];
    ws.ctr[CTR_SNAP + 3] = 0u;  // the reset stage, when there is one, puts its count here
    ws.ctr[CTR_SNAP + 4] = ws.ctr[CTR_OVERFLOW];
    ws.ctr[CTR_ITEMS] = ws.ctr[CTR_HITS] = ws.ctr[CTR_CENV] = ws.ctr[CTR_OVERFLOW] = 0u;
    ws.ctr[par ? CTR_RESET : CTR_RESET_ODD] = 0u;  // the OTHER step parity's reset counter
  }
  const int64_t e = (int64_t)blockIdx.x * TPB + threadIdx.x;
  bool want_reset = false;
  if (e < B) {
    const uint32_t latch = k > 0 ? ws.latch[e] : 0u;
    if ((latch & 0xFFu) != 0u) {
      // finished in an earlier sub-step: frozen; its outputs leave with the last sub-step
      if (k + 1 == K) {
        if (reward) reward[e] = ws.racc[e];
        if (done) done[e] = (uint8_t)(latch & 0xFFu);
        want_reset = P.auto_reset != 0;
        if (obs && !want_reset) {
          EnvRegs E;
          load_dyn(E, s, ld, e);
          store_obs(E, obs, ld, e, ((latch >> 8) & 1u) != 0u, ((latch >> 9) & 1u) != 0u);
        }
      }
    } else {
      EnvRegs E;
      load_dyn(E, s, ld, e);
      load_meta(E, s, ld, e);
      const uint32_t cw = ws.cand[e];
      want_reset = finish_and_store(E, e, s, ld, action, reward, done, obs, P, D, (cw >> F_CRASH_SHIFT) & 1u,
                                    (cw >> F_RC_SHIFT) & 1u, (cw >> F_LC_SHIFT) & 1u, ws, k, K);
    }
  }
  const uint32_t slot = queue_push(&ws.ctr[par ? CTR_RESET_ODD : CTR_RESET], want_reset);
  if (want_reset) ws.renv[slot] = (uint32_t)e;
}

// ---- K_R -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPB)
pipe_reset_kernel(float* __restrict__ s, int64_t ld, float* __restrict__ obs,
                  const __grid_constant__ CotixLunarParams P, const PipeWs ws, int par) {
  cudaGridDependencySynchronize();  // programmatic dependent launch: the previous stage is complete past here
  // eight lanes per finished env (lunar_auto_reset_coop): a reset is a chain of ~30 threefry calls,
  // and this stage is pure latency (a few hundred envs per step)
  const uint32_t n = ws.ctr[par ? CTR_RESET_ODD : CTR_RESET];
  if (blockIdx.x == 0 && threadIdx.x == 0) ws.ctr[CTR_SNAP + 3] = n;
  const uint32_t warp = (blockIdx.x * TPB + threadIdx.x) >> 5, nwarps = (gridDim.x * TPB) >> 5;
  const uint32_t grp = (threadIdx.x & 31u) >> 3, sub = threadIdx.x & 7u;
  for (uint32_t r0 = warp * 4u; r0 < n; r0 += nwarps * 4u) {
    const uint32_t r = r0 + grp;
    const bool valid = r < n;
    uint32_t e = 0u;
    EnvRegs E;
    E.key.a = E.key.b = 0u;
    E.status = 0u;
    if (valid) {
      e = ws.renv[r];
      E.key.a = __float_as_uint(s[COTIX_F_KEY0 * ld + e]);
      E.key.b = __float_as_uint(s[COTIX_F_KEY1 * ld + e]);
      E.status = __float_as_uint(s[COTIX_F_STATUS * ld + e]);
    }
    lunar_auto_reset_coop(E, P, valid);
    if (valid && sub == 0u) {
      store_env(E, s, ld, e, true);
      if (obs) store_obs(E, obs, ld, e, false, false);
    }
  }
}

// optional per-stage timing (bench.py's per-stage roofline objects): CUDA events recorded between
// the stages on the launching stream while profiling is enabled (one range, no PDL: serial times)
// stages: 0 integrate, 1 narrow, 2 contact, 3 finish, 4 reset, 5 whole step
constexpr int N_STAGES = 6;
bool g_profile = false;
cudaEvent_t g_evb[N_STAGES], g_eve[N_STAGES];
bool g_ev_made = false;
double g_stage_ms[N_STAGES];
int64_t g_stage_steps = 0;
bool g_ev_pending = false;

void profile_collect() {
  if (!g_ev_pending) return;
  if (cudaEventSynchronize(g_eve[N_STAGES - 1]) == cudaSuccess) {
    for (int k = 0; k < N_STAGES; ++k) {
      float ms = 0.0f;
      if (cudaEventElapsedTime(&ms, g_evb[k], g_eve[k]) == cudaSuccess) g_stage_ms[k] += ms;
    }
    ++g_stage_steps;
  }
  g_ev_pending = false;
}

// ---- timeline of a multi-range enqueue (measurement aid, eager launches only): an event after every
// kernel of every range; read back as milliseconds since the fork (cx_lunar_pipeline_trace_read)
constexpr int TR_STEPS = 8, TR_STAGES = 5;
bool g_trace = false;
cudaEvent_t g_tr_base = nullptr;
cudaEvent_t g_tr_ev[COTIX_MAX_CHUNKS][TR_STEPS][TR_STAGES];
int g_tr_S = 0, g_tr_T = 0;

// ---- launch configuration (cotix_lunar_pipeline_config) ------------------------------------
// The two sparse stages (narrow, contact) are bound by the latency of their slowest work item
// (about 32 us each however short the queue) and leave most of the GPU idle; the two dense stages
// are issue-bound.  Envs are independent, so the batch is cut into `split` RANGES of envs, each with
// its own queues and its own stream: the sparse stages of one range run under the dense stages of
// the others.  The ranges of a multi-step enqueue never join between steps.  Sparse kernels are
// launched at the highest priority so that they get SM slots as soon as blocks of a dense kernel
// retire instead of queueing behind its remaining blocks.
constexpr int MAX_SPLIT = COTIX_MAX_CHUNKS;
struct PipeCfg {
  int split;          // 0 = automatic
  int pdl;            // programmatic dependent launch between the stages of a range
  int prio;           // sparse stages at high launch priority
  uint32_t item_cap;  // 0 = 4 B + 4096 per range; tests force the overflow branch with a small value
  int min_range;      // automatic split: envs per range at least this many
};
PipeCfg g_cfg = {0, 1, 1, 0u, 131072};  // split 0 = automatic (see choose_split)

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
// the forced capacity (tests) can only shrink the queue: workspace sizes are those of the default
uint32_t item_capacity(int64_t B) {
  const uint32_t full = (uint32_t)(4 * B + 4096);
  return (g_cfg.item_cap && g_cfg.item_cap < full) ? g_cfg.item_cap : full;
}

// carve the workspace of one range; returns the bytes needed
size_t carve(int64_t B, void* base, PipeWs* ws) {
  const size_t cap = (size_t)(4 * B + 4096);  // layout by the default capacity
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
  float* racc = reinterpret_cast<float*>(take((size_t)B * 4));
  uint32_t* latch = reinterpret_cast<uint32_t*>(take((size_t)B * 4));
  if (ws) {
    ws->racc = racc;
    ws->latch = latch;
    ws->ctr = ctr; ws->cand = cand; ws->base = bs; ws->rec = rec; ws->item = item; ws->res = res;
    ws->hit = hit; ws->cenv = cenv; ws->renv = renv; ws->cap = item_capacity(B);
    ws->nenv = (uint32_t)B;
  }
  return off;
}

// lanes-per-warp divisor of the sparse stages for a range of n envs (see the note at the top);
// COTIX_SPARSE_SPREAD = 1 | 2 | 4 | 8 | 16 | 32 overrides (measurement aid)
int sparse_spread(int64_t n) {
  static const char* env = getenv("COTIX_SPARSE_SPREAD");
  if (env) {
    const int v = atoi(env);
    if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) return v;
  }
  // measured, 100-step contact-rich rollouts, spread 1 / 2 / 4 / 8 / 16 / 32: 1024 envs . / . / 5.56 / 5.26 / 4.93 / 4.55 ms,
  // 4096 envs 6.65 / 6.24 / 5.89 / 5.61 / 5.32 / 5.08, 16384 envs 7.02 / 6.58 / 6.30 / 6.20 / 6.24 / 7.18; steps of 65536
  // envs 67.8 / 64.9 / 63.0 / 63.8 / 69.8 us; 262144 envs 90.7 / 94.6 / 113.6 us
  return n <= 4096 ? 32 : (n <= 8192 ? 16 : (n <= 16384 ? 8 : (n <= 65536 ? 4 : 1)));
}

// the ranges of a batch: S ranges of ceil(B / S) envs rounded up to a multiple of TPB (so that every
// range starts on a 512-byte boundary of every state row), the last one takes what is left
struct RangePlan {
  int S;
  int64_t off[MAX_SPLIT + 1];
  PipeWs ws[MAX_SPLIT];
};
size_t plan_ranges(int64_t B, int S, void* workspace, RangePlan* plan) {
  int64_t len = (B + S - 1) / S;
  len = (len + TPB - 1) / TPB * TPB;
  size_t bytes = 0;
  int n = 0;
  for (int64_t o = 0; o < B && n < S; o += len, ++n) {
    const int64_t nb = (o + len <= B) ? len : B - o;
    char* base = workspace ? static_cast<char*>(workspace) + bytes : nullptr;
    bytes += carve(nb, base, plan ? &plan->ws[n] : nullptr);
    if (plan) plan->off[n] = o;
  }
  if (plan) {
    plan->S = n;
    plan->off[n] = B;
  }
  return bytes;
}
int choose_split(int64_t B) {
  // automatic = ONE range: measured on B200 at 1 M envs (profiles/r2e_split_sweep.jsonl), 1 / 2 / 4 / 8
  // ranges take 245 / 245 / 258 / 268 us per step -- co-running the stages of different ranges does not
  // raise the issue rate (profiles/r2e_timeline.txt), it only adds launches.  Ranges pay off for the
  // host-facing step, where they overlap PCIe copies with kernels (cotix_lunar_env_step_host).
  (void)B;
  int S = g_cfg.split;
  if (S > MAX_SPLIT) S = MAX_SPLIT;
  if (S < 1) S = 1;
  return S;
}

// side streams of the ranges 1 .. S-1 (range 0 runs on the caller's stream), one set per device, made
// on first use.  The fork / join events are shared by all callers on a device, so the enqueue of a
// multi-range step holds the device's mutex from fork to join.
struct PipeSide {
  cudaStream_t s[MAX_SPLIT];
  cudaEvent_t fork, join[MAX_SPLIT];
  int prio_lo, prio_hi;
  int state;  // 0 not made, 1 ok, -1 failed
  std::mutex mu;
};
PipeSide* pipe_side() {
  static PipeSide per_dev[64];
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  PipeSide& a = per_dev[dev];
  if (a.state == 0) {
    bool ok = cudaDeviceGetStreamPriorityRange(&a.prio_lo, &a.prio_hi) == cudaSuccess;
    for (int i = 1; i < MAX_SPLIT && ok; ++i)
      ok = cudaStreamCreateWithFlags(&a.s[i], cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 1; i < MAX_SPLIT && ok; ++i)
      ok = cudaEventCreateWithFlags(&a.join[i], cudaEventDisableTiming) == cudaSuccess;
    a.state = ok ? 1 : -1;
    if (!ok) cudaGetLastError();
  }
  return a.state == 1 ? &a : nullptr;
}

// which buffers step t of a multi-step enqueue uses
struct StepSeq {
  const float* sin0;   // state read by step 0
  float* sout0;        // state written by step 0
  int64_t stride;      // floats between the buffers of consecutive steps (0: the same buffers, in place)
  const float* action; // or NULL
  int64_t action_stride;
  int n_actions;       // step t uses action + (t % n_actions) * action_stride
  int n_steps;
};

}  // namespace

extern "C" int64_t cx_lunar_pipeline_workspace_bytes(int64_t B) {
  if (B <= 0) return 0;
  size_t need = 0;
  for (int S = 1; S <= MAX_SPLIT; ++S) {
    const size_t b = plan_ranges(B, S, nullptr, nullptr);
    if (b > need) need = b;
  }
  return (int64_t)need;
}

static int enqueue_steps(cudaStream_t st, int64_t B, int64_t ld, const StepSeq& q, float* reward, uint8_t* done,
                         float* obs, const CotixLunarParams* p, void* workspace, int64_t workspace_bytes,
                         int split = 0, const CxRangeHook* pre = nullptr, const CxRangeHook* post = nullptr) {
  if (B == 0 || q.n_steps <= 0) return 0;
  if (B >= (1ll << 27)) {
    cx_set_error("lunar pipeline: B must be < 2^27");
    return COTIX_EINVAL;
  }
  const int K = p->substeps < 1 ? 1 : p->substeps;
  const bool prof = g_profile && K == 1 && q.n_steps == 1;
  int S = prof ? 1 : (split > 0 ? split : choose_split(B));
  if (S > MAX_SPLIT) S = MAX_SPLIT;
  RangePlan plan;
  const size_t need = plan_ranges(B, S, workspace, &plan);
  if (!workspace || (size_t)workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255u)) {
    cx_set_error("lunar pipeline: workspace must be 256-byte aligned and >= %zu bytes", need);
    return COTIX_EINVAL;
  }
  PipeSide* side = pipe_side();
  if (!side) {
    cx_set_error("lunar pipeline: could not create the range streams");
    return COTIX_ECUDA;
  }
  LunarDerived D;
  cx_lunar_derive(p, &D);
  const int sms = cx_sm_count();
  auto qgrid = [&](int64_t max_items) {
    int64_t b = (max_items + TPB - 1) / TPB;
    const int64_t cap = (int64_t)sms * 8;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
  };
  auto qgrid_sparse = [&](int64_t max_items, int spread) {
    int64_t b = (max_items + TPB - 1) / TPB;
    const int64_t cap = (int64_t)sms * 8 * spread;
    return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
  };
  if (prof) {
    profile_collect();
    if (!g_ev_made) {
      for (int k = 0; k < N_STAGES; ++k) {
        cudaEventCreate(&g_evb[k]);
        cudaEventCreate(&g_eve[k]);
      }
      g_ev_made = true;
    }
  }
#define CX_MARK_B(k, stream) \
  if (prof) cudaEventRecord(g_evb[k], stream)
#define CX_MARK_E(k, stream) \
  if (prof) cudaEventRecord(g_eve[k], stream)
  // the chain of one range: one memset, then per step and sub-step 5 kernels, then the reset stage.
  // Stages 2..5 are launched with programmatic stream serialization: their blocks may be set up
  // while the previous stage drains, and wait in cudaGridDependencySynchronize() for its results.
  const bool trace = g_trace && !prof && K == 1 && q.n_steps <= TR_STEPS;
  if (trace) {
    if (!g_tr_base) {
      cudaEventCreate(&g_tr_base);
      for (auto& a : g_tr_ev)
        for (auto& b : a)
          for (auto& e : b) cudaEventCreate(&e);
    }
    g_tr_S = plan.S;
    g_tr_T = q.n_steps;
    cudaEventRecord(g_tr_base, st);
  }
#define CX_TRACE(stage, stream) \
  if (trace) cudaEventRecord(g_tr_ev[r][t][stage], stream)
  auto chain = [&](cudaStream_t cs, int r) -> int {
    const int64_t off = plan.off[r], n = plan.off[r + 1] - off;
    const PipeWs& ws = plan.ws[r];
    const int spread = sparse_spread(n);
    const unsigned env_blocks = (unsigned)((n + TPB - 1) / TPB);
    cudaLaunchAttribute attr[2];
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.blockDim = dim3(TPB);
    cfg.stream = cs;
    cfg.attrs = attr;
    auto set_attrs = [&](bool pdl, bool sparse) {
      int na = 0;
      if (pdl && g_cfg.pdl && !prof) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
      }
      if (g_cfg.prio && S > 1) {
        attr[na].id = cudaLaunchAttributePriority;
        attr[na].val.priority = sparse ? side->prio_hi : side->prio_lo;
        ++na;
      }
      cfg.numAttrs = na;
    };
    float* rw = reward ? reward + off : nullptr;
    uint8_t* dn = done ? done + off : nullptr;
    float* ob = obs ? obs + off : nullptr;
    if (pre && *pre) {
      const int hr = (*pre)(cs, r, off, n);
      if (hr != 0) return hr;
    }
    // counters: zeroed here once; from then on by the finish stage of every (sub-)step
    if (cudaMemsetAsync(ws.ctr, 0, CTR_WORDS * 4, cs) != cudaSuccess) return cx_check_launch("pipeline memset");
    int par = 0;       // parity of the (sub-)step: which of the two reset counters finish / reset use
    bool first = true;  // the first kernel of the chain follows a memset, not a kernel: no PDL
    for (int t = 0; t < q.n_steps; ++t) {
      const float* sin_ = q.sin0 + (int64_t)t * q.stride + off;
      float* sout = q.sout0 + (int64_t)t * q.stride + off;
      const float* act = q.action ? q.action + (int64_t)(t % q.n_actions) * q.action_stride + off : nullptr;
      CX_MARK_B(5, cs);
      // K sub-steps with the same action (cotix/_envs.py:72-103: WFE_scale forward steps per control):
      // the four step stages K times — sub-steps after the first run in place on the output state —
      // then one reset stage
      for (int k = 0; k < K; ++k) {
        CX_MARK_B(0, cs);
        set_attrs(!first, false);
        first = false;
        cfg.gridDim = dim3(env_blocks);
        cudaLaunchKernelEx(&cfg, pipe_integrate_kernel, n, k == 0 ? sin_ : (const float*)sout, sout, ld, act, *p, D, ws,
                           k, K);
        CX_MARK_E(0, cs);
        CX_TRACE(0, cs);
        CX_MARK_B(1, cs);
        set_attrs(true, true);
        cfg.gridDim = dim3(qgrid_sparse((int64_t)ws.cap * spread * TPB / NTPB, spread));
        cfg.blockDim = dim3(NTPB);
        switch (spread) {
          case 1: cudaLaunchKernelEx(&cfg, pipe_narrow_kernel<1>, *p, D, ws); break;
          case 2: cudaLaunchKernelEx(&cfg, pipe_narrow_kernel<2>, *p, D, ws); break;
          case 4: cudaLaunchKernelEx(&cfg, pipe_narrow_kernel<4>, *p, D, ws); break;
          case 8: cudaLaunchKernelEx(&cfg, pipe_narrow_kernel<8>, *p, D, ws); break;
          case 16: cudaLaunchKernelEx(&cfg, pipe_narrow_kernel<16>, *p, D, ws); break;
          default: cudaLaunchKernelEx(&cfg, pipe_narrow_kernel<32>, *p, D, ws); break;
        }
        cfg.blockDim = dim3(TPB);
        CX_MARK_E(1, cs);
        CX_TRACE(1, cs);
        CX_MARK_B(2, cs);
        cfg.gridDim = dim3(qgrid_sparse(n * spread, spread));
        switch (spread) {
          case 1: cudaLaunchKernelEx(&cfg, pipe_contact_kernel<false, 1>, sout, ld, *p, D, ws); break;
          case 2: cudaLaunchKernelEx(&cfg, pipe_contact_kernel<false, 2>, sout, ld, *p, D, ws); break;
          case 4: cudaLaunchKernelEx(&cfg, pipe_contact_kernel<false, 4>, sout, ld, *p, D, ws); break;
          case 8: cudaLaunchKernelEx(&cfg, pipe_contact_kernel<false, 8>, sout, ld, *p, D, ws); break;
          case 16: cudaLaunchKernelEx(&cfg, pipe_contact_kernel<false, 16>, sout, ld, *p, D, ws); break;
          default: cudaLaunchKernelEx(&cfg, pipe_contact_kernel<false, 32>, sout, ld, *p, D, ws); break;
        }
        // the envs that overflowed the item queue (normally none: the kernel reads the count and exits)
        cfg.gridDim = dim3((unsigned)(env_blocks < (unsigned)sms ? env_blocks : (unsigned)sms));
        cudaLaunchKernelEx(&cfg, pipe_contact_kernel<true, 1>, sout, ld, *p, D, ws);
        CX_MARK_E(2, cs);
        CX_TRACE(2, cs);
        CX_MARK_B(3, cs);
        set_attrs(true, false);
        cfg.gridDim = dim3(env_blocks);
        cudaLaunchKernelEx(&cfg, pipe_finish_kernel, n, sout, ld, act, rw, dn, ob, *p, D, ws, k, K, par);
        if (k + 1 < K) par ^= 1;  // the last sub-step's parity is the reset stage's
        CX_MARK_E(3, cs);
        CX_TRACE(3, cs);
      }
      CX_MARK_B(4, cs);
      if (p->auto_reset) {
        set_attrs(true, true);
        cfg.gridDim = dim3(qgrid(n));
        cudaLaunchKernelEx(&cfg, pipe_reset_kernel, sout, ld, ob, *p, ws, par);
      }
      par ^= 1;
      CX_MARK_E(4, cs);
      CX_TRACE(4, cs);
      CX_MARK_E(5, cs);
    }
    if (post && *post) {
      const int hr = (*post)(cs, r, off, n);
      if (hr != 0) return hr;
    }
    return cx_check_launch("lunar pipeline");
  };
#undef CX_MARK_B
#undef CX_MARK_E
#undef CX_TRACE
  int rc = 0;
  if (plan.S == 1) {
    rc = chain(st, 0);
  } else {
    std::lock_guard<std::mutex> lock(side->mu);
    if (cudaEventRecord(side->fork, st) != cudaSuccess) return cx_check_launch("pipeline fork");
    for (int r = 1; r < plan.S; ++r)
      if (cudaStreamWaitEvent(side->s[r], side->fork, 0) != cudaSuccess) return cx_check_launch("pipeline fork");
    if (pre || post) {
      // hooked (host-facing) step: ranges in natural order, so that a hook of range r may wait for an
      // event a hook of range r - 1 recorded
      for (int r = 0; r < plan.S; ++r) {
        const int rr = chain(r == 0 ? st : side->s[r], r);
        if (rr != 0 && rc == 0) rc = rr;
      }
    } else {
      // the side ranges first: the caller's stream may already be busy
      for (int r = 1; r < plan.S; ++r) {
        const int rr = chain(side->s[r], r);
        if (rr != 0 && rc == 0) rc = rr;
      }
      const int rr = chain(st, 0);
      if (rr != 0 && rc == 0) rc = rr;
    }
    // every exit path joins (an open fork would leave a stream capture unjoined)
    for (int r = 1; r < plan.S; ++r) {
      if (cudaEventRecord(side->join[r], side->s[r]) != cudaSuccess ||
          cudaStreamWaitEvent(st, side->join[r], 0) != cudaSuccess) {
        if (rc == 0) rc = cx_check_launch("pipeline join");
      }
    }
  }
  cx_count_launch_n((int64_t)(5 * K + (p->auto_reset ? 1 : 0)) * plan.S * q.n_steps);
  if (prof) g_ev_pending = true;
  return rc;
}

extern "C" int cx_launch_lunar_pipeline(cudaStream_t st, int64_t B, const float* sin_, float* sout,
                                        int64_t ld, const float* action, float* reward, uint8_t* done,
                                        float* obs, const CotixLunarParams* p, void* workspace,
                                        int64_t workspace_bytes) {
  StepSeq q{sin_, sout, 0, action, 0, 1, 1};
  return enqueue_steps(st, B, ld, q, reward, done, obs, p, workspace, workspace_bytes);
}

// n_steps steps in one enqueue: step t reads sin0 + t * stride and writes sout0 + t * stride (stride
// 0 = in place on one buffer; the tape of a rollout: sout0 = sin0 + stride), with action
// actions + (t % n_actions) * action_stride.  The ranges run their n_steps without joining.
extern "C" int cx_launch_lunar_pipeline_steps(cudaStream_t st, int64_t B, const float* sin0, float* sout0,
                                              int64_t stride, int64_t ld, const float* actions,
                                              int64_t action_stride, int n_actions, int n_steps, float* reward,
                                              uint8_t* done, float* obs, const CotixLunarParams* p,
                                              void* workspace, int64_t workspace_bytes) {
  StepSeq q{sin0, sout0, stride, actions, action_stride, n_actions < 1 ? 1 : n_actions, n_steps};
  return enqueue_steps(st, B, ld, q, reward, done, obs, p, workspace, workspace_bytes);
}

int cx_launch_lunar_pipeline_hooked(cudaStream_t st, int64_t B, float* state, int64_t ld, const float* action,
                                    float* reward, uint8_t* done, float* obs, const CotixLunarParams* p,
                                    void* workspace, int64_t workspace_bytes, int split, const CxRangeHook& pre,
                                    const CxRangeHook& post) {
  StepSeq q{state, state, 0, action, 0, 1, 1};
  return enqueue_steps(st, B, ld, q, reward, done, obs, p, workspace, workspace_bytes, split, &pre, &post);
}

// split: 0 automatic, 1..8 ranges; pdl / priorities: 0 off, 1 on; item_capacity: 0 default, > 0 forces
// the item queue of every range to that many slots (tests of the overflow branch); a negative value
// leaves a setting unchanged
extern "C" int cx_lunar_pipeline_config(int split, int pdl, int priorities, int64_t item_capacity) {
  if (split > MAX_SPLIT) return COTIX_EINVAL;
  if (split >= 0) g_cfg.split = split;
  if (pdl >= 0) g_cfg.pdl = pdl != 0;
  if (priorities >= 0) g_cfg.prio = priorities != 0;
  if (item_capacity >= 0) g_cfg.item_cap = (uint32_t)item_capacity;
  return COTIX_OK;
}
// queue counters of the last step, summed over the ranges the workspace was cut into for batch B:
// out[0..4] = items, hits, contact envs, resets, overflowed envs (DEVICE -> host copy, synchronises)
extern "C" int cx_lunar_pipeline_counters(cudaStream_t st, int64_t B, const void* workspace, int64_t* out) {
  for (int i = 0; i < 5; ++i) out[i] = 0;
  if (B <= 0 || !workspace) return COTIX_EINVAL;
  RangePlan plan;
  plan_ranges(B, (g_profile ? 1 : choose_split(B)), const_cast<void*>(workspace), &plan);
  for (int r = 0; r < plan.S; ++r) {
    uint32_t c[CTR_WORDS];
    if (cudaMemcpyAsync(c, plan.ws[r].ctr, sizeof(c), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess)
      return cx_check_launch("pipeline counters");
    for (int i = 0; i < 5; ++i) out[i] += c[CTR_SNAP + i];  // the finish stage's copy (the live counters are zeroed)
  }
  return COTIX_OK;
}

// timeline of the last traced enqueue: out[(r * T + t) * 5 + stage] = ms from the fork to the END of
// that kernel; returns S * T (0 when nothing was traced).  Synchronises.
extern "C" int cx_lunar_pipeline_trace(int enable) {
  g_trace = enable != 0;
  return COTIX_OK;
}
extern "C" int cx_lunar_pipeline_trace_read(float* out, int* S, int* T) {
  *S = g_tr_S;
  *T = g_tr_T;
  if (!g_tr_base || g_tr_S == 0) return 0;
  cudaDeviceSynchronize();
  for (int r = 0; r < g_tr_S; ++r)
    for (int t = 0; t < g_tr_T; ++t)
      for (int k = 0; k < TR_STAGES; ++k) {
        float ms = -1.0f;
        if (cudaEventElapsedTime(&ms, g_tr_base, g_tr_ev[r][t][k]) != cudaSuccess) {
          cudaGetLastError();
          ms = -1.0f;
        }
        out[(r * g_tr_T + t) * TR_STAGES + k] = ms;
      }
  return g_tr_S * g_tr_T;
}

// ---- per-stage timing (single-threaded use: bench.py) ----------------------------------------
extern "C" int cx_lunar_pipeline_profiling() { return g_profile ? 1 : 0; }
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
