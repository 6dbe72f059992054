// lunar_rollout.cu — T steps of a SMALL batch in ONE launch, the envs resident in their block.
//
// Why: below a few ten thousand envs a step is pure latency.  Stepped one launch (fused kernel) or six
// launches (wavefront pipeline) at a time, 4096 envs take 50-68 us per step although the dependent
// instruction chain of one env-step is only a few microseconds: launch gaps, queue counters in HBM, and —
// in the fused kernel — a warp that serialises the dozen different narrowphase items of its block.
// Envs never interact, so a block can own a handful of them for the whole rollout:
//
//   block = 4 warps, EPB <= 32 envs dealt to the warps (EPB / 4 each, lane = env), held in registers for
//           all T steps
//   phase A  lane = env            action, Euler + gravity + engines, exact broadphase; envs with
//                                  candidates publish their ordered world polygons and (env, pair)
//                                  items to shared memory
//   phase B  item i -> warp i % 4, lane i / 4: the items go across the warps first, so that few lanes of a
//                                  warp wait for another item's GJK / EPA branch
//   phase C  lane = env            coin flips pooled over the warp (cx_lunar_coll.cuh), impulses, key
//                                  split, joints, judge, auto-reset; the tape slot / outputs of the step
//
// Nothing but the tape (or the final state), the actions and the last step's reward / done / obs touches
// HBM; there is no counter, no workspace and no host involvement between the steps.  The per-env
// functions are those of the other two launch strategies (cx_lunar_step.cuh), called with the same
// arguments in the same order: the results are bit-identical to stepping T times.
#include <cstdio>
#include <cstdlib>

#include "cx_launch.cuh"
#include "cx_lunar_coll.cuh"
#include "cx_lunar_step.cuh"

namespace {

// threads per block: 4 warps of up to 8 envs.  Measured, 100 contact-rich steps, 8 / 4 / 2 warps per block (32 envs):
// 4096 envs 3.37 / 3.18 / 3.26 ms, 1024 envs 2.58 / 2.54 / 2.69 ms (16 warps: 3.43)
#ifndef CX_ROLLOUT_TPB
#define CX_ROLLOUT_TPB 128
#endif
constexpr int RTPB = CX_ROLLOUT_TPB;   // threads per block
constexpr int RWARPS = RTPB / 32;
constexpr int REPB = 32;    // most envs per block
constexpr int RQCAP = 256;  // items per block and step; beyond, the owning lane evaluates its own pairs
constexpr uint32_t ITEM_SKIP = 0xFFFFFFFFu;

struct RolloutShared {
  float poly[28][REPB];  // ordered world vertices: lander 0..11, right leg 12..19, left leg 20..27
  float terr[12][REPB];  // heights 0..7, knots 8..11
  uint32_t item[RQCAP];  // (env lane << 8) | pair id
  float res[4][RQCAP];   // pen.x, pen.y, cp.x, cp.y   (cp NaN = no contact)
  int qcount;
  WarpColl W[RWARPS];
};

struct BlockPolys {
  const RolloutShared& S;
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

static __device__ __noinline__ Contact rollout_item(const RolloutShared& S, int el, int pid, const CotixLunarParams& P) {
  Poly A, B;
  BlockPolys src{S, el};
  make_item_polys(src, pid, P, A, B);
  return poly_pair(A, B, mk2(P.init_dir[0], P.init_dir[1]), P.gjk_max_iters);
}

// contact source of lane's env: the block queue's results, or its own pairs on queue overflow
struct BlockContacts {
  const RolloutShared& S;
  const CotixLunarParams& P;
  uint32_t cand;
  int base, el;
  CX_DEV Contact get(int bit) const {
    if (base >= 0) {
      const int i = base + __popc(cand & ((1u << bit) - 1u));
      Contact c;
      c.pen = mk2(S.res[0][i], S.res[1][i]);
      c.cp = mk2(S.res[2][i], S.res[3][i]);
      c.status = 0u;
      return c;
    }
    return rollout_item(S, el, bit, P);
  }
  CX_DEV bool live(int bit) const {
    if (!((cand >> bit) & 1u)) return false;
    if (base >= 0) {
      const int i = base + __popc(cand & ((1u << bit) - 1u));
      const float x = S.res[2][i], y = S.res[3][i];
      return !((x != x) || (y != y));
    }
    return !contact_isnan(rollout_item(S, el, bit, P));
  }
  CX_DEV uint32_t live_bits() const {
    uint32_t live = 0u, todo = cand;
    while (todo) {
      const int bit = __ffs(todo) - 1;
      todo &= todo - 1u;
      if (this->live(bit)) live |= 1u << bit;
    }
    return live;
  }
};

#ifndef CX_ROLLOUT_MINB
#define CX_ROLLOUT_MINB (512 / CX_ROLLOUT_TPB)
#endif
__global__ void __launch_bounds__(RTPB, CX_ROLLOUT_MINB)
lunar_rollout_kernel(int64_t B, const float* __restrict__ sin0, float* __restrict__ sout0, int64_t stride, int64_t ld,
                     const float* __restrict__ actions, int64_t action_stride, int n_actions, int n_steps,
                     float* __restrict__ reward, uint8_t* __restrict__ done, float* __restrict__ obs,
                     const __grid_constant__ CotixLunarParams P, const __grid_constant__ LunarDerived D, int epb, int qcap) {
  extern __shared__ __align__(16) unsigned char rollout_smem[];  // dynamic: 16 warps need more than the 48 KB static limit
  RolloutShared& S = *reinterpret_cast<RolloutShared*>(rollout_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // the block's epb envs are dealt to its warps, epb / RWARPS each (lanes 0 .. epw-1): a warp is as fast with one
  // env as with thirty-two, and the fewer envs share a warp the fewer divergent branches it serialises
  const int epw = epb / RWARPS, el = warp * epw + lane;  // el: env slot in the block's shared arrays
  const int64_t e = (int64_t)blockIdx.x * epb + el;
  const bool mine = lane < epw && e < B;  // this thread owns env e
  EnvRegs E;
  if (mine) load_env(E, sin0, ld, e);
  bool fresh_terrain = (stride != 0) || (sin0 != sout0);  // a new slot needs the terrain rows too
#ifdef CX_ROLLOUT_CLOCKS
  long long ck[4] = {0, 0, 0, 0}, c0 = 0;
  int items_seen = 0;
#define CX_CK(i) do { const long long c1 = clock64(); ck[i] += c1 - c0; c0 = c1; } while (0)
#else
#define CX_CK(i)
#endif
  for (int t = 0; t < n_steps; ++t) {
    if (tid == 0) S.qcount = 0;
    __syncthreads();
#ifdef CX_ROLLOUT_CLOCKS
    c0 = clock64();
#endif
    // ---- phase A ------------------------------------------------------------------------------
    StepCtx cx;
    cx.cand = 0u;
    cx.base = -1;
    if (mine) {
      float am = 0.0f, as = 0.0f;
      if (actions) {
        const float* a = actions + (int64_t)(t % n_actions) * action_stride;
        am = a[e];
        as = a[ld + e];
      }
      Poly wL, wr, wl;
      integrate_and_cull<true>(E, am, as, P, D, cx, wL, wr, wl);
      if (cx.cand != 0u) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          S.poly[2 * k][el] = wL.x[k];
          S.poly[2 * k + 1][el] = wL.y[k];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          S.poly[12 + 2 * k][el] = wr.x[k];
          S.poly[13 + 2 * k][el] = wr.y[k];
          S.poly[20 + 2 * k][el] = wl.x[k];
          S.poly[21 + 2 * k][el] = wl.y[k];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) S.terr[k][el] = E.h[k];
#pragma unroll
        for (int k = 0; k < 4; ++k) S.terr[8 + k][el] = E.kn[k];
        const int cnt = __popc(cx.cand);
        const int base = atomicAdd(&S.qcount, cnt);
        if (base + cnt <= qcap) {
          cx.base = base;
          uint32_t m = cx.cand;
          int slot = base;
          while (m) {
            const int bit = __ffs(m) - 1;
            m &= m - 1u;
            S.item[slot++] = ((uint32_t)el << 8) | (uint32_t)bit;
          }
        } else {
          for (int slot = base; slot < qcap && slot < base + cnt; ++slot) S.item[slot] = ITEM_SKIP;
        }
      }
    }
    __syncthreads();
    CX_CK(0);
#ifdef CX_ROLLOUT_CLOCKS
    items_seen += S.qcount;
#endif
    // ---- phase B: items across the warps first ---------------------------------------------------
    {
      const int n = S.qcount < qcap ? S.qcount : qcap;
      for (int i = lane * RWARPS + warp; i < n; i += RTPB) {
        const uint32_t it = S.item[i];
        if (it == ITEM_SKIP) continue;
        const Contact c = rollout_item(S, (int)(it >> 8), (int)(it & 0xFFu), P);
        S.res[0][i] = c.pen.x;
        S.res[1][i] = c.pen.y;
        S.res[2][i] = c.cp.x;
        S.res[3][i] = c.cp.y;
      }
    }
    __syncthreads();
    CX_CK(1);
    // ---- phase C (every lane of a warp takes part in its pooled coin flips) -------------------------
    {
      WarpColl& W = S.W[warp];
      uint32_t m[7] = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
      bool crash = false, rc = false, lc = false, any = false;
      CoinCtx cc;
      BlockContacts bc{S, P, cx.cand, cx.base, el};
      if (mine && cx.cand != 0u) any = cell_masks(bc.live_bits(), true, m, crash, rc, lc);
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
      coll_push(W, tail, m[6] != 0u && (W.win[lane][1] >= 0 || W.win[lane][3] >= 0), (uint32_t)lane | (6u << 5),
                m[6], 0u);
      coll_rounds(W, head, tail, P);
      if (mine) {
        if (any) {
          int win[7];
#pragma unroll
          for (int k = 0; k < 7; ++k) win[k] = W.win[lane][k];
          collide_apply<BlockContacts, false>(E, P, D, bc, cc, win);
        }
        SubInfo si;
        finish_step<true>(E, P, D, cx.cs, cx.sn, cx.thr_main, cx.thr_side, crash, rc, lc, si);
        const bool last = t + 1 == n_steps;
        bool orc = si.rc, olc = si.lc;
        if (si.done != COTIX_DONE_NO && P.auto_reset) {
          lunar_auto_reset(E, P);
          orc = olc = false;
          fresh_terrain = true;
        }
        if (stride != 0) store_env(E, sout0 + (int64_t)t * stride, ld, e, true);
        else if (last) store_env(E, sout0, ld, e, fresh_terrain);
        if (last) {
          if (reward) reward[e] = si.reward;
          if (done) done[e] = (uint8_t)si.done;
          if (obs) store_obs(E, obs, ld, e, orc, olc);
        }
      }
    }
    CX_CK(2);
  }
#ifdef CX_ROLLOUT_CLOCKS
  if (tid == 0 && (blockIdx.x % 37) == 0)
    printf("block %d: A %lld B %lld C %lld cycles per step, items per step %.2f\n", blockIdx.x, ck[0] / n_steps,
           ck[1] / n_steps, ck[2] / n_steps, (double)items_seen / n_steps);
#endif
}

}  // namespace

// T steps of K = 1 sub-step each in one launch: step t reads state t (slot t of sin0 when stride != 0,
// the running state otherwise), uses actions[t % n_actions] and writes slot t of sout0 (stride != 0) or,
// after the last step, sout0.  reward / done / obs receive the last step's values — what T calls of
// cx_launch_lunar_step leave behind.
extern "C" int cx_launch_lunar_rollout_kernel(cudaStream_t st, int64_t B, const float* sin0, float* sout0, int64_t stride,
                                              int64_t ld, const float* actions, int64_t action_stride, int n_actions,
                                              int n_steps, float* reward, uint8_t* done, float* obs,
                                              const CotixLunarParams* p) {
  if (B == 0 || n_steps <= 0) return 0;
  LunarDerived D;
  cx_lunar_derive(p, &D);
  // envs per block: 32 once that fills the SMs, fewer (down to 8) for tiny batches so that every SM has work
  const int sms = cx_sm_count();
  int epb = REPB;
  while (epb > RWARPS && (B + epb - 1) / epb < (sms * 2) / 3) epb >>= 1;  // 4096 envs: 128 blocks of 32 beat 256 of 16
  if (const char* env = getenv("COTIX_ROLLOUT_EPB")) {  // measurement aid
    const int v = atoi(env);
    if (v >= RWARPS && v <= REPB && v % RWARPS == 0) epb = v;
  }
  int qcap = RQCAP;
  if (const char* env = getenv("COTIX_ROLLOUT_QCAP")) {  // tests force the queue-overflow branch with a small value
    const int v = atoi(env);
    if (v >= 0 && v < RQCAP) qcap = v;
  }
  const unsigned blocks = (unsigned)((B + epb - 1) / epb);
  static bool attr_set[64];
  int dev = 0;
  if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64 && !attr_set[dev]) {
    if (cudaFuncSetAttribute(lunar_rollout_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sizeof(RolloutShared)) != cudaSuccess)
      return cx_check_launch("lunar_rollout_kernel shared-memory attribute");
    attr_set[dev] = true;
  }
  lunar_rollout_kernel<<<blocks, RTPB, sizeof(RolloutShared), st>>>(B, sin0, sout0, stride, ld, actions, action_stride,
                                                n_actions > 0 ? n_actions : 1, n_steps, reward, done, obs, *p, D, epb, qcap);
  cx_count_launch();
  return cx_check_launch("lunar_rollout_kernel");
}
