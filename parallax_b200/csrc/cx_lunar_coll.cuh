// cx_lunar_coll.cuh — the coin flips of RandomizedCollider.resolve (cotix/_colliders.py:208-268) pooled
// over the 32 envs of a warp; shared by the contact stage of the wavefront pipeline (lunar_pipeline.cu)
// and the block-resident rollout kernel (lunar_rollout.cu).
#pragma once
#include "cx_lunar_step.cuh"

CX_DEV uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// The coin flips dominate this stage in threefry mode (a coin = 7 threefry-2x32 calls down the
// key chain split(gkey, N2)[ind2] -> split(., N1)[ind1] -> split(.)[0] -> bernoulli), and which
// cells are live differs from env to env.  Per-thread loops over 7 cells ran at 6 of 32 lanes
// (profiles/r1d_pipeline_v1.txt), so the warp pools the live cells of its 32 envs in a task
// ring in shared memory and flips ONE coin per lane per round: a tails re-queues the task at
// its next entry, a heads (or the end of the entries) retires it.  Every round is dense.
struct WarpColl {
  uint32_t gk[32][4];   // gkey[0], gkey[1] of lane's env
  uint32_t pk[32][2];   // philox key
  uint32_t st[32][2];   // steps, episode
  uint32_t tmask[256];  // task ring: live mask of the cell
  uint16_t tdesc[256];  //            env lane | cell << 5
  uint16_t tidx[256];   //            next entry, in the scan order of cell_winner
  int8_t win[32][8];
};

CX_DEV void coll_push(WarpColl& W, int& tail, bool pred, uint32_t desc, uint32_t mask, uint32_t t) {
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, pred);
  if (pred) {
    const int slot = (tail + __popc(m & lanemask_lt())) & 255;
    W.tdesc[slot] = (uint16_t)desc;
    W.tmask[slot] = mask;
    W.tidx[slot] = (uint16_t)t;
  }
  tail += __popc(m);
}

// One round: the first T = min(32, pending) tasks of the ring, m = 32 / T consecutive entries of
// each (m = 1 while the ring is full).  The extra entries are flipped speculatively: a task takes
// the FIRST heads among its m entries, exactly what flipping them one per round would find, so the
// thin rounds at the end of a warp's work (16, 8, 4, ... tasks left) collapse into one or two.
CX_DEV void coll_rounds(WarpColl& W, int& head, int& tail, const CotixLunarParams& P) {
  const int lane = threadIdx.x & 31;
  __syncwarp();
  while (head < tail) {
    const int T = min(tail - head, 32);
    const int m = 32 / T;
    const int task = lane / m, off = lane - task * m;
    const bool has = task < T;
    bool heads = false;
    uint32_t desc = 0u, mask = 0u, t0 = 0u;
    int el = 0, c = 0, q = 0, total = 0;
    if (has) {
      const int slot = (head + task) & 255;
      desc = W.tdesc[slot];
      mask = W.tmask[slot];
      t0 = W.tidx[slot];
      el = (int)(desc & 31u);
      c = (int)(desc >> 5);
      const CellDesc d = cell_desc(c);
      const int L = __popc(mask), per2 = d.reps * L;
      total = (d.b_hi - d.b_lo + 1) * per2;
      const int t = (int)t0 + off;
      if (t < total) {
        const int ind2 = d.b_hi - t / per2, rem = t % per2;
        const int r = d.reps - 1 - rem / L, j = rem % L;
        // the (L - j)-th set bit of the (at most 7-bit) live mask: live entries from the highest q down.
        // Clearing the lowest set bit L - j - 1 times is a handful of instructions; __fns was 6 % of the stage
        uint32_t mq = mask;
        for (int c2 = L - j - 1; c2 > 0; --c2) mq &= mq - 1u;
        q = __ffs(mq) - 1;
        const int ind1 = d.a0 + q + d.nq * r;
        CoinCtx cc;
        cc.mode = P.rng_mode;
        cc.p = P.coin_p;
        cc.gkey[0].a = W.gk[el][0]; cc.gkey[0].b = W.gk[el][1];
        cc.gkey[1].a = W.gk[el][2]; cc.gkey[1].b = W.gk[el][3];
        cc.pkey.a = W.pk[el][0]; cc.pkey.b = W.pk[el][1];
        cc.steps = W.st[el][0]; cc.episode = W.st[el][1];
        Key2 k2;
        k2.a = k2.b = 0u;
        if (cc.mode == COTIX_RNG_THREEFRY) k2 = tf_split(cc.gkey[d.g], (uint32_t)d.N2, (uint32_t)ind2);
        heads = coin_heads(cc, d.g, d.N1, k2, ind2, ind1);
      }
    }
    const uint32_t hb = __ballot_sync(0xFFFFFFFFu, heads);
    const uint32_t seg = (m >= 32 ? 0xFFFFFFFFu : ((1u << m) - 1u)) << (has ? task * m : 0);
    const uint32_t mine = has ? (hb & seg) : 0u;
    if (mine != 0u && lane == __ffs(mine) - 1) W.win[el][c] = (int8_t)q;
    const bool requeue = has && off == 0 && mine == 0u && (int)t0 + m < total;
    __syncwarp();
    head += T;
    coll_push(W, tail, requeue, desc, mask, t0 + (uint32_t)m);
    __syncwarp();
  }
}

