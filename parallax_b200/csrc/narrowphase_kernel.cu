// narrowphase_kernel.cu — K3: GJK/EPA narrowphase sweep, one shape pair per thread.
//
// Replaces the per-pair contact functions of cotix/_contacts.py:30-315 for a flat batch.
// HBM traffic per pair: 2 x 64 B records in, 1 + 8 + 8 (+1) B out (148 B).
// Records are read as four 16-byte vector loads per shape (a warp covers 2 KB of
// contiguous records per array); outputs are written as float2 / bytes, coalesced.
#include <mutex>

#include "../../include/cotix_b200.h"
#include "cx_launch.cuh"
#include "cx_narrow.cuh"

namespace {

struct Rec {
  float4 q[4];
};
CX_DEV Rec load_rec(const float* base, int64_t k) {
  const float4* p = reinterpret_cast<const float4*>(base) + k * 4;
  Rec r;
  r.q[0] = __ldg(p + 0);
  r.q[1] = __ldg(p + 1);
  r.q[2] = __ldg(p + 2);
  r.q[3] = __ldg(p + 3);
  return r;
}
CX_DEV int rec_type(const Rec& r) { return (int)r.q[0].x; }
CX_DEV Poly rec_poly(const Rec& r) {
  Poly p;
  int n = (int)r.q[0].y;
  p.n = n < 1 ? 1 : (n > CX_MAXV ? CX_MAXV : n);
  p.x[0] = r.q[1].x; p.y[0] = r.q[1].y; p.x[1] = r.q[1].z; p.y[1] = r.q[1].w;
  p.x[2] = r.q[2].x; p.y[2] = r.q[2].y; p.x[3] = r.q[2].z; p.y[3] = r.q[2].w;
  p.x[4] = r.q[3].x; p.y[4] = r.q[3].y; p.x[5] = r.q[3].z; p.y[5] = r.q[3].w;
  return p;
}
CX_DEV Box rec_box(const Rec& r) {
  Box b;
  b.lo = mk2(r.q[1].x, r.q[1].y);
  b.hi = mk2(r.q[1].z, r.q[1].w);
  return b;
}
CX_DEV Circ rec_circ(const Rec& r) {
  Circ c;
  c.c = mk2(r.q[1].x, r.q[1].y);
  c.r = r.q[0].z;
  return c;
}

__global__ void __launch_bounds__(128)
narrowphase_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                   uint8_t* __restrict__ exists, float* __restrict__ pen,
                   float* __restrict__ cp, uint8_t* __restrict__ status,
                   CotixNarrowParams prm) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const V2 start = mk2(prm.init_dir[0], prm.init_dir[1]);
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
    Rec ra = load_rec(a, k), rb = load_rec(b, k);
    const int ta = rec_type(ra), tb = rec_type(rb);
    Contact c = contact_none();
    uint32_t bad = 0u;
    if (ta == COTIX_SHAPE_POLYGON && tb == COTIX_SHAPE_POLYGON) {
      c = convex_vs_polygon(rec_poly(ra), rec_poly(rb), start, prm.gjk_max_iters);
    } else if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_POLYGON) {
      c = circle_vs_polygon<131>(rec_circ(ra), rec_poly(rb), start, prm.gjk_max_iters,
                                 prm.epa_iters_circle_poly);
    } else if (ta == COTIX_SHAPE_AABB && tb == COTIX_SHAPE_POLYGON) {
      c = convex_vs_polygon(rec_box(ra), rec_poly(rb), start, prm.gjk_max_iters);
    } else if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_CIRCLE) {
      c = circle_vs_circle(rec_circ(ra), rec_circ(rb));
    } else if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_AABB) {
      c = circle_vs_aabb(rec_circ(ra), rec_box(rb));
    } else if (ta == COTIX_SHAPE_AABB && tb == COTIX_SHAPE_AABB) {
      c = aabb_vs_aabb(rec_box(ra), rec_box(rb));
    } else {
      bad = 0x80u;  // not a key of _contact_funcs in this order
    }
    if (exists) exists[k] = contact_isnan(c) ? 0 : 1;
    if (pen) reinterpret_cast<float2*>(pen)[k] = make_float2(c.pen.x, c.pen.y);
    if (cp) reinterpret_cast<float2*>(cp)[k] = make_float2(c.cp.x, c.cp.y);
    if (status) status[k] = (uint8_t)(c.status | bad);
  }
}

// order_clockwise (cotix/_geometry_utils.py:60-67) on polygon records, in place.
__global__ void __launch_bounds__(128) order_clockwise_kernel(int64_t n, float* recs) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  float* r = recs + k * COTIX_SHAPE_REC_FLOATS;
  if ((int)r[0] != COTIX_SHAPE_POLYGON) return;
  int nv = (int)r[1];
  if (nv > CX_MAXV) nv = CX_MAXV;
  if (nv < 1) return;
  float vx[CX_MAXV], vy[CX_MAXV], ang[CX_MAXV];
  float sx = 0.0f, sy = 0.0f;
  for (int q = 0; q < nv; ++q) {
    vx[q] = r[4 + 2 * q];
    vy[q] = r[5 + 2 * q];
    sx = sx + vx[q];
    sy = sy + vy[q];
  }
  float mx = sx / (float)nv, my = sy / (float)nv;
  for (int q = 0; q < nv; ++q) ang[q] = cx_atan2f(vy[q] - my, vx[q] - mx);
  // stable rank: NaN last
  for (int q = 0; q < nv; ++q) {
    int rank = 0;
    float aq = ang[q];
    bool qn = aq != aq;
    for (int t = 0; t < nv; ++t) {
      float at = ang[t];
      bool tn = at != at;
      bool before;
      if (tn || qn) before = (!tn && qn) || (tn == qn && t < q);
      else before = (at < aq) || (at == aq && t < q);
      if (t != q && before) ++rank;
    }
    r[4 + 2 * rank] = vx[q];
    r[5 + 2 * rank] = vy[q];
  }
}

__global__ void __launch_bounds__(128)
support_kernel(int64_t n, const float* recs, const float* dirs, float* out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  Rec r = load_rec(recs, k);
  V2 d = mk2(dirs[2 * k], dirs[2 * k + 1]);
  V2 s;
  int t = rec_type(r);
  if (t == COTIX_SHAPE_CIRCLE) s = support(rec_circ(r), d);
  else if (t == COTIX_SHAPE_AABB) s = support(rec_box(r), d);
  else s = support(rec_poly(r), d);
  out[2 * k] = s.x;
  out[2 * k + 1] = s.y;
}

// ------------------------------------------------------------------------------------------------
// Wavefront version of the sweep (cotix_narrowphase_ws): the mixed-type kernel above runs a
// 128-iteration-capable EPA next to 5-instruction circle-circle tests in the same warp.  Here
//   sweep_classify   thread = pair      analytic pairs (circle/AABB) are finished on the spot;
//                                        polygon pairs are appended to the queue of their class
//                                        (0 polygon-polygon, 1 circle-polygon, 2 AABB-polygon)
//   sweep_gjk<C>     thread = queued pair   GJK; misses are finished, hits appended to the hit
//                                        queue of the class together with their simplex
//   sweep_epa<C>     thread = hit       EPA + contact point
// so that every warp runs one code path.  Counters live in device memory; the queue kernels use a
// fixed grid (a multiple of the SM count) and loop over the count they read.
enum { SW_Q0 = 0, SW_H0 = 3, SW_WORDS = 16 };

struct SweepWs {
  uint32_t* ctr;     // [SW_WORDS]: queue counts [0..2], hit counts [3..5]
  uint32_t* q[3];    // [n] pair indices per class
  uint32_t* h[3];    // [n] pair indices of the hits per class
  float* simplex[3];  // [6][n] per class, indexed by hit slot
  uint8_t* gstat[3];  // [n] per class: GJK status bits of the hits, by hit slot
};

CX_DEV uint32_t sw_lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
CX_DEV uint32_t sw_push(uint32_t* counter, bool pred) {  // must be reached by all 32 lanes
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, pred);
  if (m == 0u) return 0u;
  const int leader = __ffs(m) - 1;
  uint32_t b = 0u;
  if ((int)(threadIdx.x & 31u) == leader) b = atomicAdd(counter, (uint32_t)__popc(m));
  b = __shfl_sync(0xFFFFFFFFu, b, leader);
  return b + (uint32_t)__popc(m & sw_lanemask_lt());
}
// block-wide variant: ONE atomic per block and call (the queue counters are hot addresses).  Must be
// reached by all threads of a 128-thread block; scratch = 5 words of shared memory.
CX_DEV uint32_t sw_push_block(uint32_t* counter, bool pred, uint32_t* scratch) {
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, pred);
  const uint32_t w = threadIdx.x >> 5;
  if ((threadIdx.x & 31u) == 0u) scratch[w] = (uint32_t)__popc(m);
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t tot = scratch[0] + scratch[1] + scratch[2] + scratch[3];
    scratch[4] = tot ? atomicAdd(counter, tot) : 0u;
  }
  __syncthreads();
  uint32_t base = scratch[4];
  for (uint32_t j = 0; j < w; ++j) base += scratch[j];
  const uint32_t r = base + (uint32_t)__popc(m & sw_lanemask_lt());
  __syncthreads();
  return r;
}
CX_DEV void sw_store(int64_t k, const Contact& c, uint32_t extra, uint8_t* exists, float* pen, float* cp,
                     uint8_t* status) {
  if (exists) exists[k] = contact_isnan(c) ? 0 : 1;
  if (pen) reinterpret_cast<float2*>(pen)[k] = make_float2(c.pen.x, c.pen.y);
  if (cp) reinterpret_cast<float2*>(cp)[k] = make_float2(c.cp.x, c.cp.y);
  if (status) status[k] = (uint8_t)(c.status | extra);
}

// Tile of SW_TILE pairs per block iteration.  The polygon classes go to the global queues; the three
// closed-form type pairs (25 % of config 2, randomly interleaved) are first collected per formula in
// shared memory and then evaluated with the items of one formula on consecutive lanes — evaluating
// them in place costs every warp all three formulas at ~8 % lane occupancy each.
constexpr int SW_TILE = 512;
__global__ void __launch_bounds__(128)
sweep_classify_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                      uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                      uint8_t* __restrict__ status, const SweepWs ws) {
  __shared__ uint32_t s_item[6][SW_TILE];  // 0..2 closed-form formulas, 3..5 polygon classes
  __shared__ int s_cnt[6];
  __shared__ uint32_t s_base[3];
  for (int64_t t0 = (int64_t)blockIdx.x * SW_TILE; t0 < n; t0 += (int64_t)gridDim.x * SW_TILE) {
    if (threadIdx.x < 6) s_cnt[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < SW_TILE / 128; ++j) {
      const int64_t k = t0 + j * 128 + threadIdx.x;
      if (k < n) {
        // type word only: the polygon classes read their records in the GJK pass
        const int ta = (int)__ldg(a + k * COTIX_SHAPE_REC_FLOATS), tb = (int)__ldg(b + k * COTIX_SHAPE_REC_FLOATS);
        int list = -1;
        if (tb == COTIX_SHAPE_POLYGON && ta >= COTIX_SHAPE_CIRCLE && ta <= COTIX_SHAPE_POLYGON)
          list = 3 + (ta == COTIX_SHAPE_POLYGON ? 0 : (ta == COTIX_SHAPE_CIRCLE ? 1 : 2));
        else if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_CIRCLE) list = 0;
        else if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_AABB) list = 1;
        else if (ta == COTIX_SHAPE_AABB && tb == COTIX_SHAPE_AABB) list = 2;
        if (list >= 0) s_item[list][atomicAdd(&s_cnt[list], 1)] = (uint32_t)(k - t0);
        else sw_store(k, contact_none(), 0x80u, exists, pen, cp, status);  // not a key of _contact_funcs in this order
      }
    }
    __syncthreads();
    // one global atomic per class and tile (the counters are three hot addresses: one atomic per warp
    // and class made this kernel atomic-bound), then the tile's members are copied out coalesced
    if (threadIdx.x < 3 && s_cnt[3 + threadIdx.x] > 0)
      s_base[threadIdx.x] = atomicAdd(&ws.ctr[SW_Q0 + threadIdx.x], (uint32_t)s_cnt[3 + threadIdx.x]);
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 3; ++c)
      for (int i = threadIdx.x; i < s_cnt[3 + c]; i += 128) ws.q[c][s_base[c] + i] = (uint32_t)(t0 + s_item[3 + c][i]);
    const int c0 = s_cnt[0], c1 = s_cnt[1], c2 = s_cnt[2];
    for (int i = threadIdx.x; i < c0 + c1 + c2; i += 128) {
      const int cf = i < c0 ? 0 : (i < c0 + c1 ? 1 : 2);
      const int64_t k = t0 + s_item[cf][i - (cf == 0 ? 0 : (cf == 1 ? c0 : c0 + c1))];
      Contact c;
      if (cf == 0) c = circle_vs_circle(rec_circ(load_rec(a, k)), rec_circ(load_rec(b, k)));
      else if (cf == 1) c = circle_vs_aabb(rec_circ(load_rec(a, k)), rec_box(load_rec(b, k)));
      else c = aabb_vs_aabb(rec_box(load_rec(a, k)), rec_box(load_rec(b, k)));
      sw_store(k, c, 0u, exists, pen, cp, status);
    }
    __syncthreads();
  }
}

template <int CLS>
__global__ void __launch_bounds__(128)
sweep_gjk_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                 uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                 uint8_t* __restrict__ status, CotixNarrowParams prm, const SweepWs ws) {
  __shared__ uint32_t s_push[5];
  const uint32_t cnt = ws.ctr[SW_Q0 + CLS];
  const uint32_t stride = gridDim.x * blockDim.x;
  const V2 start = mk2(prm.init_dir[0], prm.init_dir[1]);
  for (uint32_t i0 = blockIdx.x * blockDim.x; i0 < cnt; i0 += stride) {  // uniform per block
    const uint32_t i = i0 + threadIdx.x;
    bool hit = false, gcap = false;
    uint32_t k = 0u;
    V2 s0 = mk2(0.0f, 0.0f), s1 = s0, s2 = s0;
    if (i < cnt) {
      k = ws.q[CLS][i];
      const Rec ra = load_rec(a, k), rb = load_rec(b, k);
      const Poly B = rec_poly(rb);
      if (CLS == 0) hit = gjk(rec_poly(ra), B, start, prm.gjk_max_iters, s0, s1, s2, gcap);
      else if (CLS == 1) hit = gjk(rec_circ(ra), B, start, prm.gjk_max_iters, s0, s1, s2, gcap);
      else hit = gjk(rec_box(ra), B, start, prm.gjk_max_iters, s0, s1, s2, gcap);
      if (!hit) {
        Contact c = contact_none();
        if (gcap) c.status |= CX_ST_GJK_CAP;
        sw_store(k, c, 0u, exists, pen, cp, status);
      }
    }
    // the simplex and GJK status travel with the hit's queue slot (coalesced), not with the pair
    const uint32_t slot = sw_push_block(&ws.ctr[SW_H0 + CLS], hit, s_push);
    if (hit) {
      ws.h[CLS][slot] = k;
      float* sx = ws.simplex[CLS];
      sx[0 * n + slot] = s0.x; sx[1 * n + slot] = s0.y;
      sx[2 * n + slot] = s1.x; sx[3 * n + slot] = s1.y;
      sx[4 * n + slot] = s2.x; sx[5 * n + slot] = s2.y;
      ws.gstat[CLS][slot] = gcap ? (uint8_t)CX_ST_GJK_CAP : (uint8_t)0;
    }
  }
}

template <int CLS>
__global__ void __launch_bounds__(128)
sweep_epa_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                 uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                 uint8_t* __restrict__ status, CotixNarrowParams prm, const SweepWs ws) {
  const uint32_t cnt = ws.ctr[SW_H0 + CLS];
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += stride) {
    const uint32_t k = ws.h[CLS][i];
    const Rec ra = load_rec(a, k), rb = load_rec(b, k);
    const Poly B = rec_poly(rb);
    const float* sx = ws.simplex[CLS];
    const V2 s0 = mk2(sx[0 * n + i], sx[1 * n + i]);
    const V2 s1 = mk2(sx[2 * n + i], sx[3 * n + i]);
    const V2 s2 = mk2(sx[4 * n + i], sx[5 * n + i]);
    Contact c = contact_none();
    c.status = ws.gstat[CLS][i];
    if (CLS == 0) convex_poly_finish(rec_poly(ra), B, s0, s1, s2, c);
    else if (CLS == 1) circle_poly_finish<131>(rec_circ(ra), B, s0, s1, s2, prm.epa_iters_circle_poly, c);
    else convex_poly_finish(rec_box(ra), B, s0, s1, s2, c);
    sw_store(k, c, 0u, exists, pen, cp, status);
  }
}

// Circle-polygon EPA in two passes.  The expansion count of a circle-polygon hit is 5 ... 100
// (the Minkowski difference has arcs), 13.5 on average, and a warp runs as long as its slowest lane:
// the one-pass kernel keeps 10.6 of 32 lanes busy.  The SHORT pass gives every hit CX_EPA_SHORT
// expansions (edge list of CX_EPA_SHORT + 3 entries); a hit that converges inside the budget is
// finished — the loop stops on its own, so the result is the one the full budget gives — and the
// others are appended to the LONG queue (the class-1 GJK queue, free by then) and run again with the
// full budget.  Identical outputs.
#ifndef CX_EPA_SHORT
#define CX_EPA_SHORT 24  // measured on BASELINE config 2: 10 / 16 / 24 / 32 / 40 / 56 -> 7.42 / 6.89 / 6.53 / 6.98 / 7.24 / 7.43 ms
#endif
enum { SW_LONG = 6 };
__global__ void __launch_bounds__(128)
sweep_epa_cp_short_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                          uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                          uint8_t* __restrict__ status, CotixNarrowParams prm, const SweepWs ws) {
  const uint32_t cnt = ws.ctr[SW_H0 + 1];
  const uint32_t stride = gridDim.x * blockDim.x;
  const int budget = prm.epa_iters_circle_poly < CX_EPA_SHORT ? prm.epa_iters_circle_poly : CX_EPA_SHORT;
  const bool whole = prm.epa_iters_circle_poly <= CX_EPA_SHORT;  // the short budget is the full one
  const float* sx = ws.simplex[1];
  // the EPA edge ring of every thread in shared memory, [entry][thread] (cx_narrow.cuh, EpaRing): 44.9 KB
  // per block, 5 blocks per SM, no local memory
  constexpr int RE = CX_EPA_SHORT + 3;
  __shared__ float s_ax[RE][128], s_ay[RE][128], s_d[RE][128];
  __shared__ uint8_t s_nx[RE][128];
  const EpaRing<128> ring{&s_ax[0][threadIdx.x], &s_ay[0][threadIdx.x], &s_d[0][threadIdx.x], &s_nx[0][threadIdx.x]};
  // (appends per warp, not per block: only ~10 % of the hits are deferred, and a block barrier per
  // iteration would make every warp wait for the slowest of the block's four)
  for (uint32_t i0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); i0 < cnt; i0 += stride) {
    const uint32_t i = i0 + (threadIdx.x & 31u);
    bool defer = false;
    if (i < cnt) {
      const uint32_t k = ws.h[1][i];
      const Circ circle = rec_circ(load_rec(a, k));
      const Poly B = rec_poly(load_rec(b, k));
      const V2 s0 = mk2(sx[0 * n + i], sx[1 * n + i]);
      const V2 s1 = mk2(sx[2 * n + i], sx[3 * n + i]);
      const V2 s2 = mk2(sx[4 * n + i], sx[5 * n + i]);
      bool ecap;
      V2 ea, eb;
      epa_edge_ring<128>(circle, B, s0, s1, s2, budget, ecap, ea, eb, ring);
      defer = ecap && !whole;
      if (!defer) {
        Contact c = contact_none();
        c.status = ws.gstat[1][i];
        if (ecap) c.status |= CX_ST_EPA_CAP;
        c.pen = closest_to_origin(ea, eb);
        c.cp = circle_poly_contact_point(circle, B);
        sw_store(k, c, 0u, exists, pen, cp, status);
      }
    }
    const uint32_t slot = sw_push(&ws.ctr[SW_LONG], defer);
    if (defer) ws.q[1][slot] = i;  // the hit's slot: pair index, simplex and status hang off it
  }
}
__global__ void __launch_bounds__(128)
sweep_epa_cp_long_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                         uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                         uint8_t* __restrict__ status, CotixNarrowParams prm, const SweepWs ws) {
  const uint32_t cnt = ws.ctr[SW_LONG];
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += stride) {
    const uint32_t h = ws.q[1][i];
    const uint32_t k = ws.h[1][h];
    const Rec ra = load_rec(a, k), rb = load_rec(b, k);
    const float* sx = ws.simplex[1];
    const V2 s0 = mk2(sx[0 * n + h], sx[1 * n + h]);
    const V2 s1 = mk2(sx[2 * n + h], sx[3 * n + h]);
    const V2 s2 = mk2(sx[4 * n + h], sx[5 * n + h]);
    Contact c = contact_none();
    c.status = ws.gstat[1][h];
    circle_poly_finish<131>(rec_circ(ra), rec_poly(rb), s0, s1, s2, prm.epa_iters_circle_poly, c);
    sw_store(k, c, 0u, exists, pen, cp, status);
  }
}

size_t sw_align(size_t x) { return (x + 255) / 256 * 256; }
size_t sweep_carve(int64_t n, void* base, SweepWs* ws) {
  size_t off = 0;
  char* p = static_cast<char*>(base);
  auto take = [&](size_t bytes) {
    char* q = p ? p + off : nullptr;
    off = sw_align(off + bytes);
    return q;
  };
  SweepWs w;
  w.ctr = reinterpret_cast<uint32_t*>(take(SW_WORDS * 4));
  for (int c = 0; c < 3; ++c) w.q[c] = reinterpret_cast<uint32_t*>(take((size_t)n * 4));
  for (int c = 0; c < 3; ++c) w.h[c] = reinterpret_cast<uint32_t*>(take((size_t)n * 4));
  for (int c = 0; c < 3; ++c) w.simplex[c] = reinterpret_cast<float*>(take((size_t)n * 6 * 4));
  for (int c = 0; c < 3; ++c) w.gstat[c] = reinterpret_cast<uint8_t*>(take((size_t)n));
  if (ws) *ws = w;
  return off;
}

}  // namespace

extern "C" int cx_launch_narrowphase(cudaStream_t st, int64_t n, const float* a, const float* b,
                                     uint8_t* exists, float* pen, float* cp, uint8_t* status,
                                     const CotixNarrowParams* p) {
  if (n == 0) return 0;
  int sms = cx_sm_count();
  int64_t blocks = (n + 127) / 128;
  // grid = a multiple of the SM count (148 on B200); grid-stride loop covers the rest
  int64_t cap = (int64_t)sms * 16 * 8;
  if (blocks > cap) blocks = cap;
  narrowphase_kernel<<<(unsigned)blocks, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p);
  cx_count_launch();
  return cx_check_launch("narrowphase_kernel");
}
extern "C" int cx_launch_order_clockwise(cudaStream_t st, int64_t n, float* recs) {
  if (n == 0) return 0;
  order_clockwise_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(n, recs);
  cx_count_launch();
  return cx_check_launch("order_clockwise_kernel");
}
extern "C" int cx_launch_support(cudaStream_t st, int64_t n, const float* recs, const float* dirs,
                                 float* out) {
  if (n == 0) return 0;
  support_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(n, recs, dirs, out);
  cx_count_launch();
  return cx_check_launch("support_kernel");
}

extern "C" int64_t cx_narrowphase_workspace_bytes(int64_t n) {
  if (n <= 0) return 0;
  return (int64_t)sweep_carve(n, nullptr, nullptr);
}
extern "C" int cx_launch_narrowphase_ws(cudaStream_t st, int64_t n, const float* a, const float* b,
                                        uint8_t* exists, float* pen, float* cp, uint8_t* status,
                                        const CotixNarrowParams* p, void* workspace, int64_t workspace_bytes) {
  if (n == 0) return 0;
  if (n >= (1ll << 32)) {
    cx_set_error("narrowphase sweep: n must be < 2^32");
    return COTIX_EINVAL;
  }
  SweepWs ws;
  const size_t need = sweep_carve(n, workspace, &ws);
  if (!workspace || (size_t)workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255u)) {
    cx_set_error("narrowphase sweep: workspace must be 256-byte aligned and >= %zu bytes", need);
    return COTIX_EINVAL;
  }
  const int sms = cx_sm_count();
  int64_t blocks = (n + 127) / 128;
  const int64_t cap = (int64_t)sms * 16;
  if (blocks > cap) blocks = cap;
  const unsigned g = (unsigned)blocks;
  if (cudaMemsetAsync(ws.ctr, 0, SW_WORDS * 4, st) != cudaSuccess) return cx_check_launch("sweep memset");
  {
    int64_t tb = (n + SW_TILE - 1) / SW_TILE;
    if (tb > cap) tb = cap;
    sweep_classify_kernel<<<(unsigned)tb, 128, 0, st>>>(n, a, b, exists, pen, cp, status, ws);
  }
  sweep_gjk_kernel<0><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_gjk_kernel<1><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_gjk_kernel<2><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_kernel<0><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_cp_short_kernel<<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_cp_long_kernel<<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_kernel<2><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  cx_count_launch_n(8);
  return cx_check_launch("narrowphase sweep");
}

// ---- no-allocation variant of the wavefront sweep (cotix_narrowphase_static) --------------------
// For callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops,
// harnesses that forbid it): the sweep's scratch (queues, hit lists, GJK simplices; 99 B per pair)
// lives in a file-scope __device__ array of a fixed, modest size, so the call allocates nothing.
// Every GPU that loads the library holds the array (about 13 MB) until the process exits.
// In HOST code the array's name is only a host-side stand-in, not a device address, so the address
// is taken once per device with cudaGetSymbolAddress (that device current) and handed to the same
// launcher the caller-workspace entry point uses.  All calls on a device share the array: the
// launcher zeroes the queue counters at the start of every call (everything else is written before
// it is read), and two calls that overlap on the GPU (different streams) race.
namespace {
constexpr int64_t SW_STATIC_MAX_PAIRS = COTIX_NARROWPHASE_STATIC_MAX_PAIRS;
// sweep_carve(n) for n a multiple of 256: 256 B of counters + (3*4 + 3*4 + 3*24 + 3*1) B per pair
constexpr size_t SW_STATIC_BYTES = 256 + (size_t)SW_STATIC_MAX_PAIRS * 99;
static_assert(SW_STATIC_MAX_PAIRS % 256 == 0, "the static scratch is sized for a multiple of 256 pairs");
static __device__ __align__(256) unsigned char g_sweep_static_scratch[SW_STATIC_BYTES];

std::mutex g_sweep_static_mu;
void* g_sweep_static_ptr[64];  // device address of the array, per device ordinal
bool g_sweep_static_known[64];

// device address of the static scratch on the CURRENT device (taken on first use per device)
int sweep_static_address(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    cx_set_error("cotix_narrowphase_static: cannot tell the current device");
    (void)cudaGetLastError();
    return COTIX_ECUDA;
  }
  std::lock_guard<std::mutex> lock(g_sweep_static_mu);
  if (!g_sweep_static_known[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_sweep_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      cx_set_error("cotix_narrowphase_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                   cudaGetErrorString(e));
      (void)cudaGetLastError();
      return COTIX_ECUDA;
    }
    g_sweep_static_ptr[dev] = p;
    g_sweep_static_known[dev] = true;
  }
  *out = g_sweep_static_ptr[dev];
  return COTIX_OK;
}
}  // namespace

extern "C" int cx_launch_narrowphase_static(cudaStream_t st, int64_t n, const float* a, const float* b,
                                            uint8_t* exists, float* pen, float* cp, uint8_t* status,
                                            const CotixNarrowParams* p) {
  if (n > SW_STATIC_MAX_PAIRS) {
    cx_set_error("cotix_narrowphase_static: n = %lld exceeds COTIX_NARROWPHASE_STATIC_MAX_PAIRS = %lld; "
                 "use cotix_narrowphase_ws with a caller-provided workspace",
                 (long long)n, (long long)SW_STATIC_MAX_PAIRS);
    return COTIX_EINVAL;
  }
  if (n == 0) return 0;
  if (sweep_carve(n, nullptr, nullptr) > SW_STATIC_BYTES) {  // cannot happen for n <= the maximum
    cx_set_error("cotix_narrowphase_static: internal scratch too small");
    return COTIX_EINVAL;
  }
  void* scratch = nullptr;
  const int rc = sweep_static_address(&scratch);
  if (rc != 0) return rc;
  return cx_launch_narrowphase_ws(st, n, a, b, exists, pen, cp, status, p, scratch, (int64_t)SW_STATIC_BYTES);
}
