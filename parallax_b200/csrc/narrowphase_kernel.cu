// narrowphase_kernel.cu — K3: GJK/EPA narrowphase sweep, one shape pair per thread.
//
// Replaces the per-pair contact functions of cotix/_contacts.py:30-315 for a flat batch.
// HBM traffic per pair: 2 x 64 B records in, 1 + 8 + 8 (+1) B out (148 B).
// Records are read as four 16-byte vector loads per shape (a warp covers 2 KB of
// contiguous records per array); outputs are written as float2 / bytes, coalesced.
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
  float* simplex;    // [6][n], indexed by pair
  uint8_t* gstat;    // [n] GJK status bits of the hits
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
CX_DEV void sw_store(int64_t k, const Contact& c, uint32_t extra, uint8_t* exists, float* pen, float* cp,
                     uint8_t* status) {
  if (exists) exists[k] = contact_isnan(c) ? 0 : 1;
  if (pen) reinterpret_cast<float2*>(pen)[k] = make_float2(c.pen.x, c.pen.y);
  if (cp) reinterpret_cast<float2*>(cp)[k] = make_float2(c.cp.x, c.cp.y);
  if (status) status[k] = (uint8_t)(c.status | extra);
}

__global__ void __launch_bounds__(128)
sweep_classify_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                      uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                      uint8_t* __restrict__ status, const SweepWs ws) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t first = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31);
  for (int64_t k0 = first; k0 < n; k0 += stride) {
    const int64_t k = k0 + (threadIdx.x & 31);
    int cls = -1;
    if (k < n) {
      // type word only: the polygon classes read their records in the GJK pass
      const int ta = (int)__ldg(a + k * COTIX_SHAPE_REC_FLOATS), tb = (int)__ldg(b + k * COTIX_SHAPE_REC_FLOATS);
      if (tb == COTIX_SHAPE_POLYGON && ta >= COTIX_SHAPE_CIRCLE && ta <= COTIX_SHAPE_POLYGON) {
        cls = ta == COTIX_SHAPE_POLYGON ? 0 : (ta == COTIX_SHAPE_CIRCLE ? 1 : 2);
      } else {
        Contact c = contact_none();
        uint32_t bad = 0u;
        if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_CIRCLE) {
          c = circle_vs_circle(rec_circ(load_rec(a, k)), rec_circ(load_rec(b, k)));
        } else if (ta == COTIX_SHAPE_CIRCLE && tb == COTIX_SHAPE_AABB) {
          c = circle_vs_aabb(rec_circ(load_rec(a, k)), rec_box(load_rec(b, k)));
        } else if (ta == COTIX_SHAPE_AABB && tb == COTIX_SHAPE_AABB) {
          c = aabb_vs_aabb(rec_box(load_rec(a, k)), rec_box(load_rec(b, k)));
        } else {
          bad = 0x80u;  // not a key of _contact_funcs in this order
        }
        sw_store(k, c, bad, exists, pen, cp, status);
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const uint32_t slot = sw_push(&ws.ctr[SW_Q0 + c], cls == c);
      if (cls == c) ws.q[c][slot] = (uint32_t)k;
    }
  }
}

template <int CLS>
__global__ void __launch_bounds__(128)
sweep_gjk_kernel(int64_t n, const float* __restrict__ a, const float* __restrict__ b,
                 uint8_t* __restrict__ exists, float* __restrict__ pen, float* __restrict__ cp,
                 uint8_t* __restrict__ status, CotixNarrowParams prm, const SweepWs ws) {
  const uint32_t cnt = ws.ctr[SW_Q0 + CLS];
  const uint32_t stride = gridDim.x * blockDim.x;
  const V2 start = mk2(prm.init_dir[0], prm.init_dir[1]);
  for (uint32_t i0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); i0 < cnt; i0 += stride) {
    const uint32_t i = i0 + (threadIdx.x & 31u);
    bool hit = false;
    uint32_t k = 0u;
    if (i < cnt) {
      k = ws.q[CLS][i];
      const Rec ra = load_rec(a, k), rb = load_rec(b, k);
      const Poly B = rec_poly(rb);
      V2 s0, s1, s2;
      bool gcap;
      if (CLS == 0) hit = gjk(rec_poly(ra), B, start, prm.gjk_max_iters, s0, s1, s2, gcap);
      else if (CLS == 1) hit = gjk(rec_circ(ra), B, start, prm.gjk_max_iters, s0, s1, s2, gcap);
      else hit = gjk(rec_box(ra), B, start, prm.gjk_max_iters, s0, s1, s2, gcap);
      if (hit) {
        ws.simplex[0 * n + k] = s0.x; ws.simplex[1 * n + k] = s0.y;
        ws.simplex[2 * n + k] = s1.x; ws.simplex[3 * n + k] = s1.y;
        ws.simplex[4 * n + k] = s2.x; ws.simplex[5 * n + k] = s2.y;
        ws.gstat[k] = gcap ? (uint8_t)CX_ST_GJK_CAP : (uint8_t)0;
      } else {
        Contact c = contact_none();
        if (gcap) c.status |= CX_ST_GJK_CAP;
        sw_store(k, c, 0u, exists, pen, cp, status);
      }
    }
    const uint32_t slot = sw_push(&ws.ctr[SW_H0 + CLS], hit);
    if (hit) ws.h[CLS][slot] = k;
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
    const V2 s0 = mk2(ws.simplex[0 * n + k], ws.simplex[1 * n + k]);
    const V2 s1 = mk2(ws.simplex[2 * n + k], ws.simplex[3 * n + k]);
    const V2 s2 = mk2(ws.simplex[4 * n + k], ws.simplex[5 * n + k]);
    Contact c = contact_none();
    c.status = ws.gstat[k];
    if (CLS == 0) convex_poly_finish(rec_poly(ra), B, s0, s1, s2, c);
    else if (CLS == 1) circle_poly_finish<131>(rec_circ(ra), B, s0, s1, s2, prm.epa_iters_circle_poly, c);
    else convex_poly_finish(rec_box(ra), B, s0, s1, s2, c);
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
  w.simplex = reinterpret_cast<float*>(take((size_t)n * 6 * 4));
  w.gstat = reinterpret_cast<uint8_t*>(take((size_t)n));
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
  sweep_classify_kernel<<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, ws);
  sweep_gjk_kernel<0><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_gjk_kernel<1><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_gjk_kernel<2><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_kernel<0><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_kernel<1><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  sweep_epa_kernel<2><<<g, 128, 0, st>>>(n, a, b, exists, pen, cp, status, *p, ws);
  cx_count_launch_n(7);
  return cx_check_launch("narrowphase sweep");
}
