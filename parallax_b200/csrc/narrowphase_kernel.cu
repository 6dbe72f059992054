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
