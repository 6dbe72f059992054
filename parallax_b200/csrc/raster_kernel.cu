// raster_kernel.cu — K6: headless rasteriser replacing the pygame window of cotix/_viz.py (no
// graphics stack on the box: the 800 x 600 "screen" is an RGB8 buffer in device memory).
//
// What it stands in for (cotix/_viz.py):
//   PyPainter.draw_line :23-26 (pygame.draw.line, width 1), draw_circle :17-21 (filled),
//   draw_polygon :28-31 (filled), next :43-49 (flip, then screen.fill((0, 0, 0))),
//   _scale_vector :33-38 world [-8, 8] x [-6, 6] -> pixels [0, 800] x [0, 600] (y flipped,
//   int() truncation), _scale_size :40-41 (int(size * 40)),
// and for a LunarLander env the draw calls of LunarLander.draw (cotix/_lunar_lander.py:220-225):
// the edges of every polygon part of every body in white (Polygon.drawEdges,
// cotix/_convex_shapes.py:189-194) and the two red landing-pad markers.
//
// Pixel rules (pygame itself is not installed anywhere here, so pixel-for-pixel agreement with
// SDL's rasteriser is UNPINNED; the rules below are this engine's definition and the CPU oracle
// restates them):
//   line     1 pixel wide, end points included, drawn from its lexicographically smaller end
//            (x, then y): along the major axis one pixel per step i, the minor coordinate is
//            m0 + sign * floor((2 i dminor + dmajor) / (2 dmajor))
//   circle   radius r >= 1 (pygame draws nothing for r < 1): pixels with dx^2 + dy^2 <= r^2
//   polygon  convex, 3..6 vertices: pixels whose edge functions are all >= 0 or all <= 0
//   primitives are painted in list order onto a black image; off-screen pixels are dropped.
//
// B200 mapping: the work is write-bound (1.44 MB per image, a few thousand lit pixels).  One
// block owns a band of 24 rows of one image: it clears the band in SHARED memory (57.6 KB),
// paints the primitives that reach the band there — in order, all threads of the block sharing
// the pixels of one primitive, every pixel in closed form so there is no serial Bresenham walk —
// and then moves the finished band to HBM with one TMA bulk copy (cp.async.bulk shared -> global),
// i.e. every byte of the image is written to HBM exactly once, by the copy engine.
#include "cx_launch.cuh"
#include "cx_lunar_step.cuh"

namespace {

constexpr int IMG_W = COTIX_IMG_W, IMG_H = COTIX_IMG_H;
#ifndef CX_BAND_ROWS
#define CX_BAND_ROWS 24
#endif
constexpr int BAND_ROWS = CX_BAND_ROWS;              // 600 = 25 * 24
constexpr int BANDS = IMG_H / BAND_ROWS;
constexpr int ROW_BYTES = IMG_W * 3;
constexpr int BAND_BYTES = BAND_ROWS * ROW_BYTES;   // 57 600, a multiple of 16
constexpr int RTPB = 256;
constexpr int PW = COTIX_PRIM_WORDS;
constexpr int LUNAR_PRIMS = 44;
constexpr int LUNAR_BANDS_PER_BLOCK = 5;             // consecutive bands painted by one block (both kernels)
static_assert(BANDS % LUNAR_BANDS_PER_BLOCK == 0, "bands per block");
constexpr int CHUNK = RTPB / (PW / 4);                // 64 records staged per pass
static_assert(IMG_H % BAND_ROWS == 0 && BAND_BYTES % 16 == 0, "band layout");

CX_DEV void put_pixel(uint8_t* band, int row0, int x, int y, uint32_t rgb) {
  if (x < 0 || x >= IMG_W || y < row0 || y >= row0 + BAND_ROWS) return;
  uint8_t* p = band + (y - row0) * ROW_BYTES + x * 3;
  p[0] = (uint8_t)(rgb >> 16);
  p[1] = (uint8_t)(rgb >> 8);
  p[2] = (uint8_t)rgb;
}
CX_DEV int imin(int a, int b) { return a < b ? a : b; }
CX_DEV int imax(int a, int b) { return a > b ? a : b; }

// minor-axis offset of step i of a line: floor((2 i dminor + dmajor) / (2 dmajor)), dmajor > 0
CX_DEV int line_offset(int i, int dminor, int dmajor) {
  if (dmajor < 16384)  // i <= dmajor, dminor <= dmajor: fits 32 bits (the on-screen case)
    return (int)((2u * (unsigned)i * (unsigned)dminor + (unsigned)dmajor) / (2u * (unsigned)dmajor));
  return (int)((2ll * i * dminor + dmajor) / (2ll * dmajor));
}

// paints primitive `q` (16 words) into the band [row0, row0 + BAND_ROWS) with the workers
// first, first + stride, ... (a warp: lane, 32)
CX_DEV void paint_prim(const int32_t* q, uint8_t* band, int row0, int first, int stride) {
  const int type = q[0];
  const uint32_t rgb = (uint32_t)q[1];
  const int r1 = row0 + BAND_ROWS - 1;
  if (type == COTIX_PRIM_LINE) {
    int x0 = q[4], y0 = q[5], x1 = q[6], y1 = q[7];
    if (x1 < x0 || (x1 == x0 && y1 < y0)) {
      int t = x0; x0 = x1; x1 = t;
      t = y0; y0 = y1; y1 = t;
    }
    if (x1 < 0 || x0 >= IMG_W || imax(y0, y1) < row0 || imin(y0, y1) > r1) return;
    const int dx = x1 - x0;                       // >= 0
    const int dy = y1 >= y0 ? y1 - y0 : y0 - y1;
    const int sy = y1 >= y0 ? 1 : -1;
    if (dx >= dy) {                               // x-major: i = x - x0
      const int lo = imax(0, -x0), hi = imin(dx, IMG_W - 1 - x0);
      for (int i = lo + first; i <= hi; i += stride) {
        const int off = dx == 0 ? 0 : line_offset(i, dy, dx);
        put_pixel(band, row0, x0 + i, y0 + sy * off, rgb);
      }
    } else {                                      // y-major: i = |y - y0|
      int lo, hi;
      if (sy > 0) { lo = imax(0, row0 - y0); hi = imin(dy, r1 - y0); }
      else        { lo = imax(0, y0 - r1);   hi = imin(dy, y0 - row0); }
      for (int i = lo + first; i <= hi; i += stride)
        put_pixel(band, row0, x0 + line_offset(i, dx, dy), y0 + sy * i, rgb);
    }
  } else if (type == COTIX_PRIM_CIRCLE) {
    const int cx = q[4], cy = q[5], r = q[2];
    if (r < 1) return;
    const int xa = imax(0, cx - r), xb = imin(IMG_W - 1, cx + r);
    const int ya = imax(row0, cy - r), yb = imin(r1, cy + r);
    if (xa > xb || ya > yb) return;
    const int w = xb - xa + 1, n = w * (yb - ya + 1);
    const long long rr = (long long)r * r;
    for (int i = first; i < n; i += stride) {
      const int x = xa + i % w, y = ya + i / w;
      const long long ddx = x - cx, ddy = y - cy;
      if (ddx * ddx + ddy * ddy <= rr) put_pixel(band, row0, x, y, rgb);
    }
  } else if (type == COTIX_PRIM_POLYGON) {
    const int nv = q[2];
    if (nv < 3 || nv > 6) return;
    int xa = q[4], xb = q[4], ya = q[5], yb = q[5];
    for (int k = 1; k < nv; ++k) {
      xa = imin(xa, q[4 + 2 * k]); xb = imax(xb, q[4 + 2 * k]);
      ya = imin(ya, q[5 + 2 * k]); yb = imax(yb, q[5 + 2 * k]);
    }
    xa = imax(xa, 0); xb = imin(xb, IMG_W - 1);
    ya = imax(ya, row0); yb = imin(yb, r1);
    if (xa > xb || ya > yb) return;
    const int w = xb - xa + 1, n = w * (yb - ya + 1);
    for (int i = first; i < n; i += stride) {
      const int x = xa + i % w, y = ya + i / w;
      bool pos = true, neg = true;
      for (int k = 0; k < nv; ++k) {
        const int j = k + 1 == nv ? 0 : k + 1;
        const long long ex = q[4 + 2 * j] - q[4 + 2 * k], ey = q[5 + 2 * j] - q[5 + 2 * k];
        const long long e = ex * (y - q[5 + 2 * k]) - ey * (x - q[4 + 2 * k]);
        pos = pos && (e >= 0);
        neg = neg && (e <= 0);
      }
      if (pos || neg) put_pixel(band, row0, x, y, rgb);
    }
  }
}

// clear -> paint in order -> one bulk copy of the band to HBM
CX_DEV void clear_band(uint8_t* band) {
  uint4* b4 = reinterpret_cast<uint4*>(band);
  for (int i = threadIdx.x; i < BAND_BYTES / 16; i += RTPB) b4[i] = make_uint4(0u, 0u, 0u, 0u);
}
// a run of primitives of one colour is a layer: its primitives are painted concurrently, one warp
// each (racing writes store the same bytes); layers are separated by a barrier.  Ends with a barrier.
CX_DEV void paint_layers(const int32_t* prims, int count, uint8_t* band, int row0) {  // count <= 64
  __shared__ uint32_t s_start[2];  // bit j: primitive j starts a layer
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < 64) {
    const int j = threadIdx.x;
    const bool st = j < count && (j == 0 || prims[j * PW + 1] != prims[(j - 1) * PW + 1]);
    const unsigned m = __ballot_sync(0xFFFFFFFFu, st);
    if (lane == 0) s_start[warp] = m;
  }
  __syncthreads();
  const unsigned long long starts = (unsigned long long)s_start[0] | ((unsigned long long)s_start[1] << 32);
  int k = 0;
  while (k < count) {
    const unsigned long long rest = starts & ~((2ull << k) - 1ull);
    const int e = rest ? __ffsll((long long)rest) - 1 : count;
    for (int j = k + warp; j < e; j += RTPB / 32) paint_prim(prims + j * PW, band, row0, lane, 32);
    k = e;
    __syncthreads();
  }
}
CX_DEV void store_band(const uint8_t* band, uint8_t* gdst) {
  // generic-proxy writes to shared memory -> visible to the async proxy (TMA) -> bulk store
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(band);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gdst), "r"(saddr), "r"((uint32_t)BAND_BYTES) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}

// _scale_vector (cotix/_viz.py:33-38) in float32, int() truncation; NaN / huge -> far off screen
CX_DEV int to_int_trunc(float v) {
  if (!(v == v)) return -(1 << 20);
  if (v > 1048576.0f) return 1 << 20;
  if (v < -1048576.0f) return -(1 << 20);
  return (int)v;
}
CX_DEV int scale_x(float x) { return to_int_trunc((x + 8.0f) / 16.0f * 800.0f); }
CX_DEV int scale_y(float y) { return to_int_trunc((-y + 6.0f) / 12.0f * 600.0f); }

CX_DEV void emit_line(int32_t* q, uint32_t rgb, float ax, float ay, float bx, float by) {
  q[0] = COTIX_PRIM_LINE;
  q[1] = (int32_t)rgb;
  q[2] = 0;
  q[3] = 0;
  q[4] = scale_x(ax);
  q[5] = scale_y(ay);
  q[6] = scale_x(bx);
  q[7] = scale_y(by);
#pragma unroll
  for (int k = 8; k < PW; ++k) q[k] = 0;
}
struct PadLines {  // the literal marker lines of cotix/_lunar_lander.py:223-224, scaled on the host
  int32_t v[2][4];
};

CX_DEV V2 world_vertex(const float v[2], float px, float py, float c, float s) {  // as world_poly
  const float ms = -s;
  return mk2(c * v[0] + ms * v[1] + px * 1.0f, s * v[0] + c * v[1] + py * 1.0f);
}
// draw call t (0..43) of LunarLander.draw for env e: 6 + 4 + 4 body edges, 7 x 4 ground edges, 2 pad
// markers.  One thread per draw call.
CX_DEV void lunar_build_prim(const float* __restrict__ state, int64_t ld, int64_t e,
                             const CotixLunarParams& P, const PadLines& pad, int t, int32_t* q) {
  if (t < 14) {  // edge k = (v_k, v_{k-1}) of body b, cotix/_convex_shapes.py:160-163
    const int b = t < 6 ? 0 : (t < 10 ? 1 : 2);
    const int n = b == 0 ? 6 : 4;
    const int k = t - (b == 0 ? 0 : (b == 1 ? 6 : 10));
    const int j = k == 0 ? n - 1 : k - 1;
    const float px = state[(b * 6 + 0) * ld + e], py = state[(b * 6 + 1) * ld + e];
    float s, c;
    cx_sincosf(state[(b * 6 + 4) * ld + e], &s, &c);
    const float (*loc)[2] = b == 0 ? P.lander_local : P.leg_local[b - 1];
    const V2 a = world_vertex(loc[k], px, py, c, s), z = world_vertex(loc[j], px, py, c, s);
    emit_line(q, 0xFFFFFFu, a.x, a.y, z.x, z.y);
  } else if (t < 42) {  // ground trapezoid g: top edge, right side, floor, left side
    const int g = (t - 14) >> 2, side = (t - 14) & 3;
    float kn[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) kn[i] = state[(COTIX_F_KNOT0 + i) * ld + e];
    const float x0 = terrain_x(kn, P, g), x1 = terrain_x(kn, P, g + 1);
    const float h0 = state[(COTIX_F_HEIGHT0 + g) * ld + e], h1 = state[(COTIX_F_HEIGHT0 + g + 1) * ld + e];
    const float fl = P.ground_floor;
    if (side == 0) emit_line(q, 0xFFFFFFu, x0, h0, x1, h1);
    else if (side == 1) emit_line(q, 0xFFFFFFu, x1, h1, x1, fl);
    else if (side == 2) emit_line(q, 0xFFFFFFu, x1, fl, x0, fl);
    else emit_line(q, 0xFFFFFFu, x0, fl, x0, h0);
  } else {
    const int l = t - 42;
    q[0] = COTIX_PRIM_LINE;
    q[1] = 0xFF0000;
    q[2] = q[3] = 0;
    q[4] = pad.v[l][0];
    q[5] = pad.v[l][1];
    q[6] = pad.v[l][2];
    q[7] = pad.v[l][3];
#pragma unroll
    for (int k = 8; k < PW; ++k) q[k] = 0;
  }
}

__global__ void __launch_bounds__(RTPB)
raster_draw_kernel(int64_t img0, const int32_t* __restrict__ prims, const int32_t* __restrict__ counts,
                   int max_prims, uint8_t* __restrict__ image) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int band = blockIdx.x;
  const int64_t img = img0 + blockIdx.y;
  int count = counts ? counts[img] : max_prims;
  if (count > max_prims) count = max_prims;
  if (count < 0) count = 0;
  int32_t* sprims = reinterpret_cast<int32_t*>(smem + BAND_BYTES);
  const int32_t* gp = prims + img * (int64_t)max_prims * PW;
  // a block paints LUNAR_BANDS_PER_BLOCK consecutive bands; the list is staged through shared memory
  // CHUNK records at a time (one 16-byte load per thread), once if it fits
  for (int b = 0; b < LUNAR_BANDS_PER_BLOCK; ++b) {
    const int bi = band * LUNAR_BANDS_PER_BLOCK + b;
    __syncthreads();  // the previous band's bulk store has read the buffer
    clear_band(smem);
    for (int c0 = 0; c0 < count; c0 += CHUNK) {
      const int n = imin(CHUNK, count - c0);
      if (b == 0 || count > CHUNK) {
        __syncthreads();
        if ((int)threadIdx.x < n * (PW / 4))
          reinterpret_cast<int4*>(sprims)[threadIdx.x] = reinterpret_cast<const int4*>(gp + (int64_t)c0 * PW)[threadIdx.x];
      }
      __syncthreads();
      paint_layers(sprims, n, smem, bi * BAND_ROWS);
    }
    store_band(smem, image + img * (int64_t)IMG_H * ROW_BYTES + (int64_t)bi * BAND_BYTES);
  }
}

__global__ void __launch_bounds__(RTPB)
lunar_draw_kernel(int64_t img0, const float* __restrict__ state, int64_t ld,
                  const int64_t* __restrict__ env_index, uint8_t* __restrict__ image,
                  const __grid_constant__ CotixLunarParams P, const PadLines pad) {
  extern __shared__ __align__(128) uint8_t smem[];
  int32_t* sprims = reinterpret_cast<int32_t*>(smem + BAND_BYTES);
  const int band = blockIdx.x;
  const int64_t img = img0 + blockIdx.y;
  const int64_t e = env_index ? env_index[img] : img;
  if (threadIdx.x < LUNAR_PRIMS) lunar_build_prim(state, ld, e, P, pad, threadIdx.x, sprims + threadIdx.x * PW);
  // the draw calls are built once per block and reused for LUNAR_BANDS_PER_BLOCK consecutive bands
  for (int b = 0; b < LUNAR_BANDS_PER_BLOCK; ++b) {
    const int bi = band * LUNAR_BANDS_PER_BLOCK + b;
    __syncthreads();  // draw calls visible; the previous band's bulk store has read the buffer
    clear_band(smem);
    __syncthreads();
    paint_layers(sprims, LUNAR_PRIMS, smem, bi * BAND_ROWS);
    store_band(smem, image + img * (int64_t)IMG_H * ROW_BYTES + (int64_t)bi * BAND_BYTES);
  }
}

__global__ void lunar_prims_kernel(int64_t n, const float* __restrict__ state, int64_t ld,
                                   const int64_t* __restrict__ env_index, int32_t* __restrict__ prims,
                                   const __grid_constant__ CotixLunarParams P, const PadLines pad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // (image, draw call)
  if (i >= n * LUNAR_PRIMS) return;
  const int64_t img = i / LUNAR_PRIMS;
  lunar_build_prim(state, ld, env_index ? env_index[img] : img, P, pad, (int)(i % LUNAR_PRIMS), prims + i * PW);
}

// Python's int((v + 8) / 16 * 800) on Python floats (float64), as LunarLander.draw passes them
int py_scale_x(double x) { return (int)((x + 8) / 16 * 800); }
int py_scale_y(double y) { return (int)((-y + 6) / 12 * 600); }
PadLines pad_lines() {
  PadLines p;
  p.v[0][0] = py_scale_x(-2);
  p.v[0][1] = py_scale_y(-1.8);
  p.v[0][2] = py_scale_x(-2);
  p.v[0][3] = py_scale_y(-1.0);
  p.v[1][0] = py_scale_x(2);
  p.v[1][1] = py_scale_y(-1.8);
  p.v[1][2] = py_scale_x(2);
  p.v[1][3] = py_scale_y(-1.0);
  return p;
}

}  // namespace

extern "C" int cx_launch_raster_draw(cudaStream_t st, int64_t n_img, const int32_t* prims,
                                     const int32_t* counts, int32_t max_prims, uint8_t* image) {
  if (n_img == 0) return 0;
  static bool once = false;
  if (!once) {
    cudaFuncSetAttribute(raster_draw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BAND_BYTES + CHUNK * PW * 4);
    once = true;
  }
  for (int64_t i0 = 0; i0 < n_img; i0 += 65535) {
    const int64_t n = n_img - i0 < 65535 ? n_img - i0 : 65535;
    raster_draw_kernel<<<dim3(BANDS / LUNAR_BANDS_PER_BLOCK, (unsigned)n), RTPB, BAND_BYTES + CHUNK * PW * 4, st>>>(i0, prims, counts, max_prims, image);
    cx_count_launch();
  }
  return cx_check_launch("raster_draw_kernel");
}

extern "C" int cx_launch_lunar_draw(cudaStream_t st, int64_t n_img, const float* state, int64_t ld,
                                    const int64_t* env_index, uint8_t* image, const CotixLunarParams* p) {
  if (n_img == 0) return 0;
  const int smem = BAND_BYTES + LUNAR_PRIMS * PW * 4;
  static bool once = false;
  if (!once) {
    cudaFuncSetAttribute(lunar_draw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    once = true;
  }
  const PadLines pad = pad_lines();
  for (int64_t i0 = 0; i0 < n_img; i0 += 65535) {
    const int64_t n = n_img - i0 < 65535 ? n_img - i0 : 65535;
    lunar_draw_kernel<<<dim3(BANDS / LUNAR_BANDS_PER_BLOCK, (unsigned)n), RTPB, smem, st>>>(i0, state, ld, env_index, image, *p, pad);
    cx_count_launch();
  }
  return cx_check_launch("lunar_draw_kernel");
}

extern "C" int cx_launch_lunar_prims(cudaStream_t st, int64_t n, const float* state, int64_t ld,
                                     const int64_t* env_index, int32_t* prims, const CotixLunarParams* p) {
  if (n == 0) return 0;
  lunar_prims_kernel<<<(unsigned)((n * LUNAR_PRIMS + 127) / 128), 128, 0, st>>>(n, state, ld, env_index, prims, *p, pad_lines());
  cx_count_launch();
  return cx_check_launch("lunar_prims_kernel");
}
