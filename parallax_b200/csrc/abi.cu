// abi.cu — the C-ABI of libcotix_b200.so (declarations + citations: include/cotix_b200.h).
// Device entry points only enqueue on the caller's stream.  There is no CPU fallback: every
// entry point fails with COTIX_ENODEV / COTIX_ECUDA when no sm_100 device is usable.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>

#include "../../include/cotix_b200.h"
#include "cx_launch.cuh"
#include "cx_lunar.cuh"

// launchers defined in the kernel translation units
extern "C" int cx_launch_narrowphase(cudaStream_t, int64_t, const float*, const float*, uint8_t*,
                                     float*, float*, uint8_t*, const CotixNarrowParams*);
extern "C" int64_t cx_narrowphase_workspace_bytes(int64_t);
extern "C" int cx_launch_narrowphase_ws(cudaStream_t, int64_t, const float*, const float*, uint8_t*, float*,
                                        float*, uint8_t*, const CotixNarrowParams*, void*, int64_t);
extern "C" int cx_launch_narrowphase_static(cudaStream_t, int64_t, const float*, const float*, uint8_t*, float*,
                                            float*, uint8_t*, const CotixNarrowParams*);
extern "C" int cx_launch_order_clockwise(cudaStream_t, int64_t, float*);
extern "C" int cx_launch_support(cudaStream_t, int64_t, const float*, const float*, float*);
extern "C" int cx_launch_lunar_step(cudaStream_t, int64_t, const float*, float*, int64_t,
                                    const float*, float*, uint8_t*, float*, const CotixLunarParams*);
extern "C" int cx_launch_lunar_reset(cudaStream_t, int64_t, const uint32_t*, float*, int64_t,
                                     const CotixLunarParams*);

namespace {
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
std::once_flag g_dev_once;
int g_sm_count = 148;
int g_dev_ok = 0;

void probe_device() {
  int dev = 0;
  cudaDeviceProp prop;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
    g_dev_ok = 0;
    return;
  }
  g_sm_count = prop.multiProcessorCount;
  g_dev_ok = (prop.major >= 10) ? 1 : 0;
}
}  // namespace

extern "C" void cx_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" int cx_sm_count(void) {
  std::call_once(g_dev_once, probe_device);
  return g_sm_count;
}
extern "C" void cx_count_launch(void) {
  // launches recorded while a stream capture is open are counted when the graph is replayed
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  (void)cs;
  g_launches.fetch_add(1, std::memory_order_relaxed);
}
extern "C" void cx_count_launch_n(int64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern "C" int cx_check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cx_set_error("%s: %s", what, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return COTIX_ECUDA;
  }
  return COTIX_OK;
}

#define CX_REQUIRE(cond, msg)            \
  do {                                   \
    if (!(cond)) {                       \
      cx_set_error("%s: %s", __func__, msg); \
      return COTIX_EINVAL;               \
    }                                    \
  } while (0)
#define CX_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t e_ = (call);                                                 \
    if (e_ != cudaSuccess) {                                                 \
      cx_set_error("%s: %s failed: %s", __func__, #call, cudaGetErrorString(e_)); \
      return COTIX_ECUDA;                                                    \
    }                                                                        \
  } while (0)

static int require_device() {
  std::call_once(g_dev_once, probe_device);
  if (!g_dev_ok) {
    cx_set_error("no usable sm_100 CUDA device (this library has no CPU fallback)");
    return COTIX_ENODEV;
  }
  return COTIX_OK;
}

// ---- small test kernels for the leaf primitives --------------------------------------
namespace {
__global__ void euler_kernel(int64_t n, float* px, float* py, const float* vx, const float* vy,
                             float* ang, const float* w, float dt) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  px[k] = px[k] + vx[k] * dt;
  py[k] = py[k] + vy[k] * dt;
  ang[k] = ang[k] + w[k] * dt;
}
__global__ void resolve_kernel(int64_t n, float* body1, float* body2, const float* pen,
                               const float* cp, float baumgarte, float baumgarte_dt) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  BodyS b[2];
  MassP m[2];
  float* src[2] = {body1 + 10 * k, body2 + 10 * k};
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    m[q].mass = src[q][0];
    m[q].inertia = src[q][1];
    b[q].px = src[q][2];
    b[q].py = src[q][3];
    b[q].vx = src[q][4];
    b[q].vy = src[q][5];
    b[q].ang = src[q][6];
    b[q].w = src[q][7];
    m[q].e = src[q][8];
    m[q].inv_mass = 1.0f / src[q][0];
    m[q].dmass = divc_plain(src[q][0]);
    m[q].dinertia = divc_plain(src[q][1]);
    m[q].mu = src[q][9];
  }
  resolve_collision(b[0], m[0], b[1], m[1], mk2(pen[2 * k], pen[2 * k + 1]),
                    mk2(cp[2 * k], cp[2 * k + 1]), baumgarte, baumgarte_dt);
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    src[q][4] = b[q].vx;
    src[q][5] = b[q].vy;
    src[q][7] = b[q].w;
  }
}
__global__ void threefry_kernel(int64_t n, uint32_t k0, uint32_t k1, const uint32_t* x0,
                                const uint32_t* x1, uint32_t* out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  uint32_t o0, o1;
  threefry2x32(k0, k1, x0[k], x1[k], o0, o1);
  out[2 * k] = o0;
  out[2 * k + 1] = o1;
}
__global__ void philox_kernel(int64_t n, uint32_t k0, uint32_t k1, const uint32_t* ctr,
                              uint32_t* out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  uint32_t o[4];
  philox4x32_10(k0, k1, ctr[4 * k], ctr[4 * k + 1], ctr[4 * k + 2], ctr[4 * k + 3], o);
  out[4 * k] = o[0];
  out[4 * k + 1] = o[1];
  out[4 * k + 2] = o[2];
  out[4 * k + 3] = o[3];
}
}  // namespace

extern "C" {

const char* cotix_last_error(void) { return g_err; }
int cotix_abi_version(void) { return COTIX_ABI_VERSION; }
int cotix_math_flavour(void) {
#ifdef COTIX_MATH_FAST
  return 1;
#else
  return 0;
#endif
}
int64_t cotix_launch_count(void) { return g_launches.load(); }

int cotix_device_info(int device, char* name, int name_len, int* sm_count, int* cc_major,
                      int* cc_minor) {
  cudaDeviceProp prop;
  CX_CUDA(cudaGetDeviceProperties(&prop, device));
  if (name && name_len > 0) {
    strncpy(name, prop.name, (size_t)name_len - 1);
    name[name_len - 1] = 0;
  }
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  return COTIX_OK;
}

int cotix_narrowphase(void* stream, int64_t n, const float* a, const float* b, uint8_t* exists,
                      float* pen, float* cp, uint8_t* status, const CotixNarrowParams* params) {
  CX_REQUIRE(n >= 0, "n < 0");
  CX_REQUIRE(params != nullptr, "params is null");
  CX_REQUIRE(n == 0 || (a && b), "null shape records");
  CX_REQUIRE(params->gjk_max_iters >= 0 && params->epa_iters_circle_poly >= 3, "bad iteration caps");
  int rc = require_device();
  if (rc) return rc;
  return cx_launch_narrowphase((cudaStream_t)stream, n, a, b, exists, pen, cp, status, params);
}

int64_t cotix_narrowphase_workspace_bytes(int64_t n) { return cx_narrowphase_workspace_bytes(n); }

int cotix_narrowphase_ws(void* stream, int64_t n, const float* a, const float* b, uint8_t* exists,
                         float* pen, float* cp, uint8_t* status, const CotixNarrowParams* params,
                         void* workspace, int64_t workspace_bytes) {
  CX_REQUIRE(n >= 0, "n < 0");
  CX_REQUIRE(params != nullptr, "params is null");
  CX_REQUIRE(n == 0 || (a && b), "null shape records");
  CX_REQUIRE(params->gjk_max_iters >= 0 && params->epa_iters_circle_poly >= 3, "bad iteration caps");
  int rc = require_device();
  if (rc) return rc;
  return cx_launch_narrowphase_ws((cudaStream_t)stream, n, a, b, exists, pen, cp, status, params, workspace,
                                  workspace_bytes);
}

int cotix_narrowphase_static(void* stream, int64_t n, const float* a, const float* b, uint8_t* exists,
                             float* pen, float* cp, uint8_t* status, const CotixNarrowParams* params) {
  CX_REQUIRE(n >= 0, "n < 0");
  CX_REQUIRE(params != nullptr, "params is null");
  CX_REQUIRE(n == 0 || (a && b), "null shape records");
  CX_REQUIRE(params->gjk_max_iters >= 0 && params->epa_iters_circle_poly >= 3, "bad iteration caps");
  int rc = require_device();
  if (rc) return rc;
  return cx_launch_narrowphase_static((cudaStream_t)stream, n, a, b, exists, pen, cp, status, params);
}

int cotix_narrowphase_host(int device, int64_t n, const float* a, const float* b, uint8_t* exists,
                           float* pen, float* cp, uint8_t* status,
                           const CotixNarrowParams* params) {
  CX_REQUIRE(n >= 0, "n < 0");
  CX_REQUIRE(params != nullptr, "params is null");
  if (n == 0) return COTIX_OK;
  CX_REQUIRE(a && b, "null shape records");
  CX_CUDA(cudaSetDevice(device));
  int rc = require_device();
  if (rc) return rc;
  cudaStream_t st;
  CX_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  const size_t rb = (size_t)n * 64;
  float *da = nullptr, *db = nullptr, *dpen = nullptr, *dcp = nullptr;
  uint8_t *dex = nullptr, *dst = nullptr;
  rc = COTIX_ECUDA;
  do {
    if (cudaMallocAsync(&da, rb, st) != cudaSuccess) break;
    if (cudaMallocAsync(&db, rb, st) != cudaSuccess) break;
    if (cudaMallocAsync(&dpen, (size_t)n * 8, st) != cudaSuccess) break;
    if (cudaMallocAsync(&dcp, (size_t)n * 8, st) != cudaSuccess) break;
    if (cudaMallocAsync(&dex, (size_t)n, st) != cudaSuccess) break;
    if (cudaMallocAsync(&dst, (size_t)n, st) != cudaSuccess) break;
    if (cudaMemcpyAsync(da, a, rb, cudaMemcpyHostToDevice, st) != cudaSuccess) break;
    if (cudaMemcpyAsync(db, b, rb, cudaMemcpyHostToDevice, st) != cudaSuccess) break;
    rc = cx_launch_narrowphase(st, n, da, db, dex, dpen, dcp, dst, params);
    if (rc) break;
    rc = COTIX_ECUDA;
    if (exists && cudaMemcpyAsync(exists, dex, (size_t)n, cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
    if (pen && cudaMemcpyAsync(pen, dpen, (size_t)n * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
    if (cp && cudaMemcpyAsync(cp, dcp, (size_t)n * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
    if (status && cudaMemcpyAsync(status, dst, (size_t)n, cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
    rc = COTIX_OK;
  } while (0);
  if (da) cudaFreeAsync(da, st);
  if (db) cudaFreeAsync(db, st);
  if (dpen) cudaFreeAsync(dpen, st);
  if (dcp) cudaFreeAsync(dcp, st);
  if (dex) cudaFreeAsync(dex, st);
  if (dst) cudaFreeAsync(dst, st);
  cudaError_t e = cudaStreamSynchronize(st);
  cudaStreamDestroy(st);
  if (rc == COTIX_ECUDA || e != cudaSuccess) {
    cx_set_error("cotix_narrowphase_host: %s", cudaGetErrorString(e != cudaSuccess ? e : cudaGetLastError()));
    return COTIX_ECUDA;
  }
  return rc;
}

int cotix_order_clockwise(void* stream, int64_t n, float* recs) {
  CX_REQUIRE(n >= 0 && (n == 0 || recs), "bad arguments");
  int rc = require_device();
  if (rc) return rc;
  return cx_launch_order_clockwise((cudaStream_t)stream, n, recs);
}

int cotix_support(void* stream, int64_t n, const float* recs, const float* dirs, float* out) {
  CX_REQUIRE(n >= 0 && (n == 0 || (recs && dirs && out)), "bad arguments");
  int rc = require_device();
  if (rc) return rc;
  return cx_launch_support((cudaStream_t)stream, n, recs, dirs, out);
}

int cotix_euler_step(void* stream, int64_t n, float* px, float* py, const float* vx,
                     const float* vy, float* ang, const float* angvel, float dt) {
  CX_REQUIRE(n >= 0 && (n == 0 || (px && py && vx && vy && ang && angvel)), "bad arguments");
  int rc = require_device();
  if (rc) return rc;
  if (n == 0) return COTIX_OK;
  euler_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, px, py, vx, vy, ang, angvel, dt);
  cx_count_launch();
  return cx_check_launch("euler_kernel");
}

int cotix_resolve_collision(void* stream, int64_t n, float* body1, float* body2, const float* pen,
                            const float* cp, float baumgarte, float baumgarte_dt) {
  CX_REQUIRE(n >= 0 && (n == 0 || (body1 && body2 && pen && cp)), "bad arguments");
  int rc = require_device();
  if (rc) return rc;
  if (n == 0) return COTIX_OK;
  resolve_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(n, body1, body2, pen, cp, baumgarte, baumgarte_dt);
  cx_count_launch();
  return cx_check_launch("resolve_kernel");
}

int cotix_threefry2x32(void* stream, int64_t n, uint32_t k0, uint32_t k1, const uint32_t* x0,
                       const uint32_t* x1, uint32_t* out) {
  CX_REQUIRE(n >= 0 && (n == 0 || (x0 && x1 && out)), "bad arguments");
  int rc = require_device();
  if (rc) return rc;
  if (n == 0) return COTIX_OK;
  threefry_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, k0, k1, x0, x1, out);
  cx_count_launch();
  return cx_check_launch("threefry_kernel");
}

int cotix_philox4x32(void* stream, int64_t n, uint32_t k0, uint32_t k1, const uint32_t* ctr,
                     uint32_t* out) {
  CX_REQUIRE(n >= 0 && (n == 0 || (ctr && out)), "bad arguments");
  int rc = require_device();
  if (rc) return rc;
  if (n == 0) return COTIX_OK;
  philox_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, k0, k1, ctr, out);
  cx_count_launch();
  return cx_check_launch("philox_kernel");
}

// ---- device memory, streams and events for hosts without a CUDA binding of their own ----------
int cotix_set_device(int device) {
  CX_CUDA(cudaSetDevice(device));
  return COTIX_OK;
}
int cotix_device_count(int* n) {
  CX_REQUIRE(n != nullptr, "n is null");
  *n = 0;
  if (cudaGetDeviceCount(n) != cudaSuccess) {
    cudaGetLastError();
    *n = 0;
  }
  return COTIX_OK;
}
int cotix_device_malloc(int64_t bytes, void** out) {
  CX_REQUIRE(out != nullptr && bytes >= 0, "bad arguments");
  *out = nullptr;
  if (bytes == 0) return COTIX_OK;
  CX_CUDA(cudaMalloc(out, (size_t)bytes));
  return COTIX_OK;
}
int cotix_device_free(void* p) {
  if (p) CX_CUDA(cudaFree(p));
  return COTIX_OK;
}
int cotix_host_malloc(int64_t bytes, void** out) {
  CX_REQUIRE(out != nullptr && bytes >= 0, "bad arguments");
  *out = nullptr;
  if (bytes == 0) return COTIX_OK;
  CX_CUDA(cudaHostAlloc(out, (size_t)bytes, cudaHostAllocDefault));
  return COTIX_OK;
}
int cotix_host_free(void* p) {
  if (p) CX_CUDA(cudaFreeHost(p));
  return COTIX_OK;
}
int cotix_memcpy(void* stream, void* dst, const void* src, int64_t bytes) {
  CX_REQUIRE(bytes >= 0 && (bytes == 0 || (dst && src)), "bad arguments");
  if (bytes) CX_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return COTIX_OK;
}
int cotix_memset(void* stream, void* dst, int value, int64_t bytes) {
  CX_REQUIRE(bytes >= 0 && (bytes == 0 || dst), "bad arguments");
  if (bytes) CX_CUDA(cudaMemsetAsync(dst, value, (size_t)bytes, (cudaStream_t)stream));
  return COTIX_OK;
}
int cotix_stream_create(void** out) {
  CX_REQUIRE(out != nullptr, "out is null");
  cudaStream_t s;
  CX_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *out = (void*)s;
  return COTIX_OK;
}
int cotix_stream_destroy(void* stream) {
  if (stream) CX_CUDA(cudaStreamDestroy((cudaStream_t)stream));
  return COTIX_OK;
}
int cotix_stream_synchronize(void* stream) {
  CX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return COTIX_OK;
}
int cotix_event_create(void** out) {
  CX_REQUIRE(out != nullptr, "out is null");
  cudaEvent_t e;
  CX_CUDA(cudaEventCreate(&e));
  *out = (void*)e;
  return COTIX_OK;
}
int cotix_event_destroy(void* event) {
  if (event) CX_CUDA(cudaEventDestroy((cudaEvent_t)event));
  return COTIX_OK;
}
int cotix_event_record(void* event, void* stream) {
  CX_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return COTIX_OK;
}
int cotix_event_elapsed_ms(void* start, void* stop, float* ms) {
  CX_REQUIRE(ms != nullptr, "ms is null");
  CX_CUDA(cudaEventSynchronize((cudaEvent_t)stop));
  CX_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
  return COTIX_OK;
}

}  // extern "C"
