// abi_lunar.cu — LunarLander entry points of the C-ABI (include/cotix_b200.h).
#include <mutex>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <limits>
#include <map>
#include <new>
#include <vector>

#include "../../include/cotix_b200.h"
#include "cx_launch.cuh"
#include "cx_pipeline.h"
#include "cx_rng.cuh"

extern "C" int cx_launch_lunar_step(cudaStream_t, int64_t, const float*, float*, int64_t,
                                    const float*, float*, uint8_t*, float*, const CotixLunarParams*);
extern "C" int cx_launch_lunar_rollout_kernel(cudaStream_t, int64_t, const float*, float*, int64_t, int64_t, const float*,
                                              int64_t, int, int, float*, uint8_t*, float*, const CotixLunarParams*);
extern "C" int cx_launch_lunar_reset(cudaStream_t, int64_t, const uint32_t*, float*, int64_t,
                                     const CotixLunarParams*);

extern "C" int cx_launch_lunar_step_bwd(cudaStream_t, int64_t, const float*, const float*, int64_t,
                                        const float*, const float*, float*, float*, uint32_t*,
                                        const float*, const float*, const CotixLunarParams*);
extern "C" int cx_launch_lunar_sweep_bwd(cudaStream_t, int64_t, int, int, const float*, int64_t, const float*, float*,
                                         float*, uint32_t*, const float*, const float*, const CotixLunarParams*);
extern "C" int cx_launch_lunar_decide(cudaStream_t, int64_t, int, const float*, int64_t, const float*, float*,
                                      uint32_t*, uint32_t*, uint32_t*, float*, uint32_t, const CotixLunarParams*,
                                      int);
extern "C" int cx_launch_raster_draw(cudaStream_t, int64_t, const int32_t*, const int32_t*, int32_t, uint8_t*);
extern "C" int cx_launch_lunar_draw(cudaStream_t, int64_t, const float*, int64_t, const int64_t*, uint8_t*,
                                    const CotixLunarParams*);
extern "C" int cx_launch_lunar_prims(cudaStream_t, int64_t, const float*, int64_t, const int64_t*, int32_t*,
                                     const CotixLunarParams*);

#define CXL_REQUIRE(cond, msg)               \
  do {                                       \
    if (!(cond)) {                           \
      cx_set_error("%s: %s", __func__, msg); \
      return COTIX_EINVAL;                   \
    }                                        \
  } while (0)
#define CXL_CUDA(call)                                                            \
  do {                                                                            \
    cudaError_t e_ = (call);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      cx_set_error("%s: %s failed: %s", __func__, #call, cudaGetErrorString(e_)); \
      return COTIX_ECUDA;                                                         \
    }                                                                             \
  } while (0)

namespace {

// ---- host-side constants of the scene (computed once, with the engine's own trig) ------
void host_order_clockwise(float (*v)[2], int n) {  // cotix/_geometry_utils.py:60-67
  float sx = 0.0f, sy = 0.0f;
  for (int i = 0; i < n; ++i) {
    sx = sx + v[i][0];
    sy = sy + v[i][1];
  }
  float mx = sx / (float)n, my = sy / (float)n;
  float ang[8];
  int idx[8];
  for (int i = 0; i < n; ++i) {
    ang[i] = cx_atan2f(v[i][1] - my, v[i][0] - mx);
    idx[i] = i;
  }
  for (int i = 1; i < n; ++i) {  // stable insertion sort
    int k = idx[i];
    int j = i - 1;
    while (j >= 0 && ang[idx[j]] > ang[k]) {
      idx[j + 1] = idx[j];
      --j;
    }
    idx[j + 1] = k;
  }
  float o[8][2];
  for (int i = 0; i < n; ++i) {
    o[i][0] = v[idx[i]][0];
    o[i][1] = v[idx[i]][1];
  }
  for (int i = 0; i < n; ++i) {
    v[i][0] = o[i][0];
    v[i][1] = o[i][1];
  }
}

// XLA's float32 erf_inv (Giles), used by jax.random.normal
float host_erf_inv(float x) {
  float w = -log1pf(-x * x);
  float p;
  if (w < 5.0f) {
    w = w - 2.5f;
    p = 2.81022636e-08f;
    p = 3.43273939e-07f + p * w;
    p = -3.5233877e-06f + p * w;
    p = -4.39150654e-06f + p * w;
    p = 0.00021858087f + p * w;
    p = -0.00125372503f + p * w;
    p = -0.00417768164f + p * w;
    p = 0.246640727f + p * w;
    p = 1.50140941f + p * w;
  } else {
    w = sqrtf(w) - 3.0f;
    p = -0.000200214257f;
    p = 0.000100950558f + p * w;
    p = 0.00134934322f + p * w;
    p = -0.00367342844f + p * w;
    p = 0.00573950773f + p * w;
    p = -0.0076224613f + p * w;
    p = 0.00943887047f + p * w;
    p = 1.00167406f + p * w;
    p = 2.83297682f + p * w;
  }
  return p * x;
}
// random_direction(PRNGKey(1)): cotix/_geometry_utils.py:37-46 via _collisions.py:287-298
void host_init_dir(float out[2]) {
  // jr.normal(key, (2,)): bits = threefry(key, [0], [1]); uniform in (-1, 1); sqrt(2)*erfinv
  uint32_t o0, o1;
  threefry2x32(0u, 1u, 0u, 1u, o0, o1);
  const float lo = nextafterf(-1.0f, 0.0f), hi = 1.0f;
  uint32_t bits[2] = {o0, o1};
  float nrm[2];
  for (int i = 0; i < 2; ++i) {
    uint32_t fb = (bits[i] >> 9) | 0x3F800000u;
    float f;
    memcpy(&f, &fb, 4);
    f = f - 1.0f;
    float u = f * (hi - lo) + lo;
    if (u < lo) u = lo;
    nrm[i] = 1.41421356237309504880f * host_erf_inv(u);
  }
  float n = sqrtf(nrm[0] * nrm[0] + nrm[1] * nrm[1]);
  out[0] = nrm[0] / n;
  out[1] = nrm[1] / n;
}

}  // namespace

extern "C" {

int cotix_lunar_default_params(CotixLunarParams* p) {
  CXL_REQUIRE(p != nullptr, "params is null");
  memset(p, 0, sizeof(*p));
  p->dt = 1e-2f;
  p->gravity_dv[0] = 0.0f;
  p->gravity_dv[1] = -0.002f;
  host_init_dir(p->init_dir);
  p->baumgarte = 0.3f;
  p->baumgarte_dt = 0.01f;
  p->coin_p = 0.5f;
  p->spring_kp = 1.0f;
  p->spring_kv0 = 0.1f;
  p->spring_kv = 0.05f;
  p->leg_damping = 0.95f;
  const float inf = std::numeric_limits<float>::infinity();
  const float props[4][4] = {{30.0f, 30.0f, 1.0f, 0.1f},
                             {1.0f, 1.0f, 1.0f, 0.1f},
                             {1.0f, 1.0f, 1.0f, 0.1f},
                             {inf, inf, 0.1f, 0.1f}};
  memcpy(p->body_props, props, sizeof(props));
  // lander: Polygon6(LANDER_POLY * 0.05), cotix/_lunar_lander.py:11-18,30
  const int LP[6][2] = {{-14, 17}, {-17, 0}, {-17, -10}, {17, -10}, {17, 0}, {14, 17}};
  for (int k = 0; k < 6; ++k) {
    p->lander_local[k][0] = (float)LP[k][0] * 0.05f;
    p->lander_local[k][1] = (float)LP[k][1] * 0.05f;
  }
  host_order_clockwise(p->lander_local, 6);
  // legs: :32-73
  float rect[4][2] = {{-2, -8}, {2, -8}, {2, 8}, {-2, 8}};
  host_order_clockwise(rect, 4);
  float s, c;
  cx_sincosf(-0.3f, &s, &c);
  for (int k = 0; k < 4; ++k) {
    float x = rect[k][0] * c + rect[k][1] * s;
    float y = rect[k][0] * (-s) + rect[k][1] * c;
    x = x * 0.05f;
    y = y * 0.05f;
    p->leg_local[1][k][0] = x;
    p->leg_local[1][k][1] = y;
    p->leg_local[0][k][0] = x * -1.0f;
    p->leg_local[0][k][1] = y * 1.0f;
  }
  const float LEG_AWAY = 24, LEG_DOWN = 8;
  p->joint_local[0][0] = LEG_AWAY * 0.05f;
  p->joint_local[0][1] = -LEG_DOWN * 0.05f;
  p->joint_local[1][0] = LEG_AWAY * 0.05f;
  p->joint_local[1][1] = (-LEG_DOWN + 8) * 0.05f;
  p->joint_local[2][0] = -LEG_AWAY * 0.05f;
  p->joint_local[2][1] = -LEG_DOWN * 0.05f;
  p->joint_local[3][0] = -LEG_AWAY * 0.05f;
  p->joint_local[3][1] = (-LEG_DOWN + 8) * 0.05f;
  p->leg_joint2_local[0] = 0.0f;
  p->leg_joint2_local[1] = 0.4f;
  p->lander_start[0] = 0.0f;
  p->lander_start[1] = 5.0f;
  p->lander_start[2] = 0.01f;
  p->leg_offset[0][0] = -LEG_AWAY * 0.05f;
  p->leg_offset[0][1] = -LEG_DOWN * 0.05f;
  p->leg_offset[1][0] = LEG_AWAY * 0.05f;
  p->leg_offset[1][1] = -LEG_DOWN * 0.05f;
  p->ground_floor = -10.0f;
  p->terrain_x_outer = 100.0f;
  p->pad_half_width = 2.0f;
  p->pad_y = -2.0f;
  p->gjk_max_iters = 32;
  p->rng_mode = COTIX_RNG_THREEFRY;
  p->substeps = 1;
  p->auto_reset = 0;
  p->judge_enabled = 0;
  p->max_episode_steps = 0;
  p->reset_perturb = 0;
  p->main_engine_dv = 0.004f;
  p->side_engine_dv = 0.0005f;
  p->side_engine_dw = 0.002f;
  p->x_limit = 12.0f;
  p->y_limit = 20.0f;
  p->land_speed = 0.05f;
  p->land_angvel = 0.05f;
  p->reset_x[0] = -6.0f;
  p->reset_x[1] = 6.0f;
  p->reset_y[0] = -1.5f;
  p->reset_y[1] = 8.0f;
  p->reset_angle[0] = -0.5f;
  p->reset_angle[1] = 0.5f;
  p->reset_v = 1.7320508f;
  p->reset_w = 0.8660254f;
  p->philox_seed[0] = 0x243F6A88u;
  p->philox_seed[1] = 0x85A308D3u;
  p->solver_loops = 1;
  return COTIX_OK;
}

static int check_params(const CotixLunarParams* p) {
  if (!p) return 0;
  if (p->rng_mode != COTIX_RNG_THREEFRY && p->rng_mode != COTIX_RNG_PHILOX) return 0;
  if (p->substeps < 1 || p->substeps > 1024) return 0;
  if (p->gjk_max_iters < 0) return 0;
  if (p->solver_loops < 1 || p->solver_loops > 64) return 0;
  return 1;
}

int cotix_lunar_reset(void* stream, int64_t B, const uint32_t* keys, float* state, int64_t ld,
                      const CotixLunarParams* p) {
  CXL_REQUIRE(B >= 0 && ld >= B, "bad batch size / leading dimension");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(B == 0 || (keys && state), "null buffer");
  return cx_launch_lunar_reset((cudaStream_t)stream, B, keys, state, ld, p);
}

int cotix_lunar_step(void* stream, int64_t B, const float* state_in, float* state_out, int64_t ld,
                     const float* action, float* reward, uint8_t* done, float* obs,
                     const CotixLunarParams* p) {
  CXL_REQUIRE(B >= 0 && ld >= B, "bad batch size / leading dimension");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(B == 0 || (state_in && state_out), "null state");
  return cx_launch_lunar_step((cudaStream_t)stream, B, state_in, state_out, ld, action, reward, done,
                              obs, p);
}

int cotix_lunar_pipeline_profile(int enable) { return cx_lunar_pipeline_profile(enable); }
int64_t cotix_lunar_pipeline_stage_ms(double* out_ms) {
  if (!out_ms) return 0;
  return cx_lunar_pipeline_stage_ms(out_ms);
}

int64_t cotix_lunar_workspace_bytes(int64_t B) { return cx_lunar_pipeline_workspace_bytes(B); }

int cotix_lunar_step_ws(void* stream, int64_t B, const float* state_in, float* state_out, int64_t ld,
                        const float* action, float* reward, uint8_t* done, float* obs,
                        const CotixLunarParams* p, void* workspace, int64_t workspace_bytes) {
  CXL_REQUIRE(B >= 0 && ld >= B, "bad batch size / leading dimension");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(B == 0 || (state_in && state_out), "null state");
  return cx_launch_lunar_pipeline((cudaStream_t)stream, B, state_in, state_out, ld, action, reward, done,
                                  obs, p, workspace, workspace_bytes);
}

// Multi-step calls without a workspace: up to COTIX_ROLLOUT_MAX_ENVS envs run all their steps inside one launch of
// the block-resident rollout kernel (csrc/lunar_rollout.cu; K = 1 sub-step per step only), larger batches one
// fused launch per step.  COTIX_ROLLOUT_KERNEL=0 keeps the per-step launches (A/B measurements, tests).
static bool use_rollout_kernel(int64_t B, int64_t n_steps, const CotixLunarParams* p) {
  static const char* env = getenv("COTIX_ROLLOUT_KERNEL");
  if (env && env[0] == '0') return false;
  return n_steps > 1 && p->substeps <= 1 && B <= COTIX_ROLLOUT_MAX_ENVS;
}

int cotix_lunar_steps_ws(void* stream, int64_t B, float* state, int64_t ld, const float* actions,
                         int32_t n_actions, int32_t n_steps, float* reward, uint8_t* done, float* obs,
                         const CotixLunarParams* p, void* workspace, int64_t workspace_bytes) {
  CXL_REQUIRE(B >= 0 && ld >= B && n_steps >= 0, "bad batch size / leading dimension / step count");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(B == 0 || state, "null state");
  CXL_REQUIRE(!actions || n_actions >= 1, "n_actions must be >= 1 when actions are given");
  if (workspace)
    return cx_launch_lunar_pipeline_steps((cudaStream_t)stream, B, state, state, 0, ld, actions, 2 * ld,
                                          n_actions, n_steps, reward, done, obs, p, workspace, workspace_bytes);
  if (use_rollout_kernel(B, n_steps, p))  // small batch: all steps in one launch, envs resident in their block
    return cx_launch_lunar_rollout_kernel((cudaStream_t)stream, B, state, state, 0, ld, actions, 2 * ld, n_actions,
                                          n_steps, reward, done, obs, p);
  for (int32_t t = 0; t < n_steps; ++t) {
    const float* a = actions ? actions + (int64_t)(t % n_actions) * 2 * ld : nullptr;
    const int rc = cx_launch_lunar_step((cudaStream_t)stream, B, state, state, ld, a, reward, done, obs, p);
    if (rc != 0) return rc;
  }
  return COTIX_OK;
}

int cotix_lunar_pipeline_config(int split, int pdl, int priorities, int64_t item_capacity) {
  const int rc = cx_lunar_pipeline_config(split, pdl, priorities, item_capacity);
  if (rc != 0) cx_set_error("cotix_lunar_pipeline_config: split must be <= %d", COTIX_MAX_CHUNKS);
  return rc;
}
int cotix_lunar_pipeline_counters(void* stream, int64_t B, const void* workspace, int64_t* out) {
  CXL_REQUIRE(B > 0 && workspace && out, "null argument");
  return cx_lunar_pipeline_counters((cudaStream_t)stream, B, workspace, out);
}

// ---- CUDA-graph capture of any sequence of device entry points ----------------------------------
struct CotixGraph {
  cudaGraphExec_t exec;
  int64_t launches;  // kernels per replay
  int nodes;
};
namespace {
thread_local int64_t g_capture_base = -1;
}
int cotix_graph_begin_capture(void* stream) {
  CXL_REQUIRE(g_capture_base < 0, "a capture is already open on this thread");
  CXL_CUDA(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeThreadLocal));
  g_capture_base = cotix_launch_count();
  return COTIX_OK;
}
int cotix_graph_end_capture(void* stream, CotixGraph** out) {
  CXL_REQUIRE(g_capture_base >= 0, "no capture open on this thread");
  cudaGraph_t g = nullptr;
  const cudaError_t ce = cudaStreamEndCapture((cudaStream_t)stream, &g);
  const int64_t captured = cotix_launch_count() - g_capture_base;
  g_capture_base = -1;
  cx_count_launch_n(-captured);  // captured, not executed
  CXL_CUDA(ce);
  CXL_REQUIRE(out != nullptr, "out is null");
  CotixGraph* cg = new (std::nothrow) CotixGraph();
  if (!cg) {
    cudaGraphDestroy(g);
    CXL_REQUIRE(false, "out of host memory");
  }
  size_t nn = 0;
  cudaGraphGetNodes(g, nullptr, &nn);
  cg->nodes = (int)nn;
  cg->launches = captured;
  const cudaError_t ie = cudaGraphInstantiate(&cg->exec, g, 0);
  cudaGraphDestroy(g);
  if (ie != cudaSuccess) {
    delete cg;
    CXL_CUDA(ie);
  }
  *out = cg;
  return COTIX_OK;
}
int cotix_graph_launch(CotixGraph* g, void* stream) {
  CXL_REQUIRE(g != nullptr, "graph is null");
  CXL_CUDA(cudaGraphLaunch(g->exec, (cudaStream_t)stream));
  cx_count_launch_n(g->launches);
  return COTIX_OK;
}
int cotix_graph_info(const CotixGraph* g, int32_t* nodes, int64_t* kernel_launches) {
  CXL_REQUIRE(g != nullptr, "graph is null");
  if (nodes) *nodes = g->nodes;
  if (kernel_launches) *kernel_launches = g->launches;
  return COTIX_OK;
}
int cotix_graph_destroy(CotixGraph* g) {
  if (!g) return COTIX_OK;
  cudaGraphExecDestroy(g->exec);
  delete g;
  return COTIX_OK;
}

// ---- differentiable rollout (K5) ----------------------------------------------------
int cotix_lunar_step_bwd(void* stream, int64_t B, const float* state_in, const float* state_next,
                         int64_t ld, const float* action, const float* g_out, float* g_in,
                         float* g_action, uint32_t* flags, const CotixLunarParams* p) {
  CXL_REQUIRE(B >= 0 && ld >= B, "bad batch size / leading dimension");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(p->substeps <= 1 && p->auto_reset == 0 && p->solver_loops == 1,
              "the adjoint needs substeps == 1, auto_reset == 0 and solver_loops == 1");
  CXL_REQUIRE(B == 0 || (state_in && g_out && g_in), "null state / cotangent");
  return cx_launch_lunar_step_bwd((cudaStream_t)stream, B, state_in, state_next, ld, action, g_out, g_in,
                                  g_action, flags, nullptr, nullptr, p);
}

int cotix_lunar_rollout(void* stream, int64_t B, int32_t T, float* tape, int64_t ld, const float* actions,
                        const CotixLunarParams* p, void* workspace, int64_t workspace_bytes) {
  CXL_REQUIRE(B >= 0 && ld >= B && T >= 0, "bad batch size / leading dimension / step count");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(B == 0 || tape, "null tape");
  const int64_t slot = (int64_t)COTIX_LUNAR_NF * ld;
  if (workspace)  // the ranges of the pipeline run their T steps without joining
    return cx_launch_lunar_pipeline_steps((cudaStream_t)stream, B, tape, tape + slot, slot, ld, actions, 2 * ld,
                                          T > 0 ? T : 1, T, nullptr, nullptr, nullptr, p, workspace,
                                          workspace_bytes);
  if (use_rollout_kernel(B, T, p))
    return cx_launch_lunar_rollout_kernel((cudaStream_t)stream, B, tape, tape + slot, slot, ld, actions, 2 * ld, T, T,
                                          nullptr, nullptr, nullptr, p);
  for (int32_t t = 0; t < T; ++t) {
    const float* a = actions ? actions + (int64_t)t * 2 * ld : nullptr;
    const int rc = cx_launch_lunar_step((cudaStream_t)stream, B, tape + t * slot, tape + (t + 1) * slot, ld, a,
                                        nullptr, nullptr, nullptr, p);
    if (rc != 0) return rc;
  }
  return 0;
}

// decision tape: 33 words per env-step; Jacobian area: room for one env-step in eight to have
// contacts (slot list + 20 x 18 matrix each) + a 256-byte header
static int64_t adj_dec_bytes(int64_t ld) { return (int64_t)COTIX_ADJ_TAPE_WORDS * ld * (int64_t)sizeof(float); }
static int64_t adj_jac_cap(int64_t ld) { return ld / 8 + 1; }
static int64_t adj_jac_bytes(int64_t ld) { return adj_jac_cap(ld) * (4 + 20 * COTIX_LUNAR_NDYN * (int64_t)sizeof(float)); }
int64_t cotix_lunar_adjoint_decisions_bytes(int64_t ld, int32_t T) {
  if (ld < 0 || T < 0) return 0;
  return (int64_t)T * adj_dec_bytes(ld);
}
int64_t cotix_lunar_adjoint_workspace_bytes(int64_t ld, int32_t T) {
  if (ld < 0 || T < 0) return 0;
  return 512 + (int64_t)T * (adj_dec_bytes(ld) + adj_jac_bytes(ld));  // header + alignment of the header
}

// Side stream of the pipelined backward pass (one per device, made on first use): the decisions and
// Jacobians of the earlier steps are formed there while the sweep of the later steps runs on the
// caller's stream.  Forked from and joined back into the caller's stream with events, so the call
// stays stream-ordered (and capturable).
struct AdjSide {
  cudaStream_t s;
  cudaEvent_t fork, done[8];
  int state;  // 0 not made, 1 ok, -1 failed
  std::mutex mu;  // held from fork to join: the stream and events are shared by all callers on the device
};
static AdjSide* adj_side() {
  static AdjSide per_dev[64];
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  AdjSide& a = per_dev[dev];
  if (a.state == 0) {
    bool ok = cudaStreamCreateWithFlags(&a.s, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < 8 && ok; ++i) ok = cudaEventCreateWithFlags(&a.done[i], cudaEventDisableTiming) == cudaSuccess;
    a.state = ok ? 1 : -1;
    if (!ok) cudaGetLastError();
  }
  return a.state == 1 ? &a : nullptr;
}
// the reverse sweep of a run of steps as ONE launch (lunar_sweep_bwd_kernel) up to this many envs;
// COTIX_SWEEP_KERNEL=0 keeps one launch per step (A/B measurements, tests)
static bool use_sweep_kernel(int64_t B) {
  static const char* env = getenv("COTIX_SWEEP_KERNEL");
  if (env && env[0] == '0') return false;
  return B <= 16384;
}
// 0: off; n > 0: on, the overlapped launches capped at n blocks per SM; n < 0: on, uncapped
static int g_adj_pipelined = -1;
void cx_lunar_adjoint_set_pipelined(int n) { g_adj_pipelined = n; }

int cotix_lunar_rollout_bwd(void* stream, int64_t B, int32_t T, const float* tape, int64_t ld,
                            const float* actions, float* g, float* g_actions, uint32_t* flags,
                            const CotixLunarParams* p, void* workspace, int64_t workspace_bytes,
                            int32_t jacobians) {
  CXL_REQUIRE(B >= 0 && ld >= B && T >= 0, "bad batch size / leading dimension / step count");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(p->substeps <= 1 && p->auto_reset == 0 && p->solver_loops == 1,
              "the adjoint needs substeps == 1, auto_reset == 0 and solver_loops == 1");
  CXL_REQUIRE(B == 0 || (tape && g), "null tape / cotangent");
  const int64_t slot = (int64_t)COTIX_LUNAR_NF * ld;
  const int64_t dslot = (int64_t)COTIX_ADJ_TAPE_WORDS * ld;
  int64_t group = 0;  // steps whose decisions (and Jacobians) fit the workspace at once
  bool jac = false;
  if (workspace && B > 0) {
    CXL_REQUIRE(((uintptr_t)workspace & 255u) == 0, "workspace must be 256-byte aligned");
    const int64_t full = adj_dec_bytes(ld) + adj_jac_bytes(ld);
    if (jacobians) {
      jac = true;
      group = (workspace_bytes - 512) / full;
      CXL_REQUIRE(workspace_bytes >= 512 + full, "workspace smaller than cotix_lunar_adjoint_workspace_bytes(ld, 1)");
    } else {
      group = workspace_bytes / adj_dec_bytes(ld);
      CXL_REQUIRE(group >= 1, "workspace smaller than cotix_lunar_adjoint_decisions_bytes(ld, 1)");
    }
    if (group > T) group = T;
    CXL_REQUIRE(B * group < (1ll << 32), "too many (step, env) items per group");
  }
  // layout: [group decision tapes][256-byte header: counter][slot list][matrices]
  float* dt = (float*)workspace;
  uint32_t* jctr = nullptr;
  uint32_t* jlist = nullptr;
  float* jM = nullptr;
  uint32_t jcap = 0;
  if (jac) {
    char* base = (char*)workspace + group * adj_dec_bytes(ld);
    base += (256 - ((uintptr_t)base & 255u)) & 255u;
    jcap = (uint32_t)(adj_jac_cap(ld) * group);
    jctr = (uint32_t*)base;
    jlist = (uint32_t*)(base + 256);
    jM = (float*)(base + 256 + (int64_t)jcap * 4);
    // header (256) + its alignment padding (< 256) are the 512 bytes the size function adds
  }
  // The whole rollout fits: cut it into NG runs of steps.  The decisions + Jacobians of all runs are
  // enqueued on the side stream, latest run first (the first with the GPU to itself, the others with
  // capped grids); the sweep of a run starts as soon as its own are done and overlaps the next one's.
  constexpr int NG = 4;
  static const char* env_overlap = getenv("COTIX_ADJ_OVERLAP");  // measurement aid: "0" turns the overlap off
  const bool overlap = g_adj_pipelined != 0 && !(env_overlap && env_overlap[0] == '0');
  static const char* env_runs = getenv("COTIX_ADJ_RUNS");  // measurement aid: "4" = the four overlapped runs
  if (jac && group >= T && T >= 1 && use_sweep_kernel(B) && !(env_runs && env_runs[0] == '4')) {
    // small batch: decisions + Jacobians of the whole rollout at full grid, then the sweep in ONE launch on the
    // caller's stream.  Measured (4096 envs x 100 steps): 3.74 ms against 3.87-4.02 with the runs overlapped on
    // the side stream -- the latency-bound sweep stretches when the Jacobian kernel shares its SMs -- and the
    // slot order of the Jacobians, hence the gradient's rounding, is the same from call to call
    int rc = cx_launch_lunar_decide((cudaStream_t)stream, B, T, tape, ld, actions, dt, flags, jctr, jlist, jM, jcap, p, 0);
    if (rc == 0)
      rc = cx_launch_lunar_sweep_bwd((cudaStream_t)stream, B, 0, T, tape, ld, actions, g, g_actions, flags, dt, jM, p);
    return rc;
  }
  AdjSide* side = (jac && overlap && group >= T && T >= 2 * NG) ? adj_side() : nullptr;
  if (side) {
    std::lock_guard<std::mutex> lock(side->mu);
    cudaStream_t st = (cudaStream_t)stream;
    const int32_t ns = (T + NG - 1) / NG;
    int32_t b[NG + 1];
    for (int i = 0; i <= NG; ++i) b[i] = (int32_t)((int64_t)i * ns < T ? (int64_t)i * ns : T);
    if (cudaEventRecord(side->fork, st) != cudaSuccess || cudaStreamWaitEvent(side->s, side->fork, 0) != cudaSuccess)
      return cx_check_launch("adjoint fork");
    const int64_t cap1 = adj_jac_cap(ld);
    int rc = 0;  // after the fork every path runs to the joins (an early return would leave a capture unjoined)
    for (int i = NG - 1; i >= 0; --i) {
      const int32_t nt = b[i + 1] - b[i];
      if (nt > 0 && rc == 0) {
        const int64_t off = cap1 * b[i];
        rc = cx_launch_lunar_decide(side->s, B, nt, tape + b[i] * slot, ld,
                                    actions ? actions + (int64_t)b[i] * 2 * ld : nullptr, dt + b[i] * dslot, flags,
                                    jctr + i, jlist + off, jM + off * 20 * COTIX_LUNAR_NDYN, (uint32_t)(cap1 * nt), p,
                                    (i == NG - 1 || g_adj_pipelined < 0) ? 0 : g_adj_pipelined);
      }
      if (cudaEventRecord(side->done[i], side->s) != cudaSuccess && rc == 0) rc = cx_check_launch("adjoint event");
    }
    for (int i = NG - 1; i >= 0; --i) {
      if (cudaStreamWaitEvent(st, side->done[i], 0) != cudaSuccess && rc == 0) rc = cx_check_launch("adjoint join");
      const int64_t off = cap1 * b[i];
      if (use_sweep_kernel(B) && rc == 0) {  // small batch: the steps of the run in one launch
        rc = cx_launch_lunar_sweep_bwd(st, B, b[i], b[i + 1], tape, ld, actions, g, g_actions, flags, dt,
                                       jM + off * 20 * COTIX_LUNAR_NDYN, p);
        continue;
      }
      for (int32_t s = b[i + 1] - 1; s >= b[i] && rc == 0; --s) {
        const float* a = actions ? actions + (int64_t)s * 2 * ld : nullptr;
        float* ga = g_actions ? g_actions + (int64_t)s * 2 * ld : nullptr;
        rc = cx_launch_lunar_step_bwd(st, B, tape + s * slot, tape + (s + 1) * slot, ld, a, g, g, ga, flags,
                                      dt + s * dslot, jM + off * 20 * COTIX_LUNAR_NDYN, p);
      }
    }
    return rc;
  }
  int32_t t = T;
  while (t > 0) {
    const int32_t nt = group > 0 ? (int32_t)(t < group ? t : group) : 1;
    const int32_t t0 = t - nt;
    if (group > 0) {
      const int rc = cx_launch_lunar_decide((cudaStream_t)stream, B, nt, tape + t0 * slot, ld,
                                            actions ? actions + (int64_t)t0 * 2 * ld : nullptr, dt, flags, jctr,
                                            jlist, jM, jcap, p, 0);
      if (rc != 0) return rc;
    }
    for (int32_t s = t - 1; s >= t0; --s) {
      const float* a = actions ? actions + (int64_t)s * 2 * ld : nullptr;
      float* ga = g_actions ? g_actions + (int64_t)s * 2 * ld : nullptr;
      const int rc = cx_launch_lunar_step_bwd((cudaStream_t)stream, B, tape + s * slot, tape + (s + 1) * slot, ld,
                                              a, g, g, ga, flags, group > 0 ? dt + (s - t0) * dslot : nullptr, jM, p);
      if (rc != 0) return rc;
    }
    t = t0;
  }
  return 0;
}

// ---- headless rasteriser (K6) -------------------------------------------------------
int cotix_raster_draw(void* stream, int64_t n_img, const int32_t* prims, const int32_t* counts,
                      int32_t max_prims, uint8_t* image) {
  CXL_REQUIRE(n_img >= 0 && max_prims >= 0 && max_prims <= 65536, "bad image / primitive count");
  CXL_REQUIRE(n_img == 0 || image, "null image");
  CXL_REQUIRE(n_img == 0 || max_prims == 0 || prims, "null primitive list");
  CXL_REQUIRE(((uintptr_t)image & 15u) == 0 && ((uintptr_t)prims & 15u) == 0, "image and prims must be 16-byte aligned");
  return cx_launch_raster_draw((cudaStream_t)stream, n_img, prims, counts, max_prims, image);
}
int cotix_lunar_draw(void* stream, int64_t n_img, const float* state, int64_t ld, const int64_t* env_index,
                     uint8_t* image, const CotixLunarParams* p) {
  CXL_REQUIRE(n_img >= 0 && ld >= 0, "bad image count / leading dimension");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(n_img == 0 || (state && image), "null state / image");
  CXL_REQUIRE(env_index || n_img <= ld, "more images than envs");
  CXL_REQUIRE(((uintptr_t)image & 15u) == 0, "image must be 16-byte aligned");
  return cx_launch_lunar_draw((cudaStream_t)stream, n_img, state, ld, env_index, image, p);
}
int cotix_lunar_prims(void* stream, int64_t n_img, const float* state, int64_t ld, const int64_t* env_index,
                      int32_t* prims, const CotixLunarParams* p) {
  CXL_REQUIRE(n_img >= 0 && ld >= 0, "bad image count / leading dimension");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_REQUIRE(n_img == 0 || (state && prims), "null state / primitive list");
  CXL_REQUIRE(env_index || n_img <= ld, "more images than envs");
  return cx_launch_lunar_prims((cudaStream_t)stream, n_img, state, ld, env_index, prims, p);
}

// ---- host-facing env handle ---------------------------------------------------------
struct HostGraph {  // one captured host step, keyed by the host buffers it copies from / to
  const float* action;
  float* obs;
  float* reward;
  uint8_t* done;
  int chunks, mode;
  cudaGraphExec_t exec;
  int64_t launches;
};

struct CotixLunarEnv;
static void drop_host_graphs(CotixLunarEnv* e);

struct CotixLunarEnv {
  int mode;
  void* ws;
  int64_t ws_bytes;
  int device;
  int64_t B;
  CotixLunarParams P;
  cudaStream_t stream;
  float* state;     // [NF][B]
  float* action;    // [2][B]
  float* obs;       // [8][B]
  float* reward;    // [B]
  uint8_t* done;    // [B]
  uint32_t* keys;   // [B][2]
  std::map<int, cudaGraphExec_t>* graphs;
  const float* graph_action;
  int64_t graph_launches;  // kernels per step in the captured graphs
  int chunks;              // host-step ranges; 0 = automatic
  int host_graphs;         // replay the pipelined host step as a CUDA graph when the buffers are pinned
  cudaStream_t s_in, s_out;
  cudaEvent_t ev_in[COTIX_MAX_CHUNKS], ev_done[COTIX_MAX_CHUNKS], ev_fork, ev_join[2];
  std::vector<HostGraph>* hgraphs;
};

int cotix_lunar_env_create(CotixLunarEnv** out, int device, int64_t B, const CotixLunarParams* p) {
  CXL_REQUIRE(out != nullptr, "out is null");
  CXL_REQUIRE(B > 0, "B <= 0");
  CXL_REQUIRE(check_params(p), "bad params");
  CXL_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  CXL_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    cx_set_error("cotix_lunar_env_create: device %d is sm_%d%d; this library is sm_100a only", device,
                 prop.major, prop.minor);
    return COTIX_ENODEV;
  }
  CotixLunarEnv* e = new (std::nothrow) CotixLunarEnv();
  CXL_REQUIRE(e != nullptr, "out of host memory");
  memset(e, 0, sizeof(*e));
  e->device = device;
  e->B = B;
  e->P = *p;
  e->graphs = new std::map<int, cudaGraphExec_t>();
  CXL_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  CXL_CUDA(cudaStreamCreateWithFlags(&e->s_in, cudaStreamNonBlocking));
  CXL_CUDA(cudaStreamCreateWithFlags(&e->s_out, cudaStreamNonBlocking));
  for (int c = 0; c < COTIX_MAX_CHUNKS; ++c) {
    CXL_CUDA(cudaEventCreateWithFlags(&e->ev_in[c], cudaEventDisableTiming));
    CXL_CUDA(cudaEventCreateWithFlags(&e->ev_done[c], cudaEventDisableTiming));
  }
  CXL_CUDA(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
  CXL_CUDA(cudaEventCreateWithFlags(&e->ev_join[0], cudaEventDisableTiming));
  CXL_CUDA(cudaEventCreateWithFlags(&e->ev_join[1], cudaEventDisableTiming));
  e->hgraphs = new std::vector<HostGraph>();
  e->host_graphs = 0;
  CXL_CUDA(cudaMalloc(&e->state, sizeof(float) * COTIX_LUNAR_NF * (size_t)B));
  CXL_CUDA(cudaMalloc(&e->action, sizeof(float) * 2 * (size_t)B));
  CXL_CUDA(cudaMalloc(&e->obs, sizeof(float) * COTIX_LUNAR_NOBS * (size_t)B));
  CXL_CUDA(cudaMalloc(&e->reward, sizeof(float) * (size_t)B));
  CXL_CUDA(cudaMalloc(&e->done, (size_t)B));
  CXL_CUDA(cudaMalloc(&e->keys, sizeof(uint32_t) * 2 * (size_t)B));
  e->mode = COTIX_MODE_AUTO;
  e->ws_bytes = cx_lunar_pipeline_workspace_bytes(B);
  CXL_CUDA(cudaMalloc(&e->ws, (size_t)e->ws_bytes));
  CXL_CUDA(cudaMemsetAsync(e->state, 0, sizeof(float) * COTIX_LUNAR_NF * (size_t)B, e->stream));
  CXL_CUDA(cudaStreamSynchronize(e->stream));
  *out = e;
  return COTIX_OK;
}

int cotix_lunar_env_destroy(CotixLunarEnv* e) {
  if (!e) return COTIX_OK;
  cudaSetDevice(e->device);
  cudaStreamSynchronize(e->stream);
  if (e->graphs) {
    for (auto& kv : *e->graphs) cudaGraphExecDestroy(kv.second);
    delete e->graphs;
  }
  cudaFree(e->state);
  cudaFree(e->action);
  cudaFree(e->obs);
  cudaFree(e->reward);
  cudaFree(e->done);
  cudaFree(e->keys);
  cudaFree(e->ws);
  for (int c = 0; c < COTIX_MAX_CHUNKS; ++c) {
    cudaEventDestroy(e->ev_in[c]);
    cudaEventDestroy(e->ev_done[c]);
  }
  cudaEventDestroy(e->ev_fork);
  cudaEventDestroy(e->ev_join[0]);
  cudaEventDestroy(e->ev_join[1]);
  drop_host_graphs(e);
  delete e->hgraphs;
  cudaStreamDestroy(e->s_in);
  cudaStreamDestroy(e->s_out);
  cudaStreamDestroy(e->stream);
  delete e;
  return COTIX_OK;
}

int cotix_lunar_env_set_mode(CotixLunarEnv* e, int mode) {
  CXL_REQUIRE(e != nullptr, "env is null");
  CXL_REQUIRE(mode == COTIX_MODE_AUTO || mode == COTIX_MODE_FUSED || mode == COTIX_MODE_PIPELINE, "bad mode");
  if (e->mode != mode && e->graphs) {  // captured graphs embed the launch strategy
    for (auto& kv : *e->graphs) cudaGraphExecDestroy(kv.second);
    e->graphs->clear();
  }
  e->mode = mode;
  return COTIX_OK;
}

int cotix_lunar_env_set_chunks(CotixLunarEnv* e, int chunks) {
  CXL_REQUIRE(e != nullptr, "env is null");
  CXL_REQUIRE(chunks >= 0 && chunks <= COTIX_MAX_CHUNKS, "chunks out of range");
  if (chunks != e->chunks) drop_host_graphs(e);
  e->chunks = chunks;
  return COTIX_OK;
}

// one step of envs [off, off + n) of the resident state on the handle's stream, by the handle's
// launch strategy (the strategy is chosen on the whole batch so that it does not vary with the
// chunking)
static int env_enqueue_range(CotixLunarEnv* e, int64_t off, int64_t n, const float* action_dev,
                             float* obs_dev) {
  cudaStream_t stream = e->stream;
  const bool pipe = e->mode == COTIX_MODE_PIPELINE ||
                    (e->mode == COTIX_MODE_AUTO && e->B >= 16384);
  float* st = e->state + off;
  const float* act = action_dev ? action_dev + off : nullptr;
  float* ob = obs_dev ? obs_dev + off : nullptr;
  if (pipe)
    return cx_launch_lunar_pipeline(stream, n, st, st, e->B, act, e->reward + off, e->done + off, ob,
                                    &e->P, e->ws, e->ws_bytes);
  return cx_launch_lunar_step(stream, n, st, st, e->B, act, e->reward + off, e->done + off, ob, &e->P);
}
static int env_enqueue_step(CotixLunarEnv* e, const float* action_dev, float* obs_dev) {
  return env_enqueue_range(e, 0, e->B, action_dev, obs_dev);
}
// number of env ranges of a host step and their length (a multiple of 32 envs = 128 bytes)
static int plan_chunks(const CotixLunarEnv* e, int64_t* len) {
  int64_t C = e->chunks;
  if (C == 0) {
    C = e->B / 262144;
    if (C > 4) C = 4;
  }
  if (C < 1) C = 1;
  int64_t L = (e->B + C - 1) / C;
  L = (L + 31) / 32 * 32;
  *len = L;
  return (int)((e->B + L - 1) / L);
}

int cotix_lunar_env_reset_host(CotixLunarEnv* e, const uint32_t* keys) {
  CXL_REQUIRE(e && keys, "null argument");
  CXL_CUDA(cudaSetDevice(e->device));
  CXL_CUDA(cudaMemcpyAsync(e->keys, keys, sizeof(uint32_t) * 2 * (size_t)e->B, cudaMemcpyHostToDevice,
                           e->stream));
  int rc = cx_launch_lunar_reset(e->stream, e->B, e->keys, e->state, e->B, &e->P);
  if (rc) return rc;
  CXL_CUDA(cudaStreamSynchronize(e->stream));
  return COTIX_OK;
}

// enqueue one host step as a software pipeline over C env ranges: s_in copies actions in, the
// handle's stream computes, s_out copies the results out; [rows][B] host arrays are copied as
// 2-D blocks of pitch B.  With `forkjoin` the side streams are forked from and joined back into
// the handle's stream (what stream capture needs).
static int enqueue_chunked(CotixLunarEnv* e, int C, int64_t len, const float* action, float* obs,
                           float* reward, uint8_t* done, bool forkjoin) {
  const size_t pitch = sizeof(float) * (size_t)e->B;
  if (forkjoin) {
    CXL_CUDA(cudaEventRecord(e->ev_fork, e->stream));
    CXL_CUDA(cudaStreamWaitEvent(e->s_in, e->ev_fork, 0));
    CXL_CUDA(cudaStreamWaitEvent(e->s_out, e->ev_fork, 0));
  }
  for (int c = 0; c < C; ++c) {
    const int64_t off = (int64_t)c * len;
    const int64_t n = (off + len <= e->B) ? len : e->B - off;
    if (action) {
      CXL_CUDA(cudaMemcpy2DAsync(e->action + off, pitch, action + off, pitch, sizeof(float) * (size_t)n, 2,
                                 cudaMemcpyHostToDevice, e->s_in));
      CXL_CUDA(cudaEventRecord(e->ev_in[c], e->s_in));
      CXL_CUDA(cudaStreamWaitEvent(e->stream, e->ev_in[c], 0));
    }
    int rc = env_enqueue_range(e, off, n, action ? e->action : nullptr, obs ? e->obs : nullptr);
    if (rc) return rc;
    CXL_CUDA(cudaEventRecord(e->ev_done[c], e->stream));
    CXL_CUDA(cudaStreamWaitEvent(e->s_out, e->ev_done[c], 0));
    if (obs)
      CXL_CUDA(cudaMemcpy2DAsync(obs + off, pitch, e->obs + off, pitch, sizeof(float) * (size_t)n,
                                 COTIX_LUNAR_NOBS, cudaMemcpyDeviceToHost, e->s_out));
    if (reward)
      CXL_CUDA(cudaMemcpyAsync(reward + off, e->reward + off, sizeof(float) * (size_t)n,
                               cudaMemcpyDeviceToHost, e->s_out));
    if (done)
      CXL_CUDA(cudaMemcpyAsync(done + off, e->done + off, (size_t)n, cudaMemcpyDeviceToHost, e->s_out));
  }
  if (forkjoin) {
    CXL_CUDA(cudaEventRecord(e->ev_join[0], e->s_in));
    CXL_CUDA(cudaEventRecord(e->ev_join[1], e->s_out));
    CXL_CUDA(cudaStreamWaitEvent(e->stream, e->ev_join[0], 0));
    CXL_CUDA(cudaStreamWaitEvent(e->stream, e->ev_join[1], 0));
  }
  return COTIX_OK;
}

// page-locked host memory (or NULL): what a captured copy node needs
static bool host_pinned(const void* p) {
  if (!p) return true;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

static void drop_host_graphs(CotixLunarEnv* e) {
  if (!e->hgraphs) return;
  for (auto& g : *e->hgraphs) cudaGraphExecDestroy(g.exec);
  e->hgraphs->clear();
}

// pipeline mode: the ranges of the wavefront pipeline ARE the chunks of the host step — H2D(action) of
// a range, its kernels and D2H(obs, reward, done) are enqueued on the range's own stream, so copies in
// both directions overlap the kernels of the other ranges (and the sparse stages of one range run
// under the dense stages of another); [rows][B] host arrays are copied as 2-D blocks of pitch B
static int enqueue_ranges(CotixLunarEnv* e, int split, const float* action, float* obs, float* reward,
                          uint8_t* done) {
  const size_t pitch = sizeof(float) * (size_t)e->B;
  // The kernels of range r wait for the kernels of range r - 1 (an event between the hooks): compute runs
  // range after range, so range 0 is done after a quarter of the compute time and its D2H copy -- the
  // bottleneck of the step -- runs under the kernels of the ranges behind it.  Left to run concurrently the
  // ranges all finish together at the end and the copies start late (1.00 -> 0.8x ms per step).
  static const char* env_serial = getenv("COTIX_E2E_SERIAL");  // measurement aid: "0" = concurrent ranges
  const bool serial = !(env_serial && env_serial[0] == '0');
  CxRangeHook pre = [&](cudaStream_t cs, int r, int64_t off, int64_t n) -> int {
    if (action)
      CXL_CUDA(cudaMemcpy2DAsync(e->action + off, pitch, action + off, pitch, sizeof(float) * (size_t)n, 2,
                                 cudaMemcpyHostToDevice, cs));
    if (serial && r > 0) CXL_CUDA(cudaStreamWaitEvent(cs, e->ev_done[r - 1], 0));
    return COTIX_OK;
  };
  CxRangeHook post = [&](cudaStream_t cs, int r, int64_t off, int64_t n) -> int {
    if (serial) CXL_CUDA(cudaEventRecord(e->ev_done[r], cs));
    if (obs)
      CXL_CUDA(cudaMemcpy2DAsync(obs + off, pitch, e->obs + off, pitch, sizeof(float) * (size_t)n,
                                 COTIX_LUNAR_NOBS, cudaMemcpyDeviceToHost, cs));
    if (reward)
      CXL_CUDA(cudaMemcpyAsync(reward + off, e->reward + off, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, cs));
    if (done) CXL_CUDA(cudaMemcpyAsync(done + off, e->done + off, (size_t)n, cudaMemcpyDeviceToHost, cs));
    return COTIX_OK;
  };
  return cx_launch_lunar_pipeline_hooked(e->stream, e->B, e->state, e->B, action ? e->action : nullptr, e->reward,
                                         e->done, obs ? e->obs : nullptr, &e->P, e->ws, e->ws_bytes, split, pre, post);
}

int cotix_lunar_env_step_host(CotixLunarEnv* e, const float* action, float* obs, float* reward,
                              uint8_t* done) {
  CXL_REQUIRE(e != nullptr, "env is null");
  CXL_CUDA(cudaSetDevice(e->device));
  const size_t B = (size_t)e->B;
  const bool pipe = e->mode == COTIX_MODE_PIPELINE || (e->mode == COTIX_MODE_AUTO && e->B >= 16384);
  int64_t len = 0;
  // pipeline mode, automatic: ranges of >= 256 K envs, at most 4 (copy / compute overlap)
  const int C = pipe ? (e->chunks > 0 ? e->chunks : plan_chunks(e, &len)) : plan_chunks(e, &len);
  if (!pipe && C == 1) {
    if (action)
      CXL_CUDA(cudaMemcpyAsync(e->action, action, sizeof(float) * 2 * B, cudaMemcpyHostToDevice, e->stream));
    int rc = env_enqueue_step(e, action ? e->action : nullptr, obs ? e->obs : nullptr);
    if (rc) return rc;
    if (obs)
      CXL_CUDA(cudaMemcpyAsync(obs, e->obs, sizeof(float) * COTIX_LUNAR_NOBS * B, cudaMemcpyDeviceToHost, e->stream));
    if (reward)
      CXL_CUDA(cudaMemcpyAsync(reward, e->reward, sizeof(float) * B, cudaMemcpyDeviceToHost, e->stream));
    if (done) CXL_CUDA(cudaMemcpyAsync(done, e->done, B, cudaMemcpyDeviceToHost, e->stream));
    CXL_CUDA(cudaStreamSynchronize(e->stream));
    return COTIX_OK;
  }
  const bool graphable = e->host_graphs && !cx_lunar_pipeline_profiling() && host_pinned(action) &&
                         host_pinned(obs) && host_pinned(reward) && host_pinned(done);
  if (!graphable) {
    int rc = pipe ? enqueue_ranges(e, C, action, obs, reward, done)
                  : enqueue_chunked(e, C, len, action, obs, reward, done, false);
    if (rc) return rc;
    if (!pipe) CXL_CUDA(cudaStreamSynchronize(e->s_out));
    CXL_CUDA(cudaStreamSynchronize(e->stream));
    return COTIX_OK;
  }
  // the whole pipelined step as one replayed CUDA graph, cached per set of host buffers (an RL
  // loop passes the same few pinned buffers every step)
  HostGraph* hit = nullptr;
  for (auto& g : *e->hgraphs)
    if (g.action == action && g.obs == obs && g.reward == reward && g.done == done && g.chunks == C &&
        g.mode == e->mode)
      hit = &g;
  if (!hit) {
    if (e->hgraphs->size() >= 8) drop_host_graphs(e);
    cudaGraph_t g;
    CXL_CUDA(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
    const int64_t before = cotix_launch_count();
    int rc = pipe ? enqueue_ranges(e, C, action, obs, reward, done)
                  : enqueue_chunked(e, C, len, action, obs, reward, done, true);
    cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
    const int64_t captured = cotix_launch_count() - before;
    cx_count_launch_n(-captured);  // captured, not executed
    if (rc) return rc;
    CXL_CUDA(ce);
    HostGraph hg{action, obs, reward, done, C, e->mode, nullptr, captured};
    CXL_CUDA(cudaGraphInstantiate(&hg.exec, g, 0));
    cudaGraphDestroy(g);
    e->hgraphs->push_back(hg);
    hit = &e->hgraphs->back();
  }
  CXL_CUDA(cudaGraphLaunch(hit->exec, e->stream));
  cx_count_launch_n(hit->launches);
  CXL_CUDA(cudaStreamSynchronize(e->stream));
  return COTIX_OK;
}

int cotix_lunar_env_set_host_graphs(CotixLunarEnv* e, int enable) {
  CXL_REQUIRE(e != nullptr, "env is null");
  e->host_graphs = enable ? 1 : 0;
  if (!enable) drop_host_graphs(e);
  return COTIX_OK;
}

int cotix_lunar_env_get_state_host(CotixLunarEnv* e, float* state) {
  CXL_REQUIRE(e && state, "null argument");
  CXL_CUDA(cudaSetDevice(e->device));
  CXL_CUDA(cudaMemcpyAsync(state, e->state, sizeof(float) * COTIX_LUNAR_NF * (size_t)e->B,
                           cudaMemcpyDeviceToHost, e->stream));
  CXL_CUDA(cudaStreamSynchronize(e->stream));
  return COTIX_OK;
}
int cotix_lunar_env_set_state_host(CotixLunarEnv* e, const float* state) {
  CXL_REQUIRE(e && state, "null argument");
  CXL_CUDA(cudaSetDevice(e->device));
  CXL_CUDA(cudaMemcpyAsync(e->state, state, sizeof(float) * COTIX_LUNAR_NF * (size_t)e->B,
                           cudaMemcpyHostToDevice, e->stream));
  CXL_CUDA(cudaStreamSynchronize(e->stream));
  return COTIX_OK;
}
float* cotix_lunar_env_state_ptr(CotixLunarEnv* e) { return e ? e->state : nullptr; }
void* cotix_lunar_env_stream(CotixLunarEnv* e) { return e ? (void*)e->stream : nullptr; }

int cotix_lunar_env_rollout_graph(CotixLunarEnv* e, int32_t n_steps, const float* action_dev) {
  CXL_REQUIRE(e != nullptr && n_steps > 0, "bad arguments");
  CXL_CUDA(cudaSetDevice(e->device));
  auto it = e->graphs->find(n_steps);
  if (it != e->graphs->end() && e->graph_action != action_dev) {
    for (auto& kv : *e->graphs) cudaGraphExecDestroy(kv.second);
    e->graphs->clear();
    it = e->graphs->end();
  }
  if (it == e->graphs->end()) {
    cudaGraph_t g;
    CXL_CUDA(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
    int rc = COTIX_OK;
    const int64_t before = cotix_launch_count();
    const bool pipe = e->mode == COTIX_MODE_PIPELINE || (e->mode == COTIX_MODE_AUTO && e->B >= 16384);
    if (pipe)  // the ranges run their n_steps without joining
      rc = cx_launch_lunar_pipeline_steps(e->stream, e->B, e->state, e->state, 0, e->B, action_dev, 0, 1, n_steps,
                                          e->reward, e->done, nullptr, &e->P, e->ws, e->ws_bytes);
    else
      for (int s = 0; s < n_steps && rc == COTIX_OK; ++s) rc = env_enqueue_step(e, action_dev, nullptr);
    cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
    const int64_t captured = cotix_launch_count() - before;
    e->graph_launches = captured / n_steps;  // kernels per step under the current strategy
    cx_count_launch_n(-captured);            // captured, not executed
    if (rc) return rc;
    CXL_CUDA(ce);
    cudaGraphExec_t ge;
    CXL_CUDA(cudaGraphInstantiate(&ge, g, 0));
    cudaGraphDestroy(g);
    (*e->graphs)[n_steps] = ge;
    e->graph_action = action_dev;
    it = e->graphs->find(n_steps);
  }
  CXL_CUDA(cudaGraphLaunch(it->second, e->stream));
  cx_count_launch_n(e->graph_launches * n_steps);
  return COTIX_OK;
}

}  // extern "C"
