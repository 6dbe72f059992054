// cotix_xla_ffi.cc — XLA FFI handlers (jax.ffi custom calls) over the C-ABI of libcotix_b200.so.
//
// The reference's only caller is a jitted JAX function (examples/test_viz.py:24-48:
// `eqx.filter_jit` of `f(env, key)` in a Python loop); this shim is how that caller reaches the
// engine: every handler unpacks XLA buffers and calls ONE entry point of include/cotix_b200.h on
// XLA's own stream.  No logic lives here, no allocation, no synchronisation — the handlers only
// enqueue, so they are command-buffer (CUDA-graph) compatible.  parallax_b200/jax_ffi.py registers
// them with jax.ffi.register_ffi_target and wraps the rollout pair in jax.custom_vjp.
//
// Build (needs the XLA FFI headers that ship with jaxlib >= 0.4.38; parallax_b200/jax_ffi.py does
// this when jax is importable):
//   g++ -O2 -std=c++17 -shared -fPIC -I$(python -c "import jax.ffi as f; print(f.include_dir())")
//       -I/usr/local/cuda/include cotix_xla_ffi.cc -o libcotix_xla_ffi.so
//       -L. -lcotix_b200 -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,'$ORIGIN'
// Without those headers (this image has neither jax nor jaxlib) the file compiles to a stub that
// only reports that the handlers are absent.
//
// Parameters travel as ONE attribute: `params` = the CotixLunarParams struct viewed as int32 words
// (every field is 4 bytes wide), so the reference-side stub fills the struct once with
// cotix_lunar_default_params and the handler never re-derives defaults.
#include <cstdint>
#include <cstring>

#include "../../include/cotix_b200.h"

#if defined(COTIX_XLA_FFI_MOCK) || __has_include("xla/ffi/api/ffi.h")
#include <cuda_runtime_api.h>

#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

namespace cotix_ffi {

using F32 = ffi::Buffer<ffi::F32>;
using U8 = ffi::Buffer<ffi::U8>;
using U32 = ffi::Buffer<ffi::U32>;
using RF32 = ffi::ResultBuffer<ffi::F32>;
using RU8 = ffi::ResultBuffer<ffi::U8>;
using RU32 = ffi::ResultBuffer<ffi::U32>;
using Words = ffi::Span<const int32_t>;

static ffi::Error fail(const char* what) {
  const char* msg = cotix_last_error();
  return ffi::Error::Internal(std::string(what) + ": " + (msg ? msg : ""));
}
static ffi::Error check(int rc, const char* what) { return rc == 0 ? ffi::Error::Success() : fail(what); }

// the C-ABI wants 256-byte aligned scratch; XLA guarantees less, so the reference-side stub passes
// 256 spare bytes and the handler aligns the pointer
struct Scratch {
  void* ptr;
  int64_t bytes;
};
static Scratch scratch(const U8& ws) {
  if (ws.size_bytes() < 256) return {nullptr, 0};
  const uintptr_t p = reinterpret_cast<uintptr_t>(ws.untyped_data());
  const uintptr_t a = (p + 255u) & ~static_cast<uintptr_t>(255u);
  return {reinterpret_cast<void*>(a), (int64_t)ws.size_bytes() - (int64_t)(a - p)};
}

static bool unpack(Words w, CotixLunarParams* p) {
  if (w.size() * sizeof(int32_t) != sizeof(CotixLunarParams)) return false;
  std::memcpy(p, w.begin(), sizeof(CotixLunarParams));
  return true;
}
#define COTIX_PARAMS(P, w)                                                                      \
  CotixLunarParams P;                                                                           \
  if (!unpack(w, &P))                                                                           \
  return ffi::Error::InvalidArgument("params must be the CotixLunarParams struct as int32 words")

// reset(key): keys u32[B, 2] -> state f32[36, B]            (LunarLander(key), _lunar_lander.py:29-143)
ffi::Error LunarResetImpl(cudaStream_t stream, U32 keys, RF32 state, Words params) {
  COTIX_PARAMS(P, params);
  const int64_t B = keys.dimensions()[0];
  return check(cotix_lunar_reset(stream, B, keys.typed_data(), state->typed_data(), B, &P), "cotix_lunar_reset");
}

// step(state, action): f32[36, B], f32[2, B], workspace u8[W] (W = 0: fused kernel)
//   -> state' f32[36, B], reward f32[B], done u8[B], obs f32[8, B]   (f(env, key), test_viz.py:25-44)
ffi::Error LunarStepImpl(cudaStream_t stream, F32 state, F32 action, U8 workspace, RF32 state_out, RF32 reward,
                         RU8 done, RF32 obs, Words params) {
  COTIX_PARAMS(P, params);
  const int64_t B = state.dimensions()[1];
  const float* act = action.element_count() ? action.typed_data() : nullptr;
  const Scratch ws = scratch(workspace);
  if (!ws.ptr)
    return check(cotix_lunar_step(stream, B, state.typed_data(), state_out->typed_data(), B, act,
                                  reward->typed_data(), done->typed_data(), obs->typed_data(), &P),
                 "cotix_lunar_step");
  return check(cotix_lunar_step_ws(stream, B, state.typed_data(), state_out->typed_data(), B, act,
                                   reward->typed_data(), done->typed_data(), obs->typed_data(), &P, ws.ptr,
                                   ws.bytes),
               "cotix_lunar_step_ws");
}

// n steps in one call: state f32[36, B], actions f32[n_actions, 2, B] -> state', reward, done, obs of
// the last step     (the loop of test_viz.py:46-48; alias operand 0 with result 0 to run in place)
ffi::Error LunarStepsImpl(cudaStream_t stream, F32 state, F32 actions, U8 workspace, RF32 state_out, RF32 reward,
                          RU8 done, RF32 obs, Words params, int32_t n_steps) {
  COTIX_PARAMS(P, params);
  const int64_t B = state.dimensions()[1];
  if (state_out->typed_data() != state.typed_data() &&
      cudaMemcpyAsync(state_out->typed_data(), state.typed_data(), state.size_bytes(), cudaMemcpyDeviceToDevice,
                      stream) != cudaSuccess)
    return ffi::Error::Internal("cotix_lunar_steps: copy of the input state failed");
  const int32_t n_act = actions.element_count() ? (int32_t)actions.dimensions()[0] : 0;
  const Scratch ws = scratch(workspace);
  return check(cotix_lunar_steps_ws(stream, B, state_out->typed_data(), B, n_act ? actions.typed_data() : nullptr,
                                    n_act, n_steps, reward->typed_data(), done->typed_data(), obs->typed_data(), &P,
                                    ws.ptr, ws.bytes),
               "cotix_lunar_steps_ws");
}

// contact functions: a, b f32[N, 16] shape records, workspace u8[W] (W = 0: single kernel)
//   -> exists u8[N], penetration f32[N, 2], contact point f32[N, 2], status u8[N]   (_contacts.py:30-315)
ffi::Error NarrowphaseImpl(cudaStream_t stream, F32 a, F32 b, U8 workspace, RU8 exists, RF32 pen, RF32 cp,
                           RU8 status, ffi::Span<const float> init_dir, int32_t gjk_max_iters,
                           int32_t epa_iters_circle_poly) {
  if (init_dir.size() != 2) return ffi::Error::InvalidArgument("init_dir must hold 2 floats");
  CotixNarrowParams np;
  np.init_dir[0] = init_dir.begin()[0];
  np.init_dir[1] = init_dir.begin()[1];
  np.gjk_max_iters = gjk_max_iters;
  np.epa_iters_circle_poly = epa_iters_circle_poly;
  const int64_t n = a.dimensions()[0];
  const Scratch ws = scratch(workspace);
  if (!ws.ptr)
    return check(cotix_narrowphase(stream, n, a.typed_data(), b.typed_data(), exists->typed_data(), pen->typed_data(),
                                   cp->typed_data(), status->typed_data(), &np),
                 "cotix_narrowphase");
  return check(cotix_narrowphase_ws(stream, n, a.typed_data(), b.typed_data(), exists->typed_data(),
                                    pen->typed_data(), cp->typed_data(), status->typed_data(), &np, ws.ptr,
                                    ws.bytes),
               "cotix_narrowphase_ws");
}

// forward rollout that keeps the tape: state0 f32[36, B], actions f32[T, 2, B] -> tape f32[T + 1, 36, B]
ffi::Error LunarRolloutImpl(cudaStream_t stream, F32 state0, F32 actions, U8 workspace, RF32 tape, Words params) {
  COTIX_PARAMS(P, params);
  const int64_t B = state0.dimensions()[1];
  const int32_t T = (int32_t)(tape->dimensions()[0] - 1);
  if (cudaMemcpyAsync(tape->typed_data(), state0.typed_data(), state0.size_bytes(), cudaMemcpyDeviceToDevice,
                      stream) != cudaSuccess)
    return ffi::Error::Internal("cotix_lunar_rollout: copy of the initial state failed");
  const Scratch ws = scratch(workspace);
  return check(cotix_lunar_rollout(stream, B, T, tape->typed_data(), B,
                                   actions.element_count() ? actions.typed_data() : nullptr, &P, ws.ptr, ws.bytes),
               "cotix_lunar_rollout");
}

// backward pass: tape, actions, g_final f32[18, B], workspace -> g0 f32[18, B], g_actions f32[T, 2, B],
// flags u32[B]                                              (jax.grad of the rollout)
ffi::Error LunarRolloutBwdImpl(cudaStream_t stream, F32 tape, F32 actions, F32 g_final, U8 workspace, RF32 g0,
                               RF32 g_actions, RU32 flags, Words params, int32_t jacobians) {
  COTIX_PARAMS(P, params);
  const int64_t B = tape.dimensions()[2];
  const int32_t T = (int32_t)(tape.dimensions()[0] - 1);
  bool ok = true;
  if (g0->typed_data() != g_final.typed_data())
    ok = cudaMemcpyAsync(g0->typed_data(), g_final.typed_data(), g_final.size_bytes(), cudaMemcpyDeviceToDevice,
                         stream) == cudaSuccess;
  ok = ok && cudaMemsetAsync(flags->typed_data(), 0, flags->size_bytes(), stream) == cudaSuccess;
  ok = ok && cudaMemsetAsync(g_actions->typed_data(), 0, g_actions->size_bytes(), stream) == cudaSuccess;
  if (!ok) return ffi::Error::Internal("cotix_lunar_rollout_bwd: preparing the outputs failed");
  const Scratch ws = scratch(workspace);
  return check(cotix_lunar_rollout_bwd(stream, B, T, tape.typed_data(), B,
                                       actions.element_count() ? actions.typed_data() : nullptr, g0->typed_data(),
                                       g_actions->typed_data(), flags->typed_data(), &P, ws.ptr, ws.bytes, jacobians),
               "cotix_lunar_rollout_bwd");
}

}  // namespace cotix_ffi

#define COTIX_STREAM Ctx<ffi::PlatformStream<cudaStream_t>>()

XLA_FFI_DEFINE_HANDLER_SYMBOL(CotixLunarReset, cotix_ffi::LunarResetImpl,
                              ffi::Ffi::Bind().COTIX_STREAM.Arg<cotix_ffi::U32>().Ret<cotix_ffi::F32>()
                                  .Attr<cotix_ffi::Words>("params"),
                              {ffi::Traits::kCmdBufferCompatible});
XLA_FFI_DEFINE_HANDLER_SYMBOL(CotixLunarStep, cotix_ffi::LunarStepImpl,
                              ffi::Ffi::Bind().COTIX_STREAM.Arg<cotix_ffi::F32>().Arg<cotix_ffi::F32>()
                                  .Arg<cotix_ffi::U8>().Ret<cotix_ffi::F32>().Ret<cotix_ffi::F32>()
                                  .Ret<cotix_ffi::U8>().Ret<cotix_ffi::F32>().Attr<cotix_ffi::Words>("params"),
                              {ffi::Traits::kCmdBufferCompatible});
XLA_FFI_DEFINE_HANDLER_SYMBOL(CotixLunarSteps, cotix_ffi::LunarStepsImpl,
                              ffi::Ffi::Bind().COTIX_STREAM.Arg<cotix_ffi::F32>().Arg<cotix_ffi::F32>()
                                  .Arg<cotix_ffi::U8>().Ret<cotix_ffi::F32>().Ret<cotix_ffi::F32>()
                                  .Ret<cotix_ffi::U8>().Ret<cotix_ffi::F32>().Attr<cotix_ffi::Words>("params")
                                  .Attr<int32_t>("n_steps"),
                              {ffi::Traits::kCmdBufferCompatible});
XLA_FFI_DEFINE_HANDLER_SYMBOL(CotixNarrowphase, cotix_ffi::NarrowphaseImpl,
                              ffi::Ffi::Bind().COTIX_STREAM.Arg<cotix_ffi::F32>().Arg<cotix_ffi::F32>()
                                  .Arg<cotix_ffi::U8>().Ret<cotix_ffi::U8>().Ret<cotix_ffi::F32>()
                                  .Ret<cotix_ffi::F32>().Ret<cotix_ffi::U8>()
                                  .Attr<ffi::Span<const float>>("init_dir").Attr<int32_t>("gjk_max_iters")
                                  .Attr<int32_t>("epa_iters_circle_poly"),
                              {ffi::Traits::kCmdBufferCompatible});
XLA_FFI_DEFINE_HANDLER_SYMBOL(CotixLunarRollout, cotix_ffi::LunarRolloutImpl,
                              ffi::Ffi::Bind().COTIX_STREAM.Arg<cotix_ffi::F32>().Arg<cotix_ffi::F32>()
                                  .Arg<cotix_ffi::U8>().Ret<cotix_ffi::F32>().Attr<cotix_ffi::Words>("params"),
                              {ffi::Traits::kCmdBufferCompatible});
XLA_FFI_DEFINE_HANDLER_SYMBOL(CotixLunarRolloutBwd, cotix_ffi::LunarRolloutBwdImpl,
                              ffi::Ffi::Bind().COTIX_STREAM.Arg<cotix_ffi::F32>().Arg<cotix_ffi::F32>()
                                  .Arg<cotix_ffi::F32>().Arg<cotix_ffi::U8>().Ret<cotix_ffi::F32>()
                                  .Ret<cotix_ffi::F32>().Ret<cotix_ffi::U32>().Attr<cotix_ffi::Words>("params")
                                  .Attr<int32_t>("jacobians"),
                              {ffi::Traits::kCmdBufferCompatible});

extern "C" int cotix_xla_ffi_available(void) { return 1; }

#else  // no XLA FFI headers on this machine

extern "C" int cotix_xla_ffi_available(void) { return 0; }

#endif
