// TEST INFRASTRUCTURE — drives the handler bodies of parallax_b200/csrc/cotix_xla_ffi.cc through
// the mock XLA FFI types (tests/xla_ffi_mock/xla/ffi/api/ffi.h) with raw device pointers, so that
// tests/test_xla_ffi_shim.py can execute them on the GPU through ctypes.
#define COTIX_XLA_FFI_MOCK 1
#include "../../parallax_b200/csrc/cotix_xla_ffi.cc"

#include <cstdio>

using namespace cotix_ffi;

static int done_(const ffi::Error& e, char* err, int errlen) {
  if (e.success()) return 0;
  if (err && errlen > 0) snprintf(err, (size_t)errlen, "%s", e.message().c_str());
  return -1;
}
static Words words(const int32_t* w, int n) { return Words(w, (size_t)n); }

extern "C" {

int shim_reset(void* stream, int64_t B, uint32_t* keys, float* state, const int32_t* w, int nw, char* err, int el) {
  return done_(LunarResetImpl((cudaStream_t)stream, U32(keys, {B, 2}), RF32(F32(state, {36, B})), words(w, nw)), err, el);
}
int shim_step(void* stream, int64_t B, float* state, float* action, void* ws, int64_t ws_bytes, float* state_out,
              float* reward, uint8_t* done, float* obs, const int32_t* w, int nw, char* err, int el) {
  F32 act = action ? F32(action, {2, B}) : F32(nullptr, {0, 2, B});
  return done_(LunarStepImpl((cudaStream_t)stream, F32(state, {36, B}), act, U8(ws, {ws_bytes}),
                             RF32(F32(state_out, {36, B})), RF32(F32(reward, {B})), RU8(U8(done, {B})),
                             RF32(F32(obs, {8, B})), words(w, nw)),
               err, el);
}
int shim_steps(void* stream, int64_t B, float* state, float* actions, int64_t n_act, void* ws, int64_t ws_bytes,
               float* state_out, float* reward, uint8_t* done, float* obs, const int32_t* w, int nw, int32_t n_steps,
               char* err, int el) {
  return done_(LunarStepsImpl((cudaStream_t)stream, F32(state, {36, B}), F32(actions, {n_act, 2, B}),
                              U8(ws, {ws_bytes}), RF32(F32(state_out, {36, B})), RF32(F32(reward, {B})),
                              RU8(U8(done, {B})), RF32(F32(obs, {8, B})), words(w, nw), n_steps),
               err, el);
}
int shim_narrowphase(void* stream, int64_t n, float* a, float* b, void* ws, int64_t ws_bytes, uint8_t* exists,
                     float* pen, float* cp, uint8_t* status, const float* init_dir, int32_t gjk, int32_t epa, char* err,
                     int el) {
  return done_(NarrowphaseImpl((cudaStream_t)stream, F32(a, {n, 16}), F32(b, {n, 16}), U8(ws, {ws_bytes}),
                               RU8(U8(exists, {n})), RF32(F32(pen, {n, 2})), RF32(F32(cp, {n, 2})), RU8(U8(status, {n})),
                               ffi::Span<const float>(init_dir, 2), gjk, epa),
               err, el);
}
int shim_rollout(void* stream, int64_t B, int64_t T, float* state0, float* actions, void* ws, int64_t ws_bytes,
                 float* tape, const int32_t* w, int nw, char* err, int el) {
  return done_(LunarRolloutImpl((cudaStream_t)stream, F32(state0, {36, B}), F32(actions, {T, 2, B}), U8(ws, {ws_bytes}),
                                RF32(F32(tape, {T + 1, 36, B})), words(w, nw)),
               err, el);
}
int shim_rollout_bwd(void* stream, int64_t B, int64_t T, float* tape, float* actions, float* g_final, void* ws,
                     int64_t ws_bytes, float* g0, float* g_actions, uint32_t* flags, const int32_t* w, int nw,
                     int32_t jacobians, char* err, int el) {
  return done_(LunarRolloutBwdImpl((cudaStream_t)stream, F32(tape, {T + 1, 36, B}), F32(actions, {T, 2, B}),
                                   F32(g_final, {18, B}), U8(ws, {ws_bytes}), RF32(F32(g0, {18, B})),
                                   RF32(F32(g_actions, {T, 2, B})), RU32(U32(flags, {B})), words(w, nw), jacobians),
               err, el);
}

}  // extern "C"
