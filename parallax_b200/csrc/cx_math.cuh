// cx_math.cuh — float32 helpers with jax.numpy semantics for the cotix B200 engine.
//
// Everything on this path is fp32 CUDA-core arithmetic (no tensor cores: nothing here is a
// dense contraction).  The translation units are compiled with -fmad=false and without
// fast-math, so +,-,*,/ and sqrtf are IEEE-754 correctly rounded and the results are
// reproducible bit for bit against a host restatement.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "cx_trig.h"

#define CX_DEV __device__ __forceinline__
#define CX_NAN __int_as_float(0x7fc00000)
#define CX_INF __int_as_float(0x7f800000)

struct V2 {
  float x, y;
};
CX_DEV V2 mk2(float x, float y) {
  V2 r;
  r.x = x;
  r.y = y;
  return r;
}
CX_DEV V2 operator+(V2 a, V2 b) { return mk2(a.x + b.x, a.y + b.y); }
CX_DEV V2 operator-(V2 a, V2 b) { return mk2(a.x - b.x, a.y - b.y); }
CX_DEV V2 operator-(V2 a) { return mk2(-a.x, -a.y); }
CX_DEV V2 operator*(V2 a, float s) { return mk2(a.x * s, a.y * s); }
CX_DEV V2 operator*(float s, V2 a) { return mk2(s * a.x, s * a.y); }
CX_DEV V2 operator/(V2 a, float s) { return mk2(a.x / s, a.y / s); }
// jnp.dot / jnp.cross / jnp.sum(a**2) / jnp.linalg.norm on 2-vectors
CX_DEV float dot2(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
CX_DEV float cross2(V2 a, V2 b) { return a.x * b.y - a.y * b.x; }
CX_DEV float sumsq2(V2 a) { return a.x * a.x + a.y * a.y; }
CX_DEV float norm2(V2 a) { return sqrtf(a.x * a.x + a.y * a.y); }
CX_DEV bool isnan2(V2 a) { return (a.x != a.x) || (a.y != a.y); }
CX_DEV V2 nan2() { return mk2(CX_NAN, CX_NAN); }
// fast_normal / perpendicular_vector: cotix/_geometry_utils.py:30-34,70-72
CX_DEV V2 perp(V2 a) { return mk2(-a.y, a.x); }
// jnp.minimum / jnp.maximum propagate NaN (fminf/fmaxf do not)
CX_DEV float jmin(float a, float b) {
  if (a != a || b != b) return CX_NAN;
  return a < b ? a : b;
}
CX_DEV float jmax(float a, float b) {
  if (a != a || b != b) return CX_NAN;
  return a > b ? a : b;
}
CX_DEV float jclip(float t, float lo, float hi) { return jmin(jmax(t, lo), hi); }
CX_DEV float jsign(float v) {
  if (v != v) return CX_NAN;
  return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f);
}
