/*
 * cx_trig.h — the engine's definition of sin/cos/atan2 in fp32.
 *
 * The reference gets jnp.cos / jnp.sin / jnp.arctan2 from XLA (its own polynomial
 * approximations; call sites cotix/_geometry_utils.py:63,84-85,107-108).  libm, CUDA's
 * libdevice and XLA all agree to ~1 ulp but not bitwise, and sin/cos/atan2 feed discrete
 * decisions (vertex order, GJK branches).  To make the engine reproducible bit for bit on
 * any IEEE-754 machine these three functions are DEFINED here using only +, -, *, / and
 * comparisons (no FMA contraction: the CUDA side is built with -fmad=false, the host side
 * with -ffp-contract=off).  Accuracy against libm over |x| <= 1e4: sin, cos <= 2 ulp, atan2 <= 3 ulp (tested in
 * tests/test_trig.py).
 *
 * Plain C, usable from nvcc (device) and gcc (host).  The CPU checker used by the tests links
 * libm by default and can switch to these definitions (its trig mode 1) so that parity tests
 * can demand exact equality.
 */
#ifndef CX_TRIG_H
#define CX_TRIG_H

#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define CX_HD __host__ __device__ __forceinline__
#else
#define CX_HD static inline
#endif

/* sin and cos of x (radians).  |x| >= 1e9, inf and NaN give NaN for both. */
CX_HD void cx_sincosf(float x, float* s_out, float* c_out) {
  float ax = fabsf(x);
  if (!(ax < 1.0e9f)) {
    float q = x - x; /* NaN for inf/NaN */
    q = q / q;       /* 0/0 = NaN for finite huge x */
    *s_out = q;
    *c_out = q;
    return;
  }
  /* Cody-Waite style reduction done in double: k = round(x * 2/pi), r = x - k*pi/2 */
  double xd = (double)x;
  double kd = rint(xd * 0.63661977236758134308);
  double t = kd * 1.57079632679489661923;
  double rd = xd - t;
  int q = ((int)kd) & 3;
  float r = (float)rd;
  float z = r * r;
  /* minimax polynomials on [-pi/4, pi/4] (Cephes single-precision coefficients) */
  float ps = -1.9515295891e-4f * z;
  ps = ps + 8.3321608736e-3f;
  ps = ps * z;
  ps = ps - 1.6666654611e-1f;
  ps = ps * z;
  ps = ps * r;
  ps = ps + r;
  float pc = 2.443315711809948e-5f * z;
  pc = pc - 1.388731625493765e-3f;
  pc = pc * z;
  pc = pc + 4.166664568298827e-2f;
  pc = pc * z;
  pc = pc * z;
  float hz = 0.5f * z;
  pc = pc - hz;
  pc = pc + 1.0f;
  float s, c;
  if (q == 0) {
    s = ps;
    c = pc;
  } else if (q == 1) {
    s = pc;
    c = -ps;
  } else if (q == 2) {
    s = -ps;
    c = -pc;
  } else {
    s = -pc;
    c = ps;
  }
  *s_out = s;
  *c_out = c;
}

/* atan of a non-negative finite-or-inf argument, result in [0, pi/2] */
CX_HD float cx_atan_pos(float x) {
  float y0;
  float t;
  if (x > 2.414213562373095f) { /* tan(3pi/8) */
    y0 = 1.5707963267948966f;
    t = -1.0f / x;
  } else if (x > 0.4142135623730950f) { /* tan(pi/8) */
    y0 = 0.7853981633974483f;
    t = (x - 1.0f) / (x + 1.0f);
  } else {
    y0 = 0.0f;
    t = x;
  }
  float z = t * t;
  float p = 8.05374449538e-2f * z;
  p = p - 1.38776856032e-1f;
  p = p * z;
  p = p + 1.99777106478e-1f;
  p = p * z;
  p = p - 3.33329491539e-1f;
  p = p * z;
  p = p * t;
  p = p + t;
  return y0 + p;
}

/* atan2(y, x) in (-pi, pi]; NaN in -> NaN out; atan2(0, 0) = 0 */
CX_HD float cx_atan2f(float y, float x) {
  if (x != x || y != y) return x + y;
  float ay = fabsf(y), ax = fabsf(x);
  float a;
  if (ax == 0.0f && ay == 0.0f) {
    a = 0.0f;
  } else if (isinf(ax) && isinf(ay)) {
    a = 0.7853981633974483f;
  } else {
    a = cx_atan_pos(ay / ax); /* ax == 0 -> inf -> pi/2 */
  }
  if (x < 0.0f) a = 3.14159265358979323846f - a;
  return (y < 0.0f) ? -a : a;
}

#endif /* CX_TRIG_H */
