"""cx_trig.h (the engine's +,-,*,/ definitions of sin/cos/atan2) against libm: sin/cos <= 2 ulp, atan2 <= 3 ulp over the
range the step uses.  CPU only (the header is plain C shared by nvcc and gcc)."""
import numpy as np


def _ulp_err(got, want):
    want = want.astype(np.float32)
    ulp = np.spacing(np.maximum(np.abs(want), np.float32(1e-30)))
    return np.abs(got.astype(np.float64) - want.astype(np.float64)) / ulp


def test_sincos_atan2_vs_libm(oracle):
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-1e4, 1e4, 200000), rng.uniform(-4, 4, 200000),
                        [0.0, -0.0, 1e-20, np.pi / 2, np.pi, -np.pi]]).astype(np.float32)
    y = rng.normal(size=x.size).astype(np.float32)
    oracle.set_trig_mode(0)
    s0, c0, a0 = oracle.trig(x, y)
    oracle.set_trig_mode(1)
    try:
        s1, c1, a1 = oracle.trig(x, y)
    finally:
        oracle.set_trig_mode(0)
    # sin/cos: absolute error near the zeros of the function is what matters; 2 ulp of 1.0 there
    assert np.max(np.abs(s1.astype(np.float64) - s0)) < 2.5e-7
    assert np.max(np.abs(c1.astype(np.float64) - c0)) < 2.5e-7
    big = np.abs(s0) > 0.1
    assert _ulp_err(s1[big], s0[big]).max() <= 2.0
    big = np.abs(c0) > 0.1
    assert _ulp_err(c1[big], c0[big]).max() <= 2.0
    assert _ulp_err(a1, a0).max() <= 3.0


def test_trig_special_values(oracle):
    oracle.set_trig_mode(1)
    try:
        x = np.float32([0.0, np.inf, -np.inf, np.nan, 2e9])
        s, c, _ = oracle.trig(x, np.zeros_like(x))
        assert s[0] == 0.0 and c[0] == 1.0
        assert np.isnan(s[1:]).all() and np.isnan(c[1:]).all()
        # atan2 quadrants and axes
        xs = np.float32([1, 0, -1, 0, -1, 0, np.nan])
        ys = np.float32([0, 1, 0, -1, -0.0, 0, 1])
        _, _, a = oracle.trig(xs, ys)
        assert a[0] == 0.0 and abs(a[1] - np.pi / 2) < 1e-6 and abs(a[2] - np.pi) < 1e-6
        assert abs(a[3] + np.pi / 2) < 1e-6 and a[5] == 0.0 and np.isnan(a[6])
    finally:
        oracle.set_trig_mode(0)
