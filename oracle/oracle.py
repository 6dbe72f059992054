"""ctypes loader for the CPU oracle (oracle/cotix_oracle.cpp).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs
import this module.  Nothing under parallax_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")


def _timing_lib_path() -> str:
    """-march=native code must be built by the machine that runs it: the file name carries a hash
    of this CPU's feature flags, so a library built on another box is never loaded here."""
    import hashlib

    flags = ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = line
                    break
    except OSError:
        pass
    return os.path.join(_HERE, "_native", f"liboracle_timing_{hashlib.md5(flags.encode()).hexdigest()[:10]}.so")

NF = 36
REC = 16


def build(force: bool = False) -> str:
    """Compile liboracle.so with the committed Makefile (g++, -ffp-contract=off)."""
    src = os.path.join(_HERE, "cotix_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


def build_timing(force: bool = False) -> str:
    """The -O3 -march=native build for the timed CPU legs of bench.py, compiled on this machine."""
    src = os.path.join(_HERE, "cotix_oracle.cpp")
    out = _timing_lib_path()
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "timing", f"TIMING_OUT={out}"] + (["-B"] if force else []))
    return out


class NarrowParams(C.Structure):
    _fields_ = [
        ("init_dir", C.c_float * 2),
        ("gjk_max_iters", C.c_int32),
        ("epa_iters_circle_poly", C.c_int32),
    ]


class LunarParams(C.Structure):
    """Mirror of CotixLunarParams (include/cotix_b200.h)."""

    _fields_ = [
        ("dt", C.c_float),
        ("gravity_dv", C.c_float * 2),
        ("init_dir", C.c_float * 2),
        ("baumgarte", C.c_float),
        ("baumgarte_dt", C.c_float),
        ("coin_p", C.c_float),
        ("spring_kp", C.c_float),
        ("spring_kv0", C.c_float),
        ("spring_kv", C.c_float),
        ("leg_damping", C.c_float),
        ("body_props", (C.c_float * 4) * 4),
        ("lander_local", (C.c_float * 2) * 6),
        ("leg_local", ((C.c_float * 2) * 4) * 2),
        ("joint_local", (C.c_float * 2) * 4),
        ("leg_joint2_local", C.c_float * 2),
        ("lander_start", C.c_float * 3),
        ("leg_offset", (C.c_float * 2) * 2),
        ("ground_floor", C.c_float),
        ("terrain_x_outer", C.c_float),
        ("pad_half_width", C.c_float),
        ("pad_y", C.c_float),
        ("gjk_max_iters", C.c_int32),
        ("rng_mode", C.c_int32),
        ("substeps", C.c_int32),
        ("auto_reset", C.c_int32),
        ("judge_enabled", C.c_int32),
        ("max_episode_steps", C.c_int32),
        ("reset_perturb", C.c_int32),
        ("main_engine_dv", C.c_float),
        ("side_engine_dv", C.c_float),
        ("side_engine_dw", C.c_float),
        ("x_limit", C.c_float),
        ("y_limit", C.c_float),
        ("land_speed", C.c_float),
        ("land_angvel", C.c_float),
        ("reset_x", C.c_float * 2),
        ("reset_y", C.c_float * 2),
        ("reset_angle", C.c_float * 2),
        ("reset_v", C.c_float),
        ("reset_w", C.c_float),
        ("philox_seed", C.c_uint32 * 2),
        ("solver_loops", C.c_int32),
    ]


_lib = None
_libs: dict = {}
_build_kind = "parity"


def select_build(kind: str) -> None:
    """"parity" (default): liboracle.so, -ffp-contract=off, the checker.  "timing": the same source
    built -O3 -march=native (liboracle_timing.so) for bench.py's timed CPU legs only."""
    global _lib, _build_kind
    assert kind in ("parity", "timing")
    _build_kind = kind
    _lib = None


def lib():
    global _lib
    if _lib is None:
        if _build_kind not in _libs:
            L = C.CDLL(build() if _build_kind == "parity" else build_timing())
            assert L.oracle_sizeof_lunar_params() == C.sizeof(LunarParams), (
                L.oracle_sizeof_lunar_params(),
                C.sizeof(LunarParams),
            )
            _libs[_build_kind] = L
        _lib = _libs[_build_kind]
    return _lib


def _p(a, ty=C.c_float):
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(ty))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def set_num_threads(n: int = 0) -> int:
    """OpenMP threads of the oracle's batched entry points; n <= 0 = every logical core.  Returns the
    count in effect.  (torchrun exports OMP_NUM_THREADS=1: the timed CPU arm sets this explicitly.)"""
    return int(lib().oracle_set_num_threads(int(n)))


def set_trig_mode(mode: int) -> None:
    """0 = libm (default), 1 = the engine's cx_trig.h definitions (for exact-equality tests)."""
    lib().oracle_set_trig_mode(int(mode))


def trig(x, y):
    """(sin x, cos x, atan2(y, x)) element-wise under the current trig mode."""
    x, y = _f32(x), _f32(y)
    s, c, a = np.zeros_like(x), np.zeros_like(x), np.zeros_like(x)
    lib().oracle_trig(C.c_int64(x.size), _p(x), _p(y), _p(s), _p(c), _p(a))
    return s, c, a


def default_lunar_params() -> LunarParams:
    p = LunarParams()
    lib().oracle_lunar_default_params(C.byref(p))
    return p


def default_narrow_params() -> NarrowParams:
    lp = default_lunar_params()
    p = NarrowParams()
    p.init_dir[0] = lp.init_dir[0]
    p.init_dir[1] = lp.init_dir[1]
    p.gjk_max_iters = 32
    p.epa_iters_circle_poly = 128
    return p


def narrowphase(a, b, params: NarrowParams | None = None, want_iters=False):
    a = _f32(a).reshape(-1, REC)
    b = _f32(b).reshape(-1, REC)
    n = a.shape[0]
    params = params or default_narrow_params()
    exists = np.zeros(n, np.uint8)
    pen = np.zeros((n, 2), np.float32)
    cp = np.zeros((n, 2), np.float32)
    status = np.zeros(n, np.uint8)
    iters = np.zeros((n, 2), np.int32)
    lib().oracle_narrowphase(
        C.c_int64(n), _p(a), _p(b), _p(exists, C.c_uint8), _p(pen), _p(cp),
        _p(status, C.c_uint8), _p(iters, C.c_int32), C.byref(params))
    if want_iters:
        return exists, pen, cp, status, iters
    return exists, pen, cp, status


def order_clockwise(recs):
    recs = _f32(recs).reshape(-1, REC).copy()
    lib().oracle_order_clockwise(C.c_int64(recs.shape[0]), _p(recs))
    return recs


def support(recs, dirs):
    recs = _f32(recs).reshape(-1, REC)
    dirs = _f32(dirs).reshape(-1, 2)
    out = np.zeros_like(dirs)
    lib().oracle_support(C.c_int64(recs.shape[0]), _p(recs), _p(dirs), _p(out))
    return out


def contains(recs, pts):
    recs = _f32(recs).reshape(-1, REC)
    pts = _f32(pts).reshape(-1, 2)
    out = np.zeros(recs.shape[0], np.uint8)
    lib().oracle_contains(C.c_int64(recs.shape[0]), _p(recs), _p(pts), _p(out, C.c_uint8))
    return out.astype(bool)


def gjk(a, b, params: NarrowParams | None = None):
    a = _f32(a).reshape(-1, REC)
    b = _f32(b).reshape(-1, REC)
    n = a.shape[0]
    params = params or default_narrow_params()
    exists = np.zeros(n, np.uint8)
    simplex = np.zeros((n, 3, 2), np.float32)
    lib().oracle_gjk(C.c_int64(n), _p(a), _p(b), C.byref(params), _p(exists, C.c_uint8), _p(simplex))
    return exists.astype(bool), simplex


def euler_step(px, py, vx, vy, ang, angvel, dt):
    px, py, ang = _f32(px).copy(), _f32(py).copy(), _f32(ang).copy()
    vx, vy, angvel = _f32(vx), _f32(vy), _f32(angvel)
    lib().oracle_euler_step(C.c_int64(px.size), _p(px), _p(py), _p(vx), _p(vy), _p(ang), _p(angvel),
                            C.c_float(dt))
    return px, py, ang


def resolve_collision(body1, body2, pen, cp, baumgarte=0.3, baumgarte_dt=0.01):
    body1 = _f32(body1).reshape(-1, 10).copy()
    body2 = _f32(body2).reshape(-1, 10).copy()
    pen = _f32(pen).reshape(-1, 2)
    cp = _f32(cp).reshape(-1, 2)
    lib().oracle_resolve_collision(C.c_int64(body1.shape[0]), _p(body1), _p(body2), _p(pen), _p(cp),
                                   C.c_float(baumgarte), C.c_float(baumgarte_dt))
    return body1, body2


# ---- RNG ----
def threefry2x32(k0, k1, x0, x1):
    x0 = np.ascontiguousarray(x0, np.uint32)
    x1 = np.ascontiguousarray(x1, np.uint32)
    out = np.zeros((x0.size, 2), np.uint32)
    lib().oracle_threefry2x32(C.c_int64(x0.size), C.c_uint32(k0), C.c_uint32(k1), _p(x0, C.c_uint32),
                              _p(x1, C.c_uint32), _p(out, C.c_uint32))
    return out


def philox4x32(k0, k1, ctr):
    ctr = np.ascontiguousarray(ctr, np.uint32).reshape(-1, 4)
    out = np.zeros_like(ctr)
    lib().oracle_philox4x32(C.c_int64(ctr.shape[0]), C.c_uint32(k0), C.c_uint32(k1),
                            _p(ctr, C.c_uint32), _p(out, C.c_uint32))
    return out


def split(key, num=2):
    out = np.zeros((num, 2), np.uint32)
    lib().oracle_split(C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), C.c_int(num), _p(out, C.c_uint32))
    return out


def random_bits(key, n):
    out = np.zeros(n, np.uint32)
    lib().oracle_random_bits(C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), C.c_int(n), _p(out, C.c_uint32))
    return out


def uniform(key, n, lo=0.0, hi=1.0):
    out = np.zeros(n, np.float32)
    lib().oracle_uniform(C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), C.c_int(n), C.c_float(lo),
                         C.c_float(hi), _p(out))
    return out


def normal(key, n):
    out = np.zeros(n, np.float32)
    lib().oracle_normal(C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), C.c_int(n), _p(out))
    return out


def bernoulli(key, p=0.5):
    return bool(lib().oracle_bernoulli(C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), C.c_float(p)))


def choice(key, probs):
    probs = _f32(probs)
    return int(lib().oracle_choice(C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), _p(probs),
                                   C.c_int(probs.size)))


# ---- LunarLander ----
def lunar_reset(keys, params: LunarParams | None = None):
    keys = np.ascontiguousarray(keys, np.uint32).reshape(-1, 2)
    B = keys.shape[0]
    params = params or default_lunar_params()
    state = np.zeros((NF, B), np.float32)
    lib().oracle_lunar_reset(C.c_int64(B), _p(keys, C.c_uint32), _p(state), C.c_int64(B), C.byref(params))
    return state


def lunar_step(state, action=None, params: LunarParams | None = None, dedupe=True, want_obs=False):
    state = _f32(state)
    assert state.shape[0] == NF
    B = state.shape[1]
    params = params or default_lunar_params()
    out = np.zeros_like(state)
    reward = np.zeros(B, np.float32)
    done = np.zeros(B, np.uint8)
    obs = np.zeros((8, B), np.float32)
    act = None if action is None else _f32(action).reshape(2, B)
    lib().oracle_lunar_step(C.c_int64(B), _p(state), _p(out), C.c_int64(B), _p(act), _p(reward),
                            _p(done, C.c_uint8), _p(obs), C.byref(params), C.c_int(1 if dedupe else 0))
    if want_obs:
        return out, reward, done, obs
    return out, reward, done


def lunar_collider_debug(state, params: LunarParams | None = None, dedupe=True):
    state = _f32(state)
    B = state.shape[1]
    params = params or default_lunar_params()
    cell = np.zeros((B, 4, 4), np.uint8)
    cand = np.zeros((B, 4, 4), np.uint8)
    chosen = np.zeros((B, 4), np.int32)
    lib().oracle_lunar_collider_debug(C.c_int64(B), _p(state), C.c_int64(B), C.byref(params),
                                      C.c_int(1 if dedupe else 0), _p(cell, C.c_uint8),
                                      _p(cand, C.c_uint8), _p(chosen, C.c_int32))
    return cell.astype(bool), cand.astype(bool), chosen


def scene_step(bodies10, part_body, part_recs, key, dt=1e-2, init_dir=None, dedupe=True, want_status=False):
    bodies10 = _f32(bodies10).reshape(-1, 10).copy()
    nb = bodies10.shape[0]
    part_body = np.ascontiguousarray(part_body, np.int32)
    part_recs = _f32(part_recs).reshape(-1, REC)
    if init_dir is None:
        lp = default_lunar_params()
        init_dir = np.array([lp.init_dir[0], lp.init_dir[1]], np.float32)
    init_dir = _f32(init_dir)
    cell = np.zeros((nb, nb), np.uint8)
    chosen = np.zeros(nb, np.int32)
    status = C.c_uint32(0)
    rc = lib().oracle_scene_step(
        C.c_int(nb), _p(bodies10), C.c_int(part_recs.shape[0]), _p(part_body, C.c_int32), _p(part_recs),
        C.c_uint32(int(key[0])), C.c_uint32(int(key[1])), C.c_float(dt), _p(init_dir), C.c_int(32),
        C.c_float(0.5), C.c_float(0.3), C.c_float(0.01), C.c_int(1 if dedupe else 0),
        _p(cell, C.c_uint8), _p(chosen, C.c_int32), C.byref(status))
    if rc != 0:
        raise RuntimeError(f"oracle_scene_step failed: {rc}")
    if want_status:
        return bodies10, cell.astype(bool), chosen, int(status.value)
    return bodies10, cell.astype(bool), chosen


# ---- record constructors (host-side helpers for tests) ----
def raster_draw(prims, counts=None):
    """prims int32 [n_img, max_prims, 16] -> uint8 [n_img, 600, 800, 3] (painter's algorithm; the pixel
    rules of parallax_b200/csrc/raster_kernel.cu restated sequentially; unpinned against pygame)."""
    prims = np.ascontiguousarray(prims, dtype=np.int32)
    n, mp = prims.shape[0], prims.shape[1]
    img = np.empty((n, 600, 800, 3), np.uint8)
    cp = None
    if counts is not None:
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        cp = counts.ctypes.data_as(C.c_void_p)
    lib().oracle_raster_draw(C.c_int64(n), prims.ctypes.data_as(C.c_void_p), cp, C.c_int32(mp),
                             img.ctypes.data_as(C.c_void_p))
    return img


def circle_rec(radius, position):
    r = np.zeros(REC, np.float32)
    r[0] = 0
    r[2] = radius
    r[4:6] = position
    return r


def aabb_rec(lower, upper):
    r = np.zeros(REC, np.float32)
    r[0] = 1
    r[4:6] = lower
    r[6:8] = upper
    return r


def polygon_rec(vertices, ordered=False):
    v = np.asarray(vertices, np.float32).reshape(-1, 2)
    r = np.zeros(REC, np.float32)
    r[0] = 2
    r[1] = v.shape[0]
    r[4:4 + 2 * v.shape[0]] = v.reshape(-1)
    if not ordered:
        r = order_clockwise(r[None])[0]
    return r
