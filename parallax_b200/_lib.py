"""Loader / builder of libcotix_b200.so — the C-ABI in include/cotix_b200.h.

The library is built in-tree with nvcc for sm_100a only (`-gencode
arch=compute_100a,code=sm_100a -lineinfo -fmad=false`).  There is no CPU fallback: if the
shared object is missing or a compute entry point is called without a B200-class device the
call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_PKG, "csrc")
# Two flavours of the same sources (COTIX_FLAVOUR = "exact" (default) | "fast"):
#   exact  libcotix_b200.so       -fmad=false, IEEE division / sqrt: bit-identical to the oracle (the parity build)
#   fast   libcotix_b200_fast.so  -fmad=true -prec-div=false -prec-sqrt=false -DCOTIX_MATH_FAST: FMA contraction
#          and 2-ulp division / sqrt; validated at north_star's 1e-5 relative per step
#          (tests/test_fast_flavour.py), NOT bit-exact; ~16 % faster on the issue-bound step
# COTIX_LIB_PATH: load another build of the library (kernel-tuning experiments)
FLAVOUR = os.environ.get("COTIX_FLAVOUR", "exact")
if FLAVOUR not in ("exact", "fast"):
    raise ValueError(f"COTIX_FLAVOUR must be 'exact' or 'fast', not {FLAVOUR!r}")
FLAVOUR_LIBS = {"exact": os.path.join(_PKG, "libcotix_b200.so"), "fast": os.path.join(_PKG, "libcotix_b200_fast.so")}
FAST_FLAGS = ["-fmad=true", "-prec-div=false", "-prec-sqrt=false", "-DCOTIX_MATH_FAST=1"]
LIB_PATH = os.environ.get("COTIX_LIB_PATH") or FLAVOUR_LIBS[FLAVOUR]
SOURCES = ["abi.cu", "abi_lunar.cu", "narrowphase_kernel.cu", "lunar_kernel.cu", "lunar_pipeline.cu", "lunar_rollout.cu",
           "scene_kernel.cu", "lunar_adjoint.cu", "raster_kernel.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-t", "8",
    "-fmad=false",            # IEEE-exact fp32: no FMA contraction (see csrc/cx_math.cuh)
    "-Xcompiler", "-fPIC", "-shared",
]


class CotixError(RuntimeError):
    pass


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise CotixError("nvcc not found; cannot build libcotix_b200.so")


def _stale(path=None) -> bool:
    path = path or LIB_PATH
    if not os.path.exists(path):
        return True
    t = os.path.getmtime(path)
    deps = [os.path.join(_CSRC, f) for f in os.listdir(_CSRC)]
    deps.append(os.path.join(_PKG, "..", "include", "cotix_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False, flavour: str | None = None) -> str:
    """Compile every CUDA translation unit into the flavour's shared library (default: the flavour
    this process loads; `build_all` makes both)."""
    out = LIB_PATH if flavour is None else FLAVOUR_LIBS[flavour]
    fast = (FLAVOUR if flavour is None else flavour) == "fast" and not os.environ.get("COTIX_LIB_PATH")
    if not force and not _stale(out):
        return out
    extra = os.environ.get("COTIX_NVCC_EXTRA", "").split()  # experiments, e.g. -DCX_BAND_ROWS=12
    if fast:
        extra = FAST_FLAGS + extra
    flags = list(NVCC_FLAGS)
    if any(x.startswith("-fmad=") for x in extra):  # the fast flavour / an experiment overrides the contraction flag
        flags = [f for f in flags if not f.startswith("-fmad=")]
    cmd = [_nvcc(), *flags, *extra, "-o", out, *[os.path.join(_CSRC, s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    env = dict(os.environ)
    env.pop("CXX", None)
    res = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if res.returncode != 0:
        raise CotixError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return out


def build_all(force: bool = False) -> list:
    """Both flavours (what __graft_entry__.build() does)."""
    return [build(force=force, flavour="exact"), build(force=force, flavour="fast")]


class NarrowParams(C.Structure):
    """CotixNarrowParams"""

    _fields_ = [
        ("init_dir", C.c_float * 2),
        ("gjk_max_iters", C.c_int32),
        ("epa_iters_circle_poly", C.c_int32),
    ]


class SceneParams(C.Structure):
    """CotixSceneParams"""

    _fields_ = [
        ("dt", C.c_float),
        ("init_dir", C.c_float * 2),
        ("gjk_max_iters", C.c_int32),
        ("coin_p", C.c_float),
        ("baumgarte", C.c_float),
        ("baumgarte_dt", C.c_float),
        ("advance_key", C.c_int32),
    ]


class LunarParams(C.Structure):
    """CotixLunarParams (include/cotix_b200.h); defaults = the reference's literals."""

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

    def copy(self) -> "LunarParams":
        q = LunarParams()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(LunarParams))
        return q


# every symbol include/cotix_b200.h declares
EXPORTS = [
    "cotix_narrowphase", "cotix_narrowphase_host", "cotix_order_clockwise",
    "cotix_narrowphase_workspace_bytes", "cotix_narrowphase_ws", "cotix_narrowphase_static",
    "cotix_lunar_default_params", "cotix_lunar_reset", "cotix_lunar_step",
    "cotix_lunar_workspace_bytes", "cotix_lunar_step_ws", "cotix_lunar_env_set_mode", "cotix_lunar_env_set_chunks", "cotix_lunar_env_set_host_graphs",
    "cotix_lunar_pipeline_profile", "cotix_lunar_pipeline_stage_ms",
    "cotix_lunar_env_create", "cotix_lunar_env_destroy", "cotix_lunar_env_reset_host",
    "cotix_lunar_env_step_host", "cotix_lunar_env_get_state_host",
    "cotix_lunar_env_set_state_host", "cotix_lunar_env_state_ptr", "cotix_lunar_env_stream",
    "cotix_lunar_env_rollout_graph", "cotix_euler_step", "cotix_resolve_collision",
    "cotix_support", "cotix_threefry2x32", "cotix_philox4x32", "cotix_last_error",
    "cotix_abi_version", "cotix_math_flavour", "cotix_launch_count", "cotix_device_info",
    "cotix_scene_default_params", "cotix_scene_create", "cotix_scene_destroy", "cotix_scene_info",
    "cotix_scene_step",
    "cotix_lunar_step_bwd", "cotix_lunar_rollout", "cotix_lunar_rollout_bwd",
    "cotix_lunar_adjoint_workspace_bytes", "cotix_lunar_adjoint_decisions_bytes", "cotix_raster_draw", "cotix_lunar_draw", "cotix_lunar_prims",
    "cotix_lunar_steps_ws", "cotix_lunar_pipeline_config", "cotix_lunar_pipeline_counters",
    "cotix_graph_begin_capture", "cotix_graph_end_capture", "cotix_graph_launch", "cotix_graph_info",
    "cotix_graph_destroy",
    "cotix_set_device", "cotix_device_count", "cotix_device_malloc", "cotix_device_free", "cotix_host_malloc",
    "cotix_host_free", "cotix_memcpy", "cotix_memset", "cotix_stream_create", "cotix_stream_destroy",
    "cotix_stream_synchronize", "cotix_event_create", "cotix_event_destroy", "cotix_event_record",
    "cotix_event_elapsed_ms",
]

_lib = None


def lib() -> C.CDLL:
    """The loaded shared library.  Raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CotixError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (nvcc, sm_100a).  There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    L.cotix_last_error.restype = C.c_char_p
    L.cotix_launch_count.restype = C.c_int64
    L.cotix_lunar_env_state_ptr.restype = C.c_void_p
    L.cotix_lunar_env_stream.restype = C.c_void_p
    vp, i64 = C.c_void_p, C.c_int64
    L.cotix_narrowphase.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, C.POINTER(NarrowParams)]
    L.cotix_narrowphase_host.argtypes = [C.c_int, i64, vp, vp, vp, vp, vp, vp, C.POINTER(NarrowParams)]
    L.cotix_narrowphase_workspace_bytes.restype = C.c_int64
    L.cotix_narrowphase_workspace_bytes.argtypes = [i64]
    L.cotix_narrowphase_ws.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, C.POINTER(NarrowParams), vp, i64]
    L.cotix_narrowphase_static.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, C.POINTER(NarrowParams)]
    L.cotix_order_clockwise.argtypes = [vp, i64, vp]
    L.cotix_support.argtypes = [vp, i64, vp, vp, vp]
    L.cotix_lunar_default_params.argtypes = [C.POINTER(LunarParams)]
    L.cotix_lunar_reset.argtypes = [vp, i64, vp, vp, i64, C.POINTER(LunarParams)]
    L.cotix_lunar_step.argtypes = [vp, i64, vp, vp, i64, vp, vp, vp, vp, C.POINTER(LunarParams)]
    L.cotix_lunar_workspace_bytes.restype = C.c_int64
    L.cotix_lunar_workspace_bytes.argtypes = [i64]
    L.cotix_lunar_step_ws.argtypes = [vp, i64, vp, vp, i64, vp, vp, vp, vp, C.POINTER(LunarParams), vp, i64]
    L.cotix_lunar_env_set_mode.argtypes = [vp, C.c_int]
    L.cotix_lunar_env_set_chunks.argtypes = [vp, C.c_int]
    L.cotix_lunar_env_set_host_graphs.argtypes = [vp, C.c_int]
    L.cotix_lunar_pipeline_profile.argtypes = [C.c_int]
    L.cotix_lunar_pipeline_stage_ms.restype = C.c_int64
    L.cotix_lunar_pipeline_stage_ms.argtypes = [C.POINTER(C.c_double)]
    L.cotix_lunar_env_create.argtypes = [C.POINTER(vp), C.c_int, i64, C.POINTER(LunarParams)]
    L.cotix_lunar_env_destroy.argtypes = [vp]
    L.cotix_lunar_env_reset_host.argtypes = [vp, vp]
    L.cotix_lunar_env_step_host.argtypes = [vp, vp, vp, vp, vp]
    L.cotix_lunar_env_get_state_host.argtypes = [vp, vp]
    L.cotix_lunar_env_set_state_host.argtypes = [vp, vp]
    L.cotix_lunar_env_state_ptr.argtypes = [vp]
    L.cotix_lunar_env_stream.argtypes = [vp]
    L.cotix_lunar_env_rollout_graph.argtypes = [vp, C.c_int32, vp]
    L.cotix_lunar_step_bwd.argtypes = [vp, i64, vp, vp, i64, vp, vp, vp, vp, vp, C.POINTER(LunarParams)]
    L.cotix_lunar_rollout.argtypes = [vp, i64, C.c_int32, vp, i64, vp, C.POINTER(LunarParams), vp, i64]
    L.cotix_lunar_rollout_bwd.argtypes = [vp, i64, C.c_int32, vp, i64, vp, vp, vp, vp, C.POINTER(LunarParams), vp, i64,
                                          C.c_int32]
    L.cotix_lunar_adjoint_workspace_bytes.restype = C.c_int64
    L.cotix_lunar_adjoint_workspace_bytes.argtypes = [i64, C.c_int32]
    L.cotix_lunar_adjoint_decisions_bytes.restype = C.c_int64
    L.cotix_lunar_adjoint_decisions_bytes.argtypes = [i64, C.c_int32]
    L.cotix_raster_draw.argtypes = [vp, i64, vp, vp, C.c_int32, vp]
    L.cotix_lunar_draw.argtypes = [vp, i64, vp, i64, vp, vp, C.POINTER(LunarParams)]
    L.cotix_lunar_prims.argtypes = [vp, i64, vp, i64, vp, vp, C.POINTER(LunarParams)]
    L.cotix_scene_default_params.argtypes = [C.POINTER(SceneParams)]
    L.cotix_scene_default_params.restype = None
    L.cotix_scene_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int32, vp, C.c_int32, vp, vp]
    L.cotix_scene_destroy.argtypes = [vp]
    L.cotix_scene_info.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64), vp]
    L.cotix_scene_step.argtypes = [vp, vp, i64, vp, i64, vp, C.POINTER(SceneParams), vp, vp, vp]
    L.cotix_euler_step.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, C.c_float]
    L.cotix_resolve_collision.argtypes = [vp, i64, vp, vp, vp, vp, C.c_float, C.c_float]
    L.cotix_threefry2x32.argtypes = [vp, i64, C.c_uint32, C.c_uint32, vp, vp, vp]
    L.cotix_philox4x32.argtypes = [vp, i64, C.c_uint32, C.c_uint32, vp, vp]
    L.cotix_device_info.argtypes = [C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int),
                                    C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.cotix_lunar_steps_ws.argtypes = [vp, i64, vp, i64, vp, C.c_int32, C.c_int32, vp, vp, vp, C.POINTER(LunarParams),
                                       vp, i64]
    L.cotix_lunar_pipeline_config.argtypes = [C.c_int, C.c_int, C.c_int, i64]
    L.cotix_lunar_pipeline_counters.argtypes = [vp, i64, vp, C.POINTER(i64)]
    L.cotix_graph_begin_capture.argtypes = [vp]
    L.cotix_graph_end_capture.argtypes = [vp, C.POINTER(vp)]
    L.cotix_graph_launch.argtypes = [vp, vp]
    L.cotix_graph_info.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(i64)]
    L.cotix_graph_destroy.argtypes = [vp]
    L.cotix_set_device.argtypes = [C.c_int]
    L.cotix_device_count.argtypes = [C.POINTER(C.c_int)]
    L.cotix_device_malloc.argtypes = [i64, C.POINTER(vp)]
    L.cotix_device_free.argtypes = [vp]
    L.cotix_host_malloc.argtypes = [i64, C.POINTER(vp)]
    L.cotix_host_free.argtypes = [vp]
    L.cotix_memcpy.argtypes = [vp, vp, vp, i64]
    L.cotix_memset.argtypes = [vp, vp, C.c_int, i64]
    L.cotix_stream_create.argtypes = [C.POINTER(vp)]
    L.cotix_stream_destroy.argtypes = [vp]
    L.cotix_stream_synchronize.argtypes = [vp]
    L.cotix_event_create.argtypes = [C.POINTER(vp)]
    L.cotix_event_destroy.argtypes = [vp]
    L.cotix_event_record.argtypes = [vp, vp]
    L.cotix_event_elapsed_ms.argtypes = [vp, vp, C.POINTER(C.c_float)]
    want = 1 if (FLAVOUR == "fast" and not os.environ.get("COTIX_LIB_PATH")) else None
    if want is not None and L.cotix_math_flavour() != want:
        raise CotixError(f"{LIB_PATH} is not the fast-math build")
    if L.cotix_abi_version() != 2:
        raise CotixError("libcotix_b200.so ABI version mismatch")
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().cotix_last_error()
        raise CotixError(f"cotix error {rc}: {msg.decode() if msg else ''}")


def launch_count() -> int:
    return int(lib().cotix_launch_count())


def default_lunar_params() -> LunarParams:
    p = LunarParams()
    check(lib().cotix_lunar_default_params(C.byref(p)))
    return p


def default_narrow_params() -> NarrowParams:
    lp = default_lunar_params()
    p = NarrowParams()
    p.init_dir[0], p.init_dir[1] = lp.init_dir[0], lp.init_dir[1]
    p.gjk_max_iters = 32
    p.epa_iters_circle_poly = 128
    return p
