"""The JAX boundary: `reset(key)` / `step(state, action)` / the n-step rollout / the contact
functions as `jax.ffi` custom calls on XLA's own stream, and `jax.grad` through the rollout via
`jax.custom_vjp` over the hand-written adjoint (the reference differentiates by tracing,
cotix/_collisions.py:99-102; a custom call is opaque to jax, so the VJP is supplied here).

This is what replaces the jitted `f(env, key)` of examples/test_viz.py:24-48 in a JAX program:

    from parallax_b200 import jax_ffi as cx
    cx.register()                                   # builds libcotix_xla_ffi.so on first use
    P = cx.default_params(judge_enabled=1, auto_reset=1)
    state = cx.reset(keys, P)                       # u32[B, 2] -> f32[36, B]
    state, reward, done, obs = jax.jit(lambda s, a: cx.step(s, a, P))(state, action)
    g_state0, g_actions = jax.grad(lambda s, a: jnp.sum(cx.rollout(s, a, P0)[1] ** 2), (0, 1))(s0, acts)

No torch anywhere on this path.  jax is imported lazily: without it (this image has neither jax
nor jaxlib, and there is no network to install them) `available()` is False, `register()` raises,
and everything else in the package works as before.  The C++ side is csrc/cotix_xla_ffi.cc
(compiled against jaxlib's XLA FFI headers, `jax.ffi.include_dir()`); tests/xla_ffi_mock holds a
minimal stand-in for those headers so that the handlers are compiled and EXECUTED on the GPU by
the test-suite even here (tests/test_xla_ffi_shim.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import _lib
from ._lib import LunarParams

_CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
SHIM_SOURCE = os.path.join(_CSRC, "cotix_xla_ffi.cc")
SHIM_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libcotix_xla_ffi.so")
TARGETS = {  # jax target name -> handler symbol in the shim
    "cotix_lunar_reset": "CotixLunarReset",
    "cotix_lunar_step": "CotixLunarStep",
    "cotix_lunar_steps": "CotixLunarSteps",
    "cotix_narrowphase": "CotixNarrowphase",
    "cotix_lunar_rollout": "CotixLunarRollout",
    "cotix_lunar_rollout_bwd": "CotixLunarRolloutBwd",
}
NF, NOBS, NDYN, REC = 36, 8, 18, 16
_registered = False


def available() -> bool:
    """True when jax (with jax.ffi) can be imported in this process."""
    try:
        import jax.ffi  # noqa: F401

        return True
    except Exception:
        return False


def _cuda_home() -> str:
    for cand in (os.environ.get("CUDA_HOME"), "/usr/local/cuda"):
        if cand and os.path.isdir(cand):
            return cand
    return "/usr/local/cuda"


def build_shim(include_dir: str | None = None, out: str | None = None, extra: list | None = None) -> str:
    """Compile csrc/cotix_xla_ffi.cc into libcotix_xla_ffi.so next to libcotix_b200.so.
    include_dir: where xla/ffi/api/ffi.h lives (default: jax.ffi.include_dir()); without any the
    file compiles to its stub (cotix_xla_ffi_available() == 0)."""
    out = out or SHIM_PATH
    if include_dir is None and available():
        import jax.ffi

        include_dir = jax.ffi.include_dir()
    _lib.build()
    cuda = _cuda_home()
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", SHIM_SOURCE, "-o", out, f"-I{cuda}/include"]
    if include_dir:
        cmd.append(f"-I{include_dir}")
    cmd += list(extra or [])
    lib_dir = os.path.dirname(_lib.LIB_PATH)
    cmd += [f"-L{lib_dir}", "-l:libcotix_b200.so", f"-L{cuda}/lib64", "-lcudart", "-Wl,-rpath,$ORIGIN",
            f"-Wl,-rpath,{lib_dir}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise _lib.CotixError("building the XLA FFI shim failed:\n" + res.stdout + res.stderr)
    return out


def register() -> None:
    """Build (if needed), load and register the handlers with jax (platform "CUDA")."""
    global _registered
    if _registered:
        return
    if not available():
        raise _lib.CotixError("jax is not importable in this process: the jax.ffi boundary needs jax + jaxlib "
                              "(there is no fallback; the ctypes mirror in parallax_b200.lunar_lander works without it)")
    import jax.ffi

    if not os.path.exists(SHIM_PATH) or os.path.getmtime(SHIM_PATH) < os.path.getmtime(SHIM_SOURCE):
        build_shim()
    so = C.CDLL(SHIM_PATH)
    if not so.cotix_xla_ffi_available():
        raise _lib.CotixError("libcotix_xla_ffi.so was built without the XLA FFI headers; rebuild with build_shim()")
    for target, symbol in TARGETS.items():
        jax.ffi.register_ffi_target(target, jax.ffi.pycapsule(getattr(so, symbol)), platform="CUDA")
    _registered = True


def default_params(**overrides) -> LunarParams:
    from . import lunar_lander as ll

    return ll.default_params(**overrides)


def params_words(P: LunarParams) -> np.ndarray:
    """The CotixLunarParams struct as int32 words: the `params` attribute of every handler."""
    return np.frombuffer(bytes(P), dtype=np.int32).copy()


def workspace_bytes(B: int) -> int:
    return int(_lib.lib().cotix_lunar_workspace_bytes(int(B)))


def _workspace(B: int, pipeline: bool | None, multi_step: bool = False):
    """The pipeline's scratch as an XLA-owned u8 operand (the handlers never allocate); a zero-sized
    operand selects the fused kernel (one step) or the block-resident rollout kernel (several steps, small
    batch).  XLA materialises the zeros once under jit."""
    import jax.numpy as jnp

    use = (B > 4736 if multi_step else B >= 16384) if pipeline is None else pipeline
    return jnp.zeros((workspace_bytes(B) + 256 if use else 0,), jnp.uint8)


def reset(keys, P: LunarParams | None = None):
    """keys u32[B, 2] (jax.random key data) -> state f32[36, B]; `LunarLander(key)` for a batch."""
    import jax
    import jax.numpy as jnp

    register()
    P = P or default_params()
    B = keys.shape[0]
    return jax.ffi.ffi_call("cotix_lunar_reset", jax.ShapeDtypeStruct((NF, B), jnp.float32))(
        keys.astype(jnp.uint32), params=params_words(P))


def step(state, action, P: LunarParams | None = None, pipeline: bool | None = None):
    """state f32[36, B], action f32[2, B] (or None) -> (state', reward f32[B], done u8[B], obs f32[8, B])."""
    import jax
    import jax.numpy as jnp

    register()
    P = P or default_params()
    B = state.shape[1]
    if action is None:
        action = jnp.zeros((0, 2, B), jnp.float32)
    out = (jax.ShapeDtypeStruct((NF, B), jnp.float32), jax.ShapeDtypeStruct((B,), jnp.float32),
           jax.ShapeDtypeStruct((B,), jnp.uint8), jax.ShapeDtypeStruct((NOBS, B), jnp.float32))
    return jax.ffi.ffi_call("cotix_lunar_step", out)(state, action, _workspace(B, pipeline), params=params_words(P))


def steps(state, actions, n_steps: int, P: LunarParams | None = None, pipeline: bool | None = None):
    """n_steps steps in ONE custom call (the loop of examples/test_viz.py:46-48): actions f32[n, 2, B]
    (step t uses actions[t % n]) or None -> (state', reward, done, obs) of the last step.  The state
    operand is aliased with the state result, so under jit with a donated argument it runs in place."""
    import jax
    import jax.numpy as jnp

    register()
    P = P or default_params()
    B = state.shape[1]
    if actions is None:
        actions = jnp.zeros((0, 2, B), jnp.float32)
    out = (jax.ShapeDtypeStruct((NF, B), jnp.float32), jax.ShapeDtypeStruct((B,), jnp.float32),
           jax.ShapeDtypeStruct((B,), jnp.uint8), jax.ShapeDtypeStruct((NOBS, B), jnp.float32))
    return jax.ffi.ffi_call("cotix_lunar_steps", out, input_output_aliases={0: 0})(
        state, actions, _workspace(B, pipeline, True), params=params_words(P), n_steps=np.int32(n_steps))


def narrowphase(a, b, wavefront: bool | None = None):
    """a, b f32[N, 16] shape records -> (exists u8[N], penetration f32[N, 2], contact point f32[N, 2],
    status u8[N]): the six contact functions of cotix/_contacts.py:30-315, dispatched per pair."""
    import jax
    import jax.numpy as jnp

    register()
    n = a.shape[0]
    npar = _lib.default_narrow_params()
    use = (n >= 65536) if wavefront is None else wavefront
    ws = jnp.zeros((int(_lib.lib().cotix_narrowphase_workspace_bytes(n)) + 256 if use else 0,), jnp.uint8)
    out = (jax.ShapeDtypeStruct((n,), jnp.uint8), jax.ShapeDtypeStruct((n, 2), jnp.float32),
           jax.ShapeDtypeStruct((n, 2), jnp.float32), jax.ShapeDtypeStruct((n,), jnp.uint8))
    return jax.ffi.ffi_call("cotix_narrowphase", out)(
        a, b, ws, init_dir=np.asarray([npar.init_dir[0], npar.init_dir[1]], np.float32),
        gjk_max_iters=np.int32(npar.gjk_max_iters), epa_iters_circle_poly=np.int32(npar.epa_iters_circle_poly))


def _adjoint_workspace(B: int, T: int, jacobians: bool):
    import jax.numpy as jnp

    L = _lib.lib()
    size = L.cotix_lunar_adjoint_workspace_bytes if jacobians else L.cotix_lunar_adjoint_decisions_bytes
    return jnp.zeros((int(size(B, T)) + 256,), jnp.uint8)


def make_rollout(P: LunarParams | None = None, pipeline: bool | None = None, jacobians: bool | None = None):
    """-> rollout(state0 f32[36, B], actions f32[T, 2, B]) -> final state f32[36, B], differentiable in
    both arguments through `jax.custom_vjp`: the forward keeps the tape (`cotix_lunar_rollout`), the
    backward is the adjoint kernel (`cotix_lunar_rollout_bwd`).  Requires substeps == 1, auto_reset == 0."""
    import jax
    import jax.numpy as jnp

    register()
    P = (P or default_params()).copy()
    words = params_words(P)

    def fwd(state0, actions):
        T, _, B = actions.shape
        tape = jax.ffi.ffi_call("cotix_lunar_rollout", jax.ShapeDtypeStruct((T + 1, NF, B), jnp.float32))(
            state0, actions, _workspace(B, pipeline, True), params=words)
        return tape[-1], (tape, actions)

    def bwd(res, g_final):  # g_final f32[36, B]: rows 0..17 (the dynamic leaves) carry a cotangent
        tape, actions = res
        T, _, B = actions.shape
        jac = (B <= 16384) if jacobians is None else bool(jacobians)
        g0, g_act, _flags = jax.ffi.ffi_call(
            "cotix_lunar_rollout_bwd",
            (jax.ShapeDtypeStruct((NDYN, B), jnp.float32), jax.ShapeDtypeStruct((T, 2, B), jnp.float32),
             jax.ShapeDtypeStruct((B,), jnp.uint32)))(
            tape, actions, g_final[:NDYN], _adjoint_workspace(B, T, jac), params=words, jacobians=np.int32(jac))
        return jnp.zeros((NF, B), jnp.float32).at[:NDYN].set(g0), g_act

    @jax.custom_vjp
    def rollout(state0, actions):
        return fwd(state0, actions)[0]

    rollout.defvjp(fwd, bwd)
    return rollout


def rollout(state0, actions, P: LunarParams | None = None):
    """Convenience: `make_rollout(P)(state0, actions)`."""
    return make_rollout(P)(state0, actions)
