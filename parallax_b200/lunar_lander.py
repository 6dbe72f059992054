"""Host-side mirror of the reference's LunarLander step for a batch of B environments.

Reference surface this stands in for:
  * `LunarLander(key)`                      cotix/_lunar_lander.py:29-143  -> `reset(keys)`
  * `f(env, key)` (the composed step)       examples/test_viz.py:25-44     -> `step(state, action)`
    = ExplicitEulerPhysics.step + gravity + RandomizedCollider.resolve + key split +
      LunarLander.step, hand-written CUDA behind `cotix_lunar_step{,_ws}`.
  * `for _ in range(n): env = f(env, key)`  examples/test_viz.py:46-48     -> `steps(...)`, `StepsGraph`
  * `jax.vjp` / `jax.grad` of that step                                    -> `step_vjp`, `rollout_vjp`
State is a `LunarState`: one [36, B] float32 device array, struct-of-arrays over the env axis
(field map in include/cotix_b200.h), with converters to the reference's per-body leaves.

Device memory comes from one of two backends (`_xp`): torch tensors, or `devmem.DeviceArray`
(ctypes over the library's own cudaMalloc wrappers) — the latter needs no torch at all, which is
what a JAX host program wants (`device="native"`).  torch is never imported unless the caller hands
in a torch tensor or a torch device string.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib, _xp
from ._lib import LunarParams

NF = 36
NOBS = 8
F_HEIGHT0, F_KNOT0, F_KEY0, F_KEY1, F_STEPS, F_STATUS, F_RETURN, F_SHAPING = 18, 26, 30, 31, 32, 33, 34, 35
RNG_THREEFRY, RNG_PHILOX = 0, 1
BODY_NAMES = ("lander", "right_leg", "left_leg")
LEAF_NAMES = ("position_x", "position_y", "velocity_x", "velocity_y", "angle", "angular_velocity")


@dataclass
class LunarState:
    """[36, B] float32 SoA state on the device (uint32 fields are bit-cast); `data` is a torch
    tensor or a `devmem.DeviceArray`."""

    data: object

    @property
    def batch(self) -> int:
        return self.data.shape[1]

    def body(self, i: int) -> dict:
        """DynamicBody leaves of body i (cotix/_bodies.py:189-206) as [B] / [B, 2] arrays (torch
        views for a torch state, host numpy copies for a native one)."""
        d = self.data
        if _xp.is_torch(d):
            import torch

            return {
                "position": torch.stack([d[i * 6 + 0], d[i * 6 + 1]], -1),
                "velocity": torch.stack([d[i * 6 + 2], d[i * 6 + 3]], -1),
                "angle": d[i * 6 + 4],
                "angular_velocity": d[i * 6 + 5],
            }
        h = d[i * 6:i * 6 + 6].numpy()
        return {"position": np.stack([h[0], h[1]], -1), "velocity": np.stack([h[2], h[3]], -1),
                "angle": h[4], "angular_velocity": h[5]}

    @property
    def key(self):
        if _xp.is_torch(self.data):
            import torch

            return self.data[F_KEY0:F_KEY1 + 1].view(torch.int32).T.contiguous()
        return np.ascontiguousarray(self.data[F_KEY0:F_KEY1 + 1].numpy().view(np.uint32).T)

    @property
    def steps(self):
        if _xp.is_torch(self.data):
            import torch

            return self.data[F_STEPS].view(torch.int32)
        return self.data[F_STEPS].view(np.uint32)

    @property
    def status(self):
        if _xp.is_torch(self.data):
            import torch

            return self.data[F_STATUS].view(torch.int32)
        return self.data[F_STATUS].view(np.uint32)

    def numpy(self) -> np.ndarray:
        return _xp.backend_of(self.data).to_numpy(self.data)

    @staticmethod
    def from_numpy(arr: np.ndarray, device="cuda") -> "LunarState":
        a = np.ascontiguousarray(arr, np.float32)
        assert a.shape[0] == NF
        return LunarState(_xp.backend_for(device).asarray(a, np.float32))

    def clone(self) -> "LunarState":
        return LunarState(self.data.clone())


def default_params(**overrides) -> LunarParams:
    """The reference's literals (SURVEY §5 'config'); keyword overrides set struct fields."""
    p = _lib.default_lunar_params()
    for k, v in overrides.items():
        cur = getattr(p, k)
        if hasattr(cur, "__len__"):
            for i, x in enumerate(v):
                cur[i] = x
        else:
            setattr(p, k, v)
    return p


def reset(keys, params: LunarParams | None = None, device="cuda") -> LunarState:
    """reset(key) for a batch: keys [B, 2] uint32 (jax PRNGKey words).  device: "cuda[:i]" / a
    torch.device (torch backend) or "native[:i]" / an int (torch-free DeviceArray backend)."""
    params = params or default_params()
    k = np.ascontiguousarray(np.asarray(keys).astype(np.uint32)).reshape(-1, 2)
    xp = _xp.backend_for(device)
    kd = xp.asarray(k, np.uint32)
    B = k.shape[0]
    state = xp.empty((NF, B), np.float32)
    with xp.device_ctx():
        _lib.check(_lib.lib().cotix_lunar_reset(xp.stream_ptr(), B, kd.data_ptr(), state.data_ptr(), B,
                                                C.byref(params)))
        if xp.name == "native":
            xp.synchronize()  # `kd` is dropped on return
    return LunarState(state)


MODE_AUTO, MODE_FUSED, MODE_PIPELINE = 0, 1, 2
PIPELINE_MIN_ENVS = 16384
_MODES = {"auto": MODE_AUTO, "fused": MODE_FUSED, "pipeline": MODE_PIPELINE}
_workspaces: dict = {}
_WORKSPACE_CACHE = 4


ROLLOUT_MAX_ENVS = 4736  # COTIX_ROLLOUT_MAX_ENVS: multi-step calls without a workspace run in one launch up to here


def _use_pipe(mode: str, B: int, n_steps: int = 1, substeps: int = 1) -> bool:
    """auto: one step -> the fused kernel below PIPELINE_MIN_ENVS; several steps in one call -> the
    block-resident rollout kernel (no workspace) up to ROLLOUT_MAX_ENVS, the pipeline beyond (measured,
    100-step contact-rich rollouts: 1024 envs 3.3 ms against 4.6, 4096 envs 3.3 against 5.1, 8192 envs
    7.6 against 6.0)."""
    m = _MODES[mode]
    if m != MODE_AUTO:
        return m == MODE_PIPELINE
    if n_steps > 1 and substeps <= 1:
        return B > ROLLOUT_MAX_ENVS
    return B >= PIPELINE_MIN_ENVS


def _workspace(xp, B: int):
    """Device scratch of the wavefront pipeline: a small LRU keyed by (backend, device, batch).  A
    cached block is only ever handed to launches on the caller's current stream of that device
    (stream order protects it); an evicted torch block stays alive in torch's caching allocator until
    the work enqueued on it has drained, an evicted native block is freed after a synchronise."""
    if not hasattr(xp, "name"):  # a torch.device (older call sites)
        xp = _xp.backend_for(xp)
    key = (xp.name, xp.device_index(), int(B))
    ws = _workspaces.pop(key, None)
    if ws is None:
        nbytes = int(_lib.lib().cotix_lunar_workspace_bytes(B))
        with xp.device_ctx():
            ws = xp.empty((nbytes + 256,), np.uint8)
        while len(_workspaces) >= _WORKSPACE_CACHE:
            old_key = next(iter(_workspaces))
            if old_key[0] == "native":
                xp.synchronize()
            _workspaces.pop(old_key)
    _workspaces[key] = ws  # most recently used last
    return ws


def _ws_args(xp, B: int):
    ws = _workspace(xp, B)
    wp = (ws.data_ptr() + 255) & ~255
    return wp, ws.numel() - (wp - ws.data_ptr())


def _ptr(x):
    return x.data_ptr() if x is not None else None


def step(state: LunarState, action=None, params: LunarParams | None = None,
         inplace: bool = False, want_obs: bool = False, mode: str = "auto"):
    """step(state, action) -> (state', reward[B], done[B]) (+ obs[8, B]).

    action: [2, B] float32 device array (main throttle, side throttle) or None for the
    reference's unactuated step.  Functional by default (a new state array), like the
    reference; `inplace=True` updates `state` (what a CUDA-graph rollout wants).
    mode: "fused" = one kernel (cotix_lunar_step), "pipeline" = the wavefront pipeline
    (cotix_lunar_step_ws), "auto" = pipeline for B >= 16384.
    Both give identical results.
    """
    params = params or default_params()
    d = state.data
    xp = _xp.backend_of(d)
    B = d.shape[1]
    ld = _xp.row_ld(d, "state")  # a column slice of a larger batch keeps the parent's row stride
    out = d if inplace else xp.empty((NF, B), np.float32, ld)
    reward = xp.empty((B,), np.float32)
    done = xp.empty((B,), np.uint8)
    obs = xp.empty((NOBS, B), np.float32, ld) if want_obs else None
    action = _xp.with_ld(xp, action, 2, B, ld, "action")
    with xp.device_ctx():
        args = (xp.stream_ptr(), B, d.data_ptr(), out.data_ptr(), ld, _ptr(action), reward.data_ptr(),
                done.data_ptr(), _ptr(obs), C.byref(params))
        if _use_pipe(mode, B):
            _lib.check(_lib.lib().cotix_lunar_step_ws(*args, *_ws_args(xp, B)))
        else:
            _lib.check(_lib.lib().cotix_lunar_step(*args))
    ns = state if inplace else LunarState(out)
    if want_obs:
        return ns, reward, done, obs
    return ns, reward, done


def _actions_ld(xp, actions, B: int, ld: int):
    """[n, 2, B] actions as a float32 array of this backend with strides (2 ld, ld, 1)."""
    if actions is None:
        return None, 0
    if isinstance(actions, _ReadyActions):
        return actions.a, actions.a.n
    actions = xp.asarray(actions, np.float32)
    assert actions.dim() == 3 and tuple(actions.shape[1:]) == (2, B), tuple(actions.shape)
    n = actions.shape[0]
    if not (actions.stride(2) == 1 and actions.stride(1) == ld and (n == 1 or actions.stride(0) == 2 * ld)):
        # one [2 n, B] buffer with row stride ld, seen as [n, 2, B]
        flat = xp.empty((2 * n, B), np.float32, ld)
        for k in range(n):
            flat[2 * k:2 * k + 2].copy_(actions[k])
        return _Actions(flat, n), n
    return actions, n


class _Actions:
    """[n, 2, B] view over a [2 n, B] buffer with row stride ld (keeps the buffer alive)."""

    def __init__(self, flat, n):
        self.flat, self.n = flat, n

    def data_ptr(self):
        return self.flat.data_ptr()


def steps(state: LunarState, actions, n_steps: int, params: LunarParams | None = None,
          want_obs: bool = False, mode: str = "auto", out: tuple | None = None):
    """n_steps steps IN PLACE on `state` in one enqueue (`cotix_lunar_steps_ws`): the loop
    `for _ in range(n): env = f(env, key)` of examples/test_viz.py:46-48.  actions: [n_actions, 2, B]
    (step t uses actions[t % n_actions]) or None.  Returns (reward, done[, obs]) of the last step.
    In pipeline mode the env ranges run their n_steps on their own streams without joining between
    steps; `StepsGraph` captures the call so that a whole rollout is one graph launch.
    out: optional preallocated (reward [B] f32, done [B] u8, obs [8, B] f32 or None)."""
    params = params or default_params()
    d = state.data
    xp = _xp.backend_of(d)
    B = d.shape[1]
    ld = _xp.row_ld(d, "state")
    if out is not None:
        reward, done, obs = out
        assert obs is None or _xp.row_ld(obs, "obs") == ld
    else:
        reward = xp.empty((B,), np.float32)
        done = xp.empty((B,), np.uint8)
        obs = xp.empty((NOBS, B), np.float32, ld) if want_obs else None
    actions, n_act = _actions_ld(xp, actions, B, ld)
    with xp.device_ctx():
        wp, wn = _ws_args(xp, B) if _use_pipe(mode, B, int(n_steps), params.substeps) else (None, 0)
        _lib.check(_lib.lib().cotix_lunar_steps_ws(
            xp.stream_ptr(), B, d.data_ptr(), ld, _ptr(actions), n_act, int(n_steps), reward.data_ptr(),
            done.data_ptr(), _ptr(obs), C.byref(params), wp, wn))
        if xp.name == "native" and isinstance(actions, _Actions):
            xp.synchronize()  # the strided copy of the actions is dropped on return
    if want_obs or (out is not None and obs is not None):
        return reward, done, obs
    return reward, done


class Graph:
    """CUDA-graph capture of device entry points (`cotix_graph_*`): what is enqueued through this
    module inside the `with` block becomes one replayable graph — the single dispatch per rollout
    that `jax.jit` gives the reference's loop (examples/test_viz.py:24,46-48).  Nothing may be
    allocated while capturing: pass preallocated buffers (`steps(..., out=...)`) and `keep()` every
    buffer the captured calls touch.  `StepsGraph` does all of that for a rollout.
    device: as for `reset` (torch or native backend)."""

    def __init__(self, device="cuda"):
        self.xp = device if hasattr(device, "stream_ptr") else _xp.backend_for(device)
        self._h = C.c_void_p()
        self._keep = []

    def __enter__(self):
        xp = self.xp
        xp.synchronize()
        if xp.name == "torch":
            t = xp.torch
            self._stream = t.cuda.Stream(xp.device)
            self._ctx = t.cuda.stream(self._stream)
            self._sptr = self._stream.cuda_stream
        else:
            from . import devmem

            self._stream = devmem.Stream(xp.device)
            self._ctx = self._stream
            self._sptr = self._stream.ptr
        self._ctx.__enter__()
        with xp.device_ctx():
            _lib.check(_lib.lib().cotix_graph_begin_capture(self._sptr))
        return self

    def __exit__(self, et, ev, tb):
        rc = _lib.lib().cotix_graph_end_capture(self._sptr, C.byref(self._h))
        self._ctx.__exit__(et, ev, tb)
        if et is None:
            _lib.check(rc)
        return False

    def keep(self, *objs):
        """Hold references to buffers a captured call reads or writes."""
        self._keep.extend(objs)

    def launch(self, stream: int | None = None):
        with self.xp.device_ctx():
            st = self.xp.stream_ptr() if stream is None else stream
            _lib.check(_lib.lib().cotix_graph_launch(self._h, st))

    def info(self) -> dict:
        n, k = C.c_int32(), C.c_int64()
        _lib.check(_lib.lib().cotix_graph_info(self._h, C.byref(n), C.byref(k)))
        return {"nodes": n.value, "kernel_launches": k.value}

    def close(self):
        if self._h:
            _lib.lib().cotix_graph_destroy(self._h)
            self._h = C.c_void_p()
        self._keep.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class StepsGraph:
    """`steps(state, actions, n_steps)` captured once and replayed: every `launch()` advances
    `state` by n_steps steps (in place) with ONE graph launch; `reward`, `done` (and `obs`) hold the
    last step's values.  Owns the buffers the graph touches, including the pipeline workspace."""

    def __init__(self, state: LunarState, actions, n_steps: int,
                 params: LunarParams | None = None, want_obs: bool = False, mode: str = "auto"):
        self.params = (params or default_params()).copy()
        d = state.data
        xp = _xp.backend_of(d)
        B = d.shape[1]
        ld = _xp.row_ld(d, "state")
        self.state, self.n_steps = state, int(n_steps)
        with xp.device_ctx():
            self.reward = xp.empty((B,), np.float32)
            self.done = xp.empty((B,), np.uint8)
            self.obs = xp.empty((NOBS, B), np.float32, ld) if want_obs else None
            self.actions, _ = _actions_ld(xp, actions, B, ld)
            ws = _workspace(xp, B) if _use_pipe(mode, B, self.n_steps, self.params.substeps) else None  # before the capture
        self.graph = Graph(xp)
        with self.graph:
            steps(state, self._actions_view(xp, B, ld), self.n_steps, self.params, mode=mode,
                  out=(self.reward, self.done, self.obs))
        self.graph.keep(d, self.actions, self.reward, self.done, self.obs, ws)

    def _actions_view(self, xp, B, ld):
        a = self.actions
        if isinstance(a, _Actions):  # already in the layout the call wants: hand `steps` a ready view
            return _ReadyActions(a)
        return a

    def launch(self):
        self.graph.launch()

    def info(self) -> dict:
        return self.graph.info()

    def close(self):
        self.graph.close()


class _ReadyActions:
    """Marks actions that are already laid out (strides 2 ld, ld, 1) so that `steps` does not copy."""

    def __init__(self, a: _Actions):
        self.a = a


def pipeline_config(split: int = -1, pdl: int = -1, priorities: int = -1, item_capacity: int = -1) -> None:
    """Launch configuration of the wavefront pipeline (`cotix_lunar_pipeline_config`; -1 = unchanged):
    split = ranges of envs run on their own streams (0 automatic), pdl, priorities, item_capacity
    (test hook for the item-queue overflow branch)."""
    _lib.check(_lib.lib().cotix_lunar_pipeline_config(int(split), int(pdl), int(priorities), int(item_capacity)))


NDYN = 18
JACOBIAN_MAX_ENVS = 16384
ADJ_REPLAY_MISMATCH, ADJ_NONFINITE, ADJ_CONTACT_SHIFT = 0x1, 0x2, 8


def step_vjp(state: LunarState, action, g_next, params: LunarParams | None = None,
             state_next: LunarState | None = None, flags=None):
    """Vector-Jacobian product of one step: what `jax.vjp(f, env)[1](g_next)` gives for the
    reference's composed step (examples/test_viz.py:25-44), computed by the hand-written adjoint
    kernel (`cotix_lunar_step_bwd`).

    state: the state BEFORE the step; g_next: [18, B] cotangent of the next state's dynamic
    leaves (row 6*body + leaf).  Returns (g_state [18, B], g_action [2, B]).  `state_next` (the
    forward result) turns on the bit-for-bit replay check; `flags` ([B] int32, zeroed by the
    caller) collects the ADJ_* bits.
    """
    params = params or default_params()
    d = state.data
    xp = _xp.backend_of(d)
    B = d.shape[1]
    ld = _xp.row_ld(d, "state")  # every [.][ld] buffer of the call shares the state's row stride
    g_next = _xp.with_ld(xp, g_next, NDYN, B, ld, "g_next")
    g_in = xp.empty((NDYN, B), np.float32, ld)
    g_act = xp.zeros((2, B), np.float32, ld)
    action = _xp.with_ld(xp, action, 2, B, ld, "action")
    sn = _xp.with_ld(xp, state_next.data, NF, B, ld, "state_next") if state_next is not None else None
    with xp.device_ctx():
        _lib.check(_lib.lib().cotix_lunar_step_bwd(
            xp.stream_ptr(), B, d.data_ptr(), _ptr(sn), ld, _ptr(action), g_next.data_ptr(), g_in.data_ptr(),
            g_act.data_ptr(), _ptr(flags), C.byref(params)))
        if xp.name == "native":
            xp.synchronize()
    return g_in, g_act


@dataclass
class Rollout:
    """Tape of a differentiable rollout: states [T+1, 36, B] (slot 0 = initial state), the
    actions [T, 2, B] it was driven with (or None) and the params."""

    tape: object
    actions: object
    params: LunarParams

    @property
    def final(self) -> LunarState:
        return LunarState(self.tape[-1])

    def state(self, t: int) -> LunarState:
        return LunarState(self.tape[t])


def rollout(state: LunarState, actions, n_steps: int, params: LunarParams | None = None, mode: str = "auto",
            tape=None) -> Rollout:
    """n_steps steps of `step`, keeping every intermediate state (`cotix_lunar_rollout`): the
    forward half of the reference's `jax.value_and_grad` over a rollout.  tape: optional
    preallocated contiguous [n_steps + 1, 36, B] buffer (what a graph capture needs)."""
    params = params or default_params()
    d = state.data
    xp = _xp.backend_of(d)
    B = d.shape[1]
    T = int(n_steps)
    if tape is None:
        tape = xp.empty((T + 1, NF, B), np.float32)
    assert tuple(tape.shape) == (T + 1, NF, B) and tape.is_contiguous()
    tape[0].copy_(d)
    if actions is not None:
        actions = xp.asarray(actions, np.float32).contiguous()
        assert tuple(actions.shape) == (T, 2, B)
    with xp.device_ctx():
        wp, wn = _ws_args(xp, B) if _use_pipe(mode, B, T, params.substeps) else (None, 0)
        _lib.check(_lib.lib().cotix_lunar_rollout(
            xp.stream_ptr(), B, T, tape.data_ptr(), B, _ptr(actions), C.byref(params), wp, wn))
    return Rollout(tape, actions, params)


def adjoint_workspace(B: int, T: int, device, decision_tape: bool | int = True, jacobians: bool | None = None):
    """Scratch of `rollout_vjp` (decision tapes, and with `jacobians` the transposed Jacobians) for a
    rollout of T steps of B envs: a uint8 array to pass as `workspace=`, or None."""
    if jacobians is None:
        jacobians = B <= JACOBIAN_MAX_ENVS
    if not decision_tape or T <= 0:
        return None
    L = _lib.lib()
    n = T if decision_tape is True else max(1, min(T, int(decision_tape)))
    size = L.cotix_lunar_adjoint_workspace_bytes if jacobians else L.cotix_lunar_adjoint_decisions_bytes
    xp = device if hasattr(device, "stream_ptr") else _xp.backend_for(device)
    with xp.device_ctx():
        return xp.empty((int(size(B, n)) + 256,), np.uint8)


def rollout_vjp(ro: Rollout, g_final, flags=None, decision_tape: bool | int = True,
                jacobians: bool | None = None, workspace=None, out: tuple | None = None):
    """Backward pass over a `Rollout` (`cotix_lunar_rollout_bwd`): g_final [18, B] = dL/d(final
    dynamic leaves).  Returns (dL/d(initial dynamic leaves) [18, B], dL/d(actions) [T, 2, B]).

    decision_tape: True = scratch for the whole rollout, an int = for that many steps at a time:
    the collider's decisions — and with `jacobians` the transposed Jacobians of the env-steps with
    contacts — are computed in batched launches over (step, env) and the T reverse launches are
    light; False = every reverse step replays its own decisions.  Without `jacobians` the results
    are bit-identical to False; with them they differ at rounding level.  jacobians=None picks them
    for B <= JACOBIAN_MAX_ENVS: they trade 18 x the contact-adjoint work, done at full occupancy, for
    short reverse launches — a win while the sweep is latency-bound (B = 4096: 9.0 -> 5.7 ms), a loss
    once it is throughput-bound (B = 65536: 23.7 -> 43 ms).
    workspace: from `adjoint_workspace` (allocated here when None); out: optional preallocated
    (g [18, B], g_actions [T, 2, B]) — g is overwritten with g_final first.  With both given the call
    allocates nothing and can be captured in a `Graph`."""
    tape = ro.tape
    xp = _xp.backend_of(tape)
    T, B = tape.shape[0] - 1, tape.shape[2]
    if jacobians is None:
        jacobians = B <= JACOBIAN_MAX_ENVS
    if out is not None:
        g, g_act = out
        assert g.is_contiguous() and g_act.is_contiguous() and tuple(g_act.shape) == (T, 2, B)
        if g.data_ptr() != g_final.data_ptr():
            g.copy_(g_final)
        g_act.zero_()
    else:
        g = xp.empty((NDYN, B), np.float32)
        g.copy_(xp.asarray(g_final, np.float32))
        g_act = xp.zeros((T, 2, B), np.float32)
    assert tuple(g.shape) == (NDYN, B)
    wp, wn, ws = None, 0, workspace
    if ws is None:
        ws = adjoint_workspace(B, T, xp, decision_tape, jacobians)
    if ws is not None:
        wp = (ws.data_ptr() + 255) & ~255
        wn = ws.numel() - (wp - ws.data_ptr())
    with xp.device_ctx():
        _lib.check(_lib.lib().cotix_lunar_rollout_bwd(
            xp.stream_ptr(), B, T, tape.data_ptr(), B, _ptr(ro.actions), g.data_ptr(), g_act.data_ptr(),
            _ptr(flags), C.byref(ro.params), wp, wn, 1 if jacobians else 0))
        if xp.name == "native" and workspace is None:
            xp.synchronize()  # the scratch is dropped on return
    return g, g_act


PIPELINE_STAGES = ("integrate", "narrow", "contact", "finish", "reset", "step")


def pipeline_profile(enable: bool) -> None:
    """Start / stop per-stage CUDA-event timing of the pipeline (measurement aid)."""
    _lib.check(_lib.lib().cotix_lunar_pipeline_profile(1 if enable else 0))


def pipeline_stage_ms() -> tuple[dict, int]:
    """({stage: average ms}, steps averaged) since profiling was enabled."""
    out = (C.c_double * len(PIPELINE_STAGES))()
    n = int(_lib.lib().cotix_lunar_pipeline_stage_ms(out))
    return {k: float(out[i]) for i, k in enumerate(PIPELINE_STAGES)}, n


def pipeline_counters(device=None, B: int | None = None) -> dict | None:
    """Queue lengths of the last pipeline step on the cached workspace of (device, B) — the most
    recently used one when B is None — summed over the env ranges (measurement aid): collision
    items, live contacts, envs with a candidate contact, auto-resets, envs that overflowed the queue."""
    want = None
    if device is not None:
        want = device if isinstance(device, (int, np.integer)) else getattr(device, "index", None)
    for (name, idx, b), ws in reversed(list(_workspaces.items())):
        if (want is None or want == idx) and (B is None or B == b):
            xp = _xp.backend_of(ws)
            wp = (ws.data_ptr() + 255) & ~255
            out = (C.c_int64 * 5)()
            with xp.device_ctx():
                _lib.check(_lib.lib().cotix_lunar_pipeline_counters(xp.stream_ptr(), b, wp, out))
            return {"items": out[0], "hits": out[1], "contact_envs": out[2], "resets": out[3], "overflow": out[4]}
    return None


class LunarLanderEnv:
    """Vector env with HOST-facing reset/step (the `e2e` path of bench.py).

    Owns the device state of B envs through the C handle (`cotix_lunar_env_*`): each
    `step(action_host)` is H2D(action) -> kernels -> D2H(obs, reward, done), overlapped over
    ranges of envs (`set_chunks`).  Host result buffers are page-locked (through torch when it is
    importable, else through `cotix_host_malloc`) and exposed as numpy arrays / torch tensors.
    """

    def __init__(self, num_envs: int, params: LunarParams | None = None, device: int = 0):
        self.B = int(num_envs)
        self.params = params or default_params()
        self.device = device
        self._h = C.c_void_p()
        _lib.check(_lib.lib().cotix_lunar_env_create(C.byref(self._h), device, self.B, C.byref(self.params)))
        if _xp.torch_available():
            import torch

            self.obs = torch.empty((NOBS, self.B), dtype=torch.float32).pin_memory()
            self.reward = torch.empty(self.B, dtype=torch.float32).pin_memory()
            self.done = torch.empty(self.B, dtype=torch.uint8).pin_memory()
            self._ptrs = (self.obs.data_ptr(), self.reward.data_ptr(), self.done.data_ptr())
        else:
            from . import devmem

            self._pinned = (devmem.PinnedArray((NOBS, self.B), np.float32), devmem.PinnedArray((self.B,), np.float32),
                            devmem.PinnedArray((self.B,), np.uint8))
            self.obs, self.reward, self.done = (p.array for p in self._pinned)
            self._ptrs = tuple(p.ptr for p in self._pinned)

    def close(self):
        if self._h:
            _lib.lib().cotix_lunar_env_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_mode(self, mode: str) -> None:
        """"auto" | "fused" | "pipeline" (see `step`)."""
        _lib.check(_lib.lib().cotix_lunar_env_set_mode(self._h, _MODES[mode]))

    def set_chunks(self, chunks: int) -> None:
        """Env ranges a host step is software-pipelined over (0 = automatic, 1 = no overlap)."""
        _lib.check(_lib.lib().cotix_lunar_env_set_chunks(self._h, int(chunks)))

    def set_host_graphs(self, enable: bool) -> None:
        """Replay the chunked host step as one CUDA graph when all host buffers are pinned."""
        _lib.check(_lib.lib().cotix_lunar_env_set_host_graphs(self._h, 1 if enable else 0))

    def reset(self, keys) -> None:
        k = np.ascontiguousarray(np.asarray(keys).astype(np.uint32)).reshape(self.B, 2)
        _lib.check(_lib.lib().cotix_lunar_env_reset_host(self._h, k.ctypes.data))

    def step(self, action_host, want_obs: bool = True):
        """action_host: [2, B] float32 host array — numpy or a (pinned, for best speed) torch tensor.
        want_obs=False leaves the observation on the device (86 % of the step's D2H bytes):
        (None, reward, done)."""
        ptr = None
        if action_host is not None:
            if isinstance(action_host, np.ndarray):
                action_host = np.ascontiguousarray(action_host, np.float32)
                assert action_host.shape == (2, self.B)
                ptr = action_host.ctypes.data
            else:
                assert tuple(action_host.shape) == (2, self.B) and action_host.is_contiguous()
                ptr = action_host.data_ptr()
        _lib.check(_lib.lib().cotix_lunar_env_step_host(self._h, ptr, self._ptrs[0] if want_obs else None,
                                                        self._ptrs[1], self._ptrs[2]))
        return (self.obs if want_obs else None), self.reward, self.done

    def get_state(self) -> np.ndarray:
        out = np.empty((NF, self.B), np.float32)
        _lib.check(_lib.lib().cotix_lunar_env_get_state_host(self._h, out.ctypes.data))
        return out

    def set_state(self, state: np.ndarray) -> None:
        s = np.ascontiguousarray(state, np.float32)
        assert s.shape == (NF, self.B)
        _lib.check(_lib.lib().cotix_lunar_env_set_state_host(self._h, s.ctypes.data))

    def rollout_graph(self, n_steps: int, action_dev=None) -> None:
        """Enqueue n_steps steps as one replayed CUDA graph on the handle's stream."""
        _lib.check(_lib.lib().cotix_lunar_env_rollout_graph(self._h, int(n_steps), _ptr(action_dev)))

    def synchronize(self) -> None:
        st = _lib.lib().cotix_lunar_env_stream(self._h)
        _lib.check(_lib.lib().cotix_stream_synchronize(st))
