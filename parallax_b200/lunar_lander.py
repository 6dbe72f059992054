"""Host-side mirror of the reference's LunarLander step for a batch of B environments.

Reference surface this stands in for:
  * `LunarLander(key)`                      cotix/_lunar_lander.py:29-143  -> `reset(keys)`
  * `f(env, key)` (the composed step)       examples/test_viz.py:25-44     -> `step(state, action)`
    = ExplicitEulerPhysics.step + gravity + RandomizedCollider.resolve + key split +
      LunarLander.step, all fused into one CUDA kernel behind `cotix_lunar_step`.
State is a `LunarState`: one [36, B] float32 device tensor, struct-of-arrays over the env axis
(field map in include/cotix_b200.h), with converters to the reference's per-body leaves.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import LunarParams

NF = 36
NOBS = 8
F_HEIGHT0, F_KNOT0, F_KEY0, F_KEY1, F_STEPS, F_STATUS, F_RETURN, F_SHAPING = 18, 26, 30, 31, 32, 33, 34, 35
RNG_THREEFRY, RNG_PHILOX = 0, 1
BODY_NAMES = ("lander", "right_leg", "left_leg")
LEAF_NAMES = ("position_x", "position_y", "velocity_x", "velocity_y", "angle", "angular_velocity")


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


@dataclass
class LunarState:
    """[36, B] float32 SoA state on the device (uint32 fields are bit-cast)."""

    data: torch.Tensor

    @property
    def batch(self) -> int:
        return self.data.shape[1]

    def body(self, i: int) -> dict:
        """DynamicBody leaves of body i (cotix/_bodies.py:189-206) as [B] / [B, 2] views."""
        d = self.data
        return {
            "position": torch.stack([d[i * 6 + 0], d[i * 6 + 1]], -1),
            "velocity": torch.stack([d[i * 6 + 2], d[i * 6 + 3]], -1),
            "angle": d[i * 6 + 4],
            "angular_velocity": d[i * 6 + 5],
        }

    @property
    def key(self) -> torch.Tensor:
        return self.data[F_KEY0:F_KEY1 + 1].view(torch.int32).T.contiguous()

    @property
    def steps(self) -> torch.Tensor:
        return self.data[F_STEPS].view(torch.int32)

    @property
    def status(self) -> torch.Tensor:
        return self.data[F_STATUS].view(torch.int32)

    def numpy(self) -> np.ndarray:
        return self.data.detach().cpu().numpy()

    @staticmethod
    def from_numpy(arr: np.ndarray, device="cuda") -> "LunarState":
        a = np.ascontiguousarray(arr, np.float32)
        assert a.shape[0] == NF
        return LunarState(torch.from_numpy(a).to(device))

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
    """reset(key) for a batch: keys [B, 2] uint32 (jax PRNGKey words)."""
    params = params or default_params()
    k = np.ascontiguousarray(np.asarray(keys).astype(np.uint32)).reshape(-1, 2)
    kd = torch.from_numpy(k.view(np.int32)).to(device)
    B = k.shape[0]
    state = torch.empty((NF, B), dtype=torch.float32, device=kd.device)
    with torch.cuda.device(kd.device):
        _lib.check(_lib.lib().cotix_lunar_reset(_stream_ptr(kd.device), B, kd.data_ptr(), state.data_ptr(),
                                                B, C.byref(params)))
    return LunarState(state)


MODE_AUTO, MODE_FUSED, MODE_PIPELINE = 0, 1, 2
PIPELINE_MIN_ENVS = 16384
_MODES = {"auto": MODE_AUTO, "fused": MODE_FUSED, "pipeline": MODE_PIPELINE}
_workspaces: "dict[tuple, torch.Tensor]" = {}
_WORKSPACE_CACHE = 4


def _workspace(device: torch.device, B: int) -> torch.Tensor:
    """Device scratch of the wavefront pipeline: a small LRU keyed by (device, batch).  A cached
    block is only ever handed to launches on the caller's current stream of that device (stream order
    protects it); eviction drops the reference and torch's caching allocator keeps the block alive
    until the work enqueued on it has drained."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), int(B))
    ws = _workspaces.pop(key, None)
    if ws is None:
        nbytes = int(_lib.lib().cotix_lunar_workspace_bytes(B))
        ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=device)
        while len(_workspaces) >= _WORKSPACE_CACHE:
            _workspaces.pop(next(iter(_workspaces)))
    _workspaces[key] = ws  # most recently used last
    return ws


def _ws_args(device: torch.device, B: int):
    ws = _workspace(device, B)
    wp = (ws.data_ptr() + 255) & ~255
    return wp, ws.numel() - (wp - ws.data_ptr())


def _row_ld(t: torch.Tensor, what: str) -> int:
    """Leading dimension of a [rows, B] buffer whose rows are contiguous (stride(1) == 1)."""
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{what}: rows must be contiguous (stride(1) == 1); call .contiguous()")
    return int(t.stride(0)) if t.shape[0] > 1 else max(int(t.stride(0)), int(t.shape[1]))


def _like_ld(rows: int, B: int, ld: int, ref: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """A [rows, B] buffer whose leading dimension is `ld` (the C-ABI takes ONE ld for every [.][ld]
    buffer of a call, so companions of a sliced state are allocated with the state's row stride)."""
    if ld == B:
        return torch.empty((rows, B), dtype=dtype, device=ref.device)
    return torch.empty_strided((rows, B), (ld, 1), dtype=dtype, device=ref.device)


def _with_ld(t: torch.Tensor | None, rows: int, B: int, ld: int, ref: torch.Tensor, what: str):
    if t is None:
        return None
    t = t.to(device=ref.device, dtype=torch.float32)
    assert tuple(t.shape) == (rows, B), f"{what}: expected shape {(rows, B)}, got {tuple(t.shape)}"
    if t.stride(1) == 1 and (rows == 1 or t.stride(0) == ld):
        return t
    out = _like_ld(rows, B, ld, ref)
    out.copy_(t)
    return out


def step(state: LunarState, action: torch.Tensor | None = None, params: LunarParams | None = None,
         inplace: bool = False, want_obs: bool = False, mode: str = "auto"):
    """step(state, action) -> (state', reward[B], done[B]) (+ obs[8, B]).

    action: [2, B] float32 device tensor (main throttle, side throttle) or None for the
    reference's unactuated step.  Functional by default (a new state tensor), like the
    reference; `inplace=True` updates `state` (what a CUDA-graph rollout wants).
    mode: "fused" = one kernel (cotix_lunar_step), "pipeline" = the wavefront pipeline
    (cotix_lunar_step_ws), "auto" = pipeline for B >= 16384.
    Both give identical results.
    """
    params = params or default_params()
    d = state.data
    B = d.shape[1]
    ld = _row_ld(d, "state")  # a column slice of a larger batch keeps the parent's row stride
    out = d if inplace else _like_ld(NF, B, ld, d)
    reward = torch.empty(B, dtype=torch.float32, device=d.device)
    done = torch.empty(B, dtype=torch.uint8, device=d.device)
    obs = _like_ld(NOBS, B, ld, d) if want_obs else None
    action = _with_ld(action, 2, B, ld, d, "action")
    m = _MODES[mode]
    use_pipe = m == MODE_PIPELINE or (m == MODE_AUTO and B >= PIPELINE_MIN_ENVS)
    with torch.cuda.device(d.device):
        args = (_stream_ptr(d.device), B, d.data_ptr(), out.data_ptr(), ld,
                action.data_ptr() if action is not None else None, reward.data_ptr(), done.data_ptr(),
                obs.data_ptr() if obs is not None else None, C.byref(params))
        if use_pipe:
            _lib.check(_lib.lib().cotix_lunar_step_ws(*args, *_ws_args(d.device, B)))
        else:
            _lib.check(_lib.lib().cotix_lunar_step(*args))
    ns = state if inplace else LunarState(out)
    if want_obs:
        return ns, reward, done, obs
    return ns, reward, done


def steps(state: LunarState, actions: torch.Tensor | None, n_steps: int, params: LunarParams | None = None,
          want_obs: bool = False, mode: str = "auto", out: tuple | None = None):
    """n_steps steps IN PLACE on `state` in one enqueue (`cotix_lunar_steps_ws`): the loop
    `for _ in range(n): env = f(env, key)` of examples/test_viz.py:46-48.  actions: [n_actions, 2, B]
    (step t uses actions[t % n_actions]) or None.  Returns (reward, done[, obs]) of the last step.
    In pipeline mode the env ranges run their n_steps on their own streams without joining between
    steps; capture the call in a `Graph` and a whole rollout is one launch."""
    params = params or default_params()
    d = state.data
    B = d.shape[1]
    ld = _row_ld(d, "state")
    if out is not None:
        reward, done, obs = out
        assert obs is None or _row_ld(obs, "obs") == ld
    else:
        reward = torch.empty(B, dtype=torch.float32, device=d.device)
        done = torch.empty(B, dtype=torch.uint8, device=d.device)
        obs = _like_ld(NOBS, B, ld, d) if want_obs else None
    n_act = 0
    if actions is not None:
        actions = actions.to(device=d.device, dtype=torch.float32)
        assert actions.dim() == 3 and tuple(actions.shape[1:]) == (2, B)
        n_act = actions.shape[0]
        if not (actions.stride(2) == 1 and actions.stride(1) == ld and actions.stride(0) == 2 * ld):
            a = torch.empty_strided((n_act, 2, B), (2 * ld, ld, 1), dtype=torch.float32, device=d.device)
            a.copy_(actions)
            actions = a
    m = _MODES[mode]
    use_pipe = m == MODE_PIPELINE or (m == MODE_AUTO and B >= PIPELINE_MIN_ENVS)
    with torch.cuda.device(d.device):
        wp, wn = _ws_args(d.device, B) if use_pipe else (None, 0)
        _lib.check(_lib.lib().cotix_lunar_steps_ws(
            _stream_ptr(d.device), B, d.data_ptr(), ld, actions.data_ptr() if actions is not None else None,
            n_act, int(n_steps), reward.data_ptr(), done.data_ptr(), obs.data_ptr() if obs is not None else None,
            C.byref(params), wp, wn))
    if want_obs or (out is not None and obs is not None):
        return reward, done, obs
    return reward, done


class Graph:
    """CUDA-graph capture of device entry points (`cotix_graph_*`): what is enqueued through this
    module inside the `with` block becomes one replayable graph — the single dispatch per rollout
    that `jax.jit` gives the reference's loop (examples/test_viz.py:24,46-48).  Nothing may be
    allocated while capturing: pass preallocated buffers (`steps(..., out=...)`) and `keep()` every
    buffer the captured calls touch.  `StepsGraph` does all of that for a rollout."""

    def __init__(self, device="cuda"):
        self.device = torch.device(device)
        self._h = C.c_void_p()
        self._keep = []

    def __enter__(self):
        torch.cuda.synchronize(self.device)
        self._stream = torch.cuda.Stream(self.device)
        self._ctx = torch.cuda.stream(self._stream)
        self._ctx.__enter__()
        _lib.check(_lib.lib().cotix_graph_begin_capture(self._stream.cuda_stream))
        return self

    def __exit__(self, et, ev, tb):
        rc = _lib.lib().cotix_graph_end_capture(self._stream.cuda_stream, C.byref(self._h))
        self._ctx.__exit__(et, ev, tb)
        if et is None:
            _lib.check(rc)
        return False

    def keep(self, *objs):
        """Hold references to buffers a captured call reads or writes."""
        self._keep.extend(objs)

    def launch(self, stream: int | None = None):
        with torch.cuda.device(self.device):
            st = _stream_ptr(self.device) if stream is None else stream
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

    def __init__(self, state: LunarState, actions: torch.Tensor | None, n_steps: int,
                 params: LunarParams | None = None, want_obs: bool = False, mode: str = "auto"):
        self.params = (params or default_params()).copy()
        d = state.data
        B = d.shape[1]
        ld = _row_ld(d, "state")
        self.state, self.n_steps = state, int(n_steps)
        self.reward = torch.empty(B, dtype=torch.float32, device=d.device)
        self.done = torch.empty(B, dtype=torch.uint8, device=d.device)
        self.obs = _like_ld(NOBS, B, ld, d) if want_obs else None
        if actions is not None:
            actions = actions.to(device=d.device, dtype=torch.float32)
            if not (actions.stride(2) == 1 and actions.stride(1) == ld and actions.stride(0) == 2 * ld):
                a = torch.empty_strided(tuple(actions.shape), (2 * ld, ld, 1), dtype=torch.float32, device=d.device)
                a.copy_(actions)
                actions = a
        self.actions = actions
        m = _MODES[mode]
        if m == MODE_PIPELINE or (m == MODE_AUTO and B >= PIPELINE_MIN_ENVS):
            ws = _workspace(d.device, B)  # allocated before the capture; the graph keeps it
        else:
            ws = None
        self.graph = Graph(d.device)
        with self.graph:
            steps(state, actions, self.n_steps, self.params, mode=mode, out=(self.reward, self.done, self.obs))
        self.graph.keep(d, actions, self.reward, self.done, self.obs, ws)

    def launch(self):
        self.graph.launch()

    def info(self) -> dict:
        return self.graph.info()

    def close(self):
        self.graph.close()


def pipeline_config(split: int = -1, pdl: int = -1, priorities: int = -1, item_capacity: int = -1) -> None:
    """Launch configuration of the wavefront pipeline (`cotix_lunar_pipeline_config`; -1 = unchanged):
    split = ranges of envs run on their own streams (0 automatic), pdl, priorities, item_capacity
    (test hook for the item-queue overflow branch)."""
    _lib.check(_lib.lib().cotix_lunar_pipeline_config(int(split), int(pdl), int(priorities), int(item_capacity)))


NDYN = 18
JACOBIAN_MAX_ENVS = 16384
ADJ_REPLAY_MISMATCH, ADJ_NONFINITE, ADJ_CONTACT_SHIFT = 0x1, 0x2, 8


def step_vjp(state: LunarState, action: torch.Tensor | None, g_next: torch.Tensor,
             params: LunarParams | None = None, state_next: LunarState | None = None,
             flags: torch.Tensor | None = None):
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
    B = d.shape[1]
    ld = _row_ld(d, "state")  # every [.][ld] buffer of the call shares the state's row stride
    g_next = _with_ld(g_next, NDYN, B, ld, d, "g_next")
    g_in = _like_ld(NDYN, B, ld, d)
    g_act = _like_ld(2, B, ld, d)
    g_act.zero_()
    action = _with_ld(action, 2, B, ld, d, "action")
    sn = None
    if state_next is not None:
        sn = _with_ld(state_next.data, NF, B, ld, d, "state_next")
    with torch.cuda.device(d.device):
        _lib.check(_lib.lib().cotix_lunar_step_bwd(
            _stream_ptr(d.device), B, d.data_ptr(), sn.data_ptr() if sn is not None else None,
            ld, action.data_ptr() if action is not None else None, g_next.data_ptr(), g_in.data_ptr(),
            g_act.data_ptr(), flags.data_ptr() if flags is not None else None, C.byref(params)))
    return g_in, g_act


@dataclass
class Rollout:
    """Tape of a differentiable rollout: states [T+1, 36, B] (slot 0 = initial state), the
    actions [T, 2, B] it was driven with (or None) and the params."""

    tape: torch.Tensor
    actions: torch.Tensor | None
    params: LunarParams

    @property
    def final(self) -> LunarState:
        return LunarState(self.tape[-1])

    def state(self, t: int) -> LunarState:
        return LunarState(self.tape[t])


def rollout(state: LunarState, actions: torch.Tensor | None, n_steps: int,
            params: LunarParams | None = None, mode: str = "auto", tape: torch.Tensor | None = None) -> Rollout:
    """n_steps steps of `step`, keeping every intermediate state (`cotix_lunar_rollout`): the
    forward half of the reference's `jax.value_and_grad` over a rollout.  tape: optional
    preallocated contiguous [n_steps + 1, 36, B] buffer (what a graph capture needs)."""
    params = params or default_params()
    d = state.data
    B = d.shape[1]
    T = int(n_steps)
    if tape is None:
        tape = torch.empty((T + 1, NF, B), dtype=torch.float32, device=d.device)
    assert tuple(tape.shape) == (T + 1, NF, B) and tape.is_contiguous()
    tape[0].copy_(d)
    if actions is not None:
        actions = actions.to(device=d.device, dtype=torch.float32).contiguous()
        assert actions.shape == (T, 2, B)
    m = _MODES[mode]
    use_pipe = m == MODE_PIPELINE or (m == MODE_AUTO and B >= PIPELINE_MIN_ENVS)
    with torch.cuda.device(d.device):
        wp, wn = _ws_args(d.device, B) if use_pipe else (None, 0)
        _lib.check(_lib.lib().cotix_lunar_rollout(
            _stream_ptr(d.device), B, T, tape.data_ptr(), B, actions.data_ptr() if actions is not None else None,
            C.byref(params), wp, wn))
    return Rollout(tape, actions, params)


def adjoint_workspace(B: int, T: int, device, decision_tape: bool | int = True, jacobians: bool | None = None):
    """Scratch of `rollout_vjp` (decision tapes, and with `jacobians` the transposed Jacobians) for a
    rollout of T steps of B envs: a uint8 tensor to pass as `workspace=`, or None."""
    if jacobians is None:
        jacobians = B <= JACOBIAN_MAX_ENVS
    if not decision_tape or T <= 0:
        return None
    L = _lib.lib()
    n = T if decision_tape is True else max(1, min(T, int(decision_tape)))
    size = L.cotix_lunar_adjoint_workspace_bytes if jacobians else L.cotix_lunar_adjoint_decisions_bytes
    return torch.empty(int(size(B, n)) + 256, dtype=torch.uint8, device=device)


def rollout_vjp(ro: Rollout, g_final: torch.Tensor, flags: torch.Tensor | None = None,
                decision_tape: bool | int = True, jacobians: bool | None = None,
                workspace: torch.Tensor | None = None, out: tuple | None = None):
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
    if jacobians is None:
        jacobians = ro.tape.shape[2] <= JACOBIAN_MAX_ENVS
    tape = ro.tape
    T, B = tape.shape[0] - 1, tape.shape[2]
    if out is not None:
        g, g_act = out
        assert g.is_contiguous() and g_act.is_contiguous() and tuple(g_act.shape) == (T, 2, B)
        if g.data_ptr() != g_final.data_ptr():
            g.copy_(g_final)
        g_act.zero_()
    else:
        g = g_final.to(device=tape.device, dtype=torch.float32).contiguous().clone()
        g_act = torch.zeros((T, 2, B), dtype=torch.float32, device=tape.device)
    assert g.shape == (NDYN, B)
    wp, wn, ws = None, 0, workspace
    if ws is None:
        ws = adjoint_workspace(B, T, tape.device, decision_tape, jacobians)
    if ws is not None:
        wp = (ws.data_ptr() + 255) & ~255
        wn = ws.numel() - (wp - ws.data_ptr())
    with torch.cuda.device(tape.device):
        _lib.check(_lib.lib().cotix_lunar_rollout_bwd(
            _stream_ptr(tape.device), B, T, tape.data_ptr(), B,
            ro.actions.data_ptr() if ro.actions is not None else None, g.data_ptr(), g_act.data_ptr(),
            flags.data_ptr() if flags is not None else None, C.byref(ro.params), wp, wn, 1 if jacobians else 0))
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


def pipeline_counters(device: torch.device | None = None, B: int | None = None) -> dict | None:
    """Queue lengths of the last pipeline step on the cached workspace of (device, B) — the most
    recently used one when B is None — summed over the env ranges (measurement aid): collision
    items, live contacts, envs with a candidate contact, auto-resets, envs that overflowed the queue."""
    for (idx, b), ws in reversed(list(_workspaces.items())):
        if (device is None or device.index in (None, idx)) and (B is None or B == b):
            wp = (ws.data_ptr() + 255) & ~255
            out = (C.c_int64 * 5)()
            with torch.cuda.device(ws.device):
                _lib.check(_lib.lib().cotix_lunar_pipeline_counters(_stream_ptr(ws.device), b, wp, out))
            return {"items": out[0], "hits": out[1], "contact_envs": out[2], "resets": out[3], "overflow": out[4]}
    return None


class LunarLanderEnv:
    """Vector env with HOST-facing reset/step (the `e2e` path of bench.py).

    Owns the device state of B envs through the C handle (`cotix_lunar_env_*`): each
    `step(action_host)` is H2D(action) -> kernels -> D2H(obs, reward, done), overlapped over
    ranges of envs (`set_chunks`).
    """

    def __init__(self, num_envs: int, params: LunarParams | None = None, device: int = 0):
        self.B = int(num_envs)
        self.params = params or default_params()
        self.device = device
        self._h = C.c_void_p()
        _lib.check(_lib.lib().cotix_lunar_env_create(C.byref(self._h), device, self.B, C.byref(self.params)))
        self.obs = torch.empty((NOBS, self.B), dtype=torch.float32).pin_memory()
        self.reward = torch.empty(self.B, dtype=torch.float32).pin_memory()
        self.done = torch.empty(self.B, dtype=torch.uint8).pin_memory()

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

    def step(self, action_host: torch.Tensor | np.ndarray | None, want_obs: bool = True):
        """action_host: [2, B] float32 host tensor (pinned for best speed).  want_obs=False leaves the
        observation on the device (86 % of the step's D2H bytes): (None, reward, done)."""
        ptr = None
        if action_host is not None:
            if isinstance(action_host, np.ndarray):
                action_host = torch.from_numpy(np.ascontiguousarray(action_host, np.float32))
            assert action_host.shape == (2, self.B) and action_host.dtype == torch.float32
            ptr = action_host.data_ptr()
        _lib.check(_lib.lib().cotix_lunar_env_step_host(self._h, ptr, self.obs.data_ptr() if want_obs else None,
                                                        self.reward.data_ptr(), self.done.data_ptr()))
        return (self.obs if want_obs else None), self.reward, self.done

    def get_state(self) -> np.ndarray:
        out = np.empty((NF, self.B), np.float32)
        _lib.check(_lib.lib().cotix_lunar_env_get_state_host(self._h, out.ctypes.data))
        return out

    def set_state(self, state: np.ndarray) -> None:
        s = np.ascontiguousarray(state, np.float32)
        assert s.shape == (NF, self.B)
        _lib.check(_lib.lib().cotix_lunar_env_set_state_host(self._h, s.ctypes.data))

    def rollout_graph(self, n_steps: int, action_dev: torch.Tensor | None = None) -> None:
        """Enqueue n_steps fused steps as one replayed CUDA graph on the handle's stream."""
        ptr = action_dev.data_ptr() if action_dev is not None else None
        _lib.check(_lib.lib().cotix_lunar_env_rollout_graph(self._h, int(n_steps), ptr))

    def synchronize(self) -> None:
        st = _lib.lib().cotix_lunar_env_stream(self._h)
        torch.cuda.ExternalStream(st, device=self.device).synchronize()
