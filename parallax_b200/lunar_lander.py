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
_workspaces: dict = {}


def _workspace(device: torch.device, B: int) -> torch.Tensor:
    """Device scratch of the wavefront pipeline, cached per (device, batch)."""
    key = (device.index, B)
    ws = _workspaces.get(key)
    if ws is None:
        nbytes = int(_lib.lib().cotix_lunar_workspace_bytes(B))
        ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=device)
        _workspaces.clear()  # one live batch shape at a time is the common case
        _workspaces[key] = ws
    return ws


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
    out = d if inplace else torch.empty_like(d)
    reward = torch.empty(B, dtype=torch.float32, device=d.device)
    done = torch.empty(B, dtype=torch.uint8, device=d.device)
    obs = torch.empty((NOBS, B), dtype=torch.float32, device=d.device) if want_obs else None
    if action is not None:
        action = action.to(device=d.device, dtype=torch.float32).contiguous()
        assert action.shape == (2, B)
    m = _MODES[mode]
    use_pipe = m == MODE_PIPELINE or (m == MODE_AUTO and B >= PIPELINE_MIN_ENVS)
    with torch.cuda.device(d.device):
        args = (_stream_ptr(d.device), B, d.data_ptr(), out.data_ptr(), d.stride(0),
                action.data_ptr() if action is not None else None, reward.data_ptr(), done.data_ptr(),
                obs.data_ptr() if obs is not None else None, C.byref(params))
        if use_pipe:
            ws = _workspace(d.device, B)
            wp = (ws.data_ptr() + 255) & ~255
            _lib.check(_lib.lib().cotix_lunar_step_ws(*args, wp, ws.numel() - (wp - ws.data_ptr())))
        else:
            _lib.check(_lib.lib().cotix_lunar_step(*args))
    ns = state if inplace else LunarState(out)
    if want_obs:
        return ns, reward, done, obs
    return ns, reward, done


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
    g_next = g_next.to(device=d.device, dtype=torch.float32).contiguous()
    assert g_next.shape == (NDYN, B)
    g_in = torch.empty_like(g_next)
    g_act = torch.zeros((2, B), dtype=torch.float32, device=d.device)
    if action is not None:
        action = action.to(device=d.device, dtype=torch.float32).contiguous()
        assert action.shape == (2, B)
    with torch.cuda.device(d.device):
        _lib.check(_lib.lib().cotix_lunar_step_bwd(
            _stream_ptr(d.device), B, d.data_ptr(), state_next.data.data_ptr() if state_next is not None else None,
            d.stride(0), action.data_ptr() if action is not None else None, g_next.data_ptr(), g_in.data_ptr(),
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
            params: LunarParams | None = None, mode: str = "auto") -> Rollout:
    """n_steps steps of `step`, keeping every intermediate state (`cotix_lunar_rollout`): the
    forward half of the reference's `jax.value_and_grad` over a rollout."""
    params = params or default_params()
    d = state.data
    B = d.shape[1]
    T = int(n_steps)
    tape = torch.empty((T + 1, NF, B), dtype=torch.float32, device=d.device)
    tape[0].copy_(d)
    if actions is not None:
        actions = actions.to(device=d.device, dtype=torch.float32).contiguous()
        assert actions.shape == (T, 2, B)
    m = _MODES[mode]
    use_pipe = m == MODE_PIPELINE or (m == MODE_AUTO and B >= PIPELINE_MIN_ENVS)
    with torch.cuda.device(d.device):
        wp, wn = None, 0
        if use_pipe:
            ws = _workspace(d.device, B)
            wp = (ws.data_ptr() + 255) & ~255
            wn = ws.numel() - (wp - ws.data_ptr())
        _lib.check(_lib.lib().cotix_lunar_rollout(
            _stream_ptr(d.device), B, T, tape.data_ptr(), B, actions.data_ptr() if actions is not None else None,
            C.byref(params), wp, wn))
    return Rollout(tape, actions, params)


def rollout_vjp(ro: Rollout, g_final: torch.Tensor, flags: torch.Tensor | None = None,
                decision_tape: bool | int = True, jacobians: bool | None = None):
    """Backward pass over a `Rollout` (`cotix_lunar_rollout_bwd`): g_final [18, B] = dL/d(final
    dynamic leaves).  Returns (dL/d(initial dynamic leaves) [18, B], dL/d(actions) [T, 2, B]).

    decision_tape: True = scratch for the whole rollout, an int = for that many steps at a time:
    the collider's decisions — and with `jacobians` the transposed Jacobians of the env-steps with
    contacts — are computed in batched launches over (step, env) and the T reverse launches are
    light; False = every reverse step replays its own decisions.  Without `jacobians` the results
    are bit-identical to False; with them they differ at rounding level.  jacobians=None picks them
    for B <= JACOBIAN_MAX_ENVS: they trade 18 x the contact-adjoint work, done at full occupancy, for
    short reverse launches — a win while the sweep is latency-bound (B = 4096: 9.0 -> 5.7 ms), a loss
    once it is throughput-bound (B = 65536: 23.7 -> 43 ms)."""
    if jacobians is None:
        jacobians = ro.tape.shape[2] <= JACOBIAN_MAX_ENVS
    tape = ro.tape
    T, B = tape.shape[0] - 1, tape.shape[2]
    g = g_final.to(device=tape.device, dtype=torch.float32).contiguous().clone()
    assert g.shape == (NDYN, B)
    g_act = torch.zeros((T, 2, B), dtype=torch.float32, device=tape.device)
    wp, wn, ws = None, 0, None
    if decision_tape and T > 0:
        L = _lib.lib()
        steps = T if decision_tape is True else max(1, min(T, int(decision_tape)))
        size = L.cotix_lunar_adjoint_workspace_bytes if jacobians else L.cotix_lunar_adjoint_decisions_bytes
        wn = int(size(B, steps))
        ws = torch.empty(wn + 256, dtype=torch.uint8, device=tape.device)
        wp = (ws.data_ptr() + 255) & ~255
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


def pipeline_counters(device: torch.device | None = None) -> dict | None:
    """Queue lengths of the last pipeline step on the cached workspace (measurement aid):
    collision items, live contacts, envs with a candidate contact, auto-resets, overflow."""
    for (idx, _B), ws in _workspaces.items():
        if device is None or device.index in (None, idx):
            off = ((ws.data_ptr() + 255) & ~255) - ws.data_ptr()
            c = ws[off:off + 32].view(torch.int32).cpu().tolist()
            return {"items": c[0], "hits": c[1], "contact_envs": c[2], "resets": c[3], "overflow": c[4]}
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

    def step(self, action_host: torch.Tensor | np.ndarray | None):
        """action_host: [2, B] float32 host tensor (pinned for best speed)."""
        ptr = None
        if action_host is not None:
            if isinstance(action_host, np.ndarray):
                action_host = torch.from_numpy(np.ascontiguousarray(action_host, np.float32))
            assert action_host.shape == (2, self.B) and action_host.dtype == torch.float32
            ptr = action_host.data_ptr()
        _lib.check(_lib.lib().cotix_lunar_env_step_host(self._h, ptr, self.obs.data_ptr(),
                                                        self.reward.data_ptr(), self.done.data_ptr()))
        return self.obs, self.reward, self.done

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
