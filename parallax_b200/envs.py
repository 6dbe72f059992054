"""Drop-in shaped wrappers: the reference's world / environment interface on top of the CUDA engine.

Reference surface mirrored here:
  * the pytree of a `LunarLander` env — `bodies[i]` are `AnyBody`s whose leaves are those of
    `DynamicBody` (cotix/_bodies.py:189-206: mass, inertia, position, velocity, angle,
    angular_velocity, elasticity, friction_coefficient) plus a `UniversalShape` of polygon parts
    (cotix/_lunar_lander.py:29-143)            -> `bodies_from_state` / `state_from_bodies`
  * `AbstractWorldState` (time, bodies) and `AbstractWorld.forward(state, control_signal, dt)`
    (cotix/_worlds.py:14-35)                   -> `LunarWorldState`, `LunarWorld.forward`
  * `AbstractJudge` (__call__, is_done, end_reward; cotix/_envs.py:9-28) and
    `AbstractEnvironment.eval(eval_period, num_NFEs, WFE_scale)` (cotix/_envs.py:37-132): per
    NFE one control evaluation, then WFE_scale forward steps of dt = eval_period / num_NFEs /
    WFE_scale with rewards accumulated and the state frozen once the judge says done
                                               -> `LunarJudge`, `LunarEnvironment.eval`
The reference leaves `AbstractWorld.forward`, the judge and the control abstract (not
functional upstream); the concrete semantics here are the engine's (include/cotix_b200.h: actions =
main / side throttle, the judge and done codes of `cotix_lunar_step`, K sub-steps per call).
Everything runs batched over B envs; the per-body leaves carry a leading env axis.
"""
from __future__ import annotations

from dataclasses import dataclass, replace

import numpy as np

from . import _xp
from . import lunar_lander as ll
from ._lib import LunarParams

LEAVES = ("mass", "inertia", "position", "velocity", "angle", "angular_velocity", "elasticity",
          "friction_coefficient")


def order_clockwise(vertices: np.ndarray) -> np.ndarray:
    """cotix/_geometry_utils.py:60-67 on [..., n, 2]: stable sort by atan2 about the centroid."""
    v = np.asarray(vertices, np.float32)
    c = v.mean(-2, keepdims=True, dtype=np.float32)
    ang = np.arctan2(v[..., 1] - c[..., 1], v[..., 0] - c[..., 0])
    idx = np.argsort(ang, axis=-1, kind="stable")
    return np.take_along_axis(v, idx[..., None], axis=-2)


def terrain_knots(state: np.ndarray, params: LunarParams) -> np.ndarray:
    """x of the 8 terrain knots per env, [8, B] (cotix/_lunar_lander.py:115-125)"""
    B = state.shape[1]
    f = np.float32
    return np.stack([np.full(B, -params.terrain_x_outer, f), state[ll.F_KNOT0], state[ll.F_KNOT0 + 1],
                     np.full(B, -params.pad_half_width, f), np.full(B, params.pad_half_width, f),
                     state[ll.F_KNOT0 + 2], state[ll.F_KNOT0 + 3], np.full(B, params.terrain_x_outer, f)])


def bodies_from_state(state, params: LunarParams | None = None) -> list[dict]:
    """The reference pytree of B LunarLander envs from the engine's SoA state ([36, B], `LunarState`
    or numpy): `bodies[0..2]` lander / right leg / left leg, `bodies[3]` the static ground.  Every
    body is a dict of the DynamicBody leaves (leading axis B) plus `shape`: the local polygon parts
    (`[parts, n, 2]`; the ground's 7 trapezoids per env, ordered as Polygon4.__init__ orders them)."""
    params = params or ll.default_params()
    s = state.numpy() if isinstance(state, ll.LunarState) else np.asarray(state, np.float32)
    B = s.shape[1]
    f = np.float32
    shapes = [np.asarray(params.lander_local, f).reshape(1, 6, 2),
              np.asarray(params.leg_local[0], f).reshape(1, 4, 2),
              np.asarray(params.leg_local[1], f).reshape(1, 4, 2)]
    bodies = []
    for b in range(3):
        m, i, e, mu = (f(params.body_props[b][k]) for k in range(4))
        bodies.append({
            "mass": np.full(B, m, f), "inertia": np.full(B, i, f),
            "position": np.stack([s[6 * b], s[6 * b + 1]], -1), "velocity": np.stack([s[6 * b + 2], s[6 * b + 3]], -1),
            "angle": s[6 * b + 4].copy(), "angular_velocity": s[6 * b + 5].copy(),
            "elasticity": np.full(B, e, f), "friction_coefficient": np.full(B, mu, f),
            "shape": shapes[b]})
    X = terrain_knots(s, params)
    Hh = s[ll.F_HEIGHT0:ll.F_HEIGHT0 + 8]
    fl = np.full(B, params.ground_floor, f)
    parts = np.stack([np.stack([np.stack([X[k], Hh[k]], -1), np.stack([X[k], fl], -1),
                                np.stack([X[k + 1], Hh[k + 1]], -1), np.stack([X[k + 1], fl], -1)], 1)
                      for k in range(7)], 1)                                   # [B, 7, 4, 2]
    m, i, e, mu = (f(params.body_props[3][k]) for k in range(4))
    z = np.zeros(B, f)
    bodies.append({"mass": np.full(B, m, f), "inertia": np.full(B, i, f), "position": np.zeros((B, 2), f),
                   "velocity": np.zeros((B, 2), f), "angle": z.copy(), "angular_velocity": z.copy(),
                   "elasticity": np.full(B, e, f), "friction_coefficient": np.full(B, mu, f),
                   "shape": order_clockwise(parts)})
    return bodies


def state_from_bodies(bodies: list[dict], keys, params: LunarParams | None = None) -> np.ndarray:
    """Inverse of `bodies_from_state`: the engine's [36, B] state from the reference pytree leaves
    (poses and velocities of bodies 0..2, the terrain from the ground's 7 parts) and the PRNG keys
    ([B, 2] uint32).  Step counter, status, return are zero; the shaping potential starts unset."""
    params = params or ll.default_params()
    f = np.float32
    B = np.asarray(bodies[0]["angle"]).shape[0]
    s = np.zeros((ll.NF, B), f)
    for b in range(3):
        d = bodies[b]
        s[6 * b], s[6 * b + 1] = np.asarray(d["position"], f).T
        s[6 * b + 2], s[6 * b + 3] = np.asarray(d["velocity"], f).T
        s[6 * b + 4], s[6 * b + 5] = np.asarray(d["angle"], f), np.asarray(d["angular_velocity"], f)
    parts = np.asarray(bodies[3]["shape"], f)                                   # [B, 7, 4, 2]
    floor = f(params.ground_floor)

    def knot(part, side):  # (x, height) of the left (0) / right (1) side of a trapezoid
        x = part[..., 0].min(-1) if side == 0 else part[..., 0].max(-1)
        on = part[..., 0] == x[..., None]
        y = np.where(on & (part[..., 1] != floor), part[..., 1], -np.inf).max(-1)
        return x, np.where(np.isfinite(y), y, floor)

    for k in range(7):
        x, h = knot(parts[:, k], 0)
        s[ll.F_HEIGHT0 + k] = h
        if k in (1, 2):
            s[ll.F_KNOT0 + k - 1] = x
        if k in (5, 6):
            s[ll.F_KNOT0 + k - 3] = x
    s[ll.F_HEIGHT0 + 7] = knot(parts[:, 6], 1)[1]
    k = np.ascontiguousarray(np.asarray(keys).astype(np.uint32)).reshape(B, 2)
    s[ll.F_KEY0], s[ll.F_KEY1] = k[:, 0].view(f), k[:, 1].view(f)
    s[ll.F_SHAPING] = np.nan
    return s


@dataclass
class LunarWorldState:
    """`AbstractWorldState` (cotix/_worlds.py:14-24): time + bodies, batched over B envs."""

    time: float
    state: ll.LunarState
    params: LunarParams

    @property
    def bodies(self) -> list[dict]:
        return bodies_from_state(self.state, self.params)


class ThrottleSignal:
    """A control signal: the [2, B] (main, side) throttle applied during a forward step."""

    def __init__(self, action):
        self.action = action


class ZeroOrderHold:
    """An `AbstractControl` (cotix/_envs.py:46-48 calls `control(state)` once per NFE and gets back a
    dense approximation + the new control): `policy(LunarWorldState) -> [2, B]` is evaluated once per
    NFE and held over the WFE_scale forward steps."""

    def __init__(self, policy):
        self.policy = policy

    def __call__(self, state: LunarWorldState):
        action = self.policy(state)
        return (lambda _state: ThrottleSignal(action)), self


class LunarJudge:
    """`AbstractJudge` (cotix/_envs.py:9-28).  The engine evaluates reward and termination inside
    the step kernel (csrc/cx_lunar_step.cuh, finish_step): per forward step it returns
    `judge(state, u) * dt + [done] * end_reward` as one number and the done code; this class only
    names that contract and decodes the codes."""

    CODES = {0: "running", 1: "crash", 2: "out_of_bounds", 3: "landed", 4: "timeout", 5: "nan"}

    def is_done(self, done):
        return (done.numpy() != 0) if hasattr(done, "is_native") else (done != 0)


class LunarWorld:
    """`AbstractWorld` (cotix/_worlds.py:27-36) for LunarLander: `forward` = one composed step of
    examples/test_viz.py:25-44 (+ engines) on the GPU."""

    def __init__(self, params: LunarParams | None = None, mode: str = "auto"):
        self.params = params or ll.default_params()
        self.mode = mode

    def forward(self, state: LunarWorldState, control_signal: ThrottleSignal | None, dt: float,
                substeps: int = 1):
        P = self.params.copy()
        P.dt = float(dt)
        P.substeps = int(substeps)
        action = control_signal.action if control_signal is not None else None
        nxt, reward, done = ll.step(state.state, action, P, mode=self.mode)
        return LunarWorldState(state.time + float(dt) * substeps, nxt, state.params), reward, done


@dataclass
class LunarEnvironment:
    """`AbstractEnvironment` (cotix/_envs.py:31-132): world, state, control, judge + `eval`."""

    world: LunarWorld
    state: LunarWorldState
    control: ZeroOrderHold
    judge: LunarJudge

    def eval(self, eval_period: float, num_NFEs: int, WFE_scale: int = 10):
        """-> (new env, reward [B]).  Per NFE: one control evaluation, then WFE_scale forward steps of
        dt = eval_period / num_NFEs / WFE_scale in ONE launch (the kernels' K sub-steps: same action,
        rewards add up, an env that finishes freezes); an env that finished in an earlier NFE stays
        frozen and earns nothing more (the `finished` carry of cotix/_envs.py:39-121)."""
        time_per_nfe = float(eval_period) / int(num_NFEs)
        dt = time_per_nfe / float(WFE_scale)
        st, control = self.state, self.control
        B = st.state.batch
        xp = _xp.backend_of(st.state.data)
        if xp.name == "torch":
            torch = xp.torch
            total = torch.zeros(B, dtype=torch.float32, device=xp.device)
            finished = torch.zeros(B, dtype=torch.bool, device=xp.device)
        else:  # torch-free: the two per-env selects of the carry run on the host
            total = np.zeros(B, np.float32)
            finished = np.zeros(B, bool)
        for _ in range(int(num_NFEs)):
            # one NFE: `dense_control_approximation_fn, new_control = control(state)`; the new
            # control is what the next NFE calls (the carry of cotix/_envs.py:40,43,117)
            dense_fn, new_control = control(st)
            nxt, reward, done = self.world.forward(st, dense_fn(st), dt, substeps=WFE_scale)
            keep = finished
            if xp.name == "torch":
                total = total + torch.where(keep, torch.zeros_like(reward), reward)
                data = torch.where(keep[None, :], st.state.data, nxt.state.data)
            else:
                total = total + np.where(keep, np.float32(0), reward.numpy())
                data = xp.asarray(np.where(keep[None, :], st.state.numpy(), nxt.state.numpy()), np.float32)
            st = LunarWorldState(nxt.time, ll.LunarState(data), st.params)
            finished = finished | self.judge.is_done(done)
            control = new_control
        # like the reference, only the state is written back: "we don't update control, since it
        # could have 'jumps' between evals" (cotix/_envs.py:126-131)
        return replace(self, state=st), total
