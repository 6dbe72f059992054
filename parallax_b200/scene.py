"""Host-side mirror of the reference's generic world step for a batch of envs that share one
static scene: `ExplicitEulerPhysics.step` + `RandomizedCollider.resolve` (+ key split), behind
`cotix_scene_step` (csrc/scene_kernel.cu).

Reference surface this stands in for:
  * `RoboCupEnv()`                               cotix/_robocup.py:14-130   -> `robocup_scene()`
  * `bodies, _ = physics.step(bodies, dt)`       cotix/_physics_solvers.py:26-33
    `bodies = collider.resolve(bodies, key)`     cotix/_colliders.py:74-351 -> `Scene.step(state, keys)`
Bodies are `AnyBody` leaves (cotix/_bodies.py:134-182); parts are the shapes of the body's
`UniversalShape` in order (cotix/_universal_shape.py).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib, _xp
from ._lib import SceneParams

REC = 16
MAX_BODIES, MAX_PARTS = 20, 28
_f = np.float32


def circle_rec(radius, position) -> np.ndarray:
    r = np.zeros(REC, _f)
    r[0], r[2] = 0, radius
    r[4:6] = position
    return r


def aabb_rec(lower, upper) -> np.ndarray:
    r = np.zeros(REC, _f)
    r[0] = 1
    r[4:6], r[6:8] = lower, upper
    return r


def polygon_rec(vertices) -> np.ndarray:
    v = np.asarray(vertices, _f).reshape(-1, 2)
    r = np.zeros(REC, _f)
    r[0], r[1] = 2, v.shape[0]
    r[4:4 + 2 * v.shape[0]] = v.reshape(-1)
    return r


@dataclass
class SceneSpec:
    """Static description + initial dynamic state of one scene."""

    names: list
    body_props: np.ndarray   # [nb, 4] mass, inertia, elasticity, friction_coefficient
    part_body: np.ndarray    # [nparts] int32
    part_recs: np.ndarray    # [nparts, 16] local shape records
    init_state: np.ndarray   # [nb, 6] px, py, vx, vy, angle, angular_velocity
    is_area: list = field(default_factory=list)
    # what env.draw paints (cotix/_robocup.py:132-150): (body, local shape record, fill colour, edge colour)
    draw_list: list = field(default_factory=list)

    @property
    def nb(self) -> int:
        return self.body_props.shape[0]

    def bodies10(self, state6=None) -> np.ndarray:
        """[nb, 10] DynamicBody leaves (cotix/_bodies.py:189-206) as the oracle takes them."""
        s = self.init_state if state6 is None else np.asarray(state6, _f)
        out = np.zeros((self.nb, 10), _f)
        out[:, 0:2] = self.body_props[:, 0:2]
        out[:, 2:8] = s
        out[:, 8:10] = self.body_props[:, 2:4]
        return out


def default_params(**overrides) -> SceneParams:
    p = SceneParams()
    _lib.lib().cotix_scene_default_params(C.byref(p))
    for k, v in overrides.items():
        cur = getattr(p, k)
        if hasattr(cur, "__len__"):
            for i, x in enumerate(v):
                cur[i] = x
        else:
            setattr(p, k, v)
    return p


def robocup_scene(n_robots: int = 0, areas_collide: bool = True, robot_radius: float = 0.09,
                  robot_mass: float = 2.0, seed: int = 3) -> SceneSpec:
    """The scene of `RoboCupEnv.__init__` (cotix/_robocup.py:14-130): field and play-area AABBs
    (`is_area=True`), two goals of three thin AABBs each, and the ball; float32 arithmetic in the
    order the reference writes it.

    As shipped the collider ignores `is_area` (cotix/_colliders.py never reads it), so the
    static AABBs overlap, `resolve_collision` divides 0 by 0 and the scene goes NaN on its first
    step — `areas_collide=True` reproduces exactly that.  `areas_collide=False`
    leaves the two area bodies without collidable parts (new).  `n_robots` adds
    circular robots after the ball (new; BASELINE config 4 uses 6 + 6), placed on two rows
    facing each other with velocities drawn from `default_rng(seed)`.
    """
    half = lambda d: np.asarray(d, _f) / _f(2)
    field_dim, play_dim, goal_dim, ball_radius = (10.4, 7.4), (9, 6), (0.2, 1), 0.022
    field_lo, field_hi = -half(field_dim), half(field_dim)
    play_lo, play_hi = -half(play_dim), half(play_dim)
    ygb_lo = np.array([play_lo[0] - _f(goal_dim[0]), -goal_dim[1] / 2], _f)
    ygb_hi = np.array([play_lo[0], goal_dim[1] / 2], _f)
    gw = 0.01
    yellow = [
        (ygb_lo, ygb_lo + np.array([gw, goal_dim[1]], _f)),
        (ygb_lo - np.array([-gw, 0], _f), ygb_lo + np.array([goal_dim[0], gw], _f)),
        (ygb_hi - np.array([goal_dim[0], gw], _f), ygb_hi),
    ]
    blue = [(np.array([-hi[0], lo[1]], _f), np.array([-lo[0], hi[1]], _f)) for lo, hi in yellow]

    inf = np.inf
    names = ["field", "play_area", "yellow_goal", "blue_goal", "ball"]
    # mass, inertia (AnyBody default 1.0), elasticity, friction_coefficient (default 1.0)
    props = [[inf, 1.0, 1.0, 1.0], [inf, 1.0, 1.0, 1.0], [inf, 1.0, 0.5, 1.0], [inf, 1.0, 0.5, 1.0],
             [0.5, 1.0, 1.0, 1.0]]
    state = [[0, 0, 0, 0, 0, 0]] * 4 + [[0.0, 0.0, 1.0, 0.01, 0.0, 10.0]]
    is_area = [True, True, False, False, False]
    part_body, recs = [], []
    if areas_collide:
        part_body += [0, 1]
        recs += [aabb_rec(field_lo, field_hi), aabb_rec(play_lo, play_hi)]
    for lo, hi in yellow:
        part_body.append(2); recs.append(aabb_rec(lo, hi))
    for lo, hi in blue:
        part_body.append(3); recs.append(aabb_rec(lo, hi))
    part_body.append(4)
    recs.append(circle_rec(_f(ball_radius) * _f(3), (0.0, 0.0)))
    rng = np.random.default_rng(seed)
    for r in range(n_robots):
        side = -1.0 if r % 2 == 0 else 1.0
        row = r // 2
        names.append(f"robot{r}")
        props.append([robot_mass, 0.5 * robot_mass * robot_radius ** 2, 0.5, 0.5])
        vx, vy = rng.normal(0.0, 1.0, 2)
        state.append([side * (1.0 + 0.9 * (row % 3)), -2.0 + 0.8 * row, -side * abs(vx), 0.3 * vy, 0.0, 0.0])
        is_area.append(False)
        part_body.append(4 + 1 + r)
        recs.append(circle_rec(robot_radius, (0.0, 0.0)))
    # RoboCupEnv.draw (cotix/_robocup.py:132-150): colours per body, edges per body; the area
    # bodies are painted whether or not they collide; robots (new) use Circle.draw's default grey
    green, white, yel, blu, red = (0, 180, 0), (255, 255, 255), (255, 255, 0), (0, 128, 255), (255, 0, 0)
    draw = [(0, aabb_rec(field_lo, field_hi), green, white), (1, aabb_rec(play_lo, play_hi), green, white)]
    draw += [(2, aabb_rec(lo, hi), yel, yel) for lo, hi in yellow]
    draw += [(3, aabb_rec(lo, hi), blu, blu) for lo, hi in blue]
    draw.append((4, circle_rec(_f(ball_radius) * _f(3), (0.0, 0.0)), red, None))
    draw += [(5 + r, circle_rec(robot_radius, (0.0, 0.0)), (128, 128, 128), None) for r in range(n_robots)]
    return SceneSpec(names, np.asarray(props, _f), np.asarray(part_body, np.int32),
                     np.stack(recs).astype(_f), np.asarray(state, _f), is_area, draw)


class Scene:
    """Device-resident static scene + the batched step.  `state` is a [nb * 6, B] float32 device
    array (rows 6b..6b+5 = px, py, vx, vy, angle, angular_velocity of body b), `keys` a [2, B]
    int32 / uint32 device array holding the key words; torch tensors or `devmem.DeviceArray`."""

    def __init__(self, spec: SceneSpec, device: int = 0):
        self.spec = spec
        self.device = device
        self._h = C.c_void_p()
        props = np.ascontiguousarray(spec.body_props, _f)
        pb = np.ascontiguousarray(spec.part_body, np.int32)
        recs = np.ascontiguousarray(spec.part_recs, _f)
        _lib.check(_lib.lib().cotix_scene_create(C.byref(self._h), device, spec.nb, props.ctypes.data,
                                                 pb.shape[0], pb.ctypes.data, recs.ctypes.data))

    def close(self):
        if self._h:
            _lib.lib().cotix_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        ng, npairs, ev = C.c_int32(), C.c_int32(), C.c_int64()
        sizes = (C.c_int32 * 9)()
        _lib.check(_lib.lib().cotix_scene_info(self._h, C.byref(ng), C.byref(npairs), C.byref(ev), sizes))
        return {"groups": ng.value, "group_sizes": list(sizes)[:ng.value], "distinct_pairs": npairs.value,
                "reference_evaluations": ev.value}

    def initial_state(self, B: int, backend: str = "cuda"):
        """[nb * 6, B] initial state on this scene's device (backend "cuda" = torch, "native")."""
        s = np.ascontiguousarray(self.spec.init_state.reshape(-1, 1).repeat(B, axis=1), np.float32)
        return _xp.backend_for(f"{backend}:{self.device}").asarray(s, np.float32)

    def step(self, state, keys, params: SceneParams | None = None, want_cells: bool = False):
        """In place on `state` and (with `advance_key`) `keys`.  Returns (cell [nb*nb, B] uint8,
        chosen [nb, B] int32, status [B] int32) when `want_cells`, else None."""
        params = params or default_params()
        nb = self.spec.nb
        xp = _xp.backend_of(state)
        assert state.shape[0] == nb * 6 and state.is_contiguous()
        B = state.shape[1]
        assert tuple(keys.shape) == (2, B) and keys.is_contiguous()
        cell = chosen = status = None
        if want_cells:
            cell = xp.empty((nb * nb, B), np.uint8)
            chosen = xp.empty((nb, B), np.int32)
            status = xp.empty((B,), np.int32)
        with xp.device_ctx():
            _lib.check(_lib.lib().cotix_scene_step(
                xp.stream_ptr(), self._h, B, state.data_ptr(), B,
                keys.data_ptr(), C.byref(params), cell.data_ptr() if want_cells else None,
                chosen.data_ptr() if want_cells else None, status.data_ptr() if want_cells else None))
        return (cell, chosen, status) if want_cells else None
