#!/usr/bin/env python
"""Generates tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN SOURCE (/root/reference/cotix,
unmodified, read-only) under `oracle/jaxlite` — a numpy float32 stand-in for the jax / equinox
API (the real jax cannot be installed here: no network).  Run in the build container only:

    python oracle/gen_golden.py            # a few minutes of pure-Python work

What is executed: cotix/_contacts.py (all six contact functions), cotix/_collisions.py (GJK, EPA),
cotix/_collision_resolution.py, cotix/_lunar_lander.py (constructor = reset, step = joints),
cotix/_physics_solvers.py, cotix/_colliders.py (RandomizedCollider.resolve), the composed
step of examples/test_viz.py:25-44, and cotix/_robocup.py (RoboCupEnv) with cotix/_bodies.py
(AnyBody) and cotix/_universal_shape.py.

Two harness choices, both documented in SURVEY A.6 / DESIGN.md §5:
  * `list(set(...))` at cotix/_colliders.py:116-120 hashes tracers by identity, so it removes
    nothing and leaves an order that differs from process to process; here `set` is bound to an
    order-preserving identity inside that module (canonical order = insertion order);
  * sin / cos / arctan2 are the engine's cx_trig.h definitions (oracle trig mode 1);
  * RoboCup only: after showing that the shipped scene raises on its first step, `eqx.error_if`
    is switched to record-and-continue (the engine's status-bit semantics).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, "jaxlite"))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.append("/root/reference")

from oracle import oracle as o  # noqa: E402

o.set_trig_mode(1)

import equinox as eqx  # noqa: E402  (jaxlite)
import jax.numpy as jnp  # noqa: E402
import jax.random as jr  # noqa: E402

import cotix._colliders as ref_colliders  # noqa: E402
import cotix._contacts as ref_contacts  # noqa: E402
from cotix._bodies import AnyBody, DynamicBody  # noqa: E402
from cotix._collision_resolution import resolve_collision  # noqa: E402
from cotix._convex_shapes import AABB, Circle, Polygon  # noqa: E402
from cotix._geometry_utils import random_direction  # noqa: E402
from cotix._lunar_lander import LunarLander  # noqa: E402
from cotix._physics_solvers import ExplicitEulerPhysics  # noqa: E402

ref_colliders.set = lambda items: list(items)  # see the module docstring

GOLD = os.path.join(ROOT, "tests", "golden")
os.makedirs(GOLD, exist_ok=True)


def f32(x):
    return np.asarray(x, np.float32)


# ---------------------------------------------------------------------------------------------
def shape_from_rec(r):
    t = int(r[0])
    if t == 0:
        return Circle(radius=jnp.array(r[2]), position=jnp.array(r[4:6]))
    if t == 1:
        return AABB(lower=jnp.array(r[4:6]), upper=jnp.array(r[6:8]))
    n = int(r[1])
    return Polygon(jnp.array(r[4:4 + 2 * n].reshape(n, 2)))


FUNCS = {(0, 0): ref_contacts.circle_vs_circle, (0, 1): ref_contacts.circle_vs_aabb,
         (1, 1): ref_contacts.aabb_vs_aabb, (0, 2): ref_contacts.circle_vs_polygon,
         (1, 2): ref_contacts.aabb_vs_polygon, (2, 2): ref_contacts.polygon_vs_polygon}


def gen_contacts(n=360, seed=7):
    from gen_inputs import random_pairs

    rng = np.random.default_rng(seed)
    a, b = random_pairs(rng, n, mix=(0.4, 0.2, 0.1, 0.1, 0.1, 0.1))
    pen = np.zeros((n, 2), np.float32)
    cp = np.zeros((n, 2), np.float32)
    for k in range(n):
        info = FUNCS[(int(a[k, 0]), int(b[k, 0]))](shape_from_rec(a[k]), shape_from_rec(b[k]))
        pen[k], cp[k] = f32(info.penetration_vector), f32(info.contact_point)
    np.savez_compressed(os.path.join(GOLD, "contacts.npz"), a=a, b=b, pen=pen, cp=cp,
                        init_dir=f32(random_direction(jr.PRNGKey(1))))
    print(f"contacts: {n} pairs, {int((~np.isnan(cp[:, 0])).sum())} hits")


def gen_resolve(n=200, seed=8):
    rng = np.random.default_rng(seed)
    b1 = rng.normal(size=(n, 10)).astype(np.float32)
    b2 = rng.normal(size=(n, 10)).astype(np.float32)
    for b in (b1, b2):
        b[:, 0] = rng.uniform(0.5, 30, n)
        b[:, 1] = rng.uniform(0.5, 30, n)
        b[:, 8] = rng.uniform(0, 1, n)
        b[:, 9] = rng.uniform(0, 1, n)
    b2[::5, 0] = np.inf
    b2[::5, 1] = np.inf
    pen = rng.normal(scale=0.1, size=(n, 2)).astype(np.float32)
    cp = rng.normal(size=(n, 2)).astype(np.float32)
    cp[::7] = np.nan
    o1, o2 = b1.copy(), b2.copy()

    def body(r):
        d = object.__new__(DynamicBody)
        for name, v in (("mass", r[0]), ("inertia", r[1]), ("position", r[2:4]), ("velocity", r[4:6]),
                        ("angle", r[6]), ("angular_velocity", r[7]), ("elasticity", r[8]),
                        ("friction_coefficient", r[9])):
            object.__setattr__(d, name, jnp.array(v))
        return d

    for k in range(n):
        info = ref_contacts.ContactInfo(penetration_vector=jnp.array(pen[k]), contact_point=jnp.array(cp[k]))
        n1, n2, _ = resolve_collision(body(b1[k]), body(b2[k]), info)
        for out, nb in ((o1, n1), (o2, n2)):
            out[k, 4:6] = f32(nb.velocity)
            out[k, 7] = f32(nb.angular_velocity)
    np.savez_compressed(os.path.join(GOLD, "resolve.npz"), b1=b1, b2=b2, pen=pen, cp=cp, o1=o1, o2=o2)
    print(f"resolve: {n} cases")


# ---------------------------------------------------------------------------------------------
def env_to_state(env, key):
    s = np.zeros(36, np.float32)
    for i in range(3):
        b = env.bodies[i]
        s[6 * i:6 * i + 2] = f32(b.position)
        s[6 * i + 2:6 * i + 4] = f32(b.velocity)
        s[6 * i + 4] = f32(b.angle)
        s[6 * i + 5] = f32(b.angular_velocity)
    parts = env.bodies[3].shape.parts  # 7 trapezoids; vertices were re-sorted by Polygon4.__init__
    xs, hs = [], []
    for p in parts:
        v = f32(p.vertices)
        top = v[v[:, 1] > -10.0 + 1e-6] if (v[:, 1] > -10.0 + 1e-6).sum() == 2 else None
        xs.append(sorted(set(float(x) for x in v[:, 0])))
        hs.append(v)
    return s, parts


def set_pose(env, state):
    """Write the dynamic leaves of an oracle-layout state vector into the reference env."""
    bodies = list(env.bodies)
    for i in range(3):
        b = bodies[i]
        b = b.set_position(jnp.array(state[6 * i:6 * i + 2])).set_velocity(jnp.array(state[6 * i + 2:6 * i + 4]))
        b = b.set_angle(jnp.array(state[6 * i + 4])).set_angular_velocity(jnp.array(state[6 * i + 5]))
        bodies[i] = b
    return eqx.tree_at(lambda e: e.bodies, env, bodies)


def dyn_of(env):
    s = np.zeros(18, np.float32)
    for i in range(3):
        b = env.bodies[i]
        s[6 * i:6 * i + 2] = f32(b.position)
        s[6 * i + 2:6 * i + 4] = f32(b.velocity)
        s[6 * i + 4] = f32(b.angle)
        s[6 * i + 5] = f32(b.angular_velocity)
    return s


def gen_reset(n=24):
    """LunarLander(key): terrain heights / knots recovered from the ground parts' vertices."""
    keys = np.stack([np.zeros(n, np.uint32), np.arange(n, dtype=np.uint32) * 7 + 1], 1)
    verts = np.zeros((n, 7, 4, 2), np.float32)
    dyn = np.zeros((n, 18), np.float32)
    local = None
    for k in range(n):
        env = LunarLander(jnp.array(keys[k]))
        for q, p in enumerate(env.bodies[3].shape.parts):
            verts[k, q] = f32(p.vertices)
        dyn[k] = dyn_of(env)
        if local is None:
            local = [f32(env.bodies[i].shape.parts[0].vertices) for i in range(3)]
    np.savez_compressed(os.path.join(GOLD, "reset.npz"), keys=keys, ground_vertices=verts, dyn=dyn,
                        lander_local=local[0], right_leg_local=local[1], left_leg_local=local[2])
    print(f"reset: {n} keys")


def gen_rollout(n_env=6, steps=24, seed=11):
    """The composed step of examples/test_viz.py:25-44 from contact-rich states."""
    from gen_inputs import lunar_keys, perturbed_lunar_state

    P = o.default_lunar_params()
    keys = lunar_keys(n_env, offset=3)
    s0 = o.lunar_reset(keys, P)
    rng = np.random.default_rng(seed)
    s0 = perturbed_lunar_state(s0, rng, P)
    # put the lander a little above its local terrain so that legs and hull touch down within the horizon
    for e in range(n_env):
        knots = [-100.0, s0[26, e], s0[27, e], -2.0, 2.0, s0[28, e], s0[29, e], 100.0]
        x = float(s0[0, e])
        k = max(0, min(6, int(np.searchsorted(knots, x)) - 1))
        t = (x - knots[k]) / (knots[k + 1] - knots[k])
        ground = (1 - t) * s0[18 + k, e] + t * s0[19 + k, e]
        dy = np.float32(ground + rng.uniform(0.3, 0.9)) - s0[1, e]
        for b in range(3):
            s0[6 * b + 1, e] += dy
            s0[6 * b + 2, e] *= 0.2
            s0[6 * b + 3, e] = -abs(s0[6 * b + 3, e]) * 0.5 - 0.5
    physics = ExplicitEulerPhysics()
    collider = ref_colliders.RandomizedCollider()
    traj = np.zeros((steps + 1, 18, n_env), np.float32)
    keytraj = np.zeros((steps + 1, 2, n_env), np.uint32)
    t0 = time.time()
    for e in range(n_env):
        env = set_pose(LunarLander(jnp.array(keys[e])), s0[:, e])
        key = jnp.array(keys[e])
        traj[0, :, e] = dyn_of(env)
        keytraj[0, :, e] = np.asarray(key)
        for t in range(steps):
            new_bodies, _ = physics.step(env.bodies, dt=1e-2)
            new_bodies = eqx.tree_at(lambda x: x[0].velocity, new_bodies,
                                     new_bodies[0].velocity + jnp.array([0.0, -0.002]))
            new_bodies = collider.resolve(new_bodies, key)
            key, _next = jr.split(key)
            env = eqx.tree_at(lambda x: x.bodies, env, new_bodies)
            env = env.step()
            traj[t + 1, :, e] = dyn_of(env)
            keytraj[t + 1, :, e] = np.asarray(key)
        print(f"rollout env {e}: {steps} steps, {time.time() - t0:.0f}s", flush=True)
    np.savez_compressed(os.path.join(GOLD, "lunar_rollout.npz"), keys=keys, state0=s0, traj=traj, keytraj=keytraj)


# ---------------------------------------------------------------------------------------------
def rec_of_shape(p):
    """Local shape record (include/cotix_b200.h) of a reference Circle / AABB / polygon part."""
    r = np.zeros(16, np.float32)
    if isinstance(p, Circle):
        r[0], r[2] = 0, f32(p.radius)
        r[4:6] = f32(p.position)
    elif isinstance(p, AABB):
        r[0] = 1
        r[4:6], r[6:8] = f32(p.lower), f32(p.upper)
    else:
        v = f32(p.vertices)
        r[0], r[1] = 2, v.shape[0]
        r[4:4 + 2 * v.shape[0]] = v.reshape(-1)
    return r


def scene_arrays(bodies):
    props = np.array([[f32(b.mass), f32(b.inertia), f32(b.elasticity), f32(b.friction_coefficient)]
                      for b in bodies], np.float32)
    part_body, recs = [], []
    for i, b in enumerate(bodies):
        for p in b.shape.parts:
            part_body.append(i)
            recs.append(rec_of_shape(p))
    return props, np.array(part_body, np.int32), np.stack(recs)


def state6(bodies):
    return np.array([[*f32(b.position), *f32(b.velocity), f32(b.angle), f32(b.angular_velocity)]
                     for b in bodies], np.float32)


def roll_scene(bodies, key, steps):
    """physics.step -> collider.resolve -> key split, the loop of examples/test_viz.py:25-44
    without the lander-specific lines."""
    physics = ExplicitEulerPhysics()
    collider = ref_colliders.RandomizedCollider()
    traj = [state6(bodies)]
    keys = [np.asarray(key, np.uint32)]
    errs = []
    for t in range(steps):
        n0 = len(eqx.ERROR_LOG)
        bodies, _ = physics.step(bodies, dt=1e-2)
        bodies = collider.resolve(bodies, key)
        key, _n = jr.split(key)
        traj.append(state6(bodies))
        keys.append(np.asarray(key, np.uint32))
        errs.append(len(eqx.ERROR_LOG) > n0)
    return np.stack(traj), np.stack(keys), np.array(errs)


def gen_robocup(steps=30):
    """(1) RoboCupEnv() exactly as shipped (cotix/_robocup.py:14-130) for 3 steps — the static
    AABBs overlap and the scene goes NaN, which is the reference's behaviour; (2) the goals, the
    ball and two circular robots built with the reference's own classes, ball driven into the
    yellow goal, robots into each other and into the ball."""
    from cotix._robocup import RoboCupEnv
    from cotix._universal_shape import UniversalShape

    t0 = time.time()
    env = RoboCupEnv()
    bodies = list(env.bodies)
    out = {}
    out["a_props"], out["a_part_body"], out["a_recs"] = scene_arrays(bodies)
    # as shipped, the first step already trips eqx.error_if at cotix/_contacts.py:105-107 (the
    # closest point bc + clip(c - bc) rounds one ulp outside a thin goal box): the reference raises
    try:
        roll_scene(bodies, jnp.array(np.array([0, 5], np.uint32)), 1)
        out["a_raises"] = np.array(False)
    except RuntimeError as ex:
        out["a_raises"] = np.array(True)
        print("as shipped, step 1 raises:", ex)
    # from here on error_if records and lets the value through (the engine reports COTIX_ST_CCP_ERR)
    eqx.ERROR_MODE = "record"
    out["a_traj"], out["a_keys"], out["a_err"] = roll_scene(bodies, jnp.array(np.array([0, 5], np.uint32)), 3)
    print(f"robocup as shipped: 3 steps, {time.time() - t0:.0f}s", flush=True)

    def robot(x, y, vx, vy, w):
        return AnyBody(shape=UniversalShape(Circle(position=jnp.zeros((2,)), radius=jnp.array(0.09))),
                       mass=jnp.array(2.0), inertia=jnp.array(0.0081), position=jnp.array([x, y]),
                       angle=jnp.array(0.0), velocity=jnp.array([vx, vy]), angular_velocity=jnp.array(w),
                       elasticity=jnp.array(0.5), friction_coefficient=jnp.array(0.5))

    ball = bodies[4].set_position(jnp.array([-4.58, 0.1])).set_velocity(jnp.array([-3.0, 0.5]))
    scene = [bodies[2], bodies[3], ball, robot(-4.40, 0.12, -4.0, 0.0, 0.0), robot(-4.18, 0.15, -6.0, -0.2, 1.0)]
    out["b_props"], out["b_part_body"], out["b_recs"] = scene_arrays(scene)
    out["b_traj"], out["b_keys"], out["b_err"] = roll_scene(scene, jnp.array(np.array([0, 9], np.uint32)), steps)
    print(f"robocup goals + ball + 2 robots: {steps} steps, {time.time() - t0:.0f}s", flush=True)
    np.savez_compressed(os.path.join(GOLD, "robocup.npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["contacts", "resolve", "reset", "rollout", "robocup"]
    if "contacts" in which:
        gen_contacts()
    if "resolve" in which:
        gen_resolve()
    if "reset" in which:
        gen_reset()
    if "rollout" in which:
        gen_rollout()
    if "robocup" in which:
        gen_robocup()
