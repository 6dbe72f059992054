"""The generic scene step (K2): RoboCup scene of cotix/_robocup.py:14-130.

CPU tests pin the oracle's generic collider and `robocup_scene()` against tests/golden/robocup.npz,
which oracle/gen_golden.py produced by executing the reference's own RoboCupEnv / AnyBody /
UniversalShape / ExplicitEulerPhysics / RandomizedCollider source.  GPU tests compare
`cotix_scene_step` with the golden rollouts and with the oracle on seeded batches, bit for bit
(NaN == NaN: the shipped scene goes NaN, and that is what must be reproduced).
"""
import ctypes as C
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "robocup.npz")


def _eq(a, b, name=""):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, name
    ok = (a == b) | (np.isnan(a) & np.isnan(b)) if a.dtype.kind == "f" else (a == b)
    assert ok.all(), f"{name}: {np.count_nonzero(~ok)} / {ok.size} mismatches, first at {np.argwhere(~ok)[:5].tolist()}"


def _bodies10(props, st6):
    nb = props.shape[0]
    b = np.zeros((nb, 10), np.float32)
    b[:, 0:2], b[:, 2:8], b[:, 8:10] = props[:, 0:2], st6, props[:, 2:4]
    return b


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


# ---------------------------------------------------------------------------------------------
# CPU
# ---------------------------------------------------------------------------------------------
def test_robocup_scene_is_the_reference_scene(gold):
    """robocup_scene() restates RoboCupEnv.__init__: same bodies, parts and float32 values as the
    leaves read off the reference's own constructor."""
    from parallax_b200 import scene as sc

    spec = sc.robocup_scene()
    _eq(spec.body_props, gold["a_props"], "body props")
    _eq(spec.part_body, gold["a_part_body"], "part -> body")
    _eq(spec.part_recs, gold["a_recs"], "part records")
    _eq(spec.init_state, gold["a_traj"][0], "initial state")
    assert spec.is_area == [True, True, False, False, False]
    # without the area parts, the remaining records are the goals and the ball of the golden scene (b)
    no_area = sc.robocup_scene(areas_collide=False)
    _eq(no_area.part_recs[:6], gold["b_recs"][:6], "goal parts")
    assert no_area.nb == 5 and list(no_area.part_body) == [2, 2, 2, 3, 3, 3, 4]


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_generic_collider_reproduces_reference_rollouts(oracle, exact_trig, gold, tag):
    """(a) RoboCupEnv as shipped, 3 steps: static bodies go NaN, eqx.error_if trips from step 2 —
    the oracle reports COTIX_ST_CCP_ERR exactly there; (b) goals + ball + 2 robots, 30 steps."""
    props, pb, recs = gold[tag + "_props"], gold[tag + "_part_body"], gold[tag + "_recs"]
    traj, keys, err = gold[tag + "_traj"], gold[tag + "_keys"], gold[tag + "_err"]
    st, key = traj[0].copy(), keys[0].copy()
    for t in range(traj.shape[0] - 1):
        out, cell, chosen, status = oracle.scene_step(_bodies10(props, st), pb, recs, key, want_status=True)
        st, key = out[:, 2:8], oracle.split(key, 2)[0]
        _eq(st, traj[t + 1], f"{tag}: state after step {t + 1}")
        _eq(key, keys[t + 1], f"{tag}: key after step {t + 1}")
        assert bool(status & 0x10) == bool(err[t]), f"{tag}: error_if at step {t + 1}"
    if tag == "a":
        assert np.isnan(traj[-1][0]).all() and np.isnan(traj[-1][3]).all()  # field and blue goal


def test_scene_analysis_matches_the_reference_grouping(cuda_lib_cpu):
    """Groups in first-appearance order, n_g entries each, distinct evaluations and the n_g^2 the
    reference's vmap-squared evaluates (cotix/_colliders.py:86-173) — host analysis only."""
    from parallax_b200 import scene as sc

    L = cuda_lib_cpu

    def info(spec):
        h = C.c_void_p()
        props = np.ascontiguousarray(spec.body_props, np.float32)
        pb = np.ascontiguousarray(spec.part_body, np.int32)
        recs = np.ascontiguousarray(spec.part_recs, np.float32)
        rc = L.cotix_scene_create(C.byref(h), -1, spec.nb, props.ctypes.data, pb.shape[0], pb.ctypes.data,
                                  recs.ctypes.data)
        if rc != 0:
            return rc
        ng, npairs, ev = C.c_int32(), C.c_int32(), C.c_int64()
        sizes = (C.c_int32 * 9)()
        assert L.cotix_scene_info(h, C.byref(ng), C.byref(npairs), C.byref(ev), sizes) == 0
        L.cotix_scene_destroy(h)
        return ng.value, list(sizes)[:ng.value], npairs.value, ev.value

    # as shipped: (AABB, AABB) 1 + 6 + 6 + 9 = 22 entries, (Circle, AABB) 8 entries
    assert info(sc.robocup_scene()) == (2, [22, 8], 32 + 8, 22 * 22 + 8 * 8)
    # areas off, 6 + 6 robots: goals 9; ball and robots vs 6 goal boxes 78; circles among themselves 78
    # entries, whose grid also holds 11 robot-vs-itself cells (i == j is evaluated, :163 only masks i < j)
    assert info(sc.robocup_scene(n_robots=12, areas_collide=False)) == (3, [9, 78, 78], 9 + 78 + 78 + 11,
                                                                        81 + 78 * 78 * 2)
    # (AABB, Circle) is not a key of _contact_funcs in that order: the reference raises
    bad = sc.SceneSpec(["ball", "box"], np.array([[1, 1, 1, 1], [1, 1, 1, 1]], np.float32), np.array([0, 1], np.int32),
                       np.stack([sc.circle_rec(1.0, (0, 0)), sc.aabb_rec((-1, -1), (1, 1))]), np.zeros((2, 6), np.float32))
    assert info(bad) == -1
    assert b"_contact_funcs" in L.cotix_last_error()


def test_scene_create_argument_checks(cuda_lib_cpu):
    """Limits and malformed descriptions are refused with COTIX_EINVAL and a message; a scene
    analysed without a device cannot be stepped."""
    from parallax_b200 import scene as sc

    L = cuda_lib_cpu

    def create(nb, props, pb, recs, device=-1):
        h = C.c_void_p()
        props = np.ascontiguousarray(props, np.float32); pb = np.ascontiguousarray(pb, np.int32)
        recs = np.ascontiguousarray(recs, np.float32)
        rc = L.cotix_scene_create(C.byref(h), device, nb, props.ctypes.data, pb.shape[0], pb.ctypes.data, recs.ctypes.data)
        return rc, h

    ball = sc.circle_rec(0.1, (0, 0))
    ok_props = np.ones((21, 4), np.float32)
    assert create(21, ok_props, [0], [ball])[0] == -1                      # more than COTIX_SCENE_MAX_BODIES
    assert create(2, ok_props, [0] * 29, [ball] * 29)[0] == -1             # more than COTIX_SCENE_MAX_PARTS
    assert create(2, ok_props, [0, 2], [ball, ball])[0] == -1              # part of a body that does not exist
    bad = ball.copy(); bad[0] = 7
    assert create(2, ok_props, [0, 1], [ball, bad])[0] == -1               # unknown shape type
    tri = sc.polygon_rec([(0, 0), (1, 0), (0, 1)]); tri[1] = 9
    assert create(2, ok_props, [0, 1], [ball, tri])[0] == -1               # polygon with 9 vertices
    # 20 single-circle bodies: 190 pairs + 18 self cells fit; the entry lists are 190 long
    rc, h = create(20, ok_props, list(range(20)), [ball] * 20)
    assert rc == 0
    ng, npairs, ev = C.c_int32(), C.c_int32(), C.c_int64()
    assert L.cotix_scene_info(h, C.byref(ng), C.byref(npairs), C.byref(ev), None) == 0
    assert (ng.value, npairs.value, ev.value) == (1, 190 + 18, 190 * 190)
    # host-only analysis: stepping it is an error, not a crash
    P = sc.SceneParams(); L.cotix_scene_default_params(C.byref(P))
    assert L.cotix_scene_step(None, h, 1, None, 1, None, C.byref(P), None, None, None) == -1
    assert L.cotix_scene_destroy(h) == 0
    assert L.cotix_scene_destroy(None) == 0


# ---------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------
def _keys_tensor(keys):
    import torch

    return torch.from_numpy(np.ascontiguousarray(np.asarray(keys, np.uint32).T).view(np.int32)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["a", "b"])
def test_cuda_scene_step_reproduces_reference_rollouts(cuda_lib, exact_trig, gold, tag):
    import torch
    from parallax_b200 import scene as sc

    props, pb, recs = gold[tag + "_props"], gold[tag + "_part_body"], gold[tag + "_recs"]
    traj, keys, err = gold[tag + "_traj"], gold[tag + "_keys"], gold[tag + "_err"]
    spec = sc.SceneSpec([f"b{i}" for i in range(props.shape[0])], props, pb, recs, traj[0])
    scene = sc.Scene(spec)
    B = 33  # the same env replicated: every lane must agree with the golden
    state = scene.initial_state(B)
    kd = _keys_tensor(np.repeat(keys[0][None], B, 0))
    for t in range(traj.shape[0] - 1):
        cell, chosen, status = scene.step(state, kd, want_cells=True)
        got = state.cpu().numpy().reshape(spec.nb, 6, B)
        for lane in (0, B - 1):
            _eq(got[:, :, lane], traj[t + 1], f"{tag}: state after step {t + 1}")
        _eq(kd.cpu().numpy().view(np.uint32)[:, 0], keys[t + 1], f"{tag}: key after step {t + 1}")
        assert bool(int(status[0].item()) & 0x10) == bool(err[t])
    scene.close()


def _random_states(spec, B, rng, spread):
    st = np.repeat(spec.init_state[:, :, None], B, axis=2).astype(np.float32)
    dyn = np.isfinite(spec.body_props[:, 0])
    for b in np.nonzero(dyn)[0]:
        st[b, 0] = rng.uniform(*spread[0], B)
        st[b, 1] = rng.uniform(*spread[1], B)
        st[b, 2:4] = rng.normal(0, 3.0, (2, B))
        st[b, 4] = rng.uniform(-3, 3, B)
        st[b, 5] = rng.normal(0, 2.0, B)
    return st


def _run_against_oracle(oracle, spec, st, keys, steps, check_cells=True):
    import torch
    from parallax_b200 import scene as sc

    nb, B = spec.nb, st.shape[2]
    scene = sc.Scene(spec)
    state = torch.from_numpy(np.ascontiguousarray(st.reshape(nb * 6, B))).cuda()
    kd = _keys_tensor(keys)
    ref, rkeys = st.copy(), np.asarray(keys, np.uint32).copy()
    hits = 0
    for t in range(steps):
        cell, chosen, status = scene.step(state, kd, want_cells=True)
        ocell = np.zeros((nb, nb, B), bool); ochosen = np.zeros((nb, B), np.int32); ostat = np.zeros(B, np.uint32)
        for e in range(B):
            out, c, ch, s = oracle.scene_step(_bodies10(spec.body_props, ref[:, :, e]), spec.part_body, spec.part_recs,
                                              rkeys[e], want_status=True)
            ref[:, :, e] = out[:, 2:8]
            rkeys[e] = oracle.split(rkeys[e], 2)[0]
            ocell[:, :, e], ochosen[:, e], ostat[e] = c, ch, s
        hits += int(ocell.sum())
        _eq(state.cpu().numpy().reshape(nb, 6, B), ref, f"state after step {t + 1}")
        _eq(kd.cpu().numpy().view(np.uint32).T, rkeys, f"keys after step {t + 1}")
        if check_cells:
            _eq(cell.cpu().numpy().reshape(nb, nb, B).astype(bool), ocell, f"cells at step {t + 1}")
            _eq(chosen.cpu().numpy(), ochosen, f"chosen at step {t + 1}")
            _eq(status.cpu().numpy().view(np.uint32), ostat, f"status at step {t + 1}")
    scene.close()
    return hits


@pytest.mark.gpu
def test_cuda_scene_step_robocup_with_robots_vs_oracle(cuda_lib, oracle, exact_trig):
    """Goals + ball + 3 robots packed in front of the yellow goal (areas off): contacts every
    step; state, keys, live cells, chosen indices and status bits equal the oracle's."""
    from parallax_b200 import scene as sc

    spec = sc.robocup_scene(n_robots=3, areas_collide=False)
    rng = np.random.default_rng(31)
    B = 384
    st = _random_states(spec, B, rng, ((-4.95, -4.2), (-0.7, 0.7)))
    keys = np.stack([np.zeros(B, np.uint32), np.arange(B, dtype=np.uint32) + 100], 1)
    hits = _run_against_oracle(oracle, spec, st, keys, steps=12)
    assert hits > 500  # the comparison saw real contacts


@pytest.mark.gpu
@pytest.mark.parametrize("areas", [False, True])
def test_cuda_scene_step_robocup_config4_vs_oracle(cuda_lib, oracle, exact_trig, areas):
    """BASELINE config 4: the RoboCup scene with 6 + 6 robots (17 bodies, 19 or 21 parts, 176 / 246
    distinct part pairs), robots and ball scattered over the yellow half so that they collide."""
    from parallax_b200 import scene as sc

    spec = sc.robocup_scene(n_robots=12, areas_collide=areas)
    rng = np.random.default_rng(34)
    B = 96
    st = _random_states(spec, B, rng, ((-4.9, -3.6), (-0.9, 0.9)))
    keys = np.stack([np.full(B, 11, np.uint32), np.arange(B, dtype=np.uint32) * 5], 1)
    hits = _run_against_oracle(oracle, spec, st, keys, steps=5)
    assert hits > 1000


@pytest.mark.gpu
def test_cuda_scene_step_shipped_robocup_batch_vs_oracle(cuda_lib, oracle, exact_trig):
    """RoboCupEnv as shipped (overlapping area AABBs, NaN statics) with the ball anywhere on the field."""
    from parallax_b200 import scene as sc

    spec = sc.robocup_scene()
    rng = np.random.default_rng(32)
    B = 256
    st = _random_states(spec, B, rng, ((-5.0, 5.0), (-3.5, 3.5)))
    keys = np.stack([np.full(B, 7, np.uint32), np.arange(B, dtype=np.uint32)], 1)
    _run_against_oracle(oracle, spec, st, keys, steps=4)


@pytest.mark.gpu
def test_cuda_scene_step_mixed_shapes_vs_oracle(cuda_lib, oracle, exact_trig):
    """Every key of _contact_funcs in one scene: Polygon6 ground, Polygon4 box, generic-free order
    (Polygon4, Polygon6), (AABB, Polygon*), (AABB, AABB), (Circle, *): bodies ordered so that each
    (higher, lower) type pair is legal."""
    from parallax_b200 import scene as sc

    hexagon = [(np.cos(a) * 3.0, np.sin(a) * 1.0 - 1.5) for a in np.linspace(0, 2 * np.pi, 7)[:-1]]
    box = [(-0.5, -0.4), (0.5, -0.4), (0.5, 0.4), (-0.5, 0.4)]
    names = ["ground6", "box4", "plank4", "crate", "crate2", "ball", "ball2"]
    props = np.array([[np.inf, np.inf, 0.3, 0.5], [2.0, 0.4, 0.6, 0.3], [1.0, 1.0, 1.0, 0.1], [1.5, 0.7, 0.4, 0.4],
                      [1.0, 0.5, 0.9, 0.2], [0.5, 0.1, 1.0, 1.0], [0.7, 0.2, 0.8, 0.6]], np.float32)
    part_body = np.array([0, 1, 2, 2, 3, 4, 5, 6], np.int32)
    recs = np.stack([sc.polygon_rec(hexagon), sc.polygon_rec(box), sc.polygon_rec([(-1, -0.1), (1, -0.1), (1, 0.1), (-1, 0.1)]),
                     sc.polygon_rec([(0.8, -0.3), (1.2, -0.3), (1.2, 0.3), (0.8, 0.3)]),
                     sc.aabb_rec((-0.4, -0.4), (0.4, 0.4)), sc.aabb_rec((-0.3, -0.5), (0.3, 0.5)),
                     sc.circle_rec(0.35, (0, 0)), sc.circle_rec(0.5, (0.1, 0.0))])
    init = np.zeros((7, 6), np.float32)
    spec = sc.SceneSpec(names, props, part_body, recs, init)
    rng = np.random.default_rng(33)
    B = 256
    st = _random_states(spec, B, rng, ((-1.6, 1.6), (-1.2, 0.8)))
    keys = np.stack([np.arange(B, dtype=np.uint32) * 3 + 1, np.arange(B, dtype=np.uint32)], 1)
    hits = _run_against_oracle(oracle, spec, st, keys, steps=6)
    assert hits > 1000


@pytest.mark.gpu
def test_cuda_scene_step_pile_of_circles_overflows_the_live_cell_list(cuda_lib, oracle, exact_trig):
    """14 circles dropped on one spot: every one of the 105 cells (i >= j) is live, more than the
    kernel's per-env list of live cells holds, so half of the envs take the table-scan path; the
    other half are spread out and have only their self contacts."""
    from parallax_b200 import scene as sc

    nb = 14
    props = np.tile(np.array([[1.0, 0.3, 0.8, 0.4]], np.float32), (nb, 1))
    props[:, 0] += np.arange(nb, dtype=np.float32) * 0.25
    recs = np.stack([sc.circle_rec(0.3 + 0.01 * i, (0, 0)) for i in range(nb)])
    spec = sc.SceneSpec([f"c{i}" for i in range(nb)], props, np.arange(nb, dtype=np.int32), recs,
                        np.zeros((nb, 6), np.float32))
    rng = np.random.default_rng(35)
    B = 64
    st = _random_states(spec, B, rng, ((-0.05, 0.05), (-0.05, 0.05)))
    far = _random_states(spec, B, rng, ((-40.0, 40.0), (-40.0, 40.0)))
    st[:, :, 1::2] = far[:, :, 1::2]
    keys = np.stack([np.full(B, 5, np.uint32), np.arange(B, dtype=np.uint32) + 9], 1)
    hits = _run_against_oracle(oracle, spec, st, keys, steps=3)
    assert hits > 32 * 100
