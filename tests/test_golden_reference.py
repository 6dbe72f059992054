"""Golden vectors produced by EXECUTING THE REFERENCE'S OWN SOURCE FILES (cotix/_contacts.py,
_collisions.py, _collision_resolution.py, _colliders.py, _lunar_lander.py, _physics_solvers.py and
the step of examples/test_viz.py:25-44) under oracle/jaxlite, a numpy float32 stand-in for the jax
API (script: oracle/gen_golden.py; fixtures: tests/golden/*.npz).  The CPU oracle — and, under
`-m gpu`, the CUDA path through the C-ABI — must reproduce them BIT FOR BIT.

The GJK start direction random_direction(PRNGKey(1)) is taken from the fixture (`init_dir`): it
goes through log1p/sqrt, where numpy and glibc differ by 2 ulp, which is exactly why it is a
parameter of the ABI and not a literal."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _eq(a, b, name=""):
    a, b = np.asarray(a), np.asarray(b)
    ok = (a == b) | (np.isnan(a) & np.isnan(b)) if a.dtype.kind == "f" else (a == b)
    assert ok.all(), f"{name}: {np.count_nonzero(~ok)} / {ok.size} mismatches at {np.argwhere(~ok)[:4].tolist()}"


def _gold(name):
    return np.load(os.path.join(GOLD, name))


def _init_dir():
    return _gold("contacts.npz")["init_dir"]


# ---- CPU: the oracle against the reference source ---------------------------------------------
def test_oracle_contacts_match_reference_source(oracle, exact_trig):
    g = _gold("contacts.npz")
    p = oracle.default_narrow_params()
    p.init_dir[0], p.init_dir[1] = float(g["init_dir"][0]), float(g["init_dir"][1])
    ex, pen, cp, _ = oracle.narrowphase(g["a"], g["b"], p)
    _eq(pen, g["pen"], "penetration_vector")
    _eq(cp, g["cp"], "contact_point")
    hits = ~np.isnan(g["cp"][:, 0])
    assert 100 < hits.sum() < 300 and (ex.astype(bool) == hits).all()
    # all six type pairs are present
    kinds = set(zip(g["a"][:, 0].astype(int).tolist(), g["b"][:, 0].astype(int).tolist()))
    assert kinds == {(0, 0), (0, 1), (1, 1), (0, 2), (1, 2), (2, 2)}
    # the start direction itself agrees with the oracle's own evaluation to 2 ulp
    mine = np.float32(oracle.default_lunar_params().init_dir[:])
    assert np.abs(mine.view(np.int32) - g["init_dir"].view(np.int32)).max() <= 2


def test_oracle_resolve_collision_matches_reference_source(oracle):
    g = _gold("resolve.npz")
    o1, o2 = oracle.resolve_collision(g["b1"], g["b2"], g["pen"], g["cp"])
    _eq(o1, g["o1"], "body1")
    _eq(o2, g["o2"], "body2")


def test_oracle_reset_matches_reference_source(oracle, exact_trig):
    """LunarLander(key): body poses, local shapes and the terrain (5 threefry sub-keys)."""
    g = _gold("reset.npz")
    P = oracle.default_lunar_params()
    s = oracle.lunar_reset(g["keys"], P)
    _eq(s[:18].T, g["dyn"], "body leaves")
    _eq(np.float32(P.lander_local), g["lander_local"], "lander vertices")
    _eq(np.float32(P.leg_local[0]), g["right_leg_local"], "right leg vertices")
    _eq(np.float32(P.leg_local[1]), g["left_leg_local"], "left leg vertices")
    for e in range(len(g["keys"])):
        kn = [-100.0, s[26, e], s[27, e], -2.0, 2.0, s[28, e], s[29, e], 100.0]
        for k in range(7):
            rec = oracle.polygon_rec([[kn[k], s[18 + k, e]], [kn[k], -10], [kn[k + 1], s[19 + k, e]], [kn[k + 1], -10]])
            _eq(rec[4:12].reshape(4, 2), g["ground_vertices"][e, k], f"ground part {k} of key {e}")


@pytest.mark.parametrize("dedupe", [True, False])
def test_oracle_rollout_matches_reference_source(oracle, exact_trig, dedupe):
    """24 composed steps (integrate, gravity, RandomizedCollider.resolve with its 306 coin keys,
    key split, joint springs) of 6 contact-rich envs: every body leaf and the PRNG key, each step."""
    g = _gold("lunar_rollout.npz")
    P = oracle.default_lunar_params()
    P.init_dir[0], P.init_dir[1] = float(_init_dir()[0]), float(_init_dir()[1])
    s = g["state0"].copy()
    _eq(s[:18], g["traj"][0], "initial state")
    for t in range(g["traj"].shape[0] - 1):
        s, _, _ = oracle.lunar_step(s, None, P, dedupe=dedupe)
        _eq(s[:18], g["traj"][t + 1], f"bodies after step {t}")
        _eq(s[30:32].view(np.uint32), g["keytraj"][t + 1], f"key after step {t}")
    cell, _, _ = oracle.lunar_collider_debug(g["state0"], P)
    assert cell.reshape(cell.shape[0], -1).sum(1).min() >= 2  # every env starts in contact


# ---- GPU: the CUDA path against the reference source --------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("sweep_mode", ["single", "pipeline"])
def test_cuda_contacts_match_reference_source(cuda_lib, sweep_mode):
    from parallax_b200 import _lib, shapes as sh

    g = _gold("contacts.npz")
    p = _lib.default_narrow_params()
    p.init_dir[0], p.init_dir[1] = float(g["init_dir"][0]), float(g["init_dir"][1])
    ci = sh.narrowphase(sh.from_records(g["a"]), sh.from_records(g["b"]), p, mode=sweep_mode)
    _eq(ci.penetration_vector.cpu().numpy(), g["pen"], "penetration_vector")
    _eq(ci.contact_point.cpu().numpy(), g["cp"], "contact_point")


@pytest.mark.gpu
def test_cuda_resolve_and_reset_match_reference_source(cuda_lib):
    import torch
    from parallax_b200 import lunar_lander as ll

    g = _gold("resolve.npz")
    d1, d2 = torch.from_numpy(g["b1"].copy()).cuda(), torch.from_numpy(g["b2"].copy()).cuda()
    dp, dc = torch.from_numpy(g["pen"]).cuda(), torch.from_numpy(g["cp"]).cuda()
    assert cuda_lib.cotix_resolve_collision(None, d1.shape[0], d1.data_ptr(), d2.data_ptr(), dp.data_ptr(),
                                            dc.data_ptr(), 0.3, 0.01) == 0
    _eq(d1.cpu().numpy(), g["o1"], "body1")
    _eq(d2.cpu().numpy(), g["o2"], "body2")
    r = _gold("reset.npz")
    s = ll.reset(r["keys"], ll.default_params()).numpy()
    _eq(s[:18].T, r["dyn"], "body leaves")
    # terrain: heights are the y of the two non-floor vertices of each trapezoid
    for e in range(len(r["keys"])):
        for k in range(7):
            ys = sorted(set(r["ground_vertices"][e, k][:, 1].tolist()) - {-10.0})
            assert set(np.float32([s[18 + k, e], s[19 + k, e]]).tolist()) >= set(np.float32(ys).tolist())


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fused", "pipeline"])
def test_cuda_rollout_matches_reference_source(cuda_lib, mode):
    from parallax_b200 import lunar_lander as ll

    g = _gold("lunar_rollout.npz")
    d = _init_dir()
    P = ll.default_params(init_dir=[float(d[0]), float(d[1])])
    st = ll.LunarState.from_numpy(g["state0"])
    for t in range(g["traj"].shape[0] - 1):
        st, _, _ = ll.step(st, None, P, mode=mode)
        s = st.numpy()
        _eq(s[:18], g["traj"][t + 1], f"bodies after step {t}")
        _eq(s[30:32].view(np.uint32), g["keytraj"][t + 1], f"key after step {t}")
