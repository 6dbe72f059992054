"""Pytree converters and the world / environment wrappers (parallax_b200/envs.py).

CPU: `bodies_from_state` reproduces the leaves of the reference's own `LunarLander(key)` pytree
(tests/golden/reset.npz, written by executing cotix/_lunar_lander.py under oracle/jaxlite) and
`state_from_bodies` inverts it.  GPU: `LunarEnvironment.eval` (the shape of
AbstractEnvironment.eval, cotix/_envs.py:37-132) against the oracle stepped by hand."""
import os

import numpy as np
import pytest

from gen_inputs import lunar_keys, perturbed_lunar_state

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_bodies_from_state_reproduces_the_reference_pytree(oracle, exact_trig):
    from parallax_b200 import envs
    from parallax_b200 import lunar_lander as ll

    g = np.load(os.path.join(GOLD, "reset.npz"))
    P = ll.default_params()
    s = oracle.lunar_reset(g["keys"], P)
    b = envs.bodies_from_state(s, P)
    assert len(b) == 4 and all(set(envs.LEAVES) <= set(x) for x in b)
    dyn = np.concatenate([np.concatenate([b[i]["position"], b[i]["velocity"], b[i]["angle"][:, None],
                                          b[i]["angular_velocity"][:, None]], 1) for i in range(3)], 1)
    assert np.array_equal(dyn, g["dyn"])                                   # DynamicBody leaves, _bodies.py:189-206
    assert np.array_equal(b[0]["shape"][0], g["lander_local"])
    assert np.array_equal(b[1]["shape"][0], g["right_leg_local"])
    assert np.array_equal(b[2]["shape"][0], g["left_leg_local"])
    assert np.array_equal(b[3]["shape"], g["ground_vertices"])             # 7 Polygon4 parts, ordered clockwise
    # masses etc. of cotix/_lunar_lander.py:78-141
    assert b[0]["mass"][0] == 30 and b[0]["inertia"][0] == 30 and b[1]["mass"][0] == 1
    assert np.isinf(b[3]["mass"][0]) and np.isinf(b[3]["inertia"][0])
    # and back: everything but the bookkeeping fields is recovered bit for bit
    s2 = envs.state_from_bodies(b, g["keys"], P)
    assert np.array_equal(s2[:32].view(np.uint32), s[:32].view(np.uint32))
    assert np.isnan(s2[35]).all() and not s2[32:35].any()


def test_converters_round_trip_on_moving_states(oracle, exact_trig):
    from parallax_b200 import envs
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    keys = lunar_keys(64, offset=100)
    s = perturbed_lunar_state(oracle.lunar_reset(keys, P), np.random.default_rng(3), P)
    for _ in range(5):
        s = oracle.lunar_step(s, None, P)[0]
    k = s[30:32].view(np.uint32).T
    s2 = envs.state_from_bodies(envs.bodies_from_state(s, P), k, P)
    assert np.array_equal(s2[:32].view(np.uint32), s[:32].view(np.uint32))


def test_order_clockwise_follows_the_reference_test():
    """test/test_geometry_utils.py:7-21: any permutation of a convex polygon's vertices is ordered
    to the same cyclic sequence."""
    from parallax_b200 import envs

    v = np.asarray([[0, 0], [0, 1], [1, 1], [1, 0]], np.float32)
    ref = envs.order_clockwise(v)
    rng = np.random.default_rng(0)
    for _ in range(20):
        assert np.array_equal(envs.order_clockwise(v[rng.permutation(4)]), ref)


@pytest.mark.gpu
def test_environment_eval_matches_the_oracle_stepped_by_hand(cuda_lib, oracle, exact_trig):
    import torch

    from parallax_b200 import envs
    from parallax_b200 import lunar_lander as ll

    o = oracle
    P = ll.default_params(judge_enabled=1, max_episode_steps=0)
    B, NFE, K, period = 512, 4, 5, 0.2
    s0 = perturbed_lunar_state(o.lunar_reset(lunar_keys(B), P), np.random.default_rng(6), P)
    calls = []

    def policy(ws):  # a state feedback law: thrust against the fall, steer against the tilt
        d = ws.state.data
        a = torch.stack([(-d[3]).clamp(0, 1), (-2.0 * d[4]).clamp(-1, 1)])
        calls.append(a.cpu().numpy())
        return a

    env = envs.LunarEnvironment(envs.LunarWorld(P, mode="fused"), envs.LunarWorldState(0.0, ll.LunarState.from_numpy(s0), P),
                                envs.ZeroOrderHold(policy), envs.LunarJudge())
    env2, reward = env.eval(period, NFE, WFE_scale=K)
    assert len(calls) == NFE and abs(env2.state.time - period) < 1e-6 and env.state.time == 0.0
    # by hand: per NFE one action, K sub-steps of dt = period / NFE / K, finished envs stay frozen
    Pk = ll.default_params(judge_enabled=1, max_episode_steps=0, substeps=K, dt=period / NFE / K)
    s, tot, fin = s0, np.zeros(B, np.float32), np.zeros(B, bool)
    for n in range(NFE):
        nxt, r, d = o.lunar_step(s, calls[n], Pk)
        tot = tot + np.where(fin, np.float32(0), r)
        s = np.where(fin[None, :], s, nxt)
        fin |= d != 0
    assert fin.sum() > 20 and (~fin).sum() > 20
    assert np.array_equal(env2.state.state.numpy(), s, equal_nan=True)
    assert np.array_equal(reward.cpu().numpy(), tot, equal_nan=True)
    # the wrapper exposes the reference pytree of the new state
    b = env2.state.bodies
    assert np.array_equal(b[0]["position"][:, 1], s[1], equal_nan=True)
    # one un-actuated forward step of the world == the composed step of examples/test_viz.py:25-44
    w = envs.LunarWorld(ll.default_params())
    ws, _, _ = w.forward(envs.LunarWorldState(0.0, ll.LunarState.from_numpy(s0), P), None, 1e-2)
    assert np.array_equal(ws.state.numpy(), o.lunar_step(s0, None, ll.default_params())[0], equal_nan=True)
