"""Properties of the oracle's LunarLander restatement that follow from the reference source
(cotix/_lunar_lander.py, cotix/_colliders.py, examples/test_viz.py:25-44).  CPU only."""
import numpy as np

from gen_inputs import lunar_keys, perturbed_lunar_state


def _eq(a, b):
    return ((a == b) | (np.isnan(a) & np.isnan(b))).all()


def test_reset_scene(oracle):
    """LunarLander.__init__ :29-143: fixed body poses, terrain structure of :109-132."""
    P = oracle.default_lunar_params()
    keys = lunar_keys(256)
    s = oracle.lunar_reset(keys, P)
    assert (s[0] == 0).all() and (s[1] == 5).all() and np.allclose(s[4], 0.01)        # lander :78-87
    assert np.allclose(s[6], -1.2) and np.allclose(s[7], 4.6) and np.allclose(s[12], 1.2)  # legs :75-76
    assert (s[2:4] == 0).all() and (s[5] == 0).all()
    h, kn = s[18:26], s[26:30]
    assert (h[3] == -2).all() and (h[4] == -2).all()                                    # pad :112-113
    assert (np.abs(h[[1, 2, 5, 6]]) <= 5).all() and (np.abs(h[[0, 7]]) <= 50).all()     # :110-114
    assert ((kn[0] >= -12) & (kn[0] < -9)).all() and ((kn[1] >= -8) & (kn[1] < -4)).all()
    assert ((kn[2] >= 4) & (kn[2] < 8)).all() and ((kn[3] >= 9) & (kn[3] < 12)).all()   # :116-125
    # k1..k5 = split(key, 5); heights = uniform(k1, (8,), -5, 5)
    for e in (0, 17, 255):
        ks = oracle.split(keys[e], 5)
        hh = oracle.uniform(ks[0], 8, -5.0, 5.0)
        assert np.array_equal(h[[1, 2, 5, 6], e], hh[[1, 2, 5, 6]])
        assert h[0, e] == hh[0] * np.float32(10) and h[7, e] == hh[7] * np.float32(10)
        assert kn[0, e] == oracle.uniform(ks[1], 1, -12.0, -9.0)[0]
        assert kn[3, e] == oracle.uniform(ks[4], 1, 9.0, 12.0)[0]
    assert np.array_equal(s[30:32].view(np.uint32).T, keys)


def test_free_fall_is_analytic(oracle):
    """Before any contact the lander only feels the per-step dv of (0, -0.002)
    (examples/test_viz.py:27-31) and the joint springs keep the legs attached."""
    P = oracle.default_lunar_params()
    s = oracle.lunar_reset(lunar_keys(8), P)
    for t in range(1, 301):
        s, rew, done = oracle.lunar_step(s, None, P)
        assert (rew == 0).all() and (done == 0).all()
    # gravity acts on the lander only (m = 30); the springs drag the two unit-mass legs along,
    # so the assembly accelerates at 0.002 * 30 / 32 per step
    g = 0.002 * 30.0 / 32.0
    assert np.allclose(s[3], -g * 300, atol=5e-3)
    assert np.allclose(s[1], 5 - g * 0.01 * 300 * 299 / 2, atol=0.05)
    assert np.abs(s[7] - (s[1] - 0.4)).max() < 0.05                    # legs follow via the springs
    assert (s[32].view(np.uint32) == 300).all()                        # step counter
    # key chain: key <- split(key)[0] each step
    k = lunar_keys(8)[3]
    for _ in range(300):
        k = oracle.split(k, 2)[0]
    assert np.array_equal(s[30:32, 3].view(np.uint32), k)


def test_touchdown_and_rebound(oracle):
    """SURVEY A.8 probe: the lander comes down onto the terrain around step ~800 and rebounds
    (Baumgarte push), no NaN over 1500 steps."""
    P = oracle.default_lunar_params()
    s = oracle.lunar_reset(lunar_keys(16), P)
    ymin = np.full(16, np.inf)
    for t in range(1500):
        s, _, _ = oracle.lunar_step(s, None, P)
        ymin = np.minimum(ymin, s[1])
    assert not np.isnan(s[:18]).any()
    assert (ymin < 3.0).all() and (ymin > -10).all()
    assert (s[3] > -1.5).all()  # the fall was arrested (free fall would be at v_y = -3)


def test_memoised_collider_equals_literal(oracle):
    """The reference evaluates 81 + 225 narrowphase calls per step for 25 distinct pairs (SURVEY
    A.6); memoising identical pairs must not change anything."""
    for mode in (0, 1):
        P = oracle.default_lunar_params()
        P.rng_mode = mode
        rng = np.random.default_rng(3 + mode)
        s = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(96), P), rng, P)
        a, b = s.copy(), s.copy()
        for _ in range(25):
            a, _, _ = oracle.lunar_step(a, None, P, dedupe=True)
            b, _, _ = oracle.lunar_step(b, None, P, dedupe=False)
        assert _eq(a, b)


def test_collider_bookkeeping(oracle):
    """all_contacts cells and the per-body choice (cotix/_colliders.py:137-140,274-295): only
    rows i > j (plus the left-leg self cell) are populated; a body picks one of its
    live cells or itself."""
    P = oracle.default_lunar_params()
    rng = np.random.default_rng(11)
    s = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(400), P), rng, P)
    cell, cand, chosen = oracle.lunar_collider_debug(s, P)
    assert cell.any()
    iu = np.triu_indices(4, 1)
    assert not cell[:, iu[0], iu[1]].any()
    diag = cell[:, np.arange(4), np.arange(4)]
    assert not diag[:, [0, 1, 3]].any()
    for e in range(400):
        for i in range(4):
            j = chosen[e, i]
            if cell[e, i].any():
                assert cell[e, i, j]
            else:
                assert j == i
    assert (cell <= cand).all()


def test_zero_action_equals_no_action_and_substeps(oracle):
    P = oracle.default_lunar_params()
    rng = np.random.default_rng(5)
    s = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(64), P), rng, P)
    a, _, _ = oracle.lunar_step(s, None, P)
    b, _, _ = oracle.lunar_step(s, np.zeros((2, 64), np.float32), P)
    assert _eq(a, b)
    P4 = oracle.default_lunar_params()
    P4.substeps = 4
    c, _, _ = oracle.lunar_step(s, None, P4)
    d = s
    for _ in range(4):
        d, _, _ = oracle.lunar_step(d, None, P)
    live = ~np.isnan(d[:18]).any(axis=0)
    assert _eq(c[:, live], d[:, live])


def test_env_logic_judge_and_auto_reset(oracle):
    """New env logic: crash / out-of-bounds / timeout end an episode, the env is
    reset in the same step with a fresh key and the episode counter advances."""
    P = oracle.default_lunar_params()
    P.judge_enabled, P.auto_reset, P.reset_perturb, P.max_episode_steps = 1, 1, 1, 50
    rng = np.random.default_rng(6)
    B = 256
    s = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B), P), rng, P)
    seen = set()
    for t in range(120):
        act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
        s, rew, done = oracle.lunar_step(s, act, P)
        seen |= set(np.unique(done).tolist())
        fin = done != 0
        assert (s[32].view(np.uint32)[fin] == 0).all()      # steps restart
        assert (s[34][fin] == 0).all()                      # return restarts
        assert (rew[done == 1] == -100).all() and (rew[done == 2] == -100).all()
    assert {0, 1, 4} <= seen
    assert (s[33].view(np.uint32) >> 8).min() >= 2          # every env went through >= 2 episodes
    assert not np.isnan(s[:18]).any()


def test_scene_step_two_circles_head_on(oracle):
    """Generic collider on a hand-made scene: two unit-mass circles approaching head-on exchange
    momentum along the normal (resolve_collision, cotix/_collision_resolution.py:76-151)."""
    bodies = np.float32([[1, 1, -0.9, 0, 1, 0.01, 0, 0, 1, 0],
                         [1, 1, 0.9, 0, -1, -0.01, 0, 0, 1, 0]])
    parts = np.stack([oracle.circle_rec(1.0, [0, 0]), oracle.circle_rec(1.0, [0, 0])])  # body-local parts
    hit = 0
    for k in range(40):
        out, cell, chosen = oracle.scene_step(bodies, [0, 1], parts, (0, k), dt=0.0)
        if cell[1, 0] and chosen[1] == 0:
            hit += 1
            assert out[0, 4] < 0 and out[1, 4] > 0      # velocities reversed along x
            assert abs(out[0, 4] + out[1, 4]) < 1e-5    # momentum conserved
        else:
            assert np.array_equal(out[:, 4], bodies[:, 4])
    assert 8 <= hit <= 32  # the coin flip (p = 0.5, cotix/_colliders.py:220) decides


def test_horizon_100_steps_trig_definitions(oracle):
    """The 100-step-horizon tolerance of north_star, on the CPU: the oracle with the engine's trig
    (cx_trig.h — what the CUDA path computes bit for bit) against the oracle with libm trig, from
    identical contact-rich states.  tests/horizon.py states and asserts the tolerance."""
    from gen_inputs import lunar_keys, perturbed_lunar_state
    from horizon import HORIZON_STEPS, check_horizon

    P = oracle.default_lunar_params()
    P.judge_enabled = 1
    B = 2048
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B), P), np.random.default_rng(161), P)
    a, b = s0.copy(), s0.copy()
    same = np.ones(B, bool)
    try:
        for _ in range(HORIZON_STEPS):
            oracle.set_trig_mode(1)
            a, _, da, oa = oracle.lunar_step(a, None, P, want_obs=True)
            oracle.set_trig_mode(0)
            b, _, db, ob = oracle.lunar_step(b, None, P, want_obs=True)
            same &= (oa[6:8] == ob[6:8]).all(axis=0) & (da == db)
    finally:
        oracle.set_trig_mode(0)
    check_horizon(a, b, same)
