"""GPU parity tests proper: the CUDA path, called through the C-ABI, against the CPU oracle on the
same seeded inputs.  Bar: bit-exact (integer/flag/index work AND fp32, because the engine
defines its own sin/cos/atan2 and is built without FMA contraction — the oracle is switched to
those trig definitions by the `exact_trig` fixture).  A second set of tests compares against
the oracle in libm mode within the north_star tolerance (1e-5 relative per step)."""
import ctypes as C

import numpy as np
import pytest

from gen_inputs import lunar_keys, perturbed_lunar_state, random_pairs

pytestmark = pytest.mark.gpu


def _eq(a, b, name=""):
    a = np.asarray(a)
    b = np.asarray(b)
    assert a.shape == b.shape, name
    ok = (a == b) | (np.isnan(a) & np.isnan(b)) if a.dtype.kind == "f" else (a == b)
    assert ok.all(), f"{name}: {np.count_nonzero(~ok)} / {ok.size} mismatches, first at {np.argwhere(~ok)[:5].tolist()}"


# ------------------------------------------------------------------------------------------
# RNG
# ------------------------------------------------------------------------------------------
def test_threefry_kat_on_device(cuda_lib, oracle):
    import torch

    # Random123 known-answer vectors for Threefry-2x32 (20 rounds)
    kats = [((0, 0), (0, 0), (0x6B200159, 0x99BA4EFE)),
            ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF), (0x1CB996FC, 0xBB002BE7)),
            ((0x13198A2E, 0x03707344), (0x243F6A88, 0x85A308D3), (0xC4923A9C, 0x483DF7A0))]
    for key, ctr, want in kats:
        x0 = torch.from_numpy(np.array([ctr[0]], np.uint32).view(np.int32)).cuda()
        x1 = torch.from_numpy(np.array([ctr[1]], np.uint32).view(np.int32)).cuda()
        out = torch.zeros(2, dtype=torch.int32).cuda()
        rc = cuda_lib.cotix_threefry2x32(None, 1, key[0], key[1], x0.data_ptr(), x1.data_ptr(), out.data_ptr())
        assert rc == 0
        got = out.cpu().numpy().view(np.uint32)
        assert tuple(int(v) for v in got) == want
    # bulk vs oracle
    rng = np.random.default_rng(1)
    x0 = rng.integers(0, 2**32, 4096, dtype=np.uint64).astype(np.uint32)
    x1 = rng.integers(0, 2**32, 4096, dtype=np.uint64).astype(np.uint32)
    d0 = torch.from_numpy(x0.view(np.int32)).cuda()
    d1 = torch.from_numpy(x1.view(np.int32)).cuda()
    out = torch.zeros((4096, 2), dtype=torch.int32).cuda()
    assert cuda_lib.cotix_threefry2x32(None, 4096, 0xDEADBEEF, 0x12345678, d0.data_ptr(), d1.data_ptr(), out.data_ptr()) == 0
    _eq(out.cpu().numpy().view(np.uint32), oracle.threefry2x32(0xDEADBEEF, 0x12345678, x0, x1), "threefry")


def test_philox_kat_on_device(cuda_lib, oracle):
    import torch

    kats = [((0, 0), (0, 0, 0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
            ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF,) * 4, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
            ((0xA4093822, 0x299F31D0), (0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344),
             (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1))]
    for key, ctr, want in kats:
        c = torch.from_numpy(np.array(ctr, np.uint32).view(np.int32)).cuda()
        out = torch.zeros(4, dtype=torch.int32).cuda()
        assert cuda_lib.cotix_philox4x32(None, 1, key[0], key[1], c.data_ptr(), out.data_ptr()) == 0
        assert tuple(int(v) for v in out.cpu().numpy().view(np.uint32)) == want
    rng = np.random.default_rng(2)
    ctr = rng.integers(0, 2**32, (1024, 4), dtype=np.uint64).astype(np.uint32)
    c = torch.from_numpy(ctr.view(np.int32)).cuda()
    out = torch.zeros((1024, 4), dtype=torch.int32).cuda()
    assert cuda_lib.cotix_philox4x32(None, 1024, 7, 9, c.data_ptr(), out.data_ptr()) == 0
    _eq(out.cpu().numpy().view(np.uint32), oracle.philox4x32(7, 9, ctr), "philox")


# ------------------------------------------------------------------------------------------
# leaf primitives
# ------------------------------------------------------------------------------------------
def test_support_and_order_clockwise(cuda_lib, oracle, exact_trig):
    import torch
    from parallax_b200 import shapes as sh

    rng = np.random.default_rng(3)
    a, b = random_pairs(rng, 4096)
    recs = np.concatenate([a, b])
    dirs = rng.normal(size=(recs.shape[0], 2)).astype(np.float32)
    dirs[::97] = 0.0          # zero direction: circle support is NaN
    dirs[5::101, 0] = np.nan  # NaN direction: polygon support is NaN
    got = sh.from_records(recs).get_support(dirs).cpu().numpy()
    _eq(got, oracle.support(recs, dirs), "support")
    # order_clockwise on shuffled polygons
    polys = recs[recs[:, 0] == 2].copy()
    for r in polys:
        n = int(r[1])
        v = r[4:4 + 2 * n].reshape(n, 2)
        r[4:4 + 2 * n] = v[rng.permutation(n)].reshape(-1)
    d = torch.from_numpy(polys).cuda()
    assert cuda_lib.cotix_order_clockwise(None, polys.shape[0], d.data_ptr()) == 0
    _eq(d.cpu().numpy(), oracle.order_clockwise(polys), "order_clockwise")


def test_euler_step_matches_reference_test(cuda_lib, oracle):
    """test/test_physics_solvers.py:9-40: a ball drifting right for 100 steps of dt=0.1."""
    import torch

    px = torch.full((1,), 0.1).cuda(); py = torch.full((1,), 0.1).cuda(); ang = torch.zeros(1).cuda()
    vx = torch.ones(1).cuda(); vy = torch.zeros(1).cuda(); w = torch.zeros(1).cuda()
    opx, opy, oang = np.float32([0.1]), np.float32([0.1]), np.float32([0.0])
    for _ in range(100):
        assert cuda_lib.cotix_euler_step(None, 1, px.data_ptr(), py.data_ptr(), vx.data_ptr(), vy.data_ptr(),
                                         ang.data_ptr(), w.data_ptr(), 0.1) == 0
        opx, opy, oang = oracle.euler_step(opx, opy, [1.0], [0.0], oang, [0.0], 0.1)
    assert px.item() > 2e-1 and abs(py.item() - 0.1) < 1e-2
    _eq(px.cpu().numpy(), opx, "px")


def test_resolve_collision(cuda_lib, oracle):
    import torch

    rng = np.random.default_rng(4)
    n = 8192
    b1 = rng.normal(size=(n, 10)).astype(np.float32)
    b2 = rng.normal(size=(n, 10)).astype(np.float32)
    for b in (b1, b2):
        b[:, 0] = rng.uniform(0.5, 30, n)   # mass
        b[:, 1] = rng.uniform(0.5, 30, n)   # inertia
        b[:, 8] = rng.uniform(0, 1, n)      # elasticity
        b[:, 9] = rng.uniform(0, 1, n)      # friction
    b2[::5, 0] = np.inf; b2[::5, 1] = np.inf  # static bodies
    pen = rng.normal(scale=0.1, size=(n, 2)).astype(np.float32)
    cp = rng.normal(size=(n, 2)).astype(np.float32)
    cp[::7] = np.nan       # no contact -> untouched
    pen[3::11] = 0.0       # zero penetration -> NaN normal, as in the reference
    d1, d2 = torch.from_numpy(b1).cuda(), torch.from_numpy(b2).cuda()
    dp, dc = torch.from_numpy(pen).cuda(), torch.from_numpy(cp).cuda()
    assert cuda_lib.cotix_resolve_collision(None, n, d1.data_ptr(), d2.data_ptr(), dp.data_ptr(), dc.data_ptr(),
                                            0.3, 0.01) == 0
    o1, o2 = oracle.resolve_collision(b1, b2, pen, cp)
    _eq(d1.cpu().numpy(), o1, "body1")
    _eq(d2.cpu().numpy(), o2, "body2")


# ------------------------------------------------------------------------------------------
# narrowphase
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("sweep_mode", ["single", "pipeline"])
@pytest.mark.parametrize("seed,n", [(10, 200_000), (11, 50_001)])
def test_narrowphase_bit_exact(cuda_lib, oracle, exact_trig, seed, n, sweep_mode):
    from parallax_b200 import shapes as sh

    rng = np.random.default_rng(seed)
    a, b = random_pairs(rng, n)
    ci = sh.narrowphase(sh.from_records(a), sh.from_records(b), mode=sweep_mode)
    ex, pen, cp, st = oracle.narrowphase(a, b)
    _eq(ci.exists.cpu().numpy(), ex.astype(bool), "exists")
    _eq(ci.penetration_vector.cpu().numpy(), pen, "pen")
    _eq(ci.contact_point.cpu().numpy(), cp, "cp")
    _eq(ci.status.cpu().numpy(), st, "status")
    assert 0.2 < ex.mean() < 0.95  # the sweep exercises both hits and misses


@pytest.mark.parametrize("sweep_mode", ["single", "pipeline"])
def test_narrowphase_edge_cases(cuda_lib, oracle, exact_trig, sweep_mode):
    from parallax_b200 import shapes as sh

    o = oracle
    sq = o.polygon_rec([[0, 0], [1, 0], [1, 1], [0, 1]])
    cases = [
        (sq, sq),                                                     # identical polygons (self pair)
        (sq, o.polygon_rec([[1, 0], [2, 0], [2, 1], [1, 1]])),        # touching along an edge
        (sq, o.polygon_rec([[1, 1], [2, 1], [2, 2], [1, 2]])),        # touching at a corner
        (sq, o.polygon_rec([[5, 5], [6, 5], [6, 6]])),                # far apart
        (sq, o.polygon_rec([[0.25, 0.25], [0.75, 0.25], [0.5, 0.75]])),  # fully contained
        (o.circle_rec(1.0, [0, 0]), o.circle_rec(2.0, [0, 0])),       # concentric circles
        (o.circle_rec(1.0, [0, 0]), o.circle_rec(1.0, [2, 0])),       # touching circles
        (o.circle_rec(0.5, [0.5, 0.5]), sq),                          # circle centred in polygon
        (o.aabb_rec([0, 0], [1, 1]), o.aabb_rec([1, 0], [2, 1])),     # touching boxes
        (o.aabb_rec([0, 0], [1, 1]), o.aabb_rec([0, 0], [1, 1])),     # identical boxes
        (o.circle_rec(1.0, [2, 2]), o.aabb_rec([0, 0], [1.5, 1.5])),  # circle near a box corner
        (sq, o.circle_rec(1.0, [0, 0])),                              # wrong order: not a dispatch key
        (o.polygon_rec([[np.nan, 0], [1, 0], [1, 1]]), sq),           # NaN vertex
    ]
    a = np.stack([c[0] for c in cases])
    b = np.stack([c[1] for c in cases])
    ci = sh.narrowphase(sh.from_records(a), sh.from_records(b), mode=sweep_mode)
    ex, pen, cp, st = o.narrowphase(a, b)
    _eq(ci.exists.cpu().numpy(), ex.astype(bool), "exists")
    _eq(ci.penetration_vector.cpu().numpy(), pen, "pen")
    _eq(ci.contact_point.cpu().numpy(), cp, "cp")
    _eq(ci.status.cpu().numpy(), st, "status")
    assert st[11] & 0x80
    # empty batch
    e = sh.narrowphase(sh.from_records(np.zeros((0, 16), np.float32)), sh.from_records(np.zeros((0, 16), np.float32)),
                       mode=sweep_mode)
    assert e.exists.numel() == 0


def test_narrowphase_host_entry_point(cuda_lib, oracle, exact_trig):
    from parallax_b200 import shapes as sh

    rng = np.random.default_rng(12)
    a, b = random_pairs(rng, 20_000)
    ex, pen, cp, st = sh.narrowphase_host(a, b)
    oex, open_, ocp, ost = oracle.narrowphase(a, b)
    _eq(ex, oex, "exists"); _eq(pen, open_, "pen"); _eq(cp, ocp, "cp"); _eq(st, ost, "status")


def test_typed_contact_functions_and_reference_cases(cuda_lib, oracle, exact_trig):
    """The reference's own parametrized circle_vs_aabb inputs (test/test_collisions.py:229-272)
    through the CUDA path: result equals the oracle and satisfies the reference's invariants
    (contact point in both shapes; moving the circle by the penetration vector separates)."""
    from parallax_b200 import shapes as sh

    from ref_cases import CIRCLE_VS_AABB_CASES

    r = np.float32([c[0] for c in CIRCLE_VS_AABB_CASES])
    pos = np.float32([c[1] for c in CIRCLE_VS_AABB_CASES])
    lo = np.float32([c[2] for c in CIRCLE_VS_AABB_CASES])
    hi = np.float32([c[3] for c in CIRCLE_VS_AABB_CASES])
    a, b = sh.Circle(r, pos), sh.AABB(lo, hi)
    info = sh.circle_vs_aabb(a, b)
    oex, open_, ocp, _ = oracle.narrowphase(a.recs.cpu().numpy(), b.recs.cpu().numpy())
    _eq(info.penetration_vector.cpu().numpy(), open_, "pen")
    _eq(info.contact_point.cpu().numpy(), ocp, "cp")
    assert info.exists.all()
    moved = a.move(info.penetration_vector)
    after = sh.circle_vs_aabb(moved, b)
    assert (after.penetration_vector.norm(dim=1) < 1e-5).all()
    with pytest.raises(TypeError):
        sh.circle_vs_circle(a, b)


# ------------------------------------------------------------------------------------------
# LunarLander
# ------------------------------------------------------------------------------------------
MODES = ["fused", "pipeline"]


def _rollout_both(oracle, ll, P, state0, steps, actions=None, check_every=1, dedupe=True, mode="fused"):
    import torch

    st = ll.LunarState.from_numpy(state0)
    ref = state0.copy()
    for t in range(steps):
        act = None if actions is None else actions[t]
        dact = None if act is None else torch.from_numpy(act).cuda()
        st, rew, done = ll.step(st, dact, P, mode=mode)
        ref, orew, odone = oracle.lunar_step(ref, act, P, dedupe=dedupe)
        if t % check_every == 0 or t == steps - 1:
            _eq(st.numpy(), ref, f"state at step {t}")
            _eq(rew.cpu().numpy(), orew, f"reward at step {t}")
            _eq(done.cpu().numpy(), odone, f"done at step {t}")
    return st, ref


def test_lunar_default_params_match_oracle(cuda_lib, oracle, exact_trig):
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    Q = oracle.default_lunar_params()
    assert bytes(P) == bytes(Q)


def test_lunar_reset_bit_exact(cuda_lib, oracle, exact_trig):
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    keys = lunar_keys(1000)
    keys[500:, 0] = ((np.arange(500, dtype=np.uint64) * 2654435761) % (2**32)).astype(np.uint32)
    _eq(ll.reset(keys, P).numpy(), oracle.lunar_reset(keys, P), "reset")
    P2 = ll.default_params(reset_perturb=1)
    _eq(ll.reset(keys, P2).numpy(), oracle.lunar_reset(keys, P2), "reset perturbed")
    P3 = ll.default_params(reset_perturb=1, rng_mode=1)
    _eq(ll.reset(keys, P3).numpy(), oracle.lunar_reset(keys, P3), "reset philox")


@pytest.mark.parametrize("mode", MODES)
def test_lunar_free_fall_and_landing_bit_exact(cuda_lib, oracle, exact_trig, mode):
    """The reference scenario itself (examples/test_viz.py:46-48): reset, then 1200 unactuated
    steps — free fall, touch-down on the pad near step ~800, rebound."""
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    s0 = oracle.lunar_reset(lunar_keys(32), P)
    st, ref = _rollout_both(oracle, ll, P, s0, 1200, check_every=50, mode=mode)
    assert ref[1].max() < 3.0  # every lander came down from y = 5 and hit the terrain


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("rng_mode", [0, 1])
def test_lunar_contact_rich_bit_exact(cuda_lib, oracle, exact_trig, rng_mode, mode):
    """BASELINE config 3 style states (perturbed poses, many start in or near the terrain)."""
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(rng_mode=rng_mode)
    rng = np.random.default_rng(20 + rng_mode)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(512), P), rng, P)
    _rollout_both(oracle, ll, P, s0, 150, check_every=10, mode=mode)


@pytest.mark.parametrize("mode", MODES)
def test_lunar_literal_oracle_small(cuda_lib, oracle, exact_trig, mode):
    """Against the LITERAL oracle (all 306 narrowphase evaluations, no memoisation)."""
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    rng = np.random.default_rng(30)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(48), P), rng, P)
    _rollout_both(oracle, ll, P, s0, 40, check_every=5, dedupe=False, mode=mode)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("rng_mode", [0, 1])
def test_lunar_env_logic_bit_exact(cuda_lib, oracle, exact_trig, rng_mode, mode):
    """Actions, judge, termination and auto-reset (new logic; oracle implements the same spec)."""
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(rng_mode=rng_mode, judge_enabled=1, auto_reset=1, reset_perturb=1,
                          max_episode_steps=90)
    rng = np.random.default_rng(40 + rng_mode)
    B, T = 384, 200
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B), P), rng, P)
    actions = rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)
    st, ref = _rollout_both(oracle, ll, P, s0, T, actions=actions, check_every=10, mode=mode)
    episodes = (ref[33].view(np.uint32) >> 8)
    assert episodes.max() >= 2  # several auto-resets happened


def test_lunar_substeps_equal_repeated_steps(cuda_lib, oracle, exact_trig):
    import torch
    from parallax_b200 import lunar_lander as ll

    P1 = ll.default_params()
    P4 = ll.default_params(substeps=4)
    rng = np.random.default_rng(50)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(256), P1), rng, P1)
    a = ll.LunarState.from_numpy(s0)
    b = ll.LunarState.from_numpy(s0)
    for _ in range(10):
        a, _, _ = ll.step(a, None, P4)
        for _ in range(4):
            b, _, _ = ll.step(b, None, P1)
    # an env that finishes (here: NaN state) freezes for the rest of a K=4 launch, whereas four
    # K=1 launches without auto-reset keep stepping it; compare the envs that never finished
    live = ~np.isnan(b.numpy()[:18]).any(axis=0)
    assert live.mean() > 0.8
    _eq(a.numpy()[:, live], b.numpy()[:, live], "K=4 vs 4 x K=1")
    ref = s0
    for _ in range(10):
        ref, _, _ = oracle.lunar_step(ref, None, P4)
    _eq(a.numpy(), ref, "K=4 vs oracle")


@pytest.mark.parametrize("mode", MODES)
def test_lunar_inplace_and_obs(cuda_lib, oracle, exact_trig, mode):
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(judge_enabled=1, auto_reset=1, max_episode_steps=20)
    rng = np.random.default_rng(60)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(128), P), rng, P)
    st = ll.LunarState.from_numpy(s0)
    ref = s0
    for _ in range(45):
        st, rew, done, obs = ll.step(st, None, P, inplace=True, want_obs=True, mode=mode)
        ref, orew, odone, oobs = oracle.lunar_step(ref, None, P, want_obs=True)
    _eq(st.numpy(), ref, "state"); _eq(obs.cpu().numpy(), oobs, "obs"); _eq(done.cpu().numpy(), odone, "done")


@pytest.mark.parametrize("mode", MODES)
def test_lunar_host_env_and_graph(cuda_lib, oracle, exact_trig, mode):
    """The HOST-facing handle (H2D action -> kernel -> D2H obs/reward/done) and CUDA-graph replay."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=64)
    B = 300
    env = ll.LunarLanderEnv(B, P)
    env.set_mode(mode)
    keys = lunar_keys(B, offset=7)
    env.reset(keys)
    ref = oracle.lunar_reset(keys, P)
    _eq(env.get_state(), ref, "reset")
    rng = np.random.default_rng(70)
    for t in range(30):
        act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
        obs, rew, done = env.step(torch.from_numpy(act))
        ref, orew, odone, oobs = oracle.lunar_step(ref, act, P, want_obs=True)
        _eq(obs.numpy(), oobs, f"obs {t}"); _eq(rew.numpy(), orew, f"rew {t}"); _eq(done.numpy(), odone, f"done {t}")
    _eq(env.get_state(), ref, "state after host steps")
    # graph replay of 16 unactuated steps, twice
    for _ in range(2):
        env.rollout_graph(16, None)
    env.synchronize()
    for _ in range(32):
        ref, _, _ = oracle.lunar_step(ref, None, P)
    _eq(env.get_state(), ref, "state after graph replay")
    env.close()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("chunks", [3, 16])
def test_lunar_host_env_chunked_overlap(cuda_lib, oracle, exact_trig, mode, chunks):
    """The software-pipelined host step (ranges of envs on three streams) gives the oracle's
    observations, rewards, dones and state; ragged last range (B is not a multiple of 32)."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=16)
    B = 1000 + 7
    env = ll.LunarLanderEnv(B, P)
    env.set_mode(mode)
    env.set_chunks(chunks)
    env.set_host_graphs(chunks == 3)  # odd steps pass pinned actions: those replay a captured graph
    keys = lunar_keys(B, offset=11)
    env.reset(keys)
    ref = oracle.lunar_reset(keys, P)
    rng = np.random.default_rng(71)
    for t in range(24):
        act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
        obs, rew, done = env.step(torch.from_numpy(act).pin_memory() if t % 2 else act)
        ref, orew, odone, oobs = oracle.lunar_step(ref, act, P, want_obs=True)
        _eq(obs.numpy(), oobs, f"obs {t}"); _eq(rew.numpy(), orew, f"rew {t}"); _eq(done.numpy(), odone, f"done {t}")
    _eq(env.get_state(), ref, "state after chunked host steps")
    with pytest.raises(Exception):
        env.set_chunks(17)
    env.close()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("rng_mode", [0, 1])
def test_lunar_substeps_env_logic_bit_exact(cuda_lib, oracle, exact_trig, rng_mode, mode):
    """K = 5 sub-steps per call with actions, judge and auto-reset (the WFE_scale loop of
    cotix/_envs.py:72-103): rewards add up over the sub-steps, an env that finishes mid-call is
    frozen and reset at the end of the call.  Fused kernel and pipeline against the oracle."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=23, substeps=5,
                          rng_mode=rng_mode)
    B = 1500
    rng = np.random.default_rng(52)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B, offset=5), P), rng, P)
    st, ref = ll.LunarState.from_numpy(s0), s0
    dones = 0
    for t in range(14):
        act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
        st, rew, done, obs = ll.step(st, torch.from_numpy(act).cuda(), P, inplace=(t % 2 == 0), want_obs=True, mode=mode)
        ref, orew, odone, oobs = oracle.lunar_step(ref, act, P, want_obs=True)
        _eq(st.numpy(), ref, f"state {t}"); _eq(rew.cpu().numpy(), orew, f"reward {t}")
        _eq(done.cpu().numpy(), odone, f"done {t}"); _eq(obs.cpu().numpy(), oobs, f"obs {t}")
        dones += int((odone != 0).sum())
    assert dones > 200  # envs did finish in the middle of calls


def test_lunar_libm_oracle_tolerance(cuda_lib, oracle):
    """Against the oracle with libm trig (the independent definition): one step from IDENTICAL
    input states agrees within 1e-5 relative (north_star), flags identical."""
    from parallax_b200 import lunar_lander as ll

    oracle.set_trig_mode(0)
    P = ll.default_params()
    rng = np.random.default_rng(80)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(4096), P), rng, P)
    st, _, _ = ll.step(ll.LunarState.from_numpy(s0), None, P)
    ref, _, _ = oracle.lunar_step(s0, None, P)
    got = st.numpy()
    num = np.abs(got[:18] - ref[:18])
    den = np.maximum(np.abs(ref[:18]), 1.0)
    bad = (num / den > 1e-5).any(axis=0)
    # a decision that flips within epsilon of its threshold changes a whole contact; allow a
    # tiny fraction of such envs and report it
    assert bad.mean() < 2e-3, f"{bad.sum()} of {bad.size} envs beyond 1e-5"
    _eq(got[30:32].view(np.uint32), ref[30:32].view(np.uint32), "key chain")


def test_pipeline_equals_fused_large_batch(cuda_lib, oracle, exact_trig):
    """BASELINE config 3 at a size where the item queue is busy: the wavefront pipeline and the
    fused kernel agree bit for bit over 60 steps with actions, judge and auto-reset (and a
    sample of envs agrees with the oracle)."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=40)
    B = 1 << 16
    rng = np.random.default_rng(90)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P).numpy(), rng, P)
    a = ll.LunarState.from_numpy(s0)
    b = ll.LunarState.from_numpy(s0)
    ref = s0[:, :512].copy()
    for t in range(60):
        act = torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).cuda()
        a, ra, da, oa = ll.step(a, act, P, inplace=(t % 2 == 0), want_obs=True, mode="fused")
        b, rb, db, ob = ll.step(b, act, P, inplace=(t % 2 == 1), want_obs=True, mode="pipeline")
        ref, rr, dr, orr = oracle.lunar_step(ref, act[:, :512].cpu().numpy(), P, want_obs=True)
        if t % 10 == 9 or t < 2:
            _eq(b.numpy(), a.numpy(), f"state {t}"); _eq(rb.cpu().numpy(), ra.cpu().numpy(), f"reward {t}")
            _eq(db.cpu().numpy(), da.cpu().numpy(), f"done {t}"); _eq(ob.cpu().numpy(), oa.cpu().numpy(), f"obs {t}")
            _eq(b.numpy()[:, :512], ref, f"oracle state {t}"); _eq(ob.cpu().numpy()[:, :512], orr, f"oracle obs {t}")
    assert (b.numpy()[33].view(np.uint32) >> 8).max() >= 2


def test_pipeline_argument_checks(cuda_lib):
    import torch
    from parallax_b200 import lunar_lander as ll

    P1 = ll.default_params()
    st = ll.reset(lunar_keys(64), P1)
    rc = cuda_lib.cotix_lunar_step_ws(None, 64, st.data.data_ptr(), st.data.data_ptr(), 64, None, None, None,
                                      None, C.byref(P1), None, 0)
    assert rc == -1 and b"workspace" in cuda_lib.cotix_last_error()
    assert cuda_lib.cotix_lunar_workspace_bytes(64) > 0


def test_missing_library_fails_loudly(monkeypatch, cuda_lib):
    import parallax_b200

    monkeypatch.setattr(parallax_b200._lib, "_lib", None)
    monkeypatch.setattr(parallax_b200._lib, "LIB_PATH", "/nonexistent/libcotix_b200.so")
    with pytest.raises(parallax_b200.CotixError):
        parallax_b200._lib.lib()
