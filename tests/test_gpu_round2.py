"""GPU parity tests added in round 2: the multi-range / multi-step / graph-replayed pipeline, the
item-queue overflow branch, sliced (ld != B) states, parity AT THE BENCHED SIZES (1 048 576 envs,
16 777 216 pairs) on strided subsets, and the 100-step-horizon tolerance against the libm-trig
oracle (north_star: "within a stated tolerance over a 100-step horizon"; stated in DESIGN.md §4)."""
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


@pytest.fixture()
def pipe_cfg():
    """Restores the pipeline's process-wide launch configuration after the test."""
    from parallax_b200 import lunar_lander as ll

    yield ll.pipeline_config
    ll.pipeline_config(split=0, pdl=1, priorities=1, item_capacity=0)


def _env_params(ll, **kw):
    base = dict(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=24)
    base.update(kw)
    return ll.default_params(**base)


@pytest.mark.parametrize("split,pdl,prio", [(1, 1, 1), (2, 1, 1), (3, 0, 1), (4, 1, 0), (7, 1, 1), (16, 0, 0)])
def test_steps_ranges_match_oracle(cuda_lib, oracle, exact_trig, pipe_cfg, split, pdl, prio):
    """cotix_lunar_steps_ws: n steps in one enqueue, the batch cut into `split` env ranges that run
    on their own streams without joining between steps (ragged last range), actions cycling —
    bit-identical to the oracle stepped one step at a time, for every launch configuration."""
    import torch
    from parallax_b200 import lunar_lander as ll

    pipe_cfg(split=split, pdl=pdl, priorities=prio)
    P = _env_params(ll)
    B, T, n_act = 5000 + 13, 9, 4
    rng = np.random.default_rng(100 + split)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B, offset=3), P), rng, P)
    acts = rng.uniform(-1, 1, size=(n_act, 2, B)).astype(np.float32)
    st = ll.LunarState.from_numpy(s0)
    rew, done, obs = ll.steps(st, torch.from_numpy(acts).cuda(), T, P, want_obs=True, mode="pipeline")
    ref = s0
    for t in range(T):
        ref, orew, odone, oobs = oracle.lunar_step(ref, acts[t % n_act], P, want_obs=True)
    _eq(st.numpy(), ref, "state"); _eq(rew.cpu().numpy(), orew, "reward"); _eq(done.cpu().numpy(), odone, "done")
    # obs of an env that was auto-reset in the last step is the reset observation in both
    _eq(obs.cpu().numpy(), oobs, "obs")
    c = ll.pipeline_counters(st.data.device, B)
    assert c["overflow"] == 0 and c["items"] > 0 and c["contact_envs"] > 0


@pytest.mark.parametrize("mode", ["fused", "pipeline"])
def test_steps_graph_replay(cuda_lib, oracle, exact_trig, pipe_cfg, mode):
    """StepsGraph: the n-step enqueue captured once through cotix_graph_* and replayed three times
    = 3 n oracle steps; the graph holds one node per kernel / memset of every range."""
    import torch
    from parallax_b200 import lunar_lander as ll

    pipe_cfg(split=3)
    P = _env_params(ll)
    B, T, n_act = 2048 + 5, 6, 3
    rng = np.random.default_rng(110)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B), P), rng, P)
    acts = rng.uniform(-1, 1, size=(n_act, 2, B)).astype(np.float32)
    st = ll.LunarState.from_numpy(s0)
    g = ll.StepsGraph(st, torch.from_numpy(acts).cuda(), T, P, want_obs=True, mode=mode)
    info = g.info()
    if mode == "pipeline":
        assert info["kernel_launches"] == 3 * T * 6 and info["nodes"] >= 3 * T * 6 + 3  # + one memset per range
    else:
        assert info["kernel_launches"] == 1  # small batch, no workspace: all T steps in one launch (lunar_rollout.cu)
    import parallax_b200
    n0 = parallax_b200.launch_count()
    for _ in range(3):
        g.launch()
    torch.cuda.synchronize()
    assert parallax_b200.launch_count() - n0 == 3 * info["kernel_launches"]
    ref = s0
    for t in range(3 * T):
        ref, orew, odone, oobs = oracle.lunar_step(ref, acts[(t % T) % n_act], P, want_obs=True)
    _eq(st.numpy(), ref, "state after 3 replays"); _eq(g.reward.cpu().numpy(), orew, "reward")
    _eq(g.done.cpu().numpy(), odone, "done"); _eq(g.obs.cpu().numpy(), oobs, "obs")
    g.close()


def test_graph_capture_abi(cuda_lib, oracle, exact_trig):
    """cotix_graph_begin_capture / end_capture / launch / info / destroy straight through the C-ABI
    around two device entry points (reset + 4 steps; a small batch without workspace runs them in one launch)."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    B = 512
    keys = lunar_keys(B, offset=9)
    kd = torch.from_numpy(keys.view(np.int32)).cuda()
    st = torch.zeros((36, B), dtype=torch.float32, device="cuda")
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    g = C.c_void_p()
    assert cuda_lib.cotix_graph_begin_capture(s.cuda_stream) == 0
    assert cuda_lib.cotix_graph_begin_capture(s.cuda_stream) == -1  # one open capture per thread
    assert cuda_lib.cotix_lunar_reset(s.cuda_stream, B, kd.data_ptr(), st.data_ptr(), B, C.byref(P)) == 0
    assert cuda_lib.cotix_lunar_steps_ws(s.cuda_stream, B, st.data_ptr(), B, None, 0, 4, None, None, None,
                                         C.byref(P), None, 0) == 0
    assert cuda_lib.cotix_graph_end_capture(s.cuda_stream, C.byref(g)) == 0
    torch.cuda.synchronize()
    assert float(st.abs().sum()) == 0.0  # captured, not executed
    n, k = C.c_int32(), C.c_int64()
    assert cuda_lib.cotix_graph_info(g, C.byref(n), C.byref(k)) == 0 and k.value == 2 and n.value == 2
    assert cuda_lib.cotix_graph_launch(g, s.cuda_stream) == 0
    s.synchronize()
    ref = oracle.lunar_reset(keys, P)
    for _ in range(4):
        ref, _, _ = oracle.lunar_step(ref, None, P)
    _eq(st.cpu().numpy(), ref, "reset + 4 steps through one graph")
    assert cuda_lib.cotix_graph_destroy(g) == 0
    assert cuda_lib.cotix_graph_end_capture(s.cuda_stream, C.byref(g)) == -1  # nothing open


def test_item_queue_overflow_branch(cuda_lib, oracle, exact_trig, pipe_cfg):
    """The LunarLander item-queue overflow branch (lunar_pipeline.cu, BASE_INLINE): with the queue
    capacity forced far below the number of candidate pairs most candidate envs evaluate their pairs
    inside the contact stage — results must stay bit-identical to the fused kernel and the oracle."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = _env_params(ll)
    B = 6000
    rng = np.random.default_rng(120)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B), P), rng, P)
    a, b, ref = ll.LunarState.from_numpy(s0), ll.LunarState.from_numpy(s0), s0
    seen = 0
    for cap in (1, 37, 150):
        pipe_cfg(split=2, item_capacity=cap)
        for t in range(4):
            act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
            ad = torch.from_numpy(act).cuda()
            a, ra, da = ll.step(a, ad, P, mode="fused")
            b, rb, db, ob = ll.step(b, ad, P, want_obs=True, mode="pipeline")
            ref, orew, odone, oobs = oracle.lunar_step(ref, act, P, want_obs=True)
            c = ll.pipeline_counters(b.data.device, B)
            assert c["overflow"] > 0, (cap, t, c)
            seen += c["overflow"]
            _eq(b.numpy(), a.numpy(), f"pipeline vs fused, cap {cap} step {t}")
            _eq(b.numpy(), ref, f"pipeline vs oracle, cap {cap} step {t}")
            _eq(rb.cpu().numpy(), orew, "reward"); _eq(db.cpu().numpy(), odone, "done"); _eq(ob.cpu().numpy(), oobs, "obs")
    assert seen > 1000


@pytest.mark.parametrize("mode", ["fused", "pipeline"])
def test_sliced_state_keeps_parent_stride(cuda_lib, oracle, exact_trig, mode):
    """A LunarState that is a column slice of a larger batch (stride(0) = parent batch): step,
    steps and step_vjp must address every [.][ld] buffer with the slice's true leading dimension
    and leave the parent's other columns untouched."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = _env_params(ll, auto_reset=0, judge_enabled=0, max_episode_steps=0)
    Bt, lo, n = 1536, 256, 777
    rng = np.random.default_rng(130)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(Bt), P), rng, P)
    parent = torch.from_numpy(s0).cuda()
    view = ll.LunarState(parent[:, lo:lo + n])
    act = rng.uniform(-1, 1, size=(2, n)).astype(np.float32)
    ad = torch.from_numpy(act).cuda()
    ref, orew, odone, oobs = oracle.lunar_step(s0[:, lo:lo + n].copy(), act, P, want_obs=True)
    # functional step: result has the slice's shape; the parent is unchanged
    ns, rew, done, obs = ll.step(view, ad, P, want_obs=True, mode=mode)
    _eq(ns.numpy(), ref, "functional"); _eq(obs.cpu().numpy(), oobs, "obs")
    _eq(parent.cpu().numpy(), s0, "parent untouched by the functional step")
    # in place: only the slice's columns change
    ll.step(view, ad, P, inplace=True, mode=mode)
    got = parent.cpu().numpy()
    _eq(got[:, lo:lo + n], ref, "in place"); _eq(got[:, :lo], s0[:, :lo], "left of the slice")
    _eq(got[:, lo + n:], s0[:, lo + n:], "right of the slice")
    # n steps in one enqueue on the slice
    ll.steps(view, ad[None], 3, P, mode=mode)
    for _ in range(3):
        ref, _, _ = oracle.lunar_step(ref, act, P)
    got = parent.cpu().numpy()
    _eq(got[:, lo:lo + n], ref, "steps in place"); _eq(got[:, :lo], s0[:, :lo], "left of the slice")
    # the adjoint on a slice equals the adjoint on a contiguous copy
    gn = torch.from_numpy(rng.normal(size=(18, n)).astype(np.float32)).cuda()
    cont = ll.LunarState(view.data.contiguous())
    g1, a1 = ll.step_vjp(view, ad, gn, P)
    g2, a2 = ll.step_vjp(cont, ad, gn, P)
    _eq(g1.cpu().numpy(), g2.cpu().numpy(), "vjp state"); _eq(a1.cpu().numpy(), a2.cpu().numpy(), "vjp action")
    with pytest.raises(ValueError):
        ll.step(ll.LunarState(parent.T[:36]), None, P)  # rows not contiguous


def test_workspace_cache_is_per_batch_shape(cuda_lib, oracle, exact_trig):
    """Alternating batch shapes keep their own cached workspace and pipeline_counters
    reads the one it is asked for."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = _env_params(ll)
    rng = np.random.default_rng(140)
    sa = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(3000), P), rng, P)
    sb = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(1000), P), rng, P)
    a, b = ll.LunarState.from_numpy(sa), ll.LunarState.from_numpy(sb)
    for t in range(3):
        a, _, _ = ll.step(a, None, P, mode="pipeline")
        wa = ll._workspace(a.data.device, 3000).data_ptr()
        b, _, _ = ll.step(b, None, P, mode="pipeline")
        assert ll._workspace(a.data.device, 3000).data_ptr() == wa  # not reallocated by the other shape
        sa, _, _ = oracle.lunar_step(sa, None, P)
        sb, _, _ = oracle.lunar_step(sb, None, P)
    _eq(a.numpy(), sa, "batch a"); _eq(b.numpy(), sb, "batch b")
    ca, cb = ll.pipeline_counters(a.data.device, 3000), ll.pipeline_counters(a.data.device, 1000)
    assert ca["items"] > cb["items"] > 0


def test_host_env_without_obs(cuda_lib, oracle, exact_trig):
    """cotix_lunar_env_step_host with obs == NULL (want_obs=False): reward / done / state as with
    obs; the multi-range host step in pipeline mode with pinned and pageable actions."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = _env_params(ll)
    B = 4000 + 3
    env = ll.LunarLanderEnv(B, P)
    env.set_mode("pipeline")
    env.set_chunks(5)
    keys = lunar_keys(B, offset=2)
    env.reset(keys)
    ref = oracle.lunar_reset(keys, P)
    rng = np.random.default_rng(150)
    for t in range(10):
        act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
        want = t % 3 != 0
        obs, rew, done = env.step(torch.from_numpy(act).pin_memory() if t % 2 else act, want_obs=want)
        ref, orew, odone, oobs = oracle.lunar_step(ref, act, P, want_obs=True)
        assert (obs is None) == (not want)
        if want:
            _eq(obs.numpy(), oobs, f"obs {t}")
        _eq(rew.numpy(), orew, f"rew {t}"); _eq(done.numpy(), odone, f"done {t}")
    _eq(env.get_state(), ref, "state")
    env.close()


# ------------------------------------------------------------------------------------------
# parity at the benched sizes
# ------------------------------------------------------------------------------------------
def test_bench_size_lunar_parity(cuda_lib, oracle, exact_trig):
    """BASELINE config 3 at bench.py's own size and workload: 1 048 576 envs x 32 steps (actions,
    judge, auto-reset) through the graph-replayed multi-range pipeline; a strided subset of 4096
    envs is checked bit for bit against the oracle (envs are independent: the oracle steps just the
    subset from the same initial state) and the item queues must not overflow."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
    B, T, n_act = 1 << 20, 32, 4
    rng = np.random.default_rng(1000)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P).numpy(), rng, P)
    acts = np.random.default_rng(1).uniform(-1, 1, size=(n_act, 2, B)).astype(np.float32)
    st = ll.LunarState.from_numpy(s0)
    g = ll.StepsGraph(st, torch.from_numpy(acts).cuda(), T // 2, P, want_obs=True)
    g.launch()
    g.launch()
    torch.cuda.synchronize()
    c = ll.pipeline_counters(st.data.device, B)
    assert c["overflow"] == 0 and c["items"] > 50000, c
    sub = np.arange(0, B, B // 4096)[:4096]
    ref = np.ascontiguousarray(s0[:, sub])
    for t in range(T):
        ref, orew, odone, oobs = oracle.lunar_step(ref, np.ascontiguousarray(acts[(t % (T // 2)) % n_act][:, sub]),
                                                   P, want_obs=True)
    got = st.numpy()
    _eq(got[:, sub], ref, "state of the strided subset")
    _eq(g.reward.cpu().numpy()[sub], orew, "reward"); _eq(g.done.cpu().numpy()[sub], odone, "done")
    _eq(g.obs.cpu().numpy()[:, sub], oobs, "obs")
    g.close()


def test_bench_size_sweep_parity(cuda_lib, oracle):
    """BASELINE config 2 at bench.py's own size: 16 777 216 pairs through the wavefront sweep; a
    strided subset of 200 000 pairs is checked bit for bit against the oracle."""
    import torch
    from parallax_b200 import shapes as sh

    n = 1 << 24
    rng = np.random.default_rng(2)
    chunk = 1 << 21
    a = torch.empty((n, 16), dtype=torch.float32, device="cuda")
    b = torch.empty((n, 16), dtype=torch.float32, device="cuda")
    sub = np.arange(0, n, 83)[:200000]
    ha_s, hb_s = [], []
    for s in range(0, n, chunk):
        ha, hb = random_pairs(rng, min(chunk, n - s))
        a[s:s + ha.shape[0]] = torch.from_numpy(ha).cuda()
        b[s:s + hb.shape[0]] = torch.from_numpy(hb).cuda()
        m = sub[(sub >= s) & (sub < s + chunk)] - s
        ha_s.append(ha[m]); hb_s.append(hb[m])
    ci = sh.narrowphase(sh.ShapeBatch(a), sh.ShapeBatch(b), mode="pipeline")
    ex, pen, cp, _ = oracle.narrowphase(np.concatenate(ha_s), np.concatenate(hb_s))
    idx = torch.from_numpy(sub).cuda()
    _eq(ci.exists[idx].cpu().numpy(), ex.astype(bool), "exists")
    _eq(ci.penetration_vector[idx].cpu().numpy(), pen, "penetration vector")
    _eq(ci.contact_point[idx].cpu().numpy(), cp, "contact point")
    assert 0.2 < float(ci.exists.float().mean()) < 0.9


# ------------------------------------------------------------------------------------------
# 100-step horizon against the independent-trig (libm) oracle
# ------------------------------------------------------------------------------------------
def test_horizon_100_steps_libm_oracle(cuda_lib, oracle):
    """north_star: "positions and velocities ... within a stated tolerance over a 100-step horizon".
    The CUDA engine (its own cx_trig.h sin/cos/atan2) and the oracle with libm trig (an independent
    definition, <= 3 ulp apart) start from IDENTICAL states and run 100 steps of the reference step
    (contact-rich perturbed states).  The stated tolerance is tests/horizon.py's."""
    from horizon import check_horizon, HORIZON_STEPS
    from parallax_b200 import lunar_lander as ll

    oracle.set_trig_mode(0)
    P = ll.default_params(judge_enabled=1)  # judge on so that leg-contact flags are reported in obs
    B = 4096
    rng = np.random.default_rng(160)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B), P), rng, P)
    st, ref = ll.LunarState.from_numpy(s0), s0
    same_flags = np.ones(B, bool)
    per_step_bad = []
    for t in range(HORIZON_STEPS):
        prev_got = st.numpy()
        st, rew, done, obs = ll.step(st, None, P, want_obs=True)
        ref, orew, odone, oobs = oracle.lunar_step(ref, None, P, want_obs=True)
        same_flags &= (obs.cpu().numpy()[6:8] == oobs[6:8]).all(axis=0) & (done.cpu().numpy() == odone)
        if t < 5:
            # per step from identical states (north_star's 1e-5): re-synchronise the first steps
            one, _, _ = oracle.lunar_step(prev_got, None, P)
            got = st.numpy()
            with np.errstate(invalid="ignore"):
                d = np.abs(got[:18] - one[:18]) / np.maximum(np.abs(one[:18]), 1.0)
            per_step_bad.append(float((np.nan_to_num(d, nan=0.0) > 1e-5).any(axis=0).mean()))
    assert max(per_step_bad) < 2e-3, per_step_bad
    print("[horizon]", check_horizon(st.numpy(), ref, same_flags))


def test_config5_nonfinite_envs_match_oracle(cuda_lib, oracle, exact_trig):
    """BASELINE config 5 (4096 envs x 100 steps, seed 4): the envs whose gradient the adjoint drops
    (COTIX_ADJ_NONFINITE) are exactly the envs the ORACLE's forward rollout drives non-finite — the
    reference's own 0/0 in resolve_collision (cotix/_collision_resolution.py:97-99,124-128: a
    zero-length penetration vector or drag velocity) — and they go non-finite at the same step."""
    import torch
    from parallax_b200 import lunar_lander as ll

    B, T = 4096, 100
    P = ll.default_params()
    rng = np.random.default_rng(4)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P).numpy(), rng, P)
    acts = rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)
    ro = ll.rollout(ll.LunarState.from_numpy(s0), torch.from_numpy(acts).cuda(), T, P, mode="fused")
    flags = torch.zeros(B, dtype=torch.int32, device="cuda")
    g = torch.zeros((18, B), device="cuda")
    g[1] = 2.0 * ro.tape[-1, 1]
    g0, ga = ll.rollout_vjp(ro, g, flags)
    f = flags.cpu().numpy()
    assert not (f & ll.ADJ_REPLAY_MISMATCH).any()
    tape = ro.tape.cpu().numpy()
    ref = s0
    first_ref = np.full(B, -1)
    for t in range(T):
        ref, _, _ = oracle.lunar_step(ref, acts[t], P)
        bad = ~np.isfinite(ref[:18]).all(axis=0)
        first_ref[(first_ref < 0) & bad] = t
        if t % 25 == 24:
            _eq(tape[t + 1], ref, f"tape slot {t + 1}")
    first_gpu = np.full(B, -1)
    for t in range(T):
        bad = ~np.isfinite(tape[t + 1, :18]).all(axis=0)
        first_gpu[(first_gpu < 0) & bad] = t
    _eq(first_gpu, first_ref, "step at which an env goes non-finite")
    dropped = (f & ll.ADJ_NONFINITE) != 0
    _eq(dropped, first_ref >= 0, "envs whose cotangent is dropped == envs the oracle drives non-finite")
    assert torch.isfinite(g0).all() and torch.isfinite(ga).all()
    print(f"[config5] {int(dropped.sum())} of {B} envs non-finite ({dropped.mean():.4f}); "
          f"first steps {np.bincount(first_ref[first_ref >= 0])[:8].tolist()} ...")


@pytest.mark.parametrize("rng_mode", [0, 1])  # threefry, philox
@pytest.mark.parametrize("B", [1, 37, 1000, 4096 + 3])
def test_rollout_kernel_matches_per_step_launches(cuda_lib, oracle, exact_trig, rng_mode, B):
    """csrc/lunar_rollout.cu (all steps of a small batch in ONE launch, envs resident in their block) against
    the same steps taken one fused launch at a time and against the oracle: state, last reward / done / obs,
    with auto-reset, actions cycling, ragged batch sizes (1 env; fewer envs per block for tiny batches)."""
    import torch
    import parallax_b200
    from parallax_b200 import lunar_lander as ll

    P = _env_params(ll, rng_mode=rng_mode)
    T, n_act = 40, 3
    rng = np.random.default_rng(900 + B)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B, offset=11), P), rng, P)
    acts = rng.uniform(-1, 1, size=(n_act, 2, B)).astype(np.float32)
    acts_d = torch.from_numpy(acts).cuda()
    st = ll.LunarState.from_numpy(s0)
    n0 = parallax_b200.launch_count()
    rew, done, obs = ll.steps(st, acts_d, T, P, want_obs=True, mode="fused")
    torch.cuda.synchronize()
    assert parallax_b200.launch_count() - n0 == 1
    st1 = ll.LunarState.from_numpy(s0)
    for t in range(T):
        _, r1, d1, o1 = ll.step(st1, acts_d[t % n_act], P, inplace=True, want_obs=True, mode="fused")
    _eq(st.numpy(), st1.numpy(), "state"); _eq(rew.cpu().numpy(), r1.cpu().numpy(), "reward")
    _eq(done.cpu().numpy(), d1.cpu().numpy(), "done"); _eq(obs.cpu().numpy(), o1.cpu().numpy(), "obs")
    if B <= 1000:
        ref = s0
        for t in range(T):
            ref, orew, odone, oobs = oracle.lunar_step(ref, acts[t % n_act], P, want_obs=True)
        _eq(st.numpy(), ref, "state vs oracle"); _eq(rew.cpu().numpy(), orew, "reward vs oracle")
        _eq(done.cpu().numpy(), odone, "done vs oracle"); _eq(obs.cpu().numpy(), oobs, "obs vs oracle")


def test_rollout_kernel_tape(cuda_lib, oracle, exact_trig):
    """The taped rollout of a small batch (cotix_lunar_rollout without a workspace = one launch of the rollout
    kernel): every slot of the tape equals the state after that many single steps, and the pipeline's tape."""
    import torch
    import parallax_b200
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    B, T = 777, 60
    rng = np.random.default_rng(77)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B, offset=5), P), rng, P)
    acts = torch.from_numpy(rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)).cuda()
    st0 = ll.LunarState.from_numpy(s0)
    n0 = parallax_b200.launch_count()
    ro = ll.rollout(st0, acts, T, P, mode="fused")
    torch.cuda.synchronize()
    assert parallax_b200.launch_count() - n0 == 1
    ro_p = ll.rollout(st0, acts, T, P, mode="pipeline")
    _eq(ro.tape.cpu().numpy(), ro_p.tape.cpu().numpy(), "tape: rollout kernel vs pipeline")
    st = ll.LunarState.from_numpy(s0)
    for t in range(T):
        ll.step(st, acts[t], P, inplace=True, mode="fused")
        if t in (0, 17, T - 1):
            _eq(ro.tape[t + 1].cpu().numpy(), st.numpy(), f"tape slot {t + 1}")


def test_rollout_kernel_queue_overflow(cuda_lib, oracle, exact_trig, monkeypatch):
    """The rollout kernel's per-block item queue overflowing (COTIX_ROLLOUT_QCAP = 3 slots for up to 32 envs):
    envs whose items do not fit evaluate their own pairs in phase C — same tape as with the full queue."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    B, T = 300, 30
    rng = np.random.default_rng(78)
    s0 = perturbed_lunar_state(oracle.lunar_reset(lunar_keys(B, offset=2), P), rng, P)
    acts = torch.from_numpy(rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)).cuda()
    st0 = ll.LunarState.from_numpy(s0)
    full = ll.rollout(st0, acts, T, P, mode="fused").tape.cpu().numpy()
    monkeypatch.setenv("COTIX_ROLLOUT_QCAP", "3")
    small = ll.rollout(st0, acts, T, P, mode="fused").tape.cpu().numpy()
    monkeypatch.setenv("COTIX_ROLLOUT_QCAP", "0")
    none = ll.rollout(st0, acts, T, P, mode="fused").tape.cpu().numpy()
    _eq(small, full, "tape with a 3-slot queue"); _eq(none, full, "tape with no queue at all")
    assert not np.array_equal(full[1, :18], full[-1, :18])


def test_rollout_kernel_long_horizon_config1(cuda_lib, oracle, exact_trig):
    """BASELINE config 1 shape through the rollout kernel: 1000 steps from reset(key), zero action, in ONE launch,
    against the same 1000 steps through the wavefront pipeline (one launch per stage and step) and, for a few envs,
    the oracle: free fall, touchdown, resting contact — state identical bit for bit at the end."""
    import torch
    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    B, T = 320, 1000
    keys = lunar_keys(B, offset=21)
    s0 = oracle.lunar_reset(keys, P)
    st = ll.LunarState.from_numpy(s0)
    ll.steps(st, None, T, P, mode="fused")  # no workspace: lunar_rollout_kernel
    sp = ll.LunarState.from_numpy(s0)
    ll.steps(sp, None, T, P, mode="pipeline")
    _eq(st.numpy(), sp.numpy(), "rollout kernel vs pipeline after 1000 steps")
    n = 6
    ref = s0[:, :n].copy()
    for _ in range(T):
        ref = oracle.lunar_step(np.ascontiguousarray(ref), None, P)[0]
    _eq(st.numpy()[:, :n], ref, "vs oracle")
    assert np.abs(st.numpy()[1] - s0[1]).max() > 1.0  # the landers did come down
