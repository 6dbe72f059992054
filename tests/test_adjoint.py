"""K5 — the hand-written adjoint of the LunarLander step (csrc/lunar_adjoint.cu, BASELINE config 5)
checked against finite differences of the CPU oracle's FORWARD step.

The oracle has no adjoint of its own: what pins the kernel is that <g_next, J d> estimated by
central differences of oracle.lunar_step along random directions d (state and action) equals
<step_vjp(g_next), d>.  The oracle is float32, so the differences are noisy (~1e-4 relative at
h = 1e-3); an env whose forward and backward one-sided differences disagree sits on a discrete
decision (a contact appearing, a coin-selected contact changing, a clip saturating) within +-h and
is excluded, as north_star excludes an epsilon band around the penetration threshold."""
import ctypes as C

import numpy as np
import pytest

from gen_inputs import lunar_keys, perturbed_lunar_state

# NaN envs (the perturbed poses put a few landers deep inside the terrain) are excluded by flag
pytestmark = [pytest.mark.gpu, pytest.mark.filterwarnings("ignore:invalid value encountered")]

H = 1e-3
KINK = 0.05  # |forward FD - backward FD| / scale above which a direction is "on a kink"


def _fd_terms(o, P, s0, act, f0, g_next, g_in, g_act, rng, n_steps=1, act_seq=None, H=H):
    """one random direction: per-env (analytic, central FD, kink measure, scale)"""
    B = s0.shape[1]
    d = rng.normal(size=(18, B)).astype(np.float32)
    da = rng.normal(size=act.shape).astype(np.float32)
    h = np.float32(H)
    sp, sm = s0.copy(), s0.copy()
    sp[:18] += h * d
    sm[:18] -= h * d
    ap, am = act + h * da, act - h * da

    def run(s, a):
        for t in range(n_steps):
            s = o.lunar_step(s, a if act_seq is None else a[t], P)[0]
        return s[:18].astype(np.float64)

    dd = (sp[:18].astype(np.float64) - sm[:18]) / (2 * H)      # the perturbation actually applied
    dda = (ap.astype(np.float64) - am) / (2 * H)
    fp, fm = run(sp, ap), run(sm, am)
    gn = g_next.astype(np.float64)
    fd_f = (gn * (fp - f0)).sum(0) / H
    fd_b = (gn * (f0 - fm)).sum(0) / H
    an = (g_in * dd).sum(0) + (g_act * dda).reshape(-1, B).sum(0)
    scale = (np.abs(gn) * (np.abs(fp - f0) + np.abs(f0 - fm))).sum(0) / (2 * H) + 1e-9
    return an, 0.5 * (fd_f + fd_b), np.abs(fd_f - fd_b) / scale, scale


def _contact_rich_state(o, ll, P, B, seed, pre_steps=5):
    rng = np.random.default_rng(seed)
    s0 = perturbed_lunar_state(o.lunar_reset(lunar_keys(B), P), rng, P)
    st = ll.LunarState.from_numpy(s0)
    for _ in range(pre_steps):  # so that resting and sliding contacts exist
        st, _, _ = ll.step(st, None, P, mode="fused")
    return st.numpy(), rng


def test_step_vjp_against_finite_differences_of_the_oracle(cuda_lib, oracle, exact_trig):
    import torch

    from parallax_b200 import lunar_lander as ll

    o = oracle
    P = ll.default_params()
    B = 4096
    s0, rng = _contact_rich_state(o, ll, P, B, seed=4)
    act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
    g_next = rng.normal(size=(18, B)).astype(np.float32)
    st, a_dev = ll.LunarState.from_numpy(s0), torch.from_numpy(act).cuda()
    nxt, _, _ = ll.step(st, a_dev, P, mode="fused")
    flags = torch.zeros(B, dtype=torch.int32, device="cuda")
    g_in, g_act = ll.step_vjp(st, a_dev, torch.from_numpy(g_next).cuda(), P, state_next=nxt, flags=flags)
    flags = flags.cpu().numpy()
    f0 = o.lunar_step(s0, act, P)[0][:18]
    assert np.array_equal(nxt.numpy()[:18], f0, equal_nan=True)            # forward parity, as everywhere
    assert int((flags & ll.ADJ_REPLAY_MISMATCH).sum()) == 0                  # the replay lands on the forward bits
    finite = (flags & ll.ADJ_NONFINITE) == 0
    ncon = (flags >> ll.ADJ_CONTACT_SHIFT) & 7
    assert (ncon[finite] >= 1).sum() >= 400, "not enough envs with a resolved contact to test the contact adjoint"
    g_in64, g_act64 = g_in.cpu().numpy().astype(np.float64), g_act.cpu().numpy().astype(np.float64)
    assert np.all(g_in64[:, ~finite] == 0)                                   # dropped cotangent of NaN envs
    f0 = f0.astype(np.float64)
    rel, kink = [], []
    for _ in range(3):
        an, fd, kk, sc = _fd_terms(o, P, s0, act, f0, g_next, g_in64, g_act64, rng)
        rel.append(np.abs(an - fd) / sc)
        kink.append(kk)
    rel, kink = np.stack(rel), np.stack(kink)
    for name, sel in (("free", finite & (ncon == 0)), ("contact", finite & (ncon >= 1))):
        r, k = rel[:, sel].ravel(), kink[:, sel].ravel()
        smooth = k < KINK
        assert smooth.mean() > 0.9, (name, smooth.mean())
        r = r[smooth]
        assert np.quantile(r, 0.5) < 2e-4, (name, np.quantile(r, 0.5))
        assert np.quantile(r, 0.99) < 1e-2, (name, np.quantile(r, 0.99))
        assert (r < 5e-2).mean() > 0.997, (name, (r < 5e-2).mean())


def test_step_vjp_is_linear_and_aliasing_safe(cuda_lib, oracle, exact_trig):
    """a VJP is linear in the cotangent; g_in may alias g_out; no action == zero action."""
    import torch

    from parallax_b200 import lunar_lander as ll

    P = ll.default_params()
    B = 1024
    s0, rng = _contact_rich_state(oracle, ll, P, B, seed=9)
    st = ll.LunarState.from_numpy(s0)
    g1 = torch.from_numpy(rng.normal(size=(18, B)).astype(np.float32)).cuda()
    g2 = torch.from_numpy(rng.normal(size=(18, B)).astype(np.float32)).cuda()
    a1, _ = ll.step_vjp(st, None, g1, P)
    a2, _ = ll.step_vjp(st, None, g2, P)
    a12, _ = ll.step_vjp(st, None, g1 + 2.0 * g2, P)
    ref = (a1 + 2.0 * a2)
    # float32 rounding inside the kernel is the only source of difference: compare per env against
    # the size of the largest term
    err = (a12 - ref).abs() / ((a1.abs() + 2.0 * a2.abs()).amax(0, keepdim=True) + 1e-6)
    assert float(err.max()) < 1e-4, float(err.max())
    zero = torch.zeros((2, B), device="cuda")
    z1, _ = ll.step_vjp(st, zero, g1, P)
    assert torch.equal(z1, a1)
    # in place: g_in == g_out buffer
    g = g1.clone()
    from parallax_b200 import _lib
    _lib.check(_lib.lib().cotix_lunar_step_bwd(torch.cuda.current_stream().cuda_stream, B, st.data.data_ptr(), None,
                                              B, None, g.data_ptr(), g.data_ptr(), None, None, C.byref(P)))
    assert torch.equal(g, a1)


def test_rollout_gradient_in_flight_config5(cuda_lib, oracle, exact_trig):
    """BASELINE config 5 shape: T = 100 steps, loss = final lander y^2; from reset(key) the lander is
    in free flight under thrust for the whole horizon, so the loss is smooth and the gradient with
    respect to the initial state and to all 100 actions must match finite differences of the oracle."""
    import torch

    from parallax_b200 import lunar_lander as ll

    o = oracle
    P = ll.default_params()
    B, T = 128, 100
    rng = np.random.default_rng(5)
    s0 = o.lunar_reset(lunar_keys(B), P)
    acts = rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)
    ro = ll.rollout(ll.LunarState.from_numpy(s0), torch.from_numpy(acts).cuda(), T, P)
    s = s0
    for t in range(T):
        s = o.lunar_step(s, acts[t], P)[0]
    assert np.array_equal(ro.final.numpy(), s, equal_nan=True)               # the tape ends on the oracle's bits
    y = ro.tape[-1, 1]
    g = torch.zeros((18, B), device="cuda")
    g[1] = 2.0 * y
    flags = torch.zeros(B, dtype=torch.int32, device="cuda")
    g0, ga = ll.rollout_vjp(ro, g, flags)
    assert int(flags.cpu().numpy().astype(np.int64).__and__(3).sum()) == 0
    g0, ga = g0.cpu().numpy().astype(np.float64), ga.cpu().numpy().astype(np.float64)
    # the loss as a function: L = y_T^2 -> use g_next = 2 y_T on the linearised output
    f0 = s[:18].astype(np.float64)
    gn = g.cpu().numpy()
    rel, ans, fds = [], [], []
    for _ in range(3):
        # a larger step than for one step: float32 rounding accumulates over the 100 steps
        an, fd, kk, sc = _fd_terms(o, P, s0, acts, f0, gn, g0, ga, rng, n_steps=T, act_seq=True, H=1e-2)
        rel.append(np.abs(an - fd) / sc)
        ans.append(an)
        fds.append(fd)
    rel, ans, fds = np.stack(rel), np.concatenate(ans), np.concatenate(fds)
    cos = float(ans @ fds / np.sqrt((ans @ ans) * (fds @ fds)))
    ratio = float(np.sqrt((ans @ ans) / (fds @ fds)))
    print(f"flight T={T}: rel err p50 {np.quantile(rel, 0.5):.2e} p90 {np.quantile(rel, 0.9):.2e} max {rel.max():.2e}; "
          f"cosine {cos:.7f} norm ratio {ratio:.5f}")
    # per direction the float32 differences are noisy where the directional derivative happens to be
    # small (the |dv| dv spring term also makes the central difference first-order accurate only);
    # over all envs and directions the two must be the same vector
    assert np.quantile(rel, 0.5) < 3e-3 and np.quantile(rel, 0.9) < 3e-2
    assert cos > 0.9999 and abs(ratio - 1.0) < 2e-3
    # translation invariance of free flight: shifting all three bodies in y shifts y_T by the same amount
    assert np.allclose(g0[1] + g0[7] + g0[13], 2.0 * f0[1], rtol=1e-3, atol=1e-4)
    # and main-engine thrust early in the rollout moves y_T more than thrust late (more time to act)
    on = (acts[:, 0] > 0.05) & (acts[:, 0] < 0.95)
    early = np.abs(ga[:10, 0])[on[:10]].mean()
    late = np.abs(ga[-10:, 0])[on[-10:]].mean()
    assert early > 2.0 * late


def test_rollout_vjp_contact_rich(cuda_lib, oracle, exact_trig):
    """contact-rich envs: fused and pipeline tapes are identical, the backward replay reproduces
    every step of the tape bit for bit, and over a 10-step horizon the gradient agrees with finite
    differences of the oracle wherever the perturbed rollouts stay on the same decisions."""
    import torch

    from parallax_b200 import lunar_lander as ll

    o = oracle
    P = ll.default_params()
    B, T = 2048, 10
    s0, rng = _contact_rich_state(o, ll, P, B, seed=11)
    acts = rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)
    a_dev = torch.from_numpy(acts).cuda()
    ro = ll.rollout(ll.LunarState.from_numpy(s0), a_dev, T, P, mode="fused")
    ro_p = ll.rollout(ll.LunarState.from_numpy(s0), a_dev, T, P, mode="pipeline")
    assert torch.equal(ro.tape.view(torch.int32), ro_p.tape.view(torch.int32))
    gn = rng.normal(size=(18, B)).astype(np.float32)
    flags = torch.zeros(B, dtype=torch.int32, device="cuda")
    g0, ga = ll.rollout_vjp(ro, torch.from_numpy(gn).cuda(), flags)
    flags = flags.cpu().numpy()
    assert int((flags & ll.ADJ_REPLAY_MISMATCH).sum()) == 0
    # default above: decision tape + transposed Jacobians of the env-steps with contacts.  The
    # decision tape alone (any group size) gives the bits of the un-taped sweep ...
    f1 = torch.zeros(B, dtype=torch.int32, device="cuda")
    r0, ra = ll.rollout_vjp(ro, torch.from_numpy(gn).cuda(), f1, decision_tape=False)
    assert np.array_equal(f1.cpu().numpy(), flags)
    for dt in (True, 3, 1):
        f2 = torch.zeros(B, dtype=torch.int32, device="cuda")
        h0, ha = ll.rollout_vjp(ro, torch.from_numpy(gn).cuda(), f2, decision_tape=dt, jacobians=False)
        assert torch.equal(h0.view(torch.int32), r0.view(torch.int32)), dt
        assert torch.equal(ha.view(torch.int32), ra.view(torch.int32)), dt
        assert np.array_equal(f2.cpu().numpy(), flags), dt
    # ... and the matrix form (any group size) the same numbers up to float32 rounding of sums taken
    # in another order: compared per env against the largest term of the gradient
    for dt in (True, 4):
        f2 = torch.zeros(B, dtype=torch.int32, device="cuda")
        h0, ha = ll.rollout_vjp(ro, torch.from_numpy(gn).cuda(), f2, decision_tape=dt)
        assert np.array_equal(f2.cpu().numpy(), flags), dt
        err = (h0 - r0).abs() / (r0.abs().amax(0, keepdim=True) + 1e-6)
        assert float(err.max()) < 2e-3 and float(err.flatten().quantile(0.99)) < 1e-5, (dt, float(err.max()))
        erra = (ha - ra).abs() / (ra.abs().amax((0, 1), keepdim=True) + 1e-9)
        assert float(erra.max()) < 2e-3, (dt, float(erra.max()))
    finite = (flags & ll.ADJ_NONFINITE) == 0
    touched = ((flags >> ll.ADJ_CONTACT_SHIFT) & 7) > 0
    assert (finite & touched).sum() >= 200
    assert torch.isfinite(g0).all() and torch.isfinite(ga).all()
    s = s0
    for t in range(T):
        s = o.lunar_step(s, acts[t], P)[0]
    f0 = s[:18].astype(np.float64)
    g0, ga = g0.cpu().numpy().astype(np.float64), ga.cpu().numpy().astype(np.float64)
    rel, kink = [], []
    for _ in range(2):
        an, fd, kk, sc = _fd_terms(o, P, s0, acts, f0, gn, g0, ga, rng, n_steps=T, act_seq=True)
        rel.append(np.abs(an - fd) / sc)
        kink.append(kk)
    rel, kink = np.stack(rel), np.stack(kink)
    sel = finite & touched
    r, k = rel[:, sel].ravel(), kink[:, sel].ravel()
    smooth = k < KINK
    print(f"contact-rich T={T}: envs {sel.sum()}, smooth fraction {smooth.mean():.3f}, rel err p50 "
          f"{np.quantile(r[smooth], 0.5):.2e} p99 {np.quantile(r[smooth], 0.99):.2e}")
    assert smooth.mean() > 0.5
    assert np.quantile(r[smooth], 0.5) < 1e-3
    assert (r[smooth] < 5e-2).mean() > 0.98


def test_adjoint_argument_checks(cuda_lib):
    import torch

    from parallax_b200 import _lib
    from parallax_b200 import lunar_lander as ll

    st = ll.reset(lunar_keys(8))
    g = torch.zeros((18, 8), device="cuda")
    for bad in (dict(substeps=2), dict(auto_reset=1)):
        with pytest.raises(_lib.CotixError):
            ll.step_vjp(st, None, g, ll.default_params(**bad))
    with pytest.raises(_lib.CotixError):
        _lib.check(_lib.lib().cotix_lunar_step_bwd(None, 8, st.data.data_ptr(), None, 8, None, None, None, None, None,
                                                  C.byref(ll.default_params())))
    # rollout_bwd: a workspace smaller than one step of decisions, or misaligned, is refused
    L = _lib.lib()
    tape = torch.zeros((3, 36, 8), device="cuda")
    tape[0] = st.data
    need = int(L.cotix_lunar_adjoint_decisions_bytes(8, 1))
    assert need == 33 * 8 * 4 and int(L.cotix_lunar_adjoint_decisions_bytes(8, 100)) == 100 * need
    full = int(L.cotix_lunar_adjoint_workspace_bytes(8, 1))
    assert full > need
    ws = torch.zeros(full + 512, dtype=torch.uint8, device="cuda")
    wp = (ws.data_ptr() + 255) & ~255
    P = ll.default_params()
    args = (None, 8, 2, tape.data_ptr(), 8, None, g.data_ptr(), None, None, C.byref(P))
    with pytest.raises(_lib.CotixError):
        _lib.check(L.cotix_lunar_rollout_bwd(*args, wp, need - 4, 0))
    with pytest.raises(_lib.CotixError):
        _lib.check(L.cotix_lunar_rollout_bwd(*args, wp + 4, need, 0))
    with pytest.raises(_lib.CotixError):
        _lib.check(L.cotix_lunar_rollout_bwd(*args, wp, need, 1))      # too small for the Jacobian area
    _lib.check(L.cotix_lunar_rollout(None, 8, 2, tape.data_ptr(), 8, None, C.byref(P), None, 0))
    _lib.check(L.cotix_lunar_rollout_bwd(*args, wp, need, 0))
    torch.cuda.synchronize()

