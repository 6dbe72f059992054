"""Exploration: the adjoint kernel against central finite differences of the CPU oracle's forward
step, per number of resolve_collision calls in the step; and timing of the differentiable rollout
(BASELINE config 5).  Usage: python scripts/exp_adjoint.py [B] [h]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gen_inputs import lunar_keys, perturbed_lunar_state  # noqa: E402

from oracle import oracle as o  # noqa: E402
from parallax_b200 import lunar_lander as ll  # noqa: E402


def fd_check(s0, act, g_next, P, h, n_dirs, rng):
    """per-env (analytic, central FD, forward FD, backward FD, scale) for n_dirs random directions"""
    B = s0.shape[1]
    st = ll.LunarState.from_numpy(s0)
    a_dev = torch.from_numpy(act).cuda()
    nxt, _, _ = ll.step(st, a_dev, P, mode="fused")
    flags = torch.zeros(B, dtype=torch.int32, device="cuda")
    g_in, g_act = ll.step_vjp(st, a_dev, torch.from_numpy(g_next).cuda(), P, state_next=nxt, flags=flags)
    g_in, g_act, flags = g_in.cpu().numpy().astype(np.float64), g_act.cpu().numpy().astype(np.float64), flags.cpu().numpy()
    f0 = o.lunar_step(s0, act, P)[0][:18].astype(np.float64)
    assert np.array_equal(nxt.numpy()[:18], f0.astype(np.float32), equal_nan=True)
    out = []
    for _ in range(n_dirs):
        d = rng.normal(size=(18, B)).astype(np.float32)
        da = rng.normal(size=(2, B)).astype(np.float32)
        sp, sm = s0.copy(), s0.copy()
        sp[:18] += np.float32(h) * d
        sm[:18] -= np.float32(h) * d
        ap, am = act + np.float32(h) * da, act - np.float32(h) * da
        # the perturbation actually applied (after rounding)
        dp = (sp[:18].astype(np.float64) - s0[:18]) / h
        dm = (s0[:18].astype(np.float64) - sm[:18]) / h
        dap = (ap.astype(np.float64) - act) / h
        dam = (act.astype(np.float64) - am) / h
        fp = o.lunar_step(sp, ap, P)[0][:18].astype(np.float64)
        fm = o.lunar_step(sm, am, P)[0][:18].astype(np.float64)
        gn = g_next.astype(np.float64)
        fd_f = (gn * (fp - f0)).sum(0) / h
        fd_b = (gn * (f0 - fm)).sum(0) / h
        fd_c = 0.5 * (fd_f + fd_b)
        an = 0.5 * ((g_in * (dp + dm)).sum(0) + (g_act * (dap + dam)).sum(0))
        scale = (np.abs(gn) * (np.abs(fp - f0) + np.abs(f0 - fm))).sum(0) / (2 * h) + 1e-12
        out.append((an, fd_c, fd_f, fd_b, scale))
    return out, flags


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    h = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-3
    o.set_trig_mode(1)
    P = ll.default_params()
    rng = np.random.default_rng(4)
    s0 = perturbed_lunar_state(o.lunar_reset(lunar_keys(B), P), rng, P)
    # a few steps so that resting / sliding contacts exist
    st = ll.LunarState.from_numpy(s0)
    for _ in range(int(os.environ.get("PRE", "5"))):
        st, _, _ = ll.step(st, None, P, mode="fused")
    s0 = st.numpy()
    act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
    g_next = rng.normal(size=(18, B)).astype(np.float32)
    res, flags = fd_check(s0, act, g_next, P, h, 4, rng)
    ncon = (flags >> 8) & 7
    print("flags: mismatch", int((flags & 1).sum()), "nonfinite", int(((flags >> 1) & 1).sum()),
          "resolves/env histogram", np.bincount(ncon, minlength=5).tolist())
    for k in range(5):
        sel = ncon == k
        if not sel.any():
            continue
        rel, kink = [], []
        for an, fc, ff, fb, sc in res:
            rel.append(np.abs(an - fc)[sel] / sc[sel])
            kink.append(np.abs(ff - fb)[sel] / sc[sel])
        rel, kink = np.concatenate(rel), np.concatenate(kink)
        smooth = kink < 0.05
        q = np.nanquantile(rel[smooth], [0.5, 0.9, 0.99, 1.0]) if smooth.any() else [np.nan] * 4
        print(f"resolves={k}: n={sel.sum()} smooth={smooth.mean():.3f} rel err (smooth) p50={q[0]:.2e} p90={q[1]:.2e} "
              f"p99={q[2]:.2e} max={q[3]:.2e}; frac(rel<2e-2 | smooth)={(rel[smooth] < 2e-2).mean():.4f}")

    if os.environ.get("FLIGHT"):
        s0 = o.lunar_reset(lunar_keys(B), P)
    # ---- config 5: B envs x 100 steps, loss = final lander y^2 --------------------------------
    T = 100
    actions = torch.from_numpy(rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)).cuda()
    st0 = ll.LunarState.from_numpy(s0)
    flags = torch.zeros(B, dtype=torch.int32, device="cuda")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for it in range(3):
        flags.zero_()
        ev[0].record()
        if os.environ.get("ADJ_PIPE"):
            from parallax_b200 import _lib
            _lib.lib().cx_lunar_adjoint_set_pipelined(int(os.environ["ADJ_PIPE"]))
        ro = ll.rollout(st0, actions, T, P, mode=os.environ.get("MODE", "fused"))
        ev[1].record()
        y = ro.tape[-1, 1]
        g = torch.zeros((18, B), device="cuda")
        g[1] = 2.0 * y
        g0, ga = ll.rollout_vjp(ro, g, flags, decision_tape=not os.environ.get("NOTAPE"))
        ev[2].record()
        torch.cuda.synchronize()
    f = flags.cpu().numpy()
    print(f"rollout B={B} T={T}: forward {ev[0].elapsed_time(ev[1]):.3f} ms, backward {ev[1].elapsed_time(ev[2]):.3f} ms; "
          f"mismatch {int((f & 1).sum())} nonfinite {int(((f >> 1) & 1).sum())} "
          f"|g0| finite {bool(torch.isfinite(g0).all())} mean|g0| {float(g0.abs().mean()):.3e} mean|ga| {float(ga.abs().mean()):.3e}")


if __name__ == "__main__":
    main()
