"""The fast-math flavour (libcotix_b200_fast.so: the same sources built -fmad=true -prec-div=false
-prec-sqrt=false) against the oracle, in a child process with COTIX_FLAVOUR=fast.  It is NOT
bit-exact; the bar is north_star's: positions and velocities within 1e-5 relative per step from
identical states, flags and termination identical except for envs within epsilon of a decision
threshold (fraction stated and asserted), and the 100-step-horizon tolerance of tests/horizon.py.
The exact flavour stays the parity build and the benched headline."""
import json
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu

PER_STEP_RTOL = 1e-5
PER_STEP_OUTLIERS = 1.5e-2    # envs beyond 1e-5 after one step: measured 0.75 %, all of them envs resolving a contact
                              # (p99 over all envs 7.5e-6, median 0)
FLAG_MISMATCH = 2e-3          # envs whose done / leg-contact flags differ after one step: measured 0
# 100-step horizon, stated for this flavour (the exact flavour's is tests/horizon.py): measured flag flips
# 0.59 %, median 4.9e-6, p90 1.6e-5, 1.9 % of the envs diverged (> 1e-2), non-finite envs identical
HORIZON = dict(flag_flips=1.5e-2, median=2e-5, p90=1e-4, diverged=4e-2, nonfinite_in_one_run_only=2e-3)


def test_fast_flavour_within_tolerance():
    code = textwrap.dedent(f"""
        import sys, json
        sys.path.insert(0, {ROOT!r}); sys.path.insert(0, {os.path.join(ROOT, 'tests')!r})
        import numpy as np, torch
        import parallax_b200
        from oracle import oracle as o
        from gen_inputs import lunar_keys, perturbed_lunar_state
        from horizon import horizon_stats, HORIZON_STEPS
        from parallax_b200 import lunar_lander as ll
        assert parallax_b200._lib.lib().cotix_math_flavour() == 1
        o.set_trig_mode(1)
        P = ll.default_params(judge_enabled=1)
        B = 8192
        rng = np.random.default_rng(321)
        s0 = perturbed_lunar_state(o.lunar_reset(lunar_keys(B), P), rng, P)
        out = {{}}
        for mode in ("fused", "pipeline"):
            act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
            st, rew, done, obs = ll.step(ll.LunarState.from_numpy(s0), torch.from_numpy(act).cuda(), P, want_obs=True, mode=mode)
            ref, orew, odone, oobs = o.lunar_step(s0, act, P, want_obs=True)
            got = st.numpy()
            with np.errstate(invalid="ignore"):
                d = np.abs(got[:18] - ref[:18]) / np.maximum(np.abs(ref[:18]), 1.0)
            both_nan = np.isnan(got[:18]) & np.isnan(ref[:18])
            d = np.where(both_nan, 0.0, np.nan_to_num(d, nan=np.inf)).max(axis=0)
            flags = (obs.cpu().numpy()[6:8] != oobs[6:8]).any(axis=0) | (done.cpu().numpy() != odone)
            out[mode] = dict(outliers=float((d > {PER_STEP_RTOL}).mean()), median=float(np.median(d)),
                             p99=float(np.percentile(d, 99)), flag_mismatch=float(flags.mean()),
                             keys_equal=bool((got[30:32].view(np.uint32) == ref[30:32].view(np.uint32)).all()))
        # 100-step horizon (the same stated tolerance as the exact flavour against libm trig)
        st, ref = ll.LunarState.from_numpy(s0[:, :4096].copy()), s0[:, :4096].copy()
        same = np.ones(4096, bool)
        for t in range(HORIZON_STEPS):
            st, rew, done, obs = ll.step(st, None, P, want_obs=True)
            ref, orew, odone, oobs = o.lunar_step(ref, None, P, want_obs=True)
            same &= (obs.cpu().numpy()[6:8] == oobs[6:8]).all(axis=0) & (done.cpu().numpy() == odone)
        out["horizon"] = horizon_stats(st.numpy(), ref, same)
        print("RESULT " + json.dumps(out))
    """)
    env = dict(os.environ, COTIX_FLAVOUR="fast")
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=1200)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("RESULT ")][-1]
    out = json.loads(line[7:])
    print("[fast flavour]", out)
    for mode in ("fused", "pipeline"):
        r = out[mode]
        assert r["keys_equal"], "the PRNG key chain is integer work: bit-exact in every flavour"
        assert r["outliers"] <= PER_STEP_OUTLIERS, r
        assert r["flag_mismatch"] <= FLAG_MISMATCH, r
        assert r["median"] <= 1e-6, r
    for k, lim in HORIZON.items():
        assert out["horizon"][k] <= lim, (k, out["horizon"])
