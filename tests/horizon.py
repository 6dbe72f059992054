"""The stated tolerance over a 100-step horizon (north_star), shared by the GPU
test (CUDA engine vs the oracle with libm trig) and the CPU test (oracle with the engine's cx_trig.h
vs the oracle with libm trig — the CUDA engine is bit-identical to the former).

Two runs that differ only in the definition of sin / cos / atan2 (<= 3 ulp apart) start from
identical contact-rich states and take 100 reference steps.  Stated tolerance:
  * the envs that go non-finite (the reference's own 0/0 in resolve_collision,
    cotix/_collision_resolution.py:97-99,124-128) are THE SAME envs in both runs;
  * at most 0.2 % of the finite envs see a reported flag (leg contact, done code) differ at any of
    the 100 steps;
  * among the rest, relative to max(|x|, 1) over all 18 dynamic leaves: median <= 5e-6, 90th
    percentile <= 5e-5, 99th percentile <= 5e-4;
  * at most 0.5 % of them are beyond 1e-2: a GJK / EPA / coin decision that flipped within epsilon of
    its threshold changes a whole contact without changing a reported flag, and the env diverges
    as a whole (chaotic contact dynamics, not rounding).
Measured (seed 160 / 161, 4096 envs): non-finite sets identical (5.5 %), flag flips 0.03 / 0.05 %,
median 1.3e-6, p90 9e-6, p99 8e-5, 0.10 % beyond 1e-2."""
import numpy as np

HORIZON_STEPS = 100
FLAG_FLIP_MAX = 2e-3
MEDIAN_MAX, P90_MAX, P99_MAX = 5e-6, 5e-5, 5e-4
DIVERGED_AT, DIVERGED_MAX = 1e-2, 5e-3


def horizon_stats(got, ref, same_flags):
    """got / ref: [36, B] states after HORIZON_STEPS steps; same_flags: [B] bool, the reported flags
    agreed at every step.  The measured figures, no assertion."""
    fa, fb = np.isfinite(got[:18]).all(axis=0), np.isfinite(ref[:18]).all(axis=0)
    fin = fa & fb
    with np.errstate(invalid="ignore"):
        d = np.abs(got[:18] - ref[:18]) / np.maximum(np.abs(ref[:18]), 1.0)
    d = np.where(fin[None], d, 0.0).max(axis=0)
    ok = same_flags & fin
    med, p90, p99 = (float(np.percentile(d[ok], q)) for q in (50, 90, 99))
    return {"nonfinite": float(1.0 - fin.mean()), "nonfinite_in_one_run_only": float((fa != fb).mean()),
            "flag_flips": float((~same_flags & fin).sum() / max(1, fin.sum())), "median": med, "p90": p90, "p99": p99,
            "diverged": float((d[ok] > DIVERGED_AT).mean()), "max": float(d[ok].max())}


def check_horizon(got, ref, same_flags, flag_flip_max=FLAG_FLIP_MAX, median_max=MEDIAN_MAX, p90_max=P90_MAX,
                  p99_max=P99_MAX, diverged_max=DIVERGED_MAX, nonfinite_mismatch_max=0.0):
    """Asserts the stated tolerance (defaults: the exact flavour's, module docstring) and returns the
    measured figures."""
    st = horizon_stats(got, ref, same_flags)
    assert st["nonfinite_in_one_run_only"] <= nonfinite_mismatch_max, st
    assert st["flag_flips"] <= flag_flip_max, f"flag flips in {st['flag_flips']:.4f} of the finite envs: {st}"
    assert st["median"] <= median_max and st["p90"] <= p90_max and st["p99"] <= p99_max, st
    assert st["diverged"] <= diverged_max, f"{st['diverged']:.4f} of the envs beyond {DIVERGED_AT}: {st}"
    return st
