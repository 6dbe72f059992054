"""Experiment: fused kernel vs wavefront pipeline on the bench workload (device-resident)."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from parallax_b200 import lunar_lander as ll
from gen_inputs import lunar_keys, perturbed_lunar_state

B = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 1 << 20
for rng_mode in (0, 1):
    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000, rng_mode=rng_mode)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P).numpy(), np.random.default_rng(1000), P)
    rng = np.random.default_rng(1)
    acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).cuda() for _ in range(4)]
    for mode in (("fused", "pipeline") if "--fused" in sys.argv else ("pipeline",)):
        st = ll.LunarState.from_numpy(s0)
        for w in range(20):
            ll.step(st, acts[w % 4], P, inplace=True, mode=mode)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 100
        e0.record()
        for k in range(n):
            ll.step(st, acts[k % 4], P, inplace=True, mode=mode)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"rng={rng_mode} {mode}: {ms*1e3:.1f} us/step  {B/ms*1e3:.3e} env-steps/s", flush=True)
        if mode == "pipeline":
            ll.pipeline_profile(True)
            for k in range(50):
                ll.step(st, acts[k % 4], P, inplace=True, mode=mode)
            torch.cuda.synchronize()
            sm, n = ll.pipeline_stage_ms()
            ll.pipeline_profile(False)
            print("   stages us:", {k: round(v * 1e3, 1) for k, v in sm.items()}, "sum", round(sum(sm.values()) * 1e3, 1), flush=True)
