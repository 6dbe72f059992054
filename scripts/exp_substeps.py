"""Experiment: K sub-steps per call, fused kernel vs wavefront pipeline (bench workload)."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from parallax_b200 import lunar_lander as ll
from gen_inputs import lunar_keys, perturbed_lunar_state

B = 1 << 20
for K in (1, 4, 10):
    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000, substeps=K)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P).numpy(), np.random.default_rng(1000), P)
    rng = np.random.default_rng(1)
    acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).cuda() for _ in range(4)]
    for mode in ("pipeline", "fused"):
        st = ll.LunarState.from_numpy(s0)
        n = max(4, 40 // K)
        for w in range(max(2, 20 // K)):
            ll.step(st, acts[w % 4], P, inplace=True, mode=mode)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(n):
            ll.step(st, acts[k % 4], P, inplace=True, mode=mode)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"K={K} {mode}: {ms * 1e3:.1f} us/call  {ms * 1e3 / K:.1f} us/sub-step  {B * K / ms * 1e3:.3e} env-steps/s", flush=True)
