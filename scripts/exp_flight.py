"""Experiment: cost of the regular (contact-free) part of the fused step kernel.
All envs start from the reference reset (lander at y=5, free fall): no candidate pairs for
the first ~700 steps, no judge, no resets -> pure phase A + C."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from parallax_b200 import lunar_lander as ll
from gen_inputs import lunar_keys

B = 1 << 20
for name, kw in [("flight threefry", dict()), ("flight philox", dict(rng_mode=1)),
                 ("flight judge philox", dict(rng_mode=1, judge_enabled=1, auto_reset=1, max_episode_steps=1000))]:
    P = ll.default_params(**kw)
    st = ll.reset(lunar_keys(B), P)
    act = torch.zeros((2, B), device="cuda")
    for _ in range(5):
        ll.step(st, act, P, inplace=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 50
    for _ in range(n):
        ll.step(st, act, P, inplace=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{name}: {ms*1e3:.1f} us/step  {B/ms/1e3:.3e} env-steps/s")
