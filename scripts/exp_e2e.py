"""Experiment: host-facing env step (H2D action -> kernels -> D2H obs/reward/done) vs the number of
env ranges the step is software-pipelined over, with the raw copy times beside it."""
import os
import sys
import time

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
import torch
from parallax_b200 import lunar_lander as ll
from gen_inputs import lunar_keys

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
rng = np.random.default_rng(1)
acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).pin_memory() for _ in range(4)]

# raw copies
d_act = torch.empty((2, B), device="cuda"); d_obs = torch.empty((8, B), device="cuda")
h_obs = torch.empty((8, B)).pin_memory()
for name, fn, nbytes in (("H2D action", lambda: d_act.copy_(acts[0], non_blocking=True), 8 * B),
                         ("D2H obs", lambda: h_obs.copy_(d_obs, non_blocking=True), 32 * B)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 20
    print(f"{name}: {dt * 1e3:.3f} ms  {nbytes / dt / 1e9:.1f} GB/s", flush=True)

env = ll.LunarLanderEnv(B, P)
env.reset(lunar_keys(B))
for w in range(30):
    env.step(acts[w % 4])
for hg, chunks in [(0, 1), (0, 4), (1, 2), (1, 4), (1, 6), (1, 8), (1, 12), (1, 16), (1, 0)]:
    env.set_host_graphs(bool(hg))
    env.set_chunks(chunks)
    for w in range(5):
        env.step(acts[w % 4])
    n = 50
    t0 = time.perf_counter()
    for k in range(n):
        env.step(acts[k % 4])
    dt = (time.perf_counter() - t0) / n
    print(f"graphs={hg} chunks={chunks}: {dt * 1e3:.3f} ms/step  {B / dt:.3e} env-steps/s", flush=True)
env.close()
