"""Experiment: device timeline of one chunked host step (COTIX_TRACE), un-graphed path."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from parallax_b200 import lunar_lander as ll
from gen_inputs import lunar_keys
B = 1 << 20
chunks = int(sys.argv[1])
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
rng = np.random.default_rng(1)
acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).pin_memory() for _ in range(4)]
env = ll.LunarLanderEnv(B, P); env.reset(lunar_keys(B)); env.set_host_graphs(False); env.set_chunks(chunks)
for w in range(40):
    env.step(acts[w % 4])
env.close()
