"""One steady-state step of the bench workload for Nsight Compute (eager launches, ONE env range):
40 warm-up steps, then 2 more.  With split = 1 a step is 6 kernels, so
  ncu --set full --clock-control none --import-source on -k regex:pipe_ --launch-skip 240 --launch-count 6 \
      -o gpurun_out/r2_step python scripts/ncu_step.py
captures exactly the 6 kernels of step 41.  usage: python scripts/ncu_step.py [envs]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gen_inputs import lunar_keys, perturbed_lunar_state  # noqa: E402
from parallax_b200 import lunar_lander as ll  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
dev = torch.device("cuda", 0)
ll.pipeline_config(split=1, pdl=0)
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), np.random.default_rng(1000), P)
acts = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, size=(4, 2, B)).astype(np.float32)).to(dev)
st = ll.LunarState.from_numpy(s0, device=dev)
for t in range(42):
    ll.step(st, acts[t % 4], P, inplace=True)
torch.cuda.synchronize()
print(ll.pipeline_counters(dev, B))
