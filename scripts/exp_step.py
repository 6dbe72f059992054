"""Step time of the bench workload (1 M envs, one graph of `steps` steps after `warm`) and the serial
per-stage times, for whichever library COTIX_LIB_PATH points at (kernel-tuning A/B).
usage: COTIX_LIB_PATH=variants/lib_x.so python scripts/exp_step.py [envs] [steps] [warm]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gen_inputs import lunar_keys, perturbed_lunar_state  # noqa: E402
from parallax_b200 import _lib  # noqa: E402
from parallax_b200 import lunar_lander as ll  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
warm = int(sys.argv[3]) if len(sys.argv) > 3 else 40
dev = torch.device("cuda", 0)
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), np.random.default_rng(1000), P)
acts = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, size=(4, 2, B)).astype(np.float32)).to(dev)
st = ll.LunarState.from_numpy(s0, device=dev)
gw, gt = ll.StepsGraph(st, acts, warm, P), ll.StepsGraph(st, acts, steps, P)
gw.launch()
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    gt.launch()
    ev1.record()
    torch.cuda.synchronize()
    best = min(best, ev0.elapsed_time(ev1) / steps)
ll.pipeline_profile(True)
for k in range(20):
    ll.step(st, acts[k % 4], P, inplace=True)
torch.cuda.synchronize()
stage, _ = ll.pipeline_stage_ms()
ll.pipeline_profile(False)
chk = float(np.nansum(np.abs(st.numpy()[:18]).astype(np.float64)))
print(json.dumps({"lib": os.path.basename(_lib.LIB_PATH), "us_per_step": 1e3 * best,
                  "stage_us": {k: round(1e3 * v, 1) for k, v in stage.items()}, "checksum": chk}), flush=True)
