"""Sweep of the pipeline's launch configuration on the bench workload (1 M envs, BASELINE config 3):
env ranges (split) x programmatic dependent launch x launch priorities, each timed as ONE replayed
CUDA graph of `steps` steps after `warm` warm-up steps from the same initial state.
usage: python scripts/exp_split.py [envs] [steps] [warm]  -> JSON lines on stdout"""
import json
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
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
warm = int(sys.argv[3]) if len(sys.argv) > 3 else 40
dev = torch.device("cuda", 0)
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), np.random.default_rng(1000), P)
acts = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, size=(4, 2, B)).astype(np.float32)).to(dev)
configs = [(1, 1, 0), (1, 0, 0), (2, 1, 1), (2, 1, 0), (3, 1, 1), (4, 1, 1), (4, 1, 0), (4, 0, 1), (6, 1, 1),
           (8, 1, 1), (8, 1, 0), (12, 1, 1), (16, 1, 1)]
ref_state = None
for split, pdl, prio in configs:
    ll.pipeline_config(split=split, pdl=pdl, priorities=prio)
    st = ll.LunarState.from_numpy(s0, device=dev)
    gw = ll.StepsGraph(st, acts, warm, P)
    gt = ll.StepsGraph(st, acts, steps, P)
    gw.launch()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    gt.launch()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / steps
    # eager (un-graphed) multi-step enqueue of the same thing, for the launch-overhead comparison
    ev0.record()
    ll.steps(st, acts, steps, P)
    ev1.record()
    torch.cuda.synchronize()
    ms_eager = ev0.elapsed_time(ev1) / steps
    final = st.numpy()
    if ref_state is None:
        ref_state = final
    same = bool(np.array_equal(final, ref_state, equal_nan=True))
    print(json.dumps({"split": split, "pdl": pdl, "prio": prio, "envs": B, "us_per_step": 1e3 * ms,
                      "us_per_step_eager": 1e3 * ms_eager, "env_steps_per_s": B / (ms * 1e-3),
                      "nodes": gt.info()["nodes"], "identical_to_first": same,
                      "counters": ll.pipeline_counters(dev, B)}), flush=True)
    gw.close()
    gt.close()
