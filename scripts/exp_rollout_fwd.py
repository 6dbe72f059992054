"""Forward rollout of config 5 (4096 envs x 100 steps, tape kept) as ONE graph: fused kernel vs the
wavefront pipeline (split 1).  usage: python scripts/exp_rollout_fwd.py [envs]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gen_inputs import lunar_keys, perturbed_lunar_state  # noqa: E402
from parallax_b200 import lunar_lander as ll  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = 100
dev = torch.device("cuda", 0)
P = ll.default_params()
rng = np.random.default_rng(4)
s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), rng, P)
acts = torch.from_numpy(rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)).to(dev)
st0 = ll.LunarState.from_numpy(s0, device=dev)
ref = None
for mode in ("fused", "pipeline"):
    tape = torch.empty((T + 1, ll.NF, B), dtype=torch.float32, device=dev)
    if mode == "pipeline":
        ll._workspace(ll._xp.backend_of(st0.data), B)
    ll.rollout(st0, acts, T, P, mode=mode, tape=tape)
    g = ll.Graph(dev)
    with g:
        ll.rollout(st0, acts, T, P, mode=mode, tape=tape)
    g.keep(tape, acts, st0.data)
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.launch(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    same = None if ref is None else bool(torch.equal(torch.nan_to_num(tape, nan=-7.0), torch.nan_to_num(ref, nan=-7.0)))
    print(f"{mode}: {np.median(ts):.3f} ms per {T}-step rollout of {B} envs; identical to fused: {same}", flush=True)
    ref = tape.clone()
