"""Timeline of a multi-range pipeline enqueue (eager launches, an event after every kernel): when
does every stage of every env range END, relative to the fork?  Shows whether the sparse stages of
one range really run under the dense stages of another.
usage: python scripts/exp_timeline.py [envs]"""
import ctypes as C
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
T = 4
dev = torch.device("cuda", 0)
L = _lib.lib()
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), np.random.default_rng(1000), P)
acts = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, size=(4, 2, B)).astype(np.float32)).to(dev)
names = "I N C F R".split()
for split, pdl, prio in [(1, 0, 0), (2, 0, 0), (2, 0, 1), (4, 0, 0), (4, 0, 1), (4, 1, 1)]:
    ll.pipeline_config(split=split, pdl=pdl, priorities=prio)
    st = ll.LunarState.from_numpy(s0, device=dev)
    ll.steps(st, acts, 40, P)
    torch.cuda.synchronize()
    L.cx_lunar_pipeline_trace(1)
    ll.steps(st, acts, T, P)
    out = (C.c_float * (16 * 8 * 5))()
    S, TT = C.c_int(), C.c_int()
    L.cx_lunar_pipeline_trace_read(out, C.byref(S), C.byref(TT))
    L.cx_lunar_pipeline_trace(0)
    print(f"== split {split} pdl {pdl} prio {prio}: end of each kernel, us since fork")
    for r in range(S.value):
        row = []
        for t in range(TT.value):
            row.append(" ".join(f"{names[k]}{1e3 * out[(r * TT.value + t) * 5 + k]:7.1f}" for k in range(5)))
        print(f"  range {r}: " + " | ".join(row))
    sys.stdout.flush()
