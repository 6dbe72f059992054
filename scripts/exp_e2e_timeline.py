"""Where a host-facing step spends its time: the end of every kernel of every env range (library trace, eager
launches) next to the wall time of the whole step and of its copies alone.
usage: python scripts/exp_e2e_timeline.py [envs] [chunks]"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from gen_inputs import lunar_keys  # noqa: E402
from parallax_b200 import _lib  # noqa: E402
from parallax_b200 import lunar_lander as ll  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
L = _lib.lib()
P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000)
rng = np.random.default_rng(1)
acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).pin_memory() for _ in range(4)]
env = ll.LunarLanderEnv(B, P)
env.reset(lunar_keys(B))
env.set_host_graphs(False)
env.set_chunks(chunks)
for w in range(30):
    env.step(acts[w % 4])
names = "I N C F R".split()
for rep in range(3):
    L.cx_lunar_pipeline_trace(1)
    t0 = time.perf_counter()
    env.step(acts[rep % 4])
    dt = time.perf_counter() - t0
    out = (C.c_float * (16 * 8 * 5))()
    S, TT = C.c_int(), C.c_int()
    L.cx_lunar_pipeline_trace_read(out, C.byref(S), C.byref(TT))
    L.cx_lunar_pipeline_trace(0)
    print(f"step wall {dt * 1e3:.3f} ms; end of each kernel, us since the fork:")
    for r in range(S.value):
        print(f"  range {r}: " + " ".join(f"{names[k]}{1e3 * out[(r * TT.value) * 5 + k]:7.1f}" for k in range(5)))
n = 20
t0 = time.perf_counter()
for k in range(n):
    env.step(acts[k % 4])
print(f"eager, untraced: {(time.perf_counter() - t0) / n * 1e3:.3f} ms/step")
env.close()
