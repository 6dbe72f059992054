"""Backward half of config 5 (4096 envs x 100 steps, loss = final lander y^2) as one graph, with a timeline
of its kernels (one event per eager launch is not available through a graph, so the eager pass is timed too).
Environment knobs of the library: COTIX_SWEEP_KERNEL=0 (one sweep launch per step), COTIX_ADJ_RUNS=1 (no runs),
COTIX_ADJ_OVERLAP=0.  usage: python scripts/exp_rollout_bwd.py [envs] [jacobians 0/1]"""
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
jac = None if len(sys.argv) <= 2 else bool(int(sys.argv[2]))
T = 100
dev = torch.device("cuda", 0)
P = ll.default_params()
rng = np.random.default_rng(4)
s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), rng, P)
acts = torch.from_numpy(rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)).to(dev)
st0 = ll.LunarState.from_numpy(s0, device=dev)
flags = torch.zeros(B, dtype=torch.int32, device=dev)
tape = torch.empty((T + 1, ll.NF, B), dtype=torch.float32, device=dev)
g = torch.zeros((18, B), device=dev)
g_act = torch.zeros((T, 2, B), device=dev)
ws = ll.adjoint_workspace(B, T, dev, jacobians=jac)
ro = ll.rollout(st0, acts, T, P, tape=tape)


def backward():
    g.zero_()
    torch.mul(tape[-1, 1], 2.0, out=g[1])
    return ll.rollout_vjp(ro, g, flags, workspace=ws, out=(g, g_act))


backward()
torch.cuda.synchronize()
gb = ll.Graph(dev)
with gb:
    backward()
gb.keep(tape, acts, g, g_act, ws, flags, st0.data)
ts = []
for _ in range(7):
    flags.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); gb.launch(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
chk = float(torch.nan_to_num(g).double().abs().sum()) + float(torch.nan_to_num(g_act).double().abs().sum())
print(f"backward: {np.median(ts):.3f} ms per {T}-step adjoint of {B} envs, graph nodes {gb.info()['nodes']}, "
      f"checksum {chk:.9e}", flush=True)
