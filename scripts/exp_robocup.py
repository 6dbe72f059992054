"""Experiment: generic scene step on BASELINE config 4 (RoboCup, 6 + 6 robots, B = 262144)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
if os.environ.get("EXP_LIB"):  # time a variant build of the library
    import parallax_b200._lib as _L
    _L.LIB_PATH = os.path.abspath(os.environ["EXP_LIB"])
from parallax_b200 import scene as sc

B = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
for areas in (False, True):
    spec = sc.robocup_scene(n_robots=12, areas_collide=areas)
    scene = sc.Scene(spec)
    print("areas_collide", areas, scene.info())
    rng = np.random.default_rng(3)
    st = np.repeat(spec.init_state[:, :, None], B, axis=2).astype(np.float32)
    dyn = np.isfinite(spec.body_props[:, 0])
    for b in np.nonzero(dyn)[0]:
        st[b, 0] += rng.normal(0, 0.3, B); st[b, 1] += rng.normal(0, 0.3, B)
        st[b, 2:4] += rng.normal(0, 1.0, (2, B))
    state = torch.from_numpy(np.ascontiguousarray(st.reshape(spec.nb * 6, B))).cuda()
    keys = torch.from_numpy(np.stack([np.zeros(B, np.uint32), np.arange(B, dtype=np.uint32)]).view(np.int32)).cuda()
    for w in range(5):
        scene.step(state, keys)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for k in range(n):
        scene.step(state, keys)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    cell, chosen, status = scene.step(state, keys, want_cells=True)
    print(f"  {ms:.3f} ms/step  {B / ms * 1e3:.3e} env-steps/s  live cells/env {cell.float().sum().item() / B:.2f}"
          f"  nan envs {torch.isnan(state).any(0).float().mean().item():.3f}", flush=True)
