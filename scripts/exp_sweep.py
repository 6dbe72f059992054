"""Experiment: mixed-type narrowphase kernel vs the wavefront sweep (BASELINE config 2)."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from parallax_b200 import shapes as sh
from gen_inputs import random_pairs

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
rng = np.random.default_rng(2)
a = torch.empty((n, 16), dtype=torch.float32, device="cuda"); b = torch.empty_like(a)
chunk = 1 << 21
for s in range(0, n, chunk):
    ha, hb = random_pairs(rng, min(chunk, n - s))
    a[s:s + len(ha)] = torch.from_numpy(ha).cuda(); b[s:s + len(hb)] = torch.from_numpy(hb).cuda()
A, B = sh.ShapeBatch(a), sh.ShapeBatch(b)
ref = None
for mode in ("single", "pipeline"):
    for _ in range(3):
        ci = sh.narrowphase(A, B, mode=mode)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ci = sh.narrowphase(A, B, mode=mode)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{mode}: {ms:.2f} ms  {n / ms * 1e3:.3e} pairs/s  hits {ci.exists.float().mean().item():.3f}", flush=True)
    cur = (ci.exists.clone(), ci.penetration_vector.clone(), ci.contact_point.clone(), ci.status.clone())
    if ref is not None:
        same = all(torch.equal(torch.nan_to_num(x.float(), nan=-7.0), torch.nan_to_num(y.float(), nan=-7.0)) for x, y in zip(ref, cur))
        print("identical to single:", same)
    ref = cur
if sh._sweep_ws:
    ws = list(sh._sweep_ws.values())[0]
    off = ((ws.data_ptr() + 255) & ~255) - ws.data_ptr()
    print("queue/hit counts:", ws[off:off + 24].view(torch.int32).cpu().numpy())
