"""Timing of the headless rasteriser: LunarLander frames per second and HBM write rate.
Usage: python scripts/exp_render.py [n_frames]"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gen_inputs import lunar_keys  # noqa: E402

from parallax_b200 import _lib, viz  # noqa: E402
from parallax_b200 import lunar_lander as ll  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
P = ll.default_params()
st = ll.reset(lunar_keys(n), P)
img = torch.empty((n, 600, 800, 3), dtype=torch.uint8, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for rep in range(3):
    ev[0].record()
    for _ in range(10):
        _lib.check(_lib.lib().cotix_lunar_draw(stream, n, st.data.data_ptr(), n, None, img.data_ptr(), C.byref(P)))
    ev[1].record()
    torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / 10
print(f"lunar_draw: {n} frames in {ms:.3f} ms = {n / ms * 1e3:.3e} frames/s, {n * 1.440144e6 / ms / 1e6:.0f} GB/s written")
prims = viz.lunar_primitives(st, params=P)
for rep in range(2):
    ev[0].record()
    for _ in range(10):
        _lib.check(_lib.lib().cotix_raster_draw(stream, n, prims.data_ptr(), None, 44, img.data_ptr()))
    ev[1].record()
    torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / 10
print(f"raster_draw (44 prims from HBM): {ms:.3f} ms = {n * 1.44e6 / ms / 1e6:.0f} GB/s written")
z = torch.empty_like(img)
ev[0].record()
for _ in range(10):
    z.zero_()
ev[1].record()
torch.cuda.synchronize()
print(f"memset of the same buffer: {ev[0].elapsed_time(ev[1]) / 10:.3f} ms")
