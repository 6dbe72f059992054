"""Re-executes a small part of the reference's own source under oracle/jaxlite (the way
oracle/gen_golden.py made tests/golden/) and checks it still reproduces the committed fixtures.
Needs /root/reference (the build container); skipped elsewhere.  CPU only."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SCRIPT = r"""
import sys, os
sys.argv = ["gen_golden.py", "none"]
import runpy
ns = runpy.run_path(os.path.join(%(root)r, "oracle", "gen_golden.py"), run_name="gen")
import numpy as np
g = np.load(os.path.join(%(root)r, "tests", "golden", "contacts.npz"))
FUNCS, shape_from_rec, f32 = ns["FUNCS"], ns["shape_from_rec"], ns["f32"]
bad = 0
for k in range(0, 360, 9):
    info = FUNCS[(int(g["a"][k, 0]), int(g["b"][k, 0]))](shape_from_rec(g["a"][k]), shape_from_rec(g["b"][k]))
    for got, want in ((f32(info.penetration_vector), g["pen"][k]), (f32(info.contact_point), g["cp"][k])):
        if not (((got == want) | (np.isnan(got) & np.isnan(want))).all()):
            bad += 1
r = np.load(os.path.join(%(root)r, "tests", "golden", "reset.npz"))
env = ns["LunarLander"](ns["jnp"].array(r["keys"][5]))
for q, p in enumerate(env.bodies[3].shape.parts):
    if not (f32(p.vertices) == r["ground_vertices"][5, q]).all():
        bad += 1
print("BAD", bad)
"""


@pytest.mark.skipif(not os.path.isdir("/root/reference/cotix"), reason="reference tree not mounted")
def test_reference_source_still_reproduces_the_fixtures(oracle):
    out = subprocess.run([sys.executable, "-c", _SCRIPT % {"root": ROOT}], capture_output=True, text=True,
                         timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "BAD 0" in out.stdout, out.stdout[-500:]
