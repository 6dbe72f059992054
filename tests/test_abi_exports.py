"""The C-ABI shared library loads without a GPU, exports every symbol include/*.h declares, and
its argument validation / error reporting works.  No compute entry point is called here."""
import ctypes as C
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        for m in re.finditer(r"\b(cotix_[a-z0-9_]+)\s*\(", src):
            names.add(m.group(1))
    return names


@pytest.fixture(scope="module")
def lib():
    import parallax_b200

    parallax_b200.build()
    return parallax_b200._lib.lib()


def test_every_declared_symbol_is_exported(lib):
    import parallax_b200

    declared = _declared_symbols()
    assert len(declared) >= 24
    assert declared == set(parallax_b200._lib.EXPORTS), declared ^ set(parallax_b200._lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/ but not exported"


def test_abi_version_and_struct_layout(lib, oracle):
    import parallax_b200

    assert lib.cotix_abi_version() == 2
    # the ctypes mirror, the oracle's mirror and the C struct agree on the size
    assert C.sizeof(parallax_b200.LunarParams) == oracle.lib().oracle_sizeof_lunar_params()
    assert C.sizeof(parallax_b200.LunarParams) == C.sizeof(oracle.LunarParams)


def test_default_params_are_the_reference_literals(lib, oracle):
    """cotix_lunar_default_params is host-only: SURVEY §5 'config' literals."""
    import parallax_b200

    p = parallax_b200._lib.default_lunar_params()
    assert abs(p.dt - 1e-2) < 1e-9 and abs(p.gravity_dv[1] + 0.002) < 1e-9          # test_viz.py:26,30
    assert abs(p.baumgarte - 0.3) < 1e-7 and abs(p.baumgarte_dt - 0.01) < 1e-9       # _collision_resolution.py:105,115
    assert p.coin_p == 0.5 and p.gjk_max_iters == 32                                  # _colliders.py:220, _collisions.py:101
    assert (p.spring_kp, round(p.spring_kv0, 6), round(p.spring_kv, 6)) == (1.0, 0.1, 0.05)  # _lunar_lander.py:189
    assert abs(p.leg_damping - 0.95) < 1e-7                                           # _lunar_lander.py:208,211
    assert [p.body_props[0][k] for k in range(4)] == pytest.approx([30, 30, 1, 0.1])
    assert p.body_props[3][0] == float("inf") and p.body_props[3][1] == float("inf")
    assert p.substeps == 1 and p.auto_reset == 0 and p.judge_enabled == 0 and p.rng_mode == 0
    # identical, byte for byte, to the oracle's own defaults
    assert bytes(p) == bytes(oracle.default_lunar_params())


def test_argument_validation_reports_errors(lib):
    import parallax_b200

    p = parallax_b200._lib.default_lunar_params()
    # bad leading dimension
    rc = lib.cotix_lunar_step(None, 8, None, None, 4, None, None, None, None, C.byref(p))
    assert rc == -1 and b"leading dimension" in lib.cotix_last_error()
    # bad params
    q = p.copy()
    q.substeps = 0
    rc = lib.cotix_lunar_reset(None, 0, None, None, 0, C.byref(q))
    assert rc == -1 and b"params" in lib.cotix_last_error()
    q = p.copy()
    q.rng_mode = 7
    assert lib.cotix_lunar_step(None, 0, None, None, 0, None, None, None, None, C.byref(q)) == -1
    # null buffers with B > 0
    assert lib.cotix_lunar_reset(None, 4, None, None, 4, C.byref(p)) == -1
    assert lib.cotix_lunar_default_params(None) == -1
    assert lib.cotix_lunar_env_create(None, 0, 4, C.byref(p)) == -1


def test_no_device_fails_loudly_not_silently(lib):
    """Without a usable sm_100 device every compute entry point returns an error code — there is
    no CPU fallback (skipped where a GPU is present: the GPU tests cover that side)."""
    import torch

    import parallax_b200

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    np_ = parallax_b200._lib.default_narrow_params()
    rc = lib.cotix_narrowphase_host(0, 4, None, None, None, None, None, None, C.byref(np_))
    assert rc < 0
    h = C.c_void_p()
    p = parallax_b200._lib.default_lunar_params()
    assert lib.cotix_lunar_env_create(C.byref(h), 0, 4, C.byref(p)) < 0
    assert lib.cotix_last_error()


def test_product_package_does_not_touch_the_oracle():
    """The oracle is test infrastructure: nothing under parallax_b200/ may import, link or call it."""
    for path in glob.glob(os.path.join(ROOT, "parallax_b200", "**", "*"), recursive=True):
        if os.path.isdir(path) or path.endswith((".so", ".pyc")):
            continue
        src = open(path, errors="ignore").read()
        for needle in ("import oracle", "from oracle", "liboracle", "oracle/", "oracle_", "cotix_oracle"):
            assert needle not in src, (path, needle)


def test_header_constants_match_the_python_mirror():
    """Routing constants the Python mirror repeats: multi-step calls up to COTIX_ROLLOUT_MAX_ENVS envs go to the
    block-resident rollout kernel (no workspace), larger ones to the pipeline."""
    import re

    from parallax_b200 import jax_ffi
    from parallax_b200 import lunar_lander as ll

    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "cotix_b200.h")).read()
    m = re.search(r"#define\s+COTIX_ROLLOUT_MAX_ENVS\s+(\d+)", hdr)
    assert m and int(m.group(1)) == ll.ROLLOUT_MAX_ENVS
    assert not ll._use_pipe("auto", ll.ROLLOUT_MAX_ENVS, n_steps=100) and ll._use_pipe("auto", ll.ROLLOUT_MAX_ENVS + 1, n_steps=100)
    assert not ll._use_pipe("auto", ll.PIPELINE_MIN_ENVS - 1) and ll._use_pipe("auto", ll.PIPELINE_MIN_ENVS)
    assert ll._use_pipe("auto", 8192, n_steps=10, substeps=4) is False  # sub-steps: the per-step routing
    assert "4736" in open(jax_ffi.__file__).read()
