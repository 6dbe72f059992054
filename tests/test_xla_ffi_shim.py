"""The XLA FFI shim (parallax_b200/csrc/cotix_xla_ffi.cc) — the jax.ffi boundary of north_star.
jax / jaxlib are not installed on either machine, so the handlers are compiled against a minimal
stand-in for the XLA FFI headers (tests/xla_ffi_mock) and their bodies are executed on the GPU with
raw device pointers; results must equal the ctypes path (which the other tests pin on the oracle).
CPU part: the shim builds as a stub without headers and as the real thing with them, and exports
every handler symbol parallax_b200/jax_ffi.py registers."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from gen_inputs import lunar_keys, perturbed_lunar_state, random_pairs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MOCK = os.path.join(ROOT, "tests", "xla_ffi_mock")


def _build_driver(tmp) -> str:
    import parallax_b200

    parallax_b200.build()
    out = os.path.join(tmp, "libshim_test.so")
    lib_dir = os.path.dirname(parallax_b200._lib.LIB_PATH)
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", f"-I{MOCK}", "-I/usr/local/cuda/include",
           os.path.join(MOCK, "driver.cc"), "-o", out, f"-L{lib_dir}", "-l:libcotix_b200.so",
           "-L/usr/local/cuda/lib64", "-lcudart", f"-Wl,-rpath,{lib_dir}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    return out


def test_shim_builds_as_stub_and_with_headers(tmp_path):
    from parallax_b200 import jax_ffi

    # without XLA FFI headers: the stub
    stub = jax_ffi.build_shim(include_dir=None, out=str(tmp_path / "stub.so")) if not jax_ffi.available() else None
    if stub:
        assert C.CDLL(stub).cotix_xla_ffi_available() == 0
    # with (mock) headers: every handler jax_ffi.register() looks up is exported
    real = jax_ffi.build_shim(include_dir=MOCK, out=str(tmp_path / "shim.so"), extra=["-DCOTIX_XLA_FFI_MOCK"])
    so = C.CDLL(real)
    assert so.cotix_xla_ffi_available() == 1
    for sym in jax_ffi.TARGETS.values():
        assert hasattr(so, sym), sym
    # the params attribute is the struct as int32 words
    P = jax_ffi.default_params(judge_enabled=1)
    w = jax_ffi.params_words(P)
    assert w.dtype == np.int32 and w.nbytes == C.sizeof(P)
    assert not jax_ffi.available() or callable(jax_ffi.register)
    if not jax_ffi.available():
        with pytest.raises(Exception):
            jax_ffi.register()


@pytest.mark.gpu
def test_shim_handlers_execute_and_match_ctypes_path(cuda_lib, tmp_path):
    import torch
    from parallax_b200 import jax_ffi
    from parallax_b200 import lunar_lander as ll
    from parallax_b200 import shapes as sh

    so = C.CDLL(_build_driver(str(tmp_path)))
    err = C.create_string_buffer(512)

    def ok(rc):
        assert rc == 0, err.value.decode()

    def eq(a, b, what):
        a, b = a.cpu().numpy(), b.cpu().numpy()
        good = (a == b) | (np.isnan(a) & np.isnan(b)) if a.dtype.kind == "f" else (a == b)
        assert good.all(), what

    st = torch.cuda.current_stream().cuda_stream
    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=12)
    w = jax_ffi.params_words(P)
    wp, nw = w.ctypes.data_as(C.POINTER(C.c_int32)), int(w.size)
    B = 20000 + 9
    keys = lunar_keys(B, offset=1)
    kd = torch.from_numpy(keys.view(np.int32)).cuda()
    state = torch.empty((36, B), device="cuda")
    ok(so.shim_reset(C.c_void_p(st), C.c_int64(B), C.c_void_p(kd.data_ptr()), C.c_void_p(state.data_ptr()), wp, nw, err, 512))
    eq(state, ll.reset(keys, P).data, "reset")
    rng = np.random.default_rng(3)
    s0 = torch.from_numpy(perturbed_lunar_state(state.cpu().numpy(), rng, P)).cuda()
    acts = torch.from_numpy(rng.uniform(-1, 1, size=(3, 2, B)).astype(np.float32)).cuda()
    ws = torch.empty(jax_ffi.workspace_bytes(B) + 256, dtype=torch.uint8, device="cuda")
    wsp = (ws.data_ptr() + 255) & ~255
    wsn = ws.numel() - (wsp - ws.data_ptr())
    for use_ws in (0, 1):
        out, rew, done, obs = (torch.empty((36, B), device="cuda"), torch.empty(B, device="cuda"),
                               torch.empty(B, dtype=torch.uint8, device="cuda"), torch.empty((8, B), device="cuda"))
        ok(so.shim_step(C.c_void_p(st), C.c_int64(B), C.c_void_p(s0.data_ptr()), C.c_void_p(acts[0].data_ptr()),
                        C.c_void_p(wsp if use_ws else 0), C.c_int64(wsn if use_ws else 0), C.c_void_p(out.data_ptr()),
                        C.c_void_p(rew.data_ptr()), C.c_void_p(done.data_ptr()), C.c_void_p(obs.data_ptr()), wp, nw, err, 512))
        ref, rr, rd, ro_ = ll.step(ll.LunarState(s0), acts[0], P, want_obs=True, mode="pipeline" if use_ws else "fused")
        eq(out, ref.data, f"step ws={use_ws}"); eq(rew, rr, "reward"); eq(done, rd, "done"); eq(obs, ro_, "obs")
    # n steps in one handler call, not aliased (the handler copies the state first)
    out = torch.empty((36, B), device="cuda")
    ok(so.shim_steps(C.c_void_p(st), C.c_int64(B), C.c_void_p(s0.data_ptr()), C.c_void_p(acts.data_ptr()), C.c_int64(3),
                     C.c_void_p(wsp), C.c_int64(wsn), C.c_void_p(out.data_ptr()), C.c_void_p(rew.data_ptr()),
                     C.c_void_p(done.data_ptr()), C.c_void_p(obs.data_ptr()), wp, nw, C.c_int32(7), err, 512))
    ref = ll.LunarState(s0.clone())
    r2, d2, o2 = ll.steps(ref, acts, 7, P, want_obs=True, mode="pipeline")
    eq(out, ref.data, "steps"); eq(rew, r2, "steps reward"); eq(obs, o2, "steps obs")
    # contact functions
    a, b = random_pairs(rng, 5000)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    ex, pen, cp, status = (torch.empty(5000, dtype=torch.uint8, device="cuda"), torch.empty((5000, 2), device="cuda"),
                           torch.empty((5000, 2), device="cuda"), torch.empty(5000, dtype=torch.uint8, device="cuda"))
    npar = jax_ffi._lib.default_narrow_params()
    idir = (C.c_float * 2)(npar.init_dir[0], npar.init_dir[1])
    ok(so.shim_narrowphase(C.c_void_p(st), C.c_int64(5000), C.c_void_p(da.data_ptr()), C.c_void_p(db.data_ptr()),
                           C.c_void_p(0), C.c_int64(0), C.c_void_p(ex.data_ptr()), C.c_void_p(pen.data_ptr()),
                           C.c_void_p(cp.data_ptr()), C.c_void_p(status.data_ptr()), idir, C.c_int32(32), C.c_int32(128), err, 512))
    ci = sh.narrowphase(sh.ShapeBatch(da), sh.ShapeBatch(db), mode="single")
    eq(ex.bool(), ci.exists, "exists"); eq(pen, ci.penetration_vector, "pen"); eq(cp, ci.contact_point, "cp")
    # rollout + adjoint = what jax.custom_vjp calls
    P0 = ll.default_params()
    w0 = jax_ffi.params_words(P0)
    w0p = w0.ctypes.data_as(C.POINTER(C.c_int32))
    Bd, T = 1024, 8
    sd = torch.from_numpy(perturbed_lunar_state(ll.reset(lunar_keys(Bd), P0).numpy(), rng, P0)).cuda()
    ad = torch.from_numpy(rng.uniform(-1, 1, size=(T, 2, Bd)).astype(np.float32)).cuda()
    tape = torch.empty((T + 1, 36, Bd), device="cuda")
    ok(so.shim_rollout(C.c_void_p(st), C.c_int64(Bd), C.c_int64(T), C.c_void_p(sd.data_ptr()), C.c_void_p(ad.data_ptr()),
                       C.c_void_p(0), C.c_int64(0), C.c_void_p(tape.data_ptr()), w0p, int(w0.size), err, 512))
    ro = ll.rollout(ll.LunarState(sd), ad, T, P0, mode="fused")
    eq(tape, ro.tape, "tape")
    gf = torch.from_numpy(rng.normal(size=(18, Bd)).astype(np.float32)).cuda()
    aws = ll.adjoint_workspace(Bd, T, sd.device, jacobians=False)  # decisions only: deterministic bits
    awp = (aws.data_ptr() + 255) & ~255
    g0, ga, fl = torch.empty((18, Bd), device="cuda"), torch.full((T, 2, Bd), 7.0, device="cuda"), torch.full((Bd,), 9, dtype=torch.int32, device="cuda")
    ok(so.shim_rollout_bwd(C.c_void_p(st), C.c_int64(Bd), C.c_int64(T), C.c_void_p(tape.data_ptr()), C.c_void_p(ad.data_ptr()),
                           C.c_void_p(gf.data_ptr()), C.c_void_p(awp), C.c_int64(aws.numel() - (awp - aws.data_ptr())),
                           C.c_void_p(g0.data_ptr()), C.c_void_p(ga.data_ptr()), C.c_void_p(fl.data_ptr()), w0p, int(w0.size),
                           C.c_int32(0), err, 512))
    flags = torch.zeros(Bd, dtype=torch.int32, device="cuda")
    g0r, gar = ll.rollout_vjp(ro, gf, flags, jacobians=False)
    eq(g0, g0r, "g0"); eq(ga, gar, "g_actions"); eq(fl, flags, "flags")
    # a bad params attribute is an error, not a crash
    assert so.shim_reset(C.c_void_p(st), C.c_int64(B), C.c_void_p(kd.data_ptr()), C.c_void_p(state.data_ptr()), wp, nw - 1, err, 512) != 0
    assert b"params" in err.value
