"""The host-side mirror must work in a process that has no torch (north_star: "called from JAX
through a thin C-ABI ... with no PyTorch").  Each test runs a child interpreter in which `import
torch` raises ImportError."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PRELUDE = """
import sys
sys.modules['torch'] = None          # `import torch` now raises ImportError
sys.path.insert(0, {root!r}); sys.path.insert(0, {tests!r})
import numpy as np
""".format(root=ROOT, tests=os.path.join(ROOT, "tests"))


def _run(body: str, timeout=900):
    code = PRELUDE + textwrap.dedent(body)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=timeout)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    return res.stdout


def test_package_imports_without_torch():
    out = _run("""
        import parallax_b200
        from parallax_b200 import lunar_lander, shapes, scene, envs, viz, sharding, devmem, jax_ffi, _xp
        assert 'torch' not in [m for m in sys.modules if sys.modules[m] is not None]
        assert not _xp.torch_available()
        assert _xp.backend_for('cuda').name == 'native' and _xp.backend_for('native:1').device == 1
        assert jax_ffi.available() in (True, False)
        P = lunar_lander.default_params()
        assert P.solver_loops == 1 and abs(P.dt - 1e-2) < 1e-9
        print('ok')
    """)
    assert "ok" in out


@pytest.mark.gpu
def test_native_backend_matches_oracle():
    """reset / step / steps / StepsGraph / rollout + adjoint / narrowphase / scene step / frames on
    `devmem.DeviceArray` memory, no torch in the process, bit-identical to the oracle."""
    out = _run("""
        from oracle import oracle as o
        from gen_inputs import lunar_keys, perturbed_lunar_state, random_pairs
        from parallax_b200 import lunar_lander as ll, shapes as sh, scene as sc, viz, devmem
        def eq(a, b, what):
            a, b = np.asarray(a), np.asarray(b)
            ok = (a == b) | (np.isnan(a) & np.isnan(b)) if a.dtype.kind == 'f' else (a == b)
            assert a.shape == b.shape and ok.all(), what
        o.set_trig_mode(1)
        P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=20)
        B = 3000 + 7
        keys = lunar_keys(B, offset=4)
        st = ll.reset(keys, P, device='native')
        assert isinstance(st.data, devmem.DeviceArray)
        ref = o.lunar_reset(keys, P)
        eq(st.numpy(), ref, 'reset')
        rng = np.random.default_rng(7)
        s0 = perturbed_lunar_state(ref, rng, P)
        st = ll.LunarState.from_numpy(s0, device='native')
        ref = s0
        for t, mode in enumerate(['fused', 'pipeline', 'fused', 'pipeline']):
            act = rng.uniform(-1, 1, size=(2, B)).astype(np.float32)
            st, rew, done, obs = ll.step(st, act, P, want_obs=True, mode=mode, inplace=(t % 2 == 1))
            ref, orew, odone, oobs = o.lunar_step(ref, act, P, want_obs=True)
            eq(st.numpy(), ref, f'step {t}'); eq(rew.numpy(), orew, 'reward'); eq(done.numpy(), odone, 'done')
            eq(obs.numpy(), oobs, 'obs')
        acts = rng.uniform(-1, 1, size=(3, 2, B)).astype(np.float32)
        ll.pipeline_config(split=3)
        rew, done = ll.steps(st, acts, 5, P, mode='pipeline')
        for t in range(5):
            ref, orew, odone = o.lunar_step(ref, acts[t % 3], P)
        eq(st.numpy(), ref, 'steps'); eq(rew.numpy(), orew, 'steps reward')
        g = ll.StepsGraph(st, acts, 4, P, want_obs=True, mode='pipeline')
        g.launch(); g.launch(); devmem.synchronize()
        for t in range(8):
            ref, orew, odone, oobs = o.lunar_step(ref, acts[(t % 4) % 3], P, want_obs=True)
        eq(st.numpy(), ref, 'graph replay'); eq(g.obs.numpy(), oobs, 'graph obs')
        assert g.info()['kernel_launches'] == 3 * 4 * 6
        g.close()
        ll.pipeline_config(split=0)
        # differentiable rollout: replay must land on the tape, gradients finite
        P0 = ll.default_params()
        T, Bd = 6, 512
        sd = perturbed_lunar_state(o.lunar_reset(lunar_keys(Bd), P0), rng, P0)
        ad = rng.uniform(-1, 1, size=(T, 2, Bd)).astype(np.float32)
        ro = ll.rollout(ll.LunarState.from_numpy(sd, device='native'), ad, T, P0, mode='fused')
        r2 = sd
        for t in range(T):
            r2 = o.lunar_step(r2, ad[t], P0)[0]
        eq(ro.final.numpy(), r2, 'rollout tape')
        flags = devmem.DeviceArray.zeros((Bd,), np.uint32)
        g0, ga = ll.rollout_vjp(ro, rng.normal(size=(18, Bd)).astype(np.float32), flags)
        f = flags.numpy()
        assert not (f & ll.ADJ_REPLAY_MISMATCH).any()
        ok = (f & ll.ADJ_NONFINITE) == 0
        assert np.isfinite(g0.numpy()[:, ok]).all() and np.isfinite(ga.numpy()[:, :, ok]).all()
        # narrowphase sweep
        a, b = random_pairs(rng, 70000)
        ci = sh.narrowphase(sh.from_records(a, device='native'), sh.from_records(b, device='native'))
        ex, pen, cp, _ = o.narrowphase(a, b)
        e2, p2, c2, _s = ci.numpy()
        eq(e2.astype(bool), ex.astype(bool), 'exists'); eq(p2, pen, 'pen'); eq(c2, cp, 'cp')
        circ = sh.Circle(np.ones(4, np.float32), np.zeros((4, 2), np.float32), device='native')
        eq(sh.circle_vs_circle(circ, circ.move(np.full((4, 2), 0.5, np.float32))).numpy()[0], np.ones(4, bool), 'typed')
        # generic scene step
        spec = sc.robocup_scene(n_robots=2, areas_collide=False)
        scene = sc.Scene(spec)
        Bs = 16
        s6 = np.repeat(spec.init_state[:, :, None], Bs, axis=2).astype(np.float32)
        s6[4:, 0] = rng.uniform(-4.9, -4.3, (spec.nb - 4, Bs)); s6[4:, 1] = rng.uniform(-0.6, 0.6, (spec.nb - 4, Bs))
        skeys = np.stack([np.zeros(Bs, np.uint32), np.arange(Bs, dtype=np.uint32)], 1)
        dstate = devmem.DeviceArray.from_numpy(np.ascontiguousarray(s6.reshape(spec.nb * 6, Bs)))
        dkeys = devmem.DeviceArray.from_numpy(np.ascontiguousarray(skeys.T))
        scene.step(dstate, dkeys)
        for e in range(Bs):
            outb, _c, _ch = o.scene_step(spec.bodies10(s6[:, :, e]), spec.part_body, spec.part_recs, skeys[e])
            s6[:, :, e] = outb[:, 2:8]
        eq(dstate.numpy().reshape(spec.nb, 6, Bs), s6, 'scene step')
        scene.close()
        # frames
        fr = viz.render_lunar(ll.LunarState.from_numpy(sd[:, :4].copy(), device='native'), params=P0)
        pr = viz.lunar_primitives(ll.LunarState.from_numpy(sd[:, :4].copy(), device='native'), params=P0)
        eq(fr.numpy(), o.raster_draw(pr.numpy()), 'frames')
        # host-facing env handle with page-locked numpy buffers
        env = ll.LunarLanderEnv(1000, P)
        k = lunar_keys(1000); env.reset(k); r3 = o.lunar_reset(k, P)
        for t in range(3):
            act = rng.uniform(-1, 1, size=(2, 1000)).astype(np.float32)
            obs, rew, done = env.step(act)
            r3, orew, odone, oobs = o.lunar_step(r3, act, P, want_obs=True)
            eq(obs, oobs, 'env obs'); eq(rew, orew, 'env reward')
        env.close()
        print('native ok')
    """)
    assert "native ok" in out
