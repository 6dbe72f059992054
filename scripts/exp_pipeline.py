"""Experiment: fused kernel vs wavefront pipeline on the bench workload (device-resident)."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
if os.environ.get("EXP_LIB"):  # time a variant build of the library (scripts only)
    import parallax_b200._lib as _L
    _L.LIB_PATH = os.path.abspath(os.environ["EXP_LIB"])
from parallax_b200 import lunar_lander as ll
from gen_inputs import lunar_keys, perturbed_lunar_state

if os.environ.get("EXP_NO_PDL"):
    from parallax_b200 import _lib as _LL
    _LL.lib().cx_lunar_pipeline_set_pdl(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 1 << 20
for rng_mode in (0, 1):
    P = ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1, max_episode_steps=1000, rng_mode=rng_mode)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P).numpy(), np.random.default_rng(1000), P)
    rng = np.random.default_rng(1)
    acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).cuda() for _ in range(4)]
    for mode in (("fused", "pipeline") if "--fused" in sys.argv else ("pipeline",)):
        st = ll.LunarState.from_numpy(s0)
        for w in range(20):
            ll.step(st, acts[w % 4], P, inplace=True, mode=mode)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 100
        e0.record()
        for k in range(n):
            ll.step(st, acts[k % 4], P, inplace=True, mode=mode)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"rng={rng_mode} {mode}: {ms*1e3:.1f} us/step  {B/ms*1e3:.3e} env-steps/s", flush=True)
        if mode == "pipeline":
            ll.pipeline_profile(True)
            for k in range(50):
                ll.step(st, acts[k % 4], P, inplace=True, mode=mode)
            torch.cuda.synchronize()
            sm, n = ll.pipeline_stage_ms()
            ll.pipeline_profile(False)
            ws = list(ll._workspaces.values())[0]
            off = ((ws.data_ptr() + 255) & ~255) - ws.data_ptr()
            ctr = ws[off:off + 64].view(torch.int32).cpu().numpy()
            print("   counters: items", ctr[0], "hits", ctr[1], "contact envs", ctr[2], "resets", ctr[3], "overflow", ctr[4])
            if "--hist" in sys.argv:
                # layout of the workspace (csrc/lunar_pipeline.cu carve()): 256-byte aligned blocks
                def al(x): return (x + 255) // 256 * 256
                cap = 4 * B + 4096
                o_ctr = 0; o_cand = al(64); o_base = o_cand + al(4 * B); o_rec = o_base + al(4 * B)
                o_item = o_rec + al(192 * B); o_res = o_item + al(4 * cap)
                n = int(ctr[0])
                item = ws[off + o_item: off + o_item + 4 * n].view(torch.int32).cpu().numpy().view(np.uint32)
                cpx = ws[off + o_res + 2 * 4 * cap: off + o_res + 2 * 4 * cap + 4 * n].view(torch.float32).cpu().numpy()
                pid = item & 31
                live = ~np.isnan(cpx)
                names = {0: "rl-L", 1: "ll-L", 9: "ll-rl", 10: "ll-ll"}
                for k in range(25):
                    nm = names.get(k, ("g%d-L" % (k - 2)) if k <= 8 else (("g%d-rl" % (k - 11)) if k <= 17 else "g%d-ll" % (k - 18)))
                    print(f"      pair {k:2d} {nm:6s} items {int((pid == k).sum()):7d} live {int(((pid == k) & live).sum()):7d}")
            print("   stages us:", {k: round(v * 1e3, 1) for k, v in sm.items()}, flush=True)
