#!/usr/bin/env python
"""bench.py — LunarLander env-steps/s (BASELINE.json metric) on N B200s, one process per GPU.

  python bench.py --gpus 1 --steps K --warmup W
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      # the CPU restatement of the reference on host cores

A "step" = one physics step of every env of the rank's batch (fresh action, judge, termination,
auto-reset).  Workload = BASELINE config 3: B = 1,048,576 envs PER GPU (weak scaling; envs are
independent, so the batch is sharded with no data-path collective), perturbed initial poses,
random actions, auto-reset on.

How `value` is timed: the K steps are ONE replayed CUDA graph (`cotix_lunar_steps_ws` captured by
`cotix_graph_*`: the single dispatch per rollout that jax.jit gives the reference's loop,
examples/test_viz.py:24,46-48), bracketed by CUDA events on the launching stream between barriers.
The host issues one call for the whole timed region, so host jitter of N processes cannot land in
the number.  `e2e` is the same metric through the host-facing env handle (`cotix_lunar_env_step_host`)
with pinned HOST buffers, H2D of the actions and D2H of obs / reward / done inside the timed region.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

ALGO_BYTES_PER_ENV_STEP = 224.0   # 136 B read + 88 B written at K=1
# algorithmic bytes per env of the pipeline stages:
#   integrate: read 18 dyn + 12 terrain + status + 2 action words, write 12 changed words + cand
#   finish   : read 18 dyn + 6 meta + 2 action + cand words, write 9 velocity + 6 meta words,
#              reward (4 B) and done (1 B)
#   narrow   : per collision item: item word + two 4-gons (2 x 32 B) read, 24-B contact written
#   contact  : per env with a candidate contact: 18 dyn + key (8 B) + cand word read, 9 velocity
#              words + status written; plus 24 B read per item
#   reset    : per reset env: key read, 36 state words written
STAGE_ALGO_BYTES = {"integrate": 4.0 * (18 + 12 + 1 + 2) + 4.0 * (12 + 1),
                    "finish": 4.0 * (18 + 6 + 2 + 1) + 4.0 * (9 + 6) + 5.0}
ALGO_BYTES_PER_PAIR = 148.0       # 2 x 64 B read + 20 B written
L2_BYTES = 126e6


def stage_algo_bytes(stage, B, ctr):
    """Algorithmic bytes of one launch of a pipeline stage."""
    if stage in STAGE_ALGO_BYTES:
        return STAGE_ALGO_BYTES[stage] * B
    if ctr is None:
        return None
    if stage == "narrow":
        return (4.0 + 64.0 + 24.0) * ctr["items"]
    if stage == "contact":
        return (72.0 + 8.0 + 4.0 + 36.0 + 4.0) * ctr["contact_envs"] + 24.0 * ctr["items"]
    if stage == "reset":
        return (8.0 + 144.0) * ctr["resets"]
    return None


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def _ncu():
    """Figures of the committed ncu --set full captures (profiles/roofline_traffic.json): DRAM bytes
    per launch, issue-slot and FP32-pipe utilisation per kernel.  Static evidence, named as such."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f)
    return {}


def _traffic(kernel):
    v = _ncu().get(kernel)
    return v.get("dram_bytes") if isinstance(v, dict) else v


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML.  The first sample is taken
    (and `ready` set) BEFORE the caller opens the timed window, then the thread sleeps one period,
    so no NVML call coincides with the window's start; samples taken between `mark_begin()` and
    `mark_end()` are the ones under the timed load."""

    def __init__(self, index: int, period=0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._halt = threading.Event()
        self.ready = threading.Event()
        self.t_begin = self.t_end = None
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            self.ready.set()
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._halt.is_set():
            try:
                t = time.perf_counter()
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((t, mhz))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self.ready.set()
            time.sleep(self.period)

    def mark_begin(self):
        self.t_begin = time.perf_counter()

    def mark_end(self):
        self.t_end = time.perf_counter()

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        inside = [m for t, m in self.samples
                  if self.t_begin is not None and self.t_end is not None and self.t_begin <= t <= self.t_end]
        allm = [m for _, m in self.samples]
        use = inside if inside else allm
        return {"sm_mhz": float(np.median(use)) if use else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(allm), "samples_in_window": len(inside),
                "note": "NVML every 2 ms from a side thread; the window is one graph launch, immune to the host"}


def _dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def bench_params(ll, args):
    return ll.default_params(judge_enabled=1, auto_reset=1, reset_perturb=1,
                             max_episode_steps=args.max_episode_steps, substeps=args.substeps,
                             rng_mode=1 if args.rng == "philox" else 0)


def initial_state_host(ll, P, B, rank, device):
    """BASELINE config 3: reset with keys (0, i) then perturb the lander pose on the host."""
    from gen_inputs import lunar_keys, perturbed_lunar_state

    keys = lunar_keys(B, offset=rank * B)
    st = ll.reset(keys, P, device=device).numpy()
    return perturbed_lunar_state(st, np.random.default_rng(1000 + rank), P)


# ---------------------------------------------------------------------------------------------
# the CPU arm: the C++ restatement of the reference (oracle/, kind "port"), -O3 -march=native build,
# OpenMP over envs with the thread count set EXPLICITLY (torchrun exports OMP_NUM_THREADS=1)
def _cpu_oracle(args):
    from oracle import oracle as o

    o.select_build("timing")
    threads = o.set_num_threads(args.ref_threads)
    return o, threads


def cpu_config3(args, o, warmup, steps):
    """Bounded sample of the bench workload (BASELINE config 3) on the host cores."""
    from gen_inputs import lunar_keys, perturbed_lunar_state

    P = o.default_lunar_params()
    P.judge_enabled, P.auto_reset, P.reset_perturb = 1, 1, 1
    P.max_episode_steps, P.substeps = args.max_episode_steps, args.substeps
    P.rng_mode = 1 if args.rng == "philox" else 0
    Bs = args.ref_envs
    st = perturbed_lunar_state(o.lunar_reset(lunar_keys(Bs), P), np.random.default_rng(1000), P)
    acts = np.random.default_rng(1).uniform(-1, 1, size=(8, 2, Bs)).astype(np.float32)
    for w in range(warmup):
        st, _, _ = o.lunar_step(st, acts[w % 8], P, dedupe=not args.ref_literal)
    t0 = time.perf_counter()
    for k in range(steps):
        st, _, _ = o.lunar_step(st, acts[k % 8], P, dedupe=not args.ref_literal)
    dt = time.perf_counter() - t0
    return Bs * steps * args.substeps / dt, dt, (
        f"{Bs} envs x {steps} steps of the bench workload per run "
        f"({'literal 306-evaluation collider' if args.ref_literal else 'memoised duplicate pairs'}, OpenMP over envs)")


def cpu_config1(o, B=1024, T=1000):
    """BASELINE config 1 as stated: B = 1024 envs x 1000 steps from reset(key), zero action, no judge
    (the loop of examples/test_viz.py:46-48) on the host cores."""
    from gen_inputs import lunar_keys

    P = o.default_lunar_params()
    st = o.lunar_reset(lunar_keys(B), P)
    t0 = time.perf_counter()
    for _ in range(T):
        st, _, _ = o.lunar_step(st, None, P)
    dt = time.perf_counter() - t0
    return B * T / dt, dt, float(st[1].mean())


def run_reference(args):
    """--impl reference: the reference's CPU path.  cotix itself needs jax + equinox (absent, no
    network), so this times the C++ restatement of it with all host threads, on a bounded sample of
    the same workload per step."""
    rank, world, _ = _dist_env()
    if rank != 0:
        return
    o, threads = _cpu_oracle(args)
    value, dt, sample = cpu_config3(args, o, args.warmup, args.steps)
    c1, c1_dt, _ = cpu_config1(o, T=args.config1_steps)
    line = {
        "impl": "reference", "metric": "LunarLander env-steps/sec", "value": value, "unit": "env-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "LunarLander 1M envs/GPU, auto-reset (BASELINE config 3); CPU arm runs a "
                               f"bounded sample of {args.ref_envs} envs per step", "envs_per_step": args.ref_envs,
                   "substeps": args.substeps, "rng": args.rng},
        "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": threads, "nproc": os.cpu_count(),
                         "kind": "port", "build": "g++ -O3 -march=native -fopenmp (oracle/Makefile timing)",
                         "sample": sample},
        "config1": {"workload": f"BASELINE config 1: 1024 envs x {args.config1_steps} steps from reset, zero action",
                    "value": c1, "unit": "env-steps/s", "seconds": c1_dt, "cores": threads},
        "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch

    import parallax_b200
    from parallax_b200 import lunar_lander as ll

    rank, world, local = _dist_env()
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist = None
        torch.cuda.set_device(0)
    dev = torch.device("cuda", local if world > 1 else 0)
    parallax_b200._lib.lib()  # fails loudly if the CUDA library is missing
    if args.split >= 0:
        ll.pipeline_config(split=args.split)

    B = args.envs
    P = bench_params(ll, args)
    state = ll.LunarState.from_numpy(initial_state_host(ll, P, B, rank, dev), device=dev)
    rng = np.random.default_rng(1 + 17 * rank)
    n_act = 4
    acts = torch.from_numpy(rng.uniform(-1, 1, size=(n_act, 2, B)).astype(np.float32)).to(dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident throughput ("value"): W warm-up steps, then K steps as ONE graph launch ---
    g_warm = ll.StepsGraph(state, acts, args.warmup, P)
    g_time = ll.StepsGraph(state, acts, args.steps, P)
    ginfo = g_time.info()
    g_warm.launch()
    sampler = ClockSampler(local if world > 1 else 0)
    sampler.start()
    sampler.ready.wait(timeout=2.0)  # first NVML query is over before the window opens
    barrier()
    n0 = parallax_b200.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.mark_begin()
    ev0.record()
    g_time.launch()
    ev1.record()
    barrier()
    sampler.mark_end()
    clocks = sampler.stop()
    launches = parallax_b200.launch_count() - n0
    ms = ev0.elapsed_time(ev1)
    reward_sum = g_time.reward.double().sum()
    if dist is not None:
        dist.all_reduce(reward_sum, op=dist.ReduceOp.SUM)  # the only collective: episode-return stats
    ms_max = allmax(ms)
    ms_ranks = [ms]
    if dist is not None:  # per-rank step time, for diagnosis of the max
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        ms_ranks = [float(x.item()) for x in g]
    total_env_steps = world * B * args.steps * args.substeps
    value = total_env_steps / (ms_max * 1e-3)
    g_warm.close()
    g_time.close()
    if args.value_only:  # the fast-math leg of the main run (child process, COTIX_FLAVOUR=fast)
        if rank == 0:
            print(json.dumps({"value": value, "ms_per_step": ms_max / args.steps, "unit": "env-steps/s",
                              "flavour": parallax_b200._lib.FLAVOUR,
                              "math_flavour": int(parallax_b200._lib.lib().cotix_math_flavour())}), flush=True)
        return

    # ---- end-to-end through the host-facing env handle ("e2e") ---------------------------
    env = ll.LunarLanderEnv(B, P, device=dev.index)
    env.set_state(state.numpy())
    host_acts = [torch.from_numpy(rng.uniform(-1, 1, size=(2, B)).astype(np.float32)).pin_memory()
                 for _ in range(n_act)]
    e2e_steps = max(3, min(args.steps, args.e2e_steps))

    def e2e_run(want_obs):
        for w in range(10):  # the first host steps touch the pinned buffers and build the range streams
            env.step(host_acts[w % n_act], want_obs=want_obs)
        barrier()
        c0 = parallax_b200.launch_count()
        t0 = time.perf_counter()
        for k in range(e2e_steps):
            env.step(host_acts[k % n_act], want_obs=want_obs)
        barrier()
        secs = allmax(time.perf_counter() - t0)
        return world * B * e2e_steps * args.substeps / secs, parallax_b200.launch_count() - c0

    e2e_value, e2e_launches = e2e_run(True)
    e2e_lean, e2e_lean_launches = e2e_run(False)
    # the ceiling of a synchronous host step: the D2H copy of one step's results alone (same bytes,
    # same pinned buffers, nothing else running) -- what the PCIe link / host memory allow
    barrier()
    dsrc = torch.empty((8 * 4 + 4 + 1) * B, dtype=torch.uint8, device=dev)
    hdst = torch.empty((8 * 4 + 4 + 1) * B, dtype=torch.uint8).pin_memory()
    hdst.copy_(dsrc, non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(10):
        hdst.copy_(dsrc, non_blocking=True)
    torch.cuda.synchronize(dev)
    d2h_ms = allmax(time.perf_counter() - t0) * 100.0
    del dsrc, hdst
    env.close()

    peak, peak_src, sm_max_mhz = _peaks()
    step_ms = ms / args.steps  # the timed region holds only the step's kernels (one graph)
    step_achieved = ALGO_BYTES_PER_ENV_STEP * B * args.substeps / (step_ms * 1e-3) / 1e9
    # per-stage durations: CUDA events between the stages on the launching stream, with the batch as
    # ONE range and no programmatic dependent launch (serial stage times, for diagnosis: in the timed
    # graph the stages of different env ranges overlap, so their sum exceeds the step time)
    pipeline = args.substeps == 1 and B >= ll.PIPELINE_MIN_ENVS
    stages, counters = None, None
    ncu = _ncu()
    if pipeline:
        ll.pipeline_profile(True)
        for k in range(min(args.steps, 50)):
            ll.step(state, acts[k % n_act], P, inplace=True)
        torch.cuda.synchronize(dev)
        stage_ms, _ = ll.pipeline_stage_ms()
        counters = ll.pipeline_counters(dev, B)
        ll.pipeline_profile(False)
        stages = {}
        for k, v in stage_ms.items():
            if k == "step" or v <= 0:
                continue
            ab = stage_algo_bytes(k, B, counters)
            prof = ncu.get(f"pipe_{k}_kernel")
            stages[f"pipe_{k}_kernel"] = {
                "ms_serial": v, "algorithmic_bytes": ab,
                "achieved": (ab / (v * 1e-3) / 1e9) if ab else None,
                "frac": (ab / (v * 1e-3) / 1e9 / peak) if ab else None,
                "ncu": prof if isinstance(prof, dict) else ({"dram_bytes": prof} if prof else None)}
        stages["serial_step_ms"] = stage_ms.get("step")
    whole = ncu.get("whole_step") if isinstance(ncu.get("whole_step"), dict) else {}
    line = {
        "metric": "LunarLander env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
        "ms_per_step_ranks": [m / args.steps for m in ms_ranks],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "LunarLander 1M envs/GPU, auto-reset (BASELINE config 3)",
                   "envs_per_gpu": B, "global_batch": world * B, "substeps": args.substeps, "rng": args.rng,
                   "max_episode_steps": args.max_episode_steps, "parallelism": f"env-shard x{world}",
                   "launch": (f"one CUDA graph of {args.steps} steps ({ginfo['nodes']} nodes, "
                              f"{ginfo['kernel_launches']} kernels): wavefront pipeline, env ranges on their own "
                              "streams" if pipeline else f"one CUDA graph of {args.steps} fused-kernel steps"),
                   "l2": f"state {36 * 4 * B / 1e6:.0f} MB + actions > L2 {L2_BYTES / 1e6:.0f} MB, no flush needed"},
        # headline roofline = the WHOLE step against SURVEY §8(d)'s 224 algorithmic bytes per env-step
        "roofline": {"bound": "hbm", "achieved": step_achieved, "peak": peak, "unit": "GB/s",
                     "frac": step_achieved / peak, "traffic": whole.get("dram_bytes"),
                     "kernel": "whole step (pipe_integrate + pipe_narrow + pipe_contact + pipe_finish + pipe_reset)",
                     "peak_source": peak_src, "algorithmic_bytes_per_env_step": ALGO_BYTES_PER_ENV_STEP,
                     "ms": step_ms,
                     # the step is issue-bound, not HBM-bound: issue-slot and FP32-pipe utilisation from the
                     # committed ncu --set full capture (static evidence; see profiles/)
                     "issue_slot": whole.get("issue_slot_pct"), "fp32_pipe": whole.get("fp32_pipe_pct"),
                     "alu_pipe": whole.get("alu_pipe_pct"), "occupancy": whole.get("occupancy_pct"),
                     "warp_instructions_per_step": whole.get("warp_instr"),
                     # what the step would take at 100 % issue-slot use: the bound that matters here
                     "issue_bound_ms": (whole["warp_instr"] / (148 * 4 * sm_max_mhz * 1e3)) if whole.get("warp_instr") else None,
                     "issue_frac": (whole["warp_instr"] / (148 * 4 * sm_max_mhz * 1e3) / step_ms) if whole.get("warp_instr") else None,
                     "ncu_source": ncu.get("_source"),
                     "stages": stages, "queue_counts": counters},
        "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": 2 * 4 * B,
                "d2h_bytes_per_step": (8 * 4 + 4 + 1) * B, "steps": e2e_steps,
                "overlap": "4 env ranges on their own streams, H2D(action) -> kernels -> D2H(obs, reward, done); compute range after range, so D2H of range r runs under the kernels of r+1..",
                "d2h_only_ms": d2h_ms, "d2h_gbs": (8 * 4 + 4 + 1) * B / (d2h_ms * 1e-3) / 1e9,
                "copy_bound_value": world * B / (d2h_ms * 1e-3),
                "note": "copy_bound_value = env-steps/s if a step cost only its D2H copy (measured, all ranks at once)",
                "reward_done_only": {"value": e2e_lean, "unit": "env-steps/s", "h2d_bytes_per_step": 2 * 4 * B,
                                     "d2h_bytes_per_step": (4 + 1) * B,
                                     "note": "obs left on the device (want_obs=False): 86 % of the D2H bytes"}},
        "gpu_launches": int(launches + e2e_launches + e2e_lean_launches),
        "clocks": clocks,
        "mean_last_reward": float(reward_sum.item()) / (world * B),
    }

    # BASELINE config 4 (RoboCup, batch 256k) runs at every N: the batch is sharded over the ranks
    line["robocup"] = robocup(args, dev, dist, rank, world, peak)
    if rank == 0 and world == 1 and not args.no_sweep:
        line["config1"] = config1(args, dev)
        line["narrowphase"] = sweep(args, dev, peak)
        line["differentiable"] = differentiable(args, dev)
        line["render"] = render(args, dev, peak)
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args)
    line["math_flavour"] = "exact (-fmad=false, IEEE division / sqrt: bit-identical to the oracle)" \
        if parallax_b200._lib.lib().cotix_math_flavour() == 0 else "fast"
    if rank == 0 and world == 1 and not args.no_sweep and parallax_b200._lib.FLAVOUR == "exact":
        line["fast_math"] = fast_math_leg(args)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def fast_math_leg(args):
    """The same device-resident measurement with libcotix_b200_fast.so (-fmad=true -prec-div=false
    -prec-sqrt=false; within 1e-5 relative per step of the exact build, tests/test_fast_flavour.py) in a
    child process.  Reported beside the headline, never as it: the exact flavour is the parity build."""
    import subprocess

    env = dict(os.environ, COTIX_FLAVOUR="fast")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    cmd = [sys.executable, os.path.abspath(__file__), "--value-only", "--steps", str(args.steps), "--warmup",
           str(args.warmup), "--envs", str(args.envs), "--rng", args.rng, "--max-episode-steps",
           str(args.max_episode_steps), "--substeps", str(args.substeps)]
    try:
        res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
        out = json.loads([l for l in res.stdout.splitlines() if l.startswith("{")][-1])
        out["build"] = "-fmad=true -prec-div=false -prec-sqrt=false (COTIX_FLAVOUR=fast); not bit-exact"
        return out
    except Exception as e:  # the leg is informative: never take the main line down
        return {"error": f"{type(e).__name__}: {e}"}


def config1(args, dev):
    """BASELINE config 1 as stated (B = 1024 envs x 1000 steps from reset(key), zero action, no judge:
    the loop of examples/test_viz.py:46-48) on one B200: `steps(state, None, 1000)` = one launch of the
    block-resident rollout kernel (a graph of 1000 fused-kernel launches took 11.9 us per step)."""
    import torch

    from gen_inputs import lunar_keys
    from parallax_b200 import lunar_lander as ll

    B, T = 1024, args.config1_steps
    P = ll.default_params()
    st = ll.reset(lunar_keys(B), P, device=dev)
    g = ll.StepsGraph(st, None, T, P)
    g.launch()  # warm-up rollout
    torch.cuda.synchronize(dev)
    st.data.copy_(ll.reset(lunar_keys(B), P, device=dev).data)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    ev0.record()
    g.launch()
    ev1.record()
    torch.cuda.synchronize(dev)
    ms = ev0.elapsed_time(ev1)
    info = g.info()
    g.close()
    return {"workload": f"BASELINE config 1: 1024 envs x {T} steps from reset, zero action", "value": B * T / (ms * 1e-3),
            "unit": "env-steps/s", "ms": ms, "us_per_step": 1e3 * ms / T, "graph_nodes": info["nodes"],
            "kernel_launches": info["kernel_launches"],
            "note": "latency-bound: 128 blocks of 8 envs, every env resident in its block for the 1000 steps; the number "
                    "to compare with the CPU arm's config1"}


def sweep(args, dev, peak):
    """BASELINE config 2: GJK/EPA narrowphase sweep over N random pairs on one B200."""
    import torch

    from gen_inputs import random_pairs
    from parallax_b200 import shapes as sh

    n = args.pairs
    rng = np.random.default_rng(2)
    chunk = 1 << 21
    a = torch.empty((n, 16), dtype=torch.float32, device=dev)
    b = torch.empty((n, 16), dtype=torch.float32, device=dev)
    for s in range(0, n, chunk):
        ha, hb = random_pairs(rng, min(chunk, n - s))
        a[s:s + ha.shape[0]] = torch.from_numpy(ha).to(dev)
        b[s:s + hb.shape[0]] = torch.from_numpy(hb).to(dev)
    A, Bt = sh.ShapeBatch(a), sh.ShapeBatch(b)
    for _ in range(3):
        ci = sh.narrowphase(A, Bt)
    torch.cuda.synchronize(dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    ev0.record()
    for _ in range(reps):
        ci = sh.narrowphase(A, Bt)
    ev1.record()
    torch.cuda.synchronize(dev)
    ms = ev0.elapsed_time(ev1) / reps
    achieved = ALGO_BYTES_PER_PAIR * n / (ms * 1e-3) / 1e9
    return {"metric": "GJK+EPA pair-tests/sec", "value": n / (ms * 1e-3), "unit": "pairs/s", "pairs": n,
            "ms_per_sweep": ms, "hit_fraction": float(ci.exists.float().mean().item()),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": _traffic("narrowphase_sweep"),
                         "algorithmic_bytes_per_pair": ALGO_BYTES_PER_PAIR,
                         "kernel": "whole sweep: sweep_classify + sweep_gjk<3> + sweep_epa<0,2> + sweep_epa_cp_short/long"}}


def robocup(args, dev, dist, rank, world, peak):
    """BASELINE config 4: the RoboCup scene with 6 + 6 robots, global batch 262144 sharded over the
    ranks, through the generic scene step (cotix_scene_step); at N = 1 the oracle's generic collider
    on a small sample beside it."""
    import torch

    from parallax_b200 import scene as sc

    Bg = args.robocup_envs
    B = (Bg + world - 1) // world
    lo = rank * B
    B = max(0, min(B, Bg - lo))
    spec = sc.robocup_scene(n_robots=12, areas_collide=False)
    scene = sc.Scene(spec, device=dev.index)
    rng = np.random.default_rng(3 + rank)
    st = np.repeat(spec.init_state[:, :, None], B, axis=2).astype(np.float32)
    for b in np.nonzero(np.isfinite(spec.body_props[:, 0]))[0]:
        st[b, 0:2] += rng.normal(0, 0.3, (2, B)).astype(np.float32)
        st[b, 2:4] += rng.normal(0, 1.0, (2, B)).astype(np.float32)
    state = torch.from_numpy(np.ascontiguousarray(st.reshape(spec.nb * 6, B))).to(dev)
    keys_h = np.stack([np.zeros(B, np.uint32), np.arange(lo, lo + B, dtype=np.uint32)])
    keys = torch.from_numpy(keys_h.view(np.int32)).to(dev)
    for _ in range(5):
        scene.step(state, keys)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    ev0.record()
    for _ in range(reps):
        scene.step(state, keys)
    ev1.record()
    torch.cuda.synchronize(dev)
    ms = ev0.elapsed_time(ev1) / reps
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    info = scene.info()
    scene.close()
    # algorithmic bytes per env-step: the dynamic state of the nb bodies read and written once
    # (6 floats each) + the key (2 words) read and written
    algo = 2.0 * (spec.nb * 6 * 4 + 8)
    out = {"metric": "RoboCup env-steps/sec", "value": Bg / (ms * 1e-3), "unit": "env-steps/s", "envs": Bg,
           "envs_per_gpu": B, "n_gpus": world, "scaling": "strong (global batch sharded)",
           "ms_per_step": ms, "bodies": spec.nb, "parts": int(spec.part_recs.shape[0]),
           "distinct_pairs_per_step": info["distinct_pairs"],
           "reference_evaluations_per_step": info["reference_evaluations"],
           "kernel": "scene_step_kernel (one env per thread)",
           "roofline": {"bound": "hbm", "algorithmic_bytes_per_env_step": algo,
                        "achieved": algo * Bg / (ms * 1e-3) / 1e9 / world, "peak": peak, "unit": "GB/s",
                        "frac": algo * Bg / (ms * 1e-3) / 1e9 / world / peak,
                        "traffic": _traffic("scene_step_kernel"),
                        "issue_slot": (_ncu().get("scene_step_kernel") or {}).get("issue_slot_pct"),
                        "alu_pipe": (_ncu().get("scene_step_kernel") or {}).get("alu_pipe_pct"),
                        "fp32_pipe": (_ncu().get("scene_step_kernel") or {}).get("fp32_pipe_pct"),
                        "ncu_source": (_ncu().get("scene_step_kernel") or {}).get("source"),
                        "note": "per GPU; bound by the integer ALU pipe (65 % busy): half of the 4.5e8 warp instructions of "
                                "a step are threefry-2x32 rounds of the collider's coin flips (jax.random parity)"}}
    if rank == 0 and world == 1 and not args.no_cpu:
        # CPU: the oracle's generic collider (memoised duplicate pairs), one core, 64 envs x 2 steps
        from oracle import oracle as o

        o.select_build("timing")
        o.set_trig_mode(1)
        n_cpu = 64
        t0 = time.perf_counter()
        for e in range(n_cpu):
            b10 = spec.bodies10(st[:, :, e])
            for _ in range(2):
                b10, _c, _ch = o.scene_step(b10, spec.part_body, spec.part_recs, keys_h[:, e])
        cpu = 2 * n_cpu / (time.perf_counter() - t0)
        o.set_trig_mode(0)
        out["cpu_port"] = {"value": cpu, "unit": "env-steps/s", "cores": 1, "sample": f"{n_cpu} envs x 2 steps"}
    return out


def differentiable(args, dev):
    """BASELINE config 5: B = 4096 envs x 100 steps, loss = final lander y^2, seed 4: forward rollout
    that keeps the tape (cotix_lunar_rollout) + the hand-written adjoint (cotix_lunar_rollout_bwd),
    captured together as ONE CUDA graph and replayed.  Beside it: what the same gradient costs by
    central finite differences of the oracle port on the host cores (2 x (18 + 200) rollouts per env,
    measured on a sample of rollouts) -- the reference's jax.grad cannot run here (no jax)."""
    import torch

    from gen_inputs import lunar_keys, perturbed_lunar_state
    from parallax_b200 import launch_count
    from parallax_b200 import lunar_lander as ll

    B, T = args.diff_envs, 100
    P = ll.default_params()
    rng = np.random.default_rng(4)
    s0 = perturbed_lunar_state(ll.reset(lunar_keys(B), P, device=dev).numpy(), rng, P)
    acts_h = rng.uniform(-1, 1, size=(T, 2, B)).astype(np.float32)
    st0 = ll.LunarState.from_numpy(s0, device=dev)
    acts = torch.from_numpy(acts_h).to(dev)
    flags = torch.zeros(B, dtype=torch.int32, device=dev)
    tape = torch.empty((T + 1, ll.NF, B), dtype=torch.float32, device=dev)
    g = torch.zeros((18, B), device=dev)
    g_act = torch.zeros((T, 2, B), device=dev)
    ws = ll.adjoint_workspace(B, T, dev)
    # small batch: the block-resident rollout kernel (one launch for the T steps), else the pipeline
    mode = "pipeline" if B > ll.ROLLOUT_MAX_ENVS else "fused"
    if mode == "pipeline":
        ll._workspace(dev, B)

    def forward():
        return ll.rollout(st0, acts, T, P, mode=mode, tape=tape)

    def backward(ro):
        g.zero_()
        torch.mul(tape[-1, 1], 2.0, out=g[1])  # d(y_T^2)/dy_T, no allocation
        return ll.rollout_vjp(ro, g, flags, workspace=ws, out=(g, g_act))

    # eager once (warm-up; also the un-graphed launch count), then one graph per half
    n0 = launch_count()
    ro = forward()
    backward(ro)
    torch.cuda.synchronize(dev)
    eager_launches = launch_count() - n0
    gf, gb = ll.Graph(dev), ll.Graph(dev)
    with gf:
        ro = forward()
    with gb:
        backward(ro)
    for gr in (gf, gb):
        gr.keep(tape, acts, g, g_act, ws, flags, st0.data)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    fwd, bwd = [], []
    reps = 5
    for it in range(2 + reps):
        flags.zero_()
        ev[0].record()
        gf.launch()
        ev[1].record()
        gb.launch()
        ev[2].record()
        torch.cuda.synchronize(dev)
        if it >= 2:
            fwd.append(ev[0].elapsed_time(ev[1]))
            bwd.append(ev[1].elapsed_time(ev[2]))
    f = flags.cpu().numpy()
    fwd_ms, bwd_ms = float(np.median(fwd)), float(np.median(bwd))
    out = {"metric": "differentiable rollouts/sec (100 steps forward + adjoint)", "unit": "rollouts/s",
           "value": B / ((fwd_ms + bwd_ms) * 1e-3), "envs": B, "steps": T, "loss": "final lander y^2",
           "forward_ms": fwd_ms, "backward_ms": bwd_ms, "kernels_per_rollout": int(eager_launches),
           "graph_launches_per_rollout": 2, "graph_nodes": [gf.info()["nodes"], gb.info()["nodes"]],
           "forward_mode": mode,
           "kernel": "forward: lunar_rollout_kernel, T steps in one launch with the envs resident in their block (the "
                     "wavefront pipeline above %d envs); backward: lunar_decide + lunar_jacobian on a side stream, "
                     "lunar_sweep_bwd_kernel per run of steps (recompute-and-reverse)" % ll.ROLLOUT_MAX_ENVS,
           "tape_bytes": int(tape.numel() * 4), "replay_mismatch_envs": int((f & 1).sum()),
           "nonfinite_envs": int(((f >> 1) & 1).sum()), "envs_with_contact": int((((f >> 8) & 7) > 0).sum()),
           "grad_finite": bool(torch.isfinite(g).all() and torch.isfinite(g_act).all())}
    gf.close()
    gb.close()
    if not args.no_cpu:
        # CPU: one forward rollout of n_cpu envs through the oracle port; a finite-difference gradient
        # of the 18 + 2 T inputs needs 2 (18 + 2 T) of them per env
        from oracle import oracle as o

        o.select_build("timing")
        o.set_num_threads(args.ref_threads)
        o.set_trig_mode(1)
        n_cpu = 256
        t0 = time.perf_counter()
        s = s0[:, :n_cpu].copy()
        for t in range(T):
            s = o.lunar_step(s, np.ascontiguousarray(acts_h[t][:, :n_cpu]), P)[0]
        cpu_rollouts = n_cpu / (time.perf_counter() - t0)
        o.set_trig_mode(0)
        out["cpu_port"] = {"value": cpu_rollouts, "unit": "forward rollouts/s", "cores": o.num_threads(),
                           "sample": f"{n_cpu} envs x {T} steps forward only",
                           "fd_gradient_rollouts_per_env": 2 * (18 + 2 * T)}
    return out


def render(args, dev, peak):
    """The headless rasteriser (K6, the stand-in for cotix/_viz.py's pygame window): env.draw of
    n_img LunarLander envs into 800 x 600 RGB8 frames in HBM.  Write-bound: algorithmic bytes =
    1.44 MB per frame (every pixel written once) + 144 B of state read."""
    import ctypes as C

    import torch

    from oracle import oracle as o
    from parallax_b200 import _lib
    from parallax_b200 import lunar_lander as ll
    from parallax_b200 import viz

    n = args.render_frames
    P = ll.default_params()
    st = ll.reset(lunar_keys_np(n), P, device=dev)
    for _ in range(3):
        viz.render_lunar(st, params=P)
    torch.cuda.synchronize(dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    idx = torch.arange(n, device=dev)
    img = torch.empty((n, 600, 800, 3), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    ev0.record()
    for _ in range(reps):
        _lib.check(_lib.lib().cotix_lunar_draw(stream, n, st.data.data_ptr(), n, idx.data_ptr(), img.data_ptr(), C.byref(P)))
    ev1.record()
    torch.cuda.synchronize(dev)
    ms = ev0.elapsed_time(ev1) / reps
    algo = n * (800 * 600 * 3 + 144)
    # CPU: the oracle's sequential painter on the same draw calls, a sample of frames, one core
    prims = viz.lunar_primitives(st, idx[:64], P).cpu().numpy()
    t0 = time.perf_counter()
    ref = o.raster_draw(prims)
    cpu = 64 / (time.perf_counter() - t0)
    same = bool(np.array_equal(img[:64].cpu().numpy(), ref))
    return {"metric": "frames/sec (800x600 RGB8, LunarLander env.draw)", "unit": "frames/s", "value": n / (ms * 1e-3),
            "frames": n, "ms": ms, "kernel": "lunar_draw_kernel (band in shared memory, TMA bulk store)",
            "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": algo / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_frame": 800 * 600 * 3 + 144,
                         "l2": f"{algo / 1e6:.0f} MB written per launch > L2"},
            "matches_oracle": same,
            "cpu_port": {"value": cpu, "unit": "frames/s", "cores": 1, "sample": "64 frames"}}


def lunar_keys_np(n):
    return np.stack([np.zeros(n, np.uint32), np.arange(n, dtype=np.uint32)], 1)


def cpu_baseline(args):
    """The oracle (a C++ port of the reference path) on the box's host cores, bounded sample."""
    o, threads = _cpu_oracle(args)
    value, _dt, sample = cpu_config3(args, o, 2, 40)
    c1, c1_dt, _ = cpu_config1(o, T=args.config1_steps)
    o.select_build("parity")
    return {"value": value, "unit": "env-steps/s", "cores": threads, "nproc": os.cpu_count(), "kind": "port",
            "build": "g++ -O3 -march=native -fopenmp (oracle/Makefile timing)", "sample": sample,
            "config1": {"value": c1, "unit": "env-steps/s", "seconds": c1_dt,
                        "workload": f"1024 envs x {args.config1_steps} steps from reset, zero action"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=1 << 20, help="envs per GPU")
    ap.add_argument("--substeps", type=int, default=1)
    ap.add_argument("--rng", default="threefry", choices=["threefry", "philox"])
    ap.add_argument("--max-episode-steps", dest="max_episode_steps", type=int, default=1000)
    ap.add_argument("--pairs", type=int, default=1 << 24)
    ap.add_argument("--robocup-envs", dest="robocup_envs", type=int, default=1 << 18)
    ap.add_argument("--render-frames", dest="render_frames", type=int, default=2048)
    ap.add_argument("--diff-envs", dest="diff_envs", type=int, default=4096)
    ap.add_argument("--e2e-steps", dest="e2e_steps", type=int, default=50)
    ap.add_argument("--config1-steps", dest="config1_steps", type=int, default=1000)
    ap.add_argument("--ref-envs", dest="ref_envs", type=int, default=16384)
    ap.add_argument("--ref-threads", dest="ref_threads", type=int, default=0, help="0 = every logical core")
    ap.add_argument("--ref-literal", dest="ref_literal", action="store_true")
    ap.add_argument("--split", type=int, default=-1, help="env ranges of the pipeline (-1 = library default)")
    ap.add_argument("--value-only", dest="value_only", action="store_true", help="device-resident leg only")
    ap.add_argument("--no-sweep", dest="no_sweep", action="store_true")
    ap.add_argument("--no-cpu", dest="no_cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
