#!/usr/bin/env python
"""profiles/roofline_traffic.json from the text summary of an `ncu --set full` capture
(profiles/summarize.py output): per kernel DRAM bytes per launch, issue-slot / FP32 (FMA) pipe / ALU
pipe utilisation, occupancy, warp instructions; `whole_step` = the pipeline's five kernels together
(time-weighted utilisations, summed traffic and instructions).  bench.py reads this file as STATIC
evidence beside its live timings.   usage: make_roofline_json.py profiles/r2f_step_ncu.txt"""
import json
import os
import re
import sys

src = sys.argv[1]
KEYS = {"gpu__time_duration.sum": "duration_us", "dram__bytes_read.sum": "rd", "dram__bytes_write.sum": "wr",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slot_pct",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "fp32_pipe_pct",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "occupancy_pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio": "lanes_per_instr",
        "smsp__inst_executed.sum": "warp_instr", "launch__registers_per_thread": "registers"}
UNIT = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
kern, cur = {}, None
for line in open(src):
    m = re.match(r"== (?:void )?(?:<unnamed>::)?(\w+)(<[\d, ]+>)?", line)
    if m:
        name = m.group(1) + (m.group(2) or "").replace(" ", "")
        cur = kern.setdefault(name, {})
        # bench.py looks the stages up by their plain names: the common instantiations (INLINE = 0, any SPREAD)
        if name.startswith("pipe_contact_kernel<0"):
            kern["pipe_contact_kernel"] = cur
        if name.startswith("pipe_narrow_kernel<"):
            kern["pipe_narrow_kernel"] = cur
        continue
    p = line.split()
    if cur is not None and len(p) >= 2 and p[0] in KEYS and KEYS[p[0]] not in cur:
        v = float(p[1].replace(",", ""))
        if KEYS[p[0]] in ("rd", "wr"):
            v *= UNIT.get(p[2] if len(p) > 2 else "byte", 1.0)
        cur[KEYS[p[0]]] = v
out_path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "roofline_traffic.json")
old = json.load(open(out_path)) if os.path.exists(out_path) else {}
out = {k: v for k, v in old.items() if not k.startswith("pipe_") and k not in ("whole_step", "_source", "_note")}
for k, v in kern.items():
    if "dram_bytes" not in v:
        v["dram_bytes"] = int(v.pop("rd", 0) + v.pop("wr", 0))
    out[k] = v
pipe = [v for k, v in kern.items() if k.startswith("pipe_") and k not in ("pipe_contact_kernel", "pipe_narrow_kernel")]
if pipe:
    T = sum(v["duration_us"] for v in pipe)
    w = lambda f: sum(v[f] * v["duration_us"] for v in pipe) / T
    instr = sum(v["warp_instr"] for v in pipe)
    out["whole_step"] = {"dram_bytes": sum(v["dram_bytes"] for v in pipe), "warp_instr": int(instr),
                         "duration_us_serialised": T, "issue_slot_pct": w("issue_slot_pct"),
                         "fp32_pipe_pct": w("fp32_pipe_pct"), "alu_pipe_pct": w("alu_pipe_pct"),
                         "occupancy_pct": w("occupancy_pct"),
                         "issue_bound_us_at_1965MHz": instr / (148 * 4 * 1965.0)}
out["_source"] = f"{src} (ncu --set full --clock-control none, one steady-state step of the bench workload, 1,048,576 envs)"
out["_note"] = ("per launch; utilisations are pct_of_peak_sustained_active; whole_step = time-weighted over the five "
                "pipeline kernels of one step (cold-cache, serialised under ncu: shares, not absolutes); "
                "narrowphase_sweep / lunar_step_kernel / scene entries: earlier captures, DRAM bytes only")
json.dump(out, open(out_path, "w"), indent=1)
print(json.dumps(out["whole_step"], indent=1))
