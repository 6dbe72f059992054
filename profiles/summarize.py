#!/usr/bin/env python
"""Summarise an Nsight Compute report (read here, on the CPU box):
   python profiles/summarize.py gpurun_out/prof.ncu-rep [kernel-regex] > profiles/<name>.txt
Prints per-launch headline metrics (duration, DRAM bytes, issue/pipe utilisation, divergence,
stall reasons) and the hottest source lines of the first matching launch."""
import csv
import io
import subprocess
import sys

RAW = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
]


def ncu(args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    pat = sys.argv[2] if len(sys.argv) > 2 else None
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    print(f"# {rep}")
    seen = set()
    for r in rows[2:]:
        name = r[ki].split("(")[0]
        if pat and pat not in name:
            continue
        print(f"\n== {name}  (launch id {r[0]})")
        for m in RAW:
            if m in hdr:
                i = hdr.index(m)
                print(f"  {m:86s} {r[i]:>16s} {units[i]}")
        if name in seen:
            continue
        seen.add(name)
        src = ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                   "--kernel-name", "regex:" + name.split("::")[-1], "--launch-count", "1"])
        agg, fn = {}, None
        for s in csv.reader(io.StringIO(src)):
            if s and s[0] == "File Path":
                fn = s[1].split("/")[-1]
            elif s and s[0] == "Line No":
                h = s
                si, ii = h.index("# Samples"), h.index("Instructions Executed")
            elif len(s) > 10 and s[0].isdigit() and s[2] == "-":
                try:
                    agg[(fn, int(s[0]))] = (int(s[si]), int(s[ii]), s[1].strip()[:70])
                except ValueError:
                    pass
        tot = sum(a[0] for a in agg.values()) or 1
        toti = sum(a[1] for a in agg.values()) or 1
        byf = {}
        for (f, _), a in agg.items():
            b = byf.setdefault(f, [0, 0])
            b[0] += a[0]
            b[1] += a[1]
        print("  -- samples / warp-instructions by file")
        for f, b in sorted(byf.items(), key=lambda kv: -kv[1][0]):
            print(f"     {f:24s} {100 * b[0] / tot:5.1f}% samples {100 * b[1] / toti:5.1f}% instr")
        print("  -- hottest source lines")
        for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:25]:
            print(f"     {100 * a[0] / tot:5.1f}% smp {100 * a[1] / toti:5.1f}% ins  {f}:{l}  {a[2]}")


if __name__ == "__main__":
    main()
