#!/usr/bin/env python
"""Condense `ncu --page raw --csv` / `--page source --csv` exports (tools/ncu_capture.sh) into the text summaries
tracked under profiles/.  usage: python tools/ncu_summarize.py <case> [<case> ...]"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
SRC = os.path.join(ROOT, "gpurun_out")

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__cycles_elapsed.avg", "smsp__cycles_active.avg", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_uniform.sum", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum",
    "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_wait_per_warp_active.pct",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__sass_inst_executed_op_shared_ld.sum",
    "sm__sass_inst_executed_op_shared_st.sum", "sm__sass_inst_executed_op_global_ld.sum",
    "sm__sass_inst_executed_op_global_st.sum",
]


def load_raw(case):
    rows = list(csv.reader(open(os.path.join(SRC, f"ncu_{case}_raw.csv"))))
    hdr, units, vals = rows[0], rows[1], rows[2]
    return {h: (u, v) for h, u, v in zip(hdr, units, vals)}


def top_source(case, n=14):
    path = os.path.join(SRC, f"ncu_{case}_source.csv")
    if not os.path.exists(path):
        return []
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    i_src, i_s, i_ex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    body = []
    for r in rows[2:]:
        if len(r) <= i_s or not r[i_s].isdigit():
            continue
        stalls = sorted(((int(r[i]) if r[i].isdigit() else 0, h) for i, h in stall_cols), reverse=True)[:2]
        body.append((int(r[i_s]), r[i_src].strip(), r[i_ex], ", ".join(f"{h[6:]}={c}" for c, h in stalls if c)))
    total = sum(b[0] for b in body) or 1
    body.sort(reverse=True)
    return [(s, 100.0 * s / total, src, ex, st) for s, src, ex, st in body[:n]], total


def main(cases):
    os.makedirs(OUT, exist_ok=True)
    for case in cases:
        d = load_raw(case)
        lines = [f"# ncu --set full --clock-control none --import-source on : case '{case}' (tools/ncu_case.py {case})",
                 f"kernel : {d['Kernel Name'][1]}",
                 f"grid {d.get('Grid Size', ('', '?'))[1]} block {d.get('Block Size', ('', '?'))[1]}  device {d.get('Device', ('', '?'))[1]} cc {d.get('CC', ('', '?'))[1]}", ""]
        for k in KEYS:
            if k in d:
                lines.append(f"{k:92s} {d[k][1]:>18s} {d[k][0]}")
        ts = top_source(case)
        if ts:
            rows, total = ts
            lines += ["", f"## hottest SASS instructions by warp-stall samples (total samples {total})",
                      f"{'samples':>8s} {'%':>6s}  {'executed':>9s}  instruction   [top stall reasons]"]
            for s, pct, src, ex, st in rows:
                lines.append(f"{s:8d} {pct:6.2f}  {ex:>9s}  {src[:90]}   [{st}]")
        with open(os.path.join(OUT, f"ncu_{case}.txt"), "w") as f:
            f.write("\n".join(lines) + "\n")
        print("wrote", f"profiles/ncu_{case}.txt")


if __name__ == "__main__":
    main(sys.argv[1:])
