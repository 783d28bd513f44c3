#!/usr/bin/env python
"""Condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the small JSON summaries
kept under profiles/: one object per captured launch with the metrics DESIGN.md argues from.

    python tools/ncu_summary.py gpurun_out/tc_r1g.ncu-rep profiles/r1c_scan_tc_ncu_full.json
"""
import csv
import io
import json
import subprocess
import sys

KEEP = [
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_tc_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_uniform.sum",
    "smsp__inst_executed.sum",
    "smsp__cycles_active.avg",
    "sm__cycles_elapsed.max",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True,
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    res = []
    for r in body:
        o = {"Kernel Name": r[col["Kernel Name"]]}
        for m in KEEP:
            if m in col and r[col[m]] != "":
                u = units[col[m]]
                o[m] = r[col[m]] + (" " + u if u else "")
        # anything that names the tensor pipe, whatever this ncu version calls it
        for h in head:
            if "pipe_tensor" in h and "Triage" not in h and "pct" in h and h not in o and \
                    r[col[h]] not in ("", "0"):
                u = units[col[h]]
                o[h] = r[col[h]] + (" " + u if u else "")
        res.append(o)
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print("wrote %d launches to %s" % (len(res), out))


if __name__ == "__main__":
    main()
