#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`) into a small JSON + markdown table.

    python tools/ncu_summary.py gpurun_out/r1_vadv_fused2.ncu-rep profiles/r1_vadv_fused2

Writes <out>.json (selected raw metrics of the first captured launch) and <out>_hot.txt (the 40
SASS instructions with the most stall samples, with their dominant stall reasons).
"""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.per_cycle_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]


def ncu(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, out = sys.argv[1], sys.argv[2]
    rows = ncu(rep, "raw")
    hdr, units, vals = rows[0], rows[1], rows[2]
    summary = {}
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS or h == "Kernel Name":
            summary[h] = {"value": v, "unit": u}
    with open(out + ".json", "w") as fh:
        json.dump(summary, fh, indent=1)
    src = ncu(rep, "source")
    if len(src) > 2:
        kname, hdr, data = src[0][1] if len(src[0]) > 1 else "", src[1], src[2:]
        ix = {h: i for i, h in enumerate(hdr)}
        stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[ix["# Samples"]]) for r in data)
        top = sorted(range(len(data)), key=lambda n: -int(data[n][ix["# Samples"]]))[:40]
        with open(out + "_hot.txt", "w") as fh:
            fh.write(f"# {kname}\n# total stall samples {tot}; line = SASS index, instruction, samples, executed, stalls\n")
            for n in sorted(top):
                r = data[n]
                st = {c[6:]: int(r[ix[c]]) for c in stall_cols if int(r[ix[c]]) > 0.002 * tot}
                fh.write(f"{n:5d} {r[ix['Source']][:70]:70s} {r[ix['# Samples']]:>6s} {r[ix['Instructions Executed']]:>9s} {st}\n")
    print(json.dumps({k: v["value"] for k, v in summary.items()}, indent=1))


if __name__ == "__main__":
    main()
