#!/usr/bin/env python
"""Condense an ``.ncu-rep`` (read with ``ncu -i … --page raw --csv``) into the markdown summary kept under profiles/.

usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep profiles/x.md ["title / what was run"]
"""
import csv
import io
import json
import os
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__cycles_elapsed.max", "SM cycles"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__cluster_size", "cluster size"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__shared_mem_per_block_dynamic", "dyn smem/block"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (smem)"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe (hmma subpipe) active %"),
    ("sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active", "uniform pipe %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "memory throughput % of peak"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM written"),
    ("dram__bytes_read.sum.per_second", "DRAM read rate"),
    ("dram__bytes_write.sum.per_second", "DRAM write rate"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "stall long_scoreboard"),
    ("smsp__average_warp_latency_issue_stalled_barrier.ratio", "stall barrier"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stalled long_scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stalled barrier / issue"),
    ("smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "stalled membar / issue"),
    ("smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio", "stalled sleeping / issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stalled wait / issue"),
    ("smsp__inst_executed.sum", "warp instructions"),
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    title = sys.argv[3] if len(sys.argv) > 3 else os.path.basename(rep)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = [f"# {title}", "", f"Source report: `{rep}` (scratch, not tracked). Extracted with "
             "`ncu -i <rep> --page raw --csv` by `scripts/ncu_summary.py`; capture flags "
             "`--set full --clock-control none --import-source on`. Durations under ncu are NOT bench numbers.", ""]
    seen = {}
    for r in rows[2:]:
        name = r[idx["Kernel Name"]]
        seen.setdefault(name, []).append(r)
    for name, rs in seen.items():
        r = rs[0]
        lines += [f"## `{name[:140]}`", "", f"instances captured: {len(rs)}", "", "| metric | value |", "|---|---|"]
        for k, label in KEYS:
            if k in idx and r[idx[k]] != "":
                lines.append(f"| {label} (`{k}`) | {r[idx[k]]} {units[idx[k]]} |")
        lines.append("")
    with open(out, "w") as f:
        f.write("\n".join(lines) + "\n")
    print(f"wrote {out} ({len(seen)} kernels)")


if __name__ == "__main__":
    main()
