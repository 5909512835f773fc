#!/usr/bin/env python
"""Summarise a `ncu --set full` report into the JSON kept under profiles/ (one object per kernel
name: the last launch of that name in the report).  usage: ncu_summary.py report.ncu-rep out.json
Reads the report through `ncu -i ... --page raw --csv` (B200_PROFILING.md); dram byte counters are
converted to MB, shared-memory sizes to KB, durations to microseconds as ncu prints them."""
import csv
import io
import json
import re
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum",
]
STALL = re.compile(r"smsp__average_warps?_issue_stalled_(\w+)_per_issue_active\.ratio$|"
                   r"smsp__average_warp_latency_issue_stalled_(\w+)\.ratio$")


def main():
    rep, out = sys.argv[1], sys.argv[2]
    text = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True,
                          capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    head, units = rows[0], rows[1]
    by_kernel = {}
    for r in rows[2:]:
        rec = dict(zip(head, r))
        name = re.sub(r"\(.*", "", rec["Kernel Name"]).split("::")[-1]
        name = re.sub(r"<.*", "", name)
        m, stalls = {}, []
        for k, v in rec.items():
            short = k.split(".", 2)[-1] if k.count(".") >= 2 and k.split(".")[1] == "TriageCompute" else k
            if short in KEEP and v not in ("", "no data"):
                m[short] = v + (" " + units[head.index(k)] if units[head.index(k)] else "")
            s = STALL.search(k)
            if s and v not in ("", "no data"):
                try:
                    stalls.append((s.group(1) or s.group(2), round(float(v), 3)))
                except ValueError:
                    pass
        stalls.sort(key=lambda t: -t[1])
        by_kernel[name] = {"kernel": name, "block": rec["Block Size"], "grid": rec["Grid Size"],
                           "metrics": m, "top_stalls_warps_per_issue": stalls[:5]}
    json.dump(list(by_kernel.values()), open(out, "w"), indent=1)
    print("wrote", out, "kernels:", ", ".join(by_kernel))


if __name__ == "__main__":
    main()
