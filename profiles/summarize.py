#!/usr/bin/env python
"""profiles/summarize.py <file.ncu-rep> — prints the metrics DESIGN.md / bench.py quote from an `ncu --set full` capture
(run here, no GPU needed: `ncu -i … --page raw --csv`)."""
import csv
import subprocess
import sys

KEYS = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio",
        "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio", "smsp__average_warp_latency_issue_stalled_barrier.ratio",
        "smsp__average_warp_latency_issue_stalled_wait.ratio", "smsp__average_warp_latency_issue_stalled_not_selected.ratio"]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr, units = rows[start], rows[start + 1]
    for r in rows[start + 2:]:
        print(f"--- launch {r[0]}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"{k} = {r[i]} {units[i]}")


def traffic(path):
    """dram__bytes_read.sum + dram__bytes_write.sum of the (single) profiled launch, in bytes"""
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr, units, r = rows[start], rows[start + 1], rows[start + 2]
    mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = hdr.index(k)
        tot += float(r[i]) * mul[units[i]]
    return int(tot)


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":            # summarize.py --traffic out.json q8_0=a.ncu-rep q4_0=b.ncu-rep ...
        import json
        res = {kv.split("=")[0]: traffic(kv.split("=")[1]) for kv in sys.argv[3:]}
        json.dump(res, open(sys.argv[2], "w"), indent=1)
        print(res)
    else:
        main(sys.argv[1])
