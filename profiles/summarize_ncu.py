#!/usr/bin/env python
"""Summarise an ncu report (read here, no GPU): `python profiles/summarize_ncu.py rep.ncu-rep title > profiles/x.md`."""
import csv, io, subprocess, sys

KEYS = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_op_read.sum",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "sm__cycles_elapsed.max", "smsp__inst_executed.sum"]


def main():
    rep, title = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt[txt.index('"ID"'):])))
    hdr, units = rows[0], rows[1]
    print(f"# {title}\n")
    for r in rows[2:]:
        print(f"## {r[hdr.index('Kernel Name')][:110]}  grid {r[hdr.index('Grid Size')]} block {r[hdr.index('Block Size')]}\n")
        for k in KEYS:
            for i, h in enumerate(hdr):
                if h == k or h.endswith("." + k):
                    print(f"- {k} = {r[i]} {units[i]}")
                    break
        print()


if __name__ == "__main__":
    main()
