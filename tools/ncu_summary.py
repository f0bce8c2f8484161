#!/usr/bin/env python
"""Developer tool: condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the handful of counters the
design notes cite: duration, pipe utilisation, issue rate, occupancy, DRAM bytes, L1/L2 hit rates, stall reasons.
  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--source KERNEL_REGEX] > profiles/xxx.txt
"""
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warp_latency_per_inst_issued.ratio",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for r in rows[2:]:
        print("=" * 100)
        print(r[idx["Kernel Name"]][:160])
        for k in KEYS:
            if k in idx:
                print(f"  {k:75s} {r[idx[k]]:>18s} {units[idx[k]]}")
        st = sorted(((float(r[idx[s]] or 0), s) for s in stalls), reverse=True)
        print("  stall reasons (warps stalled per issue-active cycle):")
        for v, s in st[:7]:
            print(f"      {s.split('stalled_')[1].split('_per_')[0]:28s} {v:8.3f}")
    if len(sys.argv) > 3 and sys.argv[2] == "--source":
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + sys.argv[3]], capture_output=True, text=True).stdout
        print(src[:200])


if __name__ == "__main__":
    main()
