#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the handful of counters DESIGN.md and
bench.py quote: duration, DRAM bytes, tensor-pipe / issue utilisation, shared-memory
wavefronts, stall-reason samples.  Usage: python scripts/ncu_summary.py rep.ncu-rep > profiles/x.txt"""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__bytes_read.sum.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max",
    "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"== {d.get('Kernel Name', '?')}  grid {d.get('Grid Size', '?')} block {d.get('Block Size', '?')}")
        for k in KEYS:
            if k in d:
                print(f"  {k:75s} {d[k]:>16s} {units[hdr.index(k)]}")
        stalls = sorted(((float(v), k) for k, v in d.items()
                         if k.startswith("smsp__pcsamp_warps_issue_stalled_") and not k.endswith("_not_issued") and v),
                        reverse=True)
        tot = sum(v for v, _ in stalls) or 1.0
        print("  stall samples: " + ", ".join(f"{k.rsplit('stalled_', 1)[1]} {100 * v / tot:.0f}%" for v, k in stalls[:8]))


if __name__ == "__main__":
    main()
