#!/usr/bin/env python
"""Condense an .ncu-rep (ncu --set full) into the few lines the roofline claims rest on.

    python tools/ncu_summary.py gpurun_out/r1_decode.ncu-rep > profiles/r1_decode_ncu.txt

Reads the report with `ncu -i ... --page raw --csv` (works without a GPU)."""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second",
    "dram__bytes_read.sum.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__cycles_elapsed.avg",
    "sm__cycles_elapsed.avg.per_second",
    "launch__registers_per_thread",
    "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem",
    "launch__shared_mem_per_block_dynamic",
    "sass__inst_executed_local_loads",
    "sass__inst_executed_local_stores",
    "smsp__inst_executed.sum",
]


def main() -> None:
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    print(f"# source: {rep}  (ncu --set full --clock-control none; per-launch values, cold-ish cache, serialised)")
    for n, r in enumerate(rows[2:]):
        get = lambda k: r[hdr.index(k)] if k in hdr else ""  # noqa: E731
        print(f"\n## launch {n}: {get('Kernel Name')}  grid {get('Grid Size')} block {get('Block Size')}")
        for k in KEYS:
            # metric columns may carry a section prefix ("FBSP.TriageCompute.<metric>")
            for h, u, v in zip(hdr, units, r):
                if (h == k or h.endswith("." + k)) and v != "":
                    print(f"{k:85s} {v:>16s} {u}")
                    break
        stalls = [(float(v), h) for h, v in zip(hdr, r)
                  if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") and v]
        stalls.sort(reverse=True)
        if stalls:
            print("top warp-stall reasons (warps stalled per issue-active cycle):")
            for v, h in stalls[:6]:
                print(f"    {h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''):40s} {v:8.3f}")


if __name__ == "__main__":
    main()
