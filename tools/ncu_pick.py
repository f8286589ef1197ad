#!/usr/bin/env python
"""Print the handful of ncu metrics DESIGN.md argues from, one column per captured launch.
  python tools/ncu_pick.py gpurun_out/x.ncu-rep [more.ncu-rep ...]"""
import csv, subprocess, sys, io
PICK = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum"]
STALL = "smsp__average_warps_issue_stalled_"
for rep in sys.argv[1:]:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    print("==", rep, [r[ki][:40] for r in data])
    for i, h in enumerate(hdr):
        if h in PICK or (h.startswith(STALL) and h.endswith("_per_issue_active.ratio") and "not_issued" not in h):
            vals = [r[i] for r in data]
            try:
                if h.startswith(STALL) and max(float(v) for v in vals) < 0.3:
                    continue
            except ValueError:
                pass
            print(f"  {h.replace(STALL, 'stall_').replace('_per_issue_active.ratio', '')[:75]:75s} {units[i]:14s} {vals}")
