#!/usr/bin/env python3
"""Print the key counters of an .ncu-rep (read offline with `ncu -i`). usage: tools/ncu_summary.py file.ncu-rep"""
import csv, io, re, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, v = rows[0], rows[-1]
pat = re.compile(r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|sm__warps_active.avg.pct_of_peak_sustained_active|launch__registers_per_thread|"
                 r"launch__block_size|launch__grid_size|launch__shared_mem_per_block_dynamic|smsp__inst_executed.sum|smsp__issue_active.avg.pct_of_peak_sustained_active|"
                 r"sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_(active|elapsed)|sm__inst_executed_pipe_fp64.sum|sm__inst_executed_pipe_lsu.sum|"
                 r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|smsp__sass_thread_inst_executed_op_(dadd|dmul|dfma)_pred_on.sum|"
                 r"l1tex__data_pipe_lsu_wavefronts_mem_shared.sum|l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed|"
                 r"smsp__cycles_active.avg|sm__cycles_elapsed.max|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|smsp__inst_executed_op_shared_(ld|st).sum|"
                 r"sm__inst_executed.sum|sm__icc_requests.*|.*icc.*|.*imc.*)$")
for a, b in zip(h, v):
    if pat.search(a):
        try:
            if float(b.replace(",", "")) == 0 and "stalled" in a: continue
        except ValueError: pass
        print(f"{a} = {b}")
