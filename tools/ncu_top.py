#!/usr/bin/env python
"""Summarise an ncu report: headline metrics + the hottest SASS instructions (by stall samples) with executed counts."""
import csv, subprocess, sys, io
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, u = r[0], r[1]
for row in r[2:]:
    print("== kernel:", row[h.index("Kernel Name")][:100])
    for k in ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
              "launch__registers_per_thread", "sm__inst_executed.avg.per_cycle_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
              "smsp__inst_executed_op_shared_atom.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_global_ld.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
              "smsp__issue_active.avg.pct", "smsp__thread_inst_executed_per_inst_executed.ratio"]:
        if k in h:
            print(f"  {k:70s} {row[h.index(k)]:>16s} {u[h.index(k)]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; data = [x for x in rows[2:] if len(x) == len(hdr) and x[0] != 'Address']
ia, isrc, ie, it, iss = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Avg. Threads Executed"), hdr.index("Warp Stall Sampling (All Samples)")
tot_s = sum(int(x[iss] or 0) for x in data); tot_i = sum(int(x[ie] or 0) for x in data)
print(f"total stall samples {tot_s}, warp instructions {tot_i}")
for x in sorted(data, key=lambda x: -int(x[iss] or 0))[:topn]:
    print(x[ia][-5:], f"{int(x[ie] or 0):>11}", x[it][:5].rjust(6), f"{100.0 * int(x[iss] or 0) / max(tot_s, 1):6.2f}%", x[isrc][:100])
