"""Readable summary of one or more .ncu-rep files (`ncu --set full`): per kernel launch the duration, DRAM bytes and
throughput, issue-slot / L1 / L2 / pipe utilisation, shared-memory bank conflicts, registers, grid.
usage: python tools/ncu_summary.py a.ncu-rep [b.ncu-rep ...] > profiles/rNN_ncu_xxx.txt"""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "sass__inst_executed_local_loads"]
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u = rows[0], rows[1]
    print(f"== {rep}")
    for r in rows[2:]:
        print("-----")
        print(f"  Kernel Name = {r[h.index('Kernel Name')].split('(')[0]}")
        for k in WANT:
            if k in h:
                i = h.index(k)
                print(f"  {k} [{u[i]}] = {r[i]}")
