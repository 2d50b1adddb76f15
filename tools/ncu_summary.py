#!/usr/bin/env python3
"""tools/ncu_summary.py REPORT.ncu-rep [header text] -- the metrics DESIGN.md quotes, one per line,
from an `ncu --set full` report (read with `ncu -i ... --page raw --csv`).  Output goes to profiles/."""
import csv
import io
import re
import subprocess
import sys

KEEP = [r"^Kernel Name$", r"^Block Size$", r"^Grid Size$", r"^gpu__time_duration\.sum$", r"^dram__bytes_(read|write)\.sum$",
        r"^dram__throughput\.avg\.pct_of_peak_sustained_elapsed$", r"^gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed$",
        r"^lts__t_sector_hit_rate\.pct$", r"^l1tex__m_xbar2l1tex_read_bytes\.sum$", r"^l1tex__m_l1tex2xbar_write_bytes\.sum$",
        r"^sm__cycles_elapsed\.max$", r"^launch__occupancy_limit_(registers|shared_mem|warps)$", r"^sm__warps_active\.avg\.pct_of_peak_sustained_active$",
        r"^(lts|l1tex|sm)__throughput\.avg\.pct_of_peak_sustained_elapsed$", r"^smsp__issue_active\.avg\.per_cycle_active$",
        r"^sm__pipe_(fma|fmaheavy|alu)_cycles_active\.avg\.pct_of_peak_sustained_elapsed$",
        r"^sm__inst_executed_pipe_(xu|lsu)\.avg\.pct_of_peak_sustained_elapsed$", r"^smsp__inst_executed\.sum$",
        r"^smsp__sass_thread_inst_executed_op_(fadd|fmul|ffma)_pred_on\.sum$",
        r"^l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum(\.pct_of_peak_sustained_elapsed)?$",
        r"^l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_ld|_op_st)?\.sum$", r"^launch__registers_per_thread$",
        r"^launch__shared_mem_per_block_dynamic$", r"^sm__warps_active\.avg\.per_cycle_active$", r"^sm__cycles_active\.avg$",
        r"^smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio$"]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    names, units, vals = rows[0], rows[1], rows[2]
    if len(sys.argv) > 2:
        print("# " + " ".join(sys.argv[2:]))
    for n, u, v in zip(names, units, vals):
        if any(re.search(k, n) for k in KEEP):
            print(f"{n:<96s}{u:<17s}{v}")


if __name__ == "__main__":
    main()
