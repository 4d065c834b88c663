"""Prints the per-kernel summary table (and top stall reasons) of an ncu report:
    python tools/ncu_summary.py gpurun_out/build.ncu-rep"""
import csv
import re
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__inst_executed_op_shared_atom.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size"]
for r in rows[2:]:
    print("==", re.sub(r"\(.*", "", r[idx["Kernel Name"]]))
    for w in want:
        if w in idx:
            print("  %-62s %s %s" % (w, r[idx[w]], units[idx[w]]))
    st = [(float(r[i].replace(",", "")), h) for i, h in enumerate(hdr)
          if re.match(r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio", h) and r[i] not in ("", "n/a")]
    print("  stalls (warps per issue):", ", ".join("%s %.2f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v)
                                                    for v, h in sorted(st, reverse=True)[:7]))
