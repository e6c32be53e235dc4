#!/usr/bin/env python
"""Text summary of one kernel of an `ncu --set full --import-source on` report: headline counters, stall samples and the
source lines with the most samples / instructions (tools/ncu_lines.py).
    python tools/ncu_summary.py <report.ncu-rep> <mangled kernel name> [lines]"""
import csv, io, os, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = sys.argv[3] if len(sys.argv) > 3 else "25"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
ix = {h: i for i, h in enumerate(hdr)}
print("ncu --set full --clock-control none --import-source on  (%s)" % os.path.basename(rep))
print("kernel:", kern)
for w in want:
    if w in ix:
        print("%-72s %s %s" % (w, vals[ix[w]], units[ix[w]]))
st = [(float(vals[i]), h.replace("smsp__pcsamp_warps_issue_stalled_", "")) for i, h in enumerate(hdr)
      if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
tot = sum(v for v, _ in st) or 1.0
print("stall samples: " + ", ".join("%s %.1f %%" % (n, 100 * v / tot) for v, n in sorted(st, reverse=True)[:8]))
print()
here = os.path.dirname(os.path.abspath(__file__))
print(subprocess.run([sys.executable, os.path.join(here, "ncu_lines.py"), rep, kern, top], capture_output=True, text=True).stdout)
