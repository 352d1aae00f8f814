#!/usr/bin/env python
"""Assemble profiles/<name>_ncu_full_summary.txt from the exports of one `ncu --set full` capture:
    python scripts/ncu_summary.py <prefix> <out> "<header line>"
<prefix>.details.txt (ncu --page details), <prefix>.source.csv (--page source --csv [--print-source sass]) and
<prefix>.raw.csv (--page raw --csv) are read; the summary keeps the details page, the top stalled SASS instructions and
the pipe-utilisation / DRAM counters of the raw page."""
import csv
import io
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
KEEP = ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_sectors_srcunit_tex_op_read.sum", "smsp__inst_executed.sum", "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second")


def main(prefix, out, header):
    parts = ["# " + header]
    parts.append(open(prefix + ".details.txt").read().rstrip())
    src = prefix + ".source.csv"
    if os.path.exists(src) and os.path.getsize(src) > 1000:
        r = subprocess.run([sys.executable, os.path.join(HERE, "ncu_top_stalls.py"), src, "25"], capture_output=True, text=True)
        parts.append("\n# top stalled SASS instructions (scripts/ncu_top_stalls.py)\n" + (r.stdout.rstrip() or r.stderr.rstrip()))
    rows = list(csv.reader(open(prefix + ".raw.csv")))
    d = dict(zip(rows[0], rows[2])) if len(rows) > 2 else {}
    u = dict(zip(rows[0], rows[1])) if len(rows) > 2 else {}
    parts.append("\n# pipe utilisation and traffic (raw page)")
    for k in KEEP:
        if k in d:
            parts.append(f"{k} {d[k]} {u.get(k, '')}")
    open(out, "w").write("\n".join(parts) + "\n")
    print("wrote", out)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3])
