#!/usr/bin/env python
"""Turn an ncu launch list (CSV from `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--csv`) of ONE ACT forward at the bench batch into profiles/<tag>_dram_traffic.json and a per-kernel summary table.

    python scripts/ncu_traffic.py gpurun_out/launches.csv --tag r02 [--commit <sha>] [--last-step N]

Rows are grouped by kernel (template arguments dropped; the LayerNorm-fused 512-wide GEMM keeps its own name, as in
bench.py's per-launch records). With --last-step N only the last N launches (= the launches of one forward) count.
"""
import argparse
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def kernel_key(name):
    base = re.sub(r"^void\s+", "", name)
    short = re.sub(r"<.*", "", base).replace("rmb::", "").strip()
    short = re.sub(r"\(.*", "", short)
    if short == "layernorm512_kernel_t":              # the two input-type instances report under bench.py's launch name
        return "layernorm512_kernel"
    if short == "gemm_tc_pair_kernel" and re.search(r"<\s*\(?int\)?512", base):
        return "gemm_tc_pair_kernel<LN>"
    return short


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--tag", default="r02")
    ap.add_argument("--commit", default="")
    ap.add_argument("--last-step", type=int, default=0)
    args = ap.parse_args()
    rows = {}
    with open(args.csv, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if "ID" not in r or "Metric Name" not in r:
            continue
        i = int(r["ID"])
        e = rows.setdefault(i, {"name": r["Kernel Name"]})
        try:
            val = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        unit = r.get("Metric Unit", "")
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6,
                 "second": 1e3, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1.0)
        e[r["Metric Name"]] = val * scale
    ids = sorted(rows)
    if args.last_step:
        ids = ids[-args.last_step:]
    kernels = {}
    for i in ids:
        e = rows[i]
        k = kernels.setdefault(kernel_key(e["name"]), {"launches": 0, "ms": 0.0, "read": 0.0, "write": 0.0})
        k["launches"] += 1
        k["ms"] += e.get("gpu__time_duration.sum", 0.0)
        k["read"] += e.get("dram__bytes_read.sum", 0.0)
        k["write"] += e.get("dram__bytes_write.sum", 0.0)
    total_ms = sum(k["ms"] for k in kernels.values())
    out = {"source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none, "
                     "python scripts/profile_step.py --batch 256 --once; the launches of the last forward",
           "commit": args.commit, "total_ms_serialised": total_ms, "kernels": {}}
    print(f"{'kernel':28s} {'launches':>8s} {'ms':>9s} {'share':>7s} {'DRAM read GB':>13s} {'DRAM write GB':>14s}")
    for name, k in sorted(kernels.items(), key=lambda kv: -kv[1]["ms"]):
        out["kernels"][name] = {"launches": k["launches"], "ms": k["ms"], "share": k["ms"] / total_ms if total_ms else 0,
                                "dram_read_bytes": k["read"], "dram_write_bytes": k["write"],
                                "dram_bytes_per_launch_mean": (k["read"] + k["write"]) / k["launches"]}
        print(f"{name:28s} {k['launches']:8d} {k['ms']:9.3f} {100 * k['ms'] / max(total_ms, 1e-9):6.1f}% {k['read'] / 1e9:13.3f} {k['write'] / 1e9:14.3f}")
    path = os.path.join(ROOT, "profiles", f"{args.tag}_dram_traffic.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print(f"-> {path}")


if __name__ == "__main__":
    sys.exit(main())
