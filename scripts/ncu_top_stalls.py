#!/usr/bin/env python
"""Top stalled SASS instructions of one kernel from `ncu -i rep --page source --csv --print-source sass` output."""
import csv
import sys


def main(path, n=25):
    rows = list(csv.reader(open(path)))
    hdr = next(r for r in rows if "# Samples" in r)
    ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    data = [r for r in rows if len(r) > isamp and r[isamp].isdigit()]
    tot = sum(int(r[isamp]) for r in data)
    print("kernel:", rows[0][1][:100] if len(rows[0]) > 1 else "?")
    print("total samples", tot, "instructions", len(data))
    reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {h: sum(int(r[hdr.index(h)] or 0) for r in data) for h in reasons}
    print("by reason:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > 0})
    order = sorted(range(len(data)), key=lambda k: -int(data[k][isamp]))[:n]
    for k in order:
        r = data[k]
        rs = {h[6:]: int(r[hdr.index(h)] or 0) for h in reasons}
        rs = {a: b for a, b in rs.items() if b > 0}
        print(f"{k:5d} {r[isamp]:>7s} {r[iex]:>9s}  {r[ia][:64]:64s} {rs}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
