#!/usr/bin/env python3
"""Opcode mix and stall hot-spots of one kernel from `ncu -i rep --page source --csv` output.
    ncu -i X.ncu-rep --page source --csv --kernel-name regex:NAME --launch-count 1 > /tmp/src.csv
    python profiles/sass_mix.py /tmp/src.csv [top]"""
import csv
import sys
from collections import Counter


def main(path, top=25):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ix = {k: i for i, k in enumerate(hdr)}
    data = [r for r in rows[h + 1:] if len(r) >= len(hdr) and r[ix["Instructions Executed"]].isdigit()]
    tot = sum(int(r[ix["Instructions Executed"]]) for r in data)
    samp = sum(int(r[ix["# Samples"]]) for r in data) or 1
    print("SASS instructions %d, warp-instructions executed %d, samples %d" % (len(data), tot, samp))
    c, s = Counter(), Counter()
    for r in data:
        toks = [t for t in r[ix["Source"]].split() if not t.startswith("@")]
        op = toks[0].split(".")[0] if toks else "?"
        c[op] += int(r[ix["Instructions Executed"]])
        s[op] += int(r[ix["# Samples"]])
    for op, n in c.most_common(top):
        print("%-10s %12d %5.1f%%   stall samples %5.1f%%" % (op, n, 100.0 * n / tot, 100.0 * s[op] / samp))
    print("-- top stall sites")
    for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:top]:
        print(r[ix["# Samples"]].rjust(7), r[ix["Instructions Executed"]].rjust(10), r[ix["Source"]][:100])


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
