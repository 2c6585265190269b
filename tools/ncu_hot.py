#!/usr/bin/env python
"""Hottest SASS instructions (by stall samples) of one kernel in an .ncu-rep, first captured launch only.
usage: python tools/ncu_hot.py rep.ncu-rep kernel_regex [n]"""
import csv
import subprocess
import sys


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel,
                          "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    first = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[first]
    col = {h: k for k, h in enumerate(hdr)}
    data = []
    for r in rows[first + 1:]:
        if r and r[0] == "Address":
            break                                   # next launch
        if len(r) == len(hdr):
            data.append(r)

    def f(r, name):
        try:
            return float(r[col[name]])
        except (ValueError, KeyError):
            return 0.0
    total = sum(f(r, "# Samples") for r in data)
    print("%d instructions, %.0f executed, %.0f samples" % (len(data), sum(f(r, "Instructions Executed") for r in data), total))
    idx = sorted(range(len(data)), key=lambda i: -f(data[i], "# Samples"))[:n]
    for i in sorted(idx):
        r = data[i]
        print("%5d %6.0f (%4.1f%%) exec %9.0f  %s" % (i, f(r, "# Samples"), 100 * f(r, "# Samples") / max(total, 1),
                                                     f(r, "Instructions Executed"), r[col["Source"]][:110]))


if __name__ == "__main__":
    main()
