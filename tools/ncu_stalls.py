#!/usr/bin/env python
"""Per-kernel warp-stall breakdown and the hottest source lines from an .ncu-rep captured with --set full --import-source on.
usage: python tools/ncu_stalls.py rep.ncu-rep kernel_regex [n_lines]"""
import csv
import io
import subprocess
import sys


def page(rep, kernel, what):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel, "--print-source", what],
                         capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    n_lines = int(sys.argv[3]) if len(sys.argv) > 3 else 12
    rows = page(rep, kernel, "sass")
    hdr = next(r for r in rows if r and r[0] == "Address")
    data = rows[rows.index(hdr) + 1:]
    col = {h: i for i, h in enumerate(hdr)}

    def f(r, name):
        try:
            return float(r[col[name]])
        except (ValueError, IndexError, KeyError):
            return 0.0
    samples = sum(f(r, "# Samples") for r in data)
    print("kernel %s: %d SASS instructions, %.0f warp instructions executed, %.0f stall samples" %
          (kernel, len(data), sum(f(r, "Instructions Executed") for r in data), samples))
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for s in sorted(stalls, key=lambda s: -sum(f(r, s) for r in data))[:8]:
        v = sum(f(r, s) for r in data)
        print("  %-26s %5.1f %%" % (s, 100 * v / max(samples, 1)))
    rows = page(rep, kernel, "cuda")
    lines = [r for r in rows if len(r) > 6 and r[0].isdigit()]
    hdr2 = next((r for r in rows if r and r[0] == "Line No"), None)
    if hdr2:
        c2 = {h: i for i, h in enumerate(hdr2)}
        si = c2.get("# Samples")
        if si is not None:
            lines.sort(key=lambda r: -float(r[si] or 0))
            print("  hottest source lines (stall samples):")
            for r in lines[:n_lines]:
                print("   %6s  L%-5s %s" % (r[si], r[0], r[1].strip()[:110]))


if __name__ == "__main__":
    main()
