#!/usr/bin/env python
"""Where a kernel's executed warp instructions and stall samples go, per source line (needs -lineinfo and
--import-source on); sums over the files of the first captured launch.
usage: python tools/ncu_exec.py rep.ncu-rep kernel_regex [n]"""
import csv
import subprocess
import sys


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel,
                          "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    lines, fname, hdr, seen_files = [], "", None, set()
    for r in rows:
        if r and r[0] == "File Name":
            if r[1] in seen_files:
                break                               # second launch
            fname = r[1]
            seen_files.add(fname)
        elif r and r[0] == "Line No":
            hdr = {h: k for k, h in enumerate(r)}
        elif hdr and r and r[0].isdigit() and len(r) > hdr.get("Instructions Executed", 99):
            try:
                ex = float(r[hdr["Instructions Executed"]] or 0)
                sm = float(r[hdr["# Samples"]] or 0)
            except ValueError:
                continue
            if ex or sm:
                lines.append((ex, sm, fname.split("/")[-1], r[0], r[1].strip()))
    tot_ex, tot_sm = sum(l[0] for l in lines), sum(l[1] for l in lines)
    print("total executed %.0f, samples %.0f" % (tot_ex, tot_sm))
    for ex, sm, f, ln, src in sorted(lines, key=lambda l: -l[0])[:n]:
        print("%5.1f%% exec %5.1f%% stall  %s:%s  %s" % (100 * ex / max(tot_ex, 1), 100 * sm / max(tot_sm, 1), f, ln, src[:100]))


if __name__ == "__main__":
    main()
