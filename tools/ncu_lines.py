#!/usr/bin/env python
"""Development tool: executed warp instructions and stall samples of one kernel per source line, by joining the SASS
page of an .ncu-rep with the line table of the kernel's cubin (for reports captured without source import).
usage: python tools/ncu_lines.py rep.ncu-rep kernel_regex file.cubin [n_lines]"""
import collections
import csv
import io
import re
import subprocess
import sys


def main():
    rep, kernel, cubin = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    line_of, cur, fn, want = {}, None, None, None
    for l in dis.splitlines():
        m = re.search(r'\.text\.(\S+)', l) if l.startswith("//---") or ".section" in l else None
        if m:
            fn = m.group(1)
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/', l)
        if m and fn and re.search(kernel, fn):
            line_of[int(m.group(1), 16)] = cur
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel, "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next(r for r in rows if r and r[0] == "Address")
    data = rows[rows.index(hdr) + 1:]
    col = {h: i for i, h in enumerate(hdr)}

    def f(r, n):
        try:
            return float(r[col[n]])
        except (ValueError, KeyError, IndexError):
            return 0.0
    base = int(data[0][col["Address"]], 16)
    acc = collections.defaultdict(lambda: [0.0, 0.0, 0])
    for r in data:
        k = line_of.get(int(r[col["Address"]], 16) - base)
        acc[k][0] += f(r, "Instructions Executed")
        acc[k][1] += f(r, "# Samples")
        acc[k][2] += 1
    tot = sum(v[0] for v in acc.values()) or 1
    tots = sum(v[1] for v in acc.values()) or 1
    print("line, executed warp instructions (share), stall samples (share), SASS instructions")
    for k, v in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%-22s %12.0f (%4.1f%%) %7.0f (%4.1f%%) %5d" % (k, v[0], 100 * v[0] / tot, v[1], 100 * v[1] / tots, v[2]))


if __name__ == "__main__":
    main()
