#!/usr/bin/env python
"""Warp instructions executed per CUDA source line of one kernel: the SASS page of an .ncu-rep (first captured launch)
joined, instruction by instruction, with the line table of the same kernel in the built library (nvdisasm -g).
usage: python tools/ncu_lines.py rep.ncu-rep kernel_regex cubin mangled_substring [n]
  e.g. cuobjdump -xelf all ffmpeg_b200/libvvcb200.so  ->  saoalf.sm_100a.cubin"""
import collections
import csv
import os
import re
import subprocess
import sys


def sass_rows(rep, kernel):
    skip = os.environ.get("NCU_SKIP")              # n-th matching launch instead of the first
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel,
                          "--print-source", "sass"] + (["--launch-skip", skip, "--launch-count", "1"] if skip else []),
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    first = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[first]
    col = {h: k for k, h in enumerate(hdr)}
    data = []
    for r in rows[first + 1:]:
        if r and r[0] == "Address":
            break
        if len(r) == len(hdr):
            data.append((r[col["Source"]].strip(), float(r[col["Instructions Executed"]] or 0), float(r[col["# Samples"]] or 0)))
    return data


def line_table(cubin, mangled):
    out = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    lines, cur, on = [], None, False
    for l in out:
        if l.startswith(".text."):
            on = mangled in l
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l) and ".byte" not in l and ".dword" not in l:
            lines.append(cur)
    return lines


def main():
    rep, kernel, cubin, mangled = sys.argv[1:5]
    n = int(sys.argv[5]) if len(sys.argv) > 5 else 40
    data, lt = sass_rows(rep, kernel), line_table(cubin, mangled)
    print("%d SASS rows in the report, %d in the library" % (len(data), len(lt)))
    if len(data) != len(lt):
        print("instruction counts differ: the library is not the profiled build")
        return
    ex, smp = collections.Counter(), collections.Counter()
    for (src, e, s), ln in zip(data, lt):
        ex[ln] += e
        smp[ln] += s
    tot = sum(ex.values())
    src_cache = {}
    for ln, e in ex.most_common(n):
        f = ln[0] if ln else "?"
        if f not in src_cache:
            try:
                src_cache[f] = open("ffmpeg_b200/csrc/" + f).read().splitlines()
            except OSError:
                src_cache[f] = []
        text = src_cache[f][ln[1] - 1].strip() if ln and ln[1] - 1 < len(src_cache[f]) else ""
        print("%5.1f%% %10.0f  smp %5.0f  %s:%d  %s" % (100 * e / tot, e, smp[ln], f, ln[1] if ln else 0, text[:100]))


if __name__ == "__main__":
    main()
