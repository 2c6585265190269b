#!/usr/bin/env python
"""Development tool: SASS instructions per source line of one kernel of a cubin (where does the code size come from?).
usage: python tools/sass_size.py file.cubin kernel_regex [n]"""
import collections
import re
import subprocess
import sys


def main():
    cubin, kernel = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    fn, cur, cnt = None, None, collections.Counter()
    for l in dis.splitlines():
        if ".section" in l:
            m = re.search(r'\.text\.(\S+)', l)
            fn = m.group(1) if m else None
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        if re.match(r'\s+/\*[0-9a-f]{4,6}\*/', l) and fn and re.search(kernel, fn):
            cnt[cur] += 1
    print("total", sum(cnt.values()))
    for k, c in cnt.most_common(top):
        print("%-28s %6d" % ("%s:%d" % k if k else "?", c))


if __name__ == "__main__":
    main()
