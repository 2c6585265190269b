#!/usr/bin/env python
"""Development tool: build an experimental variant of the library next to the product one.

    python tools/build_variant.py NAME [nvcc flags ...]        e.g.  prof -DVVC_INTRA_PROFILE

Objects go to build/NAME/, the library to build/libvvcb200_NAME.so (git-ignored, travels to the GPU box).
The perf tools take --lib build/libvvcb200_NAME.so; the product library is never touched."""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ffmpeg_b200 import build as b  # noqa: E402


def main():
    name, extra = sys.argv[1], sys.argv[2:]
    obj = os.path.join(ROOT, "build", name)
    os.makedirs(obj, exist_ok=True)
    procs, objs = [], []
    for s in sorted(glob.glob(os.path.join(b.CSRC, "*.cu"))):
        o = os.path.join(obj, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        procs.append(subprocess.Popen([b.NVCC] + b.FLAGS + extra + ["-c", s, "-o", o]))
    if any(p.wait() for p in procs):
        raise SystemExit("nvcc failed")
    so = os.path.join(ROOT, "build", "libvvcb200_%s.so" % name)
    subprocess.check_call([b.NVCC, "-shared", "-o", so] + objs + ["-Xlinker", "--no-undefined"])
    print(so)


if __name__ == "__main__":
    main()
