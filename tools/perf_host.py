#!/usr/bin/env python
"""Host-side cost of the Level-B calls of one bench picture (time spent inside each vvcb200_pic_* call, GPU running
asynchronously).  Development tool."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ffmpeg_b200 import lib  # noqa: E402
from ffmpeg_b200 import hostutil as pipeline  # noqa: E402


def main():
    import torch
    width, height = 3840, 2160
    pic = bench.make_picture(2026, width, height, 0.1)
    ctx = lib.Context(0, width, height, dpb_size=6, validate=1)
    ep = pipeline.EnginePicture(ctx, pic)
    ep.upload_refs()
    for _ in range(3):
        ep.run()
    ctx.sync()
    names = ["begin", "inter", "residual", "intra", "lmcs", "deblock", "sao_alf", "download"]
    acc = dict.fromkeys(names, 0.0)
    for n in names:
        orig = getattr(ctx, n)

        def wrapped(*a, _o=orig, _n=n, **k):
            t = time.perf_counter()
            r = _o(*a, **k)
            acc[_n] += time.perf_counter() - t
            return r
        setattr(ctx, n, wrapped)
    reps = 30
    t0 = time.perf_counter()
    for _ in range(reps):
        ep.run()
    t1 = time.perf_counter()
    ctx.sync()
    t2 = time.perf_counter()
    print("host enqueue %.3f ms / picture, with final sync %.3f ms / picture" % (1e3 * (t1 - t0) / reps, 1e3 * (t2 - t0) / reps))
    for n in names:
        print("  %-9s %.3f ms" % (n, 1e3 * acc[n] / reps))
    ctx.close()


if __name__ == "__main__":
    main()
