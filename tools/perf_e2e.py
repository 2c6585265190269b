#!/usr/bin/env python
"""Development probe: Level-B throughput of the bench picture with and without the picture download / the refined-MV
read-back, and with more DPB slots (what the copies cost on top of the kernels, and why)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ffmpeg_b200 import lib  # noqa: E402
from ffmpeg_b200 import hostutil as pipeline  # noqa: E402


def main():
    width, height = 3840, 2160
    pic = bench.make_picture(2026, width, height, 0.1)
    for slots, mv in ((4, True), (4, False), (8, True), (8, False)):
        ctx = lib.Context(0, width, height, dpb_size=2 + slots, validate=1)
        ep = pipeline.EnginePicture(ctx, pic, max_slots=slots, want_mv=mv)
        ep.upload_refs()
        for download in (True, False):
            for _ in range(4):
                ep.run(download=download)
            ctx.sync()
            t0 = time.perf_counter()
            for _ in range(60):
                ep.run(download=download)
            ctx.sync()
            dt = time.perf_counter() - t0
            print("dpb slots %d, refined mv %s, download %s: %.4f ms per picture, %.1f frames/s" % (slots, mv, download, 1e3 * dt / 60, 60 / dt))
        ctx.close()


if __name__ == "__main__":
    main()
