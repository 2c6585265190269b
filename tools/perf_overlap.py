#!/usr/bin/env python
"""Experiment: throughput of the Level-A stage chain with 1, 2 or 3 pictures in flight on separate streams (GPU only)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ffmpeg_b200 import lib  # noqa: E402
from ffmpeg_b200 import hostutil as util  # noqa: E402


def main():
    import torch
    if len(sys.argv) > 1:                       # experimental library built by tools/build_variant.py
        lib.SO = os.path.abspath(sys.argv[1])
    width, height = 3840, 2160
    pic = bench.make_picture(2026, width, height, 0.1)
    it, rs, ia, db, lf = pic["inter"], pic["residual"], pic["intra"], pic["deblock"], pic["lf"]
    ctb = pic["ctb_log2"]
    irecs = lib.intra_schedule(ia["blocks"], ia["ctu_first"], (width + 127) >> 7, (height + 127) >> 7, ctb, width, height,
                               order=lib.INTRA_ORDER_DIAGONAL)
    edge = lib.pack_deblock_edges(db["bs"], db["len_p"], db["len_q"], db["tb_w_c"], db["tb_h_c"], ctb)

    def dev(a):
        return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()

    sets = []
    for k in range(4):
        s = {}
        s["refs"] = [util.to_device_planes(r, torch) for r in it["refs"]]
        s["cur"] = util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch)
        s["tmp"] = [torch.zeros_like(t) for t in s["cur"]]
        s["res"] = [torch.zeros_like(t) for t in s["cur"]]
        s["cur_p"], s["tmp_p"], s["res_p"] = lib.device_planes(s["cur"]), lib.device_planes(s["tmp"]), lib.device_planes(s["res"])
        s["refs_p"] = [lib.device_planes(r) for r in s["refs"]]
        s["blocks"], s["levels"], s["tbs"], s["irecs"] = dev(it["blocks"]), dev(rs["levels"]), dev(rs["tbs"]), dev(irecs)
        s["prof"] = dev(it["prof"]) if len(it["prof"]) else None
        s["marks"] = torch.zeros(lib.intra_scratch_bytes(width, height), dtype=torch.uint8, device="cuda")
        s["edge"], s["qp"], s["off"] = torch.from_numpy(edge).cuda(), torch.from_numpy(db["qp"]).cuda(), torch.from_numpy(db["ctb_offsets"]).cuda()
        s["db"] = lib.deblock_tables(s["edge"], s["qp"], s["off"], db["stride4"])
        s["sao"], s["alf"], s["alfp"] = dev(lf["sao"]), dev(lf["alf"]), dev(lf["alf_params"])
        sets.append(s)

    def step(s, st):
        q = st.cuda_stream
        lib.inter_batch(s["cur_p"], s["refs_p"], s["blocks"], len(it["blocks"]), it["params"], prof=s["prof"], stream=q)
        lib.residual_batch(s["cur_p"], s["levels"], s["tbs"], len(rs["tbs"]), rs["params"], residual=s["res_p"], stream=q)
        lib.intra_picture(s["cur_p"], s["res_p"], s["irecs"], len(irecs), ia["params"], s["marks"], stream=q, shadow=s["tmp_p"])
        lib.deblock_picture(s["cur_p"], s["db"], db["params"], stream=q)
        lib.sao_picture(s["tmp_p"], s["cur_p"], s["sao"], ctb, stream=q)
        lib.alf_picture(s["cur_p"], s["tmp_p"], s["alf"], s["alfp"], ctb, stream=q)

    for n_streams in (1, 2, 3, 4):
        streams = [torch.cuda.Stream() for _ in range(n_streams)]
        for i in range(8):
            step(sets[i % n_streams], streams[i % n_streams])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 48
        e0.record()
        for st in streams:
            st.wait_event(e0)
        for i in range(steps):
            step(sets[i % n_streams], streams[i % n_streams])
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        print("%d picture(s) in flight: %.4f ms per picture, %.1f frames/s" % (n_streams, ms, 1e3 / ms))


if __name__ == "__main__":
    main()
