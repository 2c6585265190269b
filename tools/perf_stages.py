#!/usr/bin/env python
"""Stage-level timing experiments on the bench picture (GPU only; development tool, not part of the product).

    python tools/perf_stages.py [--workload 4k|1080p] [--intra-fraction F]

Times vvcb200_inter_batch on subsets of the prediction records (by tool), the residual, intra and filter stages,
each as the mean of `reps` launches bracketed by CUDA events."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ffmpeg_b200 import abi, lib  # noqa: E402
from ffmpeg_b200 import hostutil as util  # noqa: E402


def timed(torch, fn, reps=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="4k")
    ap.add_argument("--intra-fraction", type=float, default=0.1)
    ap.add_argument("--seed", type=int, default=2026)
    ap.add_argument("--lib", default=None, help="experimental library built by tools/build_variant.py")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default=None, help="inter | residual | intra | filters")
    args = ap.parse_args()
    global timed
    _timed = timed
    timed = lambda torch, fn, reps=None: _timed(torch, fn, args.reps)
    if args.lib:
        lib.SO = os.path.abspath(args.lib)
    width, height = bench.WORKLOADS[args.workload]
    pic = bench.make_picture(args.seed, width, height, args.intra_fraction)
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]

    def dev(a):
        return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()

    refs = [util.to_device_planes(r, torch) for r in it["refs"]]
    refs_p = [lib.device_planes(r) for r in refs]
    cur = util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch)
    cur_p = lib.device_planes(cur)
    prof = dev(it["prof"]) if len(it["prof"]) else None
    fwd = dev(pic["lmcs"][0]) if pic.get("lmcs") is not None else None
    b = it["blocks"]
    f = b["flags"]
    subsets = {
        "all": np.ones(len(b), bool),
        "regular": (f & (abi.PB_AFFINE | abi.PB_GPM | abi.PB_CHROMA_ONLY | abi.PB_LUMA_ONLY)) == 0,
        "regular_plain": (f & (abi.PB_AFFINE | abi.PB_GPM | abi.PB_CHROMA_ONLY | abi.PB_LUMA_ONLY | abi.PB_DMVR | abi.PB_BDOF)) == 0,
        "plain_uni": ((f & (abi.PB_AFFINE | abi.PB_GPM | abi.PB_CHROMA_ONLY | abi.PB_LUMA_ONLY | abi.PB_DMVR | abi.PB_BDOF)) == 0) & (b["pred_flag"] != 3),
        "plain_bi": ((f & (abi.PB_AFFINE | abi.PB_GPM | abi.PB_CHROMA_ONLY | abi.PB_LUMA_ONLY | abi.PB_DMVR | abi.PB_BDOF)) == 0) & (b["pred_flag"] == 3),
        "dmvr": (f & abi.PB_DMVR) != 0,
        "bdof_only": ((f & abi.PB_BDOF) != 0) & ((f & abi.PB_DMVR) == 0),
        "affine_luma": (f & abi.PB_LUMA_ONLY) != 0,
        "affine_chroma": (f & abi.PB_CHROMA_ONLY) != 0,
        "gpm": (f & abi.PB_GPM) != 0,
    }
    for name, m in subsets.items():
        sub = np.ascontiguousarray(b[m])
        if not len(sub) or args.only not in (None, "inter"):
            continue
        p = abi.InterParams.from_buffer_copy(bytes(it["params"]))
        p.n_w16, p.n_w8 = int((sub["w"] == 16).sum()), int((sub["w"] == 8).sum())
        d = dev(sub)
        ms = timed(torch, lambda: lib.inter_batch(cur_p, refs_p, d, len(sub), p, lmcs_fwd_lut=fwd, prof=prof))
        area = int((sub["w"].astype(np.int64) * sub["h"]).sum())
        print("inter %-14s %7d records %9d luma-samples  %.4f ms  (%.2f ns/record)" % (name, len(sub), area, ms, 1e6 * ms / len(sub)))

    if args.only in (None, "filters"):
        db, lf = pic["deblock"], pic["lf"]
        src = util.to_device_planes(bench.smooth_source(pic) if hasattr(bench, "smooth_source") else it["refs"][0], torch)
        src_p = lib.device_planes(src)
        edge = lib.pack_deblock_edges(db["bs"], db["len_p"], db["len_q"], db["tb_w_c"], db["tb_h_c"], pic["ctb_log2"])
        d_edge, d_qp, d_off = torch.from_numpy(edge).cuda(), torch.from_numpy(db["qp"]).cuda(), torch.from_numpy(db["ctb_offsets"]).cuda()
        tabs = lib.deblock_tables(d_edge, d_qp, d_off, db["stride4"])
        inv = dev(pic["lmcs"][1]) if pic.get("lmcs") is not None else None
        ms = timed(torch, lambda: lib.deblock_lmcs_picture(cur_p, src_p, tabs, db["params"], inv))
        print("deblock + inverse LMCS (tile pass)  %.4f ms" % ms)
        sao, alf, alfp = dev(lf["sao"]), dev(lf["alf"]), dev(lf["alf_params"])
        table = torch.zeros(lib.sao_alf_scratch_bytes(), dtype=torch.uint8, device="cuda")
        ms = timed(torch, lambda: lib.sao_alf_picture(cur_p, src_p, sao, alf, alfp, pic["ctb_log2"], scratch=table))
        print("SAO + ALF (one pass, table)         %.4f ms" % ms)
        ms = timed(torch, lambda: lib.sao_alf_picture(cur_p, src_p, sao, alf, alfp, pic["ctb_log2"], scratch=None))
        print("SAO + ALF (one pass, no table)      %.4f ms" % ms)
        ms = timed(torch, lambda: lib.sao_alf_picture(cur_p, src_p, sao, None, None, pic["ctb_log2"]))
        print("SAO alone (one pass)                %.4f ms" % ms)
        if args.only == "filters":
            return

    if args.only == "inter_after":
        # the INTER stage timed by events on the launching stream, alone and right after another stage's kernels
        sub = np.ascontiguousarray(b)
        p = abi.InterParams.from_buffer_copy(bytes(it["params"]))
        d = dev(sub)
        lf = pic["lf"]
        src = util.to_device_planes(it["refs"][0], torch)
        src_p = lib.device_planes(src)
        out2 = util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch)
        out2_p = lib.device_planes(out2)
        sao, alf, alfp = dev(lf["sao"]), dev(lf["alf"]), dev(lf["alf_params"])
        table = torch.zeros(lib.sao_alf_scratch_bytes(), dtype=torch.uint8, device="cuda")

        def inter():
            lib.inter_batch(cur_p, refs_p, d, len(sub), p, lmcs_fwd_lut=fwd, prof=prof)

        def saoalf():
            lib.sao_alf_picture(out2_p, src_p, sao, alf, alfp, pic["ctb_log2"], scratch=table)
        for label, before in (("alone", None), ("after SAO + ALF", saoalf), ("after another INTER", inter)):
            tot, n = 0.0, 20
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
            inter(); torch.cuda.synchronize()
            for e0, e1 in evs:
                if before:
                    before()
                e0.record(); inter(); e1.record()
            torch.cuda.synchronize()
            print("inter %-22s %.4f ms" % (label, sum(e0.elapsed_time(e1) for e0, e1 in evs) / n))
        return

    if args.only == "inter_rot":
        # which operand being cold (not in L2) costs the INTER stage what: rotate four copies of one operand at a time
        sub = np.ascontiguousarray(b)
        p = abi.InterParams.from_buffer_copy(bytes(it["params"]))
        R = 4
        refs_sets = [[util.to_device_planes(r, torch) for r in it["refs"]] for _ in range(R)]
        refs_ps = [[lib.device_planes(r) for r in rs_] for rs_ in refs_sets]
        curs = [util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch) for _ in range(R)]
        cur_ps = [lib.device_planes(c) for c in curs]
        blks = [dev(sub) for _ in range(R)]
        profs = [dev(it["prof"]) for _ in range(R)]
        junk = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
        for label, rr, rc, rb, flush in (("nothing rotates", 0, 0, 0, 0), ("refs rotate", 1, 0, 0, 0), ("dst rotates", 0, 1, 0, 0),
                                         ("records rotate", 0, 0, 1, 0), ("all rotate", 1, 1, 1, 0), ("L2 flushed before each", 0, 0, 0, 1)):
            k = [0]

            def fn():
                i = k[0] % R
                k[0] += 1
                lib.inter_batch(cur_ps[i if rc else 0], refs_ps[i if rr else 0], blks[i if rb else 0], len(sub), p, lmcs_fwd_lut=fwd,
                                prof=profs[i if rb else 0])
            if not flush:
                print("inter, %-24s %.4f ms" % (label, timed(torch, fn)))
            else:
                tot = 0.0
                for _ in range(10):
                    junk.add_(1)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                    tot += e0.elapsed_time(e1)
                print("inter, %-24s %.4f ms" % (label, tot / 10))
        return

    res = [torch.zeros_like(t) for t in cur]
    res_p = lib.device_planes(res)
    tbs, levels = dev(rs["tbs"]), dev(rs["levels"])
    ms = timed(torch, lambda: lib.residual_batch(cur_p, levels, tbs, len(rs["tbs"]), rs["params"], residual=res_p))
    if args.only not in (None, "residual", "intra"):
        return
    print("residual %d TBs  %.4f ms; classes %s" % (len(rs["tbs"]), ms, list(rs["params"].n_class)))
    ctb = pic["ctb_log2"]
    cw, ch = (width + (1 << ctb) - 1) >> ctb, (height + (1 << ctb) - 1) >> ctb
    marks = torch.zeros(lib.intra_scratch_bytes(width, height), dtype=torch.uint8, device="cuda")
    side = [torch.zeros_like(t) for t in cur]
    side_p = lib.device_planes(side)
    cf = ia["ctu_first"].astype(np.int64)
    ctu_of = np.repeat(np.arange(cw * ch), np.diff(cf))
    k_in = np.arange(len(ia["blocks"])) - cf[ctu_of]
    diag = (ctu_of % cw) + 2 * (ctu_of // cw)
    diag_order = ia["blocks"][np.lexsort((ctu_of // cw, k_in, diag))]
    nodep = diag_order.copy()                   # same work with every dependency removed: the kernel's throughput floor
    nodep["avail_top"] = 0
    nodep["avail_left"] = 0
    nodep["up_left"] = 0
    nodep["flags"] &= ~np.uint8(abi.IB_LMCS_SCALE)
    cases = [("scheduled", lib.intra_schedule(ia["blocks"], ia["ctu_first"], cw, ch, ctb, width, height), None),
             ("ctu diagonals", diag_order, None), ("decoding order", ia["blocks"], None), ("no dependencies", nodep, None),
             ("diagonals, marks", diag_order, -1), ("no deps, marks", nodep, -1)]
    if args.lib and "prof" in args.lib:         # occupancy experiments (profile builds read VVC_INTRA_CTAS)
        for ctas in (148, 296, 444):
            cases += [("diagonals, %d CTAs" % ctas, diag_order, ctas), ("no deps, %d CTAs" % ctas, nodep, ctas)]
    for label, recs, ctas in cases:
        shadow = None if ctas == -1 else side_p          # -1: in-place records by marks and fences
        if ctas and ctas > 0:
            os.environ["VVC_INTRA_CTAS"] = str(ctas)
        else:
            os.environ.pop("VVC_INTRA_CTAS", None)
        d = dev(recs)
        ms = timed(torch, lambda: lib.intra_picture(cur_p, res_p, d, len(recs), ia["params"], marks, shadow=shadow))
        print("intra %-15s %d records  %.4f ms" % (label, len(recs), ms))
        if hasattr(lib.load(), "vvcb200_intra_profile"):
            import ctypes as C
            buf = (C.c_ulonglong * 8)()
            lib.load().vvcb200_intra_profile(buf, 1)                      # reset
            lib.intra_picture(cur_p, res_p, d, len(recs), ia["params"], marks, shadow=shadow)
            torch.cuda.synchronize()
            lib.load().vvcb200_intra_profile(buf, 1)
            n = max(1, buf[3])
            print("   per record (cycles): ticket+load %.0f, body %.0f (of which waiting %.0f, edge gather %.0f, stage wait %.0f), "
                  "publish %.0f; pre-wait %.0f; records %d" % (buf[0] / n, buf[1] / n, buf[4] / n, buf[5] / n, buf[6] / n, buf[2] / n, buf[7] / n, buf[3]))


if __name__ == "__main__":
    main()
