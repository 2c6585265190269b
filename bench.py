#!/usr/bin/env python
"""bench.py — VVC reconstruction + in-loop-filter throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload 4k|1080p|8k|tx1080p]

One "step" = one synthetic picture through every stage the engine implements (see `config.stages`):
INTER (luma 8-tap / chroma 4-tap MC, uni/bi/weighted/BCW, DMVR, BDOF, affine + PROF, GPM) -> residual (dequant +
inverse DCT2/DST7/DCT8 + add) -> INTRA -> inverse LMCS + DEBLOCK_V + DEBLOCK_H (one pass) -> SAO + ALF + CC-ALF (one pass).
Workloads (BASELINE.json configs): 4k = config 3 (random-access picture, the one the metric is quoted on), 1080p = config 2,
8k = config 4 (all-intra 7680x4320), tx1080p = config 1 (dequant + inverse transform batch only).
`value`   : pictures/s with all descriptor batches and reference pictures resident in HBM, CUDA-event timed,
            max over ranks (N independent streams, one per GPU).
`e2e`     : the same picture through the Level-B C ABI with HOST (pinned) buffers: every step uploads all
            descriptor batches and downloads the finished picture inside the timed region.
`roofline`: the slowest kernel's algorithmic bytes / its live CUDA-event duration vs MEASURED_PEAKS.json.
`cpu_baseline` / `--impl reference`: the reference's own C (oracle/_ref; the oracle port when that is absent)
            on the box's host cores, all threads, same picture, same stage chain.
Prints exactly one JSON line on rank 0.
"""
import argparse
import json
import os
import pickle
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "4K 10-bit VVC recon+loop-filter frames/s per B200; % HBM roofline; 8-GPU agg"
WORKLOADS = {"4k": (3840, 2160), "1080p": (1920, 1080), "8k": (7680, 4320), "tx1080p": (1920, 1080)}
L2_BYTES = 126 * 1024 * 1024
STAGES = ["inter", "residual", "intra", "deblock", "sao_alf"]


def bind_near_gpu(local_rank):
    """One process per GPU: run (and, by first touch, allocate pinned host memory) on the cores of the GPU's NUMA node —
    uploads and downloads that cross the socket interconnect share it with every other rank's.  Returns the cores or None."""
    if env_int("VVCB200_BENCH_NO_BIND", 0):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w in range(words) for b in range(64) if (mask[w] >> b) & 1]
        allowed = sorted(set(cpus) & os.sched_getaffinity(0))
        if allowed and len(allowed) < len(os.sched_getaffinity(0)):
            os.sched_setaffinity(0, allowed)
            return allowed
    except Exception:
        pass
    return None


_REAL_STDOUT = None


def stdout_for_the_json_line_only():
    """Libraries print to file descriptor 1 on their own (NCCL its version line): from here on fd 1 is stderr, and
    emit_line() writes the one JSON line to what stdout was."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, line)


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


# ----------------------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------------------
def make_picture(seed, width, height, intra_fraction, lmcs=True):
    """Synthetic parsed random-access picture (SURVEY.md §8d config 2/3 shape): one CU partition of inter units (regular
    with DMVR/BDOF/BCW/WP, affine + PROF, GPM) and intra units.  Cached under /tmp (or $VVCB200_BENCH_CACHE) for re-runs."""
    from ffmpeg_b200 import abi, synth
    cache = os.path.join(os.environ.get("VVCB200_BENCH_CACHE", "/tmp"),
                         "vvcb200_bench_v12_%dx%d_%d_%d_%d.pkl" % (width, height, seed, round(intra_fraction * 1000), int(lmcs)))
    kinds = (("inter", abi.InterParams), ("residual", abi.ResidualParams), ("deblock", abi.DeblockParams),
             ("intra", abi.IntraParams), ("ref", abi.ResidualParams))
    if os.path.exists(cache):
        try:
            with open(cache, "rb") as f:
                pic = pickle.load(f)
            for k, cls in kinds:
                pic[k]["params"] = cls.from_buffer_copy(pic[k]["params"])
            return pic
        except Exception:
            pass
    pic = synth.full_picture(seed, width, height, intra_fraction=intra_fraction, lmcs=lmcs)
    try:
        dump = dict(pic)
        for k, _ in kinds:
            dump[k] = dict(pic[k], params=bytes(pic[k]["params"]))
        with open(cache + ".tmp", "wb") as f:
            pickle.dump(dump, f, protocol=4)
        os.replace(cache + ".tmp", cache)
    except OSError:
        pass
    return pic


def picture_stats(pic):
    """Samples and algorithmic bytes (SURVEY.md §8d / DESIGN.md §5) of one picture, per stage and in total."""
    S = sum(w * h for (w, h) in pic["dims"])
    S_y, S_c = pic["dims"][0][0] * pic["dims"][0][1], pic["dims"][1][0] * pic["dims"][1][1]
    tbs = pic["residual"]["tbs"]
    area = (1 << (tbs["log2_w"].astype(np.int64) + tbs["log2_h"]))
    coded = int(area.sum() + area[(tbs["flags"] & 64) != 0].sum())            # joint TBs rebuild a second plane
    b = pic["inter"]["blocks"]
    lists = np.where(b["pred_flag"] == 3, 2, 1).astype(np.int64)
    # luma + 2 x quarter-size chroma; affine units come as luma-only 4x4 and chroma-only (8x8 footprint) records
    planes_x2 = np.where(b["flags"] & 8, 2, np.where(b["flags"] & 16, 1, 3)).astype(np.int64)
    ref_samples = int((b["w"].astype(np.int64) * b["h"] * lists * planes_x2).sum() // 2)
    ib = pic["intra"]["blocks"]
    iw = np.where((ib["flags"] & 4) != 0, ib["cb_w"], ib["w"]).astype(np.int64)      # ISP records stand for the coding block
    ih = np.where((ib["flags"] & 4) != 0, ib["cb_h"], ib["h"]).astype(np.int64)
    cclm = ib["mode"] >= 81
    intra_samples = int((iw * ih * np.where(cclm, 2, 1)).sum())
    intra_bytes = int((2 * iw * ih * np.where(cclm, 2, 1)                                # samples out
                       + np.where((ib["flags"] & 16) != 0, 2 * iw * ih * np.where(cclm, 2, 1), 0)   # int16 residual in
                       + 2 * 2 * (iw + ih) * np.where(cclm, 2, 1)                        # reference edges in
                       + np.where(cclm, 8 * iw * ih, 0)).sum())                          # co-located luma for CCLM
    stage = {
        "inter": 2 * ref_samples + 2 * S,             # reference samples in, prediction out
        "residual": 2 * int(area.sum()) + 4 * coded,  # int16 levels in, destination read + write
        "intra": intra_bytes,
        "deblock": 2 * S + 2 * S,                     # one pass (inverse LMCS + both edge directions): picture in + picture out
        "sao_alf": 2 * S + 2 * S,                     # one pass (SAO + ALF + CC-ALF): picture in + picture out
    }
    a_pic = 2 * int(area.sum()) + 2 * ref_samples + 2 * S   # §8d: levels in + reference reads + final picture out
    return dict(S=S, coded_samples=coded, ref_samples=ref_samples, stage_bytes=stage, a_pic=a_pic, rho=coded / S,
                r=ref_samples / S, intra_share=intra_samples / S)


# ----------------------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if t < t0 or t > t1 + 0.1:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own C on host threads, same stage chain
# ----------------------------------------------------------------------------------------------------------
def cpu_picture(pic, threads):
    """One whole picture through the reference's stage functions, every stage spread over `threads` host threads:
    INTER sub-blocks / coding units dealt round-robin, RECON (every residual + intra prediction,
    ff_vvc_reconstruct) as a CTU-row wavefront like the reference's own scheduler, LMCS / loop filters by CTB rows with a
    join per pass.  Returns seconds, kind, planes."""
    from ffmpeg_b200 import abi, lib
    from oracle import portapi, refapi
    it, rs, ia, db, lf = pic["inter"], pic["residual"], pic["intra"], pic["deblock"], pic["lf"]
    fwd, inv = pic["lmcs"] if pic["lmcs"] is not None else (None, None)
    planes = [np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]]
    use_ref = refapi.available()

    def spread(fn, parts):
        ths = [threading.Thread(target=fn, args=(p,)) for p in parts if len(p)]
        for th in ths:
            th.start()
        for th in ths:
            th.join()

    special = abi.PB_GPM | abi.PB_AFFINE | abi.PB_LUMA_ONLY | abi.PB_CHROMA_ONLY | abi.PB_CIIP
    regular = it["blocks"][(it["blocks"]["flags"] & special) == 0] if use_ref else it["blocks"]
    blocks = [np.ascontiguousarray(regular[t::threads]) for t in range(threads)]
    cus = [np.ascontiguousarray(it["cus"][t::threads]) for t in range(threads)]
    edge = None if use_ref else lib.pack_deblock_edges(db["bs"], db["len_p"], db["len_q"], db["tb_w_c"], db["tb_h_c"],
                                                       pic["ctb_log2"])
    prof = it["prof"] if len(it["prof"]) else None
    t0 = time.perf_counter()
    if use_ref:
        spread(lambda b: refapi.inter_batch(planes, it["refs"], b, it["params"], fwd), blocks)
        spread(lambda c: refapi.inter_cus(planes, it["refs"], c, it["sb_mv"], it["params"], fwd), cus)
        refapi.recon_picture(planes, pic["ref"], ia["params"], threads=threads, inter=it, lmcs_fwd_lut=fwd)
        if inv is not None:
            refapi.lmcs_picture(planes, inv, pic["ctb_log2"], pic.get("bd", 10), threads=threads)
        refapi.deblock_picture(planes, db, threads=threads)
        refapi.sao_picture(planes, lf, threads=threads)
        refapi.alf_picture(planes, lf, threads=threads)
    else:
        spread(lambda b: portapi.inter_batch(planes, it["refs"], b, it["params"], fwd, prof=prof), blocks)
        res = [np.zeros(p.shape, np.int16) for p in planes]
        tbs = [np.ascontiguousarray(rs["tbs"][t::threads]) for t in range(threads)]
        spread(lambda tb: portapi.residual_batch(planes, rs["levels"], tb, rs["params"], residual_planes=res), tbs)
        portapi.intra_picture(planes, res, ia["blocks"], ia["params"])
        if inv is not None:
            portapi.lmcs_picture(planes, inv)
        portapi.deblock_picture(planes, edge, db["qp"], db["ctb_offsets"], db["stride4"], db["params"])
        tmp = [np.zeros_like(p) for p in planes]
        portapi.sao_picture(tmp, planes, lf)
        portapi.alf_picture(planes, tmp, lf)
    return time.perf_counter() - t0, ("reference" if use_ref else "port"), planes


def cpu_baseline(pic, budget_s=15.0, max_pics=30):
    threads = os.cpu_count() or 1
    _, _, planes = cpu_picture(pic, threads)       # warm-up (page-in, first touch); its picture is the parity reference
    n, t, kind = 0, 0.0, "port"
    while t < budget_s and n < max_pics:
        dt, kind, _ = cpu_picture(pic, threads)
        t += dt
        n += 1
    return {"value": n / t, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": "%d whole pictures through INTER + RECON (residual + intra) + LMCS + deblock + SAO + ALF; sub-blocks / CUs "
                      "dealt round-robin, RECON as a CTU-row wavefront, loop-filter CTB rows split over %d host threads "
                      "(x86 asm not built: no nasm)" % (n, threads)}, planes


def emit_host_ms(pic):
    """host time of the descriptor emitter (ffmpeg_b200/emit, compiled against the reference's headers into
    oracle/_ref/libvvcref.so) for this picture: the reference-side state the harness builds from `pic` in, the engine's
    batches out.  One host thread, fresh arrays; the reference runs these stage bodies on all its threads, CTU by CTU."""
    from ffmpeg_b200 import abi, synth
    from oracle import refapi
    if not refapi.available() or pic.get("ref") is None:
        return None
    w, h, it = pic["width"], pic["height"], pic["inter"]
    b = it["blocks"]
    plain = b[(b["flags"] & (abi.PB_AFFINE | abi.PB_GPM | abi.PB_LUMA_ONLY | abi.PB_CHROMA_ONLY)) == 0]
    cus = np.zeros(len(plain), synth.CU_DTYPE)
    for k in ("x0", "y0", "w", "h", "pred_flag", "bcw_idx", "ref_idx", "mv"):
        cus[k] = plain[k]
    cus["flags"] = plain["flags"]
    all_cus = np.concatenate([cus, it["cus"]]) if len(it["cus"]) else cus
    out = {}
    if len(all_cus):
        refapi.emit_inter_cus(w, h, all_cus, it["sb_mv"], it["params"], order=True)
        out["inter"] = round(refapi.emit_last_ms(), 3)
    has_ciip = bool(pic["ref"]["cus"]["ciip_flag"].any())
    refapi.emit_recon_picture(w, h, pic["ref"], pic["intra"]["params"], inter=it if has_ciip else None)
    out["recon"] = round(refapi.emit_last_ms(), 3)
    refapi.emit_loopfilter_ctbs(pic["lf"], w, h, pic["deblock"]["ctb_offsets"], pic["deblock"]["params"])
    out["filters"] = round(refapi.emit_last_ms(), 3)
    out["total"] = round(sum(out.values()), 3)
    out["threads"] = 1
    return out


# ----------------------------------------------------------------------------------------------------------
# workload tx1080p (BASELINE config 1): dequant + inverse transform batch only
# ----------------------------------------------------------------------------------------------------------
def tx_cpu(pic, threads):
    """The reference's itransform() over the whole TB batch, TBs dealt round-robin to `threads` host threads."""
    from oracle import portapi, refapi
    planes = [np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]]
    parts = [(np.ascontiguousarray(pic["tbs"][t::threads]), np.ascontiguousarray(pic["intra_mode"][t::threads]))
             for t in range(threads)]
    use_ref = refapi.available()

    def run(part):
        tbs, im = part
        if use_ref:
            refapi.residual_batch(planes, pic["levels"], tbs, pic["params"], im)
        else:
            portapi.residual_batch(planes, pic["levels"], tbs, pic["params"])
    ths = [threading.Thread(target=run, args=(p,)) for p in parts if len(p[0])]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    return time.perf_counter() - t0, ("reference" if use_ref else "port"), planes


def tx_config(args, width, height):
    return {"workload": "tx1080p: dequant + inverse DCT2/DST7/DCT8 (+ LFNST, TS, BDPCM, scaling lists, joint CbCr) + add over one "
                        "dense %dx%d 4:2:0 10-bit TB tiling (BASELINE config 1), seed %d" % (width, height, args.seed),
            "streams": "one independent batch stream per GPU (replicas, no collective)"}


def bench_tx(args, rank, world, local):
    from ffmpeg_b200 import synth
    width, height = WORKLOADS["tx1080p"]
    pic = synth.residual_picture(args.seed + rank, width, height)
    config = tx_config(args, width, height)
    metric = "1080p dequant + inverse transform batches/s (BASELINE config 1)"
    threads = os.cpu_count() or 1
    if args.impl == "reference":
        for _ in range(args.warmup):
            tx_cpu(pic, threads)
        t = 0.0
        for _ in range(args.steps):
            dt, kind, _ = tx_cpu(pic, threads)
            t += dt
        v = args.steps / t
        print(json.dumps({"metric": metric, "value": v, "unit": "batches/s", "n_gpus": args.gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config, "impl": "reference",
                          "cpu_baseline": {"value": v, "unit": "batches/s", "cores": threads, "kind": kind,
                                           "sample": "each step = the whole TB batch"},
                          "e2e": {"value": v, "unit": "batches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return 0
    import torch
    from ffmpeg_b200 import lib
    from ffmpeg_b200 import hostutil as util
    torch.cuda.set_device(local)
    base = [np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]]
    tbs_np, levels_np = pic["tbs"], pic["levels"]
    per_set = levels_np.nbytes + tbs_np.nbytes + 2 * sum(p.nbytes for p in base)
    rot = max(2, int(np.ceil(2.2 * L2_BYTES / per_set)))
    sets = []
    for _ in range(rot):
        d = util.to_device_planes(base, torch)
        sets.append((d, lib.device_planes(d), torch.from_numpy(levels_np).cuda(),
                     torch.from_numpy(tbs_np.view(np.uint8).reshape(-1)).cuda()))
    n = len(tbs_np)

    def step(i):
        d, dp, lv, tb = sets[i % rot]
        lib.residual_batch(dp, lv, tb, n, pic["params"])
    warmup = max(args.warmup, 3)
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lib.launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    t_host0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        step(warmup + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = lib.launch_count() - l0
    t_end, i = time.perf_counter() + 0.6, 0
    while time.perf_counter() < t_end:          # keep the load up for nvidia-smi's sampling period (not timed)
        step(i)
        i += 1
        if i % 50 == 0:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    clocks = sampler.stop(t_host0, time.perf_counter()) if sampler is not None else None
    # parity: the picture after `steps` adds differs per set; compare one fresh add on a zeroed picture
    d = util.to_device_planes(base, torch)
    lib.residual_batch(lib.device_planes(d), sets[0][2], sets[0][3], n, pic["params"])
    torch.cuda.synchronize()
    got = util.plane_md5(util.from_device_planes(d))
    area = (1 << (tbs_np["log2_w"].astype(np.int64) + tbs_np["log2_h"]))
    coded = int(area.sum() + area[(tbs_np["flags"] & 64) != 0].sum())
    algo = 2 * int(area.sum()) + 4 * coded
    peak, peak_kind = measured_peaks()
    kms = ms / args.steps
    out = {"metric": metric, "value": world * args.steps / (ms / 1e3), "unit": "batches/s", "n_gpus": world, "steps": args.steps,
           "warmup": warmup, "ms_per_step": kms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int32", "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": int(launches),
           "transform_blocks": n,
           "roofline": {"bound": "hbm", "kernel": "residual", "achieved": algo / (kms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": algo / (kms * 1e-3) / 1e9 / peak, "traffic": None, "peak_kind": peak_kind,
                        "algorithmic_bytes_per_launch": algo, "kernel_ms": kms},
           "l2": "rotating %d working sets (%.0f MB) > 126 MB L2" % (rot, rot * per_set / 1e6)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tx_cpu(pic, threads)
        n_cpu, t_cpu, kind, planes = 0, 0.0, "port", None
        while t_cpu < 10.0 and n_cpu < 40:
            dt, kind, planes = tx_cpu(pic, threads)
            t_cpu += dt
            n_cpu += 1
        ref = util.plane_md5(planes)
        out["cpu_baseline"] = {"value": n_cpu / t_cpu, "unit": "batches/s", "cores": threads, "kind": kind,
                               "sample": "%d whole batches, TBs dealt round-robin to %d threads" % (n_cpu, threads)}
        out["parity"] = {"md5_match": got == ref, "ref_md5": ref, "gpu_md5": got}
    if rank == 0:
        emit_line(out)
    return 0


# ----------------------------------------------------------------------------------------------------------
def picture_config(args, width, height, intra_fraction, lmcs):
    """What both arms print as `config`: the workload and nothing arm-specific, so that the two lines compare equal."""
    return {"workload": "%s 4:2:0 10-bit synthetic %s picture (%dx%d), CTU 128, 2 reference pictures, %.0f%% of the coding "
                        "units intra, LMCS %s, seed %d" % (args.workload, "all-intra" if intra_fraction >= 1 else "random-access",
                                                          width, height, 100 * intra_fraction, "on" if lmcs else "off", args.seed),
            "stages": ["INTER (8-tap luma / 4-tap chroma MC, uni / bi / weighted / BCW, DMVR, BDOF, affine + PROF, GPM, forward LMCS)",
                       "residual (dequant, inverse DCT2/DST7/DCT8, LFNST, TS, BDPCM, joint CbCr, scaling lists, add)",
                       "intra (planar / DC / angular + PDPC, MRL, ISP, MIP, CCLM, CIIP, luma-dependent chroma residual scaling; dependency-driven)",
                       "inverse LMCS", "DEBLOCK_V", "DEBLOCK_H", "SAO", "ALF luma + chroma + CC-ALF"],
            "streams": "one independent picture stream per GPU (replicas, no collective)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="4k", choices=sorted(WORKLOADS))
    ap.add_argument("--seed", type=int, default=2026)
    ap.add_argument("--intra-fraction", type=float, default=None, help="share of coding units that are intra (default 0.1; 8k: 1.0)")
    ap.add_argument("--lmcs", type=int, default=1, help="0: leave LMCS out of the picture (development only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.workload == "tx1080p":
        if args.impl == "reference" and rank != 0:
            return 0
        return bench_tx(args, rank, world, local)
    width, height = WORKLOADS[args.workload]
    intra_fraction = args.intra_fraction if args.intra_fraction is not None else (1.0 if args.workload == "8k" else 0.1)
    lmcs = bool(args.lmcs)
    warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    config = picture_config(args, width, height, intra_fraction, lmcs)

    if args.impl == "reference":
        if rank != 0:
            return 0
        pic = make_picture(args.seed, width, height, intra_fraction, lmcs)
        threads = os.cpu_count() or 1
        for _ in range(args.warmup):
            cpu_picture(pic, threads)
        t, kind = 0.0, "port"
        for _ in range(args.steps):
            dt, kind, _ = cpu_picture(pic, threads)
            t += dt
        v = args.steps / t
        out = {"metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config, "impl": "reference",
               "cpu": "reference C built from its own sources (no x86 asm: no nasm on this image), %d host threads" % threads,
               "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind,
                                "sample": "each step = one whole picture through the full stage chain"},
               "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0}
        emit_line(out)
        return 0

    import torch
    import torch.distributed as dist
    from ffmpeg_b200 import abi, lib
    from ffmpeg_b200 import hostutil as util          # nothing on this arm imports oracle/
    pipeline = util
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: the engine has no CPU fallback")
    bound = bind_near_gpu(local)
    torch.cuda.set_device(local)
    if world > 1:
        stdout_for_the_json_line_only()
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    pic = make_picture(args.seed + rank, width, height, intra_fraction, lmcs)
    st = picture_stats(pic)
    it, rs, ia, db, lf = pic["inter"], pic["residual"], pic["intra"], pic["deblock"], pic["lf"]
    fwd, inv = pic["lmcs"] if pic["lmcs"] is not None else (None, None)
    frame_bytes = 2 * st["S"]
    ctb = pic["ctb_log2"]
    irecs, phases = lib.intra_schedule(ia["blocks"], ia["ctu_first"], (width + (1 << ctb) - 1) >> ctb, (height + (1 << ctb) - 1) >> ctb,
                                        ctb, width, height, order=lib.INTRA_ORDER_DIAGONAL, phases=True)   # as vvcb200_pic_intra does
    iparams = lib.phased_params(ia["params"], phases)
    detail = {"transform_blocks": int(len(rs["tbs"])), "pred_blocks": int(len(it["blocks"])),
              "intra_records": int(len(irecs)), "intra_phases": {"luma": phases[0], "chroma_add_only": phases[1],
                                                                  "chroma": int(len(irecs)) - phases[0] - phases[1]},
              "intra_sample_share": round(st["intra_share"], 3),
              "coded_fraction": round(st["rho"], 3), "ref_samples_per_sample": round(st["r"], 3)}

    # ---- device-resident arm: `rot` complete copies of the picture's working set, visited round-robin -----------
    edge = lib.pack_deblock_edges(db["bs"], db["len_p"], db["len_q"], db["tb_w_c"], db["tb_h_c"], pic["ctb_log2"])

    def dev(a):
        return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()

    per_set = (len(it["refs"]) + 3) * frame_bytes + irecs.nbytes + rs["levels"].nbytes + rs["tbs"].nbytes + it["blocks"].nbytes + \
        edge.nbytes + db["qp"].nbytes
    rot = max(2, int(np.ceil(2.2 * L2_BYTES / per_set)))
    n_lanes = env_int("VVCB200_BENCH_LANES", 2)
    rot = -(-rot // n_lanes) * n_lanes               # a multiple of the lanes: a working set always meets the same lane (stream)
    detail["host_cores"] = ("rank bound to the %d cores of its GPU's NUMA node (pinned memory is then local to the GPU)" % len(bound)
                            if bound else "not bound (one NUMA node, or no NVML)")
    detail["l2"] = "rotating %d complete working sets (%.0f MB, each with its own reference pictures) > 126 MB L2" % (
        rot, rot * per_set / 1e6)
    detail["lanes"] = ("consecutive pictures alternate between two CUDA streams (independent pictures, as the B pictures of "
                       "one temporal layer): the latency-bound intra stage of one overlaps the kernels of the other; "
                       "stage_ms come from a separate single-stream pass")
    n_blocks, n_tbs = len(it["blocks"]), len(rs["tbs"])
    sets = []
    for k in range(rot):
        s = {}
        s["refs"] = [util.to_device_planes(r, torch) for r in it["refs"]]
        s["cur"] = util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch)
        s["tmp"] = [torch.zeros_like(t) for t in s["cur"]]
        s["res"] = [torch.zeros_like(t) for t in s["cur"]]
        s["cur_p"], s["tmp_p"], s["res_p"] = lib.device_planes(s["cur"]), lib.device_planes(s["tmp"]), lib.device_planes(s["res"])
        s["irecs"] = dev(irecs) if len(irecs) else None
        s["prof"] = dev(it["prof"]) if len(it["prof"]) else None
        s["marks"] = torch.zeros(lib.intra_scratch_bytes(width, height), dtype=torch.uint8, device="cuda")
        s["refs_p"] = [lib.device_planes(r) for r in s["refs"]]
        s["blocks"] = dev(it["blocks"]) if n_blocks else None
        s["levels"], s["tbs"] = dev(rs["levels"]), dev(rs["tbs"])
        s["edge"] = torch.from_numpy(edge).cuda()
        s["qp"] = torch.from_numpy(db["qp"]).cuda()
        s["off"] = torch.from_numpy(db["ctb_offsets"]).cuda()
        s["db"] = lib.deblock_tables(s["edge"], s["qp"], s["off"], db["stride4"])
        s["sao"], s["alf"], s["alfp"] = dev(lf["sao"]), dev(lf["alf"]), dev(lf["alf_params"])
        s["alf_table"] = torch.zeros(lib.sao_alf_scratch_bytes(), dtype=torch.uint8, device="cuda")
        s["fwd"] = dev(fwd) if fwd is not None else None
        s["inv"] = dev(inv) if inv is not None else None
        s["ev_luma"] = torch.cuda.Event()
        s["ev_luma"].record()                            # (torch creates the CUDA event on its first record)
        assert s["ev_luma"].cuda_event
        s["iparams"] = abi.IntraParams.from_buffer_copy(bytes(iparams))
        s["iparams"].prepared = 1
        s["iparams"].luma_done_event = s["ev_luma"].cuda_event
        sets.append(s)

    n_lanes = env_int("VVCB200_BENCH_LANES", 2)      # development: pictures in flight in the device-resident arm
    lanes = [torch.cuda.Stream() for _ in range(n_lanes)]

    # Inside a picture two things need not wait for the stage before them: the intra pre-pass (sentinels, in-place
    # predictions moved aside) needs the inter prediction only and runs beside the residual kernels; luma deblocking needs
    # final luma only and runs beside the chroma phases of the intra stage.  One helper stream per lane carries them.
    # It shortens a picture on its own a little (0.766 -> 0.749 ms); with two pictures in flight the neighbour's kernels fill
    # those gaps already (0.598 -> 0.602 ms), so the throughput loop does without.
    split_mode = env_int("VVCB200_BENCH_SPLIT", 1)   # 0 never, 1 the one-picture-at-a-time pass only, 2 always
    helpers = [torch.cuda.Stream() for _ in range(n_lanes + 1)]
    detail["in_picture_overlap"] = ("one picture at a time: intra pre-pass beside the residual kernels, luma deblocking beside the "
                                    "intra stage's chroma phases (helper stream); two pictures in flight: %s"
                                    % ("the same" if split_mode == 2 else "stage after stage"))

    def step(i, ev=None, lane=None):
        """One picture.  lane: the stream it is queued on (None: torch's current stream, with stage events)."""
        split = len(irecs) > 0 and n_blocks > 0 and (split_mode == 2 or (split_mode == 1 and lane is None))   # (as Level B: pictures with inter units)
        s = sets[i % rot]
        q = lane.cuda_stream if lane is not None else None
        main = lane if lane is not None else torch.cuda.current_stream()
        side = helpers[(i % n_lanes) if lane is not None else n_lanes]
        if n_blocks:
            lib.inter_batch(s["cur_p"], s["refs_p"], s["blocks"], n_blocks, it["params"], lmcs_fwd_lut=s["fwd"], prof=s["prof"], stream=q)
        if ev is not None:
            ev[1].record()
        if split:
            side.wait_stream(main)
            lib.intra_prepare(s["cur_p"], s["res_p"], s["irecs"], len(irecs), iparams, s["marks"], shadow=s["tmp_p"],
                              stream=side.cuda_stream)
        lib.residual_batch(s["cur_p"], s["levels"], s["tbs"], n_tbs, rs["params"], residual=s["res_p"], stream=q)
        if ev is not None:
            ev[2].record()
        if split:
            main.wait_stream(side)
            lib.intra_picture(s["cur_p"], s["res_p"], s["irecs"], len(irecs), s["iparams"], s["marks"], shadow=s["tmp_p"], stream=q)
            if ev is not None:
                ev[3].record()
            side.wait_event(s["ev_luma"])
            lib.deblock_lmcs_planes(s["tmp_p"], s["cur_p"], s["db"], db["params"], s["inv"], planes=1, stream=side.cuda_stream)
            lib.deblock_lmcs_planes(s["tmp_p"], s["cur_p"], s["db"], db["params"], s["inv"], planes=2, stream=q)
            main.wait_stream(side)
        else:
            if len(irecs):
                lib.intra_picture(s["cur_p"], s["res_p"], s["irecs"], len(irecs), iparams, s["marks"], shadow=s["tmp_p"], stream=q)
            if ev is not None:
                ev[3].record()
            lib.deblock_lmcs_picture(s["tmp_p"], s["cur_p"], s["db"], db["params"], s["inv"], stream=q)
        if ev is not None:
            ev[4].record()
        lib.sao_alf_picture(s["cur_p"], s["tmp_p"], s["sao"], s["alf"], s["alfp"], pic["ctb_log2"], scratch=s["alf_table"], stream=q)
        if ev is not None:
            ev[5].record()

    for i in range(warmup):
        step(i, lane=lanes[i % n_lanes])
    barrier()
    # stage boundaries are marked with events on the launching stream (torch's current stream), one picture at a time
    # (the first pictures after the barrier find the GPU idle and measure how fast the host submits, not how long the
    # kernels run: three extra pictures are queued first and left out)
    n_serial = max(1, min(args.steps, 10))
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(len(STAGES) + 1)] for _ in range(n_serial + 3)]
    for i in range(n_serial + 3):
        evs[i][0].record()
        step(i, evs[i])
    barrier()
    evs = evs[3:]
    stage_ms = {name: float(np.mean([evs[i][k].elapsed_time(evs[i][k + 1]) for i in range(n_serial)]))
                for k, name in enumerate(STAGES)}
    serial_ms = float(np.mean([evs[i][0].elapsed_time(evs[i][len(STAGES)]) for i in range(n_serial)]))
    detail["serial_ms_per_picture"] = round(serial_ms, 4)
    sampler = ClockSampler(local) if rank == 0 else None
    ev_start, ev_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lib.launch_count()
    t_host0 = time.perf_counter()
    ev_start.record()                                # the timed region: from here on torch's current stream ...
    for q in lanes:
        q.wait_event(ev_start)
    for i in range(args.steps):
        step(warmup + i, lane=lanes[(warmup + i) % n_lanes])
    for q in lanes:
        torch.cuda.current_stream().wait_stream(q)   # ... until both lanes have drained
    ev_end.record()
    barrier()
    t_host1 = time.perf_counter()
    launches = lib.launch_count() - l0
    ms = ev_start.elapsed_time(ev_end)
    if sampler is not None and t_host1 - t_host0 < 0.5:
        # too short for nvidia-smi's sampling period: keep the same load running a little longer (not timed)
        t_end, i = time.perf_counter() + 0.6, 0
        while time.perf_counter() < t_end:
            step(i, lane=lanes[i % n_lanes])
            i += 1
            if i % 20 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        t_host1 = time.perf_counter()
    clocks = sampler.stop(t_host0, t_host1) if sampler is not None else None
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * args.steps / (ms / 1e3)
    step(0)                                          # one more picture on the current stream: its samples for the parity check
    torch.cuda.synchronize()
    dev_md5 = util.plane_md5(util.from_device_planes(sets[0]["cur"]))
    del sets
    torch.cuda.empty_cache()

    # ---- end-to-end arm: Level-B C ABI, pinned host buffers, H2D + D2H inside the timed region ------------------
    ctx = lib.Context(local, width, height, dpb_size=len(it["refs"]) + 4,   # four slots: downloads overlap the next pictures
                      validate=1)                                         # the library's default batch checks
    ep = pipeline.EnginePicture(ctx, pic)
    ep.upload_refs()                       # reference pictures are earlier outputs of the same decoder: resident
    e2e_steps = max(60, min(args.steps, 100))         # (a pipeline of four pictures: fewer than that measures its fill and drain)

    def e2e_run(n, **kw):
        for _ in range(8):                   # twice round the context's four descriptor sets and four picture slots
            ep.run(**kw)
        ctx.sync()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            ep.run(**kw)
        ctx.sync()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return dt

    e2e_s = e2e_run(e2e_steps)
    e2e_ok = util.plane_md5(ep.out)
    e2e_same = all(util.plane_md5(o) == e2e_ok for o in ep.outs)
    # what bounds it: the same loop without the picture download, and the download alone
    no_dl_s = e2e_run(e2e_steps, download=False)
    dl_s = e2e_run(e2e_steps, only_download=True)
    h2d, d2h = ep.h2d_bytes(), ep.d2h_bytes()
    ms_full, ms_nodl, ms_dl = 1e3 * e2e_s / e2e_steps, 1e3 * no_dl_s / e2e_steps, 1e3 * dl_s / e2e_steps
    bound = max((("picture download (D2H)", ms_dl), ("uploads + kernels + host calls", ms_nodl)), key=lambda kv: kv[1])[0]
    e2e = {"value": world * e2e_steps / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "steps": e2e_steps,
           "ms_per_picture": round(ms_full, 4),
           "ms_per_picture_without_download": round(ms_nodl, 4),
           "ms_per_picture_download_alone": round(ms_dl, 4),
           "d2h_gbs_alone": round(ep.picture_bytes() / (ms_dl * 1e-3) / 1e9, 2),
           "achieved_h2d_gbs": round(h2d / (ms_full * 1e-3) / 1e9, 2), "achieved_d2h_gbs": round(d2h / (ms_full * 1e-3) / 1e9, 2),
           "bound": bound,
           "api": "vvcb200_pic_begin/inter/residual/intra/lmcs/deblock/sao_alf/download (include/vvc_b200.h), pinned host buffers, "
                  "refined motion vectors read back every picture; reference pictures stay in the context's DPB; consecutive "
                  "pictures rotate over four DPB slots and four host frames and alternate between the context's two compute "
                  "lanes, so uploads / downloads (copy streams) and the kernels of neighbouring pictures overlap"}
    ctx.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_kind = measured_peaks()
    top = max(stage_ms, key=stage_ms.get)
    # INTER (3 width classes), residual (5 block classes) and deblocking (luma, chroma) launch their grids side by side on
    # forked streams: each stage is timed, and its bytes counted, as one unit
    kernel_ms = stage_ms[top]
    algo = st["stage_bytes"][top]
    achieved = algo / (kernel_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get(args.workload, {}).get(top)
    except Exception:
        pass
    step_ms = ms / args.steps
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_kind": peak_kind,
                "algorithmic_bytes_per_launch": algo, "kernel_ms": kernel_ms,
                "stage_ms": {k: round(v, 4) for k, v in stage_ms.items()},
                "stage_frac_of_peak": {k: round(st["stage_bytes"][k] / (stage_ms[k] * 1e-3) / 1e9 / peak, 4) for k in STAGES
                                       if stage_ms[k] > 0.01},          # (a stage without work has no rate)
                "picture": {"algorithmic_bytes": st["a_pic"], "achieved": st["a_pic"] / (step_ms * 1e-3) / 1e9,
                            "frac": st["a_pic"] / (step_ms * 1e-3) / 1e9 / peak}}
    out = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
           "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int32", "data": "synthetic", "config": config, "detail": detail, "clocks": clocks, "e2e": e2e,
           "gpu_launches": int(launches), "roofline": roofline, "e2e_plane_md5": e2e_ok,
           "e2e_frames_identical": bool(e2e_same)}
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"], ref_planes = cpu_baseline(pic, budget_s=10.0 if args.workload != "8k" else 1.0)
        if args.workload != "8k":
            out["cpu_baseline"]["emit_host_ms"] = emit_host_ms(pic)
        ref_md5 = util.plane_md5(ref_planes)
        out["parity"] = {"md5_match": bool(ref_md5 == dev_md5 and ref_md5 == e2e_ok and e2e_same), "ref_md5": ref_md5,
                         "gpu_md5": dev_md5, "e2e_md5": e2e_ok,
                         "checker": "per-plane MD5 of the reference C's picture (the cpu_baseline leg) against the device-resident "
                                    "arm's and the end-to-end arm's"}
    emit_line(out)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
