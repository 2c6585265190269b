#!/usr/bin/env python
"""bench.py — VVC reconstruction + in-loop-filter throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload 4k|1080p]

One "step" = one synthetic picture through every stage the engine implements (see `config.stages`).
`value`   : pictures/s, device-resident inputs, CUDA-event timed, max over ranks (N independent streams).
`e2e`     : same picture through the Level-B C ABI with HOST (pinned) buffers, H2D + D2H inside the timed region.
`roofline`: dominant kernel's algorithmic bytes / its live CUDA-event duration vs MEASURED_PEAKS.json.
`cpu_baseline` / `--impl reference`: the reference's own C (oracle/_ref) — or the oracle port when that is
not built — on the box's host cores, all threads, same picture.
Prints exactly one JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "4K 10-bit VVC recon+loop-filter frames/s per B200; % HBM roofline; 8-GPU agg"
WORKLOADS = {"4k": (3840, 2160), "1080p": (1920, 1080)}
L2_BYTES = 126 * 1024 * 1024


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
def make_picture(seed, width, height):
    """Synthetic parsed picture (dense residual, config-1 level model).  Cached under /tmp for re-runs."""
    from ffmpeg_b200 import abi, synth
    cache = "/tmp/vvcb200_bench_%dx%d_%d.npz" % (width, height, seed)
    if os.path.exists(cache):
        z = np.load(cache)
        params = abi.ResidualParams.from_buffer_copy(z["params"].tobytes())
        return dict(tbs=z["tbs"].view(abi.TB_DTYPE).reshape(-1), levels=z["levels"], params=params,
                    intra_mode=z["intra_mode"], dims=[tuple(d) for d in z["dims"]], seed=seed)
    pic = synth.residual_picture(seed, width, height, mean_log2=4.3)
    try:
        np.savez(cache, tbs=pic["tbs"].view(np.uint8), levels=pic["levels"], intra_mode=pic["intra_mode"],
                 params=np.frombuffer(bytes(pic["params"]), np.uint8), dims=np.array(pic["dims"], np.int32))
    except OSError:
        pass
    return pic


def level_variant(levels, k):
    """Picture k of the rotation: same partition, different levels (sign pattern) — defeats L2 reuse."""
    if k == 0:
        return levels
    rng = np.random.default_rng(1000 + k)
    sign = rng.integers(0, 2, levels.size).astype(np.int16) * 2 - 1
    return (levels * sign).astype(np.int16)


def residual_algorithmic_bytes(pic):
    """SURVEY.md §8(d), stage-isolated residual: 2 B/sample int16 levels in + 2 B/sample written, plus the
    destination sample read of the in-place add (2 B) — per coded sample."""
    n = int(np.sum(1 << (pic["tbs"]["log2_w"].astype(np.int64) + pic["tbs"]["log2_h"])))
    joint = pic["tbs"]["flags"] & 64
    n_joint = int(np.sum((1 << (pic["tbs"]["log2_w"].astype(np.int64) + pic["tbs"]["log2_h"]))[joint != 0]))
    return 2 * n + 2 * (n + n_joint), n


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
# CPU arm: the reference's own C on host threads
# ----------------------------------------------------------------------------------------------------------
def cpu_run(pic, planes, threads):
    """One whole picture through the reference's residual path, TBs dealt round-robin to `threads` threads."""
    from oracle import portapi, refapi
    use_ref = refapi.available()
    kind = "reference" if use_ref else "port"
    parts = []
    for t in range(threads):
        sub = dict(pic, tbs=np.ascontiguousarray(pic["tbs"][t::threads]),
                   intra_mode=np.ascontiguousarray(pic["intra_mode"][t::threads]))
        parts.append(sub)

    def work(sub):
        if use_ref:
            refapi.residual_batch(planes, sub["levels"], sub["tbs"], sub["params"], sub["intra_mode"])
        else:
            portapi.residual_batch(planes, sub["levels"], sub["tbs"], sub["params"])
    ths = [threading.Thread(target=work, args=(p,)) for p in parts]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    return time.perf_counter() - t0, kind


def cpu_baseline(pic, budget_s=12.0):
    from tests import util
    threads = os.cpu_count() or 1
    planes = util.noise_planes(1, pic["dims"])
    cpu_run(pic, planes, threads)                      # warm-up (page-in, first-touch)
    n, t = 0, 0.0
    while t < budget_s and n < 50:
        dt, kind = cpu_run(pic, planes, threads)
        t += dt
        n += 1
    return {"value": n / t, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": "%d whole pictures (%d TBs each) through the residual stage, TBs dealt round-robin to %d host threads"
                      % (n, len(pic["tbs"]), threads)}


# ----------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="4k", choices=sorted(WORKLOADS))
    ap.add_argument("--seed", type=int, default=2026)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    width, height = WORKLOADS[args.workload]
    warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    config = {"workload": "%s 4:2:0 10-bit synthetic parsed picture (%dx%d), CTU 128" % (args.workload, width, height),
              "stages": ["dequant", "bdpcm", "lfnst", "inverse DCT2/DST7/DCT8", "add_residual", "joint CbCr"],
              "streams": "one independent picture stream per GPU (replicas, no collective)"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        pic = make_picture(args.seed, width, height)
        from tests import util
        threads = os.cpu_count() or 1
        planes = util.noise_planes(1, pic["dims"])
        for _ in range(args.warmup):
            cpu_run(pic, planes, threads)
        t, kind = 0.0, "port"
        for _ in range(args.steps):
            dt, kind = cpu_run(pic, planes, threads)
            t += dt
        v = args.steps / t
        config["cpu"] = "reference C (--disable-x86asm equivalent: no nasm on this image), %d host threads" % threads
        out = {"metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config, "impl": "reference",
               "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind,
                                "sample": "each step = one whole picture's residual stage (%d TBs)" % len(pic["tbs"])},
               "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0}
        print(json.dumps(out))
        return 0

    import torch
    import torch.distributed as dist
    from ffmpeg_b200 import abi, lib
    from tests import util
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    pic = make_picture(args.seed + rank, width, height)
    n_tb = len(pic["tbs"])
    algo_bytes, n_samples = residual_algorithmic_bytes(pic)
    frame_bytes = sum(w * h * 2 for (w, h) in pic["dims"])
    per_pic_dev = pic["levels"].nbytes + pic["tbs"].nbytes + frame_bytes
    rot = max(2, int(np.ceil(2.2 * L2_BYTES / per_pic_dev)))     # rotation set > 2x L2
    config["l2"] = "rotating %d distinct pictures (%.0f MB) > 126 MB L2 between timed iterations" % (rot, rot * per_pic_dev / 1e6)
    config["transform_blocks"] = n_tb

    # ---- device-resident arm --------------------------------------------------------------------------
    base = util.noise_planes(args.seed, pic["dims"])
    tbs_d = torch.from_numpy(pic["tbs"].view(np.uint8).reshape(-1)).cuda()
    pics_d = []
    for k in range(rot):
        planes = util.to_device_planes(base, torch)
        pics_d.append((lib.device_planes(planes), planes, torch.from_numpy(level_variant(pic["levels"], k)).cuda()))

    def step(i):
        pl, _, lv = pics_d[i % rot]
        lib.residual_batch(pl, lv, tbs_d, n_tb, pic["params"])

    for i in range(warmup):
        step(i)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lib.launch_count()
    t_host0 = time.perf_counter()
    ev0.record()
    for i in range(args.steps):
        step(warmup + i)
    ev1.record()
    barrier()
    t_host1 = time.perf_counter()
    launches = lib.launch_count() - l0
    ms = ev0.elapsed_time(ev1)
    # keep the GPU under the same load a little longer if the timed region was too short for nvidia-smi
    if sampler is not None and t_host1 - t_host0 < 0.5:
        t_end = time.perf_counter() + 0.6
        i = 0
        while time.perf_counter() < t_end:
            step(i)
            i += 1
            if i % 50 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        t_host1 = time.perf_counter()
    clocks = sampler.stop(t_host0, t_host1) if sampler is not None else None
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * args.steps / (ms / 1e3)

    # ---- end-to-end arm: Level-B C ABI, pinned host buffers, H2D + D2H inside the timed region ----------
    ctx = lib.Context(local, width, height)
    h_src, h_out = ctx.pinned_planes(), ctx.pinned_planes()
    for s, b in zip(h_src, base):
        s[...] = b
    h_levels = ctx.pinned(pic["levels"].shape, np.int16)
    h_levels[...] = pic["levels"]
    h_tbs = ctx.pinned(pic["tbs"].shape, abi.TB_DTYPE)
    h_tbs[...] = pic["tbs"]

    def e2e_step():
        ctx.begin()
        ctx.upload(h_src)
        ctx.residual(h_levels, h_tbs, pic["params"])
        ctx.download(h_out)

    e2e_steps = max(3, min(args.steps, 50))
    for _ in range(3):
        e2e_step()
    ctx.sync()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    ctx.sync()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = {"value": world * e2e_steps / e2e_s, "unit": "frames/s",
           "h2d_bytes_per_step": int(frame_bytes + pic["levels"].nbytes + pic["tbs"].nbytes),
           "d2h_bytes_per_step": int(frame_bytes), "steps": e2e_steps,
           "api": "vvcb200_pic_upload + vvcb200_pic_residual + vvcb200_pic_download (include/vvc_b200.h), pinned host buffers"}
    ctx.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_kind = measured_peaks()
    kernel_ms = ms_rank0 = ev0.elapsed_time(ev1) / args.steps        # one launch per step: the step IS the kernel
    achieved = algo_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get(args.workload)
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "residual_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_kind": peak_kind,
                "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": kernel_ms}
    out = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int32", "data": "synthetic", "config": config, "clocks": clocks, "e2e": e2e,
           "gpu_launches": int(launches), "roofline": roofline}
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(pic)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
