#!/usr/bin/env python
"""Generate tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref/libvvcref.so).

The reference ships no DSP-level golden vectors (checkasm compares variants against its own C code,
SURVEY.md §4), so the pin is: outputs of the reference's own C run here on seeded synthetic inputs.
Fixtures are small (a few hundred KB), committed, and re-checkable: `python tools/gen_golden.py --check`.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ffmpeg_b200 import abi, synth  # noqa: E402
from oracle import refapi  # noqa: E402
from tests import util  # noqa: E402


def golden_itx():
    """Every populated itx.itx[trh][trv][log2w][log2h] slot on one seeded block (checkasm vvc_itx.c:58-92 shapes)."""
    rng = np.random.default_rng(20261106)
    keys, digests = [], []
    for trh in range(3):
        for trv in range(3):
            for l2w in range(7):
                for l2h in range(7):
                    w, h = 1 << l2w, 1 << l2h
                    nzw = int(rng.integers(1, min(w, 32 if trh == 0 else 16) + 1))
                    nzh = int(rng.integers(1, min(h, 32 if trv == 0 else 16) + 1))
                    c = np.zeros((h, w), np.int32)
                    c[:nzh, :nzw] = rng.integers(-32768, 32768, (nzh, nzw))
                    flat = c.reshape(-1).copy()
                    if not refapi.itx(flat, trh, trv, l2w, l2h, nzw, nzh):
                        continue
                    keys.append((trh, trv, l2w, l2h, nzw, nzh))
                    digests.append(util.md5(flat))
    return dict(keys=np.array(keys, np.int32), md5=np.array(digests), seed=20261106)


def golden_residual():
    out = {}
    for name, seed, w, h, levels in (("cif_realistic", 11, 352, 288, "realistic"), ("wqvga_full", 12, 416, 240, "full")):
        pic = synth.residual_picture(seed, w, h, levels=levels)
        planes = util.noise_planes(seed, pic["dims"])
        res = refapi.residual_batch(planes, pic["levels"], pic["tbs"], pic["params"], pic["intra_mode"], True)
        out[name + "_plane_md5"] = np.array(util.plane_md5(planes))
        out[name + "_residual_md5"] = np.array(util.md5(res))
        out[name + "_args"] = np.array([seed, w, h, 0 if levels == "realistic" else 1], np.int32)
    return out


def golden_inter():
    out = {}
    for name, seed, w, h, n_refs in (("a", 21, 352, 288, 2), ("b", 22, 416, 240, 4)):
        pic = synth.inter_picture(seed, w, h, n_refs=n_refs)
        planes = util.noise_planes(seed, pic["dims"])
        refapi.inter_batch(planes, pic["refs"], pic["blocks"], pic["params"])
        out[name + "_plane_md5"] = np.array(util.plane_md5(planes))
        out[name + "_args"] = np.array([seed, w, h, n_refs], np.int32)
    return out


def golden_deblock():
    out = {}
    for name, seed, w, h, ctb in (("a", 31, 352, 288, 7), ("b", 32, 416, 240, 6)):
        t = synth.deblock_picture(seed, w, h, ctb_log2=ctb)
        planes = synth.smooth_planes(seed, t["dims"])
        refapi.deblock_picture(planes, t)
        out[name + "_plane_md5"] = np.array(util.plane_md5(planes))
        out[name + "_args"] = np.array([seed, w, h, ctb], np.int32)
    return out


def golden_loopfilter():
    """SAO then ALF chained (the ALF input is the SAO output), both through the reference's stage functions."""
    out = {}
    for name, seed, w, h, ctb, ns, lfase in (("a", 41, 352, 288, 7, 1, 1), ("b", 42, 416, 240, 5, 5, 0)):
        lf = synth.loopfilter_ctbs(seed, w, h, ctb_log2=ctb, n_slices=ns, lfase=lfase)
        planes = synth.smooth_planes(seed, lf["dims"])
        refapi.sao_picture(planes, lf)
        out[name + "_sao_md5"] = np.array(util.plane_md5(planes))
        refapi.alf_picture(planes, lf)
        out[name + "_alf_md5"] = np.array(util.plane_md5(planes))
        out[name + "_args"] = np.array([seed, w, h, ctb, ns, lfase], np.int32)
    return out


def golden_intra():
    """One small picture through the reference's ff_vvc_reconstruct (inter CUs predicted first), planes kept whole."""
    seed, w, h, ctb, frac = 20261107, 136, 72, 5, 0.7
    tools = ("pdpc", "mrl", "mip", "cclm", "mts", "ts", "dep_quant", "scaling", "joint", "lfnst", "bdpcm", "isp", "ciip")
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=tools, intra_fraction=frac)
    planes = [np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]]
    util.oracle_inter(planes, pic["inter"], None, True)
    refapi.recon_picture(planes, pic["ref"], pic["intra"]["params"], inter=pic["inter"])
    return dict(seed=seed, width=w, height=h, ctb_log2=ctb, tools=np.array(tools), intra_fraction=frac,
                blocks_md5=np.array(util.md5(pic["intra"]["blocks"])),
                plane0=planes[0], plane1=planes[1], plane2=planes[2])


def main():
    check = "--check" in sys.argv
    os.makedirs(util.GOLDEN, exist_ok=True)
    for name, fn in (("itx", golden_itx), ("residual", golden_residual), ("inter", golden_inter), ("deblock", golden_deblock), ("loopfilter", golden_loopfilter),
                     ("intra", golden_intra)):
        data = fn()
        path = os.path.join(util.GOLDEN, name + ".npz")
        if check:
            old = np.load(path)
            for k in data:
                assert np.array_equal(old[k], data[k]), (name, k)
            print("ok", path)
        else:
            np.savez_compressed(path, **data)
            print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
