"""GPU parity: the CUDA residual path (through the C ABI) against the oracle on identical seeded inputs.
Bit-exact (integer path): compared sample for sample and by per-plane MD5."""
import numpy as np
import pytest

from ffmpeg_b200 import abi, lib, synth
from tests import util

pytestmark = pytest.mark.gpu


def _params_for(pic):
    """ResidualParams with the class counts of exactly the TBs being submitted (subsets stay class-ordered)."""
    p = abi.ResidualParams.from_buffer_copy(bytes(pic["params"]))
    cls = abi.residual_class(pic["tbs"])
    assert np.all(np.diff(cls) >= 0)
    for c in range(4):
        p.n_class[c] = int(np.sum(cls == c))
    return p


def _run_gpu(pic, planes_np, to_residual=False, sparse=False, narrow=False):
    import torch
    pic = dict(pic, params=_params_for(pic))
    if sparse == "narrow" or narrow:
        lv, tb = lib.pack_levels(pic["levels"], pic["tbs"], sparse=bool(sparse), narrow=True)
        pic = dict(pic, levels=lv, tbs=tb, params=lib.sparse_params(pic["params"]) if sparse else pic["params"])
    elif sparse:
        lv, tb = lib.pack_levels_sparse(pic["levels"], pic["tbs"])
        pic = dict(pic, levels=lv, tbs=tb, params=lib.sparse_params(pic["params"]))
    dev = util.to_device_planes(planes_np, torch)
    tbs = torch.from_numpy(pic["tbs"].view(np.uint8).reshape(-1)).cuda()
    levels = torch.from_numpy(pic["levels"]).cuda()
    if to_residual:
        res = [torch.zeros_like(t) for t in dev]
        lib.residual_batch(None, levels, tbs, len(pic["tbs"]), pic["params"], residual=lib.device_planes(res))
        torch.cuda.synchronize()
        return [r.cpu().numpy() for r in res]
    lib.residual_batch(lib.device_planes(dev), levels, tbs, len(pic["tbs"]), pic["params"])
    torch.cuda.synchronize()
    return util.from_device_planes(dev)


@pytest.mark.parametrize("seed,w,h,levels", [(1, 416, 240, "realistic"), (2, 416, 240, "full"),
                                             (3, 832, 480, "realistic"), (4, 200, 136, "full")])
def test_residual_batch_matches_oracle(seed, w, h, levels):
    pic = synth.residual_picture(seed, w, h, levels=levels)
    base = util.noise_planes(seed, pic["dims"])
    want = [p.copy() for p in base]
    util.oracle_residual(want, pic)
    got = _run_gpu(pic, base)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs (seed %d)" % (c, seed)
    assert util.plane_md5(got) == util.plane_md5(want)
    assert util.plane_md5(_run_gpu(pic, base, sparse=True)) == util.plane_md5(want), "sparse level format"
    assert util.plane_md5(_run_gpu(pic, base, narrow=True)) == util.plane_md5(want), "int8 levels, dense boxes"
    assert util.plane_md5(_run_gpu(pic, base, sparse="narrow")) == util.plane_md5(want), "int8 levels, sparse form"


def test_residual_plane_mode_matches_oracle():
    """int16 residual-plane output: reconstructing pred + residual on the host equals the oracle's add."""
    pic = synth.residual_picture(9, 416, 240, tools=("mts", "ts", "bdpcm", "lfnst", "scaling", "dep_quant"))
    base = util.noise_planes(9, pic["dims"])
    want = [p.copy() for p in base]
    util.oracle_residual(want, pic)
    res = _run_gpu(pic, base, to_residual=True)
    for c in range(3):
        rec = np.clip(base[c].astype(np.int32) + res[c].astype(np.int32), 0, 1023).astype(np.uint16)
        assert np.array_equal(rec, want[c])


def test_empty_batch_and_single_blocks():
    import torch
    pic = synth.residual_picture(5, 64, 64)
    base = util.noise_planes(5, pic["dims"])
    dev = util.to_device_planes(base, torch)
    lib.residual_batch(lib.device_planes(dev), None, None, 0, pic["params"])        # n == 0 is a no-op
    assert all(np.array_equal(a, b) for a, b in zip(util.from_device_planes(dev), base))
    for i in range(min(8, len(pic["tbs"]))):                                          # one TB per launch
        one = dict(pic, tbs=pic["tbs"][i:i + 1], intra_mode=pic["intra_mode"][i:i + 1])
        want = [p.copy() for p in base]
        util.oracle_residual(want, one)
        got = _run_gpu(one, base)
        assert all(np.array_equal(a, b) for a, b in zip(got, want))


def test_level_a_itx_all_slots():
    """vvcb200_dsp_itx_batch over all 210 populated itx slots x 4 random (nzw, nzh), like checkasm vvc_itx.c."""
    import torch
    from oracle import portapi, refapi
    rng = np.random.default_rng(3)
    blocks, descs, offs, off = [], [], [], 0
    for trh in range(3):
        for trv in range(3):
            for l2w in range(7):
                for l2h in range(7):
                    w, h = 1 << l2w, 1 << l2h
                    legal = (trh == 0 or 2 <= l2w <= 5) and (trv == 0 or 2 <= l2h <= 5) and \
                            (w > 1 or h >= 16) and (h > 1 or w >= 16) and not (w == 1 and h == 1)
                    if (w == 1 and trh) or (h == 1 and trv) or (w == 1 and trv and l2h > 5) or (h == 1 and trh and l2w > 5):
                        legal = False
                    if not legal:
                        continue
                    for _ in range(4):
                        nzw = int(rng.integers(1, min(w, 32 if trh == 0 else 16) + 1))
                        nzh = int(rng.integers(1, min(h, 32 if trv == 0 else 16) + 1))
                        c = np.zeros((h, w), np.int32)
                        c[:nzh, :nzw] = rng.integers(-32768, 32768, (nzh, nzw))
                        blocks.append(c.reshape(-1))
                        descs.append((l2w, l2h, nzw, nzh, trh, trv))
                        offs.append(off)
                        off += w * h
    tbs = np.zeros(len(descs), abi.TB_DTYPE)
    for i, d in enumerate(descs):
        tbs[i]["log2_w"], tbs[i]["log2_h"], tbs[i]["nzw"], tbs[i]["nzh"], tbs[i]["trh"], tbs[i]["trv"] = d
    assert len(descs) == 210 * 4
    flat = np.concatenate(blocks)
    want = flat.copy()
    orc = refapi if refapi.available() else portapi
    for d, o in zip(descs, offs):
        n = 1 << (d[0] + d[1])
        blk = want[o:o + n].copy()
        orc.itx(blk, d[4], d[5], d[0], d[1], d[2], d[3])
        want[o:o + n] = blk
    dc = torch.from_numpy(flat).cuda()
    lib.dsp_itx_batch(dc, torch.from_numpy(np.array(offs, np.uint32).view(np.int32)).cuda(),
                      torch.from_numpy(tbs.view(np.uint8).reshape(-1)).cuda(), len(descs))
    torch.cuda.synchronize()
    got = dc.cpu().numpy()
    assert np.array_equal(got, want)


def test_full_size_properties_1080p():
    """1080p batch (BASELINE config 1 size): linearity/idempotence-style properties that need no oracle run:
    (a) zero levels leave the picture untouched, (b) two disjoint halves of the batch compose to the whole."""
    pic = synth.residual_picture(21, 1920, 1080, tools=("mts",), mean_log2=4.5)
    base = util.noise_planes(21, pic["dims"])
    zero = dict(pic, levels=np.zeros_like(pic["levels"]))
    got0 = _run_gpu(zero, base)
    assert all(np.array_equal(a, b) for a, b in zip(got0, base))
    whole = _run_gpu(pic, base)
    n = len(pic["tbs"])
    first = _run_gpu(dict(pic, tbs=pic["tbs"][: n // 2]), base)
    both = _run_gpu(dict(pic, tbs=pic["tbs"][n // 2:]), first)
    assert util.plane_md5(both) == util.plane_md5(whole)
    # and the oracle agrees on a bounded sample of the same batch (first 4000 TBs)
    sample = dict(pic, tbs=pic["tbs"][:4000], intra_mode=pic["intra_mode"][:4000])
    want = [p.copy() for p in base]
    util.oracle_residual(want, sample)
    got = _run_gpu(sample, base)
    assert util.plane_md5(got) == util.plane_md5(want)


def test_level_b_context_host_buffers():
    """The picture engine with HOST buffers (upload -> residual -> download) equals the oracle."""
    pic = synth.residual_picture(31, 416, 240)
    base = util.noise_planes(31, pic["dims"])
    want = [p.copy() for p in base]
    util.oracle_residual(want, pic)
    ctx = lib.Context(0, 416, 240)
    try:
        src = ctx.pinned_planes()
        out = ctx.pinned_planes()
        for s, b in zip(src, base):
            s[...] = b
        levels = ctx.pinned(pic["levels"].shape, np.int16)
        levels[...] = pic["levels"]
        tbs = ctx.pinned(pic["tbs"].shape, abi.TB_DTYPE)
        tbs[...] = pic["tbs"]
        for k in range(2):                      # twice: the second picture reuses the device buffers
            ctx.begin(k)
            ctx.upload(src)
            ctx.residual(levels, tbs, _params_for(pic))
            ctx.download(out)
            ctx.sync()
            assert util.plane_md5(out) == util.plane_md5(want)
    finally:
        ctx.close()


def test_sparse_levels_with_mostly_empty_scan_boxes():
    """A scan box filled to a quarter, as real residuals are: the sparse form is less than half the dense bytes and
    reconstructs the same picture; Level B takes the smaller form by itself."""
    pic = synth.residual_picture(12, 416, 240)
    rng = np.random.default_rng(12)
    pic = dict(pic, levels=np.where(rng.random(pic["levels"].shape) < 0.25, pic["levels"], 0).astype(np.int16))
    lv, _ = lib.pack_levels_sparse(pic["levels"], pic["tbs"])
    assert lv.nbytes < 0.5 * pic["levels"].nbytes
    base = util.noise_planes(12, pic["dims"])
    want = [p.copy() for p in base]
    util.oracle_residual(want, pic)
    assert util.plane_md5(_run_gpu(pic, base, sparse=True)) == util.plane_md5(want)
    assert util.plane_md5(_run_gpu(pic, base, sparse=False)) == util.plane_md5(want)
