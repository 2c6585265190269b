"""GPU parity: vvcb200_intra_picture (through the C ABI) against the oracle — the compiled reference's
ff_vvc_reconstruct where oracle/_ref exists, else the C port — on identical seeded pictures.  Bit-exact."""
import numpy as np
import pytest

from ffmpeg_b200 import abi, lib, synth
from tests import util

pytestmark = pytest.mark.gpu


def run_gpu(pic, schedule=True, repeat=1, shadow=False, phased=False, stale_residual=False):
    """INTER -> RESIDUAL -> INTRA on the device, Level A calls only.  shadow: hand the call a second picture, so that
    blocks rewritten in place take the sentinel protocol too (else they go by marks and fences)."""
    import torch
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]
    dev = util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch)
    dpl = lib.device_planes(dev)
    if len(it["blocks"]):
        refs = [util.to_device_planes(r, torch) for r in it["refs"]]
        blocks = torch.from_numpy(it["blocks"].view(np.uint8).reshape(-1)).cuda()
        prof = it.get("prof")
        prof = torch.from_numpy(prof.view(np.uint8).reshape(-1)).cuda() if prof is not None and len(prof) else None
        fwd = torch.from_numpy(pic["lmcs"][0]).cuda() if pic.get("lmcs") is not None else None
        lib.inter_batch(dpl, [lib.device_planes(r) for r in refs], blocks, len(it["blocks"]), it["params"],
                        lmcs_fwd_lut=fwd, prof=prof)
    # stale_residual: the planes arrive full of another picture's leftovers; only what this picture's TBs write may be read
    res = [torch.full_like(t, 0x2b67) if stale_residual else torch.zeros_like(t) for t in dev]
    rpl = lib.device_planes(res)
    if len(rs["tbs"]):
        tbs = torch.from_numpy(rs["tbs"].view(np.uint8).reshape(-1)).cuda()
        levels = torch.from_numpy(rs["levels"]).cuda()
        lib.residual_batch(dpl, levels, tbs, len(rs["tbs"]), rs["params"], residual=rpl)
    w, h, ctb = pic["width"], pic["height"], pic["ctb_log2"]
    recs = ia["blocks"]
    iparams = ia["params"]
    if phased:          # luma wavefront -> add-only batch -> chroma wavefront (what Level B runs)
        recs, counts = lib.intra_schedule(recs, ia["ctu_first"], (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb, ctb, w, h,
                                          order=lib.INTRA_ORDER_DIAGONAL if phased == "diagonal" else lib.INTRA_ORDER_DEPTH,
                                          phases=True)
        iparams = lib.phased_params(ia["params"], counts)
    elif schedule:
        recs = lib.intra_schedule(recs, ia["ctu_first"], (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb, ctb, w, h)
    scratch = torch.zeros(lib.intra_scratch_bytes(w, h), dtype=torch.uint8, device="cuda")
    side = [torch.full_like(t, 0x1234) for t in dev] if shadow else None
    spl = lib.device_planes(side) if shadow else None
    out = []
    if len(recs):
        d_recs = torch.from_numpy(recs.view(np.uint8).reshape(-1)).cuda()
        base = [t.clone() for t in dev]
        for _ in range(repeat):
            for t, b in zip(dev, base):
                t.copy_(b)
            lib.intra_picture(dpl, rpl, d_recs, len(recs), iparams, scratch, shadow=spl)
            torch.cuda.synchronize()
            out.append(util.from_device_planes(dev))
        # the per-VPDU table of phased pictures follows the marks (each area a multiple of 256 bytes)
        marks_end = (lib.intra_scratch_bytes(w, h) - 4 * ((w + 31) // 32) * ((h + 31) // 32)) // 256 * 256
        marks = scratch[256:marks_end]
        assert int((marks == 1).sum()) == 0, "no unit may be left pending after the call"
        if shadow:
            assert int((marks != 0).sum()) == 0, "with a shadow picture nothing goes by marks"
    else:
        out.append(util.from_device_planes(dev))
    return out


ALL_TOOLS = ("pdpc", "mrl", "mip", "cclm", "mts", "ts", "dep_quant", "scaling", "joint", "lfnst", "bdpcm", "isp", "ciip")
CASES = [
    (1, 208, 120, 6, ("pdpc",), 0.6),
    (2, 208, 120, 5, ("pdpc", "mrl"), 0.6),
    (3, 208, 120, 7, ("pdpc", "mrl", "mip"), 0.6),
    (4, 208, 120, 6, ("pdpc", "mrl", "mip", "cclm"), 0.6),
    (5, 416, 240, 7, ("pdpc", "mrl", "mip", "cclm", "mts"), 1.0),
    (6, 136, 72, 5, ("pdpc", "mrl", "mip", "cclm"), 0.3),
    (8, 208, 120, 6, ("isp",), 0.8),
    (9, 208, 120, 5, ("bdpcm", "ts", "lfnst"), 0.8),
    (10, 256, 192, 7, ALL_TOOLS, 0.7),
    (11, 136, 72, 5, ALL_TOOLS, 1.0),
    (12, 416, 240, 6, ALL_TOOLS, 0.4),
    (7, 832, 480, 7, ("pdpc", "mrl", "mip", "cclm", "mts"), 0.5),
]


@pytest.mark.parametrize("seed,w,h,ctb,tools,frac", CASES)
def test_intra_picture_matches_oracle(seed, w, h, ctb, tools, frac):
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=tools, intra_fraction=frac)
    want = util.oracle_recon([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], pic)
    for shadow in (False, True):
        got = run_gpu(pic, shadow=shadow)[0]
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "plane %d differs (seed %d, shadow %d)" % (c, seed, shadow)
        assert util.plane_md5(got) == util.plane_md5(want)


def test_decoding_order_and_repeat_give_the_same_picture():
    """Any topological order works (plain decoding order included), and the scratch is reusable call after call."""
    pic = synth.recon_picture(21, 416, 240, ctb_log2=6, intra_fraction=0.7)
    want = util.oracle_recon([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], pic)
    for got in run_gpu(pic, schedule=False) + run_gpu(pic, schedule=True, repeat=3) + run_gpu(pic, repeat=2, shadow=True):
        for c in range(3):
            assert np.array_equal(got[c], want[c])


def oracle_to_intra(pic):
    """INTER + RECON of a synth.full_picture() (forward-mapped inter prediction when LMCS is on)."""
    from oracle import portapi, refapi
    planes = [np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]]
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]
    fwd = pic["lmcs"][0] if pic["lmcs"] is not None else None
    util.oracle_inter(planes, it, fwd, refapi.available())
    if refapi.available():
        refapi.recon_picture(planes, pic["ref"], ia["params"], inter=it, lmcs_fwd_lut=fwd)
    else:
        res = [np.zeros(p.shape, np.int16) for p in planes]
        portapi.residual_batch(planes, rs["levels"], rs["tbs"], rs["params"], residual_planes=res)
        portapi.intra_picture(planes, res, ia["blocks"], ia["params"])
    return planes


@pytest.mark.parametrize("seed,w,h,ctb,frac", [(71, 416, 240, 6, 0.4), (72, 256, 192, 7, 0.15)])
def test_chroma_scaling_and_ciip_under_both_protocols(seed, w, h, ctb, frac):
    """Luma-dependent chroma residual scaling gives every inter unit add-only chroma records; with CIIP units they are
    the blocks rewritten in place.  Marks + fences (no shadow) and sentinels (shadow) must both match the oracle."""
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=True, intra_fraction=frac)
    want = oracle_to_intra(pic)
    flags = pic["intra"]["blocks"]["flags"]
    assert (flags & abi.IB_ADD_ONLY).any() and (flags & abi.IB_CIIP).any()
    for shadow in (False, True):
        got = run_gpu(pic, shadow=shadow)[0]
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "plane %d differs (seed %d, shadow %d)" % (c, seed, shadow)
    for shadow, phased in ((False, "diagonal"), (True, "diagonal"), (True, "depth")):
        got = run_gpu(pic, shadow=shadow, phased=phased, repeat=2)
        for g in got:
            for c in range(3):
                assert np.array_equal(g[c], want[c]), "phased: plane %d differs (seed %d, shadow %d)" % (c, seed, shadow)


@pytest.mark.parametrize("seed,w,h,ctb,tools,frac", [CASES[8], CASES[9], CASES[10]])
def test_phased_intra_matches_oracle(seed, w, h, ctb, tools, frac):
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=tools, intra_fraction=frac)
    want = util.oracle_recon([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], pic)
    for shadow in (False, True):
        got = run_gpu(pic, shadow=shadow, phased="diagonal")[0]
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "plane %d differs (seed %d, shadow %d)" % (c, seed, shadow)


def test_all_intra_1080p_property():
    """Full-size picture: no oracle run; the two orders must agree with each other bit for bit."""
    pic = synth.recon_picture(31, 1920, 1080, ctb_log2=7, intra_fraction=1.0)
    a = run_gpu(pic, schedule=False)[0]
    b = run_gpu(pic, schedule=True)[0]
    assert util.plane_md5(a) == util.plane_md5(b)
    assert util.plane_md5(run_gpu(pic, shadow=True)[0]) == util.plane_md5(a)
    assert util.plane_md5(run_gpu(pic, shadow=True, phased="diagonal")[0]) == util.plane_md5(a)


def test_empty_picture_is_a_noop():
    import torch
    dev = util.to_device_planes(util.noise_planes(1, [(64, 64), (32, 32), (32, 32)]), torch)
    before = util.from_device_planes(dev)
    ip = abi.IntraParams()
    ip.bit_depth, ip.ctb_log2_size = 10, 7
    lib.intra_picture(lib.device_planes(dev), None, None, 0, ip, None)
    assert all(np.array_equal(a, b) for a, b in zip(util.from_device_planes(dev), before))


@pytest.mark.parametrize("seed,w,h,ctb", [(61, 208, 120, 6), (62, 136, 72, 5), (63, 256, 192, 7), (64, 832, 480, 7)])
def test_dual_tree_picture_matches_oracle(seed, w, h, ctb):
    """I picture with separate luma and chroma partitions, as all-intra streams are coded."""
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=ALL_TOOLS, dual_tree=True)
    want = util.oracle_recon([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], pic)
    got = run_gpu(pic)[0]
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs (seed %d)" % (c, seed)


@pytest.mark.parametrize("seed,w,h,ctb", [(81, 416, 240, 6), (82, 832, 480, 7), (83, 136, 72, 5)])
def test_device_schedule_is_the_phased_diagonal_order_and_reconstructs_the_oracle_picture(seed, w, h, ctb):
    """vvcb200_intra_schedule_device: per phase, CTU anti-diagonals with the CTUs of a diagonal dealt record by record."""
    import torch
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=True, intra_fraction=0.35)
    ia = pic["intra"]
    recs, cf = ia["blocks"], ia["ctu_first"].astype(np.uint32)
    n, cw, ch = len(recs), (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    d_in = torch.from_numpy(recs.view(np.uint8).reshape(-1)).cuda()
    d_cf = torch.from_numpy(cf.view(np.int32)).cuda()
    d_out = torch.zeros_like(d_in)
    d_counts = torch.zeros(4, dtype=torch.int32, device="cuda")
    scratch = torch.zeros(lib.intra_schedule_scratch_bytes(n, cw, ch), dtype=torch.uint8, device="cuda")
    lib.intra_schedule_device(d_in, d_cf, n, cw, ch, d_out, d_counts, scratch)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy().view(abi.IB_DTYPE)
    counts = d_counts.cpu().numpy()
    ctu = np.repeat(np.arange(cw * ch), np.diff(cf.astype(np.int64)))
    phase = np.where(recs["c_idx"] == 0, 0, np.where((recs["flags"] & abi.IB_ADD_ONLY) != 0, 1, 2))
    k = np.zeros(n, np.int64)
    for c in range(cw * ch):
        for p in range(3):
            m = np.nonzero((ctu == c) & (phase == p))[0]
            k[m] = np.arange(len(m))
    cy, cx = ctu // cw, ctu % cw
    order = np.lexsort((cy, k, cx + 2 * cy, phase))
    assert list(counts[:3]) == [int((phase == p).sum()) for p in range(3)]
    assert got.tobytes() == recs[order].tobytes()
    # and the picture reconstructed from that order, the counts staying on the device
    want = oracle_to_intra(pic)
    dev = util.to_device_planes([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], torch)
    dpl = lib.device_planes(dev)
    it, rs = pic["inter"], pic["residual"]
    refs = [util.to_device_planes(r, torch) for r in it["refs"]]
    prof = it.get("prof")
    lib.inter_batch(dpl, [lib.device_planes(r) for r in refs], torch.from_numpy(it["blocks"].view(np.uint8).reshape(-1)).cuda(),
                    len(it["blocks"]), it["params"], lmcs_fwd_lut=torch.from_numpy(pic["lmcs"][0]).cuda(),
                    prof=torch.from_numpy(prof.view(np.uint8).reshape(-1)).cuda() if prof is not None and len(prof) else None)
    res = [torch.zeros_like(t) for t in dev]
    rpl = lib.device_planes(res)
    lib.residual_batch(dpl, torch.from_numpy(rs["levels"]).cuda(), torch.from_numpy(rs["tbs"].view(np.uint8).reshape(-1)).cuda(),
                       len(rs["tbs"]), rs["params"], residual=rpl)
    ip = abi.IntraParams.from_buffer_copy(bytes(ia["params"]))
    ip.d_phase_counts = d_counts.data_ptr()
    marks = torch.zeros(lib.intra_scratch_bytes(w, h), dtype=torch.uint8, device="cuda")
    side = [torch.zeros_like(t) for t in dev]
    lib.intra_picture(dpl, rpl, d_out, n, ip, marks, shadow=lib.device_planes(side))
    torch.cuda.synchronize()
    got_pic = util.from_device_planes(dev)
    for c in range(3):
        assert np.array_equal(got_pic[c], want[c]), "plane %d differs" % c


@pytest.mark.parametrize("seed,w,h,ctb", [(8, 208, 120, 6), (91, 416, 240, 7), (92, 136, 72, 5)])
def test_uncoded_isp_partitions_add_nothing_whatever_the_residual_planes_hold(seed, w, h, ctb):
    """ISP records carry one residual flag for the coding block; partitions without a coded TB (isp_coded) must not pick up
    what an earlier picture left in the residual planes."""
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=("isp", "pdpc", "mts", "cclm"), intra_fraction=0.9)
    isp = pic["intra"]["blocks"][(pic["intra"]["blocks"]["flags"] & abi.IB_ISP) != 0]
    assert len(isp) and (isp["isp_coded"] != 15).any(), "the picture needs partially coded ISP blocks"
    want = util.oracle_recon([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], pic)
    for shadow, phased in ((False, False), (True, "diagonal")):
        got = run_gpu(pic, shadow=shadow, phased=phased, stale_residual=True)[0]
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "plane %d differs (shadow %s)" % (c, shadow)


@pytest.mark.parametrize("shadow,phased", [(True, True), (False, True), (True, False)])
def test_pre_pass_beside_the_residual_stage_and_the_luma_done_event(shadow, phased):
    """vvcb200_intra_prepare on a second stream before / beside vvcb200_residual_batch, then vvcb200_intra_picture with
    params.prepared: same picture as the one call.  luma_done_event fires when luma is final: a copy of the luma plane taken
    on a stream that waits for that event alone, while the chroma phases still run, is the oracle's luma."""
    import torch
    seed, w, h, ctb = 73, 416, 240, 6
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=True, intra_fraction=0.4)
    want = oracle_to_intra(pic)
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]
    dev = util.to_device_planes([np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]], torch)
    dpl = lib.device_planes(dev)
    refs = [util.to_device_planes(r, torch) for r in it["refs"]]
    blocks = torch.from_numpy(it["blocks"].view(np.uint8).reshape(-1)).cuda()
    prof = torch.from_numpy(it["prof"].view(np.uint8).reshape(-1)).cuda() if len(it["prof"]) else None
    fwd = torch.from_numpy(pic["lmcs"][0]).cuda()
    res = [torch.full_like(t, 0x2b67) for t in dev]
    rpl = lib.device_planes(res)
    tbs = torch.from_numpy(rs["tbs"].view(np.uint8).reshape(-1)).cuda()
    levels = torch.from_numpy(rs["levels"]).cuda()
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    if phased:
        recs, counts = lib.intra_schedule(ia["blocks"], ia["ctu_first"], cw, ch, ctb, w, h, phases=True)
        iparams = lib.phased_params(ia["params"], counts)
    else:
        recs = lib.intra_schedule(ia["blocks"], ia["ctu_first"], cw, ch, ctb, w, h)
        iparams = abi.IntraParams.from_buffer_copy(bytes(ia["params"]))
    d_recs = torch.from_numpy(recs.view(np.uint8).reshape(-1)).cuda()
    scratch = torch.zeros(lib.intra_scratch_bytes(w, h), dtype=torch.uint8, device="cuda")
    side = [torch.full_like(t, 0x1234) for t in dev] if shadow else None
    spl = lib.device_planes(side) if shadow else None
    main, second = torch.cuda.current_stream(), torch.cuda.Stream()
    luma_copy = torch.zeros_like(dev[0])
    for _ in range(3):
        for t in dev:
            t.zero_()
        lib.inter_batch(dpl, [lib.device_planes(r) for r in refs], blocks, len(it["blocks"]), it["params"], lmcs_fwd_lut=fwd, prof=prof)
        second.wait_stream(main)                                   # the pre-pass needs the inter prediction ...
        lib.intra_prepare(dpl, rpl, d_recs, len(recs), iparams, scratch, shadow=spl, stream=second.cuda_stream)
        lib.residual_batch(dpl, levels, tbs, len(rs["tbs"]), rs["params"], residual=rpl)     # ... not the residual stage
        main.wait_stream(second)
        ev = torch.cuda.Event()
        p2 = abi.IntraParams.from_buffer_copy(bytes(iparams))
        p2.prepared = 1
        ev.record()                                                # (torch creates the CUDA event on its first record)
        p2.luma_done_event = ev.cuda_event
        assert p2.luma_done_event
        lib.intra_picture(dpl, rpl, d_recs, len(recs), p2, scratch, shadow=spl)
        second.wait_event(ev)
        with torch.cuda.stream(second):
            luma_copy.copy_(dev[0])
        torch.cuda.synchronize()
        got = util.from_device_planes(dev)
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "plane %d differs" % c
        assert np.array_equal(luma_copy.cpu().numpy().view(np.uint16), want[0])
