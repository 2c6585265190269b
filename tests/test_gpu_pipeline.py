"""GPU parity of the whole picture chain through the Level-B C ABI (host buffers in, host planes out):
INTER -> residual -> (LMCS) -> DEBLOCK_V/H -> SAO -> ALF, per-plane MD5 against the oracle's stage chain."""
import numpy as np
import pytest

from ffmpeg_b200 import lib, synth
from tests import pipeline, util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,w,h,ctb,lmcs,ns,lfase", [(1, 416, 240, 7, False, 1, 1), (2, 416, 240, 6, True, 3, 0),
                                                        (3, 832, 480, 7, False, 2, 0), (4, 256, 128, 5, True, 1, 1)])
def test_full_picture_md5(seed, w, h, ctb, lmcs, ns, lfase):
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=lmcs, n_slices=ns, lfase=lfase)
    want = pipeline.oracle_picture(pic)
    ctx = lib.Context(0, w, h, dpb_size=3)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        for _ in range(2):
            ep.run()
            ctx.sync()
            assert util.plane_md5(ep.out) == util.plane_md5(want)
    finally:
        ctx.close()


@pytest.mark.parametrize("seed,w,h,ctb,lmcs,frac", [(21, 416, 240, 7, False, 0.3), (22, 416, 240, 6, True, 0.5),
                                                    (23, 832, 480, 5, False, 1.0), (24, 1920, 1080, 7, True, 0.15)])
def test_full_picture_with_intra_cus_md5(seed, w, h, ctb, lmcs, frac):
    """One CU partition with inter and intra CUs: INTER -> residual (inter TBs in place, intra TBs to the residual
    planes) -> intra wavefront -> (LMCS) -> deblock -> SAO -> ALF, against the reference's INTER + ff_vvc_reconstruct."""
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=lmcs, intra_fraction=frac)
    want = pipeline.oracle_picture(pic)
    ctx = lib.Context(0, w, h, dpb_size=3)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        for _ in range(2):
            ep.run()
            ctx.sync()
            assert util.plane_md5(ep.out) == util.plane_md5(want)
    finally:
        ctx.close()


@pytest.mark.parametrize("seed,w,h,ctb,lmcs,ns,lfase", [(31, 416, 240, 7, False, 1, 1), (32, 832, 480, 6, True, 3, 0)])
def test_picture_from_emitted_descriptors_md5(seed, w, h, ctb, lmcs, ns, lfase):
    """the emitter (ffmpeg_b200/emit) turns the reference-side state of the picture — CTU / CodingUnit /
    TransformUnit lists, MvField table, SAOParams / ALFParams / APS list — into the engine's batches; the engine run on THOSE
    must produce what ff_vvc_predict_inter + ff_vvc_reconstruct + the loop filters produce on the same state."""
    from oracle import refapi
    if not refapi.available():
        pytest.skip("oracle/_ref not built")
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=lmcs, n_slices=ns, lfase=lfase, intra_fraction=0.3)
    want = pipeline.oracle_picture(pic)
    ctx = lib.Context(0, w, h, dpb_size=3)
    try:
        ep = pipeline.EnginePicture(ctx, pipeline.emitted_picture(pic))
        ep.upload_refs()
        ep.run()
        ctx.sync()
        assert util.plane_md5(ep.out) == util.plane_md5(want)
    finally:
        ctx.close()


def test_batches_in_an_arena_and_in_separate_allocations_give_the_same_picture():
    """vvcb200_ctx_arena_alloc: batch arrays carved out of the mirrored pinned block travel as one copy per stage call;
    arrays in allocations of their own (one copy each, the engine's own device buffers) must reconstruct the same picture,
    and the two kinds may alternate on one context."""
    w, h = 416, 240
    pic = synth.full_picture(61, w, h, ctb_log2=6, lmcs=True, intra_fraction=0.3)
    want = util.plane_md5(pipeline.oracle_picture(pic))
    ctx = lib.Context(0, w, h, dpb_size=4)
    try:
        eps = [pipeline.EnginePicture(ctx, pic, arena=True), pipeline.EnginePicture(ctx, pic, arena=False)]
        eps[0].upload_refs()
        for i in range(6):
            ep = eps[i % 2]
            ep.run()
            ctx.sync()
            assert util.plane_md5(ep.out) == want, "picture %d (%s)" % (i, "arena" if i % 2 == 0 else "separate allocations")
    finally:
        ctx.close()


def test_refined_vectors_come_back_for_the_dmvr_records():
    """Level B returns the vectors DMVR settled on (the reference's tab_dmvr_mvf, vvc_inter.c:750-762).  A batch in the
    recommended order (dmvr_first) gets them for the DMVR records only, the other entries stay as the host left them."""
    from oracle import refapi
    from ffmpeg_b200 import abi
    if not refapi.available():
        pytest.skip("oracle/_ref not built")
    w, h = 416, 240
    pic = synth.full_picture(51, w, h, ctb_log2=7)
    it = pic["inter"]
    assert it["params"].dmvr_first == 1
    dm = (it["blocks"]["flags"] & abi.PB_DMVR) != 0
    assert dm.sum() > 10
    planes = [np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]]
    want = refapi.inter_batch(planes, it["refs"], it["blocks"], it["params"], want_refined_mv=True)
    ctx = lib.Context(0, w, h, dpb_size=3)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        for r in ep.refined:
            r[...] = -77
        ep.run()
        ctx.sync()
        got = ep.refined[ep.count % 4]
        assert np.array_equal(got[dm], want[dm])
        assert np.any(want[dm] != it["blocks"]["mv"][dm])
        assert (got[~dm] == -77).all()
        assert ep.refined_bytes() == 16 * int(dm.sum())
    finally:
        ctx.close()


def test_back_to_back_pictures_overlap_safely():
    """Six pictures submitted without a sync in between, alternating between two DPB slots and two host frames: uploads,
    kernels and downloads of neighbouring pictures overlap on the context's three streams and every frame is right."""
    pic = synth.full_picture(31, 832, 480, ctb_log2=7, lmcs=True, intra_fraction=0.2)
    want = util.plane_md5(pipeline.oracle_picture(pic))
    ctx = lib.Context(0, 832, 480, dpb_size=4)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        assert ep.n_slots == 2
        ep.upload_refs()
        for rounds in range(3):
            for o in ep.outs:
                for p in o:
                    p[:] = 0
            for _ in range(6):
                ep.run()
            ctx.sync()
            for o in ep.outs:
                assert util.plane_md5(o) == want
    finally:
        ctx.close()


def test_picture_that_references_its_predecessor():
    """Consecutive pictures run in different lanes of the context (separate compute streams).  Here the second picture
    predicts from the first one's output, so its INTER stage has to wait for the other lane; submitted three times
    without a sync in between, so the first picture's slot is also rewritten while its reader may still be running."""
    import copy
    from ffmpeg_b200 import abi
    pic = synth.full_picture(33, 416, 240, ctb_log2=6, intra_fraction=0.2)
    out1 = pipeline.oracle_picture(pic)
    n_refs = len(pic["inter"]["refs"])
    pic2 = dict(pic, inter=dict(pic["inter"], refs=[[p.copy() for p in out1]] + list(pic["inter"]["refs"][1:])))
    want1, want2 = util.plane_md5(out1), util.plane_md5(pipeline.oracle_picture(pic2))
    assert want1 != want2
    ctx = lib.Context(0, 416, 240, dpb_size=n_refs + 2)
    try:
        ep1 = pipeline.EnginePicture(ctx, pic)
        ep1.upload_refs()
        eng2 = copy.copy(pic2)
        eng2["inter"] = dict(pic2["inter"], params=abi.InterParams.from_buffer_copy(bytes(pic["inter"]["params"])))
        for l in range(2):
            for i in range(16):
                if eng2["inter"]["params"].ref_slot[l][i] == 0:
                    eng2["inter"]["params"].ref_slot[l][i] = n_refs          # the slot ep1 reconstructs into
        ep2 = pipeline.EnginePicture(ctx, eng2)
        for _ in range(3):
            ep1.count, ep2.count = 0, 1                                       # ep1 -> slot n_refs, ep2 -> slot n_refs + 1
            ep1.run()
            ep2.run()
        ctx.sync()
        assert util.plane_md5(ep1.outs[0]) == want1
        assert util.plane_md5(ep2.outs[1]) == want2
    finally:
        ctx.close()


@pytest.mark.parametrize("seed,w,h,ctb,lmcs,frac", [(41, 256, 128, 6, False, 0.0), (42, 256, 128, 5, True, 0.0),
                                                    (43, 208, 120, 6, True, 0.4), (44, 416, 240, 7, False, 1.0),
                                                    (45, 832, 480, 7, True, 0.2)])
def test_full_picture_12_bit_md5(seed, w, h, ctb, lmcs, frac):
    """Every stage at bit depth 12 (16-bit containers, as in the reference's BIT_DEPTH 12 templates)."""
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=lmcs, intra_fraction=frac, bd=12)
    want = pipeline.oracle_picture(pic)
    ctx = lib.Context(0, w, h, bit_depth=12, dpb_size=4)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        for _ in range(2):
            ep.run()
            ctx.sync()
            assert util.plane_md5(ep.out) == util.plane_md5(want)
    finally:
        ctx.close()


def test_full_picture_1080p_md5():
    """BASELINE config 2 size: 1080p full picture recon + deblock + SAO + ALF."""
    pic = synth.full_picture(11, 1920, 1080)
    want = pipeline.oracle_picture(pic)
    ctx = lib.Context(0, 1920, 1080, dpb_size=3)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        ep.run()
        ctx.sync()
        assert util.plane_md5(ep.out) == util.plane_md5(want)
    finally:
        ctx.close()


def test_different_pictures_on_one_lane_do_not_see_each_others_residuals():
    """Two pictures with different CU / TB / ISP layouts alternate on the context: each lane's int16 residual picture is
    allocated once and never cleared, so a picture must read only what its own TBs wrote."""
    a = synth.full_picture(51, 416, 240, ctb_log2=6, lmcs=True, intra_fraction=0.6)
    b = synth.full_picture(52, 416, 240, ctb_log2=6, lmcs=True, intra_fraction=0.6)
    want = {id(a): util.plane_md5(pipeline.oracle_picture(a)), id(b): util.plane_md5(pipeline.oracle_picture(b))}
    ctx = lib.Context(0, 416, 240, dpb_size=6)
    try:
        ea, eb = pipeline.EnginePicture(ctx, a), pipeline.EnginePicture(ctx, b)
        ea.upload_refs()                                     # (both pictures' reference slots: a's and b's differ)
        for seq in ((ea, ea, eb, eb, ea, ea), (eb, ea, ea, eb, eb, ea)):
            for i, ep in enumerate(seq):
                ep.upload_refs()
                ep.count = i
                ep.run()
                ctx.sync()
                assert util.plane_md5(ep.out) == want[id(ep.pic)], "picture %d of the sequence" % i
    finally:
        ctx.close()


def test_export_and_windowed_download_crop_to_the_conformance_window():
    """the picture leaves either as device planes of the conformance window plus a ready event (no copy), or as
    a download of the window only; both must hold the oracle picture cropped the way AVFrame.crop_* says."""
    import torch
    pic = synth.full_picture(61, 416, 240, ctb_log2=6, intra_fraction=0.2)
    want = pipeline.oracle_picture(pic)
    win = (8, 16, 4, 12)                                    # left, right, top, bottom in luma samples
    crop = [want[0][4:240 - 12, 8:416 - 16], want[1][2:120 - 6, 4:208 - 8], want[2][2:120 - 6, 4:208 - 8]]
    ctx = lib.Context(0, 416, 240, dpb_size=4)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        ep.run(download=False)
        planes, ready = ctx.export(window=win)
        assert ready and [planes.width[c] for c in range(3)] == [392, 196, 196] and [planes.height[c] for c in range(3)] == [224, 112, 112]
        ev = torch.cuda.ExternalStream(ctx.stream())       # (the lane stream: later work on it is ordered behind the picture)
        ev.synchronize()
        got = util.read_device_planes(planes)
        for c in range(3):
            assert np.array_equal(got[c], crop[c]), "exported plane %d differs" % c
        host = [ctx.pinned(p.shape, np.uint16) for p in crop]
        ctx.download_window(host, win)
        ctx.release()                                       # nothing else is reading: the slot may be rewritten
        ctx.sync()
        for c in range(3):
            assert np.array_equal(host[c], crop[c]), "downloaded window plane %d differs" % c
        # bad windows are refused
        with pytest.raises(lib.VVCB200Error):
            ctx.export(window=(3, 0, 0, 0))
        with pytest.raises(lib.VVCB200Error):
            ctx.export(window=(208, 208, 0, 0))
    finally:
        ctx.close()
