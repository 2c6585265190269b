"""GPU parity at the sizes the numbers are quoted on: the 4K random-access bench picture through the whole Level-B chain,
and a 4K all-intra picture through the intra stage under both protocols, each against the compiled reference
(oracle/_ref; the C port where that is absent).  The pictures are bench.py's (same seeds, shared cache file), so a
bench run in the same session does not generate them again.  8K (BASELINE config 4) takes minutes to generate and is
checked by `bench.py --workload 8k`, whose line carries the same MD5 comparison (profiles/)."""
import os

import numpy as np
import pytest

import bench
from ffmpeg_b200 import lib
from oracle import refapi
from tests import pipeline, util
from tests.test_gpu_intra import oracle_to_intra, run_gpu

pytestmark = pytest.mark.gpu
THREADS = os.cpu_count() or 1


def oracle_picture_threaded(pic):
    """tests.pipeline.oracle_picture with the reference's stages spread over the host threads (bench.py's CPU arm)."""
    if not refapi.available():
        return pipeline.oracle_picture(pic)
    return bench.cpu_picture(pic, THREADS)[2]


def test_4k_bench_picture_matches_reference_md5():
    pic = bench.make_picture(2026, 3840, 2160, 0.1, True)
    want = util.plane_md5(oracle_picture_threaded(pic))
    ctx = lib.Context(0, 3840, 2160, dpb_size=len(pic["inter"]["refs"]) + 2)
    try:
        ep = pipeline.EnginePicture(ctx, pic)
        ep.upload_refs()
        for _ in range(4):                       # both lanes, every descriptor set
            ep.run()
        ctx.sync()
        for o in ep.outs:
            assert util.plane_md5(o) == want
    finally:
        ctx.close()


def test_4k_all_intra_matches_reference_under_both_protocols():
    """The sentinel / ticket protocol at full occupancy: ~1800 dependency steps, every SM polling."""
    pic = bench.make_picture(2026, 3840, 2160, 1.0, True)
    if refapi.available():
        planes = [np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]]
        refapi.recon_picture(planes, pic["ref"], pic["intra"]["params"], threads=THREADS, inter=pic["inter"],
                             lmcs_fwd_lut=pic["lmcs"][0])
        want = util.plane_md5(planes)
    else:
        want = util.plane_md5(oracle_to_intra(pic))
    for shadow, phased in ((False, False), (True, False), (True, "diagonal"), (False, "diagonal")):
        got = run_gpu(pic, shadow=shadow, phased=phased)[0]
        assert util.plane_md5(got) == want, "shadow %s, phased %s" % (shadow, phased)
