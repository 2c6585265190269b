"""CPU tests of the intra oracle: the C port (oracle/port/port_intra.c) against the compiled reference's
ff_vvc_reconstruct (oracle/_ref) on seeded pictures and against the committed golden picture; and the host-side
wavefront scheduler (vvcb200_intra_schedule), which must yield an order the port reconstructs identically from."""
import os

import numpy as np
import pytest

from ffmpeg_b200 import abi, lib, synth
from oracle import portapi, refapi
from tests import util

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
]


def _port(pic, order=None):
    planes = [np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]]
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]
    util.oracle_inter(planes, it, None, False)
    res = [np.zeros(p.shape, np.int16) for p in planes]
    portapi.residual_batch(planes, rs["levels"], rs["tbs"], rs["params"], residual_planes=res)
    portapi.intra_picture(planes, res, ia["blocks"] if order is None else order, ia["params"])
    return planes


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed,w,h,ctb,tools,frac", CASES)
def test_port_matches_reference_reconstruct(seed, w, h, ctb, tools, frac):
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=tools, intra_fraction=frac)
    want = [np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]]
    util.oracle_inter(want, pic["inter"], None, True)
    refapi.recon_picture(want, pic["ref"], pic["intra"]["params"], inter=pic["inter"])
    got = _port(pic)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c


def test_port_matches_golden():
    g = np.load(os.path.join(util.GOLDEN, "intra.npz"))
    pic = synth.recon_picture(int(g["seed"]), int(g["width"]), int(g["height"]), ctb_log2=int(g["ctb_log2"]),
                              tools=tuple(str(t) for t in g["tools"]), intra_fraction=float(g["intra_fraction"]))
    assert util.md5(pic["intra"]["blocks"]) == str(g["blocks_md5"]), "synthetic generator drifted from the fixture"
    got = _port(pic)
    for c in range(3):
        assert np.array_equal(got[c], g["plane%d" % c])


@pytest.mark.parametrize("seed,w,h,ctb", [(11, 416, 240, 5), (12, 416, 240, 6), (13, 300, 200, 7), (14, 64, 64, 6)])
@pytest.mark.parametrize("order", [lib.INTRA_ORDER_DIAGONAL, lib.INTRA_ORDER_DEPTH])
def test_schedule_is_a_topological_permutation(seed, w, h, ctb, order):
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, intra_fraction=0.8)
    ia = pic["intra"]
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    order = lib.intra_schedule(ia["blocks"], ia["ctu_first"], cw, ch, ctb, w, h, order=order)
    assert sorted(order.tobytes()[i:i + 28] for i in range(0, order.nbytes, 28)) == \
        sorted(ia["blocks"].tobytes()[i:i + 28] for i in range(0, ia["blocks"].nbytes, 28))
    a, b = _port(pic), _port(pic, order)
    for c in range(3):
        assert np.array_equal(a[c], b[c])
    if cw > 2 and ch > 1:                                   # not the decoding order: a later CTU row starts before row 0 ends
        first_row1 = next(i for i, r in enumerate(order) if r["y0"] >> (ctb - (1 if r["c_idx"] else 0)) == 1)
        last_row0 = max(i for i, r in enumerate(order) if r["y0"] >> (ctb - (1 if r["c_idx"] else 0)) == 0)
        assert first_row1 < last_row0


@pytest.mark.parametrize("order", [lib.INTRA_ORDER_DIAGONAL, lib.INTRA_ORDER_DEPTH])
def test_phased_schedule_groups_the_records_and_stays_topological(order):
    """phase_counts: luma records, then chroma add-only records, then the other chroma records, each group in the
    requested order; read front to back the array is still an order the port reconstructs identically from."""
    pic = synth.full_picture(5, 256, 192, ctb_log2=6, lmcs=True, intra_fraction=0.3)
    ia = pic["intra"]
    w, h, ctb = 256, 192, 6
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    plain = lib.intra_schedule(ia["blocks"], ia["ctu_first"], cw, ch, ctb, w, h, order=order)
    recs, (n_luma, n_add) = lib.intra_schedule(ia["blocks"], ia["ctu_first"], cw, ch, ctb, w, h, order=order, phases=True)
    add_only = (recs["flags"] & abi.IB_ADD_ONLY) != 0
    assert n_add > 0 and (recs["c_idx"][:n_luma] == 0).all()
    assert add_only[n_luma:n_luma + n_add].all() and (recs["c_idx"][n_luma:] > 0).all() and not add_only[n_luma + n_add:].any()
    phase = np.where(plain["c_idx"] == 0, 0, np.where((plain["flags"] & abi.IB_ADD_ONLY) != 0, 1, 2))
    assert recs.tobytes() == np.concatenate([plain[phase == k] for k in range(3)]).tobytes()      # a stable partition
    from oracle import portapi
    def port(order_recs):
        planes = [np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]]
        util.oracle_inter(planes, pic["inter"], pic["lmcs"][0], False)
        res = [np.zeros(p.shape, np.int16) for p in planes]
        portapi.residual_batch(planes, pic["residual"]["levels"], pic["residual"]["tbs"], pic["residual"]["params"], residual_planes=res)
        portapi.intra_picture(planes, res, order_recs, ia["params"])
        return planes
    a, b = port(ia["blocks"]), port(recs)
    for c in range(3):
        assert np.array_equal(a[c], b[c])


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed,w,h,ctb", [(61, 208, 120, 6), (62, 136, 72, 5), (63, 256, 192, 7)])
def test_port_matches_reference_on_dual_tree_pictures(seed, w, h, ctb):
    """I pictures with separate luma and chroma partitions (DUAL_TREE_LUMA / DUAL_TREE_CHROMA units): chroma blocks have
    their own availability and take their DM mode / CCLM luma from a differently partitioned luma plane."""
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, tools=ALL_TOOLS, dual_tree=True)
    want = [np.zeros((hh, ww), np.uint16) for (ww, hh) in pic["dims"]]
    refapi.recon_picture(want, pic["ref"], pic["intra"]["params"], inter=pic["inter"])
    got = _port(pic)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c
