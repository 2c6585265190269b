"""SURVEY f1 — the descriptor emitter (ffmpeg_b200/emit/vvcb200_emit.c) against the reference's own structures.
The harness (oracle/ref) turns a synthetic description into the reference's CodingUnit / PredictionUnit / TransformUnit /
MvField state, as it does for the parity tests; the emitter — product code, compiled against the reference's headers —
walks THAT state and must write the very batches the engine is tested with.  CPU only; needs oracle/_ref."""
import numpy as np
import pytest

from ffmpeg_b200 import abi, synth
from oracle import refapi

pytestmark = pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")


@pytest.mark.parametrize("seed,lmcs", [(1, True), (2, False), (3, True)])
def test_inter_emitter_writes_the_batch_of_regular_affine_and_gpm_units(seed, lmcs):
    """vvcb200_emit_inter_cu in predict_inter()'s place: <= 16x16 sub-blocks, 4x4 affine records + derive_affine_mvc chroma
    records + PROF tables, GPM records with their aux word — byte for byte the synthetic generator's batch, order included."""
    pic = synth.inter_cu_picture(seed, 416, 240, lmcs=lmcs)
    blocks, prof, p = refapi.emit_inter_cus(416, 240, pic["cus"], pic["sb_mv"], pic["params"], order=True)
    assert blocks.tobytes() == pic["blocks"].tobytes()
    assert prof.tobytes() == pic["prof"].tobytes()
    assert (p.n_w16, p.n_w8) == (pic["params"].n_w16, pic["params"].n_w8)


def test_inter_emitter_carries_dmvr_bdof_bcw_and_weighted_prediction():
    """Block-level tool mix (bi / weighted / BCW / half-pel / DMVR / BDOF): every sub-block as a one-sub-block coding unit."""
    pic = synth.inter_picture(7, 416, 240, tools=("bi", "wp", "bcw", "hpel", "dmvr", "bdof"))
    b = pic["blocks"]
    cus = np.zeros(len(b), synth.CU_DTYPE)
    for k in ("x0", "y0", "w", "h", "pred_flag", "bcw_idx", "ref_idx", "mv"):
        cus[k] = b[k]
    cus["flags"] = b["flags"]
    blocks, prof, _ = refapi.emit_inter_cus(416, 240, cus, np.zeros((0, 2, 2), np.int32), pic["params"])
    assert len(prof) == 0 and blocks.tobytes() == b.tobytes()


def _diff(a, b, names):
    """First differing record of two structured arrays, for the assertion message."""
    if len(a) != len(b):
        return "length %d != %d" % (len(a), len(b))
    for i in range(len(a)):
        if a[i].tobytes() != b[i].tobytes():
            return "record %d: %s" % (i, {k: (a[i][k].tolist(), b[i][k].tolist()) for k in names if np.any(a[i][k] != b[i][k])})
    return ""


RECON_CASES = {
    "intra tools": dict(seed=11, tools=("mrl", "mip", "cclm", "pdpc", "isp", "bdpcm", "mts", "ts", "lfnst", "joint"),
                        intra_fraction=0.7),
    "all intra dual tree": dict(seed=12, tools=("mrl", "mip", "cclm", "isp", "bdpcm", "mts", "ts", "lfnst"), dual_tree=True),
    "dep quant + scaling lists": dict(seed=20, tools=("mrl", "mip", "cclm", "mts", "ts", "lfnst", "joint", "dep_quant", "scaling"),
                                      intra_fraction=0.4),
    "ciip": dict(seed=14, tools=("mrl", "mip", "cclm", "ciip", "mts", "joint"), intra_fraction=0.3, ciip_share=0.5),
    "chroma scaling": dict(seed=15, tools=("mrl", "mip", "cclm", "isp", "ciip", "mts", "ts", "joint"), intra_fraction=0.3,
                           lmcs=True, lmcs_chroma="model"),
    "ctb 64": dict(seed=16, tools=("mrl", "mip", "cclm", "isp", "mts", "lfnst", "joint"), intra_fraction=0.5, ctb_log2=6,
                   lmcs=True, lmcs_chroma="model"),
}


@pytest.mark.parametrize("name", list(RECON_CASES))
def test_recon_emitter_writes_the_transform_blocks_and_intra_records(name):
    """vvcb200_emit_recon_ctu in ff_vvc_reconstruct()'s place: transform blocks (qp, transform pair, LFNST set from the
    reference's own derive_* functions), their levels, and the intra / CIIP / add-only records with the availability the
    reference's ReconstructedArea bookkeeping reports — byte for byte what the synthetic generator hands the engine."""
    kw = dict(RECON_CASES[name])
    seed = kw.pop("seed")
    if kw.get("lmcs_chroma") == "model":
        fwd, _ = synth.lmcs_luts(seed)
        kw["lmcs_chroma"] = synth.lmcs_model(seed, fwd)
    w, h = 416, 240
    pic = synth.recon_picture(seed, w, h, **kw)
    has_ciip = bool(pic["ref"]["cus"]["ciip_flag"].any())
    tbs, lv, ibs, first = refapi.emit_recon_picture(w, h, pic["ref"], pic["intra"]["params"],
                                                    inter=pic["inter"] if has_ciip else None)
    assert not _diff(tbs, pic["ref"]["tbs"], abi.TB_DTYPE.names)
    assert np.array_equal(lv, pic["ref"]["levels"])
    assert not _diff(ibs, pic["intra"]["blocks"], abi.IB_DTYPE.names)
    assert np.array_equal(first, pic["intra"]["ctu_first"])


@pytest.mark.parametrize("seed,ctb,frac,ns", [(40, 6, 0.3, 1), (43, 5, 0.3, 3), (44, 7, 0.2, 2)])
def test_filters_emitter_writes_the_per_unit_deblocking_state(seed, ctb, frac, ns):
    """vvcb200_emit_filters_ctu over fc->tab.mvf / tb_pos / cb_pos / coded flags: the compact motion + partition records the
    device derives boundary strengths from, and the CTB sides deblocking must not cross (slice limits, lfase 0)."""
    w, h = 416, 240
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, intra_fraction=frac)
    st = synth.deblock_state(pic)
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    slices = synth.slice_map(np.random.default_rng(seed), cw, ch, ns)
    lfase = 1 if ns == 1 else 0
    motion, part, edges = refapi.emit_deblock_units(st, w, h, ctb, slices, lfase)
    assert not _diff(motion.reshape(-1), st["motion"].reshape(-1), abi.MOTION_DTYPE.names)
    assert not _diff(part.reshape(-1), st["part"].reshape(-1), abi.PARTITION_DTYPE.names)
    assert np.array_equal(edges, (synth.alf_edge_flags(slices, cw, ch, lfase) & 3).astype(np.uint8))


@pytest.mark.parametrize("ns,lfase,tiles,lfate", [(1, 1, None, 1), (4, 0, None, 1), (3, 1, ((0, 2), (0, 1)), 0), (3, 0, ((0, 1, 3), (0, 2)), 0),
                                                   (1, 1, ((0, 2), (0, 1)), 1)])
def test_filters_emitter_writes_sao_alf_and_deblocking_parameters(ns, lfase, tiles, lfate):
    """SAOParams / ALFParams / APS list / DBParams / LADF fields -> VVCB200SaoCtb / AlfCtb / AlfParams / offsets / DeblockParams,
    with the SAO restore flags (vvc_filter.c:163-211) and the ALF edge flags (1266-1280) worked out from slice_idx, the tile
    grid and the two loop-filter-across flags."""
    w, h, ctb = 832, 480, 7
    kw = dict(tile_cols=tiles[0], tile_rows=tiles[1], lfate=lfate) if tiles else {}
    t = synth.loopfilter_ctbs(50 + ns, w, h, ctb_log2=ctb, n_slices=ns, lfase=lfase, **kw)
    d = synth.deblock_picture(60, w, h, ctb_log2=ctb, ladf=True)
    sao, alf, ap, off, edges, dp = refapi.emit_loopfilter_ctbs(t, w, h, d["ctb_offsets"], d["params"])
    assert not _diff(sao, t["sao"], abi.SAO_DTYPE.names)
    assert not _diff(alf, t["alf"], abi.ALF_CTB_DTYPE.names)
    assert ap.tobytes() == t["alf_params"].tobytes()
    assert np.array_equal(off, d["ctb_offsets"])
    assert np.array_equal(edges, t["alf"]["edges"] & 3)
    assert bytes(dp) == bytes(d["params"])


def test_refined_vectors_go_back_into_the_dmvr_table():
    """run_inter's other half: the vectors DMVR settled on (vvcb200_pic_inter's refined_mv) written into
    fc->ref->tab_dmvr_mvf as set_dmvr_info (vvc_inter.c:749-761) does — same units, same MvField."""
    pic = synth.inter_picture(9, 416, 240, tools=("bi", "dmvr", "bdof", "hpel"))
    dst = [np.zeros_like(p) for p in pic["refs"][0]]
    bad, mv = refapi.emit_store_refined_check(dst, pic["refs"], pic["blocks"], pic["params"])
    dm = (pic["blocks"]["flags"] & abi.PB_DMVR) != 0
    assert dm.sum() > 20 and np.any(mv[dm] != pic["blocks"]["mv"][dm])       # DMVR moved something
    assert bad == 0


def test_recon_emitter_into_caller_arrays():
    """VVCB200EmitBatch.fixed: the arrays are regions the caller owns (a context arena): same records, nothing allocated;
    an array that is too small is an error (VVCB200_ERR_TOO_BIG), not a write past its end."""
    w, h = 416, 240
    pic = synth.recon_picture(11, w, h, tools=("mrl", "mip", "cclm", "isp", "mts", "lfnst", "joint"), intra_fraction=0.7)
    tbs, lv, ibs, first = refapi.emit_recon_picture(w, h, pic["ref"], pic["intra"]["params"], fixed=True)
    assert tbs.tobytes() == pic["ref"]["tbs"].tobytes() and np.array_equal(lv, pic["ref"]["levels"])
    assert ibs.tobytes() == pic["intra"]["blocks"].tobytes() and np.array_equal(first, pic["intra"]["ctu_first"])
    with pytest.raises(RuntimeError, match="-2"):
        refapi.emit_recon_picture(w, h, pic["ref"], pic["intra"]["params"], fixed=True, shrink=5)
