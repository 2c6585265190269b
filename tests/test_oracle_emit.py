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
