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
