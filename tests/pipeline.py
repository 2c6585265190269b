"""The whole reconstruction + loop-filter chain of one picture, on the oracle and on the GPU engine."""
import numpy as np

from ffmpeg_b200 import lib
from oracle import portapi, refapi
from tests import util


def oracle_picture(pic):
    """Reference stage functions in stage order (vvc_thread.c:41-51): INTER, RECON (residual), LMCS, DEBLOCK_V,
    DEBLOCK_H, SAO, ALF.  Uses the compiled reference when present, else the C port."""
    planes = [np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]]
    it, rs, db, lf = pic["inter"], pic["residual"], pic["deblock"], pic["lf"]
    fwd, inv = pic["lmcs"] if pic["lmcs"] is not None else (None, None)
    edge = lib.pack_deblock_edges(db["bs"], db["len_p"], db["len_q"], db["tb_w_c"], db["tb_h_c"], pic["ctb_log2"])
    ia = pic.get("intra")
    if refapi.available():
        util.oracle_inter(planes, it, fwd, True)
        if ia is not None:        # one partition of inter and intra CUs: the reference's own RECON stage does both
            refapi.recon_picture(planes, pic["ref"], ia["params"], inter=it, lmcs_fwd_lut=fwd)
        else:
            refapi.residual_batch(planes, rs["levels"], rs["tbs"], rs["params"], rs["intra_mode"])
        if inv is not None:
            refapi.lmcs_picture(planes, inv, pic["ctb_log2"], pic.get("bd", 10))
        refapi.deblock_picture(planes, db)
        refapi.sao_picture(planes, lf)
        refapi.alf_picture(planes, lf)
        return planes
    util.oracle_inter(planes, it, fwd, False)
    if ia is not None:
        res = [np.zeros(p.shape, np.int16) for p in planes]
        portapi.residual_batch(planes, rs["levels"], rs["tbs"], rs["params"], residual_planes=res)
        portapi.intra_picture(planes, res, ia["blocks"], ia["params"])
    else:
        portapi.residual_batch(planes, rs["levels"], rs["tbs"], rs["params"])
    if inv is not None:
        portapi.lmcs_picture(planes, inv)
    portapi.deblock_picture(planes, edge, db["qp"], db["ctb_offsets"], db["stride4"], db["params"])
    tmp = [np.zeros_like(p) for p in planes]
    portapi.sao_picture(tmp, planes, lf)
    portapi.alf_picture(planes, tmp, lf)
    return planes


from ffmpeg_b200.hostutil import EnginePicture  # noqa: E402,F401  (kept importable from here)
