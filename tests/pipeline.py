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


def emitted_picture(pic):
    """SURVEY f1 end to end: the picture's batches re-made by the product's emitter (ffmpeg_b200/emit) from the reference's
    own CodingUnit / TransformUnit / MvField / fc->tab.* state, as the harness under oracle/ref builds it from `pic`.
    Returns a copy of `pic` whose inter, residual, intra and loop-filter batches are the emitter's output."""
    from ffmpeg_b200 import abi, synth
    w, h = pic["width"], pic["height"]
    it, ia, lf, db = pic["inter"], pic["intra"], pic["lf"], pic["deblock"]
    # stage INTER: regular / CIIP sub-blocks as one-sub-block coding units, then the affine / GPM coding units
    b = it["blocks"]
    plain = b[(b["flags"] & (abi.PB_AFFINE | abi.PB_GPM | abi.PB_LUMA_ONLY | abi.PB_CHROMA_ONLY)) == 0]
    cus = np.zeros(len(plain), synth.CU_DTYPE)
    for k in ("x0", "y0", "w", "h", "pred_flag", "bcw_idx", "ref_idx", "mv"):
        cus[k] = plain[k]
    cus["flags"] = plain["flags"]
    cus["sb_mv_off"] = 0
    all_cus = np.concatenate([cus, it["cus"]]) if len(it["cus"]) else cus
    blocks, prof, ip = refapi.emit_inter_cus(w, h, all_cus, it["sb_mv"], it["params"], order=True)
    # stage RECON
    has_ciip = bool(pic["ref"]["cus"]["ciip_flag"].any())
    tbs, lv, ibs, first = refapi.emit_recon_picture(w, h, pic["ref"], ia["params"], inter=it if has_ciip else None)
    order, cls = abi.tb_order(tbs)
    rp = type(pic["residual"]["params"]).from_buffer_copy(bytes(pic["residual"]["params"]))
    for c in range(4):
        rp.n_class[c] = int(np.sum(cls == c))
    # loop filters
    sao, alf, ap, off, _, dp = refapi.emit_loopfilter_ctbs(lf, w, h, db["ctb_offsets"], db["params"])
    out = dict(pic)
    out["inter"] = dict(it, blocks=blocks, prof=prof, params=ip)
    out["residual"] = dict(pic["residual"], tbs=np.ascontiguousarray(tbs[order]), levels=lv, params=rp)
    out["intra"] = dict(ia, blocks=ibs, ctu_first=first)
    out["lf"] = dict(lf, sao=sao, alf=alf, alf_params=ap)
    out["deblock"] = dict(db, ctb_offsets=off, params=dp)
    return out
