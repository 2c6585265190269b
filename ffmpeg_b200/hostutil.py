"""Host-side helpers around the C ABI that tests, bench.py and smoke() share.  Nothing here touches oracle/."""
import hashlib

import numpy as np

from . import abi, lib


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def plane_md5(planes):
    """Per-plane MD5 of the visible samples (what the reference's framemd5 hashes)."""
    return [md5(np.ascontiguousarray(p)) for p in planes]


def to_device_planes(planes_np, torch, align=64):
    """Upload numpy planes into pitch-padded torch CUDA tensors (views of width w)."""
    out = []
    for p in planes_np:
        h, w = p.shape
        stride = (w + align - 1) // align * align
        t = torch.zeros((h, stride), dtype=torch.int16, device="cuda")
        t[:, :w] = torch.from_numpy(p.view(np.int16)).cuda()
        out.append(t[:, :w])
    return out


def from_device_planes(tensors):
    return [t.cpu().numpy().view(np.uint16) for t in tensors]


class EnginePicture:
    """Pinned host copies of one picture's batches + the Level-B call sequence (HOST buffers in, HOST planes out)."""

    def __init__(self, ctx, pic, max_slots=4, want_mv=True, sparse_levels="auto", arena=True):
        it, rs, db, lf = pic["inter"], pic["residual"], pic["deblock"], pic["lf"]
        self.ctx, self.pic = ctx, pic
        self.n_refs = len(it["refs"])
        self.refs = [[ctx.pinned_copy(np.ascontiguousarray(p)) for p in r] for r in it["refs"]]
        # the picture's batches in ONE pinned block the context mirrors on the device, laid out in call order: the arrays of
        # a stage call then travel as one copy (vvcb200_ctx_arena_alloc); arena=False: one pinned allocation per array
        self.arena = None
        if arena:
            ar = self.arena = ctx.arena(self._arena_bytes(pic))
            put = ar.copy
        else:
            put = ctx.pinned_copy
        self.blocks = put(it["blocks"])
        prof = it.get("prof")
        self.prof = put(prof) if prof is not None and len(prof) else None
        luts = None if pic["lmcs"] is None else pic["lmcs"]
        lut_fwd = put(luts[0]) if luts is not None else None
        # levels travel as dense scan boxes or in the sparse form (bitmap + non-zero values), whichever is smaller for the
        # picture (the synthetic pictures fill their scan boxes to 95 %, real residuals to a third or less)
        self.res_params = rs["params"]
        levels, tbs = rs["levels"], rs["tbs"]
        if sparse_levels and len(tbs):
            # blocks whose levels fit int8 travel as int8 (VVCB200_TB_LEVELS8) in either form
            d_levels, d_tbs = lib.pack_levels(levels, tbs, sparse=False, narrow=True)
            sp_levels, sp_tbs = lib.pack_levels(levels, tbs, sparse=True, narrow=True)
            if sparse_levels is True or sp_levels.nbytes < d_levels.nbytes:   # "auto": whichever form is smaller, as an emitter would
                levels, tbs = sp_levels, sp_tbs
                self.res_params = lib.sparse_params(rs["params"])
            else:
                levels, tbs = d_levels, d_tbs
        self.levels = put(levels)
        self.tbs = put(tbs)
        ia = pic.get("intra")
        self.intra = None
        if ia is not None and len(ia["blocks"]):
            ctb = pic["ctb_log2"]
            self.intra = (put(ia["blocks"]), put(ia["ctu_first"]),
                          (pic["width"] + (1 << ctb) - 1) >> ctb, (pic["height"] + (1 << ctb) - 1) >> ctb, ia["params"])
        self.luts = None if luts is None else [lut_fwd, put(luts[1])]
        edge = lib.pack_deblock_edges(db["bs"], db["len_p"], db["len_q"], db["tb_w_c"], db["tb_h_c"], pic["ctb_log2"])
        self.edge = put(edge)
        self.qp = put(db["qp"])
        self.ctb_offsets = put(db["ctb_offsets"])
        self.sao = put(lf["sao"])
        self.alf = put(lf["alf"])
        self.alf_params = put(lf["alf_params"])
        # pictures are reconstructed into the slots after the references (ref_slot[][] already maps to 0..n_refs-1),
        # rotating over up to four slots / host frames when the context has them, as consecutive pictures of a decoder
        # do: the download of one picture then overlaps the reconstruction of the next ones (two compute lanes)
        # DMVR feedback: the vectors actually used come back every picture (the host needs them for later MV prediction)
        self.refined = [ctx.pinned((max(len(it["blocks"]), 1), 2, 2), np.int32) for _ in range(4)]
        self.want_mv = want_mv
        self.n_slots = max(1, min(max_slots, ctx.dpb_size - self.n_refs))
        self.outs = [ctx.pinned_frame() for _ in range(self.n_slots)]
        self.count = 0
        self.out = self.outs[0]

    @staticmethod
    def _arena_bytes(pic):
        """Room for every batch array of the picture (the level buffer at its dense size: the packed forms are smaller)."""
        total = 0
        for a in (pic["inter"]["blocks"], pic["inter"].get("prof"), pic["residual"]["levels"], pic["residual"]["tbs"],
                  pic["deblock"]["qp"], pic["deblock"]["ctb_offsets"], pic["lf"]["sao"], pic["lf"]["alf"], pic["lf"]["alf_params"]):
            total += 0 if a is None else a.nbytes + 256
        total += 6 * pic["deblock"]["qp"][0].nbytes + 256                       # the packed edge tables
        total += pic["residual"]["levels"].nbytes // 16 + 8 * len(pic["residual"]["tbs"]) + 1024   # sparse form's worst case
        if pic.get("intra") is not None:
            total += pic["intra"]["blocks"].nbytes + pic["intra"]["ctu_first"].nbytes + 512
        if pic["lmcs"] is not None:
            total += sum(l.nbytes + 256 for l in pic["lmcs"])
        return total + 4096

    def h2d_bytes(self):
        intra = 0 if self.intra is None else self.intra[0].nbytes + self.intra[1].nbytes
        intra += 0 if self.prof is None else self.prof.nbytes
        return intra + int(self.blocks.nbytes + self.levels.nbytes + self.tbs.nbytes + self.edge.nbytes + self.qp.nbytes +
                   self.ctb_offsets.nbytes + self.sao.nbytes + self.alf.nbytes + self.alf_params.nbytes)

    def picture_bytes(self):
        return int(sum(p.nbytes for p in self.out))

    def refined_bytes(self):
        """What comes back of the refined vectors: every record's, or — batch in the recommended order — the DMVR records'."""
        if not len(self.blocks) or not self.want_mv:
            return 0
        if self.pic["inter"]["params"].dmvr_first:
            return 16 * int(((self.pic["inter"]["blocks"]["flags"] & abi.PB_DMVR) != 0).sum())
        return int(self.refined[0].nbytes)

    def d2h_bytes(self):
        return self.picture_bytes() + self.refined_bytes()

    def upload_refs(self):
        for i, r in enumerate(self.refs):
            self.ctx.upload(r, slot=i)

    def run(self, download=True, only_download=False):
        ctx, pic = self.ctx, self.pic
        k = self.count % self.n_slots
        self.count += 1
        self.out = self.outs[k]
        if only_download:                  # the copy of a finished picture alone (what bench.py times as the D2H leg)
            ctx.download(self.out, slot=self.n_refs + k)
            return
        ctx.begin(self.n_refs + k)
        if len(self.blocks):
            ctx.inter(self.blocks, pic["inter"]["params"], None if self.luts is None else self.luts[0],
                      refined_mv=self.refined[self.count % 4] if self.want_mv else None, prof=self.prof)
        ctx.residual(self.levels, self.tbs, self.res_params)
        if self.intra is not None:
            ctx.intra(*self.intra)
        if self.luts is not None:
            ctx.lmcs(self.luts[1])
        ctx.deblock(self.edge, self.qp, self.ctb_offsets, pic["deblock"]["stride4"], pic["deblock"]["params"])
        ctx.sao_alf(self.sao, self.alf, self.alf_params, pic["ctb_log2"])
        if download:
            ctx.download(self.out)
