"""TEST INFRASTRUCTURE — plain-Python restatement of the boundary-strength / filter-length derivation
(vvc_deblock_bs and helpers, libavcodec/vvc/vvc_filter.c:308-782, max_filter_length_chroma 796-817) over the engine's
compact per-4x4 tables (VVCB200MotionUnit / VVCB200PartitionUnit), producing the packed edge bytes.  Pure-Python loops:
small pictures only.  Parity pinned: tests/test_oracle_deblock.py compares it with the compiled reference's own
vvc_deblock_bs (oracle/_ref) folded by vvcb200_pack_deblock_edges.  Never imported by the product."""
import numpy as np


def _far(a, b):
    return abs(int(a[0]) - int(b[0])) >= 8 or abs(int(a[1]) - int(b[1])) >= 8


def motion_bs(c, n):
    """boundary_strength(), vvc_filter.c:308-372; reference pictures compared through ref_id."""
    if c["pred_flag"] == 3 and n["pred_flag"] == 3:
        if c["ref_id"][0] == n["ref_id"][0] and c["ref_id"][0] == c["ref_id"][1] and n["ref_id"][0] == n["ref_id"][1]:
            return int((_far(n["mv"][0], c["mv"][0]) or _far(n["mv"][1], c["mv"][1])) and
                       (_far(n["mv"][1], c["mv"][0]) or _far(n["mv"][0], c["mv"][1])))
        if n["ref_id"][0] == c["ref_id"][0] and n["ref_id"][1] == c["ref_id"][1]:
            return int(_far(n["mv"][0], c["mv"][0]) or _far(n["mv"][1], c["mv"][1]))
        if n["ref_id"][1] == c["ref_id"][0] and n["ref_id"][0] == c["ref_id"][1]:
            return int(_far(n["mv"][1], c["mv"][0]) or _far(n["mv"][0], c["mv"][1]))
        return 1
    if c["pred_flag"] != 3 and n["pred_flag"] != 3:
        lc = 0 if c["pred_flag"] & 1 else 1
        ln = 0 if n["pred_flag"] & 1 else 1
        if c["ref_id"][lc] != n["ref_id"][ln]:
            return 1
        return int(_far(c["mv"][lc], n["mv"][ln]))
    return 1


def derive_edges(motion, part, width, height, ctb_log2, ctb_edges=None):
    """-> packed edge bytes [2][3][rows4][stride4] (bs | lenP << 2 | lenQ << 5), as vvcb200_deblock_derive_edges."""
    r4, s4 = motion.shape
    out = np.zeros((2, 3, r4, s4), np.uint8)
    ctb_w = (width + (1 << ctb_log2) - 1) >> ctb_log2
    mask = (1 << ctb_log2) - 1
    for uy in range(r4):
        for ux in range(s4):
            x, y = ux * 4, uy * 4
            if x >= width or y >= height:
                continue
            q, mq = part[uy, ux], motion[uy, ux]
            ce = 0 if ctb_edges is None else int(ctb_edges[(y >> ctb_log2) * ctb_w + (x >> ctb_log2)])
            sb_cu = mq["pred_flag"] != 0 and bool(q["cu_off_x"] & 32)
            for d in range(2):
                pos = y if d else x
                if pos == 0:
                    continue
                p, mp = (part[uy - 1, ux], motion[uy - 1, ux]) if d else (part[uy, ux - 1], motion[uy, ux - 1])
                barred = not (pos & mask) and bool(ce & (2 if d else 1))
                intra_side = mq["pred_flag"] == 0 or mp["pred_flag"] == 0 or mq["ciip_flag"] or mp["ciip_flag"]
                tb_off = int(q["tb_off"]) >> 4 if d else int(q["tb_off"]) & 15
                cu_off = int(q["cu_off_y"]) if d else int(q["cu_off_x"]) & 31
                log2_q = (int(q["tb_size"]) >> 3) & 7 if d else int(q["tb_size"]) & 7
                log2_p = (int(p["tb_size"]) >> 3) & 7 if d else int(p["tb_size"]) & 7
                has_sb = sb_cu and bool(int(q["cu_off_x"]) & (128 if d else 64))
                if tb_off == 0:
                    if not barred:
                        if (q["tb_size"] & 128) and (p["tb_size"] & 128):
                            bs = 0
                        elif intra_side:
                            bs = 2
                        elif (q["tb_size"] & 64) or (p["tb_size"] & 64):
                            bs = 1
                        elif cu_off and ((cu_off & 1) or not has_sb):
                            bs = 0
                        else:
                            bs = motion_bs(mq, mp)
                        if log2_p <= 2 or log2_q <= 2:
                            lp = lq = 1
                        else:
                            lp, lq = (7 if log2_p >= 5 else 3), (7 if log2_q >= 5 else 3)
                        if has_sb:
                            lq = min(5, lq)
                        if p["cu_off_x"] & 32:
                            lp = min(5, lp)
                        out[d, 0, uy, ux] = bs | (lp << 2) | (lq << 5)
                elif sb_cu and not (cu_off & 1):
                    i, size = tb_off << 2, 1 << log2_q
                    ln = 1 if i in (4, size - 4) else 2 if i in (8, size - 8) else 3
                    out[d, 0, uy, ux] = motion_bs(mq, mp) | (ln << 2) | (ln << 5)
                ctb_off = int(q["ctb_off"]) >> 4 if d else int(q["ctb_off"]) & 15
                if ctb_off == 0 and not (pos & 15) and not barred:
                    pcm = bool(q["cflags"] & 2) and bool(p["cflags"] & 2)
                    joint = bool(q["cflags"] & 1) or bool(p["cflags"] & 1)
                    sz_q = (int(q["ctb_size"]) >> 3) & 7 if d else int(q["ctb_size"]) & 7
                    sz_p = (int(p["ctb_size"]) >> 3) & 7 if d else int(p["ctb_size"]) & 7
                    big = sz_p >= 3 and sz_q >= 3
                    ctu_edge = bool(d) and not (y & mask)
                    for c in (1, 2):
                        cbf = bool((int(q["ctb_size"]) | int(p["ctb_size"])) & (64 if c == 1 else 128)) or joint
                        bs = 0 if pcm else 2 if intra_side else 1 if cbf else 0
                        lq = 3 if big else int(bs == 2)
                        lp = (1 if ctu_edge else 3) if big else int(bs == 2)
                        out[d, c, uy, ux] = bs | (lp << 2) | (lq << 5)
    return out
