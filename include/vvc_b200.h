/*
 * vvc_b200.h — C ABI of libvvcb200.so: the B200 (sm_100a) VVC reconstruction and
 * in-loop-filter engine that sits behind the reference decoder's VVCDSPContext surface.
 *
 * Boundary (reference file:line, relative to the FFmpeg/ffvvc tree):
 *   - the operator table this library stands in for: VVCDSPContext, libavcodec/vvc/vvcdsp.h:160-168,
 *     bound by ff_vvc_dsp_init (vvcdsp.c:228-257) with the per-arch hook at vvcdsp.c:254-256;
 *   - the stage drivers whose bodies become descriptor emitters: run_inter … run_alf,
 *     libavcodec/vvc/vvc_thread.c:432-566.
 *
 * Two levels are exported:
 *   Level A  "vvcb200_dsp_*": one batched device entry per populated VVCDSPContext member, same
 *            argument meaning as the table entry, arrays of block descriptors instead of one block.
 *            Pointers named d_* are DEVICE pointers; `stream` is a cudaStream_t passed as void*.
 *   Level B  "vvcb200_ctx_* / vvcb200_pic_*": the picture-level engine.  HOST buffers in, HOST planes
 *            out; H2D/D2H copies, streams and scratch planes are owned by the context.
 *
 * Conventions: every function returns 0 on success or a negative VVCB200_ERR_* code; nothing throws
 * across the ABI; no entry point falls back to a CPU implementation — if no CUDA device is usable the
 * call fails with VVCB200_ERR_CUDA.  Samples are 16-bit little-endian containers (8/10/12-bit content);
 * plane strides are in BYTES as in the reference (ptrdiff_t stride, vvcdsp.h:49-157).
 */
#ifndef VVC_B200_H
#define VVC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VVCB200_ABI_VERSION 2   /* 2: VVCB200IntraParams grew (prepared, luma_done_event), VVCB200InterParams.dmvr_first, VVCB200_TB_LEVELS8 */
#if defined(__GNUC__)
#define VVCB200_API __attribute__((visibility("default")))
#else
#define VVCB200_API
#endif

enum {
    VVCB200_OK          =  0,
    VVCB200_ERR_ARG     = -1,   /* bad argument (NULL, size out of range, unsupported bit depth) */
    VVCB200_ERR_CUDA    = -2,   /* CUDA runtime error; vvcb200_last_error() has the text          */
    VVCB200_ERR_NOMEM   = -3,
    VVCB200_ERR_UNSUP   = -4,   /* tool the reference snapshot does not implement either (IBC, palette, ACT, RPR) */
    VVCB200_ERR_TOO_BIG = -5,   /* exceeds a fixed capacity stated in this header                  */
};

/* enum TxType, vvcdsp.h:30-35 */
enum { VVCB200_DCT2 = 0, VVCB200_DST7 = 1, VVCB200_DCT8 = 2 };

VVCB200_API const char *vvcb200_last_error(void);
VVCB200_API int         vvcb200_abi_version(void);
/* Number of kernels this library has launched since load (all contexts, all streams). */
VVCB200_API uint64_t    vvcb200_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Residual path descriptors: dequant → (BDPCM) → (LFNST) → inverse DCT-II/DST-VII/DCT-VIII → add.
 * One record per coded TransformBlock (vvc_ctu.h:136-158); what itransform() (vvc_intra.c:432-478)
 * reads from TransformBlock/TransformUnit/CodingUnit is flattened into it by the host.
 * ---------------------------------------------------------------------------------------------- */
enum {
    VVCB200_TB_TS          = 1 << 0,  /* transform_skip_flag (tb->ts): bdShift 10, flat scaling, no inverse transform   */
    VVCB200_TB_BDPCM       = 1 << 1,  /* cu->bdpcm_flag[c_idx]: prefix sum before dequant (vvcdsp_template.c:76-95)      */
    VVCB200_TB_BDPCM_VERT  = 1 << 2,  /* intra mode == INTRA_VERT (vvc_intra.c:419-430)                                  */
    VVCB200_TB_DEP_QUANT   = 1 << 3,  /* sh_dep_quant_used_flag of the slice (vvc_intra.c:316,336)                       */
    VVCB200_TB_SCALING     = 1 << 4,  /* explicit scaling list applies to this TB (vvc_intra.c:367-369)                  */
    VVCB200_TB_INTER       = 1 << 5,  /* cu->pred_mode != MODE_INTRA — picks the Table-38 row (vvc_intra.c:342-362)       */
    VVCB200_TB_JOINT       = 1 << 6,  /* chroma TB carrying a joint CbCr residual: also rebuilt into plane 3 - c_idx      */
    VVCB200_TB_JOINT_SHIFT = 1 << 7,  /* tu->coded_flag[1] ^ tu->coded_flag[2] (vvc_intra.c:172)                          */
    VVCB200_TB_LMCS_SCALE  = 1 << 8,  /* LMCS chroma residual scaling applies (vvc_intra.c:451); scale in `lmcs_scale`    */
    VVCB200_TB_TO_RESIDUAL = 1 << 9,  /* store the residual in the int16 residual planes instead of adding it to dst: the
                                         TB belongs to an intra / CIIP block whose prediction is not there yet            */
    VVCB200_TB_LEVELS8     = 1 << 10, /* every level of this TB lies in [-128, 127] and is stored as int8: the dense box, or
                                         the non-zero list of the sparse form, two levels per int16 slot of the level buffer
                                         (padded to a whole slot; sparse: to an even number of slots).  Most blocks of most
                                         streams: halves what they upload (vvcb200_pack_levels)                            */
};

typedef struct VVCB200TransformBlock {
    uint16_t x0, y0;        /* top-left corner in samples OF THE PLANE c_idx (reference keeps luma units; host shifts)   */
    uint8_t  c_idx;         /* 0 Y, 1 Cb, 2 Cr                                                                          */
    uint8_t  log2_w, log2_h;/* 0..6 (w or h == 1 only with the other >= 16, vvcdsp.c:140-195)                            */
    uint8_t  nzw, nzh;      /* max_scan_x + 1, max_scan_y + 1 as residual parsing left them (1..32)                      */
    uint8_t  trh, trv;      /* TxType after derive_transform_type (vvc_intra.c:130-164)                                  */
    uint8_t  qp;            /* tb->qp after derive_qp (vvc_intra.c:277-310)                                              */
    uint16_t flags;         /* VVCB200_TB_*                                                                              */
    uint8_t  lfnst;         /* 0 = no LFNST; else bits1:0 = lfnst_idx (1|2), bits3:2 = lfnst_tr_set_idx, bit4 = transpose
                               (pred_mode_intra > 34 after wide-angle mapping, vvc_intra.c:76-77)                         */
    uint8_t  reserved;
    uint16_t lmcs_scale;    /* chroma_scale_coeff picked for this TB's VPDU (vvc_intra_template.c:390-428); 0 if unused  */
    uint16_t reserved2;
    uint32_t coeff_off;     /* index into the int16 level buffer of this TB's nzw*nzh levels, row-major, row stride nzw  */
} VVCB200TransformBlock;    /* 24 bytes */

/* Per-picture constants of the residual path (fields of VVCSPS/VVCPH/VVCScalingList the path reads). */
typedef struct VVCB200ResidualParams {
    int32_t  bit_depth;             /* 8, 10 or 12 (sps->bit_depth)                                     */
    int32_t  log2_transform_range;  /* 15 (sps->log2_transform_range; extended precision is not built)  */
    int32_t  joint_c_sign;          /* 1 - 2 * ph_joint_cbcr_sign_flag (vvc_intra.c:171)                */
    int32_t  levels_sparse;         /* 0: coeff_off points at the dense nzw x nzh box; 1: at the sparse form of it —
                                       (nzw * nzh + 31) / 32 bitmap words of 32 bits (bit i % 32 of word i / 32: level i
                                       of the box is non-zero), then the non-zero levels in box order; coeff_off even.
                                       What hls_residual_coding knows anyway (its significance map, vvc_cabac.c:2141-2392);
                                       vvcb200_pack_levels_sparse converts a dense buffer.                             */
    int32_t  n_class[4];            /* The batch is ordered by vvcb200_residual_class(): n_class[0] blocks of class 0, then
                                       class 1, 2, 3; the rest (class 4) follows.  Equal-class blocks share a warp.       */
    uint8_t  scaling_matrix_rec[28][64];   /* VVCScalingList.scaling_matrix_rec, vvc_ps.h:187-190      */
    uint8_t  scaling_matrix_dc_rec[14];    /* VVCScalingList.scaling_matrix_dc_rec                      */
    uint8_t  pad[2];
} VVCB200ResidualParams;

/* A picture (or any sample array triple) resident in device memory.  Alignment: the stand-alone stage entries accept any
 * 2-byte aligned planes unless they say otherwise; vvcb200_intra_picture wants planes and strides that are multiples of 8
 * bytes, the fused passes (vvcb200_deblock_lmcs_picture, vvcb200_sao_alf_picture) multiples of 16 — what
 * vvcb200_ctx_create allocates (256-byte rows). */
typedef struct VVCB200Planes {
    void     *d_plane[3];      /* device pointers, 16-bit samples */
    int32_t   stride[3];       /* bytes                           */
    int32_t   width[3];        /* samples                         */
    int32_t   height[3];       /* samples                         */
} VVCB200Planes;

/*
 * Level A, itx sub-table (vvcdsp.h:112-120).
 *
 * vvcb200_dsp_itx_batch: for each i, run itx.itx[trh][trv][log2_w][log2_h] in place on the w*h int32
 * block d_coeffs + block_off[i], with (nzw, nzh, log2_transform_range, bit_depth) as the table entry
 * takes them (vvcdsp.h:117-118).  Only flags==0 fields x0/y0/qp/... are ignored.
 */
VVCB200_API int vvcb200_dsp_itx_batch(int32_t *d_coeffs, const uint32_t *d_block_off,
                          const VVCB200TransformBlock *d_tbs, int n,
                          int log2_transform_range, int bit_depth, void *stream);

/* itx.add_residual (vvcdsp.h:113): dst = clip_pixel(dst + res) for each block; res is the w*h int32
 * block at d_res + block_off[i]; dst addressed by the TB's (c_idx, x0, y0). */
VVCB200_API int vvcb200_dsp_add_residual_batch(const VVCB200Planes *dst, const int32_t *d_res, const uint32_t *d_block_off,
                                   const VVCB200TransformBlock *d_tbs, int n, int bit_depth, void *stream);

/*
 * Fused residual stage = itransform() for a whole picture (vvc_intra.c:432-478): BDPCM, dequant, LFNST,
 * 2-D/1-D inverse transform, LMCS chroma scaling factor, add_residual, joint CbCr — one launch.
 * d_levels: int16 level buffer (see coeff_off).  The TBs of one launch must not overlap in the planes.
 * TBs flagged VVCB200_TB_TO_RESIDUAL (all TBs if `dst` is NULL) store their residual, clipped to int16, in
 * `d_residual` (same geometry as `dst`, int16 samples) instead of adding it: the wavefront intra stage adds it
 * once the prediction exists.
 */
/* Lane-group class of a transform block (host helper, no CUDA): 0..3 = 2-D transform done with 32 / 16 / 8 / 4 lanes
 * (largest side >= 32 / 16 / 8 / <= 4), 4 = generic (transform skip, BDPCM, LFNST, 1-D).  Emitters keep one list per class. */
VVCB200_API int vvcb200_residual_class(const VVCB200TransformBlock *tb);

/* Host helper (pure CPU): rewrite a dense level buffer in the sparse form.  `tbs` (n records) are
 * updated in place: coeff_off then indexes `out`.  `out` needs room for the worst case, n_dense + 2 * n + 2 * (n_dense / 32)
 * int16; returns the number of int16 written through out_len.  Typical residuals (a fifth of the scan box non-zero)
 * shrink 3-4x, which is most of what a picture uploads. */
VVCB200_API int vvcb200_pack_levels_sparse(const int16_t *dense, VVCB200TransformBlock *tbs, int n, int16_t *out,
                                           size_t out_cap, size_t *out_len);
/* The general form: sparse 0 / 1 as above; narrow != 0: blocks whose levels all fit int8 are stored that way and flagged
 * VVCB200_TB_LEVELS8 (the others stay int16).  Same worst-case room as above (dense: n_dense + n). */
VVCB200_API int vvcb200_pack_levels(const int16_t *dense, VVCB200TransformBlock *tbs, int n, int sparse, int narrow,
                                    int16_t *out, size_t out_cap, size_t *out_len);

VVCB200_API int vvcb200_residual_batch(const VVCB200Planes *dst, const VVCB200Planes *d_residual,
                           const int16_t *d_levels, const VVCB200TransformBlock *d_tbs, int n,
                           const VVCB200ResidualParams *params /* host */, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Inter prediction descriptors (stage INTER, ff_vvc_predict_inter, vvc_inter.c:899-913).
 * One record per prediction sub-block of at most 16x16 luma samples, ordered by width (see VVCB200InterParams
 * n_w16 / n_w8; an emitter simply keeps three lists): the host emitter splits larger
 * prediction units (translational MC is separable per sample, and 16x16 is the reference's own DMVR/BDOF
 * sub-block, vvc_mvs.c num_sb_x/y), and emits affine CUs as 4x4 luma + 8x8-luma-footprint chroma records.
 * ---------------------------------------------------------------------------------------------- */
enum {
    VVCB200_PB_WP          = 1 << 0,  /* explicit weighted prediction is on for the slice (vvc_inter.c:128-129,149-150) */
    VVCB200_PB_HPEL_IF     = 1 << 1,  /* mi.hpel_if_idx: luma filter set 1 (vvc_inter.c:792)                           */
    VVCB200_PB_AFFINE      = 1 << 2,  /* luma filter set 2 (vvc_inter.c:383-384)                                       */
    VVCB200_PB_LUMA_ONLY   = 1 << 3,
    VVCB200_PB_CHROMA_ONLY = 1 << 4,
    VVCB200_PB_BDOF        = 1 << 5,  /* pu->bdof_flag                                                                 */
    VVCB200_PB_DMVR        = 1 << 6,  /* pu->dmvr_flag                                                                 */
    VVCB200_PB_CIIP        = 1 << 7,  /* cu->ciip_flag: BCW is ignored (vvc_inter.c:151), blend happens in RECON       */
    VVCB200_PB_LMCS        = 1 << 8,  /* forward LMCS on the luma prediction (vvc_inter.c:888-891)                     */
    VVCB200_PB_GPM         = 1 << 9,  /* geometric partition (pred_gpm_blk, vvc_inter.c:466-521): `aux` is VVCB200_GPM_AUX  */
    VVCB200_PB_PROF0       = 1 << 10, /* pu->cb_prof_flag[0]: PROF on list 0 (affine 4x4 luma records); `aux` = table index */
    VVCB200_PB_PROF1       = 1 << 11, /* pu->cb_prof_flag[1]                                                               */
};

/* `aux` of a GPM record.  A GPM coding unit is emitted as <= 16x16 records that all carry the CU's partition and
 * size plus their own offset inside the CU (the blend weight depends on the position inside the CU).  mv[i] / ref_idx[i]
 * are gpm_mv[i] (PredictionUnit.gpm_mv, vvc_ctu.h:258); list_i says which reference list gpm_mv[i] uses. */
#define VVCB200_GPM_AUX(partition_idx, log2_cb_w, log2_cb_h, off_x, off_y, list0, list1)               \
    ((uint32_t)(partition_idx) | ((uint32_t)((log2_cb_w) - 3) << 6) | ((uint32_t)((log2_cb_h) - 3) << 8) |   \
     ((uint32_t)((off_x) >> 2) << 10) | ((uint32_t)((off_y) >> 2) << 14) | ((uint32_t)(list0) << 18) | ((uint32_t)(list1) << 19))

typedef struct VVCB200PredBlock {
    uint16_t x0, y0;        /* luma position                                                   */
    uint8_t  w, h;          /* luma size, 4/8/16 each (chroma footprint is w>>1 x h>>1)        */
    uint8_t  pred_flag;     /* PF_L0 = 1, PF_L1 = 2, PF_BI = 3 (vvc_ctu.h:222-227)             */
    uint8_t  bcw_idx;       /* MvField.bcw_idx                                                 */
    int8_t   ref_idx[2];    /* MvField.ref_idx                                                 */
    uint16_t flags;         /* VVCB200_PB_*                                                    */
    int32_t  mv[2][2];      /* MvField.mv[list] = {x, y} in 1/16 luma sample, 18-bit           */
    uint32_t aux;           /* GPM: VVCB200_GPM_AUX(...); PROF: index of the CU's table in d_prof */
} VVCB200PredBlock;         /* 32 bytes */

/* PROF offsets of one affine coding unit: PredictionUnit.diff_mv_x / diff_mv_y (vvc_ctu.h:266-267, derived by the
 * host in vvc_mvs.c:361-378), values in [-31, 31]. */
typedef struct VVCB200ProfTable {
    int8_t diff_mv_x[2][16];    /* [list][y * 4 + x] */
    int8_t diff_mv_y[2][16];
} VVCB200ProfTable;             /* 64 bytes */

/* Per-slice constants of inter prediction: PredWeightTable (vvc_ps.h:135-143) and the RefPicList -> DPB map. */
typedef struct VVCB200InterParams {
    int32_t  bit_depth;
    int32_t  log2_denom[2];         /* luma, chroma                                            */
    int32_t  n_w16, n_w8;           /* the batch is ordered by block width: n_w16 blocks of width 16, then n_w8 of
                                       width 8, then the width-4 ones (lets a warp own equal-width blocks)      */
    int32_t  num_ref_idx_active[2]; /* entries of ref_slot[list][] in use (RefPicList.nb_refs, vvcdec.h:45-50); 0 = not
                                       stated: Level B then treats all 16 as pictures this one reads                 */
    int32_t  dmvr_first;            /* the emitter's statement that the records flagged VVCB200_PB_DMVR are the first ones of
                                       the width-16 class and the first ones of the width-8 class (the recommended order,
                                       vvcb200_emit_order_blocks): vvcb200_pic_inter then returns refined vectors for those
                                       records only — the other entries of `refined_mv` are left untouched, they would repeat
                                       the vectors the host sent.  0: every entry is written                                 */
    int16_t  weight[2][3][16];      /* [list][c_idx][ref_idx]                                  */
    int16_t  offset[2][3][16];
    uint8_t  ref_slot[2][16];       /* index into the `refs` array of (list, ref_idx)          */
} VVCB200InterParams;

/*
 * Level A/B shared entry: predict every block into `dst` (fc->frame) from the reference pictures `refs`
 * (host array of n_refs plane triples with device pointers).  Covers inter.put / put_uni / put_uni_w
 * (vvcdsp.h:49-59) for luma 8-tap and chroma 4-tap, avg / w_avg (vvcdsp.h:61-66) incl. BCW, and the
 * replicate-edge fetch of emulated_edge_mc (videodsp_template.c:27-105).
 * If `d_tmp14` is non-NULL the 14-bit intermediate of list 0 (what inter.put writes, int16) is stored
 * there instead of pixels (plane geometry of `dst`) — the checkasm-style view of inter.put.
 * Blocks flagged VVCB200_PB_LMCS get lmcs.filter (vvcdsp.h:124) with the forward LUT on their luma.
 * Bi-predicted blocks of at least 8x8 flagged VVCB200_PB_DMVR run the decoder-side refinement (inter.dmvr[][] + inter.sad
 * + dmvr_mv_refine, vvc_inter.c:642-748) and those flagged VVCB200_PB_BDOF the optical-flow blend (inter.bdof_fetch_samples
 * + inter.apply_bdof); d_refined_mv receives the vectors actually used — what the reference stores in tab_dmvr_mvf
 * (vvc_inter.c:750-762) for the host's later MV derivation.
 * Affine coding units arrive as 4x4 luma-only records (filter set 2, optional PROF: inter.fetch_samples + apply_prof*,
 * vvc_inter.c:369-445) plus chroma-only records carrying the averaged vector of derive_affine_mvc (vvc_inter.c:813-826);
 * GPM coding units as VVCB200_PB_GPM records (inter.put_gpm, vvc_inter_template.c:78-98).
 */
VVCB200_API int vvcb200_inter_batch(const VVCB200Planes *dst, const VVCB200Planes *refs, int n_refs,
                                    const VVCB200PredBlock *d_blocks, int n, const VVCB200InterParams *params /* host */,
                                    const uint16_t *d_lmcs_fwd_lut /* 1 << bit_depth entries, or NULL */,
                                    const VVCB200Planes *d_tmp14, int32_t *d_refined_mv /* n x 2 x 2, or NULL */,
                                    const VVCB200ProfTable *d_prof /* PROF tables, or NULL */, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Deblocking (stages DEBLOCK_V / DEBLOCK_H, ff_vvc_deblock_vertical/horizontal, vvc_filter.c:861-1003,
 * lf.filter_luma / filter_chroma / ladf_level, vvcdsp.h:126-133).
 * Boundary strengths and filter lengths are host-type table logic over CU/TU/MV state
 * (vvc_deblock_bs, vvc_filter.c:308-782); the host derives them exactly as the reference does and hands
 * the device the per-4x4 tables below.  All tables are indexed by luma 4x4 unit:
 * t[(y >> 2) * stride4 + (x >> 2)], like fc->tab.vertical_bs & co (vvcdec.h:158-163).
 * ---------------------------------------------------------------------------------------------- */
typedef struct VVCB200DeblockParams {
    int32_t bit_depth;
    int32_t ctb_log2_size;                  /* sps->ctb_log2_size_y                                       */
    int32_t ladf_enabled;                   /* sps_ladf_enabled_flag (vvc_filter.c:835)                   */
    int32_t ladf_lowest_interval_qp_offset;
    int32_t num_ladf_intervals;             /* sps->num_ladf_intervals                                    */
    int32_t ladf_qp_offset[4];              /* sps_ladf_qp_offset[]                                       */
    int32_t ladf_interval_lower_bound[5];   /* sps->ladf_interval_lower_bound[]                           */
} VVCB200DeblockParams;

typedef struct VVCB200DeblockTables {
    const uint8_t *d_edge[2][3];    /* [0: vertical edges, 1: horizontal edges][c_idx]; one byte per unit:
                                       bs (bits 1:0) | maxFilterLengthP << 2 | maxFilterLengthQ << 5        */
    const int8_t  *d_qp[3];         /* QpY / Qp'Cb / Qp'Cr of the unit (fc->tab.qp[], vvcdec.h:131)        */
    const int8_t  *d_ctb_offsets;   /* per CTB (raster): beta_offset[3] then tc_offset[3] (DBParams)        */
    int32_t        stride4;         /* units per table row (pps->min_tu_width)                             */
    int32_t        reserved;
} VVCB200DeblockTables;

/*
 * Host helper (pure CPU, no CUDA): fold the reference's table set — vertical_bs/horizontal_bs[3],
 * vertical_p/q, horizontal_p/q (luma lengths) and the chroma TB size tables used by
 * max_filter_length_chroma (vvc_filter.c:796-817) — into the packed edge bytes above.
 * `bs[dir][c]`, `len_p[dir]`, `len_q[dir]`, `tb_w_c`, `tb_h_c`: host tables of stride4 x rows4 bytes;
 * `edge_out[dir][c]`: host output tables of the same size.  Luma coordinates; 4:2:0.
 */
VVCB200_API int vvcb200_pack_deblock_edges(uint8_t *const edge_out[2][3], const uint8_t *const bs[2][3],
                                           const uint8_t *const len_p[2], const uint8_t *const len_q[2],
                                           const uint8_t *tb_w_c, const uint8_t *tb_h_c,
                                           int stride4, int rows4, int ctb_log2_size);

/*
 * Boundary strengths and filter lengths derived ON THE DEVICE: vvc_deblock_bs and its helpers
 * (vvc_filter.c:308-782: boundary_strength, derive_max_filter_length_luma, the sub-block edge rules, the chroma cbf
 * rules) plus max_filter_length_chroma (796-817), evaluated per luma 4x4 unit from two compact per-unit tables instead of
 * the reference's fc->tab.* set.  Output: the packed edge bytes vvcb200_deblock_picture / vvcb200_deblock_lmcs_picture read
 * (exactly what vvcb200_pack_deblock_edges makes of the reference's tables).
 */
typedef struct VVCB200MotionUnit {  /* MvField (vvc_ctu.h:195-202) of one 4x4 luma unit, fc->tab.mvf                    */
    int32_t mv[2][2];               /* mv[list] = {x, y}, 1/16 sample                                                    */
    int8_t  ref_id[2];              /* a number that identifies the reference PICTURE of (list, ref_idx): the host
                                       resolves RefPicList.list[ref_idx] (a POC, possibly of another slice's lists,
                                       ff_vvc_get_ref_list) to an index into the picture's distinct references          */
    uint8_t pred_flag;              /* 0 PF_INTRA, 1 PF_L0, 2 PF_L1, 3 PF_BI (vvc_ctu.h:222-227)                          */
    uint8_t ciip_flag;
} VVCB200MotionUnit;                /* 20 bytes */

typedef struct VVCB200PartitionUnit {   /* where a 4x4 luma unit sits in its transform blocks and coding unit            */
    uint8_t tb_off;                 /* luma TB: (x - tb_pos_x0) / 4 | (y - tb_pos_y0) / 4 << 4   (fc->tab.tb_pos_x0/y0[0]) */
    uint8_t tb_size;                /* log2 tb_width | log2 tb_height << 3 | tu_coded_flag[0] << 6 | pcmf[0] << 7          */
    uint8_t cu_off_x;               /* (x - cb_pos_x) / 4 | sub-block CU (msf || iaf) << 5 | cb_width > 8 << 6 |
                                       cb_height > 8 << 7                                                                 */
    uint8_t cu_off_y;               /* (y - cb_pos_y) / 4                                                                 */
    uint8_t ctb_off;                /* chroma TB (luma coordinates): (x - tb_pos_x0[1]) / 4 | (y - tb_pos_y0[1]) / 4 << 4  */
    uint8_t ctb_size;               /* log2 chroma tb_width | log2 tb_height << 3 (chroma samples) | tu_coded_flag[1] << 6 |
                                       tu_coded_flag[2] << 7                                                              */
    uint8_t cflags;                 /* bit0 tu_joint_cbcr_residual_flag, bit1 pcmf[1]                                     */
    uint8_t reserved;
} VVCB200PartitionUnit;             /* 8 bytes */

/* d_motion, d_part: stride4 x rows4 units.  d_ctb_edges: per CTB (raster), bit0 / bit1 = the left / upper CTB side is a
 * slice or tile limit that must not be filtered across (vvc_filter.c:505-512: BOUNDARY_*_SLICE with
 * pps_loop_filter_across_slices_enabled_flag 0, BOUNDARY_*_TILE with ..._across_tiles... 0); NULL: none.
 * d_edge_out[dir][c]: stride4 x rows4 bytes each. */
VVCB200_API int vvcb200_deblock_derive_edges(uint8_t *const d_edge_out[2][3], const VVCB200MotionUnit *d_motion,
                                             const VVCB200PartitionUnit *d_part, const uint8_t *d_ctb_edges, int width,
                                             int height /* luma samples */, int stride4, int ctb_log2_size, void *stream);

/* Both deblocking passes over the whole picture, in place: all vertical edges, then all horizontal edges. */
VVCB200_API int vvcb200_deblock_picture(const VVCB200Planes *pic, const VVCB200DeblockTables *tables /* host struct */,
                                        const VVCB200DeblockParams *params /* host */, void *stream);

/*
 * The same two stages (ff_vvc_deblock_vertical + ff_vvc_deblock_horizontal, vvc_filter.c:861-1003) as ONE out-of-place
 * pass, optionally preceded by stage LMCS (ff_vvc_lmcs_filter, vvc_filter.c:1322-1332): dst = deblock(lut(src)).
 * Every 64x64 tile of a plane is staged once (128-bit loads, luma mapped through `d_inv_lut` on the way in when it is
 * non-NULL), filtered across its vertical and then its horizontal edges in shared memory, and written once.
 * What Level B runs.  `dst` and `src` must be distinct pictures of the same geometry, planes 16-byte aligned with strides
 * that are multiples of 16 bytes.  Requires what the standard guarantees: no transform block crosses a multiple of 64
 * luma samples (sps_log2_max_luma_transform_size <= 6).  Results are identical to vvcb200_lmcs_picture +
 * vvcb200_deblock_picture.
 */
VVCB200_API int vvcb200_deblock_lmcs_picture(const VVCB200Planes *dst, const VVCB200Planes *src,
                                             const VVCB200DeblockTables *tables /* host struct */,
                                             const VVCB200DeblockParams *params /* host */,
                                             const uint16_t *d_inv_lut /* 1 << bit_depth entries, or NULL */, void *stream);
/* The same pass for luma alone (planes = 1), the two chroma planes alone (2) or all three (3): luma depends on luma only
 * (its QP, its LADF levels, its edges), so the luma half may start as soon as luma is final (VVCB200IntraParams.
 * luma_done_event) while the chroma phases of the intra stage still run. */
VVCB200_API int vvcb200_deblock_lmcs_planes(const VVCB200Planes *dst, const VVCB200Planes *src,
                                            const VVCB200DeblockTables *tables /* host struct */,
                                            const VVCB200DeblockParams *params /* host */,
                                            const uint16_t *d_inv_lut /* 1 << bit_depth entries, or NULL */, int planes,
                                            void *stream);

/* ------------------------------------------------------------------------------------------------
 * SAO and ALF (stages SAO / ALF, ff_vvc_sao_filter vvc_filter.c:154-298, ff_vvc_alf_filter 1254-1319;
 * sao.band_filter / edge_filter / edge_restore vvcdsp.h:138-145, alf.classify / recon_coeff_and_clip /
 * filter[2] / filter_cc vvcdsp.h:149-157).  Both are out-of-place picture passes here: the reference keeps
 * pre-filter border copies per CTU (sao_pixel_buffer_*, alf_pixel_buffer_*, vvcdec.h:165-168) so that each
 * CTU sees unfiltered neighbours — exactly what reading a separate source plane gives.
 * ---------------------------------------------------------------------------------------------- */
enum { VVCB200_SAO_OFF = 0, VVCB200_SAO_BAND = 1, VVCB200_SAO_EDGE = 2 };  /* enum SAOType, vvc_ctu.h:92-96 */

typedef struct VVCB200SaoCtb {      /* SAOParams, vvc_ctu.h:440-451, plus the restore flags of vvc_filter.c:163-212 */
    uint8_t type_idx[3];            /* VVCB200_SAO_*                                                              */
    uint8_t band_position[3];
    uint8_t eo_class[3];            /* 0 horizontal, 1 vertical, 2 135 deg, 3 45 deg (enum SAOEOClass)            */
    uint8_t no_filter;              /* CTB sides / corners SAO must not read across because of a slice or tile
                                       boundary: bit0 vert_edge[0] (left), bit1 vert_edge[1] (right),
                                       bit2 horiz_edge[0] (top), bit3 horiz_edge[1] (bottom), bits4-7 diag_edge[0..3]
                                       (up-left, up-right, low-right, low-left)                                   */
    int16_t offset_val[3][5];       /* SaoOffsetVal, [c][0] == 0                                                  */
} VVCB200SaoCtb;                    /* 40 bytes */

/* SAO over the whole picture: dst = SAO(src); CTBs with type 0 are copied. d_ctbs: one record per CTB, raster. */
VVCB200_API int vvcb200_sao_picture(const VVCB200Planes *dst, const VVCB200Planes *src, const VVCB200SaoCtb *d_ctbs,
                                    int ctb_log2_size, int bit_depth, void *stream);

typedef struct VVCB200AlfCtb {      /* ALFParams, vvc_ctu.h:453-460, plus the CTB edge flags of vvc_filter.c:1264-1280 */
    uint8_t ctb_flag[3];            /* alf_ctb_flag[]                                                             */
    uint8_t filt_set_idx_y;         /* AlfCtbFiltSetIdxY: < 16 fixed filter set, else 16 + index into luma sets   */
    uint8_t alt_idx[2];             /* alf_ctb_filter_alt_idx[]                                                   */
    uint8_t cc_idc[2];              /* alf_ctb_cc_cb_idc, alf_ctb_cc_cr_idc                                       */
    uint8_t edges;                  /* bit0 LEFT, bit1 TOP, bit2 RIGHT, bit3 BOTTOM: slice / tile limit (picture
                                       borders are implied; setting them too is harmless)                       */
    uint8_t reserved[3];
} VVCB200AlfCtb;                    /* 12 bytes */

typedef struct VVCB200AlfParams {   /* per slice: the APS content the slice header selects (VVCALF, vvc_ps.h:163-173) */
    int32_t  bit_depth;
    int32_t  ctb_log2_size;
    int16_t  luma_coeff[7][25][12];     /* [sh_alf_aps_id_luma index][filter][coeff]                              */
    uint8_t  luma_clip_idx[7][25][12];
    int16_t  chroma_coeff[8][6];        /* APS sh_alf_aps_id_chroma                                               */
    uint8_t  chroma_clip_idx[8][6];
    int16_t  cc_coeff[2][4][7];         /* [Cb, Cr][filter][coeff] of sh_alf_cc_cb/cr_aps_id                      */
    int32_t  cc_enabled[2];             /* the CC APS exists (vvc_filter.c:1213)                                  */
} VVCB200AlfParams;

/* ALF (luma 7x7 + classification, chroma 5x5, CC-ALF) over the whole picture: dst = ALF(src); CTBs/components
 * with ALF off are copied.  `d_params`/`d_ctbs` are DEVICE pointers (the struct is larger than kernel argument space). */
VVCB200_API int vvcb200_alf_picture(const VVCB200Planes *dst, const VVCB200Planes *src, const VVCB200AlfCtb *d_ctbs,
                                    const VVCB200AlfParams *d_params, int ctb_log2_size, int bit_depth, void *stream);

/*
 * Stages SAO and ALF (ff_vvc_sao_filter vvc_filter.c:154-298 + ff_vvc_alf_filter 1254-1319) as ONE out-of-place pass:
 * dst = ALF(SAO(src)).  A CTA stages the deblocked samples of a 64x64 luma region (and its chroma) plus a ring of 4,
 * applies SAO in shared memory and runs ALF / CC-ALF on that — the SAO picture never exists in memory.  What Level B runs.
 * d_sao == NULL: SAO off; d_alf == NULL: ALF off (d_params may then be NULL too).  d_scratch: vvcb200_sao_alf_scratch_bytes()
 * bytes of device memory, 16-byte aligned, for the picture's luma coefficient / clip table (alf.recon_coeff_and_clip
 * evaluated once per (set, class, transposition) instead of once per 4x4 block); NULL: every block derives its own.
 * `src` planes 16-byte aligned with
 * 16-byte multiple strides, `dst` planes 8-byte aligned with 8-byte multiple strides, dst != src, 4:2:0.
 * Results are identical to vvcb200_sao_picture followed by vvcb200_alf_picture.
 */
VVCB200_API size_t vvcb200_sao_alf_scratch_bytes(void);
VVCB200_API int vvcb200_sao_alf_picture(const VVCB200Planes *dst, const VVCB200Planes *src, const VVCB200SaoCtb *d_sao,
                                        const VVCB200AlfCtb *d_alf, const VVCB200AlfParams *d_params, int ctb_log2_size,
                                        int bit_depth, void *d_scratch, void *stream);

/* lmcs.filter (vvcdsp.h:124) over whole luma plane in place: pix = lut[pix] (ff_vvc_lmcs_filter, vvc_filter.c:1322-1332). */
VVCB200_API int vvcb200_lmcs_picture(const VVCB200Planes *pic, const uint16_t *d_lut, int bit_depth, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Intra prediction + reconstruction of intra (and CIIP) blocks: stage RECON, ff_vvc_reconstruct
 * (vvc_intra.c:498-527) -> predict_intra (245-275) -> intra.intra_pred / intra_cclm_pred / pred_* (vvcdsp.h:98-110,
 * vvc_intra_template.c).  Intra blocks depend on their reconstructed neighbours, so they run as a wavefront over CTUs
 * (vvc_thread.c:161) after everything that has no spatial dependency (inter prediction, all inverse transforms).
 * One record per prediction block and component, in decoding order, grouped per CTU.
 * Neighbour availability is host-side bookkeeping (the ReconstructedArea lists, vvc_intra.c:574-648): the host states how
 * many neighbour samples are reconstructed; the device does the substitution / filtering / prediction arithmetic.
 * ---------------------------------------------------------------------------------------------- */
enum {
    VVCB200_IB_MIP            = 1 << 0,  /* matrix-based intra prediction (intra_mip_flag, luma or mip_chroma_direct)       */
    VVCB200_IB_MIP_TRANSPOSED = 1 << 1,  /* intra_mip_transposed_flag                                                       */
    VVCB200_IB_ISP            = 1 << 2,  /* cu->isp_split_type != ISP_NO_SPLIT (reference-sample rules, vvc_intra_template.c:474,511) */
    VVCB200_IB_BDPCM          = 1 << 3,  /* cu->bdpcm_flag[c_idx]: no PDPC (vvc_intra.c:564)                                */
    VVCB200_IB_RESIDUAL       = 1 << 4,  /* add the int16 residual plane to the prediction (the TB was coded)               */
    VVCB200_IB_CIIP           = 1 << 5,  /* CIIP blend of this (planar) prediction with the inter prediction already in dst */
    VVCB200_IB_ADD_ONLY       = 1 << 6,  /* no prediction: the block already holds its (inter) prediction; only add the
                                            residual plane.  Used for chroma blocks of inter units when the residual has
                                            to wait for reconstructed luma (VVCB200_IB_LMCS_SCALE)                         */
    VVCB200_IB_LMCS_SCALE     = 1 << 7,  /* chroma: scale the residual by the factor of the block's VPDU, derived from the
                                            reconstructed luma next to it (lmcs_derive_chroma_scale,
                                            vvc_intra_template.c:390-428)                                                  */
};

typedef struct VVCB200IntraBlock {
    uint16_t x0, y0;            /* top-left corner in samples of the plane c_idx                                          */
    uint8_t  w, h;              /* size in samples of that plane (1..64)                                                  */
    uint8_t  c_idx;             /* 0 Y, 1 Cb, 2 Cr; a CCLM record has c_idx 1 and covers both chroma planes               */
    int8_t   mode;              /* intra mode after wide-angle mapping (-14..80, vvc_intra.c:693-714); 81..83 = LT/L/T CCLM */
    uint8_t  ref_idx;           /* cu->intra_luma_ref_idx as the reference stores it (0, 1 or 2; vvc_cabac.c:1282-1290)    */
    uint8_t  flags;             /* VVCB200_IB_*                                                                           */
    uint8_t  mip_mode;          /* intra_mip_mode                                                                         */
    uint8_t  up_left;           /* lc->na.cand_up_left                                                                    */
    uint16_t avail_top;         /* reconstructed samples to the right of (x0, y0 - 1), as ff_vvc_get_top_available() would
                                   report for an unbounded request (the device takes min(needed, avail))                  */
    uint16_t avail_left;        /* same downwards from (x0 - 1, y0), ff_vvc_get_left_available()                          */
    uint8_t  cb_w, cb_h;        /* ISP: the coding block's size in luma samples (refw = cb_w + w, vvc_intra_template.c:511-514) */
    uint8_t  ciip_weight;       /* CIIP: intra weight 1..3 (ciip_derive_intra_weight, vvc_inter.c:523-543)                */
    uint8_t  res_planes;        /* CCLM record with VVCB200_IB_RESIDUAL: bit 0 = Cb has a coded residual, bit 1 = Cr     */
    /* VVCB200_IB_LMCS_SCALE: the VPDU of the coding unit's origin, in units of min(CTB size, 64) luma samples, and whether
     * the luma column left of it (bit 0) / the luma row above it (bit 1) is available (ff_vvc_get_left/top_available)     */
    uint8_t  vpdu_x, vpdu_y, vpdu_avail;
    /* ISP records with VVCB200_IB_RESIDUAL: bit k set = transform partition k of the coding block (in coding order) has a
     * coded residual (tu.coded_flag[0] of that partition's TransformUnit); the other partitions add nothing, whatever
     * the residual planes hold there — vvcb200_residual_batch only writes the TBs it is given. */
    uint8_t  isp_coded;
    uint32_t reserved2;
} VVCB200IntraBlock;            /* 28 bytes */

typedef struct VVCB200IntraParams {
    int32_t bit_depth;
    int32_t ctb_log2_size;
    int32_t chroma_vertical_collocated;   /* sps_chroma_vertical_collocated_flag (CCLM down-sampling filter)              */
    int32_t lmcs_chroma_scale;            /* sh_lmcs_used_flag && ph_chroma_residual_scale_flag: chroma records carry
                                             VVCB200_IB_LMCS_SCALE.  Must be set for the flag to be honoured: the launch
                                             picks a kernel without the scaling code when it is 0 */
    /* LMCS model for records flagged VVCB200_IB_LMCS_SCALE (VVCLMCS, vvc_ps.h:192-202) */
    uint16_t lmcs_pivot[17];
    uint16_t lmcs_chroma_scale_coeff[16];
    uint8_t  lmcs_min_bin_idx, lmcs_max_bin_idx;
    uint8_t  prepared;                    /* vvcb200_intra_prepare has been issued for this picture (and is ordered before
                                             this call): vvcb200_intra_picture then starts with the wavefronts             */
    uint8_t  pad[3];
    /* Phases (vvcb200_intra_schedule with `phase_counts`): records [0, n_luma) are the luma records, the next n_add_only
     * the chroma VVCB200_IB_ADD_ONLY records, the rest the other chroma records.  Chroma never feeds luma, and an add-only
     * record depends on luma alone, so the three groups run as luma wavefront -> plain batch -> chroma wavefront: the
     * add-only records (most records of a random-access picture with chroma residual scaling) stop being tickets of a
     * latency-bound wavefront.  Both 0: one wavefront over all records in array order. */
    int32_t  n_luma, n_add_only;
    /* Alternatively the counts in DEVICE memory, as vvcb200_intra_schedule_device leaves them (three int32: luma, add-only,
     * other chroma); non-NULL: n_luma / n_add_only are ignored and every launch reads its range there. */
    const int32_t *d_phase_counts;
    /* cudaEvent_t or NULL: recorded on the call's stream as soon as every luma sample of the picture is final (after the
     * luma wavefront; after the only wavefront when the records are not phased).  What works on luma alone — the luma
     * half of the deblocking pass — need not wait for the chroma phases. */
    void *luma_done_event;
} VVCB200IntraParams;           /* 112 bytes */

/*
 * Reconstruction of the intra blocks of one picture, in place on `pic` (which already holds the inter prediction +
 * residual of every inter block).  One launch; warps draw records from a ticket counter in array order and each
 * record waits, on per-4x4 "pending" marks in `d_scratch`, for exactly the neighbour samples it reads.  The array order
 * must therefore be topological (every record after the records it reads from): the decoding order is one, and
 * vvcb200_intra_schedule() builds better ones from per-CTU decoding-order records (see VVCB200_INTRA_ORDER_*).
 * d_residual: int16 residual planes written by vvcb200_residual_batch for TBs flagged VVCB200_TB_TO_RESIDUAL (may be
 * NULL if no record has VVCB200_IB_RESIDUAL).  Contract: a record flagged VVCB200_IB_RESIDUAL reads the planes over its
 * whole block — except ISP records, which read only the partitions named in isp_coded — so every sample it reads must
 * have been written by a TB of THIS picture; the planes are never cleared between pictures.
 * Luma-dependent chroma residual scaling (sh_lmcs_used_flag && ph_chroma_residual_scale_flag): the factor depends on
 * reconstructed luma, so the host cannot put it into VVCB200TransformBlock.lmcs_scale.  Instead every chroma TB of such
 * a picture (w * h > 4) is flagged VVCB200_TB_TO_RESIDUAL without VVCB200_TB_LMCS_SCALE, and its block gets a record
 * here flagged VVCB200_IB_RESIDUAL | VVCB200_IB_LMCS_SCALE — an intra record, or VVCB200_IB_ADD_ONLY for an inter unit.
 * d_scratch: vvcb200_intra_scratch_bytes() bytes of device memory (ticket counter + marks of the blocks that are
 * rewritten in place, CIIP and add-only); the call clears it itself.  Two calls may not share a scratch buffer
 * concurrently.  Samples of pure intra blocks need no marks: the call first fills them with 0xffff ("not yet
 * reconstructed"), so bit depths above 15 are not supported by this entry point.
 * d_shadow (may be NULL): a second picture of device memory, at least as large as `pic`, 8-byte aligned planes and
 * strides, content irrelevant and undefined afterwards.  With it, the prediction of the blocks rewritten in place is
 * moved there first and those blocks are filled with 0xffff like every other record's, so that no record needs marks
 * or memory fences — considerably faster on pictures with many CIIP units or luma-dependent chroma scaling.  Any
 * picture-sized buffer that is idle between the residual and the deblocking stage will do (the loop filters' scratch
 * picture, as vvcb200_pic_intra uses).  The results are identical either way.
 */
VVCB200_API size_t vvcb200_intra_scratch_bytes(int width, int height /* luma samples */);
enum {
    /* CTU anti-diagonals x + 2y (the wavefront of vvc_thread.c:156-184), the CTUs of a diagonal interleaved record by
     * record.  O(n) on the host; what vvcb200_pic_intra uses. */
    VVCB200_INTRA_ORDER_DIAGONAL = 0,
    /* Sorted by depth in the dependency graph, so that consecutive tickets are the records that can run now anywhere
     * in the picture.  Shortens the kernel by up to a third on pictures with scattered intra units, at ~0.1 us of
     * host time per record. */
    VVCB200_INTRA_ORDER_DEPTH    = 1,
};
/* Host helper: `in` holds the records of CTU i (raster index) at ctu_first[i] .. ctu_first[i + 1], in decoding order;
 * `out` (same length, must not alias) receives them in the requested order.  width / height: luma samples.
 * n_levels (may be NULL; VVCB200_INTRA_ORDER_DEPTH only): the length of the longest dependency chain, i.e. the number
 * of serial steps of the picture.
 * phase_counts (may be NULL): when given, `out` is additionally grouped into the three phases of
 * VVCB200IntraParams.n_luma / n_add_only (each group keeps the requested order) and receives those two counts. */
VVCB200_API int vvcb200_intra_schedule(const VVCB200IntraBlock *in, const uint32_t *ctu_first, int ctus_w, int ctus_h,
                                       int ctb_log2_size, int width, int height, int order, VVCB200IntraBlock *out,
                                       int32_t *n_levels, int32_t *phase_counts);
/*
 * The same schedule made on the device, so that the host only copies the records up as the parser emitted them:
 * d_in holds the records of CTU i (raster index) at d_ctu_first[i] .. d_ctu_first[i + 1] in decoding order; d_out (n records,
 * must not alias) receives them grouped into the three phases, each phase in the diagonal order (CTU anti-diagonals x + 2y,
 * the CTUs of a diagonal dealt record by record); d_phase_counts receives the three group sizes (pass it on in
 * VVCB200IntraParams.d_phase_counts).  d_scratch: vvcb200_intra_schedule_scratch_bytes(n, ctus_w, ctus_h) bytes.
 * What vvcb200_pic_intra uses (the host form costs ~8 ns per record on the calling thread).
 */
VVCB200_API size_t vvcb200_intra_schedule_scratch_bytes(int n, int ctus_w, int ctus_h);
VVCB200_API int vvcb200_intra_schedule_device(const VVCB200IntraBlock *d_in, const uint32_t *d_ctu_first, int n, int ctus_w,
                                              int ctus_h, VVCB200IntraBlock *d_out, int32_t *d_phase_counts, void *d_scratch,
                                              void *stream);
VVCB200_API int vvcb200_intra_picture(const VVCB200Planes *pic, const VVCB200Planes *d_residual,
                                      const VVCB200Planes *d_shadow, const VVCB200IntraBlock *d_blocks, int n,
                                      const VVCB200IntraParams *params /* host */, void *d_scratch, void *stream);
/* The first part of vvcb200_intra_picture alone (same arguments): scratch cleared, sentinels written, the predictions of
 * blocks rewritten in place moved to the shadow picture.  It needs the inter prediction but not the residual stage, so a
 * caller may issue it on a second stream beside vvcb200_residual_batch and then call vvcb200_intra_picture with
 * params->prepared = 1 (ordered after both). */
VVCB200_API int vvcb200_intra_prepare(const VVCB200Planes *pic, const VVCB200Planes *d_residual,
                                      const VVCB200Planes *d_shadow, const VVCB200IntraBlock *d_blocks, int n,
                                      const VVCB200IntraParams *params /* host */, void *d_scratch, void *stream);

/*
 * vvcb200_intra_picture without caller scratch and without any allocation: the marks live in a fixed array inside
 * the library's device image (256 KiB, pictures up to 1920x1088 luma samples; larger ones return VVCB200_ERR_ARG).
 * This entry point has no counterpart in the reference; it exists for callers that may not allocate device memory.
 * The array is per device and shared by all callers: calls that overlap in time on one device race.  The first call
 * on a device looks the array's address up; make that call before starting a stream capture.
 */
VVCB200_API int vvcb200_intra_picture_static(const VVCB200Planes *pic, const VVCB200Planes *d_residual,
                                             const VVCB200IntraBlock *d_blocks, int n,
                                             const VVCB200IntraParams *params /* host */, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Level B: the picture-level engine.  One context per GPU per stream of pictures (one decoder instance);
 * a context is not thread-safe, different contexts are independent (replicas only).
 * All work is enqueued on the context's own CUDA stream and is asynchronous with respect to the host:
 * host buffers handed to vvcb200_pic_* must stay valid and unchanged until vvcb200_ctx_sync() returns
 * (allocate them with vvcb200_host_alloc() to get true overlap; pageable memory also works, slower).
 * This replaces the bodies of run_inter … run_alf (vvc_thread.c:432-566): the host parser fills the
 * descriptor batches for one picture, the context uploads them and runs the stages as picture-level
 * passes in the dependency order of vvc_thread.c:156-184.
 * ---------------------------------------------------------------------------------------------- */
typedef struct VVCB200Ctx VVCB200Ctx;

/* chroma_format_idc: 1 (4:2:0) is what the engine builds today; others return VVCB200_ERR_UNSUP.
 * dpb_size: number of picture buffers the context owns (1..17: VVC_MAX_DPB_SIZE + 1, vvcdec.h:82-83). */
VVCB200_API int  vvcb200_ctx_create(VVCB200Ctx **ctx, int device, int width, int height,
                                    int chroma_format_idc, int bit_depth, int dpb_size);
VVCB200_API void vvcb200_ctx_destroy(VVCB200Ctx *ctx);
VVCB200_API int  vvcb200_ctx_sync(VVCB200Ctx *ctx);
/* How thoroughly vvcb200_pic_* check the order and geometry of the host batches before uploading them:
 * 0 = not at all (the reference trusts its parser too), 1 = class boundaries and every 64th record (default),
 * 2 = every record (costs ~0.5 ms of host time per 4K picture; what the tests use). */
VVCB200_API int  vvcb200_ctx_set_validation(VVCB200Ctx *ctx, int level);
/* cudaStream_t of the context (for callers that enqueue their own work or record events on it). */
VVCB200_API void *vvcb200_ctx_stream(VVCB200Ctx *ctx);
/* Device planes of DPB slot `slot` (slot < 0: the picture under reconstruction, fc->frame, vvcdec.h:96). */
VVCB200_API int  vvcb200_ctx_planes(VVCB200Ctx *ctx, int slot, VVCB200Planes *out);

/* Pinned host memory for descriptor batches and frames. */
/* A pinned host block (256-byte aligned, freed with the context) that the context mirrors on the device, once per
 * descriptor set.  Batch arrays the caller carves out of it are uploaded to the same offset of the mirror, and the arrays of
 * one stage call that lie next to each other (gaps up to 64 KiB) travel as ONE copy: lay a picture's arrays out in call
 * order, stage by stage.  It pays: with a picture download in flight every additional upload costs ~10 us of link time
 * whatever its size, and a 4K picture has fifteen.  Arrays elsewhere in host memory keep working, one copy each.  The rule
 * for host batches holds (valid and unchanged until the copy has happened: vvcb200_ctx_sync()): one region per picture
 * in flight.
 * Allocates (host + four device blocks of `bytes`): call it at set-up time, not between pictures. */
VVCB200_API void *vvcb200_ctx_arena_alloc(VVCB200Ctx *ctx, size_t bytes);
VVCB200_API void *vvcb200_host_alloc(size_t bytes);
VVCB200_API void  vvcb200_host_free(void *p);

/* Start a picture that will be reconstructed into DPB slot `slot` (its old content is dead from here on). */
VVCB200_API int vvcb200_pic_begin(VVCB200Ctx *ctx, int slot);
/* Copy host samples into DPB slot `slot` (slot < 0: current picture): a reference picture decoded elsewhere,
 * prediction samples produced elsewhere, or test input.  strides in bytes. */
VVCB200_API int vvcb200_pic_upload(VVCB200Ctx *ctx, int slot, const uint16_t *const planes[3], const int32_t strides[3]);
/* Stage INTER: upload the sub-block batch and predict into the current picture.  params->ref_slot[][] are DPB slots
 * of this context.  lmcs_fwd_lut: host, 1 << bit_depth entries, or NULL. */
VVCB200_API int vvcb200_pic_inter(VVCB200Ctx *ctx, const VVCB200PredBlock *blocks, int n,
                                  const VVCB200InterParams *params, const uint16_t *lmcs_fwd_lut,
                                  int32_t *refined_mv /* host, n x 2 x 2, valid after vvcb200_ctx_sync(); or NULL */,
                                  const VVCB200ProfTable *prof, int n_prof /* host PROF tables, or NULL / 0 */);
/* Stage RECON, residual part: upload levels + descriptors, then one fused launch (see vvcb200_residual_batch). */
VVCB200_API int vvcb200_pic_residual(VVCB200Ctx *ctx, const int16_t *levels, size_t n_levels,
                                     const VVCB200TransformBlock *tbs, int n, const VVCB200ResidualParams *params);
/* Stage RECON, intra part (after vvcb200_pic_inter and vvcb200_pic_residual of the same picture): blocks holds the
 * records of CTU i (raster index) at ctu_first[i] .. ctu_first[i + 1] in decoding order, exactly as the parser meets
 * them; the context schedules them (vvcb200_intra_schedule) and runs vvcb200_intra_picture with the residual planes
 * vvcb200_pic_residual filled for TBs flagged VVCB200_TB_TO_RESIDUAL. */
VVCB200_API int vvcb200_pic_intra(VVCB200Ctx *ctx, const VVCB200IntraBlock *blocks, const uint32_t *ctu_first,
                                  int ctus_w, int ctus_h, const VVCB200IntraParams *params);
/* Stage LMCS: inverse luma mapping of the whole picture (host LUT of 1 << bit_depth entries). */
VVCB200_API int vvcb200_pic_lmcs(VVCB200Ctx *ctx, const uint16_t *inv_lut);
/* Stages DEBLOCK_V + DEBLOCK_H.  All table pointers are HOST memory (stride4 x rows4 bytes each; ctb_offsets 6 bytes
 * per CTB); the struct is the one vvcb200_deblock_picture takes, filled with host pointers. */
VVCB200_API int vvcb200_pic_deblock(VVCB200Ctx *ctx, const VVCB200DeblockTables *host_tables,
                                    const VVCB200DeblockParams *params);
/* Stages SAO + ALF.  sao / alf: one record per CTB (raster), host memory; either may be NULL (stage skipped).
 * alf_params: host.  The filtered picture ends up in the current DPB slot. */
VVCB200_API int vvcb200_pic_sao_alf(VVCB200Ctx *ctx, const VVCB200SaoCtb *sao, const VVCB200AlfCtb *alf,
                                    const VVCB200AlfParams *alf_params, int ctb_log2_size);
/* Copy DPB slot `slot` (slot < 0: current picture) back to host memory (the AVFrame the decoder outputs). */
VVCB200_API int vvcb200_pic_download(VVCB200Ctx *ctx, int slot, uint16_t *const planes[3], const int32_t strides[3]);

/* ------------------------------------------------------------------------------------------------
 * Output path: what the decoder does with a finished picture — crop to the conformance window
 * (vvc_refs.c:188-191 sets AVFrame.crop_*, export_frame_params vvcdec.c:666-677 the visible size) and hand the
 * frame on.  Two ways out:
 *   vvcb200_pic_export          no copy: device planes of the window + an event, for a consumer on the same GPU (what an
 *                               AV_PIX_FMT_CUDA frame carries: data[] = device pointers, linesize[] = pitch, plus the
 *                               stream / event discipline of libavutil/hwcontext_cuda.c);
 *   vvcb200_pic_download_window the download of vvcb200_pic_download restricted to the window.
 * Crop values are in luma samples and must be even (4:2:0).
 * ---------------------------------------------------------------------------------------------- */
typedef struct VVCB200Window {
    int32_t crop_left, crop_right, crop_top, crop_bottom;
} VVCB200Window;

typedef struct VVCB200Export {
    VVCB200Planes planes;       /* device pointers to the window's first sample, width / height of the window, the
                                   picture's own strides                                                               */
    void         *ready;        /* cudaEvent_t owned by the context: recorded after the last operation that wrote the
                                   picture; make the consuming stream wait for it (cudaStreamWaitEvent)                */
} VVCB200Export;

/* The planes stay valid and unchanged until the slot is handed to vvcb200_pic_begin / vvcb200_pic_upload again.  A consumer
 * that may still be reading by then passes the cudaEvent_t it records after its last read to vvcb200_pic_release: the
 * next writer of the slot waits for it (the event must stay alive until that writer has been enqueued). */
VVCB200_API int vvcb200_pic_export(VVCB200Ctx *ctx, int slot, const VVCB200Window *window /* NULL: whole picture */,
                                   VVCB200Export *out);
VVCB200_API int vvcb200_pic_release(VVCB200Ctx *ctx, int slot, void *consumer_done /* cudaEvent_t or NULL */);
VVCB200_API int vvcb200_pic_download_window(VVCB200Ctx *ctx, int slot, const VVCB200Window *window,
                                            uint16_t *const planes[3], const int32_t strides[3]);

#ifdef __cplusplus
}
#endif
#endif /* VVC_B200_H */
