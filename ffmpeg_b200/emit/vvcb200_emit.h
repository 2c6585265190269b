/*
 * vvcb200_emit.h — the descriptor producer that sits in the reference decoder.
 *
 * These functions take the reference's OWN structures (VVCFrameContext, VVCLocalContext, CodingUnit, PredictionUnit,
 * TransformUnit, TransformBlock, MvField, SAOParams, ALFParams: libavcodec/vvc/vvc_ctu.h:136-460, vvcdec.h:86-200) and
 * write the batches libvvcb200.so consumes (include/vvc_b200.h).  They are the bodies the stage runners of
 * libavcodec/vvc/vvc_thread.c:432-566 get in place of their pixel work:
 *
 *   run_inter   ff_vvc_predict_inter()  ->  vvcb200_emit_inter_cu() per coding unit       (vvc_inter.c:875-913)
 *   run_recon   ff_vvc_reconstruct()    ->  vvcb200_emit_recon_ctu() per CTU              (vvc_intra.c:498-527)
 *   run_sao / run_alf                   ->  vvcb200_emit_filters() once per picture       (vvc_filter.c:154-298, 1254-1319)
 *
 * The file is compiled against the reference's headers, where they lie (the reference tree is not part of this
 * repository): oracle/Makefile builds it into oracle/_ref/libvvcref.so next to the harness that drives it in the tests.
 * Pure host C, no CUDA.  A batch grows by realloc; the caller hands the arrays to vvcb200_pic_* and frees them with
 * vvcb200_emit_free().
 */
#ifndef VVCB200_EMIT_H
#define VVCB200_EMIT_H

#include <stddef.h>
#include <stdint.h>
#include "vvc_b200.h"

struct VVCLocalContext;
struct VVCFrameContext;

typedef struct VVCB200EmitBatch {
    /* stage INTER */
    VVCB200PredBlock *blocks;          /* in emission order; sort with vvcb200_emit_order_blocks() before submitting */
    int n_blocks, cap_blocks;
    VVCB200ProfTable *prof;
    int n_prof, cap_prof;
    /* stage RECON */
    VVCB200TransformBlock *tbs;
    int n_tbs, cap_tbs;
    int16_t *levels;
    size_t n_levels, cap_levels;
    VVCB200IntraBlock *iblocks;        /* decoding order per CTU */
    int n_iblocks, cap_iblocks;
    uint32_t *ctu_first;               /* ctb_count + 1 entries once every CTU has been emitted (raster order) */
    int n_ctus, cap_ctus;
    int error;                         /* sticky: out of memory, or a unit the engine does not take */
    int fixed;                         /* the arrays and their cap_* are the caller's — regions of a vvcb200_ctx_arena_alloc
                                          block, so that a stage's arrays upload as one copy; a full array sets error to
                                          VVCB200_ERR_TOO_BIG instead of growing, vvcb200_emit_free() frees nothing */
} VVCB200EmitBatch;

void vvcb200_emit_reset(VVCB200EmitBatch *b);     /* keep the arrays, forget their content */
void vvcb200_emit_free(VVCB200EmitBatch *b);

/* The body of predict_inter() (vvc_inter.c:875-892) for lc->cu as records: regular units in sub-blocks of at most 16x16
 * with their DMVR / BDOF / BCW / weighted-prediction / half-pel flags, affine units as 4x4 luma-only records + chroma-only
 * records with derive_affine_mvc's vector + one PROF table, GPM units with VVCB200_GPM_AUX; VVCB200_PB_LMCS where the
 * reference maps the prediction forward; CIIP units flagged VVCB200_PB_CIIP (their blend happens in RECON). */
int vvcb200_emit_inter_cu(VVCB200EmitBatch *b, const struct VVCLocalContext *lc);

/* Width order + tool groups (VVCB200InterParams.n_w16 / n_w8), stable; what vvcb200_pic_inter wants. */
void vvcb200_emit_order_blocks(VVCB200EmitBatch *b, VVCB200InterParams *params);

/* The body of ff_vvc_reconstruct() (vvc_intra.c:498-527) for CTU `rs` as records: a VVCB200TransformBlock + its scan-box
 * levels per coded transform block (qp from derive_qp, transform pair from derive_transform_type, LFNST set / transposition
 * from derive_ilfnst_pred_mode_intra — the reference's own functions), and a VVCB200IntraBlock per intra prediction unit
 * and plane with the availability ff_vvc_get_top_available / ff_vvc_get_left_available report at that point of the
 * traversal, CIIP records with ciip_derive_intra_weight's weight, add-only chroma records where the residual waits for
 * the luma-dependent scale.  CTUs must come in raster order.  Defined in vvcb200_emit_recon.h, which is included at the
 * end of libavcodec/vvc/vvc_intra.c (it calls that file's static helpers). */
int vvcb200_emit_recon_ctu(VVCB200EmitBatch *b, struct VVCLocalContext *lc, int rs, int rx, int ry);

/* ---- loop filters: per-CTB parameters and the per-4x4 state the device derives boundary strengths from ---------------
 * Caller-owned arrays (pinned host memory in a decoder: they are what vvcb200_pic_deblock / vvcb200_pic_sao_alf upload). */
struct SliceContext;
typedef struct VVCB200EmitFilters {
    VVCB200SaoCtb *sao;                /* ctb_count, raster */
    VVCB200AlfCtb *alf;                /* ctb_count */
    int8_t        *ctb_offsets;        /* ctb_count x 6: beta_offset[3], tc_offset[3] (fc->tab.deblock, DBParams) */
    uint8_t       *ctb_edges;          /* ctb_count: bit0 / bit1 = left / upper CTB side not to be deblocked across */
    VVCB200MotionUnit    *motion;      /* min_pu_width x min_pu_height */
    VVCB200PartitionUnit *part;
} VVCB200EmitFilters;

/* One CTB: what ff_vvc_sao_filter (vvc_filter.c:154-212), ff_vvc_alf_filter (1254-1280) and vvc_deblock_bs (308-782) read
 * about it, flattened.  lc must have been through ff_vvc_decode_neighbour() for this CTB (boundary_flags, sc). The QP tables
 * need no conversion: VVCB200DeblockTables.d_qp[c] is a copy of fc->tab.qp[c] as it is. */
void vvcb200_emit_filters_ctu(const VVCB200EmitFilters *out, const struct VVCLocalContext *lc, int rx, int ry);
/* The APS content a slice header selects (alf_filter_luma / _chroma / _cc, vvc_filter.c:1161-1228). */
void vvcb200_emit_alf_params(VVCB200AlfParams *p, const struct VVCFrameContext *fc, const struct SliceContext *sc);
void vvcb200_emit_deblock_params(VVCB200DeblockParams *p, const struct VVCFrameContext *fc);
/* run_inter's other half (vvc_inter.c:750-762): the vectors DMVR settled on, as vvcb200_pic_inter returned them for the
 * sorted batch, written into fc->ref->tab_dmvr_mvf for the units of the flagged records. */
void vvcb200_emit_store_refined_mv(const struct VVCFrameContext *fc, const VVCB200PredBlock *blocks, int n,
                                   const int32_t *refined_mv /* n x 2 x 2 */);

/* internal: array helpers (vvcb200_emit.c) used by vvcb200_emit_recon.h */
VVCB200TransformBlock *vvcb200_emit_new_tb(VVCB200EmitBatch *b);
VVCB200IntraBlock *vvcb200_emit_new_iblock(VVCB200EmitBatch *b);
int16_t *vvcb200_emit_levels(VVCB200EmitBatch *b, int n, uint32_t *off);
void vvcb200_emit_ctu_mark(VVCB200EmitBatch *b, int rs, int end);

#endif /* VVCB200_EMIT_H */
