/*
 * oracle/ref/api_filter.c — TEST INFRASTRUCTURE.
 * Pulls the reference's loop-filter stage drivers (libavcodec/vvc/vvc_filter.c, read in place) into
 * this translation unit so that the file-static boundary-strength / tc / beta logic
 * (vvc_filter.c:308-859) is reachable from the flat test API below.
 */
#include "libavcodec/vvc/vvc_filter.c"
#include "api_common.h"
#include "vvcb200_emit.h"

#include <stdlib.h>

typedef struct FilterEnv {
    VVCFrameContext      fc;
    VVCLocalContext     *lc;
    SliceContext         sc;
    AVFrame              frame;
    VVCSPS               sps;
    VVCPPS               pps;
    H266RawSPS           rsps;
    H266RawPPS           rpps;
    H266RawPictureHeader rph;
    H266RawSliceHeader   rsh;
    RefPicList           rpl[2];
    uint16_t            *col_bd, *row_bd;
} FilterEnv;

/* Tile grid planted for the next SAO / ALF calls (test harness state, not thread-safe against concurrent re-planting):
 * col_bd[x] / row_bd[y] = first CTB column / row of the tile that holds CTB column x / row y, one entry past the end
 * included (VVCPPS.ctb_to_col_bd / ctb_to_row_bd, vvc_ps.c); n_cols == 0: one tile. */
static uint16_t g_col_bd[520], g_row_bd[520];
static int g_tile_cols, g_tile_rows, g_num_tiles = 1, g_lfate = 1;

VVCREF_EXPORT int vvcref_set_tiles(const uint16_t *col_bd, int n_cols, const uint16_t *row_bd, int n_rows, int num_tiles,
                                   int loop_filter_across_tiles)
{
    if (n_cols < 0 || n_rows < 0 || n_cols > 519 || n_rows > 519)
        return -1;
    g_tile_cols = n_cols;
    g_tile_rows = n_rows;
    for (int i = 0; i < n_cols; i++) g_col_bd[i] = col_bd[i];
    for (int i = 0; i < n_rows; i++) g_row_bd[i] = row_bd[i];
    g_num_tiles = n_cols ? num_tiles : 1;
    g_lfate = n_cols ? loop_filter_across_tiles : 1;
    return 0;
}

static FilterEnv *filter_env_new(const VVCRefPlanes *pl, int bit_depth, int ctb_log2)
{
    FilterEnv *e = calloc(1, sizeof(*e));
    if (!e)
        return NULL;
    e->lc = aligned_alloc(64, (sizeof(*e->lc) + 63) & ~(size_t)63);
    memset(e->lc, 0, sizeof(*e->lc));
    e->fc.frame   = &e->frame;
    e->fc.ps.sps  = &e->sps;
    e->fc.ps.pps  = &e->pps;
    e->fc.ps.ph.r = &e->rph;
    e->sps.r      = &e->rsps;
    e->pps.r      = &e->rpps;
    e->sc.sh.r    = &e->rsh;
    e->sc.rpl     = e->rpl;
    e->lc->fc     = &e->fc;
    e->lc->sc     = &e->sc;
    for (int c = 0; c < 3; c++) {
        e->frame.data[c]     = pl->plane[c];
        e->frame.linesize[c] = pl->stride[c];
    }
    e->sps.width  = e->pps.width  = pl->width[0];
    e->sps.height = e->pps.height = pl->height[0];
    e->sps.bit_depth          = bit_depth;
    e->sps.qp_bd_offset       = 6 * (bit_depth - 8);
    e->sps.pixel_shift        = 1;
    e->sps.hshift[1] = e->sps.hshift[2] = 1;
    e->sps.vshift[1] = e->sps.vshift[2] = 1;
    e->sps.ctb_log2_size_y    = ctb_log2;
    e->sps.ctb_size_y         = 1 << ctb_log2;
    e->sps.min_cb_log2_size_y = 2;
    e->rsps.sps_chroma_format_idc = 1;
    e->pps.ctb_width   = (pl->width[0]  + (1 << ctb_log2) - 1) >> ctb_log2;
    e->pps.ctb_height  = (pl->height[0] + (1 << ctb_log2) - 1) >> ctb_log2;
    e->pps.ctb_count   = e->pps.ctb_width * e->pps.ctb_height;
    e->pps.min_cb_width = e->pps.min_pu_width = e->pps.min_tu_width = (pl->width[0] + 3) >> 2;
    e->pps.min_cb_height = e->pps.min_pu_height = e->pps.min_tu_height = (pl->height[0] + 3) >> 2;
    /* one tile, one slice: all CTBs map to tile column/row 0 */
    e->col_bd = calloc(e->pps.ctb_width + 2, sizeof(uint16_t));
    e->row_bd = calloc(e->pps.ctb_height + 2, sizeof(uint16_t));
    if (g_tile_cols == (int)e->pps.ctb_width + 1 && g_tile_rows == (int)e->pps.ctb_height + 1) {
        memcpy(e->col_bd, g_col_bd, g_tile_cols * sizeof(uint16_t));
        memcpy(e->row_bd, g_row_bd, g_tile_rows * sizeof(uint16_t));
    }
    e->pps.ctb_to_col_bd = e->col_bd;
    e->pps.ctb_to_row_bd = e->row_bd;
    e->fc.tab.slice_idx  = calloc(e->pps.ctb_count, sizeof(int16_t));
    e->rpps.pps_loop_filter_across_slices_enabled_flag = 1;
    e->rpps.pps_loop_filter_across_tiles_enabled_flag  = 1;
    ff_vvc_dsp_init(&e->fc.vvcdsp, bit_depth);
    ff_videodsp_init(&e->fc.vdsp, bit_depth);
    return e;
}

static void filter_env_free(FilterEnv *e)
{
    free(e->fc.tab.slice_idx);
    free(e->col_bd);
    free(e->row_bd);
    free(e->lc);
    free(e);
}

/*
 * DEBLOCK_V then DEBLOCK_H over a whole picture through the reference's own ff_vvc_deblock_vertical /
 * ff_vvc_deblock_horizontal (vvc_filter.c:861-1003), CTU by CTU in raster order.  The boundary-strength and
 * filter-length tables (fc->tab.vertical_bs & co) are planted by the caller in the reference's own format;
 * the TB-position tables are filled with a value that never matches, so vvc_deblock_bs() (vvc_filter.c:756)
 * visits no TB and leaves the planted tables as they are.
 *   bs[dir][c], len_p[dir], len_q[dir], tb_w_c, tb_h_c, qp[c]: stride4 x rows4 bytes, luma 4x4 units
 *   ctb_offsets: per CTB beta_offset[3], tc_offset[3]
 *   pass_mask bit0 = vertical-edge pass, bit1 = horizontal-edge pass; CTB rows [ry0, ry1) (ry1 < 0: all) — lets the
 *   timed CPU baseline run CTB rows on several threads with a join between the passes, like the reference's
 *   own scheduler does (vvc_thread.c:156-184).
 */
VVCREF_EXPORT int vvcref_deblock_picture(const VVCRefPlanes *pl, const uint8_t *const bs[2][3],
                                         const uint8_t *const len_p[2], const uint8_t *const len_q[2],
                                         const uint8_t *tb_w_c, const uint8_t *tb_h_c, const int8_t *const qp[3],
                                         const int8_t *ctb_offsets, const VVCB200DeblockParams *p,
                                         int pass_mask, int ry0, int ry1)
{
    FilterEnv *e = filter_env_new(pl, p->bit_depth, p->ctb_log2_size);
    const size_t units = (size_t)e->pps.min_tu_width * e->pps.min_tu_height;
    uint8_t *copy[10];
    int *never = malloc(units * sizeof(int));
    DBParams *db = calloc(e->pps.ctb_count, sizeof(*db));
    int k = 0;

    for (size_t i = 0; i < units; i++)
        never[i] = -4;
    for (int i = 0; i < 2; i++)
        e->fc.tab.tb_pos_x0[i] = e->fc.tab.tb_pos_y0[i] = never;
#define DUP(src) (copy[k] = malloc(units), memcpy(copy[k], (src), units), copy[k++])
    for (int c = 0; c < 3; c++) {
        e->fc.tab.vertical_bs[c]   = DUP(bs[0][c]);
        e->fc.tab.horizontal_bs[c] = DUP(bs[1][c]);
        e->fc.tab.qp[c]            = (int8_t *)qp[c];
    }
    e->fc.tab.vertical_p   = DUP(len_p[0]);
    e->fc.tab.vertical_q   = DUP(len_q[0]);
    e->fc.tab.horizontal_p = DUP(len_p[1]);
    e->fc.tab.horizontal_q = DUP(len_q[1]);
    e->fc.tab.tb_width[CHROMA]  = (uint8_t *)tb_w_c;
    e->fc.tab.tb_height[CHROMA] = (uint8_t *)tb_h_c;
    e->fc.tab.tb_width[LUMA]    = (uint8_t *)tb_w_c;     /* not read: lengths come from the planted tables */
    e->fc.tab.tb_height[LUMA]   = (uint8_t *)tb_h_c;
    for (unsigned i = 0; i < e->pps.ctb_count; i++)
        for (int c = 0; c < 3; c++) {
            db[i].beta_offset[c] = ctb_offsets[i * 6 + c];
            db[i].tc_offset[c]   = ctb_offsets[i * 6 + 3 + c];
        }
    e->fc.tab.deblock = db;
    e->rsps.sps_ladf_enabled_flag = p->ladf_enabled;
    e->rsps.sps_ladf_lowest_interval_qp_offset = p->ladf_lowest_interval_qp_offset;
    e->sps.num_ladf_intervals = p->num_ladf_intervals;
    for (int i = 0; i < 4; i++)
        e->rsps.sps_ladf_qp_offset[i] = p->ladf_qp_offset[i];
    for (int i = 0; i < 5; i++)
        e->sps.ladf_interval_lower_bound[i] = p->ladf_interval_lower_bound[i];

    if (ry1 < 0 || ry1 > e->pps.ctb_height)
        ry1 = e->pps.ctb_height;
    for (int pass = 0; pass < 2; pass++)
        for (int ry = ry0; (pass_mask & (1 << pass)) && ry < ry1; ry++)
            for (int rx = 0; rx < e->pps.ctb_width; rx++) {
                const int x0 = rx << p->ctb_log2_size, y0 = ry << p->ctb_log2_size;
                ff_vvc_decode_neighbour(e->lc, x0, y0, rx, ry, ry * e->pps.ctb_width + rx);
                if (!pass)
                    ff_vvc_deblock_vertical(e->lc, x0, y0);
                else
                    ff_vvc_deblock_horizontal(e->lc, x0, y0);
            }
    while (k)
        free(copy[--k]);
    free(never);
    free(db);
    filter_env_free(e);
    return 0;
}

/* ---- SAO / ALF / LMCS through the reference's own stage functions ------------------------------------------ */

/* slice_idx: per CTB (raster) slice number; lfase: pps_loop_filter_across_slices_enabled_flag */
static void plant_slices(FilterEnv *e, const int16_t *slice_idx, int lfase)
{
    if (slice_idx)
        memcpy(e->fc.tab.slice_idx, slice_idx, e->pps.ctb_count * sizeof(int16_t));
    e->rpps.pps_loop_filter_across_slices_enabled_flag = lfase;
    e->rpps.num_tiles_in_pic = g_num_tiles;
    e->rpps.pps_loop_filter_across_tiles_enabled_flag = g_lfate;
}

/*
 * Stage SAO over a whole picture, in place, through ff_vvc_sao_copy_ctb_to_hv + ff_vvc_sao_filter
 * (vvc_filter.c:143-298): first every CTB saves its pre-SAO borders (what run_deblock_h does, vvc_thread.c:504-525),
 * then every CTB is filtered (run_sao).  ctbs: VVCB200SaoCtb per CTB (no_filter is NOT read: the reference
 * derives those flags itself from slice_idx / lfase).
 */
VVCREF_EXPORT size_t vvcref_sao_scratch_bytes(const VVCRefPlanes *pl, int ctb_log2)
{
    const int cw = (pl->width[0] + (1 << ctb_log2) - 1) >> ctb_log2, ch = (pl->height[0] + (1 << ctb_log2) - 1) >> ctb_log2;
    size_t n = 0;
    for (int c = 0; c < 3; c++)
        n += ((size_t)pl->width[c] * 2 * ch + (size_t)pl->height[c] * 2 * cw) * 2;
    return n;
}

/* scratch: vvcref_sao_scratch_bytes() bytes shared by all calls of one picture; phase_mask bit0 = save borders,
 * bit1 = filter; CTB rows [ry0, ry1) (ry1 < 0: all). */
VVCREF_EXPORT int vvcref_sao_picture(const VVCRefPlanes *pl, const VVCB200SaoCtb *ctbs, const int16_t *slice_idx,
                                     int lfase, int ctb_log2, int bit_depth, uint8_t *scratch, int phase_mask,
                                     int ry0, int ry1)
{
    FilterEnv *e = filter_env_new(pl, bit_depth, ctb_log2);
    SAOParams *sao = calloc(e->pps.ctb_count, sizeof(*sao));
    plant_slices(e, slice_idx, lfase);
    e->rsps.sps_sao_enabled_flag = 1;
    for (unsigned i = 0; i < e->pps.ctb_count; i++)
        for (int c = 0; c < 3; c++) {
            sao[i].type_idx[c]      = ctbs[i].type_idx[c];
            sao[i].band_position[c] = ctbs[i].band_position[c];
            sao[i].eo_class[c]      = ctbs[i].eo_class[c];
            for (int k = 0; k < 5; k++)
                sao[i].offset_val[c][k] = ctbs[i].offset_val[c][k];
        }
    e->fc.tab.sao = sao;
    for (int c = 0; c < 3; c++) {
        const int w = pl->width[c], h = pl->height[c];
        e->fc.tab.sao_pixel_buffer_h[c] = scratch;
        scratch += (size_t)w * 2 * e->pps.ctb_height * 2;
        e->fc.tab.sao_pixel_buffer_v[c] = scratch;
        scratch += (size_t)h * 2 * e->pps.ctb_width * 2;
    }
    if (ry1 < 0 || ry1 > e->pps.ctb_height)
        ry1 = e->pps.ctb_height;
    /* saving the borders of row ry also saves the bottom of row ry - 1 (vvc_filter.c:143-152) */
    for (int ry = ry0; (phase_mask & 1) && ry < ry1; ry++)
        for (int rx = 0; rx < e->pps.ctb_width; rx++)
            ff_vvc_sao_copy_ctb_to_hv(e->lc, rx, ry, ry == e->pps.ctb_height - 1);
    for (int ry = ry0; (phase_mask & 2) && ry < ry1; ry++)
        for (int rx = 0; rx < e->pps.ctb_width; rx++) {
            ff_vvc_decode_neighbour(e->lc, rx << ctb_log2, ry << ctb_log2, rx, ry, ry * e->pps.ctb_width + rx);
            ff_vvc_sao_filter(e->lc, rx << ctb_log2, ry << ctb_log2);
        }
    free(sao);
    filter_env_free(e);
    return 0;
}

/*
 * Stage ALF over a whole picture, in place, through ff_vvc_alf_copy_ctu_to_hv + ff_vvc_alf_filter
 * (vvc_filter.c:1230-1319).  `p` is the slice's resolved APS content (include/vvc_b200.h); it is unfolded into
 * the VVCALF APS list the reference indexes through the slice header (luma set k -> APS id k, chroma -> id 7 ...).
 */
VVCREF_EXPORT size_t vvcref_alf_scratch_bytes(const VVCRefPlanes *pl, int ctb_log2)
{
    const int cw = (pl->width[0] + (1 << ctb_log2) - 1) >> ctb_log2, ch = (pl->height[0] + (1 << ctb_log2) - 1) >> ctb_log2;
    size_t n = 0;
    for (int c = 0; c < 3; c++)
        n += 2 * ((size_t)pl->width[c] * (c ? ALF_BORDER_CHROMA : ALF_BORDER_LUMA) * ch * 2 + (size_t)pl->height[c] * ALF_PADDING_SIZE * cw);
    return n;
}

VVCREF_EXPORT int vvcref_alf_picture(const VVCRefPlanes *pl, const VVCB200AlfCtb *ctbs, const VVCB200AlfParams *p,
                                     const int16_t *slice_idx, int lfase, uint8_t *scratch, int phase_mask,
                                     int ry0, int ry1)
{
    FilterEnv *e = filter_env_new(pl, p->bit_depth, p->ctb_log2_size);
    ALFParams *alf = calloc(e->pps.ctb_count, sizeof(*alf));
    VVCALF *aps = calloc(8, sizeof(*aps));
    plant_slices(e, slice_idx, lfase);
    e->rsps.sps_alf_enabled_flag = 1;
    for (int k = 0; k < 7; k++) {
        memcpy(aps[k].luma_coeff, p->luma_coeff[k], sizeof(aps[k].luma_coeff));
        memcpy(aps[k].luma_clip_idx, p->luma_clip_idx[k], sizeof(aps[k].luma_clip_idx));
        e->rsh.sh_alf_aps_id_luma[k] = k;
    }
    /* APS 7 carries the chroma and the cross-component filters */
    memcpy(aps[7].chroma_coeff, p->chroma_coeff, sizeof(aps[7].chroma_coeff));
    memcpy(aps[7].chroma_clip_idx, p->chroma_clip_idx, sizeof(aps[7].chroma_clip_idx));
    for (int i = 0; i < 2; i++)
        for (int f = 0; f < 4; f++)
            for (int k = 0; k < 7; k++)
                aps[7].cc_coeff[i][f][k] = p->cc_coeff[i][f][k];
    e->rsh.sh_alf_aps_id_chroma = 7;
    e->rsh.sh_alf_cc_cb_aps_id  = p->cc_enabled[0] ? 7 : 6;
    e->rsh.sh_alf_cc_cr_aps_id  = p->cc_enabled[1] ? 7 : 6;
    for (int k = 0; k < 8; k++)
        e->fc.ps.alf_list[k] = &aps[k];
    if (!p->cc_enabled[0] || !p->cc_enabled[1])
        e->fc.ps.alf_list[6] = NULL;         /* "no such APS" (vvc_filter.c:1213); luma set 6 is then unusable */
    for (unsigned i = 0; i < e->pps.ctb_count; i++) {
        for (int c = 0; c < 3; c++)
            alf[i].ctb_flag[c] = ctbs[i].ctb_flag[c];
        alf[i].ctb_filt_set_idx_y = ctbs[i].filt_set_idx_y;
        for (int c = 0; c < 2; c++) {
            alf[i].alf_ctb_filter_alt_idx[c] = ctbs[i].alt_idx[c];
            alf[i].ctb_cc_idc[c]             = ctbs[i].cc_idc[c];
        }
    }
    e->fc.tab.alf = alf;
    for (int c = 0; c < 3; c++) {
        const int w = pl->width[c], h = pl->height[c], b = c ? ALF_BORDER_CHROMA : ALF_BORDER_LUMA;
        for (int i = 0; i < 2; i++) {
            e->fc.tab.alf_pixel_buffer_h[c][i] = scratch;
            scratch += (size_t)w * b * e->pps.ctb_height * 2;
            e->fc.tab.alf_pixel_buffer_v[c][i] = scratch;
            scratch += (size_t)h * ALF_PADDING_SIZE * e->pps.ctb_width;
        }
    }
    if (ry1 < 0 || ry1 > e->pps.ctb_height)
        ry1 = e->pps.ctb_height;
    for (int ry = ry0; (phase_mask & 1) && ry < ry1; ry++)
        for (int rx = 0; rx < e->pps.ctb_width; rx++)
            ff_vvc_alf_copy_ctu_to_hv(e->lc, rx << p->ctb_log2_size, ry << p->ctb_log2_size);
    for (int ry = ry0; (phase_mask & 2) && ry < ry1; ry++)
        for (int rx = 0; rx < e->pps.ctb_width; rx++) {
            ff_vvc_decode_neighbour(e->lc, rx << p->ctb_log2_size, ry << p->ctb_log2_size, rx, ry, ry * e->pps.ctb_width + rx);
            ff_vvc_alf_filter(e->lc, rx << p->ctb_log2_size, ry << p->ctb_log2_size);
        }
    free(aps);
    free(alf);
    filter_env_free(e);
    return 0;
}

/* Stage LMCS (inverse luma mapping of the whole picture) through ff_vvc_lmcs_filter (vvc_filter.c:1322-1332). */
VVCREF_EXPORT int vvcref_lmcs_picture(const VVCRefPlanes *pl, const uint16_t *inv_lut, int ctb_log2, int bit_depth,
                                      int ry0, int ry1)
{
    FilterEnv *e = filter_env_new(pl, bit_depth, ctb_log2);
    e->rsh.sh_lmcs_used_flag = 1;
    memcpy(e->fc.ps.lmcs.inv_lut, inv_lut, sizeof(uint16_t) << bit_depth);
    if (ry1 < 0 || ry1 > e->pps.ctb_height)
        ry1 = e->pps.ctb_height;
    for (int ry = ry0; ry < ry1; ry++)
        for (int rx = 0; rx < e->pps.ctb_width; rx++)
            ff_vvc_lmcs_filter(e->lc, rx << ctb_log2, ry << ctb_log2);
    filter_env_free(e);
    return 0;
}


/*
 * The reference's own boundary-strength / filter-length derivation (vvc_deblock_bs and its helpers, vvc_filter.c:308-782)
 * over a picture whose per-4x4 tables the caller plants in the reference's format (fc->tab.*, vvcdec.h:122-187):
 *   mvf:        MvField per 4x4 (24 bytes each, vvc_ctu.h:195-202)
 *   tb_pos_x0 / tb_pos_y0 [2], tb_width / tb_height [2]: per 4x4, luma and chroma trees (set_tb_pos, vvc_ctu.c:40-60)
 *   cb_pos_x / cb_pos_y, cb_width / cb_height: luma tree (set_cb_pos, vvc_ctu.c:1145-1166)
 *   tu_coded_flag[3], tu_joint, pcmf[2], msf, iaf: per 4x4
 *   ref_poc[2][16]: RefPicList.list of the (single) slice
 * Output: bs[dir][c], len_p[dir], len_q[dir] (stride4 x rows4 bytes, zero where the reference writes nothing).
 */
/* When set, vvcref_deblock_bs runs the product's emitter (vvcb200_emit_filters_ctu) on every CTB in place of the
 * reference's vvc_deblock_bs: the same fc->tab.* state in, the engine's compact per-4x4 records out. */
static const VVCB200EmitFilters *g_emit_filters;

VVCREF_EXPORT int vvcref_deblock_bs(int width, int height, int ctb_log2, const void *mvf,
                                    const int32_t *const tb_pos_x0[2], const int32_t *const tb_pos_y0[2],
                                    const uint8_t *const tb_width[2], const uint8_t *const tb_height[2],
                                    const int32_t *cb_pos_x, const int32_t *cb_pos_y, const uint8_t *cb_width,
                                    const uint8_t *cb_height, const uint8_t *const tu_coded_flag[3], const uint8_t *tu_joint,
                                    const uint8_t *const pcmf[2], const uint8_t *msf, const uint8_t *iaf,
                                    const int32_t *ref_poc /* [2][16] */, const int16_t *slice_idx, int lfase,
                                    uint8_t *const bs[2][3], uint8_t *const len_p[2], uint8_t *const len_q[2])
{
    VVCRefPlanes pl = { { NULL, NULL, NULL }, { 0, 0, 0 }, { width, width / 2, width / 2 }, { height, height / 2, height / 2 } };
    FilterEnv *e = filter_env_new(&pl, 10, ctb_log2);
    const size_t units = (size_t)e->pps.min_tu_width * e->pps.min_tu_height;
    plant_slices(e, slice_idx, lfase);
    e->fc.tab.mvf = (MvField *)mvf;
    for (int i = 0; i < 2; i++) {
        e->fc.tab.tb_pos_x0[i] = (int *)tb_pos_x0[i];
        e->fc.tab.tb_pos_y0[i] = (int *)tb_pos_y0[i];
        e->fc.tab.tb_width[i]  = (uint8_t *)tb_width[i];
        e->fc.tab.tb_height[i] = (uint8_t *)tb_height[i];
        e->fc.tab.pcmf[i]      = (uint8_t *)pcmf[i];
    }
    e->fc.tab.cb_pos_x[0]  = (int *)cb_pos_x;
    e->fc.tab.cb_pos_y[0]  = (int *)cb_pos_y;
    e->fc.tab.cb_width[0]  = (uint8_t *)cb_width;
    e->fc.tab.cb_height[0] = (uint8_t *)cb_height;
    for (int c = 0; c < 3; c++)
        e->fc.tab.tu_coded_flag[c] = (uint8_t *)tu_coded_flag[c];
    e->fc.tab.tu_joint_cbcr_residual_flag = (uint8_t *)tu_joint;
    e->fc.tab.msf = (uint8_t *)msf;
    e->fc.tab.iaf = (uint8_t *)iaf;
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < 16; i++)
            e->rpl[l].list[i] = ref_poc[l * 16 + i];
    if (g_emit_filters)                                     /* the picture behind each POC: DPB slot POC - 100 */
        for (int l = 0; l < 2; l++)
            for (int i = 0; i < 16; i++)
                e->rpl[l].ref[i] = &e->fc.DPB[(ref_poc[l * 16 + i] - 100) & 15];
    for (int c = 0; c < 3; c++) {
        memset(bs[0][c], 0, units);
        memset(bs[1][c], 0, units);
        e->fc.tab.vertical_bs[c]   = bs[0][c];
        e->fc.tab.horizontal_bs[c] = bs[1][c];
    }
    for (int d = 0; d < 2; d++) {
        memset(len_p[d], 0, units);
        memset(len_q[d], 0, units);
    }
    e->fc.tab.vertical_p   = len_p[0];
    e->fc.tab.vertical_q   = len_q[0];
    e->fc.tab.horizontal_p = len_p[1];
    e->fc.tab.horizontal_q = len_q[1];
    /* every slice uses the same reference picture lists here: across a slice limit the reference looks the neighbour's
     * lists up through fc->ref->rpl_tab (ff_vvc_get_ref_list, vvc_refs.c:66-74) */
    VVCFrame cur = { 0 };
    RefPicListTab **rpl_tab = calloc(e->pps.ctb_count, sizeof(*rpl_tab));
    for (unsigned i = 0; i < e->pps.ctb_count; i++)
        rpl_tab[i] = (RefPicListTab *)e->rpl;
    cur.rpl_tab = rpl_tab;
    e->fc.ref = &cur;
    for (int ry = 0; ry < (int)e->pps.ctb_height; ry++)
        for (int rx = 0; rx < (int)e->pps.ctb_width; rx++) {
            const int x0 = rx << ctb_log2, y0 = ry << ctb_log2;
            ff_vvc_decode_neighbour(e->lc, x0, y0, rx, ry, ry * e->pps.ctb_width + rx);
            if (g_emit_filters) {
                const int64_t t0 = vvcref_now_ns();
                vvcb200_emit_filters_ctu(g_emit_filters, e->lc, rx, ry);
                vvcref_emit_ns += vvcref_now_ns() - t0;
                continue;
            }
            vvc_deblock_bs(e->lc, x0, y0, 1);
            vvc_deblock_bs(e->lc, x0, y0, 0);
        }
    free(rpl_tab);
    filter_env_free(e);
    return 0;
}


/* f1 test entry: the per-4x4 state in the reference's format -> vvcb200_emit_filters_ctu -> VVCB200MotionUnit /
 * VVCB200PartitionUnit tables + the per-CTB "do not deblock across" flags.  Same inputs as vvcref_deblock_bs. */
VVCREF_EXPORT int vvcref_emit_deblock_units(int width, int height, int ctb_log2, const void *mvf,
                                            const int32_t *const tb_pos_x0[2], const int32_t *const tb_pos_y0[2],
                                            const uint8_t *const tb_width[2], const uint8_t *const tb_height[2],
                                            const int32_t *cb_pos_x, const int32_t *cb_pos_y, const uint8_t *cb_width,
                                            const uint8_t *cb_height, const uint8_t *const tu_coded_flag[3], const uint8_t *tu_joint,
                                            const uint8_t *const pcmf[2], const uint8_t *msf, const uint8_t *iaf,
                                            const int32_t *ref_poc, const int16_t *slice_idx, int lfase,
                                            VVCB200MotionUnit *motion, VVCB200PartitionUnit *part, uint8_t *ctb_edges)
{
    const size_t units = (size_t)((width + 3) >> 2) * ((height + 3) >> 2);
    uint8_t *dummy = malloc(units);
    uint8_t *const bs[2][3] = { { dummy, dummy, dummy }, { dummy, dummy, dummy } };
    uint8_t *const len[2] = { dummy, dummy };
    const VVCB200EmitFilters out = { .motion = motion, .part = part, .ctb_edges = ctb_edges };
    int ret;

    g_emit_filters = &out;
    vvcref_emit_ns = 0;
    ret = vvcref_deblock_bs(width, height, ctb_log2, mvf, tb_pos_x0, tb_pos_y0, tb_width, tb_height, cb_pos_x, cb_pos_y, cb_width,
                            cb_height, tu_coded_flag, tu_joint, pcmf, msf, iaf, ref_poc, slice_idx, lfase, bs, len, len);
    g_emit_filters = NULL;
    free(dummy);
    return ret;
}

/* f1 test entry: per-CTB SAO / ALF / deblocking parameters planted in the reference's format (SAOParams, ALFParams, the
 * VVCALF APS list behind the slice header, DBParams, the LADF fields of the SPS) exactly as vvcref_sao_picture /
 * vvcref_alf_picture / vvcref_deblock_picture plant them -> the emitter -> the engine's records, restore / edge flags
 * included (the inputs' no_filter / edges are NOT read). */
VVCREF_EXPORT int vvcref_emit_loopfilter_ctbs(int width, int height, int ctb_log2, int bit_depth,
                                              const VVCB200SaoCtb *in_sao, const VVCB200AlfCtb *in_alf,
                                              const VVCB200AlfParams *in_ap, const int16_t *slice_idx, int lfase,
                                              const int8_t *in_offsets, const VVCB200DeblockParams *in_dp,
                                              VVCB200SaoCtb *sao_out, VVCB200AlfCtb *alf_out, VVCB200AlfParams *ap_out,
                                              int8_t *offsets_out, uint8_t *ctb_edges_out, VVCB200DeblockParams *dp_out)
{
    VVCRefPlanes pl = { { NULL, NULL, NULL }, { 0, 0, 0 }, { width, width / 2, width / 2 }, { height, height / 2, height / 2 } };
    FilterEnv *e = filter_env_new(&pl, bit_depth, ctb_log2);
    const unsigned n = e->pps.ctb_count;
    SAOParams *sao = calloc(n, sizeof(*sao));
    ALFParams *alf = calloc(n, sizeof(*alf));
    DBParams *db   = calloc(n, sizeof(*db));
    VVCALF *aps    = calloc(8, sizeof(*aps));
    const VVCB200EmitFilters out = { .sao = sao_out, .alf = alf_out, .ctb_offsets = offsets_out, .ctb_edges = ctb_edges_out };

    plant_slices(e, slice_idx, lfase);
    for (unsigned i = 0; i < n; i++) {
        for (int c = 0; c < 3; c++) {
            sao[i].type_idx[c]      = in_sao[i].type_idx[c];
            sao[i].band_position[c] = in_sao[i].band_position[c];
            sao[i].eo_class[c]      = in_sao[i].eo_class[c];
            for (int k = 0; k < 5; k++)
                sao[i].offset_val[c][k] = in_sao[i].offset_val[c][k];
            alf[i].ctb_flag[c]      = in_alf[i].ctb_flag[c];
            db[i].beta_offset[c]    = in_offsets[i * 6 + c];
            db[i].tc_offset[c]      = in_offsets[i * 6 + 3 + c];
        }
        alf[i].ctb_filt_set_idx_y = in_alf[i].filt_set_idx_y;
        for (int c = 0; c < 2; c++) {
            alf[i].alf_ctb_filter_alt_idx[c] = in_alf[i].alt_idx[c];
            alf[i].ctb_cc_idc[c]             = in_alf[i].cc_idc[c];
        }
    }
    e->fc.tab.sao = sao;
    e->fc.tab.alf = alf;
    e->fc.tab.deblock = db;
    /* the APS list as vvcref_alf_picture lays it out: luma set k -> APS k, chroma + cross-component -> APS 7 */
    for (int k = 0; k < 7; k++) {
        memcpy(aps[k].luma_coeff, in_ap->luma_coeff[k], sizeof(aps[k].luma_coeff));
        memcpy(aps[k].luma_clip_idx, in_ap->luma_clip_idx[k], sizeof(aps[k].luma_clip_idx));
        e->rsh.sh_alf_aps_id_luma[k] = k;
    }
    e->rsh.sh_num_alf_aps_ids_luma = 7;
    memcpy(aps[7].chroma_coeff, in_ap->chroma_coeff, sizeof(aps[7].chroma_coeff));
    memcpy(aps[7].chroma_clip_idx, in_ap->chroma_clip_idx, sizeof(aps[7].chroma_clip_idx));
    for (int i = 0; i < 2; i++)
        for (int f = 0; f < 4; f++)
            for (int k = 0; k < 7; k++)
                aps[7].cc_coeff[i][f][k] = in_ap->cc_coeff[i][f][k];
    e->rsh.sh_alf_aps_id_chroma = 7;
    e->rsh.sh_alf_cc_cb_aps_id  = in_ap->cc_enabled[0] ? 7 : 6;
    e->rsh.sh_alf_cc_cr_aps_id  = in_ap->cc_enabled[1] ? 7 : 6;
    for (int k = 0; k < 8; k++)
        e->fc.ps.alf_list[k] = &aps[k];
    if (!in_ap->cc_enabled[0] || !in_ap->cc_enabled[1])
        e->fc.ps.alf_list[6] = NULL;
    e->rsps.sps_ladf_enabled_flag = in_dp->ladf_enabled;
    e->rsps.sps_ladf_lowest_interval_qp_offset = in_dp->ladf_lowest_interval_qp_offset;
    e->sps.num_ladf_intervals = in_dp->num_ladf_intervals;
    for (int i = 0; i < 4; i++)
        e->rsps.sps_ladf_qp_offset[i] = in_dp->ladf_qp_offset[i];
    for (int i = 0; i < 5; i++)
        e->sps.ladf_interval_lower_bound[i] = in_dp->ladf_interval_lower_bound[i];

    vvcref_emit_ns = 0;
    for (int ry = 0; ry < (int)e->pps.ctb_height; ry++)
        for (int rx = 0; rx < (int)e->pps.ctb_width; rx++) {
            int64_t t0;
            ff_vvc_decode_neighbour(e->lc, rx << ctb_log2, ry << ctb_log2, rx, ry, ry * e->pps.ctb_width + rx);
            t0 = vvcref_now_ns();
            vvcb200_emit_filters_ctu(&out, e->lc, rx, ry);
            vvcref_emit_ns += vvcref_now_ns() - t0;
        }
    {
        const int64_t t0 = vvcref_now_ns();
        vvcb200_emit_alf_params(ap_out, &e->fc, &e->sc);
        vvcb200_emit_deblock_params(dp_out, &e->fc);
        vvcref_emit_ns += vvcref_now_ns() - t0;
    }
    free(aps); free(db); free(alf); free(sao);
    filter_env_free(e);
    return 0;
}
