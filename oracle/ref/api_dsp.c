/*
 * oracle/ref/api_dsp.c — TEST INFRASTRUCTURE.
 * Flat wrappers over individual VVCDSPContext entries (libavcodec/vvc/vvcdsp.h:48-168) of the
 * reference's own C implementation, so Python tests can call them through ctypes the way
 * tests/checkasm/vvc_*.c call them through the table.
 */
#include "libavcodec/vvc/vvc_ctu.h"
#include "libavcodec/vvc/vvcdsp.h"
#include "libavcodec/vvc/vvc_data.h"
#include "api_common.h"

static VVCDSPContext ctx[3];
static int ctx_ready[3];

static VVCDSPContext *dsp(int bd)
{
    const int i = bd == 8 ? 0 : bd == 10 ? 1 : 2;
    if (!ctx_ready[i]) {
        ff_vvc_dsp_init(&ctx[i], bd);
        ctx_ready[i] = 1;
    }
    return &ctx[i];
}

/* itx.itx[trh][trv][log2w][log2h] (vvcdsp.h:117-118). Returns -1 when the slot is not populated. */
VVCREF_EXPORT int vvcref_itx(int bd, int trh, int trv, int log2w, int log2h, int *coeffs,
                             int nzw, int nzh, int log2_transform_range)
{
    void (*fn)(int *, size_t, size_t, intptr_t, intptr_t) = dsp(bd)->itx.itx[trh][trv][log2w][log2h];
    if (!fn)
        return -1;
    fn(coeffs, nzw, nzh, log2_transform_range, bd);
    return 0;
}

VVCREF_EXPORT void vvcref_add_residual(int bd, uint8_t *dst, const int *res, int w, int h, ptrdiff_t stride)
{
    dsp(bd)->itx.add_residual(dst, res, w, h, stride);
}

VVCREF_EXPORT void vvcref_add_residual_joint(int bd, uint8_t *dst, const int *res, int w, int h, ptrdiff_t stride,
                                             int c_sign, int shift)
{
    dsp(bd)->itx.add_residual_joint(dst, res, w, h, stride, c_sign, shift);
}

VVCREF_EXPORT void vvcref_transform_bdpcm(int bd, int *coeffs, int w, int h, int vertical, int log2_transform_range)
{
    dsp(bd)->itx.transform_bdpcm(coeffs, w, h, vertical, log2_transform_range);
}

/* Address of a data table exported by vvc_data.c, by name; used by tools/gen_tables.py to emit the
 * product's own constant tables and by tests to cross-check them. */
#define TBL(name) if (!strcmp(#name, id)) { *size = sizeof(name); return (const void *)name; }
#include <string.h>
VVCREF_EXPORT const void *vvcref_table(const char *id, size_t *size)
{
    TBL(ff_vvc_dct8_4x4)   TBL(ff_vvc_dct8_8x8)   TBL(ff_vvc_dct8_16x16) TBL(ff_vvc_dct8_32x32)
    TBL(ff_vvc_dst7_4x4)   TBL(ff_vvc_dst7_8x8)   TBL(ff_vvc_dst7_16x16) TBL(ff_vvc_dst7_32x32)
    TBL(ff_vvc_lfnst_4x4)  TBL(ff_vvc_lfnst_8x8)  TBL(ff_vvc_lfnst_tr_set_index)
    TBL(ff_vvc_inter_luma_filters) TBL(ff_vvc_inter_chroma_filters) TBL(ff_vvc_inter_luma_dmvr_filters)
    TBL(ff_vvc_intra_luma_filter)  TBL(ff_vvc_gpm_weights) TBL(ff_vvc_gpm_angle_idx) TBL(ff_vvc_gpm_distance_idx)
    TBL(ff_vvc_gpm_distance_lut)   TBL(ff_vvc_gpm_angle_to_mirror) TBL(ff_vvc_gpm_angle_to_weights_idx)
    TBL(ff_vvc_gpm_weights_offset_x) TBL(ff_vvc_gpm_weights_offset_y)
    TBL(ff_vvc_alf_fix_filt_coeff) TBL(ff_vvc_alf_class_to_filt_map) TBL(ff_vvc_alf_aps_class_to_filt_map)
    TBL(ff_vvc_diag_scan_x) TBL(ff_vvc_diag_scan_y)
    *size = 0;
    return NULL;
}

/* MIP weight matrices (vvc_data.c:310-1155) and the angle derivations of vvc_intra.c:662-690, for tools/gen_tables.py */
VVCREF_EXPORT const uint8_t *vvcref_mip_matrix(int size_id, int mode_id)
{
    return ff_vvc_get_mip_matrix(size_id, mode_id);
}

int ff_vvc_intra_pred_angle_derive(const int pred_mode);
int ff_vvc_intra_inv_angle_derive(const int pred_mode);
VVCREF_EXPORT int vvcref_intra_angle(int mode)     { return ff_vvc_intra_pred_angle_derive(mode); }
VVCREF_EXPORT int vvcref_intra_inv_angle(int angle) { return ff_vvc_intra_inv_angle_derive(angle); }
