/* oracle/ref/api_common.h — TEST INFRASTRUCTURE.  Shared bits of the flat (ctypes-friendly) API that
 * the tests use to drive the reference's own C code compiled into oracle/_ref/libvvcref.so. */
#ifndef ORACLE_REF_API_COMMON_H
#define ORACLE_REF_API_COMMON_H
#include <stdint.h>
#include <stddef.h>
#include "vvc_b200.h"   /* descriptor structs shared with the product (include/) */

#define VVCREF_EXPORT __attribute__((visibility("default")))

/* Host-side picture: three 16-bit planes. */
typedef struct VVCRefPlanes {
    uint8_t  *plane[3];
    int32_t   stride[3];   /* bytes */
    int32_t   width[3];
    int32_t   height[3];
} VVCRefPlanes;

/* CU-level description of an inter coding unit, for driving predict_inter() (vvc_inter.c:875-892) on affine and GPM
 * units, which the reference expands into sub-blocks itself.  Test-side only; the engine receives the expanded records. */
typedef struct VVCRefInterCU {
    int32_t x0, y0, w, h;
    int32_t kind;               /* 0 regular (one sub-block), 1 affine, 2 GPM */
    int32_t flags;              /* VVCB200_PB_WP / HPEL_IF / BDOF / DMVR / LMCS */
    int32_t pred_flag, bcw_idx;
    int32_t ref_idx[2];
    int32_t mv[2][2];           /* regular: MvField.mv; GPM: gpm_mv[i].mv[gpm_list[i]] */
    int32_t gpm_partition_idx, gpm_list[2];
    int32_t cb_prof_flag[2];
    int8_t  diff_mv_x[2][16], diff_mv_y[2][16];
    int32_t sb_mv_off;          /* affine: first entry of this CU's 4x4 sub-block vectors in `sb_mv` (raster, int32[2][2] each) */
} VVCRefInterCU;

/* CU / TU lists of a picture in decoding order, for driving ff_vvc_reconstruct() (vvc_intra.c:498-527). Test-side only. */
typedef struct VVCRefCU {
    int32_t x0, y0, w, h, ctu;          /* luma geometry, raster index of the CTU it belongs to */
    int32_t pred_mode;                  /* 0 MODE_INTER, 1 MODE_INTRA */
    int32_t intra_pred_mode_y, intra_pred_mode_c, intra_luma_ref_idx;
    int32_t mip_flag, mip_transposed, mip_mode, mip_chroma_direct;
    int32_t isp_split_type, num_intra_subpartitions;
    int32_t bdpcm_flag[3];
    int32_t ciip_flag, coded_flag;
    int32_t first_tu, n_tu;
    int32_t mts_idx, lfnst_idx;
    /* CIIP units (ciip_flag): the merge motion the reference predicts from inside RECON (vvc_intra.c:515-516) */
    int32_t inter_flags;                /* VVCB200_PB_WP / HPEL_IF / LMCS */
    int32_t pred_flag, ref_idx[2], mv[2][2];
    int32_t tree_type;                  /* 0 SINGLE_TREE, 1 DUAL_TREE_LUMA, 2 DUAL_TREE_CHROMA (vvc_ctu.h) */
} VVCRefCU;

typedef struct VVCRefTU {
    int32_t x0, y0, w, h;
    int32_t coded_flag[3];
    int32_t joint;
    int32_t tb[3];                      /* index into the VVCB200TransformBlock array, or -1 */
} VVCRefTU;

/* Host time spent inside the product's emitter (ffmpeg_b200/emit) during the last vvcref_emit_* call of this thread's
 * process, in nanoseconds — the harness's own work of building the reference-side state is not counted (glue.c). */
#include <time.h>
extern int64_t vvcref_emit_ns;
static inline int64_t vvcref_now_ns(void)
{
    struct timespec ts;
    timespec_get(&ts, TIME_UTC);             /* C11; the build is -std=c11 */
    return (int64_t)ts.tv_sec * 1000000000 + ts.tv_nsec;
}
#endif
