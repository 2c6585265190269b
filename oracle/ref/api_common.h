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
#endif
