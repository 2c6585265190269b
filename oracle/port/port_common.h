/*
 * oracle/port — TEST INFRASTRUCTURE: a plain-C CPU restatement of the reference's reconstruction
 * arithmetic, used only as the checker for the CUDA engine (tests/, smoke(), bench cpu_baseline).
 * Parity pin: every function here is compared against the reference's own compiled C
 * (oracle/_ref/libvvcref.so) in tests/test_oracle_*.py and against tests/golden/ fixtures
 * generated from it.  It is never linked into, or called by, the product library.
 */
#ifndef VVC_PORT_COMMON_H
#define VVC_PORT_COMMON_H
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include "vvc_b200.h"
#include "../../ffmpeg_b200/csrc/vvc_tables.h"   /* generated constant tables (data only) */

#define PORT_EXPORT __attribute__((visibility("default")))

typedef struct PortPlanes {      /* same layout as VVCB200Planes, host pointers */
    uint8_t  *plane[3];
    int32_t   stride[3];
    int32_t   width[3];
    int32_t   height[3];
} PortPlanes;

static inline int port_clip(int v, int lo, int hi) { return v < lo ? lo : v > hi ? hi : v; }
static inline int port_clip_intp2(int v, int p)    { return port_clip(v, -(1 << p), (1 << p) - 1); }
static inline int port_clip_pixel(int v, int bd)   { return port_clip(v, 0, (1 << bd) - 1); }
#endif
