/*
 * oracle/ref/api_inter.c — TEST INFRASTRUCTURE.
 * Pulls the reference's inter-prediction stage driver (libavcodec/vvc/vvc_inter.c, read in place) into
 * this translation unit so that its file-static helpers (DMVR search, weight derivation, edge
 * emulation: vvc_inter.c:33-110, 642-748) are reachable from the flat test API below.
 */
#include "libavcodec/vvc/vvc_inter.c"
#include "api_common.h"
