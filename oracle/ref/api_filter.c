/*
 * oracle/ref/api_filter.c — TEST INFRASTRUCTURE.
 * Pulls the reference's loop-filter stage drivers (libavcodec/vvc/vvc_filter.c, read in place) into
 * this translation unit so that the file-static boundary-strength / tc / beta logic
 * (vvc_filter.c:308-859) is reachable from the flat test API below.
 */
#include "libavcodec/vvc/vvc_filter.c"
#include "api_common.h"
