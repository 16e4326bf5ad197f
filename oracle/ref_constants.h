/* Test infrastructure (oracle/_ref build only).  Force-included (-include) ahead of the reference's
 * headers: pre-defines the include guard of src/constants.h so that the compile-time resolution
 * (constants.h:4-6) and MAX_RAY_DEPTH (constants.h:13) can be chosen per build.  All other macros
 * restate constants.h:8-16 verbatim in value. */
#ifndef H_CONSTANTS
#define H_CONSTANTS
#ifndef REF_W
#define REF_W 640
#endif
#ifndef REF_H
#define REF_H 480
#endif
#define WINDOW_WIDTH REF_W
#define WINDOW_HEIGHT REF_H
#define NR_PIXELS (WINDOW_WIDTH * WINDOW_HEIGHT)
#define EPS 0.001f
#define PI 3.14159265358979323846f
#ifndef REF_DEPTH
#define REF_DEPTH 32
#endif
#define MAX_RAY_DEPTH REF_DEPTH
#define MAX_CACHE_DEPTH 3
#endif
