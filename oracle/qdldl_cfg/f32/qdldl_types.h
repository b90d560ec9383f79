/* TEST INFRASTRUCTURE ONLY.  Build configuration (not source) for compiling the reference's
 * qdldl/src/qdldl.c: the values cmake would substitute into
 * qdldl/configure/qdldl_types.h.in for -DQDLDL_FLOAT=true -DQDLDL_LONG=false, which is the
 * configuration the reference Makefile:16 requires (linsys_t = float,
 * examples/track_iiwa_qdldl.cu:22). */
#ifndef QDLDL_TYPES_H
#define QDLDL_TYPES_H
#include <limits.h>
typedef int QDLDL_int;
typedef float QDLDL_float;
typedef unsigned char QDLDL_bool;
#define QDLDL_INT_MAX INT_MAX
#define QDLDL_FLOAT
#endif
