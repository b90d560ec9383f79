/* TEST INFRASTRUCTURE ONLY.  Build configuration (not source) for a DOUBLE-precision build
 * of the reference's qdldl/src/qdldl.c (qdldl/configure/qdldl_types.h.in with
 * QDLDL_FLOAT=false, QDLDL_LONG=false): used for the high-precision lambda* of the parity
 * tests. */
#ifndef QDLDL_TYPES_H
#define QDLDL_TYPES_H
#include <limits.h>
typedef int QDLDL_int;
typedef double QDLDL_float;
typedef unsigned char QDLDL_bool;
#define QDLDL_INT_MAX INT_MAX
#endif
