/* TEST INFRASTRUCTURE ONLY.  Type configuration for a DOUBLE-precision build of the
 * reference's qdldl/src/qdldl.c (the in-tree qdldl/include/qdldl_types.h is the
 * float/int configuration of Makefile:16).  Used with -DQDLDL_TYPES_H -include so the
 * in-tree float typedefs are skipped.  Mirrors qdldl/configure/qdldl_types.h.in. */
#ifndef ORACLE_QDLDL_F64_TYPES_H
#define ORACLE_QDLDL_F64_TYPES_H
#include <limits.h>
typedef int QDLDL_int;
typedef double QDLDL_float;
typedef unsigned char QDLDL_bool;
#define QDLDL_INT_MAX INT_MAX
#endif
