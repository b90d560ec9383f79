/* TEST INFRASTRUCTURE ONLY.  qdldl/configure/qdldl_version.h.in instantiated with the
 * version in qdldl/CMakeLists.txt:4-6. */
#ifndef QDLDL_VERSION_H_
#define QDLDL_VERSION_H_
#define QDLDL_VERSION_MAJOR 0
#define QDLDL_VERSION_MINOR 1
#define QDLDL_VERSION_PATCH 7
#endif
