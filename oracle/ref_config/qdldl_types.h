/* Stand-in for the header QDLDL's CMake generates from
 * qdldl_sources/configure/qdldl_types.h.in with DLONG=ON, DFLOAT=OFF
 * (qdldl_sources/CMakeLists.txt:60-78).  Test infrastructure only.           */
#ifndef QDLDL_TYPES_H
#define QDLDL_TYPES_H
#include <limits.h>
typedef long long     QDLDL_int;
typedef double        QDLDL_float;
typedef unsigned char QDLDL_bool;
#define QDLDL_INT_MAX LLONG_MAX
#endif
