/* Hand-written stand-in for the CMake-generated qdldl_types.h (see
 * qdldl_sources/configure/qdldl_types.h.in): DLONG ON, DFLOAT OFF.
 * TEST INFRASTRUCTURE ONLY. */
#ifndef QDLDL_TYPES_H
#define QDLDL_TYPES_H
#include <limits.h>
typedef long long     QDLDL_int;
typedef double        QDLDL_float;
typedef unsigned char QDLDL_bool;
#define QDLDL_INT_MAX LLONG_MAX
#endif
