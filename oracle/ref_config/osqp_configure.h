/* Build configuration for compiling the UNMODIFIED reference sources where they
 * lie under /root/reference into oracle/_ref/libosqp_ref.so (test infrastructure
 * only).  Written for this repo: it stands in for the header the reference's
 * CMake would generate from configure/osqp_configure.h.in, selecting the same
 * defaults (CMakeLists.txt:59-111): printing, profiling, ctrl-c, 64-bit
 * indices (c_int = long long), double precision (c_float = double).          */
#ifndef OSQP_CONFIGURE_H
#define OSQP_CONFIGURE_H
#define IS_LINUX
#define PRINTING
#define PROFILING
#define CTRLC
#define DLONG
#endif
