/* Hand-written build configuration for compiling the UNMODIFIED reference
 * (/root/reference) into oracle/_ref/libosqp_ref.so.
 *
 * TEST INFRASTRUCTURE ONLY. The reference's CMake would generate this header from
 * configure/osqp_configure.h.in with its default options (CMakeLists.txt:50-84:
 * PRINTING ON, PROFILING ON, DLONG ON, DFLOAT OFF). We do not run CMake; these
 * are the same switches, except CTRLC is OFF (no SIGINT handler in a batch run)
 * and PRINTING stays ON only because the reference's recursive_ldl.c calls
 * c_print unconditionally.
 */
#ifndef OSQP_CONFIGURE_H
#define OSQP_CONFIGURE_H
#define IS_LINUX
#define PRINTING
#define PROFILING
#define DLONG
#endif
