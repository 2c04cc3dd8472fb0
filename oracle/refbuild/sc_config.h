/* Hand-written serial (no-MPI) configuration for libsc 2.8.3, used only to
 * compile the reference's third-party topology library into oracle/_ref/.
 * TEST INFRASTRUCTURE - not part of the product. */
#ifndef _SRC_SC_CONFIG_H
#define _SRC_SC_CONFIG_H 1
#define SC_CC "gcc"
#define SC_CFLAGS "-O2"
#define SC_CPP "gcc -E"
#define SC_CPPFLAGS ""
#define SC_ENABLE_MEMALIGN 1
#define SC_ENABLE_USE_COUNTERS 1
#define SC_ENABLE_USE_REALLOC 1
#define SC_HAVE_ALIGNED_ALLOC 1
#define SC_HAVE_FSYNC 1
#define SC_HAVE_POSIX_MEMALIGN 1
#define SC_HAVE_FABS 1
#define SC_HAVE_QSORT_R 1
#define SC_HAVE_GNU_QSORT_R 1
#define SC_HAVE_SIGNAL_H 1
#define SC_HAVE_STDINT_H 1
#define SC_HAVE_STDLIB_H 1
#define SC_HAVE_STRING_H 1
#define SC_HAVE_STRTOLL 1
#define SC_HAVE_SYS_TIME_H 1
#define SC_HAVE_SYS_TYPES_H 1
#define SC_HAVE_TIME_H 1
#define SC_HAVE_UNISTD_H 1
#define SC_MEMALIGN_BYTES (SC_SIZEOF_VOID_P)
#define SC_LDFLAGS ""
#define SC_LIBS ""
#define SC_PACKAGE "libsc"
#define SC_PACKAGE_BUGREPORT ""
#define SC_PACKAGE_NAME "libsc"
#define SC_PACKAGE_STRING "libsc 2.8.3"
#define SC_PACKAGE_TARNAME "libsc"
#define SC_PACKAGE_URL ""
#define SC_PACKAGE_VERSION "2.8.3"
#define SC_SIZEOF_INT 4
#define SC_SIZEOF_UNSIGNED_INT 4
#define SC_SIZEOF_LONG 8
#define SC_SIZEOF_LONG_LONG 8
#define SC_SIZEOF_UNSIGNED_LONG 8
#define SC_SIZEOF_UNSIGNED_LONG_LONG 8
#define SC_SIZEOF_VOID_P 8
#define SC_VERSION "2.8.3"
#define SC_VERSION_MAJOR 2
#define SC_VERSION_MINOR 8
#define SC_VERSION_POINT 3
#endif
