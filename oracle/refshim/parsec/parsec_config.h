/* Hand-written stand-in for the reference's CMake-generated parsec_config.h (the values CMake would detect on this
 * x86-64 / gcc / glibc box), followed by the reference's own, non-generated parsec_config_bottom.h.  Used only by
 * `make -C oracle ref` to compile a few reference sources in place.  Not a copy of any reference file. */
#ifndef PB2_REFSHIM_PARSEC_CONFIG_H
#define PB2_REFSHIM_PARSEC_CONFIG_H
#define PARSEC_CONFIG_H_HAS_BEEN_INCLUDED 1
#include <stddef.h>
#include <stdint.h>
#include <sys/types.h>
#define BUILDING_PARSEC 1
#define PARSEC_VERSION_MAJOR 4
#define PARSEC_VERSION_MINOR 0
#define PARSEC_ARCH_X86_64 1
#define PARSEC_ATOMIC_USE_C11_ATOMICS 1
#define PARSEC_HAVE_INT128 1
#define PARSEC_SIZEOF_VOID_P 8
#define PARSEC_HAVE_STDARG_H 1
#define PARSEC_HAVE_UNISTD_H 1
#define PARSEC_HAVE_VA_COPY 1
#define PARSEC_HAVE_ATTRIBUTE_ALWAYS_INLINE 1
#define PARSEC_HAVE_ATTRIBUTE_VISIBILITY 1
#define PARSEC_HAVE_ATTRIBUTE_DEPRECATED 1
#define PARSEC_HAVE_ATTRIBUTE_FORMAT_PRINTF 1
#define PARSEC_HAVE_BUILTIN_EXPECT 1
#define PARSEC_HAVE_THREAD_LOCAL 1
#define PARSEC_HAVE_PTHREAD_GETSPECIFIC 1
#define PARSEC_HAVE_STRING_H 1
#define PARSEC_HAVE_ERRNO_H 1
#define PARSEC_HAVE_LIMITS_H 1
#define PARSEC_HAVE_STDDEF_H 1
#define PARSEC_HAVE_STDBOOL_H 1
#define PARSEC_HAVE_CTYPE_H 1
#define PARSEC_HAVE_TIMESPEC_TV_NSEC 1
#define PARSEC_HAVE_CLOCK_GETTIME 1
#include "parsec/parsec_config_bottom.h"
#endif
