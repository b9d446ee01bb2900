/* Hand-written stand-in for the reference's CMake-generated parsec_config.h: just enough for
 * compiling parsec/utils/zone_malloc.c and parsec/data_dist/matrix/grid_2Dcyclic.c in place
 * (oracle/Makefile target ref).  Not a copy of any reference file. */
#ifndef PB2_REFSHIM_PARSEC_CONFIG_H
#define PB2_REFSHIM_PARSEC_CONFIG_H
#include <stddef.h>
#include <stdint.h>
#include <sys/types.h>
#define BEGIN_C_DECLS
#define END_C_DECLS
#define PARSEC_DECLSPEC
#define __parsec_attribute_deprecated__(msg) __attribute__((deprecated(msg)))
#endif
