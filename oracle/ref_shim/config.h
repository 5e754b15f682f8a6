/* Minimal config.h so that the UNMODIFIED reference sources src/zip.c,
 * src/parallel.c and src/unzip.c compile outside the reference's autotools
 * build (autoconf/automake/gnulib are not available in this image).
 * Test infrastructure only -- see oracle/README.md. */
#ifndef GZB_REF_SHIM_CONFIG_H
#define GZB_REF_SHIM_CONFIG_H
#define _GL_ATTRIBUTE_CONST __attribute__((__const__))
#define _GL_ATTRIBUTE_PURE  __attribute__((__pure__))
#include <limits.h>
#include <stdint.h>
#endif
