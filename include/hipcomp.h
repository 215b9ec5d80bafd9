/*
 * b200comp — top-level C header of the batched lossless codec library.
 *
 * Drop-in for reference include/hipcomp.h:60-80 (version macros and the
 * hipcompType_t element-type enum).  The deprecated single-buffer
 * hipcompDecompress* prototypes of the reference (hipcomp.h:106-186) are not
 * declared: the reference ships no definitions for them
 * (src/hipcomp_api.cpp is includes only), so nothing can link against them.
 */
#ifndef HIPCOMP_H
#define HIPCOMP_H

#include "hipcomp/shared_types.h"

#include <hip/hip_runtime.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HIPCOMP_MAJOR_VERSION 2
#define HIPCOMP_MINOR_VERSION 2
#define HIPCOMP_PATCH_VERSION 0

/* Element types understood by the codecs (values are ABI). */
typedef enum hipcompType_t
{
  HIPCOMP_TYPE_CHAR = 0,      /* 1 byte  */
  HIPCOMP_TYPE_UCHAR = 1,     /* 1 byte  */
  HIPCOMP_TYPE_SHORT = 2,     /* 2 bytes */
  HIPCOMP_TYPE_USHORT = 3,    /* 2 bytes */
  HIPCOMP_TYPE_INT = 4,       /* 4 bytes */
  HIPCOMP_TYPE_UINT = 5,      /* 4 bytes */
  HIPCOMP_TYPE_LONGLONG = 6,  /* 8 bytes */
  HIPCOMP_TYPE_ULONGLONG = 7, /* 8 bytes */
  HIPCOMP_TYPE_BITS = 0xff    /* untyped byte stream */
} hipcompType_t;

#ifdef __cplusplus
}
#endif

#endif
