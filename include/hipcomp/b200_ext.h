/*
 * b200comp — extensions that have no counterpart in the reference ABI.
 * Everything a hipCOMP caller needs is in lz4.h / snappy.h / cascaded.h; these
 * helpers serve the pipeline around the codecs (host staging, multi-GPU shards).
 */
#ifndef HIPCOMP_B200_EXT_H
#define HIPCOMP_B200_EXT_H

#include "hipcomp.h"

#include <hip/hip_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Packs batch_size chunks (device_ptrs[i], device_bytes[i]) into one contiguous
 * buffer: chunk i lands at device_dst + device_dst_offsets[i].  Used to make the
 * strided output of *CompressAsync gapless before a device-to-host copy. */
hipcompStatus_t b200compGatherChunksAsync(
    const void* const* device_ptrs,
    const size_t* device_bytes,
    const size_t* device_dst_offsets,
    void* device_dst,
    size_t batch_size,
    hipStream_t stream);

/* "major.minor.patch (sm_100a)" */
const char* b200compVersionString(void);

#ifdef __cplusplus
}
#endif

#endif
