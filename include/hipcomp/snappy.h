/*
 * b200comp — batched Snappy raw-block codec, C ABI.
 *
 * Every chunk is one raw Snappy block (varint uncompressed length followed by
 * literal/copy elements).  Replaces reference include/hipcomp/snappy.h:62-193 /
 * src/lowlevel/SnappyBatch.cpp:84-245.
 */
#ifndef HIPCOMP_SNAPPY_H
#define HIPCOMP_SNAPPY_H

#include "hipcomp.h"

#include <hip/hip_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct
{
  int reserved;
} hipcompBatchedSnappyOpts_t;

static const hipcompBatchedSnappyOpts_t hipcompBatchedSnappyDefaultOpts = {0};

/* Always 0 bytes.  (reference SnappyBatch.cpp:84-102) */
hipcompStatus_t hipcompBatchedSnappyDecompressGetTempSize(
    size_t num_chunks, size_t max_uncompressed_chunk_size, size_t* temp_bytes);

/* Reads the varint preamble of each chunk (<= 5 bytes, value < 2^31, else 0).
 * (reference SnappyBatch.cpp:104-130, SnappyBatchKernels.hip:84-134) */
hipcompStatus_t hipcompBatchedSnappyGetDecompressSizeAsync(
    const void* const* device_compressed_ptrs,
    const size_t* device_compressed_bytes,
    size_t* device_uncompressed_bytes,
    size_t batch_size,
    hipStream_t stream);

/* Decompress; statuses and actual sizes may be NULL.
 * (reference SnappyBatch.cpp:132-166) */
hipcompStatus_t hipcompBatchedSnappyDecompressAsync(
    const void* const* device_compresed_ptrs,
    const size_t* device_compressed_bytes,
    const size_t* device_uncompressed_bytes,
    size_t* device_actual_uncompressed_bytes,
    size_t batch_size,
    void* const device_temp_ptr,
    const size_t temp_bytes,
    void* const* device_uncompressed_ptr,
    hipcompStatus_t* device_statuses,
    hipStream_t stream);

/* Always 0 bytes.  (reference SnappyBatch.cpp:168-187) */
hipcompStatus_t hipcompBatchedSnappyCompressGetTempSize(
    size_t batch_size,
    size_t max_chunk_size,
    hipcompBatchedSnappyOpts_t format_ops,
    size_t* temp_bytes);

/* 32 + n + n/6.  (reference SnappyBatch.cpp:72-76,189-206) */
hipcompStatus_t hipcompBatchedSnappyCompressGetMaxOutputChunkSize(
    size_t max_chunk_size,
    hipcompBatchedSnappyOpts_t format_opts,
    size_t* max_compressed_size);

/* Compress; every output buffer must hold GetMaxOutputChunkSize bytes.
 * (reference SnappyBatch.cpp:208-245) */
hipcompStatus_t hipcompBatchedSnappyCompressAsync(
    const void* const* device_uncompressed_ptr,
    const size_t* device_uncompressed_bytes,
    size_t max_uncompressed_chunk_bytes,
    size_t batch_size,
    void* device_temp_ptr,
    size_t temp_bytes,
    void* const* device_compressed_ptr,
    size_t* device_compressed_bytes,
    hipcompBatchedSnappyOpts_t format_ops,
    hipStream_t stream);

#ifdef __cplusplus
}
#endif

#endif
