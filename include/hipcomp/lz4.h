/*
 * b200comp — batched LZ4 block codec, C ABI.
 *
 * Every chunk is one raw LZ4 *block* (no frame, no size prefix).  All device_*
 * arguments are device-accessible arrays owned by the caller; calls are
 * asynchronous on `stream` and never synchronise.
 *
 * Replaces reference include/hipcomp/lz4.h:79-243 /
 * src/lowlevel/LZ4Batch.cpp:71-224 — identical symbol names, argument order
 * and return codes.
 */
#ifndef HIPCOMP_LZ4_H
#define HIPCOMP_LZ4_H

#include "hipcomp.h"

#include <hip/hip_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Options for the high-level (manager) interface: chunk size in
 * [32768, 16777216]; 65536 recommended.  (reference lz4.h:68-76) */
typedef struct
{
  size_t chunk_size;
} hipcompLZ4FormatOpts;

/* data_type selects the match granularity of the encoder: 1-, 2- or 4-byte
 * elements.  8-byte types are rejected.  (reference lz4.h:79-84) */
typedef struct
{
  hipcompType_t data_type;
} hipcompBatchedLZ4Opts_t;

static const hipcompBatchedLZ4Opts_t hipcompBatchedLZ4DefaultOpts
    = {HIPCOMP_TYPE_CHAR};

/* Workspace the caller must provide to CompressAsync:
 * batch_size * min(pow2ceil(max_chunk), 16384) * 2 bytes.
 * Errors: temp_bytes NULL or max_chunk > 16 MiB -> hipcompErrorInvalidValue.
 * (reference LZ4Batch.cpp:154-170) */
hipcompStatus_t hipcompBatchedLZ4CompressGetTempSize(
    size_t batch_size,
    size_t max_uncompressed_chunk_bytes,
    hipcompBatchedLZ4Opts_t format_opts,
    size_t* temp_bytes);

/* Worst-case compressed size of one chunk: roundUp8(n + 1 + ceil(n/255)).
 * (reference LZ4Batch.cpp:172-187, LZ4Kernels.hiph:198-202) */
hipcompStatus_t hipcompBatchedLZ4CompressGetMaxOutputChunkSize(
    size_t max_uncompressed_chunk_bytes,
    hipcompBatchedLZ4Opts_t format_opts,
    size_t* max_compressed_bytes);

/* Compress batch_size chunks.  Each output buffer must hold
 * GetMaxOutputChunkSize bytes.  (reference LZ4Batch.cpp:189-224) */
hipcompStatus_t hipcompBatchedLZ4CompressAsync(
    const void* const* device_uncompressed_ptrs,
    const size_t* device_uncompressed_bytes,
    size_t max_uncompressed_chunk_bytes,
    size_t batch_size,
    void* device_temp_ptr,
    size_t temp_bytes,
    void* const* device_compressed_ptrs,
    size_t* device_compressed_bytes,
    hipcompBatchedLZ4Opts_t format_opts,
    hipStream_t stream);

/* roundUp8(24 * num_chunks); the workspace is accepted but not needed.
 * (reference LZ4Batch.cpp:71-87) */
hipcompStatus_t hipcompBatchedLZ4DecompressGetTempSize(
    size_t num_chunks, size_t max_uncompressed_chunk_bytes, size_t* temp_bytes);

/* Decompress.  device_uncompressed_bytes[i] is the capacity of output i.
 * device_actual_uncompressed_bytes and device_statuses may be NULL.  A chunk
 * that cannot be decoded within its capacity gets status
 * hipcompErrorCannotDecompress and actual size 0; the call never faults on
 * corrupt input.  (reference LZ4Batch.cpp:89-125) */
hipcompStatus_t hipcompBatchedLZ4DecompressAsync(
    const void* const* device_compressed_ptrs,
    const size_t* device_compressed_bytes,
    const size_t* device_uncompressed_bytes,
    size_t* device_actual_uncompressed_bytes,
    size_t batch_size,
    void* const device_temp_ptr,
    size_t temp_bytes,
    void* const* device_uncompressed_ptrs,
    hipcompStatus_t* device_statuses,
    hipStream_t stream);

/* Token walk without output: uncompressed size of every chunk.
 * (reference LZ4Batch.cpp:127-152) */
hipcompStatus_t hipcompBatchedLZ4GetDecompressSizeAsync(
    const void* const* device_compressed_ptrs,
    const size_t* device_compressed_bytes,
    size_t* device_uncompressed_bytes,
    size_t batch_size,
    hipStream_t stream);

#ifdef __cplusplus
}
#endif

#endif
