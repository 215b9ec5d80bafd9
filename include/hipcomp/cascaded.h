/*
 * b200comp — batched Cascaded codec (run-length -> delta -> bit-packing), C ABI.
 *
 * The compressed partition layout is bit-exact with the reference's
 * (src/CascadedKernels.hiph); see DESIGN.md "Cascaded wire format".
 * Replaces reference include/hipcomp/cascaded.h:66-295 /
 * src/lowlevel/CascadedBatch.hip:306-462.
 */
#ifndef HIPCOMP_CASCADED_H
#define HIPCOMP_CASCADED_H

#include "hipcomp.h"

#include <hip/hip_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Legacy (non-batched) option block, kept for source compatibility. */
typedef struct
{
  int num_RLEs;
  int num_deltas;
  int use_bp;
} hipcompCascadedFormatOpts;

/* chunk_size is carried but, as in the reference, 4096 is always used
 * internally.  type gives the element width and the type byte stored in the
 * partition header.  (reference cascaded.h:90-125) */
typedef struct
{
  size_t chunk_size;
  hipcompType_t type;
  int num_RLEs;
  int num_deltas;
  int use_bp;
} hipcompBatchedCascadedOpts_t;

static const hipcompBatchedCascadedOpts_t hipcompBatchedCascadedDefaultOpts
    = {4096, HIPCOMP_TYPE_INT, 2, 1, 1};

/* Always 0 bytes.  (reference CascadedBatch.hip:306-316) */
hipcompStatus_t hipcompBatchedCascadedCompressGetTempSize(
    size_t batch_size,
    size_t max_uncompressed_chunk_bytes,
    hipcompBatchedCascadedOpts_t format_opts,
    size_t* temp_bytes);

/* roundUp4(n) + 8.  (reference CascadedBatch.hip:318-327) */
hipcompStatus_t hipcompBatchedCascadedCompressGetMaxOutputChunkSize(
    size_t max_uncompressed_chunk_bytes,
    hipcompBatchedCascadedOpts_t format_opts,
    size_t* max_compressed_bytes);

/* Compress batch_size partitions (each < 4 GiB).  Input buffers must be
 * aligned to the element type, output buffers to max(4, sizeof(type)) and at
 * least GetMaxOutputChunkSize bytes.  temp is unused.
 * (reference CascadedBatch.hip:329-357) */
hipcompStatus_t hipcompBatchedCascadedCompressAsync(
    const void* const* device_uncompressed_ptrs,
    const size_t* device_uncompressed_bytes,
    size_t max_uncompressed_chunk_bytes,
    size_t batch_size,
    void* device_temp_ptr,
    size_t temp_bytes,
    void* const* device_compressed_ptrs,
    size_t* device_compressed_bytes,
    const hipcompBatchedCascadedOpts_t format_opts,
    hipStream_t stream);

/* Always 0 bytes.  (reference CascadedBatch.hip:359-364) */
hipcompStatus_t hipcompBatchedCascadedDecompressGetTempSize(
    size_t num_chunks, size_t max_uncompressed_chunk_bytes, size_t* temp_bytes);

/* Decompress.  The element width of the whole batch is taken from the type
 * byte of partition 0 (reference behaviour, CascadedBatch.hip:166-168).
 * Statuses / actual sizes are written for every partition.
 * (reference CascadedBatch.hip:366-436) */
hipcompStatus_t hipcompBatchedCascadedDecompressAsync(
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

/* Header word 1 of each partition, or 0 when fewer than 8 compressed bytes.
 * (reference CascadedBatch.hip:262-281,438-462) */
hipcompStatus_t hipcompBatchedCascadedGetDecompressSizeAsync(
    const void* const* device_compressed_ptrs,
    const size_t* device_compressed_bytes,
    size_t* device_uncompressed_bytes,
    size_t batch_size,
    hipStream_t stream);

#ifdef __cplusplus
}
#endif

#endif
