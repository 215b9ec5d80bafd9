/*
 * b200comp — extensions that have no counterpart in the reference ABI.
 * Everything a hipCOMP caller needs is in lz4.h / snappy.h / cascaded.h; these
 * helpers serve the pipeline around the codecs (host staging, multi-GPU shards).
 */
#ifndef HIPCOMP_B200_EXT_H
#define HIPCOMP_B200_EXT_H

#include "hipcomp.h"
#include "hipcomp/lz4.h"

#include <stdint.h>

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


/* ---- Cascaded option selection per partition (the reference dropped nvCOMP's selector,
 * include/hipcomp/cascaded.h:93-99, CHANGELOG.md:113-116).  SelectOpts samples the first 1024 elements of every
 * partition and writes 4 bytes per partition {num_RLEs, num_deltas, use_bp, 0}: the option set with the smallest
 * estimated packed size among raw / bp / delta+bp / rle+bp / rle+delta+bp.  CompressWithOpts compresses every
 * partition with its own set; the partition header carries it (format unchanged), so
 * hipcompBatchedCascadedDecompressAsync - and the reference's decoder - decode the result as usual.  Output slots
 * as for hipcompBatchedCascadedCompressAsync (GetMaxOutputChunkSize bytes each). */
hipcompStatus_t b200compBatchedCascadedSelectOptsAsync(const void* const* device_uncompressed_ptrs,
                                                       const size_t* device_uncompressed_bytes, size_t batch_size,
                                                       hipcompType_t type, uint8_t* device_opts, hipStream_t stream);
hipcompStatus_t b200compBatchedCascadedCompressWithOptsAsync(const void* const* device_uncompressed_ptrs,
                                                             const size_t* device_uncompressed_bytes, size_t batch_size,
                                                             void* const* device_compressed_ptrs,
                                                             size_t* device_compressed_bytes, hipcompType_t type,
                                                             const uint8_t* device_opts, hipStream_t stream);

/* ---- LZ4 compression that allocates nothing and takes no workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments, output bytes and errors as
 * hipcompBatchedLZ4CompressAsync (lz4.h) without device_temp_ptr / temp_bytes: the per-chunk hash tables that
 * call keeps in the caller's workspace for chunks above 64 KiB live in a static device array of
 * B200COMP_LZ4_NOALLOC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until exit).
 * Limit: hipcompBatchedLZ4CompressGetTempSize(batch_size, max_uncompressed_chunk_bytes) must not exceed that
 * size, i.e. at most B200COMP_LZ4_NOALLOC_MAX_BATCH chunks per call when chunks are above 64 KiB (any batch size
 * up to 64 KiB, where the encoder keeps its tables in shared memory); larger requests return
 * hipcompErrorInvalidValue.  The call makes no cudaMalloc / cudaFree and no allocating library call.  All calls on
 * one GPU share the array: calls that overlap on the GPU (different streams, or a second thread) race; calls
 * ordered on one stream do not.  Nothing has to be cleared between calls - the encoder resets a chunk's table
 * before it uses it. */
#define B200COMP_LZ4_NOALLOC_SCRATCH_BYTES (16u << 20)
#define B200COMP_LZ4_NOALLOC_MAX_BATCH 512u /* 16 MiB / (16384-entry uint16 table per chunk) */
hipcompStatus_t b200compBatchedLZ4CompressNoAllocAsync(const void* const* device_uncompressed_ptrs,
                                                       const size_t* device_uncompressed_bytes,
                                                       size_t max_uncompressed_chunk_bytes, size_t batch_size,
                                                       void* const* device_compressed_ptrs,
                                                       size_t* device_compressed_bytes,
                                                       hipcompBatchedLZ4Opts_t format_opts, hipStream_t stream);

/* ---- Host-buffer staging pipeline (no reference counterpart: reference CHANGELOG.md:97-98 leaves
 * host <-> device staging to the caller).  One call per direction for data that lives in pinned host memory:
 * the buffer is cut into `slices` that overlap H2D copies, codec kernels and D2H traffic on three streams.
 * codec: 0 LZ4, 1 Snappy, 2 Cascaded; opts: pointer to that codec's hipcompBatched*Opts_t (NULL = defaults).
 * Host buffers must be pinned and mapped (b200compHostAlloc, or cudaHostRegister'ed by the caller): the
 * compressed chunks are packed straight into host_blob by the device.  Chunk i of the blob starts at the sum
 * of host_comp_sizes[0..i).  Calls block until the host buffers are complete. */
typedef struct b200compStager b200compStager;
hipcompStatus_t b200compHostAlloc(size_t bytes, void** out);
void b200compHostFree(void* p);
hipcompStatus_t b200compStagerCreate(int codec, size_t chunk_bytes, size_t max_bytes, const void* opts, int slices,
                                     b200compStager** out);
void b200compStagerDestroy(b200compStager* s);
hipcompStatus_t b200compStagerCompress(b200compStager* s, const void* host_raw, size_t nbytes, void* host_blob,
                                       size_t* host_comp_sizes, size_t* total_comp_bytes);
hipcompStatus_t b200compStagerDecompress(b200compStager* s, const void* host_blob, const size_t* host_comp_sizes,
                                         size_t nchunks, void* host_raw_out, size_t nbytes,
                                         hipcompStatus_t* host_statuses);
/* compress + decompress of the same buffer, slice k's decompression (from the host copy of its compressed
 * chunks) queued right behind its compression so both PCIe directions stay busy */
hipcompStatus_t b200compStagerRoundTrip(b200compStager* s, const void* host_raw, size_t nbytes, void* host_blob,
                                        size_t* host_comp_sizes, void* host_raw_out, hipcompStatus_t* host_statuses,
                                        size_t* total_comp_bytes);

/* ---- C view of the high-level managers (hipcomp.hpp classes), for C / FFI callers.
 * A manager handle wraps hipcomp::LZ4Manager / SnappyManager / CascadedManager or the
 * result of hipcomp::create_manager(); a config handle wraps CompressionConfig /
 * DecompressionConfig.  C++ exceptions are mapped to status codes at this boundary
 * (message on stderr), as the reference's LLIF does (src/Check.cpp:81-101). */
typedef struct b200compManager b200compManager;
typedef struct b200compCompConfig b200compCompConfig;
typedef struct b200compDecompConfig b200compDecompConfig;

hipcompStatus_t b200compManagerCreateLZ4(size_t uncomp_chunk_size, hipcompType_t data_type, hipStream_t stream,
                                         int device_id, b200compManager** out);
hipcompStatus_t b200compManagerCreateSnappy(size_t uncomp_chunk_size, hipStream_t stream, int device_id,
                                            b200compManager** out);
/* opts: pointer to the batched Cascaded options struct of cascaded.h */
hipcompStatus_t b200compManagerCreateCascaded(const void* opts, hipStream_t stream, int device_id,
                                              b200compManager** out);
/* hipcomp::create_manager(): sniffs the container header (blocking) */
hipcompStatus_t b200compManagerCreateFromBuffer(const void* device_comp_buffer, hipStream_t stream, int device_id,
                                                b200compManager** out);
void b200compManagerDestroy(b200compManager* m);
/* Per-chunk checksums of the container.  The HLIF container reserves a u32 per chunk for the
 * compressed and for the uncompressed bytes plus two header flags (reference BatchManager.hpp:103-122,213-257); the
 * reference never fills them and defines no function for them.  mode 1: compress() stores the CRC-32 of zlib
 * (ISO-HDLC, polynomial 0xEDB88320) of every compressed and every uncompressed chunk and sets the flags;
 * decompress() recomputes both for containers whose flags are set and reports hipcompErrorCannotDecompress through
 * the config's status on a mismatch.  mode 0 (default): arrays zero, flags clear, nothing verified - what the
 * reference writes.  Either kind of container decodes with either mode and with the reference's managers. */
hipcompStatus_t b200compManagerSetChecksumMode(b200compManager* m, int mode);

hipcompStatus_t b200compManagerConfigureCompression(b200compManager* m, size_t uncompressed_bytes,
                                                    b200compCompConfig** out, size_t* max_compressed_bytes,
                                                    size_t* num_chunks);
hipcompStatus_t b200compManagerCompress(b200compManager* m, const void* device_uncompressed,
                                        void* device_compressed, const b200compCompConfig* cfg);
hipcompStatus_t b200compManagerGetCompressedOutputSize(b200compManager* m, void* device_compressed, size_t* bytes);
hipcompStatus_t b200compManagerConfigureDecompression(b200compManager* m, const void* device_compressed,
                                                      b200compDecompConfig** out, size_t* decomp_bytes,
                                                      size_t* num_chunks);
hipcompStatus_t b200compManagerConfigureDecompressionFromComp(b200compManager* m, const b200compCompConfig* comp,
                                                              b200compDecompConfig** out, size_t* decomp_bytes,
                                                              size_t* num_chunks);
hipcompStatus_t b200compManagerDecompress(b200compManager* m, void* device_uncompressed,
                                          const void* device_compressed, const b200compDecompConfig* cfg);
hipcompStatus_t b200compManagerGetRequiredScratchBufferSize(b200compManager* m, size_t* bytes);
hipcompStatus_t b200compManagerSetScratchBuffer(b200compManager* m, void* device_scratch);
/* value of the config's pinned status word (synchronise the stream first) */
hipcompStatus_t b200compCompConfigStatus(const b200compCompConfig* cfg);
hipcompStatus_t b200compDecompConfigStatus(const b200compDecompConfig* cfg);
void b200compCompConfigDestroy(b200compCompConfig* cfg);
void b200compDecompConfigDestroy(b200compDecompConfig* cfg);

/* ---- C view of the device-wide primitives (primitives.hpp: RunLengthEncodeGPU, DeltaGPU,
 * BitPackGPU; reference src/RunLengthEncodeGPU.h, DeltaGPU.h, BitPackGPU.h).  Argument
 * meaning is that of the C++ static methods; exceptions become hipcompErrorInvalidValue. */
hipcompStatus_t b200compRleRequiredWorkspaceSize(size_t num, hipcompType_t value_type, hipcompType_t count_type,
                                                 size_t* bytes);
hipcompStatus_t b200compRleCompress(void* workspace, size_t workspace_bytes, hipcompType_t value_type,
                                    void* out_values, hipcompType_t count_type, void* out_counts,
                                    size_t* num_out_device, const void* in, size_t num, hipStream_t stream);
hipcompStatus_t b200compRleCompressDownstream(void* workspace, size_t workspace_bytes, hipcompType_t value_type,
                                              void** out_values_ptr, hipcompType_t count_type,
                                              void** out_counts_ptr, size_t* num_out_device, const void* in,
                                              const size_t* num_device, size_t max_num, hipStream_t stream);
hipcompStatus_t b200compDeltaCompress(hipcompType_t value_type, void** out_values_ptr, const void* in,
                                      const size_t* num_device, size_t max_num, hipStream_t stream);
hipcompStatus_t b200compBitPackRequiredWorkspaceSize(size_t num, hipcompType_t type, size_t* bytes);
hipcompStatus_t b200compBitPackCompress(void* workspace, size_t workspace_bytes, hipcompType_t in_type,
                                        void* const* out_ptr, const void* in, const size_t* num_device,
                                        size_t max_num, void* const* min_value_device_ptr,
                                        unsigned char* const* num_bits_device_ptr, hipStream_t stream);

/* "major.minor.patch (sm_100a)" */
const char* b200compVersionString(void);

#ifdef __cplusplus
}
#endif

#endif
