/*
 * Compatibility shim: callers of the reference ABI spell the stream type
 * hipStream_t and include <hip/hip_runtime.h>.  On B200 that is the CUDA
 * runtime.  Replaces reference include/hipcomp/hip/hip_runtime.h:28-73 (only
 * the names that appear in the public headers are mapped).
 */
#ifndef B200COMP_HIP_RUNTIME_SHIM_H
#define B200COMP_HIP_RUNTIME_SHIM_H

#ifdef __cplusplus
#include <cuda_runtime.h> /* typed cudaMalloc / cudaMallocHost overloads, as hipMalloc has in C++ */
#else
#include <cuda_runtime_api.h>
#endif

typedef cudaStream_t hipStream_t;
typedef cudaError_t hipError_t;
#define hipSuccess cudaSuccess

/* The runtime calls a hipCOMP caller makes around the library (the reference ships the same convenience map in
 * include/hipcomp/hip/hip_runtime.h:55-120 for its CUDA backend), so such code builds unchanged against this
 * tree.  Only names; everything is the CUDA runtime. */
typedef cudaEvent_t hipEvent_t;
typedef struct cudaDeviceProp hipDeviceProp_t;
typedef struct cudaPointerAttributes hipPointerAttribute_t;
#define hipMemcpyKind cudaMemcpyKind
#define hipMemcpyHostToDevice cudaMemcpyHostToDevice
#define hipMemcpyDeviceToHost cudaMemcpyDeviceToHost
#define hipMemcpyDeviceToDevice cudaMemcpyDeviceToDevice
#define hipMemcpyDefault cudaMemcpyDefault
#define hipHostMallocDefault cudaHostAllocDefault
#define hipMemoryTypeDevice cudaMemoryTypeDevice
#define hipErrorInvalidValue cudaErrorInvalidValue
#define hipMalloc cudaMalloc
#define hipMallocAsync cudaMallocAsync
#define hipMallocManaged cudaMallocManaged
#define hipHostMalloc cudaMallocHost
#define hipFree cudaFree
#define hipFreeAsync cudaFreeAsync
#define hipHostFree cudaFreeHost
#define hipMemcpy cudaMemcpy
#define hipMemcpyAsync cudaMemcpyAsync
#define hipMemset cudaMemset
#define hipMemsetAsync cudaMemsetAsync
#define hipStreamCreate cudaStreamCreate
#define hipStreamDestroy cudaStreamDestroy
#define hipStreamSynchronize cudaStreamSynchronize
#define hipDeviceSynchronize cudaDeviceSynchronize
#define hipEventCreate cudaEventCreate
#define hipEventDestroy cudaEventDestroy
#define hipEventRecord cudaEventRecord
#define hipEventSynchronize cudaEventSynchronize
#define hipEventElapsedTime cudaEventElapsedTime
#define hipGetLastError cudaGetLastError
#define hipGetErrorString cudaGetErrorString
#define hipGetDevice cudaGetDevice
#define hipSetDevice cudaSetDevice
#define hipGetDeviceProperties cudaGetDeviceProperties
#define hipDeviceGetAttribute cudaDeviceGetAttribute
#define hipPointerGetAttributes cudaPointerGetAttributes

#endif
