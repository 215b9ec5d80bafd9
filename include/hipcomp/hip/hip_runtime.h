/*
 * Compatibility shim: callers of the reference ABI spell the stream type
 * hipStream_t and include <hip/hip_runtime.h>.  On B200 that is the CUDA
 * runtime.  Replaces reference include/hipcomp/hip/hip_runtime.h:28-73 (only
 * the names that appear in the public headers are mapped).
 */
#ifndef B200COMP_HIP_RUNTIME_SHIM_H
#define B200COMP_HIP_RUNTIME_SHIM_H

#include <cuda_runtime_api.h>

typedef cudaStream_t hipStream_t;
typedef cudaError_t hipError_t;
#define hipSuccess cudaSuccess

#endif
