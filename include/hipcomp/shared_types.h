/*
 * b200comp — status codes shared by every entry point of the batched codec ABI.
 *
 * Replaces: reference include/hipcomp/shared_types.h:52-66 (same enumerator
 * names and numeric values, including the nvcomp* aliases, so existing callers
 * relink without source changes).
 */
#pragma once

typedef enum hipcompStatus_t
{
  hipcompSuccess = 0,
  hipcompErrorInvalidValue = 10,
  hipcompErrorNotSupported = 11,
  hipcompErrorCannotDecompress = 12,
  hipcompErrorCudaError = 1000,
  hipcompErrorInternal = 10000,
  /* nvCOMP 2.2 spellings kept as aliases */
  nvcompSuccess = hipcompSuccess,
  nvcompErrorInvalidValue = hipcompErrorInvalidValue,
  nvcompErrorNotSupported = hipcompErrorNotSupported,
  nvcompErrorCannotDecompress = hipcompErrorCannotDecompress,
  nvcompErrorCudaError = hipcompErrorCudaError,
  nvcompErrorInternal = hipcompErrorInternal,
} hipcompStatus_t;
