// b200comp — device-wide run-length, delta and bit-packing primitives.
//
// Same classes, static methods, argument order and meaning as the reference's
// src/RunLengthEncodeGPU.h:60-133, src/DeltaGPU.h:62-96, src/BitPackGPU.h:62-103 and the
// host/device decoder src/unpack.h:68-112 (kept out of include/ there; public here so the
// tests and callers can reach them).  Conventions shared by all three ("downstream" style,
// so that the stages chain on one stream without host synchronisation):
//   - lengths may live in device memory (`numDevice`), `maxNum` is the host-known bound;
//   - output locations may be device-resident pointers (`void** outPtr`: the kernel reads
//     the pointer from device memory at run time);
//   - `workspace` is caller-owned device memory of at least requiredWorkspaceSize() bytes;
//   - errors (bad type, workspace too small, launch failure) throw std::runtime_error.
#pragma once

#include <cstddef>
#include <cstdint>
#include <type_traits>

#include "hipcomp.h"

namespace hipcomp {

class RunLengthEncodeGPU
{
public:
  // Runs of `in[0..num)` -> outValues[k], outCounts[k], k < *numOutDevice.
  static void compress(void* workspace, size_t workspaceSize, hipcompType_t valueType, void* const outValues,
                       hipcompType_t countType, void* const outCounts, size_t* numOutDevice, const void* in,
                       const size_t num, hipStream_t stream);

  // Same, with the length read from `*numDevice` (<= maxNum) and the two output arrays found
  // through device-resident pointers.
  static void compressDownstream(void* workspace, size_t workspaceSize, hipcompType_t valueType,
                                 void** const outValuesPtr, hipcompType_t countType, void** const outCountsPtr,
                                 size_t* numOutDevice, const void* in, const size_t* numDevice, const size_t maxNum,
                                 hipStream_t stream);

  static size_t requiredWorkspaceSize(size_t num, hipcompType_t valueType, hipcompType_t countType);
};

class DeltaGPU
{
public:
  // (*outValuesPtr)[i] = in[i] - in[i-1], (*outValuesPtr)[0] = in[0], i < *numDevice.
  static void compress(void* workspace, size_t workspaceSize, hipcompType_t valueType, void** const outValuesPtr,
                       const void* inValues, const size_t* numDevice, const size_t maxNum, hipStream_t stream);

  static size_t requiredWorkspaceSize(size_t num, hipcompType_t type);
};

class BitPackGPU
{
public:
  // **minValueDevicePtr = min(in), **numBitsDevicePtr = bits of (max - min); (in[i] - min) packed
  // LSB first, numBits each, into 32-bit words (64-bit words for 8-byte types) at *outPtr.
  static void compress(void* workspace, size_t workspaceSize, hipcompType_t inType, void* const* outPtr, const void* in,
                       const size_t* numDevice, size_t maxNum, void* const* const minValueDevicePtr,
                       unsigned char* const* const numBitsDevicePtr, hipStream_t stream);

  static size_t requiredWorkspaceSize(size_t num, hipcompType_t type);
};

// Element i of a BitPackGPU stream (reference src/unpack.h:68-112).
template <typename T>
#ifdef __CUDACC__
__host__ __device__
#endif
    inline T
    unpackBytes(const void* data, const uint8_t numBits, const T minValue, const size_t i)
{
  using U = typename std::make_unsigned<T>::type;
  if (numBits == 0) return minValue;
  const uint8_t* bytes = static_cast<const uint8_t*>(data);
  const size_t first_bit = i * numBits;
  const size_t first = first_bit / 8, last = (first_bit + numBits - 1) / 8;
  const unsigned shift = (unsigned)(first_bit % 8);
  // up to 9 bytes hold one value: assemble low 64 bits, then the spill byte
  uint64_t acc = 0;
  for (size_t j = first, k = 0; j <= last && k < 8; ++j, ++k) acc |= (uint64_t)bytes[j] << (8 * k);
  uint64_t v = acc >> shift;
  if (last - first == 8) v |= (uint64_t)bytes[last] << (64 - shift);
  const uint64_t mask = numBits < 64 ? ((uint64_t)1 << numBits) - 1 : ~(uint64_t)0;
  return (T)((U)(v & mask) + (U)minValue);
}

} // namespace hipcomp
