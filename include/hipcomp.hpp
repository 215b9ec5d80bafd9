/*
 * b200comp — C++ companion of hipcomp.h (drop-in for the reference's include/hipcomp.hpp:61-203): the exception
 * type thrown by the high-level interface, the abstract one-buffer Compressor / Decompressor interfaces and the
 * element-type helper.  Interface only; the batched C API (hipcomp/*.h) and the managers (hipcomp/*.hpp) are the
 * implemented surface.
 */
#ifndef HIPCOMP_API_HPP
#define HIPCOMP_API_HPP

#include "hipcomp.h"
#include "hipcomp/lz4.h"

#include <cstdint>
#include <hip/hip_runtime.h>
#include <stdexcept>
#include <string>
#include <type_traits>

namespace hipcomp
{

/* error carrying a hipcompStatus_t (reference hipcomp.hpp:71-88) */
class HipCompException : public std::runtime_error
{
public:
  HipCompException(hipcompStatus_t err, const std::string& msg)
      : std::runtime_error(msg + " : code=" + std::to_string(err) + "."), m_err(err)
  {
  }
  hipcompStatus_t get_error() const { return m_err; }

private:
  hipcompStatus_t m_err;
};

/* one-buffer compressor interface (reference hipcomp.hpp:100-134) */
class Compressor
{
public:
  virtual ~Compressor() = default;
  virtual void configure(const size_t in_bytes, size_t* temp_bytes, size_t* out_bytes) = 0;
  virtual void compress_async(const void* in_ptr, const size_t in_bytes, void* temp_ptr, const size_t temp_bytes,
                              void* out_ptr, size_t* out_bytes, hipStream_t stream)
      = 0;
};

/* one-buffer decompressor interface (reference hipcomp.hpp:140-166) */
class Decompressor
{
public:
  virtual ~Decompressor() = default;
  virtual void configure(const void* in_ptr, const size_t in_bytes, size_t* temp_bytes, size_t* out_bytes,
                         hipStream_t stream)
      = 0;
  virtual void decompress_async(const void* in_ptr, const size_t in_bytes, void* temp_ptr, const size_t temp_bytes,
                                void* out_ptr, const size_t out_bytes, hipStream_t stream)
      = 0;
};

/* hipcompType_t of a C++ integer type (reference hipcomp.hpp:169-192) */
template <typename T>
inline hipcompType_t TypeOf()
{
  if (std::is_same<T, int8_t>::value) return HIPCOMP_TYPE_CHAR;
  if (std::is_same<T, uint8_t>::value) return HIPCOMP_TYPE_UCHAR;
  if (std::is_same<T, int16_t>::value) return HIPCOMP_TYPE_SHORT;
  if (std::is_same<T, uint16_t>::value) return HIPCOMP_TYPE_USHORT;
  if (std::is_same<T, int32_t>::value) return HIPCOMP_TYPE_INT;
  if (std::is_same<T, uint32_t>::value) return HIPCOMP_TYPE_UINT;
  if (std::is_same<T, int64_t>::value) return HIPCOMP_TYPE_LONGLONG;
  if (std::is_same<T, uint64_t>::value) return HIPCOMP_TYPE_ULONGLONG;
  throw HipCompException(hipcompErrorNotSupported, "hipcomp does not support the given type.");
}

inline void throwExceptionIfError(hipcompStatus_t error, const std::string& msg)
{
  if (error != hipcompSuccess) throw HipCompException(error, msg);
}

} // namespace hipcomp

#endif
