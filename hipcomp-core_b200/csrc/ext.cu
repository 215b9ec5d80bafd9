// b200comp — helpers around the codecs (not part of the reference ABI).
#include "lz_common.cuh"

#pragma GCC visibility push(default)
#include "hipcomp/b200_ext.h"
#pragma GCC visibility pop

namespace b200comp {

__global__ void __launch_bounds__(256)
gather_kernel(const void* const* ptrs, const size_t* bytes, const size_t* dst_off, uint8_t* dst, size_t batch)
{
  const size_t stride = (size_t)gridDim.x * 8;
  for (size_t i = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < batch; i += stride)
    warp_copy_ro(dst + dst_off[i], static_cast<const uint8_t*>(ptrs[i]), (uint32_t)bytes[i]);
}

} // namespace b200comp

using namespace b200comp;

extern "C" {

hipcompStatus_t b200compGatherChunksAsync(const void* const* device_ptrs, const size_t* device_bytes,
                                          const size_t* device_dst_offsets, void* device_dst, size_t batch_size,
                                          hipStream_t stream)
{
  static const char* fn = "b200compGatherChunksAsync()";
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_ptrs, fn);
  B200_CHECK_NOT_NULL(device_bytes, fn);
  B200_CHECK_NOT_NULL(device_dst_offsets, fn);
  B200_CHECK_NOT_NULL(device_dst, fn);
  const size_t need = (batch_size + 7) / 8;
  const size_t cap = (size_t)sm_count() * 8;
  gather_kernel<<<(int)(need < cap ? need : cap), 256, 0, stream>>>(device_ptrs, device_bytes, device_dst_offsets,
                                                                    static_cast<uint8_t*>(device_dst), batch_size);
  return check_launch(fn);
}

const char* b200compVersionString(void) { return "2.2.0 (b200comp, sm_100a)"; }

} // extern "C"
