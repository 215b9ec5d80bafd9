// b200comp — shared host/device helpers for the sm_100a codec kernels.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <cstdio>
#include <string>

#include "hipcomp.h"
#pragma GCC visibility push(default)
#include "hipcomp/cascaded.h"
#pragma GCC visibility pop

namespace b200comp {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// ---------------------------------------------------------------- host side

// Error reporting in the reference's shape: "ERROR: In <fn>(): <what>" on
// stderr, status code returned (reference src/Check.cpp:81-101).
inline hipcompStatus_t fail(hipcompStatus_t code, const char* fn, const std::string& what)
{
  std::fprintf(stderr, "ERROR: In %s: %s\n", fn, what.c_str());
  return code;
}

#define B200_CHECK_NOT_NULL(ptr, fn)                                                               \
  do {                                                                                             \
    if ((ptr) == nullptr)                                                                          \
      return ::b200comp::fail(hipcompErrorInvalidValue, fn, "'" #ptr "' must not be null.");       \
  } while (0)

// The pointer arrays of the batched API must be readable by the device (reference: HipUtils::device_pointer,
// src/HipUtils.hip:91-108, maps every argument and fails with "Memory location is not accessible by the current
// GPU").  Plain host memory is rejected here instead of faulting in the kernel; pinned / managed host memory passes.
inline bool device_accessible(const void* p)
{
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return attr.type != cudaMemoryTypeUnregistered && attr.devicePointer != nullptr;
}

#define B200_CHECK_DEVICE_PTR(ptr, fn)                                                             \
  do {                                                                                             \
    if (!::b200comp::device_accessible(ptr))                                                       \
      return ::b200comp::fail(hipcompErrorInvalidValue, fn,                                        \
                              "'" #ptr "': memory location is not accessible by the current GPU."); \
  } while (0)

// NVTX range around every C-ABI call when the library is built with -DB200COMP_NVTX (make EXTRA=-DB200COMP_NVTX):
// the calls show up by name on a timeline (the reference has no tracing).  Off by default: no cost.
#ifdef B200COMP_NVTX
}  // namespace b200comp
#include <nvtx3/nvToolsExt.h>
namespace b200comp {
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};
#define B200_TRACE(fn) ::b200comp::NvtxRange b200_trace_range_(fn)
#else
#define B200_TRACE(fn) ((void)0)
#endif

inline hipcompStatus_t check_launch(const char* fn)
{
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return fail(hipcompErrorCudaError, fn, std::string("kernel launch failed: ") + cudaGetErrorString(e));
  return hipcompSuccess;
}

// Number of SMs of the current device (cached per device id).
int sm_count();

// One-time opt-in to > 48 KiB dynamic shared memory for a kernel.
template <typename K>
inline void allow_smem(K kernel, int bytes)
{
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}


inline size_t round_up_sz(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---------------------------------------------------------------- device side

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt()
{
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

template <typename V>
__device__ __forceinline__ V warp_excl_sum(V v, V& total)
{
  V incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    V o = __shfl_up_sync(kFull, incl, d);
    if (lane_id() >= d) incl += o;
  }
  total = __shfl_sync(kFull, incl, 31);
  return incl - v;
}

template <typename V>
__device__ __forceinline__ V warp_incl_max(V v)
{
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    V o = __shfl_up_sync(kFull, v, d);
    if (lane_id() >= d) v = o > v ? o : v;
  }
  return v;
}

template <typename V>
__device__ __forceinline__ V warp_min(V v)
{
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    V o = __shfl_xor_sync(kFull, v, d);
    v = o < v ? o : v;
  }
  return v;
}

template <typename V>
__device__ __forceinline__ V warp_max(V v)
{
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    V o = __shfl_xor_sync(kFull, v, d);
    v = o > v ? o : v;
  }
  return v;
}

// streaming global accesses (data touched once: keep it out of L1)
__device__ __forceinline__ uint4 ld_stream16(const void* p)
{
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void* p, uint4 v)
{
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// 16 source bytes at a 16-, 8- or 4-byte aligned address (A = guaranteed alignment), streamed past L1
template <int A>
__device__ __forceinline__ uint4 ld_stream16_a(const uint8_t* p)
{
  uint4 r;
  if (A >= 16) {
    r = ld_stream16(p);
  } else if (A == 8) {
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.z), "=r"(r.w) : "l"(p + 8));
  } else {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(p);
    r.x = __ldg(w); r.y = __ldg(w + 1); r.z = __ldg(w + 2); r.w = __ldg(w + 3);
  }
  return r;
}

template <int A>
__device__ __forceinline__ void warp_copy_vecs(uint8_t* dst, const uint8_t* src, size_t vecs)
{
  size_t v = (size_t)lane_id();
  // four independent 16-byte loads in flight per lane
  for (; v + 96 < vecs; v += 128) {
    const uint4 a = ld_stream16_a<A>(src + 16 * v), b = ld_stream16_a<A>(src + 16 * (v + 32)),
                c = ld_stream16_a<A>(src + 16 * (v + 64)), d = ld_stream16_a<A>(src + 16 * (v + 96));
    st_stream16(dst + 16 * v, a);
    st_stream16(dst + 16 * (v + 32), b);
    st_stream16(dst + 16 * (v + 64), c);
    st_stream16(dst + 16 * (v + 96), d);
  }
  for (; v < vecs; v += 32) st_stream16(dst + 16 * v, ld_stream16_a<A>(src + 16 * v));
}

// Warp-cooperative bulk copy of read-only global memory (raw / fallback partitions).  Requires src and
// dst to be congruent modulo 4; returns false (nothing copied) otherwise.  Stores are 16 bytes wide,
// loads as wide as the source alignment allows.
__device__ __forceinline__ bool warp_copy_stream(uint8_t* dst, const uint8_t* src, size_t n)
{
  if ((reinterpret_cast<uintptr_t>(dst) ^ reinterpret_cast<uintptr_t>(src)) & 3) return false;
  const size_t lane = (size_t)lane_id();
  size_t head = (size_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
  if (head > n) head = n;
  if (lane < head) dst[lane] = __ldg(src + lane);
  dst += head;
  src += head;
  n -= head;
  const size_t vecs = n >> 4;
  const unsigned sa = (unsigned)(reinterpret_cast<uintptr_t>(src) & 15);
  if (sa == 0) warp_copy_vecs<16>(dst, src, vecs);
  else if (sa == 8) warp_copy_vecs<8>(dst, src, vecs);
  else warp_copy_vecs<4>(dst, src, vecs);
  const size_t i = (vecs << 4) + lane;
  if (i < n) dst[i] = __ldg(src + i);
  return true;
}

namespace cascaded {
// hipcompBatchedCascadedDecompressAsync for a caller that knows the batch's options (a manager, the staging
// pipeline): one kernel launch instead of the C ABI's one per element width (which reads partition 0's header
// on the device, reference CascadedBatch.hip:381-429).
hipcompStatus_t decompress_known(const hipcompBatchedCascadedOpts_t& o, const void* const* comp_ptrs,
                                 const size_t* comp_bytes, const size_t* out_caps, size_t* actual, size_t batch,
                                 void* const* out_ptrs, hipcompStatus_t* statuses, cudaStream_t stream);
} // namespace cascaded

} // namespace b200comp
