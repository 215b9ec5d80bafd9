// b200comp — shared host/device helpers for the sm_100a codec kernels.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <cstdio>
#include <string>

#include "hipcomp.h"

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

// 4, 8, 16 or 32 from an environment variable (tuning knob), else 0.
int group_lanes_from_env(const char* name);

// LZ4 / Snappy decoder variant knob (B200COMP_{LZ4,SNAPPY}_DEC_G): unset -> default,
// (LZ4: "win", the window decoder; Snappy: "round"), "pipe" -> decoupled parse/exec pipeline, "round" -> one-warp
// round decoder, "win" -> window decoder (shared-memory staged token walk),
// 4/8/16/32 -> v2 sub-warp group decoder of that width.
enum : int { kDecoderDefault = -1, kDecoderPipeline = 1, kDecoderRound = 0, kDecoderWindow = 2 };
int decoder_variant_from_env(const char* name);

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

} // namespace b200comp
