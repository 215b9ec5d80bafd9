// b200comp — warp-level byte movers shared by the LZ4 and Snappy kernels.
#pragma once

#include "common.cuh"

namespace b200comp {

// 4 bytes at an arbitrary byte address of read-only data: two aligned word
// loads + funnel shift.  Touches at most the aligned words that contain
// [p, p+4), so it never reads a byte outside the 4-byte-aligned hull of the
// requested range.
__device__ __forceinline__ uint32_t ld_unaligned32_ro(const uint8_t* p)
{
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(a & 3) * 8;
  const uint32_t lo = __ldg(w);
  if (sh == 0) return lo;
  const uint32_t hi = __ldg(w + 1);
  return __funnelshift_r(lo, hi, sh);
}

// Warp-cooperative copy of n bytes from read-only memory to `dst`.
// Short runs: one byte per lane.  Long runs: byte prologue up to a 16-byte
// aligned destination, then 16-byte stores fed by aligned word loads
// (funnel-shifted when source and destination are mutually misaligned).
__device__ __forceinline__ void warp_copy_ro(uint8_t* dst, const uint8_t* src, uint32_t n)
{
  const uint32_t lane = (uint32_t)lane_id();
  if (n < 96) {
    for (uint32_t i = lane; i < n; i += 32) dst[i] = __ldg(src + i);
    return;
  }
  const uint32_t head = (uint32_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
  if (lane < head) dst[lane] = __ldg(src + lane);
  dst += head;
  src += head;
  n -= head;
  const uint32_t vecs = n >> 4;
  const uintptr_t sa = reinterpret_cast<uintptr_t>(src);
  const uint32_t* sw = reinterpret_cast<const uint32_t*>(sa & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(sa & 3) * 8;
  uint4* dv = reinterpret_cast<uint4*>(dst);
  if (sh == 0) {
    for (uint32_t v = lane; v < vecs; v += 32) {
      const uint32_t* s = sw + 4 * v;
      uint4 r;
      r.x = __ldg(s); r.y = __ldg(s + 1); r.z = __ldg(s + 2); r.w = __ldg(s + 3);
      dv[v] = r;
    }
  } else {
    for (uint32_t v = lane; v < vecs; v += 32) {
      const uint32_t* s = sw + 4 * v;
      const uint32_t a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2), d = __ldg(s + 3), e = __ldg(s + 4);
      uint4 r;
      r.x = __funnelshift_r(a, b, sh);
      r.y = __funnelshift_r(b, c, sh);
      r.z = __funnelshift_r(c, d, sh);
      r.w = __funnelshift_r(d, e, sh);
      dv[v] = r;
    }
  }
  for (uint32_t i = (vecs << 4) + lane; i < n; i += 32) dst[i] = __ldg(src + i);
}

// Warp-cooperative LZ77 match copy inside the output buffer:
// dst[i] = dst[i - off] for i in [0, len), semantics of a sequential byte copy.
// Callers have made all earlier writes to the buffer visible (__syncwarp()).
__device__ __forceinline__ void warp_copy_match(uint8_t* dst, uint32_t off, uint32_t len)
{
  const uint32_t lane = (uint32_t)lane_id();
  const uint8_t* src = dst - off;
  if (off >= len) {
    // no overlap: plain copy (data was written by this warp: coherent loads)
    for (uint32_t i = lane; i < len; i += 32) dst[i] = src[i];
  } else if (off >= 32) {
    // overlap further than one warp-wide step: steps are ordered by a barrier
    for (uint32_t base = 0; base < len; base += 32) {
      const uint32_t i = base + lane;
      uint8_t v = 0;
      if (i < len) v = src[i];
      if (i < len) dst[i] = v;
      __syncwarp();
    }
  } else {
    // periodic pattern of `off` bytes that all precede dst
    for (uint32_t i = lane; i < len; i += 32) dst[i] = src[i % off];
  }
}

} // namespace b200comp
