// b200comp — building blocks of the "team" LZ encoders (one CTA per chunk, chunk and hash
// table resident in shared memory): mbarrier / bulk-copy (TMA 1-D) wrappers, unaligned
// shared-memory word loads, warp-cooperative shared -> global copies.
#pragma once

#include "common.cuh"

namespace b200comp {
namespace team {

constexpr uint32_t kSeg = 1024; // positions per segment (the unit a warp parses on its own)

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence()
{
  // make the initialised barriers visible to the async proxy (bulk copies complete on them)
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// blocks until the phase with the given parity has completed (try_wait suspends in hardware for a while; the
// short sleep between attempts keeps a waiting warp from eating the issue slots of the working ones)
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
  const uint32_t a = smem_addr(bar);
  for (;;) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(a), "r"(parity)
        : "memory");
    if (done) return;
    __nanosleep(64);
  }
}
// generic-proxy accesses to shared memory made so far are ordered before later async-proxy ones
__device__ __forceinline__ void fence_proxy_async()
{
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on `bar`.  Both addresses and
// the size are multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_addr(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}

// explicit shared-space loads on 32-bit shared addresses (no generic -> shared conversion in the hot loops).
// volatile: kept in program order with the barrier waits; the data they read is written before a barrier.
__device__ __forceinline__ uint32_t lds_u8(uint32_t sa)
{
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(sa));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t sa)
{
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(sa));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t sa)
{
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(sa));
  return v;
}
__device__ __forceinline__ void sts_u16(uint32_t sa, uint32_t v)
{
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(sa), "h"((unsigned short)v) : "memory");
}
// table accesses of the encoders' serial section (explicit shared-space instructions, kept in program order)
__device__ __forceinline__ uint32_t lds_u32_ordered(uint32_t sa)
{
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(sa) : "memory");
  return v;
}
__device__ __forceinline__ void red_max_u32(uint32_t sa, uint32_t v)
{
  asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(sa), "r"(v) : "memory");
}
// 4 bytes at any shared address: two aligned words + funnel shift
__device__ __forceinline__ uint32_t lds_u32_unaligned(uint32_t sa)
{
  const uint32_t al = sa & ~3u;
  return __funnelshift_r(lds_u32(al), lds_u32(al + 4u), (sa & 3u) * 8u);
}

// 4 bytes at byte offset x of a 4-byte aligned shared-memory array (two aligned words + funnel shift)
__device__ __forceinline__ uint32_t lds_u32_at(const uint32_t* base, uint32_t x)
{
  const uint32_t lo = base[x >> 2], hi = base[(x >> 2) + 1];
  return __funnelshift_r(lo, hi, (x & 3u) * 8u);
}

// length-in-0xFF-bytes field of LZ4, written by the warp; returns its size
__device__ __forceinline__ uint32_t put_lsic(uint8_t* out, uint32_t x)
{
  const uint32_t num = x / 255u + 1u;
  for (uint32_t i = (uint32_t)lane_id(); i < num; i += 32) out[i] = (i + 1 < num) ? (uint8_t)0xff : (uint8_t)(x % 255u);
  return num;
}

// Warp-cooperative copy of n bytes from shared memory (byte offset `x` of the 4-byte aligned
// array `base`) to global memory at any alignment: 16-byte stores once the destination is aligned.
// May read up to 3 bytes past the source range (callers keep that inside their allocation).
__device__ __forceinline__ void copy_s2g(uint8_t* dst, const uint32_t* base, uint32_t x, uint32_t n)
{
  const uint32_t lane = (uint32_t)lane_id();
  const uint8_t* sb = reinterpret_cast<const uint8_t*>(base);
  if (n < 64u) {
    for (uint32_t i = lane; i < n; i += 32) dst[i] = sb[x + i];
    return;
  }
  const uint32_t head = (uint32_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
  if (lane < head) dst[lane] = sb[x + lane];
  dst += head;
  x += head;
  n -= head;
  const uint32_t vecs = n >> 4;
  const uint32_t sh = (x & 3u) * 8u;
  const uint32_t* sw = base + (x >> 2);
  uint4* dv = reinterpret_cast<uint4*>(dst);
  for (uint32_t v = lane; v < vecs; v += 32) {
    const uint32_t* s = sw + 4 * v;
    const uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4];
    uint4 r;
    r.x = __funnelshift_r(a, b, sh);
    r.y = __funnelshift_r(b, c, sh);
    r.z = __funnelshift_r(c, d, sh);
    r.w = __funnelshift_r(d, e, sh);
    dv[v] = r;
  }
  for (uint32_t i = (vecs << 4) + lane; i < n; i += 32) dst[i] = sb[x + i];
}

// Stage a chunk of n bytes from global memory into shared memory so that byte i of the chunk
// lands at s_chunk[mis + i], mis = src & 15 (same 16-byte phase as the source, which makes the
// interior one bulk copy).  Called by the whole CTA; the caller then waits on `bar`.
__device__ __forceinline__ uint32_t stage_chunk(uint8_t* s_chunk, const uint8_t* src, uint32_t n, uint64_t* bar)
{
  const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15);
  uint32_t head = (16u - mis) & 15u;
  if (head > n) head = n;
  const uint32_t body = (n - head) & ~15u;
  const uint32_t tail = n - head - body;
  if (threadIdx.x == 0) {
    fence_proxy_async();
    mbar_arrive_expect_tx(bar, body);
    if (body) bulk_g2s(s_chunk + mis + head, src + head, body, bar);
  }
  if (threadIdx.x < head) s_chunk[mis + threadIdx.x] = __ldg(src + threadIdx.x);
  if (threadIdx.x >= 32 && threadIdx.x < 32 + tail) {
    const uint32_t i = head + body + (threadIdx.x - 32);
    s_chunk[mis + i] = __ldg(src + i);
  }
  return mis;
}

} // namespace team
} // namespace b200comp
