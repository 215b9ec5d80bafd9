// b200comp — per-chunk CRC-32 for the HLIF container's checksum arrays.
//
// The container reserves a u32 per chunk for the compressed and for the uncompressed bytes
// (reference src/highlevel/BatchManager.hpp:103-122,213-257) and two header flags; the reference never
// fills them (hlif_shared.hiph:124-127) and defines no function for them.  Here the function is the CRC-32
// of zlib (ISO-HDLC, reflected polynomial 0xEDB88320), so that any consumer can check a chunk with
// crc32() of zlib / Python / Go without this library.
//
// One warp per chunk.  Lane 0 takes the head of the chunk, lanes 1..31 equal slices of L bytes behind it;
// every lane runs slice-by-4 over its slice (tables in shared memory), and the 32 finished CRCs are
// combined in a 5-level tree with crc(A || B) = x^(8 |B|) * crc(A) + crc(B) in GF(2)[x] / P (what zlib's
// crc32_combine computes): every right-hand block of the tree is a whole number of slices, so one
// operator x^(8L) per chunk and one squaring per level do.  tests/test_oracle_hlif_crc.py holds the same
// arithmetic (in Python) against zlib.
#pragma once
#include "common.cuh"

namespace b200comp {
namespace crc {

constexpr uint32_t kPoly = 0xEDB88320u;
constexpr int kWarps = 8;

// x^(2^n) mod P, n = 0..31 (zlib's x2n_table)
__constant__ uint32_t kX2n[32] = {
    0x40000000u, 0x20000000u, 0x08000000u, 0x00800000u, 0x00008000u, 0xedb88320u, 0xb1e6b092u, 0xa06a2517u,
    0xed627daeu, 0x88d14467u, 0xd7bbfe6au, 0xec447f11u, 0x8e7ea170u, 0x6427800eu, 0x4d47bae0u, 0x09fe548fu,
    0x83852d0fu, 0x30362f1au, 0x7b5a9cc3u, 0x31fec169u, 0x9fec022au, 0x6c8dedc4u, 0x15d6874du, 0x5fde7a4eu,
    0xbad90e37u, 0x2e4e5eefu, 0x4eaba214u, 0xa8a472c0u, 0x429a969eu, 0x148d302au, 0xc40ba6d0u, 0xc4e22c3cu};

// a(x) * b(x) mod P, reflected bit order (bit 31 is x^0)
__device__ __forceinline__ uint32_t multmodp(uint32_t a, uint32_t b)
{
  uint32_t p = 0;
#pragma unroll 8
  for (int i = 31; i >= 0; --i) {
    p ^= b & (0u - ((a >> i) & 1u));
    b = (b >> 1) ^ (kPoly & (0u - (b & 1u)));
  }
  return p;
}

// x^(n * 2^k) mod P
__device__ __forceinline__ uint32_t x2nmodp(uint32_t n, uint32_t k)
{
  uint32_t p = 0x80000000u; // x^0
  for (; n; n >>= 1, ++k)
    if (n & 1u) p = multmodp(kX2n[k & 31u], p);
  return p;
}

// finished CRC-32 of [q, e)
__device__ __forceinline__ uint32_t crc_bytes(const uint8_t* q, const uint8_t* e, const uint32_t* t)
{
  uint32_t c = ~0u;
  while (q < e && (reinterpret_cast<uintptr_t>(q) & 3)) c = (c >> 8) ^ t[(c ^ __ldg(q++)) & 0xffu];
  for (; q + 4 <= e; q += 4) {
    c ^= __ldg(reinterpret_cast<const uint32_t*>(q));
    c = t[768 + (c & 0xffu)] ^ t[512 + ((c >> 8) & 0xffu)] ^ t[256 + ((c >> 16) & 0xffu)] ^ t[c >> 24];
  }
  while (q < e) c = (c >> 8) ^ t[(c ^ __ldg(q++)) & 0xffu];
  return ~c;
}

// CRC-32 of every chunk (ptrs[i], sizes[i]), i < n.
//   out != nullptr      : out[i] = crc
//   expected != nullptr : where *present (or present == nullptr) a chunk whose crc differs from expected[i]
//                         sets *bad = hipcompErrorCannotDecompress
__global__ void __launch_bounds__(kWarps * 32)
crc32_chunks_kernel(const void* const* ptrs, const size_t* sizes, size_t n, uint32_t* out, const uint32_t* expected,
                    const bool* present, hipcompStatus_t* bad)
{
  __shared__ uint32_t t[4 * 256];
  if (expected != nullptr && present != nullptr && !*present) return; // the container carries no checksums
  {
    uint32_t c = threadIdx.x;
#pragma unroll
    for (int k = 0; k < 8; ++k) c = (c >> 1) ^ (kPoly & (0u - (c & 1u)));
    t[threadIdx.x] = c;
    __syncthreads();
    for (int k = 1; k < 4; ++k) {
      const uint32_t prev = t[256 * (k - 1) + threadIdx.x];
      t[256 * k + threadIdx.x] = (prev >> 8) ^ t[prev & 0xffu];
      __syncthreads();
    }
  }
  const int lane = lane_id();
  const size_t stride = (size_t)gridDim.x * kWarps;
  for (size_t i = (size_t)blockIdx.x * kWarps + (threadIdx.x >> 5); i < n; i += stride) {
    const uint8_t* p = static_cast<const uint8_t*>(ptrs[i]);
    const size_t len = p ? sizes[i] : 0;
    const size_t L = (len / 32) & ~(size_t)3;
    const size_t head = len - 31 * L;
    const size_t a = lane == 0 ? 0 : head + (size_t)(lane - 1) * L;
    const size_t b = lane == 0 ? head : a + L;
    uint32_t c = crc_bytes(p + a, p + b, t);
    uint32_t op = x2nmodp((uint32_t)L, 3); // x^(8L); L < 2^29 for every chunk size the codecs accept
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t right = __shfl_down_sync(kFull, c, s);
      c = multmodp(op, c) ^ right; // meaningful on lanes that are multiples of 2s
      op = multmodp(op, op);
    }
    if (lane == 0) {
      if (out) out[i] = c;
      if (expected && c != expected[i]) *bad = hipcompErrorCannotDecompress;
    }
  }
}

} // namespace crc
} // namespace b200comp
