// b200comp — batched Cascaded codec (RLE -> delta -> bit-packing) for sm_100a.
//
// Wire format: bit-exact with reference src/CascadedKernels.hiph (see DESIGN.md
// "Cascaded wire format"); C ABI replaces src/lowlevel/CascadedBatch.hip:306-462.
//
// Kernel design (not the reference's): the reference runs one 128-thread CTA per
// partition with CUB block scans/reductions and __syncthreads() between every
// stage.  Here ONE WARP owns a partition and walks its 4096-byte chunks with
// warp-synchronous primitives only.  A chunk is staged in shared memory in a
// 16-byte XOR-swizzled layout, so that both access patterns the layers need are
// bank-conflict free AND 128 bits wide:
//   striped  (lane L takes vector L of each 512-byte tile)  - run detection,
//            compaction, deltas, scans, min/max, gathers
//   blocked  (lane g takes elements [32g, 32g+32))           - bit packing with
//            compile-time bit widths (no per-element shift bookkeeping)
// An SM keeps up to 32 independent partitions in flight with no block barrier.
// Grids are persistent (multiple of the SM count) and grid-stride over the batch.
//
// Fused paths for 4-byte elements on full, 16-byte aligned chunks (same bytes as the generic layer loop):
//   encode  num_RLEs 1 or 2, num_deltas 1: compress_chunk_rd1 - run flags from registers, only compacted
//           (value, end) pairs touch shared memory; for num_RLEs 2 the second layer is rle_encode_minmax.
//           Partitions whose first chunk turns out all distinct are sized ahead (partition_cannot_fit)
//           and stored raw at once when they cannot fit.
//   decode  num_RLEs 1: decompress_chunk_r1d1; num_RLEs 2 (library default): decompress_chunk_r2d1;
//           any deeper stack that ends in (first RLE, one delta): decompress_tail_d1 behind the generic layers.
//           All three stage the stream in shared memory, scatter deltas onto run starts of a zeroed array and
//           finish with one register prefix sum over the 1024 outputs.
#include "common.cuh"

#include <type_traits>

#pragma GCC visibility push(default)
#include "hipcomp/cascaded.h"
#include "hipcomp/b200_ext.h"
#pragma GCC visibility pop

namespace b200comp {
namespace cascaded {

constexpr int kChunkBytes = 4096;
constexpr int kWarpsPerCta = 4;
constexpr int kMaxMetaBytes = 64;
#ifndef B200_CASC_COMP_MINB
#define B200_CASC_COMP_MINB 5
#endif

// Swizzled shared-memory array of E.  Element i lives at p(i): the 16-byte vector
// index is XORed with (i / 32) & 7.  Vectors stay contiguous and aligned.
template <typename E>
struct Swz {
  static constexpr int VE = 16 / (int)sizeof(E); // elements per 16-byte vector
  static constexpr int SH = VE == 16 ? 4 : VE == 8 ? 3 : VE == 4 ? 2 : 1;
  __device__ __forceinline__ static int p(int i) { return i ^ (((i >> 5) & 7) << SH); }
};

template <typename E>
union Vec {
  uint4 q;
  E e[16 / sizeof(E)];
};

// vector that holds elements [i0, i0 + VE), i0 a multiple of VE
template <typename E>
__device__ __forceinline__ Vec<E> ld_vec(const E* arr, int i0)
{
  Vec<E> v;
  v.q = *reinterpret_cast<const uint4*>(arr + Swz<E>::p(i0));
  return v;
}
template <typename E>
__device__ __forceinline__ void st_vec(E* arr, int i0, const Vec<E>& v)
{
  *reinterpret_cast<uint4*>(arr + Swz<E>::p(i0)) = v.q;
}

__host__ __device__ constexpr int meta_size(int w, int R, int D)
{
  return ((4 + 4 * (R + 1) + w - 1) / w) * w + ((w * D + 3) / 4) * 4;
}

template <typename E>
__device__ __forceinline__ constexpr int packed_header_bytes()
{
  return sizeof(E) == 8 ? 16 : 8; // roundUp(sizeof(E)+4, max(4,sizeof(E)))
}

template <typename V>
__device__ __forceinline__ V shfl_down_t(V v, int d)
{
  return __shfl_down_sync(kFull, v, d);
}
template <typename V>
__device__ __forceinline__ V shfl_up_t(V v, int d)
{
  return __shfl_up_sync(kFull, v, d);
}
template <typename V>
__device__ __forceinline__ V shfl_t(V v, int l)
{
  return __shfl_sync(kFull, v, l);
}
// sub-word element types travel as 32-bit values
__device__ __forceinline__ uint8_t shfl_down_t(uint8_t v, int d) { return (uint8_t)__shfl_down_sync(kFull, (uint32_t)v, d); }
__device__ __forceinline__ uint16_t shfl_down_t(uint16_t v, int d) { return (uint16_t)__shfl_down_sync(kFull, (uint32_t)v, d); }
__device__ __forceinline__ uint8_t shfl_t(uint8_t v, int l) { return (uint8_t)__shfl_sync(kFull, (uint32_t)v, l); }
__device__ __forceinline__ uint16_t shfl_t(uint16_t v, int l) { return (uint16_t)__shfl_sync(kFull, (uint32_t)v, l); }

// ------------------------------------------------------------------ bit packing

// 32 values of BW bits each -> BW words, LSB first (bit i*BW of the group).
// Everything folds at compile time: per element one shift+or per word it touches.
template <int BW>
__device__ __forceinline__ void pack32_static(const uint32_t (&d)[32], uint32_t* dst, int first_word, int words)
{
#pragma unroll
  for (int t = 0; t < BW; ++t) {
    uint32_t w = 0;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const int lo = k * BW - 32 * t; // bit position of element k relative to word t
      if (lo > -BW && lo < 32) {
        if (lo >= 0) w |= d[k] << lo;
        else w |= d[k] >> (-lo);
      }
    }
    if (first_word + t < words) dst[first_word + t] = w;
  }
}

template <int MAXBW>
__device__ __forceinline__ void pack32_dispatch(uint32_t bw, const uint32_t (&d)[32], uint32_t* dst, int first_word,
                                                int words)
{
  switch (bw) {
#define B200_CASE(B)                                                                                                 \
  case B:                                                                                                            \
    if (B <= MAXBW) pack32_static<(B <= MAXBW ? B : 1)>(d, dst, first_word, words);                                  \
    break;
    B200_CASE(1) B200_CASE(2) B200_CASE(3) B200_CASE(4) B200_CASE(5) B200_CASE(6) B200_CASE(7) B200_CASE(8)
    B200_CASE(9) B200_CASE(10) B200_CASE(11) B200_CASE(12) B200_CASE(13) B200_CASE(14) B200_CASE(15) B200_CASE(16)
    B200_CASE(17) B200_CASE(18) B200_CASE(19) B200_CASE(20) B200_CASE(21) B200_CASE(22) B200_CASE(23) B200_CASE(24)
    B200_CASE(25) B200_CASE(26) B200_CASE(27) B200_CASE(28) B200_CASE(29) B200_CASE(30) B200_CASE(31) B200_CASE(32)
#undef B200_CASE
  default: break;
  }
}

// generic path for 64-bit elements: lane g packs elements [32g, 32g+32)
__device__ __forceinline__ void pack32_u64(uint32_t bw, const unsigned long long* arr, int g, int n,
                                           unsigned long long fr, uint32_t* data, int words)
{
  unsigned long long acc = 0;
  int nbits = 0;
  int wout = g * (int)bw;
  for (int k = 0; k < 32; ++k) {
    const int e = (g << 5) + k;
    const unsigned long long d = e < n ? arr[Swz<unsigned long long>::p(e)] - fr : 0ull;
    if (bw > 32) {
      acc |= (unsigned long long)(uint32_t)d << nbits;
      if (wout < words) data[wout] = (uint32_t)acc;
      ++wout;
      acc >>= 32;
      acc |= (d >> 32) << nbits;
      nbits += (int)bw - 32;
    } else {
      acc |= d << nbits;
      nbits += (int)bw;
    }
    if (nbits >= 32) {
      if (wout < words) data[wout] = (uint32_t)acc;
      ++wout;
      acc >>= 32;
      nbits -= 32;
    }
  }
  if (nbits > 0 && wout < words) data[wout] = (uint32_t)acc;
}

// Serialise arr[0..n) (swizzled smem) to `out` (global, 4-byte aligned; for raw
// 8-byte elements 8-byte aligned), raw or bit-packed.  Returns false when the
// 4-byte padded array does not fit in `avail` bytes.  Follows block_write /
// block_bitpack / get_for_bitwidth (reference CascadedKernels.hiph:394-553,646-680).
template <typename E, int N>
__device__ bool write_array(const E* arr, int n, uint8_t* out, size_t avail, bool use_bp, uint32_t& out_bytes)
{
  using S = typename std::make_signed<E>::type;
  using Z = Swz<E>;
  constexpr int VE = Z::VE;
  const int lane = lane_id();
  uint32_t* outw = reinterpret_cast<uint32_t*>(out);

  if (!use_bp) {
    const uint32_t bytes = (uint32_t)n * sizeof(E);
    const uint32_t padded = (bytes + 3u) & ~3u;
    if (padded > avail) return false;
    int done = 0; // elements stored by the vector path
    if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) {
      const int vecs = n / VE;
      uint4* ov = reinterpret_cast<uint4*>(out);
      for (int v = lane; v < vecs; v += 32) ov[v] = ld_vec<E>(arr, v * VE).q;
      done = vecs * VE;
    }
    if (sizeof(E) >= 4) {
      E* oe = reinterpret_cast<E*>(out);
      for (int i = done + lane; i < n; i += 32) oe[i] = arr[Z::p(i)];
    } else {
      constexpr int K = sizeof(E) < 4 ? 4 / (int)sizeof(E) : 1;
      const int words = padded / 4;
      for (int j = done / K + lane; j < words; j += 32) {
        uint32_t wv = 0;
#pragma unroll
        for (int c = 0; c < K; ++c) {
          const int e = j * K + c;
          const uint32_t v = e < n ? (uint32_t)arr[Z::p(e)] : 0u;
          wv |= v << (8 * sizeof(E) * c);
        }
        outw[j] = wv;
      }
    }
    out_bytes = bytes;
    return true;
  }

  // frame of reference = signed minimum; width from (max - min)
  S mn = 0, mx = 0;
  uint32_t bw = 0;
  if (n > 0) {
    S lmn = (S)arr[0], lmx = lmn; // element 0 is valid (p(0) == 0)
    for (int i0 = lane * VE; i0 < n; i0 += 32 * VE) {
      const Vec<E> v = ld_vec<E>(arr, i0);
#pragma unroll
      for (int j = 0; j < VE; ++j) {
        const S x = i0 + j < n ? (S)v.e[j] : lmn;
        lmn = x < lmn ? x : lmn;
        lmx = x > lmx ? x : lmx;
      }
    }
    mn = warp_min(lmn);
    mx = warp_max(lmx);
    if (sizeof(E) > 4) {
      const unsigned long long range = (unsigned long long)mx - (unsigned long long)mn;
      bw = 64 - __clzll((long long)range);
    } else {
      const uint32_t range = (uint32_t)(int32_t)mx - (uint32_t)(int32_t)mn;
      bw = 32 - __clz((int)range);
    }
  }
  const E fr = (E)mn;
  const int words = (int)(((uint32_t)n * bw + 31u) >> 5);
  constexpr int HDR = packed_header_bytes<E>();
  const uint32_t bytes = HDR + 4u * (uint32_t)words;
  if (bytes > avail) return false;

  if (lane == 0) {
    const uint32_t meta = (bw << 16) | (uint32_t)n;
    if (sizeof(E) == 8) {
      outw[0] = (uint32_t)((unsigned long long)fr);
      outw[1] = (uint32_t)((unsigned long long)fr >> 32);
      outw[2] = meta;
      outw[3] = 0;
    } else {
      outw[0] = (uint32_t)fr;
      outw[1] = meta;
    }
  }
  uint32_t* data = outw + HDR / 4;
  if (bw > 0) {
    // lane g packs elements [32g, 32g+32) into words [g*bw, (g+1)*bw)
    const int groups = (n + 31) >> 5;
    for (int g = lane; g < groups; g += 32) {
      if (sizeof(E) == 8 && bw > 32) {
        pack32_u64(bw, reinterpret_cast<const unsigned long long*>(arr), g, n, (unsigned long long)fr, data, words);
      } else {
        uint32_t d[32];
#pragma unroll
        for (int j = 0; j < 32 / VE; ++j) {
          const Vec<E> v = ld_vec<E>(arr, (g << 5) + j * VE);
#pragma unroll
          for (int c = 0; c < VE; ++c) {
            const int e = (g << 5) + j * VE + c;
            d[j * VE + c] = e < n ? (uint32_t)(E)(v.e[c] - fr) : 0u;
          }
        }
        constexpr int MAXBW = sizeof(E) == 1 ? 8 : sizeof(E) == 2 ? 16 : 32;
        pack32_dispatch<MAXBW>(bw, d, data, g * (int)bw, words);
      }
    }
  }
  out_bytes = bytes;
  return true;
}

// Bit-packed serialisation when the caller already knows min / max (fused paths): header + blocked
// packing with no per-element bound checks - the caller has filled arr[n .. roundUp32(n)) with `mn`, so
// the padding elements pack as zero.  Same bytes as write_array(use_bp = true).  E is 2 or 4 bytes.
template <typename E>
__device__ __forceinline__ bool write_packed_known(const E* arr, int n, int32_t mn, int32_t mx, uint8_t* out,
                                                   size_t avail, uint32_t& out_bytes)
{
  static_assert(sizeof(E) == 2 || sizeof(E) == 4, "2- or 4-byte elements");
  constexpr int VE = Swz<E>::VE;
  const int lane = lane_id();
  const uint32_t bw = 32u - (uint32_t)__clz((int)((uint32_t)mx - (uint32_t)mn));
  const int words = (int)(((uint32_t)n * bw + 31u) >> 5);
  const uint32_t bytes = 8u + 4u * (uint32_t)words;
  if (bytes > avail) return false;
  uint32_t* outw = reinterpret_cast<uint32_t*>(out);
  const E fr = (E)mn;
  if (lane == 0) {
    outw[0] = (uint32_t)fr;
    outw[1] = (bw << 16) | (uint32_t)n;
  }
  const int g = lane;
  if (bw > 0 && (g << 5) < n) {
    uint32_t d[32];
#pragma unroll
    for (int j = 0; j < 32 / VE; ++j) {
      const Vec<E> v = ld_vec<E>(arr, (g << 5) + j * VE);
      if (sizeof(E) == 2) {
        // two 16-bit lanes per SIMD subtract
        const uint32_t fr2 = (uint32_t)fr * 0x10001u;
        const uint32_t w4[4] = {v.q.x, v.q.y, v.q.z, v.q.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t x = __vsub2(w4[c], fr2);
          d[j * 8 + 2 * c] = x & 0xffffu;
          d[j * 8 + 2 * c + 1] = x >> 16;
        }
      } else {
#pragma unroll
        for (int c = 0; c < VE; ++c) d[j * VE + c] = (uint32_t)(E)(v.e[c] - fr);
      }
    }
    pack32_dispatch<sizeof(E) == 2 ? 16 : 32>(bw, d, outw + 2, g * (int)bw, words);
  }
  out_bytes = bytes;
  return true;
}

__device__ __forceinline__ uint32_t get_bits32(const uint32_t* words, int nwords, uint32_t bitpos, uint32_t nb)
{
  // nb in [1,32]
  const int idx = (int)(bitpos >> 5);
  const uint32_t sh = bitpos & 31;
  const uint32_t lo = idx < nwords ? __ldg(words + idx) : 0u;
  const uint32_t hi = idx + 1 < nwords ? __ldg(words + idx + 1) : 0u;
  const uint32_t v = __funnelshift_r(lo, hi, sh);
  return nb == 32 ? v : (v & ((1u << nb) - 1u));
}

// Read an array stored by write_array from global `in` (in_bytes long, all of
// it inside the partition) into swizzled smem `arr`.  Returns false on a
// malformed array.  Follows block_read / block_bitunpack (:563-618,702-737).
template <typename E, int N>
__device__ bool read_array(const uint8_t* in, uint32_t in_bytes, bool use_bp, E* arr, int& n_out)
{
  using Z = Swz<E>;
  const int lane = lane_id();
  if (!use_bp) {
    const uint32_t n = in_bytes / sizeof(E);
    if (n > (uint32_t)N) return false;
    const E* ie = reinterpret_cast<const E*>(in);
    for (int i = lane; i < (int)n; i += 32) arr[Z::p(i)] = __ldg(ie + i);
    n_out = (int)n;
    return true;
  }
  constexpr int HDR = packed_header_bytes<E>();
  if (in_bytes < (uint32_t)HDR) return false;
  const uint32_t* inw = reinterpret_cast<const uint32_t*>(in);
  E fr;
  uint32_t meta;
  if (sizeof(E) == 8) {
    fr = (E)(((unsigned long long)__ldg(inw + 1) << 32) | __ldg(inw));
    meta = __ldg(inw + 2);
  } else {
    fr = (E)__ldg(inw);
    meta = __ldg(inw + 1);
  }
  const uint32_t bw = meta >> 16;
  const uint32_t n = meta & 0xffffu;
  if (n > (uint32_t)N || bw > 8 * sizeof(E)) return false;
  const int words = (int)((n * bw + 31u) >> 5);
  if ((uint32_t)HDR + 4u * (uint32_t)words > ((in_bytes + 3u) & ~3u)) return false;
  const uint32_t* data = inw + HDR / 4;
  if (bw == 0) {
    for (int i = lane; i < (int)n; i += 32) arr[Z::p(i)] = fr;
  } else if (sizeof(E) == 8 && bw > 32) {
    for (int i = lane; i < (int)n; i += 32) {
      const uint32_t bitpos = (uint32_t)i * bw;
      const unsigned long long lo = get_bits32(data, words, bitpos, 32);
      const unsigned long long hi = get_bits32(data, words, bitpos + 32, bw - 32);
      arr[Z::p(i)] = (E)((lo | (hi << 32)) + fr);
    }
  } else {
    for (int i = lane; i < (int)n; i += 32) arr[Z::p(i)] = (E)((E)get_bits32(data, words, (uint32_t)i * bw, bw) + fr);
  }
  n_out = (int)n;
  return true;
}

// ------------------------------------------------------------------ layers

// In-place run-length encoding of vals[0..n): returns the run count, leaves the
// run values in vals[0..m) and the run lengths in counts[0..m).
// (reference block_rle_compress :124-241)
// Pass 1 (striped, VE elements per lane): run-end flags, warp scan of the per-lane
// flag counts, compaction of (value, end index).  Pass 2: lengths from ends.
template <typename T, int N>
__device__ int rle_encode(T* vals, int n, uint16_t* counts)
{
  using Z = Swz<T>;
  using ZC = Swz<uint16_t>;
  constexpr int VE = Z::VE;
  constexpr int TILE = 32 * VE;
  const int lane = lane_id();
  int m = 0;
  for (int base = 0; base < n; base += TILE) {
    const int i0 = base + lane * VE;
    const Vec<T> v = ld_vec<T>(vals, i0);
    T nx = shfl_down_t(v.e[0], 1);
    const int after = base + TILE < N ? base + TILE : N - 1;
    const T nb = vals[Z::p(after)]; // first element of the next tile (broadcast read)
    if (lane == 31) nx = nb;
    uint32_t flags = 0;
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      const int i = i0 + j;
      const T next = j + 1 < VE ? v.e[j + 1 < VE ? j + 1 : j] : nx;
      const bool f = i < n && (i + 1 == n || next != v.e[j]);
      flags |= (f ? 1u : 0u) << j;
    }
    int total;
    int pos = m + warp_excl_sum<int>(__popc(flags), total);
    __syncwarp(); // every lane has read its tile before anyone compacts into it
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      if (flags & (1u << j)) {
        vals[Z::p(pos)] = v.e[j];
        counts[ZC::p(pos)] = (uint16_t)(i0 + j + 1);
        ++pos;
      }
    }
    m += total;
    __syncwarp();
  }
  // ends -> lengths
  uint32_t carry = 0;
  for (int t = 0; t < m; t += 32) {
    const int r = t + lane;
    const uint32_t e = r < m ? counts[ZC::p(r)] : 0u;
    uint32_t prev = __shfl_up_sync(kFull, e, 1);
    if (lane == 0) prev = carry;
    carry = __shfl_sync(kFull, e, 31);
    if (r < m) counts[ZC::p(r)] = (uint16_t)(e - prev);
  }
  __syncwarp();
  return m;
}

// rle_encode for the second layer of the fused R = 2 path (4-byte values): the same passes, which also leave the
// signed minimum / maximum of the run values and of the run lengths (what write_array would scan for again).
template <typename T, int N>
__device__ int rle_encode_minmax(T* vals, int n, uint16_t* counts, int32_t& vmn, int32_t& vmx, int32_t& cmn,
                                 int32_t& cmx)
{
  static_assert(sizeof(T) == 4, "4-byte values");
  using Z = Swz<T>;
  using ZC = Swz<uint16_t>;
  constexpr int VE = Z::VE;
  constexpr int TILE = 32 * VE;
  const int lane = lane_id();
  int32_t lvmn = 0x7fffffff, lvmx = (int32_t)0x80000000, lcmn = 0x7fffffff, lcmx = (int32_t)0x80000000;
  int m = 0;
  for (int base = 0; base < n; base += TILE) {
    const int i0 = base + lane * VE;
    const Vec<T> v = ld_vec<T>(vals, i0);
    T nx = shfl_down_t(v.e[0], 1);
    const int after = base + TILE < N ? base + TILE : N - 1;
    const T nb = vals[Z::p(after)]; // first element of the next tile (broadcast read)
    if (lane == 31) nx = nb;
    uint32_t flags = 0;
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      const int i = i0 + j;
      const T next = j + 1 < VE ? v.e[j + 1 < VE ? j + 1 : j] : nx;
      const bool f = i < n && (i + 1 == n || next != v.e[j]);
      flags |= (f ? 1u : 0u) << j;
    }
    int total;
    int pos = m + warp_excl_sum<int>(__popc(flags), total);
    __syncwarp(); // every lane has read its tile before anyone compacts into it
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      if (flags & (1u << j)) {
        vals[Z::p(pos)] = v.e[j];
        counts[ZC::p(pos)] = (uint16_t)(i0 + j + 1);
        lvmn = min(lvmn, (int32_t)v.e[j]);
        lvmx = max(lvmx, (int32_t)v.e[j]);
        ++pos;
      }
    }
    m += total;
    __syncwarp();
  }
  // ends -> lengths, 64 runs per step (the swizzle permutes inside 64-element blocks of counts)
  uint32_t carry = 0;
  for (int u = 0; u * 64 < m; ++u) {
    const int r0 = 64 * u + lane, r1 = r0 + 32;
    const uint32_t e0 = r0 < m ? (uint32_t)counts[ZC::p(r0)] : 0u;
    const uint32_t e1 = r1 < m ? (uint32_t)counts[ZC::p(r1)] : 0u;
    uint32_t q0 = __shfl_up_sync(kFull, e0, 1);
    uint32_t q1 = __shfl_up_sync(kFull, e1, 1);
    const uint32_t l0 = __shfl_sync(kFull, e0, 31);
    if (lane == 0) {
      q0 = carry;
      q1 = l0;
    }
    carry = __shfl_sync(kFull, r1 < m ? e1 : e0, 31);
    const int32_t c0 = (int32_t)(int16_t)(e0 - q0), c1 = (int32_t)(int16_t)(e1 - q1);
    if (r0 < m) {
      counts[ZC::p(r0)] = (uint16_t)c0;
      lcmn = min(lcmn, c0);
      lcmx = max(lcmx, c0);
    }
    if (r1 < m) {
      counts[ZC::p(r1)] = (uint16_t)c1;
      lcmn = min(lcmn, c1);
      lcmx = max(lcmx, c1);
    }
  }
  __syncwarp();
  vmn = __reduce_min_sync(kFull, lvmn);
  vmx = __reduce_max_sync(kFull, lvmx);
  cmn = __reduce_min_sync(kFull, lcmn);
  cmx = __reduce_max_sync(kFull, lcmx);
  return m;
}

// In-place adjacent differences: vals[i] = vals[i+1]-vals[i] for i < n-1.
// Returns the seed vals[0].  (reference block_delta_compress :317-328)
template <typename T, int N>
__device__ T delta_encode(T* vals, int n)
{
  using Z = Swz<T>;
  constexpr int VE = Z::VE;
  constexpr int TILE = 32 * VE;
  const int lane = lane_id();
  const T seed = vals[0];
  __syncwarp();
  for (int base = 0; base < n - 1; base += TILE) {
    const int i0 = base + lane * VE;
    Vec<T> v = ld_vec<T>(vals, i0);
    T nx = shfl_down_t(v.e[0], 1);
    const int after = base + TILE < N ? base + TILE : N - 1;
    const T nb = vals[Z::p(after)];
    if (lane == 31) nx = nb;
#pragma unroll
    for (int j = 0; j < VE; ++j) v.e[j] = (T)((j + 1 < VE ? v.e[j + 1 < VE ? j + 1 : j] : nx) - v.e[j]);
    __syncwarp(); // lane 31's look-ahead read precedes the next tile's update (already read); tile is private
    st_vec<T>(vals, i0, v);
  }
  __syncwarp();
  return seed;
}

// In-place inverse delta: vals[0..n) holds differences, result has n+1 elements
// vals[e] = seed + sum_{j<e} d[j].  (reference block_delta_decompress :343-377)
// Callers guarantee n + 1 <= N.
template <typename T, int N>
__device__ void delta_decode(T* vals, int n, T seed)
{
  using Z = Swz<T>;
  constexpr int VE = Z::VE;
  constexpr int TILE = 32 * VE;
  const int lane = lane_id();
  T carry = seed;
  for (int base = 0; base <= n; base += TILE) {
    const int i0 = base + lane * VE;
    Vec<T> v = ld_vec<T>(vals, i0);
    T sum = 0;
#pragma unroll
    for (int j = 0; j < VE; ++j) {
      const T d = i0 + j < n ? v.e[j] : (T)0;
      v.e[j] = sum; // exclusive prefix inside the lane
      sum += d;
    }
    T total;
    const T off = (T)(carry + warp_excl_sum<T>(sum, total));
#pragma unroll
    for (int j = 0; j < VE; ++j) v.e[j] += off;
    st_vec<T>(vals, i0, v); // positions > n receive the total: harmless, inside the array
    carry += total;
  }
  __syncwarp();
}

// Expand runs.  counts[] is overwritten with run starts, idx[] is scratch (N entries).
//   gout != nullptr : dst is the global output (outermost layer); at most `room` elements fit.
//   gout == nullptr : the expansion replaces vals[] in place (inner layers).
// Returns the expanded length, or -1 when it exceeds N / room.
// (reference block_rle_decompress :255-305 - there one thread loops over a whole run; here every
// output element finds its run by a max-scan over scattered run indices.)
template <typename T, int N>
__device__ int rle_expand(T* vals, uint16_t* counts, int n_runs, uint16_t* idx, T* gout, size_t room)
{
  using Z = Swz<T>;
  using ZC = Swz<uint16_t>;
  constexpr int W = (int)sizeof(T);
  const int lane = lane_id();
  // 1. exclusive scan of the counts, 8 per lane
  uint32_t carry = 0;
  bool zero_run = false;
  for (int base = 0; base < n_runs; base += 256) {
    const int i0 = base + lane * 8;
    Vec<uint16_t> v = ld_vec<uint16_t>(counts, i0);
    uint32_t c[8], s[8], sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      c[j] = i0 + j < n_runs ? (uint32_t)v.e[j] : 0u;
      zero_run |= i0 + j < n_runs && c[j] == 0;
      s[j] = sum;
      sum += c[j];
    }
    uint32_t total;
    const uint32_t off = carry + warp_excl_sum<uint32_t>(sum, total);
    if (carry + total > (uint32_t)N) return -1;
    // a zero-length run never owns an output position: park it on N (ignored below)
#pragma unroll
    for (int j = 0; j < 8; ++j) v.e[j] = (uint16_t)(c[j] ? off + s[j] : (uint32_t)N);
    st_vec<uint16_t>(counts, i0, v);
    carry += total;
  }
  const int total_out = (int)carry;
  if (gout != nullptr && (size_t)total_out > room) return -1;
  // An encoder never emits empty runs.  With them a run's start can precede its own slot, which
  // the in-place expansion of an inner layer cannot honour: such a stream is reported as corrupt.
  if (gout == nullptr && __any_sync(kFull, zero_run)) return -1;
  // 2. mark the first output position of every run with the run index
  {
    Vec<uint16_t> z;
    z.q = make_uint4(0, 0, 0, 0);
    for (int i0 = lane * 8; i0 < total_out; i0 += 256) st_vec<uint16_t>(idx, i0, z);
  }
  __syncwarp();
  for (int j = lane; j < n_runs; j += 32) {
    const int s = counts[ZC::p(j)];
    if (s < total_out) idx[ZC::p(s)] = (uint16_t)j;
  }
  __syncwarp();
  // 3. inclusive max-scan of the marks, 8 per lane; the outermost layer gathers and stores here
  const bool aligned8 = gout != nullptr && (reinterpret_cast<uintptr_t>(gout) & 7) == 0;
  uint32_t cmax = 0;
  for (int base = 0; base < total_out; base += 256) {
    const int i0 = base + lane * 8;
    Vec<uint16_t> v = ld_vec<uint16_t>(idx, i0);
    uint32_t r[8], lm = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      r[j] = i0 + j < total_out ? (uint32_t)v.e[j] : 0u;
      lm = r[j] > lm ? r[j] : lm;
    }
    const uint32_t incl = warp_incl_max<uint32_t>(lm);
    uint32_t run = __shfl_up_sync(kFull, incl, 1);
    if (lane == 0) run = 0;
    run = run > cmax ? run : cmax;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      run = r[j] > run ? r[j] : run;
      r[j] = run;
    }
    const uint32_t tile_max = __shfl_sync(kFull, incl, 31);
    cmax = tile_max > cmax ? tile_max : cmax;
    if (gout != nullptr) {
      if (i0 < total_out) {
        T o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = vals[Z::p((int)r[j])];
        if (aligned8 && i0 + 8 <= total_out) {
          // 8 consecutive elements = W 8-byte words
          unsigned long long* g8 = reinterpret_cast<unsigned long long*>(gout + i0);
#pragma unroll
          for (int q = 0; q < W; ++q) {
            unsigned long long word = 0;
#pragma unroll
            for (int b = 0; b < 8 / W; ++b) word |= (unsigned long long)o[q * (8 / W) + b] << (8 * W * b);
            g8[q] = word;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (i0 + j < total_out) gout[i0 + j] = o[j];
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v.e[j] = (uint16_t)r[j];
      st_vec<uint16_t>(idx, i0, v);
    }
  }
  __syncwarp();
  if (gout == nullptr) {
    // 4. in-place gather, highest tile first: a tile only reads run values at or below itself
    constexpr int VE = Z::VE;
    constexpr int TILE = 32 * VE;
    for (int base = (total_out - 1) / TILE * TILE; base >= 0 && total_out > 0; base -= TILE) {
      const int i0 = base + lane * VE;
      Vec<T> v;
      v.q = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int j = 0; j < VE; ++j)
        if (i0 + j < total_out) v.e[j] = vals[Z::p((int)idx[ZC::p(i0 + j)])];
      __syncwarp();
      if (i0 < total_out) st_vec<T>(vals, i0, v);
      __syncwarp();
    }
  }
  return total_out;
}

// ------------------------------------------------------------------ compress

template <typename T>
struct CompSmem {
  static constexpr int N = kChunkBytes / (int)sizeof(T);
  static constexpr int BYTES = kChunkBytes + N * 2 + kMaxMetaBytes; // values, run ends/lengths, chunk metadata
};

// global -> swizzled smem, 16 bytes per lane when the source allows it
template <typename T>
__device__ __forceinline__ void load_chunk(const T* src, int n, T* A)
{
  using Z = Swz<T>;
  const int lane = lane_id();
  int done = 0;
  if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
    const int vecs = n / Z::VE;
    for (int v = lane; v < vecs; v += 32) {
      Vec<T> x;
      x.q = ld_stream16(reinterpret_cast<const uint4*>(src) + v);
      st_vec<T>(A, v * Z::VE, x);
    }
    done = vecs * Z::VE;
  }
  for (int i = done + lane; i < n; i += 32) A[Z::p(i)] = __ldg(src + i);
}

// raw copy of a partition behind the 8-byte header (fallback path, :1020-1029)
template <typename T>
__device__ void raw_copy(const T* in, size_t n, uint8_t* out)
{
  const int lane = lane_id();
  const size_t bytes = n * sizeof(T);
  const uint8_t* src = reinterpret_cast<const uint8_t*>(in);
  uint8_t* dst = out + 8;
  if (!warp_copy_stream(dst, src, bytes)) {
    T* d = reinterpret_cast<T*>(dst);
    for (size_t i = lane; i < n; i += 32) d[i] = in[i];
  }
  const size_t padded = (bytes + 3) & ~(size_t)3;
  if (lane < (int)(padded - bytes)) dst[bytes + lane] = 0;
}

// Bytes the fused encoder emits for one full chunk (metadata + arrays, the advance of its cursor) when every element
// is a run of its own and - for R = 2 - no two neighbouring deltas are equal, so that every run-length array is
// all ones (8 bytes bit-packed) and the value array holds the N - 1 adjacent differences.  0 when the chunk is not
// of that kind.  Kept out of line: it runs once per partition at most, and inlined it costs the common path
// registers and instruction cache.
template <int R>
__device__ __noinline__ uint32_t distinct_chunk_bytes(const uint4* src, bool use_bp)
{
  constexpr uint32_t N = kChunkBytes / 4;
  const int lane = lane_id();
  uint4 v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = ld_stream16(src + lane + 32 * k);
  int32_t dmn = 0x7fffffff, dmx = (int32_t)0x80000000;
  bool ok = true;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t zx = (lane == 0 && k < 7) ? v[k < 7 ? k + 1 : k].x : v[k].x;
    const uint32_t nx = __shfl_sync(kFull, zx, (lane + 1) & 31);
    const bool last = (k == 7 && lane == 31); // element N - 1: no successor, no delta
    const int32_t d0 = (int32_t)(v[k].y - v[k].x), d1 = (int32_t)(v[k].z - v[k].y), d2 = (int32_t)(v[k].w - v[k].z);
    const int32_t d3 = (int32_t)(nx - v[k].w);
    ok = ok && d0 != 0 && d1 != 0 && d2 != 0 && (last || d3 != 0);
    dmn = min(dmn, min(d0, min(d1, d2)));
    dmx = max(dmx, max(d0, max(d1, d2)));
    if (!last) {
      dmn = min(dmn, d3);
      dmx = max(dmx, d3);
    }
    if constexpr (R == 2) {
      // the delta behind d3 is the first delta of the next lane's vector (of the next tile for lane 31)
      const uint32_t zy = (lane == 0 && k < 7) ? v[k < 7 ? k + 1 : k].y : v[k].y;
      const uint32_t ny = __shfl_sync(kFull, zy, (lane + 1) & 31);
      const int32_t d4 = (int32_t)(ny - nx);
      ok = ok && d0 != d1 && d1 != d2 && (last || (d2 != d3 && d3 != d4));
    }
  }
  if (!__all_sync(kFull, ok)) return 0;
  if (!use_bp) // raw arrays: R of 2-byte run lengths (N, then N - 1 of them), N - 1 values
    return 12u + 4u * R + 2u * N + (R == 2 ? ((2u * (N - 1u) + 3u) & ~3u) : 0u) + 4u * (N - 1u);
  dmn = __reduce_min_sync(kFull, dmn);
  dmx = __reduce_max_sync(kFull, dmx);
  const uint32_t vbw = 32u - (uint32_t)__clz((int)((uint32_t)dmx - (uint32_t)dmn));
  return 12u + 4u * R + 8u * R + 8u + 4u * (((N - 1u) * vbw + 31u) >> 5);
}

// Look-ahead of the fused encoder behind a first chunk that turned out all distinct: lower bound of the partition's
// compressed size - the chunk at s4 (the output cursor stands at chunk_start) and the more_full full chunks behind
// it, exact where distinct_chunk_bytes knows, the smallest possible chunk otherwise.  Gives up when the chunk at
// s4 is not of that kind.  True when the partition cannot fit (it is stored raw).
template <int R>
__device__ __noinline__ bool partition_cannot_fit(const uint4* s4, size_t chunk_start, size_t limit, bool use_bp,
                                                  int more_full, bool ragged_tail)
{
  constexpr uint32_t kMeta = 12 + 4 * R;
  const uint32_t floor_bytes = kMeta + (use_bp ? 8u * (R + 1) : 4u); // the smallest chunk: one run
  size_t total = chunk_start + (ragged_tail ? kMeta : 0u);
  for (int j = 0; j <= more_full && total <= limit; ++j) {
    const uint32_t exact = distinct_chunk_bytes<R>(s4 + 256 * j, use_bp);
    if (j == 0 && exact == 0) return false; // not of that kind: no look-ahead
    total += exact ? exact : floor_bytes;
  }
  return total > limit;
}

// Fused fast path for the headline option set (num_RLEs = 1, num_deltas = 1, 4-byte elements, full
// 4096-byte chunk, 16-byte aligned source).  Same bytes as the generic layer loop, different data
// movement: the chunk never sits in shared memory.  Each lane keeps 8 coalesced 16-byte vectors
// (elements 128k + 4*lane + j) in registers, run-end flags come from register compares plus one
// rotating shuffle per vector, the 8 per-tile flag counts are scanned as two byte-packed words, and
// only the compacted (run value, run end) pairs go to shared memory.  One striped pass then turns
// ends into lengths and values into deltas, in place, in the swizzled layout write_array packs from.
// Returns false when the chunk does not fit (caller falls back to the raw partition).
template <typename T, int R>
__device__ bool compress_chunk_rd1(const T* src, uint8_t* out, size_t& cur, size_t limit, bool use_bp, T* A,
                                   uint16_t* C, int& runs)
{
  static_assert(sizeof(T) == 4, "fast path is for 4-byte elements");
  constexpr int N = kChunkBytes / 4;
  const int lane = lane_id();
  const size_t chunk_start = cur;
  cur += 12 + 4 * R; // meta_size(4, R, 1)
  if (cur > limit) return false;

  uint4 v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = ld_stream16(reinterpret_cast<const uint4*>(src) + lane + 32 * k);

  // run-end flags, 4 per vector
  uint32_t fl[8];
  uint32_t p0 = 0, p1 = 0; // per-tile flag counts, one byte per tile
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t z = (lane == 0 && k < 7) ? v[k < 7 ? k + 1 : k].x : v[k].x;
    const uint32_t nx = __shfl_sync(kFull, z, (lane + 1) & 31);
    uint32_t f = (v[k].x != v[k].y ? 1u : 0u) | (v[k].y != v[k].z ? 2u : 0u) | (v[k].z != v[k].w ? 4u : 0u) |
                 (v[k].w != nx ? 8u : 0u);
    if (k == 7 && lane == 31) f |= 8u; // last element of the chunk ends a run
    fl[k] = f;
    const uint32_t c = (uint32_t)__popc(f);
    if (k < 4) p0 |= c << (8 * k);
    else p1 |= c << (8 * (k - 4));
  }
  // inclusive scan over lanes of the byte-packed counts (a tile holds <= 128 flags: no carry between bytes)
  uint32_t s0 = p0, s1 = p1;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o0 = __shfl_up_sync(kFull, s0, d);
    const uint32_t o1 = __shfl_up_sync(kFull, s1, d);
    if (lane >= d) {
      s0 += o0;
      s1 += o1;
    }
  }
  const uint32_t t0 = __shfl_sync(kFull, s0, 31), t1 = __shfl_sync(kFull, s1, 31);
  const uint32_t x0 = s0 - p0, x1 = s1 - p1; // exclusive, per tile
  const T seed = (T)__shfl_sync(kFull, v[0].x, 0);

  __syncwarp(); // the previous chunk's packing reads of A / C are complete
  uint32_t base = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t ex = ((k < 4 ? x0 : x1) >> (8 * (k & 3))) & 0xffu;
    const uint32_t tt = ((k < 4 ? t0 : t1) >> (8 * (k & 3))) & 0xffu;
    T* pv = A + base + ex;
    uint16_t* pe = C + base + ex;
    const uint32_t e0 = 128u * k + 4u * (uint32_t)lane + 1u; // run end = index + 1
    const uint32_t f = fl[k];
    if (f & 1u) { *pv++ = (T)v[k].x; *pe++ = (uint16_t)e0; }
    if (f & 2u) { *pv++ = (T)v[k].y; *pe++ = (uint16_t)(e0 + 1); }
    if (f & 4u) { *pv++ = (T)v[k].z; *pe++ = (uint16_t)(e0 + 2); }
    if (f & 8u) { *pv++ = (T)v[k].w; *pe++ = (uint16_t)(e0 + 3); }
    base += tt;
  }
  const int m = (int)base; // number of runs, >= 1
  runs = m;
  __syncwarp();

  // ends -> lengths, values -> deltas, rewritten in place in swizzled order: the swizzle permutes inside a
  // 32-element block of A and a 64-element block of C, so one step covers 64 runs (two per lane)
  uint32_t carry = 0;
  int32_t cmn = 0x7fffffff, cmx = (int32_t)0x80000000, vmn = 0x7fffffff, vmx = (int32_t)0x80000000;
  for (int u = 0; u * 64 < m; ++u) {
    const int r0 = 64 * u + lane, r1 = r0 + 32;
    const uint32_t e0 = r0 < m ? (uint32_t)C[r0] : 0u;
    const uint32_t e1 = r1 < m ? (uint32_t)C[r1] : 0u;
    const T a0 = r0 < m ? A[r0] : (T)0;
    const T n0 = r0 + 1 < m ? A[r0 + 1] : (T)0;
    const T a1 = r1 < m ? A[r1] : (T)0;
    const T n1 = r1 + 1 < m ? A[r1 + 1] : (T)0;
    uint32_t q0 = __shfl_up_sync(kFull, e0, 1);
    uint32_t q1 = __shfl_up_sync(kFull, e1, 1);
    const uint32_t l0 = __shfl_sync(kFull, e0, 31);
    if (lane == 0) {
      q0 = carry;
      q1 = l0;
    }
    carry = __shfl_sync(kFull, e1, 31);
    __syncwarp();
    const int sc = (u & 3) << 4; // ((r >> 5) & 7) << 3 for r0; r1 adds 8
    const int32_t c0 = (int32_t)(int16_t)(e0 - q0), c1 = (int32_t)(int16_t)(e1 - q1);
    const int32_t d0 = (int32_t)(n0 - a0), d1 = (int32_t)(n1 - a1);
    if (r0 < m) {
      C[r0 ^ sc] = (uint16_t)c0;
      cmn = min(cmn, c0);
      cmx = max(cmx, c0);
    }
    if (r1 < m) {
      C[r1 ^ (sc | 8)] = (uint16_t)c1;
      cmn = min(cmn, c1);
      cmx = max(cmx, c1);
    }
    const int sa = ((2 * u) & 7) << 2;
    if (r0 + 1 < m) {
      A[r0 ^ sa] = (T)d0;
      vmn = min(vmn, d0);
      vmx = max(vmx, d0);
    }
    if (r1 + 1 < m) {
      A[r1 ^ (sa | 4)] = (T)d1;
      vmn = min(vmn, d1);
      vmx = max(vmx, d1);
    }
  }
  __syncwarp();

  uint32_t ob_counts, ob_vals;
  if constexpr (R == 2) {
    // num_RLEs = 2 (the library's default option set): the first RLE + delta above are shared; the
    // deltas then take the generic second run-length layer.  Same calls, same order, same bytes as the
    // layer loop of compress_partition.
    if (use_bp) {
      cmn = __reduce_min_sync(kFull, cmn);
      cmx = __reduce_max_sync(kFull, cmx);
      const int rc = (m & ~31) + lane;
      if (rc >= m && (m & 31)) C[Swz<uint16_t>::p(rc)] = (uint16_t)cmn;
      __syncwarp();
      if (!write_packed_known<uint16_t>(C, m, cmn, cmx, out + cur, limit - cur, ob_counts)) return false;
    } else {
      if (!write_array<uint16_t, N>(C, m, out + cur, limit - cur, false, ob_counts)) return false;
    }
    cur += (ob_counts + 3u) & ~3u;
    __syncwarp();
    uint32_t ob_counts2;
    if (use_bp) {
      // the second layer's scan leaves min / max of both arrays: blocked packing without a second look
      int32_t v2mn, v2mx, c2mn, c2mx;
      const int n2 = rle_encode_minmax<T, N>(A, m - 1, C, v2mn, v2mx, c2mn, c2mx);
      if (n2 == 0) v2mn = v2mx = c2mn = c2mx = 0; // empty arrays: frame 0, width 0
      const int rp = (n2 & ~31) + lane; // padding up to the next multiple of 32 packs as zero
      if (rp >= n2 && (n2 & 31)) {
        C[Swz<uint16_t>::p(rp)] = (uint16_t)c2mn;
        A[Swz<T>::p(rp)] = (T)v2mn;
      }
      __syncwarp();
      if (!write_packed_known<uint16_t>(C, n2, c2mn, c2mx, out + cur, limit - cur, ob_counts2)) return false;
      cur += (ob_counts2 + 3u) & ~3u;
      if (!write_packed_known<T>(A, n2, v2mn, v2mx, out + cur, limit - cur, ob_vals)) return false;
      cur += (ob_vals + 3u) & ~3u;
    } else {
      const int n2 = rle_encode<T, N>(A, m - 1, C);
      if (!write_array<uint16_t, N>(C, n2, out + cur, limit - cur, false, ob_counts2)) return false;
      cur += (ob_counts2 + 3u) & ~3u;
      if (!write_array<T, N>(A, n2, out + cur, limit - cur, false, ob_vals)) return false;
      cur += (ob_vals + 3u) & ~3u;
    }
    if (lane == 0) {
      uint32_t* mw = reinterpret_cast<uint32_t*>(out + chunk_start);
      mw[0] = (uint32_t)(cur - chunk_start);
      mw[1] = ob_counts;
      mw[2] = ob_counts2;
      mw[3] = ob_vals;
      mw[4] = (uint32_t)seed;
    }
    __syncwarp();
    return true;
  }
  if (use_bp) {
    cmn = __reduce_min_sync(kFull, cmn);
    cmx = __reduce_max_sync(kFull, cmx);
    vmn = __reduce_min_sync(kFull, vmn);
    vmx = __reduce_max_sync(kFull, vmx);
    if (m == 1) vmn = vmx = 0; // empty delta array: frame 0, width 0
    // padding up to the next multiple of 32 packs as zero
    {
      const int rc = (m & ~31) + lane;
      if (rc >= m && (m & 31)) C[Swz<uint16_t>::p(rc)] = (uint16_t)cmn;
      const int nv = m - 1;
      const int rv = (nv & ~31) + lane;
      if (rv >= nv && (nv & 31)) A[Swz<T>::p(rv)] = (T)vmn;
    }
    __syncwarp();
    if (!write_packed_known<uint16_t>(C, m, cmn, cmx, out + cur, limit - cur, ob_counts)) return false;
    cur += (ob_counts + 3u) & ~3u;
    if (!write_packed_known<T>(A, m - 1, vmn, vmx, out + cur, limit - cur, ob_vals)) return false;
    cur += (ob_vals + 3u) & ~3u;
  } else {
    if (!write_array<uint16_t, N>(C, m, out + cur, limit - cur, false, ob_counts)) return false;
    cur += (ob_counts + 3u) & ~3u;
    if (!write_array<T, N>(A, m - 1, out + cur, limit - cur, false, ob_vals)) return false;
    cur += (ob_vals + 3u) & ~3u;
  }
  if (lane == 0) {
    uint32_t* mw = reinterpret_cast<uint32_t*>(out + chunk_start);
    mw[0] = (uint32_t)(cur - chunk_start);
    mw[1] = ob_counts;
    mw[2] = ob_vals;
    mw[3] = (uint32_t)seed;
  }
  __syncwarp();
  return true;
}

// One warp compresses one partition.  (reference do_cascaded_compression_kernel :761-1058)
template <typename T, bool kR2>
__device__ void compress_partition(const T* in, size_t in_bytes, uint8_t* out, size_t* comp_bytes_out, int type_id,
                                   int R, int D, int bp, uint8_t* smem)
{
  constexpr int N = CompSmem<T>::N;
  constexpr int W = (int)sizeof(T);
  T* A = reinterpret_cast<T*>(smem);
  uint16_t* C = reinterpret_cast<uint16_t*>(smem + kChunkBytes);
  uint32_t* meta = reinterpret_cast<uint32_t*>(smem + kChunkBytes + N * 2);
  const int lane = lane_id();

  if (in == nullptr || in_bytes == 0) {
    if (lane == 0) *comp_bytes_out = 0;
    return;
  }
  const size_t n_total = in_bytes / W;
  const size_t limit = 8 + ((in_bytes + 3) & ~(size_t)3);
  bool use_compression = !(R == 0 && D == 0 && bp == 0);
  const int msize = meta_size(W, R, D);
  const int delta_off = ((4 + 4 * (R + 1) + W - 1) / W) * W;
  const size_t n_chunks = (n_total + N - 1) / N;
  size_t cur = 8;
  const int layers = R > D ? R : D;

  const bool fast_r1d1 = W == 4 && (kR2 ? R == 2 : R == 1) && D == 1 && (reinterpret_cast<uintptr_t>(in) & 15) == 0;

  for (size_t c = 0; c < n_chunks && use_compression; ++c) {
    if constexpr (sizeof(T) == 4) {
      if (fast_r1d1 && n_total - c * N >= (size_t)N) {
        int runs = 0;
        const bool fits = compress_chunk_rd1<T, kR2 ? 2 : 1>(in + c * N, out, cur, limit, bp != 0, A, C, runs);
        if (!fits) use_compression = false;
        // Incompressible partitions (random words): the layer code finds out that the partition does not fit
        // only after compacting and packing chunk after chunk - a chunk of random words grows by 28 bytes, so
        // the verdict falls on the partition's last chunk.  When the first chunk turns out all distinct, the
        // sizes of the chunks behind it are added up before any of them is encoded.
        else if (c == 0 && runs == N && n_total >= 2 * (size_t)N &&
                 partition_cannot_fit<kR2 ? 2 : 1>(reinterpret_cast<const uint4*>(in + N), cur, limit, bp != 0,
                                                   (int)(n_total / N - 2), (n_total % N) != 0))
          use_compression = false;
        continue;
      }
    }
    const size_t chunk_start = cur;
    cur += msize;
    if (cur > limit) { use_compression = false; break; }
    int n = (int)(n_total - c * N < (size_t)N ? n_total - c * N : (size_t)N);
    load_chunk<T>(in + c * N, n, A);
    if (lane < kMaxMetaBytes / 4) meta[lane] = 0;
    __syncwarp();

    for (int l = 0; l < layers && use_compression; ++l) {
      if (l < R) {
        n = rle_encode<T, N>(A, n, C);
        uint32_t ob;
        if (!write_array<uint16_t, N>(C, n, out + cur, limit - cur, bp != 0, ob)) { use_compression = false; break; }
        cur += (ob + 3u) & ~3u;
        if (lane == 0) meta[l + 1] = ob;
        __syncwarp();
      }
      if (l < D) {
        if (n == 0) { use_compression = false; break; } // undefined in the reference; fall back
        const T seed = delta_encode<T, N>(A, n);
        if (lane == 0) *reinterpret_cast<T*>(reinterpret_cast<uint8_t*>(meta) + delta_off + l * W) = seed;
        n -= 1;
        __syncwarp();
      }
    }
    if (!use_compression) break;

    const size_t final_pos = (cur + W - 1) / W * W;
    if (final_pos > limit) { use_compression = false; break; }
    uint32_t ob;
    if (!write_array<T, N>(A, n, out + final_pos, limit - final_pos, bp != 0, ob)) { use_compression = false; break; }
    if (final_pos != cur && lane == 0) *reinterpret_cast<uint32_t*>(out + cur) = 0;
    cur = final_pos + ((ob + 3u) & ~3u);
    const size_t next = (cur + W - 1) / W * W;
    if (next != cur && lane == 0) *reinterpret_cast<uint32_t*>(out + cur) = 0;
    cur = next;
    if (lane == 0) {
      meta[0] = (uint32_t)(cur - chunk_start);
      meta[R + 1] = ob;
    }
    __syncwarp();
    if (lane < msize / 4) reinterpret_cast<uint32_t*>(out + chunk_start)[lane] = meta[lane];
    __syncwarp();
  }

  size_t comp_bytes;
  if (use_compression) {
    comp_bytes = cur;
  } else {
    raw_copy<T>(in, n_total, out);
    comp_bytes = 8 + ((n_total * W + 3) & ~(size_t)3);
  }
  if (lane == 0) {
    uint32_t* ow = reinterpret_cast<uint32_t*>(out);
    ow[0] = use_compression ? ((uint32_t)R | ((uint32_t)D << 8) | ((uint32_t)bp << 16) | ((uint32_t)type_id << 24))
                            : ((uint32_t)type_id << 24);
    ow[1] = (uint32_t)(n_total * W);
    *comp_bytes_out = comp_bytes;
  }
}

// kR2: the instantiation launched for num_RLEs = 2, num_deltas = 1 (the library's default option set).  It is a
// separate kernel because the second run-length layer, when it shares a function with the R = 1 fast path, costs
// that path 3 % (2570 -> 2498 GB/s on C3) through register allocation alone; the host knows R at launch.
template <typename T, bool kR2>
__global__ void __launch_bounds__(kWarpsPerCta * 32, B200_CASC_COMP_MINB)
compress_kernel(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs, size_t* out_bytes,
                size_t batch, int type_id, int R, int D, int bp, const uint8_t* opts_pp)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5;
  uint8_t* smem = smem_raw + (size_t)warp * CompSmem<T>::BYTES;
  const size_t stride = (size_t)gridDim.x * kWarpsPerCta;
  for (size_t p = (size_t)blockIdx.x * kWarpsPerCta + warp; p < batch; p += stride) {
    if (opts_pp) { // options chosen per partition (b200compBatchedCascadedSelectOptsAsync): 4 bytes each
      R = opts_pp[4 * p], D = opts_pp[4 * p + 1], bp = opts_pp[4 * p + 2];
    }
    compress_partition<T, kR2>(static_cast<const T*>(in_ptrs[p]), in_bytes[p], static_cast<uint8_t*>(out_ptrs[p]),
                          out_bytes + p, type_id, R, D, bp, smem);
    __syncwarp();
  }
}

// ------------------------------------------------------------------ decompress

template <typename T>
struct DecompSmem {
  static constexpr int N = kChunkBytes / (int)sizeof(T);
  static constexpr int BYTES = kChunkBytes + 2 * (N * 2); // values, run counts/starts, run-index marks
};

// Fused fast path of the decoder for num_RLEs = 1, num_deltas = 1, bit-packed, 4-byte elements, when the
// chunk expands to exactly one full 4096-byte chunk and the destination is 16-byte aligned.
// out[i] = seed + sum of the deltas of all runs that start at or before i, so instead of decoding the run
// values, marking run starts and gathering, the deltas are scattered (added) onto the run start positions of
// a zeroed shared-memory array and ONE prefix sum over the 1024 positions - held as 8 coalesced 16-byte
// vectors per lane - produces the output, which goes to global memory straight from registers.
// Empty runs add onto the next run's start, which is what skipping them means.
// Returns false when a precondition does not hold: nothing was written and the generic path decides.
template <typename T>
__device__ bool decompress_chunk_r1d1(const uint8_t* chunk_ptr, uint32_t avail_words, T* gout, T* O, uint32_t* S,
                                      uint32_t& chunk_bytes)
{
  static_assert(sizeof(T) == 4, "fast path is for 4-byte elements");
  constexpr uint32_t N = kChunkBytes / 4;
  constexpr uint32_t kStageWords = 1024; // the C + I regions of the warp's shared memory
  const int lane = lane_id();
  // stage the compressed chunk in shared memory: one round of independent coalesced loads covers the
  // first KiB (a C3 chunk is ~500 bytes); the unpack loops below then never wait on global memory
  const uint32_t* gw = reinterpret_cast<const uint32_t*>(chunk_ptr);
  const uint32_t first = avail_words < 256u ? avail_words : 256u;
  uint32_t w[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) w[k] = (uint32_t)(lane + 32 * k) < first ? __ldg(gw + lane + 32 * k) : 0u;
  __syncwarp(); // earlier readers of S are done
#pragma unroll
  for (int k = 0; k < 8; ++k) S[lane + 32 * k] = w[k];
#pragma unroll
  for (int k = 0; k < 8; ++k) reinterpret_cast<uint4*>(O)[lane + 32 * k] = make_uint4(0, 0, 0, 0);
  __syncwarp();
  chunk_bytes = S[0];
  const uint32_t cb = S[1], vb = S[2];
  const T seed = (T)S[3];
  const uint32_t cpad = (cb + 3u) & ~3u, vpad = (vb + 3u) & ~3u;
  if (cb < 8 || vb < 8 || cb > 4096 || vb > 4096) return false;
  const uint32_t need = (16u + cpad + vpad) >> 2; // words of metadata + both arrays
  if (need > avail_words || need + 1 > kStageWords) return false; // + 1: the unpack may touch one word more
  if (need > 256u) {
    for (uint32_t i = 256u + (uint32_t)lane; i < need; i += 32) S[i] = __ldg(gw + i);
    __syncwarp();
  }
  const uint32_t* cw = S + 4;
  const uint32_t* vw = S + 4 + (cpad >> 2);
  const uint32_t cfr = cw[0], cmeta = cw[1], vfr = vw[0], vmeta = vw[1];
  const uint32_t cbw = cmeta >> 16, nc = cmeta & 0xffffu, vbw = vmeta >> 16, nv = vmeta & 0xffffu;
  if (nc > N || cbw > 16 || nv + 1 > N || vbw > 32 || nc < nv + 1) return false;
  const uint32_t cwords = (nc * cbw + 31u) >> 5, vwords = (nv * vbw + 31u) >> 5;
  if (8u + 4u * cwords > cpad || 8u + 4u * vwords > vpad) return false;
  const int m = (int)nv + 1;
  const uint32_t cmask = cbw >= 32 ? ~0u : (1u << cbw) - 1u, vmask = vbw >= 32 ? ~0u : (1u << vbw) - 1u;

  // 64 runs per step (two per lane): the two length scans travel as one word of two 16-bit sums
  uint32_t carry = 0;
  for (int u = 0; u * 64 < m; ++u) {
    const int r0 = 64 * u + lane, r1 = r0 + 32;
    uint32_t len0 = 0, len1 = 0;
    T d0 = 0, d1 = 0;
    // a value that straddles two words has its second word inside the array; otherwise the extra word
    // read is masked off (it is still inside S)
    if (r0 < m) {
      const uint32_t cp = (uint32_t)r0 * cbw, vp = (uint32_t)(r0 > 0 ? r0 - 1 : 0) * vbw;
      const uint32_t cv = __funnelshift_r(cw[2 + (cp >> 5)], cw[3 + (cp >> 5)], cp & 31) & cmask;
      const uint32_t vv = __funnelshift_r(vw[2 + (vp >> 5)], vw[3 + (vp >> 5)], vp & 31) & vmask;
      len0 = (uint32_t)(uint16_t)(cv + cfr);
      d0 = r0 == 0 ? seed : (T)(vv + vfr);
    }
    if (r1 < m) {
      const uint32_t cp = (uint32_t)r1 * cbw, vp = (uint32_t)(r1 - 1) * vbw;
      const uint32_t cv = __funnelshift_r(cw[2 + (cp >> 5)], cw[3 + (cp >> 5)], cp & 31) & cmask;
      const uint32_t vv = __funnelshift_r(vw[2 + (vp >> 5)], vw[3 + (vp >> 5)], vp & 31) & vmask;
      len1 = (uint32_t)(uint16_t)(cv + cfr);
      d1 = (T)(vv + vfr);
    }
    if (__any_sync(kFull, len0 > N || len1 > N)) return false; // keeps each 16-bit sum below 2^16
    const uint32_t packed = len0 | (len1 << 16);
    uint32_t incl = packed;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
      const uint32_t o = __shfl_up_sync(kFull, incl, dd);
      if (lane >= dd) incl += o;
    }
    const uint32_t tot = __shfl_sync(kFull, incl, 31);
    const uint32_t ex = incl - packed;
    const uint32_t start0 = carry + (ex & 0xffffu);
    const uint32_t start1 = carry + (tot & 0xffffu) + (ex >> 16);
    carry += (tot & 0xffffu) + (tot >> 16);
    if (r0 < m && start0 < N) atomicAdd(reinterpret_cast<uint32_t*>(O) + start0, (uint32_t)d0);
    if (r1 < m && start1 < N) atomicAdd(reinterpret_cast<uint32_t*>(O) + start1, (uint32_t)d1);
  }
  if (carry != N) return false;
  __syncwarp();

  const uint4* O4 = reinterpret_cast<const uint4*>(O);
  uint4 q[8];
  uint32_t s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    q[k] = O4[lane + 32 * k];
    q[k].y += q[k].x;
    q[k].z += q[k].y;
    q[k].w += q[k].z;
    s[k] = q[k].w;
  }
#pragma unroll
  for (int dd = 1; dd < 32; dd <<= 1) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t o = __shfl_up_sync(kFull, s[k], dd);
      if (lane >= dd) s[k] += o;
    }
  }
  uint32_t tile_base = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t tot = __shfl_sync(kFull, s[k], 31);
    const uint32_t ex = s[k] - q[k].w + tile_base;
    q[k].x += ex;
    q[k].y += ex;
    q[k].z += ex;
    q[k].w += ex;
    tile_base += tot;
    st_stream16(reinterpret_cast<uint4*>(gout) + lane + 32 * k, q[k]);
  }
  __syncwarp();
  return true;
}

// Tail of the decoder shared by every option set that ends in (first RLE, one delta), e.g. the library default
// num_RLEs = 2: the generic layers have left the nv deltas of the first layer in A (swizzled shared memory);
// the bit-packed first-layer run lengths are read straight from the compressed chunk.  Same scheme as
// decompress_chunk_r1d1: deltas are added onto the run starts of a zeroed array O, one register prefix sum
// produces the 1024 outputs.  O must not alias A.  Returns false before writing anything to global memory
// when a precondition fails (the generic layers then continue from the untouched A).
template <typename T, bool kSmemCounts = false>
__device__ bool decompress_tail_d1(const uint8_t* counts_ptr, uint32_t cb, const T* A, int nv, T seed, T* gout,
                                   uint32_t* O)
{
  // kSmemCounts: the caller has staged the run lengths in shared memory (plain loads instead of __ldg)
  auto ldw = [](const uint32_t* p) { return kSmemCounts ? *p : __ldg(p); };
  static_assert(sizeof(T) == 4, "4-byte elements");
  constexpr uint32_t N = kChunkBytes / 4;
  const int lane = lane_id();
  if (cb < 8 || (cb & 3)) return false;
  const uint32_t* cw = reinterpret_cast<const uint32_t*>(counts_ptr);
  const uint32_t cfr = ldw(cw), cmeta = ldw(cw + 1);
  const uint32_t cbw = cmeta >> 16, nc = cmeta & 0xffffu;
  const int m = nv + 1;
  if (nc > N || cbw > 16 || (uint32_t)m > N || nc < (uint32_t)m) return false;
  if (8u + 4u * ((nc * cbw + 31u) >> 5) > cb) return false;
  const uint32_t cmask = (1u << cbw) - 1u;
#pragma unroll
  for (int k = 0; k < 8; ++k) reinterpret_cast<uint4*>(O)[lane + 32 * k] = make_uint4(0, 0, 0, 0);
  __syncwarp();
  uint32_t carry = 0;
  for (int u = 0; u * 64 < m; ++u) {
    const int r0 = 64 * u + lane, r1 = r0 + 32;
    uint32_t len0 = 0, len1 = 0;
    T d0 = 0, d1 = 0;
    if (r0 < m) {
      const uint32_t cp = (uint32_t)r0 * cbw;
      // the word after the last packed word is inside the chunk (the next array's header follows)
      const uint32_t cv = __funnelshift_r(ldw(cw + 2 + (cp >> 5)), ldw(cw + 3 + (cp >> 5)), cp & 31) & cmask;
      len0 = (uint32_t)(uint16_t)(cv + cfr);
      d0 = r0 == 0 ? seed : A[Swz<T>::p(r0 - 1)];
    }
    if (r1 < m) {
      const uint32_t cp = (uint32_t)r1 * cbw;
      const uint32_t cv = __funnelshift_r(ldw(cw + 2 + (cp >> 5)), ldw(cw + 3 + (cp >> 5)), cp & 31) & cmask;
      len1 = (uint32_t)(uint16_t)(cv + cfr);
      d1 = A[Swz<T>::p(r1 - 1)];
    }
    if (__any_sync(kFull, len0 > N || len1 > N)) return false;
    const uint32_t packed = len0 | (len1 << 16);
    uint32_t incl = packed;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
      const uint32_t o = __shfl_up_sync(kFull, incl, dd);
      if (lane >= dd) incl += o;
    }
    const uint32_t tot = __shfl_sync(kFull, incl, 31);
    const uint32_t ex = incl - packed;
    const uint32_t start0 = carry + (ex & 0xffffu);
    const uint32_t start1 = carry + (tot & 0xffffu) + (ex >> 16);
    carry += (tot & 0xffffu) + (tot >> 16);
    if (r0 < m && start0 < N) atomicAdd(O + start0, (uint32_t)d0);
    if (r1 < m && start1 < N) atomicAdd(O + start1, (uint32_t)d1);
  }
  if (carry != N) return false;
  __syncwarp();
  const uint4* O4 = reinterpret_cast<const uint4*>(O);
  uint4 q[8];
  uint32_t sc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    q[k] = O4[lane + 32 * k];
    q[k].y += q[k].x;
    q[k].z += q[k].y;
    q[k].w += q[k].z;
    sc[k] = q[k].w;
  }
#pragma unroll
  for (int dd = 1; dd < 32; dd <<= 1) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t o = __shfl_up_sync(kFull, sc[k], dd);
      if (lane >= dd) sc[k] += o;
    }
  }
  uint32_t tile_base = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t tot = __shfl_sync(kFull, sc[k], 31);
    const uint32_t ex = sc[k] - q[k].w + tile_base;
    q[k].x += ex;
    q[k].y += ex;
    q[k].z += ex;
    q[k].w += ex;
    tile_base += tot;
    st_stream16(reinterpret_cast<uint4*>(gout) + lane + 32 * k, q[k]);
  }
  __syncwarp();
  return true;
}

// Fused decoder for num_RLEs = 2, num_deltas = 1, bit-packed, 4-byte elements (the library's default option
// set), full chunk, 16-byte aligned destination.  The second run-length layer (values2, counts2) expands to the
// deltas of the first layer; like the last stage, that expansion is a scatter plus a prefix sum: run k of layer 2
// stores values2[k] - values2[k - 1] at its start position in a zeroed array E, and the running sum of E is the
// delta array.  Runs of layer 2 are never empty in an encoder's stream (starts are distinct: plain stores); a
// stream with one is left to the generic layers, which report it.
// Shared memory: the compressed chunk is staged in Y (the C + I regions) by one round of coalesced loads, E and
// then the deltas live in X (the A region); the first layer's run lengths are then moved behind the deltas so
// that Y can become the accumulator array of decompress_tail_d1, which finishes the chunk.
// Returns false before anything is written to global memory.
template <typename T>
__device__ bool decompress_chunk_r2d1(const uint8_t* chunk_ptr, uint32_t avail_words, T* gout, T* X, uint32_t* Y,
                                      uint32_t& chunk_bytes)
{
  static_assert(sizeof(T) == 4, "4-byte elements");
  constexpr uint32_t N = kChunkBytes / 4;
  constexpr uint32_t kStageWords = 1024;
  const int lane = lane_id();
  const uint32_t* gw = reinterpret_cast<const uint32_t*>(chunk_ptr);
  if (avail_words < 5) return false;
  const uint32_t first = avail_words < 256u ? avail_words : 256u;
  uint32_t w[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) w[k] = (uint32_t)(lane + 32 * k) < first ? __ldg(gw + lane + 32 * k) : 0u;
  __syncwarp(); // earlier readers of X / Y are done
#pragma unroll
  for (int k = 0; k < 8; ++k) Y[lane + 32 * k] = w[k];
#pragma unroll
  for (int k = 0; k < 8; ++k) reinterpret_cast<uint4*>(X)[lane + 32 * k] = make_uint4(0, 0, 0, 0);
  __syncwarp();
  chunk_bytes = Y[0];
  const uint32_t cb1 = Y[1], cb2 = Y[2], vb = Y[3];
  const T seed = (T)Y[4];
  if (cb1 < 8 || cb2 < 8 || vb < 8 || ((cb1 | cb2 | vb) & 3) || cb1 > 4096 || cb2 > 4096 || vb > 4096) return false;
  const uint32_t w1 = 5, w2 = w1 + (cb1 >> 2), wv = w2 + (cb2 >> 2);
  const uint32_t need = wv + (vb >> 2); // words of metadata + the three arrays
  if (need > avail_words || need + 1 > kStageWords) return false; // + 1: the unpack may touch one word more
  if (need > 256u) {
    for (uint32_t i = 256u + (uint32_t)lane; i < need; i += 32) Y[i] = __ldg(gw + i);
    __syncwarp();
  }
  const uint32_t* c2 = Y + w2;
  const uint32_t* vw = Y + wv;
  const uint32_t c2fr = c2[0], c2meta = c2[1], vfr = vw[0], vmeta = vw[1];
  const uint32_t c2bw = c2meta >> 16, n2c = c2meta & 0xffffu, vbw = vmeta >> 16, n2 = vmeta & 0xffffu;
  if (n2c > N || c2bw > 16 || n2 > N || vbw > 32 || n2c < n2) return false;
  const uint32_t c2words = (n2c * c2bw + 31u) >> 5, vwords = (n2 * vbw + 31u) >> 5;
  if (8u + 4u * c2words > cb2 || 8u + 4u * vwords > vb) return false;
  const uint32_t c2mask = (1u << c2bw) - 1u, vmask = vbw >= 32 ? ~0u : (1u << vbw) - 1u;

  uint32_t carry = 0, prev_val = 0;
  for (uint32_t u = 0; u * 64 < n2; ++u) {
    const uint32_t r0 = 64 * u + (uint32_t)lane, r1 = r0 + 32;
    uint32_t len0 = 0, len1 = 0, val0 = 0, val1 = 0;
    // a value that straddles two words has its second word inside its array; otherwise the extra word read is
    // masked off (it is still inside Y)
    if (r0 < n2) {
      const uint32_t cp = r0 * c2bw, vp = r0 * vbw;
      const uint32_t cv = __funnelshift_r(c2[2 + (cp >> 5)], c2[3 + (cp >> 5)], cp & 31) & c2mask;
      const uint32_t vv = __funnelshift_r(vw[2 + (vp >> 5)], vw[3 + (vp >> 5)], vp & 31) & vmask;
      len0 = (uint32_t)(uint16_t)(cv + c2fr);
      val0 = vv + vfr;
    }
    if (r1 < n2) {
      const uint32_t cp = r1 * c2bw, vp = r1 * vbw;
      const uint32_t cv = __funnelshift_r(c2[2 + (cp >> 5)], c2[3 + (cp >> 5)], cp & 31) & c2mask;
      const uint32_t vv = __funnelshift_r(vw[2 + (vp >> 5)], vw[3 + (vp >> 5)], vp & 31) & vmask;
      len1 = (uint32_t)(uint16_t)(cv + c2fr);
      val1 = vv + vfr;
    }
    // an empty run (or an absurd length) is not an encoder's stream
    if (__any_sync(kFull, (r0 < n2 && (len0 == 0 || len0 > N)) || (r1 < n2 && (len1 == 0 || len1 > N)))) return false;
    const uint32_t packed = len0 | (len1 << 16);
    uint32_t incl = packed;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
      const uint32_t o = __shfl_up_sync(kFull, incl, dd);
      if (lane >= dd) incl += o;
    }
    const uint32_t tot = __shfl_sync(kFull, incl, 31);
    const uint32_t ex = incl - packed;
    const uint32_t start0 = carry + (ex & 0xffffu);
    const uint32_t start1 = carry + (tot & 0xffffu) + (ex >> 16);
    carry += (tot & 0xffffu) + (tot >> 16);
    if (carry > N - 1) return false; // the deltas of N outputs are at most N - 1
    uint32_t p0 = __shfl_up_sync(kFull, val0, 1), p1 = __shfl_up_sync(kFull, val1, 1);
    const uint32_t l0 = __shfl_sync(kFull, val0, 31);
    if (lane == 0) {
      p0 = prev_val;
      p1 = l0;
    }
    prev_val = __shfl_sync(kFull, val1, 31);
    if (r0 < n2) X[Swz<T>::p((int)start0)] = (T)(val0 - p0);
    if (r1 < n2) X[Swz<T>::p((int)start1)] = (T)(val1 - p1);
  }
  const int nv = (int)carry; // number of first-layer deltas
  __syncwarp();
  // running sum of E, 128 elements per step, in place
  uint32_t run = 0;
  for (int base = 0; base < nv; base += 128) {
    const int i0 = base + lane * 4;
    Vec<T> v = ld_vec<T>(X, i0);
    v.e[1] += v.e[0];
    v.e[2] += v.e[1];
    v.e[3] += v.e[2];
    uint32_t incl = (uint32_t)v.e[3];
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
      const uint32_t o = __shfl_up_sync(kFull, incl, dd);
      if (lane >= dd) incl += o;
    }
    const uint32_t ex = incl - (uint32_t)v.e[3] + run;
    run += __shfl_sync(kFull, incl, 31);
#pragma unroll
    for (int j = 0; j < 4; ++j) v.e[j] += (T)ex;
    st_vec<T>(X, i0, v);
  }
  // the first layer's run lengths move behind the deltas (when there is room) and Y becomes the accumulator array
  const uint32_t c1w = (cb1 >> 2) + 1; // + the word an unpack may touch behind the array
  const uint32_t at = ((uint32_t)nv + 127u) & ~127u;
  if (at + c1w > N) {
    __syncwarp();
    return decompress_tail_d1<T, false>(reinterpret_cast<const uint8_t*>(gw + w1), cb1, X, nv, seed, gout, Y);
  }
  uint32_t* c1 = reinterpret_cast<uint32_t*>(X) + at;
  for (uint32_t i = (uint32_t)lane; i < c1w; i += 32) c1[i] = Y[w1 + i];
  __syncwarp();
  return decompress_tail_d1<T, true>(reinterpret_cast<const uint8_t*>(c1), cb1, X, nv, seed, gout, Y);
}

// One warp decompresses one partition.  (reference cascaded_decompression_fcn :1106-1435)
template <typename T, bool kTail>
__device__ void decompress_partition(const uint8_t* comp, size_t comp_bytes, T* out, size_t out_cap,
                                     size_t* actual_out, hipcompStatus_t* status_out, uint8_t* smem)
{
  constexpr int N = DecompSmem<T>::N;
  constexpr int W = (int)sizeof(T);
  using Z = Swz<T>;
  T* A = reinterpret_cast<T*>(smem);
  uint16_t* C = reinterpret_cast<uint16_t*>(smem + kChunkBytes);
  uint16_t* I = reinterpret_cast<uint16_t*>(smem + kChunkBytes + N * 2);
  const int lane = lane_id();

  bool ok = true;
  size_t done = 0;
  do {
    if (comp == nullptr || comp_bytes < 8) { ok = false; break; }
    const size_t end = comp_bytes & ~(size_t)3;
    const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t*>(comp));
    const int R = w0 & 0xff, D = (w0 >> 8) & 0xff, bp = (w0 >> 16) & 0xff;
    const size_t n_total = __ldg(reinterpret_cast<const uint32_t*>(comp) + 1) / W;
    if (out_cap < n_total * W) { ok = false; break; }
    if (R == 0 && D == 0 && bp == 0) {
      if (comp_bytes < 8 + n_total * W) { ok = false; break; }
      const uint8_t* src = comp + 8;
      uint8_t* dst = reinterpret_cast<uint8_t*>(out);
      const size_t bytes = n_total * W;
      if (!warp_copy_stream(dst, src, bytes)) {
        const T* s = reinterpret_cast<const T*>(src);
        for (size_t i = lane; i < n_total; i += 32) out[i] = s[i];
      }
      done = n_total;
      break;
    }
    const int msize = meta_size(W, R, D);
    if (R > 7 || msize > kMaxMetaBytes) { ok = false; break; }
    const int delta_off = ((4 + 4 * (R + 1) + W - 1) / W) * W;
    const int layers = R > D ? R : D;
    size_t chunk = 8;
    while (chunk < end) {
      if (chunk + msize > end) { ok = false; break; }
      if constexpr (sizeof(T) == 4) {
        if (R == 1 && D == 1 && bp != 0 && n_total - done >= (size_t)N &&
            (reinterpret_cast<uintptr_t>(out + done) & 15) == 0) {
          uint32_t cbytes;
          if (decompress_chunk_r1d1<T>(comp + chunk, (uint32_t)((end - chunk) >> 2), out + done, A,
                                       reinterpret_cast<uint32_t*>(C), cbytes)) {
            done += (size_t)N;
            if (cbytes < (uint32_t)msize) { ok = false; break; }
            chunk = (chunk + (cbytes & ~3u) + W - 1) / W * W;
            continue;
          }
        }
      }
      if constexpr (kTail && sizeof(T) == 4) {
        if (R == 2 && D == 1 && bp != 0 && n_total - done >= (size_t)N &&
            (reinterpret_cast<uintptr_t>(out + done) & 15) == 0) {
          uint32_t cbytes;
          if (decompress_chunk_r2d1<T>(comp + chunk, (uint32_t)((end - chunk) >> 2), out + done, A,
                                       reinterpret_cast<uint32_t*>(C), cbytes)) {
            done += (size_t)N;
            if (cbytes < (uint32_t)msize) { ok = false; break; }
            chunk = (chunk + (cbytes & ~3u) + W - 1) / W * W;
            continue;
          }
          __syncwarp();
        }
      }
      const uint32_t* mw = reinterpret_cast<const uint32_t*>(comp + chunk);
      const uint32_t chunk_bytes = __ldg(mw);
      uint32_t off[9];
      off[0] = 0;
#pragma unroll
      for (int r = 0; r < 7; ++r)
        if (r + 1 < R) off[r + 1] = (off[r] + __ldg(mw + r + 1) + 3u) & ~3u;
      if (R > 0) {
        const uint32_t a = W > 4 ? W : 4;
        off[R] = (off[R - 1] + __ldg(mw + R) + a - 1) / a * a;
      }
      const size_t arrays = chunk + msize;
      const uint32_t final_bytes = __ldg(mw + R + 1);
      const size_t final_pos = arrays + off[R];
      if (final_pos + ((final_bytes + 3u) & ~3u) > end) { ok = false; break; }
      int n;
      if (!read_array<T, N>(comp + final_pos, final_bytes, bp != 0, A, n)) { ok = false; break; }
      __syncwarp();
      bool wrote_global = false;
      // exact reverse of the encoder's layer order R1 D1 R2 D2 ... (see DESIGN.md on
      // the reference's interleaving rule for num_deltas > num_RLEs)
      for (int l = layers - 1; l >= 0 && ok; --l) {
        if constexpr (kTail && sizeof(T) == 4) {
          // last two layers (delta, first RLE) fused for full, aligned chunks
          if (l == 0 && D == 1 && R >= 1 && bp != 0 && n + 1 <= N && n_total - done >= (size_t)N &&
              (reinterpret_cast<uintptr_t>(out + done) & 15) == 0) {
            const uint32_t cb = __ldg(mw + 1);
            const size_t cpos = arrays + off[0];
            if (cpos + ((cb + 3u) & ~3u) <= end) {
              const T seed = *reinterpret_cast<const T*>(comp + chunk + delta_off);
              __syncwarp();
              if (decompress_tail_d1<T>(comp + cpos, cb, A, n, seed, out + done, reinterpret_cast<uint32_t*>(C))) {
                n = N;
                wrote_global = true;
                break;
              }
              __syncwarp();
            }
          }
        }
        if (l < D) {
          if (n + 1 > N) { ok = false; break; }
          const T seed = *reinterpret_cast<const T*>(comp + chunk + delta_off + l * W);
          delta_decode<T, N>(A, n, seed);
          n += 1;
        }
        if (l < R) {
          const uint32_t cb = __ldg(mw + l + 1);
          const size_t cpos = arrays + off[l];
          int nc;
          if (cpos + ((cb + 3u) & ~3u) > end || !read_array<uint16_t, N>(comp + cpos, cb, bp != 0, C, nc) || nc < n) {
            ok = false;
            break;
          }
          __syncwarp();
          // the expansion of the outermost layer goes straight to the output buffer
          const bool last = (l == 0);
          const int m = rle_expand<T, N>(A, C, n, I, last ? out + done : nullptr, last ? n_total - done : 0);
          if (m < 0) { ok = false; break; }
          wrote_global = last;
          n = m;
        }
      }
      if (!ok) break;
      if (done + (size_t)n > n_total) { ok = false; break; }
      if (!wrote_global) {
        T* o = out + done;
        int st = 0;
        if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) {
          const int vecs = n / Z::VE;
          for (int v = lane; v < vecs; v += 32) reinterpret_cast<uint4*>(o)[v] = ld_vec<T>(A, v * Z::VE).q;
          st = vecs * Z::VE;
        }
        for (int i = st + lane; i < n; i += 32) o[i] = A[Z::p(i)];
      }
      done += (size_t)n;
      if (chunk_bytes < (uint32_t)msize) { ok = false; break; }
      chunk = (chunk + (chunk_bytes & ~3u) + W - 1) / W * W;
      __syncwarp();
    }
    if (ok && done != n_total) ok = false;
  } while (false);

  if (lane == 0) {
    if (actual_out) *actual_out = ok ? done * W : 0;
    if (status_out) *status_out = ok ? hipcompSuccess : hipcompErrorCannotDecompress;
  }
}
// The element width of the whole batch comes from partition 0's type byte, as in
// the reference (CascadedBatch.hip:156-168).  One launch per width; the three
// that do not match return after one load.  For 4-byte elements there are two instantiations, selected the same
// way from partition 0's header: kTail (num_RLEs >= 2, num_deltas = 1 - the library's default option set) fuses
// the last two layers behind the generic ones; keeping it out of the other instantiation keeps that one's
// register allocation (80 regs, no spills) for the headline R = 1 fast path.  Both decode any partition.
template <typename T, bool kTail>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
decompress_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, void* const* out_ptrs,
                  const size_t* out_caps, size_t* actual, hipcompStatus_t* statuses, size_t batch, int known_type,
                  int known_tail)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  // known_type >= 0: the caller (a manager, the staging pipeline) knows the batch's options and launched only
  // this instantiation; else the C ABI's rule: partition 0's header decides for the batch
  const uint8_t* p0 = known_type >= 0 ? nullptr : static_cast<const uint8_t*>(comp_ptrs[0]);
  const int type_id = known_type >= 0 ? known_type : (p0 ? p0[3] : -1);
  int width;
  switch (type_id) {
  case HIPCOMP_TYPE_CHAR: case HIPCOMP_TYPE_UCHAR: width = 1; break;
  case HIPCOMP_TYPE_SHORT: case HIPCOMP_TYPE_USHORT: width = 2; break;
  case HIPCOMP_TYPE_INT: case HIPCOMP_TYPE_UINT: width = 4; break;
  case HIPCOMP_TYPE_LONGLONG: case HIPCOMP_TYPE_ULONGLONG: width = 8; break;
  default: width = 1; break; // unknown type byte: the 1-byte instantiation reports the errors
  }
  if (width != (int)sizeof(T)) return;
  if (sizeof(T) == 4) {
    // num_RLEs, num_deltas, use_bp of partition 0
    const bool tail_batch = known_type >= 0 ? known_tail != 0 : (p0[0] >= 2 && p0[1] == 1 && p0[2] != 0);
    if (tail_batch != kTail) return;
  }
  const int warp = threadIdx.x >> 5;
  uint8_t* smem = smem_raw + (size_t)warp * DecompSmem<T>::BYTES;
  const size_t stride = (size_t)gridDim.x * kWarpsPerCta;
  for (size_t p = (size_t)blockIdx.x * kWarpsPerCta + warp; p < batch; p += stride) {
    const bool known = type_id >= 0 && type_id <= 7;
    decompress_partition<T, kTail>(known ? static_cast<const uint8_t*>(comp_ptrs[p]) : nullptr, comp_bytes[p],
                            static_cast<T*>(out_ptrs[p]), out_caps[p], actual ? actual + p : nullptr,
                            statuses ? statuses + p : nullptr, smem);
    __syncwarp();
  }
}

__global__ void size_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, size_t* out, size_t batch)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < batch; i += (size_t)gridDim.x * blockDim.x) {
    size_t v = 0;
    if (comp_bytes[i] >= 8 && comp_ptrs[i] != nullptr) v = static_cast<const uint32_t*>(comp_ptrs[i])[1];
    out[i] = v;
  }
}

// ------------------------------------------------------------------ option selector
// The reference's cascaded.h:93-99 / CHANGELOG.md:113-116 dropped nvCOMP's selector; the format still carries the
// option set in every partition header, so a per-partition choice decodes everywhere.  One warp samples the first
// kSelSample elements of a partition and estimates, from six statistics, the packed size of the option sets
// {raw, bp, delta+bp, rle+bp, rle+delta+bp}; the smallest wins, ties go to the fewer layers.  Pure integer
// arithmetic: tests/test_gpu_cascaded_select.py restates it in numpy and demands the same choices.
constexpr int kSelSample = 1024;

__device__ __forceinline__ uint32_t range_bits(unsigned long long span)
{
  return span == 0 ? 0u : 64u - (uint32_t)__clzll((long long)span);
}

template <typename T>
__global__ void __launch_bounds__(128)
select_kernel(const void* const* in_ptrs, const size_t* in_bytes, size_t batch, uint8_t* opts_pp)
{
  using S = typename std::make_signed<T>::type;
  const int lane = lane_id();
  const size_t stride = (size_t)gridDim.x * 4;
  for (size_t p = (size_t)blockIdx.x * 4 + (threadIdx.x >> 5); p < batch; p += stride) {
    const T* v = static_cast<const T*>(in_ptrs[p]);
    const size_t n_total = in_bytes[p] / sizeof(T);
    const int m = (int)(n_total < (size_t)kSelSample ? n_total : (size_t)kSelSample);
    unsigned long long vmin = ~0ull, vmax = 0;
    long long dmin = LLONG_MAX, dmax = LLONG_MIN, rmin = LLONG_MAX, rmax = LLONG_MIN; // all deltas / non-zero deltas
    uint32_t nruns = 0, maxrun = 0; // maxrun: lane-local maximum of the completed runs, reduced below
    uint32_t open_len = 0;          // length so far of the run that reaches the current position (uniform)
    for (int i0 = 0; i0 < m; i0 += 32) {
      const int i = i0 + lane;
      const bool in = i < m;
      const T x = in ? v[i] : (T)0;
      const T prev = (in && i > 0) ? v[i - 1] : x;
      if (in) {
        vmin = min(vmin, (unsigned long long)x);
        vmax = max(vmax, (unsigned long long)x);
      }
      const bool start = in && (i == 0 || x != prev);
      if (in && i > 0) {
        const long long d = (long long)(S)(T)(x - prev);
        dmin = min(dmin, d), dmax = max(dmax, d);
        if (start) rmin = min(rmin, d), rmax = max(rmax, d);
      }
      const unsigned sm = __ballot_sync(kFull, start);
      const int cnt = min(32, m - i0);
      nruns += (uint32_t)__popc(sm);
      if (start) {
        // the run this lane starts ends at the next start of the group (runs that leave the group stay open)
        const unsigned above = sm & (0xFFFFFFFEu << lane);
        if (above) maxrun = max(maxrun, (uint32_t)(__ffs((int)above) - 1 - lane));
        // the first start of the group closes the run that was open
        if ((sm & ((1u << lane) - 1u)) == 0u) maxrun = max(maxrun, open_len + (uint32_t)lane);
      }
      open_len = sm ? (uint32_t)(cnt - (31 - __clz((int)sm))) : open_len + (uint32_t)cnt;
    }
    maxrun = max(maxrun, open_len);
    maxrun = __reduce_max_sync(kFull, maxrun);
    // 64-bit reductions by shuffle
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      vmin = min(vmin, (unsigned long long)__shfl_xor_sync(kFull, (long long)vmin, o));
      vmax = max(vmax, (unsigned long long)__shfl_xor_sync(kFull, (long long)vmax, o));
      dmin = min(dmin, __shfl_xor_sync(kFull, dmin, o));
      dmax = max(dmax, __shfl_xor_sync(kFull, dmax, o));
      rmin = min(rmin, __shfl_xor_sync(kFull, rmin, o));
      rmax = max(rmax, __shfl_xor_sync(kFull, rmax, o));
    }
    if (lane == 0) {
      uint8_t R = 0, D = 0, bp = 0;
      if (m >= 2) {
        const unsigned long long W8 = 8ull * sizeof(T);
        const unsigned long long vb = range_bits(vmax - vmin);
        const unsigned long long db = range_bits((unsigned long long)(dmax - dmin));
        const unsigned long long rb = rmin > rmax ? 0ull : range_bits((unsigned long long)(rmax - rmin));
        const unsigned long long lb = range_bits(maxrun > 0 ? maxrun - 1u : 0u);
        const unsigned long long mm = (unsigned long long)m, nr = nruns;
        unsigned long long best = mm * W8; // raw
        const unsigned long long s_bp = mm * vb + 64, s_d = (mm - 1) * db + W8 + 64, s_r = nr * (vb + lb) + 128,
                                 s_rd = (nr > 0 ? nr - 1 : 0) * rb + nr * lb + W8 + 128;
        if (s_bp < best) best = s_bp, R = 0, D = 0, bp = 1;
        if (s_d < best) best = s_d, R = 0, D = 1, bp = 1;
        if (s_r < best) best = s_r, R = 1, D = 0, bp = 1;
        if (s_rd < best) best = s_rd, R = 1, D = 1, bp = 1;
      }
      opts_pp[4 * p] = R, opts_pp[4 * p + 1] = D, opts_pp[4 * p + 2] = bp, opts_pp[4 * p + 3] = 0;
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------ launch

static int grid_for(size_t batch, int ctas_per_sm)
{
  const size_t need = (batch + kWarpsPerCta - 1) / kWarpsPerCta;
  const size_t cap = (size_t)sm_count() * ctas_per_sm;
  return (int)(need < cap ? (need ? need : 1) : cap);
}

template <typename T>
static hipcompStatus_t launch_compress(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs,
                                       size_t* out_bytes, size_t batch, hipcompBatchedCascadedOpts_t o,
                                       cudaStream_t stream, const uint8_t* opts_pp = nullptr)
{
  constexpr int smem = CompSmem<T>::BYTES * kWarpsPerCta;
  const int per_sm = (227 * 1024) / (smem + 1024);
  const int grid = grid_for(batch, per_sm < 16 ? per_sm : 16);
  if (sizeof(T) == 4 && o.num_RLEs == 2 && o.num_deltas == 1 && opts_pp == nullptr) {
    allow_smem(compress_kernel<T, true>, smem); // per device, so not cached in a static
    compress_kernel<T, true><<<grid, kWarpsPerCta * 32, smem, stream>>>(
        in_ptrs, in_bytes, out_ptrs, out_bytes, batch, (int)o.type, o.num_RLEs, o.num_deltas, o.use_bp, nullptr);
  } else {
    allow_smem(compress_kernel<T, false>, smem);
    compress_kernel<T, false><<<grid, kWarpsPerCta * 32, smem, stream>>>(
        in_ptrs, in_bytes, out_ptrs, out_bytes, batch, (int)o.type, o.num_RLEs, o.num_deltas, o.use_bp, opts_pp);
  }
  return check_launch("hipcompBatchedCascadedCompressAsync()");
}

template <typename T>
static hipcompStatus_t launch_decompress(const void* const* comp_ptrs, const size_t* comp_bytes,
                                         void* const* out_ptrs, const size_t* out_caps, size_t* actual,
                                         hipcompStatus_t* statuses, size_t batch, cudaStream_t stream)
{
  constexpr int smem = DecompSmem<T>::BYTES * kWarpsPerCta;
  allow_smem(decompress_kernel<T, false>, smem);
  const int per_sm = (227 * 1024) / (smem + 1024);
  const int grid = grid_for(batch, per_sm < 16 ? per_sm : 16);
  decompress_kernel<T, false><<<grid, kWarpsPerCta * 32, smem, stream>>>(comp_ptrs, comp_bytes, out_ptrs, out_caps,
                                                                        actual, statuses, batch, -1, 0);
  if (sizeof(T) == 4) {
    allow_smem(decompress_kernel<T, true>, smem);
    decompress_kernel<T, true><<<grid, kWarpsPerCta * 32, smem, stream>>>(comp_ptrs, comp_bytes, out_ptrs, out_caps,
                                                                         actual, statuses, batch, -1, 0);
  }
  return check_launch("hipcompBatchedCascadedDecompressAsync()");
}

// One launch: the instantiation the options name (callers that know the batch's options: managers, staging).
template <typename T, bool kTail>
static hipcompStatus_t launch_decompress_known(const void* const* comp_ptrs, const size_t* comp_bytes,
                                               void* const* out_ptrs, const size_t* out_caps, size_t* actual,
                                               hipcompStatus_t* statuses, size_t batch, int type, cudaStream_t stream)
{
  constexpr int smem = DecompSmem<T>::BYTES * kWarpsPerCta;
  allow_smem(decompress_kernel<T, kTail>, smem);
  const int per_sm = (227 * 1024) / (smem + 1024);
  const int grid = grid_for(batch, per_sm < 16 ? per_sm : 16);
  decompress_kernel<T, kTail><<<grid, kWarpsPerCta * 32, smem, stream>>>(comp_ptrs, comp_bytes, out_ptrs, out_caps, actual,
                                                                        statuses, batch, type, kTail ? 1 : 0);
  return check_launch("cascaded::decompress_known()");
}

hipcompStatus_t decompress_known(const hipcompBatchedCascadedOpts_t& o, const void* const* comp_ptrs,
                                 const size_t* comp_bytes, const size_t* out_caps, size_t* actual, size_t batch,
                                 void* const* out_ptrs, hipcompStatus_t* statuses, cudaStream_t stream)
{
  if (batch == 0) return hipcompSuccess;
  const int t = (int)o.type;
  switch (t) {
  case HIPCOMP_TYPE_CHAR: case HIPCOMP_TYPE_UCHAR:
    return launch_decompress_known<uint8_t, false>(comp_ptrs, comp_bytes, out_ptrs, out_caps, actual, statuses, batch, t, stream);
  case HIPCOMP_TYPE_SHORT: case HIPCOMP_TYPE_USHORT:
    return launch_decompress_known<uint16_t, false>(comp_ptrs, comp_bytes, out_ptrs, out_caps, actual, statuses, batch, t, stream);
  case HIPCOMP_TYPE_INT: case HIPCOMP_TYPE_UINT:
    if (o.num_RLEs >= 2 && o.num_deltas == 1 && o.use_bp != 0)
      return launch_decompress_known<uint32_t, true>(comp_ptrs, comp_bytes, out_ptrs, out_caps, actual, statuses, batch, t, stream);
    return launch_decompress_known<uint32_t, false>(comp_ptrs, comp_bytes, out_ptrs, out_caps, actual, statuses, batch, t, stream);
  case HIPCOMP_TYPE_LONGLONG: case HIPCOMP_TYPE_ULONGLONG:
    return launch_decompress_known<unsigned long long, false>(comp_ptrs, comp_bytes, out_ptrs, out_caps, actual, statuses, batch, t, stream);
  default:
    return fail(hipcompErrorInvalidValue, "cascaded::decompress_known()", "unsupported element type");
  }
}

} // namespace cascaded
} // namespace b200comp

using namespace b200comp;

extern "C" {

hipcompStatus_t hipcompBatchedCascadedCompressGetTempSize(size_t, size_t, hipcompBatchedCascadedOpts_t,
                                                          size_t* temp_bytes)
{
  B200_CHECK_NOT_NULL(temp_bytes, "hipcompBatchedCascadedCompressGetTempSize()");
  *temp_bytes = 0;
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedCascadedCompressGetMaxOutputChunkSize(size_t max_uncompressed_chunk_bytes,
                                                                    hipcompBatchedCascadedOpts_t,
                                                                    size_t* max_compressed_bytes)
{
  B200_CHECK_NOT_NULL(max_compressed_bytes, "hipcompBatchedCascadedCompressGetMaxOutputChunkSize()");
  *max_compressed_bytes = round_up_sz(max_uncompressed_chunk_bytes, 4) + 8;
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedCascadedCompressAsync(const void* const* device_uncompressed_ptrs,
                                                    const size_t* device_uncompressed_bytes, size_t, size_t batch_size,
                                                    void*, size_t, void* const* device_compressed_ptrs,
                                                    size_t* device_compressed_bytes,
                                                    const hipcompBatchedCascadedOpts_t o, hipStream_t stream)
{
  static const char* fn = "hipcompBatchedCascadedCompressAsync()";
  B200_TRACE(fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_bytes, fn);
  if (o.num_RLEs < 0 || o.num_RLEs > 7 || o.num_deltas < 0 || o.num_deltas > 255 || o.use_bp < 0 || o.use_bp > 1)
    return fail(hipcompErrorInvalidValue, fn, "unsupported cascaded options (num_RLEs in [0,7], use_bp in {0,1})");
  int w;
  switch (o.type) {
  case HIPCOMP_TYPE_CHAR: case HIPCOMP_TYPE_UCHAR: w = 1; break;
  case HIPCOMP_TYPE_SHORT: case HIPCOMP_TYPE_USHORT: w = 2; break;
  case HIPCOMP_TYPE_INT: case HIPCOMP_TYPE_UINT: w = 4; break;
  case HIPCOMP_TYPE_LONGLONG: case HIPCOMP_TYPE_ULONGLONG: w = 8; break;
  default: return fail(hipcompErrorInvalidValue, fn, "Unsupported input data type");
  }
  if (cascaded::meta_size(w, o.num_RLEs, o.num_deltas) > cascaded::kMaxMetaBytes)
    return fail(hipcompErrorInvalidValue, fn, "chunk metadata for these options exceeds 64 bytes");
  switch (w) {
  case 1: return cascaded::launch_compress<uint8_t>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream);
  case 2: return cascaded::launch_compress<uint16_t>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream);
  case 4: return cascaded::launch_compress<uint32_t>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream);
  default: return cascaded::launch_compress<unsigned long long>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream);
  }
}

hipcompStatus_t hipcompBatchedCascadedDecompressGetTempSize(size_t, size_t, size_t* temp_bytes)
{
  B200_CHECK_NOT_NULL(temp_bytes, "hipcompBatchedCascadedDecompressGetTempSize()");
  *temp_bytes = 0;
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedCascadedDecompressAsync(const void* const* device_compressed_ptrs,
                                                      const size_t* device_compressed_bytes,
                                                      const size_t* device_uncompressed_bytes,
                                                      size_t* device_actual_uncompressed_bytes, size_t batch_size,
                                                      void* const, size_t, void* const* device_uncompressed_ptrs,
                                                      hipcompStatus_t* device_statuses, hipStream_t stream)
{
  static const char* fn = "hipcompBatchedCascadedDecompressAsync()";
  B200_TRACE(fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_ptrs, fn);
  hipcompStatus_t s;
  s = cascaded::launch_decompress<uint32_t>(device_compressed_ptrs, device_compressed_bytes, device_uncompressed_ptrs, device_uncompressed_bytes, device_actual_uncompressed_bytes, device_statuses, batch_size, stream);
  if (s != hipcompSuccess) return s;
  s = cascaded::launch_decompress<uint8_t>(device_compressed_ptrs, device_compressed_bytes, device_uncompressed_ptrs, device_uncompressed_bytes, device_actual_uncompressed_bytes, device_statuses, batch_size, stream);
  if (s != hipcompSuccess) return s;
  s = cascaded::launch_decompress<uint16_t>(device_compressed_ptrs, device_compressed_bytes, device_uncompressed_ptrs, device_uncompressed_bytes, device_actual_uncompressed_bytes, device_statuses, batch_size, stream);
  if (s != hipcompSuccess) return s;
  return cascaded::launch_decompress<unsigned long long>(device_compressed_ptrs, device_compressed_bytes, device_uncompressed_ptrs, device_uncompressed_bytes, device_actual_uncompressed_bytes, device_statuses, batch_size, stream);
}

hipcompStatus_t hipcompBatchedCascadedGetDecompressSizeAsync(const void* const* device_compressed_ptrs,
                                                             const size_t* device_compressed_bytes,
                                                             size_t* device_uncompressed_bytes, size_t batch_size,
                                                             hipStream_t stream)
{
  static const char* fn = "hipcompBatchedCascadedGetDecompressSizeAsync()";
  B200_TRACE(fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  const size_t need = (batch_size + 255) / 256;
  const size_t cap = (size_t)sm_count() * 8;
  cascaded::size_kernel<<<(int)(need < cap ? need : cap), 256, 0, stream>>>(device_compressed_ptrs, device_compressed_bytes,
                                                                            device_uncompressed_bytes, batch_size);
  return check_launch(fn);
}


/* ---- extension: per-partition option selection (include/hipcomp/b200_ext.h) */
hipcompStatus_t b200compBatchedCascadedSelectOptsAsync(const void* const* device_uncompressed_ptrs,
                                                       const size_t* device_uncompressed_bytes, size_t batch_size,
                                                       hipcompType_t type, uint8_t* device_opts, hipStream_t stream)
{
  static const char* fn = "b200compBatchedCascadedSelectOptsAsync()";
  B200_TRACE(fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_opts, fn);
  const size_t need = (batch_size + 3) / 4;
  const size_t cap = (size_t)sm_count() * 16;
  const int grid = (int)(need < cap ? need : cap);
  switch (type) {
  case HIPCOMP_TYPE_CHAR: case HIPCOMP_TYPE_UCHAR:
    cascaded::select_kernel<uint8_t><<<grid, 128, 0, stream>>>(device_uncompressed_ptrs, device_uncompressed_bytes, batch_size, device_opts);
    break;
  case HIPCOMP_TYPE_SHORT: case HIPCOMP_TYPE_USHORT:
    cascaded::select_kernel<uint16_t><<<grid, 128, 0, stream>>>(device_uncompressed_ptrs, device_uncompressed_bytes, batch_size, device_opts);
    break;
  case HIPCOMP_TYPE_INT: case HIPCOMP_TYPE_UINT:
    cascaded::select_kernel<uint32_t><<<grid, 128, 0, stream>>>(device_uncompressed_ptrs, device_uncompressed_bytes, batch_size, device_opts);
    break;
  case HIPCOMP_TYPE_LONGLONG: case HIPCOMP_TYPE_ULONGLONG:
    cascaded::select_kernel<unsigned long long><<<grid, 128, 0, stream>>>(device_uncompressed_ptrs, device_uncompressed_bytes, batch_size, device_opts);
    break;
  default:
    return fail(hipcompErrorInvalidValue, fn, "unsupported element type");
  }
  return check_launch(fn);
}

hipcompStatus_t b200compBatchedCascadedCompressWithOptsAsync(const void* const* device_uncompressed_ptrs,
                                                             const size_t* device_uncompressed_bytes, size_t batch_size,
                                                             void* const* device_compressed_ptrs,
                                                             size_t* device_compressed_bytes, hipcompType_t type,
                                                             const uint8_t* device_opts, hipStream_t stream)
{
  static const char* fn = "b200compBatchedCascadedCompressWithOptsAsync()";
  B200_TRACE(fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_opts, fn);
  hipcompBatchedCascadedOpts_t o = hipcompBatchedCascadedDefaultOpts;
  o.type = type;
  o.num_RLEs = 1, o.num_deltas = 1, o.use_bp = 1; // overridden per partition by device_opts
  switch (type) {
  case HIPCOMP_TYPE_CHAR: case HIPCOMP_TYPE_UCHAR:
    return cascaded::launch_compress<uint8_t>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream, device_opts);
  case HIPCOMP_TYPE_SHORT: case HIPCOMP_TYPE_USHORT:
    return cascaded::launch_compress<uint16_t>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream, device_opts);
  case HIPCOMP_TYPE_INT: case HIPCOMP_TYPE_UINT:
    return cascaded::launch_compress<uint32_t>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream, device_opts);
  case HIPCOMP_TYPE_LONGLONG: case HIPCOMP_TYPE_ULONGLONG:
    return cascaded::launch_compress<unsigned long long>(device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes, batch_size, o, stream, device_opts);
  default:
    return fail(hipcompErrorInvalidValue, fn, "unsupported element type");
  }
}

} // extern "C"
