// b200comp — batched Cascaded codec (RLE -> delta -> bit-packing) for sm_100a.
//
// Wire format: bit-exact with reference src/CascadedKernels.hiph (see DESIGN.md
// "Cascaded wire format"); C ABI replaces src/lowlevel/CascadedBatch.hip:306-462.
//
// Kernel design (not the reference's): the reference runs one 128-thread CTA per
// partition with CUB block scans/reductions and __syncthreads() between every
// stage.  Here ONE WARP owns a partition and walks its 4096-byte chunks with
// warp-synchronous primitives only (ballot/popc run detection, shuffle scans,
// per-lane 32-element bit-packing against a bank-conflict-free padded shared
// memory layout), so an SM keeps 16-32 independent partitions in flight with no
// block-wide barrier anywhere.  Grids are persistent (multiple of the SM count)
// and grid-stride over the batch.
#include "common.cuh"

#include <type_traits>

#pragma GCC visibility push(default)
#include "hipcomp/cascaded.h"
#pragma GCC visibility pop

namespace b200comp {
namespace cascaded {

constexpr int kChunkBytes = 4096;
constexpr int kWarpsPerCta = 4;
constexpr int kMaxMetaBytes = 64;

// Padded shared-memory array of N elements: one 4-byte (8 for 64-bit elements)
// pad per 32 elements makes both the striped pattern (lane L reads 32r+L) and
// the blocked pattern (lane L reads 32L+k) conflict free.
template <typename E, int N>
struct PArr {
  static constexpr int PAD = sizeof(E) >= 4 ? 1 : 4 / (int)sizeof(E);
  static constexpr int SIZE = N + (N / 32) * PAD;
  static constexpr int BYTES = SIZE * (int)sizeof(E);
  __device__ __forceinline__ static int p(int i) { return i + (i >> 5) * PAD; }
};

__host__ __device__ constexpr int meta_size(int w, int R, int D)
{
  return ((4 + 4 * (R + 1) + w - 1) / w) * w + ((w * D + 3) / 4) * 4;
}

template <typename E>
__device__ __forceinline__ constexpr int packed_header_bytes()
{
  return sizeof(E) == 8 ? 16 : 8; // roundUp(sizeof(E)+4, max(4,sizeof(E)))
}

// ------------------------------------------------------------------ bit I/O

// Serialise arr[0..n) (padded smem) to `out` (global, 4-byte aligned; for raw
// 8-byte elements 8-byte aligned), raw or bit-packed.  Returns false when the
// 4-byte padded array does not fit in `avail` bytes.  Follows block_write /
// block_bitpack / get_for_bitwidth (reference CascadedKernels.hiph:394-553,646-680).
template <typename E, int N>
__device__ bool write_array(const E* arr, int n, uint8_t* out, size_t avail, bool use_bp, uint32_t& out_bytes)
{
  using P = PArr<E, N>;
  using S = typename std::make_signed<E>::type;
  const int lane = lane_id();
  uint32_t* outw = reinterpret_cast<uint32_t*>(out);

  if (!use_bp) {
    const uint32_t bytes = (uint32_t)n * sizeof(E);
    const uint32_t padded = (bytes + 3u) & ~3u;
    if (padded > avail) return false;
    if (sizeof(E) >= 4) {
      E* oe = reinterpret_cast<E*>(out);
      for (int i = lane; i < n; i += 32) oe[i] = arr[P::p(i)];
    } else {
      constexpr int K = 4 / (int)sizeof(E);
      const int words = padded / 4;
      for (int j = lane; j < words; j += 32) {
        uint32_t wv = 0;
#pragma unroll
        for (int c = 0; c < K; ++c) {
          const int e = j * K + c;
          const uint32_t v = e < n ? (uint32_t)arr[P::p(e)] : 0u;
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
    S lmn = (S)arr[P::p(lane < n ? lane : 0)], lmx = lmn;
    for (int i = lane + 32; i < n; i += 32) {
      const S v = (S)arr[P::p(i)];
      lmn = v < lmn ? v : lmn;
      lmx = v > lmx ? v : lmx;
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
      unsigned long long acc = 0;
      int nbits = 0;
      int wout = g * (int)bw;
#pragma unroll 4
      for (int k = 0; k < 32; ++k) {
        const int e = (g << 5) + k;
        const E d = e < n ? (E)(arr[P::p(e)] - fr) : (E)0;
        if (sizeof(E) == 8 && bw > 32) {
          acc |= (unsigned long long)(uint32_t)d << nbits;
          nbits += 32;
          if (wout < words) data[wout] = (uint32_t)acc;
          ++wout;
          acc >>= 32;
          nbits -= 32;
          acc |= (unsigned long long)((unsigned long long)d >> 32) << nbits;
          nbits += (int)bw - 32;
        } else {
          acc |= (unsigned long long)d << nbits;
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
  const unsigned long long both = ((unsigned long long)hi << 32) | lo;
  const uint32_t v = (uint32_t)(both >> sh);
  return nb == 32 ? v : (v & ((1u << nb) - 1u));
}

// Read an array stored by write_array from global `in` (in_bytes long, all of
// it inside the partition) into padded smem `arr`.  Returns false on a
// malformed array.  Follows block_read / block_bitunpack (:563-618,702-737).
template <typename E, int N>
__device__ bool read_array(const uint8_t* in, uint32_t in_bytes, bool use_bp, E* arr, int& n_out)
{
  using P = PArr<E, N>;
  const int lane = lane_id();
  if (!use_bp) {
    const uint32_t n = in_bytes / sizeof(E);
    if (n > (uint32_t)N) return false;
    const E* ie = reinterpret_cast<const E*>(in);
    for (int i = lane; i < (int)n; i += 32) arr[P::p(i)] = __ldg(ie + i);
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
  for (int i = lane; i < (int)n; i += 32) {
    E d = 0;
    if (bw > 0) {
      const uint32_t bitpos = (uint32_t)i * bw;
      if (sizeof(E) == 8 && bw > 32) {
        const unsigned long long lo = get_bits32(data, words, bitpos, 32);
        const unsigned long long hi = get_bits32(data, words, bitpos + 32, bw - 32);
        d = (E)(lo | (hi << 32));
      } else {
        d = (E)get_bits32(data, words, bitpos, bw);
      }
    }
    arr[P::p(i)] = (E)(d + fr);
  }
  n_out = (int)n;
  return true;
}

// ------------------------------------------------------------------ layers

// In-place run-length encoding of vals[0..n): returns the run count, leaves the
// run values in vals[0..m) and the run lengths in counts[0..m).
// (reference block_rle_compress :124-241)
template <typename T, int N>
__device__ int rle_encode(T* vals, int n, uint16_t* counts)
{
  using P = PArr<T, N>;
  using PC = PArr<uint16_t, N>;
  const int lane = lane_id();
  const unsigned lt = lanemask_lt();
  int m = 0, prev_end = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    const T cur = i < n ? vals[P::p(i)] : (T)0;
    const T nxt = i + 1 < n ? vals[P::p(i + 1)] : (T)0;
    const bool flag = i < n && (i + 1 == n || nxt != cur);
    const unsigned mask = __ballot_sync(kFull, flag);
    __syncwarp();
    if (flag) {
      const unsigned below = mask & lt;
      const int pos = m + __popc(below);
      const int pe = below ? base + 32 - __clz((int)below) : prev_end;
      vals[P::p(pos)] = cur;
      counts[PC::p(pos)] = (uint16_t)(i + 1 - pe);
    }
    if (mask) prev_end = base + 32 - __clz((int)mask);
    m += __popc(mask);
    __syncwarp();
  }
  return m;
}

// In-place adjacent differences: vals[i] = vals[i+1]-vals[i] for i < n-1.
// Returns the seed vals[0].  (reference block_delta_compress :317-328)
template <typename T, int N>
__device__ T delta_encode(T* vals, int n)
{
  using P = PArr<T, N>;
  const int lane = lane_id();
  const T seed = vals[P::p(0)];
  __syncwarp();
  for (int base = 0; base < n - 1; base += 32) {
    const int i = base + lane;
    T d = 0;
    if (i < n - 1) d = (T)(vals[P::p(i + 1)] - vals[P::p(i)]);
    __syncwarp();
    if (i < n - 1) vals[P::p(i)] = d;
    __syncwarp();
  }
  return seed;
}

// In-place inverse delta: vals[0..n) holds differences, result has n+1 elements
// vals[e] = seed + sum_{j<e} d[j].  (reference block_delta_decompress :343-377)
template <typename T, int N>
__device__ void delta_decode(T* vals, int n, T seed)
{
  using P = PArr<T, N>;
  const int lane = lane_id();
  T carry = seed;
  for (int base = 0; base < n; base += 1024) {
    const int lo = base + (lane << 5);
    T sum = 0;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int e = lo + k;
      if (e < n) sum += vals[P::p(e)];
    }
    T total;
    T run = (T)(carry + warp_excl_sum<T>(sum, total));
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int e = lo + k;
      if (e < n) {
        const T d = vals[P::p(e)];
        vals[P::p(e)] = run;
        run += d;
      }
    }
    carry += total;
  }
  __syncwarp();
  if (lane == 0) vals[P::p(n)] = carry;
  __syncwarp();
}

// Expand runs: dst[i] = vals[run(i)].  counts[] is overwritten with run starts,
// idx[] is scratch.  Returns the expanded length or -1 when it exceeds N.
// dst is padded smem when dst_global is false, else a plain global array.
// (reference block_rle_decompress :255-305 — there one thread loops over a whole
// run; here every output element is produced by its own lane.)
template <typename T, int N>
__device__ int rle_decode(const T* vals, uint16_t* counts, int n_runs, uint16_t* idx, T* dst, bool dst_global)
{
  using P = PArr<T, N>;
  using PC = PArr<uint16_t, N>;
  const int lane = lane_id();
  // 1. exclusive scan of the counts (blocked, 32 runs per lane per pass)
  uint32_t carry = 0;
  for (int base = 0; base < n_runs; base += 1024) {
    const int lo = base + (lane << 5);
    uint32_t sum = 0;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int e = lo + k;
      if (e < n_runs) sum += counts[PC::p(e)];
    }
    uint32_t total;
    uint32_t run = carry + warp_excl_sum<uint32_t>(sum, total);
    if (carry + total > (uint32_t)N) return -1;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int e = lo + k;
      if (e < n_runs) {
        const uint32_t c = counts[PC::p(e)];
        // a zero-length run never owns an output position: park it on N (ignored)
        counts[PC::p(e)] = (uint16_t)(c ? run : (uint32_t)N);
        run += c;
      }
    }
    carry += total;
  }
  const int total_out = (int)carry;
  __syncwarp();
  // 2. mark the first output position of every run with the run index
  for (int i = lane; i < total_out; i += 32) idx[PC::p(i)] = 0;
  __syncwarp();
  for (int j = lane; j < n_runs; j += 32) {
    const int s = counts[PC::p(j)];
    if (s < total_out) idx[PC::p(s)] = (uint16_t)j;
  }
  __syncwarp();
  // 3. inclusive max-scan of the marks (blocked)
  uint32_t cmax = 0;
  for (int base = 0; base < total_out; base += 1024) {
    const int lo = base + (lane << 5);
    uint32_t lm = 0;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int e = lo + k;
      if (e < total_out) {
        const uint32_t v = idx[PC::p(e)];
        lm = v > lm ? v : lm;
      }
    }
    const uint32_t incl = warp_incl_max<uint32_t>(lm);
    uint32_t run = __shfl_up_sync(kFull, incl, 1);
    if (lane == 0) run = 0;
    run = run > cmax ? run : cmax;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int e = lo + k;
      if (e < total_out) {
        const uint32_t v = idx[PC::p(e)];
        run = v > run ? v : run;
        idx[PC::p(e)] = (uint16_t)run;
      }
    }
    const uint32_t pass_max = __shfl_sync(kFull, incl, 31);
    cmax = pass_max > cmax ? pass_max : cmax;
  }
  __syncwarp();
  // 4. gather (striped: coalesced when dst is global)
  for (int i = lane; i < total_out; i += 32) {
    const T v = vals[P::p(idx[PC::p(i)])];
    if (dst_global)
      dst[i] = v;
    else
      dst[P::p(i)] = v;
  }
  __syncwarp();
  return total_out;
}

// ------------------------------------------------------------------ compress

template <typename T>
struct CompSmem {
  static constexpr int N = kChunkBytes / (int)sizeof(T);
  static constexpr int BYTES = PArr<T, N>::BYTES + PArr<uint16_t, N>::BYTES + kMaxMetaBytes;
};

__device__ void warp_zero_words(uint32_t* p, int nwords)
{
  for (int i = lane_id(); i < nwords; i += 32) p[i] = 0;
}

// raw copy of a partition behind the 8-byte header (fallback path, :1020-1029)
template <typename T>
__device__ void raw_copy(const T* in, size_t n, uint8_t* out)
{
  const int lane = lane_id();
  const size_t bytes = n * sizeof(T);
  const uint8_t* src = reinterpret_cast<const uint8_t*>(in);
  uint8_t* dst = out + 8;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    const size_t v = bytes / 16;
    for (size_t i = lane; i < v; i += 32) st_stream16(dst + 16 * i, ld_stream16(src + 16 * i));
    for (size_t i = v * 16 + lane; i < bytes; i += 32) dst[i] = src[i];
  } else {
    T* d = reinterpret_cast<T*>(dst);
    for (size_t i = lane; i < n; i += 32) d[i] = in[i];
  }
  const size_t padded = (bytes + 3) & ~(size_t)3;
  if (lane < (int)(padded - bytes)) dst[bytes + lane] = 0;
}

// One warp compresses one partition.  (reference do_cascaded_compression_kernel :761-1058)
template <typename T>
__device__ void compress_partition(const T* in, size_t in_bytes, uint8_t* out, size_t* comp_bytes_out, int type_id,
                                   int R, int D, int bp, uint8_t* smem)
{
  constexpr int N = CompSmem<T>::N;
  constexpr int W = (int)sizeof(T);
  using P = PArr<T, N>;
  T* A = reinterpret_cast<T*>(smem);
  uint16_t* C = reinterpret_cast<uint16_t*>(smem + P::BYTES);
  uint32_t* meta = reinterpret_cast<uint32_t*>(smem + P::BYTES + PArr<uint16_t, N>::BYTES);
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

  for (size_t c = 0; c < n_chunks && use_compression; ++c) {
    const size_t chunk_start = cur;
    cur += msize;
    if (cur > limit) { use_compression = false; break; }
    int n = (int)(n_total - c * N < (size_t)N ? n_total - c * N : (size_t)N);
    const T* src = in + c * N;
    for (int i = lane; i < n; i += 32) A[P::p(i)] = __ldg(src + i);
    warp_zero_words(meta, kMaxMetaBytes / 4);
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

template <typename T>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
compress_kernel(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs, size_t* out_bytes,
                size_t batch, int type_id, int R, int D, int bp)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5;
  uint8_t* smem = smem_raw + (size_t)warp * CompSmem<T>::BYTES;
  const size_t stride = (size_t)gridDim.x * kWarpsPerCta;
  for (size_t p = (size_t)blockIdx.x * kWarpsPerCta + warp; p < batch; p += stride) {
    compress_partition<T>(static_cast<const T*>(in_ptrs[p]), in_bytes[p], static_cast<uint8_t*>(out_ptrs[p]),
                          out_bytes + p, type_id, R, D, bp, smem);
    __syncwarp();
  }
}

// ------------------------------------------------------------------ decompress

template <typename T>
struct DecompSmem {
  static constexpr int N = kChunkBytes / (int)sizeof(T);
  static constexpr int BYTES = 2 * PArr<T, N>::BYTES + 2 * PArr<uint16_t, N>::BYTES;
};

// One warp decompresses one partition.  (reference cascaded_decompression_fcn :1106-1435)
template <typename T>
__device__ void decompress_partition(const uint8_t* comp, size_t comp_bytes, T* out, size_t out_cap,
                                     size_t* actual_out, hipcompStatus_t* status_out, uint8_t* smem)
{
  constexpr int N = DecompSmem<T>::N;
  constexpr int W = (int)sizeof(T);
  using P = PArr<T, N>;
  using PC = PArr<uint16_t, N>;
  T* A = reinterpret_cast<T*>(smem);
  T* B = reinterpret_cast<T*>(smem + P::BYTES);
  uint16_t* C = reinterpret_cast<uint16_t*>(smem + 2 * P::BYTES);
  uint16_t* I = reinterpret_cast<uint16_t*>(smem + 2 * P::BYTES + PC::BYTES);
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
      if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
        const size_t v = bytes / 16;
        for (size_t i = lane; i < v; i += 32) st_stream16(dst + 16 * i, ld_stream16(src + 16 * i));
        for (size_t i = v * 16 + lane; i < bytes; i += 32) dst[i] = src[i];
      } else {
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
      T* cur = A;
      T* oth = B;
      if (!read_array<T, N>(comp + final_pos, final_bytes, bp != 0, cur, n)) { ok = false; break; }
      __syncwarp();
      bool wrote_global = false;
      // exact reverse of the encoder's layer order R1 D1 R2 D2 ... (see DESIGN.md on
      // the reference's interleaving rule for num_deltas > num_RLEs)
      for (int l = layers - 1; l >= 0 && ok; --l) {
        if (l < D) {
          if (n + 1 > N) { ok = false; break; }
          const T seed = *reinterpret_cast<const T*>(comp + chunk + delta_off + l * W);
          delta_decode<T, N>(cur, n, seed);
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
          const bool last = (l == 0);
          if (last) {
            // the expansion of the outermost layer goes straight to the output
            // buffer, if the partition has room for it
            uint32_t sum = 0;
            for (int j = lane; j < n; j += 32) sum += C[PC::p(j)];
            uint32_t tot;
            warp_excl_sum<uint32_t>(sum, tot);
            if (tot > (uint32_t)N || done + tot > n_total) { ok = false; break; }
          }
          const int m = rle_decode<T, N>(cur, C, n, I, last ? out + done : oth, last);
          if (m < 0) { ok = false; break; }
          if (last) {
            wrote_global = true;
          } else {
            T* t = cur; cur = oth; oth = t;
          }
          n = m;
        }
      }
      if (!ok) break;
      if (done + (size_t)n > n_total) { ok = false; break; }
      if (!wrote_global) {
        T* o = out + done;
        for (int i = lane; i < n; i += 32) o[i] = cur[P::p(i)];
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
// that do not match return after one load.
template <typename T>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
decompress_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, void* const* out_ptrs,
                  const size_t* out_caps, size_t* actual, hipcompStatus_t* statuses, size_t batch)
{
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const uint8_t* p0 = static_cast<const uint8_t*>(comp_ptrs[0]);
  const int type_id = p0 ? p0[3] : -1;
  int width;
  switch (type_id) {
  case HIPCOMP_TYPE_CHAR: case HIPCOMP_TYPE_UCHAR: width = 1; break;
  case HIPCOMP_TYPE_SHORT: case HIPCOMP_TYPE_USHORT: width = 2; break;
  case HIPCOMP_TYPE_INT: case HIPCOMP_TYPE_UINT: width = 4; break;
  case HIPCOMP_TYPE_LONGLONG: case HIPCOMP_TYPE_ULONGLONG: width = 8; break;
  default: width = 1; break; // unknown type byte: the 1-byte instantiation reports the errors
  }
  if (width != (int)sizeof(T)) return;
  const int warp = threadIdx.x >> 5;
  uint8_t* smem = smem_raw + (size_t)warp * DecompSmem<T>::BYTES;
  const size_t stride = (size_t)gridDim.x * kWarpsPerCta;
  for (size_t p = (size_t)blockIdx.x * kWarpsPerCta + warp; p < batch; p += stride) {
    const bool known = type_id >= 0 && type_id <= 7;
    decompress_partition<T>(known ? static_cast<const uint8_t*>(comp_ptrs[p]) : nullptr, comp_bytes[p],
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
                                       cudaStream_t stream)
{
  constexpr int smem = CompSmem<T>::BYTES * kWarpsPerCta;
  allow_smem(compress_kernel<T>, smem); // per device, so not cached in a static
  const int per_sm = (227 * 1024) / (smem + 1024);
  compress_kernel<T><<<grid_for(batch, per_sm < 16 ? per_sm : 16), kWarpsPerCta * 32, smem, stream>>>(
      in_ptrs, in_bytes, out_ptrs, out_bytes, batch, (int)o.type, o.num_RLEs, o.num_deltas, o.use_bp);
  return check_launch("hipcompBatchedCascadedCompressAsync()");
}

template <typename T>
static hipcompStatus_t launch_decompress(const void* const* comp_ptrs, const size_t* comp_bytes,
                                         void* const* out_ptrs, const size_t* out_caps, size_t* actual,
                                         hipcompStatus_t* statuses, size_t batch, cudaStream_t stream)
{
  constexpr int smem = DecompSmem<T>::BYTES * kWarpsPerCta;
  allow_smem(decompress_kernel<T>, smem);
  const int per_sm = (227 * 1024) / (smem + 1024);
  decompress_kernel<T><<<grid_for(batch, per_sm < 16 ? per_sm : 16), kWarpsPerCta * 32, smem, stream>>>(
      comp_ptrs, comp_bytes, out_ptrs, out_caps, actual, statuses, batch);
  return check_launch("hipcompBatchedCascadedDecompressAsync()");
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
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
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
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
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

} // extern "C"
