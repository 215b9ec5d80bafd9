// b200comp — device-wide RLE / delta / bit-packing (include/hipcomp/primitives.hpp).
//
// Replaces reference src/RunLengthEncodeGPU.hip:166-633 (hipCUB DeviceRunLengthEncode, or
// rleInit/rleReduce/rleFinalize + DeviceScan), src/DeltaGPU.hip:79-169 and
// src/BitPackGPU.hip:184-611.  All three are streaming passes bounded by HBM bandwidth:
//   RLE     tiles of 4096 elements: (1) per-tile run-start count and last start position,
//           (2) one-CTA scan of the tile aggregates (sum / running max), (3) emit: every
//           run END writes its value and length at the scanned index.  The second read of
//           the input comes from L2 for inputs up to ~100 MB.  No library calls.
//   Delta   one pass, 16-byte vectors when the buffers allow.
//   BitPack (1) min/max reduction with a last-block-done fold (one launch), (2) packing:
//           a thread packs 32 values into numBits words with shifts that fold at compile
//           time (switch on numBits), tiles staged through padded shared memory so that
//           both the global reads and the global writes are coalesced.
#include "common.cuh"
#include "workspace.hpp"

#pragma GCC visibility push(default)
#include "hipcomp/b200_ext.h"
#include "hipcomp/primitives.hpp"
#pragma GCC visibility pop

#include <stdexcept>
#include <string>

namespace b200comp {
namespace prim {

template <int BYTES> struct UIntOf;
template <> struct UIntOf<1> { using type = uint8_t; };
template <> struct UIntOf<2> { using type = uint16_t; };
template <> struct UIntOf<4> { using type = uint32_t; };
template <> struct UIntOf<8> { using type = unsigned long long; };

static int type_bytes(hipcompType_t t)
{
  switch (t) {
  case HIPCOMP_TYPE_CHAR:
  case HIPCOMP_TYPE_UCHAR:
  case HIPCOMP_TYPE_BITS: return 1;
  case HIPCOMP_TYPE_SHORT:
  case HIPCOMP_TYPE_USHORT: return 2;
  case HIPCOMP_TYPE_INT:
  case HIPCOMP_TYPE_UINT: return 4;
  case HIPCOMP_TYPE_LONGLONG:
  case HIPCOMP_TYPE_ULONGLONG: return 8;
  default: throw std::runtime_error("Unknown type: " + std::to_string((int)t));
  }
}
static bool type_signed(hipcompType_t t)
{
  return t == HIPCOMP_TYPE_CHAR || t == HIPCOMP_TYPE_SHORT || t == HIPCOMP_TYPE_INT || t == HIPCOMP_TYPE_LONGLONG;
}

static void launch_ok(const char* what)
{
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) throw std::runtime_error(std::string("Failed to launch ") + what + ": " + cudaGetErrorString(e));
}

// ------------------------------------------------------------------ RLE
constexpr int kRleThreads = 256;
constexpr int kRleItems = 16;
constexpr int kRleTile = kRleThreads * kRleItems;

struct RleAgg {
  unsigned long long count;      // run starts in the tile
  unsigned long long last_start; // position of the last run start in the tile (0 if none)
};

template <typename V>
__device__ __forceinline__ void rle_load(const V* in, size_t base, size_t n, V (&x)[kRleItems], V& prev, V& next)
{
  // thread-blocked: this thread owns [base, base + kRleItems)
#pragma unroll
  for (int k = 0; k < kRleItems; ++k) x[k] = base + k < n ? in[base + k] : V(0);
  prev = base > 0 && base - 1 < n ? in[base - 1] : V(0);
  next = base + kRleItems < n ? in[base + kRleItems] : V(0);
}

template <typename V>
__global__ void __launch_bounds__(kRleThreads)
rle_tile_kernel(const V* in, const size_t* num_dev, size_t num_host, RleAgg* agg)
{
  const size_t n = num_dev ? *num_dev : num_host;
  const size_t tile0 = (size_t)blockIdx.x * kRleTile;
  if (tile0 >= n) {
    if (threadIdx.x == 0) agg[blockIdx.x] = RleAgg{0, 0};
    return;
  }
  const size_t base = tile0 + (size_t)threadIdx.x * kRleItems;
  V x[kRleItems], prev, next;
  rle_load(in, base, n, x, prev, next);
  unsigned cnt = 0;
  unsigned long long last = 0;
#pragma unroll
  for (int k = 0; k < kRleItems; ++k) {
    const size_t i = base + k;
    const bool start = i < n && (i == 0 || x[k] != (k ? x[k - 1] : prev));
    cnt += start;
    if (start) last = i;
  }
  __shared__ unsigned s_cnt[kRleThreads / 32];
  __shared__ unsigned long long s_last[kRleThreads / 32];
  unsigned c = cnt;
  unsigned long long l = last;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    c += __shfl_xor_sync(kFull, c, d);
    const unsigned long long o = __shfl_xor_sync(kFull, l, d);
    l = o > l ? o : l;
  }
  if ((threadIdx.x & 31) == 0) {
    s_cnt[threadIdx.x >> 5] = c;
    s_last[threadIdx.x >> 5] = l;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long tc = 0, tl = 0;
    for (int w = 0; w < kRleThreads / 32; ++w) {
      tc += s_cnt[w];
      tl = s_last[w] > tl ? s_last[w] : tl;
    }
    agg[blockIdx.x] = RleAgg{tc, tl};
  }
}

// exclusive scan over the tile aggregates (sum of counts, running max of last starts)
__global__ void __launch_bounds__(1024)
rle_scan_kernel(RleAgg* agg, size_t tiles, size_t* num_out)
{
  __shared__ unsigned long long w_cnt[32], w_last[32];
  __shared__ unsigned long long carry_cnt, carry_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_cnt = carry_last = 0;
  __syncthreads();
  for (size_t t0 = 0; t0 < tiles; t0 += 1024) {
    const size_t t = t0 + threadIdx.x;
    RleAgg a{0, 0};
    if (t < tiles) a = agg[t];
    unsigned long long ic = a.count, il = a.last_start;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long oc = __shfl_up_sync(kFull, ic, d), ol = __shfl_up_sync(kFull, il, d);
      if (lane >= d) {
        ic += oc;
        il = ol > il ? ol : il;
      }
    }
    if (lane == 31) {
      w_cnt[warp] = ic;
      w_last[warp] = il;
    }
    __syncthreads();
    const unsigned long long base_c = carry_cnt, base_l = carry_last;
    unsigned long long pc = 0, pl = 0; // aggregates of the warps in front of this one
    for (int w = 0; w < warp; ++w) {
      pc += w_cnt[w];
      pl = w_last[w] > pl ? w_last[w] : pl;
    }
    // exclusive values for this tile
    unsigned long long ec = __shfl_up_sync(kFull, ic, 1), el = __shfl_up_sync(kFull, il, 1);
    if (lane == 0) ec = el = 0;
    ec += pc + base_c;
    el = el > pl ? el : pl;
    el = el > base_l ? el : base_l;
    if (t < tiles) agg[t] = RleAgg{ec, el};
    __syncthreads();
    if (threadIdx.x == 1023) {
      carry_cnt = base_c + pc + ic;
      const unsigned long long m = il > pl ? il : pl;
      carry_last = m > base_l ? m : base_l;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *num_out = (size_t)carry_cnt;
}

template <typename V, typename C>
__global__ void __launch_bounds__(kRleThreads)
rle_emit_kernel(const V* in, const size_t* num_dev, size_t num_host, const RleAgg* agg, V* out_vals, C* out_cnts,
                void** out_vals_ptr, void** out_cnts_ptr)
{
  const size_t n = num_dev ? *num_dev : num_host;
  const size_t tile0 = (size_t)blockIdx.x * kRleTile;
  if (tile0 >= n) return;
  V* vals = out_vals_ptr ? static_cast<V*>(*out_vals_ptr) : out_vals;
  C* cnts = out_cnts_ptr ? static_cast<C*>(*out_cnts_ptr) : out_cnts;
  const size_t base = tile0 + (size_t)threadIdx.x * kRleItems;
  V x[kRleItems], prev, next;
  rle_load(in, base, n, x, prev, next);
  unsigned cnt = 0;
  unsigned long long last = 0;
  unsigned flags = 0;
#pragma unroll
  for (int k = 0; k < kRleItems; ++k) {
    const size_t i = base + k;
    const bool start = i < n && (i == 0 || x[k] != (k ? x[k - 1] : prev));
    cnt += start;
    flags |= (unsigned)start << k;
    if (start) last = i;
  }
  // block-wide exclusive (sum, max) over threads
  __shared__ unsigned s_cnt[kRleThreads / 32];
  __shared__ unsigned long long s_last[kRleThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned ic = cnt;
  unsigned long long il = last;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned oc = __shfl_up_sync(kFull, ic, d);
    const unsigned long long ol = __shfl_up_sync(kFull, il, d);
    if (lane >= d) {
      ic += oc;
      il = ol > il ? ol : il;
    }
  }
  if (lane == 31) {
    s_cnt[warp] = ic;
    s_last[warp] = il;
  }
  __syncthreads();
  const RleAgg a = agg[blockIdx.x];
  unsigned long long idx = a.count, cur = a.last_start;
  for (int w = 0; w < warp; ++w) {
    idx += s_cnt[w];
    cur = s_last[w] > cur ? s_last[w] : cur;
  }
  unsigned ec = __shfl_up_sync(kFull, ic, 1);
  unsigned long long el = __shfl_up_sync(kFull, il, 1);
  if (lane == 0) {
    ec = 0;
    el = 0;
  }
  idx += ec;           // run starts in front of this thread's first element
  cur = el > cur ? el : cur; // start of the run that is open there
#pragma unroll
  for (int k = 0; k < kRleItems; ++k) {
    const size_t i = base + k;
    if (i >= n) break;
    if ((flags >> k) & 1u) {
      cur = i;
      ++idx;
    }
    const bool end = i + 1 == n || x[k] != (k + 1 < kRleItems ? x[k + 1] : next);
    if (end) {
      vals[idx - 1] = x[k];
      cnts[idx - 1] = (C)(i - cur + 1);
    }
  }
}

static size_t rle_tiles(size_t num) { return (num + kRleTile - 1) / kRleTile; }

template <typename V, typename C>
static void rle_run(void* workspace, size_t workspace_bytes, const void* in, const size_t* num_dev, size_t num, void* vals,
                    void* cnts, void** vals_ptr, void** cnts_ptr, size_t* num_out, cudaStream_t stream)
{
  const size_t tiles = rle_tiles(num);
  Workspace ws(workspace, workspace_bytes, alignof(RleAgg));
  RleAgg* agg = ws.reserve<RleAgg>(tiles ? tiles : 1);
  if (tiles)
    rle_tile_kernel<V><<<(unsigned)tiles, kRleThreads, 0, stream>>>(static_cast<const V*>(in), num_dev, num, agg);
  rle_scan_kernel<<<1, 1024, 0, stream>>>(agg, tiles, num_out);
  if (tiles)
    rle_emit_kernel<V, C><<<(unsigned)tiles, kRleThreads, 0, stream>>>(static_cast<const V*>(in), num_dev, num, agg,
                                                                      static_cast<V*>(vals), static_cast<C*>(cnts), vals_ptr,
                                                                      cnts_ptr);
  launch_ok("RLE kernels");
}

template <typename V>
static void rle_by_count(int count_bytes, void* workspace, size_t wb, const void* in, const size_t* nd, size_t num,
                         void* vals, void* cnts, void** vp, void** cp, size_t* num_out, cudaStream_t s)
{
  switch (count_bytes) {
  case 1: rle_run<V, uint8_t>(workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  case 2: rle_run<V, uint16_t>(workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  case 4: rle_run<V, uint32_t>(workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  default: rle_run<V, unsigned long long>(workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  }
}

static void rle_dispatch(hipcompType_t vt, hipcompType_t ct, void* workspace, size_t wb, const void* in, const size_t* nd,
                         size_t num, void* vals, void* cnts, void** vp, void** cp, size_t* num_out, cudaStream_t s)
{
  const size_t need = hipcomp::RunLengthEncodeGPU::requiredWorkspaceSize(num, vt, ct);
  if (wb < need)
    throw std::runtime_error("Invalid workspace size: " + std::to_string(wb) + ", need at least " + std::to_string(need));
  const int cb = type_bytes(ct);
  switch (type_bytes(vt)) {
  case 1: rle_by_count<uint8_t>(cb, workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  case 2: rle_by_count<uint16_t>(cb, workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  case 4: rle_by_count<uint32_t>(cb, workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  default: rle_by_count<unsigned long long>(cb, workspace, wb, in, nd, num, vals, cnts, vp, cp, num_out, s); break;
  }
}

// ------------------------------------------------------------------ Delta
constexpr int kDeltaThreads = 256;

template <typename V>
__global__ void __launch_bounds__(kDeltaThreads)
delta_kernel(V* const* out_ptr, const V* in, const size_t* num_dev)
{
  constexpr int VE = 16 / (int)sizeof(V);
  const size_t n = *num_dev;
  V* out = *out_ptr;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec_ok) {
    const size_t vecs = n / VE;
    for (size_t v = tid; v < vecs; v += stride) {
      union {
        uint4 q;
        V e[VE];
      } a, r;
      a.q = ld_stream16(reinterpret_cast<const uint4*>(in) + v);
      V prev = v ? in[v * VE - 1] : V(0);
#pragma unroll
      for (int k = 0; k < VE; ++k) {
        r.e[k] = (V)(a.e[k] - prev);
        prev = a.e[k];
      }
      st_stream16(reinterpret_cast<uint4*>(out) + v, r.q);
    }
    for (size_t i = vecs * VE + tid; i < n; i += stride) out[i] = (V)(in[i] - (i ? in[i - 1] : V(0)));
  } else {
    for (size_t i = tid; i < n; i += stride) out[i] = (V)(in[i] - (i ? in[i - 1] : V(0)));
  }
}

template <typename V>
static void delta_run(void** out_ptr, const void* in, const size_t* num_dev, size_t max_num, cudaStream_t s)
{
  constexpr int VE = 16 / (int)sizeof(V);
  const size_t need = (max_num / VE + kDeltaThreads) / kDeltaThreads;
  const size_t cap = (size_t)sm_count() * 8;
  delta_kernel<V><<<(unsigned)(need < cap ? need : cap), kDeltaThreads, 0, s>>>(reinterpret_cast<V* const*>(out_ptr),
                                                                                 static_cast<const V*>(in), num_dev);
  launch_ok("delta_kernel");
}

// ------------------------------------------------------------------ BitPack
constexpr int kBpThreads = 256;

struct BpPartial {
  unsigned long long mn, mx; // order-preserving unsigned keys
};

// signed values are compared through a sign-bit flip so one unsigned reduction serves all types
template <typename U, bool SIGNED>
__device__ __forceinline__ unsigned long long bp_key(U v)
{
  constexpr unsigned long long flip = SIGNED ? (1ull << (8 * sizeof(U) - 1)) : 0ull;
  return (unsigned long long)v ^ flip;
}

template <typename U, bool SIGNED>
__global__ void __launch_bounds__(kBpThreads)
bp_minmax_kernel(const U* in, const size_t* num_dev, BpPartial* partial, unsigned* ticket, U* const* min_out,
                 unsigned char* const* bits_out)
{
  const size_t n = *num_dev;
  unsigned long long mn = ~0ull, mx = 0ull;
  constexpr int VE = 16 / (int)sizeof(U);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t done = 0;
  if ((reinterpret_cast<uintptr_t>(in) & 15) == 0) {
    const size_t vecs = n / VE;
    for (size_t v = tid; v < vecs; v += stride) {
      union {
        uint4 q;
        U e[VE];
      } a;
      a.q = __ldg(reinterpret_cast<const uint4*>(in) + v);
#pragma unroll
      for (int k = 0; k < VE; ++k) {
        const unsigned long long key = bp_key<U, SIGNED>(a.e[k]);
        mn = key < mn ? key : mn;
        mx = key > mx ? key : mx;
      }
    }
    done = vecs * VE;
  }
  for (size_t i = done + tid; i < n; i += stride) {
    const unsigned long long key = bp_key<U, SIGNED>(in[i]);
    mn = key < mn ? key : mn;
    mx = key > mx ? key : mx;
  }
  mn = warp_min(mn);
  mx = warp_max(mx);
  __shared__ unsigned long long s_mn[kBpThreads / 32], s_mx[kBpThreads / 32];
  __shared__ bool is_last;
  if ((threadIdx.x & 31) == 0) {
    s_mn[threadIdx.x >> 5] = mn;
    s_mx[threadIdx.x >> 5] = mx;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kBpThreads / 32; ++w) {
      mn = s_mn[w] < mn ? s_mn[w] : mn;
      mx = s_mx[w] > mx ? s_mx[w] : mx;
    }
    partial[blockIdx.x] = BpPartial{mn, mx};
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  // the last CTA to finish folds the partials (reference bitPackConfigFinalizeKernel)
  __threadfence();
  mn = ~0ull;
  mx = 0ull;
  for (unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
    const volatile unsigned long long* pv = reinterpret_cast<const volatile unsigned long long*>(partial + b);
    const BpPartial p{pv[0], pv[1]};
    mn = p.mn < mn ? p.mn : mn;
    mx = p.mx > mx ? p.mx : mx;
  }
  mn = warp_min(mn);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) {
    s_mn[threadIdx.x >> 5] = mn;
    s_mx[threadIdx.x >> 5] = mx;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kBpThreads / 32; ++w) {
      mn = s_mn[w] < mn ? s_mn[w] : mn;
      mx = s_mx[w] > mx ? s_mx[w] : mx;
    }
    if (n == 0) mn = mx = bp_key<U, SIGNED>(U(0));
    const unsigned long long range = mx - mn; // keys are order preserving: max - min of the values
    **bits_out = (unsigned char)(range ? 64 - __clzll((long long)range) : 0);
    constexpr unsigned long long flip = SIGNED ? (1ull << (8 * sizeof(U) - 1)) : 0ull;
    **min_out = (U)(mn ^ flip);
    *ticket = 0; // ready for the next call on this workspace
  }
}

// 32 values of BW bits -> BW words (LSB first); shifts fold at compile time
template <int BW>
__device__ __forceinline__ void bp_pack32(const uint32_t (&d)[32], uint32_t* dst)
{
#pragma unroll
  for (int t = 0; t < BW; ++t) {
    uint32_t w = 0;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const int lo = k * BW - 32 * t;
      if (lo > -BW && lo < 32) {
        if (lo >= 0) w |= d[k] << lo;
        else w |= d[k] >> (-lo);
      }
    }
    dst[t + ((t) >> 5)] = w; // padded smem row (see bp_pack_kernel)
  }
}

constexpr int kBpGroups = 128;           // groups of 32 values per CTA tile
constexpr int kBpTile = kBpGroups * 32;  // 4096 values

// Pads an index so that rows of 32 words start on successive banks.
__device__ __forceinline__ int pad32(int i) { return i + (i >> 5); }

template <typename U>
__global__ void __launch_bounds__(kBpGroups)
bp_pack_kernel(const U* in, const size_t* num_dev, U* const* min_ptr, const unsigned char* const* bits_ptr,
               void* const* out_ptr)
{
  extern __shared__ uint32_t smem[];
  const size_t n = *num_dev;
  const int bw = **bits_ptr;
  if (bw == 0 || n == 0) return;
  const U mn = **min_ptr;
  uint32_t* out = static_cast<uint32_t*>(*out_ptr);
  // total words: whole 32-bit words, rounded to 64-bit words for 8-byte types (reference OUT type)
  const unsigned long long total_bits = (unsigned long long)n * bw;
  const size_t total_words = sizeof(U) == 8 ? (size_t)((total_bits + 63) / 64) * 2 : (size_t)((total_bits + 31) / 32);
  const size_t tiles = (n + kBpTile - 1) / kBpTile;
  const int lo_bw = bw > 32 ? 32 : bw, hi_bw = bw > 32 ? bw - 32 : 0;
  uint32_t* s_in = smem;                        // pad32(kBpTile) words: low halves (or whole values)
  uint32_t* s_hi = s_in + pad32(kBpTile);       // 8-byte types: high halves
  uint32_t* s_out = sizeof(U) == 8 ? s_hi + pad32(kBpTile) : s_hi; // kBpGroups rows of 65 words
  constexpr int kRow = 65;
  for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const size_t e0 = tile * kBpTile;
    const int cnt = (int)(n - e0 < (size_t)kBpTile ? n - e0 : (size_t)kBpTile);
    __syncthreads();
    for (int i = threadIdx.x; i < kBpTile; i += kBpGroups) {
      unsigned long long d = 0;
      if (i < cnt) d = (unsigned long long)(U)(in[e0 + i] - mn);
      s_in[pad32(i)] = (uint32_t)d;
      if (sizeof(U) == 8) s_hi[pad32(i)] = (uint32_t)(d >> 32);
    }
    __syncthreads();
    // thread g packs values [32g, 32g+32) of the tile into bw words of its row
    const int g = threadIdx.x;
    uint32_t* row = s_out + g * kRow;
    if (bw <= 32) {
      uint32_t d[32];
#pragma unroll
      for (int k = 0; k < 32; ++k) d[k] = s_in[pad32(g * 32 + k)];
      switch (bw) {
#define B200_CASE(B) case B: bp_pack32<B>(d, row); break;
        B200_CASE(1) B200_CASE(2) B200_CASE(3) B200_CASE(4) B200_CASE(5) B200_CASE(6) B200_CASE(7) B200_CASE(8)
        B200_CASE(9) B200_CASE(10) B200_CASE(11) B200_CASE(12) B200_CASE(13) B200_CASE(14) B200_CASE(15) B200_CASE(16)
        B200_CASE(17) B200_CASE(18) B200_CASE(19) B200_CASE(20) B200_CASE(21) B200_CASE(22) B200_CASE(23) B200_CASE(24)
        B200_CASE(25) B200_CASE(26) B200_CASE(27) B200_CASE(28) B200_CASE(29) B200_CASE(30) B200_CASE(31) B200_CASE(32)
#undef B200_CASE
      default: break;
      }
    } else {
      // 33..64 bits: bit-serial accumulator over the 32 values
      unsigned long long acc = 0;
      int nbits = 0, w = 0;
      for (int k = 0; k < 32; ++k) {
        const uint32_t lo = s_in[pad32(g * 32 + k)], hi = s_hi[pad32(g * 32 + k)];
        acc |= (unsigned long long)lo << nbits; // nbits < 32
        row[w + (w >> 5)] = (uint32_t)acc;
        ++w;
        acc >>= 32;
        acc |= (unsigned long long)hi << nbits;
        nbits += hi_bw;
        if (nbits >= 32) {
          row[w + (w >> 5)] = (uint32_t)acc;
          ++w;
          acc >>= 32;
          nbits -= 32;
        }
      }
    }
    (void)lo_bw;
    __syncthreads();
    // the tile's words are contiguous in the output: group g owns words [g*bw, (g+1)*bw)
    const size_t w0 = tile * (size_t)kBpGroups * bw;
    const int tile_words = kBpGroups * bw;
    for (int j = threadIdx.x; j < tile_words; j += kBpGroups) {
      const size_t wi = w0 + j;
      if (wi < total_words) {
        const int gg = j / bw, t = j - gg * bw;
        out[wi] = s_out[gg * kRow + t + (t >> 5)];
      }
    }
  }
}

static size_t bp_grid(size_t max_num)
{
  const size_t need = (max_num + kBpThreads * 64 - 1) / (kBpThreads * 64);
  const size_t cap = (size_t)sm_count() * 4;
  return need < cap ? (need ? need : 1) : cap;
}

template <typename U, bool SIGNED>
static void bp_run(void* workspace, size_t wb, void* const* out_ptr, const void* in, const size_t* num_dev, size_t max_num,
                   void* const* min_ptr, unsigned char* const* bits_ptr, cudaStream_t s)
{
  Workspace ws(workspace, wb, 16);
  unsigned* ticket = ws.reserve<unsigned>(4);
  const size_t grid = bp_grid(max_num);
  BpPartial* partial = ws.reserve<BpPartial>(grid);
  cudaMemsetAsync(ticket, 0, 16, s);
  bp_minmax_kernel<U, SIGNED><<<(unsigned)grid, kBpThreads, 0, s>>>(static_cast<const U*>(in), num_dev, partial, ticket,
                                                                   reinterpret_cast<U* const*>(min_ptr), bits_ptr);
  const size_t tiles = (max_num + kBpTile - 1) / kBpTile;
  const size_t cap = (size_t)sm_count() * 8;
  const int smem = (int)sizeof(uint32_t) * ((sizeof(U) == 8 ? 2 : 1) * (kBpTile + kBpTile / 32) + kBpGroups * 65);
  allow_smem(bp_pack_kernel<U>, smem);
  bp_pack_kernel<U><<<(unsigned)(tiles < cap ? (tiles ? tiles : 1) : cap), kBpGroups, smem, s>>>(
      static_cast<const U*>(in), num_dev, reinterpret_cast<U* const*>(min_ptr), bits_ptr, out_ptr);
  launch_ok("bit packing kernels");
}

} // namespace prim
} // namespace b200comp

// ------------------------------------------------------------------ public classes
namespace hipcomp {
using namespace b200comp;
using namespace b200comp::prim;

void RunLengthEncodeGPU::compress(void* workspace, size_t workspaceSize, hipcompType_t valueType, void* const outValues,
                                  hipcompType_t countType, void* const outCounts, size_t* numOutDevice, const void* in,
                                  const size_t num, hipStream_t stream)
{
  rle_dispatch(valueType, countType, workspace, workspaceSize, in, nullptr, num, outValues, outCounts, nullptr, nullptr,
               numOutDevice, stream);
}

void RunLengthEncodeGPU::compressDownstream(void* workspace, size_t workspaceSize, hipcompType_t valueType,
                                            void** const outValuesPtr, hipcompType_t countType, void** const outCountsPtr,
                                            size_t* numOutDevice, const void* in, const size_t* numDevice,
                                            const size_t maxNum, hipStream_t stream)
{
  rle_dispatch(valueType, countType, workspace, workspaceSize, in, numDevice, maxNum, nullptr, nullptr, outValuesPtr,
               outCountsPtr, numOutDevice, stream);
}

size_t RunLengthEncodeGPU::requiredWorkspaceSize(size_t num, hipcompType_t valueType, hipcompType_t countType)
{
  (void)type_bytes(valueType);
  (void)type_bytes(countType);
  const size_t tiles = rle_tiles(num);
  return Workspace::footprint<RleAgg>(tiles ? tiles : 1, alignof(RleAgg));
}

void DeltaGPU::compress(void*, size_t, hipcompType_t valueType, void** const outValuesPtr, const void* inValues,
                        const size_t* numDevice, const size_t maxNum, hipStream_t stream)
{
  switch (type_bytes(valueType)) {
  case 1: delta_run<uint8_t>(outValuesPtr, inValues, numDevice, maxNum, stream); break;
  case 2: delta_run<uint16_t>(outValuesPtr, inValues, numDevice, maxNum, stream); break;
  case 4: delta_run<uint32_t>(outValuesPtr, inValues, numDevice, maxNum, stream); break;
  default: delta_run<unsigned long long>(outValuesPtr, inValues, numDevice, maxNum, stream); break;
  }
}

size_t DeltaGPU::requiredWorkspaceSize(size_t, hipcompType_t) { return 0; }

void BitPackGPU::compress(void* workspace, size_t workspaceSize, hipcompType_t inType, void* const* outPtr, const void* in,
                          const size_t* numDevice, size_t maxNum, void* const* const minValueDevicePtr,
                          unsigned char* const* const numBitsDevicePtr, hipStream_t stream)
{
  const size_t need = requiredWorkspaceSize(maxNum, inType);
  if (workspaceSize < need)
    throw std::runtime_error("Insufficient workspace size: " + std::to_string(workspaceSize) + ", need " + std::to_string(need));
  const bool sg = type_signed(inType);
#define B200_BP(U)                                                                                                   \
  do {                                                                                                               \
    if (sg) bp_run<U, true>(workspace, workspaceSize, outPtr, in, numDevice, maxNum, minValueDevicePtr, numBitsDevicePtr, stream); \
    else bp_run<U, false>(workspace, workspaceSize, outPtr, in, numDevice, maxNum, minValueDevicePtr, numBitsDevicePtr, stream);   \
  } while (0)
  switch (type_bytes(inType)) {
  case 1: B200_BP(uint8_t); break;
  case 2: B200_BP(uint16_t); break;
  case 4: B200_BP(uint32_t); break;
  default: B200_BP(unsigned long long); break;
  }
#undef B200_BP
}

size_t BitPackGPU::requiredWorkspaceSize(size_t num, hipcompType_t type)
{
  // never below the reference's figure (BitPackGPU.hip:595-603) so existing callers' buffers stay valid
  const size_t blocks = (num + 255) / 256;
  const size_t ref = (size_t)type_bytes(type) * (blocks < 4096 ? blocks : 4096) * sizeof(uint64_t) * 2;
  const size_t mine = Workspace::footprint<unsigned>(4) + Workspace::footprint<BpPartial>(bp_grid(num));
  return ref > mine ? ref : mine;
}

} // namespace hipcomp

// ------------------------------------------------------------------ C view (b200_ext.h)
namespace {
template <typename F>
hipcompStatus_t guarded(const char* fn, F&& f)
{
  try {
    f();
    return hipcompSuccess;
  } catch (const std::exception& e) {
    return b200comp::fail(hipcompErrorInvalidValue, fn, e.what());
  }
}
} // namespace

extern "C" {

hipcompStatus_t b200compRleRequiredWorkspaceSize(size_t num, hipcompType_t value_type, hipcompType_t count_type, size_t* bytes)
{
  B200_CHECK_NOT_NULL(bytes, "b200compRleRequiredWorkspaceSize()");
  return guarded("b200compRleRequiredWorkspaceSize()",
                 [&] { *bytes = hipcomp::RunLengthEncodeGPU::requiredWorkspaceSize(num, value_type, count_type); });
}
hipcompStatus_t b200compRleCompress(void* workspace, size_t workspace_bytes, hipcompType_t value_type, void* out_values,
                                    hipcompType_t count_type, void* out_counts, size_t* num_out_device, const void* in,
                                    size_t num, hipStream_t stream)
{
  return guarded("b200compRleCompress()", [&] {
    hipcomp::RunLengthEncodeGPU::compress(workspace, workspace_bytes, value_type, out_values, count_type, out_counts,
                                          num_out_device, in, num, stream);
  });
}
hipcompStatus_t b200compRleCompressDownstream(void* workspace, size_t workspace_bytes, hipcompType_t value_type,
                                              void** out_values_ptr, hipcompType_t count_type, void** out_counts_ptr,
                                              size_t* num_out_device, const void* in, const size_t* num_device,
                                              size_t max_num, hipStream_t stream)
{
  return guarded("b200compRleCompressDownstream()", [&] {
    hipcomp::RunLengthEncodeGPU::compressDownstream(workspace, workspace_bytes, value_type, out_values_ptr, count_type,
                                                    out_counts_ptr, num_out_device, in, num_device, max_num, stream);
  });
}
hipcompStatus_t b200compDeltaCompress(hipcompType_t value_type, void** out_values_ptr, const void* in,
                                      const size_t* num_device, size_t max_num, hipStream_t stream)
{
  return guarded("b200compDeltaCompress()",
                 [&] { hipcomp::DeltaGPU::compress(nullptr, 0, value_type, out_values_ptr, in, num_device, max_num, stream); });
}
hipcompStatus_t b200compBitPackRequiredWorkspaceSize(size_t num, hipcompType_t type, size_t* bytes)
{
  B200_CHECK_NOT_NULL(bytes, "b200compBitPackRequiredWorkspaceSize()");
  return guarded("b200compBitPackRequiredWorkspaceSize()", [&] { *bytes = hipcomp::BitPackGPU::requiredWorkspaceSize(num, type); });
}
hipcompStatus_t b200compBitPackCompress(void* workspace, size_t workspace_bytes, hipcompType_t in_type, void* const* out_ptr,
                                        const void* in, const size_t* num_device, size_t max_num,
                                        void* const* min_value_device_ptr, unsigned char* const* num_bits_device_ptr,
                                        hipStream_t stream)
{
  return guarded("b200compBitPackCompress()", [&] {
    hipcomp::BitPackGPU::compress(workspace, workspace_bytes, in_type, out_ptr, in, num_device, max_num, min_value_device_ptr,
                                  num_bits_device_ptr, stream);
  });
}

} // extern "C"
