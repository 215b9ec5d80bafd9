// b200comp — high-level interface: LZ4Manager / SnappyManager / CascadedManager + create_manager().
//
// Replaces reference src/highlevel/{ManagerBase,BatchManager,LZ4Manager,SnappyManager,
// CascadedManager}.hpp, {LZ4,Snappy,Cascaded}HlifKernels.hip, CompressionConfigs.cpp,
// PinnedPtrs.hpp, hipcompManagerFactory.cpp and src/hipcomp_common_deps/hlif_shared.hiph.
//
// Container bytes are the reference's: CommonHeader(64) | FormatSpecHeader |
// pad to 8 | u64 offsets[n] | u64 sizes[n] | u32 comp_checksums[n] | u32 decomp_checksums[n] |
// chunk data.  What changes is how the data region is produced.  The reference runs one
// persistent kernel in which every CTA compresses into a private scratch slot, claims
// output space with an atomicAdd and copies the slot over (chunks land in completion
// order).  Here the buffer is cut into WAVES of W chunks; each wave is
//     setup   (pointer / size arrays of the wave, in scratch)
//     encode  (the same batched kernels the C ABI launches, chunk i -> scratch slot i,
//              compressed sizes written straight into the container's sizes[] array)
//     place   (one-CTA exclusive scan of the sizes -> offsets[], running cursor in
//              CommonHeader::comp_data_size; first wave also fills the header)
//     gather  (slot i -> data + offsets[i], 16-byte vector copies)
// so chunks land in index order (deterministic container bytes), no atomics, and every
// kernel improvement of the low-level path is inherited.  Decompression is setup -> the
// batched decoder -> a status fold into the config's pinned status word.
#include "common.cuh"
#include "workspace.hpp"
#include "crc32.cuh"

#pragma GCC visibility push(default)
#include "hipcomp/b200_ext.h"
#include "hipcomp/cascaded.h"
#include "hipcomp/hipcompManagerFactory.hpp"
#include "hipcomp/lz4.h"
#include "hipcomp/snappy.h"
#pragma GCC visibility pop

#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace hipcomp {

// ------------------------------------------------------------------ wire structs
// reference src/hipcomp_common_deps/hlif_shared_types.hpp:58-84
enum FormatType : uint8_t { kFmtLZ4 = 0, kFmtSnappy = 1, kFmtANS = 2, kFmtGDeflate = 3, kFmtCascaded = 4, kFmtBitcomp = 5 };

struct CommonHeader {
  uint32_t magic_number;
  uint8_t major_version;
  uint8_t minor_version;
  uint8_t format;
  uint64_t comp_data_size;
  uint64_t decomp_data_size;
  uint64_t num_chunks;
  bool include_chunk_starts;
  uint32_t full_comp_buffer_checksum;
  uint32_t decomp_buffer_checksum;
  bool include_per_chunk_comp_buffer_checksums;
  bool include_per_chunk_decomp_buffer_checksums;
  uint64_t uncomp_chunk_size;
  uint32_t comp_data_offset;
};
static_assert(sizeof(CommonHeader) == 64, "CommonHeader is a wire format");
static_assert(offsetof(CommonHeader, comp_data_size) == 8 && offsetof(CommonHeader, num_chunks) == 24
                  && offsetof(CommonHeader, uncomp_chunk_size) == 48 && offsetof(CommonHeader, comp_data_offset) == 56,
              "CommonHeader layout");

static void cu_check(cudaError_t e, const char* what)
{
  if (e != cudaSuccess) {
    cudaGetLastError();
    throw std::runtime_error(std::string("hipcomp HLIF: ") + what + ": " + cudaGetErrorString(e));
  }
}
static void st_check(hipcompStatus_t s, const char* what)
{
  if (s != hipcompSuccess) throw std::runtime_error(std::string("hipcomp HLIF: ") + what + " failed with status " + std::to_string((int)s));
}

// ------------------------------------------------------------------ pinned status pool
// reference src/highlevel/PinnedPtrs.hpp:61-211: pinned host words handed out one at a
// time, first block of 10, then blocks of 100; a handle returns its word on destruction.
template <typename T>
struct PinnedPtrPool {
  static constexpr size_t kFirst = 10, kNext = 100;
  std::vector<T*> blocks;
  std::vector<T*> free_list;

  PinnedPtrPool() { grow(kFirst); }
  PinnedPtrPool(const PinnedPtrPool&) = delete;
  PinnedPtrPool& operator=(const PinnedPtrPool&) = delete;
  ~PinnedPtrPool()
  {
    for (T* b : blocks) cudaFreeHost(b);
  }
  void grow(size_t n)
  {
    T* b = nullptr;
    cu_check(cudaHostAlloc(reinterpret_cast<void**>(&b), n * sizeof(T), cudaHostAllocDefault), "cudaHostAlloc(status pool)");
    blocks.push_back(b);
    for (size_t i = 0; i < n; ++i) free_list.push_back(b + i);
  }
  T* take()
  {
    if (free_list.empty()) grow(kNext);
    T* p = free_list.back();
    free_list.pop_back();
    return p;
  }
  void give(T* p) { free_list.push_back(p); }
  size_t available() const { return free_list.size(); }
  size_t capacity() const { return blocks.empty() ? 0 : kFirst + (blocks.size() - 1) * kNext; }
};

struct StatusHandle {
  PinnedPtrPool<hipcompStatus_t>& pool;
  hipcompStatus_t* ptr;
  explicit StatusHandle(PinnedPtrPool<hipcompStatus_t>& p) : pool(p), ptr(p.take()) { *ptr = hipcompSuccess; }
  StatusHandle(const StatusHandle&) = delete;
  ~StatusHandle() { pool.give(ptr); }
};

// ------------------------------------------------------------------ configs
// reference src/highlevel/CompressionConfigs.{hpp,cpp}:52-164
struct CompressionConfig::CompressionConfigImpl {
  StatusHandle status;
  explicit CompressionConfigImpl(PinnedPtrPool<hipcompStatus_t>& pool) : status(pool) {}
};

CompressionConfig::CompressionConfig(PinnedPtrPool<hipcompStatus_t>& pool, size_t uncompressed_buffer_size)
    : impl(std::make_shared<CompressionConfigImpl>(pool)),
      uncompressed_buffer_size(uncompressed_buffer_size),
      max_compressed_buffer_size(0),
      num_chunks(0)
{
}
hipcompStatus_t* CompressionConfig::get_status() const { return impl->status.ptr; }
CompressionConfig::~CompressionConfig() {}
CompressionConfig::CompressionConfig(CompressionConfig&& o)
    : impl(std::move(o.impl)),
      uncompressed_buffer_size(o.uncompressed_buffer_size),
      max_compressed_buffer_size(o.max_compressed_buffer_size),
      num_chunks(o.num_chunks)
{
}
CompressionConfig::CompressionConfig(const CompressionConfig& o)
    : impl(o.impl),
      uncompressed_buffer_size(o.uncompressed_buffer_size),
      max_compressed_buffer_size(o.max_compressed_buffer_size),
      num_chunks(o.num_chunks)
{
}
CompressionConfig& CompressionConfig::operator=(const CompressionConfig& o)
{
  impl = o.impl;
  uncompressed_buffer_size = o.uncompressed_buffer_size;
  max_compressed_buffer_size = o.max_compressed_buffer_size;
  num_chunks = o.num_chunks;
  return *this;
}
CompressionConfig& CompressionConfig::operator=(CompressionConfig&& o)
{
  impl = std::move(o.impl);
  uncompressed_buffer_size = o.uncompressed_buffer_size;
  max_compressed_buffer_size = o.max_compressed_buffer_size;
  num_chunks = o.num_chunks;
  return *this;
}

struct DecompressionConfig::DecompressionConfigImpl {
  StatusHandle status;
  explicit DecompressionConfigImpl(PinnedPtrPool<hipcompStatus_t>& pool) : status(pool) {}
};

DecompressionConfig::DecompressionConfig(PinnedPtrPool<hipcompStatus_t>& pool)
    : impl(std::make_shared<DecompressionConfigImpl>(pool)), decomp_data_size(0), num_chunks(0)
{
}
hipcompStatus_t* DecompressionConfig::get_status() const { return impl->status.ptr; }
DecompressionConfig::~DecompressionConfig() {}
DecompressionConfig::DecompressionConfig(DecompressionConfig&& o)
    : impl(std::move(o.impl)), decomp_data_size(o.decomp_data_size), num_chunks(o.num_chunks)
{
}
DecompressionConfig::DecompressionConfig(const DecompressionConfig& o)
    : impl(o.impl), decomp_data_size(o.decomp_data_size), num_chunks(o.num_chunks)
{
}
DecompressionConfig& DecompressionConfig::operator=(const DecompressionConfig& o)
{
  impl = o.impl;
  decomp_data_size = o.decomp_data_size;
  num_chunks = o.num_chunks;
  return *this;
}
DecompressionConfig& DecompressionConfig::operator=(DecompressionConfig&& o)
{
  impl = std::move(o.impl);
  decomp_data_size = o.decomp_data_size;
  num_chunks = o.num_chunks;
  return *this;
}

// ------------------------------------------------------------------ kernels
namespace {

using b200comp::kFull;

// pointer / size arrays of one compression wave
__global__ void __launch_bounds__(256)
setup_compress_kernel(const uint8_t* decomp, size_t decomp_bytes, size_t chunk, size_t wave_start, size_t n,
                      uint8_t* slots, size_t slot_stride, const void** in_ptrs, size_t* in_bytes, void** out_ptrs)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t begin = (wave_start + i) * chunk;
    const size_t left = decomp_bytes - begin;
    in_ptrs[i] = decomp + begin;
    in_bytes[i] = left < chunk ? left : chunk;
    out_ptrs[i] = slots + i * slot_stride;
  }
}

struct HeaderFill {
  CommonHeader* header; // null: not the first wave
  uint8_t format;
  uint64_t decomp_data_size, num_chunks, uncomp_chunk_size;
  uint32_t comp_data_offset;
  bool checksums; // the per-chunk CRC-32 arrays are filled (b200compManagerSetChecksumMode)
};

// offsets[i] = cursor + exclusive sum of the (padded) sizes of the wave; cursor += total.
// One CTA: the scan is a few microseconds next to the encode of the same wave.
constexpr int kPlaceItems = 8;                                  // chunks per lane per round
constexpr int kPlaceStage = 32 * kPlaceItems + 32;              // stage elements per warp (one pad per 8)
constexpr size_t kPlaceSmem = (size_t)32 * kPlaceStage * sizeof(unsigned long long);
__global__ void __launch_bounds__(1024)
place_kernel(const size_t* sizes, size_t* offsets, size_t n, unsigned long long* cursor, size_t align, HeaderFill hf)
{
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long base_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    base_s = *cursor;
    if (hf.header) {
      CommonHeader* h = hf.header;
      h->magic_number = 0;
      h->major_version = 2;
      h->minor_version = 2;
      h->format = hf.format;
      h->decomp_data_size = hf.decomp_data_size;
      h->num_chunks = hf.num_chunks;
      h->include_chunk_starts = true;
      h->full_comp_buffer_checksum = 0;
      h->decomp_buffer_checksum = 0;
      h->include_per_chunk_comp_buffer_checksums = hf.checksums;
      h->include_per_chunk_decomp_buffer_checksums = hf.checksums;
      h->uncomp_chunk_size = hf.uncomp_chunk_size;
      h->comp_data_offset = hf.comp_data_offset;
    }
  }
  __syncthreads();
  unsigned long long base = base_s;
  // Each warp owns 256 consecutive chunks per round.  Sizes are loaded coalesced, transposed through a padded
  // shared-memory stage so that every lane sums 8 consecutive chunks in registers, scanned once per round
  // (warp shuffle + 32 warp totals), and the offsets go back through the same stage to coalesced stores.
  // A 256 Ki-chunk wave (1 GiB of 4 KiB cascaded chunks) is 32 rounds.
  constexpr int kItems = kPlaceItems;
  const bool pow2 = (align & (align - 1)) == 0; // 1, 4 or 8 in practice: no 64-bit division per chunk
  extern __shared__ unsigned long long place_stage[];
  unsigned long long* stage = place_stage + (size_t)warp * kPlaceStage;
  for (size_t tile = 0; tile < n; tile += 1024 * kItems) {
    const size_t w0 = tile + (size_t)warp * (32 * kItems);
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      const size_t i = w0 + k * 32 + lane;
      unsigned long long x = 0;
      if (i < n) x = pow2 ? (sizes[i] + align - 1) & ~(unsigned long long)(align - 1) : (sizes[i] + align - 1) / align * align;
      const int e = k * 32 + lane;
      stage[e + (e >> 3)] = x;
    }
    __syncwarp();
    unsigned long long v[kItems];
    unsigned long long sum = 0;
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      v[k] = stage[lane * (kItems + 1) + k];
      sum += v[k];
    }
    unsigned long long incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long o = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = warp_tot[lane], wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(kFull, wi, d);
        if (lane >= d) wi += o;
      }
      warp_tot[lane] = wi - w; // exclusive
      if (lane == 31) base_s = base + wi;
    }
    __syncthreads();
    unsigned long long run = base + warp_tot[warp] + (incl - sum);
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      stage[lane * (kItems + 1) + k] = run;
      run += v[k];
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      const size_t i = w0 + k * 32 + lane;
      const int e = k * 32 + lane;
      if (i < n) offsets[i] = stage[e + (e >> 3)];
    }
    base = base_s;
    __syncthreads();
  }
  if (threadIdx.x == 0) *cursor = base;
}

// Two-level placement for large waves (one CTA scanning 256 Ki sizes is latency-bound at ~0.2 ms):
// per-tile sums -> place_kernel over the tile sums (cursor + header, as above) -> per-tile scan.
constexpr int kPlaceTile = 256 * kPlaceItems; // chunks per CTA

__device__ __forceinline__ unsigned long long padded_size(size_t x, size_t align, bool pow2)
{
  return pow2 ? (x + align - 1) & ~(unsigned long long)(align - 1) : (x + align - 1) / align * align;
}

__global__ void __launch_bounds__(256)
place_sum_kernel(const size_t* sizes, size_t n, size_t align, size_t* tile_sum)
{
  __shared__ unsigned long long wsum[8];
  const bool pow2 = (align & (align - 1)) == 0;
  const size_t t0 = (size_t)blockIdx.x * kPlaceTile;
  unsigned long long s = 0;
#pragma unroll
  for (int k = 0; k < kPlaceItems; ++k) {
    const size_t i = t0 + (size_t)k * 256 + threadIdx.x;
    if (i < n) s += padded_size(sizes[i], align, pow2);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(kFull, s, d);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += wsum[w];
    tile_sum[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256)
place_scatter_kernel(const size_t* sizes, size_t* offsets, size_t n, size_t align, const size_t* tile_base)
{
  __shared__ unsigned long long stage_all[8 * kPlaceStage];
  __shared__ unsigned long long warp_tot[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool pow2 = (align & (align - 1)) == 0;
  unsigned long long* stage = stage_all + warp * kPlaceStage;
  const size_t w0 = (size_t)blockIdx.x * kPlaceTile + (size_t)warp * (32 * kPlaceItems);
#pragma unroll
  for (int k = 0; k < kPlaceItems; ++k) {
    const size_t i = w0 + k * 32 + lane;
    const int e = k * 32 + lane;
    stage[e + (e >> 3)] = i < n ? padded_size(sizes[i], align, pow2) : 0ull;
  }
  __syncwarp();
  unsigned long long v[kPlaceItems];
  unsigned long long sum = 0;
#pragma unroll
  for (int k = 0; k < kPlaceItems; ++k) {
    v[k] = stage[lane * (kPlaceItems + 1) + k];
    sum += v[k];
  }
  unsigned long long incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  unsigned long long run = tile_base[blockIdx.x] + (incl - sum);
  for (int w = 0; w < warp; ++w) run += warp_tot[w];
#pragma unroll
  for (int k = 0; k < kPlaceItems; ++k) {
    stage[lane * (kPlaceItems + 1) + k] = run;
    run += v[k];
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < kPlaceItems; ++k) {
    const size_t i = w0 + k * 32 + lane;
    const int e = k * 32 + lane;
    if (i < n) offsets[i] = stage[e + (e >> 3)];
  }
}

// Chunk g of the container -> (stream pointer, stream size, output pointer, output capacity).  The container is
// untrusted: a chunk that lies beyond the declared decompressed size, or whose stream does not fit the declared
// compressed data, is handed to the decoder as an empty stream with no output room and flags the config's status.
__global__ void __launch_bounds__(256)
setup_decompress_kernel(const uint8_t* data, const size_t* offsets, const size_t* sizes, const uint64_t* comp_data_size_ptr,
                        uint8_t* decomp, size_t decomp_bytes, size_t chunk, size_t wave_start, size_t n,
                        const void** comp_ptrs, size_t* comp_sizes, void** out_ptrs, size_t* out_caps,
                        hipcompStatus_t* status)
{
  const size_t comp_data_size = (size_t)*comp_data_size_ptr; // header word of the (device-resident) container
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t g = wave_start + i;
    const size_t begin = g * chunk;
    const size_t left = begin < decomp_bytes ? decomp_bytes - begin : 0;
    const size_t off = offsets[g], sz = sizes[g];
    const bool bad = left == 0 || off > comp_data_size || sz > comp_data_size - off;
    comp_ptrs[i] = data + (bad ? 0 : off);
    comp_sizes[i] = bad ? 0 : sz;
    out_ptrs[i] = decomp + (bad ? 0 : begin);
    out_caps[i] = bad ? 0 : (left < chunk ? left : chunk);
    if (bad) *status = hipcompErrorCannotDecompress;
  }
}

// any failed chunk -> the config's pinned status word (reference hlif_shared.hiph:332-337)
__global__ void __launch_bounds__(256)
fold_status_kernel(const hipcompStatus_t* statuses, size_t n, hipcompStatus_t* out)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const hipcompStatus_t s = statuses[i];
    if (s != hipcompSuccess) *out = s;
  }
}

inline int grid_for(size_t n, int block)
{
  const size_t need = (n + block - 1) / block;
  const size_t cap = (size_t)b200comp::sm_count() * 8;
  return (int)(need < cap ? (need ? need : 1) : cap);
}

template <typename T>
inline T* align8(T* p)
{
  return reinterpret_cast<T*>((reinterpret_cast<uintptr_t>(p) + 7) & ~(uintptr_t)7);
}

} // namespace

// ------------------------------------------------------------------ the manager
// One class plays the reference's ManagerBase<F> + BatchManager<F>
// (src/highlevel/ManagerBase.hpp:80-324, BatchManager.hpp:71-281); a codec supplies the
// format header and the two batched calls.
namespace {

// Uncompressed bytes per wave: the scratch buffer (compressed-chunk slots) scales with it, while the
// setup / placement / gather launches around a wave cost ~0.1 ms: 128 MiB for the LZ codecs (a wave takes
// ~1 ms), 512 MiB for Cascaded, whose kernels need ~0.2 ms for that (measured: CascadedManager compress
// 1881 GB/s with 1 GiB waves, 1424 with 128 MiB).
constexpr size_t kWaveBytes = (size_t)128 << 20;
// chunks per decompression wave (36 bytes of scratch each): enough to fill the decoders' grid
constexpr size_t kDecompWaveChunks = (size_t)1 << 16;

class BatchManager : public hipcompManagerBase
{
public:
  BatchManager(size_t uncomp_chunk_size, hipStream_t stream, int device_id)
      : chunk_(uncomp_chunk_size), stream_(stream), device_id_(device_id)
  {
    if (chunk_ == 0) throw std::invalid_argument("hipcomp HLIF: uncompressed chunk size must be > 0");
    cu_check(cudaHostAlloc(reinterpret_cast<void**>(&header_host_), sizeof(CommonHeader) + 64, cudaHostAllocDefault),
             "cudaHostAlloc(header)");
    cu_check(cudaFuncSetAttribute(place_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPlaceSmem),
             "cudaFuncSetAttribute(place_kernel)");
  }
  ~BatchManager() override
  {
    scratch_.release();
    cudaFreeHost(header_host_);
  }
  BatchManager(const BatchManager&) = delete;
  BatchManager& operator=(const BatchManager&) = delete;

  // ---- codec hooks
  virtual uint8_t format() const = 0;
  virtual const void* format_header() const = 0;
  virtual size_t format_header_bytes() const = 0;
  virtual size_t max_comp_chunk() const = 0;
  virtual size_t wave_bytes() const { return kWaveBytes; }
  virtual size_t comp_temp_bytes(size_t n) const = 0;
  virtual size_t chunk_align() const { return 1; }
  virtual void batch_compress(const void* const* in_ptrs, const size_t* in_bytes, size_t n, void* temp, size_t temp_bytes,
                              void* const* out_ptrs, size_t* out_bytes) = 0;
  virtual void batch_decompress(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps,
                                size_t* actual, size_t n, void* const* out_ptrs, hipcompStatus_t* statuses) = 0;

  // ---- hipcompManagerBase
  CompressionConfig configure_compression(const size_t decomp_buffer_size) final
  {
    CompressionConfig cfg{status_pool_, decomp_buffer_size};
    cfg.num_chunks = (decomp_buffer_size + chunk_ - 1) / chunk_;
    // true layout (the reference budgets 4 B per size entry, BatchManager.hpp:213-224)
    cfg.max_compressed_buffer_size = sizeof(CommonHeader) + format_header_bytes() + 8 + cfg.num_chunks * (8 + 8 + 4 + 4)
                                     + cfg.num_chunks * slot_stride_;
    return cfg;
  }

  void compress(const uint8_t* decomp_buffer, uint8_t* comp_buffer, const CompressionConfig& cfg) final
  {
    ensure_scratch();
    const size_t n = cfg.num_chunks;
    CommonHeader* header = reinterpret_cast<CommonHeader*>(comp_buffer);
    uint8_t* fmt = comp_buffer + sizeof(CommonHeader);
    cu_check(cudaMemcpyAsync(fmt, format_header(), format_header_bytes(), cudaMemcpyHostToDevice, stream_), "copy format header");
    cu_check(cudaMemsetAsync(&header->comp_data_size, 0, sizeof(uint64_t), stream_), "reset comp_data_size");
    size_t* offsets = align8(reinterpret_cast<size_t*>(fmt + format_header_bytes()));
    size_t* sizes = offsets + n;
    // two u32 arrays: CRC-32 of every compressed / uncompressed chunk when checksums are on, else zero (as the
    // reference leaves them)
    uint32_t* comp_crc = reinterpret_cast<uint32_t*>(sizes + n);
    uint32_t* decomp_crc = comp_crc + n;
    uint8_t* data = reinterpret_cast<uint8_t*>(sizes + n) + 8 * n;
    const bool crcs = checksum_mode_ != 0;
    if (!crcs) cu_check(cudaMemsetAsync(sizes + n, 0, 8 * n, stream_), "zero checksums");

    HeaderFill hf{header, format(), cfg.uncompressed_buffer_size, n, chunk_, (uint32_t)(data - comp_buffer), crcs};
    unsigned long long* cursor = reinterpret_cast<unsigned long long*>(&header->comp_data_size);
    if (n == 0) {
      place_kernel<<<1, 1024, kPlaceSmem, stream_>>>(sizes, offsets, 0, cursor, 1, hf);
      cu_check(cudaGetLastError(), "place_kernel");
      return;
    }
    scratch_.rewind();
    uint8_t* slots = scratch_.reserve<uint8_t>(wave_ * slot_stride_);
    const size_t temp_bytes = comp_temp_bytes(wave_);
    uint8_t* temp = scratch_.reserve<uint8_t>(temp_bytes);
    const void** in_ptrs = scratch_.reserve<const void*>(wave_);
    size_t* in_bytes = scratch_.reserve<size_t>(wave_);
    void** out_ptrs = scratch_.reserve<void*>(wave_);
    size_t* tile_sum = scratch_.reserve<size_t>(wave_ / kPlaceTile + 1);
    size_t* tile_base = scratch_.reserve<size_t>(wave_ / kPlaceTile + 1);
    for (size_t start = 0; start < n; start += wave_) {
      const size_t m = n - start < wave_ ? n - start : wave_;
      setup_compress_kernel<<<grid_for(m, 256), 256, 0, stream_>>>(decomp_buffer, cfg.uncompressed_buffer_size, chunk_, start,
                                                                  m, slots, slot_stride_, in_ptrs, in_bytes, out_ptrs);
      batch_compress(in_ptrs, in_bytes, m, temp, temp_bytes, out_ptrs, sizes + start);
      if (crcs) {
        const int g = crc_grid(m);
        b200comp::crc::crc32_chunks_kernel<<<g, b200comp::crc::kWarps * 32, 0, stream_>>>(
            in_ptrs, in_bytes, m, decomp_crc + start, nullptr, nullptr, nullptr);
        b200comp::crc::crc32_chunks_kernel<<<g, b200comp::crc::kWarps * 32, 0, stream_>>>(
            out_ptrs, sizes + start, m, comp_crc + start, nullptr, nullptr, nullptr);
      }
      if (m <= 8192) {
        place_kernel<<<1, 1024, kPlaceSmem, stream_>>>(sizes + start, offsets + start, m, cursor, chunk_align(), hf);
      } else {
        const size_t tiles = (m + kPlaceTile - 1) / kPlaceTile;
        place_sum_kernel<<<(int)tiles, 256, 0, stream_>>>(sizes + start, m, chunk_align(), tile_sum);
        place_kernel<<<1, 1024, kPlaceSmem, stream_>>>(tile_sum, tile_base, tiles, cursor, 1, hf);
        place_scatter_kernel<<<(int)tiles, 256, 0, stream_>>>(sizes + start, offsets + start, m, chunk_align(), tile_base);
      }
      hf.header = nullptr;
      st_check(b200compGatherChunksAsync(out_ptrs, sizes + start, offsets + start, data, m, stream_), "gather");
    }
    cu_check(cudaGetLastError(), "compress launches");
    (void)cfg.get_status(); // encoders cannot fail per chunk: status stays hipcompSuccess
  }

  DecompressionConfig configure_decompression(const uint8_t* comp_buffer) final
  {
    // The reference copies two header words asynchronously into the config it then returns by
    // value (ManagerBase.hpp:164-168) and asks the caller to synchronise; here the header is read
    // with one blocking copy so the returned config is complete.
    DecompressionConfig cfg{status_pool_};
    cu_check(cudaMemcpyAsync(header_host_, comp_buffer, sizeof(CommonHeader), cudaMemcpyDeviceToHost, stream_), "read header");
    cu_check(cudaStreamSynchronize(stream_), "sync");
    cfg.decomp_data_size = header_host_->decomp_data_size;
    cfg.num_chunks = (uint32_t)header_host_->num_chunks;
    // the container is untrusted: chunk size and chunk count must agree with this manager and with each other
    const uint64_t want = (header_host_->decomp_data_size + chunk_ - 1) / chunk_;
    if (header_host_->uncomp_chunk_size != chunk_ || header_host_->num_chunks != want) {
      *cfg.get_status() = hipcompErrorCannotDecompress;
      cfg.decomp_data_size = 0;
      cfg.num_chunks = 0;
    }
    return cfg;
  }

  DecompressionConfig configure_decompression(const CompressionConfig& comp_config) final
  {
    DecompressionConfig cfg{status_pool_};
    cfg.decomp_data_size = comp_config.uncompressed_buffer_size;
    cfg.num_chunks = (uint32_t)comp_config.num_chunks;
    return cfg;
  }

  void decompress(uint8_t* decomp_buffer, const uint8_t* comp_buffer, const DecompressionConfig& cfg) final
  {
    ensure_scratch();
    const size_t n = cfg.num_chunks;
    if (n == 0) return;
    const uint8_t* fmt = comp_buffer + sizeof(CommonHeader);
    const size_t* offsets = align8(reinterpret_cast<const size_t*>(fmt + format_header_bytes()));
    const size_t* sizes = offsets + n;
    const uint32_t* comp_crc = reinterpret_cast<const uint32_t*>(sizes + n);
    const uint32_t* decomp_crc = comp_crc + n;
    const uint8_t* data = reinterpret_cast<const uint8_t*>(sizes + n) + 8 * n;
    const CommonHeader* header = reinterpret_cast<const CommonHeader*>(comp_buffer);
    const bool crcs = checksum_mode_ != 0; // verified where the container's header says it carries them
    scratch_.rewind();
    const void** comp_ptrs = scratch_.reserve<const void*>(dwave_);
    size_t* comp_sizes = scratch_.reserve<size_t>(dwave_);
    void** out_ptrs = scratch_.reserve<void*>(dwave_);
    size_t* out_caps = scratch_.reserve<size_t>(dwave_);
    size_t* actual = scratch_.reserve<size_t>(dwave_);
    hipcompStatus_t* statuses = scratch_.reserve<hipcompStatus_t>(dwave_);
    for (size_t start = 0; start < n; start += dwave_) {
      const size_t m = n - start < dwave_ ? n - start : dwave_;
      setup_decompress_kernel<<<grid_for(m, 256), 256, 0, stream_>>>(
          data, offsets, sizes, &reinterpret_cast<const CommonHeader*>(comp_buffer)->comp_data_size, decomp_buffer, cfg.decomp_data_size, chunk_, start, m, comp_ptrs, comp_sizes,
          out_ptrs, out_caps, cfg.get_status());
      if (crcs)
        b200comp::crc::crc32_chunks_kernel<<<crc_grid(m), b200comp::crc::kWarps * 32, 0, stream_>>>(
            comp_ptrs, comp_sizes, m, nullptr, comp_crc + start, &header->include_per_chunk_comp_buffer_checksums,
            cfg.get_status());
      batch_decompress(comp_ptrs, comp_sizes, out_caps, actual, m, out_ptrs, statuses);
      fold_status_kernel<<<grid_for(m, 256), 256, 0, stream_>>>(statuses, m, cfg.get_status());
      if (crcs)
        b200comp::crc::crc32_chunks_kernel<<<crc_grid(m), b200comp::crc::kWarps * 32, 0, stream_>>>(
            out_ptrs, actual, m, nullptr, decomp_crc + start, &header->include_per_chunk_decomp_buffer_checksums,
            cfg.get_status());
    }
    cu_check(cudaGetLastError(), "decompress launches");
  }

  // 0: the arrays stay zero and are ignored (the reference's behaviour); 1: filled on compress, verified on
  // decompress (include/hipcomp/b200_ext.h, b200compManagerSetChecksumMode)
  void set_checksum_mode(int mode) { checksum_mode_ = mode; }

  void set_scratch_buffer(uint8_t* new_scratch_buffer) final { scratch_.adopt(new_scratch_buffer, scratch_bytes_); }
  size_t get_required_scratch_buffer_size() final { return scratch_bytes_; }

  size_t get_compressed_output_size(uint8_t* comp_buffer) final
  {
    cu_check(cudaMemcpy(header_host_, comp_buffer, sizeof(CommonHeader), cudaMemcpyDefault), "read header");
    return header_host_->comp_data_size + header_host_->comp_data_offset;
  }

protected:
  // Called at the end of the concrete constructor (reference BatchManager::finish_init).
  void finish_init()
  {
    slot_stride_ = b200comp::round_up_sz(max_comp_chunk(), 16);
    size_t w = wave_bytes() / chunk_;
    if (w < 64) w = 64;
    if (w > (1u << 20)) w = 1u << 20;
    wave_ = w;
    dwave_ = kDecompWaveChunks;
    using W = b200comp::Workspace;
    const size_t comp_need = W::footprint<uint8_t>(wave_ * slot_stride_) + W::footprint<uint8_t>(comp_temp_bytes(wave_))
                             + 3 * W::footprint<size_t>(wave_) + 2 * W::footprint<size_t>(wave_ / kPlaceTile + 1);
    const size_t decomp_need = 5 * W::footprint<size_t>(dwave_) + W::footprint<hipcompStatus_t>(dwave_);
    scratch_bytes_ = comp_need > decomp_need ? comp_need : decomp_need;
  }

  size_t chunk_;
  hipStream_t stream_;
  int device_id_;

private:
  void ensure_scratch()
  {
    if (!scratch_.valid()) scratch_.acquire(scratch_bytes_, stream_);
  }

  static int crc_grid(size_t m)
  {
    const size_t need = (m + b200comp::crc::kWarps - 1) / b200comp::crc::kWarps;
    const size_t cap = (size_t)b200comp::sm_count() * 8;
    return (int)(need < cap ? need : cap);
  }

  int checksum_mode_ = 0;
  CommonHeader* header_host_ = nullptr;
  PinnedPtrPool<hipcompStatus_t> status_pool_;
  b200comp::Workspace scratch_;
  size_t scratch_bytes_ = 0, slot_stride_ = 0, wave_ = 0, dwave_ = 0;
};

// ---- LZ4 (reference src/highlevel/LZ4Manager.hpp:67-174)
class LZ4BatchManager final : public BatchManager
{
public:
  LZ4BatchManager(size_t chunk, hipcompType_t data_type, hipStream_t stream, int device_id)
      : BatchManager(chunk, stream, device_id)
  {
    spec_.data_type = data_type;
    opts_.data_type = data_type;
    st_check(hipcompBatchedLZ4CompressGetMaxOutputChunkSize(chunk, opts_, &max_comp_), "LZ4 GetMaxOutputChunkSize");
    finish_init();
  }
  uint8_t format() const override { return kFmtLZ4; }
  const void* format_header() const override { return &spec_; }
  size_t format_header_bytes() const override { return sizeof(LZ4FormatSpecHeader); }
  size_t max_comp_chunk() const override { return max_comp_; }
  size_t comp_temp_bytes(size_t n) const override
  {
    size_t t = 0;
    st_check(hipcompBatchedLZ4CompressGetTempSize(n, chunk_, opts_, &t), "LZ4 CompressGetTempSize");
    return t;
  }
  void batch_compress(const void* const* in_ptrs, const size_t* in_bytes, size_t n, void* temp, size_t temp_bytes,
                      void* const* out_ptrs, size_t* out_bytes) override
  {
    st_check(hipcompBatchedLZ4CompressAsync(in_ptrs, in_bytes, chunk_, n, temp, temp_bytes, out_ptrs, out_bytes, opts_, stream_),
             "LZ4 CompressAsync");
  }
  void batch_decompress(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps, size_t* actual,
                        size_t n, void* const* out_ptrs, hipcompStatus_t* statuses) override
  {
    st_check(hipcompBatchedLZ4DecompressAsync(comp_ptrs, comp_bytes, out_caps, actual, n, nullptr, 0, out_ptrs, statuses, stream_),
             "LZ4 DecompressAsync");
  }

private:
  LZ4FormatSpecHeader spec_{};
  hipcompBatchedLZ4Opts_t opts_{};
  size_t max_comp_ = 0;
};

// ---- Snappy (reference src/highlevel/SnappyManager.hpp:65-151)
class SnappyBatchManager final : public BatchManager
{
public:
  SnappyBatchManager(size_t chunk, hipStream_t stream, int device_id) : BatchManager(chunk, stream, device_id)
  {
    st_check(hipcompBatchedSnappyCompressGetMaxOutputChunkSize(chunk, hipcompBatchedSnappyDefaultOpts, &max_comp_),
             "Snappy GetMaxOutputChunkSize");
    finish_init();
  }
  uint8_t format() const override { return kFmtSnappy; }
  const void* format_header() const override { return &spec_; }
  size_t format_header_bytes() const override { return sizeof(SnappyFormatSpecHeader); } // empty struct: 1 byte
  size_t max_comp_chunk() const override { return max_comp_; }
  size_t comp_temp_bytes(size_t) const override { return 0; }
  void batch_compress(const void* const* in_ptrs, const size_t* in_bytes, size_t n, void* temp, size_t temp_bytes,
                      void* const* out_ptrs, size_t* out_bytes) override
  {
    st_check(hipcompBatchedSnappyCompressAsync(in_ptrs, in_bytes, chunk_, n, temp, temp_bytes, out_ptrs, out_bytes,
                                               hipcompBatchedSnappyDefaultOpts, stream_),
             "Snappy CompressAsync");
  }
  void batch_decompress(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps, size_t* actual,
                        size_t n, void* const* out_ptrs, hipcompStatus_t* statuses) override
  {
    st_check(hipcompBatchedSnappyDecompressAsync(comp_ptrs, comp_bytes, out_caps, actual, n, nullptr, 0, out_ptrs, statuses,
                                                 stream_),
             "Snappy DecompressAsync");
  }

private:
  SnappyFormatSpecHeader spec_{};
  size_t max_comp_ = 0;
};

// ---- Cascaded (reference src/highlevel/CascadedManager.hpp:65-165): one manager chunk
//      (options.chunk_size bytes) is one batched-Cascaded partition.
class CascadedBatchManager final : public BatchManager
{
public:
  CascadedBatchManager(const hipcompBatchedCascadedOpts_t& options, hipStream_t stream, int device_id)
      : BatchManager(options.chunk_size, stream, device_id)
  {
    std::memset(&spec_, 0, sizeof(spec_)); // padding bytes of the options struct go on the wire
    spec_.options.chunk_size = options.chunk_size;
    spec_.options.type = options.type;
    spec_.options.num_RLEs = options.num_RLEs;
    spec_.options.num_deltas = options.num_deltas;
    spec_.options.use_bp = options.use_bp;
    st_check(hipcompBatchedCascadedCompressGetMaxOutputChunkSize(chunk_, spec_.options, &max_comp_),
             "Cascaded GetMaxOutputChunkSize");
    finish_init();
  }
  uint8_t format() const override { return kFmtCascaded; }
  const void* format_header() const override { return &spec_; }
  size_t format_header_bytes() const override { return sizeof(CascadedFormatSpecHeader); }
  size_t max_comp_chunk() const override { return max_comp_; }
  size_t wave_bytes() const override { return (size_t)512 << 20; }
  size_t comp_temp_bytes(size_t) const override { return 0; }
  size_t chunk_align() const override { return 8; } // partitions are read as words of the element type
  void batch_compress(const void* const* in_ptrs, const size_t* in_bytes, size_t n, void* temp, size_t temp_bytes,
                      void* const* out_ptrs, size_t* out_bytes) override
  {
    st_check(hipcompBatchedCascadedCompressAsync(in_ptrs, in_bytes, chunk_, n, temp, temp_bytes, out_ptrs, out_bytes,
                                                 spec_.options, stream_),
             "Cascaded CompressAsync");
  }
  void batch_decompress(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps, size_t* actual,
                        size_t n, void* const* out_ptrs, hipcompStatus_t* statuses) override
  {
    // the manager knows the element type and the layers: one launch (the C ABI launches one per width)
    st_check(b200comp::cascaded::decompress_known(spec_.options, comp_ptrs, comp_bytes, out_caps, actual, n, out_ptrs,
                                                  statuses, stream_),
             "Cascaded DecompressAsync");
  }

private:
  CascadedFormatSpecHeader spec_;
  size_t max_comp_ = 0;
};

} // namespace

// behind b200compManagerSetChecksumMode (hlif_capi.cu): reaches the batch manager inside a PimplManager
// (whose `impl` is protected, as in the reference's class)
namespace {
struct PimplAccess : PimplManager {
  static hipcompManagerBase* get(PimplManager* p) { return (p->*(&PimplAccess::impl)).get(); }
};
} // namespace
bool b200_set_checksum_mode(hipcompManagerBase* m, int mode)
{
  if (auto* p = dynamic_cast<PimplManager*>(m)) m = PimplAccess::get(p);
  if (auto* b = dynamic_cast<BatchManager*>(m)) {
    b->set_checksum_mode(mode);
    return true;
  }
  return false;
}

LZ4Manager::LZ4Manager(size_t uncomp_chunk_size, hipcompType_t data_type, hipStream_t user_stream, const int device_id)
{
  impl = std::make_unique<LZ4BatchManager>(uncomp_chunk_size, data_type, user_stream, device_id);
}
LZ4Manager::~LZ4Manager() {}

SnappyManager::SnappyManager(size_t uncomp_chunk_size, hipStream_t user_stream, int device_id)
{
  impl = std::make_unique<SnappyBatchManager>(uncomp_chunk_size, user_stream, device_id);
}
SnappyManager::~SnappyManager() {}

CascadedManager::CascadedManager(const hipcompBatchedCascadedOpts_t& options, hipStream_t user_stream, int device_id)
{
  impl = std::make_unique<CascadedBatchManager>(options, user_stream, device_id);
}
CascadedManager::~CascadedManager() {}

// reference src/highlevel/hipcompManagerFactory.cpp:65-148
std::shared_ptr<hipcompManagerBase> create_manager(const uint8_t* comp_buffer, hipStream_t stream, const int device_id)
{
  struct {
    CommonHeader common;
    uint8_t spec[32];
  } host;
  cu_check(cudaMemcpyAsync(&host, comp_buffer, sizeof(CommonHeader) + sizeof(CascadedFormatSpecHeader), cudaMemcpyDeviceToHost,
                           stream),
           "read container header");
  cu_check(cudaStreamSynchronize(stream), "sync");
  switch (host.common.format) {
  case kFmtLZ4: {
    LZ4FormatSpecHeader spec;
    std::memcpy(&spec, host.spec, sizeof(spec));
    return std::make_shared<LZ4Manager>(host.common.uncomp_chunk_size, spec.data_type, stream, device_id);
  }
  case kFmtSnappy:
    return std::make_shared<SnappyManager>(host.common.uncomp_chunk_size, stream, device_id);
  case kFmtCascaded: {
    CascadedFormatSpecHeader spec;
    std::memcpy(&spec, host.spec, sizeof(spec));
    return std::make_shared<CascadedManager>(spec.options, stream, device_id);
  }
  case kFmtANS:
  case kFmtGDeflate:
  case kFmtBitcomp:
    throw std::runtime_error("hipcomp create_manager: ANS / GDeflate / Bitcomp containers are not supported by this build");
  default:
    throw std::runtime_error("hipcomp create_manager: unknown container format " + std::to_string((int)host.common.format));
  }
}
} // namespace hipcomp
