// b200comp — host-buffer staging pipeline around the batched codecs (the reference has
// none: its CHANGELOG.md:97-98 leaves host <-> device staging to the caller).
//
// A caller with data in (pinned) host memory gets one call per direction.  The buffer is cut into slices that
// travel down three streams, so the H2D copy of one slice, the codec kernels of another and the D2H traffic of a
// third overlap (PCIe is full duplex).  The compressed slice leaves the device through the gather kernel, which
// packs the chunks straight into the caller's mapped host blob (zero-copy stores: no intermediate device blob,
// no host-side size round trip before the copy can be sized); the only host waits are one per call at the end
// (compress / decompress) or one per slice, other slices being queued behind it (round trip).
#include "lz_common.cuh"

#include <algorithm>
#include <vector>

#pragma GCC visibility push(default)
#include "hipcomp/b200_ext.h"
#include "hipcomp/cascaded.h"
#include "hipcomp/lz4.h"
#include "hipcomp/snappy.h"
#pragma GCC visibility pop

namespace b200comp {

constexpr int kStagerStreams = 3;

// in_ptrs / sizes of the regular layout: chunk i = [i * chunk, min((i + 1) * chunk, nbytes)) of `base`
__global__ void stager_layout_kernel(uint8_t* base, size_t chunk, size_t nbytes, size_t lo, size_t n, void** ptrs,
                                     size_t* sizes)
{
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t off = (lo + i) * chunk;
    ptrs[lo + i] = base + off;
    sizes[lo + i] = off + chunk <= nbytes ? chunk : nbytes - off;
  }
}

// offsets[lo + i] = *running + sum(sizes[lo .. lo + i)); *running += sum(sizes[lo .. lo + n)).  One CTA; the
// slices of a call run it in order (event chain), so *running is the byte position of the slice in the blob.
__global__ void __launch_bounds__(1024) stager_scan_kernel(const size_t* sizes, size_t lo, size_t n, size_t* offsets,
                                                           unsigned long long* running)
{
  __shared__ unsigned long long s_warp[32];
  __shared__ unsigned long long s_base, s_tile;
  if (threadIdx.x == 0) s_base = *running;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (size_t t0 = 0; t0 < n; t0 += 1024) {
    const size_t i = t0 + threadIdx.x;
    const unsigned long long v = i < n ? sizes[lo + i] : 0ull;
    unsigned long long incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long o = __shfl_up_sync(kFull, incl, d);
      if ((int)lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const unsigned long long w = s_warp[lane];
      unsigned long long wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(kFull, wi, d);
        if ((int)lane >= d) wi += o;
      }
      s_warp[lane] = wi - w; // exclusive sum of the warp totals
      if (lane == 31) s_tile = wi;
    }
    __syncthreads();
    if (i < n) offsets[lo + i] = s_base + s_warp[warp] + incl - v;
    __syncthreads();
    if (threadIdx.x == 0) s_base += s_tile;
    __syncthreads();
  }
  if (threadIdx.x == 0) *running = s_base;
}

// comp_ptrs[lo + i] = blob + offsets[lo + i]
__global__ void stager_blob_ptrs_kernel(uint8_t* blob, const size_t* offsets, size_t lo, size_t n, void** ptrs)
{
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    ptrs[lo + i] = blob + offsets[lo + i];
}

} // namespace b200comp

using namespace b200comp;

struct b200compStager {
  int codec = 0;
  size_t chunk = 0, max_bytes = 0, max_chunks = 0, stride = 0;
  int slices = 4, device = 0;
  hipcompBatchedLZ4Opts_t lz4_opts;
  hipcompBatchedSnappyOpts_t snappy_opts;
  hipcompBatchedCascadedOpts_t casc_opts;
  cudaStream_t streams[2 * kStagerStreams] = {}; // [0, 3): compress direction, [3, 6): decompress direction
  std::vector<cudaEvent_t> slice_done, scan_done;
  uint8_t *d_raw = nullptr, *d_back = nullptr, *d_comp = nullptr, *d_blob = nullptr, *d_temp = nullptr, *d_dtemp = nullptr;
  void **d_raw_ptrs = nullptr, **d_back_ptrs = nullptr, **d_comp_ptrs = nullptr, **d_blob_ptrs = nullptr;
  size_t *d_back_sizes = nullptr, *d_raw_sizes = nullptr, *d_comp_sizes = nullptr, *d_offsets = nullptr, *d_actual = nullptr,
         *d_blob_sizes = nullptr;
  hipcompStatus_t* d_status = nullptr;
  unsigned long long* d_running = nullptr;
  size_t temp_region = 0, dtemp_bytes = 0; // per stream
  size_t* h_offsets = nullptr; // pinned scratch: blob offsets computed on the host for decompression
};

namespace {

const char* kFn = "b200compStager";

#define STAGER_CUDA(call)                                                                          \
  do {                                                                                             \
    const cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess)                                                                         \
      return fail(hipcompErrorCudaError, kFn, std::string(#call ": ") + cudaGetErrorString(e_));   \
  } while (0)
#define STAGER_OK(call)                                                                            \
  do {                                                                                             \
    const hipcompStatus_t s_ = (call);                                                             \
    if (s_ != hipcompSuccess) return s_;                                                           \
  } while (0)

hipcompStatus_t codec_sizes(const b200compStager* s, size_t n, size_t* temp, size_t* dtemp, size_t* max_out)
{
  switch (s->codec) {
  case 0:
    STAGER_OK(hipcompBatchedLZ4CompressGetTempSize(n, s->chunk, s->lz4_opts, temp));
    STAGER_OK(hipcompBatchedLZ4DecompressGetTempSize(n, s->chunk, dtemp));
    return hipcompBatchedLZ4CompressGetMaxOutputChunkSize(s->chunk, s->lz4_opts, max_out);
  case 1:
    STAGER_OK(hipcompBatchedSnappyCompressGetTempSize(n, s->chunk, s->snappy_opts, temp));
    STAGER_OK(hipcompBatchedSnappyDecompressGetTempSize(n, s->chunk, dtemp));
    return hipcompBatchedSnappyCompressGetMaxOutputChunkSize(s->chunk, s->snappy_opts, max_out);
  default:
    STAGER_OK(hipcompBatchedCascadedCompressGetTempSize(n, s->chunk, s->casc_opts, temp));
    STAGER_OK(hipcompBatchedCascadedDecompressGetTempSize(n, s->chunk, dtemp));
    return hipcompBatchedCascadedCompressGetMaxOutputChunkSize(s->chunk, s->casc_opts, max_out);
  }
}

hipcompStatus_t codec_compress(b200compStager* s, size_t lo, size_t n, int which, cudaStream_t st)
{
  void* temp = s->d_temp + (size_t)which * s->temp_region;
  const size_t temp_bytes = s->temp_region;
  const void* const* in = s->d_raw_ptrs + lo;
  void* const* out = s->d_comp_ptrs + lo;
  switch (s->codec) {
  case 0:
    return hipcompBatchedLZ4CompressAsync(in, s->d_raw_sizes + lo, s->chunk, n, temp, temp_bytes, out, s->d_comp_sizes + lo,
                                          s->lz4_opts, st);
  case 1:
    return hipcompBatchedSnappyCompressAsync(in, s->d_raw_sizes + lo, s->chunk, n, temp, temp_bytes, out,
                                             s->d_comp_sizes + lo, s->snappy_opts, st);
  default:
    return hipcompBatchedCascadedCompressAsync(in, s->d_raw_sizes + lo, s->chunk, n, temp, temp_bytes, out,
                                               s->d_comp_sizes + lo, s->casc_opts, st);
  }
}

hipcompStatus_t codec_decompress(b200compStager* s, size_t lo, size_t n, int which, cudaStream_t st)
{
  void* dtemp = s->d_dtemp + (size_t)which * s->dtemp_bytes;
  const void* const* in = s->d_blob_ptrs + lo;
  void* const* out = s->d_back_ptrs + lo;
  switch (s->codec) {
  case 0:
    return hipcompBatchedLZ4DecompressAsync(in, s->d_blob_sizes + lo, s->d_back_sizes + lo, s->d_actual + lo, n, dtemp,
                                            s->dtemp_bytes, out, s->d_status + lo, st);
  case 1:
    return hipcompBatchedSnappyDecompressAsync(in, s->d_blob_sizes + lo, s->d_back_sizes + lo, s->d_actual + lo, n, dtemp,
                                               s->dtemp_bytes, out, s->d_status + lo, st);
  default:
    return cascaded::decompress_known(s->casc_opts, in, s->d_blob_sizes + lo, s->d_back_sizes + lo, s->d_actual + lo, n, out,
                                      s->d_status + lo, st);
  }
}

inline int grid_for(size_t n) { return (int)std::min<size_t>((n + 255) / 256, 1024); }

// slice k of a call over nchunks chunks
inline void slice_range(const b200compStager* s, size_t nchunks, int k, size_t* lo, size_t* hi)
{
  *lo = nchunks * (size_t)k / (size_t)s->slices;
  *hi = nchunks * (size_t)(k + 1) / (size_t)s->slices;
}

// H2D raw slice -> CompressAsync -> offsets -> chunks packed into the mapped host blob, sizes to the host
hipcompStatus_t enqueue_compress_slice(b200compStager* s, int k, const uint8_t* h_raw, size_t nbytes, size_t nchunks,
                                       uint8_t* h_blob, size_t* h_sizes)
{
  size_t lo, hi;
  slice_range(s, nchunks, k, &lo, &hi);
  const size_t n = hi - lo;
  cudaStream_t st = s->streams[k % kStagerStreams];
  if (n) {
    const size_t b0 = lo * s->chunk, b1 = std::min(hi * s->chunk, nbytes);
    STAGER_CUDA(cudaMemcpyAsync(s->d_raw + b0, h_raw + b0, b1 - b0, cudaMemcpyHostToDevice, st));
    stager_layout_kernel<<<grid_for(n), 256, 0, st>>>(s->d_raw, s->chunk, nbytes, lo, n, s->d_raw_ptrs, s->d_raw_sizes);
    STAGER_OK(codec_compress(s, lo, n, k % kStagerStreams, st));
  }
  if (k > 0) STAGER_CUDA(cudaStreamWaitEvent(st, s->scan_done[k - 1], 0));
  stager_scan_kernel<<<1, 1024, 0, st>>>(s->d_comp_sizes, lo, n, s->d_offsets, s->d_running);
  STAGER_CUDA(cudaEventRecord(s->scan_done[k], st));
  if (n) {
    STAGER_OK(b200compGatherChunksAsync(s->d_comp_ptrs + lo, s->d_comp_sizes + lo, s->d_offsets + lo, h_blob, n, st));
    STAGER_CUDA(cudaMemcpyAsync(h_sizes + lo, s->d_comp_sizes + lo, n * sizeof(size_t), cudaMemcpyDeviceToHost, st));
  }
  STAGER_CUDA(cudaEventRecord(s->slice_done[k], st));
  return check_launch(kFn);
}

// H2D blob slice + sizes -> DecompressAsync -> D2H raw slice + statuses.  blob offsets of the slice's chunks are
// in s->h_offsets (host prefix sum of h_sizes).
hipcompStatus_t enqueue_decompress_slice(b200compStager* s, int k, const uint8_t* h_blob, const size_t* h_sizes,
                                         size_t nbytes, size_t nchunks, uint8_t* h_out, hipcompStatus_t* h_status)
{
  size_t lo, hi;
  slice_range(s, nchunks, k, &lo, &hi);
  const size_t n = hi - lo;
  if (!n) return hipcompSuccess;
  // its own stream set: the H2D copy of a later slice's input must not queue behind this slice's D2H copy
  cudaStream_t st = s->streams[kStagerStreams + k % kStagerStreams];
  const size_t c0 = s->h_offsets[lo], c1 = s->h_offsets[hi];
  STAGER_CUDA(cudaMemcpyAsync(s->d_blob + c0, h_blob + c0, c1 - c0, cudaMemcpyHostToDevice, st));
  STAGER_CUDA(cudaMemcpyAsync(s->d_blob_sizes + lo, h_sizes + lo, n * sizeof(size_t), cudaMemcpyHostToDevice, st));
  STAGER_CUDA(cudaMemcpyAsync(s->d_offsets + lo, s->h_offsets + lo, n * sizeof(size_t), cudaMemcpyHostToDevice, st));
  stager_blob_ptrs_kernel<<<grid_for(n), 256, 0, st>>>(s->d_blob, s->d_offsets, lo, n, s->d_blob_ptrs);
  stager_layout_kernel<<<grid_for(n), 256, 0, st>>>(s->d_back, s->chunk, nbytes, lo, n, s->d_back_ptrs, s->d_back_sizes);
  STAGER_OK(codec_decompress(s, lo, n, k % kStagerStreams, st));
  const size_t b0 = lo * s->chunk, b1 = std::min(hi * s->chunk, nbytes);
  STAGER_CUDA(cudaMemcpyAsync(h_out + b0, s->d_back + b0, b1 - b0, cudaMemcpyDeviceToHost, st));
  if (h_status)
    STAGER_CUDA(cudaMemcpyAsync(h_status + lo, s->d_status + lo, n * sizeof(hipcompStatus_t), cudaMemcpyDeviceToHost, st));
  return check_launch(kFn);
}

hipcompStatus_t check_call(const b200compStager* s, size_t nbytes)
{
  if (!s) return fail(hipcompErrorInvalidValue, kFn, "null stager");
  if (nbytes > s->max_bytes) return fail(hipcompErrorInvalidValue, kFn, "buffer larger than the stager's max_bytes");
  STAGER_CUDA(cudaSetDevice(s->device));
  return hipcompSuccess;
}

} // namespace

extern "C" {

hipcompStatus_t b200compHostAlloc(size_t bytes, void** out)
{
  B200_CHECK_NOT_NULL(out, "b200compHostAlloc()");
  const cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable | cudaHostAllocMapped);
  if (e != cudaSuccess) return fail(hipcompErrorCudaError, "b200compHostAlloc()", cudaGetErrorString(e));
  return hipcompSuccess;
}

void b200compHostFree(void* p)
{
  if (p) cudaFreeHost(p);
}

void b200compStagerDestroy(b200compStager* s)
{
  if (!s) return;
  cudaSetDevice(s->device);
  for (auto st : s->streams)
    if (st) cudaStreamSynchronize(st);
  for (void* p : {(void*)s->d_raw, (void*)s->d_back, (void*)s->d_comp, (void*)s->d_blob, (void*)s->d_temp, (void*)s->d_dtemp,
                  (void*)s->d_raw_ptrs, (void*)s->d_back_ptrs, (void*)s->d_comp_ptrs, (void*)s->d_blob_ptrs,
                  (void*)s->d_raw_sizes, (void*)s->d_back_sizes, (void*)s->d_comp_sizes, (void*)s->d_offsets,
                  (void*)s->d_actual, (void*)s->d_blob_sizes, (void*)s->d_status, (void*)s->d_running})
    if (p) cudaFree(p);
  if (s->h_offsets) cudaFreeHost(s->h_offsets);
  for (auto e : s->slice_done) cudaEventDestroy(e);
  for (auto e : s->scan_done) cudaEventDestroy(e);
  for (auto st : s->streams)
    if (st) cudaStreamDestroy(st);
  delete s;
}

hipcompStatus_t b200compStagerCreate(int codec, size_t chunk_bytes, size_t max_bytes, const void* opts, int slices,
                                     b200compStager** out)
{
  B200_CHECK_NOT_NULL(out, kFn);
  if (codec < 0 || codec > 2 || chunk_bytes == 0 || max_bytes == 0)
    return fail(hipcompErrorInvalidValue, kFn, "codec must be 0 (LZ4), 1 (Snappy) or 2 (Cascaded); sizes must be positive");
  b200compStager* s = new b200compStager;
  s->codec = codec;
  s->chunk = chunk_bytes;
  s->max_bytes = max_bytes;
  s->max_chunks = (max_bytes + chunk_bytes - 1) / chunk_bytes;
  s->slices = (int)std::max<size_t>(1, std::min<size_t>(slices > 0 ? slices : 8, s->max_chunks));
  s->lz4_opts = hipcompBatchedLZ4DefaultOpts;
  s->snappy_opts = hipcompBatchedSnappyDefaultOpts;
  s->casc_opts = hipcompBatchedCascadedDefaultOpts;
  if (opts) {
    if (codec == 0) s->lz4_opts = *static_cast<const hipcompBatchedLZ4Opts_t*>(opts);
    if (codec == 1) s->snappy_opts = *static_cast<const hipcompBatchedSnappyOpts_t*>(opts);
    if (codec == 2) s->casc_opts = *static_cast<const hipcompBatchedCascadedOpts_t*>(opts);
  }
  cudaGetDevice(&s->device);
  *out = nullptr;
  const size_t slice_chunks = (s->max_chunks + s->slices - 1) / s->slices + 1;
  size_t max_out = 0;
  hipcompStatus_t rc = codec_sizes(s, slice_chunks, &s->temp_region, &s->dtemp_bytes, &max_out);
  if (rc != hipcompSuccess) {
    delete s;
    return rc;
  }
  s->temp_region = (s->temp_region + 255) / 256 * 256 + 256;
  s->dtemp_bytes = (s->dtemp_bytes + 255) / 256 * 256 + 256;
  s->stride = (max_out + 15) / 16 * 16;
  const size_t nc = s->max_chunks;
  const size_t raw_bytes = nc * chunk_bytes + 64;
  bool ok = true;
  auto dev = [&](auto** p, size_t bytes) { ok = ok && cudaMalloc(reinterpret_cast<void**>(p), bytes ? bytes : 16) == cudaSuccess; };
  dev(&s->d_raw, raw_bytes);
  dev(&s->d_back, raw_bytes);
  dev(&s->d_comp, nc * s->stride);
  dev(&s->d_blob, nc * s->stride + 64);
  dev(&s->d_temp, kStagerStreams * s->temp_region);
  dev(&s->d_dtemp, kStagerStreams * s->dtemp_bytes);
  dev(&s->d_raw_ptrs, nc * sizeof(void*));
  dev(&s->d_back_ptrs, nc * sizeof(void*));
  dev(&s->d_comp_ptrs, nc * sizeof(void*));
  dev(&s->d_blob_ptrs, nc * sizeof(void*));
  dev(&s->d_raw_sizes, nc * sizeof(size_t));
  dev(&s->d_back_sizes, nc * sizeof(size_t));
  dev(&s->d_comp_sizes, nc * sizeof(size_t));
  dev(&s->d_offsets, (nc + 1) * sizeof(size_t));
  dev(&s->d_actual, nc * sizeof(size_t));
  dev(&s->d_blob_sizes, nc * sizeof(size_t));
  dev(&s->d_status, nc * sizeof(hipcompStatus_t));
  dev(&s->d_running, sizeof(unsigned long long));
  ok = ok && cudaHostAlloc(reinterpret_cast<void**>(&s->h_offsets), (nc + 1) * sizeof(size_t), cudaHostAllocPortable) == cudaSuccess;
  for (auto& st : s->streams) ok = ok && cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess;
  s->slice_done.resize(s->slices);
  s->scan_done.resize(s->slices);
  for (auto& e : s->slice_done) ok = ok && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
  for (auto& e : s->scan_done) ok = ok && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    const cudaError_t e = cudaGetLastError();
    b200compStagerDestroy(s);
    return fail(hipcompErrorCudaError, kFn, std::string("allocation failed: ") + cudaGetErrorString(e));
  }
  // strided compress output slots
  stager_layout_kernel<<<grid_for(nc), 256, 0, s->streams[0]>>>(s->d_comp, s->stride, nc * s->stride, 0, nc, s->d_comp_ptrs,
                                                                s->d_actual /* scratch: slot sizes, unused */);
  cudaStreamSynchronize(s->streams[0]);
  *out = s;
  return check_launch(kFn);
}

static hipcompStatus_t begin_call(b200compStager* s)
{
  STAGER_CUDA(cudaMemsetAsync(s->d_running, 0, sizeof(unsigned long long), s->streams[0]));
  cudaEvent_t e = s->scan_done[0];
  STAGER_CUDA(cudaEventRecord(e, s->streams[0]));
  for (int i = 1; i < 2 * kStagerStreams; ++i) STAGER_CUDA(cudaStreamWaitEvent(s->streams[i], e, 0));
  return hipcompSuccess;
}

static hipcompStatus_t end_call(b200compStager* s)
{
  for (auto st : s->streams) STAGER_CUDA(cudaStreamSynchronize(st));
  return hipcompSuccess;
}

static void host_offsets(b200compStager* s, const size_t* h_sizes, size_t lo, size_t hi)
{
  size_t acc = s->h_offsets[lo];
  for (size_t i = lo; i < hi; ++i) {
    s->h_offsets[i] = acc;
    acc += h_sizes[i];
  }
  s->h_offsets[hi] = acc;
}

hipcompStatus_t b200compStagerCompress(b200compStager* s, const void* host_raw, size_t nbytes, void* host_blob,
                                       size_t* host_comp_sizes, size_t* total_comp_bytes)
{
  STAGER_OK(check_call(s, nbytes));
  B200_CHECK_NOT_NULL(host_raw, kFn);
  B200_CHECK_NOT_NULL(host_blob, kFn);
  B200_CHECK_NOT_NULL(host_comp_sizes, kFn);
  const size_t nchunks = (nbytes + s->chunk - 1) / s->chunk;
  STAGER_OK(begin_call(s));
  for (int k = 0; k < s->slices; ++k)
    STAGER_OK(enqueue_compress_slice(s, k, static_cast<const uint8_t*>(host_raw), nbytes, nchunks,
                                     static_cast<uint8_t*>(host_blob), host_comp_sizes));
  STAGER_OK(end_call(s));
  if (total_comp_bytes) {
    size_t t = 0;
    for (size_t i = 0; i < nchunks; ++i) t += host_comp_sizes[i];
    *total_comp_bytes = t;
  }
  return hipcompSuccess;
}

hipcompStatus_t b200compStagerDecompress(b200compStager* s, const void* host_blob, const size_t* host_comp_sizes,
                                         size_t nchunks, void* host_raw_out, size_t nbytes,
                                         hipcompStatus_t* host_statuses)
{
  STAGER_OK(check_call(s, nbytes));
  B200_CHECK_NOT_NULL(host_blob, kFn);
  B200_CHECK_NOT_NULL(host_comp_sizes, kFn);
  B200_CHECK_NOT_NULL(host_raw_out, kFn);
  if (nchunks != (nbytes + s->chunk - 1) / s->chunk)
    return fail(hipcompErrorInvalidValue, kFn, "nchunks does not match nbytes / chunk_bytes");
  s->h_offsets[0] = 0;
  for (int k = 0; k < s->slices; ++k) {
    size_t lo, hi;
    slice_range(s, nchunks, k, &lo, &hi);
    host_offsets(s, host_comp_sizes, lo, hi);
    if (s->h_offsets[hi] > nchunks * s->stride) return fail(hipcompErrorInvalidValue, kFn, "compressed sizes exceed the stager's blob");
    STAGER_OK(enqueue_decompress_slice(s, k, static_cast<const uint8_t*>(host_blob), host_comp_sizes, nbytes, nchunks,
                                       static_cast<uint8_t*>(host_raw_out), host_statuses));
  }
  return end_call(s);
}

hipcompStatus_t b200compStagerRoundTrip(b200compStager* s, const void* host_raw, size_t nbytes, void* host_blob,
                                        size_t* host_comp_sizes, void* host_raw_out, hipcompStatus_t* host_statuses,
                                        size_t* total_comp_bytes)
{
  STAGER_OK(check_call(s, nbytes));
  B200_CHECK_NOT_NULL(host_raw, kFn);
  B200_CHECK_NOT_NULL(host_blob, kFn);
  B200_CHECK_NOT_NULL(host_comp_sizes, kFn);
  B200_CHECK_NOT_NULL(host_raw_out, kFn);
  const size_t nchunks = (nbytes + s->chunk - 1) / s->chunk;
  STAGER_OK(begin_call(s));
  s->h_offsets[0] = 0;
  const int depth = std::min(s->slices, kStagerStreams) - 1; // slices in flight in front of the host
  for (int k = 0; k < s->slices + depth; ++k) {
    if (k < s->slices)
      STAGER_OK(enqueue_compress_slice(s, k, static_cast<const uint8_t*>(host_raw), nbytes, nchunks,
                                       static_cast<uint8_t*>(host_blob), host_comp_sizes));
    const int j = k - depth;
    if (j >= 0) {
      // the compressed slice (and its sizes) is on the host once its event has fired; later slices are queued
      STAGER_CUDA(cudaEventSynchronize(s->slice_done[j]));
      size_t lo, hi;
      slice_range(s, nchunks, j, &lo, &hi);
      host_offsets(s, host_comp_sizes, lo, hi);
      STAGER_OK(enqueue_decompress_slice(s, j, static_cast<const uint8_t*>(host_blob), host_comp_sizes, nbytes, nchunks,
                                         static_cast<uint8_t*>(host_raw_out), host_statuses));
    }
  }
  STAGER_OK(end_call(s));
  if (total_comp_bytes) *total_comp_bytes = s->h_offsets[nchunks];
  return hipcompSuccess;
}

} // extern "C"
