// b200comp — stream-ordered workspace (replaces the reference's TempSpaceBroker,
// src/TempSpaceBroker.h:59-113 / .cpp:64-109).
//
// The reference carves typed, aligned regions out of one caller-provided buffer with a
// bump pointer and throws when the buffer is exhausted.  `Workspace` keeps that contract
// (reserve<T>(n) / next() / spaceLeft(), same alignment rule: alignof(T), and at least
// `min_align`), and adds what the reference leaves to its callers: the backing buffer can
// be OWNED by the workspace and obtained from the device's stream-ordered memory pool
// (cudaMallocAsync / cudaFreeAsync on the caller's stream), so a manager that allocates
// its scratch on first use never synchronises the device and the memory is recycled by
// a PRIVATE per-device pool across calls (its release threshold is raised so freed scratch stays
// cached; the application's default pool is never touched).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <stdexcept>
#include <string>

namespace b200comp {

class Workspace
{
public:
  // Non-owning view over [ptr, ptr + bytes): the TempSpaceBroker use case.
  Workspace(void* ptr, size_t bytes, size_t min_align = 16)
      : base_(static_cast<uint8_t*>(ptr)), cur_(static_cast<uint8_t*>(ptr)), bytes_(bytes), min_align_(min_align)
  {
  }

  // Owning: `bytes` from the stream-ordered pool of the current device.
  Workspace(size_t bytes, cudaStream_t stream, size_t min_align = 16) : bytes_(bytes), min_align_(min_align)
  {
    acquire(bytes, stream);
  }

  Workspace() = default;
  Workspace(const Workspace&) = delete;
  Workspace& operator=(const Workspace&) = delete;

  ~Workspace() { release(); }

  // (Re)allocate an owned backing buffer on `stream`; any previous owned buffer is freed
  // in stream order first.
  void acquire(size_t bytes, cudaStream_t stream)
  {
    release();
    void* p = nullptr;
    cudaMemPool_t pool = private_pool();
    cudaError_t e = pool ? cudaMallocFromPoolAsync(&p, bytes ? bytes : 1, pool, stream)
                         : cudaMallocAsync(&p, bytes ? bytes : 1, stream);
    if (e != cudaSuccess) {
      cudaGetLastError();
      throw std::runtime_error(std::string("Workspace: cudaMallocAsync(") + std::to_string(bytes)
                               + ") failed: " + cudaGetErrorString(e));
    }
    base_ = cur_ = static_cast<uint8_t*>(p);
    bytes_ = bytes;
    stream_ = stream;
    owned_ = true;
  }

  // Adopt a caller-owned buffer (set_scratch_buffer): frees an owned one in stream order.
  void adopt(void* ptr, size_t bytes)
  {
    release();
    base_ = cur_ = static_cast<uint8_t*>(ptr);
    bytes_ = bytes;
    owned_ = false;
  }

  void release()
  {
    if (owned_ && base_) cudaFreeAsync(base_, stream_);
    base_ = cur_ = nullptr;
    owned_ = false;
  }

  // Start carving again from the beginning (one call = one layout).
  void rewind() { cur_ = base_; }

  // n objects of T, aligned to max(alignof(T), min_align).  Throws std::runtime_error
  // when the buffer cannot hold them (reference TempSpaceBroker.cpp:87-109).
  template <typename T>
  T* reserve(size_t n)
  {
    const size_t a = alignof(T) > min_align_ ? alignof(T) : min_align_;
    const uintptr_t p = (reinterpret_cast<uintptr_t>(cur_) + a - 1) / a * a;
    const size_t need = n * sizeof(T);
    const uintptr_t end = reinterpret_cast<uintptr_t>(base_) + bytes_;
    if (base_ == nullptr || p > end || need > end - p)
      throw std::runtime_error("Workspace: temp space exhausted: need " + std::to_string(need) + " bytes, "
                               + std::to_string(spaceLeft()) + " of " + std::to_string(bytes_) + " left");
    cur_ = reinterpret_cast<uint8_t*>(p + need);
    return reinterpret_cast<T*>(p);
  }

  void* next() const { return cur_; }
  size_t spaceLeft() const { return base_ ? bytes_ - (size_t)(cur_ - base_) : 0; }
  size_t size() const { return bytes_; }
  void* data() const { return base_; }
  bool owned() const { return owned_; }
  bool valid() const { return base_ != nullptr; }

  // Bytes reserve<T>(n) consumes in the worst case (for sizing a layout up front).
  template <typename T>
  static size_t footprint(size_t n, size_t min_align = 16)
  {
    const size_t a = alignof(T) > min_align ? alignof(T) : min_align;
    return n * sizeof(T) + a;
  }

private:
  // The library's own stream-ordered pool on the current device (created on first use, kept for the life of
  // the process).  Freed scratch stays cached in it; the host application's default pool keeps its settings.
  static cudaMemPool_t private_pool()
  {
    constexpr int kMaxDevices = 64;
    static cudaMemPool_t pools[kMaxDevices] = {};
    static bool tried[kMaxDevices] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
    if (!tried[dev]) {
      tried[dev] = true;
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = dev;
      cudaMemPool_t pool = nullptr;
      if (cudaMemPoolCreate(&pool, &props) == cudaSuccess) {
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        pools[dev] = pool;
      }
      cudaGetLastError();
    }
    return pools[dev];
  }

  uint8_t* base_ = nullptr;
  uint8_t* cur_ = nullptr;
  size_t bytes_ = 0;
  size_t min_align_ = 16;
  cudaStream_t stream_ = nullptr;
  bool owned_ = false;
};

} // namespace b200comp
