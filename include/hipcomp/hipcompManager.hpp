// b200comp — high-level interface (HLIF): one call compresses / decompresses a whole device
// buffer into / from a self-describing container.
//
// Same classes, members and call semantics as the reference's
// include/hipcomp/hipcompManager.hpp:66-301 (CompressionConfig :75-103, DecompressionConfig :108-136,
// hipcompManagerBase :146-233, PimplManager :238-301), so code written against the reference
// compiles unchanged.  The container layout is the reference's.
#pragma once

#include <cstddef>
#include <cstdint>
#include <memory>

#include "hipcomp.h"

namespace hipcomp {

template <typename T>
struct PinnedPtrPool;

// Result of configure_compression(): sizes the caller needs to allocate the output, plus a
// pinned-host status word the compression kernels write on error.
struct CompressionConfig {
private:
  struct CompressionConfigImpl;
  std::shared_ptr<CompressionConfigImpl> impl;

public:
  size_t uncompressed_buffer_size;
  size_t max_compressed_buffer_size;
  size_t num_chunks;

  CompressionConfig(PinnedPtrPool<hipcompStatus_t>& pool, size_t uncompressed_buffer_size);
  hipcompStatus_t* get_status() const;

  CompressionConfig(CompressionConfig&& other);
  CompressionConfig(const CompressionConfig& other);
  CompressionConfig& operator=(CompressionConfig&& other);
  CompressionConfig& operator=(const CompressionConfig& other);
  ~CompressionConfig();
};

// Result of configure_decompression(): output size and chunk count (filled asynchronously on the
// manager's stream when configured from a compressed buffer — synchronise before reading).
struct DecompressionConfig {
private:
  struct DecompressionConfigImpl;
  std::shared_ptr<DecompressionConfigImpl> impl;

public:
  size_t decomp_data_size;
  uint32_t num_chunks;

  DecompressionConfig(PinnedPtrPool<hipcompStatus_t>& pool);
  hipcompStatus_t* get_status() const;

  DecompressionConfig(DecompressionConfig&& other);
  DecompressionConfig(const DecompressionConfig& other);
  DecompressionConfig& operator=(DecompressionConfig&& other);
  DecompressionConfig& operator=(const DecompressionConfig& other);
  ~DecompressionConfig();
};

struct hipcompManagerBase {
  virtual CompressionConfig configure_compression(const size_t decomp_buffer_size) = 0;
  virtual void compress(const uint8_t* decomp_buffer, uint8_t* comp_buffer, const CompressionConfig& comp_config) = 0;
  virtual DecompressionConfig configure_decompression(const uint8_t* comp_buffer) = 0;
  virtual DecompressionConfig configure_decompression(const CompressionConfig& comp_config) = 0;
  virtual void decompress(uint8_t* decomp_buffer, const uint8_t* comp_buffer,
                          const DecompressionConfig& decomp_config) = 0;
  virtual void set_scratch_buffer(uint8_t* new_scratch_buffer) = 0;
  virtual size_t get_required_scratch_buffer_size() = 0;
  virtual size_t get_compressed_output_size(uint8_t* comp_buffer) = 0;
  virtual ~hipcompManagerBase() = default;
};

// The concrete managers (LZ4Manager, SnappyManager, CascadedManager) derive from this and only set
// `impl`; every call forwards.
struct PimplManager : hipcompManagerBase {
protected:
  std::unique_ptr<hipcompManagerBase> impl;

public:
  PimplManager() : impl(nullptr) {}
  PimplManager(const PimplManager&) = delete;
  PimplManager& operator=(const PimplManager&) = delete;
  virtual ~PimplManager() {}

  virtual CompressionConfig configure_compression(const size_t decomp_buffer_size)
  {
    return impl->configure_compression(decomp_buffer_size);
  }
  virtual void compress(const uint8_t* decomp_buffer, uint8_t* comp_buffer, const CompressionConfig& comp_config)
  {
    impl->compress(decomp_buffer, comp_buffer, comp_config);
  }
  virtual DecompressionConfig configure_decompression(const uint8_t* comp_buffer)
  {
    return impl->configure_decompression(comp_buffer);
  }
  virtual DecompressionConfig configure_decompression(const CompressionConfig& comp_config)
  {
    return impl->configure_decompression(comp_config);
  }
  virtual void decompress(uint8_t* decomp_buffer, const uint8_t* comp_buffer, const DecompressionConfig& decomp_config)
  {
    impl->decompress(decomp_buffer, comp_buffer, decomp_config);
  }
  virtual void set_scratch_buffer(uint8_t* new_scratch_buffer) { impl->set_scratch_buffer(new_scratch_buffer); }
  virtual size_t get_required_scratch_buffer_size() { return impl->get_required_scratch_buffer_size(); }
  virtual size_t get_compressed_output_size(uint8_t* comp_buffer)
  {
    return impl->get_compressed_output_size(comp_buffer);
  }
};

} // namespace hipcomp
