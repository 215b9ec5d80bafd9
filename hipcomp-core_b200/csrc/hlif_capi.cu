// b200comp — C view of the HLIF managers (include/hipcomp/b200_ext.h), for FFI callers and
// the Python test harness.  Thin: every function forwards to the C++ classes of hlif.cu.
#include "common.cuh"

#pragma GCC visibility push(default)
#include "hipcomp/b200_ext.h"
#include "hipcomp/hipcompManagerFactory.hpp"
#pragma GCC visibility pop

#include <exception>
#include <memory>
#include <stdexcept>

namespace hipcomp {
bool b200_set_checksum_mode(hipcompManagerBase* m, int mode); // hlif.cu
}

struct b200compManager {
  std::shared_ptr<hipcomp::hipcompManagerBase> m;
};
struct b200compCompConfig {
  hipcomp::CompressionConfig c;
};
struct b200compDecompConfig {
  hipcomp::DecompressionConfig c;
};

namespace {
template <typename F>
hipcompStatus_t guarded(const char* fn, F&& f)
{
  try {
    f();
    return hipcompSuccess;
  } catch (const std::invalid_argument& e) {
    return b200comp::fail(hipcompErrorInvalidValue, fn, e.what());
  } catch (const std::exception& e) {
    return b200comp::fail(hipcompErrorInternal, fn, e.what());
  }
}
} // namespace

extern "C" {

hipcompStatus_t b200compManagerCreateLZ4(size_t chunk, hipcompType_t type, hipStream_t stream, int dev, b200compManager** out)
{
  B200_CHECK_NOT_NULL(out, "b200compManagerCreateLZ4()");
  return guarded("b200compManagerCreateLZ4()", [&] { *out = new b200compManager{std::make_shared<hipcomp::LZ4Manager>(chunk, type, stream, dev)}; });
}
hipcompStatus_t b200compManagerCreateSnappy(size_t chunk, hipStream_t stream, int dev, b200compManager** out)
{
  B200_CHECK_NOT_NULL(out, "b200compManagerCreateSnappy()");
  return guarded("b200compManagerCreateSnappy()", [&] { *out = new b200compManager{std::make_shared<hipcomp::SnappyManager>(chunk, stream, dev)}; });
}
hipcompStatus_t b200compManagerCreateCascaded(const void* opts, hipStream_t stream, int dev, b200compManager** out)
{
  B200_CHECK_NOT_NULL(out, "b200compManagerCreateCascaded()");
  B200_CHECK_NOT_NULL(opts, "b200compManagerCreateCascaded()");
  return guarded("b200compManagerCreateCascaded()", [&] {
    *out = new b200compManager{std::make_shared<hipcomp::CascadedManager>(*static_cast<const hipcompBatchedCascadedOpts_t*>(opts), stream, dev)};
  });
}
hipcompStatus_t b200compManagerCreateFromBuffer(const void* buf, hipStream_t stream, int dev, b200compManager** out)
{
  B200_CHECK_NOT_NULL(out, "b200compManagerCreateFromBuffer()");
  B200_CHECK_NOT_NULL(buf, "b200compManagerCreateFromBuffer()");
  return guarded("b200compManagerCreateFromBuffer()",
                 [&] { *out = new b200compManager{hipcomp::create_manager(static_cast<const uint8_t*>(buf), stream, dev)}; });
}
void b200compManagerDestroy(b200compManager* m) { delete m; }

hipcompStatus_t b200compManagerSetChecksumMode(b200compManager* m, int mode)
{
  static const char* fn = "b200compManagerSetChecksumMode()";
  B200_CHECK_NOT_NULL(m, fn);
  if (mode != 0 && mode != 1) return b200comp::fail(hipcompErrorInvalidValue, fn, "mode must be 0 (none) or 1 (CRC-32 per chunk)");
  if (!hipcomp::b200_set_checksum_mode(m->m.get(), mode))
    return b200comp::fail(hipcompErrorNotSupported, fn, "not a manager of this library");
  return hipcompSuccess;
}

hipcompStatus_t b200compManagerConfigureCompression(b200compManager* m, size_t bytes, b200compCompConfig** out,
                                                    size_t* max_compressed_bytes, size_t* num_chunks)
{
  static const char* fn = "b200compManagerConfigureCompression()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(out, fn);
  return guarded(fn, [&] {
    *out = new b200compCompConfig{m->m->configure_compression(bytes)};
    if (max_compressed_bytes) *max_compressed_bytes = (*out)->c.max_compressed_buffer_size;
    if (num_chunks) *num_chunks = (*out)->c.num_chunks;
  });
}
hipcompStatus_t b200compManagerCompress(b200compManager* m, const void* in, void* out, const b200compCompConfig* cfg)
{
  static const char* fn = "b200compManagerCompress()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(cfg, fn);
  return guarded(fn, [&] { m->m->compress(static_cast<const uint8_t*>(in), static_cast<uint8_t*>(out), cfg->c); });
}
hipcompStatus_t b200compManagerGetCompressedOutputSize(b200compManager* m, void* comp, size_t* bytes)
{
  static const char* fn = "b200compManagerGetCompressedOutputSize()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(bytes, fn);
  return guarded(fn, [&] { *bytes = m->m->get_compressed_output_size(static_cast<uint8_t*>(comp)); });
}
hipcompStatus_t b200compManagerConfigureDecompression(b200compManager* m, const void* comp, b200compDecompConfig** out,
                                                      size_t* decomp_bytes, size_t* num_chunks)
{
  static const char* fn = "b200compManagerConfigureDecompression()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(out, fn);
  return guarded(fn, [&] {
    *out = new b200compDecompConfig{m->m->configure_decompression(static_cast<const uint8_t*>(comp))};
    if (decomp_bytes) *decomp_bytes = (*out)->c.decomp_data_size;
    if (num_chunks) *num_chunks = (*out)->c.num_chunks;
  });
}
hipcompStatus_t b200compManagerConfigureDecompressionFromComp(b200compManager* m, const b200compCompConfig* comp,
                                                              b200compDecompConfig** out, size_t* decomp_bytes,
                                                              size_t* num_chunks)
{
  static const char* fn = "b200compManagerConfigureDecompressionFromComp()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(comp, fn);
  B200_CHECK_NOT_NULL(out, fn);
  return guarded(fn, [&] {
    *out = new b200compDecompConfig{m->m->configure_decompression(comp->c)};
    if (decomp_bytes) *decomp_bytes = (*out)->c.decomp_data_size;
    if (num_chunks) *num_chunks = (*out)->c.num_chunks;
  });
}
hipcompStatus_t b200compManagerDecompress(b200compManager* m, void* out, const void* comp, const b200compDecompConfig* cfg)
{
  static const char* fn = "b200compManagerDecompress()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(cfg, fn);
  return guarded(fn, [&] { m->m->decompress(static_cast<uint8_t*>(out), static_cast<const uint8_t*>(comp), cfg->c); });
}
hipcompStatus_t b200compManagerGetRequiredScratchBufferSize(b200compManager* m, size_t* bytes)
{
  static const char* fn = "b200compManagerGetRequiredScratchBufferSize()";
  B200_CHECK_NOT_NULL(m, fn);
  B200_CHECK_NOT_NULL(bytes, fn);
  return guarded(fn, [&] { *bytes = m->m->get_required_scratch_buffer_size(); });
}
hipcompStatus_t b200compManagerSetScratchBuffer(b200compManager* m, void* scratch)
{
  static const char* fn = "b200compManagerSetScratchBuffer()";
  B200_CHECK_NOT_NULL(m, fn);
  return guarded(fn, [&] { m->m->set_scratch_buffer(static_cast<uint8_t*>(scratch)); });
}
hipcompStatus_t b200compCompConfigStatus(const b200compCompConfig* cfg) { return cfg ? *cfg->c.get_status() : hipcompErrorInvalidValue; }
hipcompStatus_t b200compDecompConfigStatus(const b200compDecompConfig* cfg) { return cfg ? *cfg->c.get_status() : hipcompErrorInvalidValue; }
void b200compCompConfigDestroy(b200compCompConfig* cfg) { delete cfg; }
void b200compDecompConfigDestroy(b200compDecompConfig* cfg) { delete cfg; }

} // extern "C"
