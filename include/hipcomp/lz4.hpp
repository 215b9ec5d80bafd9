// b200comp — LZ4 high-level manager (replaces reference include/hipcomp/lz4.hpp:58-67).
#pragma once

#include <memory>

#include "hipcompManager.hpp"

namespace hipcomp {

// Written after the CommonHeader in every LZ4 HLIF container.
struct LZ4FormatSpecHeader {
  hipcompType_t data_type;
};

// Splits a buffer into `uncomp_chunk_size` chunks, LZ4-compresses them as one batch and writes the
// HLIF container.  `data_type` selects the match granularity (char / short / int).
struct LZ4Manager : PimplManager {
  LZ4Manager(size_t uncomp_chunk_size, hipcompType_t data_type, hipStream_t user_stream = 0,
             const int device_id = 0);
  ~LZ4Manager();
};

} // namespace hipcomp
