// b200comp — Snappy high-level manager (replaces reference include/hipcomp/snappy.hpp:38-51).
#pragma once

#include <memory>

#include "hipcompManager.hpp"

namespace hipcomp {

struct SnappyFormatSpecHeader {
  // no options
};

struct SnappyManager : PimplManager {
  SnappyManager(size_t uncomp_chunk_size, hipStream_t user_stream = 0, int device_id = 0);
  ~SnappyManager();
};

} // namespace hipcomp
