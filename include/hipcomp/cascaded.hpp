// b200comp — Cascaded high-level manager (replaces reference include/hipcomp/cascaded.hpp:55-66).
#pragma once

#include "cascaded.h"
#include "hipcompManager.hpp"

namespace hipcomp {

struct CascadedFormatSpecHeader {
  hipcompBatchedCascadedOpts_t options;
};

// Each `options.chunk_size` slice of the input becomes one batched-Cascaded partition.
struct CascadedManager : PimplManager {
  CascadedManager(const hipcompBatchedCascadedOpts_t& options = hipcompBatchedCascadedDefaultOpts,
                  hipStream_t user_stream = 0, int device_id = 0);
  virtual ~CascadedManager();
};

} // namespace hipcomp
