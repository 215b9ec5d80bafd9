// b200comp — manager factory (replaces reference include/hipcomp/hipcompManagerFactory.hpp and
// src/highlevel/hipcompManagerFactory.cpp:65-148): reads the CommonHeader of an HLIF container
// (blocking D2H copy of header + format header) and builds the manager able to decode it.
// Formats: LZ4 (0), Snappy (1), Cascaded (4); ANS / GDeflate / Bitcomp containers throw, as the
// reference's stubs do.
#pragma once

#include <memory>

#include "cascaded.hpp"
#include "hipcompManager.hpp"
#include "lz4.hpp"
#include "snappy.hpp"

namespace hipcomp {

std::shared_ptr<hipcompManagerBase> create_manager(const uint8_t* comp_buffer, hipStream_t stream = 0,
                                                   const int device_id = 0);

} // namespace hipcomp
