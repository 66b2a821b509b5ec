// "cuda" variants of the srsRAN LDPC decoder / rate dematcher factories: host-side drop-in above the C ABI of
// include/ldpc_cuda.h. Compiled against the (read-only) srsRAN Project 23.5 headers; it mirrors
//   create_ldpc_decoder_factory_sw / create_ldpc_rate_dematcher_factory_sw
//   (include/srsran/phy/upper/channel_coding/channel_coding_factories.h:52-77,
//    lib/phy/upper/channel_coding/channel_coding_factories.cpp:257-271)
// and returns the same std::shared_ptr<..._factory> types, so that pusch_decoder_factory_sw_configuration
// (include/srsran/phy/upper/channel_processors/channel_processor_factories.h:237-244), pusch_decoder_impl and
// ldpc_decoder_benchmark run unmodified on top of the B200 kernels.
#pragma once

#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"

#include <memory>

namespace srsran {

/// Creates an LDPC decoder factory whose decoders run on CUDA device \c device (sm_100 only).
/// Returns \c nullptr when no usable device exists - exactly what create_ldpc_decoder_factory_sw does for an unsupported
/// variant (channel_coding_factories.cpp:115), so callers keep their report_fatal_error_if_not checks. There is no CPU
/// fallback.
std::shared_ptr<ldpc_decoder_factory> create_ldpc_decoder_factory_cuda(int device = 0);

/// Creates an LDPC rate dematcher factory whose dematchers run on CUDA device \c device (sm_100 only).
std::shared_ptr<ldpc_rate_dematcher_factory> create_ldpc_rate_dematcher_factory_cuda(int device = 0);

} // namespace srsran
