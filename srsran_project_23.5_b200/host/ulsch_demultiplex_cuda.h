// "cuda" variant of the UL-SCH demultiplexer: the reference's ulsch_demultiplex interface
//   (include/srsran/phy/upper/channel_processors/ulsch_demultiplex.h:37-103; reference implementation
//    lib/phy/upper/channel_processors/ulsch_demultiplex_impl.cpp:316-453, created by
//    ulsch_demultiplex_factory::create, channel_processor_factories.h)
// above ldpc_cuda_ulsch_demultiplex / ldpc_cuda_ulsch_placeholders (include/ldpc_cuda.h), so that pusch_processor_impl
// (lib/phy/upper/channel_processors/pusch_processor_impl.cpp:259) runs unmodified on top of the B200 kernel. The batched,
// device-resident entry (ldpc_cuda_ulsch_demultiplex_batch_device) is what keeps the LLRs of a PUSCH with UCI in HBM
// between the descrambler and the rate dematcher.
#pragma once

#include "srsran/phy/upper/channel_processors/ulsch_demultiplex.h"

#include <memory>

namespace srsran {

/// Creates a demultiplexer that runs on CUDA device \c device (sm_100 only). Returns \c nullptr when no usable device
/// exists; there is no CPU fallback. get_placeholders is host arithmetic (closed form, no device work).
std::unique_ptr<ulsch_demultiplex> create_ulsch_demultiplex_cuda(int device = 0);

} // namespace srsran
