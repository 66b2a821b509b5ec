// "cuda" variant of the srsRAN soft demodulator: host-side drop-in above ldpc_cuda_demodulate_soft / ldpc_cuda_descramble
// (include/ldpc_cuda.h). Compiled against the (read-only) srsRAN Project 23.5 headers; mirrors
//   demodulation_mapper::demodulate_soft   include/srsran/phy/upper/channel_modulation/demodulation_mapper.h:50-70
//   (created by channel_modulation_factory::create_demodulation_mapper, channel_modulation_factories.h:32-41;
//    reference implementation lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:36-106)
// so that pusch_demodulator_impl (lib/phy/upper/channel_processors/pusch_demodulator_impl.cpp:88-98) runs unmodified on
// top of the B200 kernel. The batched, device-resident entry (ldpc_cuda_demodulate_batch_device) additionally fuses the
// descrambling of pusch_demodulator_impl::descramble (:100-152) and leaves the LLRs in HBM for the fused decoder.
#pragma once

#include "srsran/phy/upper/channel_modulation/demodulation_mapper.h"
#include "srsran/phy/upper/channel_processors/pusch_demodulator.h"

#include <memory>

namespace srsran {

/// Creates a soft demodulator that runs on CUDA device \c device (sm_100 only). Returns \c nullptr when no usable device
/// exists; there is no CPU fallback. Results equal the reference's generic per-symbol code bit for bit.
std::unique_ptr<demodulation_mapper> create_demodulation_mapper_cuda(int device = 0);

/// pusch_demodulator_impl::descramble without repetition placeholders (c_init = rnti * 2^15 + n_id) on the device.
/// \c handle_owner must be a mapper created by create_demodulation_mapper_cuda (it owns the device context).
void descramble_cuda(demodulation_mapper&             handle_owner,
                     span<log_likelihood_ratio>       out,
                     span<const log_likelihood_ratio> in,
                     unsigned                         c_init);

/// pusch_demodulator_impl::descramble (lib/phy/upper/channel_processors/pusch_demodulator_impl.cpp:100-152) as a whole:
/// c_init = rnti * 2^15 + n_id and the repetition placeholders of \c config.placeholders for \c config.modulation and
/// \c config.nof_tx_layers, on the device.
void descramble_cuda(demodulation_mapper&                    handle_owner,
                     span<log_likelihood_ratio>              out,
                     span<const log_likelihood_ratio>        in,
                     const pusch_demodulator::configuration& config);

} // namespace srsran
