// "cuda" PUSCH decoder: the reference's pusch_decoder interface
//   (include/srsran/phy/upper/channel_processors/pusch_decoder.h:37-75)
// implemented with ONE batched, fused call per transport block (ldpc_cuda_dematch_decode_batch: rate dematch + HARQ
// combine in an HBM-resident soft-buffer arena -> decode -> CRC early stop) instead of the per-codeblock loop of
//   pusch_decoder_impl::decode (lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:121-225).
// What that loop does around the per-codeblock calls stays here, on the host, with the same observable behaviour:
// codeblock CRC flags and decoded messages live in the caller's rx_softbuffer, codeblocks already decoded are combined
// but not decoded again, the transport block is assembled from the messages, checked with CRC24A when it has several
// codeblocks, and all codeblock flags are reset when that check fails. Differences by design:
//   * the soft bits are NOT kept in rx_softbuffer::get_codeblock_soft_bits but in device memory, keyed by the
//     rx_softbuffer object (25 KB per codeblock never cross PCIe again);
//   * no 52-codeblock limit (MAX_NOF_SEGMENTS): geometry comes from ldpc_cuda_segment;
//   * TB assembly copies exactly nof_data_bits per codeblock (the reference's srsvec::copy_offset drops the last
//     (nof_data_bits % 8) bits of the first codeblock, lib/srsvec/bit.cpp:333-358 - never triggered by 3GPP TB sizes).
#pragma once

#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include "srsran/phy/upper/channel_processors/channel_processor_factories.h"

#include <cstdint>
#include <memory>

namespace srsran {

struct pusch_decoder_cuda_configuration {
  /// CRC calculators for the transport-block CRC24A (any reference CRC factory).
  std::shared_ptr<crc_calculator_factory> crc_factory;
  /// CUDA device ordinal (sm_100 only).
  int device = 0;
  /// Capacity of the device soft-buffer arena in codeblocks (25 344 bytes each). When it is exhausted the least recently
  /// used HARQ process loses its soft bits, which is what a soft-buffer pool expiry does.
  unsigned arena_codeblocks = 4096;
  /// Largest number of codeblocks of one transport block.
  unsigned max_codeblocks_per_tb = 256;
  /// 0: every decoder owns a device handle and launches its transport block alone. > 0: the decoders of this factory
  /// share a SLOT AGGREGATOR (slot_aggregator.h): transport blocks of decoder threads that are inside decode() at the same
  /// time - the reference runs one PUSCH per executor thread, uplink_processor_impl.cpp:155-171 - are staged into one
  /// page-locked block by their own threads and launched together, up to this many codeblocks per launch. decode() stays
  /// synchronous for every caller.
  unsigned aggregate_codeblocks = 0;
  /// Staging blocks of the aggregator: 2 = one launch in flight while the next batch collects, 3 = two in flight.
  unsigned aggregate_blocks = 3;
};

/// Returns nullptr when no usable device exists (no CPU fallback), like the reference's factories for an unsupported type.
std::shared_ptr<pusch_decoder_factory> create_pusch_decoder_factory_cuda(const pusch_decoder_cuda_configuration& config);

/// What the slot aggregator of a factory has done so far (all zero for a factory without one).
struct pusch_decoder_cuda_aggregator_stats {
  uint64_t nof_launches;
  uint64_t nof_transport_blocks;
  uint64_t max_transport_blocks_per_launch;
  /// Nanoseconds summed over all transport blocks: waiting for a staging block, copying in, from staged until done,
  /// copying out; launch_ns is summed over the launches (time inside the batched C-ABI call).
  uint64_t join_wait_ns, stage_ns, follow_ns, drain_ns, launch_ns;
};
pusch_decoder_cuda_aggregator_stats get_pusch_decoder_cuda_aggregator_stats(pusch_decoder_factory& factory);

/// HARQ processes whose device soft bits were dropped because a decoder's PRIVATE arena (used with the reference's host
/// pool) was full, since the process started. With create_rx_softbuffer_pool_cuda nothing is ever dropped: reservations are
/// refused like the reference's, and rx_softbuffer_pool_cuda_stats counts them. A decoder keys its private arena by the
/// caller's rx_softbuffer object: when the host pool hands the same object to another (rnti, harq) with the same codeblock
/// count and the first PDU says new_data = false, the old soft bits are combined - as the reference does with the host
/// memory of a re-used buffer, which it never clears either (rx_softbuffer_codeblock_pool.h).
uint64_t pusch_decoder_cuda_private_arena_evictions();

} // namespace srsran
