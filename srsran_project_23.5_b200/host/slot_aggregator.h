// Internal to the "cuda" PUSCH decoder: the slot aggregator.
//
// The reference decodes one PUSCH per executor thread (lib/phy/upper/uplink_processor_impl.cpp:155-171 hands every PDU to
// its own pusch_processor / pusch_decoder), so K decoder objects call decode() at about the same time, each with ONE
// transport block. On a GPU a launch per transport block leaves most of the 148 SMs idle (a 12-codeblock block occupies
// 6). The aggregator sits between K pusch_decoder_cuda objects of one factory and the C ABI:
//
//   * every caller copies its LLRs (pageable memory of the caller) into a slice of a page-locked staging block and adds
//     its codeblock descriptors to the block's batch - in parallel, on the callers' own threads;
//   * the first caller that finds the block unled becomes its LEADER (no extra thread, no hand-off for K = 1): it closes
//     the block, opens the next one for late arrivals, waits for the copies into its own block to finish and issues ONE
//     ldpc_cuda_dematch_decode_tb_batch for all transport blocks present (one kernel launch per (BG, Z) group + one
//     assembly launch);
//   * the other callers sleep until the leader has marked their block done, then copy their own results out of the
//     staging block; the block returns to the ring when the last of them has left.
//
// With nof_blocks = 3 two launches can be in flight while a third block fills (the copies of one overlap the kernels of
// the other); transport blocks that arrive while the device is busy batch up on their own, which is what turns K callers
// into device-sized launches.
//
// A batch is homogeneous in what the handle holds per call (soft-buffer arena, iteration limit); a caller that differs
// waits for the next block.
#pragma once

#include "../../include/ldpc_cuda.h"
#include "device_arena_index.h"

#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <vector>

namespace srsran {
namespace cuda_detail {

/// One transport block handed in by a decoder thread. The caller fills the first group of fields.
struct slot_request {
  unsigned                 nof_cbs = 0;
  const ldpc_cuda_cb_desc* descs   = nullptr; ///< in_offset relative to llrs, out_offset relative to cb_out
  const int8_t*            llrs    = nullptr;
  uint64_t                 nof_llrs = 0;
  uint8_t*                 cb_out   = nullptr; ///< in/out: messages of the codeblocks (prefilled with the stored ones)
  uint64_t                 cb_out_bytes = 0;
  ldpc_cuda_cb_result*     results      = nullptr;
  ldpc_cuda_tb_desc        tbd          = {};      ///< first_codeblock / tb_offset are set by the aggregator
  uint8_t*                 tb_out       = nullptr;
  uint64_t                 tb_out_bytes = 0;
  ldpc_cuda_tb_result*     tbr          = nullptr;
  unsigned                 iterations   = 6;
  bool                     new_data     = true;
  const void*              softbuffer_key = nullptr; ///< the caller's rx_softbuffer object (private arena only)
  void*                    pool_arena     = nullptr; ///< arena of a rx_softbuffer_pool_cuda; nullptr: the private arena
  uint64_t                 pool_arena_bytes = 0;     ///< (descs[].soft_offset are then relative to the process' slots)

  // aggregator state
  uint64_t in_off = 0, out_off = 0, tb_off = 0;
  unsigned first_cb = 0, tb_index = 0, first_slot = 0;
  bool     zero_slots = false;
  bool     done       = false;
};

struct slot_aggregator_stats {
  uint64_t nof_launches         = 0; ///< batched calls issued
  uint64_t nof_transport_blocks = 0; ///< transport blocks they carried
  uint64_t max_tbs_per_launch   = 0;
  // where the callers' time went, nanoseconds summed over all transport blocks (launch_ns: over all launches)
  uint64_t join_wait_ns = 0; ///< waiting for a place in the open block
  uint64_t stage_ns     = 0; ///< copying LLRs / descriptors into the staging block
  uint64_t launch_ns    = 0; ///< inside the batched C-ABI call (leaders only)
  uint64_t follow_ns    = 0; ///< from staged until done, leading included
  uint64_t drain_ns     = 0; ///< copying the results out
};

class slot_aggregator
{
public:
  /// \param max_cbs        codeblocks one launch can carry (all transport blocks present).
  /// \param arena_cbs      capacity of the private device arena, codeblocks.
  /// \param nof_blocks     staging blocks in the ring: 2 = one launch in flight, 3 = two.
  /// Returns nullptr when no usable device exists.
  static std::shared_ptr<slot_aggregator> create(int device, unsigned max_cbs, unsigned arena_cbs, unsigned nof_blocks);
  ~slot_aggregator();

  /// Decodes one transport block; blocks until its results are in the caller's buffers. Thread safe.
  void decode(slot_request& r);

  slot_aggregator_stats stats();

private:
  enum class phase { idle, open, running, draining };

  struct block {
    ldpc_cuda_handle* h = nullptr;
    // page-locked staging
    int8_t*              in      = nullptr;
    uint8_t*             out     = nullptr;
    uint8_t*             tb      = nullptr;
    ldpc_cuda_cb_desc*   descs   = nullptr;
    ldpc_cuda_cb_result* results = nullptr;
    ldpc_cuda_tb_desc*   tbds    = nullptr;
    ldpc_cuda_tb_result* tbrs    = nullptr;
    // batch under construction / in flight
    phase                      state = phase::idle;
    uint64_t                   gen   = 0;
    uint64_t                   in_fill = 0, out_fill = 0, tb_fill = 0;
    unsigned                   ncbs = 0, writers = 0, readers = 0;
    std::vector<slot_request*> reqs;
    void*                      arena      = nullptr; // what this batch works on
    uint64_t                   arena_bytes = 0;
    unsigned                   iterations  = 0;
    // what the handle currently holds
    void*    attached        = nullptr;
    unsigned set_iterations  = 0;
  };

  slot_aggregator() = default;
  bool lead(std::unique_lock<std::mutex>& lk, block& b);
  void launch(block& b);

  int      device = 0;
  unsigned cap_cbs = 0, arena_cbs = 0;
  uint64_t cap_in = 0, cap_out = 0, cap_tb = 0;

  std::mutex              m;
  std::condition_variable cv;
  std::vector<block>      blocks;
  unsigned                open_idx = 0;
  uint64_t                next_gen = 1;
  void*                   private_arena = nullptr; // created on first use
  arena_index             index{0};
  slot_aggregator_stats   counters;
};

} // namespace cuda_detail
} // namespace srsran
