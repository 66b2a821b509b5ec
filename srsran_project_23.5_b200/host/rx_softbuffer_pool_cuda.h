// Device-resident HARQ soft-buffer pool: the reference's rx_softbuffer_pool interface
//   (include/srsran/phy/upper/rx_softbuffer_pool.h:33-100)
// with the observable behaviour of rx_softbuffer_pool_impl / rx_softbuffer_impl / rx_softbuffer_codeblock_pool
//   (lib/phy/upper/rx_softbuffer_pool_impl.cpp:27-76, rx_softbuffer_impl.h:34-262, rx_softbuffer_codeblock_pool.h:35-117):
// reservation by (rnti, harq_ack_id), available -> reserved -> locked -> released, expiry in slots, new codeblock entries
// when the number of codeblocks changes (lowest free entries first), memory zero-initialised once and never re-zeroed.
// The implementation is a single reservation table under one mutex with an O(1) lowest-free-slot bitmap
// (rx_softbuffer_pool_cuda.cpp). The soft bits live in ONE device arena (25 344 bytes per codeblock entry) shared by every
// pusch_decoder_cuda of that device; the host side of an entry keeps only the CRC flag and the decoded message bits.
// get_codeblock_soft_bits therefore returns an empty span: this pool is for pusch_decoder_cuda, which finds the device
// location of each codeblock through rx_softbuffer_device (below). No 52-codeblock limit per transport block.
#pragma once

#include "srsran/phy/upper/rx_softbuffer_pool.h"

#include <cstdint>
#include <memory>

namespace srsran {

/// What pusch_decoder_cuda needs from a soft buffer whose soft bits are on the device.
class rx_softbuffer_device
{
public:
  virtual ~rx_softbuffer_device() = default;
  /// Device pointer and size in bytes of the arena all entries live in.
  virtual void*    arena() const       = 0;
  virtual uint64_t arena_bytes() const = 0;
  /// Byte offset inside the arena of the soft bits of codeblock \c codeblock_id of this buffer.
  virtual uint64_t soft_offset(unsigned codeblock_id) const = 0;
};

struct rx_softbuffer_pool_cuda_config {
  /// Same meaning as rx_softbuffer_pool_config (rx_softbuffer_pool.h:89-98); max_codeblock_size is fixed at 25 344.
  unsigned max_softbuffers      = 0;
  unsigned max_nof_codeblocks   = 0;
  unsigned expire_timeout_slots = 0;
  /// CUDA device ordinal; a negative value creates the pool without device memory (host logic only, for CPU tests).
  int device = 0;
};

/// Returns nullptr if the device arena cannot be allocated.
std::unique_ptr<rx_softbuffer_pool> create_rx_softbuffer_pool_cuda(const rx_softbuffer_pool_cuda_config& config);

/// What happened to reservations since the pool was created (the reference pool keeps no such record: a reservation that
/// expires, or that is refused, is invisible there).
struct rx_softbuffer_pool_cuda_stats {
  uint64_t reservations;      ///< successful reserve_softbuffer calls
  uint64_t refused_in_use;    ///< the buffer of that identifier was locked by another holder
  uint64_t refused_no_buffer; ///< no buffer of that identifier and none available
  uint64_t refused_no_slots;  ///< not enough free codeblock entries
  uint64_t expired;           ///< reservations whose entries run_slot took back after the timeout (HARQ state lost)
  uint64_t released;          ///< reservations freed by run_slot after release()
};

class rx_softbuffer_pool_cuda_introspection
{
public:
  virtual ~rx_softbuffer_pool_cuda_introspection()   = default;
  virtual rx_softbuffer_pool_cuda_stats stats() const = 0;
};

/// False when \c pool was not created by create_rx_softbuffer_pool_cuda.
bool get_rx_softbuffer_pool_cuda_stats(const rx_softbuffer_pool& pool, rx_softbuffer_pool_cuda_stats& out);

} // namespace srsran
