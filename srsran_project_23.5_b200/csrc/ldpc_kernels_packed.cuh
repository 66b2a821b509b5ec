// Packed (2 codeblocks per thread, half2 arithmetic) decoder kernel - placeholder until the kernel lands.
#pragma once
#include "ldpc_device.cuh"
namespace ldpc_b200 {
enum { PACKED_SCALE_NONE = 0 };
static inline int packed_scale_variant(const unsigned char*) { return PACKED_SCALE_NONE; }
static inline int launch_packed_decode(const GroupParams&, int, const CbWork*, unsigned, const int8_t*, uint8_t*,
                                       CbResult*, int, cudaStream_t)
{
  return 1;
}
}  // namespace ldpc_b200
