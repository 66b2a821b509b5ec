// Shared device-side definitions of the B200 LDPC receive chain (graphs in constant memory, per-codeblock work
// descriptors, exact LLR arithmetic). Reference semantics are cited per function; paths are relative to the srsRAN
// Project 23.5 tree.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "ldpc_bg_tables.h"

namespace ldpc_b200 {

constexpr int LLR_MAX = 120;  // include/srsran/phy/upper/log_likelihood_ratio.h:234
constexpr int LLR_INF = 127;  // include/srsran/phy/upper/log_likelihood_ratio.h:239
constexpr int MAX_Z = 384;
constexpr int MAX_ROW_DEGREE = 19;  // BG1 rows 0-3 (ldpc_luts_impl.cpp:4383-4519)

// Base graph in CSR form for one lifting set, shifts already reduced modulo Z by the host for the launch.
struct GraphView {
  int M;        // check rows
  int K;        // information nodes bg_K (22 / 10)
  int n_full;   // 68 / 52
  int n_edges;  // 316 / 197
};

// Constant-memory copy of both base graphs (all 8 lifting sets). 2 * (47*2 + 316 + 8*316*2) bytes < 12 KB.
struct GraphTables {
  uint16_t row_ptr[2][LDPC_BG1_M + 1];
  uint8_t col[2][LDPC_BG1_EDGES];
  uint16_t shift[2][LDPC_NOF_LIFTING_SETS][LDPC_BG1_EDGES];
};

// Work item for the decode kernels: one codeblock. Mirrors ldpc_cuda_cb_desc after host validation.
struct CbWork {
  unsigned long long in_offset;    // bytes from the LLR base (decode-only) or soft-buffer arena base (fused)
  unsigned long long out_offset;   // bytes from the output base
  unsigned long long rm_offset;    // bytes from the rate-matched input base (dematch)
  unsigned in_length;              // LLRs handed to the decoder
  unsigned rm_length;              // E
  unsigned Nref;
  unsigned short nof_filler_bits;
  unsigned char crc_poly;          // 0 none, 1 CRC16, 2 CRC24A, 3 CRC24B
  unsigned char early_stop;
  unsigned char rv;
  unsigned char mod_order;
  unsigned char new_data;
  unsigned char skip_decode;
  unsigned result_index;           // index into the caller's result / descriptor order
  unsigned reserved;
};

struct CbResult {  // == ldpc_cuda_cb_result
  unsigned short iterations;
  unsigned char crc_ok;
  unsigned char flags;
};

// Launch-wide parameters of a (base graph, lifting size) group.
struct GroupParams {
  int bg;            // 1 / 2
  int Z;
  int lifting_set;   // 0..7
  int max_iterations;
  unsigned char scale_lut[LLR_MAX + 1];  // scaled magnitude for m in [0,120] (generic / AVX / NEON arithmetic)
  unsigned char pad[3];
};

// CRC constants for on-device early stop: byte table and x^(8 j) mod P, j in [0, 1056].
struct CrcTables {
  unsigned table[4][256];
  unsigned pow8[4][1057];
  unsigned poly[4];
  unsigned order[4];
};

// ---------------------------------------------------------------------------------------------------------------------
// Exact LLR arithmetic (lib/phy/upper/log_likelihood_ratio.cpp:38-85) specialised to the two uses the decoder makes.

// v2c = soft - c2v (ldpc_decoder_generic.cpp:30-44 -> operator-, log_likelihood_ratio.h:111). c2v is never infinite
// (it is a scaled minimum <= 120), so: infinite soft passes through, otherwise saturate at +-120; the a == -b special
// case (result 0) coincides with the plain difference.
__device__ __forceinline__ int llr_v2c(int soft, int c2v)
{
  int d = soft - c2v;
  d = max(-LLR_MAX, min(LLR_MAX, d));
  return (abs(soft) > LLR_MAX) ? soft : d;
}

// soft = promotion_sum(c2v, v2c) (ldpc_decoder_generic.cpp:118-130, log_likelihood_ratio.cpp:72-85): an infinite v2c
// passes through, otherwise |sum| > 120 is promoted to +-127.
__device__ __forceinline__ int llr_soft_update(int c2v, int v2c)
{
  int s = c2v + v2c;
  int p = (s > LLR_MAX) ? LLR_INF : ((s < -LLR_MAX) ? -LLR_INF : s);
  return (abs(v2c) > LLR_MAX) ? v2c : p;
}

// Saturating LLR sum of the rate dematcher (combine_softbits, ldpc_rate_dematcher_impl.cpp:113-123 -> operator+=,
// log_likelihood_ratio.cpp:55-70), full special-value rules.
__device__ __forceinline__ int llr_add(int a, int b)
{
  if (a == -b) {
    return 0;
  }
  if (abs(a) > LLR_MAX) {
    return a;
  }
  if (abs(b) > LLR_MAX) {
    return b;
  }
  return max(-LLR_MAX, min(LLR_MAX, a + b));
}

}  // namespace ldpc_b200
