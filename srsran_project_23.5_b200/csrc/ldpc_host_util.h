// Host-side helpers of the C ABI that need no GPU: validation, geometry, scaling / CRC table construction.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>

#include "../../include/ldpc_cuda.h"
#include "ldpc_bg_tables.h"

namespace ldpc_b200 {

// Lifting set index (0..7) of a lifting size, -1 if Z is none of the 51 sizes. Table look-up: the host entries validate
// tens of thousands of descriptors per call.
inline int lifting_set_of(int Z)
{
  struct Table {
    signed char set[385];
    Table()
    {
      for (signed char& v : set) {
        v = -1;
      }
      for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
        set[LDPC_LIFTING_SIZES[i]] = (signed char)LDPC_LIFTING_SET[i];
      }
    }
  };
  static const Table table;
  return (Z >= 0 && Z <= 384) ? table.set[Z] : -1;
}

inline unsigned bg_K(int bg) { return bg == 2 ? 10 : 22; }
inline unsigned bg_N_short(int bg) { return bg == 2 ? 50 : 66; }
inline unsigned bg_M(int bg) { return bg == 2 ? LDPC_BG2_M : LDPC_BG1_M; }

// Scaled check-to-variable magnitude for m in [0,120]:
//  GENERIC ldpc_decoder_generic.cpp:80-89  (int8)std::round(float(m) * sf)
//  AVX     avx2_support.h:65-106           sf >= .9999 ? m : (m * uint16(sf * 65536)) >> 16
//  NEON    neon_support.h:73-86            sf >= .9999 ? m : (m * uint8(sf * 256)) >> 8
inline void build_scale_lut(float sf, unsigned mode, unsigned char lut[121])
{
  for (int m = 0; m <= 120; ++m) {
    int v;
    switch (mode) {
      case LDPC_CUDA_SCALE_AVX:
        v = (sf >= .9999) ? m : (int)(((unsigned)m * (unsigned)(uint16_t)(sf * 65536.0f)) >> 16);
        break;
      case LDPC_CUDA_SCALE_NEON:
        v = (sf >= .9999) ? m : (int)(((unsigned)m * (unsigned)(uint8_t)(sf * 256.0f)) >> 8);
        break;
      default:
        v = (int)std::round((float)m * sf);
        break;
    }
    lut[m] = (unsigned char)v;
  }
}

inline void crc_params(int code, unsigned& poly, unsigned& order)
{
  switch (code) {
    case LDPC_CUDA_CRC16:
      poly = 0x11021, order = 16;
      break;
    case LDPC_CUDA_CRC24A:
      poly = 0x1864cfb, order = 24;
      break;
    case LDPC_CUDA_CRC24B:
      poly = 0x1800063, order = 24;
      break;
    default:
      poly = 0, order = 0;
      break;
  }
}

// Byte table of crc_calculator_lut_impl.cpp:41-61 and x^(8 j) mod P.
inline void build_crc_tables(int code, unsigned table[256], unsigned* pow8, unsigned npow)
{
  unsigned poly, order;
  crc_params(code, poly, order);
  if (order == 0) {
    std::memset(table, 0, 256 * sizeof(unsigned));
    for (unsigned j = 0; j != npow; ++j) {
      pow8[j] = 0;
    }
    return;
  }
  const unsigned top = 1U << order, mask = top - 1;
  for (unsigned i = 0; i != 256; ++i) {
    unsigned r = i << (order - 8);
    for (int k = 0; k != 8; ++k) {
      r <<= 1;
      if (r & top) {
        r ^= poly;
      }
    }
    table[i] = r & mask;
  }
  unsigned p = 1;
  for (unsigned j = 0; j != npow; ++j) {
    pow8[j] = p;
    for (int k = 0; k != 8; ++k) {
      p <<= 1;
      if (p & top) {
        p ^= poly;
      }
    }
    p &= mask;
  }
}

// Argument checks = the reference's asserts (ldpc_decoder_impl.cpp:33-83, ldpc_rate_dematcher_impl.cpp:43-100).
inline int validate_decode_desc(const ldpc_cuda_cb_desc& d)
{
  if ((d.base_graph != 1 && d.base_graph != 2) || lifting_set_of(d.lifting_size) < 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  const unsigned Z = d.lifting_size, K = bg_K(d.base_graph) * Z;
  if (d.crc_poly > LDPC_CUDA_CRC24B || d.nof_filler_bits >= K) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (d.in_length < K + 2 * Z || d.in_length > bg_N_short(d.base_graph) * Z) {
    return LDPC_CUDA_ERR_ARG;
  }
  return LDPC_CUDA_OK;
}

inline int validate_dematch_desc(const ldpc_cuda_cb_desc& d)
{
  if ((d.base_graph != 1 && d.base_graph != 2) || lifting_set_of(d.lifting_size) < 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  const unsigned Z = d.lifting_size, sys = (bg_K(d.base_graph) - 2) * Z, N = bg_N_short(d.base_graph) * Z;
  if (d.rv > 3 || d.nof_filler_bits >= sys) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (!(d.mod_order == 1 || d.mod_order == 2 || d.mod_order == 4 || d.mod_order == 6 || d.mod_order == 8) ||
      d.rm_length % d.mod_order != 0 || d.rm_length > 8448u * 35u || d.Nref > 25344u) {
    return LDPC_CUDA_ERR_ARG;
  }
  // The reference walks off the circular buffer when it ends inside the systematic part (unsigned underflow at
  // ldpc_rate_dematcher_impl.cpp:171); 3GPP never configures that, so it is rejected here.
  const unsigned Ncb = (d.Nref > 0 && d.Nref < N) ? d.Nref : N;
  if (Ncb <= sys) {
    return LDPC_CUDA_ERR_ARG;
  }
  return LDPC_CUDA_OK;
}

}  // namespace ldpc_b200
