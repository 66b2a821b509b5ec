// Generic layered normalised min-sum decoder kernel: any base graph, any of the 51 lifting sizes, any scaling table.
// One thread block per codeblock, one thread per lifted lane (check lane t of every layer), int32 arithmetic.
// It is the general GPU path (small / odd Z, non-default scaling) and the in-tree cross-check of the packed kernel.
//
// Formulation (exactly equivalent to lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:116-135 with the generic
// kernels of ldpc_decoder_generic.cpp): the reference keeps every check-to-variable message (46 x 27 x Z bytes); here
// each (layer, lane) keeps the compressed form {scaled min1, scaled min2, argmin edge, sign bit per edge}, from which
// the message on edge e is  sign_e ? -(e == argmin ? q2 : q1) : (e == argmin ? q2 : q1)  - exact because the scaling
// is applied after the min1/min2 selection (ldpc_decoder_generic.cpp:106-114). "First visit" of a layer
// (is_check_to_var_initialised, ldpc_decoder_impl.cpp:181-185) is the all-zero state (c2v = 0 => v2c = soft).
#pragma once

#include "ldpc_device.cuh"

namespace ldpc_b200 {

__constant__ GraphTables c_graphs;
__constant__ CrcTables c_crc;

struct LayerState {  // 8 bytes per (layer, lane)
  unsigned char q1;   // scaled first minimum
  unsigned char q2;   // scaled second minimum
  unsigned char idx;  // edge (adjacency order) that holds the first minimum
  unsigned char pad;
  unsigned signs;     // bit e: check-to-variable message on edge e is negative
};

// CRC of the first nbits of an MSB-first packed bit string held in shared memory, computed by one warp.
// crc_calculator_lut_impl.cpp:121-153: byte-table walk, initial value 0, non-byte tail == bit-serial remainder.
// Lane k walks its own contiguous chunk of whole bytes; chunk remainders are advanced by x^(bits that follow) mod P
// (24-step shift-and-add) and XOR-combined - valid because the CRC has no initial value or final XOR (linear).
__device__ __forceinline__ unsigned crc_mulmod(unsigned a, unsigned b, unsigned poly, unsigned order)
{
  // (a(x) * b(x)) mod P, operands below x^order.
  unsigned r = 0;
  for (int i = (int)order - 1; i >= 0; --i) {
    r <<= 1;
    if (r >> order) {
      r ^= poly;
    }
    if ((b >> i) & 1U) {
      r ^= a;
    }
  }
  return r;
}

__device__ unsigned crc_warp(const unsigned char* bits, unsigned nbits, int poly_code, const unsigned* s_table)
{
  const unsigned lane = threadIdx.x & 31;
  const unsigned poly = c_crc.poly[poly_code], order = c_crc.order[poly_code];
  const unsigned mask = (1U << order) - 1;
  const unsigned nbytes = nbits >> 3, tail = nbits & 7;
  const unsigned chunk = (nbytes + 31) / 32;
  unsigned b0 = min(lane * chunk, nbytes), b1 = min(b0 + chunk, nbytes);
  unsigned crc = 0;
  for (unsigned i = b0; i < b1; ++i) {
    unsigned idx = ((crc >> (order - 8)) & 0xffU) ^ bits[i];
    crc = ((crc << 8) ^ s_table[idx]) & mask;
  }
  // advance by the bytes that follow this chunk, then by the tail bits
  if (b1 > b0) {
    crc = crc_mulmod(crc, c_crc.pow8[poly_code][nbytes - b1], poly, order);
    for (unsigned k = 0; k < tail; ++k) {
      crc <<= 1;
      if (crc >> order) {
        crc ^= poly;
      }
    }
  } else {
    crc = 0;
  }
  if (lane == 31 && tail) {
    // tail bits of the message itself, bit-serial
    unsigned byte = bits[nbytes];
    unsigned t = 0;
    for (unsigned k = 0; k < tail; ++k) {
      unsigned bit = (byte >> (7 - k)) & 1U;
      t ^= bit << (order - 1);
      t <<= 1;
      if (t >> order) {
        t ^= poly;
      }
    }
    crc ^= t & mask;
  }
  for (int o = 16; o; o >>= 1) {
    crc ^= __shfl_xor_sync(0xffffffffu, crc, o);
  }
  return crc & mask;
}

// Shared memory: soft[n_full * Z] int8 | hard bits[ceil(K Z / 8)] | shift[edges] u16 | crc table 256 u32 |
// state[M * Z] LayerState (8-byte aligned first in the buffer).
__global__ void __launch_bounds__(384) ldpc_decode_scalar_kernel(GroupParams gp,
                                                                  const CbWork* __restrict__ work,
                                                                  unsigned nof_cbs,
                                                                  const int8_t* __restrict__ llr_base,
                                                                  uint8_t* __restrict__ out_base,
                                                                  CbResult* __restrict__ results)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int g = gp.bg - 1;
  const int Z = gp.Z;
  const int M = g ? LDPC_BG2_M : LDPC_BG1_M;
  const int K = g ? 10 : 22;
  const int n_full = K + M;
  const int n_edges = g ? LDPC_BG2_EDGES : LDPC_BG1_EDGES;
  const unsigned msg_len = (unsigned)(K * Z);
  const unsigned msg_bytes = (msg_len + 7) / 8;

  LayerState* state = reinterpret_cast<LayerState*>(smem);
  unsigned* s_crc_table = reinterpret_cast<unsigned*>(smem + (size_t)M * Z * sizeof(LayerState));
  unsigned short* s_shift = reinterpret_cast<unsigned short*>(s_crc_table + 256);
  signed char* soft = reinterpret_cast<signed char*>(s_shift + ((n_edges + 7) & ~7));
  unsigned char* hard = reinterpret_cast<unsigned char*>(soft + (((size_t)n_full * Z + 15) & ~(size_t)15));
  __shared__ unsigned s_last;
  __shared__ unsigned s_crc_ok;

  const int tid = threadIdx.x;
  for (int e = tid; e < n_edges; e += blockDim.x) {
    s_shift[e] = c_graphs.shift[g][gp.lifting_set][e] % Z;
  }

  for (unsigned cb = blockIdx.x; cb < nof_cbs; cb += gridDim.x) {
    const CbWork w = work[cb];
    const int8_t* in = llr_base + w.in_offset;
    uint8_t* out = out_base + w.out_offset;
    CbResult res = {0, 0, 0};
    if (w.skip_decode) {
      continue;
    }
    if (tid == 0) {
      s_last = 0;
    }
    __syncthreads();
    // load_soft_bits (ldpc_decoder_impl.cpp:148-173): two punctured nodes = 0, then the input; everything else 0.
    unsigned my_last = 0;
    for (unsigned i = tid; i < (unsigned)(n_full * Z); i += blockDim.x) {
      int v = 0;
      if (i >= 2u * Z && i - 2u * Z < w.in_length) {
        v = in[i - 2u * Z];
        if (v != 0) {
          my_last = i - 2u * Z + 1;
        }
      }
      soft[i] = (signed char)v;
    }
    for (unsigned i = tid; i < (unsigned)(M * Z); i += blockDim.x) {
      state[i] = LayerState{0, 0, 0, 0, 0};
    }
    if (w.crc_poly) {
      for (int i = tid; i < 256; i += blockDim.x) {
        s_crc_table[i] = c_crc.table[w.crc_poly][i];
      }
    }
    atomicMax(&s_last, my_last);
    __syncthreads();
    const unsigned input_size = s_last;  // one past the last non-zero LLR (ldpc_decoder_impl.cpp:86-94)
    if (input_size == 0) {
      if (!w.crc_poly) {  // output.one()
        for (unsigned i = tid; i < msg_bytes; i += blockDim.x) {
          out[i] = (i == msg_bytes - 1 && (msg_len & 7)) ? (unsigned char)(0xff << (8 - (msg_len & 7))) : 0xff;
        }
      }
      if (tid == 0) {
        res.flags = 1;
        results[w.result_index] = res;
      }
      __syncthreads();
      continue;
    }
    // ldpc_decoder_impl.cpp:105-114
    unsigned cb_len = max(input_size + 2u * Z, msg_len + 4u * Z);
    cb_len = ((cb_len + Z - 1) / Z) * Z;
    const int nof_layers = (int)(cb_len / Z) - K;
    const unsigned nsig = msg_len - w.nof_filler_bits;
    const bool check_each = w.crc_poly && w.early_stop;

    int it = 0;
    bool ok = false;
    for (; it < gp.max_iterations && !ok; ++it) {
      for (int layer = 0; layer < nof_layers; ++layer) {
        if (tid < Z) {
          const int e0 = c_graphs.row_ptr[g][layer], e1 = c_graphs.row_ptr[g][layer + 1];
          const LayerState st = state[layer * Z + tid];
          // pass 1 (analyze_var_to_check_msgs, ldpc_decoder_generic.cpp:53-78) in the check-lane frame
          int min1 = LLR_MAX, min2 = LLR_MAX, idx = 0;
          unsigned sgn = 0, vsign = 0;
          for (int e = e0; e < e1; ++e) {
            const int k = e - e0;
            int j = tid + s_shift[e];
            j -= (j >= Z) ? Z : 0;
            const int x = soft[c_graphs.col[g][e] * Z + j];
            int c = (k == st.idx) ? st.q2 : st.q1;
            c = ((st.signs >> k) & 1U) ? -c : c;
            const int v = llr_v2c(x, c);
            const int a = abs(v);
            const bool is_min = a < min1;
            const int new2 = is_min ? min1 : a;
            min2 = (a < min2) ? new2 : min2;
            idx = is_min ? k : idx;
            min1 = is_min ? a : min1;
            const unsigned neg = (v < 0) ? 1U : 0U;
            sgn ^= neg;
            vsign |= neg << k;
          }
          // scaling (scale_llr, ldpc_decoder_generic.cpp:80-89) through the host-built table
          const int q1 = gp.scale_lut[min1], q2 = gp.scale_lut[min2];
          // pass 2 (compute_check_to_var_msgs :93-116) + soft update (compute_soft_bits :118-130)
          unsigned nsigns = 0;
          for (int e = e0; e < e1; ++e) {
            const int k = e - e0;
            int j = tid + s_shift[e];
            j -= (j >= Z) ? Z : 0;
            const int addr = c_graphs.col[g][e] * Z + j;
            const int x = soft[addr];
            int c = (k == st.idx) ? st.q2 : st.q1;
            c = ((st.signs >> k) & 1U) ? -c : c;
            const int v = llr_v2c(x, c);
            const unsigned neg = (sgn ^ (vsign >> k)) & 1U;
            int cn = (k == idx) ? q2 : q1;
            cn = neg ? -cn : cn;
            nsigns |= neg << k;
            soft[addr] = (signed char)llr_soft_update(cn, v);
          }
          state[layer * Z + tid] = LayerState{(unsigned char)q1, (unsigned char)q2, (unsigned char)idx, 0, nsigns};
        }
        __syncthreads();
      }
      if (check_each || it + 1 == gp.max_iterations) {
        // get_hard_bits (ldpc_decoder_generic.cpp:132-138): bit = soft <= 0, MSB first.
        for (unsigned i = tid; i < msg_bytes; i += blockDim.x) {
          unsigned byte = 0;
          for (unsigned k = 0; k < 8; ++k) {
            const unsigned b = 8 * i + k;
            if (b < msg_len && soft[b] <= 0) {
              byte |= 0x80U >> k;
            }
          }
          hard[i] = (unsigned char)byte;
        }
        __syncthreads();
        if (w.crc_poly) {
          if (tid < 32) {
            const unsigned c = crc_warp(hard, nsig, w.crc_poly, s_crc_table);
            if (tid == 0) {
              s_crc_ok = (c == 0);
            }
          }
          __syncthreads();
          ok = s_crc_ok != 0;
        }
      }
    }
    // Output: the reference rewrites the bits after every checked iteration, so the last computed bits stand.
    // bit_buffer::insert keeps the bits of the last byte beyond the message as the caller left them
    for (unsigned i = tid; i < msg_bytes; i += blockDim.x) {
      unsigned char v = hard[i];
      if (i == msg_bytes - 1 && (msg_len & 7)) {
        const unsigned char keep = (unsigned char)(0xffU >> (msg_len & 7));
        v = (unsigned char)((v & ~keep) | (out[i] & keep));
      }
      out[i] = v;
    }
    if (tid == 0) {
      res.iterations = (unsigned short)it;
      res.crc_ok = ok ? 1 : 0;
      results[w.result_index] = res;
    }
    __syncthreads();
  }
}

static inline size_t scalar_kernel_smem_bytes(int bg, int Z)
{
  const int M = (bg == 2) ? LDPC_BG2_M : LDPC_BG1_M;
  const int K = (bg == 2) ? 10 : 22;
  const int n_edges = (bg == 2) ? LDPC_BG2_EDGES : LDPC_BG1_EDGES;
  size_t s = (size_t)M * Z * sizeof(LayerState) + 256 * 4 + (size_t)((n_edges + 7) & ~7) * 2;
  s += (((size_t)(K + M) * Z + 15) & ~(size_t)15);
  s += (((size_t)K * Z + 7) / 8 + 15) & ~(size_t)15;
  return s;
}

}  // namespace ldpc_b200
