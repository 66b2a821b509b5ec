/* TEST INFRASTRUCTURE ONLY - see ldpc_oracle.h. Plain-C restatement of the reference PUSCH LDPC receive chain.
 *
 * Every function cites the reference file:line it follows (paths relative to the reference root). The structure is
 * deliberately the reference's own (explicit per-edge check-to-variable storage, two passes per layer), NOT the compressed
 * formulation the CUDA kernels use, so that the two are independent statements of the same arithmetic.
 */
#include "ldpc_oracle.h"
#include "ldpc_bg_tables.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define LLR_MAX 120 /* include/srsran/phy/upper/log_likelihood_ratio.h:234 */
#define LLR_INF 127 /* include/srsran/phy/upper/log_likelihood_ratio.h:239 */

/* ---------------------------------------------------------------------------------------------------------------------
 * LLR arithmetic: lib/phy/upper/log_likelihood_ratio.cpp:38-85.
 */
static int llr_isinf(int v)
{
  return abs(v) > LLR_MAX;
}

/* tackle_special_sums, log_likelihood_ratio.cpp:38-53. Returns 1 and sets *r when the sum is special. */
static int special_sum(int a, int b, int* r)
{
  if (a == -b) {
    *r = 0;
    return 1;
  }
  if (llr_isinf(a)) {
    *r = a;
    return 1;
  }
  if (llr_isinf(b)) {
    *r = b;
    return 1;
  }
  return 0;
}

/* operator+=, log_likelihood_ratio.cpp:55-70. */
int oracle_llr_add(int a, int b)
{
  int r;
  if (special_sum(a, b, &r)) {
    return r;
  }
  int t = a + b;
  if (abs(t) > LLR_MAX) {
    return (t > 0) ? LLR_MAX : -LLR_MAX;
  }
  return t;
}

/* operator-, log_likelihood_ratio.h:111: a + (-b). */
int oracle_llr_sub(int a, int b)
{
  return oracle_llr_add(a, -b);
}

/* promotion_sum, log_likelihood_ratio.cpp:72-85. */
int oracle_llr_promotion_sum(int a, int b)
{
  int r;
  if (special_sum(a, b, &r)) {
    return r;
  }
  int t = a + b;
  if (abs(t) > LLR_MAX) {
    return (t > 0) ? LLR_INF : -LLR_INF;
  }
  return t;
}

/* hard_decision, log_likelihood_ratio.cpp:225-248 + to_hard_bit log_likelihood_ratio.h:86: bit = (llr <= 0), MSB first.
 * The bits are written with bit_buffer::insert (adt/bit_buffer.h:95-140), which keeps the bits of the last byte that lie
 * beyond the n-th bit as the caller left them. */
void oracle_hard_decision(const int8_t* soft, unsigned n, uint8_t* out_packed)
{
  unsigned nbytes = (n + 7) / 8;
  for (unsigned i = 0; i != nbytes; ++i) {
    uint8_t  byte = 0;
    unsigned lim  = (n - 8 * i < 8) ? (n - 8 * i) : 8;
    for (unsigned k = 0; k != lim; ++k) {
      if (soft[8 * i + k] <= 0) {
        byte |= (uint8_t)(0x80U >> k);
      }
    }
    if (lim < 8) {
      uint8_t keep = (uint8_t)(0xffU >> lim);
      byte         = (uint8_t)(byte | (out_packed[i] & keep));
    }
    out_packed[i] = byte;
  }
}

/* ---------------------------------------------------------------------------------------------------------------------
 * CRC: lib/phy/upper/channel_coding/crc_calculator_lut_impl.cpp:33-39 (polynomials), :121-153 (calculate over a
 * bit_buffer). The byte-LUT walk followed by reversecrcbit() (crc_calculator_lut_impl.h:88-100) is the plain MSB-first
 * bit-serial remainder over exactly nbits bits, initial value 0, no final XOR.
 */
unsigned oracle_crc_packed(int poly_code, const uint8_t* packed, unsigned nbits)
{
  unsigned poly, order;
  switch (poly_code) {
    case 1:
      poly  = 0x11021;
      order = 16;
      break;
    case 2:
      poly  = 0x1864cfb;
      order = 24;
      break;
    case 3:
      poly  = 0x1800063;
      order = 24;
      break;
    default:
      return 0xffffffffU;
  }
  unsigned top = 1U << order;
  unsigned crc = 0;
  for (unsigned i = 0; i != nbits; ++i) {
    unsigned bit = (packed[i / 8] >> (7 - (i % 8))) & 1U;
    crc ^= bit << (order - 1);
    crc <<= 1;
    if (crc & top) {
      crc ^= poly;
    }
  }
  return crc & (top - 1);
}

/* ---------------------------------------------------------------------------------------------------------------------
 * Check-to-variable scaling.
 *  mode 0: generic, ldpc_decoder_generic.cpp:80-89: (int8)std::round(float(m) * sf).
 *  mode 1: AVX2 / AVX-512, avx2_support.h:65-106: (m * uint16(sf * 65536)) >> 16 (sf >= .9999 bypasses).
 *  mode 2: NEON, neon_support.h:73-86: (m * uint8(sf * 256)) >> 8.
 * m is a magnitude in [0, 120] (minima are initialised to LLR_MAX, ldpc_decoder_impl.cpp:239-240).
 */
int oracle_scale(int m, float sf, int mode)
{
  switch (mode) {
    case 0:
      return (int)roundf((float)m * sf);
    case 1:
      if (sf >= .9999) {
        return m;
      }
      return (int)(((unsigned)m * (unsigned)(uint16_t)(sf * 65536.0f)) >> 16);
    case 2:
      if (sf >= .9999) {
        return m;
      }
      return (int)(((unsigned)m * (unsigned)(uint8_t)(sf * 256.0f)) >> 8);
    default:
      return m;
  }
}

/* ---------------------------------------------------------------------------------------------------------------------
 * Graphs: ldpc_graph_impl.h:39-49 (dimensions), ldpc_luts_impl.cpp:4521-4544 (get_graph, "% ls").
 */
typedef struct {
  int             M, N_full, N_short, K;
  const uint16_t* row_ptr;
  const uint8_t*  col;
  const uint16_t* shift; /* for the lifting set of Z */
} graph_t;

static int lifting_set_of(int Z)
{
  for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
    if (LDPC_LIFTING_SIZES[i] == Z) {
      return LDPC_LIFTING_SET[i];
    }
  }
  return -1;
}

static int get_graph(int bg, int Z, graph_t* g)
{
  int set = lifting_set_of(Z);
  if (set < 0) {
    return -1;
  }
  if (bg == 1) {
    g->M       = LDPC_BG1_M;
    g->N_full  = LDPC_BG1_N_FULL;
    g->row_ptr = LDPC_BG1_ROW_PTR;
    g->col     = LDPC_BG1_COL;
    g->shift   = LDPC_BG1_SHIFT[set];
  } else if (bg == 2) {
    g->M       = LDPC_BG2_M;
    g->N_full  = LDPC_BG2_N_FULL;
    g->row_ptr = LDPC_BG2_ROW_PTR;
    g->col     = LDPC_BG2_COL;
    g->shift   = LDPC_BG2_SHIFT[set];
  } else {
    return -1;
  }
  g->N_short = g->N_full - 2;
  g->K       = g->N_full - g->M;
  return 0;
}

int oracle_graph_shift(int bg, int Z, int m, int n)
{
  graph_t g;
  if (get_graph(bg, Z, &g) != 0 || m < 0 || m >= g.M) {
    return -1;
  }
  for (unsigned e = g.row_ptr[m]; e != g.row_ptr[m + 1]; ++e) {
    if (g.col[e] == n) {
      return g.shift[e] % Z;
    }
  }
  return -1;
}

/* ---------------------------------------------------------------------------------------------------------------------
 * Decoder: ldpc_decoder_impl.cpp:33-297 with the generic kernels of ldpc_decoder_generic.cpp:30-138.
 */
int oracle_decode(int           bg,
                  int           Z,
                  unsigned      nof_filler_bits,
                  int           crc_poly_code,
                  unsigned      max_iterations,
                  float         scaling_factor,
                  int           scale_mode,
                  const int8_t* llrs,
                  unsigned      nof_llrs,
                  uint8_t*      out_packed)
{
  graph_t g;
  /* init(), ldpc_decoder_impl.cpp:33-58. */
  if (get_graph(bg, Z, &g) != 0) {
    return -1;
  }
  if (max_iterations == 0 || !(scaling_factor > 0 && scaling_factor < 1)) {
    return -1;
  }
  const unsigned uZ                   = (unsigned)Z;
  const unsigned hrr                  = (unsigned)g.K + 4; /* bg_N_high_rate */
  const unsigned message_length       = (unsigned)g.K * uZ;
  const unsigned max_input_length     = (unsigned)g.N_short * uZ;
  const unsigned min_input_length     = message_length + 2 * uZ;
  const unsigned nof_significant_bits = message_length - nof_filler_bits;
  /* decode() size checks, :69-83. */
  if (nof_llrs > max_input_length || nof_llrs < min_input_length || nof_filler_bits >= message_length) {
    return -2;
  }

  /* Trailing-zero trim, :86-94. */
  unsigned last = nof_llrs;
  while (last > 0 && llrs[last - 1] == 0) {
    --last;
  }
  if (last == 0) {
    if (crc_poly_code == 0) {
      memset(out_packed, 0xff, (message_length + 7) / 8); /* output.one() */
      if (message_length % 8) {
        /* bit_buffer::one() only sets the valid bits of the last word. */
        out_packed[message_length / 8] = (uint8_t)(0xff << (8 - message_length % 8));
      }
    }
    return 0;
  }
  unsigned input_size = last;

  /* Storage laid out as in ldpc_decoder_impl.h:175,204-213 with node_size_byte == Z (generic). */
  const unsigned slots = hrr + 1; /* hrr nodes + one extension node per layer (MAX_CHECK_NODE_DEGREE = 27 for BG1) */
  int8_t*        soft  = (int8_t*)calloc((size_t)g.N_full * uZ, 1);
  int8_t*        c2v   = (int8_t*)calloc((size_t)g.M * slots * uZ, 1);
  int8_t*        v2c   = (int8_t*)calloc((size_t)slots * uZ, 1);
  uint8_t*       init  = (uint8_t*)calloc((size_t)g.M, 1);
  int8_t*        min1  = (int8_t*)malloc(uZ);
  int8_t*        min2  = (int8_t*)malloc(uZ);
  uint8_t*       idx   = (uint8_t*)malloc(uZ);
  uint8_t*       sgn   = (uint8_t*)malloc(uZ);

  /* load_soft_bits(), :148-173: two punctured nodes = 0, then all nof_llrs inputs (not trimmed). */
  memcpy(soft + 2 * uZ, llrs, nof_llrs);

  /* :105-114. */
  unsigned codeblock_length = input_size + 2 * uZ;
  if (codeblock_length < message_length + 4 * uZ) {
    codeblock_length = message_length + 4 * uZ;
  }
  if (codeblock_length % uZ != 0) {
    codeblock_length = (codeblock_length / uZ + 1) * uZ;
  }
  unsigned nof_layers = codeblock_length / uZ - (unsigned)g.K;

  int result = 0;
  for (unsigned it = 0; it != max_iterations; ++it) {
    for (unsigned layer = 0; layer != nof_layers; ++layer) {
      int8_t* c2v_l = c2v + (size_t)layer * slots * uZ;

      /* update_variable_to_check_messages(), :175-200 + compute_var_to_check_msgs generic :30-44. */
      for (unsigned i = 0; i != hrr * uZ; ++i) {
        v2c[i] = init[layer] ? (int8_t)oracle_llr_sub(soft[i], c2v_l[i]) : soft[i];
      }
      if (layer >= 4) {
        unsigned off = (hrr + layer - 4) * uZ;
        for (unsigned j = 0; j != uZ; ++j) {
          v2c[hrr * uZ + j] = init[layer] ? (int8_t)oracle_llr_sub(soft[off + j], c2v_l[hrr * uZ + j]) : soft[off + j];
        }
      }

      /* update_check_to_variable_messages(), :220-297. */
      memset(min1, LLR_MAX, uZ);
      memset(min2, LLR_MAX, uZ);
      memset(idx, 0, uZ);
      memset(sgn, 0, uZ);
      unsigned e0 = g.row_ptr[layer], e1 = g.row_ptr[layer + 1];
      /* pass 1: analyze_var_to_check_msgs generic :53-78. */
      for (unsigned e = e0; e != e1; ++e) {
        unsigned var_node = e - e0;
        unsigned shift    = g.shift[e] % uZ;
        unsigned base     = (g.col[e] < hrr ? g.col[e] : hrr) * uZ;
        for (unsigned j = 0; j != uZ; ++j) {
          int      v      = v2c[base + j];
          int      a      = abs(v);
          unsigned t      = (j + uZ - shift) % uZ;
          int      is_min = a < min1[t];
          int      new2   = is_min ? min1[t] : a;
          if (a < min2[t]) {
            min2[t] = (int8_t)new2;
          }
          if (is_min) {
            min1[t] = (int8_t)a;
            idx[t]  = (uint8_t)var_node;
          }
          sgn[t] ^= (v >= 0) ? 0U : 1U;
        }
      }
      /* pass 2: compute_check_to_var_msgs generic :93-116. */
      for (unsigned e = e0; e != e1; ++e) {
        unsigned var_node = e - e0;
        unsigned shift    = g.shift[e] % uZ;
        unsigned base     = (g.col[e] < hrr ? g.col[e] : hrr) * uZ;
        for (unsigned j = 0; j != uZ; ++j) {
          unsigned t = (j + uZ - shift) % uZ;
          int      m = (var_node != idx[t]) ? min1[t] : min2[t];
          if (!llr_isinf(m)) {
            m = oracle_scale(m, scaling_factor, scale_mode);
          }
          unsigned final_sign = sgn[t] ^ ((v2c[base + j] >= 0) ? 0U : 1U);
          c2v_l[base + j]     = (int8_t)(final_sign ? -m : m);
        }
      }
      init[layer] = 1;

      /* update_soft_bits(), :202-218 + compute_soft_bits generic :118-130. */
      for (unsigned e = e0; e != e1; ++e) {
        unsigned base  = (g.col[e] < hrr ? g.col[e] : hrr) * uZ;
        unsigned sbase = (unsigned)g.col[e] * uZ;
        for (unsigned j = 0; j != uZ; ++j) {
          soft[sbase + j] = (int8_t)oracle_llr_promotion_sum(c2v_l[base + j], v2c[base + j]);
        }
      }
    }

    /* Early stop, :125-133. */
    if (crc_poly_code != 0) {
      oracle_hard_decision(soft, message_length, out_packed);
      if (oracle_crc_packed(crc_poly_code, out_packed, nof_significant_bits) == 0) {
        result = (int)(it + 1);
        break;
      }
    }
  }
  /* :138-145. */
  if (crc_poly_code == 0) {
    oracle_hard_decision(soft, message_length, out_packed);
  }

  free(soft);
  free(c2v);
  free(v2c);
  free(init);
  free(min1);
  free(min2);
  free(idx);
  free(sgn);
  return result;
}

/* ---------------------------------------------------------------------------------------------------------------------
 * Rate dematcher: ldpc_rate_dematcher_impl.cpp:43-110 (rate_dematch), :125-198 (allot_llrs), :113-123
 * (combine_softbits), :200-254 (deinterleave).
 */
int oracle_rate_dematch(int8_t*       out,
                        unsigned      out_len,
                        const int8_t* in,
                        unsigned      in_len,
                        int           new_data,
                        int           rv,
                        int           Qm,
                        unsigned      Nref,
                        unsigned      nof_filler_bits)
{
  unsigned block_length = out_len;
  /* :59-63 */
  unsigned buffer_length = (Nref > 0) ? (Nref < block_length ? Nref : block_length) : block_length;
  /* :75-96 */
  double   factors1[4] = {0, 17, 33, 56};
  double   factors2[4] = {0, 13, 25, 43};
  double*  factor;
  unsigned n_short, bg_K;
  if (block_length % 66 == 0) {
    factor  = factors1;
    n_short = 66;
    bg_K    = 22;
  } else if (block_length % 50 == 0) {
    factor  = factors2;
    n_short = 50;
    bg_K    = 10;
  } else {
    return -1;
  }
  unsigned Z = block_length / n_short;
  if (lifting_set_of((int)Z) < 0 || rv < 0 || rv > 3) {
    return -1;
  }
  if (!(Qm == 1 || Qm == 2 || Qm == 4 || Qm == 6 || Qm == 8) || in_len % (unsigned)Qm != 0) {
    return -1;
  }
  unsigned nof_systematic_bits = (bg_K - 2) * Z;
  if (nof_filler_bits >= nof_systematic_bits) {
    return -1;
  }
  /* :101-102 */
  double   tmp      = (factor[rv] * buffer_length) / block_length;
  unsigned shift_k0 = (unsigned)((uint16_t)floor(tmp)) * Z;

  /* :104-110 deinterleave, :200-210: aux[K*j + i] = in[i*Qm + j]. */
  int8_t*       aux = NULL;
  const int8_t* src = in;
  if (Qm > 1 && in_len > 0) {
    aux        = (int8_t*)malloc(in_len);
    unsigned K = in_len / (unsigned)Qm;
    for (unsigned idx = 0, i = 0; i != K; ++i) {
      for (unsigned j = 0; j != (unsigned)Qm; ++j, ++idx) {
        aux[K * j + i] = in[idx];
      }
    }
    src = aux;
  }

  /* allot_llrs(), :125-198, literally. */
  unsigned nof_info_bits = nof_systematic_bits - nof_filler_bits;
  int      is_copy_mode  = new_data;
  unsigned tmp_idx       = shift_k0;
  unsigned left          = in_len;
  while (left != 0) {
    if (tmp_idx < nof_info_bits) {
      unsigned n = nof_info_bits - tmp_idx;
      if (n > left) {
        n = left;
      }
      if (is_copy_mode) {
        memset(out, 0, tmp_idx);
        memcpy(out + tmp_idx, src, n);
      } else {
        for (unsigned i = 0; i != n; ++i) {
          out[tmp_idx + i] = (int8_t)oracle_llr_add(out[tmp_idx + i], src[i]);
        }
      }
      tmp_idx += n;
      src += n;
      left -= n;
    } else if (is_copy_mode) {
      memset(out, 0, nof_info_bits);
    }
    if (is_copy_mode) {
      memset(out + nof_info_bits, LLR_INF, nof_filler_bits);
    }
    if (tmp_idx < nof_systematic_bits) {
      tmp_idx = nof_systematic_bits;
    }
    unsigned n = buffer_length - tmp_idx;
    if (n > left) {
      n = left;
    }
    if (is_copy_mode) {
      memcpy(out + tmp_idx, src, n);
    } else {
      for (unsigned i = 0; i != n; ++i) {
        out[tmp_idx + i] = (int8_t)oracle_llr_add(out[tmp_idx + i], src[i]);
      }
    }
    tmp_idx = (tmp_idx + n) % buffer_length;
    src += n;
    left -= n;
    if (left != 0) {
      is_copy_mode = 0;
    }
  }
  if (is_copy_mode && tmp_idx != 0) {
    /* out.last(buffer_length - tmp_idx): the LAST bytes of the whole N-long span, :195-197. */
    unsigned n = buffer_length - tmp_idx;
    memset(out + (block_length - n), 0, n);
  }
  free(aux);
  return 0;
}

/* ---------------------------------------------------------------------------------------------------------------------
 * RX segmentation geometry: ldpc_segmenter_impl.cpp:226-308, compute_rm_length :56-65, helpers
 * include/srsran/phy/upper/channel_coding/ldpc/ldpc.h:128-207.
 */
int oracle_segment_rx(unsigned  tbs,
                      int       bg,
                      int       Qm,
                      unsigned  nof_layers,
                      unsigned  nof_ch_symbols,
                      unsigned* meta_out,
                      unsigned  max_cbs)
{
  if ((bg != 1 && bg != 2) || nof_layers < 1 || nof_layers > 4 || nof_ch_symbols % nof_layers != 0 || tbs == 0) {
    return -1;
  }
  unsigned tb_crc   = (tbs <= 3824) ? 16 : 24;    /* ldpc.h:128-133 */
  unsigned in_bits  = tbs + tb_crc;               /* nof_tb_bits_in */
  unsigned max_seg  = (bg == 1) ? 8448 : 3840;    /* ldpc.h:140-151 */
  unsigned C        = (in_bits <= max_seg) ? 1 : (in_bits + (max_seg - 24) - 1) / (max_seg - 24);
  unsigned out_bits = in_bits + ((C > 1) ? 24 * C : 0);
  /* compute_lifting_size, ldpc.h:157-193 */
  unsigned ref_length = 22;
  if (bg == 2) {
    if (in_bits > 640) {
      ref_length = 10;
    } else if (in_bits > 560) {
      ref_length = 9;
    } else if (in_bits > 192) {
      ref_length = 8;
    } else {
      ref_length = 6;
    }
  }
  unsigned Z = 0;
  for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
    if ((unsigned)LDPC_LIFTING_SIZES[i] * C * ref_length >= out_bits) {
      Z = LDPC_LIFTING_SIZES[i];
      break;
    }
  }
  if (Z == 0) {
    return -1;
  }
  unsigned seg_len  = ((bg == 1) ? 22 : 10) * Z; /* compute_codeblock_size ldpc.h:199-207 */
  unsigned crc_bits = (C > 1) ? 24 : 0;
  unsigned max_info = (out_bits + C - 1) / C - crc_bits;
  unsigned fillers  = seg_len - (max_info + crc_bits);
  unsigned sym_pl   = nof_ch_symbols / nof_layers;
  unsigned nshort   = C - (sym_pl % C);
  unsigned inv_rate = (bg == 1) ? 3 : 5;
  if (meta_out != NULL) {
    if (C > max_cbs) {
      return -2;
    }
    unsigned cw_offset = 0;
    for (unsigned i = 0; i != C; ++i) {
      unsigned t  = (i < nshort) ? (sym_pl / C) : ((sym_pl + C - 1) / C);
      unsigned rm = t * nof_layers * (unsigned)Qm;
      unsigned* m = meta_out + 6 * i;
      m[0]        = Z;
      m[1]        = seg_len * inv_rate;
      m[2]        = rm;
      m[3]        = fillers;
      m[4]        = cw_offset;
      m[5]        = (C == 1) ? tb_crc : crc_bits;
      cw_offset += rm;
    }
  }
  return (int)C;
}
