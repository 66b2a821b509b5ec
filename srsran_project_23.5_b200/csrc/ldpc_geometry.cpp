// Host-only geometry helpers of the C ABI (no CUDA): TB -> codeblock segmentation and graph lookups.
#include "ldpc_host_util.h"

using namespace ldpc_b200;

extern "C" {

// ldpc_segmenter_impl.cpp:226-308 (RX segment), :56-65 (compute_rm_length), ldpc.h:128-207 (helpers) and
// select_crc, pusch_decoder_impl.cpp:44-56. The reference's 52-codeblock cap (codeblock_metadata.h:88) is lifted.
int ldpc_cuda_segment(uint32_t tbs, int bg, int rv, int Qm, uint32_t Nref, uint32_t nof_layers, uint32_t nof_ch_symbols,
                      ldpc_cuda_cb_desc* descs, uint32_t max_cbs, uint32_t* nof_crc_bits_out)
{
  if ((bg != 1 && bg != 2) || rv < 0 || rv > 3 || nof_layers < 1 || nof_layers > 4 || tbs == 0 ||
      nof_ch_symbols % nof_layers != 0 || !(Qm == 1 || Qm == 2 || Qm == 4 || Qm == 6 || Qm == 8)) {
    return LDPC_CUDA_ERR_ARG;
  }
  const uint32_t tb_crc = (tbs <= 3824) ? 16 : 24;
  const uint32_t in_bits = tbs + tb_crc;
  const uint32_t max_seg = (bg == 1) ? 8448 : 3840;
  const uint32_t C = (in_bits <= max_seg) ? 1 : (in_bits + (max_seg - 24) - 1) / (max_seg - 24);
  const uint32_t out_bits = in_bits + ((C > 1) ? 24 * C : 0);
  uint32_t ref_length = 22;
  if (bg == 2) {
    ref_length = (in_bits > 640) ? 10 : (in_bits > 560) ? 9 : (in_bits > 192) ? 8 : 6;
  }
  uint32_t Z = 0;
  for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
    if ((uint32_t)LDPC_LIFTING_SIZES[i] * C * ref_length >= out_bits) {
      Z = LDPC_LIFTING_SIZES[i];
      break;
    }
  }
  if (Z == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (descs == nullptr || C > max_cbs) {
    return (int)C;
  }
  const uint32_t seg_len = bg_K(bg) * Z;
  const uint32_t crc_bits = (C > 1) ? 24 : 0;
  const uint32_t max_info = (out_bits + C - 1) / C - crc_bits;
  const uint32_t fillers = seg_len - (max_info + crc_bits);
  const uint32_t sym_pl = nof_ch_symbols / nof_layers;
  const uint32_t nshort = C - (sym_pl % C);
  uint64_t cw_offset = 0;
  for (uint32_t i = 0; i != C; ++i) {
    const uint32_t t = (i < nshort) ? (sym_pl / C) : ((sym_pl + C - 1) / C);
    ldpc_cuda_cb_desc& d = descs[i];
    std::memset(&d, 0, sizeof(d));
    d.base_graph = (uint8_t)bg;
    d.lifting_size = (uint16_t)Z;
    d.nof_filler_bits = (uint16_t)fillers;
    d.rv = (uint8_t)rv;
    d.mod_order = (uint8_t)Qm;
    d.Nref = Nref;
    d.rm_length = t * nof_layers * (uint32_t)Qm;
    d.in_length = bg_N_short(bg) * Z;
    d.crc_poly = (C > 1) ? LDPC_CUDA_CRC24B : (tbs > 3824 ? LDPC_CUDA_CRC24A : LDPC_CUDA_CRC16);
    d.early_stop = 1;
    d.new_data = 1;
    d.in_offset = cw_offset;
    cw_offset += d.rm_length;
    if (nof_crc_bits_out) {
      nof_crc_bits_out[i] = (C == 1) ? tb_crc : crc_bits;
    }
  }
  return (int)C;
}

int ldpc_cuda_graph_shift(int bg, int Z, int m, int n)
{
  const int set = lifting_set_of(Z);
  if (set < 0 || (bg != 1 && bg != 2) || m < 0 || m >= (int)bg_M(bg)) {
    return -1;
  }
  const uint16_t* row_ptr = (bg == 1) ? LDPC_BG1_ROW_PTR : LDPC_BG2_ROW_PTR;
  const uint8_t* col = (bg == 1) ? LDPC_BG1_COL : LDPC_BG2_COL;
  const uint16_t* shift = (bg == 1) ? LDPC_BG1_SHIFT[set] : LDPC_BG2_SHIFT[set];
  for (unsigned e = row_ptr[m]; e != row_ptr[m + 1]; ++e) {
    if (col[e] == n) {
      return shift[e] % Z;
    }
  }
  return -1;
}

}  // extern "C"
