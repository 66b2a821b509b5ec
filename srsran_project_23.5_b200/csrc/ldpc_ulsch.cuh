// UL-SCH demultiplexing on the device: splitting the descrambled soft bits of a PUSCH codeword into
// SCH data, HARQ-ACK, CSI part 1 and CSI part 2 (TS 38.212 6.2.7), so that the LLRs of a PUSCH that carries UCI go from the
// demodulator to the rate dematcher without leaving HBM.
//
// Replaces ulsch_demultiplex_impl::demultiplex / get_placeholders
// (lib/phy/upper/channel_processors/ulsch_demultiplex_impl.cpp:316-453, generic walk :76-312). The reference walks the
// resource elements of every OFDM symbol one by one with running counters. Here the walk is inverted into a CLOSED FORM per
// resource element: inside one symbol every field takes "every d-th of the elements still eligible for it, n times", so the
// class of element i and its rank inside its class follow from a few integer divisions (ulsch_classify), and one thread per
// element copies its nof_bits_per_re soft bits to where they belong. The per-symbol strides and counts (which depend on what
// earlier symbols consumed) are 14 small records built on the host (ulsch_plan).
#pragma once

#include <cstdint>
#include <vector>

#include "ldpc_device.cuh"

namespace ldpc_b200 {

// One OFDM symbol of one codeword. All positions are in resource elements (x nof_bits_per_re soft bits).
struct UlschSymbol {
  unsigned in_base;    // first input element of the symbol
  unsigned n_re;       // elements of the symbol (DM-RS symbols: the ones left for data)
  unsigned rvd_d, rvd_n;    // reserved elements: every rvd_d-th element, rvd_n of them (0: none)
  unsigned ack_d, ack_n;    // HARQ-ACK: every ack_d-th of (reserved elements | all elements), ack_n of them
  unsigned csi1_d, csi1_n;  // CSI part 1: every csi1_d-th of the elements that are neither reserved nor HARQ-ACK
  unsigned csi2_d, csi2_n;  // CSI part 2: every csi2_d-th of the elements not taken by HARQ-ACK (unreserved case) / CSI 1
  unsigned sch_base, ack_base, csi1_base, csi2_base;  // first output element of the symbol in every field
  unsigned ack_on_reserved;  // HARQ-ACK punctures reserved elements (G_rvd != 0): SCH / CSI 2 get zeros there
  unsigned job;              // index of the codeword
};

struct UlschJob {  // one codeword: byte offsets of its buffers, bits per element
  unsigned long long in_offset, sch_offset, ack_offset, csi1_offset, csi2_offset;
  unsigned bits_per_re;
  unsigned reserved;
};

enum UlschField { ULSCH_SCH = 0, ULSCH_ACK = 1, ULSCH_CSI1 = 2, ULSCH_CSI2 = 3 };

struct UlschWhere {
  int      field;   // where the element's soft bits go (UlschField)
  unsigned rank;    // index of the element inside that field, within the symbol
  int      zero_field;  // -1, or the field (SCH / CSI 2) that receives an all-zero element at zero_rank instead of input
  unsigned zero_rank;
};

// multiples of d below i among the first n multiples: |{k < n : k d < i}|
__host__ __device__ inline unsigned ulsch_taken_before(unsigned i, unsigned d, unsigned n)
{
  return (n == 0) ? 0u : min((i + d - 1u) / d, n);
}
// i is one of the first n multiples of d
__host__ __device__ inline bool ulsch_takes(unsigned i, unsigned d, unsigned n) { return n != 0 && i % d == 0 && i / d < n; }

// Class and rank of element i of a symbol without DM-RS. Mirrors the order of the tests in
// ulsch_demultiplex_impl.cpp:246-300: reserved?, HARQ-ACK, CSI part 1 (never on a reserved element), CSI part 2, SCH.
__host__ __device__ inline UlschWhere ulsch_classify(const UlschSymbol& s, unsigned i)
{
  UlschWhere w = {ULSCH_SCH, 0u, -1, 0u};
  const bool     reserved   = ulsch_takes(i, s.rvd_d, s.rvd_n);
  const unsigned rvd_before = ulsch_taken_before(i, s.rvd_d, s.rvd_n);
  bool           ack        = false;
  unsigned       ack_before = 0;
  if (s.ack_on_reserved) {
    // the counter of the HARQ-ACK stride advances on reserved elements only (:259)
    ack        = reserved && ulsch_takes(i / s.rvd_d, s.ack_d, s.ack_n);
    ack_before = ulsch_taken_before(rvd_before, s.ack_d, s.ack_n);
  } else {
    ack        = ulsch_takes(i, s.ack_d, s.ack_n);
    ack_before = ulsch_taken_before(i, s.ack_d, s.ack_n);
  }
  if (ack && !s.ack_on_reserved) {
    w.field = ULSCH_ACK, w.rank = ack_before;
    return w;  // the element is gone (:277 continue)
  }
  // elements CSI part 1 may use: not reserved, not HARQ-ACK
  const unsigned gone1       = s.ack_on_reserved ? rvd_before : ack_before;
  const unsigned elig_before = i - gone1;
  const bool     csi1        = !reserved && ulsch_takes(elig_before, s.csi1_d, s.csi1_n);
  const unsigned csi1_before = ulsch_taken_before(elig_before, s.csi1_d, s.csi1_n);
  // elements that reach the CSI part 2 test: everything but unreserved-case HARQ-ACK and CSI part 1
  const unsigned reach_before = i - (s.ack_on_reserved ? 0u : ack_before) - csi1_before;
  const bool     csi2         = !csi1 && ulsch_takes(reach_before, s.csi2_d, s.csi2_n);
  const unsigned csi2_before  = ulsch_taken_before(reach_before, s.csi2_d, s.csi2_n);
  const unsigned sch_before   = reach_before - csi2_before;
  if (ack) {  // punctured reserved element: the input goes to HARQ-ACK, the field underneath gets zeros (:262-266)
    w.field = ULSCH_ACK, w.rank = ack_before;
    w.zero_field = csi2 ? ULSCH_CSI2 : ULSCH_SCH;
    w.zero_rank  = csi2 ? csi2_before : sch_before;
    return w;
  }
  if (csi1) {
    w.field = ULSCH_CSI1, w.rank = csi1_before;
  } else if (csi2) {
    w.field = ULSCH_CSI2, w.rank = csi2_before;
  } else {
    w.field = ULSCH_SCH, w.rank = sch_before;
  }
  return w;
}

// What the host knows about one codeword (== ldpc_cuda_ulsch_desc without the buffer offsets).
struct UlschGeometry {
  unsigned mod_order, nof_layers, nof_prb, start_symbol, nof_symbols, dmrs_type, nof_cdm_groups_without_data;
  unsigned dmrs_symbol_mask, nof_harq_ack_rvd, nof_enc_harq_ack, nof_enc_csi1, nof_enc_csi2;
};

struct UlschTotals {
  unsigned in_re, sch_re, ack_re, csi1_re, csi2_re;
};

// The per-symbol records of one codeword (TS 38.212 6.2.7 steps as the reference performs them per symbol,
// ulsch_demultiplex_impl.cpp:121-236). Returns false when the configuration is inconsistent (where the reference asserts):
// no DM-RS symbol, fields that do not fit, bits left over.
inline bool ulsch_plan(const UlschGeometry& g_in, unsigned job, std::vector<UlschSymbol>& out, UlschTotals& tot)
{
  UlschGeometry g = g_in;
  if (g.nof_enc_harq_ack == 0 && g.nof_enc_csi1 == 0 && g.nof_enc_csi2 == 0) {
    g.nof_harq_ack_rvd = 0;  // no UCI field at all: the reference copies the input through (:324-331), nothing is reserved
  }
  const unsigned bpr = g.mod_order * g.nof_layers;
  if (bpr == 0 || g.nof_prb == 0 || g.start_symbol + g.nof_symbols > 14 || (g.dmrs_type != 1 && g.dmrs_type != 2) ||
      g.nof_cdm_groups_without_data < 1 || g.nof_cdm_groups_without_data > (g.dmrs_type == 1 ? 2u : 3u)) {
    return false;
  }
  int first_dmrs = -1, l1 = -1, l1_csi = -1;
  for (int l = 0; l != 14; ++l) {
    const bool d = (g.dmrs_symbol_mask >> l) & 1u;
    if (d && first_dmrs < 0) {
      first_dmrs = l;
    }
    if (!d && first_dmrs >= 0 && l1 < 0) {
      l1 = l;  // first symbol without DM-RS behind the first DM-RS symbol (:33-46)
    }
    if (!d && l1_csi < 0) {
      l1_csi = l;  // first symbol without DM-RS (:48-55)
    }
  }
  if (first_dmrs < 0 || l1 < 0 || l1_csi < 0) {
    return false;
  }
  const unsigned re_dmrs_per_prb = g.nof_cdm_groups_without_data * (g.dmrs_type == 1 ? 6u : 4u);
  if (re_dmrs_per_prb > 12) {
    return false;
  }
  const unsigned n_dmrs_sym = (12u - re_dmrs_per_prb) * g.nof_prb, n_sc = 12u * g.nof_prb;
  auto share = [bpr](unsigned remaining, unsigned avail_re, unsigned& d, unsigned& n) {
    // "every d-th of avail_re elements, n of them" for `remaining` bits (:157-161 and alike)
    d = 1, n = avail_re;
    if (remaining < avail_re * bpr) {
      d = (avail_re * bpr) / remaining;
      n = (remaining + bpr - 1) / bpr;
    }
  };
  unsigned done_rvd = 0, done_ack = 0, done_csi1 = 0, done_csi2 = 0;
  tot = UlschTotals{0, 0, 0, 0, 0};
  for (unsigned l = g.start_symbol; l != g.start_symbol + g.nof_symbols; ++l) {
    UlschSymbol s = {};
    s.job = job, s.in_base = tot.in_re;
    s.sch_base = tot.sch_re, s.ack_base = tot.ack_re, s.csi1_base = tot.csi1_re, s.csi2_base = tot.csi2_re;
    s.ack_on_reserved = g.nof_harq_ack_rvd != 0;
    if ((g.dmrs_symbol_mask >> l) & 1u) {
      s.n_re = n_dmrs_sym;  // data only (:126-131)
      tot.in_re += s.n_re, tot.sch_re += s.n_re;
      if (s.n_re != 0) {
        out.push_back(s);
      }
      continue;
    }
    s.n_re            = n_sc;
    unsigned free_re  = n_sc;  // M_uci_sc
    unsigned rvd_re   = 0;     // M_uci_rvd
    if ((int)l >= l1) {
      const unsigned left_rvd = g.nof_harq_ack_rvd - done_rvd, left_ack = g.nof_enc_harq_ack - done_ack;
      if (g.nof_harq_ack_rvd != 0 && left_rvd != 0) {
        share(left_rvd, free_re, s.rvd_d, s.rvd_n);
        rvd_re = s.rvd_n;
        if (left_ack != 0) {
          share(left_ack, rvd_re, s.ack_d, s.ack_n);
        }
      } else if (left_ack != 0) {
        share(left_ack, free_re, s.ack_d, s.ack_n);
        free_re -= s.ack_n;
        // with reserved bits configured but used up, the reference still walks in the "reserved" mode, in which no
        // element is reserved any more and so no HARQ-ACK element can be placed (:255-267): it then asserts. Refuse.
        if (s.ack_on_reserved) {
          return false;
        }
      }
    }
    if ((int)l >= l1_csi) {
      const unsigned left1 = g.nof_enc_csi1 - done_csi1, left2 = g.nof_enc_csi2 - done_csi2;
      if (free_re > rvd_re && left1 != 0) {
        share(left1, free_re - rvd_re, s.csi1_d, s.csi1_n);
        free_re -= s.csi1_n;
      }
      if (free_re > 0 && left2 != 0) {
        share(left2, free_re, s.csi2_d, s.csi2_n);
        free_re -= s.csi2_n;
      }
    }
    done_rvd += s.rvd_n * bpr, done_ack += s.ack_n * bpr, done_csi1 += s.csi1_n * bpr, done_csi2 += s.csi2_n * bpr;
    tot.in_re += n_sc;
    tot.ack_re += s.ack_n, tot.csi1_re += s.csi1_n, tot.csi2_re += s.csi2_n;
    tot.sch_re += free_re;  // includes the zeroed elements under punctured reserved ones
    out.push_back(s);
  }
  // "all bits have been processed" (:304-311)
  return done_rvd == g.nof_harq_ack_rvd && done_ack == g.nof_enc_harq_ack && done_csi1 == g.nof_enc_csi1 &&
         done_csi2 == g.nof_enc_csi2;
}

// grid (blocks of 256 elements, symbols of all codewords). One thread = one resource element.
__global__ void __launch_bounds__(256) ldpc_ulsch_demux_kernel(const UlschSymbol* __restrict__ symbols,
                                                               const UlschJob* __restrict__ jobs,
                                                               const int8_t* __restrict__ in, int8_t* __restrict__ sch,
                                                               int8_t* __restrict__ uci)
{
  const UlschSymbol s = symbols[blockIdx.y];
  const unsigned    i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n_re) {
    return;
  }
  const UlschJob j   = jobs[s.job];
  const unsigned bpr = j.bits_per_re;
  UlschWhere     w   = {ULSCH_SCH, i, -1, 0u};  // DM-RS symbol: data only
  if (s.rvd_n | s.ack_n | s.csi1_n | s.csi2_n) {
    w = ulsch_classify(s, i);
  }
  auto dest = [&](int field, unsigned rank) -> int8_t* {
    switch (field) {
      case ULSCH_ACK:
        return uci + j.ack_offset + (size_t)(s.ack_base + rank) * bpr;
      case ULSCH_CSI1:
        return uci + j.csi1_offset + (size_t)(s.csi1_base + rank) * bpr;
      case ULSCH_CSI2:
        return uci + j.csi2_offset + (size_t)(s.csi2_base + rank) * bpr;
      default:
        return sch + j.sch_offset + (size_t)(s.sch_base + rank) * bpr;
    }
  };
  const int8_t* src = in + j.in_offset + (size_t)(s.in_base + i) * bpr;
  int8_t*       dst = dest(w.field, w.rank);
  for (unsigned b = 0; b != bpr; ++b) {
    dst[b] = src[b];
  }
  if (w.zero_field >= 0) {
    int8_t* z = dest(w.zero_field, w.zero_rank);
    for (unsigned b = 0; b != bpr; ++b) {
      z[b] = 0;
    }
  }
}

}  // namespace ldpc_b200
