// Downlink mirror of the receive chain: LDPC encoder and rate matcher on the device, used by the
// product's own tools (bench.py builds its synthetic codewords with them) and as the transmit half of loop-back tests.
//
// Encoder: same contract as ldpc_encoder::encode (include/srsran/phy/upper/channel_coding/ldpc/ldpc_encoder.h:41-46,
// behaviour of lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp:33-150): input bg_K*Z unpacked bits (one per
// byte, filler bits = 254 encoded as zeros), output the codeblock with the 2Z punctured systematic bits removed
// (N_short*Z bytes, fillers written back as 254). Method (TS 38.212 5.3.2): H = [A B 0; C D I] with B the 4x4
// double-diagonal core - p0 from the sum of the four core rows, p1..p3 by back-substitution, extension parities by
// direct row evaluation. One block per codeblock, thread t owns lane t of every node; the circulant products are
// shared-memory rotations.
//
// Rate matcher: ldpc_rate_matcher::rate_match (ldpc_rate_matcher.h:41-45, ldpc_rate_matcher_impl.cpp:43-200): bit
// selection from the circular buffer from k0(rv) skipping the filler range, then bit interleaving
// out[i*Qm + j] = sel[(E/Qm)*j + i] - evaluated in closed form per output bit with the geometry of the dematcher.
#pragma once

#include "ldpc_dematch.cuh"
#include "ldpc_device.cuh"

namespace ldpc_b200 {

constexpr int ENC_FILLER_BIT = 254;  // include/srsran/phy/upper/channel_coding/ldpc/ldpc.h:113
constexpr int ENC_THREADS = 384;

// work[cb]: in_offset = message (bg_K*Z bytes), out_offset = codeblock (N_short*Z bytes), reserved = lifting set,
// crc_poly field carries the base graph (1 / 2), in_length the lifting size.
__global__ void __launch_bounds__(ENC_THREADS) ldpc_encode_kernel(const CbWork* __restrict__ work, unsigned nof_cbs,
                                                                   const uint8_t* __restrict__ msg_base,
                                                                   uint8_t* __restrict__ out_base)
{
  extern __shared__ uint8_t enc_smem[];
  __shared__ int s_odd, s_fail;
  const int t = threadIdx.x;
  for (unsigned cb = blockIdx.x; cb < nof_cbs; cb += gridDim.x) {
    const CbWork w = work[cb];
    const int bg = w.crc_poly, Z = (int)w.in_length, set = (int)w.reserved, g = bg - 1;
    const int M = (bg == 2) ? LDPC_BG2_M : LDPC_BG1_M, K = (bg == 2) ? 10 : 22, n_full = K + M;
    const uint8_t* msg = msg_base + w.in_offset;
    uint8_t* out = out_base + w.out_offset;
    uint8_t* cw = enc_smem;                    // [n_full][Z]
    uint8_t* lam = enc_smem + n_full * MAX_Z;  // [4][Z]
    const uint16_t* row_ptr = c_graphs.row_ptr[g];
    const uint8_t* col = c_graphs.col[g];
    const uint16_t* shift = c_graphs.shift[g][set];
    __syncthreads();  // the previous codeblock of this block is done with the shared arrays
    if (t < Z) {
      for (int c = 0; c < K; ++c) {
        const int b = msg[c * Z + t];
        cw[c * Z + t] = (b == ENC_FILLER_BIT) ? 0 : (uint8_t)(b & 1);
      }
    }
    if (t == 0) {
      // column K has three entries in the core rows; two carry the same shift and cancel in the sum of the four rows
      int sh[4], cnt = 0;
      for (int m = 0; m < 4; ++m) {
        for (int e = row_ptr[m]; e < row_ptr[m + 1]; ++e) {
          if (col[e] == K && cnt < 4) {
            sh[cnt++] = shift[e] % Z;
          }
        }
      }
      int odd = -1;
      if (cnt == 3) {
        odd = (sh[0] == sh[1]) ? sh[2] : ((sh[0] == sh[2]) ? sh[1] : ((sh[1] == sh[2]) ? sh[0] : -1));
      }
      s_odd = odd;
      s_fail = odd < 0;
    }
    __syncthreads();
    if (t < Z) {
      // lambda_m = sum_j A_mj s_j for the four core rows, and their sum
      int sum = 0;
      for (int m = 0; m < 4; ++m) {
        int acc = 0;
        for (int e = row_ptr[m]; e < row_ptr[m + 1]; ++e) {
          const int c = col[e];
          if (c < K) {
            int j = t + (int)(shift[e] % Z);
            j -= (j >= Z) ? Z : 0;
            acc ^= cw[c * Z + j];
          }
        }
        lam[m * Z + t] = (uint8_t)acc;
        sum ^= acc;
      }
      // P^odd p0 = sum  =>  p0[(t + odd) mod Z] = sum[t]
      int j = t + max(s_odd, 0);
      j -= (j >= Z) ? Z : 0;
      cw[K * Z + j] = (uint8_t)sum;
    }
    __syncthreads();
    // back-substitution: repeatedly take a core row with exactly one unknown parity node among K+1..K+3
    unsigned known = 1u;  // bit i: node K + i known (uniform across the block)
    for (int pass = 0; pass < 3; ++pass) {
      int row = -1, unknown = -1;
      for (int m = 0; m < 4 && row < 0; ++m) {
        int nunk = 0, u = -1;
        for (int e = row_ptr[m]; e < row_ptr[m + 1]; ++e) {
          const int c = col[e];
          if (c >= K && c < K + 4 && !((known >> (c - K)) & 1u)) {
            u = c - K, ++nunk;
          }
        }
        if (nunk == 1) {
          row = m, unknown = u;
        }
      }
      if (row < 0) {
        if (t == 0) {
          s_fail = 1;
        }
        break;
      }
      if (t < Z) {
        int acc = lam[row * Z + t];
        for (int e = row_ptr[row]; e < row_ptr[row + 1]; ++e) {
          const int c = col[e];
          if (c >= K && c < K + 4 && ((known >> (c - K)) & 1u)) {
            int j = t + (int)(shift[e] % Z);
            j -= (j >= Z) ? Z : 0;
            acc ^= cw[c * Z + j];
          }
        }
        cw[(K + unknown) * Z + t] = (uint8_t)acc;
      }
      known |= 1u << unknown;
      __syncthreads();
    }
    if (t < Z) {
      // extension rows: p_{K+m} = sum of the row's other blocks (its own column is an identity); they only read nodes
      // below K + 4, so no ordering between rows is needed
      for (int m = 4; m < M; ++m) {
        int acc = 0;
        for (int e = row_ptr[m]; e < row_ptr[m + 1]; ++e) {
          const int c = col[e];
          if (c != K + m) {
            int j = t + (int)(shift[e] % Z);
            j -= (j >= Z) ? Z : 0;
            acc ^= cw[c * Z + j];
          }
        }
        cw[(K + m) * Z + t] = (uint8_t)acc;
      }
      const int fail = s_fail;
      for (int c = 2; c < n_full; ++c) {
        int b = cw[c * Z + t];
        if (c < K && msg[c * Z + t] == ENC_FILLER_BIT) {
          b = ENC_FILLER_BIT;
        }
        out[(c - 2) * Z + t] = fail ? (uint8_t)0xff : (uint8_t)b;  // 0xff marks an internal table inconsistency
      }
    }
  }
}

// work[cb]: in_offset = codeblock (N_short*Z bytes), out_offset = E output bits (one per byte); geometry fields as for
// the dematcher (rm_length = E, rv, mod_order, Nref, nof_filler_bits); crc_poly carries the base graph, in_length Z.
__global__ void __launch_bounds__(256) ldpc_rate_match_kernel(const CbWork* __restrict__ work, unsigned nof_cbs,
                                                               const uint8_t* __restrict__ cw_base,
                                                               uint8_t* __restrict__ out_base)
{
  for (unsigned cb = blockIdx.x; cb < nof_cbs; cb += gridDim.x) {
    const CbWork w = work[cb];
    const DematchGeom g = make_dematch_geom(w.crc_poly, w.in_length, w.nof_filler_bits, w.rv, w.mod_order, w.Nref,
                                            w.rm_length, 1);
    const uint8_t* cw = cw_base + w.in_offset;
    uint8_t* out = out_base + w.out_offset;
    const unsigned Kq = g.E / g.Qm;
    for (unsigned o = threadIdx.x; o < g.E; o += blockDim.x) {
      const unsigned i = o / g.Qm, j = o - i * g.Qm;
      const unsigned e = j * Kq + i;           // rank in selection order
      const unsigned L = (g.Ls0 + e) % g.V;    // linear index (fillers removed) inside the circular buffer
      const unsigned p = L < g.info ? L : L + (g.sys - g.info);
      out[o] = cw[p];
    }
  }
}

}  // namespace ldpc_b200
