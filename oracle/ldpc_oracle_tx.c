/* TEST INFRASTRUCTURE ONLY - see ldpc_oracle.h. Transmit-side restatement (LDPC encoder + rate matcher) used solely to
 * make test signals (valid codewords) where the compiled reference is not available. Checked bit-for-bit against the
 * reference encoder / rate matcher in tests/test_oracle_vs_reference.py.
 */
#include "ldpc_oracle.h"
#include "ldpc_bg_tables.h"

#include <stdlib.h>
#include <string.h>

#define FILLER_BIT 254 /* include/srsran/phy/upper/channel_coding/ldpc/ldpc.h:113 */

typedef struct {
  int             M, N_full, K;
  const uint16_t* row_ptr;
  const uint8_t*  col;
  const uint16_t* shift;
} txgraph_t;

static int tx_graph(int bg, int Z, txgraph_t* g)
{
  int set = -1;
  for (int i = 0; i != LDPC_NOF_LIFTING_SIZES; ++i) {
    if (LDPC_LIFTING_SIZES[i] == Z) {
      set = LDPC_LIFTING_SET[i];
    }
  }
  if (set < 0) {
    return -1;
  }
  if (bg == 1) {
    g->M = LDPC_BG1_M, g->N_full = LDPC_BG1_N_FULL;
    g->row_ptr = LDPC_BG1_ROW_PTR, g->col = LDPC_BG1_COL, g->shift = LDPC_BG1_SHIFT[set];
  } else if (bg == 2) {
    g->M = LDPC_BG2_M, g->N_full = LDPC_BG2_N_FULL;
    g->row_ptr = LDPC_BG2_ROW_PTR, g->col = LDPC_BG2_COL, g->shift = LDPC_BG2_SHIFT[set];
  } else {
    return -1;
  }
  g->K = g->N_full - g->M;
  return 0;
}

/* acc[t] ^= node[(t + shift) mod Z]: one circulant block of H times a variable node (TS 38.212 5.3.2). */
static void xor_rot(uint8_t* acc, const uint8_t* node, unsigned shift, unsigned Z)
{
  for (unsigned t = 0; t != Z; ++t) {
    acc[t] ^= node[(t + shift) % Z];
  }
}

/* Systematic encoder, same contract as ldpc_encoder::encode (include/srsran/phy/upper/channel_coding/ldpc/
 * ldpc_encoder.h, behaviour of lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp): input K = bg_K*Z unpacked bits
 * (filler bits = 254, encoded as zeros), output the first out_len bits of [msg(2Z..K) | parity], i.e. the codeblock
 * with the 2Z punctured systematic bits removed; filler positions are written back as 254.
 * Method: H = [A B 0; C D I] with B the 4x4 double-diagonal core: p0 from the sum of the four core rows, p1..p3 by
 * back-substitution, extension parities by direct row evaluation. */
int oracle_encode(int bg, int Z, const uint8_t* msg_bits, uint8_t* out_bits, unsigned out_len)
{
  txgraph_t g;
  if (tx_graph(bg, Z, &g) != 0) {
    return -1;
  }
  const unsigned uZ = (unsigned)Z, K = (unsigned)g.K;
  if (out_len > (unsigned)(g.N_full - 2) * uZ) {
    return -1;
  }
  uint8_t* cw  = (uint8_t*)calloc((size_t)g.N_full * uZ, 1);
  uint8_t* lam = (uint8_t*)calloc(4 * (size_t)uZ, 1);
  uint8_t* sum = (uint8_t*)calloc(uZ, 1);
  for (unsigned i = 0; i != K * uZ; ++i) {
    cw[i] = (msg_bits[i] == FILLER_BIT) ? 0 : (msg_bits[i] & 1);
  }
  /* lambda_i = sum_j A_ij s_j for the four core rows. */
  int      p0_shift[4] = {-1, -1, -1, -1};
  for (unsigned m = 0; m != 4; ++m) {
    for (unsigned e = g.row_ptr[m]; e != g.row_ptr[m + 1]; ++e) {
      unsigned c = g.col[e];
      if (c < K) {
        xor_rot(lam + m * uZ, cw + c * uZ, g.shift[e] % uZ, uZ);
      } else if (c == K) {
        p0_shift[m] = (int)(g.shift[e] % uZ);
      } else if (g.shift[e] % uZ != 0) {
        goto fail; /* core columns K+1..K+3 are identities in both base graphs */
      }
    }
  }
  for (unsigned m = 0; m != 4; ++m) {
    for (unsigned t = 0; t != uZ; ++t) {
      sum[t] ^= lam[m * uZ + t];
    }
  }
  /* Column K has three entries in the core; two carry the same shift and cancel in the row sum. */
  {
    int cnt = 0, sh[4], rows[4];
    for (int m = 0; m != 4; ++m) {
      if (p0_shift[m] >= 0) {
        sh[cnt] = p0_shift[m], rows[cnt] = m, ++cnt;
      }
    }
    (void)rows;
    if (cnt != 3) {
      goto fail;
    }
    int odd;
    if (sh[0] == sh[1]) {
      odd = sh[2];
    } else if (sh[0] == sh[2]) {
      odd = sh[1];
    } else if (sh[1] == sh[2]) {
      odd = sh[0];
    } else {
      goto fail;
    }
    /* P^odd p0 = sum  =>  p0[(t + odd) mod Z] = sum[t]. */
    uint8_t* p0 = cw + K * uZ;
    for (unsigned t = 0; t != uZ; ++t) {
      p0[(t + (unsigned)odd) % uZ] = sum[t];
    }
  }
  /* Back-substitution: repeatedly take a core row with exactly one unknown parity node among K+1..K+3. */
  {
    int known[4] = {1, 0, 0, 0};
    for (int pass = 0; pass != 3; ++pass) {
      int progressed = 0;
      for (unsigned m = 0; m != 4 && !progressed; ++m) {
        int unknown = -1, nunk = 0;
        for (unsigned e = g.row_ptr[m]; e != g.row_ptr[m + 1]; ++e) {
          unsigned c = g.col[e];
          if (c >= K && c < K + 4 && !known[c - K]) {
            unknown = (int)(c - K), ++nunk;
          }
        }
        if (nunk != 1) {
          continue;
        }
        uint8_t* dst = cw + (K + (unsigned)unknown) * uZ;
        memcpy(dst, lam + m * uZ, uZ);
        for (unsigned e = g.row_ptr[m]; e != g.row_ptr[m + 1]; ++e) {
          unsigned c = g.col[e];
          if (c >= K && c < K + 4 && known[c - K]) {
            xor_rot(dst, cw + c * uZ, g.shift[e] % uZ, uZ);
          }
        }
        known[unknown] = 1;
        progressed     = 1;
      }
      if (!progressed) {
        goto fail;
      }
    }
  }
  /* Extension rows: p_{K+m} = sum of the row's other blocks (its own column is an identity). */
  for (unsigned m = 4; m != (unsigned)g.M; ++m) {
    uint8_t* dst = cw + (K + m) * uZ;
    for (unsigned e = g.row_ptr[m]; e != g.row_ptr[m + 1]; ++e) {
      unsigned c = g.col[e];
      if (c != K + m) {
        xor_rot(dst, cw + c * uZ, g.shift[e] % uZ, uZ);
      }
    }
  }
  memcpy(out_bits, cw + 2 * uZ, out_len);
  for (unsigned i = 2 * uZ; i != K * uZ && (i - 2 * uZ) < out_len; ++i) {
    if (msg_bits[i] == FILLER_BIT) {
      out_bits[i - 2 * uZ] = FILLER_BIT;
    }
  }
  free(cw), free(lam), free(sum);
  return 0;
fail:
  free(cw), free(lam), free(sum);
  return -3;
}

/* Rate matcher (bit selection + interleaving), lib/phy/upper/channel_coding/ldpc/ldpc_rate_matcher_impl.cpp:43-200:
 * walk the circular buffer of Ncb bits from k0(rv), skipping the filler range, then out[i*Qm + j] = sel[(E/Qm)*j + i]. */
int oracle_rate_match(uint8_t*       out,
                      unsigned       E,
                      const uint8_t* in,
                      unsigned       N,
                      int            rv,
                      int            Qm,
                      unsigned       Nref,
                      unsigned       nof_filler_bits)
{
  unsigned Ncb = (Nref > 0) ? (Nref < N ? Nref : N) : N;
  double   f1[4] = {0, 17, 33, 56}, f2[4] = {0, 13, 25, 43};
  double*  f;
  unsigned n_short, bg_K;
  if (N % 66 == 0) {
    f = f1, n_short = 66, bg_K = 22;
  } else if (N % 50 == 0) {
    f = f2, n_short = 50, bg_K = 10;
  } else {
    return -1;
  }
  if (rv < 0 || rv > 3 || Qm < 1 || E % (unsigned)Qm != 0) {
    return -1;
  }
  unsigned Z     = N / n_short;
  unsigned sys   = (bg_K - 2) * Z;
  unsigned info  = sys - nof_filler_bits;
  unsigned k0    = (unsigned)((uint16_t)((f[rv] * Ncb) / N)) * Z;
  uint8_t* sel   = (uint8_t*)malloc(E ? E : 1);
  unsigned idx   = k0;
  for (unsigned o = 0; o != E;) {
    if (idx >= info && idx < sys) {
      idx = sys % Ncb;
      continue;
    }
    sel[o++] = in[idx];
    idx      = (idx + 1) % Ncb;
  }
  if (Qm == 1) {
    memcpy(out, sel, E);
  } else {
    unsigned Kq = E / (unsigned)Qm;
    for (unsigned o = 0, i = 0; i != Kq; ++i) {
      for (unsigned j = 0; j != (unsigned)Qm; ++j, ++o) {
        out[o] = sel[Kq * j + i];
      }
    }
  }
  free(sel);
  return 0;
}
