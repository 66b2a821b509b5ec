/*
 * TEST INFRASTRUCTURE ONLY (parity oracle): plain-C restatement of the reference's soft demodulator and LLR descrambler,
 * the upstream producer of the LDPC receive chain's input. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg may load it; the product never does.
 *
 * Follows (paths relative to the srsRAN Project 23.5 tree), generic per-symbol code paths:
 *   demodulation_mapper_impl::demodulate_soft      lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:36-106
 *   demodulate_soft_QPSK   (generic tail)          .../demodulation_mapper_qpsk.cpp:120-168
 *   demodulate_soft_QAM16  (generic tail)          .../demodulation_mapper_qam16.cpp:180-246
 *   demodulate_soft_QAM64  (generic tail)          .../demodulation_mapper_qam64.cpp:266-332
 *   demodulate_soft_QAM256 (generic tail)          .../demodulation_mapper_qam256.cpp:279-350
 *   interval_function / compute_interval_idx       .../demodulation_mapper_intervals.h:36-64
 *   log_likelihood_ratio::quantize                 lib/phy/upper/log_likelihood_ratio.cpp:87-96
 *   pseudo_random_generator (TS 38.211 5.2.1)      lib/phy/upper/sequence_generators/pseudo_random_generator_impl.cpp
 *   pusch_demodulator_impl::descramble             lib/phy/upper/channel_processors/pusch_demodulator_impl.cpp:100-152
 *                                                  (no repetition placeholders: pure data)
 *
 * Floating point: single precision, one rounding per written operation, EXCEPT the multiply-adds the reference build
 * contracts (gcc -O3 -mfma, -ffp-contract=fast: `slope * value + intercept` and `0.8F - c * |x|`), written as fmaf()
 * here. This file must be compiled with -ffp-contract=off. Pinned against the compiled reference's generic path
 * bit-for-bit (tests/test_oracle_vs_reference.py); the reference's AVX2 kernels differ from its own generic code by at
 * most one LLR step (reciprocal-multiply instead of divide, round-half-even instead of half-away), reported by the test.
 *
 * The slope / intercept tables are not transcribed: they are derived from the TS 38.211 5.1 bit mapping (nearest
 * constellation level with the bit 0 / 1 from the middle of each interval), with the reference's float expressions
 * (integer * d, integer / 21 or / 85).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#define LLR_MAX_F 120.0f

static int8_t quantize(float value, float range_limit)
{
  float clipped = value;
  if (fabsf(value) > range_limit) {
    clipped = copysignf(range_limit, value);
  }
  return (int8_t)roundf(clipped / range_limit * LLR_MAX_F);
}

/* PAM level (odd integer) of the bits (b[0] = sign bit ... b[nb-1] = finest) per TS 38.211 5.1.3-5.1.6. */
static int pam_level(const int* b, int nb)
{
  int step = 2;
  int v    = 2 - (1 - 2 * b[nb - 1]); /* innermost bracket: 1 or 3; the outer brackets are built from the inside out */
  for (int k = nb - 2; k >= 1; --k) {
    step *= 2;
    v = step - (1 - 2 * b[k]) * v;
  }
  return (1 - 2 * b[0]) * v;
}

typedef struct {
  int   nof_intervals;
  float width;
  float slope[16];
  float intercept[16];
} interval_table;

/* Table of axis bit k (0 = sign bit) for a constellation of L levels per axis (8: 64QAM, 16: 256QAM). */
static void build_table(int L, int k, interval_table* t)
{
  const int   nb   = (L == 8) ? 3 : 4;
  const float d    = 1.0f / sqrtf((L == 8) ? 42.0f : 170.0f);
  const float den  = (L == 8) ? 21.0f : 85.0f;
  const int   fine = (k == nb - 1);
  t->nof_intervals = fine ? L / 2 : L;
  t->width         = (float)(fine ? 4 : 2) * d;
  for (int i = 0; i != t->nof_intervals; ++i) {
    /* twice the middle of the interval in units of d (kept integer) */
    const int mid2 = fine ? 2 * (4 * (i - L / 4) + 2) : 2 * (2 * (i - L / 2) + 1);
    int best[2] = {0, 0}, dist[2] = {1 << 30, 1 << 30};
    for (int code = 0; code != L; ++code) {
      int b[4];
      for (int j = 0; j != nb; ++j) {
        b[j] = (code >> j) & 1;
      }
      const int a = pam_level(b, nb), dd = abs(2 * a - mid2);
      if (dd < dist[b[k]]) {
        dist[b[k]] = dd, best[b[k]] = a;
      }
    }
    const int a0 = best[0], a1 = best[1];
    t->slope[i]     = (float)(2 * (a0 - a1)) * d;
    t->intercept[i] = (float)((a1 * a1 - a0 * a0) / 2) / den;
  }
}

static float interval_function(float value, float rcp_noise, const interval_table* t)
{
  int idx = (int)floorf(value / t->width) + t->nof_intervals / 2;
  idx     = idx < 0 ? 0 : (idx > t->nof_intervals - 1 ? t->nof_intervals - 1 : idx);
  float l = fmaf(t->slope[idx], value, t->intercept[idx]);
  l *= rcp_noise;
  return l;
}

static int8_t demod_bpsk(float re, float im, float noise)
{
  if (!(noise > 0)) {
    return 0;
  }
  const float l = 2.0f * 1.41421356237309504880f * (re + im) / noise;
  return quantize(l, 24.0f);
}

static int8_t demod_qpsk(float x, float noise)
{
  if (!(noise > 0)) {
    return 0;
  }
  const float l = 2.0f * 1.41421356237309504880f * x / noise;
  return quantize(l, 24.0f); /* QPSK and BPSK quantise over +-24, the QAMs over +-20 (demodulation_mapper_qpsk.cpp:35) */
}

static int8_t demod_qam16_01(float x, float noise)
{
  const float s10 = 1.0f / sqrtf(10.0f);
  if (!(noise > 0)) {
    return 0;
  }
  float l = 4 * s10 * x;
  if (fabsf(x) > 2 * s10) {
    l = 2 * l - copysignf(0.8f, x);
  }
  l /= noise;
  return quantize(l, 20.0f);
}

static int8_t demod_qam16_23(float x, float noise)
{
  const float s10 = 1.0f / sqrtf(10.0f);
  if (!(noise > 0)) {
    return 0;
  }
  float l = fmaf(-(4 * s10), fabsf(x), 0.8f);
  l /= noise;
  return quantize(l, 20.0f);
}

void oracle_demodulate_soft(int mod_order, int pi2, unsigned n, const float* sym, const float* noise, int8_t* llr)
{
  interval_table t[4];
  const int L = (mod_order == 6) ? 8 : 16, nb = mod_order / 2;
  if (mod_order == 6 || mod_order == 8) {
    for (int k = 0; k != nb; ++k) {
      build_table(L, k, &t[k]);
    }
  }
  for (unsigned i = 0; i != n; ++i) {
    const float re = sym[2 * i], im = sym[2 * i + 1], nv = noise[i];
    switch (mod_order) {
      case 1:
        if (pi2 && (i & 1)) {
          llr[i] = demod_bpsk(im, -re, nv);
        } else {
          llr[i] = demod_bpsk(re, im, nv);
        }
        break;
      case 2:
        llr[2 * i]     = demod_qpsk(re, nv);
        llr[2 * i + 1] = demod_qpsk(im, nv);
        break;
      case 4:
        llr[4 * i]     = demod_qam16_01(re, nv);
        llr[4 * i + 1] = demod_qam16_01(im, nv);
        llr[4 * i + 2] = demod_qam16_23(re, nv);
        llr[4 * i + 3] = demod_qam16_23(im, nv);
        break;
      default: {
        float rcp = 0;
        if (nv > 0) {
          rcp = 1 / nv;
        }
        for (int k = 0; k != nb; ++k) {
          llr[(unsigned)mod_order * i + 2 * k]     = quantize(interval_function(re, rcp, &t[k]), 20.0f);
          llr[(unsigned)mod_order * i + 2 * k + 1] = quantize(interval_function(im, rcp, &t[k]), 20.0f);
        }
      }
    }
  }
}

/* TS 38.211 5.2.1: c(n) = x1(n + 1600) ^ x2(n + 1600); x1(n+31) = x1(n+3) ^ x1(n), x1(0) = 1;
 * x2(n+31) = x2(n+3) ^ x2(n+2) ^ x2(n+1) ^ x2(n), x2(i) = bit i of c_init. Bit serial. */
void oracle_gold_sequence(unsigned c_init, unsigned n, uint8_t* c)
{
  uint32_t x1 = 1, x2 = c_init & 0x7fffffffu;
  for (unsigned i = 0; i != 1600 + n; ++i) {
    if (i >= 1600) {
      c[i - 1600] = (uint8_t)((x1 ^ x2) & 1u);
    }
    const uint32_t f1 = ((x1 >> 3) ^ x1) & 1u, f2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1u;
    x1 = (x1 >> 1) | (f1 << 30);
    x2 = (x2 >> 1) | (f2 << 30);
  }
}

/* apply_xor on soft bits: the LLR changes sign where the chip is 1 (two's complement negation of the int8). */
void oracle_descramble(unsigned c_init, unsigned n, const int8_t* in, int8_t* out)
{
  uint32_t x1 = 1, x2 = c_init & 0x7fffffffu;
  for (unsigned i = 0; i != 1600 + n; ++i) {
    if (i >= 1600) {
      const unsigned k = i - 1600;
      out[k] = ((x1 ^ x2) & 1u) ? (int8_t)(-in[k]) : in[k];
    }
    const uint32_t f1 = ((x1 >> 3) ^ x1) & 1u, f2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1u;
    x1 = (x1 >> 1) | (f1 << 30);
    x2 = (x2 >> 1) | (f2 << 30);
  }
}

/* pusch_demodulator_impl::descramble with the repetition placeholders of UCI multiplexing (pusch_demodulator_impl.cpp:
 * 100-152; placeholder positions from ulsch_placeholder_list::for_each, ulsch_placeholder_list.h:73-95): for every listed
 * RE and layer the soft bit y = Qm (layers re + layer) + 1 repeats the scrambling chip of the bit before it, the
 * Qm - 2 bits behind it (x placeholders) are copied as they are, and the sequence keeps running underneath (the chips of
 * those positions are discarded). Walks the sequence exactly like the reference: segment by segment. */
void oracle_descramble_placeholders(unsigned c_init, unsigned n, const int8_t* in, int8_t* out, unsigned mod_order,
                                    unsigned nof_layers, unsigned nof_re, const uint16_t* re_indexes)
{
  uint8_t* c = (uint8_t*)malloc(n + 16);
  oracle_gold_sequence(c_init, n + 16, c);
  const unsigned nx = mod_order > 2 ? mod_order - 2 : 0;
  unsigned pos = 0, chip = 0; /* next element, next chip of the sequence */
  for (unsigned r = 0; r != nof_re; ++r) {
    for (unsigned layer = 0; layer != nof_layers; ++layer) {
      const unsigned y = mod_order * (nof_layers * (unsigned)re_indexes[r] + layer) + 1;
      /* plain descrambling up to the element before the placeholder (exclusive) */
      while (pos + 1 < y) {
        out[pos] = c[chip] ? (int8_t)(-in[pos]) : in[pos];
        ++pos, ++chip;
      }
      const int toggle = c[chip]; /* mask of the next element; then 1 + nx chips are discarded */
      chip += 2 + nx;
      for (unsigned k = 0; k != 2; ++k, ++pos) {
        out[pos] = toggle ? (int8_t)(-in[pos]) : in[pos];
      }
      for (unsigned k = 0; k != nx; ++k, ++pos) {
        out[pos] = in[pos];
      }
    }
  }
  for (; pos < n; ++pos, ++chip) {
    out[pos] = c[chip] ? (int8_t)(-in[pos]) : in[pos];
  }
  free(c);
}
