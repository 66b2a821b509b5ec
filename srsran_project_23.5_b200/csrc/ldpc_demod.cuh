// Soft demodulator + LLR descrambler: the upstream producer of the LDPC receive chain's input, so that
// the int8 LLRs are born in HBM instead of crossing PCIe.
//
// Replaces (paths relative to the srsRAN Project 23.5 tree)
//   demodulation_mapper::demodulate_soft            include/srsran/phy/upper/channel_modulation/demodulation_mapper.h
//                                                   lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:36-106
//     generic per-symbol forms                      demodulation_mapper_qpsk.cpp:120-128, _qam16.cpp:180-206,
//                                                   _qam64.cpp:266-332, _qam256.cpp:279-350, _intervals.h:36-64
//   log_likelihood_ratio::quantize                  lib/phy/upper/log_likelihood_ratio.cpp:87-96
//   pusch_demodulator_impl::descramble (no repetition placeholders) + pseudo_random_generator::apply_xor
//                                                   lib/phy/upper/channel_processors/pusch_demodulator_impl.cpp:100-152
//                                                   lib/phy/upper/sequence_generators/pseudo_random_generator_impl.cpp:384-470
//
// Floating point: every operation is an explicit round-to-nearest intrinsic (no compiler contraction) in the order of the
// reference's generic code; the two multiply-adds the reference build contracts (gcc -mfma) are __fmaf_rn. Result:
// bit-exact with the reference's generic path (and within one LLR step of its AVX2 kernels, which differ from its own
// generic code): tests/test_gpu_demod.py.
//
// Work split: a block of 256 threads produces 1, 2, 4 or 8 spans of 8192 consecutive LLRs of one job (the launch picks the
// span count from the amount of work). Phase 1 (descrambling only): each warp jumps the two 31-bit LFSRs of TS 38.211
// 5.2.1 to its first chip with GF(2) matrix powers A^(2^j) applied co-operatively (one row per lane, ballot); lane L takes
// the state after L more products with A^256 and produces its eight 32-chip words alone, into shared memory. Phase 2, per
// modulation: 256QAM one symbol (8 LLRs, 8-byte store) per thread and step, four in flight; 64QAM one symbol (6 LLRs);
// 16QAM / QPSK 8 LLRs; BPSK 4 LLRs; existing LLRs (descramble only) 16 per thread and step. HBM-bound by design: 1 byte
// out + 12 / Qm bytes in per LLR; measured figures in profiles/r01_ncu_demod_summary.txt and profiles/r01_sweep_demod.md.
#pragma once

#include "ldpc_device.cuh"

namespace ldpc_b200 {

struct DemodJob {
  unsigned long long sym_offset;  // first symbol (complex float) / noise variance of the job; LLR index for mod_order 0
  unsigned long long llr_offset;  // first output LLR, bytes from the output base
  unsigned nof_llrs;
  unsigned c_init;                // scrambling initial state: rnti * 2^15 + n_id (pusch_demodulator_impl.cpp:105)
  unsigned char mod_order;        // 1, 2, 4, 6, 8; 0: the input already is int8 LLRs (descramble only)
  unsigned char pi2;              // pi/2-BPSK (mod_order 1)
  unsigned char descramble;
  unsigned char reserved0;
  unsigned reserved1;
};

// Piecewise-linear max-log tables of 64QAM (m = 0) and 256QAM (m = 1), per axis bit k (0 = sign bit); filled by the host
// from the TS 38.211 5.1 bit mapping (build_demod_tables in ldpc_api.cu) with the reference's float expressions.
struct DemodTables {
  float width[2][4];
  float rcp_width[2][4];     // hi = RN(1 / width)
  float rcp_width_lo[2][4];  // lo = RN(1 / width - hi)
  int nof_intervals[2][4];
  float slope[2][4][16];
  float intercept[2][4][16];
  float qam16_gain;       // 4 / sqrt(10)
  float qam16_threshold;  // 2 / sqrt(10)
};
__constant__ DemodTables c_demod;

// Rows of A^(2^j) for the x1 (index 0) and x2 (index 1) shift registers, state bit i = x(n + i).
__device__ unsigned g_prg_pow[2][32][32];

// 32-chip words of the x1 sequence from chip Nc = 1600 on (word w = chips 1600 + 32 w ... + 31, chip i = bit i): x1 starts
// from a fixed state, so its first 2^21 chips are computed once on the host (256 KB, L2-resident).
constexpr unsigned DEMOD_X1_WORDS = 1u << 16;
__device__ __align__(16) unsigned g_x1_words[DEMOD_X1_WORDS];

constexpr int DEMOD_THREADS = 256;
constexpr int DEMOD_CHUNK = DEMOD_THREADS * 32;  // LLRs per block and span
#ifndef DEMOD_MIN_BLOCKS
#define DEMOD_MIN_BLOCKS 8
#endif
#ifndef DEMOD_BATCH
#define DEMOD_BATCH 4
#endif
constexpr unsigned DEMOD_QAM256_BATCH = DEMOD_BATCH;  // symbols in flight per thread; divides 32 / 8 steps per span
constexpr int DEMOD_MAX_SPANS = 8;               // spans per block: the launch picks 1, 2, 4 or 8 from the amount of work
constexpr float DEMOD_2SQRT2 = 2.0f * 1.41421356237309504880f;

// a / b correctly rounded with rb = RN(1 / b) (Markstein: q0 = RN(a rb), r = a - b q0 exactly, q = RN(q0 + r rb)), without
// the reciprocal approximation and range checks of a general division. Falls back to the IEEE division when the
// intermediate leaves the range in which the residual is exact.
__device__ __forceinline__ float demod_div(float a, float b, float rb)
{
  const float q0 = __fmul_rn(a, rb);
  const float r = __fmaf_rn(-b, q0, a);
  const float q = __fmaf_rn(r, rb, q0);
  return (fabsf(q0) < 1e30f && fabsf(rb) > 1e-30f) ? q : __fdiv_rn(a, b);
}

// Same without the fallback, for operands known to stay in range.
__device__ __forceinline__ float demod_div_markstein(float a, float b, float rb)
{
  const float q0 = __fmul_rn(a, rb);
  return __fmaf_rn(__fmaf_rn(-b, q0, a), rb, q0);
}

// a / d correctly rounded for a CONSTANT divisor in two operations (Brisebarre-Muller: 1 / d = hi + lo with
// hi = RN(1 / d), lo = RN(1 / d - hi); q = RN(a hi + RN(a lo))). Unlike Markstein's sequence this is not correct for
// every divisor: it is verified exhaustively, over all 2^23 significands, for the four divisors used here (quantiser
// ranges 20 and 24, coarse interval widths of 64QAM and 256QAM) by tests/test_demod_constants.py.
__device__ __forceinline__ float demod_div_const(float a, float hi, float lo)
{
  return __fmaf_rn(a, hi, __fmul_rn(a, lo));
}

// log_likelihood_ratio::quantize (log_likelihood_ratio.cpp:87-96): clip to +-RANGE, / RANGE * 120, std::round (half away
// from zero), int8. Every step is odd in the value, so magnitude and sign are carried apart: the magnitude of 4 LLRs is
// formed here, the sign bits join the scrambling chips in one byte mask (demod_quantize4).
//   y  = min(|value|, RANGE)                      (= |clipped|)
//   a  = RN(y / RANGE)                            (two-operation division by a constant, demod_div_const)
//   b  = RN(a * 120) * 2^17 = RN(a * 15728640)    (scaling by a power of two commutes with the rounding)
//   |q| = floor(b / 2^17 + 0.5) = bits 16-22 of RZ(b + 2^16 + 2^24): the sum lies in [2^24, 2^25), unit in the last place
//         2, so one round-toward-zero addition leaves floor((b + 2^16) / 2) in the mantissa
// Returns the bit pattern whose byte 2 is 0x80 | |q| (|q| = 0..120; bit 7 is the odd exponent's last bit and serves as a
// guard bit for the byte-wise negation below).
// NANSAFE: any value; a NaN gives |q| = 0 like the float -> int conversion of the reference's platform. Otherwise the
// caller guarantees |value| < 1e30 (no NaN, no infinity) and the clip costs no instruction.
template <int RANGE, bool NANSAFE>
__device__ __forceinline__ unsigned demod_qmag(float value)
{
  constexpr float range_limit = (float)RANGE, hi = 1.0f / (float)RANGE,
                  lo = (float)(1.0 / (double)RANGE - (double)(1.0f / (float)RANGE));
  float a;
  if (NANSAFE) {
    const float y = fminf(fmaxf(fabsf(value), 0.0f), range_limit);  // fmaxf drops a NaN
    a = demod_div_const(y, hi, lo);
  } else {
    // min(RN(|value| / RANGE), 1) = RN(min(|value|, RANGE) / RANGE) (rounding is monotone): the clip rides on the last
    // multiply-add of the division as its saturation modifier. Needs |value| * hi finite (|value| < 1e30).
    const float y = fabsf(value);
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(a) : "f"(y), "f"(hi), "f"(__fmul_rn(y, lo)));
  }
  const float b = __fmul_rn(a, 15728640.0f);
  return __float_as_uint(__fadd_rz(b, 16842752.0f));
}

// byte 0 = 0xff if a < 0 else 0, byte 1 the same for b: prmt with the selector's replicate-sign bit (bit 3 of a nibble),
// which __byte_perm does not document.
__device__ __forceinline__ unsigned demod_sign_bytes(float a, float b)
{
  unsigned r;
  asm("prmt.b32 %0, %1, %2, 0x00fb;" : "=r"(r) : "r"(__float_as_uint(a)), "r"(__float_as_uint(b)));
  return r;
}

// bit u of chips4 -> bit 8 u (u = 0..3)
__device__ __forceinline__ unsigned demod_chip_lsb(unsigned chips4) { return ((chips4 & 0xfu) * 0x00204081u) & 0x01010101u; }

// Four LLR values -> four int8 (byte u = LLR u), descrambled with chip u = bit u of chips4. The bytes whose value is
// negative XOR whose chip is 1 are negated in two's complement; zero stays zero, so the sign of a zero or NaN value does
// not matter. Per byte, with g = 0x80 | |q| and n = 1 where the byte is negated: g - n never borrows from the next byte
// (g >= 0x80); XOR with 0x7f where negated (0x80 + |q| - 1 -> 0x100 - |q|, and 0x7f -> 0 for |q| = 0), with 0x80 where kept.
template <int RANGE, bool NANSAFE>
__device__ __forceinline__ unsigned demod_quantize4(float l0, float l1, float l2, float l3, unsigned chips4)
{
  const unsigned u0 = demod_qmag<RANGE, NANSAFE>(l0), u1 = demod_qmag<RANGE, NANSAFE>(l1),
                 u2 = demod_qmag<RANGE, NANSAFE>(l2), u3 = demod_qmag<RANGE, NANSAFE>(l3);
  // byte 2 of each magnitude word; sign bit of each value replicated over a byte (PRMT selector bit 3)
  const unsigned mag = __byte_perm(__byte_perm(u0, u1, 0x0062), __byte_perm(u2, u3, 0x0062), 0x5410);
  const unsigned sgn = __byte_perm(demod_sign_bytes(l0, l1), demod_sign_bytes(l2, l3), 0x5410);
  const unsigned chip_lsb = demod_chip_lsb(chips4);
  const unsigned m = sgn ^ (chip_lsb * 0xffu);
  const unsigned one = (sgn & 0x01010101u) ^ chip_lsb;
  return (mag - one) ^ m ^ 0x80808080u;
}

// Descrambling of four int8 LLRs that already exist (mod_order 0; any byte value): (v ^ 0xff) + 1 per byte without
// carries across bytes.
__device__ __forceinline__ unsigned demod_negate4(unsigned packed, unsigned chips4)
{
  const unsigned one = demod_chip_lsb(chips4);
  const unsigned x = packed ^ (one * 0xffu);
  return ((x & 0x7f7f7f7fu) + one) ^ (x & 0x80808080u);
}

// Per-block copy of the job's interval tables (divergent indices: shared memory instead of the constant cache);
// (slope, intercept) pairs: one 8-byte load per LLR, conflict-free (16 entries = 32 banks, equal indices broadcast).
// Every axis bit is indexed with the interval of the COARSE bits: the finest bit of the reference uses intervals exactly
// twice as wide and half as many (quotient halved exactly, floor(q / 2) = floor(q) >> 1, clamp(i, 0, L - 1) >> 1 =
// clamp(i >> 1, 0, L / 2 - 1)), so its line holds every entry twice.
struct DemodSmem {
  float2 line[4][16];
};

// (int)floorf(value / width) + L / 2 of the reference (demodulation_mapper_intervals.h:36-64) for the coarse width,
// clamped to 0 .. L - 1, wrapping like the reference's platform: quotients beyond the int range convert to INT_MIN there
// (interval 0 for every bit). Here the conversion saturates, INT_MAX + L / 2 wraps to a negative number and the clamp
// gives interval 0 as well; an infinite or NaN value leaves the Markstein sequence as NaN, which is sent to INT_MIN
// explicitly unless the caller knows the value is finite (FINITE).
template <int NB, bool FINITE>
__device__ __forceinline__ int demod_interval_index(float value)
{
  constexpr int L = 1 << NB, m = NB - 3;
  const float q = demod_div_const(value, c_demod.rcp_width[m][0], c_demod.rcp_width_lo[m][0]);
  const int i = (FINITE || q == q) ? __float2int_rd(q) : (int)0x80000000;
  return max(0, min(L - 1, (int)((unsigned)i + (unsigned)(L / 2))));
}

// interval_function of one axis value for all NB axis bits of 64QAM (NB = 3) / 256QAM (NB = 4): out[2 k] = LLR of axis
// bit k before quantisation.
template <int NB, bool FINITE>
__device__ __forceinline__ void demod_axis(const DemodSmem& tab, float x, float rcp_noise, float* out)
{
  const int ic = demod_interval_index<NB, FINITE>(x);
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    const float2 t = tab.line[k][ic];
    out[2 * k] = __fmul_rn(__fmaf_rn(t.x, x, t.y), rcp_noise);
  }
}

// Four consecutive LLRs j .. j + 3 (j a multiple of 4) of a job of modulation order Q <= 4, descrambled with the low four
// bits of chips4, one per byte. valid = number of LLRs that exist (1..4); the others come out as 0. Straight-line per
// modulation: the symbol, its noise variance and the reciprocal are fetched once.
template <int Q>
__device__ __forceinline__ unsigned demod_four(const DemodJob& job, const DemodSmem& tab, unsigned j, unsigned valid,
                                               unsigned chips4, const float2* __restrict__ sym,
                                               const float* __restrict__ noise)
{
  float l[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  if (Q == 4) {
    const unsigned s = j >> 2;
    const float2 z = __ldg(sym + s);
    const float nv = __ldg(noise + s);
    if (nv > 0) {  // false for NaN
      const float rcp = __fdiv_rn(1.0f, nv);
      const float g = c_demod.qam16_gain, th = c_demod.qam16_threshold;
      float l0 = __fmul_rn(g, z.x), l1 = __fmul_rn(g, z.y);
      if (fabsf(z.x) > th) {
        l0 = __fsub_rn(__fmul_rn(2.0f, l0), copysignf(0.8f, z.x));
      }
      if (fabsf(z.y) > th) {
        l1 = __fsub_rn(__fmul_rn(2.0f, l1), copysignf(0.8f, z.y));
      }
      l[0] = demod_div(l0, nv, rcp);
      l[1] = demod_div(l1, nv, rcp);
      l[2] = demod_div(__fmaf_rn(-g, fabsf(z.x), 0.8f), nv, rcp);
      l[3] = demod_div(__fmaf_rn(-g, fabsf(z.y), 0.8f), nv, rcp);
    }
    return demod_quantize4<20, true>(l[0], l[1], l[2], l[3], chips4);
  }
  if (Q == 2) {
#pragma unroll
    for (unsigned h = 0; h < 2; ++h) {
      if (2 * h < valid) {
        const float2 z = __ldg(sym + (j >> 1) + h);
        const float nv = __ldg(noise + (j >> 1) + h);
        if (nv > 0) {
          const float rcp = __fdiv_rn(1.0f, nv);
          l[2 * h] = demod_div(__fmul_rn(DEMOD_2SQRT2, z.x), nv, rcp);
          l[2 * h + 1] = demod_div(__fmul_rn(DEMOD_2SQRT2, z.y), nv, rcp);
        }
      }
    }
    return demod_quantize4<24, true>(l[0], l[1], l[2], l[3], chips4);
  }
  if (Q == 1) {
#pragma unroll
    for (unsigned u = 0; u < 4; ++u) {
      if (u < valid) {
        const float2 z = __ldg(sym + j + u);
        const float nv = __ldg(noise + j + u);
        if (nv > 0) {
          const float sum = (job.pi2 && (u & 1u)) ? __fadd_rn(z.y, -z.x) : __fadd_rn(z.x, z.y);
          l[u] = demod_div(__fmul_rn(DEMOD_2SQRT2, sum), nv, __fdiv_rn(1.0f, nv));
        }
      }
    }
    return demod_quantize4<24, true>(l[0], l[1], l[2], l[3], chips4);
  }
  return 0u;
}

// The eight LLRs of a 256QAM symbol (demodulation_mapper_qam256.cpp:279-350: LLR 2 k = axis bit k of the real part,
// LLR 2 k + 1 of the imaginary part), descrambled with the low eight bits of chips8. rcp = RN(1 / noise variance) or 0.
// TAME: both parts and rcp are below 1e14 in magnitude, so every LLR value stays below 1e30 (|slope| < 3, |intercept| < 8)
// and neither NaN nor infinity can arise: the NaN guards of the interval index and the quantiser are left out.
template <bool TAME>
__device__ __forceinline__ uint2 demod_symbol_qam256(const DemodSmem& tab, float2 z, float rcp, unsigned chips8)
{
  float l[8];
  demod_axis<4, TAME>(tab, z.x, rcp, l);
  demod_axis<4, TAME>(tab, z.y, rcp, l + 1);
  uint2 r;
  r.x = demod_quantize4<20, !TAME>(l[0], l[1], l[2], l[3], chips8);
  r.y = demod_quantize4<20, !TAME>(l[4], l[5], l[6], l[7], chips8 >> 4);
  return r;
}

// x1 / x2 advanced by k <= 28 chips: the 28 new bits x(n+31..n+58) follow from the state in one step.
__device__ __forceinline__ unsigned prg_advance(unsigned s, unsigned k, bool is_x2)
{
  const unsigned nb = is_x2 ? ((s >> 3) ^ (s >> 2) ^ (s >> 1) ^ s) : ((s >> 3) ^ s);
  const unsigned long long w = (unsigned long long)s | ((unsigned long long)(nb & 0x0fffffffu) << 31);
  return (unsigned)(w >> k) & 0x7fffffffu;
}

// Phase 2 of a block for every modulation but 256QAM: 4 consecutive LLRs (one packed store) per thread and step.
template <int Q>
__device__ __forceinline__ void demod_loop4(const DemodJob& job, const DemodSmem& tab, const unsigned* chips, unsigned j0,
                                            unsigned spans, const float2* __restrict__ sym,
                                            const float* __restrict__ noise, const int8_t* in, int8_t* dst)
{
  const unsigned t = threadIdx.x;
  const bool aligned = ((reinterpret_cast<unsigned long long>(dst) & 3ull) == 0);
  const bool in_aligned = ((reinterpret_cast<unsigned long long>(in) & 3ull) == 0);
#pragma unroll 2
  for (unsigned it = 0; it < spans * (32u / 4u); ++it) {
    const unsigned local = 4u * (it * DEMOD_THREADS + t);  // first of this thread's 4 LLRs inside the chunk
    const unsigned j = j0 + local;
    if (j >= job.nof_llrs) {
      break;
    }
    const unsigned cw = chips[local >> 5] >> (local & 31u);
    const unsigned valid = min(4u, job.nof_llrs - j);
    unsigned packed;
    if (Q != 0) {
      packed = demod_four<Q>(job, tab, j, valid, cw, sym, noise);
    } else {
      packed = 0;
      if (in_aligned && valid == 4) {
        packed = *reinterpret_cast<const unsigned*>(in + j);
      } else {
        for (unsigned u = 0; u < valid; ++u) {
          packed |= ((unsigned)in[j + u] & 0xffu) << (8u * u);
        }
      }
      packed = demod_negate4(packed, cw);
    }
    if (aligned && valid == 4) {
      *reinterpret_cast<unsigned*>(dst + j) = packed;
    } else {
      for (unsigned u = 0; u < valid; ++u) {
        dst[j + u] = (int8_t)(packed >> (8u * u));
      }
    }
  }
}

// Phase 2 for QPSK / 16QAM: 8 consecutive LLRs (4 or 2 symbols, one 8-byte store) per thread and step. When all
// of them are "tame" (|re|, |im|, |noise variance| and its reciprocal below 1e14: nothing can overflow, no NaN) the
// divisions by the noise variance are plain Markstein sequences and the quantiser clips for free; a noise variance <= 0
// gives rcp = 0, hence q0 = 0 and LLR 0 as in the reference. Anything else, and the ragged end of a job, goes through
// demod_four.
template <int Q>
__device__ __forceinline__ void demod_loop8(const DemodJob& job, const DemodSmem& tab, const unsigned* chips, unsigned j0,
                                            unsigned spans, const float2* __restrict__ sym,
                                            const float* __restrict__ noise, int8_t* dst)
{
  constexpr unsigned NS = 8u / Q;  // symbols per step
  const unsigned t = threadIdx.x;
  const unsigned long long dst_bits = reinterpret_cast<unsigned long long>(dst);
#pragma unroll 2
  for (unsigned it = 0; it < spans * (32u / 8u); ++it) {
    const unsigned local = 8u * (it * DEMOD_THREADS + t);
    const unsigned j = j0 + local;
    if (j >= job.nof_llrs) {
      break;
    }
    const unsigned cw = chips[local >> 5] >> (local & 31u);
    const bool full = j + 8u <= job.nof_llrs;
    float2 z[NS];
    float nv[NS], rcp[NS];
    bool tame = full;
    if (full) {
#pragma unroll
      for (unsigned i = 0; i < NS; ++i) {
        z[i] = __ldg(sym + j / Q + i);
        nv[i] = __ldg(noise + j / Q + i);
      }
#pragma unroll
      for (unsigned i = 0; i < NS; ++i) {
        rcp[i] = (nv[i] > 0) ? __frcp_rn(nv[i]) : 0.0f;
        tame = tame && fabsf(z[i].x) < 1e14f && fabsf(z[i].y) < 1e14f && fabsf(nv[i]) < 1e14f && rcp[i] < 1e14f;
      }
    }
    uint2 w;
    if (tame) {
      float l[8];
#pragma unroll
      for (unsigned i = 0; i < NS; ++i) {
        if (Q == 4) {
          const float g = c_demod.qam16_gain, th = c_demod.qam16_threshold;
          float l0 = __fmul_rn(g, z[i].x), l1 = __fmul_rn(g, z[i].y);
          if (fabsf(z[i].x) > th) {
            l0 = __fsub_rn(__fmul_rn(2.0f, l0), copysignf(0.8f, z[i].x));
          }
          if (fabsf(z[i].y) > th) {
            l1 = __fsub_rn(__fmul_rn(2.0f, l1), copysignf(0.8f, z[i].y));
          }
          l[4 * i] = demod_div_markstein(l0, nv[i], rcp[i]);
          l[4 * i + 1] = demod_div_markstein(l1, nv[i], rcp[i]);
          l[4 * i + 2] = demod_div_markstein(__fmaf_rn(-g, fabsf(z[i].x), 0.8f), nv[i], rcp[i]);
          l[4 * i + 3] = demod_div_markstein(__fmaf_rn(-g, fabsf(z[i].y), 0.8f), nv[i], rcp[i]);
        } else if (Q == 2) {
          l[2 * i] = demod_div_markstein(__fmul_rn(DEMOD_2SQRT2, z[i].x), nv[i], rcp[i]);
          l[2 * i + 1] = demod_div_markstein(__fmul_rn(DEMOD_2SQRT2, z[i].y), nv[i], rcp[i]);
        } else {
          const float sum = (job.pi2 && (i & 1u)) ? __fadd_rn(z[i].y, -z[i].x) : __fadd_rn(z[i].x, z[i].y);
          l[i] = demod_div_markstein(__fmul_rn(DEMOD_2SQRT2, sum), nv[i], rcp[i]);
        }
      }
      constexpr int RANGE = (Q == 4) ? 20 : 24;
      w.x = demod_quantize4<RANGE, false>(l[0], l[1], l[2], l[3], cw);
      w.y = demod_quantize4<RANGE, false>(l[4], l[5], l[6], l[7], cw >> 4);
    } else {
      const unsigned left = job.nof_llrs - j;
      w.x = demod_four<Q>(job, tab, j, min(4u, left), cw, sym, noise);
      w.y = (left > 4u) ? demod_four<Q>(job, tab, j + 4u, min(4u, left - 4u), cw >> 4, sym, noise) : 0u;
    }
    if (full && (dst_bits & 7ull) == 0) {
      *reinterpret_cast<uint2*>(dst + j) = w;
    } else if (full && (dst_bits & 3ull) == 0) {
      reinterpret_cast<unsigned*>(dst + j)[0] = w.x;
      reinterpret_cast<unsigned*>(dst + j)[1] = w.y;
    } else {
      for (unsigned u = 0; u < 8u && j + u < job.nof_llrs; ++u) {
        dst[j + u] = (int8_t)((u < 4u ? w.x : w.y) >> (8u * (u & 3u)));
      }
    }
  }
}

// Phase 2 for 64QAM: one symbol = 6 LLRs per thread and step. The block's LLR range [j0, end) is not a multiple of 6, so
// the symbols that straddle its two ends are computed by both neighbouring blocks and each writes its own bytes (pairs of
// LLRs never straddle: every bound is even). chips[-1] and chips[256 spans] exist as padding (their content only reaches
// bytes that are not written).
__device__ __forceinline__ void demod_loop_qam64(const DemodJob& job, const DemodSmem& tab, const unsigned* chips,
                                                 unsigned j0, unsigned spans, const float2* __restrict__ sym,
                                                 const float* __restrict__ noise, int8_t* dst)
{
  const unsigned end = min(j0 + DEMOD_CHUNK * spans, job.nof_llrs);
  const unsigned s_last = (end - 1u) / 6u;
  const bool even = ((reinterpret_cast<unsigned long long>(dst) & 1ull) == 0);
#pragma unroll 2
  for (unsigned s = j0 / 6u + threadIdx.x; s <= s_last; s += DEMOD_THREADS) {
    const float2 z = __ldg(sym + s);
    const float nv = __ldg(noise + s);
    const float rcp = (nv > 0) ? __frcp_rn(nv) : 0.0f;  // not > 0 covers NaN
    const int local = (int)(6u * s) - (int)j0;          // >= -5
    const unsigned wi = (unsigned)(local + 32) >> 5, sh = (unsigned)(local + 32) & 31u;
    const unsigned cw = __funnelshift_r(chips[(int)wi - 1], chips[wi], sh);
    float l[8];
    l[6] = l[7] = 0.0f;
    unsigned w0, w1;
    if (fabsf(z.x) < 1e14f && fabsf(z.y) < 1e14f && rcp < 1e14f) {
      demod_axis<3, true>(tab, z.x, rcp, l);
      demod_axis<3, true>(tab, z.y, rcp, l + 1);
      w0 = demod_quantize4<20, false>(l[0], l[1], l[2], l[3], cw);
      w1 = demod_quantize4<20, false>(l[4], l[5], l[6], l[7], cw >> 4);
    } else {
      demod_axis<3, false>(tab, z.x, rcp, l);
      demod_axis<3, false>(tab, z.y, rcp, l + 1);
      w0 = demod_quantize4<20, true>(l[0], l[1], l[2], l[3], cw);
      w1 = demod_quantize4<20, true>(l[4], l[5], l[6], l[7], cw >> 4);
    }
#pragma unroll
    for (unsigned u = 0; u < 6u; u += 2u) {
      const unsigned j = 6u * s + u;
      if (j >= j0 && j < end) {
        const unsigned pair = ((u < 4u ? w0 : w1) >> (8u * (u & 3u))) & 0xffffu;
        if (even) {
          *reinterpret_cast<unsigned short*>(dst + j) = (unsigned short)pair;
        } else {
          dst[j] = (int8_t)pair;
          dst[j + 1] = (int8_t)(pair >> 8);
        }
      }
    }
  }
}

// Phase 1 of a warp: the 32-chip words of its spans x 1024 chips starting at chip j0 + first of the job, into
// chips[(first / 32) ...]. A warp owns spans x 1024 consecutive chips. The state of the registers at chip 1600 + first chip
// of the warp (Nc = 1600, TS 38.211 5.2.1) is reached with the matrix powers A^(2^j) (one row per lane, ballot = matrix x
// vector). The chips then form groups of 256 (4 per span, at most 32): lane L owns group L; its start state is reached by L
// co-operative products with A^256 (the rows stay in registers, every lane keeps the state of its own step); from there
// the lane produces its eight words alone, advancing by 28 + 4 chips in between. (Earlier versions: every lane stepping
// from the warp's start, up to 36 steps, was a third of the kernel's instructions; groups of 4 lanes per span with 8
// products per span still a quarter: profiles/r01_ncu_demod_summary.txt.) X1_TABLE: only x2 is generated, the words of x1
// come from g_x1_words.
template <bool X1_TABLE>
__device__ __forceinline__ void demod_chip_words(const DemodJob& job, unsigned* chips, unsigned j0, unsigned first,
                                                 unsigned spans, unsigned lane)
{
  const unsigned n = 1600u + j0 + first;
  unsigned s1 = 1u, s2 = job.c_init & 0x7fffffffu;
  for (unsigned jb = 0; jb < 32; ++jb) {
    if ((n >> jb) & 1u) {
      if (!X1_TABLE) {
        s1 = __ballot_sync(0xffffffffu, __popc(g_prg_pow[0][jb][lane] & s1) & 1);
      }
      s2 = __ballot_sync(0xffffffffu, __popc(g_prg_pow[1][jb][lane] & s2) & 1);
    }
  }
  const unsigned left = (j0 + first < job.nof_llrs) ? job.nof_llrs - (j0 + first) : 0u;
  const unsigned groups = min(4u * spans, (left + 255u) / 256u);
  const unsigned q1 = X1_TABLE ? 0u : g_prg_pow[0][8][lane], q2 = g_prg_pow[1][8][lane];
  unsigned m1 = s1, m2 = s2;
  for (unsigned g = 1; g < groups; ++g) {
    if (!X1_TABLE) {
      s1 = __ballot_sync(0xffffffffu, __popc(q1 & s1) & 1);
    }
    s2 = __ballot_sync(0xffffffffu, __popc(q2 & s2) & 1);
    if (lane == g) {
      m1 = s1, m2 = s2;
    }
  }
  if (lane < groups) {
    uint4* dst_words = reinterpret_cast<uint4*>(&chips[first / 32u + 8u * lane]);
    const uint4* x1_words = reinterpret_cast<const uint4*>(&g_x1_words[X1_TABLE ? (j0 + first) / 32u + 8u * lane : 0u]);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      unsigned w[4];
      uint4 x1 = make_uint4(0u, 0u, 0u, 0u);
      if (X1_TABLE) {
        x1 = __ldg(x1_words + h);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned b2 = ((m2 >> 3) ^ (m2 >> 2) ^ (m2 >> 1) ^ m2) & 1u;
        if (X1_TABLE) {
          w[k] = (k == 0 ? x1.x : k == 1 ? x1.y : k == 2 ? x1.z : x1.w) ^ (m2 | (b2 << 31));
        } else {
          const unsigned b1 = ((m1 >> 3) ^ m1) & 1u;
          w[k] = (m1 ^ m2) | ((b1 ^ b2) << 31);  // chip i of the word = bit i
        }
        if (h == 0 || k < 3) {
          if (!X1_TABLE) {
            m1 = prg_advance(prg_advance(m1, 28u, false), 4u, false);
          }
          m2 = prg_advance(prg_advance(m2, 28u, true), 4u, true);
        }
      }
      dst_words[h] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

// grid (chunks, jobs). in_llrs: input of descramble-only jobs.
__global__ void __launch_bounds__(DEMOD_THREADS, DEMOD_MIN_BLOCKS) ldpc_demod_kernel(const DemodJob* __restrict__ jobs,
                                                                   const float2* __restrict__ symbols,
                                                                   const float* __restrict__ noise_vars,
                                                                   const int8_t* __restrict__ in_llrs,
                                                                   int8_t* __restrict__ out, unsigned spans)
{
  const DemodJob job = jobs[blockIdx.y];
  const unsigned chunk = DEMOD_CHUNK * spans;  // LLRs of this block: spans (1..DEMOD_MAX_SPANS) x 1024 per warp
  const unsigned j0 = blockIdx.x * chunk;
  if (j0 >= job.nof_llrs) {
    return;
  }
  const unsigned t = threadIdx.x, lane = t & 31u;
  __shared__ __align__(16) unsigned chips_padded[DEMOD_THREADS * DEMOD_MAX_SPANS + 8];
  unsigned* const chips = chips_padded + 4;  // one word of padding on either side for the 64QAM loop, 16-byte alignment kept
  __shared__ DemodSmem tab;
  if (job.mod_order >= 6 && t < 64) {
    const int m = job.mod_order == 8 ? 1 : 0;
    const unsigned k = t >> 4, i = (k == (m ? 3u : 2u)) ? ((t & 15u) >> 1) : (t & 15u);  // finest bit: every entry twice
    tab.line[k][t & 15] = make_float2(c_demod.slope[m][k][i], c_demod.intercept[m][k][i]);
  }
  if (job.descramble) {
    // x1 does not depend on c_init: while the warp's chips lie inside the precomputed part of its sequence, its words
    // are read from a table and only x2 is generated
    const unsigned first = (t >> 5) * 1024u * spans;
    if ((j0 + first) / 32u + 32u * spans <= DEMOD_X1_WORDS) {
      demod_chip_words<true>(job, chips, j0, first, spans, lane);
    } else {
      demod_chip_words<false>(job, chips, j0, first, spans, lane);
    }
  } else {
    for (unsigned k = t; k < DEMOD_THREADS * spans; k += DEMOD_THREADS) {
      chips[k] = 0u;  // no descrambling: every chip 0
    }
  }
  __syncthreads();
  const float2* sym = symbols + job.sym_offset;
  const float* noise = noise_vars + job.sym_offset;
  const int8_t* in = in_llrs + job.sym_offset;
  int8_t* dst = out + job.llr_offset;
  const unsigned long long dst_bits = reinterpret_cast<unsigned long long>(dst);
  if (job.mod_order == 8) {
    // one symbol = 8 LLRs = one 8-byte store per thread and step (nof_llrs is a multiple of 8)
    // and span: DEMOD_QAM256_BATCH symbols in flight per thread (loads first, one address computation)
#pragma unroll 1
    for (unsigned it = 0; it < spans * (32u / 8u); it += DEMOD_QAM256_BATCH) {
      const unsigned local = 8u * (it * DEMOD_THREADS + t);
      const unsigned j = j0 + local;
      if (j >= job.nof_llrs) {
        break;
      }
      const float2* ps = sym + (j >> 3);
      const float* pn = noise + (j >> 3);
      float2 z[DEMOD_QAM256_BATCH];
      float nv[DEMOD_QAM256_BATCH];
#pragma unroll
      for (unsigned i = 0; i < DEMOD_QAM256_BATCH; ++i) {
        const bool ok = j + 8u * DEMOD_THREADS * i < job.nof_llrs;
        z[i] = ok ? __ldg(ps + DEMOD_THREADS * i) : make_float2(0.0f, 0.0f);
        nv[i] = ok ? __ldg(pn + DEMOD_THREADS * i) : 0.0f;
      }
      // rcp = RN(1 / noise variance) or 0 (not > 0 covers NaN); symbols outside the batch count as (0, 0) / 0
      float rcp[DEMOD_QAM256_BATCH];
      bool tame = true;
#pragma unroll
      for (unsigned i = 0; i < DEMOD_QAM256_BATCH; ++i) {
        rcp[i] = (nv[i] > 0) ? __frcp_rn(nv[i]) : 0.0f;
        tame = tame && fabsf(z[i].x) < 1e14f && fabsf(z[i].y) < 1e14f && rcp[i] < 1e14f;
      }
      uint2 w[DEMOD_QAM256_BATCH];
      if (tame) {
#pragma unroll
        for (unsigned i = 0; i < DEMOD_QAM256_BATCH; ++i) {
          const unsigned li = local + 8u * DEMOD_THREADS * i;
          w[i] = demod_symbol_qam256<true>(tab, z[i], rcp[i], chips[li >> 5] >> (li & 31u));
        }
      } else {
#pragma unroll
        for (unsigned i = 0; i < DEMOD_QAM256_BATCH; ++i) {
          const unsigned li = local + 8u * DEMOD_THREADS * i;
          w[i] = demod_symbol_qam256<false>(tab, z[i], rcp[i], chips[li >> 5] >> (li & 31u));
        }
      }
#pragma unroll
      for (unsigned i = 0; i < DEMOD_QAM256_BATCH; ++i) {
        const unsigned ji = j + 8u * DEMOD_THREADS * i;
        if (ji < job.nof_llrs) {
          if ((dst_bits & 7ull) == 0) {
            *reinterpret_cast<uint2*>(dst + ji) = w[i];
          } else if ((dst_bits & 3ull) == 0) {
            reinterpret_cast<unsigned*>(dst + ji)[0] = w[i].x;
            reinterpret_cast<unsigned*>(dst + ji)[1] = w[i].y;
          } else {
            for (unsigned u = 0; u < 4; ++u) {
              dst[ji + u] = (int8_t)(w[i].x >> (8u * u));
              dst[ji + 4 + u] = (int8_t)(w[i].y >> (8u * u));
            }
          }
        }
      }
    }
    return;
  }
  switch (job.mod_order) {  // uniform over the block
    case 6: demod_loop_qam64(job, tab, chips, j0, spans, sym, noise, dst); break;
    case 4: demod_loop8<4>(job, tab, chips, j0, spans, sym, noise, dst); break;
    case 2: demod_loop8<2>(job, tab, chips, j0, spans, sym, noise, dst); break;
    case 1: demod_loop4<1>(job, tab, chips, j0, spans, sym, noise, in, dst); break;  // 13 B of input per LLR: HBM-bound as it is
    default: {
      // descrambling of existing LLRs: 16 per thread and step when both buffers allow 16-byte accesses
      const unsigned long long in_bits = reinterpret_cast<unsigned long long>(in);
      if (((dst_bits | in_bits) & 15ull) == 0) {
        const unsigned t = threadIdx.x;
#pragma unroll 2
        for (unsigned it = 0; it < spans * (32u / 16u); ++it) {
          const unsigned local = 16u * (it * DEMOD_THREADS + t);
          const unsigned j = j0 + local;
          if (j + 16u > job.nof_llrs) {
            // ragged end: at most one such thread per job
            for (unsigned u = 0; j + u < job.nof_llrs; ++u) {
              const int8_t v = in[j + u];
              dst[j + u] = ((chips[(local + u) >> 5] >> ((local + u) & 31u)) & 1u) ? (int8_t)(-v) : v;
            }
            break;
          }
          const uint4 v = *reinterpret_cast<const uint4*>(in + j);
          const unsigned cw = chips[local >> 5] >> (local & 31u);
          uint4 r;
          r.x = demod_negate4(v.x, cw);
          r.y = demod_negate4(v.y, cw >> 4);
          r.z = demod_negate4(v.z, cw >> 8);
          r.w = demod_negate4(v.w, cw >> 12);
          *reinterpret_cast<uint4*>(dst + j) = r;
        }
      } else {
        demod_loop4<0>(job, tab, chips, j0, spans, sym, noise, in, dst);
      }
    }
  }
}

// UCI repetition placeholders (pusch_demodulator_impl.cpp:100-152) as an amendment of LLRs that were descrambled as plain
// data: a placeholder at soft bit y must carry the chip of bit y - 1 instead of its own, and the nx bits behind it no chip
// at all. One warp per placeholder: jump both registers to chip y - 1 (the state holds the next 31 chips), lane 0 flips
// bit y if chips y - 1 and y differ, lane k (1..nx) flips bit y + k back if its chip was 1.
struct PlaceholderWork {
  unsigned long long llr_offset;
  unsigned y;
  unsigned c_init;
  unsigned nof_llrs;
  unsigned nx;
};

__global__ void __launch_bounds__(128) ldpc_demod_placeholder_kernel(const PlaceholderWork* __restrict__ work, unsigned n,
                                                                     int8_t* llrs)
{
  const unsigned idx = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
  if (idx >= n) {
    return;
  }
  const PlaceholderWork w = work[idx];
  const unsigned pos = 1600u + w.y - 1u;
  unsigned s1 = 1u, s2 = w.c_init & 0x7fffffffu;
  for (unsigned jb = 0; jb < 32; ++jb) {
    if ((pos >> jb) & 1u) {
      const unsigned r1 = g_prg_pow[0][jb][lane], r2 = g_prg_pow[1][jb][lane];
      s1 = __ballot_sync(0xffffffffu, __popc(r1 & s1) & 1);
      s2 = __ballot_sync(0xffffffffu, __popc(r2 & s2) & 1);
    }
  }
  const unsigned chips = s1 ^ s2;  // bit k = chip of soft bit y - 1 + k, k < 31
  const unsigned q = w.y + lane;
  if (lane <= w.nx && q < w.nof_llrs) {
    const unsigned flip = (lane == 0) ? ((chips ^ (chips >> 1)) & 1u) : ((chips >> (lane + 1u)) & 1u);
    if (flip) {
      int8_t* p = llrs + w.llr_offset + q;
      *p = (int8_t)(-*p);
    }
  }
}

}  // namespace ldpc_b200
