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
// Work split: a block of 256 threads produces 8192 consecutive LLRs of one job. Phase 1 (descrambling only): each warp
// jumps the two 31-bit LFSRs of TS 38.211 5.2.1 to its first chip with GF(2) matrix powers A^(2^j) applied
// co-operatively (one row per lane, ballot), groups of 4 lanes are reached with A^128 the same way, each lane then advances
// by at most 96 chips, 28 at a time, and writes its 32-chip word to shared memory. Phase 2: a thread computes 4 consecutive LLRs per step (one packed 32-bit store, coalesced),
// 8 steps. HBM-bound by design: 1 byte out + 12 / Qm bytes in per LLR.
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
  float rcp_width[2][4];  // RN(1 / width)
  int nof_intervals[2][4];
  float slope[2][4][16];
  float intercept[2][4][16];
  float qam16_gain;       // 4 / sqrt(10)
  float qam16_threshold;  // 2 / sqrt(10)
};
__constant__ DemodTables c_demod;

// Rows of A^(2^j) for the x1 (index 0) and x2 (index 1) shift registers, state bit i = x(n + i).
__device__ unsigned g_prg_pow[2][32][32];

constexpr int DEMOD_THREADS = 256;
constexpr int DEMOD_CHUNK = DEMOD_THREADS * 32;  // LLRs per block
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

// Same for a constant divisor and a bounded dividend (interval width, quantiser range): no fallback needed.
__device__ __forceinline__ float demod_div_const(float a, float b, float rb)
{
  const float q0 = __fmul_rn(a, rb);
  return __fmaf_rn(__fmaf_rn(-b, q0, a), rb, q0);
}

// log_likelihood_ratio::quantize (log_likelihood_ratio.cpp:87-96): clip, / range * 120, std::round (half away from zero).
// RCP = RN(1 / RANGE). Rounding: nearest-even through the 1.5 * 2^23 magic constant (the integer falls out of the
// mantissa), ties corrected to "away".
template <int RANGE>
__device__ __forceinline__ int demod_quantize(float value)
{
  constexpr float range_limit = (float)RANGE, rcp = 1.0f / (float)RANGE;
  float clipped = value;
  if (fabsf(value) > range_limit) {
    clipped = copysignf(range_limit, value);
  }
  const float x = __fmul_rn(demod_div_const(clipped, range_limit, rcp), 120.0f);
  const float magic = 12582912.0f;
  float t = __fadd_rn(x, magic);
  const float diff = __fsub_rn(x, __fsub_rn(t, magic));
  if (fabsf(diff) == 0.5f) {
    t = __fadd_rn(__fadd_rn(x, copysignf(0.5f, x)), magic);
  }
  const int i = __float_as_int(t) - 0x4B400000;
  return (x == x) ? i : 0;  // NaN -> 0, as the conversion of the reference's platform
}

// Per-block copy of the job's interval tables (divergent indices: shared memory instead of the constant cache).
struct DemodSmem {
  float slope[4][16];
  float intercept[4][16];
};

// (symbols beyond +-1e30 would overflow the quotient; the reference's float -> int conversion is undefined there too)
__device__ __forceinline__ float demod_interval(const DemodSmem& tab, float value, float rcp_noise, int m, int k)
{
  const int n = c_demod.nof_intervals[m][k];
  int idx = __float2int_rd(demod_div_const(value, c_demod.width[m][k], c_demod.rcp_width[m][k])) + n / 2;
  idx = max(0, min(n - 1, idx));
  const float l = __fmaf_rn(tab.slope[k][idx], value, tab.intercept[k][idx]);
  return __fmul_rn(l, rcp_noise);
}

// Per-symbol state shared by the LLRs of a symbol.
struct DemodSym {
  unsigned index;
  float2 z;
  float nv, rcp;  // rcp = RN(1 / nv) when nv > 0, else 0
};

// LLR j of a job (before descrambling).
__device__ __forceinline__ int demod_llr(const DemodJob& job, const DemodSmem& tab, unsigned j, DemodSym& cur,
                                         const float2* __restrict__ sym, const float* __restrict__ noise)
{
  const unsigned q = job.mod_order;
  unsigned s, b;
  switch (q) {
    case 1: s = j, b = 0; break;
    case 2: s = j >> 1, b = j & 1u; break;
    case 4: s = j >> 2, b = j & 3u; break;
    case 6: s = j / 6u, b = j - 6u * s; break;
    default: s = j >> 3, b = j & 7u; break;
  }
  if (s != cur.index) {
    cur.index = s;
    cur.z = __ldg(sym + s);
    cur.nv = __ldg(noise + s);
    cur.rcp = (cur.nv > 0) ? __fdiv_rn(1.0f, cur.nv) : 0.0f;
  }
  const float2 z = cur.z;
  const float nv = cur.nv;
  const float x = (b & 1u) ? z.y : z.x;
  if (q >= 6) {
    return demod_quantize<20>(demod_interval(tab, x, cur.rcp, q == 8 ? 1 : 0, (int)(b >> 1)));
  }
  if (!(nv > 0)) {  // also false for NaN
    return 0;
  }
  if (q == 1) {
    const float sum = (job.pi2 && (j & 1u)) ? __fadd_rn(z.y, -z.x) : __fadd_rn(z.x, z.y);
    return demod_quantize<24>(demod_div(__fmul_rn(DEMOD_2SQRT2, sum), nv, cur.rcp));
  }
  if (q == 2) {
    return demod_quantize<24>(demod_div(__fmul_rn(DEMOD_2SQRT2, x), nv, cur.rcp));
  }
  float l;
  if (b < 2) {
    l = __fmul_rn(c_demod.qam16_gain, x);
    if (fabsf(x) > c_demod.qam16_threshold) {
      l = __fsub_rn(__fmul_rn(2.0f, l), copysignf(0.8f, x));
    }
  } else {
    l = __fmaf_rn(-c_demod.qam16_gain, fabsf(x), 0.8f);
  }
  return demod_quantize<20>(demod_div(l, nv, cur.rcp));
}

// Four consecutive LLRs j .. j + 3 (j a multiple of 4) of a job of modulation order Q, before descrambling, one per byte.
// valid = number of LLRs that exist (1..4). Straight-line per modulation: the symbol, its noise variance and the
// reciprocal are fetched once.
template <int Q>
__device__ __forceinline__ unsigned demod_four(const DemodJob& job, const DemodSmem& tab, unsigned j, unsigned valid,
                                               const float2* __restrict__ sym, const float* __restrict__ noise)
{
  int v[4] = {0, 0, 0, 0};
  if (Q == 8 || Q == 4) {
    const unsigned s = (Q == 8) ? (j >> 3) : (j >> 2);
    const float2 z = __ldg(sym + s);
    const float nv = __ldg(noise + s);
    const bool ok = nv > 0;  // false for NaN
    const float rcp = ok ? __fdiv_rn(1.0f, nv) : 0.0f;
    if (Q == 8) {
      const int k0 = (int)((j >> 1) & 2u);  // axis-bit level of LLRs 0,1; LLRs 2,3 are one level finer
      v[0] = demod_quantize<20>(demod_interval(tab, z.x, rcp, 1, k0));
      v[1] = demod_quantize<20>(demod_interval(tab, z.y, rcp, 1, k0));
      v[2] = demod_quantize<20>(demod_interval(tab, z.x, rcp, 1, k0 + 1));
      v[3] = demod_quantize<20>(demod_interval(tab, z.y, rcp, 1, k0 + 1));
    } else if (ok) {
      const float g = c_demod.qam16_gain, th = c_demod.qam16_threshold;
      float l0 = __fmul_rn(g, z.x), l1 = __fmul_rn(g, z.y);
      if (fabsf(z.x) > th) {
        l0 = __fsub_rn(__fmul_rn(2.0f, l0), copysignf(0.8f, z.x));
      }
      if (fabsf(z.y) > th) {
        l1 = __fsub_rn(__fmul_rn(2.0f, l1), copysignf(0.8f, z.y));
      }
      v[0] = demod_quantize<20>(demod_div(l0, nv, rcp));
      v[1] = demod_quantize<20>(demod_div(l1, nv, rcp));
      v[2] = demod_quantize<20>(demod_div(__fmaf_rn(-g, fabsf(z.x), 0.8f), nv, rcp));
      v[3] = demod_quantize<20>(demod_div(__fmaf_rn(-g, fabsf(z.y), 0.8f), nv, rcp));
    }
  } else if (Q == 2) {
#pragma unroll
    for (unsigned h = 0; h < 2; ++h) {
      if (2 * h < valid) {
        const float2 z = __ldg(sym + (j >> 1) + h);
        const float nv = __ldg(noise + (j >> 1) + h);
        if (nv > 0) {
          const float rcp = __fdiv_rn(1.0f, nv);
          v[2 * h] = demod_quantize<24>(demod_div(__fmul_rn(DEMOD_2SQRT2, z.x), nv, rcp));
          v[2 * h + 1] = demod_quantize<24>(demod_div(__fmul_rn(DEMOD_2SQRT2, z.y), nv, rcp));
        }
      }
    }
  } else if (Q == 1) {
#pragma unroll
    for (unsigned u = 0; u < 4; ++u) {
      if (u < valid) {
        const float2 z = __ldg(sym + j + u);
        const float nv = __ldg(noise + j + u);
        if (nv > 0) {
          const float sum = (job.pi2 && (u & 1u)) ? __fadd_rn(z.y, -z.x) : __fadd_rn(z.x, z.y);
          v[u] = demod_quantize<24>(demod_div(__fmul_rn(DEMOD_2SQRT2, sum), nv, __fdiv_rn(1.0f, nv)));
        }
      }
    }
  } else {  // 64QAM: the four LLRs straddle symbols
    DemodSym cur;
    cur.index = 0xffffffffu;
#pragma unroll
    for (unsigned u = 0; u < 4; ++u) {
      if (u < valid) {
        v[u] = demod_llr(job, tab, j + u, cur, sym, noise);
      }
    }
  }
  return ((unsigned)v[0] & 0xffu) | (((unsigned)v[1] & 0xffu) << 8) | (((unsigned)v[2] & 0xffu) << 16) |
         (((unsigned)v[3] & 0xffu) << 24);
}

// x1 / x2 advanced by k <= 28 chips: the 28 new bits x(n+31..n+58) follow from the state in one step.
__device__ __forceinline__ unsigned prg_advance(unsigned s, unsigned k, bool is_x2)
{
  const unsigned nb = is_x2 ? ((s >> 3) ^ (s >> 2) ^ (s >> 1) ^ s) : ((s >> 3) ^ s);
  const unsigned long long w = (unsigned long long)s | ((unsigned long long)(nb & 0x0fffffffu) << 31);
  return (unsigned)(w >> k) & 0x7fffffffu;
}

// grid (chunks, jobs). in_llrs: input of descramble-only jobs.
__global__ void __launch_bounds__(DEMOD_THREADS) ldpc_demod_kernel(const DemodJob* __restrict__ jobs,
                                                                   const float2* __restrict__ symbols,
                                                                   const float* __restrict__ noise_vars,
                                                                   const int8_t* __restrict__ in_llrs,
                                                                   int8_t* __restrict__ out)
{
  const DemodJob job = jobs[blockIdx.y];
  const unsigned j0 = blockIdx.x * DEMOD_CHUNK;
  if (j0 >= job.nof_llrs) {
    return;
  }
  const unsigned t = threadIdx.x, lane = t & 31u;
  __shared__ unsigned chips[DEMOD_THREADS];
  __shared__ DemodSmem tab;
  if (job.mod_order >= 6 && t < 64) {
    const int m = job.mod_order == 8 ? 1 : 0;
    tab.slope[t >> 4][t & 15] = c_demod.slope[m][t >> 4][t & 15];
    tab.intercept[t >> 4][t & 15] = c_demod.intercept[m][t >> 4][t & 15];
  }
  if (job.descramble) {
    // state of both registers at chip 1600 + first chip of the warp (Nc = 1600, TS 38.211 5.2.1)
    const unsigned n = 1600u + j0 + (t & ~31u) * 32u;
    unsigned s1 = 1u, s2 = job.c_init & 0x7fffffffu;
    for (unsigned jb = 0; jb < 32; ++jb) {
      if ((n >> jb) & 1u) {
        const unsigned r1 = g_prg_pow[0][jb][lane], r2 = g_prg_pow[1][jb][lane];
        s1 = __ballot_sync(0xffffffffu, __popc(r1 & s1) & 1);
        s2 = __ballot_sync(0xffffffffu, __popc(r2 & s2) & 1);
      }
    }
    // Lane L needs the state 32 L chips further. Groups of 4 lanes: the start of group g follows from the start of group
    // g - 1 by one more co-operative product with A^128 (its rows stay in registers); inside a group a lane advances by
    // 0, 32, 64 or 96 chips in steps of 28. (Advancing every lane on its own from the warp's start took up to 36 steps
    // and was a third of the kernel's instructions: profiles/r01_ncu_demod_summary.txt.)
    {
      const unsigned q1 = g_prg_pow[0][7][lane], q2 = g_prg_pow[1][7][lane];
      unsigned m1 = s1, m2 = s2;
#pragma unroll
      for (unsigned g = 1; g < 8; ++g) {
        s1 = __ballot_sync(0xffffffffu, __popc(q1 & s1) & 1);
        s2 = __ballot_sync(0xffffffffu, __popc(q2 & s2) & 1);
        if ((lane >> 2) == g) {
          m1 = s1, m2 = s2;
        }
      }
      s1 = m1, s2 = m2;
    }
    const unsigned ahead = 32u * (lane & 3u), full = ahead / 28u, rest = ahead - 28u * full;
    for (unsigned i = 0; i < full; ++i) {
      s1 = ((s1 >> 28) | (((s1 >> 3) ^ s1) << 3)) & 0x7fffffffu;
      s2 = ((s2 >> 28) | (((s2 >> 3) ^ (s2 >> 2) ^ (s2 >> 1) ^ s2) << 3)) & 0x7fffffffu;
    }
    s1 = prg_advance(s1, rest, false);
    s2 = prg_advance(s2, rest, true);
    const unsigned b1 = ((s1 >> 3) ^ s1) & 1u, b2 = ((s2 >> 3) ^ (s2 >> 2) ^ (s2 >> 1) ^ s2) & 1u;
    chips[t] = (s1 ^ s2) | ((b1 ^ b2) << 31);  // chip i of the word = bit i
  }
  __syncthreads();
  const float2* sym = symbols + job.sym_offset;
  const float* noise = noise_vars + job.sym_offset;
  const int8_t* in = in_llrs + job.sym_offset;
  int8_t* dst = out + job.llr_offset;
  const bool aligned = ((reinterpret_cast<unsigned long long>(dst) & 3ull) == 0);
#pragma unroll 1
  for (unsigned it = 0; it < 32u / 4u; ++it) {
    const unsigned local = 4u * (it * DEMOD_THREADS + t);  // first of this thread's 4 LLRs inside the chunk
    const unsigned j = j0 + local;
    if (j >= job.nof_llrs) {
      break;
    }
    const unsigned cw = job.descramble ? (chips[local >> 5] >> (local & 31u)) : 0u;
    const unsigned valid = min(4u, job.nof_llrs - j);
    unsigned packed;
    switch (job.mod_order) {  // uniform over the block
      case 8: packed = demod_four<8>(job, tab, j, valid, sym, noise); break;
      case 6: packed = demod_four<6>(job, tab, j, valid, sym, noise); break;
      case 4: packed = demod_four<4>(job, tab, j, valid, sym, noise); break;
      case 2: packed = demod_four<2>(job, tab, j, valid, sym, noise); break;
      case 1: packed = demod_four<1>(job, tab, j, valid, sym, noise); break;
      default:
        packed = 0;
        for (unsigned u = 0; u < valid; ++u) {
          packed |= ((unsigned)in[j + u] & 0xffu) << (8u * u);
        }
    }
    // descrambling: two's complement negation of the bytes whose chip is 1 (per byte: (v ^ 0xff) + 1)
    const unsigned m = (cw & 1u) * 0xffu | ((cw >> 1) & 1u) * 0xff00u | ((cw >> 2) & 1u) * 0xff0000u |
                       ((cw >> 3) & 1u) * 0xff000000u;
    const unsigned x = packed ^ m, one = m & 0x01010101u;
    packed = ((x & 0x7f7f7f7fu) + one) ^ (x & 0x80808080u);  // byte-wise x + one without carries across bytes
    if (aligned && j + 3 < job.nof_llrs) {
      *reinterpret_cast<unsigned*>(dst + j) = packed;
    } else {
      for (unsigned u = 0; u < 4 && j + u < job.nof_llrs; ++u) {
        dst[j + u] = (int8_t)(packed >> (8u * u));
      }
    }
  }
}

}  // namespace ldpc_b200
