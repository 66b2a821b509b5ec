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
// co-operatively (one row per lane, ballot), each lane then advances by 32 * lane chips 28 at a time and writes its 32-chip
// word to shared memory. Phase 2: a thread computes 4 consecutive LLRs per step (one packed 32-bit store, coalesced),
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

__device__ __forceinline__ int demod_quantize(float value, float range_limit)
{
  float clipped = value;
  if (fabsf(value) > range_limit) {
    clipped = copysignf(range_limit, value);
  }
  return (int)roundf(__fmul_rn(__fdiv_rn(clipped, range_limit), 120.0f));
}

__device__ __forceinline__ float demod_interval(float value, float rcp_noise, int m, int k)
{
  const int n = c_demod.nof_intervals[m][k];
  int idx = (int)floorf(__fdiv_rn(value, c_demod.width[m][k])) + n / 2;
  idx = max(0, min(n - 1, idx));
  const float l = __fmaf_rn(c_demod.slope[m][k][idx], value, c_demod.intercept[m][k][idx]);
  return __fmul_rn(l, rcp_noise);
}

// LLR j of a job (before descrambling).
__device__ __forceinline__ int demod_llr(const DemodJob& job, unsigned j, const float2* __restrict__ sym,
                                         const float* __restrict__ noise)
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
  const float2 z = __ldg(sym + s);
  const float nv = __ldg(noise + s);
  const float x = (b & 1u) ? z.y : z.x;
  if (q == 1) {
    if (!(nv > 0)) {
      return 0;
    }
    const float sum = (job.pi2 && (j & 1u)) ? __fadd_rn(z.y, -z.x) : __fadd_rn(z.x, z.y);
    return demod_quantize(__fdiv_rn(__fmul_rn(DEMOD_2SQRT2, sum), nv), 24.0f);
  }
  if (q == 2) {
    if (!(nv > 0)) {
      return 0;
    }
    return demod_quantize(__fdiv_rn(__fmul_rn(DEMOD_2SQRT2, x), nv), 24.0f);
  }
  if (q == 4) {
    if (!(nv > 0)) {
      return 0;
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
    return demod_quantize(__fdiv_rn(l, nv), 20.0f);
  }
  float rcp = 0.0f;
  if (nv > 0) {
    rcp = __fdiv_rn(1.0f, nv);
  }
  return demod_quantize(demod_interval(x, rcp, q == 8 ? 1 : 0, (int)(b >> 1)), 20.0f);
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
    // this lane's word starts 32 * lane chips further
    unsigned left = 32u * lane;
    while (left) {
      const unsigned k = min(left, 28u);
      s1 = prg_advance(s1, k, false);
      s2 = prg_advance(s2, k, true);
      left -= k;
    }
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
    unsigned packed = 0;
#pragma unroll
    for (unsigned u = 0; u < 4; ++u) {
      int v = 0;
      if (j + u < job.nof_llrs) {
        v = job.mod_order ? demod_llr(job, j + u, sym, noise) : (int)in[j + u];
        v = ((cw >> u) & 1u) ? -v : v;
      }
      packed |= ((unsigned)v & 0xffu) << (8u * u);
    }
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
