// Packed layered min-sum decoder for sm_100a: TWO codeblocks per thread block, one thread per lifted lane, every value
// a half2 (low half = codeblock A, high half = codeblock B). Both codeblocks share (base graph, lifting size), so the
// circulant rotations, addresses and control flow are identical for the two halves and every arithmetic instruction
// does two lanes of work (HFMA2 on the FMA pipe, HMNMX2 / HSET2 / LOP3 on the ALU pipe - profiles/r01_pipes_microbench.md).
//
// Exactness: all LLR values are integers of magnitude <= 336 -> exact in fp16; "infinity" (+-127, sticky) is the fp16
// +-inf, which the arithmetic below propagates without special cases. The arithmetic is the reference's
// (ldpc_decoder_impl.cpp:116-135, ldpc_decoder_generic.cpp:30-130, log_likelihood_ratio.cpp:38-85):
//   v2c  u = soft - c_old                     (inf stays inf; the +-120 clamp is applied where it matters, below)
//   min1/min2 over |u| with both initialised to 120 (values >= 120 never win, exactly as the strict compares of
//        ldpc_decoder_generic.cpp:64-77), sign parity = XOR of the fp16 sign bits
//   c2v  c_new = sign * (|u| == min1 ? Q(min2) : Q(min1)) - equal to the reference's argmin-index rule because ties
//        imply min2 == min1
//   soft n = promote(c_new + clamp(u, +-120)), frozen (u = +-inf) lanes keep u
// Per (row, lane) the state kept between iterations is {Q1s, Ds = Q2s - Q1s, idx, W}: T-signed scaled minima, the
// argmin edge as a number, and the sign bits of u as an integer-valued fp16 (bit k = edge k) - all built on the FMA pipe.
// Thread-private state lives partly in registers, partly in shared memory (word-major, conflict-free); the hrr soft
// values (bg_K + 4 nodes) live in shared memory; extension-node soft values (one private lane per row) live in an
// L2-resident global scratch and are prefetched one row ahead.
#pragma once

#include <cuda_fp16.h>
#include <utility>

#include "ldpc_device.cuh"

namespace ldpc_b200 {

enum { PACKED_SCALE_NONE = 0, PACKED_SCALE_GENERIC_08 = 1, PACKED_SCALE_AVX_08 = 2 };

struct PackedParams {
  const CbWork* work;
  const int8_t* llr_base;
  uint8_t* out_base;
  CbResult* results;
  unsigned* ext_scratch;  // [gridDim.x][M - 4][zpad] half2 words
  unsigned nof_cbs;
  int Z;
  int Z4;      // 4 * Z (byte stride of one node in the soft array)
  int zpad;    // blockDim.x
  int max_iterations;
  unsigned short shift4[LDPC_BG1_EDGES];  // 4 * (shift % Z) per edge
};

template <int BG>
struct Graph;
template <>
struct Graph<1> {
  static constexpr int M = LDPC_BG1_M, K = 22, E = LDPC_BG1_EDGES, HRR = 26;
  static constexpr int row_begin(int r) { return LDPC_BG1_ROW_PTR[r]; }
  static constexpr int col(int e) { return LDPC_BG1_COL[e]; }
};
template <>
struct Graph<2> {
  static constexpr int M = LDPC_BG2_M, K = 10, E = LDPC_BG2_EDGES, HRR = 14;
  static constexpr int row_begin(int r) { return LDPC_BG2_ROW_PTR[r]; }
  static constexpr int col(int e) { return LDPC_BG2_COL[e]; }
};

// State words per row: Q1s, Ds, idx, W0 (+ W1 when the row has more than 10 edges).
template <int BG>
struct StateLayout {
  using G = Graph<BG>;
  static constexpr int degree(int r) { return G::row_begin(r + 1) - G::row_begin(r); }
  static constexpr int words(int r) { return degree(r) > 10 ? 5 : 4; }
  static constexpr int base(int r)
  {
    int b = 0;
    for (int i = 0; i < r; ++i) {
      b += words(i);
    }
    return b;
  }
  static constexpr int total = base(G::M);
};

// ---- half2 helpers on raw bit patterns -------------------------------------------------------------------------------
__device__ __forceinline__ __half2 as_h2(unsigned x) { return *reinterpret_cast<__half2*>(&x); }
__device__ __forceinline__ unsigned as_u(__half2 x) { return *reinterpret_cast<unsigned*>(&x); }
__device__ __forceinline__ unsigned h2c(float v)
{
  return as_u(__float2half2_rn(v));
}
__device__ __forceinline__ unsigned hfma(unsigned a, unsigned b, unsigned c) { return as_u(__hfma2(as_h2(a), as_h2(b), as_h2(c))); }
__device__ __forceinline__ unsigned hfma_relu(unsigned a, unsigned b, unsigned c)
{
  return as_u(__hfma2_relu(as_h2(a), as_h2(b), as_h2(c)));
}
__device__ __forceinline__ unsigned hfma_sat(unsigned a, unsigned b, unsigned c)
{
  return as_u(__hfma2_sat(as_h2(a), as_h2(b), as_h2(c)));
}
__device__ __forceinline__ unsigned hadd(unsigned a, unsigned b) { return as_u(__hadd2(as_h2(a), as_h2(b))); }
__device__ __forceinline__ unsigned hsub(unsigned a, unsigned b) { return as_u(__hsub2(as_h2(a), as_h2(b))); }
__device__ __forceinline__ unsigned hmul(unsigned a, unsigned b) { return as_u(__hmul2(as_h2(a), as_h2(b))); }
__device__ __forceinline__ unsigned habs(unsigned a) { return as_u(__habs2(as_h2(a))); }
__device__ __forceinline__ unsigned hmin(unsigned a, unsigned b) { return as_u(__hmin2(as_h2(a), as_h2(b))); }
__device__ __forceinline__ unsigned hmax(unsigned a, unsigned b) { return as_u(__hmax2(as_h2(a), as_h2(b))); }
__device__ __forceinline__ unsigned heq01(unsigned a, unsigned b) { return as_u(__heq2(as_h2(a), as_h2(b))); }

constexpr unsigned H2_SIGN = 0x80008000u;

// Scaled magnitude Q(m), m integer-valued half2 in [0,120].
//  GENERIC 0.8: round(0.8 m) = m - round(m / 5)       (0.2 m is never at a tie, so one FMA rounding is exact)
//  AVX     0.8: (m * 52428) >> 16 = m - floor(m / 5) - [m > 0], floor(m/5) = round((m - 2) / 5)
// The magic constant 1536 keeps the sum in [1024, 2048) where the fp16 ulp is 1. Verified against the host table for
// every m by packed_scale_selfcheck() when the handle is created.
template <int SCALE>
__device__ __forceinline__ unsigned scale_q(unsigned m)
{
  const unsigned magic = h2c(1536.f), fifth = h2c(0.2f);
  if (SCALE == PACKED_SCALE_GENERIC_08) {
    unsigned r = hsub(hfma(m, fifth, magic), magic);
    return hsub(m, r);
  } else {
    unsigned r = hsub(hfma(hsub(m, h2c(2.f)), fifth, magic), magic);
    unsigned nz = hfma_sat(m, h2c(1.f), h2c(0.f));
    return hsub(hsub(m, r), nz);
  }
}

// ---- per-thread context ----------------------------------------------------------------------------------------------
template <int BG, int NREG>
struct PackedCtx {
  unsigned char* soft;  // shared, HRR nodes of Z half2 words (byte pointer: offsets are kept in bytes)
  unsigned* sst;        // shared state, already offset by tid; word w lives at sst[(w - NREG) * zpad]
  unsigned* ext;        // global extension-node scratch of this block, offset by tid; row r at ext[(r - 4) * zpad]
  unsigned rs[NREG > 0 ? NREG : 1];
  unsigned t4;          // 4 * tid
  unsigned Z4;
  unsigned zpad;
  unsigned ext_pref;    // extension soft value of the next row, prefetched while the current row is processed
  int nof_layers;
};

template <int BG, int NREG, int W>
__device__ __forceinline__ unsigned state_load(const PackedCtx<BG, NREG>& c)
{
  if constexpr (W < NREG) {
    return c.rs[W];
  } else {
    return c.sst[(W - NREG) * c.zpad];
  }
}
template <int BG, int NREG, int W>
__device__ __forceinline__ void state_store(PackedCtx<BG, NREG>& c, unsigned v)
{
  if constexpr (W < NREG) {
    c.rs[W] = v;
  } else {
    c.sst[(W - NREG) * c.zpad] = v;
  }
}

// Values that live across the two passes of one row.
template <int D>
struct RowRegs {
  unsigned u[D];    // variable-to-check messages (unclamped)
  unsigned off[D];  // byte offset of the soft value inside the shared soft array (hrr edges)
  unsigned M1, M2, S;
};

template <int BG, int NREG, int R, int K>
__device__ __forceinline__ void edge_pass1(PackedCtx<BG, NREG>& c, const PackedParams& p, RowRegs<StateLayout<BG>::degree(R)>& rr,
                                           unsigned Q1s, unsigned Ds, unsigned idx, unsigned W0, unsigned W1, unsigned xe)
{
  using G = Graph<BG>;
  constexpr int E = G::row_begin(R) + K;
  constexpr int COL = G::col(E);
  unsigned x;
  if constexpr (COL < G::HRR) {
    const unsigned a = c.t4 + p.shift4[E];
    const unsigned b = a - c.Z4;
    rr.off[K] = min(a, b) + COL * c.Z4;  // (lane + shift) mod Z, in bytes, plus the node base
    x = *reinterpret_cast<const unsigned*>(c.soft + rr.off[K]);
  } else {
    static_assert(COL == G::K + R, "extension column");
    x = xe;
  }
  // previous check-to-variable message of this edge from the compressed state
  const unsigned ind = heq01(idx, h2c((float)K));
  const unsigned mag = hfma(ind, Ds, Q1s);
  const unsigned wbits = (K < 10) ? W0 : W1;
  const unsigned sh = wbits * (1u << (15 - (K < 10 ? K : K - 10)));
  const unsigned cold = mag ^ (sh & H2_SIGN);
  const unsigned u = hsub(x, cold);
  rr.u[K] = u;
  const unsigned a = habs(u);
  const unsigned t1 = hmax(rr.M1, a);
  rr.M2 = hmin(rr.M2, t1);
  rr.M1 = hmin(rr.M1, a);
  rr.S ^= u;
}

template <int BG, int NREG, int R, int K>
__device__ __forceinline__ void edge_pass2(PackedCtx<BG, NREG>& c, RowRegs<StateLayout<BG>::degree(R)>& rr, unsigned Q2s,
                                           unsigned nD, unsigned& idxn, unsigned& Wn0, unsigned& Wn1, unsigned xe)
{
  using G = Graph<BG>;
  constexpr int E = G::row_begin(R) + K;
  constexpr int COL = G::col(E);
  const unsigned one = h2c(1.f);
  const unsigned u = rr.u[K];
  // 0 where |u| == min1 (this edge carries the minimum), 1 elsewhere
  const unsigned indp = hfma_sat(habs(u), one, rr.M1 ^ H2_SIGN);
  const unsigned cm = hfma(indp, nD, Q2s);
  const unsigned cn = cm ^ (u & H2_SIGN);
  if constexpr (K > 0) {
    idxn = hfma(indp, h2c(-(float)K), idxn);
  }
  const unsigned sg = hfma_sat(u, h2c(-1.f), h2c(0.f));  // 1 where u < 0
  if constexpr (K < 10) {
    Wn0 = hfma(sg, h2c((float)(1 << K)), Wn0);
  } else {
    Wn1 = hfma(sg, h2c((float)(1 << (K - 10))), Wn1);
  }
  // soft = promotion_sum(c_new, clamp(u, +-120)); frozen lanes (u = +-inf) keep u
  const unsigned v = hmax(hmin(u, h2c(120.f)), h2c(-120.f));
  const unsigned s = hadd(v, cn);
  const unsigned r = hfma_relu(habs(s), one, h2c(-120.f));
  const unsigned z = hmul(r, s);
  const unsigned n1 = hfma(z, h2c(65504.f), s);
  const unsigned ru = hfma_relu(u, u, h2c(-46656.f));
  const unsigned n = hfma(ru, u, n1);
  if constexpr (COL < G::HRR) {
    *reinterpret_cast<unsigned*>(c.soft + rr.off[K]) = n;
  } else {
    if (n != xe) {
      c.ext[(R - 4) * c.zpad] = n;
    }
  }
}

template <int BG, int NREG, int SCALE, int R, int... Ks>
__device__ __forceinline__ void do_row(PackedCtx<BG, NREG>& c, const PackedParams& p, std::integer_sequence<int, Ks...>)
{
  using G = Graph<BG>;
  using L = StateLayout<BG>;
  constexpr int D = L::degree(R), SB = L::base(R);
  const unsigned Q1s = state_load<BG, NREG, SB>(c), Ds = state_load<BG, NREG, SB + 1>(c);
  const unsigned idx = state_load<BG, NREG, SB + 2>(c), W0 = state_load<BG, NREG, SB + 3>(c);
  unsigned W1 = 0;
  if constexpr (D > 10) {
    W1 = state_load<BG, NREG, SB + 4>(c);
  }
  const unsigned xe = c.ext_pref;
  if constexpr (R + 1 >= 4 && R + 1 < G::M) {
    if (R + 1 < c.nof_layers) {
      c.ext_pref = c.ext[(R + 1 - 4) * c.zpad];
    }
  }
  RowRegs<D> rr;
  rr.M1 = h2c(120.f), rr.M2 = h2c(120.f), rr.S = 0;
  (edge_pass1<BG, NREG, R, Ks>(c, p, rr, Q1s, Ds, idx, W0, W1, xe), ...);
  const unsigned T = rr.S & H2_SIGN;
  const unsigned Q1n = scale_q<SCALE>(rr.M1) ^ T, Q2n = scale_q<SCALE>(rr.M2) ^ T;
  const unsigned Dn = hsub(Q2n, Q1n);
  unsigned idxn = h2c((float)(D * (D - 1) / 2)), Wn0 = h2c(1024.f), Wn1 = h2c(1024.f);
  (edge_pass2<BG, NREG, R, Ks>(c, rr, Q2n, Dn ^ H2_SIGN, idxn, Wn0, Wn1, xe), ...);
  state_store<BG, NREG, SB>(c, Q1n);
  state_store<BG, NREG, SB + 1>(c, Dn);
  state_store<BG, NREG, SB + 2>(c, idxn);
  state_store<BG, NREG, SB + 3>(c, Wn0);
  if constexpr (D > 10) {
    state_store<BG, NREG, SB + 4>(c, Wn1);
  }
}

template <int BG, int NREG, int SCALE, int... Rs>
__device__ __forceinline__ void do_iteration(PackedCtx<BG, NREG>& c, const PackedParams& p, bool active,
                                             std::integer_sequence<int, Rs...>)
{
  (
      [&] {
        if (Rs < c.nof_layers) {
          if (active) {
            do_row<BG, NREG, SCALE, Rs>(c, p, std::make_integer_sequence<int, StateLayout<BG>::degree(Rs)>{});
          }
          __syncthreads();
        }
      }(),
      ...);
}

// ---- kernel ----------------------------------------------------------------------------------------------------------
// Shared memory: soft[HRR * Z] words | state[(total - NREG) * zpad] words | hard[2][msg_bytes16] | crc tables 2 x 256 |
// small flags. One thread block processes pairs of codeblocks (2p, 2p + 1) of the group's work array.
template <int BG, int NREG>
static inline size_t packed_smem_bytes(int Z, int zpad)
{
  using G = Graph<BG>;
  const size_t msg_bytes16 = ((size_t)(G::K * Z + 7) / 8 + 15) & ~(size_t)15;
  return (size_t)G::HRR * Z * 4 + (size_t)(StateLayout<BG>::total - NREG) * zpad * 4 + 2 * msg_bytes16 + 2 * 256 * 4 + 64;
}

template <int BG, int NREG, int SCALE>
__global__ void __launch_bounds__(384, 1) ldpc_decode_packed_kernel(const __grid_constant__ PackedParams p)
{
  using G = Graph<BG>;
  using L = StateLayout<BG>;
  extern __shared__ __align__(16) unsigned char smem[];
  const int Z = p.Z, zpad = p.zpad, tid = threadIdx.x;
  const bool active = tid < Z;
  const unsigned msg_len = (unsigned)(G::K * Z), msg_bytes = (msg_len + 7) / 8, msg_bytes16 = (msg_bytes + 15) & ~15u;

  unsigned* soft_w = reinterpret_cast<unsigned*>(smem);
  unsigned* state_w = soft_w + G::HRR * Z;
  unsigned char* hard = reinterpret_cast<unsigned char*>(state_w + (size_t)(L::total - NREG) * zpad);
  unsigned* s_crc_table = reinterpret_cast<unsigned*>(hard + 2 * msg_bytes16);
  unsigned* s_flags = s_crc_table + 512;  // [0..1] last non-zero + 1 per codeblock, [2..3] crc ok

  PackedCtx<BG, NREG> c;
  c.soft = smem;
  c.sst = state_w + tid;
  c.ext = p.ext_scratch + (size_t)blockIdx.x * (G::M - 4) * zpad + tid;
  c.t4 = 4u * tid;
  c.Z4 = (unsigned)p.Z4;
  c.zpad = (unsigned)zpad;
  c.ext_pref = 0;

  for (unsigned pair = blockIdx.x; 2 * pair < p.nof_cbs; pair += gridDim.x) {
    const bool hasB = 2 * pair + 1 < p.nof_cbs;
    const CbWork wA = p.work[2 * pair];
    const CbWork wB = p.work[hasB ? 2 * pair + 1 : 2 * pair];
    const bool useA = !wA.skip_decode, useB = hasB && !wB.skip_decode;
    if (!useA && !useB) {
      continue;
    }
    if (tid < 4) {
      s_flags[tid] = 0;
    }
    __syncthreads();
    // load_soft_bits (ldpc_decoder_impl.cpp:148-173) for both codeblocks, int8 -> fp16, +-127 -> +-inf
    const int8_t* inA = p.llr_base + wA.in_offset;
    const int8_t* inB = p.llr_base + wB.in_offset;
    unsigned lastA = 0, lastB = 0;
    if (active) {
      for (int n = 0; n < G::K + G::M; ++n) {
        int a = 0, b = 0;
        if (n >= 2) {
          const unsigned i = (unsigned)(n - 2) * Z + tid;
          if (useA && i < wA.in_length) {
            a = inA[i];
            lastA = a ? i + 1 : lastA;
          }
          if (useB && i < wB.in_length) {
            b = inB[i];
            lastB = b ? i + 1 : lastB;
          }
        }
        float fa = (float)a, fb = (float)b;
        fa = (a > LLR_MAX) ? __int_as_float(0x7f800000) : ((a < -LLR_MAX) ? __int_as_float(0xff800000) : fa);
        fb = (b > LLR_MAX) ? __int_as_float(0x7f800000) : ((b < -LLR_MAX) ? __int_as_float(0xff800000) : fb);
        const unsigned v = as_u(__floats2half2_rn(fa, fb));
        if (n < G::HRR) {
          soft_w[n * Z + tid] = v;
        } else {
          c.ext[(n - G::HRR) * zpad] = v;
        }
      }
    }
#pragma unroll
    for (int w = 0; w < NREG; ++w) {
      c.rs[w] = 0;
    }
    for (int w = 0; w < L::total - NREG; ++w) {
      state_w[w * zpad + tid] = 0;
    }
    for (int i = tid; i < 512; i += zpad) {
      s_crc_table[i] = c_crc.table[i < 256 ? wA.crc_poly : wB.crc_poly][i & 255];
    }
    atomicMax(&s_flags[0], lastA);
    atomicMax(&s_flags[1], lastB);
    __syncthreads();
    const unsigned inszA = s_flags[0], inszB = s_flags[1];
    // per-codeblock status: the all-zero shortcut (ldpc_decoder_impl.cpp:88-94) finishes a codeblock immediately
    bool doneA = !useA, doneB = !useB;
    CbResult resA = {0, 0, 0}, resB = {0, 0, 0};
    auto zero_shortcut = [&](const CbWork& w, CbResult& res) {
      if (!w.crc_poly) {
        uint8_t* out = p.out_base + w.out_offset;
        for (unsigned i = tid; i < msg_bytes; i += zpad) {
          out[i] = (i == msg_bytes - 1 && (msg_len & 7)) ? (unsigned char)(0xff << (8 - (msg_len & 7))) : 0xff;
        }
      }
      res.flags = 1;
      if (tid == 0) {
        p.results[w.result_index] = res;
      }
    };
    if (useA && inszA == 0) {
      zero_shortcut(wA, resA);
      doneA = true;
    }
    if (useB && inszB == 0) {
      zero_shortcut(wB, resB);
      doneB = true;
    }
    // number of layers (ldpc_decoder_impl.cpp:105-114); the pair runs the larger count - surplus layers of the
    // shorter codeblock see an all-zero extension node and leave its high-rate nodes untouched
    auto layers_of = [&](unsigned insz) {
      unsigned cb_len = max(insz + 2u * Z, msg_len + 4u * Z);
      cb_len = ((cb_len + Z - 1) / Z) * Z;
      return (int)(cb_len / Z) - G::K;
    };
    c.nof_layers = max(doneA ? 4 : layers_of(inszA), doneB ? 4 : layers_of(inszB));

    int it = 0;
    while (it < p.max_iterations && !(doneA && doneB)) {
      do_iteration<BG, NREG, SCALE>(c, p, active, std::make_integer_sequence<int, G::M>{});
      ++it;
      const bool last = (it == p.max_iterations);
      const bool chkA = !doneA && ((wA.crc_poly && wA.early_stop) || last);
      const bool chkB = !doneB && ((wB.crc_poly && wB.early_stop) || last);
      if (chkA || chkB) {
        // get_hard_bits (ldpc_decoder_generic.cpp:132-138): bit = soft <= 0, MSB first, for both halves at once
        for (unsigned i = tid; i < msg_bytes; i += zpad) {
          unsigned ba = 0, bb = 0;
          for (unsigned k = 0; k < 8; ++k) {
            const unsigned b = 8 * i + k;
            if (b < msg_len) {
              const unsigned v = soft_w[b];
              ba |= (((v & 0x7fffu) == 0 || (v & 0x8000u)) ? 0x80u : 0u) >> k;
              bb |= (((v & 0x7fff0000u) == 0 || (v & 0x80000000u)) ? 0x80u : 0u) >> k;
            }
          }
          if (chkA) {
            hard[i] = (unsigned char)ba;
          }
          if (chkB) {
            hard[msg_bytes16 + i] = (unsigned char)bb;
          }
        }
        __syncthreads();
        const int warp = tid >> 5;
        if (warp == 0 && chkA && wA.crc_poly) {
          const unsigned r = crc_warp(hard, msg_len - wA.nof_filler_bits, wA.crc_poly, s_crc_table);
          if (tid == 0) {
            s_flags[2] = (r == 0);
          }
        }
        if (warp == 1 && chkB && wB.crc_poly) {
          const unsigned r = crc_warp(hard + msg_bytes16, msg_len - wB.nof_filler_bits, wB.crc_poly, s_crc_table + 256);
          if ((tid & 31) == 0) {
            s_flags[3] = (r == 0);
          }
        }
        __syncthreads();
        auto finish = [&](const CbWork& w, CbResult& res, bool& done, const unsigned char* bits, bool ok) {
          if (ok || last) {
            uint8_t* out = p.out_base + w.out_offset;
            for (unsigned i = tid; i < msg_bytes; i += zpad) {
              out[i] = bits[i];
            }
            res.iterations = (unsigned short)it;
            res.crc_ok = ok ? 1 : 0;
            if (tid == 0) {
              p.results[w.result_index] = res;
            }
            done = true;
          }
        };
        if (chkA) {
          finish(wA, resA, doneA, hard, wA.crc_poly && s_flags[2]);
        }
        if (chkB) {
          finish(wB, resB, doneB, hard + msg_bytes16, wB.crc_poly && s_flags[3]);
        }
        __syncthreads();
      }
    }
  }
}

// ---- host side -------------------------------------------------------------------------------------------------------
// Which closed-form scaling reproduces the host table (built from the reference formulas by build_scale_lut)?
static inline int packed_scale_variant(const unsigned char* lut)
{
  bool gen = true, avx = true;
  for (int m = 0; m <= LLR_MAX; ++m) {
    const int r5 = (2 * m + 5) / 10;  // round(m / 5), never at a tie
    gen = gen && (lut[m] == m - r5);
    avx = avx && (lut[m] == m - m / 5 - (m > 0 ? 1 : 0));
  }
  return gen ? PACKED_SCALE_GENERIC_08 : (avx ? PACKED_SCALE_AVX_08 : PACKED_SCALE_NONE);
}

template <int SCALE>
__global__ void packed_scale_selfcheck_kernel(unsigned char* out)
{
  const int m = threadIdx.x;
  if (m <= LLR_MAX) {
    const unsigned q = scale_q<SCALE>(h2c((float)m));
    out[m] = (unsigned char)__half2float(__low2half(as_h2(q)));
  }
}

// Runs the fp16 closed form for every magnitude on the device and compares with the host table. 0 = identical.
static inline int packed_scale_selfcheck(int variant, const unsigned char* lut, cudaStream_t stream)
{
  unsigned char* d = nullptr;
  unsigned char h[LLR_MAX + 1];
  if (cudaMalloc(&d, LLR_MAX + 1) != cudaSuccess) {
    return -1;
  }
  if (variant == PACKED_SCALE_GENERIC_08) {
    packed_scale_selfcheck_kernel<PACKED_SCALE_GENERIC_08><<<1, 128, 0, stream>>>(d);
  } else {
    packed_scale_selfcheck_kernel<PACKED_SCALE_AVX_08><<<1, 128, 0, stream>>>(d);
  }
  cudaError_t e = cudaMemcpyAsync(h, d, LLR_MAX + 1, cudaMemcpyDeviceToHost, stream);
  e = (e == cudaSuccess) ? cudaStreamSynchronize(stream) : e;
  cudaFree(d);
  if (e != cudaSuccess) {
    return -1;
  }
  for (int m = 0; m <= LLR_MAX; ++m) {
    if (h[m] != lut[m]) {
      return 1;
    }
  }
  return 0;
}

constexpr int PACKED_NREG_BG1 = 68;
constexpr int PACKED_NREG_BG2 = 40;
constexpr size_t PACKED_SMEM_LIMIT = 232448;  // 227 KB opt-in maximum per block on sm_100

template <int BG, int NREG, int SCALE>
static inline int packed_launch_variant(const PackedParams& p, int zpad, unsigned n_pairs, int sm_count, cudaStream_t stream)
{
  const size_t smem = packed_smem_bytes<BG, NREG>(p.Z, zpad);
  auto kernel = ldpc_decode_packed_kernel<BG, NREG, SCALE>;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PACKED_SMEM_LIMIT) != cudaSuccess) {
      return -1;
    }
    attr_set = true;
  }
  int ctas_per_sm = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kernel, zpad, smem) != cudaSuccess || ctas_per_sm < 1) {
    return -1;
  }
  const unsigned grid = std::min<unsigned>(n_pairs, (unsigned)(sm_count * ctas_per_sm));
  kernel<<<grid, zpad, smem, stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// Size in words of the extension-node scratch the packed kernel needs (independent of the batch).
static inline size_t packed_ext_scratch_words(int sm_count) { return (size_t)sm_count * 8 * LDPC_BG1_M * MAX_Z; }

// Returns 0 when launched, > 0 when the configuration is not covered (caller uses the general kernel), < 0 on CUDA error.
static inline int launch_packed_decode(const GroupParams& gp, int scale_variant, const CbWork* d_work, unsigned n,
                                       const int8_t* d_llr, uint8_t* d_out, CbResult* d_res, unsigned* ext_scratch,
                                       int sm_count, cudaStream_t stream)
{
  if (scale_variant == PACKED_SCALE_NONE || ext_scratch == nullptr || gp.Z < 8) {
    return 1;
  }
  PackedParams p;
  std::memset(&p, 0, sizeof(p));
  p.work = d_work, p.llr_base = d_llr, p.out_base = d_out, p.results = d_res, p.ext_scratch = ext_scratch;
  p.nof_cbs = n, p.Z = gp.Z, p.Z4 = 4 * gp.Z, p.max_iterations = gp.max_iterations;
  const int zpad = std::max(64, ((gp.Z + 31) / 32) * 32);
  p.zpad = zpad;
  const int n_edges = (gp.bg == 2) ? LDPC_BG2_EDGES : LDPC_BG1_EDGES;
  for (int e = 0; e != n_edges; ++e) {
    const unsigned s = (gp.bg == 2) ? LDPC_BG2_SHIFT[gp.lifting_set][e] : LDPC_BG1_SHIFT[gp.lifting_set][e];
    p.shift4[e] = (unsigned short)(4 * (s % (unsigned)gp.Z));
  }
  const unsigned n_pairs = (n + 1) / 2;
  const bool gen = scale_variant == PACKED_SCALE_GENERIC_08;
  if (gp.bg == 1) {
    const bool small = packed_smem_bytes<1, 0>(gp.Z, zpad) <= PACKED_SMEM_LIMIT / 2;
    if (small) {
      return gen ? packed_launch_variant<1, 0, PACKED_SCALE_GENERIC_08>(p, zpad, n_pairs, sm_count, stream)
                 : packed_launch_variant<1, 0, PACKED_SCALE_AVX_08>(p, zpad, n_pairs, sm_count, stream);
    }
    if (packed_smem_bytes<1, PACKED_NREG_BG1>(gp.Z, zpad) > PACKED_SMEM_LIMIT) {
      return 1;
    }
    return gen ? packed_launch_variant<1, PACKED_NREG_BG1, PACKED_SCALE_GENERIC_08>(p, zpad, n_pairs, sm_count, stream)
               : packed_launch_variant<1, PACKED_NREG_BG1, PACKED_SCALE_AVX_08>(p, zpad, n_pairs, sm_count, stream);
  }
  const bool small = packed_smem_bytes<2, 0>(gp.Z, zpad) <= PACKED_SMEM_LIMIT / 2;
  if (small) {
    return gen ? packed_launch_variant<2, 0, PACKED_SCALE_GENERIC_08>(p, zpad, n_pairs, sm_count, stream)
               : packed_launch_variant<2, 0, PACKED_SCALE_AVX_08>(p, zpad, n_pairs, sm_count, stream);
  }
  if (packed_smem_bytes<2, PACKED_NREG_BG2>(gp.Z, zpad) > PACKED_SMEM_LIMIT) {
    return 1;
  }
  return gen ? packed_launch_variant<2, PACKED_NREG_BG2, PACKED_SCALE_GENERIC_08>(p, zpad, n_pairs, sm_count, stream)
             : packed_launch_variant<2, PACKED_NREG_BG2, PACKED_SCALE_AVX_08>(p, zpad, n_pairs, sm_count, stream);
}

}  // namespace ldpc_b200
