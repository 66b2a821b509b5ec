// Rate dematching (bit de-interleave + circular-buffer de-selection + HARQ soft combining) as a closed form per soft-buffer
// position, so that every position is produced by one independent thread with coalesced soft-buffer traffic.
//
// Reference: ldpc_rate_dematcher_impl.cpp:43-110 (rate_dematch), :125-198 (allot_llrs), :113-123 (combine_softbits),
// :200-254 (deinterleave). allot_llrs walks the circular buffer [0, Ncb) from k0 in "laps": a lap runs from the current
// index to the end of the buffer, skipping the filler range [info, sys). The first lap copies when new_data is set, every
// later lap (and every lap of a retransmission) saturating-adds. The literal range operations of the reference - and
// therefore which bytes keep their PREVIOUS contents - are reproduced exactly (SURVEY.md 8a row R3, quirks i-v):
//   new data: out[0, min(k0, info)) = 0 (or out[0, info) = 0 when k0 >= info); fillers = +127; positions reached in the
//   first lap are copied; positions before the start are first reached in lap 1 and ADDED to what is there (0 if zeroed
//   above, stale otherwise); if the input ends inside the first lap, the LAST (Ncb - idx_end) bytes of the N-long span
//   are zeroed (not [idx_end, Ncb)); bytes never named keep their contents.
#pragma once

#include "ldpc_device.cuh"

namespace ldpc_b200 {

struct DematchGeom {
  unsigned N;      // N_short * Z, soft-buffer length
  unsigned Ncb;    // circular buffer length
  unsigned info;   // (bg_K - 2) Z - F : first filler position
  unsigned sys;    // (bg_K - 2) Z     : first parity position
  unsigned k0;     // start position
  unsigned E;      // rate-matched length
  unsigned Qm;     // bits per symbol
  unsigned new_data;
  // derived
  unsigned s0;     // effective start (k0, or sys when k0 falls in the filler range)
  unsigned V;      // valid positions per full lap
  unsigned Ls0;    // linear rank of s0
  unsigned zero_from;  // new data, input ended in lap 0: [zero_from, N) is zeroed; N when nothing is
};

__host__ __device__ inline unsigned dematch_linear(const DematchGeom& g, unsigned p)
{
  return (p < g.info) ? p : p - (g.sys - g.info);
}

// k0 = floor(c_rv * Ncb / N) * Z (ldpc_rate_dematcher_impl.cpp:101-102).
__host__ __device__ inline DematchGeom make_dematch_geom(int bg, unsigned Z, unsigned F, unsigned rv, unsigned Qm,
                                                         unsigned Nref, unsigned E, unsigned new_data)
{
  DematchGeom g;
  const unsigned n_short = (bg == 2) ? 50 : 66, K = (bg == 2) ? 10 : 22;
  g.N = n_short * Z;
  g.Ncb = (Nref > 0) ? (Nref < g.N ? Nref : g.N) : g.N;
  g.sys = (K - 2) * Z;
  g.info = g.sys - F;
  // The reference computes floor(c_rv * Ncb / N) in double (ldpc_rate_dematcher_impl.cpp:101-102). The integer quotient is
  // the same number: c_rv * Ncb < 2^21 is exact, and a non-integer quotient of integers with a divisor <= 25 344 is at
  // least 1 / 25 344 away from an integer, far more than the rounding error of one double division of values below 64.
  const unsigned f1[4] = {0, 17, 33, 56}, f2[4] = {0, 13, 25, 43};
  const unsigned factor = (bg == 2) ? f2[rv & 3] : f1[rv & 3];
  g.k0 = ((factor * g.Ncb) / g.N) * Z;
  g.E = E;
  g.Qm = Qm;
  g.new_data = new_data;
  g.s0 = (g.k0 >= g.info && g.k0 < g.sys) ? g.sys : g.k0;
  g.V = g.info + (g.Ncb - g.sys);
  g.Ls0 = dematch_linear(g, g.s0);
  g.zero_from = g.N;
  if (new_data) {
    const unsigned lap0 = g.V - g.Ls0;  // positions from s0 to the end of the buffer
    if (E < lap0 || E == 0) {
      // index left by steps A-D of the first loop pass (allot_llrs :137-191)
      unsigned end;
      if (E == 0) {
        end = g.k0;  // loop never ran: zero(out.last(Ncb - k0)) when k0 != 0
      } else if (g.s0 < g.info) {
        const unsigned n = (g.info - g.s0 < E) ? (g.info - g.s0) : E;  // systematic chunk
        end = g.sys + (E - n);                                        // then jump to sys and take the rest
      } else {
        end = g.s0 + E;
      }
      end %= g.Ncb;
      if (end != 0) {
        g.zero_from = g.N - (g.Ncb - end);
      }
    }
  }
  return g;
}

// De-interleaved input element x (deinterleave_llrs :200-210): aux[(E/Qm) j + i] = in[i Qm + j].
__device__ __forceinline__ int dematch_input(const DematchGeom& g, const int8_t* __restrict__ in, unsigned x)
{
  if (g.Qm == 1) {
    return in[x];
  }
  // x < E <= 295 680 < 2^24 and Kq = E / Qm are exact in float: the truncated quotient estimate is off by at most one
  const unsigned Kq = g.E / g.Qm;
  unsigned j = (unsigned)(__uint2float_rz(x) * __frcp_rz(__uint2float_rz(Kq)));
  int i = (int)x - (int)(j * Kq);
  if (i < 0) {
    i += (int)Kq, --j;
  } else if (i >= (int)Kq) {
    i -= (int)Kq, ++j;
  }
  return in[(unsigned)i * g.Qm + j];
}

// Where soft-buffer position p sits in the walk: its rank r in the stream of de-interleaved input elements (the first
// element that lands on it; later laps add V each), whether it lies before the start position (first reached in lap 1)
// and whether the walk reaches it at all.
struct DematchRank {
  unsigned r;
  bool before_start;
  bool live;  // not a filler and inside the circular buffer
};

__device__ __forceinline__ DematchRank dematch_rank(const DematchGeom& g, unsigned p)
{
  DematchRank d;
  const bool is_filler = (p >= g.info && p < g.sys);
  d.live = !is_filler && p < g.Ncb;
  const unsigned L = dematch_linear(g, p);
  d.before_start = L < g.Ls0;
  d.r = d.before_start ? (L + g.V - g.Ls0) : (L - g.Ls0);
  return d;
}

// Final content of soft-buffer position p given its previous content and x0 = input element d.r (ignored when d.r >= E).
// Split from dematch_rank so that a caller can issue the loads of many positions before combining any of them.
__device__ __forceinline__ int dematch_combine(const DematchGeom& g, const int8_t* __restrict__ in, unsigned p, int prev,
                                               const DematchRank& d, int x0)
{
  int v = prev;
  if (g.new_data && g.E != 0) {
    if (p >= g.info && p < g.sys) {
      return LLR_INF;
    }
    // zeroing done by the first loop pass before it copies
    if (p < g.info && (g.k0 >= g.info || p < g.k0)) {
      v = 0;
    }
  }
  if (d.live && d.r < g.E) {
    v = (g.new_data && !d.before_start) ? x0 : llr_add(v, x0);  // copy in the first lap, combine otherwise
    for (unsigned r = d.r + g.V; r < g.E; r += g.V) {
      v = llr_add(v, dematch_input(g, in, r));
    }
  }
  if (p >= g.zero_from) {
    v = 0;
  }
  return v;
}

__device__ __forceinline__ int dematch_position(const DematchGeom& g, const int8_t* __restrict__ in, unsigned p, int prev)
{
  const DematchRank d = dematch_rank(g, p);
  const int x0 = (d.live && d.r < g.E) ? dematch_input(g, in, d.r) : 0;
  return dematch_combine(g, in, p, prev, d, x0);
}

// ---- the same dematcher as three data-parallel phases over a staging copy of the soft buffer (fused decoder prologue) ----
// A (dematch_stage_in): every position takes its previous content with the new-data initialisation applied (the zero /
//    filler ranges the first loop pass of allot_llrs writes before it copies) and the trailing zero range - the walk
//    never reaches that range when it exists, so zeroing it first is the same as zeroing it last;
// B: every de-interleaved input element is copied (first lap of new data) or saturating-added to the position it lands
//    on. E <= V (dematch_scatter_once): a position receives at most one element, so the elements are taken in INPUT
//    order - coalesced loads, symbol index i and bit index j known without a division - and no ordering is needed.
//    Otherwise (dematch_scatter_lap) lap by lap: at most one element per position and lap, laps in order, so the result
//    equals the serial walk;
// C (dematch_stage_out): the combined soft buffer goes back to the HBM arena.
__device__ __forceinline__ int dematch_init(const DematchGeom& g, unsigned p, int prev)
{
  if (p >= g.zero_from) {
    return 0;
  }
  if (g.new_data && g.E != 0) {
    if (p >= g.info && p < g.sys) {
      return LLR_INF;
    }
    if (p < g.info && (g.k0 >= g.info || p < g.k0)) {
      return 0;
    }
  }
  return prev;
}

// New data from the start of the buffer that covers the whole message and ends inside the first lap (the usual first
// transmission): [0, info) and [sys, end) are copied, fillers are +127, [end, N) is zeroed - nothing of the previous
// content survives, so the arena is not read at all.
__device__ __forceinline__ bool dematch_is_fresh(const DematchGeom& g)
{
  return g.new_data && g.k0 == 0 && g.Ncb == g.N && g.E >= g.info && g.E < g.V;
}

// stage[lo, hi) = value by the threads of one group; 16-byte stores on the aligned interior (stage is 16-byte aligned).
__device__ __forceinline__ void dematch_fill(int8_t* stage, unsigned lo, unsigned hi, int value, unsigned t, unsigned stride)
{
  if (lo >= hi) {
    return;
  }
  unsigned a = (lo + 15u) & ~15u;
  a = a < hi ? a : hi;
  unsigned b = hi & ~15u;
  b = b > a ? b : a;
  for (unsigned p = lo + t; p < a; p += stride) {
    stage[p] = (int8_t)value;
  }
  const unsigned w = (unsigned)(value & 0xff) * 0x01010101u;
  const uint4 v = make_uint4(w, w, w, w);
  for (unsigned p = a + 16u * t; p < b; p += 16u * stride) {
    *reinterpret_cast<uint4*>(stage + p) = v;
  }
  for (unsigned p = b + t; p < hi; p += stride) {
    stage[p] = (int8_t)value;
  }
}

// Phase A for one codeblock.
__device__ __noinline__ void dematch_stage_in(const DematchGeom& g, const int8_t* __restrict__ soft, int8_t* stage,
                                                 unsigned t, unsigned stride)
{
  if (dematch_is_fresh(g)) {
    dematch_fill(stage, g.info, g.sys, LLR_INF, t, stride);
    dematch_fill(stage, g.zero_from, g.N, 0, t, stride);
    return;
  }
  const bool plain = !(g.new_data && g.E != 0) && g.zero_from >= g.N;  // dematch_init is the identity
  if (plain && (reinterpret_cast<uintptr_t>(soft) & 15u) == 0) {
    constexpr int VB = 5;  // 66 * 384 / 16 / 384 = 4.1 vectors per thread
    const unsigned nvec = g.N / 16u;
    for (unsigned v0 = t; v0 < nvec; v0 += VB * stride) {
      uint4 x[VB];
#pragma unroll
      for (int k = 0; k < VB; ++k) {
        const unsigned v = v0 + k * stride;
        if (v < nvec) {
          x[k] = __ldg(reinterpret_cast<const uint4*>(soft) + v);
        }
      }
#pragma unroll
      for (int k = 0; k < VB; ++k) {
        const unsigned v = v0 + k * stride;
        if (v < nvec) {
          reinterpret_cast<uint4*>(stage)[v] = x[k];
        }
      }
    }
    for (unsigned p = nvec * 16u + t; p < g.N; p += stride) {
      stage[p] = soft[p];
    }
    return;
  }
  constexpr int UB = 11;
  for (unsigned p0 = t; p0 < g.N; p0 += UB * stride) {
    int x[UB];
#pragma unroll
    for (int k = 0; k < UB; ++k) {
      const unsigned p = p0 + k * stride;
      x[k] = (p < g.N) ? (int)soft[p] : 0;
    }
#pragma unroll
    for (int k = 0; k < UB; ++k) {
      const unsigned p = p0 + k * stride;
      if (p < g.N) {
        stage[p] = (int8_t)dematch_init(g, p, x[k]);
      }
    }
  }
}

// Phase C for one codeblock (the conversion to the decoder's format reads the same staging copy).
__device__ __noinline__ void dematch_stage_out(const DematchGeom& g, const int8_t* stage, int8_t* __restrict__ soft,
                                                  unsigned t, unsigned stride)
{
  unsigned done = 0;
  if ((reinterpret_cast<uintptr_t>(soft) & 15u) == 0) {
    const unsigned nvec = g.N / 16u;
    for (unsigned v = t; v < nvec; v += stride) {
      reinterpret_cast<uint4*>(soft)[v] = reinterpret_cast<const uint4*>(stage)[v];
    }
    done = nvec * 16u;
  }
  for (unsigned p = done + t; p < g.N; p += stride) {
    soft[p] = stage[p];
  }
}

__device__ __forceinline__ unsigned dematch_nof_laps(const DematchGeom& g)
{
  return g.E ? (g.Ls0 + g.E + g.V - 1) / g.V : 0u;
}

// E <= V and Qm in {1, 2, 4, 6, 8}: handled by dematch_scatter_once.
__device__ __forceinline__ bool dematch_single_pass(const DematchGeom& g)
{
  const bool qm_ok = g.Qm == 1 || g.Qm == 2 || g.Qm == 4 || g.Qm == 6 || g.Qm == 8;
  return g.E != 0 && g.E <= g.V && qm_ok && (g.E % g.Qm) == 0;
}

// Symbol i carries the elements e = j * (E / Qm) + i, j < Qm, of the de-interleaved stream (deinterleave_llrs :200-210).
template <int QM>
__device__ __forceinline__ void dematch_scatter_once_q(const DematchGeom& g, const int8_t* __restrict__ in, int8_t* stage,
                                                       unsigned t, unsigned stride)
{
  constexpr int BI = (QM >= 4) ? 2 : ((QM == 2) ? 4 : 8);  // symbols per batch: 8 .. 16 LLRs in flight per thread
  const unsigned Kq = g.E / QM;
  const unsigned gap = g.sys - g.info;
  const unsigned wrap = g.V - g.Ls0;  // elements from here on are in the second lap (positions before the start)
  const bool vec = (QM == 2 || QM == 4 || QM == 8) && (reinterpret_cast<uintptr_t>(in) % QM) == 0;
  for (unsigned i0 = t; i0 < Kq; i0 += BI * stride) {
    int x[BI][QM];
#pragma unroll
    for (int b = 0; b < BI; ++b) {
      const unsigned i = i0 + b * stride;
      if (i < Kq) {
        if (vec) {
          unsigned lo = 0, hi = 0;
          if (QM == 2) {
            lo = __ldg(reinterpret_cast<const unsigned short*>(in) + i);
          } else if (QM == 4) {
            lo = __ldg(reinterpret_cast<const unsigned*>(in) + i);
          } else {
            const uint2 w = __ldg(reinterpret_cast<const uint2*>(in) + i);
            lo = w.x, hi = w.y;
          }
#pragma unroll
          for (int j = 0; j < QM; ++j) {
            x[b][j] = (int)(int8_t)(((j < 4) ? lo : hi) >> (8 * (j & 3)));
          }
        } else {
#pragma unroll
          for (int j = 0; j < QM; ++j) {
            x[b][j] = in[i * QM + j];
          }
        }
      }
    }
#pragma unroll
    for (int b = 0; b < BI; ++b) {
      const unsigned i = i0 + b * stride;
      if (i < Kq) {
#pragma unroll
        for (int j = 0; j < QM; ++j) {
          const unsigned e = (unsigned)j * Kq + i;
          const bool first = e < wrap;
          const unsigned L = first ? g.Ls0 + e : e - wrap;
          const unsigned p = L < g.info ? L : L + gap;
          if (g.new_data && first) {
            stage[p] = (int8_t)x[b][j];
          } else {
            stage[p] = (int8_t)llr_add((int)stage[p], x[b][j]);
          }
        }
      }
    }
  }
}

__device__ __noinline__ void dematch_scatter_once(const DematchGeom& g, const int8_t* __restrict__ in, int8_t* stage,
                                                     unsigned t, unsigned stride)
{
  switch (g.Qm) {
    case 1:
      dematch_scatter_once_q<1>(g, in, stage, t, stride);
      break;
    case 2:
      dematch_scatter_once_q<2>(g, in, stage, t, stride);
      break;
    case 4:
      dematch_scatter_once_q<4>(g, in, stage, t, stride);
      break;
    case 6:
      dematch_scatter_once_q<6>(g, in, stage, t, stride);
      break;
    default:
      dematch_scatter_once_q<8>(g, in, stage, t, stride);
      break;
  }
}

// ---- word-wise scatter: the usual first transmission ---------------------------------------------------------------------
// New data whose E elements all land in the first lap (nothing is combined, every position is hit at most once), with
// E / Qm, the start rank and the filler range all multiples of 4: four consecutive symbols (4 Qm input bytes = Qm words)
// are transposed in registers into Qm words of four consecutive elements each and stored with Qm 32-bit stores. ~6
// instructions per word instead of ~20 branchy ones per byte: the byte-wise scatter was 20 % of the fused kernel's time on
// full-slot transport blocks (tools/phase_timing.py).
__device__ __forceinline__ bool dematch_wordwise(const DematchGeom& g, const int8_t* in)
{
  const bool qm_ok = g.Qm == 2 || g.Qm == 4 || g.Qm == 6 || g.Qm == 8;
  if (!qm_ok || !g.new_data || g.E == 0 || g.E % (4u * g.Qm) != 0) {
    return false;
  }
  return g.E <= g.V - g.Ls0 && ((g.Ls0 | g.info | (g.sys - g.info)) & 3u) == 0 && (reinterpret_cast<uintptr_t>(in) & 3u) == 0;
}

template <int QM>
__device__ __forceinline__ void dematch_quad_load(const int8_t* __restrict__ in, unsigned q, unsigned (&w)[QM])
{
  const unsigned* src = reinterpret_cast<const unsigned*>(in) + q * QM;
#pragma unroll
  for (int j = 0; j < QM; ++j) {
    w[j] = __ldg(src + j);
  }
}

// Symbols 4q .. 4q + 3: input byte s QM + j (symbol s, bit j) is element j (E / QM) + 4q + s of the de-interleaved stream.
template <int QM>
__device__ __forceinline__ void dematch_quad_store(const DematchGeom& g, unsigned q, const unsigned (&w)[QM], int8_t* stage)
{
  const unsigned Kq = g.E / QM, gap = g.sys - g.info;
#pragma unroll
  for (int j = 0; j < QM; ++j) {
    const int k0 = j, k1 = QM + j, k2 = 2 * QM + j, k3 = 3 * QM + j;
    const unsigned lo = __byte_perm(w[k0 >> 2], w[k1 >> 2], (unsigned)((k0 & 3) | ((4 + (k1 & 3)) << 4)));
    const unsigned hi = __byte_perm(w[k2 >> 2], w[k3 >> 2], (unsigned)((k2 & 3) | ((4 + (k3 & 3)) << 4)));
    const unsigned L = g.Ls0 + (unsigned)j * Kq + 4u * q;
    const unsigned p = L < g.info ? L : L + gap;
    *reinterpret_cast<unsigned*>(stage + p) = __byte_perm(lo, hi, 0x5410);
  }
}

// Both codeblocks of a pair (doB = false: only A): all loads of both are in flight before the first store.
template <int QM>
__device__ __forceinline__ void dematch_scatter_words_q(const DematchGeom& gA, const int8_t* __restrict__ inA, int8_t* stA,
                                                        bool doB, const DematchGeom& gB, const int8_t* __restrict__ inB,
                                                        int8_t* stB, unsigned t, unsigned stride)
{
  const unsigned nA = gA.E / (4u * QM), nB = doB ? gB.E / (4u * QM) : 0u;
  for (unsigned q = t; q < max(nA, nB); q += stride) {
    unsigned a[QM], b[QM];
    if (q < nA) {
      dematch_quad_load<QM>(inA, q, a);
    }
    if (q < nB) {
      dematch_quad_load<QM>(inB, q, b);
    }
    if (q < nA) {
      dematch_quad_store<QM>(gA, q, a, stA);
    }
    if (q < nB) {
      dematch_quad_store<QM>(gB, q, b, stB);
    }
  }
}

// gA (and gB when doB) satisfy dematch_wordwise and share the modulation order.
__device__ __noinline__ void dematch_scatter_words(const DematchGeom& gA, const int8_t* __restrict__ inA, int8_t* stA, bool doB,
                                                      const DematchGeom& gB, const int8_t* __restrict__ inB, int8_t* stB,
                                                      unsigned t, unsigned stride)
{
  switch (gA.Qm) {
    case 2:
      dematch_scatter_words_q<2>(gA, inA, stA, doB, gB, inB, stB, t, stride);
      break;
    case 4:
      dematch_scatter_words_q<4>(gA, inA, stA, doB, gB, inB, stB, t, stride);
      break;
    case 6:
      dematch_scatter_words_q<6>(gA, inA, stA, doB, gB, inB, stB, t, stride);
      break;
    default:
      dematch_scatter_words_q<8>(gA, inA, stA, doB, gB, inB, stB, t, stride);
      break;
  }
}

// Elements of lap `lap` handled by thread t of `stride`; loads are issued BATCH at a time before they are combined.
__device__ __noinline__ void dematch_scatter_lap(const DematchGeom& g, const int8_t* __restrict__ in, int8_t* stage,
                                                    unsigned lap, unsigned t, unsigned stride)
{
  constexpr int BATCH = 8;
  const unsigned first = lap * g.V;  // Ls0 + r runs over [first, first + V) in this lap
  const unsigned lo = first > g.Ls0 ? first - g.Ls0 : 0u;
  const unsigned end = first + g.V - g.Ls0;
  const unsigned hi = end < g.E ? end : g.E;
  const bool copy = g.new_data && lap == 0;
  for (unsigned r0 = lo + t; r0 < hi; r0 += BATCH * stride) {
    int x[BATCH];
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      const unsigned r = r0 + k * stride;
      x[k] = (r < hi) ? dematch_input(g, in, r) : 0;
    }
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      const unsigned r = r0 + k * stride;
      if (r < hi) {
        const unsigned L = g.Ls0 + r - first;
        const unsigned p = L < g.info ? L : L + (g.sys - g.info);
        stage[p] = (int8_t)(copy ? x[k] : llr_add((int)stage[p], x[k]));
      }
    }
  }
}

// Standalone kernel: one block per codeblock, threads stride over the N positions (coalesced read-modify-write).
__global__ void __launch_bounds__(256) ldpc_dematch_kernel(int bg, int Z, const CbWork* __restrict__ work, unsigned nof_cbs,
                                                            const int8_t* __restrict__ rm_base, int8_t* __restrict__ soft_base)
{
  for (unsigned cb = blockIdx.x; cb < nof_cbs; cb += gridDim.x) {
    const CbWork w = work[cb];
    const DematchGeom g = make_dematch_geom(bg, (unsigned)Z, w.nof_filler_bits, w.rv, w.mod_order, w.Nref, w.rm_length,
                                            w.new_data);
    const int8_t* in = rm_base + w.rm_offset;
    int8_t* soft = soft_base + w.in_offset;
    for (unsigned p = threadIdx.x; p < g.N; p += blockDim.x) {
      const int prev = soft[p];
      const int v = dematch_position(g, in, p, prev);
      if (v != prev) {
        soft[p] = (int8_t)v;
      }
    }
  }
}

}  // namespace ldpc_b200
