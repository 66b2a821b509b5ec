// Layered normalised min-sum decoder for sm_100a, "TMEM-message" formulation.
//
// Work split: one thread per lifted lane, TWO codeblocks per thread (half2: low half = codeblock A, high half = B; both
// share base graph and lifting size so addresses, rotations and control flow are common). A thread block holds
// G = 384 / roundup32(Z) independent codeblock pairs ("groups"), each with its own named barrier (phases outside the row
// loop) and its own mbarrier (the row barrier of the layered schedule, split into arrive and wait).
//
// Where the data live (per SM, Z = 384: everything on chip, nothing but the input LLRs and output bits touches HBM):
//   * a-posteriori soft values of the high-rate nodes (bg_K + 4 columns)            shared memory, half2 per lane
//   * soft value and check-to-variable message of the extension nodes (one private
//     lane per row, shift 0)                                                        shared memory, thread-private words
//   * check-to-variable messages of all other edges (the reference's check_to_var
//     array, ldpc_decoder_impl.h:204-210), one biased byte per codeblock and edge   TENSOR MEMORY (tcgen05.ld / .st,
//                                                                                   32x32b shape: one TMEM lane per
//                                                                                   thread, <= 170 private columns)
// Keeping the messages explicitly (instead of the compressed min1/min2/index/sign state of the first packed kernel)
// removes the per-edge reconstruction and the index/sign bookkeeping from the instruction stream, and moving them to
// TMEM frees shared memory for the extension nodes (no L2 round trips inside an iteration).
//
// Arithmetic (bit-exact with lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:116-135 + ldpc_decoder_generic.cpp
// :30-130 + lib/phy/upper/log_likelihood_ratio.cpp:38-85). Soft values, variable-to-check messages and the scaled minima
// live in the SCALED domain X = 544 x: 544 x is exact in fp16 for every integer |x| <= 120 (17 x < 2048), 544 * 120 =
// 65280 is finite and 544 * 121 = 65824 rounds to infinity, so that "|sum| > 120 -> +-127 (sticky)" of promotion_sum IS the
// fp16 overflow of one addition; the LLR "infinity" (+-127) is the fp16 infinity.
//   v2c   U = SOFT - 544 c_old (one FMA; overflows to +-inf when |soft - c_old| > 120);
//         V = sign(U) min(|U|, T), T = fma(|SOFT|, 2^-12, 65280) = 65280 for a finite SOFT (the clamp at +-120 of operator-)
//         and +inf for an infinite one (abs(a) > 120 -> a), in one FMA + one HMNMX2.XORSIGN
//   min1/min2 over |V|, capped at 65280 (the reference initialises both to 120 and compares strictly, so values >= 120
//         never win); computed as a tree of sorted pairs with HMNMX2 / VIMNMX3.U16x2 on the (order-preserving) bit
//         patterns of the non-negative halves
//   c2v   C_new = sign * (|V| == MIN1 ? Q(MIN2) : Q(MIN1))  == the reference's argmin-index rule, because a tie for
//         the minimum implies min2 == min1
//   soft  SOFT' = V + C_new (promotion and stickiness by overflow / inf arithmetic)
// Messages are stored unscaled as biased bytes: 1152 + c = fma(C, 1/544, 1152) (the rounding absorbs the inexact 1/544).
// Row structure (degrees, columns, shifts) comes from per-launch tables in the kernel parameters (constant bank): an
// iteration is a list of steps (one row, or two rows without a common column), one 16-byte descriptor each, read with
// uniform loads - descriptors, edge records, TMEM addresses and the class dispatch all stay on the uniform datapath. The
// row bodies are instantiated per number of high-rate edges so that the per-edge loops unroll
// while the code stays small enough for the instruction cache (the first packed kernel unrolled all 316 edges and was
// instruction-fetch bound, profiles/r01_ncu_packed_v1_summary.txt).
#pragma once

#include <atomic>

#include <cuda_fp16.h>

#include "ldpc_device.cuh"
#include "ldpc_dematch.cuh"

namespace ldpc_b200 {

enum { TM_SCALE_NONE = 0, TM_SCALE_GENERIC_08 = 1, TM_SCALE_AVX_08 = 2, TM_SCALE_LUT = 3 };
// Kernel flavour: GENERAL (any Z, <= 384 threads), FULL (Z = 384 folded at compile time), WIDE (decode-only blocks of up to
// 512 threads: 128 registers, 128 TMEM columns per warp, overflow message words in shared memory).
enum { TM_MODE_GENERAL = 0, TM_MODE_FULL = 1, TM_MODE_WIDE = 2 };

constexpr int TM_THREADS = 384;       // threads of a block whose lifting size fills it (Z = 384), and of every fused block
constexpr int TM_THREADS_WIDE = 512;  // decode-only blocks of smaller lifting sizes hold floor(512 / roundup32(Z)) groups
constexpr int TM_COLS_PER_WARP = 170; // 512 TMEM columns / 3 warps per lane quadrant (<= 12 warps)
constexpr int TM_COLS_PER_WARP_WIDE = 128;  // 4 warps per lane quadrant (13 .. 16 warps); messages beyond go to shared memory

struct __align__(8) TmEdge {
  unsigned s4;  // 4 * (shift % Z): rotation in bytes
  unsigned cb;  // byte offset of the column's soft values inside the group's soft array: col * 4 * Z
};

constexpr unsigned TM_CRC_POW_LEN = 22 * 384;  // longest message in bits
// x^e mod P for CRC16 / CRC24A / CRC24B (poly codes 1..3), e < TM_CRC_POW_LEN: the parallel CRC check of the decoder.
// Filled once per device by upload_constants (ldpc_api.cu).
__device__ __align__(32) unsigned g_crc_pow[3][TM_CRC_POW_LEN];

// Development aid (build with -DTM_PHASE_TIMING, tools/phase_timing.py): cycles thread 0 of every group spends in the
// phases of a pair, summed over the launch.
#ifdef TM_PHASE_TIMING
__device__ unsigned long long g_tm_phase[16];
#define TM_PHASE(i)                                                                 \
  do {                                                                              \
    if (t == 0) {                                                                   \
      const long long now_ = clock64();                                             \
      atomicAdd(&g_tm_phase[i], (unsigned long long)(now_ - phase_t0_));            \
      phase_t0_ = now_;                                                             \
    }                                                                               \
  } while (0)
#else
#define TM_PHASE(i) do { } while (0)
#endif

struct TmemParams {
  const CbWork* work;
  const int8_t* llr_base;   // decoder input LLRs; fused: the soft-buffer arena (same memory as soft_rw)
  const int8_t* rm_base;    // fused: rate-matched LLRs of the batch
  int8_t* soft_rw;          // fused: the soft-buffer arena, updated in place by the prologue
  uint8_t* out_base;
  CbResult* results;
  unsigned* pair_counter;  // dynamic distribution of codeblock pairs over the groups of all blocks
  unsigned nof_cbs;
  int Z;
  int Z4;           // 4 * Z
  int zpad;         // threads per group = roundup32(Z)
  int groups;       // codeblock pairs processed concurrently by one block
  int max_iterations;
  unsigned group_stride;  // bytes of shared memory per group
  // per row: class | first entry in edge[] << 8 | first message word << 20. class = 2 * (number of high-rate edges) + (1 if
  // the row has an extension node); message word: TMEM word relative to the warp's column block (see tm_cap)
  unsigned rowdesc[LDPC_BG1_M + 2];
  // The schedule of one iteration: a step is one row, or two consecutive rows without a common high-rate column.
  //   x = class | rows of the step << 8                               (pair class: 16 * DA + DB, DA >= DB)
  //   y = first edge | first message word << 16 of row A, z = the same of row B
  //   w = byte offset of the extension node's word of row A | of row B << 16 (behind the thread's word of row 4)
  // One 16-byte uniform load per step, fetched one step ahead; a zero sentinel follows the last step.
  uint4 steps[LDPC_BG1_M + 1];
  // per number of layers L: steps that cover rows [0, L) | rows they cover << 8. A layer limit inside a pair runs the whole
  // pair: the surplus row sees an all-zero extension node (the kernel clears its word) and leaves every soft value as it is.
  unsigned short layer_steps[LDPC_BG1_M + 2];
  unsigned short tm_words;             // TMEM words used per thread
  unsigned short tm_cap;               // TMEM words a thread owns (rowtm >= tm_cap: overflow word rowtm - tm_cap in shared memory)
  unsigned short ovf_words;            // overflow words per thread
  unsigned ovf_stride;                 // bytes between consecutive overflow words of one thread (4 x threads per block)
  unsigned char scale_lut[LLR_MAX + 1 + 3];  // TM_SCALE_LUT: scaled magnitude for m in [0, 120] (any factor, any rounding mode)
  TmEdge edge[LDPC_BG1_EDGES];         // high-rate edges only, row after row (extension edges are implicit)
};

// ---- small helpers on raw half2 bit patterns -------------------------------------------------------------------------
namespace tm {

__device__ __forceinline__ __half2 h2(unsigned x) { return *reinterpret_cast<__half2*>(&x); }
__device__ __forceinline__ unsigned un(__half2 x) { return *reinterpret_cast<unsigned*>(&x); }
__device__ __forceinline__ unsigned hc(float v) { return un(__float2half2_rn(v)); }
__device__ __forceinline__ unsigned hadd(unsigned a, unsigned b) { return un(__hadd2(h2(a), h2(b))); }
__device__ __forceinline__ unsigned hsub(unsigned a, unsigned b) { return un(__hsub2(h2(a), h2(b))); }
__device__ __forceinline__ unsigned hfma(unsigned a, unsigned b, unsigned c) { return un(__hfma2(h2(a), h2(b), h2(c))); }
__device__ __forceinline__ unsigned hfma_sat(unsigned a, unsigned b, unsigned c) { return un(__hfma2_sat(h2(a), h2(b), h2(c))); }
__device__ __forceinline__ unsigned hmin(unsigned a, unsigned b) { return un(__hmin2(h2(a), h2(b))); }
__device__ __forceinline__ unsigned hmax(unsigned a, unsigned b) { return un(__hmax2(h2(a), h2(b))); }
// |a| inside the consumer (HMNMX2 / HFMA2 take |R| operands for free)
__device__ __forceinline__ unsigned hmin_abs(unsigned a, unsigned b) { return un(__hmin2(__habs2(h2(a)), __habs2(h2(b)))); }
__device__ __forceinline__ unsigned hmax_abs(unsigned a, unsigned b) { return un(__hmax2(__habs2(h2(a)), __habs2(h2(b)))); }
__device__ __forceinline__ unsigned hadd_sat_abs(unsigned a, unsigned b) { return un(__hadd2_sat(__habs2(h2(a)), h2(b))); }
__device__ __forceinline__ unsigned hfma_relu_abs(unsigned a, unsigned b, unsigned c)
{
  return un(__hfma2_relu(__habs2(h2(a)), h2(b), h2(c)));
}
__device__ __forceinline__ unsigned hfma_relu(unsigned a, unsigned b, unsigned c) { return un(__hfma2_relu(h2(a), h2(b), h2(c))); }
__device__ __forceinline__ unsigned hfma_abs(unsigned a, unsigned b, unsigned c) { return un(__hfma2(__habs2(h2(a)), h2(b), h2(c))); }
__device__ __forceinline__ unsigned hmul(unsigned a, unsigned b) { return un(__hmul2(h2(a), h2(b))); }
// integer min / max on the two non-negative 16-bit halves (fp16 bit patterns of non-negative values order like integers)
__device__ __forceinline__ unsigned umin2(unsigned a, unsigned b)
{
  unsigned r;
  asm("min.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ unsigned umax2(unsigned a, unsigned b)
{
  unsigned r;
  asm("max.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ unsigned umin3(unsigned a, unsigned b, unsigned c) { return umin2(umin2(a, b), c); }

__device__ __forceinline__ unsigned lds32(unsigned addr)
{
  unsigned r;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts32(unsigned addr, unsigned v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// FULL: the block is one group of 384 lanes, all of them live (Z = 384): barrier 1 with immediate operands.
template <bool FULL = false>
__device__ __forceinline__ void group_barrier(int id, int nthreads)
{
  if (FULL || nthreads == 384) {  // one group per block (no operand register to rebuild per row)
    asm volatile("bar.sync 1, 384;" ::: "memory");
  } else {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
  }
}

// ---- the row barrier, split ------------------------------------------------------------------------------------------
// A row may read the soft values only after every warp of the group has stored the previous row's. With bar.sync a warp
// that is ahead waits until the LAST warp has reached its own barrier instruction, i.e. has also finished the dispatch and
// address arithmetic of the next row. An mbarrier separates the two halves: a warp arrives as soon as its stores are
// out (one elected lane, after __syncwarp), works out the next row, and only then waits - for the arrivals, not for the
// other warps' progress beyond them.
__device__ __forceinline__ void row_arrive(unsigned bar, unsigned t4)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void row_wait(unsigned bar, unsigned& parity)
{
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "ROW_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra ROW_DONE;\n"
      "bra ROW_WAIT;\n"
      "ROW_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
  parity ^= 1u;
}

// ---- tensor memory as thread-private message storage (shape 32x32b: thread i of the warp <-> TMEM lane base + i) -----
__device__ __forceinline__ void tm_ld8(unsigned ta, unsigned* r)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(ta));
}
__device__ __forceinline__ void tm_ld2(unsigned ta, unsigned* r)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(ta));
}
__device__ __forceinline__ void tm_ld4(unsigned ta, unsigned* r)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(ta));
}
__device__ __forceinline__ void tm_ld1(unsigned ta, unsigned* r)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r[0]) : "r"(ta));
}
template <int W>
__device__ __forceinline__ void tm_load(unsigned ta, unsigned* r)
{
  if constexpr (W >= 8) {
    tm_ld8(ta, r);
    tm_load<W - 8>(ta + 8, r + 8);
  } else if constexpr (W >= 4) {
    tm_ld4(ta, r);
    tm_load<W - 4>(ta + 4, r + 4);
  } else if constexpr (W >= 2) {
    tm_ld2(ta, r);
    tm_load<W - 2>(ta + 2, r + 2);
  } else if constexpr (W == 1) {
    tm_ld1(ta, r);
  }
}
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_st1(unsigned ta, const unsigned* r)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(ta), "r"(r[0]) : "memory");
}
__device__ __forceinline__ void tm_st2(unsigned ta, const unsigned* r)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(ta), "r"(r[0]), "r"(r[1]) : "memory");
}
__device__ __forceinline__ void tm_st4(unsigned ta, const unsigned* r)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(ta), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3])
               : "memory");
}
__device__ __forceinline__ void tm_st8(unsigned ta, const unsigned* r)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(ta), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
template <int W>
__device__ __forceinline__ void tm_store(unsigned ta, const unsigned* r)
{
  if constexpr (W >= 8) {
    tm_st8(ta, r);
    tm_store<W - 8>(ta + 8, r + 8);
  } else if constexpr (W >= 4) {
    tm_st4(ta, r);
    tm_store<W - 4>(ta + 4, r + 4);
  } else if constexpr (W >= 2) {
    tm_st2(ta, r);
    tm_store<W - 2>(ta + 2, r + 2);
  } else if constexpr (W == 1) {
    tm_st1(ta, r);
  }
}

constexpr unsigned SIGN2 = 0x80008000u;

// sign(a) * min(|a|, |b|) for b > 0 (one HMNMX2.XORSIGN)
__device__ __forceinline__ unsigned hmin_xorsign_abs(unsigned a, unsigned b)
{
  unsigned r;
  asm("min.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

constexpr float TM_S = 544.f;          // scaled domain: X = 544 x (see the header comment)
constexpr float TM_CAP = 65280.f;      // 544 * 120

// Scaled magnitude 544 Q(m) of M = 544 m, m integer in [0,120] (closed forms checked against the host table for every m
// by tm_scale_selfcheck() when the algorithm is set):
//  GENERIC 0.8: round(0.8 m) = m - round(m / 5)   (0.2 m is never at a tie: the relative error 2^-11 of the fp16 constant
//               0.2 / 544 moves m / 5 by less than 0.012, and one FMA rounding against the magic constant is exact)
//  AVX     0.8: (m * 52428) >> 16 = m - floor(m / 5) - [m > 0], floor(m / 5) = round((m - 2) / 5)
// The magic constant 1536 keeps the sum in [1024, 2048) where the fp16 ulp is 1.
//  LUT     any other table (scaling factor != 0.8, the NEON variant's rounding): m = M / 544 read off the mantissa of
//          fma(M, 1/544, 1024) = 1024 + m (ulp 1), one 16-bit shared-memory load per half of 544 * lut[m] as fp16 bits
template <int SCALE>
__device__ __forceinline__ unsigned scale_q(unsigned M, unsigned lut = 0, unsigned fifth = 0x0e060e06u)
{
  // |M| throughout (an operand modifier): min1 arrives with the sign parity of the row in its sign bits
  const unsigned magic = hc(1536.f), nS = hc(-TM_S);
  if (SCALE == TM_SCALE_LUT) {
    const unsigned idx = hfma_abs(M, hc(1.f / TM_S), hc(1024.f));
    unsigned qa, qb;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(qa) : "r"(lut + 2u * (idx & 0x3ffu)));
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(qb) : "r"(lut + 2u * ((idx >> 16) & 0x3ffu)));
    return qa | (qb << 16);
  } else if (SCALE == TM_SCALE_GENERIC_08) {
    const unsigned r = hsub(hfma_abs(M, fifth, magic), magic);
    return un(__hfma2(h2(r), h2(nS), __habs2(h2(M))));
  } else {
    const unsigned r = hsub(hfma(un(__hsub2(__habs2(h2(M)), h2(hc(2.f * TM_S)))), fifth, magic), magic);
    const unsigned nz = un(__hfma2_sat(__habs2(h2(M)), h2(hc(1.f)), h2(hc(0.f))));
    return hfma(nz, nS, un(__hfma2(h2(r), h2(nS), __habs2(h2(M)))));
  }
}

// (min1, min2) of a set kept as a sorted pair of non-negative bit patterns.
struct MinPair {
  unsigned m1, m2;
};
__device__ __forceinline__ MinPair merge(MinPair a, MinPair b)
{
  MinPair r;
  // m1 carries the sign parity of its set in its sign bits (XORSIGN), m2 is a non-negative bit pattern (fp16 order ==
  // integer order)
  r.m1 = hmin_xorsign_abs(a.m1, b.m1);
  r.m2 = umin3(hmax_abs(a.m1, b.m1), a.m2, b.m2);
  return r;
}
// Tree reduction over N sorted pairs.
template <int N>
__device__ __forceinline__ MinPair reduce_pairs(MinPair* p)
{
  if constexpr (N == 1) {
    return p[0];
  } else {
    constexpr int H = N / 2;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      p[i] = merge(p[2 * i], p[2 * i + 1]);
    }
    if constexpr (N & 1) {
      p[H] = p[N - 1];
      return reduce_pairs<H + 1>(p);
    } else {
      return reduce_pairs<H>(p);
    }
  }
}

// Per-thread context of the row loop.
struct Ctx {
  unsigned soft_base;  // shared address of the group's high-rate soft array (uniform per warp)
  unsigned ext_soft;   // shared address of this thread's extension-node word of row 4 (row r: + (r - 4) * zpad4)
  unsigned t4;         // 4 * lane
  unsigned Z4;
  unsigned zpad4;
  unsigned tm_addr;    // TMEM address of this warp's column block (lane base in the upper half)
  unsigned tm_cap;     // words of that block; message words beyond it live in shared memory:
  unsigned ovf;        //   shared address of this thread's overflow word 0, word i at + i * ovf_stride
  unsigned ovf_stride;
  unsigned k64, k12, kfifth, kinv;  // constants the row bodies want in registers (see the kernel)
  unsigned row_bar;    // shared address of the group's row mbarrier
  unsigned row_parity; // phase the next row_wait waits for
  unsigned lut;        // TM_SCALE_LUT: shared address of the scaling table (fp16 bit patterns of 544 * lut[m])
  bool active;         // lane < Z
};

// One check row: D high-rate edges (rotated accesses to shared soft values, messages in cw[]) plus, when EXT, the row's
// private extension node. Split into the part that does not touch shared soft values (pre: runs while slower warps finish
// the previous row), the loads and the arithmetic, so that two rows without a common column can be interleaved between
// one pair of barriers (pair_body). e0 / r / tmw are warp-uniform.
template <int D, bool EXT, int SCALE, int MODE>
struct Row {
  static constexpr bool FULL = MODE == TM_MODE_FULL, WIDE = MODE == TM_MODE_WIDE;
  static constexpr int NV = D + (EXT ? 1 : 0);  // values entering the check
  static constexpr int W = (D + 1) / 2;         // message words of the row
  unsigned addr[D], u[NV], cw[W];
  unsigned xaddr, tma;

  __device__ __forceinline__ void pre(const Ctx& c, const TmemParams& p, unsigned xoff, int e0, unsigned tmw)
  {
    // previous messages of the row (thread-private TMEM columns); the latency overlaps the shared-memory loads. Blocks of
    // more than 12 warps own 128 columns per warp: the last rows' messages are thread-private shared-memory words then.
    if (WIDE && tmw >= c.tm_cap) {
      tma = 0x80000000u | (c.ovf + (tmw - c.tm_cap) * c.ovf_stride);
#pragma unroll
      for (int i = 0; i < W; ++i) {
        cw[i] = lds32((tma & 0x7fffffffu) + (unsigned)i * c.ovf_stride);
      }
    } else {
      tma = c.tm_addr + tmw;
      tm_load<W>(tma, cw);
    }
    // addresses: lane j of column n talks to check lane (j - shift) mod Z (ldpc_decoder_generic.cpp:66)
#pragma unroll
    for (int k = 0; k < D; ++k) {
      const TmEdge ed = p.edge[e0 + k];
      const unsigned a = c.t4 + ed.s4;
      addr[k] = min(a, a - (FULL ? 4u * MAX_Z : c.Z4)) + (ed.cb + c.soft_base);
    }
    // The row's extension node has degree one: its variable-to-check message is its channel LLR until the soft value
    // LLR + c2v is promoted to +-inf (|sum| > 120), and +-inf for ever after (soft - c2v_old = LLR exactly otherwise:
    // log_likelihood_ratio.cpp:38-85). So the node is ONE thread-private word holding that message - no soft value, no
    // stored check-to-variable message.
    xaddr = 0;
    if constexpr (EXT) {
      xaddr = c.ext_soft + xoff;
      u[D] = lds32(xaddr);  // thread-private: no ordering with other threads needed
    }
  }

  __device__ __forceinline__ void load()
  {
#pragma unroll
    for (int k = 0; k < D; ++k) {
      u[k] = lds32(addr[k]);
    }
  }

  // after load() and tm_wait_ld()
  __device__ __forceinline__ void compute(const Ctx& c)
  {
    // previous messages: byte b = 128 + c_old; 0x6400 | b is the fp16 number 1152 + c_old
    const unsigned cap = hc(TM_CAP), nS = hc(-TM_S);
#pragma unroll
    for (int k = 0; k < D; ++k) {
      const unsigned tb = __byte_perm(cw[k >> 1], c.k64, (k & 1) ? 0x4342 : 0x4140);
      // variable-to-check message (compute_var_to_check_msgs): difference (overflowing beyond +-120), then the clamp of
      // operator- for a finite soft value / the pass-through of an infinite one
      const unsigned d = hfma(hsub(tb, hc(1152.f)), nS, u[k]);
      const unsigned lim = hfma_abs(u[k], c.k12, cap);  // 65280, or +inf for an infinite soft value
      u[k] = hmin_xorsign_abs(d, lim);
    }
    // analyze_var_to_check_msgs (ldpc_decoder_generic.cpp:53-78): min1, min2 (capped at 120), sign parity
    // Sorted pairs; the sign bits of m1 accumulate the sign parity of the values on the way (min.xorsign.abs), the cap
    // (120, 120) enters as the partner of an odd last value or as a final clamp.
    constexpr int NP = (NV + 1) / 2;
    MinPair pr[NP];
#pragma unroll
    for (int i = 0; i < NV / 2; ++i) {
      pr[i].m1 = hmin_xorsign_abs(u[2 * i], u[2 * i + 1]);
      pr[i].m2 = hmax_abs(u[2 * i], u[2 * i + 1]);
    }
    if constexpr (NV & 1) {
      pr[NP - 1].m1 = hmin_xorsign_abs(u[NV - 1], cap);
      pr[NP - 1].m2 = cap;
    }
    MinPair mm = reduce_pairs<NP>(pr);
    if constexpr (!(NV & 1)) {
      mm.m1 = hmin_xorsign_abs(mm.m1, cap);
      mm.m2 = hmin(mm.m2, cap);
    }
    const unsigned T = mm.m1 & SIGN2;
    // scale_llr (ldpc_decoder_generic.cpp:80-89), sign parity folded in
    const unsigned Q1 = scale_q<SCALE>(mm.m1, c.lut, c.kfifth), Q2 = scale_q<SCALE>(mm.m2, c.lut, c.kfifth);
    const unsigned Q2s = Q2 ^ T;
    const unsigned nD = hsub(Q1, Q2) ^ T;
    const unsigned nM1 = mm.m1 | SIGN2;
    // compute_check_to_var_msgs (:93-116) + compute_soft_bits (:118-130)
    unsigned tn[D + 1];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const unsigned uk = u[k];
      const unsigned indp = hadd_sat_abs(uk, nM1);  // 0 where |v2c| == min1, 1 elsewhere (differences are 0 or >= 544)
      const unsigned cm = hfma(indp, nD, Q2s);
      const unsigned cn = cm ^ (uk & SIGN2);
      // promotion_sum: |sum| > 120 overflows to +-inf, an infinite v2c stays infinite
      const unsigned n = hadd(uk, cn);
      if (k < D) {
        if (FULL || c.active) {
          sts32(addr[k], n);
        }
        tn[k] = hfma(cn, c.kinv, hc(1152.f));
      } else {
        // the message becomes +-inf once the soft value is promoted (sign of the sum = sign of the message then)
        const unsigned prom = __heq2_mask(__habs2(h2(n)), h2(0x7c007c00u));
        sts32(xaddr, (n & prom) | (uk & ~prom));
      }
    }
    if constexpr (D & 1) {
      tn[D] = tn[D - 1];
    }
    unsigned wv[W];
#pragma unroll
    for (int i = 0; i < W; ++i) {
      wv[i] = __byte_perm(tn[2 * i], tn[2 * i + 1], 0x6420);
    }
    if (WIDE && (tma & 0x80000000u)) {
#pragma unroll
      for (int i = 0; i < W; ++i) {
        sts32((tma & 0x7fffffffu) + (unsigned)i * c.ovf_stride, wv[i]);
      }
    } else {
      tm_store<W>(tma, wv);
    }
  }
};

template <int D, bool EXT, int SCALE, int MODE>
__device__ __forceinline__ void row_body(Ctx& c, const TmemParams& p, unsigned xoff, int e0, unsigned tmw, int bar_id, int zpad)
{
  Row<D, EXT, SCALE, MODE> R;
  R.pre(c, p, xoff, e0, tmw);
  // The layered schedule orders rows through the shared soft values only: everything above (row descriptor, dispatch,
  // message load, addresses, extension node) runs while slower warps finish the previous row.
  row_wait(c.row_bar, c.row_parity);
  R.load();
  tm_wait_ld();
  R.compute(c);
  row_arrive(c.row_bar, c.t4);
}

// Two consecutive rows WITHOUT a common high-rate column (their extension columns are private anyway): they read and
// write disjoint soft values, so running them side by side between one pair of barriers gives exactly the result of
// running them one after the other - two independent dependency chains per thread, half the barriers and half the loop /
// dispatch work for these (short) rows. In both 5G base graphs most rows of the extension part are pairwise orthogonal
// to their successor (BG1: 14 pairs among rows 16-45, BG2: 13).
template <int DA, int DB, int SCALE, int MODE>
__device__ __forceinline__ void pair_body(Ctx& c, const TmemParams& p, unsigned xoffA, int e0A, unsigned tmwA, unsigned xoffB, int e0B,
                                          unsigned tmwB, int bar_id, int zpad)
{
  Row<DA, true, SCALE, MODE> A;
  Row<DB, true, SCALE, MODE> B;
  A.pre(c, p, xoffA, e0A, tmwA);
  B.pre(c, p, xoffB, e0B, tmwB);
  row_wait(c.row_bar, c.row_parity);
  A.load();
  B.load();
  tm_wait_ld();
  A.compute(c);
  B.compute(c);
  row_arrive(c.row_bar, c.t4);
}

// One step of the schedule (TmemParams::steps): single-row classes are < 40, pair classes (16 * DA + DB) >= 50.
constexpr bool tm_row_used(int bg, int d, bool ext)
{
  return bg == 1 ? (ext ? (d == 2 || (d >= 5 && d <= 9)) : d == 19) : (ext ? (d >= 3 && d <= 5) : (d == 8 || d == 10));
}
constexpr bool tm_pair_used(int bg, int da, int db)
{
  return bg == 1 ? ((da == 5 && (db == 5 || db == 4)) || (da == 4 && (db == 4 || db == 3)))
                 : ((da == 4 && (db == 3 || db == 2)) || (da == 3 && (db == 3 || db == 2)));
}

// Only the row shapes the base graph has are compiled into its kernels (tm_build_tables checks the schedule against the
// same predicates).
template <int BG, int SCALE, int MODE>
__device__ __forceinline__ void step_dispatch(Ctx& c, const TmemParams& p, int bar_id, int zpad, const uint4 st)
{
  const int cls = st.x & 0xffu;
  const unsigned xoffA = st.w & 0xffffu, xoffB = st.w >> 16;
  const int e0A = st.y & 0xffffu, e0B = st.z & 0xffffu;
  const unsigned tmwA = st.y >> 16, tmwB = st.z >> 16;
  switch (cls) {
#define LDPC_TM_CASE(D, EXT)                        \
  case 2 * (D) + (EXT):                             \
    if constexpr (tm_row_used(BG, D, (EXT) != 0)) {  \
      row_body<D, (EXT) != 0, SCALE, MODE>(c, p, xoffA, e0A, tmwA, bar_id, zpad); \
    }                                               \
    break;
#define LDPC_TM_PAIR(DA, DB)                                                                   \
  case 16 * (DA) + (DB):                                                                       \
    if constexpr (tm_pair_used(BG, DA, DB)) {                                                  \
      pair_body<DA, DB, SCALE, MODE>(c, p, xoffA, e0A, tmwA, xoffB, e0B, tmwB, bar_id, zpad);  \
    }                                                                                          \
    break;
    LDPC_TM_CASE(2, 1)
    LDPC_TM_CASE(3, 1)
    LDPC_TM_CASE(4, 1)
    LDPC_TM_CASE(5, 1)
    LDPC_TM_CASE(6, 1)
    LDPC_TM_CASE(7, 1)
    LDPC_TM_CASE(8, 1)
    LDPC_TM_CASE(9, 1)
    LDPC_TM_CASE(8, 0)
    LDPC_TM_CASE(10, 0)
    LDPC_TM_CASE(19, 0)
    LDPC_TM_PAIR(5, 5)
    LDPC_TM_PAIR(5, 4)
    LDPC_TM_PAIR(4, 4)
    LDPC_TM_PAIR(4, 3)
    LDPC_TM_PAIR(4, 2)
    LDPC_TM_PAIR(3, 3)
    LDPC_TM_PAIR(3, 2)
#undef LDPC_TM_PAIR
#undef LDPC_TM_CASE
    default:
      break;
  }
}

// The fused kernels' dispatch. Runs the step `st` (number s). The rows of the high-rate core (no extension node: BG1 4 x 19 edges, BG2 8 / 10) come in runs of
// one class: a run stays inside the class' code and skips the dispatch - at code rates close to 1 these rows are most of an
// iteration (`pusch_slot` +1.2 %; at R = 1/3 the loop-carried state costs 2 %, so the decode-only kernels keep the plain
// dispatch above). `next` is the prefetched descriptor of step s + 1 on entry and on return.
template <int BG, int SCALE, int MODE>
__device__ __forceinline__ void step_dispatch_runs(Ctx& c, const TmemParams& p, int bar_id, int zpad, uint4 st, uint4& next, int& s,
                                              int nof_steps)
{
  const unsigned cls = st.x & 0xffu;
#define LDPC_TM_ONCE(BODY)                                       \
  {                                                              \
    const unsigned xoffA = st.w & 0xffffu, xoffB = st.w >> 16;   \
    const int e0A = st.y & 0xffffu, e0B = st.z & 0xffffu;        \
    const unsigned tmwA = st.y >> 16, tmwB = st.z >> 16;         \
    (void)xoffB, (void)e0B, (void)tmwB;                          \
    BODY;                                                        \
    ++s;                                                         \
  }
#define LDPC_TM_RUN(BODY)                                        \
  for (;;) {                                                     \
    const unsigned xoffA = st.w & 0xffffu, xoffB = st.w >> 16;   \
    const int e0A = st.y & 0xffffu, e0B = st.z & 0xffffu;        \
    const unsigned tmwA = st.y >> 16, tmwB = st.z >> 16;         \
    (void)xoffB, (void)e0B, (void)tmwB;                          \
    BODY;                                                        \
    ++s;                                                         \
    if (s >= nof_steps || (next.x & 0xffu) != cls) {             \
      break;                                                     \
    }                                                            \
    st = next;                                                   \
    next = p.steps[s + 1];                                       \
  }
  switch (cls) {
#define LDPC_TM_CASE(D, EXT)                        \
  case 2 * (D) + (EXT):                             \
    if constexpr (tm_row_used(BG, D, (EXT) != 0)) {  \
      if constexpr ((EXT) == 0) {                    \
        LDPC_TM_RUN((row_body<D, false, SCALE, MODE>(c, p, xoffA, e0A, tmwA, bar_id, zpad))) \
      } else {                                       \
        LDPC_TM_ONCE((row_body<D, true, SCALE, MODE>(c, p, xoffA, e0A, tmwA, bar_id, zpad))) \
      }                                              \
    }                                               \
    break;
#define LDPC_TM_PAIR(DA, DB)                                                                   \
  case 16 * (DA) + (DB):                                                                       \
    if constexpr (tm_pair_used(BG, DA, DB)) {                                                  \
      LDPC_TM_ONCE((pair_body<DA, DB, SCALE, MODE>(c, p, xoffA, e0A, tmwA, xoffB, e0B, tmwB, bar_id, zpad)))  \
    }                                                                                          \
    break;
    LDPC_TM_CASE(2, 1)
    LDPC_TM_CASE(3, 1)
    LDPC_TM_CASE(4, 1)
    LDPC_TM_CASE(5, 1)
    LDPC_TM_CASE(6, 1)
    LDPC_TM_CASE(7, 1)
    LDPC_TM_CASE(8, 1)
    LDPC_TM_CASE(9, 1)
    LDPC_TM_CASE(8, 0)
    LDPC_TM_CASE(10, 0)
    LDPC_TM_CASE(19, 0)
    LDPC_TM_PAIR(5, 5)
    LDPC_TM_PAIR(5, 4)
    LDPC_TM_PAIR(4, 4)
    LDPC_TM_PAIR(4, 3)
    LDPC_TM_PAIR(4, 2)
    LDPC_TM_PAIR(3, 3)
    LDPC_TM_PAIR(3, 2)
#undef LDPC_TM_PAIR
#undef LDPC_TM_CASE
#undef LDPC_TM_RUN
#undef LDPC_TM_ONCE
    default:
      ++s;  // not reached: tm_build_tables only emits classes the kernel has
      break;
  }
}

// Byte j of a and of b (selector 0x0400 + 0x0101 j) as a scaled half2: byte ^ 0x80 under a 0x64 high byte is the fp16
// number 1024 + (value + 128); the multiplication by 544 overflows to +-inf exactly for |value| > 120 (the "infinite" LLR
// +-127, log_likelihood_ratio.h:239).
__device__ __forceinline__ unsigned convert_bytes(unsigned a, unsigned b, unsigned selector)
{
  const unsigned w = (__byte_perm(a, b, selector) & 0x00ff00ffu) ^ 0x64806480u;
  return hmul(hadd(w, hc(-1152.f)), hc(TM_S));
}

// 16 LLRs at in + i0 (16-byte aligned) of a codeblock with len LLRs; positions at or beyond len read as zero.
__device__ __forceinline__ uint4 load16(const int8_t* in, unsigned i0, unsigned len)
{
  uint4 r = make_uint4(0u, 0u, 0u, 0u);
  if (i0 + 16u <= len) {
    r = __ldg(reinterpret_cast<const uint4*>(in + i0));
  } else if (i0 < len) {
    unsigned w[4] = {0u, 0u, 0u, 0u};
    for (unsigned k = 0; i0 + k < len; ++k) {
      w[k >> 2] |= (unsigned)(unsigned char)__ldg(in + i0 + k) << (8u * (k & 3u));
    }
    r = make_uint4(w[0], w[1], w[2], w[3]);
  }
  return r;
}

// Warp-uniform value that the compiler can also see as uniform (the result of a warp reduction lives in a uniform
// register), so that loops and branches on it use the uniform datapath. v >= 0.
__device__ __forceinline__ int uniform(int v) { return (int)__reduce_max_sync(0xffffffffu, (unsigned)v); }

}  // namespace tm

// Shared memory: [crc tables 3 x 256 words][group 0][group 1]...; group = soft[HRR * Z] | ext_soft[(M-4) * zpad] |
// hard[2][msg_bytes16] | flags[16]. ext_soft holds one word per extension node (see row_body).
struct TmLayout {
  unsigned soft, ext_soft, hard, flags, stage, stage_stride, total;
};
// fused: + stage[2][N_short * Z] bytes, the int8 soft buffers of the pair while the dematcher prologue combines them
__host__ __device__ inline TmLayout tm_layout(int bg, int Z, int zpad, bool fused)
{
  const unsigned hrr = (bg == 2) ? 14 : 26, M = (bg == 2) ? LDPC_BG2_M : LDPC_BG1_M, K = (bg == 2) ? 10 : 22;
  const unsigned msg_bytes16 = (((unsigned)(K * Z + 7) / 8) + 15) & ~15u;
  TmLayout l;
  l.soft = 0;
  l.ext_soft = l.soft + hrr * Z * 4;
  l.hard = l.ext_soft + (M - 4) * zpad * 4;
  l.flags = l.hard + 2 * msg_bytes16;
  l.stage = (l.flags + 16 * 4 + 15) & ~15u;
  l.stage_stride = ((((bg == 2) ? 50u : 66u) * Z) + 15) & ~15u;
  l.total = l.stage + (fused ? 2 * l.stage_stride : 0);
  return l;
}
constexpr unsigned TM_CRC_TABLE_BYTES = 3 * 256 * 4 + 16;  // + the TMEM base address written by tcgen05.alloc

// FUSED: the prologue is the rate dematcher (ldpc_rate_dematcher_impl.cpp:43-254): every soft-buffer position is combined
// with the new transmission in the HBM arena and handed to the decoder in the same pass (no separate kernel, the soft
// buffer is read once and written once).
// FULL: Z = 384 (one group of 12 warps, every lane live) with Z, the barrier operands and the all-lanes-active predicate
// folded at compile time - the configuration BASELINE.json's metric is quoted on.
// WIDE: decode-only blocks of up to 512 threads for lifting sizes that leave room for more groups.
// Register cap: the fused kernels need their 168; for the decode-only kernels a cap of 120 makes ptxas settle on 86 registers
// with a schedule that measures 1 % faster than the 111 it takes when left alone (profiles/r02_kernel_experiments.md).
template <int BG, int SCALE, bool FUSED, int MODE>
__global__ void __launch_bounds__(MODE == TM_MODE_WIDE ? TM_THREADS_WIDE : TM_THREADS, 1) __maxnreg__(FUSED ? 168 : 120)
    ldpc_decode_tmem_kernel(const __grid_constant__ TmemParams p)
{
  constexpr bool FULL = MODE == TM_MODE_FULL;
  constexpr int K = (BG == 2) ? 10 : 22, M = (BG == 2) ? LDPC_BG2_M : LDPC_BG1_M, HRR = K + 4;
  extern __shared__ __align__(16) unsigned char smem[];
  const int Z = FULL ? MAX_Z : p.Z, zpad = FULL ? MAX_Z : p.zpad, tid = threadIdx.x;
  const int warp = tid >> 5;
  const int grp = FULL ? 0 : tm::uniform(tid / zpad);
  const int t = tid - grp * zpad;
  const int wig = tm::uniform(t >> 5);   // warp index inside the group
  const int bar_id = 1 + grp;
  const unsigned msg_len = (unsigned)(K * Z), msg_bytes = (msg_len + 7) / 8, msg_bytes16 = (msg_bytes + 15) & ~15u;
  const TmLayout lay = tm_layout(BG, Z, zpad, FUSED);

  // tensor memory: the whole 512 columns (one block per SM), released at the end
  if (warp == 0) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smem + 3 * 256 * 4);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (SCALE == TM_SCALE_LUT && tid <= LLR_MAX) {  // the first 3 KB of shared memory are free for it
    reinterpret_cast<unsigned short*>(smem)[tid] = __half_as_ushort(__float2half_rn(tm::TM_S * (float)p.scale_lut[tid]));
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");

  unsigned char* gsm = smem + TM_CRC_TABLE_BYTES + (size_t)grp * p.group_stride;
  unsigned* soft_w = reinterpret_cast<unsigned*>(gsm + lay.soft);
  unsigned* ext_soft_w = reinterpret_cast<unsigned*>(gsm + lay.ext_soft);
  unsigned char* hard = gsm + lay.hard;
  // 0,1: last non-zero + 1; 4: pair; 8..11: CRC remainder accumulators of (A, B) for even / odd iterations
  volatile unsigned* flags = reinterpret_cast<volatile unsigned*>(gsm + lay.flags);

  tm::Ctx c;
  c.soft_base = (unsigned)tm::uniform((int)(unsigned)__cvta_generic_to_shared(soft_w));
  c.ext_soft = (unsigned)__cvta_generic_to_shared(ext_soft_w + t);
  c.t4 = 4u * (unsigned)t;
  c.Z4 = (unsigned)p.Z4;
  c.zpad4 = 4u * (unsigned)zpad;
  const unsigned tmem_base = *reinterpret_cast<volatile unsigned*>(smem + 3 * 256 * 4);
  c.tm_addr = (unsigned)tm::uniform((int)(tmem_base + (((unsigned)(warp & 3) * 32u) << 16) + (unsigned)(warp >> 2) * (unsigned)p.tm_cap));
  c.tm_cap = p.tm_cap;
  c.ovf_stride = p.ovf_stride;
  // overflow message words (blocks of more than 12 warps): behind the groups' buffers, word i of thread tid at
  // (i * threads + tid) * 4 - consecutive threads, consecutive banks
  c.ovf = (unsigned)__cvta_generic_to_shared(smem + TM_CRC_TABLE_BYTES + (size_t)p.groups * p.group_stride) + 4u * (unsigned)tid;
  c.lut = (unsigned)__cvta_generic_to_shared(smem);
  c.active = FULL || t < Z;
  c.row_bar = (unsigned)tm::uniform((int)(unsigned)__cvta_generic_to_shared(gsm + lay.flags + 56));  // flags[14..15]
  c.row_parity = 0;
  if (t == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(c.row_bar), "r"(zpad) : "memory");
  }
  {
    // Constants that are the register operand of a PRMT / HFMA2 in every edge: derived from a kernel parameter so that the
    // compiler keeps them in registers instead of rebuilding them (IMAD.MOV + PRMT) in front of every use.
    const unsigned zero = (unsigned)p.max_iterations >> 31;
    c.k64 = 0x64646464u | zero;
    c.k12 = tm::hc(0.000244140625f) | zero;
    c.kfifth = tm::hc(0.2f / tm::TM_S) | zero;
    c.kinv = tm::hc(1.f / tm::TM_S) | zero;
  }

  // The pair after the current one is claimed one pair early: the atomic's round trip runs under this pair's loads instead
  // of in front of a barrier, and the next pair's inputs (fused: rate-matched LLRs and previous soft bits) are pulled into
  // L2 while this pair is decoded.
  constexpr bool PREFETCH = true;
  if (PREFETCH && t == 0) {
    flags[5] = atomicAdd(p.pair_counter, 1u);
  }
#ifdef TM_PHASE_TIMING
  long long phase_t0_ = clock64();
#endif
  for (;;) {
    // next pair of codeblocks for this group
    if (t == 0) {
      unsigned cur;
      if (PREFETCH) {
        cur = flags[5];  // claimed while the previous pair was staged in; the claim after that follows below
      } else {
        cur = atomicAdd(p.pair_counter, 1u);
        flags[5] = 0x7fffffffu;
      }
      flags[4] = cur;
      flags[0] = 0, flags[1] = 0, flags[8] = 0, flags[9] = 0, flags[10] = 0, flags[11] = 0;
    }
    tm::group_barrier<FULL>(bar_id, zpad);
    const unsigned pair = (unsigned)tm::uniform((int)flags[4]);
    if (2 * pair >= p.nof_cbs) {
      break;
    }
    unsigned claimed_early = 0;
    if (!FUSED && t == 0) {
      claimed_early = atomicAdd(p.pair_counter, 1u);  // consumed behind the loads (fused: issued under the staging phase)
    }
    const bool hasB = 2 * pair + 1 < p.nof_cbs;
    const CbWork wA = p.work[2 * pair];
    const CbWork wB = p.work[hasB ? 2 * pair + 1 : 2 * pair];
    const bool useA = !wA.skip_decode, useB = hasB && !wB.skip_decode;
    if (!FUSED && !useA && !useB) {
      if (t == 0) {
        flags[5] = claimed_early;
      }
      tm::group_barrier<FULL>(bar_id, zpad);
      continue;
    }
    // load_soft_bits (ldpc_decoder_impl.cpp:148-173) for both codeblocks: int8 -> fp16, +-127 -> +-inf; punctured and
    // missing nodes 0. Messages start at zero (first visit of a layer, ldpc_decoder_impl.cpp:181-185).
    const int8_t* inA = p.llr_base + wA.in_offset;
    const int8_t* inB = p.llr_base + wB.in_offset;
    unsigned lastA = 0, lastB = 0;
    auto convert = [](int a, int b) { return tm::convert_bytes((unsigned)a, (unsigned)b, 0x0400); };
    auto put = [&](int n, unsigned v) {
      if (n < HRR) {
        soft_w[n * Z + t] = v;
      } else {
        ext_soft_w[(n - HRR) * zpad + t] = v;
      }
    };
    // Loads are issued in batches of U columns before any of them is consumed: the prologue is pure DRAM latency
    // (profiles/r01_ncu_tmem_full_summary.txt: one-at-a-time loads were 12 % of all stall samples).
    constexpr int U = (BG == 2) ? 10 : 11;
    static_assert((K + M - 2) % U == 0, "batches must tile the stored columns");
    if (c.active) {
      soft_w[t] = 0, soft_w[Z + t] = 0;  // punctured columns
    }
    if constexpr (!FUSED) {
      const unsigned lenA = useA ? wA.in_length : 0u, lenB = useB ? wB.in_length : 0u;
      const bool vec16 = tm::uniform(((Z & 15) == 0 && ((reinterpret_cast<size_t>(inA) | reinterpret_cast<size_t>(inB)) & 15) == 0) ? 1 : 0);
      if (vec16) {
        // Z is a multiple of 16 and both inputs are 16-byte aligned: a thread takes 16 consecutive LLRs (16 lanes of one
        // column) of both codeblocks with two 16-byte loads and writes their 16 half2 words with four 16-byte stores.
        // Input byte 16 v .. 16 v + 15 = lanes 16 (v mod Z/16) .. of stored column v / (Z/16). The trailing-zero scan only
        // has to be exact to the vector: the number of layers depends on the input size rounded up to a multiple of Z.
        const unsigned V = (unsigned)Z >> 4, nvec = (unsigned)(K + M - 2) * V;
        constexpr int UV = 3;
#pragma unroll 1
        for (unsigned v0 = (unsigned)t; v0 < nvec; v0 += UV * (unsigned)zpad) {
          uint4 a[UV], b[UV];
#pragma unroll
          for (int j = 0; j < UV; ++j) {
            const unsigned v = v0 + (unsigned)(j * zpad), i0 = 16u * v;
            a[j] = tm::load16(inA, i0, v < nvec ? lenA : 0u);
            b[j] = tm::load16(inB, i0, v < nvec ? lenB : 0u);
          }
#pragma unroll
          for (int j = 0; j < UV; ++j) {
            const unsigned v = v0 + (unsigned)(j * zpad);
            if (v < nvec) {
              lastA = ((a[j].x | a[j].y) | (a[j].z | a[j].w)) ? 16u * (v + 1) : lastA;
              lastB = ((b[j].x | b[j].y) | (b[j].z | b[j].w)) ? 16u * (v + 1) : lastB;
              const unsigned n = 2u + v / V, lane0 = 16u * (v - (n - 2u) * V);
              unsigned* dst = (n < (unsigned)HRR) ? soft_w + n * (unsigned)Z + lane0 : ext_soft_w + (n - HRR) * (unsigned)zpad + lane0;
              const unsigned aw[4] = {a[j].x, a[j].y, a[j].z, a[j].w}, bw[4] = {b[j].x, b[j].y, b[j].z, b[j].w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint4 o;
                o.x = tm::convert_bytes(aw[q], bw[q], 0x0400);
                o.y = tm::convert_bytes(aw[q], bw[q], 0x0501);
                o.z = tm::convert_bytes(aw[q], bw[q], 0x0602);
                o.w = tm::convert_bytes(aw[q], bw[q], 0x0703);
                *reinterpret_cast<uint4*>(dst + 4 * q) = o;
              }
            }
          }
        }
      } else if (c.active) {
#pragma unroll 1
        for (int n0 = 2; n0 < K + M; n0 += U) {
          int a[U], b[U];
#pragma unroll
          for (int k = 0; k < U; ++k) {
            const unsigned i = (unsigned)(n0 - 2 + k) * Z + t;
            a[k] = (i < lenA) ? (int)__ldg(inA + i) : 0;
            b[k] = (i < lenB) ? (int)__ldg(inB + i) : 0;
          }
#pragma unroll
          for (int k = 0; k < U; ++k) {
            const int n = n0 + k;
            const unsigned i = (unsigned)(n - 2) * Z + t;
            lastA = a[k] ? i + 1 : lastA;
            lastB = b[k] ? i + 1 : lastB;
            put(n, convert(a[k], b[k]));
          }
        }
      }
    } else {
      // Rate dematching + HARQ combining as the load (ldpc_dematch.cuh, phases A-C), on a shared-memory copy of the two
      // soft buffers; codeblocks with skip_decode are combined too. One pass over the arena: read once, written once.
      const DematchGeom gA = make_dematch_geom(BG, (unsigned)Z, wA.nof_filler_bits, wA.rv, wA.mod_order, wA.Nref,
                                               wA.rm_length, wA.new_data);
      const DematchGeom gB = make_dematch_geom(BG, (unsigned)Z, wB.nof_filler_bits, wB.rv, wB.mod_order, wB.Nref,
                                               wB.rm_length, wB.new_data);
      const int8_t* rmA = p.rm_base + wA.rm_offset;
      const int8_t* rmB = p.rm_base + wB.rm_offset;
      int8_t* softA = p.soft_rw + wA.in_offset;
      int8_t* softB = p.soft_rw + wB.in_offset;
      int8_t* stA = reinterpret_cast<int8_t*>(gsm + lay.stage);
      int8_t* stB = stA + lay.stage_stride;
      TM_PHASE(0);  // claim + work records + geometry
      // The claim of the pair after this one: the atomic's round trip to L2 (~1 us) runs under the staging and the
      // scatter's own loads instead of in front of the barrier every thread waits at (it was 3 % of the pair's time).
      unsigned claimed = 0;
      if (PREFETCH && t == 0) {
        claimed = (2 * pair < p.nof_cbs) ? atomicAdd(p.pair_counter, 1u) : pair;
      }
      // phase A: staging copy of the previous content (not read at all for a fresh first transmission)
      dematch_stage_in(gA, softA, stA, (unsigned)t, (unsigned)zpad);
      if (hasB) {
        dematch_stage_in(gB, softB, stB, (unsigned)t, (unsigned)zpad);
      }
      tm::group_barrier<FULL>(bar_id, zpad);
      TM_PHASE(1);  // stage in
      // phase B: one pass in input order when no position is hit twice, else lap by lap
      const bool onceA = dematch_single_pass(gA), onceB = hasB && dematch_single_pass(gB);
      // first transmissions with word-aligned geometry (the usual case) take the word-wise scatter, both codeblocks at once
      const bool wordA = onceA && dematch_wordwise(gA, rmA), wordB = onceB && dematch_wordwise(gB, rmB);
      const bool pairW = wordA && wordB && gA.Qm == gB.Qm;
      if (wordA) {
        dematch_scatter_words(gA, rmA, stA, pairW, gB, rmB, stB, (unsigned)t, (unsigned)zpad);
      } else if (onceA) {
        dematch_scatter_once(gA, rmA, stA, (unsigned)t, (unsigned)zpad);
      }
      TM_PHASE(11);  // scatter of codeblock A (thread 0's own share; both codeblocks in the word-wise case)
      if (wordB && !pairW) {
        dematch_scatter_words(gB, rmB, stB, false, gB, rmB, stB, (unsigned)t, (unsigned)zpad);
      } else if (onceB && !pairW) {
        dematch_scatter_once(gB, rmB, stB, (unsigned)t, (unsigned)zpad);
      }
      TM_PHASE(12);  // scatter of codeblock B
      const unsigned lapsA = onceA ? 0u : dematch_nof_laps(gA), lapsB = (!hasB || onceB) ? 0u : dematch_nof_laps(gB);
      for (unsigned lap = 0; lap < max(lapsA, lapsB); ++lap) {
        if (lap < lapsA) {
          dematch_scatter_lap(gA, rmA, stA, lap, (unsigned)t, (unsigned)zpad);
        }
        if (lap < lapsB) {
          dematch_scatter_lap(gB, rmB, stB, lap, (unsigned)t, (unsigned)zpad);
        }
        if (lap + 1 < max(lapsA, lapsB)) {
          tm::group_barrier<FULL>(bar_id, zpad);
        }
      }
      if (PREFETCH && t == 0) {
        flags[5] = claimed;
      }
      tm::group_barrier<FULL>(bar_id, zpad);
      TM_PHASE(2);  // scatter / combine
      // phase C: back to the arena, and on to the decoder
      dematch_stage_out(gA, stA, softA, (unsigned)t, (unsigned)zpad);
      if (hasB) {
        dematch_stage_out(gB, stB, softB, (unsigned)t, (unsigned)zpad);
      }
      // columns past the last position that can be non-zero are never visited by the decoder (its number of layers
      // follows from the last non-zero LLR): they are not converted
      const unsigned nzA = useA ? (dematch_is_fresh(gA) ? gA.zero_from : gA.N) : 0u;
      const unsigned nzB = useB ? (dematch_is_fresh(gB) ? gB.zero_from : gB.N) : 0u;
      const int n_hi = min(K + M, max(HRR, (int)((max(nzA, nzB) + Z - 1) / Z) + 2));
      if (tm::uniform((Z & 15) == 0 ? 1 : 0)) {
        // as in the decode-only load: a thread converts 16 consecutive lanes of one column of both codeblocks (two
        // 16-byte shared-memory loads, four 16-byte stores) instead of one lane of every column
        const unsigned V = (unsigned)Z >> 4, nvec = (unsigned)(n_hi - 2) * V;
        for (unsigned v = (unsigned)t; v < nvec; v += (unsigned)zpad) {
          const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
          const uint4 a = useA ? *reinterpret_cast<const uint4*>(stA + 16u * v) : zero4;
          const uint4 b = useB ? *reinterpret_cast<const uint4*>(stB + 16u * v) : zero4;
          lastA = ((a.x | a.y) | (a.z | a.w)) ? 16u * (v + 1) : lastA;
          lastB = ((b.x | b.y) | (b.z | b.w)) ? 16u * (v + 1) : lastB;
          const unsigned n = 2u + v / V, lane0 = 16u * (v - (n - 2u) * V);
          unsigned* dst = (n < (unsigned)HRR) ? soft_w + n * (unsigned)Z + lane0 : ext_soft_w + (n - HRR) * (unsigned)zpad + lane0;
          const unsigned aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 o;
            o.x = tm::convert_bytes(aw[q], bw[q], 0x0400);
            o.y = tm::convert_bytes(aw[q], bw[q], 0x0501);
            o.z = tm::convert_bytes(aw[q], bw[q], 0x0602);
            o.w = tm::convert_bytes(aw[q], bw[q], 0x0703);
            *reinterpret_cast<uint4*>(dst + 4 * q) = o;
          }
        }
      } else if (c.active) {
#pragma unroll 6
        for (int n = 2; n < n_hi; ++n) {
          const unsigned i = (unsigned)(n - 2) * Z + t;
          const int a = useA ? (int)stA[i] : 0;
          const int b = useB ? (int)stB[i] : 0;
          lastA = a ? i + 1 : lastA;
          lastB = b ? i + 1 : lastB;
          put(n, convert(a, b));
        }
      }
    }
    TM_PHASE(3);  // stage out + conversion (decode-only: the whole load)
    if (FUSED && !useA && !useB) {  // combined only (both CRCs already passed in an earlier transmission)
      tm::group_barrier<FULL>(bar_id, zpad);
      continue;
    }
    if (lastA) {
      atomicMax(const_cast<unsigned*>(&flags[0]), lastA);
    }
    if (lastB) {
      atomicMax(const_cast<unsigned*>(&flags[1]), lastB);
    }
    if (!FUSED && t == 0) {
      flags[5] = claimed_early;
    }
    tm::group_barrier<FULL>(bar_id, zpad);
    const unsigned inszA = flags[0], inszB = flags[1];  // one past the last non-zero LLR (ldpc_decoder_impl.cpp:86-94)
    bool doneA = !useA, doneB = !useB;
    CbResult resA = {0, 0, 0}, resB = {0, 0, 0};
    auto zero_shortcut = [&](const CbWork& w, CbResult& res) {
      if (!w.crc_poly) {  // output.one()
        uint8_t* out = p.out_base + w.out_offset;
        for (unsigned i = t; i < msg_bytes; i += zpad) {
          out[i] = (i == msg_bytes - 1 && (msg_len & 7)) ? (unsigned char)(0xff << (8 - (msg_len & 7))) : 0xff;
        }
      }
      res.flags = 1;
      if (t == 0) {
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
    // number of layers (ldpc_decoder_impl.cpp:105-114); the pair runs the larger count - surplus layers of the shorter
    // codeblock see an all-zero extension node, which leaves its high-rate nodes untouched
    auto layers_of = [&](unsigned insz) {
      unsigned cb_len = max(insz + 2u * Z, msg_len + 4u * Z);
      cb_len = ((cb_len + Z - 1) / Z) * Z;
      return (int)(cb_len / Z) - K;
    };
    const int nof_layers = tm::uniform(max(doneA ? 4 : layers_of(inszA), doneB ? 4 : layers_of(inszB)));
    // The schedule runs whole steps: when the layer limit falls between the two rows of a pair the second row runs too. Its
    // extension node lies behind the last non-zero LLR of both codeblocks: with that (thread-private) word at zero the
    // row's minimum is zero, every message to a high-rate node is zero and no soft value changes - the same argument that
    // lets the two codeblocks of a pair run the larger of their layer counts.
    const unsigned lsteps = p.layer_steps[nof_layers];
    const int nof_steps = (int)(lsteps & 0xffu), layers_run = (int)(lsteps >> 8);
    if (layers_run > nof_layers && c.active) {
      ext_soft_w[(nof_layers - 4) * zpad + t] = 0;
    }
    {
      // all messages of the layers that will run start at zero (biased bytes)
      unsigned init[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        init[i] = 0x80808080u;
      }
      // first message word behind the layers that run: a TMEM word, or tm_cap + an overflow word
      const int first_free = (layers_run < M) ? (int)(p.rowdesc[layers_run] >> 20) : (int)p.tm_cap + (int)p.ovf_words;
      const int words = min(first_free, (int)p.tm_words);
      for (int w = 0; w < words; w += 8) {
        tm::tm_st8(c.tm_addr + w, init);
      }
      for (int w = 0; w < first_free - (int)p.tm_cap; ++w) {
        tm::sts32(c.ovf + (unsigned)w * c.ovf_stride, 0x80808080u);
      }
      tm::tm_wait_st();
    }
    const unsigned pair_next = PREFETCH ? (unsigned)tm::uniform((int)flags[5]) : 0x7fffffffu;
    if (PREFETCH && 2 * pair_next < p.nof_cbs) {
      // inputs of the next pair -> L2, one 128-byte line per thread and step; the previous soft bits only when they
      // will be read (dematch_is_fresh covers the usual new-data case)
      const unsigned cbn = min(2u, p.nof_cbs - 2 * pair_next);
      for (unsigned h = 0; h < cbn; ++h) {
        const CbWork wn = p.work[2 * pair_next + h];
        if constexpr (!FUSED) {
          const char* in = reinterpret_cast<const char*>(p.llr_base + wn.in_offset);
          for (unsigned o = 128u * t; o < (wn.skip_decode ? 0u : wn.in_length); o += 128u * zpad) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(in + o));
          }
          continue;
        }
        const char* rm = reinterpret_cast<const char*>(p.rm_base + wn.rm_offset);
        for (unsigned o = 128u * t; o < wn.rm_length; o += 128u * zpad) {
          asm volatile("prefetch.global.L2 [%0];" ::"l"(rm + o));
        }
        if (!wn.new_data) {
          const char* in = reinterpret_cast<const char*>(p.llr_base + wn.in_offset);
          for (unsigned o = 128u * t; o < (unsigned)((K + M - 2) * Z); o += 128u * zpad) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(in + o));
          }
        }
      }
    }

    TM_PHASE(4);  // input size, layers, TMEM reset, prefetch of the next pair
    int it = 0;
    while (tm::uniform((it < p.max_iterations && !(doneA && doneB)) ? 1 : 0)) {
      // the barrier that orders a row after its predecessor is inside the bodies
      tm::row_arrive(c.row_bar, c.t4);  // every row waits for its predecessor's arrivals: these stand in for the first one's
      uint4 next = p.steps[0];
      for (int s = 0; s < nof_steps;) {
        const uint4 st = next;
        next = p.steps[s + 1];  // the descriptor of the step after this one travels under this step's arithmetic
        if constexpr (FUSED) {
          tm::step_dispatch_runs<BG, SCALE, MODE>(c, p, bar_id, zpad, st, next, s, nof_steps);
        } else {
          tm::step_dispatch<BG, SCALE, MODE>(c, p, bar_id, zpad, st);
          ++s;
        }
      }
      tm::row_wait(c.row_bar, c.row_parity);  // the last row's arrivals
      tm::tm_wait_st();
      TM_PHASE(5);  // row passes of one iteration
      ++it;
      const bool last = (it == p.max_iterations);
      const bool chkA = !doneA && ((wA.crc_poly && wA.early_stop) || last);
      const bool chkB = !doneB && ((wB.crc_poly && wB.early_stop) || last);
      if (chkA || chkB) {
        // CRC of the hard decisions (ldpc_decoder_impl.cpp:125-133) without serialising on one warp: the remainder of
        // the message polynomial is linear in its bits, so every thread XORs x^(n-1-i) mod P (table crc_pow) for the set
        // bits it owns and the group XOR-reduces the partial remainders. remainder == 0 <=> crc->calculate(...) == 0
        // (x is invertible modulo the generator polynomials, all of which have a constant term).
        const bool crcA = chkA && wA.crc_poly, crcB = chkB && wB.crc_poly;
        const unsigned acc_a = 8 + 2 * (it & 1), acc_b = acc_a + 1;
        if (crcA || crcB) {
          const unsigned nbA = msg_len - wA.nof_filler_bits, nbB = msg_len - wB.nof_filler_bits;
          const unsigned* tabA = g_crc_pow[crcA ? wA.crc_poly - 1 : 0];
          const unsigned* tabB = g_crc_pow[crcB ? wB.crc_poly - 1 : 0];
          unsigned ra = 0, rb = 0;
          constexpr int CU = (K == 22) ? 11 : 10;
          // Usual case: both codeblocks (or the only one still being checked) use the same polynomial and message length,
          // so one table entry serves both halves of the word; the hard bits of both halves are formed with three
          // integer operations and turned into AND masks. (The general form below costs about twice the instructions and
          // was 27 % of the fused kernel's instructions on full-slot transport blocks.)
          const bool one_table = !crcA || !crcB || (wA.crc_poly == wB.crc_poly && nbA == nbB);
          if (one_table) {
            if (c.active) {
              const unsigned nb = crcA ? nbA : nbB;
              const unsigned* top = (crcA ? tabA : tabB) + (nb - 1 - (unsigned)t);  // entry of this lane in column 0
#pragma unroll 1
              for (int n0 = 0; n0 < K; n0 += CU) {
                unsigned v[CU], e[CU];
#pragma unroll
                for (int u = 0; u < CU; ++u) {
                  v[u] = soft_w[(unsigned)(n0 + u) * Z + t];
                }
#pragma unroll
                for (int u = 0; u < CU; ++u) {
                  const unsigned col = (unsigned)(n0 + u) * Z;
                  e[u] = (col + t < nb) ? __ldg(top - col) : 0u;
                }
#pragma unroll
                for (int u = 0; u < CU; ++u) {
                  // get_hard_bits (ldpc_decoder_generic.cpp:132-138): bit = soft <= 0, i.e. sign bit set or magnitude zero;
                  // adding 0x7fff to the 15 magnitude bits carries into bit 15 exactly when they are not all zero
                  const unsigned nz = (v[u] & 0x7fff7fffu) + 0x7fff7fffu;
                  const unsigned hb = (v[u] | ~nz) & 0x80008000u;
                  ra ^= e[u] & (unsigned)((int)(hb << 16) >> 31);
                  rb ^= e[u] & (unsigned)((int)hb >> 31);
                }
              }
            }
          } else if (c.active) {
            // all table loads are issued before the first is consumed (no branch on the bit: the entry is masked)
#pragma unroll 1
            for (int n0 = 0; n0 < K; n0 += CU) {
              unsigned v[CU], ea[CU], eb[CU];
#pragma unroll
              for (int u = 0; u < CU; ++u) {
                v[u] = soft_w[(unsigned)(n0 + u) * Z + t];
              }
#pragma unroll
              for (int u = 0; u < CU; ++u) {
                const unsigned i = (unsigned)(n0 + u) * Z + t;
                ea[u] = __ldg(tabA + (i < nbA ? nbA - 1 - i : 0));
                eb[u] = __ldg(tabB + (i < nbB ? nbB - 1 - i : 0));
              }
#pragma unroll
              for (int u = 0; u < CU; ++u) {
                const unsigned i = (unsigned)(n0 + u) * Z + t;
                // get_hard_bits (ldpc_decoder_generic.cpp:132-138): bit = soft <= 0
                const bool ba = (v[u] & 0x7fffu) == 0 || (v[u] & 0x8000u);
                const bool bb = (v[u] & 0x7fff0000u) == 0 || (v[u] & 0x80000000u);
                ra ^= (crcA && ba && i < nbA) ? ea[u] : 0u;
                rb ^= (crcB && bb && i < nbB) ? eb[u] : 0u;
              }
            }
          }
          TM_PHASE(6);  // check: sweep over the hard bits
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            ra ^= __shfl_xor_sync(0xffffffffu, ra, o);
            rb ^= __shfl_xor_sync(0xffffffffu, rb, o);
          }
          if ((t & 31) == 0) {
            if (ra) {
              atomicXor(const_cast<unsigned*>(&flags[acc_a]), ra);
            }
            if (rb) {
              atomicXor(const_cast<unsigned*>(&flags[acc_b]), rb);
            }
          }
          if (t == 0) {  // the accumulators of the next iteration (nobody touches them before the barriers below)
            flags[acc_a ^ 2] = 0, flags[acc_b ^ 2] = 0;
          }
        }
        tm::group_barrier<FULL>(bar_id, zpad);
        TM_PHASE(8);  // check: reduction + barrier
        const bool okA = crcA && flags[acc_a] == 0, okB = crcB && flags[acc_b] == 0;
        // the message bytes are only formed for a codeblock that finishes now
        if (((chkA && (okA || last)) || (chkB && (okB || last))) && (FULL || (Z & 31) == 0)) {
          // Z a multiple of 32: lane l of warp w holds message bit n Z + 32 w + l of column n, so one ballot per column and
          // codeblock is four message bytes (bit-reversed: the first bit of a byte is its MSB), stored by one lane. (The
          // byte-wise loop below reads the soft values with a stride of 8 words - an 8-way bank conflict per load - and
          // was 7 % of the fused kernel's time on full-slot transport blocks.)
          // Columns go in batches: all loads of a batch are in flight together, and lane u of the warp keeps and stores the
          // two words of the batch's column u (one store instruction per codeblock and batch).
          const unsigned lane = (unsigned)t & 31u;
          constexpr int CB = (K == 22) ? 11 : 10;
#pragma unroll 1
          for (int n0 = 0; n0 < K; n0 += CB) {
            unsigned v[CB];
#pragma unroll
            for (int u = 0; u < CB; ++u) {
              v[u] = soft_w[(unsigned)(n0 + u) * Z + t];
            }
            unsigned wa = 0, wb = 0;
#pragma unroll
            for (int u = 0; u < CB; ++u) {
              // get_hard_bits as in the CRC sweep: sign bit set or magnitude zero
              const unsigned nz = (v[u] & 0x7fff7fffu) + 0x7fff7fffu;
              const unsigned hb = (v[u] | ~nz) & 0x80008000u;
              const unsigned ba = __ballot_sync(0xffffffffu, (hb & 0x8000u) != 0);
              const unsigned bb = __ballot_sync(0xffffffffu, (hb & 0x80000000u) != 0);
              wa = (lane == (unsigned)u) ? ba : wa;
              wb = (lane == (unsigned)u) ? bb : wb;
            }
            if (lane < (unsigned)CB) {
              // byte offset of the four bytes of (column n0 + lane, this warp's 32 lanes): a multiple of 4
              const unsigned o = ((unsigned)(n0 + (int)lane) * Z + ((unsigned)t - lane)) >> 3;
              *reinterpret_cast<unsigned*>(hard + o) = __byte_perm(__brev(wa), 0u, 0x0123);
              *reinterpret_cast<unsigned*>(hard + msg_bytes16 + o) = __byte_perm(__brev(wb), 0u, 0x0123);
            }
          }
          tm::group_barrier<FULL>(bar_id, zpad);  // the bytes are read by other threads than the ones that stored them
        } else if ((chkA && (okA || last)) || (chkB && (okB || last))) {
          for (unsigned i = t; i < msg_bytes; i += zpad) {
            unsigned ba = 0, bb = 0;
#pragma unroll
            for (unsigned k = 0; k < 8; ++k) {
              const unsigned b = 8 * i + k;
              if (b < msg_len) {
                const unsigned v = soft_w[b];
                ba |= (((v & 0x7fffu) == 0 || (v & 0x8000u)) ? 0x80u : 0u) >> k;
                bb |= (((v & 0x7fff0000u) == 0 || (v & 0x80000000u)) ? 0x80u : 0u) >> k;
              }
            }
            hard[i] = (unsigned char)ba;
            hard[msg_bytes16 + i] = (unsigned char)bb;
          }
        }
        TM_PHASE(9);  // finish: message bytes
        auto finish = [&](const CbWork& w, CbResult& res, bool& done, const unsigned char* bits, bool ok) {
          if (ok || last) {
            uint8_t* out = p.out_base + w.out_offset;
            // bit_buffer::insert keeps the bits of the last byte beyond the message as the caller left them
            for (unsigned i = t; i < msg_bytes; i += zpad) {
              unsigned char v = bits[i];
              if (i == msg_bytes - 1 && (msg_len & 7)) {
                const unsigned char keep = (unsigned char)(0xffU >> (msg_len & 7));
                v = (unsigned char)((v & ~keep) | (out[i] & keep));
              }
              out[i] = v;
            }
            res.iterations = (unsigned short)it;
            res.crc_ok = ok ? 1 : 0;
            if (t == 0) {
              p.results[w.result_index] = res;
            }
            done = true;
          }
        };
        if (chkA) {
          finish(wA, resA, doneA, hard, okA);
        }
        if (chkB) {
          finish(wB, resB, doneB, hard + msg_bytes16, okB);
        }
        tm::group_barrier<FULL>(bar_id, zpad);
      } else if (t == 0) {
        // No check after this iteration (one half decodes without early stop, the other has finished): the accumulators
        // the next iteration would use may still hold the remainder of an earlier check of the same parity.
        flags[8 + 2 * ((it + 1) & 1)] = 0, flags[9 + 2 * ((it + 1) & 1)] = 0;
      }
      TM_PHASE(10);  // finish: output + barrier
    }
    tm::group_barrier<FULL>(bar_id, zpad);  // flags / hard bits are reused by the next pair
    TM_PHASE(7);
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ---- host side -------------------------------------------------------------------------------------------------------
static inline int tm_scale_variant(const unsigned char* lut)
{
  bool gen = true, avx = true;
  for (int m = 0; m <= LLR_MAX; ++m) {
    const int r5 = (2 * m + 5) / 10;  // round(m / 5), never at a tie
    gen = gen && (lut[m] == m - r5);
    avx = avx && (lut[m] == m - m / 5 - (m > 0 ? 1 : 0));
  }
  // any other table with lut[m] <= m (a scaling factor below 1): table look-up inside the same kernel
  bool shrinking = true;
  for (int m = 0; m <= LLR_MAX; ++m) {
    shrinking = shrinking && lut[m] <= m;
  }
  return gen ? TM_SCALE_GENERIC_08 : (avx ? TM_SCALE_AVX_08 : (shrinking ? TM_SCALE_LUT : TM_SCALE_NONE));
}

template <int SCALE>
__global__ void tm_scale_selfcheck_kernel(unsigned char* out)
{
  const int m = threadIdx.x;
  if (m <= LLR_MAX) {
    const unsigned q = tm::scale_q<SCALE>(tm::hc(tm::TM_S * (float)m));
    out[m] = (unsigned char)(__half2float(__low2half(tm::h2(q))) * (1.f / tm::TM_S) + 0.25f);
  }
}

// Runs the fp16 closed form for every magnitude on the device and compares with the host table. 0 = identical.
static inline int tm_scale_selfcheck(int variant, const unsigned char* lut, cudaStream_t stream)
{
  if (variant == TM_SCALE_LUT) {
    return 0;  // a table look-up of the very table: nothing to prove
  }
  unsigned char* d = nullptr;
  unsigned char h[LLR_MAX + 1];
  if (cudaMalloc(&d, LLR_MAX + 1) != cudaSuccess) {
    return -1;
  }
  if (variant == TM_SCALE_GENERIC_08) {
    tm_scale_selfcheck_kernel<TM_SCALE_GENERIC_08><<<1, 128, 0, stream>>>(d);
  } else {
    tm_scale_selfcheck_kernel<TM_SCALE_AVX_08><<<1, 128, 0, stream>>>(d);
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

// Fills the row / edge tables of a (base graph, lifting size) group. Returns false when a row shape has no kernel case.
// tm_cap: message words a thread owns in TMEM (a multiple of 8 is used); rows that do not fit get shared-memory words.
static inline bool tm_build_tables(int bg, int Z, int lifting_set, TmemParams& p, unsigned tm_cap = TM_COLS_PER_WARP)
{
  const int M = (bg == 2) ? LDPC_BG2_M : LDPC_BG1_M, K = (bg == 2) ? 10 : 22, hrr = K + 4;
  const uint16_t* row_ptr = (bg == 2) ? LDPC_BG2_ROW_PTR : LDPC_BG1_ROW_PTR;
  const uint8_t* col = (bg == 2) ? LDPC_BG2_COL : LDPC_BG1_COL;
  const uint16_t* shift = (bg == 2) ? LDPC_BG2_SHIFT[lifting_set] : LDPC_BG1_SHIFT[lifting_set];
  unsigned ne = 0, tmw = 0, ovf = 0;
  bool spilled = false;  // once a row went to shared memory all later ones do (the reset loop relies on the order)
  p.tm_cap = (unsigned short)tm_cap;
  for (int r = 0; r != M; ++r) {
    unsigned d = 0, ext = 0;
    const unsigned row_edge = ne;
    for (int e = row_ptr[r]; e != row_ptr[r + 1]; ++e) {
      if (col[e] < hrr) {
        p.edge[ne].s4 = 4u * (shift[e] % (unsigned)Z);
        p.edge[ne].cb = 4u * (unsigned)Z * col[e];
        ++ne, ++d;
      } else {
        // the row's private extension node: last edge, identity circulant
        if (col[e] != K + r || shift[e] % (unsigned)Z != 0 || e + 1 != row_ptr[r + 1] || r < 4) {
          return false;
        }
        ext = 1;
      }
    }
    const bool ok = ext ? (d >= 2 && d <= 9) : (d == 8 || d == 10 || d == 19);
    if (!ok) {
      return false;
    }
    const unsigned w = (d + 1) / 2;
    unsigned row_tm;
    spilled = spilled || tmw + w > (tm_cap & ~7u);
    if (spilled) {
      row_tm = tm_cap + ovf;
      ovf += w;
    } else {
      row_tm = tmw;
      tmw += w;
    }
    p.rowdesc[r] = (2 * d + ext) | (row_edge << 8) | (row_tm << 20);
  }
  p.tm_words = (unsigned short)tmw;
  p.ovf_words = (unsigned short)ovf;
  // pairs of consecutive rows (both with an extension node) that share no high-rate column, chosen greedily in row order;
  // A is the row with the larger high-rate degree (the order inside a pair does not matter)
  auto step_words = [&](int r, unsigned& y) { y = ((p.rowdesc[r] >> 8) & 0xfffu) | ((p.rowdesc[r] >> 20) << 16); };
  const unsigned zpad4 = 4u * (((unsigned)Z + 31u) / 32u) * 32u;
  auto ext_offset = [&](int r) { return r >= 4 ? (unsigned)(r - 4) * zpad4 : 0u; };  // < 2^16: 41 * 1536
  unsigned ns = 0;
  p.layer_steps[0] = 0;
  for (int r = 0; r < M;) {
    bool disjoint = r >= 4 && r + 1 < M;
    for (int a = row_ptr[r]; disjoint && a != row_ptr[r + 1]; ++a) {
      for (int b = row_ptr[r + 1]; b != row_ptr[r + 2] && disjoint; ++b) {
        disjoint = !(col[a] == col[b] && col[a] < hrr);
      }
    }
    const int d0 = (p.rowdesc[r] & 0xffu) >> 1, d1 = (r + 1 < M) ? (int)((p.rowdesc[r + 1] & 0xffu) >> 1) : 0;
    const int da = d0 >= d1 ? d0 : d1, db = d0 >= d1 ? d1 : d0;
    uint4& st = p.steps[ns];
    if (disjoint && tm::tm_pair_used(bg, da, db)) {
      const int rA = d0 >= d1 ? r : r + 1, rB = d0 >= d1 ? r + 1 : r;
      st.x = (unsigned)(16 * da + db) | (2u << 8);
      step_words(rA, st.y);
      step_words(rB, st.z);
      st.w = ext_offset(rA) | (ext_offset(rB) << 16);
      p.layer_steps[r + 1] = p.layer_steps[r + 2] = (unsigned short)((ns + 1) | ((unsigned)(r + 2) << 8));
      r += 2;
    } else {
      if (!tm::tm_row_used(bg, d0, r >= 4)) {
        return false;  // a row shape the kernels of this base graph were not compiled for
      }
      st.x = (p.rowdesc[r] & 0xffu) | (1u << 8);
      step_words(r, st.y);
      st.z = 0, st.w = ext_offset(r);
      p.layer_steps[r + 1] = (unsigned short)((ns + 1) | ((unsigned)(r + 1) << 8));
      r += 1;
    }
    ++ns;
  }
  for (unsigned i = ns; i <= (unsigned)LDPC_BG1_M; ++i) {
    p.steps[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  return ((tmw + 7) & ~7u) <= tm_cap;
}

// Builds that only want the kernels or the table builder (a single instantiation for a SASS listing, the schedule check of
// tests/test_tm_schedule.py) define LDPC_TM_DEVICE_ONLY and stop here: the launchers below instantiate every kernel.
#ifndef LDPC_TM_DEVICE_ONLY
constexpr size_t TM_SMEM_LIMIT = 232448;      // 227 KB opt-in maximum per block on sm_100
constexpr size_t TM_SMEM_MIN = 120 * 1024;    // more than half an SM: one block per SM owns all 512 TMEM columns

// cudaFuncSetAttribute applies to the CURRENT device only: the opt-in is tracked per device (a process may hold handles on
// several GPUs, and several host threads may launch at once).
constexpr int TM_MAX_DEVICES = 64;

template <int BG, int SCALE, bool FUSED, int MODE>
static inline int tm_launch_full(const TmemParams& p, size_t smem, unsigned grid, int device, cudaStream_t stream)
{
  auto kernel = ldpc_decode_tmem_kernel<BG, SCALE, FUSED, MODE>;
  static std::atomic<bool> attr_set[TM_MAX_DEVICES];
  if (device < 0 || device >= TM_MAX_DEVICES || !attr_set[device].load(std::memory_order_acquire)) {
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TM_SMEM_LIMIT);
    if (e != cudaSuccess) {
      std::fprintf(stderr, "ldpc_cuda: cudaFuncSetAttribute failed: %s\n", cudaGetErrorString(e));
      return -1;
    }
    if (device >= 0 && device < TM_MAX_DEVICES) {
      attr_set[device].store(true, std::memory_order_release);
    }
  }
  kernel<<<grid, p.groups * p.zpad, smem, stream>>>(p);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    std::fprintf(stderr, "ldpc_cuda: packed decode launch failed: %s\n", cudaGetErrorString(e));
    return -1;
  }
  return 0;
}

template <int BG, int SCALE, bool FUSED>
static inline int tm_launch_variant(const TmemParams& p, size_t smem, unsigned grid, int device, cudaStream_t stream)
{
  if (p.Z == MAX_Z && p.groups == 1) {
    return tm_launch_full<BG, SCALE, FUSED, TM_MODE_FULL>(p, smem, grid, device, stream);
  }
  if constexpr (!FUSED) {
    if (p.groups * p.zpad > TM_THREADS) {
      return tm_launch_full<BG, SCALE, false, TM_MODE_WIDE>(p, smem, grid, device, stream);
    }
  }
  return tm_launch_full<BG, SCALE, FUSED, TM_MODE_GENERAL>(p, smem, grid, device, stream);
}

template <int BG, int SCALE>
static inline int tm_launch_fused_or_not(bool fused, const TmemParams& p, size_t smem, unsigned grid, int device,
                                         cudaStream_t stream)
{
  return fused ? tm_launch_variant<BG, SCALE, true>(p, smem, grid, device, stream)
               : tm_launch_variant<BG, SCALE, false>(p, smem, grid, device, stream);
}

// Returns 0 when launched, > 0 when the configuration is not covered (caller uses the general kernel), < 0 on CUDA error.
// d_rm != nullptr selects the fused kernel: d_llr is then the soft-buffer arena, combined in place with the rate-matched
// LLRs at d_rm before decoding.
static inline int launch_tmem_decode(const GroupParams& gp, int scale_variant, const CbWork* d_work, unsigned n,
                                     const int8_t* d_llr, const int8_t* d_rm, uint8_t* d_out, CbResult* d_res,
                                     unsigned* d_pair_counter, int sm_count, int device, cudaStream_t stream)
{
  if (scale_variant == TM_SCALE_NONE || d_pair_counter == nullptr) {
    return 1;
  }
  TmemParams p;
  std::memset(&p, 0, sizeof(p));
  p.work = d_work, p.llr_base = d_llr, p.out_base = d_out, p.results = d_res, p.pair_counter = d_pair_counter;
  p.rm_base = d_rm, p.soft_rw = const_cast<int8_t*>(d_llr);
  const bool fused = d_rm != nullptr;
  p.nof_cbs = n, p.Z = gp.Z, p.Z4 = 4 * gp.Z, p.max_iterations = gp.max_iterations;
  p.zpad = ((gp.Z + 31) / 32) * 32;
  // Decode-only blocks run up to 512 threads (the kernel fits 128 registers): floor(512 / roundup32(Z)) groups, e.g. 4
  // instead of 3 at Z = 128 and 2 instead of 1 for Z in (192, 256]. Above 12 warps a warp owns 128 TMEM columns instead
  // of 170, and the message words of the last rows go to shared memory. The fused kernel needs 168 registers: 384.
  // (Single-warp groups, Z <= 32, lose 4 - 12 % with 15 instead of 12 groups: they stay at 384.)
  const int max_threads = (fused || gp.Z == MAX_Z || p.zpad < 64) ? TM_THREADS : TM_THREADS_WIDE;
  p.groups = std::min(max_threads / p.zpad, 15);  // named barriers 1 .. 15
  const TmLayout lay = tm_layout(gp.bg, gp.Z, p.zpad, fused);
  p.group_stride = lay.total;
  const unsigned n_pairs = (n + 1) / 2;
  if ((unsigned)p.groups > n_pairs) {
    p.groups = (int)n_pairs;
  }
  size_t smem = 0;
  for (;; --p.groups) {
    const unsigned threads = (unsigned)(p.groups * p.zpad);
    const unsigned cap = threads > (unsigned)TM_THREADS ? TM_COLS_PER_WARP_WIDE : TM_COLS_PER_WARP;
    if (!tm_build_tables(gp.bg, gp.Z, gp.lifting_set, p, cap)) {
      return 1;
    }
    p.ovf_stride = 4u * threads;
    smem = TM_CRC_TABLE_BYTES + (size_t)p.groups * lay.total + (size_t)p.ovf_words * p.ovf_stride;
    if (smem <= TM_SMEM_LIMIT || p.groups == 1) {
      break;
    }
  }
  if (smem > TM_SMEM_LIMIT) {
    return 1;
  }
  smem = std::max(smem, TM_SMEM_MIN);
  const unsigned grid = std::min<unsigned>((n_pairs + p.groups - 1) / p.groups, (unsigned)sm_count);
  if (cudaMemsetAsync(d_pair_counter, 0, sizeof(unsigned), stream) != cudaSuccess) {
    return -1;
  }
  if (scale_variant == TM_SCALE_LUT) {
    std::memcpy(p.scale_lut, gp.scale_lut, LLR_MAX + 1);
    return gp.bg == 1 ? tm_launch_fused_or_not<1, TM_SCALE_LUT>(fused, p, smem, grid, device, stream)
                      : tm_launch_fused_or_not<2, TM_SCALE_LUT>(fused, p, smem, grid, device, stream);
  }
  const bool gen = scale_variant == TM_SCALE_GENERIC_08;
  if (gp.bg == 1) {
    return gen ? tm_launch_fused_or_not<1, TM_SCALE_GENERIC_08>(fused, p, smem, grid, device, stream)
               : tm_launch_fused_or_not<1, TM_SCALE_AVX_08>(fused, p, smem, grid, device, stream);
  }
  return gen ? tm_launch_fused_or_not<2, TM_SCALE_GENERIC_08>(fused, p, smem, grid, device, stream)
             : tm_launch_fused_or_not<2, TM_SCALE_AVX_08>(fused, p, smem, grid, device, stream);
}
#endif  // LDPC_TM_DEVICE_ONLY

}  // namespace ldpc_b200
