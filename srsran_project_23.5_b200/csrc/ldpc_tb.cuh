// Transport-block assembly and TB CRC24A on the device.
//
// Replaces the part of pusch_decoder_impl::decode that follows the per-codeblock loop
// (lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:156-222): the nof_data_bits = K - crc - fillers leading bits
// of every decoded codeblock message are concatenated MSB first into the transport block (+ its 24 CRC bits when the
// block was segmented, :196-200), the CRC24A remainder of the result is formed (:216, crc_calculator::calculate_byte over
// whole bytes, init 0, no final xor) and tb_crc_ok is set as :205-222 does: one codeblock -> its own CRC flag; several
// -> every codeblock flag set AND remainder == 0.
//
// Work split: one thread per 4 OUTPUT bytes (gather: conflict-free stores; a byte takes its bits from at most two
// codeblocks). The CRC is linear in the message bits: a thread XORs x^e mod P for the set bits of its bytes (table
// g_crc_pow of the decoder kernel), a block reduces its 1024 bytes, multiplies the partial remainder by
// x^(bits behind the block) mod P - two table look-ups and two 24-bit carry-less multiplications - and stores it; a
// second small kernel (one block per transport block) XORs the partial remainders, counts the failed codeblocks and
// writes the public result. No atomics, no fences (a first version that accumulated with same-address atomics and a
// last-block-done counter took 612 us for 128 full-slot blocks; see profiles/). HBM traffic = the bits themselves, once.
#pragma once

#include "ldpc_device.cuh"

namespace ldpc_b200 {

struct TbWork {
  unsigned long long tb_offset;  // bytes from the transport-block output base
  unsigned first_cb;             // index of the block's first entry in the TbCb array
  unsigned nof_cbs;              // C
  unsigned data_bits;            // bits taken from each codeblock message: K - nof_crc_bits - nof_filler_bits
  unsigned nbytes;               // bytes written: (tbs + (C > 1 ? 24 : 0)) / 8
  unsigned chunk_base;           // index of its first per-chunk partial remainder
  unsigned reserved;
};

struct TbCb {
  unsigned long long src_offset;  // bytes of the codeblock's message from the message base
  unsigned already_ok;            // CRC flag set by an earlier transmission (skip_decode)
  unsigned result_index;          // index of its CbResult
};

struct TbResult {  // == ldpc_cuda_tb_result
  unsigned crc_remainder;
  unsigned short nof_cb_failed;
  unsigned char tb_crc_ok;
  unsigned char reserved;
};

constexpr int TB_THREADS = 256;
constexpr int TB_BYTES_PER_THREAD = 4;
constexpr int TB_CHUNK = TB_THREADS * TB_BYTES_PER_THREAD;  // output bytes per block; 8 * TB_CHUNK <= TM_CRC_POW_LEN
constexpr unsigned TB_POW_CHUNKS = 1024;                   // x^(8 TB_CHUNK k) mod P for k < 1024: blocks up to 1 MB
__device__ unsigned g_tb_chunk_pow[TB_POW_CHUNKS];
static_assert(8 * TB_CHUNK + 24 <= (int)TM_CRC_POW_LEN, "in-chunk exponents come from g_crc_pow");
constexpr unsigned TB_CRC24A_POLY = 0x1864CFBu;  // crc_calculator.h:33-39
constexpr unsigned TB_CRC24A_CODE = 2;           // ldpc_cuda_crc / g_crc_pow row TB_CRC24A_CODE - 1

// a(x) * b(x) mod P for 24-bit remainders.
__device__ __forceinline__ unsigned tb_mulmod24(unsigned a, unsigned b)
{
  unsigned r = 0;
#pragma unroll 1
  for (int i = 23; i >= 0; --i) {
    r <<= 1;
    r ^= (r & 0x1000000u) ? TB_CRC24A_POLY : 0u;
    r ^= ((b >> i) & 1u) ? a : 0u;
  }
  return r;
}

// x^e mod P by square and multiply.
__device__ __forceinline__ unsigned tb_xpow24(unsigned e)
{
  unsigned result = 1, base = 2;
  while (e) {
    if (e & 1u) {
      result = tb_mulmod24(result, base);
    }
    base = tb_mulmod24(base, base);
    e >>= 1;
  }
  return result;
}

// grid (chunks of TB_CHUNK output bytes, transport blocks). partial: one weighted remainder per chunk.
__global__ void __launch_bounds__(TB_THREADS) ldpc_tb_assemble_kernel(const TbWork* __restrict__ tbs,
                                                                      const TbCb* __restrict__ cbs,
                                                                      const uint8_t* __restrict__ msgs,
                                                                      uint8_t* __restrict__ tb_out,
                                                                      unsigned* __restrict__ partial)
{
  const unsigned tb = blockIdx.y;
  const TbWork w = tbs[tb];
  const unsigned j0 = blockIdx.x * TB_CHUNK;
  if (j0 >= w.nbytes) {
    return;
  }
  const unsigned L = min((unsigned)TB_CHUNK, w.nbytes - j0);
  const unsigned i = threadIdx.x;
  __shared__ unsigned red[TB_THREADS / 32];
  unsigned r = 0;
  // codeblock and bit offset of this thread's first byte; its next bytes lie 8 TB_THREADS bits further each (one division
  // per thread, then a step that wraps at most once for codeblocks of 2048 data bits or more)
  unsigned cb = (8u * (j0 + i)) / w.data_bits;
  unsigned o = 8u * (j0 + i) - cb * w.data_bits;
#pragma unroll
  for (int q = 0; q < TB_BYTES_PER_THREAD; ++q) {
    const unsigned ib = q * TB_THREADS + i;  // byte of the chunk: consecutive lanes, consecutive bytes
    if (q != 0) {
      o += 8u * TB_THREADS;
      while (o >= w.data_bits) {
        o -= w.data_bits;
        ++cb;
      }
    }
    if (ib < L) {
      const uint8_t* s = msgs + cbs[w.first_cb + cb].src_offset;
      const unsigned two = ((unsigned)s[o >> 3] << 8) | (unsigned)s[(o >> 3) + 1];  // >= 16 message bits follow the data
      unsigned v = (two >> (8 - (o & 7u))) & 0xffu;
      const unsigned avail = w.data_bits - o;
      if (avail < 8) {  // the byte straddles two codeblocks
        v &= 0xff00u >> avail;
        if (cb + 1 < w.nof_cbs) {
          v |= (unsigned)msgs[cbs[w.first_cb + cb + 1].src_offset] >> avail;
        }
      }
      tb_out[w.tb_offset + j0 + ib] = (uint8_t)v;
      // bit b of the byte (LSB = 0) carries x^(8 (L - 1 - ib) + b) inside the chunk: 8 consecutive table entries, two
      // 16-byte loads (the first version issued 8 scalar loads per byte with 128-byte lane strides and was bound by the
      // L1 data pipe at 97 %: profiles/r01_ncu_tb_assemble_summary.txt)
      const uint4* pw = reinterpret_cast<const uint4*>(g_crc_pow[TB_CRC24A_CODE - 1] + 8u * (L - 1 - ib));
      const uint4 e0 = __ldg(pw), e1 = __ldg(pw + 1);
      r ^= ((v & 1u) ? e0.x : 0u) ^ ((v & 2u) ? e0.y : 0u) ^ ((v & 4u) ? e0.z : 0u) ^ ((v & 8u) ? e0.w : 0u);
      r ^= ((v & 16u) ? e1.x : 0u) ^ ((v & 32u) ? e1.y : 0u) ^ ((v & 64u) ? e1.z : 0u) ^ ((v & 128u) ? e1.w : 0u);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    r ^= __shfl_xor_sync(0xffffffffu, r, o);
  }
  if ((i & 31u) == 0) {
    red[i >> 5] = r;
  }
  __syncthreads();
  const unsigned nof_chunks = (w.nbytes + TB_CHUNK - 1) / TB_CHUNK;
  if (i == 0) {
    unsigned acc = 0;
#pragma unroll
    for (int k = 0; k < TB_THREADS / 32; ++k) {
      acc ^= red[k];
    }
    // x^(transport-block bits behind this chunk + 24): crc_calculator's checksum is the remainder of m(x) x^24, so that
    // crc_remainder equals crc_calculator::calculate_byte of the assembled bytes and not only its zero test. The last
    // chunk has 24 bits behind it; chunk c < last has 8 TB_CHUNK (last - c - 1) + 8 (bytes of the last chunk) + 24.
    if (acc != 0) {
      const unsigned last = nof_chunks - 1, c = blockIdx.x;
      const unsigned* pw1 = g_crc_pow[TB_CRC24A_CODE - 1];
      if (c == last) {
        acc = tb_mulmod24(acc, pw1[24]);
      } else {
        const unsigned tail = 8u * (w.nbytes - last * TB_CHUNK) + 24u, k = last - c - 1;
        acc = tb_mulmod24(acc, pw1[tail]);
        acc = tb_mulmod24(acc, k < TB_POW_CHUNKS ? g_tb_chunk_pow[k] : tb_xpow24(8u * TB_CHUNK * k));
      }
    }
    partial[w.chunk_base + blockIdx.x] = acc;
  }
}

// One block per transport block: remainder = XOR of its chunks' weighted remainders; codeblock flags
// (pusch_decoder_impl.cpp:188-195) and the verdict (:205-222).
__global__ void __launch_bounds__(128) ldpc_tb_finish_kernel(const TbWork* __restrict__ tbs, const TbCb* __restrict__ cbs,
                                                             const CbResult* __restrict__ results,
                                                             const unsigned* __restrict__ partial,
                                                             TbResult* __restrict__ tb_results)
{
  const unsigned tb = blockIdx.x, i = threadIdx.x;
  const TbWork w = tbs[tb];
  const unsigned nof_chunks = (w.nbytes + TB_CHUNK - 1) / TB_CHUNK;
  unsigned rem = 0, bad = 0;
  for (unsigned k = i; k < nof_chunks; k += 128) {
    rem ^= partial[w.chunk_base + k];
  }
  for (unsigned k = i; k < w.nof_cbs; k += 128) {
    const TbCb c = cbs[w.first_cb + k];
    bad += (c.already_ok || results[c.result_index].crc_ok) ? 0u : 1u;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    rem ^= __shfl_xor_sync(0xffffffffu, rem, o);
    bad += __shfl_xor_sync(0xffffffffu, bad, o);
  }
  __shared__ unsigned s_rem[4], s_bad[4];
  if ((i & 31u) == 0) {
    s_rem[i >> 5] = rem, s_bad[i >> 5] = bad;
  }
  __syncthreads();
  if (i == 0) {
    rem = s_rem[0] ^ s_rem[1] ^ s_rem[2] ^ s_rem[3];
    bad = s_bad[0] + s_bad[1] + s_bad[2] + s_bad[3];
    TbResult out;
    out.crc_remainder = rem;
    out.nof_cb_failed = (unsigned short)min(bad, 0xffffu);
    out.tb_crc_ok = (bad == 0 && (w.nof_cbs == 1 || rem == 0)) ? 1 : 0;
    out.reserved = 0;
    tb_results[tb] = out;
  }
}

}  // namespace ldpc_b200
