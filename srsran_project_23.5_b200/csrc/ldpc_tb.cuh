// Transport-block assembly and TB CRC24A on the device.
//
// Replaces the part of pusch_decoder_impl::decode that follows the per-codeblock loop
// (lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:156-222): the nof_data_bits = K - crc - fillers leading bits
// of every decoded codeblock message are concatenated MSB first into the transport block (+ its 24 CRC bits when the
// block was segmented, :196-200), the CRC24A remainder of the result is formed (:216, crc_calculator::calculate_byte over
// whole bytes, init 0, no final xor) and tb_crc_ok is set as :205-222 does: one codeblock -> its own CRC flag; several
// -> every codeblock flag set AND remainder == 0.
//
// Work split: one thread per OUTPUT byte (gather: conflict-free coalesced stores; a byte takes its bits from at most two
// codeblocks). The CRC is linear in the message bits: a thread XORs x^e mod P for the set bits of its byte (table
// g_crc_pow of the decoder kernel), a block reduces its 256 bytes, multiplies the partial remainder by
// x^(bits behind the block) mod P and XORs it into the transport block's accumulator; the last block of a transport
// block to finish counts the failed codeblocks and writes the public result. HBM traffic = the bits themselves, once.
#pragma once

#include "ldpc_device.cuh"

namespace ldpc_b200 {

struct TbWork {
  unsigned long long tb_offset;  // bytes from the transport-block output base
  unsigned first_cb;             // index of the block's first entry in the TbCb array
  unsigned nof_cbs;              // C
  unsigned data_bits;            // bits taken from each codeblock message: K - nof_crc_bits - nof_filler_bits
  unsigned nbytes;               // bytes written: (tbs + (C > 1 ? 24 : 0)) / 8
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

// grid (chunks of 256 output bytes, transport blocks). scratch: {remainder accumulator, finished-chunk counter} per
// transport block, zero on entry.
__global__ void __launch_bounds__(TB_THREADS) ldpc_tb_assemble_kernel(const TbWork* __restrict__ tbs,
                                                                      const TbCb* __restrict__ cbs,
                                                                      const uint8_t* __restrict__ msgs,
                                                                      const CbResult* __restrict__ results,
                                                                      uint8_t* __restrict__ tb_out,
                                                                      TbResult* __restrict__ tb_results,
                                                                      unsigned* scratch)
{
  const unsigned tb = blockIdx.y;
  const TbWork w = tbs[tb];
  const unsigned j0 = blockIdx.x * TB_THREADS;
  if (j0 >= w.nbytes) {
    return;
  }
  const unsigned L = min((unsigned)TB_THREADS, w.nbytes - j0);
  const unsigned i = threadIdx.x;
  __shared__ unsigned red[TB_THREADS / 32];
  __shared__ unsigned last_flag;
  unsigned r = 0;
  if (i < L) {
    const unsigned p = 8u * (j0 + i);    // first transport-block bit of this byte
    const unsigned cb = p / w.data_bits;  // codeblock it starts in
    const unsigned o = p - cb * w.data_bits;
    const uint8_t* s = msgs + cbs[w.first_cb + cb].src_offset;
    const unsigned two = ((unsigned)s[o >> 3] << 8) | (unsigned)s[(o >> 3) + 1];  // message has >= 16 bits behind the data
    unsigned v = (two >> (8 - (o & 7u))) & 0xffu;
    const unsigned avail = w.data_bits - o;
    if (avail < 8) {  // the byte straddles two codeblocks
      v &= 0xff00u >> avail;
      if (cb + 1 < w.nof_cbs) {
        v |= (unsigned)msgs[cbs[w.first_cb + cb + 1].src_offset] >> avail;
      }
    }
    tb_out[w.tb_offset + j0 + i] = (uint8_t)v;
    // bit b of the byte (LSB = 0) carries x^(8 (L - 1 - i) + b) inside the chunk
    const unsigned* pw = g_crc_pow[TB_CRC24A_CODE - 1] + 8u * (L - 1 - i);
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const unsigned e = __ldg(pw + b);
      r ^= ((v >> b) & 1u) ? e : 0u;
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
  const unsigned nof_chunks = (w.nbytes + TB_THREADS - 1) / TB_THREADS;
  if (i == 0) {
    unsigned acc = 0;
#pragma unroll
    for (int k = 0; k < TB_THREADS / 32; ++k) {
      acc ^= red[k];
    }
    // transport-block bits behind this chunk, + 24: crc_calculator's checksum is the remainder of m(x) x^24, so that
    // crc_remainder equals crc_calculator::calculate_byte of the assembled bytes and not only its zero test
    const unsigned behind = 8u * (w.nbytes - j0 - L) + 24u;
    if (acc != 0) {
      acc = tb_mulmod24(acc, tb_xpow24(behind));
      atomicXor(&scratch[2 * tb], acc);
    }
    __threadfence();
    last_flag = (atomicAdd(&scratch[2 * tb + 1], 1u) == nof_chunks - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (!last_flag) {
    return;
  }
  // last chunk of the transport block to finish: codeblock flags (pusch_decoder_impl.cpp:188-195) and the verdict
  __threadfence();
  __shared__ unsigned failed;
  if (i == 0) {
    failed = 0;
  }
  __syncthreads();
  unsigned mine = 0;
  for (unsigned k = i; k < w.nof_cbs; k += TB_THREADS) {
    const TbCb c = cbs[w.first_cb + k];
    mine += (c.already_ok || results[c.result_index].crc_ok) ? 0u : 1u;
  }
  if (mine) {
    atomicAdd(&failed, mine);
  }
  __syncthreads();
  if (i == 0) {
    const unsigned rem = atomicXor(&scratch[2 * tb], 0u);
    TbResult out;
    out.crc_remainder = rem;
    out.nof_cb_failed = (unsigned short)min(failed, 0xffffu);
    out.tb_crc_ok = (failed == 0 && (w.nof_cbs == 1 || rem == 0)) ? 1 : 0;
    out.reserved = 0;
    tb_results[tb] = out;
  }
}

}  // namespace ldpc_b200
