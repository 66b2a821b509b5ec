/* TEST INFRASTRUCTURE ONLY. Plain-C restatement ("port") of the reference algorithm for the PUSCH LDPC receive chain.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this; the product
 * library (libsrsran_ldpc_cuda.so) never links or calls it and has no CPU fallback.
 *
 * Parity pinned: tests/test_oracle_vs_reference.py checks every function below bit-for-bit against the reference's own
 * compiled code (oracle/_ref/libsrsran_ref.so) in the build container, and tests/golden/ holds fixtures generated from the
 * reference (tools/gen_golden.py) that are re-checked wherever /root/reference is absent.
 *
 * Conventions shared with include/ldpc_cuda.h: bg in {1,2}; crc poly code 0 none, 1 CRC16, 2 CRC24A, 3 CRC24B;
 * scale mode 0 generic round, 1 AVX floor(x*uint16(sf*2^16)>>16), 2 NEON floor(x*uint8(sf*2^8)>>8).
 */
#ifndef LDPC_ORACLE_H
#define LDPC_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

int      oracle_llr_add(int a, int b);
int      oracle_llr_sub(int a, int b);
int      oracle_llr_promotion_sum(int a, int b);
void     oracle_hard_decision(const int8_t* soft, unsigned n, uint8_t* out_packed);
unsigned oracle_crc_packed(int poly_code, const uint8_t* packed, unsigned nbits);
int      oracle_scale(int magnitude, float scaling_factor, int scale_mode);

/* Shift of edge (row m, column n) for lifting size Z, or -1 if there is no edge / bad arguments. */
int oracle_graph_shift(int bg, int Z, int m, int n);

/* Returns iterations (>= 1) when the reference would return a value, 0 for nullopt, < 0 for invalid arguments (where the
 * reference asserts). out_packed: ceil(bg_K * Z / 8) bytes, MSB first; left untouched where the reference leaves it. */
int oracle_decode(int           bg,
                  int           Z,
                  unsigned      nof_filler_bits,
                  int           crc_poly_code,
                  unsigned      max_iterations,
                  float         scaling_factor,
                  int           scale_mode,
                  const int8_t* llrs,
                  unsigned      nof_llrs,
                  uint8_t*      out_packed);

/* out: in/out soft buffer of out_len = N_short * Z LLRs. Returns 0, or < 0 for invalid arguments. */
int oracle_rate_dematch(int8_t*       out,
                        unsigned      out_len,
                        const int8_t* in,
                        unsigned      in_len,
                        int           new_data,
                        int           rv,
                        int           Qm,
                        unsigned      Nref,
                        unsigned      nof_filler_bits);

/* TB -> codeblock geometry (RX segmenter). meta_out: per CB 6 unsigned {Z, full_length, rm_length, nof_filler_bits,
 * cw_offset, nof_crc_bits}; returns the number of codeblocks C (no 52-CB cap), or < 0. meta_out may be NULL. */
int oracle_segment_rx(unsigned  tbs,
                      int       bg,
                      int       Qm,
                      unsigned  nof_layers,
                      unsigned  nof_ch_symbols,
                      unsigned* meta_out,
                      unsigned  max_cbs);

/* Test-signal generators (transmit side), ldpc_oracle_tx.c. Bits are unpacked, one per byte, filler bits = 254. */
int oracle_encode(int bg, int Z, const uint8_t* msg_bits, uint8_t* out_bits, unsigned out_len);
int oracle_rate_match(uint8_t*       out,
                      unsigned       E,
                      const uint8_t* in,
                      unsigned       N,
                      int            rv,
                      int            Qm,
                      unsigned       Nref,
                      unsigned       nof_filler_bits);

#ifdef __cplusplus
}
#endif
#endif
