/*
 * ldpc_cuda.h - C ABI of the B200 (sm_100a) PUSCH LDPC receive chain: rate dematching with HARQ soft combining,
 * layered normalised min-sum decoding of 3GPP BG1/BG2 over all 51 lifting sizes, CRC16/24A/24B early stop.
 *
 * This is the drop-in boundary. The entry points are exactly what a "cuda" variant of the reference's factories binds
 * (paths relative to the srsRAN Project 23.5 tree):
 *
 *   ldpc_cuda_decode           <->  ldpc_decoder::decode
 *                                   include/srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h:44-74
 *                                   (lib/phy/upper/channel_coding/ldpc/ldpc_decoder_impl.cpp:60-146)
 *   ldpc_cuda_rate_dematch     <->  ldpc_rate_dematcher::rate_dematch
 *                                   include/srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h:52-55
 *                                   (lib/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher_impl.cpp:43-198)
 *   ldpc_cuda_create/destroy   <->  ldpc_decoder_factory::create / ldpc_rate_dematcher_factory::create
 *                                   include/srsran/phy/upper/channel_coding/channel_coding_factories.h:52-77
 *   ldpc_cuda_decode_batch*    <->  the per-codeblock loop of pusch_decoder_impl::decode
 *                                   lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:156-201
 *                                   (rate_dematch :182, decode_cblk :94-118/:186, CRC flags :188-195) - new batched entry
 *   ldpc_cuda_softbuffer_*     <->  rx_softbuffer::get_codeblock_soft_bits / get_codeblocks_crc storage
 *                                   include/srsran/phy/upper/rx_softbuffer.h:42-72 (HBM-resident equivalent)
 *   ldpc_cuda_tb_assemble*     <->  transport-block assembly + TB CRC24A + verdict of pusch_decoder_impl::decode
 *                                   lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:156-222 (:196-200 bit copy,
 *                                   :205-222 CRC check) - on the device, so a slot's transport blocks never leave HBM
 *   ldpc_cuda_demodulate*      <->  demodulation_mapper::demodulate_soft
 *                                   include/srsran/phy/upper/channel_modulation/demodulation_mapper.h
 *                                   (lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:36-106) followed by
 *                                   pusch_demodulator_impl::descramble
 *                                   lib/phy/upper/channel_processors/pusch_demodulator_impl.cpp:100-152 - the producer of
 *                                   the LLRs, so that they are born in device memory
 *   ldpc_cuda_ulsch_*          <->  ulsch_demultiplex::demultiplex / get_placeholders
 *                                   lib/phy/upper/channel_processors/ulsch_demultiplex_impl.cpp:316-453
 *   ldpc_cuda_segment          <->  ldpc_segmenter_rx::segment geometry
 *                                   lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_impl.cpp:226-308 (no 52-CB cap)
 *
 * Conventions
 *   - Plain C, no CUDA or torch types. "_d" suffix = device pointer on the handle's device; everything else is host
 *     memory owned by the caller, not retained after the call returns.
 *   - LLRs are int8 in {-127} u [-120,120] u {127} (log_likelihood_ratio, log_likelihood_ratio.h:234,239).
 *   - Hard bits are packed MSB first: bit i -> byte i/8, mask 0x80 >> (i%8) (bit_buffer, adt/bit_buffer.h:95-139).
 *   - Error convention: the reference asserts/aborts on invalid arguments and has no error codes
 *     (support/srsran_assert.h:60-105). The C ABI returns 0 on success and a negative ldpc_cuda_status otherwise; the C++
 *     adaptor (srsran_project_23.5_b200/host) turns a non-zero status into the reference's assert/abort. A decode that does
 *     not pass its CRC is NOT an error: it is reported as data (iterations == 0 <-> nullopt).
 *   - There is no CPU fallback: every compute entry point fails with LDPC_CUDA_ERR_CUDA if no sm_100 device is usable.
 *   - A handle is not thread-safe (same as one reference decoder object); use one handle per host thread / per GPU.
 */
#ifndef LDPC_CUDA_H
#define LDPC_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LDPC_CUDA_ABI_VERSION 1

typedef struct ldpc_cuda_handle ldpc_cuda_handle;

typedef enum {
  LDPC_CUDA_OK            = 0,
  LDPC_CUDA_ERR_ARG       = -1, /* invalid argument (where the reference would assert) */
  LDPC_CUDA_ERR_CUDA      = -2, /* CUDA runtime error, no device, or not an sm_100 device */
  LDPC_CUDA_ERR_CAPACITY  = -3, /* batch larger than the handle was created for */
  LDPC_CUDA_ERR_UNSUPPORTED = -4
} ldpc_cuda_status;

/* CRC polynomials the PUSCH decoder can hand to the LDPC decoder (pusch_decoder_impl.cpp:44-56 select_crc). */
typedef enum { LDPC_CUDA_CRC_NONE = 0, LDPC_CUDA_CRC16 = 1, LDPC_CUDA_CRC24A = 2, LDPC_CUDA_CRC24B = 3 } ldpc_cuda_crc;

/* Check-to-variable scaling arithmetic. GENERIC is the parity target (ldpc_decoder_generic.cpp:80-89:
 * round(float(m)*sf)); AVX reproduces avx2_support.h:82-91 / avx512_support.h:82-91 ((m*uint16(sf*65536))>>16);
 * NEON reproduces neon_support.h:73-86 ((m*uint8(sf*256))>>8). */
typedef enum { LDPC_CUDA_SCALE_GENERIC = 0, LDPC_CUDA_SCALE_AVX = 1, LDPC_CUDA_SCALE_NEON = 2 } ldpc_cuda_scale_mode;

/* Decoder algorithm parameters = ldpc_decoder::configuration::algorithm_details (ldpc_decoder.h:46-51) + scale mode. */
typedef struct {
  uint32_t max_iterations; /* > 0, default 6 */
  float    scaling_factor; /* in (0,1), default 0.8f */
  uint32_t scale_mode;     /* ldpc_cuda_scale_mode */
  uint32_t reserved;
} ldpc_cuda_algorithm;

/* One codeblock = codeblock_metadata (include/srsran/phy/upper/codeblock_metadata.h:38-73) reduced to what the path
 * reads, plus the buffer locations for the batched entries. */
typedef struct {
  uint8_t  base_graph;      /* 1 or 2                                   tb_common.base_graph   */
  uint8_t  crc_poly;        /* ldpc_cuda_crc used for early stop / check; NONE = run max_iterations, no check */
  uint8_t  rv;              /* 0..3                                     tb_common.rv       (dematch) */
  uint8_t  mod_order;       /* bits per symbol Qm: 1, 2, 4, 6, 8        tb_common.mod      (dematch) */
  uint16_t lifting_size;    /* Z, one of the 51 sizes                   tb_common.lifting_size */
  uint16_t nof_filler_bits; /* F                                        cb_specific.nof_filler_bits */
  uint32_t Nref;            /* LBRM buffer limit, 0 = none              tb_common.Nref     (dematch) */
  uint32_t rm_length;       /* E: rate-matched LLRs of this codeblock   cb_specific.rm_length (dematch) */
  uint32_t in_length;       /* decode-only entries: number of input LLRs, (bg_K+2)Z <= in_length <= N_short*Z */
  uint8_t  new_data;        /* 1: first transmission (copy), 0: retransmission (combine)   (dematch) */
  uint8_t  skip_decode;     /* 1: CRC already OK from an earlier transmission - dematch/combine only
                               (pusch_decoder_impl.cpp:184) */
  uint8_t  early_stop;      /* 1: check the CRC after every iteration; 0: run max_iterations then check once
                               (pusch_decoder_impl.cpp:104-117). Ignored when crc_poly == NONE. */
  uint8_t  reserved;
  uint64_t in_offset;       /* byte offset of this codeblock's input LLRs inside the batch input buffer */
  uint64_t soft_offset;     /* byte offset of its N_short*Z soft buffer inside the soft-buffer arena (fused entries) */
  uint64_t out_offset;      /* byte offset of its ceil(bg_K*Z/8) output bytes inside the batch output buffer */
} ldpc_cuda_cb_desc;

/* Per-codeblock result = optional<unsigned> of ldpc_decoder::decode + the CRC flag the PUSCH decoder keeps. */
typedef struct {
  uint16_t iterations; /* iterations run; with a CRC: value of the optional when crc_ok, else max_iterations run */
  uint8_t  crc_ok;     /* 1 iff the optional has a value */
  uint8_t  flags;      /* bit0: all-zero input shortcut taken (ldpc_decoder_impl.cpp:88-94) */
} ldpc_cuda_cb_result;

/* ---------------------------------------------------------------------------------------------------------------------
 * Life cycle.
 */
int         ldpc_cuda_abi_version(void);
const char* ldpc_cuda_status_string(int status);
/* Number of usable sm_100 devices (0 when there is none / no driver). Never fails. */
int         ldpc_cuda_device_count(void);

/* flags for ldpc_cuda_create */
#define LDPC_CUDA_FLAG_FORCE_SCALAR 1u /* always use the general one-lane-per-thread kernel (cross-check / debugging) */
#define LDPC_CUDA_FLAG_COPY_ONLY 2u    /* measurement aid: the host entries perform every copy of a call (same chunks, same
                                          streams) but launch no decode kernel, so outputs and results are undefined; gives
                                          the copy-only ceiling of the end-to-end figures (bench.py) */

/* device: CUDA ordinal. max_codeblocks: largest batch a single call may carry (device scratch and pinned staging are
 * sized for it: max_codeblocks * (25344 + 1056 + descriptors) bytes). */
int  ldpc_cuda_create(ldpc_cuda_handle** handle, int device, uint32_t max_codeblocks, uint32_t flags);
void ldpc_cuda_destroy(ldpc_cuda_handle* handle);
int  ldpc_cuda_set_algorithm(ldpc_cuda_handle* handle, const ldpc_cuda_algorithm* alg);
/* The handle's CUDA stream as an opaque pointer (cudaStream_t) so callers can order their own work after it. */
void* ldpc_cuda_stream(ldpc_cuda_handle* handle);
int   ldpc_cuda_synchronize(ldpc_cuda_handle* handle);
/* Number of kernels this handle has launched since creation (bench.py's gpu_launches claim). */
uint64_t ldpc_cuda_launch_count(const ldpc_cuda_handle* handle);
/* How many of those were launches of the general (one codeblock per block, any configuration) decoder kernel instead of the
 * tensor-memory kernel: 0 unless LDPC_CUDA_FLAG_FORCE_SCALAR is set or the configuration is not covered (tests assert it). */
uint64_t ldpc_cuda_general_kernel_launch_count(const ldpc_cuda_handle* handle);

/* ---------------------------------------------------------------------------------------------------------------------
 * Drop-in, one codeblock per call, host buffers, synchronous (what the reference interfaces need).
 */

/* ldpc_decoder::decode. llrs: desc->in_length int8. out_bits: ceil(bg_K*Z/8) bytes, in/out (left untouched where the
 * reference leaves it untouched: all-zero input with a CRC). *iterations = N >= 1 when the CRC passed after N iterations
 * (optional has a value), 0 otherwise (nullopt). Uses desc->{base_graph,lifting_size,nof_filler_bits,crc_poly,in_length};
 * early stop is always on when crc_poly != NONE, as in the reference interface. */
int ldpc_cuda_decode(ldpc_cuda_handle*        handle,
                     const ldpc_cuda_cb_desc* desc,
                     const int8_t*            llrs,
                     uint8_t*                 out_bits,
                     uint32_t*                iterations);

/* ldpc_rate_dematcher::rate_dematch. soft_inout: N_short*Z int8, persistent HARQ buffer (in/out). in: desc->rm_length
 * int8. Uses desc->{base_graph,lifting_size,nof_filler_bits,rv,mod_order,Nref,rm_length,new_data}. */
int ldpc_cuda_rate_dematch(ldpc_cuda_handle* handle, const ldpc_cuda_cb_desc* desc, int8_t* soft_inout, const int8_t* in);

/* ---------------------------------------------------------------------------------------------------------------------
 * Batched entries. All codeblocks of a call may differ in every descriptor field; internally they are grouped by
 * (base_graph, lifting_size) and each group is one kernel launch.
 */

/* Decode-only, HOST buffers: H2D of descriptors + LLRs, kernels, D2H of bits + results, all inside the call. */
int ldpc_cuda_decode_batch(ldpc_cuda_handle*        handle,
                           uint32_t                 nof_codeblocks,
                           const ldpc_cuda_cb_desc* descs,
                           const int8_t*            llrs,     /* base; codeblock i at llrs + descs[i].in_offset */
                           uint8_t*                 out_bits, /* base; codeblock i at out_bits + descs[i].out_offset */
                           ldpc_cuda_cb_result*     results);

/* Decode-only, DEVICE buffers, asynchronous on the handle's stream. descs are host memory (copied at enqueue);
 * results_d / out_bits_d / llrs_d are device memory. Call ldpc_cuda_synchronize (or use the stream) before reading. */
int ldpc_cuda_decode_batch_device(ldpc_cuda_handle*        handle,
                                  uint32_t                 nof_codeblocks,
                                  const ldpc_cuda_cb_desc* descs,
                                  const int8_t*            llrs_d,
                                  uint8_t*                 out_bits_d,
                                  ldpc_cuda_cb_result*     results_d);

/* Fused rate-dematch (+HARQ combine) + decode + CRC, soft buffers resident in a device arena (see below).
 * HOST input/output: rm_llrs = rate-matched LLRs (codeblock i: descs[i].rm_length bytes at in_offset). For each
 * codeblock: dematch/combine into arena[soft_offset .. +N_short*Z) exactly as rate_dematch would, then (unless
 * skip_decode) decode from that soft buffer. */
int ldpc_cuda_dematch_decode_batch(ldpc_cuda_handle*        handle,
                                   uint32_t                 nof_codeblocks,
                                   const ldpc_cuda_cb_desc* descs,
                                   const int8_t*            rm_llrs,
                                   uint8_t*                 out_bits,
                                   ldpc_cuda_cb_result*     results);
/* Same, DEVICE input/output, asynchronous on the handle's stream. */
int ldpc_cuda_dematch_decode_batch_device(ldpc_cuda_handle*        handle,
                                          uint32_t                 nof_codeblocks,
                                          const ldpc_cuda_cb_desc* descs,
                                          const int8_t*            rm_llrs_d,
                                          uint8_t*                 out_bits_d,
                                          ldpc_cuda_cb_result*     results_d);

/* ---------------------------------------------------------------------------------------------------------------------
 * Transport-block assembly + TB CRC on the device (what pusch_decoder_impl::decode does after its codeblock loop,
 * pusch_decoder_impl.cpp:156-222).
 */

/* One transport block = a run of consecutive descriptors of the same call (ldpc_cuda_segment order). */
typedef struct {
  uint32_t first_codeblock; /* index of its first codeblock in descs / results */
  uint32_t nof_codeblocks;  /* C >= 1 */
  uint32_t tbs_bits;        /* transport block size in bits, a multiple of 8 (transport_block is a byte span) */
  uint32_t nof_cb_crc_bits; /* cb_specific.nof_crc_bits of its codeblocks: 24 when C > 1, else 16 or 24 (ldpc_cuda_segment) */
  uint64_t tb_offset;       /* byte offset of the assembled block inside the transport-block output buffer; the kernel
                               writes tbs_bits/8 bytes, plus the 3 bytes of the TB CRC24A when C > 1 */
} ldpc_cuda_tb_desc;

/* pusch_decoder_result::tb_crc_ok and what it was derived from. */
typedef struct {
  uint32_t crc_remainder;  /* CRC24A remainder over the assembled bytes (meaningful when C > 1; 0 = check passed) */
  uint16_t nof_cb_failed;  /* codeblocks whose CRC flag is not set (neither by this call nor by skip_decode) */
  uint8_t  tb_crc_ok;      /* C == 1: the codeblock's CRC flag; C > 1: no failed codeblock and crc_remainder == 0 */
  uint8_t  reserved;
} ldpc_cuda_tb_result;

/* DEVICE buffers, asynchronous on the handle's stream (enqueue it behind the decode call that produced out_bits_d and
 * results_d). descs / tbs are host memory, copied at enqueue; descs are the descriptors of that decode call
 * (out_offset, skip_decode, base_graph, lifting_size, nof_filler_bits are read). Unlike the reference, which leaves the
 * caller's transport_block untouched unless every codeblock passed, the assembled bits are always written: consumers
 * look at tb_crc_ok. */
int ldpc_cuda_tb_assemble_device(ldpc_cuda_handle*          handle,
                                 uint32_t                   nof_tbs,
                                 const ldpc_cuda_tb_desc*   tbs,
                                 uint32_t                   nof_codeblocks,
                                 const ldpc_cuda_cb_desc*   descs,
                                 const uint8_t*             out_bits_d,
                                 const ldpc_cuda_cb_result* results_d,
                                 uint8_t*                   tb_out_d,
                                 ldpc_cuda_tb_result*       tb_results_d);

/* ldpc_cuda_tb_assemble_device for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 * Same arguments, same results, same kernels; the work records and per-chunk CRC remainders live in one fixed array in
 * device memory (file-scope `static __device__`, about 100 KB, held by every GPU that loads the library until exit) and
 * the records are staged through a fixed host array, so the call performs no cudaMalloc / cudaFree / cudaHostAlloc.
 * Limits per call (beyond them: LDPC_CUDA_ERR_CAPACITY and a message on stderr, nothing is enqueued):
 * LDPC_CUDA_NOALLOC_MAX_TBS transport blocks, LDPC_CUDA_NOALLOC_MAX_CBS codeblocks (= nof_codeblocks, the length of
 * descs), LDPC_CUDA_NOALLOC_MAX_CHUNKS chunks of 1024 assembled bytes summed over the blocks (a 16-cell slot of full-band
 * blocks needs 16 / 2432 / 2512). The array is shared by all handles of a device and keeps the last call's contents (the
 * remainders a call reads are cleared first): calls on one stream are ordered by it; calls whose GPU work may overlap
 * (different handles or streams of one device) race on it and must be serialised by the caller. The first call on a
 * device resolves the array's address (cudaGetSymbolAddress); make it once before capturing a graph. */
#define LDPC_CUDA_NOALLOC_MAX_TBS 64u
#define LDPC_CUDA_NOALLOC_MAX_CBS 4096u
#define LDPC_CUDA_NOALLOC_MAX_CHUNKS 8192u
int ldpc_cuda_tb_assemble_device_noalloc(ldpc_cuda_handle*          handle,
                                         uint32_t                   nof_tbs,
                                         const ldpc_cuda_tb_desc*   tbs,
                                         uint32_t                   nof_codeblocks,
                                         const ldpc_cuda_cb_desc*   descs,
                                         const uint8_t*             out_bits_d,
                                         const ldpc_cuda_cb_result* results_d,
                                         uint8_t*                   tb_out_d,
                                         ldpc_cuda_tb_result*       tb_results_d);

/* ldpc_cuda_dematch_decode_batch followed by the assembly of its transport blocks in the same call, HOST buffers:
 * out_bits / results as in ldpc_cuda_dematch_decode_batch; tb_out / tb_results receive the assembled blocks.
 * out_bits may be NULL when no codeblock has skip_decode set: the codeblock messages then stay in device memory and only
 * the assembled blocks and the verdicts are copied back (less PCIe traffic in both directions). */
int ldpc_cuda_dematch_decode_tb_batch(ldpc_cuda_handle*        handle,
                                      uint32_t                 nof_codeblocks,
                                      const ldpc_cuda_cb_desc* descs,
                                      const int8_t*            rm_llrs,
                                      uint8_t*                 out_bits,
                                      ldpc_cuda_cb_result*     results,
                                      uint32_t                 nof_tbs,
                                      const ldpc_cuda_tb_desc* tbs,
                                      uint8_t*                 tb_out,
                                      ldpc_cuda_tb_result*     tb_results);

/* ---------------------------------------------------------------------------------------------------------------------
 * Upstream producer of the LLRs: soft demodulation of equalised symbols + descrambling, on the device
 * (pusch_demodulator_impl::demodulate after the equaliser, pusch_demodulator_impl.cpp:88-98: demodulate_soft, descramble).
 * Results are bit-exact with the reference's generic per-symbol code (its AVX2 kernels differ from that by one LLR step on
 * a few values per million).
 */

/* One job = one codeword (one PUSCH transmission). Symbols are complex floats (re, im interleaved = cf_t), one noise
 * variance per symbol; noise_var <= 0 or NaN gives LLR 0 as in the reference. */
typedef struct {
  uint8_t  mod_order;     /* bits per symbol: 1 (BPSK), 2, 4, 6, 8; 0 = the input already is int8 LLRs (descramble only) */
  uint8_t  pi2_bpsk;      /* with mod_order 1: pi/2-BPSK (modulation_scheme::PI_2_BPSK) */
  uint8_t  descramble;    /* 1: apply the scrambling sequence of c_init (LLR sign flips), 0: demodulate only */
  uint8_t  reserved;
  uint32_t nof_symbols;   /* symbols of the job; LLRs written = nof_symbols * mod_order (mod_order 0: number of LLRs) */
  uint32_t c_init;        /* rnti * 2^15 + n_id (pusch_demodulator_impl.cpp:105) */
  uint32_t reserved2;
  uint64_t symbol_offset; /* index of its first symbol / noise variance in the input arrays (mod_order 0: first input LLR) */
  uint64_t llr_offset;    /* byte offset of its first LLR in the output buffer */
} ldpc_cuda_demod_desc;

/* DEVICE buffers, asynchronous on the handle's stream. in_llrs_d may be NULL unless a job has mod_order 0. The output
 * can be handed to ldpc_cuda_dematch_decode_batch_device as rm_llrs_d without leaving the device. */
int ldpc_cuda_demodulate_batch_device(ldpc_cuda_handle*           handle,
                                      uint32_t                    nof_jobs,
                                      const ldpc_cuda_demod_desc* jobs,
                                      const float*                symbols_d,
                                      const float*                noise_vars_d,
                                      const int8_t*               in_llrs_d,
                                      int8_t*                     llrs_d);

/* demodulation_mapper::demodulate_soft, HOST buffers, synchronous. llrs: nof_symbols * mod_order int8. */
int ldpc_cuda_demodulate_soft(ldpc_cuda_handle* handle,
                              int               mod_order,
                              int               pi2_bpsk,
                              uint32_t          nof_symbols,
                              const float*      symbols,
                              const float*      noise_vars,
                              int8_t*           llrs);

/* pusch_demodulator_impl::descramble without repetition placeholders, HOST buffers, synchronous (in may equal out). */
int ldpc_cuda_descramble(ldpc_cuda_handle* handle, uint32_t c_init, uint32_t nof_llrs, const int8_t* in, int8_t* out);

/* UCI repetition placeholders of pusch_demodulator_impl::descramble (lib/phy/upper/channel_processors/
 * pusch_demodulator_impl.cpp:100-152, positions from ulsch_placeholder_list::for_each, include/srsran/phy/upper/
 * channel_processors/ulsch_placeholder_list.h:73-95): for every listed RE and layer, soft bit
 * y = mod_order * (nof_layers * re + layer) + 1 repeats the scrambling chip of the bit before it and the mod_order - 2 bits
 * behind it are not descrambled. One descriptor per codeword that carries placeholders. */
typedef struct ldpc_cuda_placeholder_desc {
  uint64_t llr_offset; /* first LLR of the codeword, bytes from llrs_d */
  uint32_t c_init;     /* rnti * 2^15 + n_id */
  uint32_t nof_llrs;   /* LLRs of the codeword (placeholder bits beyond it are ignored) */
  uint32_t first_re;   /* index of the codeword's first entry in re_indexes */
  uint16_t nof_re;     /* number of placeholder REs (config.placeholders) */
  uint8_t  mod_order;  /* 1, 2, 4, 6, 8 */
  uint8_t  nof_layers; /* config.nof_tx_layers */
} ldpc_cuda_placeholder_desc; /* 24 bytes */

/* Amends, in HBM and asynchronously on the handle's stream, codewords that were descrambled as plain data
 * (ldpc_cuda_demodulate_batch_device with descramble = 1) so that they equal the output of the reference's descramble with
 * config.placeholders. jobs and re_indexes (HOST arrays) are consumed before the call returns. */
int ldpc_cuda_descramble_placeholders_device(ldpc_cuda_handle*                 handle,
                                             uint32_t                          nof_jobs,
                                             const ldpc_cuda_placeholder_desc* jobs,
                                             const uint16_t*                   re_indexes,
                                             uint32_t                          nof_re_indexes,
                                             int8_t*                           llrs_d);

/* pusch_demodulator_impl::descramble WITH config.placeholders, HOST buffers, synchronous (in may equal out):
 * ldpc_cuda_descramble followed by the placeholder amendment before the LLRs leave the device. */
int ldpc_cuda_descramble_uci(ldpc_cuda_handle* handle,
                             uint32_t          c_init,
                             uint32_t          nof_llrs,
                             int               mod_order,
                             uint32_t          nof_layers,
                             uint32_t          nof_re,
                             const uint16_t*   re_indexes,
                             const int8_t*     in,
                             int8_t*           out);

/* ---------------------------------------------------------------------------------------------------------------------
 * UL-SCH demultiplexing (TS 38.212 6.2.7) on the device: ulsch_demultiplex::demultiplex / get_placeholders
 * (include/srsran/phy/upper/channel_processors/ulsch_demultiplex.h:37-103,
 *  lib/phy/upper/channel_processors/ulsch_demultiplex_impl.cpp:316-453; called at pusch_processor_impl.cpp:259) - between
 * the descrambler and the rate dematcher, so that a PUSCH with HARQ-ACK / CSI multiplexed keeps its LLRs in device memory.
 */
typedef struct {
  uint8_t  mod_order;                   /* bits per symbol of config.modulation: 1, 2, 4, 6, 8 */
  uint8_t  nof_layers;                  /* config.nof_layers */
  uint8_t  start_symbol_index;          /* config.start_symbol_index */
  uint8_t  nof_symbols;                 /* config.nof_symbols */
  uint8_t  dmrs_type;                   /* config.dmrs: 1 or 2 */
  uint8_t  nof_cdm_groups_without_data; /* config.nof_cdm_groups_without_data */
  uint16_t dmrs_symbol_mask;            /* config.dmrs_symbol_mask, bit l = OFDM symbol l carries DM-RS */
  uint16_t nof_prb;                     /* config.nof_prb */
  uint16_t reserved;
  uint32_t nof_harq_ack_rvd;            /* config.nof_harq_ack_rvd (G_rvd^ACK) */
  uint32_t nof_enc_harq_ack_bits;       /* harq_ack.size()  = message_information::nof_enc_harq_ack_bits */
  uint32_t nof_enc_csi_part1_bits;      /* csi_part1.size() */
  uint32_t nof_enc_csi_part2_bits;      /* csi_part2.size() */
  uint64_t in_offset;                   /* byte offset of the codeword's soft bits in the input buffer */
  uint64_t sch_offset;                  /* byte offset of its SCH data in the SCH output buffer */
  uint64_t harq_ack_offset;             /* byte offsets of the three UCI fields in the UCI output buffer */
  uint64_t csi_part1_offset;
  uint64_t csi_part2_offset;
} ldpc_cuda_ulsch_desc; /* 72 bytes */

/* Sizes that follow from a descriptor: soft bits read (= input.size()), SCH data soft bits written (= sch_data.size()).
 * Returns LDPC_CUDA_ERR_ARG for a configuration the reference asserts on (no DM-RS symbol, fields that do not fit). */
int ldpc_cuda_ulsch_sizes(const ldpc_cuda_ulsch_desc* desc, uint32_t* nof_input_bits, uint32_t* nof_sch_bits);

/* DEVICE buffers, asynchronous on the handle's stream; descs are host memory, consumed before the call returns. sch_d may
 * be handed to ldpc_cuda_dematch_decode_batch_device as rm_llrs_d. */
int ldpc_cuda_ulsch_demultiplex_batch_device(ldpc_cuda_handle*           handle,
                                             uint32_t                    nof_codewords,
                                             const ldpc_cuda_ulsch_desc* descs,
                                             const int8_t*               in_llrs_d,
                                             int8_t*                     sch_d,
                                             int8_t*                     uci_d);

/* ulsch_demultiplex::demultiplex, HOST buffers, synchronous (the offsets in desc are ignored). */
int ldpc_cuda_ulsch_demultiplex(ldpc_cuda_handle*           handle,
                                const ldpc_cuda_ulsch_desc* desc,
                                const int8_t*               input,
                                int8_t*                     sch_data,
                                int8_t*                     harq_ack,
                                int8_t*                     csi_part1,
                                int8_t*                     csi_part2);

/* ulsch_demultiplex::get_placeholders (host only, no GPU needed): resource-element indexes (in units of elements of the
 * codeword, as ulsch_placeholder_list holds them) of the fields that carry exactly ONE information bit. nof_*_bits = O_ACK,
 * O_CSI-1, O_CSI-2. Returns the number of indexes (written up to max_indexes) or a negative status. */
int ldpc_cuda_ulsch_placeholders(const ldpc_cuda_ulsch_desc* desc,
                                 uint32_t                    nof_harq_ack_bits,
                                 uint32_t                    nof_csi_part1_bits,
                                 uint32_t                    nof_csi_part2_bits,
                                 uint16_t*                   re_indexes,
                                 uint32_t                    max_indexes);

/* ---------------------------------------------------------------------------------------------------------------------
 * HBM-resident soft-buffer arena (storage behind rx_softbuffer; zero-initialised once, like the reference pool,
 * rx_softbuffer_codeblock_pool.h). Offsets in ldpc_cuda_cb_desc::soft_offset index into it.
 */
int ldpc_cuda_softbuffer_reserve(ldpc_cuda_handle* handle, uint64_t nof_bytes); /* (re)allocates and zeroes */
int ldpc_cuda_softbuffer_zero(ldpc_cuda_handle* handle, uint64_t offset, uint64_t nof_bytes);
int ldpc_cuda_softbuffer_read(ldpc_cuda_handle* handle, uint64_t offset, int8_t* dst, uint64_t nof_bytes);
int ldpc_cuda_softbuffer_write(ldpc_cuda_handle* handle, uint64_t offset, const int8_t* src, uint64_t nof_bytes);
/* Device address of the arena (NULL before reserve). */
void* ldpc_cuda_softbuffer_ptr(ldpc_cuda_handle* handle);

/* A soft-buffer arena that outlives / is shared by several handles of one device (= the memory of one
 * rx_softbuffer_pool, lib/phy/upper/rx_softbuffer_codeblock_pool.h:35-117, which all PUSCH decoder threads share).
 * arena_create allocates nof_bytes of device memory, zero-initialised; softbuffer_attach makes a handle use it (not owned:
 * destroying the handle or reserving a private arena leaves it alone). Handles working on the same arena must use
 * disjoint soft_offset ranges at any one time, as the reference's decoders use disjoint codeblock entries. */
int ldpc_cuda_arena_create(int device, uint64_t nof_bytes, void** arena_d);
int ldpc_cuda_arena_destroy(int device, void* arena_d);
int ldpc_cuda_softbuffer_attach(ldpc_cuda_handle* handle, void* arena_d, uint64_t nof_bytes);

/* Sizes the handle's device staging buffers of the HOST entries ahead of time (they otherwise grow on demand, i.e. with a
 * cudaMalloc inside the first call that needs more): in_bytes / out_bytes / tb_bytes = the largest input, codeblock-message
 * and transport-block byte ranges one call will bring. No counterpart in the reference. */
int ldpc_cuda_reserve_staging(ldpc_cuda_handle* handle, uint64_t in_bytes, uint64_t out_bytes, uint64_t tb_bytes);

/* Page-locked host memory for callers that stage their buffers (the slot aggregator of host/pusch_decoder_cuda.cpp copies
 * the LLRs of every PUSCH decoder thread into one such block): the host entries above then copy at PCIe speed instead of
 * the driver's bounce-buffer speed for pageable memory. Portable across devices. No counterpart in the reference.
 * LDPC_CUDA_HOST_WRITE_COMBINED: for buffers the CPU only WRITES (inputs): the stores bypass the CPU caches, so the copy
 * engine never has to snoop lines out of the caches of the cores that filled the buffer (measured on the B200 host: a
 * 22 MB input filled by 16 threads copies 2.4x slower from cached memory); CPU reads of such memory are very slow. */
#define LDPC_CUDA_HOST_WRITE_COMBINED 1u
int ldpc_cuda_host_alloc(uint64_t nof_bytes, uint32_t flags, void** ptr);
int ldpc_cuda_host_free(void* ptr);

/* ---------------------------------------------------------------------------------------------------------------------
 * Several GPUs of one box from ONE process (BASELINE.json north_star: "a slot's transport blocks are sharded
 * round-robin across the 8 GPUs of one box by host streams, with no NCCL on the decode path"). Replaces nothing in the
 * reference, whose unit of concurrency is one decoder object per uplink thread (apps/gnb/gnb_appconfig.h:367
 * nof_ul_threads, lib/phy/upper/uplink_processor_impl.cpp:155-171); what makes the split legal is that the codeblock loop
 * of pusch_decoder_impl::decode (lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:156-201) carries nothing from one
 * codeblock to the next. A pool owns one handle + one host thread per device; its calls are synchronous like the
 * single-device host entries and take the same HOST buffers and descriptors: every device copies its share straight from /
 * to the caller's buffers (pitched copies for interleaved shares), nothing is exchanged between devices.
 * A pool is not thread-safe (one caller at a time).
 */
typedef struct ldpc_cuda_pool ldpc_cuda_pool;

/* devices: distinct CUDA ordinals (sm_100). max_codeblocks_per_device: as ldpc_cuda_create. */
int      ldpc_cuda_pool_create(ldpc_cuda_pool** pool, const int* devices, uint32_t nof_devices,
                               uint32_t max_codeblocks_per_device, uint32_t flags);
void     ldpc_cuda_pool_destroy(ldpc_cuda_pool* pool);
uint32_t ldpc_cuda_pool_nof_devices(const ldpc_cuda_pool* pool);
/* The handle of device number index (0 .. nof_devices - 1), e.g. for ldpc_cuda_softbuffer_read; owned by the pool. */
ldpc_cuda_handle* ldpc_cuda_pool_handle(ldpc_cuda_pool* pool, uint32_t index);
int      ldpc_cuda_pool_set_algorithm(ldpc_cuda_pool* pool, const ldpc_cuda_algorithm* alg);
/* A soft-buffer arena of nof_bytes_per_device on EVERY device (zeroed); forgets where HARQ processes live. soft_offset of
 * a descriptor is an offset into the arena of whichever device the transport block is routed to, so a caller lays out ONE
 * arena and every device mirrors it (180 GB per GPU: cfg 5 needs 1 GB). */
int      ldpc_cuda_pool_softbuffer_reserve(ldpc_cuda_pool* pool, uint64_t nof_bytes_per_device);
uint64_t ldpc_cuda_pool_launch_count(const ldpc_cuda_pool* pool);
/* Codeblocks each device has processed since creation (counts[nof_counts >= nof_devices]). */
int      ldpc_cuda_pool_codeblocks_per_device(const ldpc_cuda_pool* pool, uint64_t* counts, uint32_t nof_counts);
/* Device number that holds the soft buffers of HARQ process harq_key, -1 when the pool has not seen it. */
int      ldpc_cuda_pool_device_of(const ldpc_cuda_pool* pool, uint64_t harq_key);

/* ldpc_cuda_decode_batch over all devices: codeblock i runs on device i mod nof_devices (SURVEY.md 8d cfg 4). */
int ldpc_cuda_pool_decode_batch(ldpc_cuda_pool*          pool,
                                uint32_t                 nof_codeblocks,
                                const ldpc_cuda_cb_desc* descs,
                                const int8_t*            llrs,
                                uint8_t*                 out_bits,
                                ldpc_cuda_cb_result*     results);

/* The same split with inputs and outputs in DEVICE memory (LLRs written there by the demodulator, ldpc_cuda_demodulate_*):
 * asynchronous, kernels only. llrs_d[g] / out_bits_d[g] are buffers on device number g addressed with the callers' offsets
 * (descs[i].in_offset / out_offset for the codeblocks i = g, g + G, ... it runs); results_d[g][k] receives the result of
 * codeblock g + k * G. Wait with ldpc_cuda_pool_synchronize. */
int ldpc_cuda_pool_decode_batch_device(ldpc_cuda_pool*             pool,
                                       uint32_t                    nof_codeblocks,
                                       const ldpc_cuda_cb_desc*    descs,
                                       const int8_t* const*        llrs_d,
                                       uint8_t* const*             out_bits_d,
                                       ldpc_cuda_cb_result* const* results_d);
int ldpc_cuda_pool_synchronize(ldpc_cuda_pool* pool);

/* ldpc_cuda_dematch_decode_tb_batch over all devices: transport blocks stay whole (a block is assembled and CRC-checked
 * where its codeblocks ran). harq_keys[t] names the HARQ process of block t (any injective code of (cell, rnti, harq_id)):
 * a block whose codeblocks all have new_data = 1 goes to the next device in turn and that device becomes the home of the
 * process; every later block of the process goes to its home, where its soft buffers are. */
int ldpc_cuda_pool_dematch_decode_tb_batch(ldpc_cuda_pool*          pool,
                                           uint32_t                 nof_codeblocks,
                                           const ldpc_cuda_cb_desc* descs,
                                           const int8_t*            rm_llrs,
                                           uint8_t*                 out_bits, /* may be NULL, see the single-device entry */
                                           ldpc_cuda_cb_result*     results,
                                           uint32_t                 nof_tbs,
                                           const ldpc_cuda_tb_desc* tbs,
                                           const uint64_t*          harq_keys,
                                           uint8_t*                 tb_out,
                                           ldpc_cuda_tb_result*     tb_results);

/* ---------------------------------------------------------------------------------------------------------------------
 * Transmit side (also how bench.py builds its synthetic codewords without leaving the device).
 * Bits are unpacked, one per byte, as in the reference's encoder chain; filler bits are 254
 * (include/srsran/phy/upper/channel_coding/ldpc/ldpc.h:113).
 *   encode     = ldpc_encoder::encode (ldpc_encoder.h:41-46, ldpc_encoder_impl.cpp:33-150): bg_K*Z message bits at
 *                in_offset -> N_short*Z codeblock bits at out_offset (the 2Z punctured bits are not part of it).
 *   rate_match = ldpc_rate_matcher::rate_match (ldpc_rate_matcher.h:41-45, ldpc_rate_matcher_impl.cpp:43-200):
 *                N_short*Z codeblock bits at in_offset -> rm_length bits at out_offset (rv, mod_order, Nref, fillers).
 */
int ldpc_cuda_encode_batch_device(ldpc_cuda_handle*        handle,
                                  uint32_t                 nof_codeblocks,
                                  const ldpc_cuda_cb_desc* descs,
                                  const uint8_t*           msg_bits_d,
                                  uint8_t*                 codeblocks_d);
int ldpc_cuda_rate_match_batch_device(ldpc_cuda_handle*        handle,
                                      uint32_t                 nof_codeblocks,
                                      const ldpc_cuda_cb_desc* descs,
                                      const uint8_t*           codeblocks_d,
                                      uint8_t*                 out_bits_d);
int ldpc_cuda_encode(ldpc_cuda_handle* handle, const ldpc_cuda_cb_desc* desc, const uint8_t* msg_bits, uint8_t* codeblock);
int ldpc_cuda_rate_match(ldpc_cuda_handle*        handle,
                         const ldpc_cuda_cb_desc* desc,
                         const uint8_t*           codeblock,
                         uint8_t*                 out_bits);

/* ---------------------------------------------------------------------------------------------------------------------
 * Host-side geometry helpers (no GPU needed).
 */

/* TB -> codeblock geometry of ldpc_segmenter_rx::segment. Fills up to max_cbs descriptors (base_graph, lifting_size,
 * nof_filler_bits, rm_length, mod_order, rv, Nref, crc_poly per pusch_decoder_impl.cpp:44-56, in_offset = cw_offset);
 * nof_crc_bits_out (optional) receives cb_specific.nof_crc_bits. Returns the number of codeblocks C (> 0) or a negative
 * status; when C > max_cbs nothing is written and C is still returned. */
int ldpc_cuda_segment(uint32_t           tbs_bits,
                      int                base_graph,
                      int                rv,
                      int                mod_order,
                      uint32_t           Nref,
                      uint32_t           nof_layers,
                      uint32_t           nof_ch_symbols,
                      ldpc_cuda_cb_desc* descs,
                      uint32_t           max_cbs,
                      uint32_t*          nof_crc_bits_out);

/* Circulant shift of base-graph edge (row m, column n) for lifting size Z, or -1 when there is no edge. */
int ldpc_cuda_graph_shift(int base_graph, int lifting_size, int m, int n);

#ifdef __cplusplus
}
#endif
#endif /* LDPC_CUDA_H */
