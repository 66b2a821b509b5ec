// TEST INFRASTRUCTURE ONLY - not part of the product, never linked into libsrsran_ldpc_cuda.so.
//
// C-ABI shim (for ctypes) around the UNMODIFIED reference implementation of the PUSCH LDPC receive chain.
// It is compiled together with the reference's own sources where they lie under /root/reference (see
// oracle/Makefile) into oracle/_ref/libsrsran_ref.so and is used
//   * by tests/ to pin the plain-C restatement (oracle/ldpc_oracle.c) and the CUDA path bit-for-bit,
//   * by tools/ to dump the 3GPP base-graph tables and to generate tests/golden fixtures,
//   * by bench.py (--impl reference / cpu_baseline) as the timed multi-threaded CPU baseline.
//
// Every function goes through the reference's public interfaces:
//   create_ldpc_decoder_factory_sw / create_ldpc_rate_dematcher_factory_sw / create_ldpc_encoder_factory_sw /
//   create_ldpc_rate_matcher_factory_sw / create_crc_calculator_factory_sw / create_ldpc_segmenter_*_factory_sw
//   (include/srsran/phy/upper/channel_coding/channel_coding_factories.h:43-105) and pusch_decoder_impl
//   (lib/phy/upper/channel_processors/pusch_decoder_impl.h) with the real rx_softbuffer_pool.

#include "phy/upper/channel_coding/ldpc/ldpc_graph_impl.h"
#include "phy/upper/channel_coding/ldpc/ldpc_luts_impl.h"
#include "phy/upper/channel_processors/pusch_decoder_impl.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include "srsran/phy/upper/channel_processors/channel_processor_factories.h"
#include "srsran/phy/upper/log_likelihood_ratio.h"
#include "srsran/phy/upper/rx_softbuffer_pool.h"
#include "srsran/phy/upper/unique_rx_softbuffer.h"
#include "srsran/ran/slot_point.h"
#include "srsran/srsvec/bit.h"
#include "srsran/support/cpu_features.h"

#ifdef DROPIN_WITH_CUDA_FACTORIES
// Second build of this shim (oracle/_ref/libdropin.so, see oracle/Makefile): the same test entry points, but the
// variant name "cuda" selects the product's factories (srsran_project_23.5_b200/host/ldpc_cuda_factories.h), so that the
// reference's UNMODIFIED pusch_decoder_impl / interfaces run on top of the B200 kernels. Used by tests/test_gpu_dropin.py.
#include "../srsran_project_23.5_b200/host/ldpc_cuda_factories.h"
#include "../srsran_project_23.5_b200/host/pusch_decoder_cuda.h"
#include "../srsran_project_23.5_b200/host/rx_softbuffer_pool_cuda.h"
#include "../srsran_project_23.5_b200/host/demodulation_mapper_cuda.h"
#include "../srsran_project_23.5_b200/host/ulsch_demultiplex_cuda.h"
#endif

#include "phy/upper/channel_modulation/demodulation_mapper_impl.h"
#include "phy/upper/sequence_generators/pseudo_random_generator_impl.h"
#include "srsran/phy/upper/channel_modulation/evm_calculator.h"
#include "srsran/phy/upper/channel_processors/pusch_demodulator.h"
#include "srsran/phy/upper/equalization/channel_equalizer.h"
// pusch_demodulator_impl::descramble (the placeholder-aware descrambler) is a private member: the test shim opens the class
// for this one header, the reference's code itself is untouched.
#define private public
#include "phy/upper/channel_processors/pusch_demodulator_impl.h"
#include "phy/upper/channel_processors/ulsch_demultiplex_impl.h"
#undef private
#include <algorithm>
#include <atomic>
#include <map>
#include <chrono>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

using namespace srsran;

namespace {

std::shared_ptr<ldpc_decoder_factory> decoder_factory_of(const char* type)
{
#ifdef DROPIN_WITH_CUDA_FACTORIES
  if (std::strcmp(type, "cuda") == 0) {
    return create_ldpc_decoder_factory_cuda(0);
  }
#endif
  return create_ldpc_decoder_factory_sw(type);
}

std::shared_ptr<ldpc_rate_dematcher_factory> dematcher_factory_of(const char* type)
{
#ifdef DROPIN_WITH_CUDA_FACTORIES
  if (std::strcmp(type, "cuda") == 0) {
    return create_ldpc_rate_dematcher_factory_cuda(0);
  }
#endif
  return create_ldpc_rate_dematcher_factory_sw(type);
}

ldpc_base_graph_type to_bg(int bg)
{
  return bg == 2 ? ldpc_base_graph_type::BG2 : ldpc_base_graph_type::BG1;
}

// poly codes shared with include/ldpc_cuda.h: 0 = none, 1 = CRC16, 2 = CRC24A, 3 = CRC24B.
bool to_poly(int code, crc_generator_poly& poly)
{
  switch (code) {
    case 1:
      poly = crc_generator_poly::CRC16;
      return true;
    case 2:
      poly = crc_generator_poly::CRC24A;
      return true;
    case 3:
      poly = crc_generator_poly::CRC24B;
      return true;
    default:
      return false;
  }
}

codeblock_metadata make_meta(int bg, int Z, int rv, int Qm, unsigned Nref, unsigned fillers, unsigned crc_bits)
{
  codeblock_metadata m            = {};
  m.tb_common.base_graph          = to_bg(bg);
  m.tb_common.lifting_size        = static_cast<ldpc::lifting_size_t>(Z);
  m.tb_common.rv                  = rv;
  m.tb_common.mod                 = static_cast<modulation_scheme>(Qm);
  m.tb_common.Nref                = Nref;
  m.cb_specific.nof_filler_bits   = fillers;
  m.cb_specific.nof_crc_bits      = crc_bits;
  m.cb_specific.full_length       = (bg == 2 ? 50 : 66) * Z;
  return m;
}

struct decoder_handle {
  std::unique_ptr<ldpc_decoder>   dec;
  std::unique_ptr<crc_calculator> crc[4];
};

// Plain rx_softbuffer for transport blocks with more codeblocks than the reference pool supports (MAX_NOF_SEGMENTS = 52):
// CRC flags and decoded message bits only - enough for a pusch_decoder that keeps its soft bits elsewhere.
class big_softbuffer : public rx_softbuffer
{
public:
  explicit big_softbuffer(unsigned nof_cbs) : crcs(new bool[nof_cbs]()), n(nof_cbs)
  {
    for (unsigned i = 0; i != nof_cbs; ++i) {
      bits.emplace_back(std::make_unique<dynamic_bit_buffer>(22 * 384));
      bits.back()->zero();
    }
  }
  ~big_softbuffer() override { delete[] crcs; }
  unsigned   get_nof_codeblocks() const override { return n; }
  void       reset_codeblocks_crc() override { std::fill(crcs, crcs + n, false); }
  span<bool> get_codeblocks_crc() override { return span<bool>(crcs, n); }
  span<log_likelihood_ratio> get_codeblock_soft_bits(unsigned, unsigned) override { return {}; }
  bit_buffer get_codeblock_data_bits(unsigned cb, unsigned data_size) override { return bits[cb]->first(data_size); }

private:
  bool*                                            crcs;
  unsigned                                         n;
  std::vector<std::unique_ptr<dynamic_bit_buffer>> bits;
};

struct pusch_handle {
  std::unique_ptr<pusch_decoder>                     dec;
  std::unique_ptr<rx_softbuffer_pool>                pool;
  bool                                               batched  = false;
  bool                                               big_pool = false; // the pool itself takes > 52 codeblocks
  std::map<unsigned, std::unique_ptr<big_softbuffer>> big; // key (rnti << 8) | harq, transport blocks above 52 codeblocks
};

} // namespace

extern "C" {

int ref_cpu_supports(const char* type)
{
  std::string t(type);
  if (t == "generic") {
    return 1;
  }
  if (t == "avx2") {
    return cpu_supports_feature(cpu_feature::avx2) ? 1 : 0;
  }
  if (t == "avx512") {
    return (cpu_supports_feature(cpu_feature::avx512f) && cpu_supports_feature(cpu_feature::avx512bw)) ? 1 : 0;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Graphs (ldpc_luts_impl.cpp: get_graph :4521-4544, get_adjacency_matrix; ldpc.h:41-104 lifting sizes).
int ref_lifting_sizes(int* out51)
{
  int n = 0;
  for (auto ls : ldpc::all_lifting_sizes) {
    out51[n++] = static_cast<int>(ls);
  }
  return n;
}

int ref_lifting_index(int Z)
{
  return ldpc::get_lifting_index(static_cast<ldpc::lifting_size_t>(Z));
}

// pcm: [M][N_full] row-major, 0xffff = no edge, else shift % Z.  adj: [M][20], 0xffff-terminated.
int ref_graph_dump(int bg, int Z, uint16_t* pcm, uint16_t* adj)
{
  ldpc::BG_matrix_t                  g = ldpc::get_graph(to_bg(bg), static_cast<ldpc::lifting_size_t>(Z));
  const ldpc::BG_adjacency_matrix_t* a = ldpc::get_adjacency_matrix(to_bg(bg));
  unsigned                           M = (bg == 2) ? ldpc::BG2_M : ldpc::BG1_M;
  unsigned                           N = (bg == 2) ? ldpc::BG2_N_FULL : ldpc::BG1_N_FULL;
  for (unsigned m = 0; m != M; ++m) {
    for (unsigned n = 0; n != N; ++n) {
      pcm[m * N + n] = g[m][n];
    }
    for (unsigned e = 0; e != ldpc::MAX_BG_CHECK_EDGES; ++e) {
      adj[m * ldpc::MAX_BG_CHECK_EDGES + e] = (*a)[m][e];
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// LLR arithmetic (log_likelihood_ratio.cpp:38-85) and hard decision (:225-248).
int ref_llr_add(int a, int b)
{
  return (log_likelihood_ratio(a) + log_likelihood_ratio(b)).to_value_type();
}
int ref_llr_sub(int a, int b)
{
  return (log_likelihood_ratio(a) - log_likelihood_ratio(b)).to_value_type();
}
int ref_llr_promotion_sum(int a, int b)
{
  return log_likelihood_ratio::promotion_sum(log_likelihood_ratio(a), log_likelihood_ratio(b)).to_value_type();
}
int ref_llr_quantize(float v, float range)
{
  return log_likelihood_ratio::quantize(v, range).to_value_type();
}
void ref_hard_decision(const int8_t* soft, unsigned n, uint8_t* out_packed)
{
  dynamic_bit_buffer bits(n);
  hard_decision(bits, span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(soft), n));
  std::memcpy(out_packed, bits.get_buffer().data(), (n + 7) / 8);
}

// ---------------------------------------------------------------------------------------------------------------
// CRC (crc_calculator_lut_impl.cpp / crc_calculator_clmul_impl.cpp). impl in {"auto","lut","clmul"}.
// Returns the checksum of the first nbits of the MSB-first packed buffer, or 0xffffffff on bad arguments.
unsigned ref_crc_packed(const char* impl, int poly_code, const uint8_t* packed, unsigned nbits)
{
  crc_generator_poly poly;
  if (!to_poly(poly_code, poly)) {
    return 0xffffffffU;
  }
  auto factory = create_crc_calculator_factory_sw(impl);
  if (!factory) {
    return 0xffffffffU;
  }
  auto               crc = factory->create(poly);
  dynamic_bit_buffer bits(nbits);
  std::memcpy(bits.get_buffer().data(), packed, (nbits + 7) / 8);
  return crc->calculate(bits);
}

// ---------------------------------------------------------------------------------------------------------------
// Decoder (ldpc_decoder.h:44-74). One handle = one reference decoder object (stateful, not thread-safe).
void* ref_decoder_create(const char* type)
{
  auto factory = decoder_factory_of(type);
  if (!factory) {
    return nullptr;
  }
  auto dec = factory->create();
  if (!dec) {
    return nullptr;
  }
  auto* h       = new decoder_handle;
  h->dec        = std::move(dec);
  auto crc_fact = create_crc_calculator_factory_sw("auto");
  h->crc[1]     = crc_fact->create(crc_generator_poly::CRC16);
  h->crc[2]     = crc_fact->create(crc_generator_poly::CRC24A);
  h->crc[3]     = crc_fact->create(crc_generator_poly::CRC24B);
  return h;
}

void ref_decoder_destroy(void* handle)
{
  delete static_cast<decoder_handle*>(handle);
}

// Returns the number of iterations if the reference returned a value, 0 if it returned nullopt.
// out_packed must hold ceil(K/8) bytes, K = bg_K * Z; it is pre-filled by the caller (the reference leaves it
// untouched on some paths, e.g. all-zero input with a CRC calculator).
int ref_decoder_decode(void*         handle,
                       int           bg,
                       int           Z,
                       unsigned      nof_filler_bits,
                       int           crc_poly_code,
                       unsigned      max_iterations,
                       float         scaling_factor,
                       const int8_t* llrs,
                       unsigned      nof_llrs,
                       uint8_t*      out_packed)
{
  auto*    h = static_cast<decoder_handle*>(handle);
  unsigned K = ((bg == 2) ? 10 : 22) * Z;

  dynamic_bit_buffer out(K);
  std::memcpy(out.get_buffer().data(), out_packed, (K + 7) / 8);

  ldpc_decoder::configuration cfg = {};
  cfg.block_conf = make_meta(bg, Z, 0, 1, 0, nof_filler_bits, (crc_poly_code == 1) ? 16 : 24);
  cfg.algorithm_conf.max_iterations = max_iterations;
  cfg.algorithm_conf.scaling_factor = scaling_factor;

  crc_calculator* crc = (crc_poly_code >= 1 && crc_poly_code <= 3) ? h->crc[crc_poly_code].get() : nullptr;

  optional<unsigned> r = h->dec->decode(
      out, span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(llrs), nof_llrs), crc, cfg);
  std::memcpy(out_packed, out.get_buffer().data(), (K + 7) / 8);
  return r.has_value() ? static_cast<int>(r.value()) : 0;
}

// Multi-threaded batch decode used as the timed CPU baseline (SURVEY.md 8(d) "CPU baseline timing"): nthreads
// workers, one reference decoder object each, disjoint contiguous slices of the batch. Returns wall seconds of the
// decode region only (objects are created before the clock starts). iters[i] = iterations or 0.
double ref_decoder_decode_batch_mt(const char*   type,
                                   unsigned      nthreads,
                                   unsigned      ncb,
                                   int           bg,
                                   int           Z,
                                   unsigned      nof_filler_bits,
                                   int           crc_poly_code,
                                   unsigned      max_iterations,
                                   float         scaling_factor,
                                   const int8_t* llrs,
                                   unsigned      llr_stride,
                                   unsigned      nof_llrs,
                                   uint8_t*      out_packed,
                                   unsigned      out_stride,
                                   int*          iters)
{
  if (nthreads == 0) {
    nthreads = 1;
  }
  std::vector<void*> handles(nthreads);
  for (auto& h : handles) {
    h = ref_decoder_create(type);
    if (h == nullptr) {
      return -1.0;
    }
  }
  std::atomic<unsigned>    ready{0};
  std::atomic<bool>        go{false};
  std::vector<std::thread> workers;
  for (unsigned t = 0; t != nthreads; ++t) {
    workers.emplace_back([&, t]() {
      unsigned begin = static_cast<unsigned>((static_cast<uint64_t>(ncb) * t) / nthreads);
      unsigned end   = static_cast<unsigned>((static_cast<uint64_t>(ncb) * (t + 1)) / nthreads);
      ready.fetch_add(1);
      while (!go.load(std::memory_order_acquire)) {
      }
      for (unsigned i = begin; i != end; ++i) {
        iters[i] = ref_decoder_decode(handles[t],
                                      bg,
                                      Z,
                                      nof_filler_bits,
                                      crc_poly_code,
                                      max_iterations,
                                      scaling_factor,
                                      llrs + static_cast<size_t>(i) * llr_stride,
                                      nof_llrs,
                                      out_packed + static_cast<size_t>(i) * out_stride);
      }
    });
  }
  while (ready.load() != nthreads) {
  }
  auto t0 = std::chrono::steady_clock::now();
  go.store(true, std::memory_order_release);
  for (auto& w : workers) {
    w.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  for (auto& h : handles) {
    ref_decoder_destroy(h);
  }
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---------------------------------------------------------------------------------------------------------------
// Rate dematcher (ldpc_rate_dematcher.h:52-55). out is the persistent in/out soft buffer of N_short * Z LLRs.
void* ref_dematcher_create(const char* type)
{
  auto factory = dematcher_factory_of(type);
  if (!factory) {
    return nullptr;
  }
  auto d = factory->create();
  return d ? d.release() : nullptr;
}

void ref_dematcher_destroy(void* handle)
{
  delete static_cast<ldpc_rate_dematcher*>(handle);
}

void ref_dematcher_run(void*         handle,
                       int8_t*       out,
                       unsigned      out_len,
                       const int8_t* in,
                       unsigned      in_len,
                       int           new_data,
                       int           bg,
                       int           Z,
                       int           rv,
                       int           Qm,
                       unsigned      Nref,
                       unsigned      nof_filler_bits)
{
  auto*              d    = static_cast<ldpc_rate_dematcher*>(handle);
  codeblock_metadata meta = make_meta(bg, Z, rv, Qm, Nref, nof_filler_bits, 24);
  d->rate_dematch(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(out), out_len),
                  span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), in_len),
                  new_data != 0,
                  meta);
}

// What pusch_decoder_impl::decode does per codeblock of a first transmission (pusch_decoder_impl.cpp:156-201): rate dematch
// into the codeblock's soft buffer, then decode with CRC early stop - for a batch of codeblocks of one geometry, nthreads
// workers with one dematcher + one decoder object each on disjoint slices (the reference runs one pusch_decoder per uplink
// thread). Codeblock i: rm_length rate-matched LLRs at rm + i * rm_stride; every worker keeps ONE soft buffer, as new data
// overwrites it. Returns the wall seconds of the region; iters[i] = iterations or 0.
double ref_dematch_decode_batch_mt(const char*   dematcher_type,
                                   const char*   decoder_type,
                                   unsigned      nthreads,
                                   unsigned      ncb,
                                   int           bg,
                                   int           Z,
                                   int           rv,
                                   int           Qm,
                                   unsigned      nof_filler_bits,
                                   int           crc_poly_code,
                                   unsigned      max_iterations,
                                   const int8_t* rm,
                                   unsigned      rm_stride,
                                   unsigned      rm_length,
                                   uint8_t*      out_packed,
                                   unsigned      out_stride,
                                   int*          iters)
{
  nthreads = nthreads == 0 ? 1 : nthreads;
  const unsigned     N = ((bg == 2) ? 50 : 66) * Z;
  std::vector<void*> dec(nthreads), dem(nthreads);
  for (unsigned t = 0; t != nthreads; ++t) {
    dec[t] = ref_decoder_create(decoder_type);
    dem[t] = ref_dematcher_create(dematcher_type);
    if (dec[t] == nullptr || dem[t] == nullptr) {
      return -1.0;
    }
  }
  std::atomic<unsigned>    ready{0};
  std::atomic<bool>        go{false};
  std::vector<std::thread> workers;
  for (unsigned t = 0; t != nthreads; ++t) {
    workers.emplace_back([&, t]() {
      const unsigned      begin = static_cast<unsigned>((static_cast<uint64_t>(ncb) * t) / nthreads);
      const unsigned      end   = static_cast<unsigned>((static_cast<uint64_t>(ncb) * (t + 1)) / nthreads);
      std::vector<int8_t> soft(N, 0);
      ready.fetch_add(1);
      while (!go.load(std::memory_order_acquire)) {
      }
      for (unsigned i = begin; i != end; ++i) {
        ref_dematcher_run(dem[t], soft.data(), N, rm + static_cast<size_t>(i) * rm_stride, rm_length, 1, bg, Z, rv, Qm, 0,
                          nof_filler_bits);
        iters[i] = ref_decoder_decode(dec[t], bg, Z, nof_filler_bits, crc_poly_code, max_iterations, 0.8f, soft.data(), N,
                                      out_packed + static_cast<size_t>(i) * out_stride);
      }
    });
  }
  while (ready.load() != nthreads) {
  }
  const auto t0 = std::chrono::steady_clock::now();
  go.store(true, std::memory_order_release);
  for (auto& w : workers) {
    w.join();
  }
  const auto t1 = std::chrono::steady_clock::now();
  for (unsigned t = 0; t != nthreads; ++t) {
    ref_decoder_destroy(dec[t]);
    ref_dematcher_destroy(dem[t]);
  }
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---------------------------------------------------------------------------------------------------------------
// Transmit side, used only to make test signals: encoder (ldpc_encoder.h) and rate matcher (ldpc_rate_matcher.h).
// Bits are unpacked, one per byte; filler bits in the message are 254 (ldpc.h:113).
int ref_encode(const char* type, int bg, int Z, const uint8_t* msg_bits, uint8_t* out_bits, unsigned out_len)
{
  auto factory = create_ldpc_encoder_factory_sw(type);
  if (!factory) {
    return -1;
  }
  auto enc = factory->create();
  if (!enc) {
    return -1;
  }
  unsigned                               K   = ((bg == 2) ? 10 : 22) * Z;
  codeblock_metadata::tb_common_metadata cfg = make_meta(bg, Z, 0, 1, 0, 0, 24).tb_common;
  enc->encode(span<uint8_t>(out_bits, out_len), span<const uint8_t>(msg_bits, K), cfg);
  return 0;
}

int ref_rate_match(uint8_t*       out,
                   unsigned       E,
                   const uint8_t* in,
                   unsigned       N,
                   int            bg,
                   int            Z,
                   int            rv,
                   int            Qm,
                   unsigned       Nref,
                   unsigned       nof_filler_bits)
{
  auto               rm   = create_ldpc_rate_matcher_factory_sw()->create();
  codeblock_metadata meta = make_meta(bg, Z, rv, Qm, Nref, nof_filler_bits, 24);
  rm->rate_match(span<uint8_t>(out, E), span<const uint8_t>(in, N), meta);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Segmenter, RX geometry (ldpc_segmenter_impl.cpp:226-308). Caller must make sure the TB has <= 52 codeblocks
// (MAX_NOF_SEGMENTS, codeblock_metadata.h:88): beyond that the reference overflows a static_vector.
// meta_out: per CB 6 unsigned {Z, full_length, rm_length, nof_filler_bits, cw_offset, nof_crc_bits}.
int ref_segment_rx(unsigned  tbs,
                   int       bg,
                   int       rv,
                   int       Qm,
                   unsigned  Nref,
                   unsigned  nof_layers,
                   unsigned  nof_ch_symbols,
                   unsigned* meta_out)
{
  if (ldpc::compute_nof_codeblocks(units::bits(tbs), to_bg(bg)) > MAX_NOF_SEGMENTS) {
    return -1;
  }
  auto                              seg = create_ldpc_segmenter_rx_factory_sw()->create();
  std::vector<log_likelihood_ratio> llrs(static_cast<size_t>(nof_ch_symbols) * Qm);
  segmenter_config                  cfg = {};
  cfg.base_graph                        = to_bg(bg);
  cfg.rv                                = rv;
  cfg.mod                               = static_cast<modulation_scheme>(Qm);
  cfg.Nref                              = Nref;
  cfg.nof_layers                        = nof_layers;
  cfg.nof_ch_symbols                    = nof_ch_symbols;
  static_vector<described_rx_codeblock, MAX_NOF_SEGMENTS> cbs;
  seg->segment(cbs, llrs, tbs, cfg);
  for (unsigned i = 0; i != cbs.size(); ++i) {
    const codeblock_metadata& m = cbs[i].second;
    meta_out[6 * i + 0]         = m.tb_common.lifting_size;
    meta_out[6 * i + 1]         = m.cb_specific.full_length;
    meta_out[6 * i + 2]         = m.cb_specific.rm_length;
    meta_out[6 * i + 3]         = m.cb_specific.nof_filler_bits;
    meta_out[6 * i + 4]         = m.cb_specific.cw_offset;
    meta_out[6 * i + 5]         = m.cb_specific.nof_crc_bits;
  }
  return static_cast<int>(cbs.size());
}

// Segmenter, TX side (ldpc_segmenter_impl.cpp:89-222): TB bytes -> C segments of K unpacked bits (fillers = 254),
// TB CRC, CB CRCs and zero padding attached. segs_out: [C][K] bytes. Returns C (<= 52) or -1.
int ref_segment_tx(const uint8_t* tb,
                   unsigned       tb_bytes,
                   int            bg,
                   int            rv,
                   int            Qm,
                   unsigned       Nref,
                   unsigned       nof_layers,
                   unsigned       nof_ch_symbols,
                   uint8_t*       segs_out,
                   unsigned       seg_stride)
{
  if (ldpc::compute_nof_codeblocks(units::bits(tb_bytes * 8), to_bg(bg)) > MAX_NOF_SEGMENTS) {
    return -1;
  }
  auto             seg = create_ldpc_segmenter_tx_factory_sw(create_crc_calculator_factory_sw("auto"))->create();
  segmenter_config cfg = {};
  cfg.base_graph       = to_bg(bg);
  cfg.rv               = rv;
  cfg.mod              = static_cast<modulation_scheme>(Qm);
  cfg.Nref             = Nref;
  cfg.nof_layers       = nof_layers;
  cfg.nof_ch_symbols   = nof_ch_symbols;
  static_vector<described_segment, MAX_NOF_SEGMENTS> segments;
  seg->segment(segments, span<const uint8_t>(tb, tb_bytes), cfg);
  for (unsigned i = 0; i != segments.size(); ++i) {
    const bit_buffer& data = segments[i].get_data();
    unsigned          K    = data.size();
    unsigned          F    = segments[i].get_metadata().cb_specific.nof_filler_bits;
    uint8_t*          dst  = segs_out + static_cast<size_t>(i) * seg_stride;
    for (unsigned b = 0; b != K; ++b) {
      dst[b] = data.extract(b, 1);
    }
    // The TX chain marks filler bits for the encoder (pdsch_encoder_impl.cpp does the same substitution).
    for (unsigned b = K - F; b != K; ++b) {
      dst[b] = ldpc::FILLER_BIT;
    }
  }
  return static_cast<int>(segments.size());
}

// ---------------------------------------------------------------------------------------------------------------
// pusch_decoder_impl + the real soft-buffer pool (pusch_decoder_impl.cpp:121-225, rx_softbuffer_pool_impl.cpp).
void* ref_pusch_create(const char* decoder_type,
                       const char* dematcher_type,
                       unsigned    max_softbuffers,
                       unsigned    max_nof_codeblocks,
                       unsigned    expire_timeout_slots)
{
  auto crc_f = create_crc_calculator_factory_sw("auto");
#ifdef DROPIN_WITH_CUDA_FACTORIES
  if (std::strcmp(decoder_type, "cuda-batched") == 0) {
    // the product's pusch_decoder (one fused batched call per transport block) behind the reference interface
    pusch_decoder_cuda_configuration pc = {};
    pc.crc_factory                      = crc_f;
    pc.arena_codeblocks                 = std::max(256u, max_softbuffers * max_nof_codeblocks);
    auto fac                            = create_pusch_decoder_factory_cuda(pc);
    auto dec                            = fac ? fac->create() : nullptr;
    if (!dec) {
      return nullptr;
    }
    auto* h                      = new pusch_handle;
    h->dec                       = std::move(dec);
    h->batched                   = true;
    if (std::strcmp(dematcher_type, "cuda-pool") == 0) {
      // device-resident HARQ pool (soft bits in one HBM arena shared by the decoders, reference FSM on the host)
      rx_softbuffer_pool_cuda_config cc = {};
      cc.max_softbuffers                = max_softbuffers;
      cc.max_nof_codeblocks             = max_nof_codeblocks;
      cc.expire_timeout_slots           = expire_timeout_slots;
      h->pool                           = create_rx_softbuffer_pool_cuda(cc);
      h->big_pool                       = true;
      if (!h->pool) {
        delete h;
        return nullptr;
      }
      return h;
    }
    rx_softbuffer_pool_config rc = {};
    rc.max_codeblock_size        = ldpc::MAX_CODEBLOCK_SIZE;
    rc.max_softbuffers           = max_softbuffers;
    rc.max_nof_codeblocks        = max_nof_codeblocks;
    rc.expire_timeout_slots      = expire_timeout_slots;
    h->pool                      = create_rx_softbuffer_pool(rc);
    return h;
  }
#endif
  auto dec_f   = decoder_factory_of(decoder_type);
  auto dem_f   = dematcher_factory_of(dematcher_type);
  auto seg_f   = create_ldpc_segmenter_rx_factory_sw();
  auto decoder = dec_f ? dec_f->create() : nullptr;
  auto demat   = dem_f ? dem_f->create() : nullptr;
  if (!decoder || !demat) {
    return nullptr;
  }
  pusch_decoder_impl::sch_crc crcs;
  crcs.crc16  = crc_f->create(crc_generator_poly::CRC16);
  crcs.crc24A = crc_f->create(crc_generator_poly::CRC24A);
  crcs.crc24B = crc_f->create(crc_generator_poly::CRC24B);

  auto* h = new pusch_handle;
  h->dec  = std::make_unique<pusch_decoder_impl>(seg_f->create(), std::move(demat), std::move(decoder), std::move(crcs));
  rx_softbuffer_pool_config pc = {};
  pc.max_codeblock_size        = ldpc::MAX_CODEBLOCK_SIZE;
  pc.max_softbuffers           = max_softbuffers;
  pc.max_nof_codeblocks        = max_nof_codeblocks;
  pc.expire_timeout_slots      = expire_timeout_slots;
  h->pool                      = create_rx_softbuffer_pool(pc);
  return h;
}

void ref_pusch_destroy(void* handle)
{
  delete static_cast<pusch_handle*>(handle);
}

// stats_out: {tb_crc_ok, nof_codeblocks_total, nof_observations, min_iters, max_iters, mean_iters*1000}.
// Returns 0, or -1 if no soft buffer could be reserved. release_on_ok mirrors uplink_processor_impl.cpp:168-170.
int ref_pusch_decode(void*         handle,
                     unsigned      slot_count,
                     unsigned      rnti,
                     unsigned      harq_id,
                     uint8_t*      tb_out,
                     unsigned      tb_bytes,
                     const int8_t* llrs,
                     unsigned      nof_llrs,
                     int           bg,
                     int           rv,
                     int           Qm,
                     unsigned      Nref,
                     unsigned      nof_layers,
                     unsigned      nof_ldpc_iterations,
                     int           use_early_stop,
                     int           new_data,
                     int           release_on_ok,
                     unsigned*     stats_out)
{
  auto*    h       = static_cast<pusch_handle*>(handle);
  unsigned nof_cbs = ldpc::compute_nof_codeblocks(units::bits(tb_bytes * 8), to_bg(bg));
  slot_point               slot(1, slot_count % (1024 * 20));
  rx_softbuffer_identifier id  = {static_cast<uint16_t>(rnti), static_cast<uint8_t>(harq_id)};
  const bool               big = nof_cbs > MAX_NOF_SEGMENTS && !h->big_pool;
  if (big && !h->batched) {
    return -2; // the reference pusch_decoder_impl and pool stop at 52 codeblocks
  }
  unique_rx_softbuffer buf = big ? unique_rx_softbuffer() : h->pool->reserve_softbuffer(slot, id, nof_cbs);
  rx_softbuffer*       sb  = nullptr;
  if (big) {
    auto& slot_sb = h->big[(rnti << 8) | harq_id];
    if (!slot_sb || slot_sb->get_nof_codeblocks() != nof_cbs) {
      slot_sb = std::make_unique<big_softbuffer>(nof_cbs);
    }
    sb = slot_sb.get();
  } else {
    if (!buf.is_valid()) {
      return -1;
    }
    sb = &buf.get();
  }
  pusch_decoder::configuration cfg = {};
  cfg.segmenter_cfg.base_graph     = to_bg(bg);
  cfg.segmenter_cfg.rv             = rv;
  cfg.segmenter_cfg.mod            = static_cast<modulation_scheme>(Qm);
  cfg.segmenter_cfg.Nref           = Nref;
  cfg.segmenter_cfg.nof_layers     = nof_layers;
  cfg.segmenter_cfg.nof_ch_symbols = nof_llrs / Qm;
  cfg.nof_ldpc_iterations          = nof_ldpc_iterations;
  cfg.use_early_stop               = use_early_stop != 0;
  cfg.new_data                     = new_data != 0;

  pusch_decoder_result stats = {};
  h->dec->decode(span<uint8_t>(tb_out, tb_bytes),
                 stats,
                 sb,
                 span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(llrs), nof_llrs),
                 cfg);
  stats_out[0] = stats.tb_crc_ok ? 1 : 0;
  stats_out[1] = stats.nof_codeblocks_total;
  stats_out[2] = stats.ldpc_decoder_stats.get_nof_observations();
  if (stats.ldpc_decoder_stats.get_nof_observations() > 0) {
    stats_out[3] = stats.ldpc_decoder_stats.get_min();
    stats_out[4] = stats.ldpc_decoder_stats.get_max();
    stats_out[5] = static_cast<unsigned>(stats.ldpc_decoder_stats.get_mean() * 1000.0 + 0.5);
  } else {
    stats_out[3] = stats_out[4] = stats_out[5] = 0;
  }
  if (stats.tb_crc_ok && release_on_ok && buf.is_valid()) {
    buf.release();
  }
  h->pool->run_slot(slot);
  return 0;
}

// K pusch_decoder objects of ONE factory on K threads, as uplink_processor_impl.cpp:155-171 runs them (one PUSCH per
// executor thread), sharing one soft-buffer pool. Thread k decodes `reps` transport blocks, each sent nof_tx times
// (rv 0, 2, 3, 1; new_data on the first) as (rnti = k + 1, harq = rep % 8).
//   decoder_type: "generic" | "avx2" | "avx512" (the reference's pusch_decoder_impl) | "cuda-batched" (one launch per
//                 decode()) | "cuda-aggregated" (the factory's slot aggregator, `aggregate_blocks` staging blocks)
//   pool_type:    "sw" (the reference's host pool) | "cuda-pool" (the product's device pool, takes > 52 codeblocks)
//   llrs:         [K][reps][nof_tx][nof_llrs];  tb_out: [K][reps][tb_bytes] (of the last transmission);
//   warm_reps:    untimed repetitions per thread before the clock starts;
//   stats_out:    [K][reps][nof_tx][6] as ref_pusch_decode;  agg_out[8]: launches, transport blocks, max TBs per launch,
//                 then join_wait / stage / follow / drain / launch nanoseconds (pusch_decoder_cuda_aggregator_stats).
// Returns the wall-clock seconds from a common start to the last thread's end, < 0 on failure.
double ref_pusch_concurrent(const char*   decoder_type,
                            const char*   pool_type,
                            unsigned      K,
                            unsigned      reps,
                            unsigned      nof_tx,
                            unsigned      tb_bytes,
                            const int8_t* llrs,
                            unsigned      nof_llrs,
                            int           bg,
                            int           Qm,
                            unsigned      nof_layers,
                            unsigned      nof_ldpc_iterations,
                            int           use_early_stop,
                            unsigned      aggregate_blocks,
                            unsigned      warm_reps,
                            uint8_t*      tb_out,
                            unsigned*     stats_out,
                            uint64_t*     agg_out)
{
  const unsigned nof_cbs = ldpc::compute_nof_codeblocks(units::bits(tb_bytes * 8), to_bg(bg));
  auto           crc_f   = create_crc_calculator_factory_sw("auto");
  std::vector<std::unique_ptr<pusch_decoder>> decs(K);
  std::shared_ptr<pusch_decoder_factory>      cuda_factory;
  const std::string                           type(decoder_type);
  if (type == "cuda-batched" || type == "cuda-aggregated") {
#ifdef DROPIN_WITH_CUDA_FACTORIES
    pusch_decoder_cuda_configuration pc = {};
    pc.crc_factory                      = crc_f;
    pc.arena_codeblocks                 = std::max(256u, K * 8 * nof_cbs);
    pc.max_codeblocks_per_tb            = std::max(256u, nof_cbs);
    if (type == "cuda-aggregated") {
      pc.aggregate_codeblocks = std::max(1024u, K * nof_cbs);
      pc.aggregate_blocks     = aggregate_blocks;
    }
    cuda_factory = create_pusch_decoder_factory_cuda(pc);
    for (auto& d : decs) {
      d = cuda_factory ? cuda_factory->create() : nullptr;
    }
#endif
  } else {
    auto dec_f = decoder_factory_of(decoder_type);
    auto dem_f = dematcher_factory_of(decoder_type);
    auto seg_f = create_ldpc_segmenter_rx_factory_sw();
    for (auto& d : decs) {
      auto decoder = dec_f ? dec_f->create() : nullptr;
      auto demat   = dem_f ? dem_f->create() : nullptr;
      if (!decoder || !demat) {
        return -1.0;
      }
      pusch_decoder_impl::sch_crc crcs;
      crcs.crc16  = crc_f->create(crc_generator_poly::CRC16);
      crcs.crc24A = crc_f->create(crc_generator_poly::CRC24A);
      crcs.crc24B = crc_f->create(crc_generator_poly::CRC24B);
      d = std::make_unique<pusch_decoder_impl>(seg_f->create(), std::move(demat), std::move(decoder), std::move(crcs));
    }
  }
  for (auto& d : decs) {
    if (!d) {
      return -1.0;
    }
  }
  std::unique_ptr<rx_softbuffer_pool> pool;
  if (std::strcmp(pool_type, "cuda-pool") == 0) {
#ifdef DROPIN_WITH_CUDA_FACTORIES
    rx_softbuffer_pool_cuda_config cc = {};
    cc.max_softbuffers                = K * 8;
    cc.max_nof_codeblocks             = K * 8 * nof_cbs;
    cc.expire_timeout_slots           = 100000;
    pool                              = create_rx_softbuffer_pool_cuda(cc);
#endif
  } else {
    if (nof_cbs > MAX_NOF_SEGMENTS) {
      return -2.0;
    }
    rx_softbuffer_pool_config rc = {};
    rc.max_codeblock_size        = ldpc::MAX_CODEBLOCK_SIZE;
    rc.max_softbuffers           = K * 8;
    rc.max_nof_codeblocks        = K * 8 * nof_cbs;
    rc.expire_timeout_slots      = 100000;
    pool                         = create_rx_softbuffer_pool(rc);
  }
  if (!pool) {
    return -1.0;
  }
  static const unsigned rvs[4] = {0, 2, 3, 1};
  std::atomic<unsigned> arrived{0};
  std::atomic<int>      failed{0};
  std::vector<std::chrono::steady_clock::time_point> t_begin(K), t_end(K);
  std::atomic<unsigned> warmed{0};
  auto worker = [&](unsigned k) {
    arrived.fetch_add(1);
    while (arrived.load() < K) { // common start
    }
    // untimed repetitions first (they decode the first transport blocks, which are decoded again below): lazy
    // allocations and the first launches of every handle are out of the way when the clock starts
    for (unsigned w = 0; w != warm_reps + reps; ++w) {
      if (w == warm_reps) {
        warmed.fetch_add(1);
        while (warmed.load() < K && !failed.load()) {
        }
        t_begin[k] = std::chrono::steady_clock::now();
      }
      const unsigned rep = w < warm_reps ? w % reps : w - warm_reps;
      rx_softbuffer_identifier id  = {static_cast<uint16_t>(k + 1), static_cast<uint8_t>(rep % 8)};
      slot_point               slot(1, rep % 10240);
      unique_rx_softbuffer     buf = pool->reserve_softbuffer(slot, id, nof_cbs);
      if (!buf.is_valid()) {
        failed.store(1);
        break;
      }
      bool released = false;
      for (unsigned t = 0; t != nof_tx; ++t) {
        const size_t                 item = (static_cast<size_t>(k) * reps + rep) * nof_tx + t;
        pusch_decoder::configuration cfg  = {};
        cfg.segmenter_cfg.base_graph      = to_bg(bg);
        cfg.segmenter_cfg.rv              = rvs[t % 4];
        cfg.segmenter_cfg.mod             = static_cast<modulation_scheme>(Qm);
        cfg.segmenter_cfg.Nref            = 0;
        cfg.segmenter_cfg.nof_layers      = nof_layers;
        cfg.segmenter_cfg.nof_ch_symbols  = nof_llrs / Qm;
        cfg.nof_ldpc_iterations           = nof_ldpc_iterations;
        cfg.use_early_stop                = use_early_stop != 0;
        cfg.new_data                      = t == 0;
        pusch_decoder_result stats        = {};
        if (released) {
          // the block passed in an earlier transmission and its buffer was released (uplink_processor_impl.cpp:168-170)
          std::memset(stats_out + item * 6, 0xff, 6 * sizeof(unsigned));
          continue;
        }
        decs[k]->decode(span<uint8_t>(tb_out + (static_cast<size_t>(k) * reps + rep) * tb_bytes, tb_bytes),
                        stats,
                        &buf.get(),
                        span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(llrs) + item * nof_llrs,
                                                         nof_llrs),
                        cfg);
        unsigned* so = stats_out + item * 6;
        so[0]        = stats.tb_crc_ok ? 1 : 0;
        so[1]        = stats.nof_codeblocks_total;
        so[2]        = stats.ldpc_decoder_stats.get_nof_observations();
        so[3] = so[4] = so[5] = 0;
        if (so[2] > 0) {
          so[3] = stats.ldpc_decoder_stats.get_min();
          so[4] = stats.ldpc_decoder_stats.get_max();
          so[5] = static_cast<unsigned>(stats.ldpc_decoder_stats.get_mean() * 1000.0 + 0.5);
        }
        if (stats.tb_crc_ok) {
          buf.release();
          released = true;
        }
      }
      if (!released) {
        buf.release();
      }
    }
    t_end[k] = std::chrono::steady_clock::now();
  };
  std::vector<std::thread> threads;
  for (unsigned k = 0; k != K; ++k) {
    threads.emplace_back(worker, k);
  }
  for (auto& t : threads) {
    t.join();
  }
  if (failed.load()) {
    return -3.0;
  }
  std::memset(agg_out, 0, 8 * sizeof(uint64_t));
#ifdef DROPIN_WITH_CUDA_FACTORIES
  if (cuda_factory) {
    const pusch_decoder_cuda_aggregator_stats as = get_pusch_decoder_cuda_aggregator_stats(*cuda_factory);
    agg_out[0] = as.nof_launches, agg_out[1] = as.nof_transport_blocks, agg_out[2] = as.max_transport_blocks_per_launch;
    agg_out[3] = as.join_wait_ns, agg_out[4] = as.stage_ns, agg_out[5] = as.follow_ns, agg_out[6] = as.drain_ns;
    agg_out[7] = as.launch_ns;
  }
#endif
  const auto first = *std::min_element(t_begin.begin(), t_begin.end());
  const auto last  = *std::max_element(t_end.begin(), t_end.end());
  return std::chrono::duration<double>(last - first).count();
}

// ---- soft-buffer pool state machine, reference ("sw") or the product's device pool without device memory ("cuda-host") -----
struct pool_handle {
  std::unique_ptr<rx_softbuffer_pool>                pool;
  std::vector<std::unique_ptr<unique_rx_softbuffer>> held;
};

void* ref_pool_create(const char* type, unsigned max_softbuffers, unsigned max_nof_codeblocks, unsigned expire_timeout_slots)
{
  auto* h = new pool_handle;
  if (std::strcmp(type, "sw") == 0) {
    rx_softbuffer_pool_config rc = {};
    rc.max_codeblock_size        = ldpc::MAX_CODEBLOCK_SIZE;
    rc.max_softbuffers           = max_softbuffers;
    rc.max_nof_codeblocks        = max_nof_codeblocks;
    rc.expire_timeout_slots      = expire_timeout_slots;
    h->pool                      = create_rx_softbuffer_pool(rc);
  }
#ifdef DROPIN_WITH_CUDA_FACTORIES
  if (std::strcmp(type, "cuda-host") == 0 || std::strcmp(type, "cuda") == 0) {
    rx_softbuffer_pool_cuda_config cc = {};
    cc.max_softbuffers                = max_softbuffers;
    cc.max_nof_codeblocks             = max_nof_codeblocks;
    cc.expire_timeout_slots           = expire_timeout_slots;
    cc.device                         = std::strcmp(type, "cuda") == 0 ? 0 : -1;
    h->pool                           = create_rx_softbuffer_pool_cuda(cc);
  }
#endif
  if (!h->pool) {
    delete h;
    return nullptr;
  }
  return h;
}

void ref_pool_destroy(void* handle)
{
  auto* h = static_cast<pool_handle*>(handle);
  h->held.clear();
  delete h;
}

// Returns a token (>= 0) for a valid, locked soft buffer, -1 for an invalid reservation.
int ref_pool_reserve(void* handle, unsigned slot_count, unsigned rnti, unsigned harq_id, unsigned nof_codeblocks)
{
  auto*                    h = static_cast<pool_handle*>(handle);
  slot_point               slot(1, slot_count % (1024 * 20));
  rx_softbuffer_identifier id = {static_cast<uint16_t>(rnti), static_cast<uint8_t>(harq_id)};
  auto buf = std::make_unique<unique_rx_softbuffer>(h->pool->reserve_softbuffer(slot, id, nof_codeblocks));
  if (!buf->is_valid()) {
    return -1;
  }
  for (size_t i = 0; i != h->held.size(); ++i) {
    if (!h->held[i]) {
      h->held[i] = std::move(buf);
      return static_cast<int>(i);
    }
  }
  h->held.push_back(std::move(buf));
  return static_cast<int>(h->held.size() - 1);
}

int ref_pool_nof_codeblocks(void* handle, int token)
{
  return static_cast<int>(static_cast<pool_handle*>(handle)->held[token]->get().get_nof_codeblocks());
}

// value < 0: read the CRC flag of codeblock cb; otherwise set it. Returns the (new) flag.
int ref_pool_crc(void* handle, int token, unsigned cb, int value)
{
  span<bool> crc = static_cast<pool_handle*>(handle)->held[token]->get().get_codeblocks_crc();
  if (value >= 0) {
    crc[cb] = value != 0;
  }
  return crc[cb] ? 1 : 0;
}

// First byte of the stored message of codeblock cb (value < 0: read, else write): data bits persist with the entry.
int ref_pool_data_byte(void* handle, int token, unsigned cb, int value)
{
  bit_buffer bits = static_cast<pool_handle*>(handle)->held[token]->get().get_codeblock_data_bits(cb, 64);
  if (value >= 0) {
    bits.get_buffer()[0] = static_cast<uint8_t>(value);
  }
  return bits.get_buffer()[0];
}

void ref_pool_unlock(void* handle, int token)
{
  static_cast<pool_handle*>(handle)->held[token].reset();
}

void ref_pool_release(void* handle, int token)
{
  auto* h = static_cast<pool_handle*>(handle);
  h->held[token]->release();
  h->held[token].reset();
}

void ref_pool_run_slot(void* handle, unsigned slot_count)
{
  static_cast<pool_handle*>(handle)->pool->run_slot(slot_point(1, slot_count % (1024 * 20)));
}

// Counters of the product's pool: out[6] = reservations, refused_in_use, refused_no_buffer, refused_no_slots, expired,
// released. Returns 0, or -1 for a pool without them (the reference's).
int ref_pool_stats(void* handle, unsigned long long* out)
{
#ifdef DROPIN_WITH_CUDA_FACTORIES
  rx_softbuffer_pool_cuda_stats st = {};
  if (get_rx_softbuffer_pool_cuda_stats(*static_cast<pool_handle*>(handle)->pool, st)) {
    out[0] = st.reservations, out[1] = st.refused_in_use, out[2] = st.refused_no_buffer, out[3] = st.refused_no_slots;
    out[4] = st.expired, out[5] = st.released;
    return 0;
  }
#endif
  (void)handle, (void)out;
  return -1;
}

// Mean time in nanoseconds of reserve_softbuffer(nof_codeblocks) + release + run_slot on a pool whose first `occupied`
// codeblock entries are taken by another reservation (how far a first-fit scan would have to walk).
double ref_pool_time_reserve(void* handle, unsigned occupied, unsigned nof_codeblocks, unsigned repetitions)
{
  auto*      h = static_cast<pool_handle*>(handle);
  slot_point slot(1, 0);
  auto       blocker = h->pool->reserve_softbuffer(slot, {1, 0}, occupied);
  if (occupied != 0 && !blocker.is_valid()) {
    return -1.0;
  }
  const auto t0 = std::chrono::steady_clock::now();
  for (unsigned i = 0; i != repetitions; ++i) {
    auto buf = h->pool->reserve_softbuffer(slot, {2, 0}, nof_codeblocks);
    if (!buf.is_valid()) {
      return -1.0;
    }
    buf.release();
    h->pool->run_slot(slot);
  }
  const auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double, std::nano>(t1 - t0).count() / repetitions;
}


// ---- UL-SCH demultiplexer (ulsch_demultiplex.h:37-103, ulsch_demultiplex_impl.cpp:316-453) ------------------------------
// cfg[9] = mod_order, nof_layers, nof_prb, start_symbol, nof_symbols, dmrs_type (1/2), nof_cdm_groups_without_data,
// dmrs_symbol_mask, nof_harq_ack_rvd.
static ulsch_demultiplex::configuration ulsch_config_of(const unsigned* cfg)
{
  ulsch_demultiplex::configuration c;
  switch (cfg[0]) {
    case 1: c.modulation = modulation_scheme::BPSK; break;
    case 2: c.modulation = modulation_scheme::QPSK; break;
    case 4: c.modulation = modulation_scheme::QAM16; break;
    case 6: c.modulation = modulation_scheme::QAM64; break;
    default: c.modulation = modulation_scheme::QAM256; break;
  }
  c.nof_layers                  = cfg[1];
  c.nof_prb                     = cfg[2];
  c.start_symbol_index          = cfg[3];
  c.nof_symbols                 = cfg[4];
  c.dmrs                        = cfg[5] == 1 ? dmrs_type::TYPE1 : dmrs_type::TYPE2;
  c.nof_cdm_groups_without_data = cfg[6];
  c.dmrs_symbol_mask            = symbol_slot_mask(14);
  for (unsigned l = 0; l != 14; ++l) {
    if ((cfg[7] >> l) & 1u) {
      c.dmrs_symbol_mask.set(l);
    }
  }
  c.nof_harq_ack_rvd = cfg[8];
  return c;
}

static void ulsch_run(ulsch_demultiplex& demux, const unsigned* cfg, const int8_t* in, unsigned in_len, int8_t* sch,
                      unsigned sch_len, int8_t* ack, unsigned ack_len, int8_t* csi1, unsigned csi1_len, int8_t* csi2,
                      unsigned csi2_len)
{
  auto sp = [](int8_t* p, unsigned n) { return span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(p), n); };
  demux.demultiplex(sp(sch, sch_len), sp(ack, ack_len), sp(csi1, csi1_len), sp(csi2, csi2_len),
                    span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), in_len),
                    ulsch_config_of(cfg));
}

// info[6] = O_ack, G_ack, O_csi1, G_csi1, O_csi2, G_csi2. Returns the number of placeholder elements.
static int ulsch_list(ulsch_demultiplex& demux, const unsigned* cfg, const unsigned* info, uint16_t* list, unsigned list_max)
{
  ulsch_demultiplex::message_information mi;
  mi.nof_harq_ack_bits = info[0], mi.nof_enc_harq_ack_bits = info[1];
  mi.nof_csi_part1_bits = info[2], mi.nof_enc_csi_part1_bits = info[3];
  mi.nof_csi_part2_bits = info[4], mi.nof_enc_csi_part2_bits = info[5];
  ulsch_placeholder_list pl = demux.get_placeholders(mi, ulsch_config_of(cfg));
  unsigned               n  = 0;
  pl.for_each(modulation_scheme::QPSK, 1, [&](unsigned i_bit, unsigned) {
    // QPSK, one layer: the callback receives bit index 2 * re + 1 (ulsch_placeholder_list.h:73-95)
    if (n < list_max) {
      list[n] = static_cast<uint16_t>((i_bit - 1) / 2);
    }
    ++n;
  });
  return static_cast<int>(n);
}

void ref_ulsch_demultiplex(const unsigned* cfg, const int8_t* in, unsigned in_len, int8_t* sch, unsigned sch_len, int8_t* ack,
                           unsigned ack_len, int8_t* csi1, unsigned csi1_len, int8_t* csi2, unsigned csi2_len)
{
  ulsch_demultiplex_impl demux;
  ulsch_run(demux, cfg, in, in_len, sch, sch_len, ack, ack_len, csi1, csi1_len, csi2, csi2_len);
}

int ref_ulsch_placeholders(const unsigned* cfg, const unsigned* info, uint16_t* list, unsigned list_max)
{
  ulsch_demultiplex_impl demux;
  return ulsch_list(demux, cfg, info, list, list_max);
}

#ifdef DROPIN_WITH_CUDA_FACTORIES
// The product's "cuda" demultiplexer behind the same reference interface (host/ulsch_demultiplex_cuda.h).
int dropin_ulsch_demultiplex(const unsigned* cfg, const int8_t* in, unsigned in_len, int8_t* sch, unsigned sch_len, int8_t* ack,
                             unsigned ack_len, int8_t* csi1, unsigned csi1_len, int8_t* csi2, unsigned csi2_len)
{
  std::unique_ptr<ulsch_demultiplex> demux = create_ulsch_demultiplex_cuda(0);
  if (!demux) {
    return -1;
  }
  ulsch_run(*demux, cfg, in, in_len, sch, sch_len, ack, ack_len, csi1, csi1_len, csi2, csi2_len);
  return 0;
}

int dropin_ulsch_placeholders(const unsigned* cfg, const unsigned* info, uint16_t* list, unsigned list_max)
{
  std::unique_ptr<ulsch_demultiplex> demux = create_ulsch_demultiplex_cuda(0);
  if (!demux) {
    return -1;
  }
  return ulsch_list(*demux, cfg, info, list, list_max);
}
#endif

// ---- upstream producer of the LLRs: soft demodulator + descrambler ---------------------------------
// demodulation_mapper_impl::demodulate_soft (lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:74-106).
// mod_order 1 = BPSK, 1 + pi2 = pi/2-BPSK, 2/4/6/8 = QPSK/16/64/256QAM. batch > 0 feeds the mapper `batch` symbols at a
// time: below the width of its AVX2 kernels (16/8/16/4 symbols) only the generic per-symbol code runs.
int ref_demodulate_soft(int mod_order, int pi2, unsigned nof_symbols, const float* symbols, const float* noise_vars,
                        int8_t* llrs, unsigned batch)
{
  modulation_scheme mod;
  switch (mod_order) {
    case 1: mod = pi2 ? modulation_scheme::PI_2_BPSK : modulation_scheme::BPSK; break;
    case 2: mod = modulation_scheme::QPSK; break;
    case 4: mod = modulation_scheme::QAM16; break;
    case 6: mod = modulation_scheme::QAM64; break;
    case 8: mod = modulation_scheme::QAM256; break;
    default: return -1;
  }
  // batch == 0xffffffff (drop-in library only): the product's "cuda" mapper behind the same reference interface
  std::unique_ptr<demodulation_mapper> demod_ptr;
#ifdef DROPIN_WITH_CUDA_FACTORIES
  if (batch == 0xffffffffu) {
    demod_ptr = create_demodulation_mapper_cuda(0);
    if (!demod_ptr) {
      return -2;
    }
    batch = 0;
  }
#endif
  if (!demod_ptr) {
    demod_ptr = std::make_unique<demodulation_mapper_impl>();
  }
  demodulation_mapper& demod = *demod_ptr;
  const cf_t* sym = reinterpret_cast<const cf_t*>(symbols);
  log_likelihood_ratio* out = reinterpret_cast<log_likelihood_ratio*>(llrs);
  if (batch == 0) {
    batch = nof_symbols;
  }
  for (unsigned b = 0; b < nof_symbols; b += batch) {
    const unsigned n = std::min(batch, nof_symbols - b);
    demod.demodulate_soft(span<log_likelihood_ratio>(out + (size_t)b * mod_order, (size_t)n * mod_order),
                          span<const cf_t>(sym + b, n), span<const float>(noise_vars + b, n), mod);
  }
  return 0;
}

// pusch_demodulator_impl::descramble without placeholders (pusch_demodulator_impl.cpp:100-152):
// pseudo_random_generator::init(c_init) + apply_xor(span<llr>, span<const llr>).
void ref_descramble(unsigned c_init, unsigned n, const int8_t* in, int8_t* out)
{
  pseudo_random_generator_impl prg;
  prg.init(c_init);
  prg.apply_xor(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(out), n),
                span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), n));
}

namespace {
class unused_equalizer : public channel_equalizer
{
public:
  void equalize(re_list&, noise_var_list&, const re_list&, const ch_est_list&, span<const float>, float) override
  {
    std::abort();
  }
};
} // namespace

// pusch_demodulator_impl::descramble of the reference with UCI placeholders (pusch_demodulator_impl.cpp:100-152).
void ref_descramble_placeholders(unsigned        rnti,
                                 unsigned        n_id,
                                 unsigned        n,
                                 const int8_t*   in,
                                 int8_t*         out,
                                 unsigned        mod_order,
                                 unsigned        nof_layers,
                                 unsigned        nof_re,
                                 const uint16_t* re_indexes)
{
  static pusch_demodulator_impl demod(std::make_unique<unused_equalizer>(),
                                      std::make_unique<demodulation_mapper_impl>(),
                                      nullptr,
                                      std::make_unique<pseudo_random_generator_impl>());
  pusch_demodulator::configuration config = {};
  config.rnti                             = static_cast<uint16_t>(rnti);
  config.n_id                             = n_id;
  config.modulation                       = static_cast<modulation_scheme>(mod_order);
  config.nof_tx_layers                    = nof_layers;
  for (unsigned i = 0; i != nof_re; ++i) {
    config.placeholders.push_back(re_indexes[i]);
  }
  demod.descramble(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(out), n),
                   span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), n),
                   config);
}

#ifdef DROPIN_WITH_CUDA_FACTORIES
// The same call through the product's adaptor descramble_cuda(mapper, out, in, config) (drop-in library only).
int dropin_descramble_placeholders(unsigned        rnti,
                                   unsigned        n_id,
                                   unsigned        n,
                                   const int8_t*   in,
                                   int8_t*         out,
                                   unsigned        mod_order,
                                   unsigned        nof_layers,
                                   unsigned        nof_re,
                                   const uint16_t* re_indexes)
{
  std::unique_ptr<demodulation_mapper> mapper = create_demodulation_mapper_cuda(0);
  if (!mapper) {
    return -2;
  }
  pusch_demodulator::configuration config = {};
  config.rnti                             = static_cast<uint16_t>(rnti);
  config.n_id                             = n_id;
  config.modulation                       = static_cast<modulation_scheme>(mod_order);
  config.nof_tx_layers                    = nof_layers;
  for (unsigned i = 0; i != nof_re; ++i) {
    config.placeholders.push_back(re_indexes[i]);
  }
  descramble_cuda(*mapper,
                  span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(out), n),
                  span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), n),
                  config);
  return 0;
}
#endif

} // extern "C"
