// See pusch_decoder_cuda.h.
#include "pusch_decoder_cuda.h"

#include "../../include/ldpc_cuda.h"
#include "device_arena_index.h"
#include "rx_softbuffer_pool_cuda.h"
#include "slot_aggregator.h"
#include "srsran/phy/upper/channel_processors/pusch_decoder.h"
#include "srsran/ran/modulation_scheme.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <unordered_map>
#include <vector>

using namespace srsran;
using srsran::cuda_detail::arena_index;
using srsran::cuda_detail::CB_SLOT_BYTES;

namespace {

constexpr unsigned LONG_CRC_BITS = 24;

void check(int status, const char* what)
{
  if (status != LDPC_CUDA_OK) {
    std::fprintf(stderr, "pusch_decoder_cuda fatal: %s: %s\n", what, ldpc_cuda_status_string(status));
    std::abort();
  }
}

void require(bool cond, const char* what)
{
  if (!cond) {
    std::fprintf(stderr, "pusch_decoder_cuda fatal: %s\n", what);
    std::abort();
  }
}

class pusch_decoder_cuda : public pusch_decoder
{
public:
  /// Either its own handle h_ (one launch per decode()), or - h_ == nullptr - the factory's slot aggregator.
  pusch_decoder_cuda(ldpc_cuda_handle*                             h_,
                     std::shared_ptr<cuda_detail::slot_aggregator> agg_,
                     unsigned                                      arena_cbs_,
                     unsigned                                      max_cbs) :
    h(h_), agg(std::move(agg_)), arena_cbs(arena_cbs_), index(arena_cbs_), descs(max_cbs), results(max_cbs), crc_bits(max_cbs)
  {
  }
  ~pusch_decoder_cuda() override
  {
    if (h != nullptr) {
      ldpc_cuda_destroy(h);
    }
  }

  // pusch_decoder interface (pusch_decoder.h:66-74); reference implementation pusch_decoder_impl.cpp:121-225.
  void decode(span<uint8_t>                    transport_block,
              pusch_decoder_result&            stats,
              rx_softbuffer*                   soft_codeword,
              span<const log_likelihood_ratio> llrs,
              const configuration&             cfg) override
  {
    const unsigned tb_size = static_cast<unsigned>(transport_block.size()) * 8;
    const int      bg      = (cfg.segmenter_cfg.base_graph == ldpc_base_graph_type::BG2) ? 2 : 1;
    const int      Qm      = static_cast<int>(get_bits_per_symbol(cfg.segmenter_cfg.mod));
    // ldpc_segmenter_rx::segment geometry (ldpc_segmenter_impl.cpp:226-308), CRC choice of pusch_decoder_impl.cpp:44-56
    const int C = ldpc_cuda_segment(tb_size,
                                    bg,
                                    static_cast<int>(cfg.segmenter_cfg.rv),
                                    Qm,
                                    cfg.segmenter_cfg.Nref,
                                    cfg.segmenter_cfg.nof_layers,
                                    cfg.segmenter_cfg.nof_ch_symbols,
                                    descs.data(),
                                    static_cast<uint32_t>(descs.size()),
                                    crc_bits.data());
    require(C > 0 && static_cast<size_t>(C) <= descs.size(), "invalid segmentation or too many codeblocks");
    const unsigned nof_cbs = static_cast<unsigned>(C);
    require(nof_cbs == soft_codeword->get_nof_codeblocks(), "wrong number of codeblocks in the soft buffer");
    require(descs[nof_cbs - 1].in_offset + descs[nof_cbs - 1].rm_length <= llrs.size(), "not enough LLRs");

    span<bool> cb_crcs = soft_codeword->get_codeblocks_crc();
    if (cfg.new_data) {
      std::fill(cb_crcs.begin(), cb_crcs.end(), false);
    }

    const unsigned Z = descs[0].lifting_size, K = (bg == 2 ? 10u : 22u) * Z, kb = (K + 7) / 8;
    // Where the soft bits of this HARQ process live: in the device arena of a rx_softbuffer_pool_cuda (shared by every
    // decoder of the device, reservation / expiry handled by the pool), or - with the reference's host pool - in this
    // decoder's private arena, indexed by the rx_softbuffer object with least-recently-used eviction.
    const auto* dev_sb = dynamic_cast<const rx_softbuffer_device*>(soft_codeword);
    unsigned    first_slot = 0;
    if (dev_sb != nullptr && dev_sb->arena() == nullptr) {
      dev_sb = nullptr;
    }
    if (agg) {
      // the aggregator attaches the arena (the pool's, or its own private one shared by the factory's decoders)
    } else if (dev_sb != nullptr) {
      if (attached != dev_sb->arena()) {
        check(ldpc_cuda_softbuffer_attach(h, dev_sb->arena(), dev_sb->arena_bytes()), "ldpc_cuda_softbuffer_attach");
        attached      = dev_sb->arena();
        private_arena = false; // attaching released the decoder's own arena
      }
    } else {
      if (attached != nullptr || !private_arena) {
        check(ldpc_cuda_softbuffer_reserve(h, static_cast<uint64_t>(arena_cbs) * CB_SLOT_BYTES),
              "ldpc_cuda_softbuffer_reserve");
        attached      = nullptr;
        private_arena = true;
        index         = arena_index(arena_cbs);
      }
      bool        fresh = false;
      const auto& slot  = index.get(soft_codeword, nof_cbs, fresh);
      first_slot        = slot.first;
      if (fresh && !cfg.new_data) {
        // the soft bits of this process were evicted (or never seen): combine into zeros
        check(ldpc_cuda_softbuffer_zero(h, static_cast<uint64_t>(slot.first) * CB_SLOT_BYTES,
                                        static_cast<uint64_t>(nof_cbs) * CB_SLOT_BYTES),
              "ldpc_cuda_softbuffer_zero");
      }
    }
    cb_out.resize(static_cast<size_t>(nof_cbs) * kb);
    for (unsigned r = 0; r != nof_cbs; ++r) {
      ldpc_cuda_cb_desc& d = descs[r];
      d.early_stop         = cfg.use_early_stop ? 1 : 0;
      d.new_data           = cfg.new_data ? 1 : 0;
      d.skip_decode        = cb_crcs[r] ? 1 : 0; // decoded in an earlier transmission: combine only (:179-184)
      d.soft_offset        = dev_sb ? dev_sb->soft_offset(r) : static_cast<uint64_t>(first_slot + r) * CB_SLOT_BYTES;
      d.out_offset         = static_cast<uint64_t>(r) * kb;
      // bits a decode does not write (all-zero input, padding of the last byte) keep the stored message's content
      bit_buffer message = soft_codeword->get_codeblock_data_bits(r, K);
      std::memcpy(&cb_out[d.out_offset], message.get_buffer().data(), kb);
    }
    // One call: rate dematch + combine + decode + codeblock CRCs, then - still on the device - the transport block
    // assembly, its CRC24A and the verdict (pusch_decoder_impl.cpp:156-222).
    const unsigned tb_and_crc_size = tb_size + (nof_cbs > 1 ? LONG_CRC_BITS : 0);
    ldpc_cuda_tb_desc tbd = {0, nof_cbs, tb_size, crc_bits[0], 0};
    ldpc_cuda_tb_result tbr = {};
    tb_tmp.assign((tb_and_crc_size + 7) / 8, 0);
    if (agg) {
      // ... issued by the slot aggregator together with the transport blocks of the decoders that are in decode() now
      cuda_detail::slot_request rq;
      rq.nof_cbs          = nof_cbs;
      rq.descs            = descs.data();
      rq.llrs             = reinterpret_cast<const int8_t*>(llrs.data());
      rq.nof_llrs         = descs[nof_cbs - 1].in_offset + descs[nof_cbs - 1].rm_length;
      rq.cb_out           = cb_out.data();
      rq.cb_out_bytes     = cb_out.size();
      rq.results          = results.data();
      rq.tbd              = tbd;
      rq.tb_out           = tb_tmp.data();
      rq.tb_out_bytes     = tb_tmp.size();
      rq.tbr              = &tbr;
      rq.iterations       = cfg.nof_ldpc_iterations;
      rq.new_data         = cfg.new_data;
      rq.softbuffer_key   = soft_codeword;
      rq.pool_arena       = dev_sb ? dev_sb->arena() : nullptr;
      rq.pool_arena_bytes = dev_sb ? dev_sb->arena_bytes() : 0;
      agg->decode(rq);
    } else {
      if (cfg.nof_ldpc_iterations != iterations) {
        ldpc_cuda_algorithm alg = {cfg.nof_ldpc_iterations, 0.8f, LDPC_CUDA_SCALE_GENERIC, 0};
        check(ldpc_cuda_set_algorithm(h, &alg), "ldpc_cuda_set_algorithm");
        iterations = cfg.nof_ldpc_iterations;
      }
      check(ldpc_cuda_dematch_decode_tb_batch(h,
                                              nof_cbs,
                                              descs.data(),
                                              reinterpret_cast<const int8_t*>(llrs.data()),
                                              cb_out.data(),
                                              results.data(),
                                              1,
                                              &tbd,
                                              tb_tmp.data(),
                                              &tbr),
            "ldpc_cuda_dematch_decode_tb_batch");
    }

    // codeblock bookkeeping of the soft buffer (pusch_decoder_impl.cpp:179-195): messages and CRC flags
    stats.nof_codeblocks_total = nof_cbs;
    stats.ldpc_decoder_stats.reset();
    for (unsigned r = 0; r != nof_cbs; ++r) {
      if (!cb_crcs[r]) {
        bit_buffer message = soft_codeword->get_codeblock_data_bits(r, K);
        std::memcpy(message.get_buffer().data(), &cb_out[static_cast<size_t>(r) * kb], kb);
        if (results[r].crc_ok) {
          cb_crcs[r] = true;
          stats.ldpc_decoder_stats.update(results[r].iterations);
        } else {
          stats.ldpc_decoder_stats.update(cfg.nof_ldpc_iterations);
        }
      }
    }
    const bool all_cbs_ok = std::all_of(cb_crcs.begin(), cb_crcs.end(), [](bool a) { return a; });
    require(all_cbs_ok == (tbr.nof_cb_failed == 0), "device and host codeblock CRC flags disagree");

    stats.tb_crc_ok = false;
    if (nof_cbs == 1) {
      stats.tb_crc_ok = cb_crcs[0];
      if (stats.tb_crc_ok) {
        std::memcpy(transport_block.data(), tb_tmp.data(), transport_block.size());
      }
    } else if (all_cbs_ok) {
      std::memcpy(transport_block.data(), tb_tmp.data(), transport_block.size());
      if (tbr.tb_crc_ok) {
        stats.tb_crc_ok = true;
      } else {
        soft_codeword->reset_codeblocks_crc(); // at least one codeblock is a false positive
      }
    }
  }

private:
  ldpc_cuda_handle*                             h;
  std::shared_ptr<cuda_detail::slot_aggregator> agg;
  unsigned                         arena_cbs;
  arena_index                      index;
  const void*                      attached      = nullptr; // arena of a rx_softbuffer_pool_cuda in use
  bool                             private_arena = false;   // the decoder's own arena exists (allocated on first use)
  std::vector<ldpc_cuda_cb_desc>   descs;
  std::vector<ldpc_cuda_cb_result> results;
  std::vector<uint32_t>            crc_bits;
  std::vector<uint8_t>             cb_out;
  std::vector<uint8_t>             tb_tmp;
  unsigned                         iterations = 0;
};

class pusch_decoder_factory_cuda : public pusch_decoder_factory
{
public:
  explicit pusch_decoder_factory_cuda(const pusch_decoder_cuda_configuration& cfg_) : cfg(cfg_) {}

  std::unique_ptr<pusch_decoder> create() override
  {
    // the transport block CRC24A is computed on the device: cfg.crc_factory is not used by this decoder
    if (cfg.aggregate_codeblocks != 0) {
      std::lock_guard<std::mutex> lk(m);
      if (!agg) {
        agg = cuda_detail::slot_aggregator::create(cfg.device, std::max(cfg.aggregate_codeblocks, cfg.max_codeblocks_per_tb),
                                                   cfg.arena_codeblocks, cfg.aggregate_blocks);
      }
      if (!agg) {
        return nullptr;
      }
      return std::make_unique<pusch_decoder_cuda>(nullptr, agg, cfg.arena_codeblocks, cfg.max_codeblocks_per_tb);
    }
    ldpc_cuda_handle* h = nullptr;
    if (ldpc_cuda_create(&h, cfg.device, cfg.max_codeblocks_per_tb, 0) != LDPC_CUDA_OK) {
      return nullptr;
    }
    return std::make_unique<pusch_decoder_cuda>(h, nullptr, cfg.arena_codeblocks, cfg.max_codeblocks_per_tb);
  }

  cuda_detail::slot_aggregator_stats aggregator_stats()
  {
    std::lock_guard<std::mutex> lk(m);
    return agg ? agg->stats() : cuda_detail::slot_aggregator_stats();
  }

private:
  pusch_decoder_cuda_configuration              cfg;
  std::mutex                                    m;
  std::shared_ptr<cuda_detail::slot_aggregator> agg;
};

} // namespace

std::shared_ptr<pusch_decoder_factory>
srsran::create_pusch_decoder_factory_cuda(const pusch_decoder_cuda_configuration& config)
{
  if (!config.crc_factory || ldpc_cuda_device_count() <= config.device || config.device < 0 ||
      config.arena_codeblocks < config.max_codeblocks_per_tb) {
    return nullptr;
  }
  return std::make_shared<pusch_decoder_factory_cuda>(config);
}

pusch_decoder_cuda_aggregator_stats srsran::get_pusch_decoder_cuda_aggregator_stats(pusch_decoder_factory& factory)
{
  pusch_decoder_cuda_aggregator_stats out = {};
  if (auto* f = dynamic_cast<pusch_decoder_factory_cuda*>(&factory)) {
    const cuda_detail::slot_aggregator_stats s = f->aggregator_stats();
    out.nof_launches                           = s.nof_launches;
    out.nof_transport_blocks                   = s.nof_transport_blocks;
    out.max_transport_blocks_per_launch        = s.max_tbs_per_launch;
    out.join_wait_ns = s.join_wait_ns, out.stage_ns = s.stage_ns, out.follow_ns = s.follow_ns, out.drain_ns = s.drain_ns;
    out.launch_ns = s.launch_ns;
  }
  return out;
}

uint64_t srsran::pusch_decoder_cuda_private_arena_evictions()
{
  return cuda_detail::private_arena_evictions().load(std::memory_order_relaxed);
}
