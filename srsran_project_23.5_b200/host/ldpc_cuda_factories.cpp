// See ldpc_cuda_factories.h. One ldpc_cuda handle (stream + scratch) per decoder / dematcher object, like one
// reference object per pusch_decoder (not thread-safe, one instance per UL worker thread).
#include "ldpc_cuda_factories.h"

#include "../../include/ldpc_cuda.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h"

#include <cstdio>
#include <cstdlib>

using namespace srsran;

namespace {

// The reference's error convention for invalid arguments is assert + abort (support/srsran_assert.h:60-105); the C ABI
// reports a status instead, which becomes a fatal error here.
void check(int status, const char* what)
{
  if (status != LDPC_CUDA_OK) {
    std::fprintf(stderr, "ldpc_cuda fatal: %s: %s\n", what, ldpc_cuda_status_string(status));
    std::abort();
  }
}

unsigned bg_nodes_short(unsigned bg)
{
  return bg == 2 ? 50 : 66;
}

void fill_common(ldpc_cuda_cb_desc& d, const codeblock_metadata& m)
{
  d                 = {};
  d.base_graph      = (m.tb_common.base_graph == ldpc_base_graph_type::BG2) ? 2 : 1;
  d.lifting_size    = static_cast<uint16_t>(m.tb_common.lifting_size);
  d.nof_filler_bits = static_cast<uint16_t>(m.cb_specific.nof_filler_bits);
}

class ldpc_decoder_cuda : public ldpc_decoder
{
public:
  explicit ldpc_decoder_cuda(ldpc_cuda_handle* h_) : h(h_) {}
  ~ldpc_decoder_cuda() override { ldpc_cuda_destroy(h); }

  // ldpc_decoder interface (ldpc_decoder.h:69-74; reference implementation ldpc_decoder_impl.cpp:60-146).
  optional<unsigned>
  decode(bit_buffer& output, span<const log_likelihood_ratio> input, crc_calculator* crc, const configuration& cfg) override
  {
    ldpc_cuda_cb_desc d;
    fill_common(d, cfg.block_conf);
    d.in_length = static_cast<uint32_t>(input.size());
    d.crc_poly  = LDPC_CUDA_CRC_NONE;
    if (crc != nullptr) {
      switch (crc->get_generator_poly()) {
        case crc_generator_poly::CRC16:
          d.crc_poly = LDPC_CUDA_CRC16;
          break;
        case crc_generator_poly::CRC24A:
          d.crc_poly = LDPC_CUDA_CRC24A;
          break;
        case crc_generator_poly::CRC24B:
          d.crc_poly = LDPC_CUDA_CRC24B;
          break;
        default:
          // ldpc_decoder_impl::init asserts 16 or 24 CRC bits (ldpc_decoder_impl.cpp:50-53)
          check(LDPC_CUDA_ERR_ARG, "decode: CRC polynomial is not CRC16/CRC24A/CRC24B");
      }
    }
    if (cfg.algorithm_conf.max_iterations != alg.max_iterations || cfg.algorithm_conf.scaling_factor != alg.scaling_factor) {
      alg.max_iterations = cfg.algorithm_conf.max_iterations;
      alg.scaling_factor = cfg.algorithm_conf.scaling_factor;
      alg.scale_mode     = LDPC_CUDA_SCALE_GENERIC;  // the parity target is the reference's generic decoder
      check(ldpc_cuda_set_algorithm(h, &alg), "set_algorithm");
    }
    // output.size() == bg_K * Z is asserted by the reference (ldpc_decoder_impl.cpp:69-72)
    const unsigned msg_bits = (d.base_graph == 2 ? 10U : 22U) * d.lifting_size;
    if (output.size() != msg_bits) {
      check(LDPC_CUDA_ERR_ARG, "decode: output size is not bg_K * lifting_size");
    }
    uint32_t iterations = 0;
    check(ldpc_cuda_decode(h,
                           &d,
                           reinterpret_cast<const int8_t*>(input.data()),
                           output.get_buffer().data(),
                           &iterations),
          "ldpc_cuda_decode");
    if (iterations == 0) {
      return {};
    }
    return iterations;
  }

private:
  ldpc_cuda_handle*   h;
  ldpc_cuda_algorithm alg = {0, 0.F, LDPC_CUDA_SCALE_GENERIC, 0};
};

class ldpc_rate_dematcher_cuda : public ldpc_rate_dematcher
{
public:
  explicit ldpc_rate_dematcher_cuda(ldpc_cuda_handle* h_) : h(h_) {}
  ~ldpc_rate_dematcher_cuda() override { ldpc_cuda_destroy(h); }

  // ldpc_rate_dematcher interface (ldpc_rate_dematcher.h:52-55; reference ldpc_rate_dematcher_impl.cpp:43-110).
  void rate_dematch(span<log_likelihood_ratio>       output,
                    span<const log_likelihood_ratio> input,
                    bool                             new_data,
                    const codeblock_metadata&        cfg) override
  {
    ldpc_cuda_cb_desc d;
    fill_common(d, cfg);
    // The reference infers base graph and lifting size from the soft-buffer length (ldpc_rate_dematcher_impl.cpp:78-96).
    const unsigned N = static_cast<unsigned>(output.size());
    if (N % 66 == 0) {
      d.base_graph = 1;
    } else if (N % 50 == 0) {
      d.base_graph = 2;
    } else {
      check(LDPC_CUDA_ERR_ARG, "rate_dematch: soft-buffer length is not 66 Z or 50 Z");
    }
    d.lifting_size = static_cast<uint16_t>(N / bg_nodes_short(d.base_graph));
    d.rv           = static_cast<uint8_t>(cfg.tb_common.rv);
    d.mod_order    = static_cast<uint8_t>(get_bits_per_symbol(cfg.tb_common.mod));
    d.Nref         = cfg.tb_common.Nref;
    d.rm_length    = static_cast<uint32_t>(input.size());
    d.new_data     = new_data ? 1 : 0;
    check(ldpc_cuda_rate_dematch(h,
                                 &d,
                                 reinterpret_cast<int8_t*>(output.data()),
                                 reinterpret_cast<const int8_t*>(input.data())),
          "ldpc_cuda_rate_dematch");
  }

private:
  ldpc_cuda_handle* h;
};

class ldpc_decoder_factory_cuda : public ldpc_decoder_factory
{
public:
  explicit ldpc_decoder_factory_cuda(int device_) : device(device_) {}
  std::unique_ptr<ldpc_decoder> create() override
  {
    ldpc_cuda_handle* h = nullptr;
    if (ldpc_cuda_create(&h, device, 1, 0) != LDPC_CUDA_OK) {
      return nullptr;
    }
    return std::make_unique<ldpc_decoder_cuda>(h);
  }

private:
  int device;
};

class ldpc_rate_dematcher_factory_cuda : public ldpc_rate_dematcher_factory
{
public:
  explicit ldpc_rate_dematcher_factory_cuda(int device_) : device(device_) {}
  std::unique_ptr<ldpc_rate_dematcher> create() override
  {
    ldpc_cuda_handle* h = nullptr;
    if (ldpc_cuda_create(&h, device, 1, 0) != LDPC_CUDA_OK) {
      return nullptr;
    }
    return std::make_unique<ldpc_rate_dematcher_cuda>(h);
  }

private:
  int device;
};

} // namespace

std::shared_ptr<ldpc_decoder_factory> srsran::create_ldpc_decoder_factory_cuda(int device)
{
  if (device < 0 || device >= ldpc_cuda_device_count()) {
    return nullptr;
  }
  return std::make_shared<ldpc_decoder_factory_cuda>(device);
}

std::shared_ptr<ldpc_rate_dematcher_factory> srsran::create_ldpc_rate_dematcher_factory_cuda(int device)
{
  if (device < 0 || device >= ldpc_cuda_device_count()) {
    return nullptr;
  }
  return std::make_shared<ldpc_rate_dematcher_factory_cuda>(device);
}
