// See demodulation_mapper_cuda.h.
#include "demodulation_mapper_cuda.h"
#include <vector>

#include "../../include/ldpc_cuda.h"
#include "srsran/ran/modulation_scheme.h"

#include <cstdio>
#include <cstdlib>

using namespace srsran;

namespace {

// Invalid arguments and device errors end like srsran_assert (support/srsran_assert.h:60-105): message + abort.
void check(int status, const char* what)
{
  if (status != LDPC_CUDA_OK) {
    std::fprintf(stderr, "demodulation_mapper_cuda fatal: %s: %s\n", what, ldpc_cuda_status_string(status));
    std::abort();
  }
}

class demodulation_mapper_cuda : public demodulation_mapper
{
public:
  explicit demodulation_mapper_cuda(ldpc_cuda_handle* h_) : h(h_) {}
  ~demodulation_mapper_cuda() override { ldpc_cuda_destroy(h); }

  // demodulation_mapper interface (demodulation_mapper.h:66-69; reference demodulation_mapper_impl.cpp:74-106).
  void demodulate_soft(span<log_likelihood_ratio> llrs,
                       span<const cf_t>           symbols,
                       span<const float>          noise_vars,
                       modulation_scheme          mod) override
  {
    if (symbols.size() != noise_vars.size() || symbols.size() * get_bits_per_symbol(mod) != llrs.size()) {
      check(LDPC_CUDA_ERR_ARG, "demodulate_soft: incompatible lengths");
    }
    check(ldpc_cuda_demodulate_soft(h,
                                    static_cast<int>(get_bits_per_symbol(mod)),
                                    mod == modulation_scheme::PI_2_BPSK ? 1 : 0,
                                    static_cast<uint32_t>(symbols.size()),
                                    reinterpret_cast<const float*>(symbols.data()),
                                    noise_vars.data(),
                                    reinterpret_cast<int8_t*>(llrs.data())),
          "ldpc_cuda_demodulate_soft");
  }

  ldpc_cuda_handle* handle() { return h; }

private:
  ldpc_cuda_handle* h;
};

} // namespace

std::unique_ptr<demodulation_mapper> srsran::create_demodulation_mapper_cuda(int device)
{
  ldpc_cuda_handle* h = nullptr;
  if (device < 0 || ldpc_cuda_device_count() <= device || ldpc_cuda_create(&h, device, 1, 0) != LDPC_CUDA_OK) {
    return nullptr;
  }
  return std::make_unique<demodulation_mapper_cuda>(h);
}

void srsran::descramble_cuda(demodulation_mapper&             handle_owner,
                             span<log_likelihood_ratio>       out,
                             span<const log_likelihood_ratio> in,
                             unsigned                         c_init)
{
  auto* owner = dynamic_cast<demodulation_mapper_cuda*>(&handle_owner);
  if (owner == nullptr || in.size() != out.size()) {
    check(LDPC_CUDA_ERR_ARG, "descramble_cuda: not a cuda mapper or incompatible lengths");
  }
  check(ldpc_cuda_descramble(owner->handle(),
                             c_init,
                             static_cast<uint32_t>(in.size()),
                             reinterpret_cast<const int8_t*>(in.data()),
                             reinterpret_cast<int8_t*>(out.data())),
        "ldpc_cuda_descramble");
}

void srsran::descramble_cuda(demodulation_mapper&                    handle_owner,
                             span<log_likelihood_ratio>              out,
                             span<const log_likelihood_ratio>        in,
                             const pusch_demodulator::configuration& config)
{
  auto* owner = dynamic_cast<demodulation_mapper_cuda*>(&handle_owner);
  if (owner == nullptr || in.size() != out.size()) {
    check(LDPC_CUDA_ERR_ARG, "descramble_cuda: not a cuda mapper or incompatible lengths");
  }
  // The list only exposes its entries through for_each (ulsch_placeholder_list.h:73-95), as soft-bit positions
  // Qm (layers re + layer) + 1; the RE indexes are recovered from the entries of layer 0.
  const unsigned            qm     = get_bits_per_symbol(config.modulation);
  const unsigned            layers = config.nof_tx_layers;
  std::vector<uint16_t>     re_indexes;
  config.placeholders.for_each(config.modulation, layers, [&](unsigned i_bit, unsigned /* nof_x */) {
    const unsigned symbol = (i_bit - 1) / qm;
    if (symbol % layers == 0) {
      re_indexes.push_back(static_cast<uint16_t>(symbol / layers));
    }
  });
  check(ldpc_cuda_descramble_uci(owner->handle(),
                                 config.rnti * 32768U + config.n_id,
                                 static_cast<uint32_t>(in.size()),
                                 static_cast<int>(qm),
                                 layers,
                                 static_cast<uint32_t>(re_indexes.size()),
                                 re_indexes.data(),
                                 reinterpret_cast<const int8_t*>(in.data()),
                                 reinterpret_cast<int8_t*>(out.data())),
        "ldpc_cuda_descramble_uci");
}
