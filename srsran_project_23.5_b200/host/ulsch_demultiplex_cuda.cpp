// See ulsch_demultiplex_cuda.h.
#include "ulsch_demultiplex_cuda.h"

#include "../../include/ldpc_cuda.h"
#include "srsran/ran/modulation_scheme.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace srsran;

namespace {

[[noreturn]] void fatal(const char* what, int status)
{
  std::fprintf(stderr, "ulsch_demultiplex_cuda fatal: %s: %s\n", what, ldpc_cuda_status_string(status));
  std::abort();
}

ldpc_cuda_ulsch_desc to_desc(const ulsch_demultiplex::configuration& c, size_t g_ack, size_t g_csi1, size_t g_csi2)
{
  ldpc_cuda_ulsch_desc d;
  std::memset(&d, 0, sizeof(d));
  d.mod_order                   = static_cast<uint8_t>(get_bits_per_symbol(c.modulation));
  d.nof_layers                  = static_cast<uint8_t>(c.nof_layers);
  d.start_symbol_index          = static_cast<uint8_t>(c.start_symbol_index);
  d.nof_symbols                 = static_cast<uint8_t>(c.nof_symbols);
  d.dmrs_type                   = c.dmrs == dmrs_type::TYPE1 ? 1 : 2;
  d.nof_cdm_groups_without_data = static_cast<uint8_t>(c.nof_cdm_groups_without_data);
  for (unsigned l = 0; l != c.dmrs_symbol_mask.size() && l != 14; ++l) {
    d.dmrs_symbol_mask |= c.dmrs_symbol_mask.test(l) ? static_cast<uint16_t>(1u << l) : 0;
  }
  d.nof_prb                = static_cast<uint16_t>(c.nof_prb);
  d.nof_harq_ack_rvd       = c.nof_harq_ack_rvd;
  d.nof_enc_harq_ack_bits  = static_cast<uint32_t>(g_ack);
  d.nof_enc_csi_part1_bits = static_cast<uint32_t>(g_csi1);
  d.nof_enc_csi_part2_bits = static_cast<uint32_t>(g_csi2);
  return d;
}

int8_t* raw(span<log_likelihood_ratio> s)
{
  return reinterpret_cast<int8_t*>(s.data());
}

class ulsch_demultiplex_cuda : public ulsch_demultiplex
{
public:
  explicit ulsch_demultiplex_cuda(ldpc_cuda_handle* h_) : h(h_) {}
  ~ulsch_demultiplex_cuda() override { ldpc_cuda_destroy(h); }

  void demultiplex(span<log_likelihood_ratio>       sch_data,
                   span<log_likelihood_ratio>       harq_ack,
                   span<log_likelihood_ratio>       csi_part1,
                   span<log_likelihood_ratio>       csi_part2,
                   span<const log_likelihood_ratio> input,
                   const configuration&             config) override
  {
    const ldpc_cuda_ulsch_desc d = to_desc(config, harq_ack.size(), csi_part1.size(), csi_part2.size());
    uint32_t                   in_bits = 0, sch_bits = 0;
    int                        rc      = ldpc_cuda_ulsch_sizes(&d, &in_bits, &sch_bits);
    // the size assertions of the reference (ulsch_demultiplex_impl.cpp:324-330, 386-391)
    if (rc != LDPC_CUDA_OK || in_bits != input.size() || sch_bits != sch_data.size()) {
      fatal("configuration, field sizes and soft-bit counts do not fit", rc != LDPC_CUDA_OK ? rc : LDPC_CUDA_ERR_ARG);
    }
    rc = ldpc_cuda_ulsch_demultiplex(h, &d, reinterpret_cast<const int8_t*>(input.data()), raw(sch_data), raw(harq_ack),
                                     raw(csi_part1), raw(csi_part2));
    if (rc != LDPC_CUDA_OK) {
      fatal("ldpc_cuda_ulsch_demultiplex", rc);
    }
  }

  ulsch_placeholder_list get_placeholders(const message_information& info, const configuration& config) override
  {
    const ldpc_cuda_ulsch_desc d =
        to_desc(config, info.nof_enc_harq_ack_bits, info.nof_enc_csi_part1_bits, info.nof_enc_csi_part2_bits);
    std::vector<uint16_t> re(static_cast<size_t>(ulsch_placeholder_list::MAX_NOF_PLACEHOLDERS));
    const int n = ldpc_cuda_ulsch_placeholders(&d, info.nof_harq_ack_bits, info.nof_csi_part1_bits, info.nof_csi_part2_bits,
                                               re.data(), static_cast<uint32_t>(re.size()));
    if (n < 0 || static_cast<size_t>(n) > re.size()) {
      fatal("ldpc_cuda_ulsch_placeholders", n < 0 ? n : LDPC_CUDA_ERR_CAPACITY);
    }
    ulsch_placeholder_list list;
    for (int i = 0; i != n; ++i) {
      list.push_back(re[i]);
    }
    return list;
  }

private:
  ldpc_cuda_handle* h;
};

} // namespace

std::unique_ptr<ulsch_demultiplex> srsran::create_ulsch_demultiplex_cuda(int device)
{
  ldpc_cuda_handle* h = nullptr;
  if (ldpc_cuda_create(&h, device, 1, 0) != LDPC_CUDA_OK) {
    return nullptr;
  }
  return std::make_unique<ulsch_demultiplex_cuda>(h);
}
