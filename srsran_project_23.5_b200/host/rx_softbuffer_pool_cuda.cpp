// See rx_softbuffer_pool_cuda.h.
#include "rx_softbuffer_pool_cuda.h"

#include "../../include/ldpc_cuda.h"
#include "srsran/adt/bit_buffer.h"
#include "srsran/phy/upper/unique_rx_softbuffer.h"
#include "srsran/ran/slot_point.h"

#include <mutex>
#include <vector>

using namespace srsran;

namespace {

constexpr unsigned CB_SLOT_BYTES    = 66 * 384; // ldpc::MAX_CODEBLOCK_SIZE soft bits, one byte each
constexpr unsigned CB_MESSAGE_BITS  = 22 * 384; // largest message (divide_ceil(max_codeblock_size, 3))
constexpr unsigned UNRESERVED_CB_ID = UINT32_MAX;

// rx_softbuffer_codeblock_pool.h:35-117 with the soft bits on the device: entry i owns arena bytes [i, i + 1) * 25 344.
class codeblock_pool
{
public:
  explicit codeblock_pool(unsigned nof_codeblocks) : reserved(nof_codeblocks, false)
  {
    data_bits.reserve(nof_codeblocks);
    for (unsigned i = 0; i != nof_codeblocks; ++i) {
      data_bits.emplace_back(std::make_unique<dynamic_bit_buffer>(CB_MESSAGE_BITS));
      data_bits.back()->zero();
    }
  }

  unsigned reserve()
  {
    for (unsigned cb_id = 0; cb_id != reserved.size(); ++cb_id) { // first available, like the reference
      if (!reserved[cb_id]) {
        reserved[cb_id] = true;
        return cb_id;
      }
    }
    return UNRESERVED_CB_ID;
  }

  void free(unsigned cb_id)
  {
    if (cb_id != UNRESERVED_CB_ID) {
      reserved[cb_id] = false;
    }
  }

  bit_buffer& get_data_bits(unsigned cb_id) { return *data_bits[cb_id]; }

private:
  std::vector<bool>                                reserved;
  std::vector<std::unique_ptr<dynamic_bit_buffer>> data_bits;
};

// rx_softbuffer_impl.h:34-262.
class softbuffer_cuda : public unique_rx_softbuffer::softbuffer, public rx_softbuffer_device
{
public:
  softbuffer_cuda(codeblock_pool& pool_, void* arena_, uint64_t arena_bytes_) :
    pool(pool_), arena_d(arena_), arena_size(arena_bytes_)
  {
  }

  bool reserve(const rx_softbuffer_identifier& id, const slot_point& expire_slot, unsigned nof_codeblocks)
  {
    std::unique_lock<std::mutex> lock(fsm_mutex);
    if (current_state == state::locked) {
      return false;
    }
    reservation_id          = id;
    reservation_expire_slot = expire_slot;
    if (nof_codeblocks == codeblock_ids.size()) {
      current_state = state::reserved;
      return true;
    }
    free_codeblocks();
    crc.resize(nof_codeblocks); // keeps the flags of the indices that remain, as static_vector::resize does
    codeblock_ids.assign(nof_codeblocks, UNRESERVED_CB_ID);
    for (unsigned& cb_id : codeblock_ids) {
      cb_id = pool.reserve();
      if (cb_id == UNRESERVED_CB_ID) {
        free_codeblocks();
        return false;
      }
    }
    current_state = state::reserved;
    return true;
  }

  void run_slot(const slot_point& slot)
  {
    std::unique_lock<std::mutex> lock(fsm_mutex);
    const bool                   is_released = current_state == state::released;
    const bool is_expired = (current_state == state::reserved) && (reservation_expire_slot <= slot);
    if (is_released || is_expired) {
      free_codeblocks();
    }
  }

  bool match_id(const rx_softbuffer_identifier& identifier) const { return reservation_id == identifier; }

  bool is_available() const
  {
    std::unique_lock<std::mutex> lock(fsm_mutex);
    return (current_state == state::available) || (current_state == state::released);
  }

  // rx_softbuffer
  unsigned get_nof_codeblocks() const override { return static_cast<unsigned>(crc.size()); }
  void     reset_codeblocks_crc() override
  {
    for (unsigned i = 0; i != crc.size(); ++i) {
      crc[i] = false;
    }
  }
  span<bool> get_codeblocks_crc() override { return span<bool>(crc.data(), crc.size()); }
  span<log_likelihood_ratio> get_codeblock_soft_bits(unsigned /*codeblock_id*/, unsigned /*codeblock_size*/) override
  {
    return {}; // the soft bits are device-resident: see rx_softbuffer_device
  }
  bit_buffer get_codeblock_data_bits(unsigned codeblock_id, unsigned data_size) override
  {
    return pool.get_data_bits(codeblock_ids[codeblock_id]).first(data_size);
  }

  // rx_softbuffer_device
  void*    arena() const override { return arena_d; }
  uint64_t arena_bytes() const override { return arena_size; }
  uint64_t soft_offset(unsigned codeblock_id) const override
  {
    return static_cast<uint64_t>(codeblock_ids[codeblock_id]) * CB_SLOT_BYTES;
  }

  // unique_rx_softbuffer::softbuffer
  void lock() override
  {
    std::unique_lock<std::mutex> lock(fsm_mutex);
    current_state = state::locked;
  }
  void unlock() override
  {
    std::unique_lock<std::mutex> lock(fsm_mutex);
    if (current_state == state::locked) {
      current_state = state::reserved;
    }
  }
  void release() override
  {
    std::unique_lock<std::mutex> lock(fsm_mutex);
    current_state = state::released;
  }

private:
  enum class state { available = 0, reserved, locked, released };

  // bool storage that can hand out span<bool> (std::vector<bool> cannot)
  struct flags {
    std::unique_ptr<bool[]> v;
    size_t                  n = 0;
    void                    resize(size_t m)
    {
      std::unique_ptr<bool[]> w(new bool[m]());
      for (size_t i = 0; i != (m < n ? m : n); ++i) {
        w[i] = v[i];
      }
      v = std::move(w);
      n = m;
    }
    size_t size() const { return n; }
    bool*  data() { return v.get(); }
    bool&  operator[](size_t i) { return v[i]; }
  };

  void free_codeblocks()
  {
    for (unsigned cb_id : codeblock_ids) {
      pool.free(cb_id);
    }
    codeblock_ids.clear();
    current_state = state::available;
  }

  mutable std::mutex       fsm_mutex;
  state                    current_state  = state::available;
  rx_softbuffer_identifier reservation_id = {};
  slot_point               reservation_expire_slot;
  codeblock_pool&          pool;
  void*                    arena_d;
  uint64_t                 arena_size;
  flags                    crc;
  std::vector<unsigned>    codeblock_ids;
};

// rx_softbuffer_pool_impl.cpp:27-76.
class rx_softbuffer_pool_cuda : public rx_softbuffer_pool
{
public:
  rx_softbuffer_pool_cuda(const rx_softbuffer_pool_cuda_config& cfg, void* arena_, uint64_t arena_bytes_) :
    device(cfg.device), arena_d(arena_), cb_pool(cfg.max_nof_codeblocks), expire_timeout_slots(cfg.expire_timeout_slots)
  {
    for (unsigned i = 0; i != cfg.max_softbuffers; ++i) {
      buffers.emplace_back(std::make_unique<softbuffer_cuda>(cb_pool, arena_, arena_bytes_));
    }
  }

  ~rx_softbuffer_pool_cuda() override
  {
    buffers.clear();
    if (arena_d != nullptr) {
      ldpc_cuda_arena_destroy(device, arena_d);
    }
  }

  unique_rx_softbuffer
  reserve_softbuffer(const slot_point& slot, const rx_softbuffer_identifier& id, unsigned nof_codeblocks) override
  {
    std::unique_lock<std::mutex> lock(mutex);
    slot_point                   expire_slot = slot + expire_timeout_slots;
    for (auto& buffer : buffers) { // same identifier first
      if (buffer->match_id(id)) {
        if (!buffer->reserve(id, expire_slot, nof_codeblocks)) {
          return unique_rx_softbuffer();
        }
        return unique_rx_softbuffer(*buffer);
      }
    }
    for (auto& buffer : buffers) { // else the first available
      if (buffer->is_available()) {
        if (!buffer->reserve(id, expire_slot, nof_codeblocks)) {
          return unique_rx_softbuffer();
        }
        return unique_rx_softbuffer(*buffer);
      }
    }
    return unique_rx_softbuffer();
  }

  void run_slot(const slot_point& slot) override
  {
    std::unique_lock<std::mutex> lock(mutex);
    for (auto& buffer : buffers) {
      buffer->run_slot(slot);
    }
  }

private:
  int                                           device;
  void*                                         arena_d;
  std::mutex                                    mutex;
  codeblock_pool                                cb_pool;
  std::vector<std::unique_ptr<softbuffer_cuda>> buffers;
  unsigned                                      expire_timeout_slots;
};

} // namespace

std::unique_ptr<rx_softbuffer_pool> srsran::create_rx_softbuffer_pool_cuda(const rx_softbuffer_pool_cuda_config& config)
{
  void*          arena = nullptr;
  const uint64_t bytes = static_cast<uint64_t>(config.max_nof_codeblocks) * CB_SLOT_BYTES;
  if (config.device >= 0) {
    if (config.max_nof_codeblocks == 0 || ldpc_cuda_arena_create(config.device, bytes, &arena) != LDPC_CUDA_OK) {
      return nullptr;
    }
  }
  return std::make_unique<rx_softbuffer_pool_cuda>(config, arena, arena ? bytes : 0);
}
