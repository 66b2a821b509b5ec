// Device-resident HARQ soft-buffer pool (see rx_softbuffer_pool_cuda.h).
//
// Design: ONE table of reservations under ONE mutex. A reservation ("ledger line") records who holds it, until when, in
// which phase it is and which codeblock slots of the device arena it owns; the objects handed to the caller are thin
// handles (index + back pointer) without state of their own. Codeblock slots come from a two-level bitmap that always
// yields the LOWEST free slot in O(1) (two count-trailing-zeros), so the placement - and with it the contents a caller
// can observe in re-used slots - equals the first-fit order of rx_softbuffer_codeblock_pool.h:35-117 without its linear
// scan: reserving a 152-codeblock buffer in a 38 912-slot pool takes a few microseconds.
//
// Observable behaviour follows rx_softbuffer_pool_impl.cpp:27-76 / rx_softbuffer_impl.h:112-262 (checked operation by
// operation against the compiled reference in tests/test_softbuffer_pool.py):
//   - reserve: the line that carries the identifier is used if there is one (refused while its holder is active), else the
//     lowest idle or dropped line; a line whose codeblock count is unchanged keeps its slots, flags and stored messages;
//     otherwise it returns its slots and takes new ones, lowest first, all or nothing;
//   - a holder going out of scope puts the line back to "held"; release() marks it "dropped";
//   - run_slot returns the slots of dropped lines and of held lines whose deadline has passed.
// The two state assertions of the reference (lock only from "held", release only from "held" / "active") abort with the
// same message. Expiries, releases and refused reservations are counted (rx_softbuffer_pool_cuda_stats).
#include "rx_softbuffer_pool_cuda.h"

#include "../../include/ldpc_cuda.h"
#include "srsran/adt/bit_buffer.h"
#include "srsran/phy/upper/unique_rx_softbuffer.h"
#include "srsran/ran/slot_point.h"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

using namespace srsran;

namespace {

constexpr unsigned SLOT_BYTES   = 66 * 384; // soft bits of the largest codeblock, one byte each (ldpc::MAX_CODEBLOCK_SIZE)
constexpr unsigned MESSAGE_BITS = 22 * 384; // largest decoded message kept for retransmissions
constexpr unsigned NO_SLOT      = UINT32_MAX;

[[noreturn]] void fatal(const char* what)
{
  std::fprintf(stderr, "rx_softbuffer_pool_cuda: %s\n", what);
  std::fflush(stderr);
  std::abort();
}

// Lowest-free-index allocator over n slots: bit set = free. One summary bit per 64-slot word.
class lowest_free
{
public:
  explicit lowest_free(unsigned n) : leaf((n + 63) / 64, 0), top((leaf.size() + 63) / 64, 0)
  {
    for (unsigned i = 0; i != n; ++i) {
      give(i);
    }
  }

  unsigned take()
  {
    for (size_t t = 0; t != top.size(); ++t) { // one word covers 4096 slots
      if (top[t] == 0) {
        continue;
      }
      const size_t   w = 64 * t + static_cast<size_t>(__builtin_ctzll(top[t]));
      const unsigned b = static_cast<unsigned>(__builtin_ctzll(leaf[w]));
      leaf[w] &= leaf[w] - 1;
      if (leaf[w] == 0) {
        top[t] &= ~(1ull << (w % 64));
      }
      return static_cast<unsigned>(64 * w) + b;
    }
    return NO_SLOT;
  }

  void give(unsigned i)
  {
    leaf[i / 64] |= 1ull << (i % 64);
    top[i / 4096] |= 1ull << ((i / 64) % 64);
  }

private:
  std::vector<uint64_t> leaf, top;
};

enum class phase : uint8_t { idle, held, active, dropped };

struct ledger_line {
  rx_softbuffer_identifier who = {};
  slot_point               deadline;
  phase                    ph = phase::idle;
  std::vector<unsigned>    slots;         // codeblock slots owned (empty after the slots went back)
  std::unique_ptr<bool[]>  crc;           // one flag per codeblock; survives re-sizing for the indices that remain
  unsigned                 nof_flags = 0; // what get_nof_codeblocks() reports (kept when the slots go back)
};

class pool_core;

// What the caller holds: no state, everything lives in the ledger.
class buffer_handle final : public unique_rx_softbuffer::softbuffer, public rx_softbuffer_device
{
public:
  buffer_handle(pool_core& core_, unsigned line_) : core(core_), line(line_) {}

  // rx_softbuffer
  unsigned                   get_nof_codeblocks() const override;
  void                       reset_codeblocks_crc() override;
  span<bool>                 get_codeblocks_crc() override;
  span<log_likelihood_ratio> get_codeblock_soft_bits(unsigned, unsigned) override { return {}; } // device-resident
  bit_buffer                 get_codeblock_data_bits(unsigned codeblock_id, unsigned data_size) override;
  // rx_softbuffer_device
  void*    arena() const override;
  uint64_t arena_bytes() const override;
  uint64_t soft_offset(unsigned codeblock_id) const override;
  // unique_rx_softbuffer::softbuffer
  void lock() override;
  void unlock() override;
  void release() override;

private:
  pool_core& core;
  unsigned   line;
};

class pool_core final : public rx_softbuffer_pool, public rx_softbuffer_pool_cuda_introspection
{
public:
  pool_core(const rx_softbuffer_pool_cuda_config& cfg, void* arena_, uint64_t arena_bytes_) :
    device(cfg.device),
    arena_d(arena_),
    arena_size(arena_bytes_),
    timeout_slots(cfg.expire_timeout_slots),
    free_slots(cfg.max_nof_codeblocks),
    ledger(cfg.max_softbuffers)
  {
    messages.reserve(cfg.max_nof_codeblocks);
    for (unsigned i = 0; i != cfg.max_nof_codeblocks; ++i) {
      messages.emplace_back(std::make_unique<dynamic_bit_buffer>(MESSAGE_BITS));
      messages.back()->zero();
    }
    handles.reserve(cfg.max_softbuffers);
    for (unsigned i = 0; i != cfg.max_softbuffers; ++i) {
      handles.emplace_back(*this, i);
    }
  }

  ~pool_core() override
  {
    if (arena_d != nullptr) {
      ldpc_cuda_arena_destroy(device, arena_d);
    }
  }

  unique_rx_softbuffer
  reserve_softbuffer(const slot_point& slot, const rx_softbuffer_identifier& id, unsigned nof_codeblocks) override
  {
    unsigned pick = NO_SLOT;
    {
      std::lock_guard<std::mutex> guard(mtx);
      // the line that carries the identifier, else the lowest line nobody holds
      for (unsigned i = 0; i != ledger.size() && pick == NO_SLOT; ++i) {
        pick = (ledger[i].who == id) ? i : NO_SLOT;
      }
      for (unsigned i = 0; i != ledger.size() && pick == NO_SLOT; ++i) {
        pick = (ledger[i].ph == phase::idle || ledger[i].ph == phase::dropped) ? i : NO_SLOT;
      }
      if (pick == NO_SLOT) {
        ++counters.refused_no_buffer;
        return unique_rx_softbuffer();
      }
      if (!assign(ledger[pick], id, slot + timeout_slots, nof_codeblocks)) {
        return unique_rx_softbuffer();
      }
      ++counters.reservations;
    }
    return unique_rx_softbuffer(handles[pick]); // locks the line (takes the mutex again)
  }

  void run_slot(const slot_point& slot) override
  {
    std::lock_guard<std::mutex> guard(mtx);
    for (ledger_line& l : ledger) {
      if (l.ph == phase::dropped) {
        ++counters.released;
        give_back(l);
      } else if (l.ph == phase::held && l.deadline <= slot) {
        ++counters.expired;
        give_back(l);
      }
    }
  }

  rx_softbuffer_pool_cuda_stats stats() const override
  {
    std::lock_guard<std::mutex> guard(mtx);
    return counters;
  }

  // -- what the handles forward to -----------------------------------------------------------------------------------
  void activate(unsigned i)
  {
    std::lock_guard<std::mutex> guard(mtx);
    if (ledger[i].ph != phase::held) {
      fatal("It is only possible to reserve in reserved state.");
    }
    ledger[i].ph = phase::active;
  }
  void deactivate(unsigned i)
  {
    std::lock_guard<std::mutex> guard(mtx);
    if (ledger[i].ph == phase::active) {
      ledger[i].ph = phase::held;
    }
  }
  void drop(unsigned i)
  {
    std::lock_guard<std::mutex> guard(mtx);
    if (ledger[i].ph != phase::held && ledger[i].ph != phase::active) {
      fatal("It is only possible to reserve in reserved state.");
    }
    ledger[i].ph = phase::dropped;
  }
  ledger_line& at(unsigned i) { return ledger[i]; }
  bit_buffer&  message(unsigned slot) { return *messages[slot]; }
  void*        arena() const { return arena_d; }
  uint64_t     arena_bytes() const { return arena_size; }

private:
  void give_back(ledger_line& l)
  {
    for (unsigned s : l.slots) {
      free_slots.give(s);
    }
    l.slots.clear();
    l.ph = phase::idle;
  }

  // mtx held. The identifier and the deadline are recorded even when the reservation then fails for lack of slots.
  bool assign(ledger_line& l, const rx_softbuffer_identifier& id, const slot_point& deadline, unsigned n)
  {
    if (l.ph == phase::active) {
      ++counters.refused_in_use;
      return false;
    }
    l.who      = id;
    l.deadline = deadline;
    if (n != l.slots.size()) {
      give_back(l);
      std::unique_ptr<bool[]> flags(new bool[n]());
      for (unsigned i = 0; i != n && i != l.nof_flags; ++i) {
        flags[i] = l.crc[i];
      }
      l.crc       = std::move(flags);
      l.nof_flags = n;
      l.slots.reserve(n);
      for (unsigned i = 0; i != n; ++i) {
        const unsigned s = free_slots.take();
        if (s == NO_SLOT) {
          ++counters.refused_no_slots;
          give_back(l);
          return false;
        }
        l.slots.push_back(s);
      }
    }
    l.ph = phase::held;
    return true;
  }

  int                                              device;
  void*                                            arena_d;
  uint64_t                                         arena_size;
  unsigned                                         timeout_slots;
  mutable std::mutex                               mtx;
  lowest_free                                      free_slots;
  std::vector<ledger_line>                         ledger;
  std::vector<buffer_handle>                       handles;
  std::vector<std::unique_ptr<dynamic_bit_buffer>> messages;
  rx_softbuffer_pool_cuda_stats                    counters = {};
};

unsigned buffer_handle::get_nof_codeblocks() const
{
  return core.at(line).nof_flags;
}
void buffer_handle::reset_codeblocks_crc()
{
  ledger_line& l = core.at(line);
  std::fill(l.crc.get(), l.crc.get() + l.nof_flags, false);
}
span<bool> buffer_handle::get_codeblocks_crc()
{
  ledger_line& l = core.at(line);
  return span<bool>(l.crc.get(), l.nof_flags);
}
bit_buffer buffer_handle::get_codeblock_data_bits(unsigned codeblock_id, unsigned data_size)
{
  ledger_line& l = core.at(line);
  if (codeblock_id >= l.slots.size() || data_size > MESSAGE_BITS) {
    fatal("codeblock index or data size out of range");
  }
  return core.message(l.slots[codeblock_id]).first(data_size);
}
void* buffer_handle::arena() const
{
  return core.arena();
}
uint64_t buffer_handle::arena_bytes() const
{
  return core.arena_bytes();
}
uint64_t buffer_handle::soft_offset(unsigned codeblock_id) const
{
  const ledger_line& l = core.at(line);
  if (codeblock_id >= l.slots.size()) {
    fatal("codeblock index out of range");
  }
  return static_cast<uint64_t>(l.slots[codeblock_id]) * SLOT_BYTES;
}
void buffer_handle::lock()
{
  core.activate(line);
}
void buffer_handle::unlock()
{
  core.deactivate(line);
}
void buffer_handle::release()
{
  core.drop(line);
}

} // namespace

std::unique_ptr<rx_softbuffer_pool> srsran::create_rx_softbuffer_pool_cuda(const rx_softbuffer_pool_cuda_config& config)
{
  void*          arena = nullptr;
  const uint64_t bytes = static_cast<uint64_t>(config.max_nof_codeblocks) * SLOT_BYTES;
  if (config.device >= 0) {
    if (config.max_nof_codeblocks == 0 || ldpc_cuda_arena_create(config.device, bytes, &arena) != LDPC_CUDA_OK) {
      return nullptr;
    }
  }
  return std::make_unique<pool_core>(config, arena, arena ? bytes : 0);
}

bool srsran::get_rx_softbuffer_pool_cuda_stats(const rx_softbuffer_pool& pool, rx_softbuffer_pool_cuda_stats& out)
{
  const auto* p = dynamic_cast<const rx_softbuffer_pool_cuda_introspection*>(&pool);
  if (p == nullptr) {
    return false;
  }
  out = p->stats();
  return true;
}
