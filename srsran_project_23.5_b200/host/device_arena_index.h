// Internal to the "cuda" PUSCH decoder (pusch_decoder_cuda.cpp, slot_aggregator.cpp): where the soft bits of a HARQ
// process live in a decoder's PRIVATE device arena, i.e. when the caller uses the reference's host rx_softbuffer_pool
// (lib/phy/upper/rx_softbuffer_pool_impl.h) rather than create_rx_softbuffer_pool_cuda. Not thread safe: the owner locks.
#pragma once

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <unordered_map>

namespace srsran {
namespace cuda_detail {

constexpr unsigned CB_SLOT_BYTES = 66 * 384; // largest soft buffer of one codeblock (ldpc::MAX_CODEBLOCK_SIZE)

inline std::atomic<uint64_t>& private_arena_evictions()
{
  static std::atomic<uint64_t> counter{0};
  return counter;
}

inline void arena_require(bool cond, const char* what)
{
  if (!cond) {
    std::fprintf(stderr, "pusch_decoder_cuda fatal: %s\n", what);
    std::abort();
  }
}

// Device soft-buffer arena in units of one codeblock slot, keyed by the caller's rx_softbuffer object.
class arena_index
{
public:
  explicit arena_index(unsigned nof_slots) { free_ranges[0] = nof_slots; }

  struct entry {
    unsigned first = 0, count = 0;
    uint64_t last_use = 0;
    unsigned in_flight = 0; // transport blocks of this process inside a launch: never chosen as a victim
  };

  // Returns the entry of key; fresh is set when it had to be (re)allocated - its slots then hold foreign data.
  entry& get(const void* key, unsigned count, bool& fresh)
  {
    ++tick;
    auto it = table.find(key);
    if (it != table.end() && it->second.count == count) {
      it->second.last_use = tick;
      fresh               = false;
      return it->second;
    }
    if (it != table.end()) {
      release(it->second);
      table.erase(it);
    }
    unsigned first = 0;
    while (!allocate(count, first)) {
      // evict the least recently used HARQ process (equivalent to a soft-buffer expiry)
      auto victim = table.end();
      for (auto v = table.begin(); v != table.end(); ++v) {
        if (v->second.in_flight == 0 && (victim == table.end() || v->second.last_use < victim->second.last_use)) {
          victim = v;
        }
      }
      arena_require(victim != table.end(), "transport block larger than the device soft-buffer arena");
      // Not silent: the reference's pool refuses the reservation when it is out of entries (rx_softbuffer_impl.h:141-150);
      // this private arena instead drops the oldest HARQ process, whose next retransmission combines into zeros.
      const uint64_t n = private_arena_evictions().fetch_add(1, std::memory_order_relaxed);
      if (n == 0) {
        std::fprintf(stderr,
                     "pusch_decoder_cuda: private soft-buffer arena full, dropping the soft bits of the least recently used "
                     "HARQ process (further drops are counted only: pusch_decoder_cuda_private_arena_evictions()); raise "
                     "arena_codeblocks or use create_rx_softbuffer_pool_cuda\n");
      }
      release(victim->second);
      table.erase(victim);
    }
    entry& e   = table[key];
    e.first    = first;
    e.count    = count;
    e.last_use = tick;
    fresh      = true;
    return e;
  }

  // Brackets a launch that works on the slots of key (slot aggregator: several launches may be in flight).
  void pin(const void* key, int delta)
  {
    auto it = table.find(key);
    if (it != table.end()) {
      it->second.in_flight = static_cast<unsigned>(static_cast<int>(it->second.in_flight) + delta);
    }
  }

private:
  bool allocate(unsigned count, unsigned& first)
  {
    for (auto it = free_ranges.begin(); it != free_ranges.end(); ++it) {
      if (it->second >= count) {
        first              = it->first;
        const unsigned rem = it->second - count;
        free_ranges.erase(it);
        if (rem != 0) {
          free_ranges[first + count] = rem;
        }
        return true;
      }
    }
    return false;
  }

  void release(const entry& e)
  {
    auto it = free_ranges.emplace(e.first, e.count).first;
    auto nx = std::next(it);
    if (nx != free_ranges.end() && it->first + it->second == nx->first) {
      it->second += nx->second;
      free_ranges.erase(nx);
    }
    if (it != free_ranges.begin()) {
      auto pv = std::prev(it);
      if (pv->first + pv->second == it->first) {
        pv->second += it->second;
        free_ranges.erase(it);
      }
    }
  }

  std::unordered_map<const void*, entry> table;
  std::map<unsigned, unsigned>           free_ranges; // first slot -> number of slots
  uint64_t                               tick = 0;
};

} // namespace cuda_detail
} // namespace srsran
