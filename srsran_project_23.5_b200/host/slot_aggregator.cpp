// See slot_aggregator.h.
#include "slot_aggregator.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

using namespace srsran::cuda_detail;

namespace {

constexpr uint64_t ALIGN = 64;

uint64_t round_up(uint64_t v)
{
  return (v + ALIGN - 1) / ALIGN * ALIGN;
}

void check(int status, const char* what)
{
  if (status != LDPC_CUDA_OK) {
    std::fprintf(stderr, "pusch_decoder_cuda (slot aggregator) fatal: %s: %s\n", what, ldpc_cuda_status_string(status));
    std::abort();
  }
}

uint64_t now_ns()
{
  return static_cast<uint64_t>(
      std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count());
}

// Copy into a staging block that the copy engine reads next and the CPU never: non-temporal stores, so that the lines do
// not stay (dirty) in the caches of the many cores that fill one block - the device's reads would have to snoop them out
// (measured on the B200 host: a 22 MB block filled by 16 threads with plain stores copies to the device 2.4x slower
// than the same block filled by one). dst is 16-byte aligned.
void stream_copy(void* dst, const void* src, size_t n)
{
#if defined(__SSE2__)
  auto*       d    = static_cast<__m128i*>(dst);
  const auto* s    = static_cast<const __m128i*>(src);
  size_t      vecs = n / 64;
  for (size_t i = 0; i != vecs; ++i) {
    const __m128i a = _mm_loadu_si128(s + 4 * i), b = _mm_loadu_si128(s + 4 * i + 1);
    const __m128i c = _mm_loadu_si128(s + 4 * i + 2), e = _mm_loadu_si128(s + 4 * i + 3);
    _mm_stream_si128(d + 4 * i, a);
    _mm_stream_si128(d + 4 * i + 1, b);
    _mm_stream_si128(d + 4 * i + 2, c);
    _mm_stream_si128(d + 4 * i + 3, e);
  }
  std::memcpy(static_cast<char*>(dst) + vecs * 64, static_cast<const char*>(src) + vecs * 64, n - vecs * 64);
  _mm_sfence();
#else
  std::memcpy(dst, src, n);
#endif
}

template <typename T>
T* pinned(uint64_t count, uint32_t flags = 0)
{
  void* p = nullptr;
  check(ldpc_cuda_host_alloc(count * sizeof(T), flags, &p), "ldpc_cuda_host_alloc");
  return static_cast<T*>(p);
}

} // namespace

std::shared_ptr<slot_aggregator> slot_aggregator::create(int device_, unsigned max_cbs, unsigned arena_cbs_, unsigned nof_blocks)
{
  if (max_cbs == 0 || nof_blocks < 2 || ldpc_cuda_device_count() <= device_ || device_ < 0) {
    return nullptr;
  }
  std::shared_ptr<slot_aggregator> a(new slot_aggregator);
  a->device    = device_;
  a->cap_cbs   = max_cbs;
  a->arena_cbs = arena_cbs_;
  a->index     = arena_index(arena_cbs_);
  // a codeblock brings at most 8448 * 35 rate-matched LLRs (the ABI's limit), typically 1-3 x 25 344: room for the
  // typical case at full occupancy, and for any single transport block the handle can take
  a->cap_in  = std::max<uint64_t>(static_cast<uint64_t>(max_cbs) * 3 * CB_SLOT_BYTES, 64u << 20);
  a->cap_out = static_cast<uint64_t>(max_cbs) * round_up(1056);
  a->cap_tb  = a->cap_out + 4096;
  a->blocks.resize(nof_blocks);
  for (block& b : a->blocks) {
    if (ldpc_cuda_create(&b.h, device_, max_cbs, 0) != LDPC_CUDA_OK) {
      return nullptr; // the destructor releases what exists
    }
    // no cudaMalloc inside a decode() later on
    check(ldpc_cuda_reserve_staging(b.h, a->cap_in, a->cap_out, a->cap_tb), "ldpc_cuda_reserve_staging");
    b.in      = pinned<int8_t>(a->cap_in);
    b.out     = pinned<uint8_t>(a->cap_out);
    b.tb      = pinned<uint8_t>(a->cap_tb);
    b.descs   = pinned<ldpc_cuda_cb_desc>(max_cbs);
    b.results = pinned<ldpc_cuda_cb_result>(max_cbs);
    b.tbds    = pinned<ldpc_cuda_tb_desc>(max_cbs);
    b.tbrs    = pinned<ldpc_cuda_tb_result>(max_cbs);
  }
  a->blocks[0].state = phase::open;
  a->blocks[0].gen   = a->next_gen++;
  return a;
}

slot_aggregator::~slot_aggregator()
{
  for (block& b : blocks) {
    if (b.h != nullptr) {
      ldpc_cuda_destroy(b.h);
    }
    ldpc_cuda_host_free(b.in);
    ldpc_cuda_host_free(b.out);
    ldpc_cuda_host_free(b.tb);
    ldpc_cuda_host_free(b.descs);
    ldpc_cuda_host_free(b.results);
    ldpc_cuda_host_free(b.tbds);
    ldpc_cuda_host_free(b.tbrs);
  }
  if (private_arena != nullptr) {
    ldpc_cuda_arena_destroy(device, private_arena);
  }
}

slot_aggregator_stats slot_aggregator::stats()
{
  std::lock_guard<std::mutex> lk(m);
  return counters;
}

void slot_aggregator::decode(slot_request& r)
{
  const uint64_t need_in = round_up(r.nof_llrs), need_out = round_up(r.cb_out_bytes), need_tb = round_up(r.tb_out_bytes);
  if (r.nof_cbs == 0 || r.nof_cbs > cap_cbs || need_in > cap_in || need_out > cap_out || need_tb > cap_tb) {
    std::fprintf(stderr, "pusch_decoder_cuda (slot aggregator) fatal: transport block larger than a launch can carry\n");
    std::abort();
  }
  const uint64_t               t0 = now_ns();
  std::unique_lock<std::mutex> lk(m);
  if (r.pool_arena == nullptr && private_arena == nullptr) {
    check(ldpc_cuda_arena_create(device, static_cast<uint64_t>(arena_cbs) * CB_SLOT_BYTES, &private_arena),
          "ldpc_cuda_arena_create");
  }
  void* const    arena       = r.pool_arena != nullptr ? r.pool_arena : private_arena;
  const uint64_t arena_bytes = r.pool_arena != nullptr ? r.pool_arena_bytes : static_cast<uint64_t>(arena_cbs) * CB_SLOT_BYTES;

  // 1. a place in the open block
  block* b = nullptr;
  for (;;) {
    b = &blocks[open_idx];
    const bool same_kind = b->reqs.empty() || (b->arena == arena && b->iterations == r.iterations);
    if (same_kind && b->ncbs + r.nof_cbs <= cap_cbs && b->in_fill + need_in <= cap_in && b->out_fill + need_out <= cap_out &&
        b->tb_fill + need_tb <= cap_tb) {
      break;
    }
    cv.wait(lk); // its members lead it as soon as a block of the ring is free; the next open block is empty
  }
  r.in_off   = b->in_fill;
  r.out_off  = b->out_fill;
  r.tb_off   = b->tb_fill;
  r.first_cb = b->ncbs;
  r.tb_index = static_cast<unsigned>(b->reqs.size());
  r.done     = false;
  b->in_fill += need_in;
  b->out_fill += need_out;
  b->tb_fill += need_tb;
  b->ncbs += r.nof_cbs;
  b->arena       = arena;
  b->arena_bytes = arena_bytes;
  b->iterations  = r.iterations;
  b->reqs.push_back(&r);
  ++b->writers;
  const uint64_t my_gen = b->gen;
  r.zero_slots = false;
  if (r.pool_arena == nullptr) {
    bool        fresh = false;
    const auto& e     = index.get(r.softbuffer_key, r.nof_cbs, fresh);
    r.first_slot      = e.first;
    r.zero_slots      = fresh && !r.new_data; // evicted or never seen: combine into zeros
    index.pin(r.softbuffer_key, +1);
  }

  // 2. fill the slice (callers in parallel)
  lk.unlock();
  const uint64_t t1 = now_ns();
  stream_copy(b->in + r.in_off, r.llrs, r.nof_llrs);
  stream_copy(b->out + r.out_off, r.cb_out, r.cb_out_bytes);
  for (unsigned i = 0; i != r.nof_cbs; ++i) {
    ldpc_cuda_cb_desc d = r.descs[i];
    d.in_offset += r.in_off;
    d.out_offset += r.out_off;
    if (r.pool_arena == nullptr) {
      d.soft_offset = static_cast<uint64_t>(r.first_slot + i) * CB_SLOT_BYTES;
    }
    b->descs[r.first_cb + i] = d;
  }
  ldpc_cuda_tb_desc t = r.tbd;
  t.first_codeblock   = r.first_cb;
  t.tb_offset         = r.tb_off;
  b->tbds[r.tb_index] = t;
  const uint64_t t2 = now_ns();
  lk.lock();
  if (--b->writers == 0) {
    cv.notify_all();
  }

  // 3. lead the block, or sleep until somebody has
  while (!r.done) {
    if (b->state == phase::open && b->gen == my_gen && lead(lk, *b)) {
      continue;
    }
    if (!r.done) {
      cv.wait(lk);
    }
  }

  // 4. take the results out of the staging block
  lk.unlock();
  const uint64_t t3 = now_ns();
  std::memcpy(r.cb_out, b->out + r.out_off, r.cb_out_bytes);
  std::memcpy(r.results, b->results + r.first_cb, sizeof(ldpc_cuda_cb_result) * r.nof_cbs);
  std::memcpy(r.tb_out, b->tb + r.tb_off, r.tb_out_bytes);
  *r.tbr = b->tbrs[r.tb_index];
  const uint64_t t4 = now_ns();
  lk.lock();
  counters.join_wait_ns += t1 - t0;
  counters.stage_ns += t2 - t1;
  counters.follow_ns += t3 - t2;
  counters.drain_ns += t4 - t3;
  if (r.pool_arena == nullptr) {
    index.pin(r.softbuffer_key, -1);
  }
  if (--b->readers == 0) {
    b->state = phase::idle;
    cv.notify_all();
  }
}

// Called with the lock held by a member of the open block b. Returns false (nothing done) when no block of the ring is
// free to take over as the open one: b then stays open and keeps collecting.
bool slot_aggregator::lead(std::unique_lock<std::mutex>& lk, block& b)
{
  unsigned next = 0;
  for (; next != blocks.size() && blocks[next].state != phase::idle; ++next) {
  }
  if (next == blocks.size()) {
    return false;
  }
  block& o   = blocks[next];
  o.state    = phase::open;
  o.gen      = next_gen++;
  o.in_fill = o.out_fill = o.tb_fill = 0;
  o.ncbs                             = 0;
  o.reqs.clear();
  open_idx = next;
  b.state  = phase::running;
  cv.notify_all(); // callers that did not fit
  cv.wait(lk, [&b] { return b.writers == 0; });
  counters.nof_launches += 1;
  counters.nof_transport_blocks += b.reqs.size();
  counters.max_tbs_per_launch = std::max<uint64_t>(counters.max_tbs_per_launch, b.reqs.size());
  lk.unlock();
  const uint64_t t0 = now_ns();
  launch(b);
  const uint64_t t1 = now_ns();
  lk.lock();
  counters.launch_ns += t1 - t0;
  for (slot_request* q : b.reqs) {
    q->done = true;
  }
  b.readers = static_cast<unsigned>(b.reqs.size());
  b.state   = phase::draining;
  cv.notify_all();
  return true;
}

// One batched call for every transport block of the block (pusch_decoder_impl.cpp:156-222 for each of them).
void slot_aggregator::launch(block& b)
{
  if (b.attached != b.arena) {
    check(ldpc_cuda_softbuffer_attach(b.h, b.arena, b.arena_bytes), "ldpc_cuda_softbuffer_attach");
    b.attached = b.arena;
  }
  if (b.set_iterations != b.iterations) {
    ldpc_cuda_algorithm alg = {b.iterations, 0.8f, LDPC_CUDA_SCALE_GENERIC, 0};
    check(ldpc_cuda_set_algorithm(b.h, &alg), "ldpc_cuda_set_algorithm");
    b.set_iterations = b.iterations;
  }
  for (const slot_request* q : b.reqs) {
    if (q->zero_slots) {
      check(ldpc_cuda_softbuffer_zero(b.h, static_cast<uint64_t>(q->first_slot) * CB_SLOT_BYTES,
                                      static_cast<uint64_t>(q->nof_cbs) * CB_SLOT_BYTES),
            "ldpc_cuda_softbuffer_zero");
    }
  }
  check(ldpc_cuda_dematch_decode_tb_batch(b.h, b.ncbs, b.descs, b.in, b.out, b.results, static_cast<uint32_t>(b.reqs.size()),
                                          b.tbds, b.tb, b.tbrs),
        "ldpc_cuda_dematch_decode_tb_batch");
}
