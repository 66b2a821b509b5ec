// Single-process scheduler over several devices (include/ldpc_cuda.h, ldpc_cuda_pool_*): one handle and one host thread per
// GPU of the box, a call's codeblocks (or transport blocks) sharded over them, no collective on the data path.
//
// Why it looks like this: codeblocks are independent (the loop body of pusch_decoder_impl::decode,
// lib/phy/upper/channel_processors/pusch_decoder_impl.cpp:156-201, shares nothing between codeblocks but the transport-block
// offsets), so a slot's work splits without any exchange between devices. The reference's unit of concurrency is one
// decoder per uplink thread (apps/gnb/gnb_appconfig.h:367 nof_ul_threads); here it is one worker thread per GPU, each
// driving the ordinary single-device host entry on its share, so that the copy / decode / copy pipelines of all devices run
// at once. Decode-only batches go round-robin by codeblock (i -> device i mod G, SURVEY.md 8d cfg 4) straight out of the
// caller's buffers (pitched copies, no repacking). Transport blocks stay whole: a new block goes to the next device in
// turn, a retransmission to the device that holds the soft buffer of its HARQ process (the table below), and the block is
// assembled and CRC-checked where its codeblocks ran.
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/ldpc_cuda.h"

namespace {

// One worker per device: runs the jobs handed to it in order, on its own thread (cudaSetDevice is per thread inside the
// single-device entries).
class device_worker
{
public:
  device_worker() : th([this] { loop(); }) {}
  ~device_worker()
  {
    {
      std::lock_guard<std::mutex> g(m);
      quit = true;
    }
    cv.notify_all();
    th.join();
  }
  void post(std::function<int()> f)
  {
    {
      std::lock_guard<std::mutex> g(m);
      job     = std::move(f);
      pending = true;
      done    = false;
    }
    cv.notify_all();
  }
  int wait()
  {
    std::unique_lock<std::mutex> g(m);
    cv.wait(g, [this] { return done; });
    return status;
  }

private:
  void loop()
  {
    std::unique_lock<std::mutex> g(m);
    for (;;) {
      cv.wait(g, [this] { return pending || quit; });
      if (quit) {
        return;
      }
      pending                = false;
      std::function<int()> f = std::move(job);
      g.unlock();
      const int rc = f();
      g.lock();
      status = rc;
      done   = true;
      cv.notify_all();
    }
  }

  std::mutex              m;
  std::condition_variable cv;
  std::function<int()>    job;
  bool                    pending = false, done = true, quit = false;
  int                     status = LDPC_CUDA_OK;
  std::thread             th; // last: the thread starts in the constructor
};

struct shard {
  std::vector<ldpc_cuda_cb_desc>   descs;
  std::vector<ldpc_cuda_cb_result> results;
  std::vector<uint32_t>            origin; // index of every local codeblock in the caller's arrays
  std::vector<ldpc_cuda_tb_desc>   tbs;
  std::vector<ldpc_cuda_tb_result> tb_results;
  std::vector<uint32_t>            tb_origin;
};

} // namespace

struct ldpc_cuda_pool {
  std::vector<ldpc_cuda_handle*>              handles;
  std::vector<int>                            devices;
  std::vector<std::unique_ptr<device_worker>> workers;
  std::vector<shard>                          shards;
  std::unordered_map<uint64_t, uint32_t>      home; // HARQ process key -> device index that holds its soft buffers
  uint32_t                                    next_device = 0;
  std::vector<uint64_t>                       codeblocks_done; // per device, since creation
};

namespace {

// Runs job(g) on every device that has work, concurrently; first error wins.
int run_on_devices(ldpc_cuda_pool* p, const std::function<bool(uint32_t)>& has_work, const std::function<int(uint32_t)>& job)
{
  const uint32_t G = static_cast<uint32_t>(p->handles.size());
  for (uint32_t g = 0; g != G; ++g) {
    if (has_work(g)) {
      p->workers[g]->post([&job, g] { return job(g); });
    }
  }
  int rc = LDPC_CUDA_OK;
  for (uint32_t g = 0; g != G; ++g) {
    if (has_work(g)) {
      const int r = p->workers[g]->wait();
      rc          = (rc == LDPC_CUDA_OK) ? r : rc;
    }
  }
  return rc;
}

} // namespace

extern "C" {

int ldpc_cuda_pool_create(ldpc_cuda_pool** pool, const int* devices, uint32_t nof_devices, uint32_t max_codeblocks_per_device,
                          uint32_t flags)
{
  if (pool == nullptr || devices == nullptr || nof_devices == 0 || nof_devices > 64 || max_codeblocks_per_device == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  *pool = nullptr;
  for (uint32_t a = 0; a != nof_devices; ++a) {
    for (uint32_t b = a + 1; b != nof_devices; ++b) {
      if (devices[a] == devices[b]) {
        return LDPC_CUDA_ERR_ARG;
      }
    }
  }
  ldpc_cuda_pool* p = new (std::nothrow) ldpc_cuda_pool;
  if (p == nullptr) {
    return LDPC_CUDA_ERR_CUDA;
  }
  for (uint32_t g = 0; g != nof_devices; ++g) {
    ldpc_cuda_handle* h  = nullptr;
    const int         rc = ldpc_cuda_create(&h, devices[g], max_codeblocks_per_device, flags);
    if (rc != LDPC_CUDA_OK) {
      ldpc_cuda_pool_destroy(p);
      return rc;
    }
    p->handles.push_back(h);
    p->devices.push_back(devices[g]);
  }
  for (uint32_t g = 0; g != nof_devices; ++g) {
    p->workers.emplace_back(new device_worker);
  }
  p->shards.resize(nof_devices);
  p->codeblocks_done.assign(nof_devices, 0);
  *pool = p;
  return LDPC_CUDA_OK;
}

void ldpc_cuda_pool_destroy(ldpc_cuda_pool* p)
{
  if (p == nullptr) {
    return;
  }
  p->workers.clear(); // joins the threads
  for (ldpc_cuda_handle* h : p->handles) {
    ldpc_cuda_destroy(h);
  }
  delete p;
}

uint32_t ldpc_cuda_pool_nof_devices(const ldpc_cuda_pool* p) { return p ? static_cast<uint32_t>(p->handles.size()) : 0; }

ldpc_cuda_handle* ldpc_cuda_pool_handle(ldpc_cuda_pool* p, uint32_t index)
{
  return (p != nullptr && index < p->handles.size()) ? p->handles[index] : nullptr;
}

int ldpc_cuda_pool_set_algorithm(ldpc_cuda_pool* p, const ldpc_cuda_algorithm* alg)
{
  if (p == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  for (ldpc_cuda_handle* h : p->handles) {
    const int rc = ldpc_cuda_set_algorithm(h, alg);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
  }
  return LDPC_CUDA_OK;
}

int ldpc_cuda_pool_softbuffer_reserve(ldpc_cuda_pool* p, uint64_t nof_bytes_per_device)
{
  if (p == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  p->home.clear();
  for (ldpc_cuda_handle* h : p->handles) {
    const int rc = ldpc_cuda_softbuffer_reserve(h, nof_bytes_per_device);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
  }
  return LDPC_CUDA_OK;
}

uint64_t ldpc_cuda_pool_launch_count(const ldpc_cuda_pool* p)
{
  uint64_t n = 0;
  if (p != nullptr) {
    for (const ldpc_cuda_handle* h : p->handles) {
      n += ldpc_cuda_launch_count(h);
    }
  }
  return n;
}

int ldpc_cuda_pool_codeblocks_per_device(const ldpc_cuda_pool* p, uint64_t* counts, uint32_t nof_counts)
{
  if (p == nullptr || counts == nullptr || nof_counts < p->handles.size()) {
    return LDPC_CUDA_ERR_ARG;
  }
  std::copy(p->codeblocks_done.begin(), p->codeblocks_done.end(), counts);
  return LDPC_CUDA_OK;
}

int ldpc_cuda_pool_device_of(const ldpc_cuda_pool* p, uint64_t harq_key)
{
  if (p == nullptr) {
    return -1;
  }
  const auto it = p->home.find(harq_key);
  return it == p->home.end() ? -1 : static_cast<int>(it->second);
}

// ---- decode-only batch: codeblock i -> device i mod G ------------------------------------------------------------------
int ldpc_cuda_pool_decode_batch(ldpc_cuda_pool* p, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* llrs,
                                uint8_t* out_bits, ldpc_cuda_cb_result* results)
{
  if (p == nullptr || descs == nullptr || llrs == nullptr || out_bits == nullptr || results == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  const uint32_t G = static_cast<uint32_t>(p->handles.size());
  // Every device thread picks its own codeblocks (i = g, g + G, ...) out of the caller's list, reads them from, and writes
  // its messages to, the caller's buffers at the caller's offsets, and puts its results where they belong: nothing of the
  // split runs serially on the calling thread.
  const int rc = run_on_devices(
      p, [n](uint32_t g) { return g < n; },
      [p, n, G, descs, llrs, out_bits, results](uint32_t g) {
        shard& s = p->shards[g];
        s.descs.clear();
        for (uint32_t i = g; i < n; i += G) {
          s.descs.push_back(descs[i]);
        }
        s.results.assign(s.descs.size(), ldpc_cuda_cb_result{0, 0, 0});
        const int r = ldpc_cuda_decode_batch(p->handles[g], static_cast<uint32_t>(s.descs.size()), s.descs.data(), llrs,
                                             out_bits, s.results.data());
        if (r == LDPC_CUDA_OK) {
          for (size_t k = 0; k != s.results.size(); ++k) {
            results[g + k * G] = s.results[k];
          }
        }
        return r;
      });
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  for (uint32_t g = 0; g != G && g < n; ++g) {
    p->codeblocks_done[g] += (n - g + G - 1) / G;
  }
  return LDPC_CUDA_OK;
}

// ---- decode-only batch, inputs already in device memory (e.g. written there by the demodulator): kernels only -------------
int ldpc_cuda_pool_decode_batch_device(ldpc_cuda_pool* p, uint32_t n, const ldpc_cuda_cb_desc* descs,
                                       const int8_t* const* llrs_d, uint8_t* const* out_bits_d,
                                       ldpc_cuda_cb_result* const* results_d)
{
  if (p == nullptr || descs == nullptr || llrs_d == nullptr || out_bits_d == nullptr || results_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  const uint32_t G = static_cast<uint32_t>(p->handles.size());
  for (uint32_t g = 0; g != G && g < n; ++g) {
    if (llrs_d[g] == nullptr || out_bits_d[g] == nullptr || results_d[g] == nullptr) {
      return LDPC_CUDA_ERR_ARG;
    }
  }
  const int rc = run_on_devices(
      p, [n](uint32_t g) { return g < n; },
      [p, n, G, descs, llrs_d, out_bits_d, results_d](uint32_t g) {
        shard& s = p->shards[g];  // built by the device's own thread, like everything else of the split
        s.descs.clear();
        for (uint32_t i = g; i < n; i += G) {
          s.descs.push_back(descs[i]);
        }
        return ldpc_cuda_decode_batch_device(p->handles[g], static_cast<uint32_t>(s.descs.size()), s.descs.data(), llrs_d[g],
                                             out_bits_d[g], results_d[g]);
      });
  for (uint32_t g = 0; g != G && g < n && rc == LDPC_CUDA_OK; ++g) {
    p->codeblocks_done[g] += (n - g + G - 1) / G;
  }
  return rc;
}

int ldpc_cuda_pool_synchronize(ldpc_cuda_pool* p)
{
  if (p == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  int rc = LDPC_CUDA_OK;
  for (ldpc_cuda_handle* h : p->handles) {
    const int r = ldpc_cuda_synchronize(h);
    rc          = (rc == LDPC_CUDA_OK) ? r : rc;
  }
  return rc;
}

// ---- fused transport-block batch: whole blocks per device, HARQ affinity ------------------------------------------------
int ldpc_cuda_pool_dematch_decode_tb_batch(ldpc_cuda_pool* p, uint32_t n, const ldpc_cuda_cb_desc* descs,
                                           const int8_t* rm_llrs, uint8_t* out_bits, ldpc_cuda_cb_result* results,
                                           uint32_t nof_tbs, const ldpc_cuda_tb_desc* tbs, const uint64_t* harq_keys,
                                           uint8_t* tb_out, ldpc_cuda_tb_result* tb_results)
{
  if (p == nullptr || descs == nullptr || rm_llrs == nullptr || results == nullptr || tbs == nullptr || harq_keys == nullptr ||
      tb_out == nullptr || tb_results == nullptr || nof_tbs == 0 || n == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  const uint32_t G = static_cast<uint32_t>(p->handles.size());
  for (shard& s : p->shards) {
    s.descs.clear(), s.origin.clear(), s.tbs.clear(), s.tb_origin.clear();
  }
  for (uint32_t t = 0; t != nof_tbs; ++t) {
    const ldpc_cuda_tb_desc& tb = tbs[t];
    if (tb.nof_codeblocks == 0 || tb.first_codeblock >= n || tb.nof_codeblocks > n - tb.first_codeblock) {
      return LDPC_CUDA_ERR_ARG;
    }
    // A block whose codeblocks are all first transmissions starts a process: next device in turn. Anything else goes where
    // the process lives; a retransmission of an unknown process (its first transmission never came through this pool) is
    // treated like a new one - it combines into the zero-initialised arena of its new home, as a fresh reservation of the
    // reference pool would (rx_softbuffer_codeblock_pool.h).
    bool fresh = true;
    for (uint32_t r = 0; r != tb.nof_codeblocks; ++r) {
      fresh = fresh && descs[tb.first_codeblock + r].new_data != 0;
    }
    uint32_t   g  = 0;
    const auto it = p->home.find(harq_keys[t]);
    if (it != p->home.end() && !fresh) {
      g = it->second;
    } else {
      g                     = p->next_device;
      p->next_device        = (p->next_device + 1) % G;
      p->home[harq_keys[t]] = g;
    }
    shard&            s  = p->shards[g];
    ldpc_cuda_tb_desc lt = tb;
    lt.first_codeblock   = static_cast<uint32_t>(s.descs.size());
    s.tbs.push_back(lt);
    s.tb_origin.push_back(t);
    for (uint32_t r = 0; r != tb.nof_codeblocks; ++r) {
      s.descs.push_back(descs[tb.first_codeblock + r]);
      s.origin.push_back(tb.first_codeblock + r);
    }
  }
  for (shard& s : p->shards) {
    s.results.assign(s.descs.size(), ldpc_cuda_cb_result{0, 0, 0});
    s.tb_results.assign(s.tbs.size(), ldpc_cuda_tb_result{0, 0, 0, 0});
  }
  const int rc = run_on_devices(
      p, [p](uint32_t g) { return !p->shards[g].tbs.empty(); },
      [p, rm_llrs, out_bits, tb_out](uint32_t g) {
        shard& s = p->shards[g];
        return ldpc_cuda_dematch_decode_tb_batch(p->handles[g], static_cast<uint32_t>(s.descs.size()), s.descs.data(), rm_llrs,
                                                 out_bits, s.results.data(), static_cast<uint32_t>(s.tbs.size()), s.tbs.data(),
                                                 tb_out, s.tb_results.data());
      });
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  for (uint32_t g = 0; g != G; ++g) {
    const shard& s = p->shards[g];
    for (size_t k = 0; k != s.origin.size(); ++k) {
      results[s.origin[k]] = s.results[k];
    }
    for (size_t k = 0; k != s.tb_origin.size(); ++k) {
      tb_results[s.tb_origin[k]] = s.tb_results[k];
    }
    p->codeblocks_done[g] += s.origin.size();
  }
  return LDPC_CUDA_OK;
}

} // extern "C"
