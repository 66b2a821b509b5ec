// Staging of PAGEABLE caller memory for the host entries (ldpc_api.cu).
//
// cudaMemcpyAsync from pageable memory goes through the driver's bounce buffers on the calling thread: 9 GB/s on the B200
// host against 48 GB/s from page-locked memory (bench.py e2e: 3.0 against 16.1 Gb/s). The reference's callers hand over
// pageable memory (pusch_processor_impl keeps its LLRs in ordinary containers), so the host entries stage such inputs
// themselves: a few worker threads copy chunk c + 1 into one of three page-locked bounce buffers with non-temporal stores
// (the lines must not stay dirty in the workers' caches, where the copy engine would have to snoop them out one by one:
// profiles/r02_pusch_callers.md) while the copy engine moves chunk c from another.
#pragma once

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <cuda_runtime.h>

namespace ldpc_b200 {

// Non-temporal on x86 (any alignment: the bytes in front of the first 16-byte boundary of dst are copied plainly), plain
// memcpy elsewhere.
inline void stream_copy(void* dst, const void* src, size_t n)
{
#if defined(__SSE2__)
  const size_t head = std::min<size_t>(n, (16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
  std::memcpy(dst, src, head);
  dst = static_cast<char*>(dst) + head, src = static_cast<const char*>(src) + head, n -= head;
  auto*        d    = static_cast<__m128i*>(dst);
  const auto*  s    = static_cast<const __m128i*>(src);
  const size_t vecs = n / 64;
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

class HostStager
{
public:
  static constexpr unsigned NOF_BUFFERS = 3;

  HostStager()
  {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned n  = std::min(8u, std::max(1u, hw / 2));
    for (unsigned i = 0; i + 1 < n; ++i) {  // the calling thread is the n-th worker
      workers.emplace_back([this] { run(); });
    }
  }

  ~HostStager()
  {
    {
      std::lock_guard<std::mutex> lk(m);
      quit = true;
    }
    cv.notify_all();
    for (std::thread& t : workers) {
      t.join();
    }
    for (unsigned i = 0; i != NOF_BUFFERS; ++i) {
      cudaFreeHost(buf[i]);
      if (ev[i]) {
        cudaEventDestroy(ev[i]);
      }
    }
  }

  HostStager(const HostStager&)            = delete;
  HostStager& operator=(const HostStager&) = delete;

  // A bounce buffer of at least `bytes` that no copy reads any more. Blocks until the copy engine has finished with the
  // buffer's previous contents. nullptr on a CUDA error.
  void* acquire(size_t bytes, unsigned& slot)
  {
    slot = next++ % NOF_BUFFERS;
    if (ev[slot] == nullptr && cudaEventCreateWithFlags(&ev[slot], cudaEventDisableTiming) != cudaSuccess) {
      return nullptr;
    }
    if (used[slot] && cudaEventSynchronize(ev[slot]) != cudaSuccess) {
      return nullptr;
    }
    if (bytes > cap[slot]) {
      cudaFreeHost(buf[slot]);
      buf[slot] = nullptr, cap[slot] = 0;
      const size_t want = bytes + (bytes >> 3) + 4096;
      if (cudaHostAlloc(&buf[slot], want, cudaHostAllocPortable) != cudaSuccess) {
        return nullptr;
      }
      cap[slot] = want;
    }
    return buf[slot];
  }

  // To be called after the asynchronous copy out of the buffer has been enqueued on `stream`.
  cudaError_t release(unsigned slot, cudaStream_t stream)
  {
    used[slot] = true;
    return cudaEventRecord(ev[slot], stream);
  }

  // dst (page-locked, 16-byte aligned) <- src (pageable), split over the workers and the calling thread.
  void copy(void* dst, const void* src, size_t bytes)
  {
    const size_t nparts = workers.size() + 1;
    const size_t part   = ((bytes + nparts - 1) / nparts + 4095) & ~static_cast<size_t>(4095);
    if (workers.empty() || bytes < (1u << 20)) {
      stream_copy(dst, src, bytes);
      return;
    }
    {
      std::lock_guard<std::mutex> lk(m);
      job_dst = static_cast<char*>(dst), job_src = static_cast<const char*>(src);
      job_bytes = bytes, job_part = part;
      job_next = 1;  // part 0 is the caller's
      job_parts = (bytes + part - 1) / part;
      job_done  = 0;
      ++job_id;
    }
    cv.notify_all();
    stream_copy(dst, src, std::min(part, bytes));
    std::unique_lock<std::mutex> lk(m);
    ++job_done;
    // take parts the workers have not picked up yet, then wait for the ones in flight
    while (job_next < job_parts) {
      const size_t k = job_next++;
      lk.unlock();
      do_part(k);
      lk.lock();
      ++job_done;
    }
    done_cv.wait(lk, [this] { return job_done == job_parts; });
  }

private:
  void do_part(size_t k)
  {
    const size_t off = k * job_part;
    stream_copy(job_dst + off, job_src + off, std::min(job_part, job_bytes - off));
  }

  void run()
  {
    uint64_t                     seen = 0;
    std::unique_lock<std::mutex> lk(m);
    for (;;) {
      cv.wait(lk, [&] { return quit || (job_id != seen && job_next < job_parts); });
      if (quit) {
        return;
      }
      while (job_next < job_parts) {
        const size_t k = job_next++;
        lk.unlock();
        do_part(k);
        lk.lock();
        if (++job_done == job_parts) {
          done_cv.notify_all();
        }
      }
      seen = job_id;
    }
  }

  std::vector<std::thread> workers;
  std::mutex               m;
  std::condition_variable  cv, done_cv;
  bool                     quit = false;
  char*                    job_dst   = nullptr;
  const char*              job_src   = nullptr;
  size_t                   job_bytes = 0, job_part = 0, job_next = 0, job_parts = 0, job_done = 0;
  uint64_t                 job_id = 0;

  void*       buf[NOF_BUFFERS]  = {};
  size_t      cap[NOF_BUFFERS]  = {};
  cudaEvent_t ev[NOF_BUFFERS]   = {};
  bool        used[NOF_BUFFERS] = {};
  unsigned    next              = 0;
};

// true for host memory the CUDA runtime does not know (neither cudaHostAlloc'ed nor cudaHostRegister'ed).
inline bool is_pageable(const void* p)
{
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;  // unknown: let the runtime deal with it
  }
  return a.type == cudaMemoryTypeUnregistered;
}

}  // namespace ldpc_b200
