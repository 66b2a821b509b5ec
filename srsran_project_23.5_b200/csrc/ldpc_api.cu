// C ABI of the B200 LDPC receive chain (include/ldpc_cuda.h): handle life cycle, batching/grouping, launches.
// No CPU fallback: every compute entry point needs an sm_100 device and fails with LDPC_CUDA_ERR_CUDA otherwise.
#include <algorithm>
#include <cstdio>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <mutex>
#include <new>
#include <vector>

#include "ldpc_host_util.h"
#include "ldpc_host_stager.h"
#include "ldpc_kernels_scalar.cuh"
#include "ldpc_dematch.cuh"
#include "ldpc_kernels_tmem.cuh"  // needs c_crc / crc_warp from the scalar header
#include "ldpc_encode.cuh"
#include "ldpc_tb.cuh"
#include "ldpc_demod.cuh"
#include "ldpc_ulsch.cuh"
#include <cmath>

using namespace ldpc_b200;

#define CUDA_TRY(expr)                                                                                       \
  do {                                                                                                       \
    cudaError_t err__ = (expr);                                                                              \
    if (err__ != cudaSuccess) {                                                                              \
      std::fprintf(stderr, "ldpc_cuda: %s failed: %s (%s:%d)\n", #expr, cudaGetErrorString(err__), __FILE__, \
                   __LINE__);                                                                                \
      return LDPC_CUDA_ERR_CUDA;                                                                             \
    }                                                                                                        \
  } while (0)

#ifndef LDPC_MAX_CHUNKS
#define LDPC_MAX_CHUNKS 32  // e2e 15.4-15.5 (16 chunks) -> 15.9 Gb/s (32) -> 15.6 (48): tools/ab_e2e.sh
#endif
constexpr unsigned MAX_CHUNKS = LDPC_MAX_CHUNKS;
static_assert(sizeof(DemodJob) == sizeof(ldpc_cuda_demod_desc) && sizeof(DemodJob) == 32, "DemodJob mirrors the ABI struct");
static_assert(sizeof(ldpc_cuda_placeholder_desc) == 24 && sizeof(PlaceholderWork) == 24, "placeholder descriptors");
static_assert(sizeof(ldpc_cuda_ulsch_desc) == 72, "ldpc_cuda_ulsch_desc layout");
static_assert(sizeof(TbResult) == sizeof(ldpc_cuda_tb_result) && sizeof(ldpc_cuda_tb_result) == 8, "TbResult mirrors the ABI struct");

struct ldpc_cuda_handle {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  uint32_t max_cbs = 0;
  uint32_t flags = 0;
  ldpc_cuda_algorithm alg = {6, 0.8f, LDPC_CUDA_SCALE_GENERIC, 0};
  unsigned char scale_lut[121];
  int packed_scale = TM_SCALE_NONE;  // closed-form scaling usable by the packed (TMEM-message) kernel
  CbWork* d_work = nullptr;
  CbWork* h_work[2] = {nullptr, nullptr};  // pinned, double-buffered: staging of call k overlaps the kernels of call k-1
  cudaEvent_t ev_work[2] = {nullptr, nullptr};
  unsigned work_slot = 0;
  // host-buffer entries: input copies, kernels and output copies of consecutive chunks overlap
  cudaStream_t stream2 = nullptr, h2d_stream = nullptr, d2h_stream = nullptr;
  cudaEvent_t ev_in[MAX_CHUNKS] = {}, ev_done[MAX_CHUNKS] = {}, ev_misc = nullptr;
  CbResult* d_results = nullptr;
  int8_t* d_in = nullptr;
  size_t d_in_cap = 0;
  uint8_t* d_out = nullptr;
  size_t d_out_cap = 0;
  int8_t* arena = nullptr;
  size_t arena_bytes = 0;
  bool arena_owned = false;  // false: attached, belongs to a pool shared between handles
  unsigned* pair_counter = nullptr;  // packed kernel: dynamic work distribution
  uint64_t launches = 0;
  uint64_t general_launches = 0;  // of those: the general decoder kernel (fallback)
  // transport-block assembly: work records + accumulators, and the output staging of the host entry
  uint8_t* d_tbwork = nullptr;
  size_t d_tbwork_cap = 0;
  uint8_t* d_tb_out = nullptr;
  size_t d_tb_out_cap = 0;
  uint8_t* h_tbwork[2] = {nullptr, nullptr};  // pinned, double-buffered like h_work: enqueueing never waits for the GPU
  size_t h_tbwork_cap[2] = {0, 0};
  cudaEvent_t ev_tbwork[2] = {nullptr, nullptr};
  unsigned tbwork_slot = 0;
  // demodulator jobs (same double-buffered pinned staging) and the device staging of its host entries
  uint8_t* d_demod = nullptr;
  size_t d_demod_cap = 0;
  uint8_t* h_demod[2] = {nullptr, nullptr};
  size_t h_demod_cap[2] = {0, 0};
  cudaEvent_t ev_demod[2] = {nullptr, nullptr};
  unsigned demod_slot = 0;
  std::vector<uint32_t> order;
  std::vector<uint32_t> group_begin;
  // pageable caller buffers of the host entries: bounce buffers + copy workers (created on first use), and a page-locked
  // mirror of the output range
  HostStager* stager = nullptr;
  uint8_t* out_mirror = nullptr;
  size_t out_mirror_cap = 0;
};

namespace {

int usable_device(int device)
{
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    return 0;
  }
  return prop.major == 10;
}

int upload_constants_locked();

// The tables in __constant__ / __device__ memory are per device and immutable: they are uploaded by the first handle
// created on a device and never again (later uploads would rewrite them under kernels of other handles, and the static
// host staging below is not re-entrant).
int upload_constants(int device)
{
  static std::mutex mtx;
  static bool done[TM_MAX_DEVICES] = {};
  std::lock_guard<std::mutex> lock(mtx);
  if (device >= 0 && device < TM_MAX_DEVICES && done[device]) {
    return LDPC_CUDA_OK;
  }
  const int rc = upload_constants_locked();
  if (rc == LDPC_CUDA_OK) {
    CUDA_TRY(cudaDeviceSynchronize());
    if (device >= 0 && device < TM_MAX_DEVICES) {
      done[device] = true;
    }
  }
  return rc;
}

int upload_constants_locked()
{
  static GraphTables gt;
  std::memset(&gt, 0, sizeof(gt));
  std::memcpy(gt.row_ptr[0], LDPC_BG1_ROW_PTR, sizeof(LDPC_BG1_ROW_PTR));
  std::memcpy(gt.row_ptr[1], LDPC_BG2_ROW_PTR, sizeof(LDPC_BG2_ROW_PTR));
  std::memcpy(gt.col[0], LDPC_BG1_COL, sizeof(LDPC_BG1_COL));
  std::memcpy(gt.col[1], LDPC_BG2_COL, sizeof(LDPC_BG2_COL));
  for (int s = 0; s != LDPC_NOF_LIFTING_SETS; ++s) {
    std::memcpy(gt.shift[0][s], LDPC_BG1_SHIFT[s], sizeof(LDPC_BG1_SHIFT[s]));
    std::memcpy(gt.shift[1][s], LDPC_BG2_SHIFT[s], sizeof(LDPC_BG2_SHIFT[s]));
  }
  CUDA_TRY(cudaMemcpyToSymbol(c_graphs, &gt, sizeof(gt)));
  static CrcTables ct;
  std::memset(&ct, 0, sizeof(ct));
  for (int code = 1; code <= 3; ++code) {
    build_crc_tables(code, ct.table[code], ct.pow8[code], 1057);
    crc_params(code, ct.poly[code], ct.order[code]);
  }
  CUDA_TRY(cudaMemcpyToSymbol(c_crc, &ct, sizeof(ct)));
  // x^e mod P per polynomial (parallel CRC check of the TMEM decoder kernel)
  static std::vector<unsigned> pw;
  pw.assign(3 * TM_CRC_POW_LEN, 0);
  for (int code = 1; code <= 3; ++code) {
    const unsigned top = 1U << ct.order[code];
    unsigned       r   = 1;
    for (unsigned e = 0; e != TM_CRC_POW_LEN; ++e) {
      pw[(code - 1) * TM_CRC_POW_LEN + e] = r;
      r <<= 1;
      if (r & top) {
        r ^= ct.poly[code];
      }
    }
  }
  CUDA_TRY(cudaMemcpyToSymbol(g_crc_pow, pw.data(), pw.size() * sizeof(unsigned)));
  // demodulator: interval tables of 64QAM / 256QAM from the TS 38.211 5.1.5 / 5.1.6 bit mapping, scrambler jump matrices
  {
    static DemodTables dt;
    std::memset(&dt, 0, sizeof(dt));
    for (int m = 0; m != 2; ++m) {
      const int L = m ? 16 : 8, nb = m ? 4 : 3;
      const float d = 1.0f / std::sqrt(m ? 170.0f : 42.0f), den = m ? 85.0f : 21.0f;
      // PAM level of every axis bit pattern: Re = (1 - 2 b0) [L/2 - (1 - 2 b1) [L/4 - ...]] (odd integers, unit 1/sqrt(norm))
      int level[16];
      for (int code = 0; code != L; ++code) {
        int v = 2 - (1 - 2 * ((code >> (nb - 1)) & 1)), step = 2;
        for (int k = nb - 2; k >= 1; --k) {
          step *= 2;
          v = step - (1 - 2 * ((code >> k) & 1)) * v;
        }
        level[code] = (1 - 2 * (code & 1)) * v;
      }
      for (int k = 0; k != nb; ++k) {
        const bool fine = (k == nb - 1);  // the finest bit uses half as many intervals, twice as wide (as the reference)
        const int n = fine ? L / 2 : L;
        dt.nof_intervals[m][k] = n;
        dt.width[m][k] = (float)(fine ? 4 : 2) * d;
        dt.rcp_width[m][k] = 1.0f / dt.width[m][k];
        dt.rcp_width_lo[m][k] = (float)(1.0 / (double)dt.width[m][k] - (double)dt.rcp_width[m][k]);
        for (int i = 0; i != n; ++i) {
          // nearest level with bit k = 0 / 1 from the middle of interval i (distances doubled to stay integer)
          const int mid2 = fine ? 2 * (4 * (i - L / 4) + 2) : 2 * (2 * (i - L / 2) + 1);
          int best[2] = {0, 0}, dist[2] = {1 << 30, 1 << 30};
          for (int code = 0; code != L; ++code) {
            const int bit = (code >> k) & 1, dd = std::abs(2 * level[code] - mid2);
            if (dd < dist[bit]) {
              dist[bit] = dd, best[bit] = level[code];
            }
          }
          // max-log LLR on the interval: ((x - a1)^2 - (x - a0)^2) / norm = slope * x + intercept
          dt.slope[m][k][i] = (float)(2 * (best[0] - best[1])) * d;
          dt.intercept[m][k][i] = (float)((best[1] * best[1] - best[0] * best[0]) / 2) / den;
        }
      }
    }
    const float s10 = 1.0f / std::sqrt(10.0f);
    dt.qam16_gain = 4 * s10;
    dt.qam16_threshold = 2 * s10;
    CUDA_TRY(cudaMemcpyToSymbol(c_demod, &dt, sizeof(dt)));
    // x1 words from chip Nc = 1600 on (TS 38.211 5.2.1: x1(n + 31) = x1(n + 3) ^ x1(n), x1(0) = 1, x1(1..30) = 0)
    {
      static std::vector<unsigned> x1w(DEMOD_X1_WORDS);
      uint32_t x1 = 1u;
      for (unsigned n = 0; n != 1600u; ++n) {
        x1 = (x1 >> 1) | ((((x1 >> 3) ^ x1) & 1u) << 30);
      }
      for (unsigned w = 0; w != DEMOD_X1_WORDS; ++w) {
        unsigned word = 0;
        for (unsigned b = 0; b != 32u; ++b) {
          word |= (x1 & 1u) << b;
          x1 = (x1 >> 1) | ((((x1 >> 3) ^ x1) & 1u) << 30);
        }
        x1w[w] = word;
      }
      CUDA_TRY(cudaMemcpyToSymbol(g_x1_words, x1w.data(), sizeof(unsigned) * DEMOD_X1_WORDS));
    }
    // A^(2^j) of the two shift registers (state bit i = x(n + i); row i = state bits whose XOR is the new bit i)
    static unsigned pw[2][32][32];
    std::memset(pw, 0, sizeof(pw));
    for (int q = 0; q != 2; ++q) {
      for (int i = 0; i != 30; ++i) {
        pw[q][0][i] = 1u << (i + 1);
      }
      pw[q][0][30] = q ? 0xFu : 0x9u;  // x2(n+31) = x2(n+3)^x2(n+2)^x2(n+1)^x2(n); x1(n+31) = x1(n+3)^x1(n)
      for (int j = 1; j != 32; ++j) {
        for (int i = 0; i != 31; ++i) {
          unsigned row = 0;
          for (int k = 0; k != 31; ++k) {
            if ((pw[q][j - 1][i] >> k) & 1u) {
              row ^= pw[q][j - 1][k];
            }
          }
          pw[q][j][i] = row;
        }
      }
    }
    CUDA_TRY(cudaMemcpyToSymbol(g_prg_pow, pw, sizeof(pw)));
  }
  // x^(8 TB_CHUNK k) mod P_24A (transport-block CRC: weight of a whole chunk, ldpc_tb.cuh)
  {
    static std::vector<unsigned> cp;
    cp.assign(TB_POW_CHUNKS, 0);
    const unsigned top = 1U << 24;
    unsigned r = 1;
    for (unsigned k = 0; k != TB_POW_CHUNKS; ++k) {
      cp[k] = r;
      for (unsigned e = 0; e != 8u * TB_CHUNK; ++e) {
        r <<= 1;
        if (r & top) {
          r ^= TB_CRC24A_POLY;
        }
      }
    }
    CUDA_TRY(cudaMemcpyToSymbol(g_tb_chunk_pow, cp.data(), cp.size() * sizeof(unsigned)));
  }
  return LDPC_CUDA_OK;
}

int ensure_capacity(ldpc_cuda_handle* h, size_t in_bytes, size_t out_bytes)
{
  if (in_bytes > h->d_in_cap) {
    if (h->d_in) {
      CUDA_TRY(cudaFree(h->d_in));
      h->d_in = nullptr, h->d_in_cap = 0;
    }
    size_t cap = in_bytes + (in_bytes >> 3) + 4096;
    CUDA_TRY(cudaMalloc(&h->d_in, cap));
    h->d_in_cap = cap;
  }
  if (out_bytes > h->d_out_cap) {
    if (h->d_out) {
      CUDA_TRY(cudaFree(h->d_out));
      h->d_out = nullptr, h->d_out_cap = 0;
    }
    size_t cap = out_bytes + (out_bytes >> 3) + 4096;
    CUDA_TRY(cudaMalloc(&h->d_out, cap));
    h->d_out_cap = cap;
  }
  return LDPC_CUDA_OK;
}

// Sort codeblock indices by (base graph, lifting size); every group becomes one launch.
void group_codeblocks(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs)
{
  h->order.resize(n);
  for (uint32_t i = 0; i != n; ++i) {
    h->order[i] = i;
  }
  auto key = [&](uint32_t i) { return (unsigned)descs[i].base_graph * 1024u + descs[i].lifting_size; };
  bool sorted = true;
  for (uint32_t i = 1; i < n && sorted; ++i) {
    sorted = key(i - 1) <= key(i);
  }
  if (!sorted) {
    std::stable_sort(h->order.begin(), h->order.end(), [&](uint32_t a, uint32_t b) { return key(a) < key(b); });
  }
  h->group_begin.clear();
  for (uint32_t i = 0; i != n; ++i) {
    if (i == 0 || key(h->order[i]) != key(h->order[i - 1])) {
      h->group_begin.push_back(i);
    }
  }
  h->group_begin.push_back(n);
}

GroupParams make_group(const ldpc_cuda_handle* h, int bg, int Z)
{
  GroupParams gp;
  std::memset(&gp, 0, sizeof(gp));
  gp.bg = bg;
  gp.Z = Z;
  gp.lifting_set = lifting_set_of(Z);
  gp.max_iterations = (int)h->alg.max_iterations;
  std::memcpy(gp.scale_lut, h->scale_lut, sizeof(gp.scale_lut));
  return gp;
}

int launch_dematch_group(ldpc_cuda_handle* h, int bg, int Z, const CbWork* d_work, uint32_t n, const int8_t* d_rm,
                         cudaStream_t stream);

// d_rm != nullptr: fused entry - d_llr is the soft-buffer arena and the group is rate-dematched / combined first, inside
// the decoder's prologue (fast kernel) or by the stand-alone dematch kernel in front of the general decoder.
int launch_decode_group(ldpc_cuda_handle* h, int bg, int Z, const CbWork* d_work, uint32_t n, const int8_t* d_llr,
                        const int8_t* d_rm, uint8_t* d_out, CbResult* d_res, cudaStream_t stream, unsigned* pair_counter)
{
  if (h->flags & LDPC_CUDA_FLAG_COPY_ONLY) {
    return LDPC_CUDA_OK;  // measurement aid: the host entries move their bytes, nothing is decoded
  }
  GroupParams gp = make_group(h, bg, Z);
  if (!(h->flags & LDPC_CUDA_FLAG_FORCE_SCALAR) && h->packed_scale != TM_SCALE_NONE) {
    int rc = launch_tmem_decode(gp, h->packed_scale, d_work, n, d_llr, d_rm, d_out, d_res, pair_counter, h->sm_count,
                                h->device, stream);
    if (rc == 0) {
      h->launches += 1;
      return LDPC_CUDA_OK;
    }
    if (rc < 0) {
      return LDPC_CUDA_ERR_CUDA;
    }
    // rc > 0: configuration not covered by the packed kernel -> general kernel below
  }
  if (d_rm != nullptr) {
    int rc = launch_dematch_group(h, bg, Z, d_work, n, d_rm, stream);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
  }
  const size_t smem = scalar_kernel_smem_bytes(bg, Z);
  static std::atomic<bool> attr_set[TM_MAX_DEVICES];  // per device: cudaFuncSetAttribute acts on the current one only
  if (h->device >= TM_MAX_DEVICES || !attr_set[h->device].load(std::memory_order_acquire)) {
    CUDA_TRY(cudaFuncSetAttribute(ldpc_decode_scalar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    if (h->device < TM_MAX_DEVICES) {
      attr_set[h->device].store(true, std::memory_order_release);
    }
  }
  const int threads = std::max(32, ((Z + 31) / 32) * 32);
  const unsigned grid = std::min<unsigned>(n, (unsigned)h->sm_count * 8u);
  ldpc_decode_scalar_kernel<<<grid, threads, smem, stream>>>(gp, d_work, n, d_llr, d_out, d_res);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  h->general_launches += 1;
  return LDPC_CUDA_OK;
}

int launch_dematch_group(ldpc_cuda_handle* h, int bg, int Z, const CbWork* d_work, uint32_t n, const int8_t* d_rm,
                         cudaStream_t stream)
{
  const unsigned grid = std::min<unsigned>(n, (unsigned)h->sm_count * 16u);
  ldpc_dematch_kernel<<<grid, 256, 0, stream>>>(bg, Z, d_work, n, d_rm, h->arena);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  return LDPC_CUDA_OK;
}

// Checks every descriptor of a call and sorts the codeblocks into (BG, Z) groups (h->order, h->group_begin).
int validate_and_group(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, bool fused)
{
  if (n > h->max_cbs) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  for (uint32_t i = 0; i != n; ++i) {
    const ldpc_cuda_cb_desc& d = descs[i];
    int rc = fused ? validate_dematch_desc(d) : validate_decode_desc(d);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
    if (fused) {
      const uint64_t N = (uint64_t)bg_N_short(d.base_graph) * d.lifting_size;
      if (d.crc_poly > LDPC_CUDA_CRC24B || d.soft_offset + N > h->arena_bytes ||
          d.nof_filler_bits >= bg_K(d.base_graph) * d.lifting_size) {
        return LDPC_CUDA_ERR_ARG;
      }
    }
  }
  group_codeblocks(h, n, descs);
  return LDPC_CUDA_OK;
}

// Work records [b, e) in launch order into the pinned array of `slot`. fused: in_offset = soft_offset, in_length = N.
void fill_work(ldpc_cuda_handle* h, unsigned slot, uint32_t b, uint32_t e, const ldpc_cuda_cb_desc* descs, bool fused)
{
  CbWork* h_work = h->h_work[slot];
  for (uint32_t k = b; k != e; ++k) {
    const uint32_t i = h->order[k];
    const ldpc_cuda_cb_desc& d = descs[i];
    CbWork w = {};
    w.in_offset = fused ? d.soft_offset : d.in_offset;
    w.out_offset = d.out_offset;
    w.rm_offset = d.in_offset;
    w.in_length = fused ? bg_N_short(d.base_graph) * d.lifting_size : d.in_length;
    w.rm_length = d.rm_length;
    w.Nref = d.Nref;
    w.nof_filler_bits = d.nof_filler_bits;
    w.crc_poly = d.crc_poly;
    w.early_stop = d.early_stop;
    w.rv = d.rv;
    w.mod_order = d.mod_order;
    w.new_data = d.new_data;
    w.skip_decode = fused ? d.skip_decode : 0;
    w.result_index = i;
    h_work[k] = w;
  }
}

// The pinned work array of the next slot, free again: it was last read by the upload of the call before the previous one.
int next_work_slot(ldpc_cuda_handle* h, unsigned& slot)
{
  slot = (h->work_slot++) & 1u;
  CUDA_TRY(cudaEventSynchronize(h->ev_work[slot]));
  return LDPC_CUDA_OK;
}

// Validates, groups, fills the whole work array and uploads it on the handle's stream (device-buffer entries).
int stage_work(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, bool fused)
{
  int rc = validate_and_group(h, n, descs, fused);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  unsigned slot = 0;
  rc = next_work_slot(h, slot);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  fill_work(h, slot, 0, n, descs, fused);
  CUDA_TRY(cudaMemcpyAsync(h->d_work, h->h_work[slot], sizeof(CbWork) * n, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaEventRecord(h->ev_work[slot], h->stream));
  return LDPC_CUDA_OK;
}

int run_groups(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, bool fused, const int8_t* d_in,
               uint8_t* d_out, CbResult* d_res)
{
  CUDA_TRY(cudaMemsetAsync(d_res, 0, sizeof(CbResult) * n, h->stream));
  for (size_t g = 0; g + 1 < h->group_begin.size(); ++g) {
    const uint32_t b = h->group_begin[g], e = h->group_begin[g + 1];
    const ldpc_cuda_cb_desc& d0 = descs[h->order[b]];
    const int rc = launch_decode_group(h, d0.base_graph, d0.lifting_size, h->d_work + b, e - b, fused ? h->arena : d_in,
                                       fused ? d_in : nullptr, d_out, d_res, h->stream, h->pair_counter);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
  }
  return LDPC_CUDA_OK;
}

// Work records of the transmit-side kernels (ldpc_encode.cuh): crc_poly carries the base graph, in_length the lifting
// size, reserved the lifting set. rate_match: geometry fields as for the dematcher.
int stage_tx_work(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, bool rate_match)
{
  if (n > h->max_cbs) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  for (uint32_t i = 0; i != n; ++i) {
    const ldpc_cuda_cb_desc& d = descs[i];
    if ((d.base_graph != 1 && d.base_graph != 2) || lifting_set_of(d.lifting_size) < 0 ||
        d.nof_filler_bits >= (bg_K(d.base_graph) - 2) * d.lifting_size) {
      return LDPC_CUDA_ERR_ARG;
    }
    if (rate_match) {
      const int rc = validate_dematch_desc(d);
      if (rc != LDPC_CUDA_OK) {
        return rc;
      }
    }
  }
  const unsigned slot = (h->work_slot++) & 1u;
  CUDA_TRY(cudaEventSynchronize(h->ev_work[slot]));
  CbWork* h_work = h->h_work[slot];
  for (uint32_t i = 0; i != n; ++i) {
    const ldpc_cuda_cb_desc& d = descs[i];
    CbWork& w = h_work[i];
    std::memset(&w, 0, sizeof(w));
    w.in_offset = d.in_offset;
    w.out_offset = d.out_offset;
    w.in_length = d.lifting_size;
    w.crc_poly = d.base_graph;
    w.reserved = (unsigned)lifting_set_of(d.lifting_size);
    w.nof_filler_bits = d.nof_filler_bits;
    w.rm_length = d.rm_length;
    w.Nref = d.Nref;
    w.rv = d.rv;
    w.mod_order = d.mod_order;
    w.result_index = i;
  }
  CUDA_TRY(cudaMemcpyAsync(h->d_work, h_work, sizeof(CbWork) * n, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaEventRecord(h->ev_work[slot], h->stream));
  return LDPC_CUDA_OK;
}

// Transport-block assembly (ldpc_tb.cuh) in two steps, so that the host entries can upload the work records of ALL blocks
// of a call up front (a copy enqueued behind a stream dependency reaches the copy engine only when the dependency has
// resolved, i.e. behind every input copy submitted in the meantime: the first group of a 16-cell slot was assembled 2 ms
// late) and launch the kernels group by group behind the chunks that complete them.
struct TbStaged {
  const TbWork* d_tw = nullptr;
  const TbCb* d_tc = nullptr;
  unsigned* d_partial = nullptr;
  unsigned slot = 0;
};

// Validates the geometry and uploads the work records of transport blocks [0, nof_tbs) on `stream`; h->ev_tbwork[slot] is
// recorded behind the upload.
int tb_stage(ldpc_cuda_handle* h, cudaStream_t stream, uint32_t nof_tbs, const ldpc_cuda_tb_desc* tbs, uint32_t n,
             const ldpc_cuda_cb_desc* descs, TbStaged& staged)
{
  if (nof_tbs > 65535u) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  // device buffer: work records (uploaded) | one partial remainder per chunk of every transport block (device only)
  const size_t off_cb = sizeof(TbWork) * nof_tbs, bytes = off_cb + sizeof(TbCb) * n;
  uint64_t total_chunks = 0;
  for (uint32_t t = 0; t != nof_tbs; ++t) {
    total_chunks += ((uint64_t)tbs[t].tbs_bits / 8 + 3 + TB_CHUNK - 1) / TB_CHUNK;
  }
  const size_t off_partial = (bytes + 15) & ~(size_t)15, dev_bytes = off_partial + sizeof(unsigned) * total_chunks;
  const unsigned slot = (h->tbwork_slot++) & 1u;
  if (h->ev_tbwork[slot] == nullptr) {
    CUDA_TRY(cudaEventCreateWithFlags(&h->ev_tbwork[slot], cudaEventDisableTiming));
  } else {
    CUDA_TRY(cudaEventSynchronize(h->ev_tbwork[slot]));  // the copy that last read this staging buffer has finished
  }
  if (bytes > h->h_tbwork_cap[slot]) {
    if (h->h_tbwork[slot]) {
      CUDA_TRY(cudaFreeHost(h->h_tbwork[slot]));
      h->h_tbwork[slot] = nullptr, h->h_tbwork_cap[slot] = 0;
    }
    const size_t cap = bytes + (bytes >> 2) + 4096;
    CUDA_TRY(cudaHostAlloc(&h->h_tbwork[slot], cap, cudaHostAllocDefault));
    h->h_tbwork_cap[slot] = cap;
  }
  uint8_t* stage = h->h_tbwork[slot];
  std::memset(stage, 0, bytes);
  TbWork* tw = reinterpret_cast<TbWork*>(stage);
  TbCb* tc = reinterpret_cast<TbCb*>(stage + off_cb);
  unsigned chunk_base = 0;
  for (uint32_t t = 0; t != nof_tbs; ++t) {
    const ldpc_cuda_tb_desc& tb = tbs[t];
    if (tb.nof_codeblocks == 0 || tb.first_codeblock > n || tb.nof_codeblocks > n - tb.first_codeblock ||
        tb.tbs_bits == 0 || tb.tbs_bits % 8 != 0 || (tb.nof_cb_crc_bits != 16 && tb.nof_cb_crc_bits != 24) ||
        (tb.nof_codeblocks > 1 && tb.nof_cb_crc_bits != 24)) {
      return LDPC_CUDA_ERR_ARG;
    }
    const ldpc_cuda_cb_desc& d0 = descs[tb.first_codeblock];
    if ((d0.base_graph != 1 && d0.base_graph != 2) || lifting_set_of(d0.lifting_size) < 0) {
      return LDPC_CUDA_ERR_ARG;
    }
    const unsigned K = bg_K(d0.base_graph) * d0.lifting_size;
    if (K <= tb.nof_cb_crc_bits + d0.nof_filler_bits) {
      return LDPC_CUDA_ERR_ARG;
    }
    const unsigned data_bits = K - tb.nof_cb_crc_bits - d0.nof_filler_bits;
    const uint64_t total = (uint64_t)tb.tbs_bits + (tb.nof_codeblocks > 1 ? 24u : 0u);
    // "All TB bits should be filled at this point" (pusch_decoder_impl.cpp:203)
    if ((uint64_t)data_bits * tb.nof_codeblocks < total || total / 8 > 0x7fffffffu) {
      return LDPC_CUDA_ERR_ARG;
    }
    for (uint32_t r = 0; r != tb.nof_codeblocks; ++r) {
      const ldpc_cuda_cb_desc& d = descs[tb.first_codeblock + r];
      if (d.base_graph != d0.base_graph || d.lifting_size != d0.lifting_size || d.nof_filler_bits != d0.nof_filler_bits) {
        return LDPC_CUDA_ERR_ARG;
      }
      TbCb& c = tc[tb.first_codeblock + r];
      c.src_offset = d.out_offset;
      c.already_ok = d.skip_decode ? 1u : 0u;
      c.result_index = tb.first_codeblock + r;
    }
    tw[t].tb_offset = tb.tb_offset;
    tw[t].first_cb = tb.first_codeblock;
    tw[t].nof_cbs = tb.nof_codeblocks;
    tw[t].data_bits = data_bits;
    tw[t].nbytes = (unsigned)(total / 8);
    tw[t].chunk_base = chunk_base;
    chunk_base += (tw[t].nbytes + TB_CHUNK - 1) / TB_CHUNK;
  }
  if (dev_bytes > h->d_tbwork_cap) {
    // the previous buffer may still be read by an earlier launch: wait for the handle's work before releasing it
    if (h->d_tbwork) {
      CUDA_TRY(cudaDeviceSynchronize());
      CUDA_TRY(cudaFree(h->d_tbwork));
      h->d_tbwork = nullptr, h->d_tbwork_cap = 0;
    }
    const size_t cap = dev_bytes + (dev_bytes >> 2) + 4096;
    CUDA_TRY(cudaMalloc(&h->d_tbwork, cap));
    h->d_tbwork_cap = cap;
  }
  CUDA_TRY(cudaMemcpyAsync(h->d_tbwork, stage, bytes, cudaMemcpyHostToDevice, stream));
  CUDA_TRY(cudaEventRecord(h->ev_tbwork[slot], stream));
  staged.d_tw = reinterpret_cast<const TbWork*>(h->d_tbwork);
  staged.d_tc = reinterpret_cast<const TbCb*>(h->d_tbwork + off_cb);
  staged.d_partial = reinterpret_cast<unsigned*>(h->d_tbwork + off_partial);
  staged.slot = slot;
  return LDPC_CUDA_OK;
}

// Assembles the staged transport blocks [t0, t1) on `stream` (which is ordered behind the upload of tb_stage).
// d_tb_results: the entry of block 0.
int tb_launch(ldpc_cuda_handle* h, cudaStream_t stream, const TbStaged& staged, const ldpc_cuda_tb_desc* tbs, uint32_t t0,
              uint32_t t1, const uint8_t* d_msgs, const CbResult* d_results, uint8_t* d_tb_out, TbResult* d_tb_results)
{
  if (t0 >= t1) {
    return LDPC_CUDA_OK;
  }
  unsigned max_chunks = 0;
  for (uint32_t t = t0; t != t1; ++t) {
    const unsigned nbytes = tbs[t].tbs_bits / 8 + (tbs[t].nof_codeblocks > 1 ? 3u : 0u);
    max_chunks = std::max(max_chunks, (nbytes + TB_CHUNK - 1) / TB_CHUNK);
  }
  const dim3 grid(max_chunks, t1 - t0);
  ldpc_tb_assemble_kernel<<<grid, TB_THREADS, 0, stream>>>(staged.d_tw + t0, staged.d_tc, d_msgs, d_tb_out, staged.d_partial);
  CUDA_TRY(cudaGetLastError());
  ldpc_tb_finish_kernel<<<t1 - t0, 128, 0, stream>>>(staged.d_tw + t0, staged.d_tc, d_results, staged.d_partial,
                                                    d_tb_results + t0);
  CUDA_TRY(cudaGetLastError());
  h->launches += 2;
  return LDPC_CUDA_OK;
}

// Both steps on one stream (device-buffer entry).
int enqueue_tb_assemble(ldpc_cuda_handle* h, cudaStream_t stream, uint32_t nof_tbs, const ldpc_cuda_tb_desc* tbs,
                        uint32_t n, const ldpc_cuda_cb_desc* descs, const uint8_t* d_msgs, const CbResult* d_results,
                        uint8_t* d_tb_out, TbResult* d_tb_results)
{
  TbStaged staged;
  const int rc = tb_stage(h, stream, nof_tbs, tbs, n, descs, staged);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  return tb_launch(h, stream, staged, tbs, 0, nof_tbs, d_msgs, d_results, d_tb_out, d_tb_results);
}

// Pinned double-buffered staging + device buffer of the demodulator's job lists: waits until the copy that last used the
// chosen slot has been consumed, grows the buffers if needed. The caller fills h->h_demod[slot], copies it to h->d_demod on
// its stream and records h->ev_demod[slot] behind the copy.
int demod_stage(ldpc_cuda_handle* h, size_t bytes, unsigned& slot_out)
{
  const unsigned slot = (h->demod_slot++) & 1u;
  slot_out = slot;
  if (h->ev_demod[slot] == nullptr) {
    CUDA_TRY(cudaEventCreateWithFlags(&h->ev_demod[slot], cudaEventDisableTiming));
  } else {
    CUDA_TRY(cudaEventSynchronize(h->ev_demod[slot]));
  }
  if (bytes > h->h_demod_cap[slot]) {
    if (h->h_demod[slot]) {
      CUDA_TRY(cudaFreeHost(h->h_demod[slot]));
      h->h_demod[slot] = nullptr, h->h_demod_cap[slot] = 0;
    }
    CUDA_TRY(cudaHostAlloc(&h->h_demod[slot], 2 * bytes + 4096, cudaHostAllocDefault));
    h->h_demod_cap[slot] = 2 * bytes + 4096;
  }
  if (bytes > h->d_demod_cap) {
    if (h->d_demod) {
      CUDA_TRY(cudaDeviceSynchronize());
      CUDA_TRY(cudaFree(h->d_demod));
      h->d_demod = nullptr, h->d_demod_cap = 0;
    }
    CUDA_TRY(cudaMalloc(&h->d_demod, 2 * bytes + 4096));
    h->d_demod_cap = 2 * bytes + 4096;
  }
  return LDPC_CUDA_OK;
}

// Demodulator / descrambler jobs (ldpc_demod.cuh): validation, upload through pinned double-buffered staging, launch.
int enqueue_demod(ldpc_cuda_handle* h, cudaStream_t stream, uint32_t nof_jobs, const ldpc_cuda_demod_desc* jobs,
                  const float* d_symbols, const float* d_noise, const int8_t* d_in_llrs, int8_t* d_out)
{
  if (nof_jobs > 65535u) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  unsigned max_chunks = 0;
  uint64_t total_chunks = 0;
  for (uint32_t i = 0; i != nof_jobs; ++i) {
    const ldpc_cuda_demod_desc& j = jobs[i];
    const bool mod_ok = j.mod_order == 0 || j.mod_order == 1 || j.mod_order == 2 || j.mod_order == 4 || j.mod_order == 6 ||
                        j.mod_order == 8;
    const uint64_t nllr = (uint64_t)j.nof_symbols * (j.mod_order ? j.mod_order : 1u);
    if (!mod_ok || (j.pi2_bpsk && j.mod_order != 1) || nllr > 0x7fffffffu || (j.mod_order == 0 && d_in_llrs == nullptr) ||
        (j.mod_order != 0 && (d_symbols == nullptr || d_noise == nullptr))) {
      return LDPC_CUDA_ERR_ARG;
    }
    max_chunks = std::max(max_chunks, (unsigned)((nllr + DEMOD_CHUNK - 1) / DEMOD_CHUNK));
    total_chunks += (nllr + DEMOD_CHUNK - 1) / DEMOD_CHUNK;
  }
  if (max_chunks == 0) {
    return LDPC_CUDA_OK;
  }
  // Spans of 8192 LLRs per block: more of them amortise the scrambler's jump-ahead, fewer keep a small launch spread over
  // the SMs (8 resident blocks each).
  unsigned spans = DEMOD_MAX_SPANS;
  while (spans > 1 && total_chunks / spans < 8u * (unsigned)h->sm_count) {
    spans /= 2;
  }
  max_chunks = (max_chunks + spans - 1) / spans;
  const size_t bytes = sizeof(DemodJob) * nof_jobs;
  unsigned slot = 0;
  int rc_stage = demod_stage(h, bytes, slot);
  if (rc_stage != LDPC_CUDA_OK) {
    return rc_stage;
  }
  DemodJob* hj = reinterpret_cast<DemodJob*>(h->h_demod[slot]);
  for (uint32_t i = 0; i != nof_jobs; ++i) {
    const ldpc_cuda_demod_desc& j = jobs[i];
    DemodJob& w = hj[i];
    std::memset(&w, 0, sizeof(w));
    w.sym_offset = j.symbol_offset;
    w.llr_offset = j.llr_offset;
    w.nof_llrs = j.nof_symbols * (j.mod_order ? j.mod_order : 1u);
    w.c_init = j.c_init;
    w.mod_order = j.mod_order;
    w.pi2 = j.pi2_bpsk ? 1 : 0;
    w.descramble = j.descramble ? 1 : 0;
  }
  CUDA_TRY(cudaMemcpyAsync(h->d_demod, hj, bytes, cudaMemcpyHostToDevice, stream));
  CUDA_TRY(cudaEventRecord(h->ev_demod[slot], stream));
  const dim3 grid(max_chunks, nof_jobs);
  ldpc_demod_kernel<<<grid, DEMOD_THREADS, 0, stream>>>(reinterpret_cast<const DemodJob*>(h->d_demod),
                                                         reinterpret_cast<const float2*>(d_symbols), d_noise, d_in_llrs,
                                                         d_out, spans);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  return LDPC_CUDA_OK;
}

// UCI repetition placeholders (ldpc_demod.cuh, ldpc_demod_placeholder_kernel): one work item per (RE, layer).
int enqueue_placeholders(ldpc_cuda_handle* h, cudaStream_t stream, uint32_t nof_jobs,
                         const ldpc_cuda_placeholder_desc* jobs, const uint16_t* re_indexes, uint32_t nof_re_indexes,
                         int8_t* d_llrs)
{
  size_t total = 0;
  for (uint32_t i = 0; i != nof_jobs; ++i) {
    const ldpc_cuda_placeholder_desc& j = jobs[i];
    const bool mod_ok = j.mod_order == 1 || j.mod_order == 2 || j.mod_order == 4 || j.mod_order == 6 || j.mod_order == 8;
    if (!mod_ok || j.nof_layers == 0 || j.nof_layers > 8 || (uint64_t)j.first_re + j.nof_re > nof_re_indexes ||
        (j.nof_re != 0 && re_indexes == nullptr)) {
      return LDPC_CUDA_ERR_ARG;
    }
    total += (size_t)j.nof_re * j.nof_layers;
  }
  if (total == 0) {
    return LDPC_CUDA_OK;
  }
  if (total > 0x7fffffffu) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  const size_t bytes = sizeof(PlaceholderWork) * total;
  unsigned slot = 0;
  int rc = demod_stage(h, bytes, slot);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  PlaceholderWork* hw = reinterpret_cast<PlaceholderWork*>(h->h_demod[slot]);
  size_t k = 0;
  for (uint32_t i = 0; i != nof_jobs; ++i) {
    const ldpc_cuda_placeholder_desc& j = jobs[i];
    for (uint32_t r = 0; r != j.nof_re; ++r) {
      for (uint32_t layer = 0; layer != j.nof_layers; ++layer) {
        PlaceholderWork& w = hw[k++];
        // ulsch_placeholder_list::for_each (ulsch_placeholder_list.h:73-95): the y placeholder is the second bit of the symbol
        w.llr_offset = j.llr_offset;
        w.y = (uint32_t)j.mod_order * ((uint32_t)j.nof_layers * re_indexes[j.first_re + r] + layer) + 1u;
        w.c_init = j.c_init;
        w.nof_llrs = j.nof_llrs;
        w.nx = j.mod_order > 2 ? j.mod_order - 2u : 0u;
      }
    }
  }
  CUDA_TRY(cudaMemcpyAsync(h->d_demod, hw, bytes, cudaMemcpyHostToDevice, stream));
  CUDA_TRY(cudaEventRecord(h->ev_demod[slot], stream));
  const unsigned warps_per_block = 4;
  const unsigned grid = (unsigned)((total + warps_per_block - 1) / warps_per_block);
  ldpc_demod_placeholder_kernel<<<grid, 32 * warps_per_block, 0, stream>>>(
      reinterpret_cast<const PlaceholderWork*>(h->d_demod), (unsigned)total, d_llrs);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  return LDPC_CUDA_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
extern "C" {

int ldpc_cuda_abi_version(void) { return LDPC_CUDA_ABI_VERSION; }

const char* ldpc_cuda_status_string(int status)
{
  switch (status) {
    case LDPC_CUDA_OK:
      return "ok";
    case LDPC_CUDA_ERR_ARG:
      return "invalid argument";
    case LDPC_CUDA_ERR_CUDA:
      return "CUDA error or no sm_100 device (there is no CPU fallback)";
    case LDPC_CUDA_ERR_CAPACITY:
      return "batch exceeds the capacity the handle was created with";
    case LDPC_CUDA_ERR_UNSUPPORTED:
      return "unsupported";
    default:
      return "unknown status";
  }
}

int ldpc_cuda_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  int usable = 0;
  for (int d = 0; d != n; ++d) {
    usable += usable_device(d);
  }
  return usable;
}

static int create_resources(ldpc_cuda_handle* h);

int ldpc_cuda_create(ldpc_cuda_handle** handle, int device, uint32_t max_codeblocks, uint32_t flags)
{
  if (handle == nullptr || max_codeblocks == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  *handle = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n || !usable_device(device)) {
    cudaGetLastError();
    return LDPC_CUDA_ERR_CUDA;
  }
  CUDA_TRY(cudaSetDevice(device));
  ldpc_cuda_handle* h = new (std::nothrow) ldpc_cuda_handle;
  if (h == nullptr) {
    return LDPC_CUDA_ERR_CUDA;
  }
  h->device = device;
  h->max_cbs = max_codeblocks;
  h->flags = flags;
  int rc = create_resources(h);
  if (rc == LDPC_CUDA_OK) {
    ldpc_cuda_algorithm alg = {6, 0.8f, LDPC_CUDA_SCALE_GENERIC, 0};
    rc = ldpc_cuda_set_algorithm(h, &alg);
  }
  if (rc != LDPC_CUDA_OK) {
    ldpc_cuda_destroy(h);  // releases whatever was created before the failure
    return rc;
  }
  *handle = h;
  return LDPC_CUDA_OK;
}

// Streams, events and buffers of a new handle (the caller destroys the handle when this fails half-way).
static int create_resources(ldpc_cuda_handle* h)
{
  const uint32_t max_codeblocks = h->max_cbs;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, h->device));
  h->sm_count = prop.multiProcessorCount;
  int rc = upload_constants(h->device);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  CUDA_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  CUDA_TRY(cudaMalloc(&h->d_work, sizeof(CbWork) * max_codeblocks));
  CUDA_TRY(cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
  CUDA_TRY(cudaStreamCreateWithFlags(&h->h2d_stream, cudaStreamNonBlocking));
  // The copy-out stream also runs the transport-block assembly kernels of the host entries. The decode kernels own whole SMs
  // (384 threads x 168 registers = the register file), and the chunks' kernels follow each other without a gap, so at
  // equal priority an assembly kernel only got SMs when the LAST chunk had drained: the blocks of a 16-cell slot came back
  // 0.8 ms after the last decode. At the highest priority its blocks are placed first whenever a decode block retires.
  int prio_lo = 0, prio_hi = 0;
  CUDA_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  CUDA_TRY(cudaStreamCreateWithPriority(&h->d2h_stream, cudaStreamNonBlocking, prio_hi));
  for (int i = 0; i != 2; ++i) {
    CUDA_TRY(cudaMallocHost(&h->h_work[i], sizeof(CbWork) * max_codeblocks));
    CUDA_TRY(cudaEventCreateWithFlags(&h->ev_work[i], cudaEventDisableTiming));
  }
  for (unsigned i = 0; i != MAX_CHUNKS; ++i) {
    CUDA_TRY(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
  }
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_misc, cudaEventDisableTiming));
  CUDA_TRY(cudaMalloc(&h->d_results, sizeof(CbResult) * max_codeblocks));
  CUDA_TRY(cudaMalloc(&h->pair_counter, 64));
  return LDPC_CUDA_OK;
}

void ldpc_cuda_destroy(ldpc_cuda_handle* h)
{
  if (h == nullptr) {
    return;
  }
  cudaSetDevice(h->device);
  if (h->stream) {
    cudaStreamSynchronize(h->stream);
  }
  cudaFree(h->d_work);
  for (cudaStream_t st : {h->stream2, h->h2d_stream, h->d2h_stream}) {
    if (st) {
      cudaStreamSynchronize(st);
      cudaStreamDestroy(st);
    }
  }
  for (int i = 0; i != 2; ++i) {
    cudaFreeHost(h->h_work[i]);
    if (h->ev_work[i]) {
      cudaEventDestroy(h->ev_work[i]);
    }
  }
  for (unsigned i = 0; i != MAX_CHUNKS; ++i) {
    if (h->ev_in[i]) {
      cudaEventDestroy(h->ev_in[i]);
    }
    if (h->ev_done[i]) {
      cudaEventDestroy(h->ev_done[i]);
    }
  }
  if (h->ev_misc) {
    cudaEventDestroy(h->ev_misc);
  }
  delete h->stager;
  cudaFreeHost(h->out_mirror);
  cudaFree(h->d_results);
  cudaFree(h->d_in);
  cudaFree(h->d_out);
  if (h->arena_owned) {
    cudaFree(h->arena);
  }
  cudaFree(h->pair_counter);
  cudaFree(h->d_tbwork);
  cudaFree(h->d_demod);
  for (int i = 0; i != 2; ++i) {
    cudaFreeHost(h->h_demod[i]);
    if (h->ev_demod[i]) {
      cudaEventDestroy(h->ev_demod[i]);
    }
  }
  for (int i = 0; i != 2; ++i) {
    cudaFreeHost(h->h_tbwork[i]);
    if (h->ev_tbwork[i]) {
      cudaEventDestroy(h->ev_tbwork[i]);
    }
  }
  cudaFree(h->d_tb_out);
  if (h->stream) {
    cudaStreamDestroy(h->stream);
  }
  delete h;
}

int ldpc_cuda_set_algorithm(ldpc_cuda_handle* h, const ldpc_cuda_algorithm* alg)
{
  // asserts of ldpc_decoder_impl::init, ldpc_decoder_impl.cpp:46-50
  if (h == nullptr || alg == nullptr || alg->max_iterations == 0 || alg->max_iterations > 65535 ||
      !(alg->scaling_factor > 0.f && alg->scaling_factor < 1.f) || alg->scale_mode > LDPC_CUDA_SCALE_NEON) {
    return LDPC_CUDA_ERR_ARG;
  }
  h->alg = *alg;
  build_scale_lut(alg->scaling_factor, alg->scale_mode, h->scale_lut);
  h->packed_scale = tm_scale_variant(h->scale_lut);
  if (h->packed_scale != TM_SCALE_NONE) {
    // the fp16 closed form must reproduce the reference table for every magnitude, or the general kernel is used
    CUDA_TRY(cudaSetDevice(h->device));
    const int chk = tm_scale_selfcheck(h->packed_scale, h->scale_lut, h->stream);
    if (chk < 0) {
      return LDPC_CUDA_ERR_CUDA;
    }
    if (chk > 0) {
      std::fprintf(stderr, "ldpc_cuda: packed scaling self-check failed, using the general kernel\n");
      h->packed_scale = TM_SCALE_NONE;
    }
  }
  return LDPC_CUDA_OK;
}

void* ldpc_cuda_stream(ldpc_cuda_handle* h) { return h ? (void*)h->stream : nullptr; }

int ldpc_cuda_synchronize(ldpc_cuda_handle* h)
{
  if (h == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

uint64_t ldpc_cuda_launch_count(const ldpc_cuda_handle* h) { return h ? h->launches : 0; }
uint64_t ldpc_cuda_general_kernel_launch_count(const ldpc_cuda_handle* h) { return h ? h->general_launches : 0; }

// ---- batched, device buffers ----------------------------------------------------------------------------------------
int ldpc_cuda_decode_batch_device(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* llrs_d,
                                  uint8_t* out_bits_d, ldpc_cuda_cb_result* results_d)
{
  if (h == nullptr || descs == nullptr || llrs_d == nullptr || out_bits_d == nullptr || results_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  int rc = stage_work(h, n, descs, false);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  return run_groups(h, n, descs, false, llrs_d, out_bits_d, reinterpret_cast<CbResult*>(results_d));
}

int ldpc_cuda_dematch_decode_batch_device(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs,
                                          const int8_t* rm_llrs_d, uint8_t* out_bits_d, ldpc_cuda_cb_result* results_d)
{
  if (h == nullptr || descs == nullptr || rm_llrs_d == nullptr || out_bits_d == nullptr || results_d == nullptr ||
      h->arena == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  int rc = stage_work(h, n, descs, true);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  return run_groups(h, n, descs, true, rm_llrs_d, out_bits_d, reinterpret_cast<CbResult*>(results_d));
}

// ---- batched, host buffers ------------------------------------------------------------------------------------------
// A chunk = consecutive codeblocks (in launch order) of one (BG, Z) group whose input / output byte ranges are copied and
// decoded as a unit.
struct Chunk {
  uint32_t b, e;            // range in launch order
  uint64_t in_lo, in_hi;    // host input bytes (hull)
  uint64_t out_lo, out_hi;  // host output bytes
  std::vector<std::pair<uint64_t, uint64_t>> in_runs;  // contiguous input pieces to copy; empty: the hull
  // Contiguous output pieces; empty: the hull [out_lo, out_hi). More than one piece when the chunk's messages are not
  // adjacent in the caller's buffer (e.g. the transport blocks in between belong to another device of a pool): the bytes
  // in between are then neither read nor written.
  std::vector<std::pair<uint64_t, uint64_t>> out_runs;
  // Equally long pieces at a constant stride (a batch sharded round-robin over several devices: codeblock i, i + G, ...):
  // ONE pitched copy instead of thousands of small ones. rows == 0: not in this form.
  uint64_t in_stride = 0, in_width = 0, in_rows = 0;
  uint64_t out_stride = 0, out_width = 0, out_rows = 0;  // outputs likewise; never the hull then (it belongs to others too)
};

// Pitched host <-> device copy with the same pitch on both sides (the device buffers mirror the host layout).
static cudaError_t copy_rows(void* dst, const void* src, uint64_t stride, uint64_t width, uint64_t rows, cudaMemcpyKind kind,
                             cudaStream_t stream)
{
  return cudaMemcpy2DAsync(dst, stride, src, stride, width, rows, kind, stream);
}

static inline void cb_ranges(const ldpc_cuda_cb_desc& d, bool fused, uint64_t& in_end, uint64_t& out_end)
{
  in_end = d.in_offset + (fused ? d.rm_length : d.in_length);
  out_end = d.out_offset + ((uint64_t)bg_K(d.base_graph) * d.lifting_size + 7) / 8;
}

// Recognises a chunk whose codeblocks lie at a constant stride in the caller's buffers (input pieces of equal length,
// output messages of equal length, gaps in between that belong to somebody else) and records the pitched form.
static void strided_form(const ldpc_cuda_handle* h, const ldpc_cuda_cb_desc* descs, bool fused, Chunk& c)
{
  const uint32_t n = c.e - c.b;
  if (n < 2) {
    return;
  }
  const ldpc_cuda_cb_desc& d0 = descs[h->order[c.b]];
  const ldpc_cuda_cb_desc& d1 = descs[h->order[c.b + 1]];
  uint64_t ie0, oe0;
  cb_ranges(d0, fused, ie0, oe0);
  const uint64_t iw = ie0 - d0.in_offset, ow = oe0 - d0.out_offset;
  if (d1.in_offset <= d0.in_offset || d1.out_offset <= d0.out_offset) {
    return;
  }
  const uint64_t is = d1.in_offset - d0.in_offset, os = d1.out_offset - d0.out_offset;
  bool in_ok = iw != 0 && is > iw, out_ok = os > ow;  // dense layouts keep the plain range copies
  for (uint32_t k = 1; k != n && (in_ok || out_ok); ++k) {
    const ldpc_cuda_cb_desc& d = descs[h->order[c.b + k]];
    uint64_t ie, oe;
    cb_ranges(d, fused, ie, oe);
    in_ok = in_ok && d.in_offset == d0.in_offset + k * is && ie - d.in_offset == iw;
    out_ok = out_ok && d.out_offset == d0.out_offset + k * os && oe - d.out_offset == ow;
  }
  if (in_ok) {
    c.in_stride = is, c.in_width = iw, c.in_rows = n;
    c.in_runs.clear();
    c.in_runs.push_back({d0.in_offset, d0.in_offset + iw});  // counted as one copy
  }
  if (out_ok) {
    c.out_stride = os, c.out_width = ow, c.out_rows = n;
  }
}

// Splits the staged groups into at most MAX_CHUNKS chunks of about equal input size. Chunks are only formed when the
// caller's OUTPUT layout is monotone in launch order (every codeblock's message lies behind its predecessor's), so that
// the output byte ranges of different chunks are disjoint; otherwise each group is one chunk over the whole span. The
// INPUT may lie anywhere (e.g. retransmissions and new transport blocks in different host buffers' regions): a chunk
// copies the contiguous runs of its codeblocks' inputs one by one.
static void plan_chunks(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, bool fused,
                        std::vector<Chunk>& chunks)
{
  chunks.clear();
  bool monotone = true;
  uint64_t out_prev = 0, in_lo = ~0ull, in_hi = 0, out_lo = ~0ull, out_hi = 0, total_in = 0;
  for (uint32_t k = 0; k != n; ++k) {
    const ldpc_cuda_cb_desc& d = descs[h->order[k]];
    uint64_t ie, oe;
    cb_ranges(d, fused, ie, oe);
    monotone = monotone && (k == 0 || d.out_offset >= out_prev);
    out_prev = oe;
    total_in += ie - d.in_offset;
    in_lo = std::min(in_lo, d.in_offset), in_hi = std::max(in_hi, ie);
    out_lo = std::min(out_lo, d.out_offset), out_hi = std::max(out_hi, oe);
  }
  const size_t nof_groups = h->group_begin.size() - 1;
  constexpr uint64_t MIN_CHUNK_BYTES = 8ull << 20;
  constexpr uint64_t RUN_GAP = 4096;
  constexpr size_t MAX_RUNS = 4096;  // copies per call; beyond that the layout is too scattered to be worth it
  auto single = [&]() {
    chunks.clear();
    chunks.push_back({0, n, in_lo, in_hi, out_lo, out_hi, {}});  // one copy in, all launches, one copy out
    Chunk& c = chunks.back();
    strided_form(h, descs, fused, c);
    c.in_runs.clear();
    // output pieces in launch order (exactly adjacent messages merge); too scattered: the hull, as for one device alone
    if (c.out_rows == 0) {
      for (uint32_t k = 0; k != n && c.out_runs.size() <= MAX_RUNS; ++k) {
        const ldpc_cuda_cb_desc& d = descs[h->order[k]];
        uint64_t ie, oe;
        cb_ranges(d, fused, ie, oe);
        if (!c.out_runs.empty() && d.out_offset == c.out_runs.back().second) {
          c.out_runs.back().second = oe;
        } else {
          c.out_runs.push_back({d.out_offset, oe});
        }
      }
      if (c.out_runs.size() == 1 || c.out_runs.size() > MAX_RUNS) {
        c.out_runs.clear();
      }
    }
  };
  if (!monotone || 2 * nof_groups > MAX_CHUNKS || total_in < 2 * MIN_CHUNK_BYTES) {
    single();
    return;
  }
  // a chunk closes when it reaches the target size or at the end of its group: at most MAX_CHUNKS chunks
  const uint64_t by_size = MAX_CHUNKS - nof_groups;
  const uint64_t target = std::max<uint64_t>(MIN_CHUNK_BYTES, (total_in + by_size - 1) / by_size);
  size_t runs = 0;
  for (size_t g = 0; g != nof_groups; ++g) {
    uint32_t b = h->group_begin[g];
    const uint32_t ge = h->group_begin[g + 1];
    while (b != ge) {
      Chunk c = {b, b, ~0ull, 0, descs[h->order[b]].out_offset, 0, {}};
      uint64_t bytes = 0;
      do {
        const ldpc_cuda_cb_desc& d = descs[h->order[c.e]];
        uint64_t ie;
        cb_ranges(d, fused, ie, c.out_hi);
        if (ie > d.in_offset) {
          // continues the current run (small gaps, e.g. alignment padding between codeblocks, are copied along)
          if (!c.in_runs.empty() && d.in_offset >= c.in_runs.back().first && d.in_offset <= c.in_runs.back().second + RUN_GAP) {
            c.in_runs.back().second = std::max(c.in_runs.back().second, ie);
          } else {
            c.in_runs.push_back({d.in_offset, ie});
          }
          c.in_lo = std::min(c.in_lo, d.in_offset), c.in_hi = std::max(c.in_hi, ie);
          bytes += ie - d.in_offset;
        }
        if (!c.out_runs.empty() && d.out_offset == c.out_runs.back().second) {
          c.out_runs.back().second = c.out_hi;
        } else {
          c.out_runs.push_back({d.out_offset, c.out_hi});
        }
        ++c.e;
      } while (c.e != ge && (bytes < target || chunks.size() + 1 == MAX_CHUNKS));
      if (c.in_hi < c.in_lo) {
        c.in_lo = c.in_hi = in_lo;  // nothing to copy
      }
      strided_form(h, descs, fused, c);
      if (c.out_rows != 0 || c.out_runs.size() == 1) {
        c.out_runs.clear();  // one pitched copy / the hull is exact
      }
      runs += c.in_runs.size() + c.out_runs.size();
      chunks.push_back(c);
      b = c.e;
    }
  }
  if (runs > MAX_RUNS) {
    single();
  }
}

static int batch_host_enqueue(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* in,
                              uint8_t* out, ldpc_cuda_cb_result* results, bool fused, uint32_t nof_tbs,
                              const ldpc_cuda_tb_desc* tbs, uint8_t* tb_out, ldpc_cuda_tb_result* tb_results);

// Every exit, also the error exits taken in the middle of the chunk loop, leaves the four streams idle: copies already
// enqueued into the caller's buffers (or out of the pinned work slot) have finished when the caller sees the status.
static int batch_host(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* in, uint8_t* out,
                      ldpc_cuda_cb_result* results, bool fused, uint32_t nof_tbs = 0, const ldpc_cuda_tb_desc* tbs = nullptr,
                      uint8_t* tb_out = nullptr, ldpc_cuda_tb_result* tb_results = nullptr)
{
  const int rc = batch_host_enqueue(h, n, descs, in, out, results, fused, nof_tbs, tbs, tb_out, tb_results);
  if (rc != LDPC_CUDA_OK && h != nullptr && h->stream != nullptr) {
    cudaSetDevice(h->device);
    for (cudaStream_t st : {h->h2d_stream, h->stream, h->stream2, h->d2h_stream}) {
      cudaStreamSynchronize(st);
    }
  }
  return rc;
}

static int batch_host_enqueue(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* in,
                              uint8_t* out, ldpc_cuda_cb_result* results, bool fused, uint32_t nof_tbs,
                              const ldpc_cuda_tb_desc* tbs, uint8_t* tb_out, ldpc_cuda_tb_result* tb_results)
{
  if (nof_tbs != 0 && (tbs == nullptr || tb_out == nullptr || tb_results == nullptr)) {
    return LDPC_CUDA_ERR_ARG;
  }
  // out == nullptr (transport-block variant only): the codeblock messages stay in device memory, only the assembled blocks
  // and the verdicts travel back. A codeblock with skip_decode needs its message from the caller, so it excludes that.
  const bool keep_messages = (out == nullptr);
  if (h == nullptr || descs == nullptr || in == nullptr || (keep_messages && nof_tbs == 0) || results == nullptr ||
      (fused && h->arena == nullptr)) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  if (keep_messages) {
    for (uint32_t i = 0; i != n; ++i) {
      if (descs[i].skip_decode) {
        return LDPC_CUDA_ERR_ARG;
      }
    }
  }
  CUDA_TRY(cudaSetDevice(h->device));
  // LDPC_CUDA_TRACE=1: one line per host-entry call on stderr with where its wall-clock time went
  static const bool trace = std::getenv("LDPC_CUDA_TRACE") != nullptr;
  // LDPC_CUDA_TRACE=2: additionally the device-side timeline of the chunks (when each chunk's input copy and kernel ended)
  static const bool trace_chunks = trace && std::atoi(std::getenv("LDPC_CUDA_TRACE")) >= 2;
  std::vector<cudaEvent_t> tl;  // [0] start, then (copy end, kernel end) per chunk
  auto mark = [&](cudaStream_t st) {
    if (trace_chunks) {
      cudaEvent_t e = nullptr;
      cudaEventCreate(&e);
      cudaEventRecord(e, st);
      tl.push_back(e);
    }
  };
  const auto t_enter = std::chrono::steady_clock::now();
  // The work records are filled and uploaded chunk by chunk in front of the chunk's input copy (same stream), so that the
  // host's share (25 ns per codeblock: 0.5 ms for a 16-cell slot, 1.7 ms for 65 536 codeblocks) hides behind the copies
  // and kernels of the chunks before. (Uploading everything first cost that time up front; issuing input copies before a
  // single upload delayed every kernel behind them, profiles/r01_icache_microbench.md.)
  int rc = validate_and_group(h, n, descs, fused);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  unsigned work_slot = 0;
  rc = next_work_slot(h, work_slot);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  const auto t_staged = std::chrono::steady_clock::now();
  std::vector<Chunk> chunks;
  plan_chunks(h, n, descs, fused, chunks);
  if (chunks.empty() || chunks.size() > MAX_CHUNKS) {
    return LDPC_CUDA_ERR_ARG;
  }
  const auto t_planned = std::chrono::steady_clock::now();
  uint64_t in_lo = ~0ull, in_hi = 0, out_hi = 0;
  const uint64_t out_lo = chunks.front().out_lo;
  for (const Chunk& c : chunks) {
    in_lo = std::min(in_lo, c.in_lo), in_hi = std::max(in_hi, c.in_hi), out_hi = std::max(out_hi, c.out_hi);
  }
  rc = ensure_capacity(h, in_hi - in_lo, out_hi - out_lo);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  // Output bytes / bits a codeblock does not write (skip_decode, all-zero input with a CRC, the bits of the last byte
  // beyond a message that is not a whole number of bytes) must keep the caller's contents.
  bool preserve = false;
  for (uint32_t i = 0; i != n && !preserve && !keep_messages; ++i) {
    preserve = descs[i].skip_decode || descs[i].crc_poly != LDPC_CUDA_CRC_NONE ||
               (bg_K(descs[i].base_graph) * descs[i].lifting_size) % 8 != 0;
  }
  const int8_t* d_in0 = h->d_in - in_lo;   // device address of host input byte 0
  uint8_t* d_out0 = h->d_out - out_lo;     // device address of host output byte 0
  // Pageable caller memory (what the reference's callers have): inputs go through the handle's page-locked bounce buffers,
  // filled by a few threads with non-temporal stores while the copy engine drains the previous piece
  // (ldpc_host_stager.h); outputs land in a page-locked mirror that is copied to the caller's buffer once at the end.
  // A pageable cudaMemcpyAsync would instead run at the driver's bounce-buffer speed on this thread and stall the
  // pipeline of chunks at every copy.
  const bool stage_in = !(h->flags & LDPC_CUDA_FLAG_COPY_ONLY) && is_pageable(in + in_lo);
  const bool stage_out = !keep_messages && !(h->flags & LDPC_CUDA_FLAG_COPY_ONLY) && (out_hi - out_lo) <= (1ull << 30) &&
                         is_pageable(out + out_lo);
  if ((stage_in || stage_out) && h->stager == nullptr) {
    h->stager = new (std::nothrow) HostStager;
    if (h->stager == nullptr) {
      return LDPC_CUDA_ERR_CUDA;
    }
  }
  if (stage_out && out_hi - out_lo > h->out_mirror_cap) {
    cudaFreeHost(h->out_mirror);
    h->out_mirror = nullptr, h->out_mirror_cap = 0;
    const size_t want = (size_t)(out_hi - out_lo) + ((out_hi - out_lo) >> 3) + 4096;
    CUDA_TRY(cudaHostAlloc(&h->out_mirror, want, cudaHostAllocPortable));
    h->out_mirror_cap = want;
  }
  uint8_t* const out_h = stage_out ? h->out_mirror - out_lo : out;  // where device -> host copies of messages land
  std::vector<std::pair<uint64_t, uint64_t>> mirrored;              // output pieces to hand on to the caller at the end
  constexpr uint64_t STAGE_PIECE = 16ull << 20;
  // host bytes [lo, lo + len) of the input -> their place in the device staging
  auto copy_in = [&](uint64_t lo, uint64_t len) -> cudaError_t {
    if (!stage_in || len < (256u << 10)) {
      return cudaMemcpyAsync(h->d_in + (lo - in_lo), in + lo, len, cudaMemcpyHostToDevice, h->h2d_stream);
    }
    for (uint64_t o = 0; o < len; o += STAGE_PIECE) {
      const uint64_t piece = std::min(STAGE_PIECE, len - o);
      unsigned slot = 0;
      void* b = h->stager->acquire(piece, slot);
      if (b == nullptr) {
        return cudaErrorMemoryAllocation;
      }
      h->stager->copy(b, in + lo + o, piece);
      cudaError_t e = cudaMemcpyAsync(h->d_in + (lo + o - in_lo), b, piece, cudaMemcpyHostToDevice, h->h2d_stream);
      if (e == cudaSuccess) {
        e = h->stager->release(slot, h->h2d_stream);
      }
      if (e != cudaSuccess) {
        return e;
      }
    }
    return cudaSuccess;
  };
  // the caller's messages (bits a decode does not write must survive) -> device; through the mirror when pageable
  auto copy_preserved = [&](uint64_t lo, uint64_t len) -> cudaError_t {
    if (stage_out) {
      h->stager->copy(out_h + lo, out + lo, len);
    }
    return cudaMemcpyAsync(h->d_out + (lo - out_lo), out_h + lo, len, cudaMemcpyHostToDevice, h->h2d_stream);
  };
  auto copy_out = [&](uint64_t lo, uint64_t len) -> cudaError_t {
    if (stage_out) {
      mirrored.push_back({lo, lo + len});
    }
    return cudaMemcpyAsync(out_h + lo, h->d_out + (lo - out_lo), len, cudaMemcpyDeviceToHost, h->d2h_stream);
  };
  // work list upload + result reset on the main stream come first; the copy streams start behind earlier work
  CUDA_TRY(cudaMemsetAsync(h->d_results, 0, sizeof(CbResult) * n, h->stream));
  if (keep_messages) {
    // messages a codeblock does not write (all-zero input with a CRC) read as zeros in the assembly
    CUDA_TRY(cudaMemsetAsync(h->d_out, 0, out_hi - out_lo, h->stream));
  }
  CUDA_TRY(cudaEventRecord(h->ev_misc, h->stream));
  CUDA_TRY(cudaStreamWaitEvent(h->h2d_stream, h->ev_misc, 0));
  CUDA_TRY(cudaStreamWaitEvent(h->stream2, h->ev_misc, 0));
  // Transport blocks (if any): assembled from the messages in device memory on the copy-out stream and copied back as whole
  // blocks. When the blocks' codeblocks are decoded in block order, groups of blocks are assembled and sent back as soon as
  // the chunk holding their last codeblock has been decoded, so that only the last group remains behind the last chunk.
  TbResult* d_tb_res = nullptr;
  uint8_t* d_tb_bytes = nullptr;
  TbStaged tb_staged;
  uint64_t tb_lo = ~0ull;
  uint32_t tb_next = 0;                 // first block not yet assembled
  std::vector<uint32_t> tb_ready;       // chunk after which block t is complete (empty: assemble everything at the end)
  if (nof_tbs != 0) {
    uint64_t tb_hi = 0;
    for (uint32_t t = 0; t != nof_tbs; ++t) {
      tb_lo = std::min<uint64_t>(tb_lo, tbs[t].tb_offset);
      tb_hi = std::max<uint64_t>(tb_hi, tbs[t].tb_offset + tbs[t].tbs_bits / 8 + (tbs[t].nof_codeblocks > 1 ? 3 : 0));
    }
    const size_t need = (size_t)(tb_hi - tb_lo) + sizeof(TbResult) * nof_tbs + 16;
    if (need > h->d_tb_out_cap) {
      if (h->d_tb_out) {
        CUDA_TRY(cudaDeviceSynchronize());
        CUDA_TRY(cudaFree(h->d_tb_out));
        h->d_tb_out = nullptr, h->d_tb_out_cap = 0;
      }
      CUDA_TRY(cudaMalloc(&h->d_tb_out, need + (need >> 2)));
      h->d_tb_out_cap = need + (need >> 2);
    }
    d_tb_res = reinterpret_cast<TbResult*>(h->d_tb_out);  // results first (aligned), blocks behind them
    d_tb_bytes = h->d_tb_out + ((sizeof(TbResult) * nof_tbs + 15) & ~(size_t)15);
    // work records of every block now, in front of the input copies; the kernels follow group by group (tb_stage)
    rc = tb_stage(h, h->h2d_stream, nof_tbs, tbs, n, descs, tb_staged);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
    CUDA_TRY(cudaStreamWaitEvent(h->d2h_stream, h->ev_tbwork[tb_staged.slot], 0));
    if (chunks.size() > 1) {
      std::vector<uint32_t> chunk_of(n);  // by caller index
      for (size_t ci = 0; ci != chunks.size(); ++ci) {
        for (uint32_t k = chunks[ci].b; k != chunks[ci].e; ++k) {
          chunk_of[h->order[k]] = (uint32_t)ci;
        }
      }
      tb_ready.resize(nof_tbs);
      bool in_order = true;
      for (uint32_t t = 0; t != nof_tbs && in_order; ++t) {
        const ldpc_cuda_tb_desc& tb = tbs[t];
        if (tb.nof_codeblocks == 0 || tb.first_codeblock >= n || tb.nof_codeblocks > n - tb.first_codeblock) {
          in_order = false;  // left to the argument checks of the assembly
          break;
        }
        uint32_t last = 0;
        for (uint32_t r = 0; r != tb.nof_codeblocks; ++r) {
          last = std::max(last, chunk_of[tb.first_codeblock + r]);
        }
        tb_ready[t] = last;
        // groups are copied back as one byte range: blocks must be complete, and laid out, in index order
        in_order = t == 0 || (tb_ready[t] >= tb_ready[t - 1] && tb.tb_offset >= tbs[t - 1].tb_offset);
      }
      if (!in_order) {
        tb_ready.clear();
      }
    }
  }
  // assembles blocks [t0, t1) on the copy-out stream (which has waited for their chunks) and sends them back
  std::vector<uint32_t> tb_calls;  // trace: transport blocks per assembly call
  std::vector<cudaEvent_t> tl_tb;  // trace: end of each assembly call's copy back
  auto assemble_tbs = [&](uint32_t t0, uint32_t t1) -> int {
    if (t0 >= t1) {
      return LDPC_CUDA_OK;
    }
    tb_calls.push_back(t1 - t0);
    int arc = tb_launch(h, h->d2h_stream, tb_staged, tbs, t0, t1, d_out0, h->d_results, d_tb_bytes - tb_lo, d_tb_res);
    if (arc != LDPC_CUDA_OK) {
      cudaStreamSynchronize(h->d2h_stream);
      cudaStreamSynchronize(h->stream2);
      cudaStreamSynchronize(h->stream);
      return arc;
    }
    // block by block (exactly adjacent blocks as one copy): the bytes between two blocks of this call may belong to
    // somebody else, e.g. to the blocks another device of a pool assembles into the same buffer
    // blocks of one size at a constant stride (the usual layout of a slot: aligned offsets leave gaps): ONE pitched copy
    // instead of one copy per block (5 us each: 37 full-slot blocks took 300 us instead of 120)
    if (t1 - t0 >= 3) {
      const uint64_t w0 = tbs[t0].tbs_bits / 8 + (tbs[t0].nof_codeblocks > 1 ? 3 : 0);
      const uint64_t stride = tbs[t0 + 1].tb_offset - tbs[t0].tb_offset;
      bool uniform = tbs[t0 + 1].tb_offset > tbs[t0].tb_offset && stride >= w0;
      for (uint32_t t = t0 + 1; t != t1 && uniform; ++t) {
        uniform = tbs[t].tb_offset == tbs[t0].tb_offset + (t - t0) * stride &&
                  tbs[t].tbs_bits / 8 + (tbs[t].nof_codeblocks > 1 ? 3 : 0) == w0;
      }
      if (uniform && stride != w0) {
        CUDA_TRY(cudaMemcpy2DAsync(tb_out + tbs[t0].tb_offset, stride, d_tb_bytes + (tbs[t0].tb_offset - tb_lo), stride, w0,
                                   t1 - t0, cudaMemcpyDeviceToHost, h->d2h_stream));
        tb_next = t1;
        if (trace_chunks) {
          cudaEvent_t e = nullptr;
          cudaEventCreate(&e);
          cudaEventRecord(e, h->d2h_stream);
          tl_tb.push_back(e);
        }
        return LDPC_CUDA_OK;
      }
    }
    uint64_t lo = 0, hi = 0;
    for (uint32_t t = t0; t != t1; ++t) {
      const uint64_t b = tbs[t].tb_offset, e = b + tbs[t].tbs_bits / 8 + (tbs[t].nof_codeblocks > 1 ? 3 : 0);
      if (hi != lo && b == hi) {
        hi = e;
        continue;
      }
      if (hi != lo) {
        CUDA_TRY(cudaMemcpyAsync(tb_out + lo, d_tb_bytes + (lo - tb_lo), hi - lo, cudaMemcpyDeviceToHost, h->d2h_stream));
      }
      lo = b, hi = e;
    }
    if (hi != lo) {
      CUDA_TRY(cudaMemcpyAsync(tb_out + lo, d_tb_bytes + (lo - tb_lo), hi - lo, cudaMemcpyDeviceToHost, h->d2h_stream));
    }
    tb_next = t1;
    if (trace_chunks) {
      cudaEvent_t e = nullptr;
      cudaEventCreate(&e);
      cudaEventRecord(e, h->d2h_stream);
      tl_tb.push_back(e);
    }
    return LDPC_CUDA_OK;
  };
  const uint32_t tb_group = std::max<uint32_t>(1, (nof_tbs + 3) / 4);  // at most ~4 early groups
  mark(h->h2d_stream);
  for (size_t ci = 0; ci != chunks.size(); ++ci) {
    const Chunk& c = chunks[ci];
    fill_work(h, work_slot, c.b, c.e, descs, fused);
    CUDA_TRY(cudaMemcpyAsync(h->d_work + c.b, h->h_work[work_slot] + c.b, sizeof(CbWork) * (c.e - c.b), cudaMemcpyHostToDevice,
                             h->h2d_stream));
    if (ci + 1 == chunks.size()) {
      CUDA_TRY(cudaEventRecord(h->ev_work[work_slot], h->h2d_stream));
    }
    if (c.in_rows != 0) {
      const uint64_t first = descs[h->order[c.b]].in_offset;
      CUDA_TRY(copy_rows(h->d_in + (first - in_lo), in + first, c.in_stride, c.in_width, c.in_rows, cudaMemcpyHostToDevice,
                         h->h2d_stream));
    } else if (c.in_runs.empty()) {
      if (c.in_hi > c.in_lo) {
        CUDA_TRY(copy_in(c.in_lo, c.in_hi - c.in_lo));
      }
    } else {
      for (const auto& run : c.in_runs) {
        CUDA_TRY(copy_in(run.first, run.second - run.first));
      }
    }
    if (preserve) {
      if (c.out_rows != 0) {
        const uint64_t first = descs[h->order[c.b]].out_offset;
        CUDA_TRY(copy_rows(h->d_out + (first - out_lo), out + first, c.out_stride, c.out_width, c.out_rows,
                           cudaMemcpyHostToDevice, h->h2d_stream));
      } else if (!c.out_runs.empty()) {
        for (const auto& run : c.out_runs) {
          CUDA_TRY(copy_preserved(run.first, run.second - run.first));
        }
      } else {
        CUDA_TRY(copy_preserved(c.out_lo, c.out_hi - c.out_lo));
      }
    }
    CUDA_TRY(cudaEventRecord(h->ev_in[ci], h->h2d_stream));
    mark(h->h2d_stream);
    // kernels of consecutive chunks alternate between two streams (and two work counters) so that the blocks of chunk
    // c + 1 fill the SMs as the last blocks of chunk c retire
    cudaStream_t ks = (ci & 1) ? h->stream2 : h->stream;
    CUDA_TRY(cudaStreamWaitEvent(ks, h->ev_in[ci], 0));
    // a chunk spans several groups only in the single-chunk fallback
    for (size_t g = 0; g + 1 < h->group_begin.size(); ++g) {
      const uint32_t b = std::max(c.b, h->group_begin[g]), e = std::min(c.e, h->group_begin[g + 1]);
      if (b >= e) {
        continue;
      }
      const ldpc_cuda_cb_desc& dg = descs[h->order[b]];
      rc = launch_decode_group(h, dg.base_graph, dg.lifting_size, h->d_work + b, e - b, fused ? h->arena : d_in0,
                               fused ? d_in0 : nullptr, d_out0, h->d_results, ks, h->pair_counter + (ci & 1));
      if (rc != LDPC_CUDA_OK) {
        return rc;
      }
    }
    CUDA_TRY(cudaEventRecord(h->ev_done[ci], ks));
    mark(ks);
    CUDA_TRY(cudaStreamWaitEvent(h->d2h_stream, h->ev_done[ci], 0));
    if (!keep_messages) {
      if (c.out_rows != 0) {
        const uint64_t first = descs[h->order[c.b]].out_offset;
        CUDA_TRY(copy_rows(out + first, h->d_out + (first - out_lo), c.out_stride, c.out_width, c.out_rows,
                           cudaMemcpyDeviceToHost, h->d2h_stream));
      } else if (!c.out_runs.empty()) {
        for (const auto& run : c.out_runs) {
          CUDA_TRY(copy_out(run.first, run.second - run.first));
        }
      } else {
        CUDA_TRY(copy_out(c.out_lo, c.out_hi - c.out_lo));
      }
    }
    if (!tb_ready.empty() && ci + 1 != chunks.size()) {
      uint32_t t1 = tb_next;
      while (t1 != nof_tbs && tb_ready[t1] <= ci) {
        ++t1;
      }
      if (t1 - tb_next >= tb_group) {
        rc = assemble_tbs(tb_next, t1);
        if (rc != LDPC_CUDA_OK) {
          return rc;
        }
      }
    }
  }
  if (nof_tbs != 0) {
    // transport blocks not assembled behind their own chunk (see above) are assembled once every chunk has been decoded
    // (the copy-out stream already waits for all of them)
    rc = assemble_tbs(tb_next, nof_tbs);
    if (rc != LDPC_CUDA_OK) {
      return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(tb_results, d_tb_res, sizeof(TbResult) * nof_tbs, cudaMemcpyDeviceToHost, h->d2h_stream));
  }
  CUDA_TRY(cudaMemcpyAsync(results, h->d_results, sizeof(CbResult) * n, cudaMemcpyDeviceToHost, h->d2h_stream));
  mark(h->d2h_stream);
  const auto t_enqueued = std::chrono::steady_clock::now();
  CUDA_TRY(cudaStreamSynchronize(h->d2h_stream));
  // everything of this call has finished (the last copy waited for every kernel); later calls may use any stream
  CUDA_TRY(cudaStreamSynchronize(h->stream2));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  for (const auto& piece : mirrored) {
    h->stager->copy(out + piece.first, out_h + piece.first, piece.second - piece.first);
  }
  if (trace) {
    const auto t_done = std::chrono::steady_clock::now();
    auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::micro>(b - a).count();
    };
    size_t runs = 0;
    for (const Chunk& c : chunks) {
      runs += c.in_runs.size() + c.out_runs.size();
    }
    std::fprintf(stderr,
                 "ldpc_cuda trace: %u codeblocks, %u transport blocks, %zu chunks (%zu copy runs), in %llu B, preserve %d: "
                 "validate + group %.0f us, plan %.0f us, enqueue %.0f us, wait %.0f us\n",
                 n, nof_tbs, chunks.size(), runs, (unsigned long long)(in_hi - in_lo), (int)preserve, us(t_enter, t_staged),
                 us(t_staged, t_planned), us(t_planned, t_enqueued), us(t_enqueued, t_done));
    if (trace_chunks && tl.size() == 2 + 2 * chunks.size()) {
      float end = 0;
      cudaEventElapsedTime(&end, tl[0], tl.back());
      std::fprintf(stderr, "ldpc_cuda trace: last result copied back at %.0f us; transport blocks per assembly call:", end * 1e3f);
      for (size_t k = 0; k != tb_calls.size(); ++k) {
        float at = 0;
        if (k < tl_tb.size()) {
          cudaEventElapsedTime(&at, tl[0], tl_tb[k]);
        }
        std::fprintf(stderr, " %u (back at %.0f us)", tb_calls[k], at * 1e3f);
      }
      std::fprintf(stderr, "; chunk (codeblocks): input copied / decoded at us:");
      for (size_t ci = 0; ci != chunks.size(); ++ci) {
        float a = 0, b = 0;
        cudaEventElapsedTime(&a, tl[0], tl[1 + 2 * ci]);
        cudaEventElapsedTime(&b, tl[0], tl[2 + 2 * ci]);
        std::fprintf(stderr, " %zu (%u): %.0f / %.0f;", ci, chunks[ci].e - chunks[ci].b, a * 1e3f, b * 1e3f);
      }
      std::fprintf(stderr, "\n");
    }
    for (cudaEvent_t e : tl) {
      cudaEventDestroy(e);
    }
    for (cudaEvent_t e : tl_tb) {
      cudaEventDestroy(e);
    }
  }
  return LDPC_CUDA_OK;
}

int ldpc_cuda_decode_batch(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* llrs,
                           uint8_t* out_bits, ldpc_cuda_cb_result* results)
{
  return batch_host(h, n, descs, llrs, out_bits, results, false);
}

int ldpc_cuda_dematch_decode_batch(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* rm_llrs,
                                   uint8_t* out_bits, ldpc_cuda_cb_result* results)
{
  return batch_host(h, n, descs, rm_llrs, out_bits, results, true);
}

// ---- transport-block assembly + TB CRC24A on the device (pusch_decoder_impl.cpp:156-222) --------------------------------
int ldpc_cuda_dematch_decode_tb_batch(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const int8_t* rm_llrs,
                                      uint8_t* out_bits, ldpc_cuda_cb_result* results, uint32_t nof_tbs,
                                      const ldpc_cuda_tb_desc* tbs, uint8_t* tb_out, ldpc_cuda_tb_result* tb_results)
{
  if (nof_tbs == 0 || n == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  return batch_host(h, n, descs, rm_llrs, out_bits, results, true, nof_tbs, tbs, tb_out, tb_results);
}

int ldpc_cuda_tb_assemble_device(ldpc_cuda_handle* h, uint32_t nof_tbs, const ldpc_cuda_tb_desc* tbs, uint32_t n,
                                 const ldpc_cuda_cb_desc* descs, const uint8_t* out_bits_d,
                                 const ldpc_cuda_cb_result* results_d, uint8_t* tb_out_d, ldpc_cuda_tb_result* tb_results_d)
{
  if (h == nullptr || tbs == nullptr || descs == nullptr || out_bits_d == nullptr || results_d == nullptr ||
      tb_out_d == nullptr || tb_results_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_tbs == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  return enqueue_tb_assemble(h, h->stream, nof_tbs, tbs, n, descs, out_bits_d, reinterpret_cast<const CbResult*>(results_d),
                             tb_out_d, reinterpret_cast<TbResult*>(tb_results_d));
}

// ---- transport-block assembly without any allocation -------------------------------------------------------------------
// Same kernels as ldpc_cuda_tb_assemble_device, but the work records and the per-chunk partial remainders live in ONE
// file-scope device array of fixed size instead of the handle's growable cudaMalloc'd buffer, and the records are staged
// through a fixed host array instead of pinned memory that grows on demand: no cudaMalloc / cudaFree / cudaHostAlloc on
// any call. Every handle of a device shares the array, and it keeps the previous call's contents.
static_assert(LDPC_CUDA_NOALLOC_MAX_TBS <= 65535u, "transport blocks index blockIdx.y");
constexpr size_t NOALLOC_OFF_CB = sizeof(TbWork) * LDPC_CUDA_NOALLOC_MAX_TBS;
constexpr size_t NOALLOC_OFF_PARTIAL = (NOALLOC_OFF_CB + sizeof(TbCb) * LDPC_CUDA_NOALLOC_MAX_CBS + 255) & ~(size_t)255;
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_OFF_PARTIAL + sizeof(unsigned) * LDPC_CUDA_NOALLOC_MAX_CHUNKS;
static __device__ __align__(256) unsigned char g_tb_noalloc_scratch[NOALLOC_SCRATCH_BYTES];  // ~100 KB per GPU
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex s_noalloc_mutex;                                 // guards the two host-side objects below
static unsigned char* s_noalloc_dev[NOALLOC_MAX_DEVICES] = {};     // device address of the array, per device
static unsigned char s_noalloc_stage[NOALLOC_OFF_PARTIAL];         // host staging of the work records (pageable)

// In host code `g_tb_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address a kernel can
// use is resolved once per device, with that device current.
static int noalloc_scratch_of(int device, unsigned char** ptr)
{
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  if (s_noalloc_dev[device] == nullptr) {
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_tb_noalloc_scratch));
    s_noalloc_dev[device] = static_cast<unsigned char*>(p);
  }
  *ptr = s_noalloc_dev[device];
  return LDPC_CUDA_OK;
}

int ldpc_cuda_tb_assemble_device_noalloc(ldpc_cuda_handle* h, uint32_t nof_tbs, const ldpc_cuda_tb_desc* tbs, uint32_t n,
                                         const ldpc_cuda_cb_desc* descs, const uint8_t* out_bits_d,
                                         const ldpc_cuda_cb_result* results_d, uint8_t* tb_out_d,
                                         ldpc_cuda_tb_result* tb_results_d)
{
  if (h == nullptr || tbs == nullptr || descs == nullptr || out_bits_d == nullptr || results_d == nullptr ||
      tb_out_d == nullptr || tb_results_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_tbs == 0) {
    return LDPC_CUDA_OK;
  }
  uint64_t total_chunks = 0;
  for (uint32_t t = 0; t != nof_tbs; ++t) {
    total_chunks += ((uint64_t)tbs[t].tbs_bits / 8 + 3 + TB_CHUNK - 1) / TB_CHUNK;
  }
  if (nof_tbs > LDPC_CUDA_NOALLOC_MAX_TBS || n > LDPC_CUDA_NOALLOC_MAX_CBS || total_chunks > LDPC_CUDA_NOALLOC_MAX_CHUNKS) {
    std::fprintf(stderr,
                 "ldpc_cuda_tb_assemble_device_noalloc: %u transport blocks / %u codeblocks / %llu chunks of %d bytes exceed "
                 "the fixed scratch (%u / %u / %u): use ldpc_cuda_tb_assemble_device or split the call\n",
                 nof_tbs, n, (unsigned long long)total_chunks, TB_CHUNK, LDPC_CUDA_NOALLOC_MAX_TBS,
                 LDPC_CUDA_NOALLOC_MAX_CBS, LDPC_CUDA_NOALLOC_MAX_CHUNKS);
    return LDPC_CUDA_ERR_CAPACITY;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  std::lock_guard<std::mutex> lock(s_noalloc_mutex);
  unsigned char* scratch = nullptr;
  int rc = noalloc_scratch_of(h->device, &scratch);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  // work records, validated like ldpc_cuda_tb_assemble_device's
  const size_t bytes = NOALLOC_OFF_CB + sizeof(TbCb) * n;
  std::memset(s_noalloc_stage, 0, bytes);
  TbWork* tw = reinterpret_cast<TbWork*>(s_noalloc_stage);
  TbCb* tc = reinterpret_cast<TbCb*>(s_noalloc_stage + NOALLOC_OFF_CB);
  unsigned max_chunks = 0, chunk_base = 0;
  for (uint32_t t = 0; t != nof_tbs; ++t) {
    const ldpc_cuda_tb_desc& tb = tbs[t];
    if (tb.nof_codeblocks == 0 || tb.first_codeblock > n || tb.nof_codeblocks > n - tb.first_codeblock ||
        tb.tbs_bits == 0 || tb.tbs_bits % 8 != 0 || (tb.nof_cb_crc_bits != 16 && tb.nof_cb_crc_bits != 24) ||
        (tb.nof_codeblocks > 1 && tb.nof_cb_crc_bits != 24)) {
      return LDPC_CUDA_ERR_ARG;
    }
    const ldpc_cuda_cb_desc& d0 = descs[tb.first_codeblock];
    if ((d0.base_graph != 1 && d0.base_graph != 2) || lifting_set_of(d0.lifting_size) < 0) {
      return LDPC_CUDA_ERR_ARG;
    }
    const unsigned K = bg_K(d0.base_graph) * d0.lifting_size;
    if (K <= tb.nof_cb_crc_bits + d0.nof_filler_bits) {
      return LDPC_CUDA_ERR_ARG;
    }
    const unsigned data_bits = K - tb.nof_cb_crc_bits - d0.nof_filler_bits;
    const uint64_t total = (uint64_t)tb.tbs_bits + (tb.nof_codeblocks > 1 ? 24u : 0u);
    if ((uint64_t)data_bits * tb.nof_codeblocks < total || total / 8 > 0x7fffffffu) {
      return LDPC_CUDA_ERR_ARG;
    }
    for (uint32_t r = 0; r != tb.nof_codeblocks; ++r) {
      const ldpc_cuda_cb_desc& d = descs[tb.first_codeblock + r];
      if (d.base_graph != d0.base_graph || d.lifting_size != d0.lifting_size || d.nof_filler_bits != d0.nof_filler_bits) {
        return LDPC_CUDA_ERR_ARG;
      }
      TbCb& c = tc[tb.first_codeblock + r];
      c.src_offset = d.out_offset;
      c.already_ok = d.skip_decode ? 1u : 0u;
      c.result_index = tb.first_codeblock + r;
    }
    tw[t].tb_offset = tb.tb_offset;
    tw[t].first_cb = tb.first_codeblock;
    tw[t].nof_cbs = tb.nof_codeblocks;
    tw[t].data_bits = data_bits;
    tw[t].nbytes = (unsigned)(total / 8);
    tw[t].chunk_base = chunk_base;
    const unsigned chunks = (tw[t].nbytes + TB_CHUNK - 1) / TB_CHUNK;
    chunk_base += chunks;
    max_chunks = std::max(max_chunks, chunks);
  }
  // Stream order keeps this behind the kernels of an earlier call on the same stream that still read the array. The
  // source is pageable memory: the copy has left s_noalloc_stage when the call returns.
  CUDA_TRY(cudaMemcpyAsync(scratch, s_noalloc_stage, bytes, cudaMemcpyHostToDevice, h->stream));
  unsigned* d_partial = reinterpret_cast<unsigned*>(scratch + NOALLOC_OFF_PARTIAL);
  // the array holds the last call's remainders: the ones this call reads start from zero
  CUDA_TRY(cudaMemsetAsync(d_partial, 0, sizeof(unsigned) * chunk_base, h->stream));
  const TbWork* d_tw = reinterpret_cast<const TbWork*>(scratch);
  const TbCb* d_tc = reinterpret_cast<const TbCb*>(scratch + NOALLOC_OFF_CB);
  const dim3 grid(max_chunks, nof_tbs);
  ldpc_tb_assemble_kernel<<<grid, TB_THREADS, 0, h->stream>>>(d_tw, d_tc, out_bits_d, tb_out_d, d_partial);
  CUDA_TRY(cudaGetLastError());
  ldpc_tb_finish_kernel<<<nof_tbs, 128, 0, h->stream>>>(d_tw, d_tc, reinterpret_cast<const CbResult*>(results_d), d_partial,
                                                        reinterpret_cast<TbResult*>(tb_results_d));
  CUDA_TRY(cudaGetLastError());
  h->launches += 2;
  return LDPC_CUDA_OK;
}

// ---- soft demodulator + descrambler (demodulation_mapper::demodulate_soft, pusch_demodulator_impl::descramble) -----------
int ldpc_cuda_demodulate_batch_device(ldpc_cuda_handle* h, uint32_t nof_jobs, const ldpc_cuda_demod_desc* jobs,
                                      const float* symbols_d, const float* noise_vars_d, const int8_t* in_llrs_d,
                                      int8_t* llrs_d)
{
  if (h == nullptr || jobs == nullptr || llrs_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_jobs == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  return enqueue_demod(h, h->stream, nof_jobs, jobs, symbols_d, noise_vars_d, in_llrs_d, llrs_d);
}

// Host-buffer mirrors: inputs through the handle's input staging, LLRs through its output staging.
static int demod_host(ldpc_cuda_handle* h, const ldpc_cuda_demod_desc& job, const void* in0, size_t in0_bytes,
                      const void* in1, size_t in1_bytes, int8_t* out, size_t out_bytes)
{
  CUDA_TRY(cudaSetDevice(h->device));
  const size_t off1 = (in0_bytes + 255) & ~(size_t)255;
  int rc = ensure_capacity(h, off1 + in1_bytes, out_bytes);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  CUDA_TRY(cudaMemcpyAsync(h->d_in, in0, in0_bytes, cudaMemcpyHostToDevice, h->stream));
  if (in1_bytes) {
    CUDA_TRY(cudaMemcpyAsync(h->d_in + off1, in1, in1_bytes, cudaMemcpyHostToDevice, h->stream));
  }
  rc = enqueue_demod(h, h->stream, 1, &job, reinterpret_cast<const float*>(h->d_in),
                     reinterpret_cast<const float*>(h->d_in + off1), h->d_in, reinterpret_cast<int8_t*>(h->d_out));
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  CUDA_TRY(cudaMemcpyAsync(out, h->d_out, out_bytes, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_demodulate_soft(ldpc_cuda_handle* h, int mod_order, int pi2_bpsk, uint32_t nof_symbols, const float* symbols,
                              const float* noise_vars, int8_t* llrs)
{
  if (h == nullptr || symbols == nullptr || noise_vars == nullptr || llrs == nullptr || mod_order < 1 || mod_order > 8) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_symbols == 0) {
    return LDPC_CUDA_OK;
  }
  ldpc_cuda_demod_desc job = {};
  job.mod_order = (uint8_t)mod_order, job.pi2_bpsk = pi2_bpsk ? 1 : 0, job.nof_symbols = nof_symbols;
  return demod_host(h, job, symbols, sizeof(float) * 2 * (size_t)nof_symbols, noise_vars, sizeof(float) * (size_t)nof_symbols,
                    llrs, (size_t)nof_symbols * mod_order);
}

int ldpc_cuda_descramble(ldpc_cuda_handle* h, uint32_t c_init, uint32_t nof_llrs, const int8_t* in, int8_t* out)
{
  if (h == nullptr || in == nullptr || out == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_llrs == 0) {
    return LDPC_CUDA_OK;
  }
  ldpc_cuda_demod_desc job = {};
  job.mod_order = 0, job.descramble = 1, job.nof_symbols = nof_llrs, job.c_init = c_init;
  return demod_host(h, job, in, nof_llrs, nullptr, 0, out, nof_llrs);
}

int ldpc_cuda_descramble_uci(ldpc_cuda_handle* h, uint32_t c_init, uint32_t nof_llrs, int mod_order, uint32_t nof_layers,
                             uint32_t nof_re, const uint16_t* re_indexes, const int8_t* in, int8_t* out)
{
  if (h == nullptr || in == nullptr || out == nullptr || nof_re > 0xffffu || nof_layers == 0 || nof_layers > 8 ||
      (nof_re != 0 && re_indexes == nullptr)) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_llrs == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  int rc = ensure_capacity(h, nof_llrs, nof_llrs);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  ldpc_cuda_demod_desc job = {};
  job.mod_order = 0, job.descramble = 1, job.nof_symbols = nof_llrs, job.c_init = c_init;
  ldpc_cuda_placeholder_desc ph = {};
  ph.c_init = c_init, ph.nof_llrs = nof_llrs, ph.nof_re = (uint16_t)nof_re, ph.mod_order = (uint8_t)mod_order;
  ph.nof_layers = (uint8_t)nof_layers;
  CUDA_TRY(cudaMemcpyAsync(h->d_in, in, nof_llrs, cudaMemcpyHostToDevice, h->stream));
  rc = enqueue_demod(h, h->stream, 1, &job, nullptr, nullptr, h->d_in, reinterpret_cast<int8_t*>(h->d_out));
  if (rc == LDPC_CUDA_OK) {
    rc = enqueue_placeholders(h, h->stream, 1, &ph, re_indexes, nof_re, reinterpret_cast<int8_t*>(h->d_out));
  }
  if (rc != LDPC_CUDA_OK) {
    cudaStreamSynchronize(h->stream);
    return rc;
  }
  CUDA_TRY(cudaMemcpyAsync(out, h->d_out, nof_llrs, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_descramble_placeholders_device(ldpc_cuda_handle* h, uint32_t nof_jobs, const ldpc_cuda_placeholder_desc* jobs,
                                             const uint16_t* re_indexes, uint32_t nof_re_indexes, int8_t* llrs_d)
{
  if (h == nullptr || (nof_jobs != 0 && (jobs == nullptr || llrs_d == nullptr))) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (nof_jobs == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  return enqueue_placeholders(h, h->stream, nof_jobs, jobs, re_indexes, nof_re_indexes, llrs_d);
}

// ---- UL-SCH demultiplexing (ulsch_demultiplex_impl.cpp:316-453) ------------------------------------------------------------
static UlschGeometry ulsch_geometry(const ldpc_cuda_ulsch_desc& d)
{
  UlschGeometry g;
  g.mod_order = d.mod_order, g.nof_layers = d.nof_layers, g.nof_prb = d.nof_prb, g.start_symbol = d.start_symbol_index;
  g.nof_symbols = d.nof_symbols, g.dmrs_type = d.dmrs_type, g.nof_cdm_groups_without_data = d.nof_cdm_groups_without_data;
  g.dmrs_symbol_mask = d.dmrs_symbol_mask, g.nof_harq_ack_rvd = d.nof_harq_ack_rvd;
  g.nof_enc_harq_ack = d.nof_enc_harq_ack_bits, g.nof_enc_csi1 = d.nof_enc_csi_part1_bits;
  g.nof_enc_csi2 = d.nof_enc_csi_part2_bits;
  return g;
}

static bool ulsch_valid(const ldpc_cuda_ulsch_desc& d)
{
  const bool mod_ok = d.mod_order == 1 || d.mod_order == 2 || d.mod_order == 4 || d.mod_order == 6 || d.mod_order == 8;
  return mod_ok && d.nof_layers >= 1 && d.nof_layers <= 4 && d.nof_prb >= 1 && d.nof_prb <= 275 && d.nof_symbols >= 1;
}

int ldpc_cuda_ulsch_sizes(const ldpc_cuda_ulsch_desc* desc, uint32_t* nof_input_bits, uint32_t* nof_sch_bits)
{
  if (desc == nullptr || !ulsch_valid(*desc)) {
    return LDPC_CUDA_ERR_ARG;
  }
  std::vector<UlschSymbol> plan;
  UlschTotals tot;
  if (!ulsch_plan(ulsch_geometry(*desc), 0, plan, tot)) {
    return LDPC_CUDA_ERR_ARG;
  }
  const unsigned bpr = (unsigned)desc->mod_order * desc->nof_layers;
  if (nof_input_bits) {
    *nof_input_bits = tot.in_re * bpr;
  }
  if (nof_sch_bits) {
    *nof_sch_bits = tot.sch_re * bpr;
  }
  return LDPC_CUDA_OK;
}

int ldpc_cuda_ulsch_placeholders(const ldpc_cuda_ulsch_desc* desc, uint32_t nof_harq_ack_bits, uint32_t nof_csi_part1_bits,
                                 uint32_t nof_csi_part2_bits, uint16_t* re_indexes, uint32_t max_indexes)
{
  if (desc == nullptr || !ulsch_valid(*desc) || (re_indexes == nullptr && max_indexes != 0)) {
    return LDPC_CUDA_ERR_ARG;
  }
  // no placeholders below QPSK, nor when no field carries exactly one bit (ulsch_demultiplex_impl.cpp:399-408)
  if (desc->mod_order < 2 || (nof_harq_ack_bits != 1 && nof_csi_part1_bits != 1 && nof_csi_part2_bits != 1)) {
    return 0;
  }
  std::vector<UlschSymbol> plan;
  UlschTotals tot;
  if (!ulsch_plan(ulsch_geometry(*desc), 0, plan, tot)) {
    return LDPC_CUDA_ERR_ARG;
  }
  // An element's index in the placeholder list counts the elements that consume input before it, which is every element:
  // the zero-filled element underneath a punctured reserved one does not advance the counter (:413-445).
  uint32_t count = 0;
  for (const UlschSymbol& s : plan) {
    if (!(s.rvd_n | s.ack_n | s.csi1_n | s.csi2_n)) {
      continue;
    }
    for (unsigned i = 0; i != s.n_re; ++i) {
      const UlschWhere w = ulsch_classify(s, i);
      const bool one = (w.field == ULSCH_ACK && nof_harq_ack_bits == 1) || (w.field == ULSCH_CSI1 && nof_csi_part1_bits == 1) ||
                       (w.field == ULSCH_CSI2 && nof_csi_part2_bits == 1);
      if (one) {
        if (count < max_indexes) {
          re_indexes[count] = (uint16_t)(s.in_base + i);
        }
        ++count;
      }
    }
  }
  return (int)count;
}

// Plans every codeword, uploads symbols + jobs through the demodulator's pinned staging and launches the kernel.
static int enqueue_ulsch(ldpc_cuda_handle* h, cudaStream_t stream, uint32_t n, const ldpc_cuda_ulsch_desc* descs,
                         const int8_t* d_in, int8_t* d_sch, int8_t* d_uci)
{
  std::vector<UlschSymbol> plan;
  std::vector<UlschJob> jobs(n);
  unsigned max_re = 0;
  for (uint32_t k = 0; k != n; ++k) {
    const ldpc_cuda_ulsch_desc& d = descs[k];
    UlschTotals tot;
    if (!ulsch_valid(d) || !ulsch_plan(ulsch_geometry(d), k, plan, tot)) {
      return LDPC_CUDA_ERR_ARG;
    }
    jobs[k].in_offset = d.in_offset, jobs[k].sch_offset = d.sch_offset, jobs[k].ack_offset = d.harq_ack_offset;
    jobs[k].csi1_offset = d.csi_part1_offset, jobs[k].csi2_offset = d.csi_part2_offset;
    jobs[k].bits_per_re = (unsigned)d.mod_order * d.nof_layers, jobs[k].reserved = 0;
  }
  if (plan.empty()) {
    return LDPC_CUDA_OK;
  }
  if (plan.size() > 65535u) {
    return LDPC_CUDA_ERR_CAPACITY;
  }
  for (const UlschSymbol& s : plan) {
    max_re = std::max(max_re, s.n_re);
  }
  const size_t sym_bytes = sizeof(UlschSymbol) * plan.size(), off_jobs = (sym_bytes + 15) & ~(size_t)15;
  const size_t bytes = off_jobs + sizeof(UlschJob) * n;
  unsigned slot = 0;
  const int rc = demod_stage(h, bytes, slot);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  std::memcpy(h->h_demod[slot], plan.data(), sym_bytes);
  std::memcpy(h->h_demod[slot] + off_jobs, jobs.data(), sizeof(UlschJob) * n);
  CUDA_TRY(cudaMemcpyAsync(h->d_demod, h->h_demod[slot], bytes, cudaMemcpyHostToDevice, stream));
  CUDA_TRY(cudaEventRecord(h->ev_demod[slot], stream));
  const dim3 grid((max_re + 255) / 256, (unsigned)plan.size());
  ldpc_ulsch_demux_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const UlschSymbol*>(h->d_demod),
                                                     reinterpret_cast<const UlschJob*>(h->d_demod + off_jobs), d_in, d_sch,
                                                     d_uci);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  return LDPC_CUDA_OK;
}

int ldpc_cuda_ulsch_demultiplex_batch_device(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_ulsch_desc* descs,
                                             const int8_t* in_llrs_d, int8_t* sch_d, int8_t* uci_d)
{
  if (h == nullptr || descs == nullptr || in_llrs_d == nullptr || sch_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  for (uint32_t k = 0; k != n; ++k) {
    if (uci_d == nullptr && (descs[k].nof_enc_harq_ack_bits | descs[k].nof_enc_csi_part1_bits | descs[k].nof_enc_csi_part2_bits)) {
      return LDPC_CUDA_ERR_ARG;
    }
  }
  CUDA_TRY(cudaSetDevice(h->device));
  return enqueue_ulsch(h, h->stream, n, descs, in_llrs_d, sch_d, uci_d);
}

int ldpc_cuda_ulsch_demultiplex(ldpc_cuda_handle* h, const ldpc_cuda_ulsch_desc* desc, const int8_t* input, int8_t* sch_data,
                                int8_t* harq_ack, int8_t* csi_part1, int8_t* csi_part2)
{
  if (h == nullptr || desc == nullptr || input == nullptr || sch_data == nullptr ||
      (desc->nof_enc_harq_ack_bits != 0 && harq_ack == nullptr) || (desc->nof_enc_csi_part1_bits != 0 && csi_part1 == nullptr) ||
      (desc->nof_enc_csi_part2_bits != 0 && csi_part2 == nullptr)) {
    return LDPC_CUDA_ERR_ARG;
  }
  uint32_t in_bits = 0, sch_bits = 0;
  int rc = ldpc_cuda_ulsch_sizes(desc, &in_bits, &sch_bits);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  // device staging: input | then outputs SCH, HARQ-ACK, CSI 1, CSI 2 one after the other
  const size_t a = desc->nof_enc_harq_ack_bits, c1 = desc->nof_enc_csi_part1_bits, c2 = desc->nof_enc_csi_part2_bits;
  rc = ensure_capacity(h, in_bits, sch_bits + a + c1 + c2 + 16);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  ldpc_cuda_ulsch_desc d = *desc;
  d.in_offset = 0, d.sch_offset = 0, d.harq_ack_offset = sch_bits, d.csi_part1_offset = sch_bits + a;
  d.csi_part2_offset = sch_bits + a + c1;
  int8_t* d_out = reinterpret_cast<int8_t*>(h->d_out);
  CUDA_TRY(cudaMemcpyAsync(h->d_in, input, in_bits, cudaMemcpyHostToDevice, h->stream));
  rc = enqueue_ulsch(h, h->stream, 1, &d, h->d_in, d_out, d_out);
  if (rc != LDPC_CUDA_OK) {
    cudaStreamSynchronize(h->stream);
    return rc;
  }
  CUDA_TRY(cudaMemcpyAsync(sch_data, d_out, sch_bits, cudaMemcpyDeviceToHost, h->stream));
  if (a) {
    CUDA_TRY(cudaMemcpyAsync(harq_ack, d_out + d.harq_ack_offset, a, cudaMemcpyDeviceToHost, h->stream));
  }
  if (c1) {
    CUDA_TRY(cudaMemcpyAsync(csi_part1, d_out + d.csi_part1_offset, c1, cudaMemcpyDeviceToHost, h->stream));
  }
  if (c2) {
    CUDA_TRY(cudaMemcpyAsync(csi_part2, d_out + d.csi_part2_offset, c2, cudaMemcpyDeviceToHost, h->stream));
  }
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

// ---- drop-in single codeblock ---------------------------------------------------------------------------------------
int ldpc_cuda_decode(ldpc_cuda_handle* h, const ldpc_cuda_cb_desc* desc, const int8_t* llrs, uint8_t* out_bits,
                     uint32_t* iterations)
{
  if (desc == nullptr || iterations == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  ldpc_cuda_cb_desc d = *desc;
  d.in_offset = 0, d.out_offset = 0, d.skip_decode = 0;
  d.early_stop = 1;  // ldpc_decoder::decode with a CRC calculator always stops early (ldpc_decoder_impl.cpp:125-133)
  ldpc_cuda_cb_result r = {0, 0, 0};
  int rc = batch_host(h, 1, &d, llrs, out_bits, &r, false);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  *iterations = r.crc_ok ? r.iterations : 0;
  return LDPC_CUDA_OK;
}

int ldpc_cuda_rate_dematch(ldpc_cuda_handle* h, const ldpc_cuda_cb_desc* desc, int8_t* soft_inout, const int8_t* in)
{
  if (h == nullptr || desc == nullptr || soft_inout == nullptr || (in == nullptr && desc->rm_length != 0)) {
    return LDPC_CUDA_ERR_ARG;
  }
  int rc = validate_dematch_desc(*desc);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  const size_t N = (size_t)bg_N_short(desc->base_graph) * desc->lifting_size;
  rc = ensure_capacity(h, N + desc->rm_length + 64, 16);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  int8_t* d_soft = h->d_in;
  int8_t* d_rm = h->d_in + ((N + 63) & ~(size_t)63);
  rc = ensure_capacity(h, ((N + 63) & ~(size_t)63) + desc->rm_length + 64, 16);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  d_soft = h->d_in, d_rm = h->d_in + ((N + 63) & ~(size_t)63);
  CbWork& w = h->h_work[0][0];  // the stream is idle (synchronised above), both slots are free
  std::memset(&w, 0, sizeof(w));
  w.rm_length = desc->rm_length, w.Nref = desc->Nref, w.nof_filler_bits = desc->nof_filler_bits;
  w.rv = desc->rv, w.mod_order = desc->mod_order, w.new_data = desc->new_data;
  CUDA_TRY(cudaMemcpyAsync(h->d_work, h->h_work[0], sizeof(CbWork), cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaMemcpyAsync(d_soft, soft_inout, N, cudaMemcpyHostToDevice, h->stream));
  if (desc->rm_length) {
    CUDA_TRY(cudaMemcpyAsync(d_rm, in, desc->rm_length, cudaMemcpyHostToDevice, h->stream));
  }
  ldpc_dematch_kernel<<<1, 256, 0, h->stream>>>(desc->base_graph, desc->lifting_size, h->d_work, 1, d_rm, d_soft);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  CUDA_TRY(cudaMemcpyAsync(soft_inout, d_soft, N, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

// ---- HBM-resident soft-buffer arena ---------------------------------------------------------------------------------
int ldpc_cuda_softbuffer_reserve(ldpc_cuda_handle* h, uint64_t nof_bytes)
{
  if (h == nullptr || nof_bytes == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  if (h->arena && h->arena_owned) {
    CUDA_TRY(cudaFree(h->arena));
  }
  h->arena = nullptr, h->arena_bytes = 0, h->arena_owned = false;
  CUDA_TRY(cudaMalloc(&h->arena, nof_bytes));
  h->arena_bytes = nof_bytes;
  h->arena_owned = true;
  // zero-initialised once, like the reference pool (rx_softbuffer_codeblock_pool.h)
  CUDA_TRY(cudaMemsetAsync(h->arena, 0, nof_bytes, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_softbuffer_zero(ldpc_cuda_handle* h, uint64_t offset, uint64_t nof_bytes)
{
  if (h == nullptr || h->arena == nullptr || offset + nof_bytes > h->arena_bytes) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaMemsetAsync(h->arena + offset, 0, nof_bytes, h->stream));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_softbuffer_read(ldpc_cuda_handle* h, uint64_t offset, int8_t* dst, uint64_t nof_bytes)
{
  if (h == nullptr || h->arena == nullptr || dst == nullptr || offset + nof_bytes > h->arena_bytes) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaMemcpyAsync(dst, h->arena + offset, nof_bytes, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_softbuffer_write(ldpc_cuda_handle* h, uint64_t offset, const int8_t* src, uint64_t nof_bytes)
{
  if (h == nullptr || h->arena == nullptr || src == nullptr || offset + nof_bytes > h->arena_bytes) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaMemcpyAsync(h->arena + offset, src, nof_bytes, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

void* ldpc_cuda_softbuffer_ptr(ldpc_cuda_handle* h) { return h ? h->arena : nullptr; }

int ldpc_cuda_arena_create(int device, uint64_t nof_bytes, void** arena_d)
{
  if (arena_d == nullptr || nof_bytes == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  *arena_d = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n || !usable_device(device)) {
    cudaGetLastError();
    return LDPC_CUDA_ERR_CUDA;
  }
  CUDA_TRY(cudaSetDevice(device));
  void* ptr = nullptr;
  CUDA_TRY(cudaMalloc(&ptr, nof_bytes));
  // zero-initialised once, like the reference pool (rx_softbuffer_codeblock_pool.h)
  CUDA_TRY(cudaMemset(ptr, 0, nof_bytes));
  *arena_d = ptr;
  return LDPC_CUDA_OK;
}

int ldpc_cuda_reserve_staging(ldpc_cuda_handle* h, uint64_t in_bytes, uint64_t out_bytes, uint64_t tb_bytes)
{
  if (h == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaDeviceSynchronize());  // the buffers may be in use by earlier asynchronous calls
  const int rc = ensure_capacity(h, in_bytes, out_bytes);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  const size_t need = (size_t)tb_bytes + sizeof(TbResult) * h->max_cbs + 16;
  if (tb_bytes != 0 && need > h->d_tb_out_cap) {
    if (h->d_tb_out) {
      CUDA_TRY(cudaFree(h->d_tb_out));
      h->d_tb_out = nullptr, h->d_tb_out_cap = 0;
    }
    CUDA_TRY(cudaMalloc(&h->d_tb_out, need));
    h->d_tb_out_cap = need;
  }
  return LDPC_CUDA_OK;
}

#ifdef TM_PHASE_TIMING
// Development builds only (tools/phase_timing.py): reads and clears the per-phase cycle counters of the decoder kernel.
extern "C" int ldpc_cuda_debug_phase_cycles(unsigned long long* out16)
{
  unsigned long long zero[16] = {};
  if (cudaMemcpyFromSymbol(out16, g_tm_phase, sizeof(zero)) != cudaSuccess ||
      cudaMemcpyToSymbol(g_tm_phase, zero, sizeof(zero)) != cudaSuccess) {
    return LDPC_CUDA_ERR_CUDA;
  }
  return LDPC_CUDA_OK;
}
#endif

int ldpc_cuda_host_alloc(uint64_t nof_bytes, uint32_t flags, void** ptr)
{
  if (ptr == nullptr || nof_bytes == 0 || (flags & ~LDPC_CUDA_HOST_WRITE_COMBINED) != 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  *ptr = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return LDPC_CUDA_ERR_CUDA;
  }
  void* p = nullptr;
  CUDA_TRY(cudaHostAlloc(&p, nof_bytes,
                         cudaHostAllocPortable | ((flags & LDPC_CUDA_HOST_WRITE_COMBINED) ? cudaHostAllocWriteCombined : 0)));
  *ptr = p;
  return LDPC_CUDA_OK;
}

int ldpc_cuda_host_free(void* ptr)
{
  if (ptr != nullptr) {
    CUDA_TRY(cudaFreeHost(ptr));
  }
  return LDPC_CUDA_OK;
}

int ldpc_cuda_arena_destroy(int device, void* arena_d)
{
  if (arena_d == nullptr) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(device));
  CUDA_TRY(cudaDeviceSynchronize());
  CUDA_TRY(cudaFree(arena_d));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_softbuffer_attach(ldpc_cuda_handle* h, void* arena_d, uint64_t nof_bytes)
{
  if (h == nullptr || arena_d == nullptr || nof_bytes == 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  if (h->arena && h->arena_owned) {
    CUDA_TRY(cudaFree(h->arena));
  }
  h->arena = static_cast<int8_t*>(arena_d);
  h->arena_bytes = nof_bytes;
  h->arena_owned = false;
  return LDPC_CUDA_OK;
}

// ---- transmit side (LDPC encoder + rate matcher), device buffers ------------------------------------------------------
int ldpc_cuda_encode_batch_device(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs, const uint8_t* msg_bits_d,
                                  uint8_t* codeblocks_d)
{
  if (h == nullptr || descs == nullptr || msg_bits_d == nullptr || codeblocks_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  const int rc = stage_tx_work(h, n, descs, false);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  const size_t smem = (size_t)(LDPC_BG1_N_FULL + 4) * MAX_Z;
  const unsigned grid = std::min<unsigned>(n, (unsigned)h->sm_count * 4u);
  ldpc_encode_kernel<<<grid, ENC_THREADS, smem, h->stream>>>(h->d_work, n, msg_bits_d, codeblocks_d);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  return LDPC_CUDA_OK;
}

int ldpc_cuda_rate_match_batch_device(ldpc_cuda_handle* h, uint32_t n, const ldpc_cuda_cb_desc* descs,
                                      const uint8_t* codeblocks_d, uint8_t* out_bits_d)
{
  if (h == nullptr || descs == nullptr || codeblocks_d == nullptr || out_bits_d == nullptr) {
    return LDPC_CUDA_ERR_ARG;
  }
  if (n == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  const int rc = stage_tx_work(h, n, descs, true);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  const unsigned grid = std::min<unsigned>(n, (unsigned)h->sm_count * 16u);
  ldpc_rate_match_kernel<<<grid, 256, 0, h->stream>>>(h->d_work, n, codeblocks_d, out_bits_d);
  CUDA_TRY(cudaGetLastError());
  h->launches += 1;
  return LDPC_CUDA_OK;
}

// ---- transmit side, one codeblock, host buffers (mirrors of ldpc_encoder::encode / ldpc_rate_matcher::rate_match) -------
static int tx_single(ldpc_cuda_handle* h, const ldpc_cuda_cb_desc* desc, const uint8_t* in, uint8_t* out, bool rate_match)
{
  if (h == nullptr || desc == nullptr || in == nullptr || out == nullptr ||
      (desc->base_graph != 1 && desc->base_graph != 2) || lifting_set_of(desc->lifting_size) < 0) {
    return LDPC_CUDA_ERR_ARG;
  }
  const size_t Z = desc->lifting_size, N = bg_N_short(desc->base_graph) * Z, K = bg_K(desc->base_graph) * Z;
  const size_t in_bytes = rate_match ? N : K, out_bytes = rate_match ? desc->rm_length : N;
  if (out_bytes == 0) {
    return LDPC_CUDA_OK;
  }
  CUDA_TRY(cudaSetDevice(h->device));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  int rc = ensure_capacity(h, in_bytes, out_bytes);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  ldpc_cuda_cb_desc d = *desc;
  d.in_offset = 0, d.out_offset = 0;
  CUDA_TRY(cudaMemcpyAsync(h->d_in, in, in_bytes, cudaMemcpyHostToDevice, h->stream));
  rc = rate_match ? ldpc_cuda_rate_match_batch_device(h, 1, &d, reinterpret_cast<const uint8_t*>(h->d_in), h->d_out)
                  : ldpc_cuda_encode_batch_device(h, 1, &d, reinterpret_cast<const uint8_t*>(h->d_in), h->d_out);
  if (rc != LDPC_CUDA_OK) {
    return rc;
  }
  CUDA_TRY(cudaMemcpyAsync(out, h->d_out, out_bytes, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return LDPC_CUDA_OK;
}

int ldpc_cuda_encode(ldpc_cuda_handle* h, const ldpc_cuda_cb_desc* desc, const uint8_t* msg_bits, uint8_t* codeblock)
{
  return tx_single(h, desc, msg_bits, codeblock, false);
}

int ldpc_cuda_rate_match(ldpc_cuda_handle* h, const ldpc_cuda_cb_desc* desc, const uint8_t* codeblock, uint8_t* out_bits)
{
  return tx_single(h, desc, codeblock, out_bits, true);
}

}  // extern "C"
