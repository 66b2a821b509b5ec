// Can TMEM (256 KB/SM) serve as per-thread private scratch for a 12-warp CTA? Each warp owns TMEM lanes
// 32*(warp%4)..+31; warps w, w+4, w+8 share those lanes and split the 512 columns 170/170/170.
// Checks correctness (no aliasing) and measures tcgen05.ld / tcgen05.st x4 throughput. Not part of the product.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tm_ld4(unsigned taddr, unsigned& a, unsigned& b, unsigned& c, unsigned& d)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr));
}
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_st4(unsigned taddr, unsigned a, unsigned b, unsigned c, unsigned d)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__global__ void __launch_bounds__(384, 1) k_tmem(unsigned* errors, long long* cycles, unsigned* sink)
{
  __shared__ unsigned s_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) {
    unsigned dst = (unsigned)__cvta_generic_to_shared(&s_base);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned base = s_base;
  const unsigned my = base + (((unsigned)(warp & 3) * 32u) << 16) + (unsigned)(warp >> 2) * 168u;
  // correctness: 42 x4 stores, then read back
  for (int c = 0; c < 168; c += 4) {
    unsigned v = tid * 1000u + c;
    tm_st4(my + c, v, v + 1, v + 2, v + 3);
  }
  tm_wait_st();
  __syncthreads();
  unsigned bad = 0;
  for (int c = 0; c < 168; c += 4) {
    unsigned a, b, cc, d, v = tid * 1000u + c;
    tm_ld4(my + c, a, b, cc, d);
    tm_wait_ld();
    bad += (a != v) + (b != v + 1) + (cc != v + 2) + (d != v + 3);
  }
  if (bad) atomicAdd(errors, bad);
  __syncthreads();
  // throughput: ld x4 -> modify -> st x4 over the thread's columns, 64 passes
  unsigned acc = 0;
  long long t0 = clock64();
  for (int pass = 0; pass < 64; ++pass) {
    for (int c = 0; c < 168; c += 4) {
      unsigned a, b, cc, d;
      tm_ld4(my + c, a, b, cc, d);
      tm_wait_ld();
      acc += a ^ d;
      tm_st4(my + c, a + 1, b + 1, cc + 1, d + 1);
    }
  }
  tm_wait_st();
  long long t1 = clock64();
  // load-only latency chain
  long long t2 = clock64();
  unsigned addr = my;
  for (int i = 0; i < 256; ++i) {
    unsigned a, b, cc, d;
    tm_ld4(addr, a, b, cc, d);
    tm_wait_ld();
    addr = my + ((a + acc) & 4);  // dependent
  }
  long long t3 = clock64();
  sink[blockIdx.x * 384 + tid] = acc + addr;
  if (tid == 0) { cycles[2 * blockIdx.x] = t1 - t0; cycles[2 * blockIdx.x + 1] = t3 - t2; }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512));
  }
}

int main()
{
  unsigned *err, *sink; long long* cyc;
  cudaMalloc(&err, 4); cudaMemset(err, 0, 4); cudaMalloc(&cyc, 148 * 16); cudaMalloc(&sink, 148 * 384 * 4);
  k_tmem<<<148, 384>>>(err, cyc, sink);
  cudaError_t e = cudaDeviceSynchronize();
  unsigned herr; long long h[296];
  cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost); cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("status=%s errors=%u\n", cudaGetErrorString(e), herr);
  double words = 64.0 * 168 * 384;  // words loaded (and stored) per CTA
  printf("ld+st x4 loop: %lld cycles per CTA -> %.1f B/clk/SM loaded + same stored; %.1f cycles per (ld x4 + wait + st x4) per warp-pass\n",
         h[0], words * 4 / h[0], (double)h[0] / (64 * 42));
  printf("dependent ld x4 + wait latency: %.1f cycles\n", (double)h[1] / 256);
  return 0;
}
