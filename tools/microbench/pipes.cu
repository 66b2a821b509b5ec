// Microbenchmark: per-SM issue throughput (thread-ops / clk / SM) of the packed-integer / half2 / logic instructions
// an LDPC min-sum kernel can be built from, and of shared-memory accesses, on sm_100a. Not part of the product.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run: ./pipes
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 512
#define NACC 8

template <int OP>
__device__ __forceinline__ unsigned op(unsigned x, unsigned y, unsigned z)
{
  if (OP == 0) return __vadd2(x, y);                                  // VIADD.16x2 ?
  if (OP == 1) return __vmins2(x, y);                                 // VIMNMX.S16x2
  if (OP == 2) return __vimin3_s16x2(x, y, z);                        // VIMNMX3.S16x2
  if (OP == 3) return __viaddmin_s16x2(x, y, z);                      // VIADDMNMX.S16x2
  if (OP == 4) { __half2 r = __hadd2(*(__half2*)&x, *(__half2*)&y); return *(unsigned*)&r; }   // HADD2
  if (OP == 5) { __half2 r = __hmin2(*(__half2*)&x, *(__half2*)&y); return *(unsigned*)&r; }   // HMNMX2
  if (OP == 6) return __hlt2_mask(*(__half2*)&x, *(__half2*)&y) ^ z;  // HSET2 + LOP3
  if (OP == 7) return (x & y) ^ z;                                    // LOP3
  if (OP == 8) return __byte_perm(x, y, z);                           // PRMT
  if (OP == 9) return __funnelshift_l(x, y, z);                       // SHF
  if (OP == 10) return x * y + z;                                     // IMAD
  if (OP == 11) return x + y + z;                                     // IADD3
  if (OP == 12) return __vabsdiffu4(x, y);                            // VABSDIFF4
  if (OP == 13) { __half2 r = __hfma2(*(__half2*)&x, *(__half2*)&y, *(__half2*)&z); return *(unsigned*)&r; } // HFMA2
  if (OP == 14) return __vminu4(x, y);                                // emulated 8x4 min
  if (OP == 15) return __vsubss4(x, y);                               // emulated 8x4 saturating sub
  if (OP == 16) return max((int)x, (int)y);                           // IMNMX / VIMNMX
  if (OP == 17) return __vmaxs2(x, y) + z;                            // VIMNMX.S16x2 + IADD (two pipes?)
  if (OP == 18) { __half2 r = __hadd2(*(__half2*)&x, *(__half2*)&y); return (*(unsigned*)&r) ^ z; } // HADD2 + LOP3 mix
  if (OP == 19) { __half2 r = __hadd2(*(__half2*)&x, *(__half2*)&y); return __vmins2(*(unsigned*)&r, z); } // HADD2+VIMNMX
  if (OP == 20) { __half2 r = __hmin2(*(__half2*)&x, *(__half2*)&y); return (*(unsigned*)&r) ^ z; } // HMNMX2 + LOP3
  if (OP == 21) { __half2 r = __hmin2(*(__half2*)&x, *(__half2*)&y); __half2 s = __hadd2(r, *(__half2*)&z); return *(unsigned*)&s; } // HMNMX2+HADD2
  if (OP == 22) return __vadd2(x, y) ^ z;                             // VIADD.16x2 + LOP3
  if (OP == 23) return __vadd2(x, y) * z;                             // VIADD.16x2 + IMAD
  return x;
}

template <int OP>
__global__ void __launch_bounds__(512) k_alu(const unsigned* in, unsigned* out, long long* cycles)
{
  unsigned acc[NACC];
  unsigned y = in[threadIdx.x], z = in[threadIdx.x + 512];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = in[threadIdx.x + 32 * i] + i;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 4
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i] = op<OP>(acc[i], acc[(i + 1) % NACC] ^ y, z);
  }
  long long t1 = clock64();
  unsigned s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s ^= acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// Shared memory: MODE 0 LDS.32 conflict-free, 1 LDS.64, 2 LDS.128, 3 LDS.U8 consecutive bytes, 4 STS.32, 5 STS.U8,
// 6 SHFL, 7 LDS.32 + STS.32 pair, 8 LDS.U16
template <int MODE>
__global__ void __launch_bounds__(512) k_smem(const unsigned* in, unsigned* out, long long* cycles)
{
  __shared__ __align__(16) unsigned sm[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = in[i & 1023];
  __syncthreads();
  unsigned acc = 0, idx = threadIdx.x;
  unsigned off = in[0] & 1;  // runtime zero-ish to keep address math alive
  long long t0 = clock64();
#pragma unroll 8
  for (int it = 0; it < ITERS; ++it) {
    unsigned a = (idx + it * 32 * (1 + off)) & 1023;
    if (MODE == 0) acc ^= sm[a];
    if (MODE == 1) { uint2 v = ((uint2*)sm)[a]; acc ^= v.x ^ v.y; }
    if (MODE == 2) { uint4 v = ((uint4*)sm)[a & 1023]; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    if (MODE == 3) acc ^= ((unsigned char*)sm)[a + off];
    if (MODE == 4) sm[a] = acc + it;
    if (MODE == 5) ((unsigned char*)sm)[a + off] = (unsigned char)(acc + it);
    if (MODE == 6) acc ^= __shfl_sync(0xffffffffu, acc + it, (threadIdx.x + 1) & 31);
    if (MODE == 7) { unsigned v = sm[a]; sm[(a + 512) & 1023] = v + it; acc ^= v; }
    if (MODE == 8) acc ^= ((unsigned short*)sm)[a + off];
  }
  long long t1 = clock64();
  __syncthreads();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc ^ sm[threadIdx.x];
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static const char* alu_names[] = {"VIADD.16x2(__vadd2)", "VIMNMX.S16x2", "VIMNMX3.S16x2", "VIADDMNMX.S16x2", "HADD2", "HMNMX2",
  "HSET2+LOP3", "LOP3", "PRMT", "SHF", "IMAD", "IADD3", "VABSDIFF4", "HFMA2", "vminu4(emul)", "vsubss4(emul)", "IMNMX.S32",
  "VIMNMX.S16x2+IADD", "HADD2+LOP3", "HADD2+VIMNMX.S16x2", "HMNMX2+LOP3", "HMNMX2+HADD2", "VIADD.16x2+LOP3", "VIADD.16x2+IMAD"};
static const char* smem_names[] = {"LDS.32", "LDS.64", "LDS.128", "LDS.U8", "STS.32", "STS.U8", "SHFL", "LDS.32+STS.32", "LDS.U16"};

template <int OP> void run_alu(unsigned* in, unsigned* out, long long* cyc, int sms, int ctas_per_sm)
{
  int grid = sms * ctas_per_sm;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_alu<OP><<<grid, 512>>>(in, out, cyc);
  cudaEventRecord(e0);
  k_alu<OP><<<grid, 512>>>(in, out, cyc);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[2048]; cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; ++i) avg += h[i]; avg /= grid;
  double ops_per_cta = (double)ITERS * NACC * 512;
  // per-SM throughput while ctas_per_sm CTAs are co-resident
  printf("%-24s ctas/SM=%d  %8.1f thread-ops/clk/SM (kernel-inner clock)   %7.3f ms\n", alu_names[OP], ctas_per_sm,
         ops_per_cta * ctas_per_sm / avg, ms);
}
template <int MODE> void run_smem(unsigned* in, unsigned* out, long long* cyc, int sms, int ctas_per_sm)
{
  int grid = sms * ctas_per_sm;
  k_smem<MODE><<<grid, 512>>>(in, out, cyc);
  k_smem<MODE><<<grid, 512>>>(in, out, cyc);
  cudaDeviceSynchronize();
  long long h[2048]; cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; ++i) avg += h[i]; avg /= grid;
  printf("%-24s ctas/SM=%d  %8.1f thread-accesses/clk/SM\n", smem_names[MODE], ctas_per_sm, (double)ITERS * 512 * ctas_per_sm / avg);
}

template <int OP> struct AluLoop { static void go(unsigned* in, unsigned* out, long long* cyc, int sms) {
  run_alu<OP>(in, out, cyc, sms, 1); run_alu<OP>(in, out, cyc, sms, 2); AluLoop<OP + 1>::go(in, out, cyc, sms); } };
template <> struct AluLoop<24> { static void go(unsigned*, unsigned*, long long*, int) {} };
template <int M> struct SmemLoop { static void go(unsigned* in, unsigned* out, long long* cyc, int sms) {
  run_smem<M>(in, out, cyc, sms, 2); SmemLoop<M + 1>::go(in, out, cyc, sms); } };
template <> struct SmemLoop<9> { static void go(unsigned*, unsigned*, long long*, int) {} };

int main()
{
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount;
  printf("device %s, %d SMs, clock %d kHz\n", p.name, sms, p.clockRate);
  unsigned *in, *out; long long* cyc;
  cudaMalloc(&in, 8192 * 4); cudaMalloc(&out, sms * 4 * 512 * 4); cudaMalloc(&cyc, 2048 * 8);
  unsigned h[8192]; for (int i = 0; i < 8192; ++i) h[i] = 0x3c003c00u ^ (i * 2654435761u & 0x03ff03ffu);
  h[0] &= ~1u;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  AluLoop<0>::go(in, out, cyc, sms);
  SmemLoop<0>::go(in, out, cyc, sms);
  printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
