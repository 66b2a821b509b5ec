// Instruction-cache capacity probe: a straight-line body of N HFMA2/LOP3 instructions (4 independent chains) looped many
// times by 12 warps per SM (the decoder's occupancy). Reports issue rate vs code size. Not part of the product.
#include <cstdio>
#include <cuda_runtime.h>
#define I8 "fma.rn.f16x2 %0, %0, %8, %9;\n\tlop3.b32 %1, %1, %8, %9, 0x96;\n\tfma.rn.f16x2 %2, %2, %8, %9;\n\tlop3.b32 %3, %3, %8, %9, 0x96;\n\tfma.rn.f16x2 %4, %4, %8, %9;\n\tlop3.b32 %5, %5, %8, %9, 0x96;\n\tfma.rn.f16x2 %6, %6, %8, %9;\n\tlop3.b32 %7, %7, %8, %9, 0x96;\n\t"
#define I16 I8 I8
#define I64 I16 I16 I16 I16
#define I256 I64 I64 I64 I64
#define I1024 I256 I256 I256 I256
#define BLK1024 if (SYNC && (i & 3) == 3) __syncthreads(); asm volatile(I64 : "+r"(a), "+r"(b), "+r"(c), "+r"(d), "+r"(e), "+r"(f), "+r"(g), "+r"(h) : "r"(k1), "r"(k2));
template <int NK, int TH, bool SYNC>  // body = NK * 64 instructions, optional barrier every 256
__global__ void __launch_bounds__(TH, 1) k(unsigned* out, int reps, long long* cyc)
{
  unsigned a = threadIdx.x, b = a * 3 + 1, c = a ^ 0x55, d = a + 7, e = a + 9, f = a * 5, g = a ^ 0x77, h = a + 100, k1 = a | 0x3c003c00, k2 = a >> 3;
  long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
#pragma unroll
    for (int i = 0; i < NK; ++i) {
      BLK1024
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * TH + threadIdx.x] = a ^ b ^ c ^ d ^ e ^ f ^ g ^ h;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int NK, int TH, bool SYNC>
void run(unsigned* out, long long* cyc)
{
  int reps = 32768 / NK;
  k<NK, TH, SYNC><<<148, TH>>>(out, reps, cyc);
  cudaDeviceSynchronize();
  k<NK, TH, SYNC><<<148, TH>>>(out, reps, cyc);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double instr = (double)reps * NK * 64 * (TH / 32);  // warp-instructions per SM
  printf("sync %d threads %d body %6d instr = %6.1f KB: %.3f warp-instr/clk/SM (%s)\n", (int)SYNC, TH, NK * 64, NK * 1.0, instr / h[0], cudaGetErrorString(e));
}
int main()
{
  unsigned* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  run<24,384,true>(out, cyc); run<40,384,true>(out, cyc); run<48,384,true>(out, cyc); run<56,384,true>(out, cyc); run<64,384,true>(out, cyc); run<72,384,true>(out, cyc); run<80,384,true>(out, cyc); run<96,384,true>(out, cyc); run<112,384,true>(out, cyc); run<128,384,true>(out, cyc); run<144,384,true>(out, cyc); run<160,384,true>(out, cyc); run<192,384,true>(out, cyc); run<224,384,true>(out, cyc); run<256,384,true>(out, cyc); run<320,384,true>(out, cyc); run<384,384,true>(out, cyc);
  return 0;
}
