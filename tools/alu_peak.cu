// Register-only throughput microbenchmark of the integer ops the MAP kernels are made of (sm_100a).
// Prints lane-ops per clock per SM for each op, using clock64() deltas on a full-chip launch.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
template <int OP> __device__ __forceinline__ unsigned op(unsigned a, unsigned b, unsigned c)
{
  if (OP == 0) return (unsigned)__viaddmin_s32_relu((int)a, (int)b, 65535);       // VIADDMNMX.RELU (32-bit biased sat add)
  // max(a, b) with a loop-invariant b is idempotent and a + b folds into a + ITERS * b: the compiler removed both loops and the
  // rows printed 2020.6 and 14973.0 "lane-ops/clk/SM". Volatile PTX keeps every instruction (checked in the SASS).
  if (OP == 1) { unsigned r; asm volatile("max.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }   // VIMNMX
  if (OP == 2) return __vadd2(a, b);                                               // VIADD.16x2
  if (OP == 3) return __vmaxs2(a, b);                                              // VIMNMX.S16x2
  if (OP == 4) return __viaddmax_s16x2(a, b, c);                                   // VIADDMNMX.S16x2
  if (OP == 5) return a * 3u + b;                                                  // IMAD (fma pipe)
  if (OP == 6) return (unsigned)__vimax3_s32((int)a, (int)b, (int)c);              // VIMNMX3
  if (OP == 7) { unsigned r; asm volatile("add.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }   // IADD3 (or IMAD.IADD, as ptxas picks)
  if (OP == 8) return __vimax3_s16x2(a, b, c);                                     // VIMNMX3.S16x2
  return __byte_perm(a, b, 0x5410);                                                // PRMT
}
template <int OP> __global__ void k(unsigned *out, unsigned seed, long long *cyc)
{
  unsigned v[8];
  for (int i = 0; i < 8; i++) v[i] = seed + threadIdx.x * 8 + i;
  unsigned b = seed ^ 0x1234u, c = seed * 7u;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = op<OP>(v[i], b, c);
  }
  long long t1 = clock64();
  unsigned s = 0;
  for (int i = 0; i < 8; i++) s ^= v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
// mixed: alternate an ALU-pipe op with an FMA-pipe op to see whether they dual-issue
__global__ void kmix(unsigned *out, unsigned seed, long long *cyc)
{
  unsigned v[8];
  for (int i = 0; i < 8; i++) v[i] = seed + threadIdx.x * 8 + i;
  unsigned b = seed ^ 0x1234u;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++) v[i] = (unsigned)__viaddmin_s32_relu((int)v[i], (int)b, 65535);
#pragma unroll
    for (int i = 4; i < 8; i++) v[i] = v[i] * 3u + b;
  }
  long long t1 = clock64();
  unsigned s = 0;
  for (int i = 0; i < 8; i++) s ^= v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char *name, int sms, unsigned *d, long long *dc)
{
  const int T = 1024;
  k<OP><<<sms, T>>>(d, 1, dc);
  k<OP><<<sms, T>>>(d, 2, dc);
  cudaDeviceSynchronize();
  long long h[1024], mx = 0;
  cudaMemcpy(h, dc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  for (int i = 0; i < sms; i++) if (h[i] > mx) mx = h[i];
  printf("%-32s %7.1f lane-ops/clk/SM\n", name, (double)T * 8 * ITERS / (double)mx);
}
int main()
{
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount;
  unsigned *d; long long *dc;
  cudaMalloc(&d, sizeof(unsigned) * sms * 1024); cudaMalloc(&dc, sizeof(long long) * 1024);
  printf("%s, %d SMs, %d MHz max\n", p.name, sms, p.clockRate / 1000);
  run<0>("VIADDMNMX.RELU", sms, d, dc); run<1>("max.s32 (2 per VIMNMX3)", sms, d, dc); run<2>("VIADD.16x2", sms, d, dc);
  run<3>("vmaxs2 (2 per VIMNMX3.S16x2)", sms, d, dc); run<4>("VIADDMNMX.S16x2", sms, d, dc); run<5>("IMAD", sms, d, dc);
  run<6>("VIMNMX3", sms, d, dc); run<7>("add.u32 (IADD3)", sms, d, dc); run<8>("VIMNMX3.S16x2", sms, d, dc); run<9>("PRMT", sms, d, dc);
  kmix<<<sms, 1024>>>(d, 1, dc); kmix<<<sms, 1024>>>(d, 2, dc); cudaDeviceSynchronize();
  long long h[1024], mx = 0; cudaMemcpy(h, dc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  for (int i = 0; i < sms; i++) if (h[i] > mx) mx = h[i];
  printf("%-32s %7.1f lane-ops/clk/SM\n", "VIADDMNMX+IMAD 1:1", 1024.0 * 8 * ITERS / (double)mx);
  return 0;
}
