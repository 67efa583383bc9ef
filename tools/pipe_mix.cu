// Issue-rate microbenchmark for the instruction mixes the MAP kernel could be made of (sm_100a): cycles per
// warp-instruction per SM sub-partition with 1 and 8 warps per scheduler. Register-only, 8 independent chains.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
#define A_VADD(d, a, b)      asm volatile("add.s16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b))
#define A_ADDMAXS(d, a, b, c) d = __viaddmax_s16x2(a, b, c)
#define A_ADDMAXU(d, a, b, c) d = __viaddmax_u16x2(a, b, c)
#define A_MAXU(d, a, b)      d = __vmaxu2(a, b)
#define A_MAXS(d, a, b)      d = __vmaxs2(a, b)
#define F_IMAD(d, a, one, b) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b))
#define A_IADD(d, a, b)      asm volatile("add.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b))
#define A_LOP(d, a, b, c)    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c))
#define A_PRMT(d, a, b)      asm volatile("prmt.b32 %0, %1, %2, 0x9999;" : "=r"(d) : "r"(a), "r"(b))
#define A_SHF(d, a, b)       asm volatile("shf.l.wrap.b32 %0, %1, %2, 1;" : "=r"(d) : "r"(a), "r"(b))

template <int MODE> __global__ void k(unsigned *out, unsigned seed, unsigned one, long long *cyc)
{
  unsigned v[8], w[8];
  for (int i = 0; i < 8; i++) { v[i] = seed + threadIdx.x * 8 + i; w[i] = v[i] * 3u; }
  unsigned b = seed ^ 0x1234u, c = seed * 7u;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) A_VADD(v[i], v[i], b);
      if (MODE == 1) A_ADDMAXS(v[i], v[i], b, c);
      if (MODE == 2) A_MAXU(v[i], v[i], b);
      if (MODE == 3) F_IMAD(v[i], v[i], one, b);
      if (MODE == 4) A_IADD(v[i], v[i], b);
      if (MODE == 5) A_LOP(v[i], v[i], b, c);
      if (MODE == 6) A_PRMT(v[i], v[i], b);
      if (MODE == 7) A_SHF(v[i], v[i], b);
      if (MODE == 8) { F_IMAD(w[i], v[i], one, b); A_MAXU(v[i], w[i], c); }                       // add on FMA + max on ALU   (2 instr)
      if (MODE == 9) { F_IMAD(w[i], w[i], one, b); A_ADDMAXS(v[i], v[i], b, c); }                 // independent IMAD || fused (2 instr)
      if (MODE == 10) { F_IMAD(w[i], v[i], one, b); A_ADDMAXU(v[i], v[i], c, w[i]); }             // IMAD feeding a fused op   (2 instr)
      if (MODE == 11) { F_IMAD(w[i], v[i], one, b); A_MAXU(w[i], w[i], c); A_ADDMAXU(v[i], v[i], b, w[i]); }   // 1 FMA : 2 ALU (3 instr)
      if (MODE == 12) A_ADDMAXU(v[i], v[i], b, c);
      if (MODE == 13) { A_VADD(w[i], v[i], b); A_MAXS(v[i], w[i], c); }                           // today's unfused pair on ALU only
    }
  }
  long long t1 = clock64();
  unsigned s = 0;
  for (int i = 0; i < 8; i++) s ^= v[i] ^ w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
static const int NINSTR[14] = {1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 3, 1, 2};
static const char *NAME[14] = {"VIADD.16x2", "VIADDMNMX.S16x2", "VIMNMX.U16x2", "IMAD (add on fma pipe)", "IADD3 / add.u32", "LOP3", "PRMT", "SHF",
  "IMAD -> VIMNMX.U16x2", "IMAD || VIADDMNMX.S16x2", "IMAD -> VIADDMNMX.U16x2", "IMAD -> VIMNMX.U -> VIADDMNMX.U", "VIADDMNMX.U16x2", "VIADD.16x2 -> VIMNMX.S16x2"};
template <int MODE> void run(int sms, unsigned *d, long long *dc)
{
  for (int T : {128, 1024}) {
    k<MODE><<<sms, T>>>(d, 1, 1, dc);
    k<MODE><<<sms, T>>>(d, 2, 1, dc);
    cudaDeviceSynchronize();
    long long h[1024], mx = 0;
    cudaMemcpy(h, dc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    for (int i = 0; i < sms; i++) if (h[i] > mx) mx = h[i];
    const double winstr = (double)(T / 32 / 4) * 8 * ITERS * NINSTR[MODE];       // warp-instructions per sub-partition
    printf("%-34s %2d warp/SMSP  %.3f cycles per warp-instruction  (%.2f per group of %d)\n", NAME[MODE], T / 128, mx / winstr, mx / winstr * NINSTR[MODE], NINSTR[MODE]);
  }
}
int main()
{
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int sms = p.multiProcessorCount;
  unsigned *d; long long *dc;
  cudaMalloc(&d, sizeof(unsigned) * sms * 1024); cudaMalloc(&dc, sizeof(long long) * 1024);
  printf("%s, %d SMs\n", p.name, sms);
  run<0>(sms, d, dc); run<1>(sms, d, dc); run<12>(sms, d, dc); run<2>(sms, d, dc); run<3>(sms, d, dc); run<4>(sms, d, dc); run<5>(sms, d, dc);
  run<6>(sms, d, dc); run<7>(sms, d, dc); run<13>(sms, d, dc); run<8>(sms, d, dc); run<9>(sms, d, dc); run<10>(sms, d, dc); run<11>(sms, d, dc);
  return 0;
}
