/*
 * fec_tdec_packed.cu — the fast path of the windowed max-log-MAP turbo decoder (classes P=16 and P=8).
 *
 * Replaces tdec_win{avx16,sse16}_dec + run_tdec_iteration_16bit + the CRC early-stop loop of decode_tb_cb
 * (turbodecoder_win.h:310-586, turbodecoder_iter.h:73-147, sch.c:361-391) for every half-iteration whose
 * inputs satisfy the range condition below; code blocks that do not are handed, untouched, to the exact
 * 32-bit kernel in fec_kernels.cu (same results by construction, it saturates like the reference).
 *
 * Arithmetic: two sub-blocks (lanes 2j, 2j+1) of one code block per thread, all 8 trellis states in 8
 * registers of packed int16x2; adds are VIADD.16x2 / VIADDMNMX.S16x2 (max(a+b,c), one instruction), which
 * WRAP where the reference SATURATES. They are identical as long as no intermediate leaves int16.
 * Range condition (proof in DESIGN.md section 3.3): with G = max|x| + max|y| over the half-iteration
 * (x = systematic + apriori, y = parity, tails included), every intermediate of the reference is bounded
 * by 10000 + 10 G, so G <= FEC_PK_GMAX = 2000 excludes saturation. G of the NEXT half-iteration is known
 * when the current one ends (max|new extrinsic| is tracked), so the decision is taken before any work is lost.
 *
 * Memory: the extrinsic/apriori array E of every resident code block lives in shared memory (pair rows of
 * Lw words), the interleaver address table of decoder 2 once per CTA; beta checkpoints every 8 steps live in
 * TENSOR MEMORY (tcgen05.st/ld, 32 lanes x 8 columns per checkpoint, lane-private), the 7 recomputed betas
 * of a chunk in registers. Systematic/parity streams are read from HBM/L2 as 16-byte words (4 steps).
 * One warp = 32/(P/2) code blocks in lockstep; warps are independent (no CTA barrier in the loop).
 */
#include "fec_kernels.cuh"

#define PK_CH      8                 /* checkpoint spacing = chunk length (trellis steps) */
#define PK_NEG     0xD8F0D8F0u       /* (-10000, -10000) */
#define PK_WARPS   FEC_PK_WARPS

#ifndef FEC_PK_TMEM
#define FEC_PK_TMEM 1
#endif
/* tuning switches (A/B-measured on B200, see profiles/): */
#ifndef PK_KEEP_ADDR
#define PK_KEEP_ADDR 1   /* keep the 16 E addresses of a chunk in registers instead of recomputing them at write-back */
#endif

/* ---------------------------------------------------------------- packed helpers ---- */
__device__ __forceinline__ uint32_t pk_add(uint32_t a, uint32_t b) { return __vadd2(a, b); }
__device__ __forceinline__ uint32_t pk_sub(uint32_t a, uint32_t b) { return __vsub2(a, b); }
__device__ __forceinline__ uint32_t pk_max(uint32_t a, uint32_t b) { return __vmaxs2(a, b); }
__device__ __forceinline__ uint32_t pk_min(uint32_t a, uint32_t b) { return __vmins2(a, b); }
__device__ __forceinline__ uint32_t pk_addmax(uint32_t a, uint32_t b, uint32_t c) { return __viaddmax_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t pk_prmt(uint32_t a, uint32_t b, uint32_t sel)
{
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

__device__ __forceinline__ uint32_t pk_lds_u16(uint32_t addr)
{
  uint32_t r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}
__device__ __forceinline__ void pk_sts_u16(uint32_t addr, uint32_t val)
{
  asm volatile("st.shared.u16 [%0], %1;" :: "r"(addr), "r"(val) : "memory");
}

/* scheduling fence: ptxas does not move memory instructions across a warp barrier, which keeps the prefetch
 * loads where they are issued in the source (it otherwise sinks them to the register rotation at the loop end) */
#ifndef PK_OPT_DESC
#define PK_OPT_DESC 1
#endif
#ifndef PK_EPI_NOINLINE
#define PK_EPI_NOINLINE 1
#endif
#if PK_EPI_NOINLINE
#define PK_EPI_INLINE __noinline__
#else
#define PK_EPI_INLINE __forceinline__
#endif
#ifndef PK_LATE_ZERO
#define PK_LATE_ZERO 1
#endif
#ifndef PK_TAIL_EARLY
#define PK_TAIL_EARLY 1
#endif
#ifndef PK_NORM8
#define PK_NORM8 0
#endif
#ifndef PK_DIAG_NOLOAD
#define PK_DIAG_NOLOAD 0
#endif
#ifndef PK_TIMELINE
#define PK_TIMELINE 0
#endif
#ifndef PK_PF_BETA
#define PK_PF_BETA 0
#endif
#ifndef PK_PF_L1
#define PK_PF_L1 1
#endif
#ifndef PK_CRC64
#define PK_CRC64 1       /* CRC syndrome words by 8-byte instead of 16-byte loads */
#endif
#ifndef PK_PINMODE
#define PK_PINMODE 1
#endif
#if PK_PINMODE == 1
#define PK_PIN() asm volatile("bar.warp.sync 0xffffffff;" ::: "memory")
#else
#define PK_PIN() do { } while (0)
#endif

/* global loads as volatile asm: ptxas otherwise sinks the prefetch loads down to their first use (register
 * pressure heuristic), which collapses the software-pipelining distance to zero */
__device__ __forceinline__ uint4 pk_ldg128(const void *p)
{
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 pk_ldg64(const void *p)
{
  uint2 r;
  asm volatile("ld.global.nc.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t pk_ldg32(const void *p)
{
  uint32_t r;
  asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

/* turbodecoder_win.h:244-261: subtract state 0 from states 1..7, state 0 := 0 (every even k != 0) */
__device__ __forceinline__ void pk_norm(uint32_t *o)
{
#pragma unroll
  for (int i = 1; i < 8; i++) o[i] = pk_sub(o[i], o[0]);
  o[0] = 0;
}

/* ---- run-time range certificate (TRK variants) ------------------------------------------------
 * Beyond G = FEC_PK_GMAX the a-priori proof no longer excludes saturation, but real traffic stays far from it
 * (the reference's own demodulator at 64QAM: peak |intermediate| ~12000). The TRK variant of a half-iteration
 * therefore (1) normalises exactly where the reference does (after every even step k != 0, no normalisation at
 * the lane shifts), so that every state metric, branch sum and LLR sum of the main passes equals the reference's
 * value as long as nothing overflowed, and (2) records the extremes of the state metrics after every
 * normalisation and at the start of the two main passes, plus the extremes of the two LLR maxima. The
 * certificate evaluated from them at the end of the half-iteration (pk_certificate) proves by induction over
 * the normalisation windows that no add of the reference saturated and none of ours wrapped. */
struct PkTrk { uint32_t hi, lo; };
/* extremes of a NORMALISED state (o[0] == 0 is covered by the accumulators' initial value 0) */
__device__ __forceinline__ void pk_trk7(PkTrk &t, const uint32_t *o)
{
  t.hi = __vimax3_s16x2(t.hi, o[1], o[2]); t.hi = __vimax3_s16x2(t.hi, o[3], o[4]);
  t.hi = __vimax3_s16x2(t.hi, o[5], o[6]); t.hi = __vmaxs2(t.hi, o[7]);
  t.lo = __vimin3_s16x2(t.lo, o[1], o[2]); t.lo = __vimin3_s16x2(t.lo, o[3], o[4]);
  t.lo = __vimin3_s16x2(t.lo, o[5], o[6]); t.lo = __vmins2(t.lo, o[7]);
}
__device__ __forceinline__ void pk_trk8(PkTrk &t, const uint32_t *o)
{
  pk_trk7(t, o);
  t.hi = __vmaxs2(t.hi, o[0]); t.lo = __vmins2(t.lo, o[0]);
}
/* normalise + record, at the reference's normalisation points */
__device__ __forceinline__ void pk_norm_trk(uint32_t *o, PkTrk &t) { pk_norm(o); pk_trk7(t, o); }

/* backward step (turbodecoder_win.h:397-418), 4 adds + 8 add-max */
__device__ __forceinline__ void pk_beta(uint32_t *o, uint32_t x, uint32_t y)
{
  const uint32_t xy = pk_add(x, y);
  uint32_t n[8];
  n[0] = pk_addmax(o[4], xy, o[0]);
  n[1] = pk_addmax(o[0], xy, o[4]);
  n[2] = pk_addmax(o[1], x, pk_add(o[5], y));
  n[3] = pk_addmax(o[1], y, pk_add(o[5], x));
  n[4] = pk_addmax(o[2], y, pk_add(o[6], x));
  n[5] = pk_addmax(o[2], x, pk_add(o[6], y));
  n[6] = pk_addmax(o[3], xy, o[7]);
  n[7] = pk_addmax(o[7], xy, o[3]);
#pragma unroll
  for (int i = 0; i < 8; i++) o[i] = n[i];
}

/* forward branches (turbodecoder_win.h:521-539): m = bit-0 hypothesis, n = bit-1 hypothesis */
__device__ __forceinline__ void pk_alpha_branches(const uint32_t *o, uint32_t x, uint32_t y, uint32_t *m, uint32_t *n)
{
  const uint32_t xy = pk_add(x, y);
  m[0] = o[0];             m[1] = pk_add(o[3], y); m[2] = pk_add(o[4], y); m[3] = o[7];
  m[4] = o[1];             m[5] = pk_add(o[2], y); m[6] = pk_add(o[5], y); m[7] = o[6];
  n[0] = pk_add(o[1], xy); n[1] = pk_add(o[2], x); n[2] = pk_add(o[5], x); n[3] = pk_add(o[6], xy);
  n[4] = pk_add(o[0], xy); n[5] = pk_add(o[3], x); n[6] = pk_add(o[4], x); n[7] = pk_add(o[7], xy);
}

/* ---- mixed-form arithmetic of the a-priori-range kernel (PK_MIX) --------------------------------------------------
 * The packed add / add-max instructions exist on the ALU pipe only and take two issue cycles per warp (64 lanes per clock
 * and SM, profiles/r1_pipe_mix.txt); a plain 32-bit add issues at full rate or goes to the otherwise idle FMA pipe as IMAD.
 * Two int16 lanes (v_hi, v_lo) therefore travel in one of three forms:
 *   P form  two's complement halves, as in memory                       (branch metrics for the fused add-max, LLRs, extrinsics)
 *   I form  the integer v_hi * 65536 + v_lo                              (branch metrics for plain 32-bit adds)
 *   S form  I + 0x80008000: both halves offset-binary                    (state metrics alpha / beta)
 * While every lane stays inside int16 — the range condition this kernel runs under anyway — S + I is a 32-bit add giving S,
 * S (+) P a packed add giving S (VIADDMNMX.U16x2 fuses it with the unsigned per-lane max), S (+) S a packed add giving P
 * (the two offsets cancel mod 2^16: the LLR sums alpha-branch + beta, reduced by the signed fused add-max as before), and
 * S - S a 32-bit subtract giving I (normalisation). Same integers in every lane as the P-form code, so still bit-exact with
 * the reference; per trellis step about 20 packed adds leave the ALU pipe. The certificate (TRK) variant keeps the P form:
 * its extremes are tracked with signed packed min / max. */
#ifndef PK_MIX
#define PK_MIX 0         /* measured on B200 (tools/experiments/README.md): bit-exact, ALU-pipe cycles per step 150 -> 118, but 6 % SLOWER —
                            with one warp per scheduler the loops are issue-bound and the form conversions add instructions */
#endif
#define PK_C 0x80008000u
__device__ __forceinline__ uint32_t pk_to_i(uint32_t p) { return p - 2u * (p & 0x8000u); }
__device__ __forceinline__ uint32_t pk_addmax_u(uint32_t a, uint32_t b, uint32_t c) { return __viaddmax_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t pk_max_u(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }

struct PkGam { uint32_t xP, yP, xyP, xI, yI; };
__device__ __forceinline__ PkGam pk_gammas(uint32_t x, uint32_t y)
{
  PkGam g;
  g.xP = x; g.yP = y; g.xyP = pk_add(x, y); g.xI = pk_to_i(x); g.yI = pk_to_i(y);
  return g;
}
/* S-form state: subtract state 0 from all, state 0 := 0 */
__device__ __forceinline__ void pk_norm_s(uint32_t *o)
{
  const uint32_t c = PK_C - o[0];
#pragma unroll
  for (int i = 1; i < 8; i++) o[i] += c;
  o[0] = PK_C;
}
/* backward step on S-form states: 4 plain adds + 8 fused unsigned add-max */
__device__ __forceinline__ void pk_beta_s(uint32_t *o, const PkGam &g)
{
  uint32_t n[8];
  n[0] = pk_addmax_u(o[4], g.xyP, o[0]);
  n[1] = pk_addmax_u(o[0], g.xyP, o[4]);
  n[2] = pk_addmax_u(o[1], g.xP, o[5] + g.yI);
  n[3] = pk_addmax_u(o[1], g.yP, o[5] + g.xI);
  n[4] = pk_addmax_u(o[2], g.yP, o[6] + g.xI);
  n[5] = pk_addmax_u(o[2], g.xP, o[6] + g.yI);
  n[6] = pk_addmax_u(o[3], g.xyP, o[7]);
  n[7] = pk_addmax_u(o[7], g.xyP, o[3]);
#pragma unroll
  for (int i = 0; i < 8; i++) o[i] = n[i];
}
/* forward branches on S-form states with I-form branch metrics: 12 plain adds */
__device__ __forceinline__ void pk_alpha_branches_s(const uint32_t *o, uint32_t xI, uint32_t yI, uint32_t *m, uint32_t *n)
{
  const uint32_t xyI = xI + yI;
  m[0] = o[0];       m[1] = o[3] + yI; m[2] = o[4] + yI; m[3] = o[7];
  m[4] = o[1];       m[5] = o[2] + yI; m[6] = o[5] + yI; m[7] = o[6];
  n[0] = o[1] + xyI; n[1] = o[2] + xI; n[2] = o[5] + xI; n[3] = o[6] + xyI;
  n[4] = o[0] + xyI; n[5] = o[3] + xI; n[6] = o[4] + xI; n[7] = o[7] + xyI;
}

/* ---------------------------------------------------------------- tensor-memory checkpoint store ---- */
#if FEC_PK_TMEM
__device__ __forceinline__ void pk_ck_store(uint32_t taddr, const uint32_t *o)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               :: "r"(taddr), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]) : "memory");
}
__device__ __forceinline__ void pk_ck_store_fence() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void pk_ck_load_issue(uint32_t taddr, uint32_t *o)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3]), "=r"(o[4]), "=r"(o[5]), "=r"(o[6]), "=r"(o[7]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void pk_ck_load_wait(uint32_t *o)
{
  /* the registers are tied to the wait so that no use can be scheduled above it */
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(o[0]), "+r"(o[1]), "+r"(o[2]), "+r"(o[3]), "+r"(o[4]), "+r"(o[5]), "+r"(o[6]), "+r"(o[7]) :: "memory");
}
#endif

/* ---------------------------------------------------------------- per-thread view of one half-iteration ---- */
struct PkView {
  const uint32_t *Srow;    /* systematic stream of the code block + 4*j words (HBM), read only when par == 0; step k of this
                              thread's pair sits at word (k>>2)*4*TPC + (k&3) from here (fec_tables.h) */
  const uint32_t *Yrow;    /* parity stream of this decoder, same addressing */
  const uint32_t *Ynext;   /* parity stream of the other decoder (prefetch only) */
  const uint2    *Rrow;    /* CRC syndrome words of this decoder and polynomial + 2*j; step k at (k>>1)*2*TPC + (k&1) */
  int             tpc4;    /* 4*TPC */
  const uint16_t *Tlo, *Thi; /* smem: decoder-2 byte offsets into E of lanes 2j / 2j+1 at step k */
  uint32_t        Eb;      /* shared-memory byte address of this code block's extrinsic array E */
  uint32_t        ident;   /* decoder-1 byte offset of lane 2j at step 0: 4*j*Lw (lane 2j+1: +2; step k: +4k) */
  int             par;     /* 0: decoder 1 (natural order), 1: decoder 2 (interleaved order) */
};

/* raw global loads of one 8-step chunk [k0, k0+8). FULL: all 8 steps exist; else only steps i >= i0 (bottom chunk) */
struct PkRaw { uint32_t s[8], y[8]; };

template <bool FULL, bool VEC>
__device__ __forceinline__ void pk_load_raw(const PkView &v, int k0, int i0, PkRaw &r)
{
#if PK_DIAG_NOLOAD      /* timing diagnosis only: what would the kernel cost if the stream loads were free (results are garbage) */
  if (FULL && VEC) {
#pragma unroll
    for (int i = 0; i < 8; i++) { r.y[i] = (uint32_t)((k0 + i) & 3) * 0x00010001u; r.s[i] = (uint32_t)((k0 + 2 * i) & 3) * 0x00010001u; }
    return;
  }
#endif
  if (FULL && VEC) {
    const int w0 = (k0 >> 2) * v.tpc4;
    const uint4 y0 = pk_ldg128(v.Yrow + w0), y1 = pk_ldg128(v.Yrow + w0 + v.tpc4);
    r.y[0] = y0.x; r.y[1] = y0.y; r.y[2] = y0.z; r.y[3] = y0.w; r.y[4] = y1.x; r.y[5] = y1.y; r.y[6] = y1.z; r.y[7] = y1.w;
    uint4 s0 = make_uint4(0, 0, 0, 0), s1 = s0;
    if (!v.par) { s0 = pk_ldg128(v.Srow + w0); s1 = pk_ldg128(v.Srow + w0 + v.tpc4); }
    r.s[0] = s0.x; r.s[1] = s0.y; r.s[2] = s0.z; r.s[3] = s0.w; r.s[4] = s1.x; r.s[5] = s1.y; r.s[6] = s1.z; r.s[7] = s1.w;
  } else {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      r.s[i] = 0; r.y[i] = 0;
      if (FULL || i >= i0) {
        const int k = k0 + i, w = (k >> 2) * v.tpc4 + (k & 3);
        r.y[i] = pk_ldg32(v.Yrow + w); if (!v.par) r.s[i] = pk_ldg32(v.Srow + w);
      }
    }
  }
}

/* the lines pk_load_raw<true, true> will read for chunk k0, asked into L1 ahead of time (no register is tied up) */
__device__ __forceinline__ void pk_prefetch_raw(const PkView &v, int k0)
{
  const int w0 = (k0 >> 2) * v.tpc4;
  asm volatile("prefetch.global.L1 [%0];" :: "l"(v.Yrow + w0));
  asm volatile("prefetch.global.L1 [%0];" :: "l"(v.Yrow + w0 + v.tpc4));
  if (!v.par) {
    asm volatile("prefetch.global.L1 [%0];" :: "l"(v.Srow + w0));
    asm volatile("prefetch.global.L1 [%0];" :: "l"(v.Srow + w0 + v.tpc4));
  }
}

/* CRC syndrome words of the chunk. Step k sits at uint2 index (k>>1)*2*TPC + (k&1): with k0 even (VEC: L % 4 == 0, so every
 * chunk base is) the 8 steps are 4 pairs 128 bytes apart — one base pointer per chunk and constant offsets, instead of
 * one 64-bit address computation per load (5 instructions each in the SASS of the first version) */
template <bool FULL, bool VEC, int TPC>
__device__ __forceinline__ void pk_load_crc(const PkView &v, int k0, int i0, uint2 *rr)
{
  if (FULL && VEC) {
    const uint2 *base = v.Rrow + (k0 >> 1) * (2 * TPC);
#pragma unroll
    for (int i = 0; i < 8; i++) rr[i] = pk_ldg64(base + (i >> 1) * (2 * TPC) + (i & 1));
  } else {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      rr[i] = make_uint2(0, 0);
      if (FULL || i >= i0) { const int k = k0 + i; rr[i] = pk_ldg64(v.Rrow + (k >> 1) * (2 * TPC) + (k & 1)); }
    }
  }
}

/* E byte addresses of the two lanes at step k (decoder 2: table in shared memory; decoder 1: own pair row) */
__device__ __forceinline__ void pk_eaddr(const PkView &v, int k, uint32_t &alo, uint32_t &ahi)
{
  uint32_t ol, oh;
  if (v.par) { ol = v.Tlo[k]; oh = v.Thi[k]; } else { ol = v.ident + 4u * (uint32_t)k; oh = ol + 2u; }
  alo = v.Eb + ol; ahi = v.Eb + oh;
}

/* apriori gather + x = systematic + apriori for the chunk (addresses are recomputed at write-back time:
 * two more LSU/FMA-pipe instructions per step instead of 16 live registers) */
template <bool FULL>
__device__ __forceinline__ void pk_finish_chunk(const PkView &v, int k0, int i0, const PkRaw &r, uint32_t *xs, uint32_t *es, uint32_t *ka = nullptr)
{
#pragma unroll
  for (int i = 0; i < 8; i++) {
    xs[i] = 0; es[i] = 0;
    if (FULL || i >= i0) {
      uint32_t alo, ahi;
      pk_eaddr(v, k0 + i, alo, ahi);
      if (PK_KEEP_ADDR && ka) { ka[2 * i] = alo; ka[2 * i + 1] = ahi; }
      const uint32_t e = pk_prmt(pk_lds_u16(alo), pk_lds_u16(ahi), 0x5410u);
      es[i] = e;
      xs[i] = pk_add(r.s[i], e);
    }
  }
}

/* Normalisation schedule of THIS kernel: every 4th step (after steps i = 4 and i = 0 going backward, i = 3 and
 * i = 7 going forward) instead of the reference's every 2nd. Outputs (LLR = difference of two maxima over the
 * same alpha/beta offsets) do not depend on when the common offset is removed as long as nothing overflows;
 * with this schedule every sum stays within max(13 G, 10000 + 10 G). */

/* 8 backward steps of the main pass (i = 7..0); the checkpoint is the NORMALISED state after step i = 0 */
template <bool FULL, bool MIX>
__device__ __forceinline__ void pk_beta_chunk(uint32_t *o, const uint32_t *xs, const uint32_t *ys, int i0, uint32_t ck_addr)
{
#pragma unroll
  for (int i = 7; i >= 0; i--) {
    if (FULL || i >= i0) {
      if (MIX) { pk_beta_s(o, pk_gammas(xs[i], ys[i])); if ((i & 3) == 0) pk_norm_s(o); }
      else { pk_beta(o, xs[i], ys[i]); if (PK_NORM8 ? i == 0 : (i & 3) == 0) pk_norm(o); }
      if (FULL && i == 0) pk_ck_store(ck_addr, o);
    }
  }
}
/* TRK variant: the reference's schedule. Step i of the chunk is trellis step k = k0 + i, kpar = k0 & 1, knz = (k0 != 0);
 * the checkpoint is the state BEFORE normalisation, as the reference stores beta (turbodecoder_win.h:419-432) */
__device__ __forceinline__ void pk_beta_chunk_trk(uint32_t *o, const uint32_t *xs, const uint32_t *ys, uint32_t ck_addr,
                                                  int kpar, bool knz, PkTrk &t)
{
#pragma unroll
  for (int i = 7; i >= 0; i--) {
    pk_beta(o, xs[i], ys[i]);
    if (i == 0) pk_ck_store(ck_addr, o);
    if ((((i + kpar) & 1) == 0) && (i != 0 || knz)) pk_norm_trk(o, t);
  }
}

struct PkAcc { uint32_t crc, emx, emn, wlo, whi; uint32_t vmx, vmn, rb; PkTrk ta; };   /* vmx.. only used by the TRK variant */

/* one chunk of the forward pass: recompute the inner betas from the checkpoint `ck` (= normalised beta at k0+8),
 * then alpha recursion + LLR + extrinsic write-back + CRC syndrome + decision bits for steps i = i0..7 */
template <bool FULL, bool SHIFT, bool TRK, bool MIX>
__device__ __forceinline__ void pk_alpha_chunk(const PkView &v, int k0, uint32_t *o, const uint32_t *ck, int i0,
                                               const uint32_t *xs, const uint32_t *ys, const uint32_t *es,
                                               const uint2 *rr, PkAcc &acc, const uint32_t *ka, int kpar = 0, bool top = false)
{
  uint32_t b[8], bb[8][8], m[8], n[8], xi[8], yi[8];
#pragma unroll
  for (int s = 0; s < 8; s++) b[s] = ck[s];
  if (TRK) {
    /* ck = beta at k0+8 before normalisation; the reference normalises it if k0+8 is even, except at the top of the
     * sub-block (k0+8 == L), where the recursion starts from the shifted warm-up state as it is */
    if (!top && !(kpar & 1)) pk_norm(b);
  }
#pragma unroll
  for (int i = 7; i >= 1; i--) {
    if (FULL || i >= i0 + 1) {
      if (MIX) {
        const PkGam g = pk_gammas(xs[i], ys[i]);
        xi[i] = g.xI; yi[i] = g.yI;                     /* the forward pass below needs the I forms only */
        pk_beta_s(b, g);
        if (i == 4) pk_norm_s(b);
#pragma unroll
        for (int s = 0; s < 8; s++) bb[i][s] = b[s];
        continue;
      }
      pk_beta(b, xs[i], ys[i]);
      if (TRK) {
#pragma unroll
        for (int s = 0; s < 8; s++) bb[i][s] = b[s];                      /* beta[k0+i] as the reference stores it */
        if (((i + kpar) & 1) == 0) pk_norm(b);
      } else {
        if (!PK_NORM8 && i == 4) pk_norm(b);
#pragma unroll
        for (int s = 0; s < 8; s++) bb[i][s] = b[s];
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (FULL || i >= i0) {
      if (MIX) {
        if (i == 0 || (!FULL && i == i0)) { xi[i] = pk_to_i(xs[i]); yi[i] = pk_to_i(ys[i]); }   /* the step the backward loop did not visit */
        pk_alpha_branches_s(o, xi[i], yi[i], m, n);
      } else pk_alpha_branches(o, xs[i], ys[i], m, n);
      const uint32_t *bn = (i == 7) ? ck : bb[(i + 1) & 7];
      uint32_t m0 = pk_add(bn[0], m[0]), m1 = pk_add(bn[0], n[0]);
#pragma unroll
      for (int s = 1; s < 8; s++) { m0 = pk_addmax(bn[s], m[s], m0); m1 = pk_addmax(bn[s], n[s], m1); }
      if (TRK) { acc.vmx = __vimax3_s16x2(acc.vmx, m0, m1); acc.vmn = __vimin3_s16x2(acc.vmn, m0, m1); }
      /* LLR r = m1 - m0, next apriori en = r - (what this decoder was given). A packed subtract is NOT + add + add 1; the sign
       * test wants r - 1 = m1 + ~m0 anyway, so r itself is never formed (unless it has to be halved): en = (r - 1) + ~es + 2 */
      uint32_t rm1, en;
      if (SHIFT) {
        uint32_t r = pk_sub(m1, m0);
        r = ((r >> 1) & 0x7FFF7FFFu) | (r & 0x80008000u);                  /* divide_output, sse16 only (:61,565) */
        en = pk_sub(r, es[i]);
        rm1 = pk_add(r, 0xFFFFFFFFu);
      } else {
        rm1 = pk_add(m1, ~m0);
        en = pk_add(pk_add(rm1, ~es[i]), 0x00020002u);
      }
      uint32_t alo, ahi;
      if (PK_KEEP_ADDR) { alo = ka[2 * i]; ahi = ka[2 * i + 1]; } else pk_eaddr(v, k0 + i, alo, ahi);
      pk_sts_u16(alo, en);
      pk_sts_u16(ahi, en >> 16);
      acc.emx = pk_max(acc.emx, en); acc.emn = pk_min(acc.emn, en);
      /* decision: bit 1 <=> LLR > 0 <=> r - 1 >= 0. The masks below are all-ones for a ZERO bit (sign of r - 1: one packed add
       * with an immediate instead of the two-instruction negation); the CRC takes them inverted inside its LOP3, the decision
       * words collect w <- 2 w + mask (an IMAD on the otherwise idle FMA pipe instead of a funnel shift on the ALU pipe):
       * after the 32 steps of a word w = -(inverted bits), so the decisions are ~(-w) = w - 1 (taken at the stores) */
      const uint32_t zlo = pk_prmt(rm1, 0u, 0x9999u), zhi = pk_prmt(rm1, 0u, 0xBBBBu);
      acc.crc ^= (~zlo & rr[i].x) ^ (~zhi & rr[i].y);
      acc.wlo = acc.wlo * 2u + zlo; acc.whi = acc.whi * 2u + zhi;
#pragma unroll
      for (int s = 0; s < 8; s++) o[s] = MIX ? pk_max_u(m[s], n[s]) : pk_max(m[s], n[s]);
      if (TRK) { if ((((i + kpar) & 1) == 0) && (k0 + i != 0)) pk_norm_trk(o, acc.ta); }
      else if (PK_NORM8 && !MIX ? i == 7 : (i & 3) == 3) { if (MIX) pk_norm_s(o); else pk_norm(o); }
    } else { acc.wlo = acc.wlo * 2u + 0xFFFFFFFFu; acc.whi = acc.whi * 2u + 0xFFFFFFFFu; }   /* a zero bit for a step that does not exist */
  }
}

/* first 40 steps of a stream (pair-interleaved layout: 40 * TPC words, contiguous) into L1, one or two lines per thread */
template <int TPC>
__device__ __forceinline__ void pk_prefetch_warm(const uint32_t *y, const uint32_t *s, int j)
{
  constexpr int NL = (40 * TPC * 4 + 127) / 128;
#pragma unroll
  for (int l = 0; l < (NL + TPC - 1) / TPC; l++) {
    const int line = j + l * TPC;
    if (line < NL) {
      asm volatile("prefetch.global.L1 [%0];" :: "l"(y + line * 32));
      if (s) asm volatile("prefetch.global.L1 [%0];" :: "l"(s + line * 32));
    }
  }
}

/* One half-iteration of one constituent decoder for the two lanes of this thread.
 * TPC = threads per code block (P/2); VEC: L % 4 == 0 (16-byte loads). Decisions go to bits_lo/bits_hi as
 * MSB-first words; bit k of a lane sits at bit index k + (8*nck - L) of its row. */
template <int TPC, bool VEC, bool SHIFT, bool TRK>
__device__ __forceinline__ void pk_half_iteration(const PkView &v, int L, uint32_t tbase, unsigned gmask, int j,
                                                   const int16_t *tail, uint32_t *bits_lo, uint32_t *bits_hi, PkAcc &acc,
                                                   uint32_t *Ezero, int nEz)
{
  const int nfull = L >> 3, rem = L & 7;
  const int kpar = VEC ? 0 : (L & 1);               /* parity of every chunk base k0 (chunks are 8 steps, aligned to L) */
  constexpr bool MIX = PK_MIX && !TRK;              /* S / I / P forms (see pk_beta_s) instead of P form throughout */
  uint32_t o[8], xs[8], es[8], ka[16];
  PkRaw raw, nraw;
  PkTrk tb;                                         /* TRK: extremes of the beta recursion (warm-up + main pass) */
  tb.hi = 0; tb.lo = 0;
  if (TRK) { acc.ta.hi = 0; acc.ta.lo = 0; acc.vmx = 0x80008000u; acc.vmn = 0x7FFF7FFFu; }

#if PK_TAIL_EARLY
  /* the 3 tail steps' (x, y) pairs, 12 bytes behind the streams (4-byte aligned): asked for now, used after the beta warm-up
   * (loaded where they are used, every half-iteration waited out a global-memory latency there: 0.6 % of the kernel) */
  uint32_t tw[3];
#pragma unroll
  for (int k = 0; k < 3; k++) tw[k] = pk_ldg32(reinterpret_cast<const uint32_t *>(tail) + k);
#endif
  /* pull this half-iteration's rows (first touch comes from HBM) and the next one's towards L2 */
  for (int off = j * 28; off < L * TPC; off += TPC * 32) {      /* each of the TPC threads takes every TPC-th 128-byte line */
    asm volatile("prefetch.global.L2 [%0];" :: "l"(v.Yrow + off));
    asm volatile("prefetch.global.L2 [%0];" :: "l"(v.Ynext + off));
    if (!v.par) asm volatile("prefetch.global.L2 [%0];" :: "l"(v.Srow + off));
  }
  /* a fresh code block starts from apriori = 0; zeroed here so that the HBM fetches above are already under way */
  if (nEz) {
    for (int i = j; i < nEz; i += TPC) Ezero[i] = 0;
    __syncwarp(gmask);
  }

  /* ---- beta warm-up over the first 40 symbols of both sub-blocks (turbodecoder_win.h:321-384) ---- */
#pragma unroll
  for (int i = 0; i < 8; i++) o[i] = MIX ? PK_C : 0u;   /* the reference starts from 8 equal metrics (-10000): only differences matter */
  /* every loop below has the shape  { issue the loads of the NEXT chunk; PK_PIN; work on this chunk; rotate }  and
   * starts with a load-only iteration: no load is in flight when a loop is entered, so the only scoreboard
   * wait on global loads sits at the rotation, a whole chunk after their issue */
#pragma unroll 1
  for (int c = 5; c >= 0; c--) {
    if (c > 0) pk_load_raw<true, true>(v, 8 * (c - 1), 0, nraw);
    PK_PIN();
    if (c < 5) {
      pk_finish_chunk<true>(v, 8 * c, 0, raw, xs, es);
#pragma unroll
      for (int i = 7; i >= 0; i--) {
        if (MIX) { pk_beta_s(o, pk_gammas(xs[i], raw.y[i])); if ((i & 3) == 0) pk_norm_s(o); continue; }
        pk_beta(o, xs[i], raw.y[i]);
        if (TRK) { if (!(i & 1) && (i != 0 || c != 0)) pk_norm_trk(o, tb); }     /* even k != 0 (:244-261) */
        else if ((i & 3) == 0) pk_norm(o);
      }
    }
    raw = nraw;
  }
  /* lane shift: sub-block d starts from the estimate of sub-block d+1 (:347-355); the last sub-block of the
   * code block starts from the known end state through the 3 tail steps with WRAPPING adds (:263-307) */
  {
    int t[8];
    t[0] = 0;
#pragma unroll
    for (int i = 1; i < 8; i++) t[i] = -10000;
#pragma unroll
    for (int k = 2; k >= 0; k--) {
#if PK_TAIL_EARLY
      const int tx = (short)(tw[k] & 0xFFFFu), ty = (short)(tw[k] >> 16);
#else
      const int tx = tail[2 * k], ty = tail[2 * k + 1];
#endif
      const int txy = (short)(tx + ty);
      int nn[8];
      nn[0] = max((int)(short)(t[4] + txy), t[0]);
      nn[1] = max(t[4], (int)(short)(t[0] + txy));
      nn[2] = max((int)(short)(t[5] + ty), (int)(short)(t[1] + tx));
      nn[3] = max((int)(short)(t[5] + tx), (int)(short)(t[1] + ty));
      nn[4] = max((int)(short)(t[6] + tx), (int)(short)(t[2] + ty));
      nn[5] = max((int)(short)(t[6] + ty), (int)(short)(t[2] + tx));
      nn[6] = max(t[7], (int)(short)(t[3] + txy));
      nn[7] = max((int)(short)(t[7] + txy), t[3]);
#pragma unroll
      for (int i = 0; i < 8; i++) t[i] = nn[i];
    }
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t w = __shfl_down_sync(gmask, o[i], 1, TPC);
      o[i] = (j == TPC - 1) ? ((o[i] >> 16) | ((uint32_t)(t[i] + (MIX ? 32768 : 0)) << 16)) : pk_prmt(o[i], w, 0x5432u);
    }
    if (TRK) pk_trk8(tb, o);                          /* the reference starts the main pass from this state as it is */
    else if (MIX) pk_norm_s(o);
    else pk_norm(o);
  }
  /* ---- beta main pass, chunks aligned from the top; checkpoint c = normalised beta at k = L-8c ---- */
  pk_ck_store(tbase, o);
  {
    /* two chunks (16 steps) per iteration, the loads of the next 16 steps issued first: a single beta chunk is
     * shorter than an L2 hit, and the register rotation at the end of the body has to wait for those loads */
    const int c0 = nfull & 1;                        /* chunk c covers steps [L - 8(c+1), L - 8c) */
    PkRaw rawB, nrawB;
    if (c0) {
      pk_load_raw<true, VEC>(v, L - PK_CH, 0, raw);
      PK_PIN();
      pk_finish_chunk<true>(v, L - PK_CH, 0, raw, xs, es);
      if (TRK) pk_beta_chunk_trk(o, xs, raw.y, tbase + 8, kpar, L != PK_CH, tb);
      else pk_beta_chunk<true, MIX>(o, xs, raw.y, 0, tbase + 8);
    }
#pragma unroll 1
    for (int c = c0 - 2; c < nfull; c += 2) {
      const int k0 = L - PK_CH * (c + 1);
      if (c + 2 < nfull) { pk_load_raw<true, VEC>(v, k0 - 2 * PK_CH, 0, nraw); pk_load_raw<true, VEC>(v, k0 - 3 * PK_CH, 0, nrawB); }
#if PK_PF_BETA
      /* the L2 round trip under load is longer than the 16 steps the register loads are issued ahead (3.7 % of the kernel's
       * warp-cycles waited at the rotation below): the lines of the iteration after the next go to L1 now. Measured: decode
       * phase -1 % with early stop, +1.3 % with all 8 half-iterations, pipelined step unchanged — off (tools/experiments/README.md) */
      if (VEC && c + 4 < nfull) { pk_prefetch_raw(v, k0 - 4 * PK_CH); pk_prefetch_raw(v, k0 - 5 * PK_CH); }
#endif
      PK_PIN();
      if (c >= c0) {
        pk_finish_chunk<true>(v, k0, 0, raw, xs, es);
        if (TRK) pk_beta_chunk_trk(o, xs, raw.y, tbase + 8 * (c + 1), kpar, true, tb);
        else pk_beta_chunk<true, MIX>(o, xs, raw.y, 0, tbase + 8 * (c + 1));
        pk_finish_chunk<true>(v, k0 - PK_CH, 0, rawB, xs, es);
        if (TRK) pk_beta_chunk_trk(o, xs, rawB.y, tbase + 8 * (c + 2), kpar, k0 != PK_CH, tb);
        else pk_beta_chunk<true, MIX>(o, xs, rawB.y, 0, tbase + 8 * (c + 2));
      }
      raw = nraw; rawB = nrawB;
    }
    /* the bottom partial chunk only serves the (unused) end of the recursion: nothing to do */
  }
  pk_ck_store_fence();
  if (TRK) acc.rb = __vmaxs2(tb.hi, ~tb.lo);        /* max(hi, -lo - 1) per lane; the certificate adds the 1 back */
  /* ---- alpha warm-up over the last 40 symbols (:459-506), normalisation by the warm-up counter ---- */
  {
    uint32_t m[8], n[8];
#pragma unroll
    for (int i = 0; i < 8; i++) o[i] = MIX ? PK_C : 0u;
#pragma unroll 1
    for (int c = -1; c < 5; c++) {
      if (c < 4) pk_load_raw<true, VEC>(v, L - 40 + 8 * (c + 1), 0, nraw);
      PK_PIN();
      if (c >= 0) {
        pk_finish_chunk<true>(v, L - 40 + 8 * c, 0, raw, xs, es);
#pragma unroll
        for (int i = 0; i < 8; i++) {
          if (MIX) {
            pk_alpha_branches_s(o, pk_to_i(xs[i]), pk_to_i(raw.y[i]), m, n);
#pragma unroll
            for (int s = 0; s < 8; s++) o[s] = pk_max_u(m[s], n[s]);
            if ((i & 3) == 3) pk_norm_s(o);
            continue;
          }
          pk_alpha_branches(o, xs[i], raw.y[i], m, n);
#pragma unroll
          for (int s = 0; s < 8; s++) o[s] = pk_max(m[s], n[s]);
          if (TRK) { if (!(i & 1) && (i != 0 || c != 0)) pk_norm_trk(o, acc.ta); }   /* warm-up counter even, != 0 (:459-506) */
          else if ((i & 3) == 3) pk_norm(o);
        }
      }
      raw = nraw;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint32_t w = __shfl_up_sync(gmask, o[i], 1, TPC);
    const uint32_t first = MIX ? ((i == 0) ? 0x8000u : 0x58F0u) : ((i == 0) ? 0u : 0xD8F0u);   /* 0 / -10000, offset-binary in S form */
    o[i] = (j == 0) ? (first | (o[i] << 16)) : pk_prmt(w, o[i], 0x5432u);
  }
  if (TRK) {
    /* start state of the main pass as it is; the known start state (0, -10000 x 7) of sub-block 0 is covered by the
     * certificate's constant term, not by the measurement */
    uint32_t tv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) tv[i] = (j == 0) ? (o[i] & 0xFFFF0000u) : o[i];
    pk_trk8(acc.ta, tv);
  } else if (MIX) pk_norm_s(o);
  else pk_norm(o);
  /* ---- alpha main pass, ascending k: bottom partial chunk (if any), then the full chunks ---- */
  acc.wlo = 0; acc.whi = 0;
  int wi = 0, cc = 0;
  uint2 rr[8];
  uint32_t ck[8];
  if (rem) {
    const int k0 = rem - PK_CH, i0 = PK_CH - rem;
    pk_ck_load_issue(tbase + 8 * nfull, ck);
    pk_load_raw<false, VEC>(v, k0, i0, raw);
    pk_load_crc<false, VEC, TPC>(v, k0, i0, rr);
    PK_PIN();
    pk_finish_chunk<false>(v, k0, i0, raw, xs, es, ka);
    pk_ck_load_wait(ck);
    pk_alpha_chunk<false, SHIFT, TRK, MIX>(v, k0, o, ck, i0, xs, raw.y, es, rr, acc, ka, kpar, false);
    cc = 1;
  }
#pragma unroll 1
  for (int c = nfull; c >= 0; c--) {                 /* iteration c works on steps [k0, k0+8), k0 = rem + 8(nfull-1-c) */
    const int k0 = rem + PK_CH * (nfull - 1 - c);
    if (c > 0) pk_load_raw<true, VEC>(v, k0 + PK_CH, 0, nraw);
    if (c < nfull) { pk_ck_load_issue(tbase + 8 * c, ck); pk_load_crc<true, VEC, TPC>(v, k0, 0, rr); }
    PK_PIN();
    if (c < nfull) {
      pk_finish_chunk<true>(v, k0, 0, raw, xs, es, ka);
      pk_ck_load_wait(ck);
      pk_alpha_chunk<true, SHIFT, TRK, MIX>(v, k0, o, ck, 0, xs, raw.y, es, rr, acc, ka, kpar, c == 0);
      if ((cc & 3) == 3) { bits_lo[wi] = acc.wlo - 1u; bits_hi[wi] = acc.whi - 1u; wi++; }
      cc++;
    }
    raw = nraw;
  }
  if (cc & 3) { const int sh = 32 - 8 * (cc & 3); bits_lo[wi] = (acc.wlo - 1u) << sh; bits_hi[wi] = (acc.whi - 1u) << sh; }
#if PK_PF_L1
  /* the next half-iteration starts with the first 40 steps of the other decoder's rows (10 lines per stream):
   * pull them into L1 while the CRC is reduced and the slot decides how to go on */
  pk_prefetch_warm<TPC>(v.Ynext - 4 * j, v.par ? v.Srow - 4 * j : nullptr, j);
#endif
}

/* ---------------------------------------------------------------- output of a finished code block ---- */
/* n bits (1..32) of bit row `row` (MSB-first words) starting at bit k, left-aligned in the result */
__device__ __forceinline__ uint32_t pk_row_bits(const uint32_t *row, int k)
{
  const int wi = k >> 5, off = k & 31;
  return __funnelshift_l(row[wi + 1], row[wi], off);
}

/* natural-order hard decisions -> K/8 bytes MSB first (turbodecoder_win.h:676-716). rows: [P][nw+1] words,
 * row d holds bits d*L .. d*L+L-1 starting at bit index `pad`. Called by the TPC threads of one code block. */
template <int TPC>
__device__ __forceinline__ void pk_write_output(const uint32_t *__restrict__ rows, int nw1, int pad, int K, int L, int j, uint8_t *__restrict__ out)
{
  const int nwords = K >> 5;
  uint32_t *out32 = reinterpret_cast<uint32_t *>(out);
  int d = (j << 5) / L, k = (j << 5) - d * L;
#pragma unroll 4
  for (int ow = j; ow < nwords; ow += TPC) {
    const int avail = L - k;
    uint32_t w = pk_row_bits(rows + d * nw1, k + pad);
    if (avail < 32) w = (w & ~(0xFFFFFFFFu >> avail)) | (pk_row_bits(rows + (d + 1) * nw1, pad) >> avail);
    out32[ow] = __byte_perm(w, 0, 0x0123);
    k += 32 * TPC;
    while (k >= L) { k -= L; d++; }
  }
  for (int ob = (nwords << 2) + j; ob < (K >> 3); ob += TPC) {      /* K/8 not a multiple of 4: last 1..3 bytes */
    const int p = ob << 3, d = p / L, k = p - d * L, avail = L - k;
    uint32_t w = pk_row_bits(rows + d * nw1, k + pad);
    if (avail < 8) w = (w & ~(0xFFFFFFFFu >> avail)) | (pk_row_bits(rows + (d + 1) * nw1, pad) >> avail);
    out[ob] = (uint8_t)(w >> 24);
  }
}

/* A finished code block: natural-order bit rows (decoder 2 leaves them in interleaved order) and the output bytes.
 * Kept out of line so that this cold code does not take part in the register allocation of the trellis loops. */
template <int TPC>
__device__ PK_EPI_INLINE void pk_finish_block(int par, uint32_t Eb, const uint16_t *Tlo, const uint16_t *Thi, const uint32_t *erow,
                                              uint32_t *blo, uint32_t *bhi, const uint32_t *rows, int nw1, int pad, int K, int L, int j,
                                              unsigned gmask, uint8_t *out)
{
    if (par) {
      /* decisions are in interleaved order: spread them to natural positions through the (now dead) E array
       * (a halfword whose bit 15 is the decision), then rebuild the natural bit rows from bits 15 / 31 of the pair
       * words. Both loops work on 8 steps at a time, all loads first: with one warp per scheduler a load-use
       * pair per step costs a full shared-memory latency. */
      if ((L & 3) == 0) {
        for (int k0 = 0; k0 < L; k0 += 8) {
          const int kb = k0 + pad, wi = kb >> 5, sh = kb & 31;
          const bool two = k0 + 4 < L;
          const uint2 tl0 = *reinterpret_cast<const uint2 *>(Tlo + k0), th0 = *reinterpret_cast<const uint2 *>(Thi + k0);
          uint2 tl1 = tl0, th1 = th0;
          if (two) { tl1 = *reinterpret_cast<const uint2 *>(Tlo + k0 + 4); th1 = *reinterpret_cast<const uint2 *>(Thi + k0 + 4); }
          const uint32_t fl = __funnelshift_l(blo[wi + 1], blo[wi], sh), fh = __funnelshift_l(bhi[wi + 1], bhi[wi], sh);   /* steps k0.. at bits 31.. */
          const uint32_t ol[8] = {tl0.x & 0xFFFFu, tl0.x >> 16, tl0.y & 0xFFFFu, tl0.y >> 16, tl1.x & 0xFFFFu, tl1.x >> 16, tl1.y & 0xFFFFu, tl1.y >> 16};
          const uint32_t oh[8] = {th0.x & 0xFFFFu, th0.x >> 16, th0.y & 0xFFFFu, th0.y >> 16, th1.x & 0xFFFFu, th1.x >> 16, th1.y & 0xFFFFu, th1.y >> 16};
#pragma unroll
          for (int i = 0; i < 8; i++) {
            if (i < 4 || two) {
              pk_sts_u16(Eb + ol[i], fl >> (16 - i));
              pk_sts_u16(Eb + oh[i], fh >> (16 - i));
            }
          }
        }
      } else {
        for (int k = 0; k < L; k++) {
          const int kb = k + pad;
          pk_sts_u16(Eb + Tlo[k], 0u - ((blo[kb >> 5] >> (31 - (kb & 31))) & 1u));
          pk_sts_u16(Eb + Thi[k], 0u - ((bhi[kb >> 5] >> (31 - (kb & 31))) & 1u));
        }
      }
      __syncwarp(gmask);
      /* 16 steps at a time: bits 15 / 31 of word i go to bits 15-i / 31-i of u (both lanes with one shift + one
       * LOP3), then the two 16-bit fields are appended to the row words */
      uint32_t wlo = 0, whi = 0;
      int fill = pad;                                   /* bits already in the current row word */
      int wi = 0;
      for (int k0 = 0; k0 < L; k0 += 16) {
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 16; i++) w[i] = (k0 + i < L) ? erow[k0 + i] : 0u;
        uint32_t u = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) u |= (w[i] >> i) & (0x80008000u >> i);
        const int n = min(16, L - k0);                  /* valid steps in u: the top n bits of each half */
        const uint32_t flo = u << 16, fhi = u & 0xFFFF0000u;        /* left-aligned 16-bit fields */
        /* append n bits (left-aligned in f) to the MSB-first accumulators */
        if (fill + n < 32) {
          wlo |= flo >> fill; whi |= fhi >> fill; fill += n;
        } else {
          wlo |= flo >> fill; whi |= fhi >> fill;
          blo[wi] = wlo; bhi[wi] = whi; wi++;
          const int used = 32 - fill;                   /* bits of the field that went into the finished word */
          wlo = used < 16 ? flo << used : 0u; whi = used < 16 ? fhi << used : 0u;
          fill = n - used;
        }
      }
      if (fill) { blo[wi] = wlo; bhi[wi] = whi; }
    }
    __syncwarp(gmask);
    pk_write_output<TPC>(rows, nw1, pad, K, L, j, out);
}

/* A code block that leaves the a-priori range in mid-decode continues in the certificate kernel where it stands: its state
 * is the extrinsic array E (shared memory) and the half-iteration count. E goes to a scratch slot indexed like the hand-over
 * list entry; blocks beyond the scratch capacity restart from their soft buffer (cont = 0). Cold path, kept out of line. */
#ifndef PK_HO_INLINE
#define PK_HO_INLINE 0
#endif
#ifndef PK_HO_REFILL_CALL
#define PK_HO_REFILL_CALL 0    /* 1: the refill loop also goes through pk_hand_over — measured 3 % slower on the headline workload
                                  (MAP kernel 1.455 vs 1.412 ms, tools/experiments/ab_handover.sh): a second call site in the
                                  refill loop costs the hot path registers although it never runs there */
#endif
#ifndef PK_HO_CONT
#define PK_HO_CONT 1
#endif
#if PK_HO_INLINE
#define PK_HO_ATTR __forceinline__
#else
#define PK_HO_ATTR __noinline__
#endif
template <int TPC>
__device__ PK_HO_ATTR void pk_hand_over(unsigned int *trk_count, uint32_t *trk_list, uint32_t *trk_cont, uint32_t *econt, uint32_t econt_cap,
                                          uint32_t cb, const uint32_t *Eslot, int Es, uint32_t n1, uint32_t G, int j, int g, unsigned gmask)
{
  uint32_t idx = 0;
  if (j == 0) idx = atomicAdd(trk_count, 1u);
  idx = __shfl_sync(gmask, idx, g * TPC);
  if (j == 0) trk_list[idx] = cb;
  if (n1 && idx < econt_cap) {
    uint32_t *dst = econt + (size_t)idx * Es;
    for (int i = j; i < Es; i += TPC) dst[i] = Eslot[i];
    if (j == 0) trk_cont[idx] = n1 | (G << 8);
  } else if (j == 0) trk_cont[idx] = 0;
}

/* ---------------------------------------------------------------- run-time range certificate ---- */
/* Inputs: packed per-lane extremes of one code block's half-iteration in the reference's frame (TRK schedule):
 * [alo, ahi] alpha states after every normalisation and at the start of the main pass, rb = max(bhi, -blo - 1) for
 * beta, [vmn, vmx] the two LLR maxima m0/m1 of every step; G = max|x| + max|y| (a priori, as for FEC_PK_GMAX).
 * With Ra, Rb the state magnitudes: a state entering a normalisation window is within R, a window is at most 3 steps,
 * every step adds at most G in magnitude, so
 *   states and branch sums stay within R + 3G, the differences o[i] - o[0] of the normalisation within 2R + 6G   (C1)
 *   LLR sums alpha-branch + beta stay within (Ra + 3G) + (Rb + 2G)                                               (C2)
 *   sub-block 0 starts from (0, -10000 x 7): within 10000 + 3G for its first three steps, LLR sums there within
 *   10000 + 3G + Rb + 2G; the reference's warm-ups start from -10000: 10000 + 3G                                 (C3)
 *   the LLR m1 - m0 stays within vmx - vmn                                                                      (C4)
 * If all four are <= 32767, induction over the windows (the first state is exact, so its measured extremes are exact,
 * so its window cannot overflow, so the next state is exact ...) shows that no saturating add of the reference clipped
 * and no wrapping add of this kernel wrapped: both computed the same integers. */
__device__ __forceinline__ bool pk_certificate(uint32_t ahi, uint32_t alo, uint32_t rb, uint32_t vmx, uint32_t vmn, int G)
{
  const int a_hi = max((int)(short)(ahi & 0xFFFFu), (int)(short)(ahi >> 16));
  const int a_lo = min((int)(short)(alo & 0xFFFFu), (int)(short)(alo >> 16));
  const int Ra = max(a_hi, -a_lo);
  const int Rb = max((int)(short)(rb & 0xFFFFu), (int)(short)(rb >> 16)) + 1;
  const int v0 = (int)(short)(vmx & 0xFFFFu) - (int)(short)(vmn & 0xFFFFu), v1 = (int)(short)(vmx >> 16) - (int)(short)(vmn >> 16);
  const int R = max(Ra, Rb);
  return 2 * R + 6 * G <= 32767 && Ra + Rb + 5 * G <= 32767 && 10000 + Rb + 5 * G <= 32767 && max(v0, v1) <= 32767;
}

#if PK_TIMELINE
/* diagnosis build only (tools/experiments/timeline.py): start / queue-empty / end time of every CTA of the first kernel */
__device__ unsigned long long pk_tl_buf[6 * 16384];
__device__ unsigned int pk_tl_n;
__device__ __forceinline__ unsigned long long pk_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
extern "C" int fec_cuda_debug_timeline(unsigned long long *out, unsigned int cap)
{
  unsigned int n = 0;
  cudaMemcpyFromSymbol(&n, pk_tl_n, sizeof(n));
  if (n > 16384) n = 16384;
  if (n > cap) n = cap;
  cudaMemcpyFromSymbol(out, pk_tl_buf, sizeof(unsigned long long) * 6 * n);
  unsigned int z = 0;
  cudaMemcpyToSymbol(pk_tl_n, &z, sizeof(z));
  return (int)n;
}
#endif

/* ---------------------------------------------------------------- the kernel ---- */
/* shared memory: [Tlo, Thi: TPC*Lw halfwords each][E: nslots x (TPC*Lw+1) words][bit rows: nslots x P x nw1 words] */
size_t fec_pk_smem_bytes(const fec_geom_t *g)
{
  const size_t tpc = g->P / 2, nslots = (size_t)FEC_PK_WARPS * (32 / tpc), nw1 = (g->L + 31) / 32 + 1;
  return 4 * (tpc * g->Lw + nslots * (tpc * g->Lw + 1) + nslots * g->P * nw1);
}

/* TRK = false: the common kernel, code blocks within the a-priori range FEC_PK_GMAX; what leaves it goes to trk_list.
 * TRK = true: works through that list with the run-time range certificate; what the certificate cannot cover goes to
 * exact_list (the 32-bit kernel). Both restart a handed-over code block from its soft buffer. */
template <int P, bool VEC, bool TRK>
__global__ void __launch_bounds__(FEC_PK_WARPS * 32, 1) fec_tdec_pk_kernel(const FecDecodeArgs a)
{
  if (TRK) { if (*a.list_count == 0) return; }
#if PK_TIMELINE
  unsigned long long tl0 = pk_gtime(), tl1 = 0;
#endif
  constexpr int TPC = P / 2, CPW = 32 / TPC;
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(16) FecCbDesc s_desc[FEC_PK_WARPS * (32 / (P / 2))];   /* descriptor of the code block in each slot */
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane / TPC, j = lane % TPC;
  const int K = a.g.K, L = a.g.L, Lp = a.g.Lp, Lw = a.g.Lw, SS = a.g.SS;
  const int nw1 = (L + 31) / 32 + 1, Es = TPC * Lw + 1, nslots = FEC_PK_WARPS * CPW, slot = warp * CPW + g;
  const int pad = (8 - (L & 7)) & 7;
  uint16_t *Tlo = reinterpret_cast<uint16_t *>(smem), *Thi = Tlo + TPC * Lw;
  uint32_t *Eslot = reinterpret_cast<uint32_t *>(smem) + TPC * Lw + slot * Es;
  uint32_t *rows = reinterpret_cast<uint32_t *>(smem) + TPC * Lw + nslots * Es + slot * (P * nw1);
  const unsigned gmask = (TPC == 32) ? 0xFFFFFFFFu : (((1u << TPC) - 1u) << (g * TPC));

  /* the decoder-2 address table of this K into shared memory, 4 independent loads in flight per thread (one load per
   * iteration had every thread wait out 23 global-memory latencies before the CTA's first code block) */
  for (int r = 0; r < TPC; r++)
    for (int k0 = tid; k0 < L; k0 += 4 * FEC_PK_WARPS * 32) {
      uint32_t w[4];
#pragma unroll
      for (int u = 0; u < 4; u++) { const int k = k0 + u * FEC_PK_WARPS * 32; w[u] = k < L ? a.pair_addr[r * Lp + k] : 0u; }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int k = k0 + u * FEC_PK_WARPS * 32;
        if (k < L) { Tlo[r * Lw + k] = (uint16_t)((w[u] & 0xFFFFu) * 2u); Thi[r * Lw + k] = (uint16_t)((w[u] >> 16) * 2u); }
      }
    }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"((uint32_t)__cvta_generic_to_shared(&s_tmem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = s_tmem + ((uint32_t)(warp * 32) << 16);

  if (j == 0) s_desc[slot] = a.cbs[0];
  __syncwarp();
  int cb = -1, n = 0;
  uint32_t M = 0, G = 0;                         /* G: range figure max|x| + max|y| of the NEXT half-iteration of this slot */
  bool exhausted = false, fresh = false;
  for (;;) {
    /* ---- refill this group's slot from the work queue (group-uniform control flow) ---- */
    while (cb < 0 && !exhausted) {
      unsigned q = 0;
      if (j == 0) q = atomicAdd(a.queue, 1u);
      q = __shfl_sync(gmask, q, g * TPC);
      if (q >= (TRK ? *a.list_count : a.ncb)) {
        exhausted = true;
#if PK_TIMELINE
        if (!tl1) tl1 = pk_gtime();
#endif
        break;
      }
      const unsigned qi = q;
      if (TRK) q = a.list[q];
#if PK_OPT_DESC
      {
        /* descriptor (32 bytes, lanes 0/1 of the group) and max|soft| are fetched side by side */
        const uint4 *src = reinterpret_cast<const uint4 *>(a.cbs + q);
        uint4 dq = make_uint4(0, 0, 0, 0);
        if (j < 2) dq = pk_ldg128(src + j);
        M = a.soft_max[a.status_off + q];
        if (j < 2) reinterpret_cast<uint4 *>(&s_desc[slot])[j] = dq;
      }
#else
      M = a.soft_max[a.cbs[q].status];
#endif
      if (2 * M > FEC_PK_GMAX_TRK) {                /* first half-iteration already outside what the certificate can cover */
        if (j == 0) a.exact_list[atomicAdd(a.exact_count, 1u)] = q;
        continue;
      }
      if (!TRK && 2 * M + M / 2 > FEC_PK_GMAX) {    /* outside the a-priori range now or (decoder 1: G = 2M + max|extrinsic|) certain to
                                                       leave it with the second iteration: certificate kernel from the start */
#if PK_HO_REFILL_CALL
        pk_hand_over<TPC>(a.trk_count, a.trk_list, a.trk_cont, a.econt, a.econt_cap, q, Eslot, Es, 0u, 0u, j, g, gmask);
#else
        if (j == 0) { const uint32_t idx = atomicAdd(a.trk_count, 1u); a.trk_list[idx] = q; a.trk_cont[idx] = 0; }
#endif
        continue;
      }
      cb = (int)q; n = 0; G = 2 * M;
      fresh = true;
      if (TRK) {
        const uint32_t cont = a.trk_cont[qi];
        if (cont) {                                   /* handed over in mid-decode: go on from its extrinsics */
          const uint32_t *src = a.econt + (size_t)qi * Es;
          for (int i = j; i < Es; i += TPC) Eslot[i] = src[i];
          n = (int)(cont & 0xFFu); G = cont >> 8;
          fresh = false;
        }
      }
    }
    __syncwarp();
    if (!__any_sync(0xFFFFFFFFu, cb >= 0)) break;

    /* ---- one half-iteration for every slot of the warp; idle slots run on code block 0's inputs and their
     *      own scratch E (keeps the warp convergent for the .aligned tensor-memory accesses), results dropped ---- */
    const bool active = cb >= 0;
#if PK_OPT_DESC
    __syncwarp();
    const FecCbDesc desc = s_desc[slot];             /* idle slots keep their last (or code block 0's) descriptor */
#else
    const FecCbDesc desc = a.cbs[active ? cb : 0];
#endif
    const int par = n & 1;
    PkView v;
    v.par = par;
    v.tpc4 = 4 * TPC;
    v.Srow = reinterpret_cast<const uint32_t *>(desc.soft) + 4 * j;
    v.Yrow = reinterpret_cast<const uint32_t *>(desc.soft + (1 + par) * SS) + 4 * j;
    v.Ynext = reinterpret_cast<const uint32_t *>(desc.soft + (2 - par) * SS) + 4 * j;
    v.Rrow = a.pair_crc[desc.crc_sel] + (size_t)par * TPC * Lp + 2 * j;
    v.Tlo = Tlo + j * Lw; v.Thi = Thi + j * Lw;
    v.Eb = (uint32_t)__cvta_generic_to_shared(Eslot);
    v.ident = (uint32_t)(4 * j * Lw);
    PkAcc acc;
    acc.crc = 0; acc.emx = 0x80008000u; acc.emn = 0x7FFF7FFFu;
    uint32_t *blo = rows + (2 * j) * nw1, *bhi = rows + (2 * j + 1) * nw1;
    pk_half_iteration<TPC, VEC, P == 8, TRK>(v, L, tbase, gmask, j, desc.soft + 3 * SS + (par ? 6 : 0), blo, bhi, acc,
                                              Eslot, fresh ? Es : 0);
    fresh = false;
    bool cert = true;
    if (TRK) {
      uint32_t ahi = acc.ta.hi, alo = acc.ta.lo, rb = acc.rb, vmx = acc.vmx, vmn = acc.vmn;
#pragma unroll
      for (int s = TPC / 2; s > 0; s >>= 1) {
        ahi = pk_max(ahi, __shfl_xor_sync(0xFFFFFFFFu, ahi, s, TPC)); alo = pk_min(alo, __shfl_xor_sync(0xFFFFFFFFu, alo, s, TPC));
        rb = pk_max(rb, __shfl_xor_sync(0xFFFFFFFFu, rb, s, TPC));
        vmx = pk_max(vmx, __shfl_xor_sync(0xFFFFFFFFu, vmx, s, TPC)); vmn = pk_min(vmn, __shfl_xor_sync(0xFFFFFFFFu, vmn, s, TPC));
      }
      if (G > FEC_PK_GMAX) cert = pk_certificate(ahi, alo, rb, vmx, vmn, (int)G);
    }
    uint32_t crc = acc.crc, emx = acc.emx, emn = acc.emn;
#pragma unroll
    for (int s = TPC / 2; s > 0; s >>= 1) {
      crc ^= __shfl_xor_sync(0xFFFFFFFFu, crc, s, TPC);
      emx = pk_max(emx, __shfl_xor_sync(0xFFFFFFFFu, emx, s, TPC));
      emn = pk_min(emn, __shfl_xor_sync(0xFFFFFFFFu, emn, s, TPC));
    }
    __syncwarp();
    if (active && !cert) {
      /* the certificate does not cover this half-iteration: its results are not used, the exact kernel re-decodes the block */
      if (j == 0) a.exact_list[atomicAdd(a.exact_count, 1u)] = (uint32_t)cb;
      cb = -1;
    } else if (active) {
      const int n1 = n + 1, crc_ok = (crc == 0);
      const bool done = (crc_ok && !(desc.flags & FEC_F_NO_EARLY)) || (n1 >= (int)desc.max_its);
      if (done) {
        pk_finish_block<TPC>(par, v.Eb, v.Tlo, v.Thi, Eslot + j * Lw, blo, bhi, rows, nw1, pad, K, L, j, gmask, desc.out);
        if (j == 0) { a.status[desc.status].its = (uint32_t)n1; a.status[desc.status].crc_ok = (uint32_t)crc_ok; }
        __syncwarp(gmask);
        cb = -1;
      } else {
        /* range condition of the NEXT half-iteration: |x| <= max|new extrinsic| (+ max|systematic| for decoder 1) */
        const int e0 = (short)(emx & 0xFFFFu), e1 = (short)(emx >> 16), e2 = -(int)(short)(emn & 0xFFFFu), e3 = -(int)(short)(emn >> 16);
        const uint32_t emax = (uint32_t)max(max(e0, e1), max(max(e2, e3), 0));
        const uint32_t bx = (n1 & 1) ? emax : emax + M;
        G = max(bx, M) + M;
        if (G > FEC_PK_GMAX_TRK) {
          if (j == 0) a.exact_list[atomicAdd(a.exact_count, 1u)] = (uint32_t)cb;   /* re-decoded from scratch by the exact kernel */
          cb = -1;
        } else if (!TRK && G > FEC_PK_GMAX) {
#if PK_HO_CONT
          pk_hand_over<TPC>(a.trk_count, a.trk_list, a.trk_cont, a.econt, a.econt_cap, (uint32_t)cb, Eslot, Es, (uint32_t)n1, G, j, g, gmask);   /* the certificate kernel goes on from here */
#else
          if (j == 0) { const uint32_t idx = atomicAdd(a.trk_count, 1u); a.trk_list[idx] = (uint32_t)cb; a.trk_cont[idx] = 0; }
#endif
          cb = -1;
        } else n = n1;
      }
    }
    __syncwarp();
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(s_tmem) : "memory");
#if PK_TIMELINE
  if (!TRK && tid == 0) {
    unsigned smid; asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    const unsigned i = atomicAdd(&pk_tl_n, 1u);
    if (i < 16384) {
      unsigned long long *r = pk_tl_buf + 6 * (size_t)i;
      r[0] = (unsigned long long)(size_t)a.queue; r[1] = smid; r[2] = tl0; r[3] = tl1; r[4] = pk_gtime(); r[5] = blockIdx.x;
    }
  }
#endif
}

template <int P, bool TRK>
static cudaError_t pk_launch(const FecDecodeArgs &a, int grid, size_t smem, cudaStream_t st)
{
  if ((a.g.L & 3) == 0) fec_tdec_pk_kernel<P, true, TRK><<<grid, FEC_PK_WARPS * 32, smem, st>>>(a);
  else                  fec_tdec_pk_kernel<P, false, TRK><<<grid, FEC_PK_WARPS * 32, smem, st>>>(a);
  return cudaGetLastError();
}

cudaError_t fec_launch_decode_packed(const FecDecodeArgs &a, int sm_count, cudaStream_t st, int stages)
{
  if (a.ncb == 0) return cudaSuccess;
  const size_t smem = fec_pk_smem_bytes(&a.g);
  const int tpc = a.g.P / 2, per_cta = FEC_PK_WARPS * (32 / tpc);
  int grid = (int)((a.ncb + per_cta - 1) / per_cta);
  if (grid > sm_count) grid = sm_count;
  cudaError_t e = cudaSuccess;
  if (stages & FEC_ST_A) e = a.g.P == 16 ? pk_launch<16, false>(a, grid, smem, st) : pk_launch<8, false>(a, grid, smem, st);
  if (e != cudaSuccess || !(stages & FEC_ST_B)) return e;
  /* the certificate kernel works through what the first one handed over; its CTAs leave at once when the list is empty */
  FecDecodeArgs b = a;
  b.list = a.trk_list; b.list_count = a.trk_count; b.queue = a.trk_queue;
  return a.g.P == 16 ? pk_launch<16, true>(b, grid, smem, st) : pk_launch<8, true>(b, grid, smem, st);
}

template <int P, bool VEC, bool TRK> static cudaError_t pk_attr()
{
  /* opt in to everything the SM has: 227 KB per block minus this kernel's static (and system-reserved) shared memory */
  cudaFuncAttributes fa;
  cudaError_t e = cudaFuncGetAttributes(&fa, fec_tdec_pk_kernel<P, VEC, TRK>);
  if (e != cudaSuccess) return e;
  int dev = 0, optin = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return cudaFuncSetAttribute(fec_tdec_pk_kernel<P, VEC, TRK>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
}

cudaError_t fec_pk_init(void)
{
  cudaError_t e;
  if ((e = pk_attr<16, true, false>()) != cudaSuccess) return e;
  if ((e = pk_attr<16, false, false>()) != cudaSuccess) return e;
  if ((e = pk_attr<8, true, false>()) != cudaSuccess) return e;
  if ((e = pk_attr<8, false, false>()) != cudaSuccess) return e;
  if ((e = pk_attr<16, true, true>()) != cudaSuccess) return e;
  if ((e = pk_attr<16, false, true>()) != cudaSuccess) return e;
  if ((e = pk_attr<8, true, true>()) != cudaSuccess) return e;
  return pk_attr<8, false, true>();
}
