/*
 * fec_kernels.cu — hand-written sm_100a kernels of the transport-block decode path.
 *
 *   fec_tdec_win_kernel<P>   windowed max-log-MAP turbo decoder with in-kernel half-iteration driver,
 *                            QPP (de)interleaving, hard decision, fused CRC and early stop
 *                            (replaces tdec_win{sse16,avx16}_dec + run_tdec_iteration_16bit + the CRC loop
 *                             of decode_tb_cb: turbodecoder_win.h:310-586, turbodecoder_iter.h:73-147, sch.c:361-391)
 *   fec_tdec_small_kernel    non-windowed wrapping decoder for K<=400 (turbodecoder_sse.c:105-353)
 *   fec_rm_rx_kernel         rate de-matching + HARQ combine as a gather (rm_turbo.c:378-430)
 *   fec_extract_kernel       natural [s p p'] input -> device layout (turbodecoder_win.h:634-674)
 *   fec_tb_assemble_kernel   code-block concatenation + CRC24A (sch.c:396-418, 478-485)
 *
 * Arithmetic contract (bit-exact with the reference's AVX2 build): see DESIGN.md section 3.
 * State metrics are kept as BIASED 32-bit values u = v + 32768 in [0,65535] so that the reference's
 * saturating int16 add is ONE instruction: VIADDMNMX.RELU (max(min(u+g,65535),0)).
 */
#include "fec_kernels.cuh"
#include <stdio.h>

#ifndef FEC_WIN8_TMEM
#define FEC_WIN8_TMEM 1      /* int8 two-lane decoder: beta checkpoints in tensor memory (0: shared memory) */
#endif
#define FEC_WIN8_TCOLS 128   /* tensor-memory columns per CTA: (nck + S_CH - 1) * 4 <= 108 for every K of the int8 classes */
#define BIAS   32768
#define NEGINF (BIAS - 10000)
#define S_CH   FEC_WIN_CHUNK

__device__ __forceinline__ int sadd(int u, int g) { return __viaddmin_s32_relu(u, g, 65535); }
__device__ __forceinline__ int sat16(int v) { return max(min(v, 32767), -32768); }
/* B8 = true: the int8 decoders of the 8-bit LLR path (tdec_win{sse8,avx8}, turbodecoder_win.h:97-178): the same biased
 * representation with bias 128 and the clamp at 255, i.e. _mm_adds_epi8 in one instruction */
template <bool B8> __device__ __forceinline__ int saddT(int u, int g) { return __viaddmin_s32_relu(u, g, B8 ? 255 : 65535); }
template <bool B8> __device__ __forceinline__ int satT(int v) { return B8 ? max(min(v, 127), -128) : max(min(v, 32767), -32768); }
#define BIAS_T(B8)   ((B8) ? 128 : BIAS)
#define NEGINF_T(B8) ((B8) ? 128 : NEGINF)          /* INF = 0 in the int8 decoders (turbodecoder_win.h:128,160) */
__device__ __forceinline__ int max8(const int *v)
{
  return max(__vimax3_s32(v[0], v[1], v[2]), __vimax3_s32(__vimax3_s32(v[3], v[4], v[5]), v[6], v[7]));
}

/* pack 8 biased metrics as 8 true int16 (u ^ 0x8000 is the two's-complement int16 of u-32768); int8 decoders: the biased
 * values 0..255 as they are */
template <bool B8>
__device__ __forceinline__ uint4 pack8(const int *o)
{
  const uint32_t x = B8 ? 0u : 0x80008000u;
  uint4 r;
  r.x = __byte_perm(o[0], o[1], 0x5410) ^ x;
  r.y = __byte_perm(o[2], o[3], 0x5410) ^ x;
  r.z = __byte_perm(o[4], o[5], 0x5410) ^ x;
  r.w = __byte_perm(o[6], o[7], 0x5410) ^ x;
  return r;
}
/* sign-extend the low int16 of a word: PRMT with the sign-replicate selector bit (__byte_perm masks it off) */
__device__ __forceinline__ int sext_lo(uint32_t w)
{
  int r;
  asm("prmt.b32 %0, %1, 0, 0x9910;" : "=r"(r) : "r"(w));
  return r;
}
/* unpack to TRUE signed values */
template <bool B8>
__device__ __forceinline__ void unpack8_true(uint4 p, int *b)
{
  if (B8) {
    b[0] = (int)(p.x & 0xFFFFu) - 128; b[1] = (int)(p.x >> 16) - 128;
    b[2] = (int)(p.y & 0xFFFFu) - 128; b[3] = (int)(p.y >> 16) - 128;
    b[4] = (int)(p.z & 0xFFFFu) - 128; b[5] = (int)(p.z >> 16) - 128;
    b[6] = (int)(p.w & 0xFFFFu) - 128; b[7] = (int)(p.w >> 16) - 128;
  } else {
    b[0] = sext_lo(p.x); b[1] = ((int)p.x) >> 16;
    b[2] = sext_lo(p.y); b[3] = ((int)p.y) >> 16;
    b[4] = sext_lo(p.z); b[5] = ((int)p.z) >> 16;
    b[6] = sext_lo(p.w); b[7] = ((int)p.w) >> 16;
  }
}

/* turbodecoder_win.h:244-261. 16-bit (normalize_period 2): every even k != 0, subtract state 0, saturating.
 * int8 (normalize_max, normalize_period 1): every k != 0, subtract the maximum of the 8 states from all of them */
template <bool B8>
__device__ __forceinline__ void win_norm(int k, int *o)
{
  if (B8) {
    if (k != 0) {
      const int t = 128 - max8(o);
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = saddT<true>(o[i], t);
    }
  } else if (((k & 1) == 0) && k != 0) {
    const int t = BIAS - o[0];
#pragma unroll
    for (int i = 1; i < 8; i++) o[i] = sadd(o[i], t);
    o[0] = BIAS;
  }
}

/* backward step (turbodecoder_win.h:397-418); x,y,xy are TRUE signed branch metrics */
template <bool B8>
__device__ __forceinline__ void win_beta_step(int *o, int x, int y)
{
  const int xy = satT<B8>(x + y);
  int n[8];
  n[0] = max(saddT<B8>(o[4], xy), o[0]);
  n[1] = max(o[4], saddT<B8>(o[0], xy));
  n[2] = max(saddT<B8>(o[5], y), saddT<B8>(o[1], x));
  n[3] = max(saddT<B8>(o[5], x), saddT<B8>(o[1], y));
  n[4] = max(saddT<B8>(o[6], x), saddT<B8>(o[2], y));
  n[5] = max(saddT<B8>(o[6], y), saddT<B8>(o[2], x));
  n[6] = max(o[7], saddT<B8>(o[3], xy));
  n[7] = max(saddT<B8>(o[7], xy), o[3]);
#pragma unroll
  for (int i = 0; i < 8; i++) o[i] = n[i];
}

/* forward branches (turbodecoder_win.h:521-539): m = bit 0 hypothesis, n = bit 1 hypothesis */
template <bool B8>
__device__ __forceinline__ void win_alpha_branches(const int *o, int x, int y, int *m, int *n)
{
  const int xy = satT<B8>(x + y);
  m[0] = o[0];                  m[1] = saddT<B8>(o[3], y); m[2] = saddT<B8>(o[4], y); m[3] = o[7];
  m[4] = o[1];                  m[5] = saddT<B8>(o[2], y); m[6] = saddT<B8>(o[5], y); m[7] = o[6];
  n[0] = saddT<B8>(o[1], xy);   n[1] = saddT<B8>(o[2], x); n[2] = saddT<B8>(o[5], x); n[3] = saddT<B8>(o[6], xy);
  n[4] = saddT<B8>(o[0], xy);   n[5] = saddT<B8>(o[3], x); n[6] = saddT<B8>(o[4], x); n[7] = saddT<B8>(o[7], xy);
}

struct WinLane {           /* per-lane view of one half-iteration */
  const int16_t  *S;       /* systematic stream of this lane (ignored when smask == 0) */
  const int16_t  *Y;       /* parity stream of this lane */
  const uint32_t *T;       /* step table of this lane and parity */
  void           *E;       /* extrinsic array of this code block (shared memory): int16, int8 in the int8 decoders */
  int             smask;   /* -1: x = sat(S + E) (decoder 1);  0: x = E (decoder 2) */
  int             tpc, dj, dh; /* pairs per code block, pair index and half of this lane */
};

/* int16 offset of (lane d, step k) inside one stream of the device layout (fec_tables.h) */
__device__ __forceinline__ int win_widx(const WinLane &w, int k) { return 2 * ((((k >> 2) * w.tpc + w.dj) << 2) + (k & 3)) + w.dh; }

template <bool B8> __device__ __forceinline__ int win_e(const WinLane &w, uint32_t idx)
{
  return B8 ? (int)static_cast<const int8_t *>(w.E)[idx] : (int)static_cast<const int16_t *>(w.E)[idx];
}
template <bool B8> __device__ __forceinline__ void win_e_store(const WinLane &w, uint32_t idx, int v)
{
  if (B8) static_cast<int8_t *>(w.E)[idx] = (int8_t)v; else static_cast<int16_t *>(w.E)[idx] = (int16_t)v;
}

/* beta checkpoint / chunk entries: 8 state metrics of one lane. 16-bit decoders: 8 x int16 (16 bytes); int8 decoders: the
 * biased values 0..255 as 8 bytes, which with the int8 extrinsic array halves the CTA's shared memory (4 CTAs per SM) */
template <bool B8> struct WinCk;
template <> struct WinCk<false> {
  typedef uint4 type;
  static __device__ __forceinline__ uint4 pack(const int *o) { return pack8<false>(o); }
  static __device__ __forceinline__ void unpack_true(uint4 p, int *b) { unpack8_true<false>(p, b); }
};
template <> struct WinCk<true> {
  typedef uint2 type;
  static __device__ __forceinline__ uint2 pack(const int *o)
  {
    uint2 r;
    r.x = __byte_perm(__byte_perm(o[0], o[1], 0x4040), __byte_perm(o[2], o[3], 0x4040), 0x5410);
    r.y = __byte_perm(__byte_perm(o[4], o[5], 0x4040), __byte_perm(o[6], o[7], 0x4040), 0x5410);
    return r;
  }
  static __device__ __forceinline__ void unpack_true(uint2 p, int *b)
  {
    b[0] = (int)(p.x & 0xFFu) - 128; b[1] = (int)((p.x >> 8) & 0xFFu) - 128; b[2] = (int)((p.x >> 16) & 0xFFu) - 128; b[3] = (int)(p.x >> 24) - 128;
    b[4] = (int)(p.y & 0xFFu) - 128; b[5] = (int)((p.y >> 8) & 0xFFu) - 128; b[6] = (int)((p.y >> 16) & 0xFFu) - 128; b[7] = (int)(p.y >> 24) - 128;
  }
};

/* Four trellis steps of a lane in three 16-byte loads: the device layout keeps steps 4b..4b+3 of a lane PAIR in one 16-byte
 * group per stream, and the step table is contiguous per lane. The loops below walk the sub-block in such groups and fetch
 * group b-1 / b+1 while group b is computed: with one sub-block per thread and a handful of warps per SM the kernel is bound
 * by the latency of these loads, not by their bandwidth (measured: ~970 cycles per step with one load triple per step). */
struct WinBlk { uint4 s, y, t; };

__device__ __forceinline__ WinBlk win_fetch(const WinLane &w, int kb)
{
  WinBlk b;
  const int g = kb * w.tpc + w.dj;
  b.s = w.smask ? __ldg(reinterpret_cast<const uint4 *>(w.S) + g) : make_uint4(0, 0, 0, 0);
  b.y = __ldg(reinterpret_cast<const uint4 *>(w.Y) + g);
  b.t = __ldg(reinterpret_cast<const uint4 *>(w.T) + kb);
  return b;
}
__device__ __forceinline__ uint32_t win_comp(const uint4 &v, int j) { return j == 0 ? v.x : j == 1 ? v.y : j == 2 ? v.z : v.w; }
__device__ __forceinline__ int win_half(uint32_t v, int dh) { return (int)(short)(dh ? v >> 16 : v); }

/* k = khi-1 down to klo: f(k, s, y, t) with s, y the TRUE systematic (0 for decoder 2) and parity values, t the step-table word */
template <typename F>
__device__ __forceinline__ void win_for_desc(const WinLane &w, int khi, int klo, F &&f)
{
  if (khi <= klo) return;
  const int b_lo = klo >> 2;
  int kb = (khi - 1) >> 2;
  WinBlk cur = win_fetch(w, kb), nxt = cur;
  for (; kb >= b_lo; kb--) {
    if (kb > b_lo) nxt = win_fetch(w, kb - 1);
#pragma unroll
    for (int j = 3; j >= 0; j--) {
      const int k = 4 * kb + j;
      if (k < khi && k >= klo) f(k, win_half(win_comp(cur.s, j), w.dh), win_half(win_comp(cur.y, j), w.dh), win_comp(cur.t, j));
    }
    cur = nxt;
  }
}
/* k = klo up to khi-1 */
template <typename F>
__device__ __forceinline__ void win_for_asc(const WinLane &w, int klo, int khi, F &&f)
{
  if (khi <= klo) return;
  const int b_hi = (khi - 1) >> 2;
  int kb = klo >> 2;
  WinBlk cur = win_fetch(w, kb), nxt = cur;
  for (; kb <= b_hi; kb++) {
    if (kb < b_hi) nxt = win_fetch(w, kb + 1);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int k = 4 * kb + j;
      if (k < khi && k >= klo) f(k, win_half(win_comp(cur.s, j), w.dh), win_half(win_comp(cur.y, j), w.dh), win_comp(cur.t, j));
    }
    cur = nxt;
  }
}

size_t fec_win_smem_bytes(const fec_geom_t *g, uint32_t slots, int llr8)
{
  const size_t nthr = (size_t)slots * g->P;
  const size_t nck = (g->L + S_CH - 1) / S_CH, nw = (g->L + 31) / 32;
  const size_t esz = llr8 ? 1 : 2, cksz = llr8 == 2 ? 0 : llr8 ? 8 : 16;    /* llr8 == 2: checkpoints and chunk buffer in tensor memory */
  size_t b = 0;
  b += (((size_t)slots * g->P * g->LpE * esz) + 15) & ~(size_t)15; /* E */
  b += nck * nthr * cksz;                                          /* beta checkpoints */
  b += (size_t)(S_CH - 1) * nthr * cksz;                           /* beta chunk buffer */
  b += nw * nthr * 4;                                              /* decision words per lane */
  if (llr8 != 2) b += (size_t)slots * ((g->K + 31) / 32) * 4;      /* output assembly (tensor-memory variant: inside E) */
  return b;
}

uint32_t fec_win_pick_slots(const fec_geom_t *g, int llr8)
{
  uint32_t s = FEC_WIN_THREADS / g->P;
  if (llr8) {
    /* the int8 kernel is latency-bound: take the CTA size that keeps most code blocks resident per SM (shared memory
     * 227 KB with 1 KB reserved per CTA; 128 registers per thread -> 512 threads) */
    uint32_t best = 1, best_cbs = 0;
    s = FEC_WIN_THREADS / (g->P / 2);             /* two lanes per thread (fec_tdec_win8p_kernel) */
    if (FEC_WIN8_TMEM) return s;                   /* tensor-memory variant: full 128-thread CTAs (whole warps, 4 lane quarters) */
    for (uint32_t t = s; t >= 1; t--) {
      uint32_t ctas = (uint32_t)((227 * 1024) / (fec_win_smem_bytes(g, t, 1) + 1024));
      if (ctas * t * (g->P / 2) > 512) ctas = 512 / (t * (g->P / 2));
      if (ctas * t > best_cbs) { best_cbs = ctas * t; best = t; }
    }
    return best;
  }
  while (s > 1 && fec_win_smem_bytes(g, s, llr8) > 220 * 1024) s--;
  return s;
}

template <int P, bool B8>
__global__ void __launch_bounds__(FEC_WIN_THREADS, B8 ? 4 : 1) fec_tdec_win_kernel(const FecDecodeArgs a)
{
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_cb[FEC_WIN_THREADS / 8], s_n[FEC_WIN_THREADS / 8], s_fresh[FEC_WIN_THREADS / 8];
  __shared__ int s_empty;

  const int tid = threadIdx.x, nthr = blockDim.x;
  const int slot = tid / P, d = tid % P, slots = nthr / P;
  const int K = a.g.K, L = a.g.L, Lp = a.g.Lp, LpE = a.g.LpE, SS = a.g.SS;
  const int nck = (L + S_CH - 1) / S_CH, nw = (L + 31) / 32, kw = (K + 31) / 32;
  const unsigned lane_mask = (P == 32) ? 0xFFFFFFFFu : (((1u << P) - 1u) << ((tid & 31) & ~(P - 1)));

  typedef typename WinCk<B8>::type ck_t;
  constexpr int ESZ = B8 ? 1 : 2;
  size_t off = 0;
  unsigned char *E = smem + (size_t)slot * P * LpE * ESZ;
  off += (((size_t)slots * P * LpE * ESZ) + 15) & ~(size_t)15;
  ck_t *ckpt = (ck_t *)(smem + off);  off += (size_t)nck * nthr * sizeof(ck_t);
  ck_t *chunk = (ck_t *)(smem + off); off += (size_t)(S_CH - 1) * nthr * sizeof(ck_t);
  uint32_t *bitsw = (uint32_t *)(smem + off); off += (size_t)nw * nthr * 4;
  uint32_t *outw = (uint32_t *)(smem + off) + (size_t)slot * kw;

  if (d == 0) { s_cb[slot] = -1; s_fresh[slot] = 0; }
  if (tid == 0) s_empty = 0;
  __syncthreads();

  for (;;) {
    /* ---- refill idle slots from the work queue ---- */
    if (d == 0 && s_cb[slot] < 0 && !s_empty) {
      const unsigned q = atomicAdd(a.queue, 1u);
      const unsigned total = a.list ? *a.list_count : a.ncb;
      if (q < total) { s_cb[slot] = a.list ? (int)a.list[q] : (int)q; s_n[slot] = 0; s_fresh[slot] = 1; }
      else s_empty = 1;
    }
    __syncthreads();
    const int cb = s_cb[slot];
    if (!__syncthreads_or(cb >= 0)) break;
    if (cb >= 0 && s_fresh[slot]) {               /* apriori = 0 before the first half-iteration */
      uint32_t *e32 = (uint32_t *)E;
      for (int i = d; i < P * LpE * ESZ / 4; i += P) e32[i] = 0;
    }
    __syncthreads();
    if (d == 0) s_fresh[slot] = 0;

    int done = 0;
    if (cb >= 0) {
      const FecCbDesc desc = a.cbs[cb];
      const int n = s_n[slot], par = n & 1;
      WinLane w;
      w.S = desc.soft;
      w.Y = desc.soft + (1 + par) * SS;
      w.tpc = P / 2; w.dj = d >> 1; w.dh = d & 1;
      w.T = a.step_tab + par * SS + d * Lp;
      w.E = E;
      w.smask = par ? 0 : -1;
      const int16_t *tail = desc.soft + 3 * SS + (par ? 6 : 0);   /* x at [0,2,4], parity at [1,3,5] */
      const uint32_t *R = a.crcpow[desc.crc_sel];

      int o[8];
      /* ---- beta warm-up over the first 40 symbols of this sub-block (turbodecoder_win.h:321-384) ---- */
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = NEGINF_T(B8);
      win_for_desc(w, 40, 0, [&](int k, int sv, int yv, uint32_t t) {
        win_beta_step<B8>(o, satT<B8>(sv + win_e<B8>(w, t & 0xFFFFu)), yv); win_norm<B8>(k, o);
      });
      /* lane shift: sub-block d starts from the estimate of sub-block d+1 (:347-355) */
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = __shfl_down_sync(lane_mask, o[i], 1, P);
      if (d == P - 1) {   /* known end state through the 3 tail steps, WRAPPING adds (:263-307) */
        int t[8];
        t[0] = 0;
#pragma unroll
        for (int i = 1; i < 8; i++) t[i] = B8 ? 0 : -10000;
        for (int k = 2; k >= 0; k--) {
          const int tx = tail[2 * k], ty = tail[2 * k + 1];
          /* int8: the scalar add of the tail trellis saturates upwards only, z > 127 ? 127 : (int8_t) z (:236-243) */
#define TADD(a, b) (B8 ? (((a) + (b)) > 127 ? 127 : (int)(signed char)((a) + (b))) : (int)(short)((a) + (b)))
          const int txy = TADD(tx, ty);
          int nn[8];
          nn[0] = max(TADD(t[4], txy), t[0]);
          nn[1] = max(t[4], TADD(t[0], txy));
          nn[2] = max(TADD(t[5], ty), TADD(t[1], tx));
          nn[3] = max(TADD(t[5], tx), TADD(t[1], ty));
          nn[4] = max(TADD(t[6], tx), TADD(t[2], ty));
          nn[5] = max(TADD(t[6], ty), TADD(t[2], tx));
          nn[6] = max(t[7], TADD(t[3], txy));
          nn[7] = max(TADD(t[7], txy), t[3]);
#undef TADD
#pragma unroll
          for (int i = 0; i < 8; i++) t[i] = nn[i];
        }
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = t[i] + BIAS_T(B8);
      }
      /* ---- beta main pass, checkpoint every S_CH steps (store is pre-normalisation, :419-432) ---- */
      ckpt[tid] = WinCk<B8>::pack(o);
      {
        int c = 1, knext = L - S_CH;
        win_for_desc(w, L, 0, [&](int k, int sv, int yv, uint32_t t) {
          win_beta_step<B8>(o, satT<B8>(sv + win_e<B8>(w, t & 0xFFFFu)), yv);
          if (k == knext && c < nck) { ckpt[c * nthr + tid] = WinCk<B8>::pack(o); c++; knext -= S_CH; }
          win_norm<B8>(k, o);
        });
      }
      /* ---- alpha warm-up over the last 40 symbols (:459-506) ---- */
      int m[8], nb[8];
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = NEGINF_T(B8);
      win_for_asc(w, L - 40, L, [&](int k, int sv, int yv, uint32_t t) {
        win_alpha_branches<B8>(o, satT<B8>(sv + win_e<B8>(w, t & 0xFFFFu)), yv, m, nb);
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = max(m[i], nb[i]);
        win_norm<B8>(k - (L - 40), o);
      });
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = __shfl_up_sync(lane_mask, o[i], 1, P);
      if (d == 0) {
        o[0] = BIAS_T(B8);
#pragma unroll
        for (int i = 1; i < 8; i++) o[i] = NEGINF_T(B8);
      }
      /* ---- alpha main pass: per chunk, recompute beta from its checkpoint, then alpha + LLR ---- */
      uint32_t crc = 0, wbits = 0, pend = 0;   /* CRC table word of step k is XORed in at step k+1: the lookup is off the critical path */
      for (int c = nck - 1; c >= 0; c--) {
        const int khi = L - c * S_CH, klo = max(0, khi - S_CH);
        const ck_t ck = ckpt[c * nthr + tid];
        {
          int b[8];
          WinCk<B8>::unpack_true(ck, b);
#pragma unroll
          for (int i = 0; i < 8; i++) b[i] += BIAS_T(B8);
          if (c > 0) win_norm<B8>(khi, b);
          win_for_desc(w, khi, klo + 1, [&](int k, int sv, int yv, uint32_t t) {
            win_beta_step<B8>(b, satT<B8>(sv + win_e<B8>(w, t & 0xFFFFu)), yv);
            chunk[(k - klo - 1) * nthr + tid] = WinCk<B8>::pack(b);
            win_norm<B8>(k, b);
          });
        }
        win_for_asc(w, klo, khi, [&](int k, int sv, int yv, uint32_t t) {
          const int ea = t & 0xFFFFu, nat = t >> 16;
          const int e = win_e<B8>(w, ea);
          win_alpha_branches<B8>(o, satT<B8>(sv + e), yv, m, nb);
          int b[8], v0[8], v1[8];
          WinCk<B8>::unpack_true((k == khi - 1) ? ck : chunk[(k - klo) * nthr + tid], b);
#pragma unroll
          for (int i = 0; i < 8; i++) { v0[i] = saddT<B8>(m[i], b[i]); v1[i] = saddT<B8>(nb[i], b[i]); }
          int r = satT<B8>(max8(v1) - max8(v0));
          if (P == 8 || B8) r >>= 1;                 /* divide_output: sse16 and the int8 decoders (turbodecoder_win.h:61,127,159,565) */
          if (B8) {
            /* srslte_vec_sub_bbb (vector_simd.c:165-192): saturating in blocks of 32 array elements, the K mod 32 elements
             * behind the last block by the scalar wrapping loop; the arrays are in sub-block order of the NATURAL bit
             * (element k*P + d), so those are the bits at step L-1 of every lane when K mod 32 == 16 */
            const bool wrap = (K & 31) && ((par ? nat - (nat / L) * L : k) == L - 1);
            win_e_store<B8>(w, ea, wrap ? (int)(signed char)(r - e) : satT<true>(r - e));
          } else
          win_e_store<B8>(w, ea, r - e);             /* next apriori = LLR - (what this decoder was given), wrapping */
          if (a.dbg_llr) a.dbg_llr[(size_t)desc.status * FEC_MAX_K + nat] = (int16_t)r;
          crc ^= pend;
          pend = r > 0 ? __ldg(R + (K - 1 - nat)) : 0u;
          wbits |= (uint32_t)(r > 0) << (k & 31);
          if ((k & 31) == 31 || k == L - 1) { bitsw[(k >> 5) * nthr + tid] = wbits; wbits = 0; }
#pragma unroll
          for (int i = 0; i < 8; i++) o[i] = max(m[i], nb[i]);
          win_norm<B8>(k, o);
        });
      }
      /* ---- CRC over the K hard decisions of this half-iteration: XOR-reduce over the P lanes ---- */
      crc ^= pend;
#pragma unroll
      for (int s = P / 2; s > 0; s >>= 1) crc ^= __shfl_xor_sync(lane_mask, crc, s, P);
      const int n1 = n + 1;
      const int crc_ok = (crc == 0);
      done = (crc_ok && !(desc.flags & FEC_F_NO_EARLY)) || (n1 >= (int)desc.max_its);
      if (done) {
        /* hard bits -> bytes, natural order, MSB first (turbodecoder_win.h:676-716) */
        for (int i = d; i < kw; i += P) outw[i] = 0;
        __syncwarp(lane_mask);
        for (int k = 0; k < L; k++) {
          if ((bitsw[(k >> 5) * nthr + tid] >> (k & 31)) & 1u) {
            const int nat = __ldg(w.T + k) >> 16;
            atomicOr(&outw[nat >> 5], 1u << (8 * ((nat >> 3) & 3) + 7 - (nat & 7)));
          }
        }
        __syncwarp(lane_mask);
        const uint8_t *ob = (const uint8_t *)outw;
        for (int i = d; i < K / 8; i += P) desc.out[i] = ob[i];
        if (d == 0) { a.status[desc.status].its = (uint32_t)n1; a.status[desc.status].crc_ok = (uint32_t)crc_ok; }
      }
      if (d == 0) s_n[slot] = n1;
    }
    __syncthreads();
    if (d == 0 && done) s_cb[slot] = -1;
  }
}

/* =====================================================================================================
 * int8 decoders, two sub-blocks per thread: the same arithmetic as fec_tdec_win_kernel<P, true> with the lane pair (2j, 2j+1)
 * of the device layout in the two 16-bit halves of a register. State metrics are biased by 128 per half; the saturating
 * int8 add of the reference is VIADDMNMX.S16x2.RELU with the clamp 0x00FF00FF (max(min(u + g, 255), 0) per half), so one
 * instruction serves two lanes: half the instruction count of the one-lane kernel, which is bound by exactly that.
 * Shared memory per code block is unchanged (int8 extrinsics, 8 bytes per lane and checkpoint).
 * ===================================================================================================== */
#define P2_CLAMP 0x00FF00FFu
#define P2_BIAS  0x00800080u
__device__ __forceinline__ uint32_t p2_sadd(uint32_t u, uint32_t g) { return __viaddmin_s16x2_relu(u, g, P2_CLAMP); }
__device__ __forceinline__ uint32_t p2_sat8(uint32_t v) { return __vmins2(__vmaxs2(v, 0xFF80FF80u), 0x007F007Fu); }
__device__ __forceinline__ uint32_t p2_addsat8(uint32_t a, uint32_t b) { return __vmins2(__viaddmax_s16x2(a, b, 0xFF80FF80u), 0x007F007Fu); }
__device__ __forceinline__ uint32_t p2_max8(const uint32_t *v)
{
  return __vmaxs2(__vimax3_s16x2(v[0], v[1], v[2]), __vimax3_s16x2(__vimax3_s16x2(v[3], v[4], v[5]), v[6], v[7]));
}
__device__ __forceinline__ void p2_norm(int k, uint32_t *o)
{
  if (k != 0) {
    const uint32_t t = __vsub2(P2_BIAS, p2_max8(o));
#pragma unroll
    for (int i = 0; i < 8; i++) o[i] = p2_sadd(o[i], t);
  }
}
__device__ __forceinline__ void p2_beta_step(uint32_t *o, uint32_t x, uint32_t y)
{
  const uint32_t xy = p2_addsat8(x, y);
  uint32_t n[8];
  n[0] = __vmaxs2(p2_sadd(o[4], xy), o[0]);
  n[1] = __vmaxs2(o[4], p2_sadd(o[0], xy));
  n[2] = __vmaxs2(p2_sadd(o[5], y), p2_sadd(o[1], x));
  n[3] = __vmaxs2(p2_sadd(o[5], x), p2_sadd(o[1], y));
  n[4] = __vmaxs2(p2_sadd(o[6], x), p2_sadd(o[2], y));
  n[5] = __vmaxs2(p2_sadd(o[6], y), p2_sadd(o[2], x));
  n[6] = __vmaxs2(o[7], p2_sadd(o[3], xy));
  n[7] = __vmaxs2(p2_sadd(o[7], xy), o[3]);
#pragma unroll
  for (int i = 0; i < 8; i++) o[i] = n[i];
}
__device__ __forceinline__ void p2_alpha_branches(const uint32_t *o, uint32_t x, uint32_t y, uint32_t *m, uint32_t *n)
{
  const uint32_t xy = p2_addsat8(x, y);
  m[0] = o[0];              m[1] = p2_sadd(o[3], y); m[2] = p2_sadd(o[4], y); m[3] = o[7];
  m[4] = o[1];              m[5] = p2_sadd(o[2], y); m[6] = p2_sadd(o[5], y); m[7] = o[6];
  n[0] = p2_sadd(o[1], xy); n[1] = p2_sadd(o[2], x); n[2] = p2_sadd(o[5], x); n[3] = p2_sadd(o[6], xy);
  n[4] = p2_sadd(o[0], xy); n[5] = p2_sadd(o[3], x); n[6] = p2_sadd(o[4], x); n[7] = p2_sadd(o[7], xy);
}
/* 8 packed state words (two biased bytes each) <-> 16 bytes: word i/2 = [o_i.lo, o_i.hi, o_i+1.lo, o_i+1.hi] */
__device__ __forceinline__ uint4 p2_pack(const uint32_t *o)
{
  return make_uint4(__byte_perm(o[0], o[1], 0x6420), __byte_perm(o[2], o[3], 0x6420), __byte_perm(o[4], o[5], 0x6420), __byte_perm(o[6], o[7], 0x6420));
}
__device__ __forceinline__ void p2_unpack(uint4 p, uint32_t *b)     /* biased */
{
  b[0] = __byte_perm(p.x, 0, 0x4140); b[1] = __byte_perm(p.x, 0, 0x4342);
  b[2] = __byte_perm(p.y, 0, 0x4140); b[3] = __byte_perm(p.y, 0, 0x4342);
  b[4] = __byte_perm(p.z, 0, 0x4140); b[5] = __byte_perm(p.z, 0, 0x4342);
  b[6] = __byte_perm(p.w, 0, 0x4140); b[7] = __byte_perm(p.w, 0, 0x4342);
}

struct P2Lane {
  const uint4 *S, *Y;        /* systematic / parity stream of the code block (16-byte groups of 4 steps per lane pair) */
  const uint4 *T0, *T1;      /* step tables of lanes 2j and 2j+1 */
  const int8_t *E;           /* extrinsic array (shared memory) */
  int smask, tpc, j;
};
struct P2Blk { uint4 s, y, t0, t1; };
__device__ __forceinline__ P2Blk p2_fetch(const P2Lane &w, int kb)
{
  P2Blk b;
  const int g = kb * w.tpc + w.j;
  b.s = w.smask ? __ldg(w.S + g) : make_uint4(0, 0, 0, 0);
  b.y = __ldg(w.Y + g);
  b.t0 = __ldg(w.T0 + kb);
  b.t1 = __ldg(w.T1 + kb);
  return b;
}
/* f(k, s, y, t0, t1): packed true systematic and parity values of the two lanes, their step-table words */
template <typename F>
__device__ __forceinline__ void p2_for_desc(const P2Lane &w, int khi, int klo, F &&f)
{
  if (khi <= klo) return;
  const int b_lo = klo >> 2;
  int kb = (khi - 1) >> 2;
  P2Blk cur = p2_fetch(w, kb), nxt = cur;
  for (; kb >= b_lo; kb--) {
    if (kb > b_lo) nxt = p2_fetch(w, kb - 1);
#pragma unroll
    for (int q = 3; q >= 0; q--) {
      const int k = 4 * kb + q;
      if (k < khi && k >= klo) f(k, win_comp(cur.s, q), win_comp(cur.y, q), win_comp(cur.t0, q), win_comp(cur.t1, q));
    }
    cur = nxt;
  }
}
template <typename F>
__device__ __forceinline__ void p2_for_asc(const P2Lane &w, int klo, int khi, F &&f)
{
  if (khi <= klo) return;
  const int b_hi = (khi - 1) >> 2;
  int kb = klo >> 2;
  P2Blk cur = p2_fetch(w, kb), nxt = cur;
  for (; kb <= b_hi; kb++) {
    if (kb < b_hi) nxt = p2_fetch(w, kb + 1);
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int k = 4 * kb + q;
      if (k < khi && k >= klo) f(k, win_comp(cur.s, q), win_comp(cur.y, q), win_comp(cur.t0, q), win_comp(cur.t1, q));
    }
    cur = nxt;
  }
}
/* x of the two lanes: sat8(systematic + a-priori), the a-priori values gathered from the int8 extrinsic array */
__device__ __forceinline__ uint32_t p2_x(const P2Lane &w, uint32_t s, uint32_t t0, uint32_t t1, uint32_t &e)
{
  e = __byte_perm((uint32_t)(int)w.E[t0 & 0xFFFFu], (uint32_t)(int)w.E[t1 & 0xFFFFu], 0x5410);
  return p2_addsat8(s, e);
}

/* TM = true: beta checkpoints and the chunk buffer live in TENSOR MEMORY (tcgen05.st / ld, 32x32b.x4: 4 columns per entry,
 * private to the thread's lane) instead of shared memory: shared memory per code block halves (extrinsics + decision bits
 * only), so more code blocks are resident per SM, which is what this latency-bound kernel needs. The tcgen05 accesses are
 * warp-collective (.aligned): idle slots run on code block 0's inputs with their results dropped, so that every warp of the
 * CTA stays convergent. */
__device__ __forceinline__ void p2_tm_store(uint32_t taddr, uint4 v)
{
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" :: "r"(taddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void p2_tm_store_fence() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void p2_tm_load_issue(uint32_t taddr, uint4 &v)
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void p2_tm_load_wait(uint4 &v)
{   /* the registers are tied to the wait so that no use can be scheduled above it */
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w) :: "memory");
}
__device__ __forceinline__ uint4 p2_tm_load(uint32_t taddr)
{
  uint4 v;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w) :: "memory");
  return v;
}

template <int P, bool TM>
__global__ void __launch_bounds__(FEC_WIN_THREADS, 4) fec_tdec_win8p_kernel(const FecDecodeArgs a)
{
  constexpr int TPC = P / 2;
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_cb[FEC_WIN_THREADS / 8], s_n[FEC_WIN_THREADS / 8], s_fresh[FEC_WIN_THREADS / 8];
  __shared__ int s_empty;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, nthr = blockDim.x;
  const int slot = tid / TPC, j = tid % TPC, slots = nthr / TPC, nlane = slots * P;
  const int K = a.g.K, L = a.g.L, Lp = a.g.Lp, LpE = a.g.LpE, SS = a.g.SS;
  const int nck = (L + S_CH - 1) / S_CH, nw = (L + 31) / 32, kw = (K + 31) / 32;
  const unsigned lane_mask = (((1u << TPC) - 1u) << ((tid & 31) & ~(TPC - 1)));

  size_t off = 0;
  int8_t *E = reinterpret_cast<int8_t *>(smem) + (size_t)slot * P * LpE;
  off += (((size_t)slots * P * LpE) + 15) & ~(size_t)15;
  uint4 *ckpt = (uint4 *)(smem + off);  if (!TM) off += (size_t)nck * nthr * 16;
  uint4 *chunk = (uint4 *)(smem + off); if (!TM) off += (size_t)(S_CH - 1) * nthr * 16;
  uint32_t *bitsw = (uint32_t *)(smem + off); off += (size_t)nw * nlane * 4;
  /* output assembly words: a finished code block no longer needs its extrinsics, so in the tensor-memory variant they share
   * the (larger) extrinsic array of the slot, which is zeroed again for the next code block */
  uint32_t *outw = TM ? (uint32_t *)E : (uint32_t *)(smem + off) + (size_t)slot * kw;
  const int l0 = slot * P + 2 * j;                /* lane ids of this thread inside the CTA: l0, l0 + 1 */

  if (j == 0) { s_cb[slot] = -1; s_fresh[slot] = 0; }
  if (tid == 0) s_empty = 0;
  uint32_t tck = 0, tch = 0;                     /* tensor-memory addresses of this thread's checkpoint / chunk entries */
  if (TM) {
    if ((tid >> 5) == 0) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "n"(FEC_WIN8_TCOLS) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (TM) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    tck = s_tmem + ((uint32_t)((tid >> 5) & 3) << 21);      /* lane field: 32 * (warp % 4) */
    tch = tck + 4u * (uint32_t)nck;
  }

  for (;;) {
    if (j == 0 && s_cb[slot] < 0 && !s_empty) {
      const unsigned q = atomicAdd(a.queue, 1u);
      const unsigned total = a.list ? *a.list_count : a.ncb;
      if (q < total) { s_cb[slot] = a.list ? (int)a.list[q] : (int)q; s_n[slot] = 0; s_fresh[slot] = 1; }
      else s_empty = 1;
    }
    __syncthreads();
    const int cb = s_cb[slot];
    if (!__syncthreads_or(cb >= 0)) break;
    if (cb >= 0 && s_fresh[slot]) {
      uint32_t *e32 = (uint32_t *)E;
      for (int i = j; i < P * LpE / 4; i += TPC) e32[i] = 0;
    }
    __syncthreads();
    if (j == 0) s_fresh[slot] = 0;

    int done = 0;
    const bool active = cb >= 0;
    if (TM || active) {
      const FecCbDesc desc = a.cbs[active ? cb : 0];
      const int n = s_n[slot], par = n & 1;
      P2Lane w;
      w.S = reinterpret_cast<const uint4 *>(desc.soft);
      w.Y = reinterpret_cast<const uint4 *>(desc.soft + (1 + par) * SS);
      w.T0 = reinterpret_cast<const uint4 *>(a.step_tab + par * SS + (2 * j) * Lp);
      w.T1 = reinterpret_cast<const uint4 *>(a.step_tab + par * SS + (2 * j + 1) * Lp);
      w.E = E; w.smask = par ? 0 : -1; w.tpc = TPC; w.j = j;
      const int16_t *tail = desc.soft + 3 * SS + (par ? 6 : 0);
      const uint32_t *R = a.crcpow[desc.crc_sel];
      const uint32_t *T0s = a.step_tab + par * SS + (2 * j) * Lp, *T1s = T0s + Lp;

      uint32_t o[8], e;
      /* ---- beta warm-up (turbodecoder_win.h:321-384), INF = 0: all states start at 0 ---- */
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = P2_BIAS;
      p2_for_desc(w, 40, 0, [&](int k, uint32_t s, uint32_t y, uint32_t t0, uint32_t t1) {
        p2_beta_step(o, p2_x(w, s, t0, t1, e), y); p2_norm(k, o);
      });
      /* lane shift: lane d starts from the estimate of lane d+1: low half <- own high half, high half <- neighbour's low half */
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = __byte_perm(o[i], __shfl_down_sync(lane_mask, o[i], 1, TPC), 0x5432);
      if (j == TPC - 1) {   /* last lane: known end state through the 3 tail steps (:263-307), scalar add saturating upwards only (:236-243) */
        int t[8];
#pragma unroll
        for (int i = 0; i < 8; i++) t[i] = 0;
        for (int k = 2; k >= 0; k--) {
          const int tx = tail[2 * k], ty = tail[2 * k + 1];
#define TADD(a, b) (((a) + (b)) > 127 ? 127 : (int)(signed char)((a) + (b)))
          const int txy = TADD(tx, ty);
          int nn[8];
          nn[0] = max(TADD(t[4], txy), t[0]);
          nn[1] = max(t[4], TADD(t[0], txy));
          nn[2] = max(TADD(t[5], ty), TADD(t[1], tx));
          nn[3] = max(TADD(t[5], tx), TADD(t[1], ty));
          nn[4] = max(TADD(t[6], tx), TADD(t[2], ty));
          nn[5] = max(TADD(t[6], ty), TADD(t[2], tx));
          nn[6] = max(t[7], TADD(t[3], txy));
          nn[7] = max(TADD(t[7], txy), t[3]);
#undef TADD
#pragma unroll
          for (int i = 0; i < 8; i++) t[i] = nn[i];
        }
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = (o[i] & 0xFFFFu) | ((uint32_t)(t[i] + 128) << 16);
      }
      /* ---- beta main pass, checkpoint every S_CH steps (stored before normalisation, :419-432) ---- */
      if (TM) p2_tm_store(tck, p2_pack(o)); else ckpt[tid] = p2_pack(o);
      {
        int c = 1, knext = L - S_CH;
        p2_for_desc(w, L, 0, [&](int k, uint32_t s, uint32_t y, uint32_t t0, uint32_t t1) {
          p2_beta_step(o, p2_x(w, s, t0, t1, e), y);
          if (k == knext && c < nck) { if (TM) p2_tm_store(tck + 4u * c, p2_pack(o)); else ckpt[c * nthr + tid] = p2_pack(o); c++; knext -= S_CH; }
          p2_norm(k, o);
        });
      }
      if (TM) p2_tm_store_fence();
      /* ---- alpha warm-up over the last 40 symbols (:459-506) ---- */
      uint32_t m[8], nb[8];
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = P2_BIAS;
      p2_for_asc(w, L - 40, L, [&](int k, uint32_t s, uint32_t y, uint32_t t0, uint32_t t1) {
        p2_alpha_branches(o, p2_x(w, s, t0, t1, e), y, m, nb);
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = __vmaxs2(m[i], nb[i]);
        p2_norm(k - (L - 40), o);
      });
      /* lane d starts from the estimate of lane d-1: high half <- own low half, low half <- neighbour's high half; lane 0: known state */
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = __byte_perm(o[i], __shfl_up_sync(lane_mask, o[i], 1, TPC), 0x1076);
      if (j == 0) {
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = (o[i] & 0xFFFF0000u) | 0x0080u;
      }
      /* ---- alpha main pass: per chunk, recompute beta from its checkpoint, then alpha + LLR ---- */
      uint32_t crc = 0, wb0 = 0, wb1 = 0, pend0 = 0, pend1 = 0;   /* CRC table words are XORed in one step late: the lookup is off the critical path */
      for (int c = nck - 1; c >= 0; c--) {
        const int khi = L - c * S_CH, klo = max(0, khi - S_CH);
        const uint4 ck = TM ? p2_tm_load(tck + 4u * c) : ckpt[c * nthr + tid];
        {
          uint32_t b[8];
          p2_unpack(ck, b);
          if (c > 0) p2_norm(khi, b);
          p2_for_desc(w, khi, klo + 1, [&](int k, uint32_t s, uint32_t y, uint32_t t0, uint32_t t1) {
            p2_beta_step(b, p2_x(w, s, t0, t1, e), y);
            if (TM) p2_tm_store(tch + 4u * (k - klo - 1), p2_pack(b)); else chunk[(k - klo - 1) * nthr + tid] = p2_pack(b);
            p2_norm(k, b);
          });
          if (TM) p2_tm_store_fence();
        }
        uint4 pref = ck;                           /* tensor memory: the chunk entry of the next step is fetched one step ahead */
        if (TM && klo < khi - 1) p2_tm_load_issue(tch, pref);
        p2_for_asc(w, klo, khi, [&](int k, uint32_t s, uint32_t y, uint32_t t0, uint32_t t1) {
          const int ea0 = t0 & 0xFFFFu, ea1 = t1 & 0xFFFFu, nat0 = t0 >> 16, nat1 = t1 >> 16;
          p2_alpha_branches(o, p2_x(w, s, t0, t1, e), y, m, nb);
          uint32_t b[8], v0[8], v1[8];
          if (TM) {
            if (k == khi - 1) p2_unpack(ck, b);
            else {
              p2_tm_load_wait(pref);
              p2_unpack(pref, b);
              if (k + 1 < khi - 1) p2_tm_load_issue(tch + 4u * (k + 1 - klo), pref);
            }
          } else p2_unpack((k == khi - 1) ? ck : chunk[(k - klo) * nthr + tid], b);
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const uint32_t bt = __vsub2(b[i], P2_BIAS);          /* true beta */
            v0[i] = p2_sadd(m[i], bt); v1[i] = p2_sadd(nb[i], bt);
          }
          uint32_t r = p2_sat8(__vsub2(p2_max8(v1), p2_max8(v0)));
          r = ((r >> 1) & 0x7FFF7FFFu) | (r & 0x80008000u);       /* divide_output: arithmetic >> 1 per lane (:127,159,565) */
          const uint32_t dif = __vsub2(r, e);
          uint32_t ne = p2_sat8(dif);
          if (K & 31) {
            /* srslte_vec_sub_bbb: the K mod 32 array elements behind the last block of 32 wrap (see fec_tdec_win_kernel) */
            const bool w0 = (par ? nat0 - (nat0 / L) * L : k) == L - 1, w1 = (par ? nat1 - (nat1 / L) * L : k) == L - 1;
            const uint32_t wr = __vsub2((dif & 0x00FF00FFu) ^ 0x00800080u, 0x00800080u);
            const uint32_t msk = (w0 ? 0x0000FFFFu : 0u) | (w1 ? 0xFFFF0000u : 0u);
            ne = (wr & msk) | (ne & ~msk);
          }
          E[ea0] = (int8_t)(ne & 0xFFu); E[ea1] = (int8_t)((ne >> 16) & 0xFFu);
          const int r0 = (short)(r & 0xFFFFu), r1 = (short)(r >> 16);
          if (a.dbg_llr && active) { a.dbg_llr[(size_t)desc.status * FEC_MAX_K + nat0] = (int16_t)r0; a.dbg_llr[(size_t)desc.status * FEC_MAX_K + nat1] = (int16_t)r1; }
          crc ^= pend0 ^ pend1;
          pend0 = r0 > 0 ? __ldg(R + (K - 1 - nat0)) : 0u; pend1 = r1 > 0 ? __ldg(R + (K - 1 - nat1)) : 0u;
          wb0 |= (uint32_t)(r0 > 0) << (k & 31); wb1 |= (uint32_t)(r1 > 0) << (k & 31);
          if ((k & 31) == 31 || k == L - 1) { bitsw[(k >> 5) * nlane + l0] = wb0; bitsw[(k >> 5) * nlane + l0 + 1] = wb1; wb0 = wb1 = 0; }
#pragma unroll
          for (int i = 0; i < 8; i++) o[i] = __vmaxs2(m[i], nb[i]);
          p2_norm(k, o);
        });
      }
      crc ^= pend0 ^ pend1;
#pragma unroll
      for (int s = TPC / 2; s > 0; s >>= 1) crc ^= __shfl_xor_sync(lane_mask, crc, s, TPC);
      const int n1 = n + 1;
      const int crc_ok = (crc == 0);
      done = active && ((crc_ok && !(desc.flags & FEC_F_NO_EARLY)) || (n1 >= (int)desc.max_its));
      if (done) {
        for (int i = j; i < kw; i += TPC) outw[i] = 0;
        __syncwarp(lane_mask);
        /* natural positions from the step tables, 4 steps per load (the rows are Lp >= L words, 16-byte aligned) and 4 loads in
         * flight: one dependent 4-byte load per step made this stage as long as a third of a half-iteration */
        for (int h = 0; h < 2; h++) {
          const uint4 *T4 = reinterpret_cast<const uint4 *>(h ? T1s : T0s);
          const int nq = (L + 3) >> 2;
          for (int q0 = 0; q0 < nq; q0 += 4) {
            uint4 t[4];
#pragma unroll
            for (int u = 0; u < 4; u++) t[u] = (q0 + u < nq) ? __ldg(T4 + q0 + u) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int u = 0; u < 4; u++) {
#pragma unroll
              for (int c = 0; c < 4; c++) {
                const int k = 4 * (q0 + u) + c;
                if (k < L && ((bitsw[(k >> 5) * nlane + l0 + h] >> (k & 31)) & 1u)) {
                  const int nat = (int)(win_comp(t[u], c) >> 16);
                  atomicOr(&outw[nat >> 5], 1u << (8 * ((nat >> 3) & 3) + 7 - (nat & 7)));
                }
              }
            }
          }
        }
        __syncwarp(lane_mask);
        const uint8_t *ob = (const uint8_t *)outw;
        for (int i = j; i < K / 8; i += TPC) desc.out[i] = ob[i];
        if (j == 0) { a.status[desc.status].its = (uint32_t)n1; a.status[desc.status].crc_ok = (uint32_t)crc_ok; }
      }
      if (j == 0) s_n[slot] = n1;
    }
    __syncthreads();
    if (j == 0 && done) s_cb[slot] = -1;
  }
  if (TM) {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if ((tid >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(s_tmem), "n"(FEC_WIN8_TCOLS) : "memory");
  }
}

/* =====================================================================================================
 * Non-windowed decoder for K <= 400 (AUTO class 0): one thread per code block, wrapping int16
 * arithmetic (turbodecoder_sse.c). alpha is stored for every step, beta runs backward with the LLR.
 * ===================================================================================================== */
#define W16(v) ((int)(short)(v))

__global__ void fec_tdec_small_kernel(const FecDecodeArgs a)
{
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_empty;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int K = a.g.K;
  uint4 *A = (uint4 *)smem;                                        /* [(K+1)][nthr] alpha store */
  int16_t *E = (int16_t *)(smem + (size_t)(K + 1) * nthr * 16);    /* [K][nthr] */
  uint8_t *ob = (uint8_t *)(E + (size_t)K * nthr);                 /* [K/8][nthr] */
  const uint32_t *T0 = a.step_tab;
  int cb = -1, n = 0;
  if (tid == 0) s_empty = 0;
  __syncthreads();

  for (;;) {
    if (cb < 0 && !s_empty) {
      const unsigned q = atomicAdd(a.queue, 1u);
      const unsigned total = a.list ? *a.list_count : a.ncb;
      if (q < total) { cb = a.list ? (int)a.list[q] : (int)q; n = 0; for (int i = 0; i < K; i++) E[i * nthr + tid] = 0; }
      else s_empty = 1;
    }
    if (!__syncthreads_or(cb >= 0)) break;
    if (cb >= 0) {
      const FecCbDesc desc = a.cbs[cb];
      const int par = n & 1, smask = par ? 0 : -1;
      const int16_t *S = desc.soft, *Y = desc.soft + (1 + par) * a.g.SS;
      const uint32_t *T = T0 + par * a.g.SS;
      const int16_t *tail = desc.soft + 3 * a.g.SS + (par ? 6 : 0);
      const uint32_t *R = a.crcpow[desc.crc_sel];
      for (int i = 0; i < K / 8; i++) ob[i * nthr + tid] = 0;

      /* forward recursion (turbodecoder_sse.c:211-297) */
      int al[8];
      al[0] = 0;
#pragma unroll
      for (int i = 1; i < 8; i++) al[i] = -10000;
      {
        int tmp[8];
#pragma unroll
        for (int i = 0; i < 8; i++) tmp[i] = al[i] + BIAS;
        A[tid] = pack8<false>(tmp);
      }
      for (int k = 0; k < K; k++) {
        const uint32_t t = __ldg(T + k);
        const int x = W16(((int)__ldg(S + k) & smask) + E[(t & 0xFFFFu) * nthr + tid]);
        const int p = (int)__ldg(Y + k);
        const int g1 = W16(x + p) >> 1, g0 = W16(x - p) >> 1;
        int nx[8];
        /* alpha'[j] = max(alpha[SP[j]] + g, alpha[SN[j]] - g), g = g1 for states {0,1,6,7}, g0 for {2,3,4,5} */
        nx[0] = max(W16(al[1] + g1), W16(al[0] - g1));
        nx[1] = max(W16(al[2] + g0), W16(al[3] - g0));
        nx[2] = max(W16(al[5] + g0), W16(al[4] - g0));
        nx[3] = max(W16(al[6] + g1), W16(al[7] - g1));
        nx[4] = max(W16(al[0] + g1), W16(al[1] - g1));
        nx[5] = max(W16(al[3] + g0), W16(al[2] - g0));
        nx[6] = max(W16(al[4] + g0), W16(al[5] - g0));
        nx[7] = max(W16(al[7] + g1), W16(al[6] - g1));
        int tmp[8];
#pragma unroll
        for (int i = 0; i < 8; i++) { al[i] = nx[i]; tmp[i] = nx[i] + BIAS; }
        A[(k + 1) * nthr + tid] = pack8<false>(tmp);                       /* stored before normalisation */
        if ((k & 3) == 3) {
          const int z = al[0];
#pragma unroll
          for (int i = 0; i < 8; i++) al[i] = W16(al[i] - z);
        }
      }
      /* backward recursion + LLR (turbodecoder_sse.c:105-206) */
      int be[8];
      be[0] = 0;
#pragma unroll
      for (int i = 1; i < 8; i++) be[i] = -10000;
      uint32_t crc = 0;
      int cnt = 0;
      for (int k = K + 2; k >= 0; k--) {
        int g0, g1, ea = 0, nat = 0, e = 0;
        if (k >= K) {
          const int tx = tail[2 * (k - K)], tp = tail[2 * (k - K) + 1];
          g0 = W16((tx - tp) / 2); g1 = W16((tx + tp) / 2);
        } else {
          const uint32_t t = __ldg(T + k);
          ea = t & 0xFFFFu; nat = t >> 16;
          e = E[ea * nthr + tid];
          const int x = W16(((int)__ldg(S + k) & smask) + e);
          const int p = (int)__ldg(Y + k);
          g1 = W16(x + p) >> 1; g0 = W16(x - p) >> 1;
        }
        /* sp[j] = (beta+g)[BP[j]], sn[j] = (beta-g)[BN[j]], g = g1 for states {0,3,4,7}, g0 for {1,2,5,6} */
        int sp[8], sn[8];
        sp[0] = W16(be[4] + g1); sn[0] = W16(be[0] - g1);
        sp[1] = W16(be[0] + g1); sn[1] = W16(be[4] - g1);
        sp[2] = W16(be[1] + g0); sn[2] = W16(be[5] - g0);
        sp[3] = W16(be[5] + g0); sn[3] = W16(be[1] - g0);
        sp[4] = W16(be[6] + g0); sn[4] = W16(be[2] - g0);
        sp[5] = W16(be[2] + g0); sn[5] = W16(be[6] - g0);
        sp[6] = W16(be[3] + g1); sn[6] = W16(be[7] - g1);
        sp[7] = W16(be[7] + g1); sn[7] = W16(be[3] - g1);
#pragma unroll
        for (int i = 0; i < 8; i++) be[i] = max(sp[i], sn[i]);
        if (k < K) {
          int av[8];
          unpack8_true<false>(A[k * nthr + tid], av);
          int mp = W16(sp[0] + av[0]), mn = W16(sn[0] + av[0]);
#pragma unroll
          for (int i = 1; i < 8; i++) { mp = max(mp, W16(sp[i] + av[i])); mn = max(mn, W16(sn[i] + av[i])); }
          const int r = W16(mp - mn);                              /* hMax() returns 0x7FFF-max: :97-102,178 */
          E[ea * nthr + tid] = (int16_t)(r - e);
          if (a.dbg_llr) a.dbg_llr[(size_t)desc.status * FEC_MAX_K + nat] = (int16_t)r;
          if (r > 0) { crc ^= __ldg(R + (K - 1 - nat)); ob[(nat >> 3) * nthr + tid] |= (uint8_t)(0x80u >> (nat & 7)); }
          if (++cnt == 4) {
            cnt = 0;
            const int z = be[0];
#pragma unroll
            for (int i = 0; i < 8; i++) be[i] = W16(be[i] - z);
          }
        }
      }
      n++;
      const int crc_ok = (crc == 0);
      if ((crc_ok && !(desc.flags & FEC_F_NO_EARLY)) || n >= (int)desc.max_its) {
        for (int i = 0; i < K / 8; i++) desc.out[i] = ob[i * nthr + tid];
        a.status[desc.status].its = (uint32_t)n;
        a.status[desc.status].crc_ok = (uint32_t)crc_ok;
        cb = -1;
      }
    }
  }
}

static size_t fec_small_smem_per_cb(const fec_geom_t *g) { return (size_t)(g->K + 1) * 16 + (size_t)g->K * 2 + g->K / 8; }

cudaError_t fec_launch_decode(const FecDecodeArgs &a, int sm_count, cudaStream_t st, int stages)
{
  if (a.ncb == 0) return cudaSuccess;
  if (a.g.P == 0) {
    int nthr = (int)((200 * 1024) / fec_small_smem_per_cb(&a.g));
    if (nthr > 32) nthr = 32;
    size_t smem = fec_small_smem_per_cb(&a.g) * nthr + 16;
    int grid = (int)((a.ncb + nthr - 1) / nthr);
    if (grid > sm_count * 4) grid = sm_count * 4;
    fec_tdec_small_kernel<<<grid, nthr, smem, st>>>(a);
    return cudaGetLastError();
  }
  if (a.exact_list && !a.dbg_llr) {
    /* fast path first; whatever it hands over (range condition) is decoded by the exact kernel */
    cudaError_t e = fec_launch_decode_packed(a, sm_count, st, stages);
    if (e != cudaSuccess || !(stages & FEC_ST_C)) return e;
    FecDecodeArgs b = a;
    b.list = a.exact_list; b.list_count = a.exact_count; b.queue = a.queue + 1;
    const uint32_t slots = b.slots;
    const size_t smem = fec_win_smem_bytes(&b.g, slots, 0);
    const int grid = sm_count;                     /* persistent; CTAs find the list empty and exit in the common case */
    if (b.g.P == 16) fec_tdec_win_kernel<16, false><<<grid, slots * 16, smem, st>>>(b);
    else             fec_tdec_win_kernel<8, false><<<grid, slots * 8, smem, st>>>(b);
    return cudaGetLastError();
  }
  const uint32_t slots = a.slots;
  const size_t smem = fec_win_smem_bytes(&a.g, slots, a.llr8 ? 1 + FEC_WIN8_TMEM : 0);
  int grid = (int)((a.ncb + slots - 1) / slots);
  int per_sm = (int)(((a.llr8 ? 227 : 220) * 1024) / (smem + 1024));
  if (a.llr8 && FEC_WIN8_TMEM && per_sm > 512 / FEC_WIN8_TCOLS) per_sm = 512 / FEC_WIN8_TCOLS;
  const int cap = sm_count * (per_sm < 1 ? 1 : per_sm);
  if (grid > cap) grid = cap;
  if (a.llr8 && a.g.P == 32)      fec_tdec_win8p_kernel<32, FEC_WIN8_TMEM != 0><<<grid, slots * 16, smem, st>>>(a);
  else if (a.llr8 && a.g.P == 16) fec_tdec_win8p_kernel<16, FEC_WIN8_TMEM != 0><<<grid, slots * 8, smem, st>>>(a);
  else if (a.g.P == 16)           fec_tdec_win_kernel<16, false><<<grid, slots * 16, smem, st>>>(a);
  else                            fec_tdec_win_kernel<8, false><<<grid, slots * 8, smem, st>>>(a);
  return cudaGetLastError();
}

/* =====================================================================================================
 * Rate de-matching + HARQ combining as a GATHER over soft-buffer positions (rm_turbo.c:378-430 computes
 * the same sums as a scatter: soft[deinter[i % N]] += e[i]).
 * ===================================================================================================== */
__device__ __forceinline__ void fec_softmax_commit(uint32_t *dst, int amax)
{
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) amax = max(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, s));
  if ((threadIdx.x & 31) == 0 && amax > 0) atomicMax(dst, (uint32_t)amax);
}

/* One CTA per code block. The received LLRs are staged in shared memory in CIRCULAR-BUFFER order (entry p of the
 * NULL-free buffer <- e[i], p = (i + k0 offset of the rv) mod N; i, i+N, i+2N... are the repeats of one entry and are
 * summed in the reference's order; entries never transmitted are 0): a coalesced read and a conflict-free write.
 * The slot is then produced 16 bytes per thread: the per-K rank table (device offset -> p) turns the sub-block
 * de-interleaving into a gather from shared memory, added to the old content for a retransmission, and written
 * coalesced together with max |entry|. HBM traffic = the algorithmic minimum (2E bytes in, the slot out / in+out). */
#ifndef RM_UNROLL
#define RM_UNROLL 2
#endif
constexpr int kRmUnroll = RM_UNROLL;
#ifndef RM_STREAM_ST
#define RM_STREAM_ST 1
#endif
/* the slot leaves with a streaming store (evict-first in L2): it is read again only when its code block's turn comes in the
 * decoder, long after 1 GB of other slots has gone by — until then it should not push the running decoder's rows out of L2 */
__device__ __forceinline__ void rm_store_slot(uint4 *p, uint4 v)
{
#if RM_STREAM_ST
  asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
#else
  *p = v;
#endif
}
__device__ __forceinline__ uint32_t rm_lds_u16(uint32_t addr)
{
  uint32_t r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}

/* Phase 1 (staging) by the bulk-copy engine: the tile is e[] IN THE ORDER RECEIVED (entry i = e[i], entries E..N zero; the per-rv
 * table d.rank maps every soft position to 2i), so the whole slice arrives with ONE cp.async.bulk (UBLKCP) signalled on an
 * mbarrier while the threads zero the tail: no LDG / STS instruction is spent on staging. Bulk copies want 16-byte aligned
 * addresses and sizes: the tile starts (e & 15) bytes into shared memory so that source and destination are congruent, the
 * 16-byte aligned interior goes by bulk copy and the ragged head / tail (< 8 LLRs each) by plain loads. */
__global__ void __launch_bounds__(FEC_RM_THREADS) fec_rm_rx_kernel(const FecRmDesc *descs, uint32_t *soft_max)
{
  extern __shared__ __align__(16) unsigned char rm_smem[];
  __shared__ __align__(8) unsigned long long rm_mbar;
  const FecRmDesc d = descs[blockIdx.x];
  const int tid = threadIdx.x;
  const size_t ea = reinterpret_cast<size_t>(d.e);
  const uint32_t delta = (uint32_t)(ea & 15);
  int16_t *rm_tile = reinterpret_cast<int16_t *>(rm_smem + delta);
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(rm_smem), mb = (uint32_t)__cvta_generic_to_shared(&rm_mbar);
  if (d.E <= d.N) {
    /* the common case (no repetition) */
    const size_t a0 = (ea + 15) & ~(size_t)15, a1 = (ea + 2 * (size_t)d.E) & ~(size_t)15;
    const bool bulk = a1 > a0;
    const uint32_t nb = bulk ? (uint32_t)(a1 - a0) : 0u;
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mb));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      if (bulk) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"(nb) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(sbase + delta + (uint32_t)(a0 - ea)), "l"(a0), "r"(nb), "r"(mb) : "memory");
      }
    }
    /* ragged head and tail of the slice */
    const uint32_t h = bulk ? (uint32_t)(a0 - ea) >> 1 : d.E, t0 = bulk ? (uint32_t)(a1 - ea) >> 1 : d.E;
    if ((uint32_t)tid < h) rm_tile[tid] = __ldg(d.e + tid);
    if (t0 + tid < d.E) rm_tile[t0 + tid] = __ldg(d.e + t0 + tid);
    /* entries never transmitted (E .. N; entry N is what the table's "no coded bit here" offsets point at) all read the ONE
     * zero behind the slice: the gather clamps its offsets to it, so the tile is E + 1 entries, not N + 1 (at the bench shape
     * 14 KB instead of 35 KB: more CTAs per SM, and no zero-fill pass) */
    if (tid == 0) rm_tile[d.E] = 0;
    __syncthreads();                                   /* the barrier object is initialised for everybody */
    if (bulk) {
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(mb) : "memory");
    }
  } else {
    for (uint32_t i = tid; i < d.N; i += FEC_RM_THREADS) {
      int acc = 0;
      for (uint32_t j = i; j < d.E; j += d.N) acc += (int)__ldg(d.e + j);    /* repeats accumulate when E > 3K+12 */
      rm_tile[i] = (int16_t)acc;
    }
    if (tid == 0) rm_tile[d.N] = 0;
    __syncthreads();
  }
  /* the rank table holds byte offsets into the tile: per soft value one 16-bit shared-memory load, per pair one PRMT, one
   * packed wrapping add (rm_turbo.c:411-420 adds int16 with wrap-around) when combining, and a packed min / max for max|soft| */
  const uint4 *rank4 = reinterpret_cast<const uint4 *>(d.rank);
  uint4 *soft4 = reinterpret_cast<uint4 *>(d.soft);
  const uint32_t tb = (uint32_t)__cvta_generic_to_shared(rm_tile);
  const uint32_t lim = 2u * min(d.E, d.N) * 0x00010001u;        /* byte offset of the zero entry, in both halves */
  uint32_t mx = 0, mn = 0;
  if (d.new_data) {
#pragma unroll kRmUnroll
    for (int v = tid; v < (int)d.n_out / 8; v += FEC_RM_THREADS) {
      const uint4 rk = __ldg(rank4 + v);
      const uint32_t rw[4] = {__vminu2(rk.x, lim), __vminu2(rk.y, lim), __vminu2(rk.z, lim), __vminu2(rk.w, lim)};
      uint32_t ow[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        ow[k] = __byte_perm(rm_lds_u16(tb + (rw[k] & 0xFFFFu)), rm_lds_u16(tb + (rw[k] >> 16)), 0x5410);
        mx = __vmaxs2(mx, ow[k]); mn = __vmins2(mn, ow[k]);
      }
      rm_store_slot(soft4 + v, make_uint4(ow[0], ow[1], ow[2], ow[3]));
    }
  } else {
#pragma unroll 2
    for (int v = tid; v < (int)d.n_out / 8; v += FEC_RM_THREADS) {
      const uint4 rk = __ldg(rank4 + v);
      const uint4 old = soft4[v];
      const uint32_t rw[4] = {__vminu2(rk.x, lim), __vminu2(rk.y, lim), __vminu2(rk.z, lim), __vminu2(rk.w, lim)};
      uint32_t ow[4] = {old.x, old.y, old.z, old.w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        ow[k] = __vadd2(ow[k], __byte_perm(rm_lds_u16(tb + (rw[k] & 0xFFFFu)), rm_lds_u16(tb + (rw[k] >> 16)), 0x5410));
        mx = __vmaxs2(mx, ow[k]); mn = __vmins2(mn, ow[k]);
      }
      rm_store_slot(soft4 + v, make_uint4(ow[0], ow[1], ow[2], ow[3]));
    }
  }
  const int hi = max((int)(short)(mx & 0xFFFFu), (int)(short)(mx >> 16)), lo = min((int)(short)(mn & 0xFFFFu), (int)(short)(mn >> 16));
  fec_softmax_commit(soft_max + d.slot, max(hi, -lo));
}

/* 8-bit LLR path (srslte_rm_turbo_rx_lut_8bit, rm_turbo.c:432-469,543-631): int8 LLRs in, int8 wrapping accumulation.
 * The soft buffer keeps its int16 slots; an entry holds the int8 value the reference's buffer would hold, sign-extended
 * (wrapping int8 addition is addition mod 256, so sums may be formed wider and reduced once). Halves the ingest, not the
 * soft-buffer traffic; same gather structure as above without the tuning. */
__device__ __forceinline__ uint32_t rm_sext8x2(uint32_t v)
{
  return __vsub2((v & 0x00FF00FFu) ^ 0x00800080u, 0x00800080u);
}

__device__ __forceinline__ uint32_t rm_lds_s8(uint32_t addr)
{
  uint32_t r;
  asm volatile("ld.shared.s8 %0, [%1];" : "=r"(r) : "r"(addr));
  return r;
}

/* Same structure as fec_rm_rx_kernel on an int8 tile: the slice e[0..E) in the order received behind one bulk copy, one zero
 * byte behind it, the per-(K, rv) table holding byte offsets into it (ktab_rank halves them). Repeats (E > N) are summed
 * mod 256 into the tile, which is all the wrapping int8 accumulation keeps of them. */
__global__ void __launch_bounds__(FEC_RM_THREADS) fec_rm_rx8_kernel(const FecRmDesc *descs, uint32_t *soft_max)
{
  extern __shared__ __align__(16) unsigned char rm_smem[];
  __shared__ __align__(8) unsigned long long rm_mbar;
  const FecRmDesc d = descs[blockIdx.x];
  const int8_t *e8 = reinterpret_cast<const int8_t *>(d.e);
  const int tid = threadIdx.x;
  const size_t ea = reinterpret_cast<size_t>(e8);
  const uint32_t delta = (uint32_t)(ea & 15);
  int8_t *tile = reinterpret_cast<int8_t *>(rm_smem + delta);
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(rm_smem), mb = (uint32_t)__cvta_generic_to_shared(&rm_mbar);
  if (d.E <= d.N) {
    const size_t a0 = (ea + 15) & ~(size_t)15, a1 = (ea + (size_t)d.E) & ~(size_t)15;
    const bool bulk = a1 > a0;
    const uint32_t nb = bulk ? (uint32_t)(a1 - a0) : 0u;
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mb));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      if (bulk) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"(nb) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(sbase + delta + (uint32_t)(a0 - ea)), "l"(a0), "r"(nb), "r"(mb) : "memory");
      }
      tile[d.E] = 0;
    }
    const uint32_t h = bulk ? (uint32_t)(a0 - ea) : d.E, t0 = bulk ? (uint32_t)(a1 - ea) : d.E;      /* ragged head / tail: < 16 LLRs each */
    for (uint32_t i = tid; i < h; i += FEC_RM_THREADS) tile[i] = __ldg(e8 + i);
    for (uint32_t i = t0 + tid; i < d.E; i += FEC_RM_THREADS) tile[i] = __ldg(e8 + i);
    __syncthreads();
    if (bulk) {
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(mb) : "memory");
    }
  } else {
    for (uint32_t i = tid; i < d.N; i += FEC_RM_THREADS) {
      int acc = 0;
      for (uint32_t j = i; j < d.E; j += d.N) acc += (int)__ldg(e8 + j);    /* repeats accumulate when E > 3K+12 */
      tile[i] = (int8_t)acc;
    }
    if (tid == 0) tile[d.N] = 0;
    __syncthreads();
  }
  const uint4 *rank4 = reinterpret_cast<const uint4 *>(d.rank);
  uint4 *soft4 = reinterpret_cast<uint4 *>(d.soft);
  const uint32_t tb = (uint32_t)__cvta_generic_to_shared(tile);
  const uint32_t lim = min(d.E, d.N) * 0x00010001u;           /* offset of the zero byte, in both halves */
  uint32_t mx = 0, mn = 0;
#pragma unroll 2
  for (int v = tid; v < (int)d.n_out / 8; v += FEC_RM_THREADS) {
    const uint4 rk = __ldg(rank4 + v);
    uint4 old = make_uint4(0, 0, 0, 0);
    if (!d.new_data) old = soft4[v];
    const uint32_t rw[4] = {__vminu2(rk.x, lim), __vminu2(rk.y, lim), __vminu2(rk.z, lim), __vminu2(rk.w, lim)};
    uint32_t ow[4] = {old.x, old.y, old.z, old.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t in = __byte_perm(rm_lds_s8(tb + (rw[k] & 0xFFFFu)), rm_lds_s8(tb + (rw[k] >> 16)), 0x5410);
      ow[k] = d.new_data ? in : rm_sext8x2(__vadd2(ow[k], in));
      mx = __vmaxs2(mx, ow[k]); mn = __vmins2(mn, ow[k]);
    }
    rm_store_slot(soft4 + v, make_uint4(ow[0], ow[1], ow[2], ow[3]));
  }
  const int hi = max((int)(short)(mx & 0xFFFFu), (int)(short)(mx >> 16)), lo = min((int)(short)(mn & 0xFFFFu), (int)(short)(mn >> 16));
  fec_softmax_commit(soft_max + d.slot, max(hi, -lo));
}

/* max_tile: the largest min(E, N) + 1 of the launch (0: not known, the slot size bounds it) */
cudaError_t fec_launch_rm(const FecRmDesc *descs, uint32_t *soft_max, uint32_t n, uint32_t max_n_out, cudaStream_t st, int llr8, uint32_t max_tile)
{
  if (n == 0) return cudaSuccess;
  if (max_tile == 0 || max_tile > max_n_out) max_tile = max_n_out;
  if (llr8) {
    fec_rm_rx8_kernel<<<n, FEC_RM_THREADS, (size_t)max_tile + 32, st>>>(descs, soft_max);       /* one byte per LLR */
    return cudaGetLastError();
  }
  fec_rm_rx_kernel<<<n, FEC_RM_THREADS, (size_t)max_tile * 2 + 32, st>>>(descs, soft_max);   /* the tile + up to 14 bytes of shift */
  return cudaGetLastError();
}

/* natural [s p p'] x K + 12 tail  ->  device layout (srslte_tdec_* level input, force_not_sb path) */
__global__ void __launch_bounds__(256) fec_extract_kernel(const FecExtractDesc *descs, uint32_t *soft_max)
{
  const FecExtractDesc d = descs[blockIdx.y];
  const int n_out = 3 * d.g.SS + 16;
  int amax = 0;
  for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < n_out; o += gridDim.x * blockDim.x) {
    int16_t v = 0;
    if (o >= 3 * d.g.SS) { if (o < 3 * d.g.SS + 12) v = d.in[3 * d.g.K + (o - 3 * d.g.SS)]; }
    else {
      const int s = o / d.g.SS, r = o % d.g.SS;
      int lane, k;
      if (d.g.P) { const int wd = r >> 1, tpc = d.g.P / 2; lane = 2 * ((wd >> 2) % tpc) + (r & 1); k = ((wd >> 2) / tpc) * 4 + (wd & 3); }
      else { lane = 0; k = r; }
      if (k < d.g.L) v = d.in[3 * (lane * d.g.L + k) + s];
    }
    d.soft[o] = v;
    amax = max(amax, v < 0 ? -(int)v : (int)v);
  }
  fec_softmax_commit(soft_max + d.slot, amax);
}

cudaError_t fec_launch_extract(const FecExtractDesc *descs, uint32_t *soft_max, uint32_t n, cudaStream_t st)
{
  if (n == 0) return cudaSuccess;
  fec_extract_kernel<<<dim3(8, n), 256, 0, st>>>(descs, soft_max);
  return cudaGetLastError();
}

/* (stream s, natural position x) or tail entry t >= 3K  ->  offsets in the device and the reference soft-buffer layouts */
__device__ __forceinline__ void fec_soft_pair(const fec_geom_t &g, int t, int sb, int &od, int &orf)
{
  if (t >= 3 * g.K) { od = 3 * g.SS + (t - 3 * g.K); orf = (sb && g.P) ? 3 * (g.K + 32) + (t - 3 * g.K) : t; }
  else {
    const int s = t % 3, x = t / 3, dl = x / g.L, kk = x % g.L;
    od = g.P ? s * g.SS + 2 * ((((kk >> 2) * (g.P / 2) + (dl >> 1)) << 2) + (kk & 3)) + (dl & 1) : s * g.SS + kk;
    orf = (sb && g.P) ? s * (g.K + 32) + (x % g.L) * g.P + x / g.L : t;
  }
}

/* device layout -> reference soft-buffer layout (18600 int16), for parity checks and srslte_rm_turbo_rx_lut.
 * keep != 0: entries of `ref` that no code-block position maps to are left as they are (the reference never
 * touches them either), else they are zeroed. Single CTA. */
__global__ void fec_soft_to_ref_kernel(const int16_t *dev, int16_t *ref, fec_geom_t g, int sb, int keep)
{
  if (!keep) {
    for (int i = threadIdx.x; i < 18600; i += blockDim.x) ref[i] = 0;
    __syncthreads();
  }
  for (int t = threadIdx.x; t < 3 * g.K + 12; t += blockDim.x) {
    int od, orf;
    fec_soft_pair(g, t, sb, od, orf);
    ref[orf] = dev[od];
  }
}

/* reference soft-buffer layout -> device layout (a caller-owned host soft buffer entering srslte_rm_turbo_rx_lut) */
__global__ void fec_ref_to_soft_kernel(const int16_t *ref, int16_t *dev, fec_geom_t g, int sb)
{
  for (int i = threadIdx.x; i < 3 * g.SS + 16; i += blockDim.x) dev[i] = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < 3 * g.K + 12; t += blockDim.x) {
    int od, orf;
    fec_soft_pair(g, t, sb, od, orf);
    dev[od] = ref[orf];
  }
}

cudaError_t fec_launch_soft_to_ref(const int16_t *soft_dev, int16_t *ref_dev, fec_geom_t g, int sb_layout, int keep, cudaStream_t st)
{
  fec_soft_to_ref_kernel<<<1, 1024, 0, st>>>(soft_dev, ref_dev, g, sb_layout, keep);
  return cudaGetLastError();
}

cudaError_t fec_launch_ref_to_soft(const int16_t *ref_dev, int16_t *soft_dev, fec_geom_t g, int sb_layout, cudaStream_t st)
{
  fec_ref_to_soft_kernel<<<1, 1024, 0, st>>>(ref_dev, soft_dev, g, sb_layout);
  return cudaGetLastError();
}

/* =====================================================================================================
 * Transport-block assembly: concatenate code-block payloads exactly as decode_tb_cb's sequential writes
 * do (sch.c:365,396-401: CB i writes K/8 bytes at i*rlen/8, later CBs overwrite its CRC bytes), refresh
 * the decoded-CB cache (sch.c:410-417) and check CRC24A (sch.c:478-485). One CTA per transport block.
 * ===================================================================================================== */
__global__ void __launch_bounds__(256) fec_tb_assemble_kernel(const FecTbDesc *descs, const FecCbStatus *status,
                                                              const uint32_t *r24a)
{
  const FecTbDesc &d = descs[blockIdx.x];
  __shared__ uint32_t s_red[8], s_tab[256];
  __shared__ int s_ok;
  const int tid = threadIdx.x;
  {
    uint32_t c = (uint32_t)tid << 16;              /* (tid * x^24) mod g, g = 0x1864CFB */
#pragma unroll
    for (int i = 0; i < 8; i++) c = ((c << 1) ^ ((c & 0x800000u) ? 0x864CFBu : 0u)) & 0xFFFFFFu;
    s_tab[tid] = c;
  }
  if (tid == 0) {
    int ok = 1;
    for (uint32_t c = 0; c < d.C; c++) if (d.st_idx[c] >= 0 && !status[d.st_idx[c]].crc_ok) ok = 0;
    s_ok = ok && d.all_ok;
  }
  if (tid < 3) d.data[d.tbs / 8 + tid] = 0;
  __syncthreads();
  for (uint32_t c = 0; c < d.C; c++) {
    const int fresh = d.st_idx[c] >= 0;
    const uint32_t nby = fresh ? d.K[c] / 8 : d.rlen[c] / 8, base = c * d.rlen[c] / 8;
    for (uint32_t i = tid; i < nby; i += blockDim.x) d.data[base + i] = d.cb_bytes[c][i];
    /* "If TB CRC failed, save correct CB for next retransmission" (sch.c:409-417): the cache is refreshed only while some
     * code block of the transport block is still wrong — a (pointless) retransmission after an ACK finds it as it was */
    if (fresh && !s_ok && status[d.st_idx[c]].crc_ok)
      for (uint32_t i = tid; i < d.rlen[c] / 8; i += blockDim.x) d.cache[c][i] = d.cb_bytes[c][i];
    __syncthreads();
  }
  int ret = -1;
  if (s_ok) {
    /* CRC-24A of the tbs bits (crc.c:144-158): every thread runs the byte-table CRC over its own contiguous run of bytes (one
     * shared-memory lookup per byte), then moves its 24-bit remainder to its place, remainder(x) * x^(bits behind the run) mod g,
     * through the power table r24a[m] = x^(m+24) mod g (one lookup per set remainder bit); the XOR of all is the CRC */
    uint32_t acc = 0;
    const uint32_t nbytes = d.tbs / 8, per = (nbytes + blockDim.x - 1) / blockDim.x;
    const uint32_t b0 = min(nbytes, tid * per), b1 = min(nbytes, b0 + per);
    if (b1 > b0) {
      uint32_t c = 0;
      for (uint32_t b = b0; b < b1; b++) c = ((c << 8) ^ s_tab[((c >> 16) ^ d.data[b]) & 0xFFu]) & 0xFFFFFFu;
      const uint32_t nshift = 8 * (nbytes - b1);
      if (nshift == 0) acc = c;
      else
        while (c) {
          const uint32_t hb = 31 - __clz(c), e = hb + nshift;
          acc ^= e >= 24 ? r24a[e - 24] : (1u << e);
          c &= ~(1u << hb);
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc ^= __shfl_xor_sync(0xFFFFFFFFu, acc, s);
    if ((tid & 31) == 0) s_red[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
      uint32_t par_rx = 0;
      for (int i = 0; i < 8; i++) par_rx ^= s_red[i];
      const uint32_t par_tx = ((uint32_t)d.data[nbytes] << 16) | ((uint32_t)d.data[nbytes + 1] << 8) | d.data[nbytes + 2];
      ret = (par_rx == par_tx && par_rx) ? 0 : -1;
    }
  }
  if (tid == 0) *d.ret = ret;
}

cudaError_t fec_launch_tb_assemble(const FecTbDesc *descs, const FecCbStatus *status, const uint32_t *r24a,
                                   uint32_t n, cudaStream_t st)
{
  if (n == 0) return cudaSuccess;
  fec_tb_assemble_kernel<<<n, 256, 0, st>>>(descs, status, r24a);
  return cudaGetLastError();
}

cudaError_t fec_kernels_init(void)
{
  cudaError_t e;
  e = fec_pk_init();
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fec_tdec_win_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  if (e != cudaSuccess) return e;
  /* fec_tdec_win_kernel<P, true> (one lane per thread) is the first form of the int8 decoders; the launcher uses the
   * two-lane fec_tdec_win8p_kernel, so it is not instantiated */
  e = cudaFuncSetAttribute(fec_tdec_win8p_kernel<32, FEC_WIN8_TMEM != 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fec_tdec_win8p_kernel<16, FEC_WIN8_TMEM != 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fec_tdec_win_kernel<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fec_tdec_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  if (e != cudaSuccess) return e;
  return cudaSuccess;
}
