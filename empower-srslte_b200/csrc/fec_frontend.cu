/*
 * fec_frontend.cu — the stages between the equaliser and decode_tb, fused into one pass over the equalised symbols
 * (SURVEY.md 8f rows f1 and f2):
 *
 *   soft demodulation    srslte_demod_soft_demodulate_s   demod_soft.c:437-456 (SSE bodies + scalar tails)
 *   scrambling sequence  srslte_sequence_LTE_pr           sequence.c:54-86 (36.211 7.2 Gold sequence)
 *   descrambling         srslte_scrambling_s_offset       scrambling.c:49-52
 *   CSI weighting        csi_correction (DL)              pdsch.c:675-776
 *   channel de-interleaver + RI/ACK handling (UL)         sch.c:860-881, 903-916; pusch.c:617-657
 *
 * The transport-block decoder then reads the int16 LLRs this kernel leaves in HBM; the host sends 8 bytes per
 * resource element instead of 2 bytes per LLR (12 at 64QAM). Results are bit-identical with the reference's AVX2 build,
 * quirks included: round-to-nearest in the SIMD bodies but truncation in the scalar tails, the CSI weights of a
 * QPSK / 64QAM symbol pair partly swapped by the blend in pdsch.c:724-726,745-752, and g[0] overwritten by the last
 * RI position (RI positions get interleaver index 0, sch.c:559-560).
 *
 * Elementwise, HBM-bound work: 8 B read + 2*Qm B written per resource element; one CTA per 2048 symbols, the scrambling
 * bits of the tile generated in shared memory from a jump table (no sequential 1600-step warm-up), the DL output
 * staged in shared memory so that global stores are full 16-byte lines.
 */
#include "fec_kernels.cuh"
#include <math.h>

#ifndef FE_SEG
#define FE_SEG      2048          /* symbols per CTA */
#endif
#define FE_THREADS  256
#define FE_BLK_BITS 256           /* scrambling bits produced per jump-table entry */

/* ---- 36.211 7.2 sequence: jump tables ---------------------------------------------------------------------------
 * x2(1600 + 256 b + k), k < 31, is a GF(2)-linear function of c_init: one 31-bit mask per (b, k); x1 does not depend
 * on c_init: one 31-bit state per b. From a state, 32 bits per step come out of word-parallel LFSR updates. */
extern "C" void fec_gen_gold_tables(uint32_t *x2_masks /* [FEC_FE_NBLK][32] */, uint32_t *x1_states /* [FEC_FE_NBLK] */)
{
  const uint32_t total = 1600 + FEC_FE_NBLK * FE_BLK_BITS + 31;
  uint32_t *m = (uint32_t *)malloc(sizeof(uint32_t) * (total + 31));
  uint8_t *x1 = (uint8_t *)calloc(total + 31, 1);
  for (int k = 0; k < 31; k++) m[k] = 1u << k;
  x1[0] = 1;
  for (uint32_t n = 0; n < total; n++) {
    m[n + 31] = m[n + 3] ^ m[n + 2] ^ m[n + 1] ^ m[n];
    x1[n + 31] = x1[n + 3] ^ x1[n];
  }
  for (uint32_t b = 0; b < FEC_FE_NBLK; b++) {
    uint32_t s1 = 0;
    for (int k = 0; k < 31; k++) {
      x2_masks[b * 32 + k] = m[1600 + b * FE_BLK_BITS + k];
      s1 |= (uint32_t)x1[1600 + b * FE_BLK_BITS + k] << k;
    }
    x2_masks[b * 32 + 31] = 0;
    x1_states[b] = s1;
  }
  free(m); free(x1);
}

/* s: bits 0..30 = x(n .. n+30). Returns x(n .. n+31) and advances the state by 32. */
__device__ __forceinline__ uint32_t fe_x2_next32(uint32_t &s)
{
  unsigned long long v = s;
  v |= ((v ^ (v >> 1) ^ (v >> 2) ^ (v >> 3)) & 0x0FFFFFFFull) << 31;
  v |= (((v >> 28) ^ (v >> 29) ^ (v >> 30) ^ (v >> 31)) & 0x1Full) << 59;
  s = (uint32_t)(v >> 32) & 0x7FFFFFFFu;
  return (uint32_t)v;
}
__device__ __forceinline__ uint32_t fe_x1_next32(uint32_t &s)
{
  unsigned long long v = s;
  v |= ((v ^ (v >> 3)) & 0x0FFFFFFFull) << 31;
  v |= (((v >> 28) ^ (v >> 31)) & 0x1Full) << 59;
  s = (uint32_t)(v >> 32) & 0x7FFFFFFFu;
  return (uint32_t)v;
}
__device__ __forceinline__ void fe_block_states(const uint32_t *x2_masks, const uint32_t *x1_states, uint32_t blk, uint32_t c_init,
                                                uint32_t &s1, uint32_t &s2)
{
  s2 = 0;
#pragma unroll
  for (int k = 0; k < 31; k++) s2 |= (uint32_t)(__popc(x2_masks[blk * 32 + k] & c_init) & 1) << k;
  s1 = x1_states[blk];
}
/* single bit c(n): used for the one out-of-tile bit of the g[0] quirk */
__device__ uint32_t fe_seq_bit(const uint32_t *x2_masks, const uint32_t *x1_states, uint32_t c_init, uint32_t n)
{
  uint32_t s1, s2, w = 0;
  fe_block_states(x2_masks, x1_states, n / FE_BLK_BITS, c_init, s1, s2);
  for (uint32_t i = 0; i <= (n % FE_BLK_BITS) / 32; i++) w = fe_x1_next32(s1) ^ fe_x2_next32(s2);
  return (w >> (n & 31)) & 1;
}

/* ---- conversions with the reference's x86 semantics ---- */
/* cvt(t)ps2dq + packssdw: NaN / out of int32 range -> 0x80000000 -> -32768 */
__device__ __forceinline__ int fe_sat16_rn(float v)
{
  if (!(v >= -2147483648.0f && v < 2147483648.0f)) return -32768;
  return max(-32768, min(32767, __float2int_rn(v)));
}
__device__ __forceinline__ int fe_sat16_rz(float v)
{
  if (!(v >= -2147483648.0f && v < 2147483648.0f)) return -32768;
  return max(-32768, min(32767, __float2int_rz(v)));
}
/* C conversion to short in the scalar tails: cvttsd2si to 32 bits, low half kept */
__device__ __forceinline__ int fe_c16(double v)
{
  if (!(v > -2147483649.0 && v < 2147483648.0)) return 0;
  return (int)(short)__double2int_rz(v);
}
__device__ __forceinline__ int fe_wrap16(int v) { return (int)(short)v; }
__device__ __forceinline__ int fe_abs16(int v) { return fe_wrap16(v < 0 ? -v : v); }     /* pabsw */

struct FeConst { float qpsk_scale; double sqrt2, off16, off64a, off64b; int ioff16, ioff64a, ioff64b; };

/* LLRs of one symbol, reference order; `body` = the symbol lies in the SIMD part of the reference's loop */
template <int mod>
__device__ __forceinline__ void fe_demod(float re, float im, bool body, bool body_im, const FeConst &c, int *v)
{
  if (mod == 6) {
    if (body) {
      const int yr = fe_sat16_rn(__fmul_rn(re, -700.0f)), yi = fe_sat16_rn(__fmul_rn(im, -700.0f));
      const int ar = fe_wrap16(fe_abs16(yr) - c.ioff64a), ai = fe_wrap16(fe_abs16(yi) - c.ioff64a);
      v[0] = yr; v[1] = yi; v[2] = ar; v[3] = ai;
      v[4] = fe_wrap16(fe_abs16(ar) - c.ioff64b); v[5] = fe_wrap16(fe_abs16(ai) - c.ioff64b);
    } else {
      const int yr = fe_c16((double)__fmul_rn(700.0f, re)), yi = fe_c16((double)__fmul_rn(700.0f, im));
      v[0] = fe_c16(-(double)yr); v[1] = fe_c16(-(double)yi);
      v[2] = fe_c16((double)abs(yr) - c.off64a); v[3] = fe_c16((double)abs(yi) - c.off64a);
      v[4] = fe_c16((double)abs(v[2]) - c.off64b); v[5] = fe_c16((double)abs(v[3]) - c.off64b);
    }
  } else if (mod == 4) {
    if (body) {
      const int yr = fe_sat16_rn(__fmul_rn(re, -400.0f)), yi = fe_sat16_rn(__fmul_rn(im, -400.0f));
      v[0] = yr; v[1] = yi;
      v[2] = fe_wrap16(fe_abs16(yr) - c.ioff16); v[3] = fe_wrap16(fe_abs16(yi) - c.ioff16);
    } else {
      const int yr = fe_c16((double)__fmul_rn(400.0f, re)), yi = fe_c16((double)__fmul_rn(400.0f, im));
      v[0] = fe_wrap16(-yr); v[1] = fe_wrap16(-yi);
      v[2] = fe_c16((double)abs(yr) - c.off16); v[3] = fe_c16((double)abs(yi) - c.off16);
    }
  } else if (mod == 2) {
    /* srslte_vec_convert_fi: 16 floats per AVX2 step, so re and im of one symbol may fall on different sides */
    const float pr = __fmul_rn(re, c.qpsk_scale), pi = __fmul_rn(im, c.qpsk_scale);
    v[0] = body ? fe_sat16_rz(pr) : fe_c16((double)pr);
    v[1] = body_im ? fe_sat16_rz(pi) : fe_c16((double)pi);
  } else {
    const float s = __fmul_rn(-100.0f, __fadd_rn(re, im));
    v[0] = fe_c16((double)s / c.sqrt2);
  }
}

/* t[i] for a run-time i without spilling the array to local memory */
template <int MOD>
__device__ __forceinline__ int fe_pick(const int *t, uint32_t i)
{
  int r = t[0];
#pragma unroll
  for (int k = 1; k < MOD; k++) if (i == (uint32_t)k) r = t[k];
  return r;
}

__device__ __forceinline__ int fe_mulhi16(int x, int w) { return (x * w) >> 16; }

/* position of `key` in the ascending list, and whether it is present */
__device__ __forceinline__ uint32_t fe_lower_bound(const uint32_t *a, uint32_t n, uint32_t key, bool &found)
{
  uint32_t lo = 0, hi = n;
  while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (a[mid] < key) lo = mid + 1; else hi = mid; }
  found = lo < n && a[lo] == key;
  return lo;
}

/* max of the CSI weights per transport block (srslte_vec_max_fi, pdsch.c:699-703) */
__global__ void __launch_bounds__(FE_THREADS) fec_fe_csimax_kernel(const FecFeDesc *descs, float *csi_max)
{
  const FecFeDesc d = descs[blockIdx.x];
  __shared__ float red[FE_THREADS / 32];
  float m = -INFINITY;
  if (d.csi) for (uint32_t i = threadIdx.x; i < d.nof_re; i += FE_THREADS) { const float v = d.csi[i]; if (v > m) m = v; }
  for (int o = 16; o; o >>= 1) { const float v = __shfl_xor_sync(0xFFFFFFFFu, m, o); if (v > m) m = v; }
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < FE_THREADS / 32; i++) if (red[i] > m) m = red[i];
    csi_max[blockIdx.x] = d.nof_re ? m : 1.0f;
  }
}

/* ---- UL tile: the channel de-interleaver is a transpose (q is column-major: symbol c*rows + r; g is row-major:
 * (r*cols + c)*Qm + k, sch.c:550-568), so a CTA takes FE_SEG/cols ROWS of all columns: per column a run of consecutive
 * symbols is read (coalesced), the LLRs go through a shared-memory tile in g order and leave as 16-byte stores. The few
 * bits at and behind the first HARQ-ACK / RI position (they sit in the bottom rows) take the per-bit path: ACK zeroing
 * (sch.c:913-916), RI removal (binary search in the ascending position lists), scattered 2-byte stores. ---- */
#define FE_UL_MAXCOLS 14   /* columns of the channel interleaver: 12 / 10 (normal / extended CP), one less with SRS; bound for the shared-memory run table */
#define FE_UL_RT 170       /* rows per tile at most: 255 + 6 * 170 bits fit FE_UL_RUNW words */
#define FE_UL_RUNW 40      /* scrambling words per column run: 255 + 6*FE_SEG/cols bits at most for cols >= 9 ... see fe_tile_ul */
template <int MOD>
__device__ __forceinline__ void fe_tile_ul(const FecFeDesc &d, uint32_t tile_idx, const uint32_t *x2_masks,
                                           const uint32_t *x1_states, const FeConst &kc, uint32_t *cw, int16_t *tile)
{
  constexpr uint32_t mod = MOD;
  const uint32_t cols = d.cols, rows = d.rows;
  const uint32_t rt = min((uint32_t)FE_SEG / cols, (uint32_t)FE_UL_RT), r0 = tile_idx * rt;
  const uint32_t nr = min(rt, rows - r0), nsym = nr * cols;
  const uint32_t total_bits = d.nof_re * mod;

  /* ---- scrambling bits of the cols runs of this tile: run c covers q bits [(c*rows + r0)*mod, +nr*mod); word w of the run
   *      holds c(256*(first block) + 32 w + b) at bit b ---- */
  const uint32_t run_blks = (255 + nr * mod + FE_BLK_BITS - 1) / FE_BLK_BITS;        /* <= FE_UL_RUNW / 8 */
  for (uint32_t j = threadIdx.x; j < cols * run_blks; j += FE_THREADS) {
    const uint32_t c = j / run_blks, i = j - c * run_blks;
    const uint32_t blk = ((c * rows + r0) * mod) / FE_BLK_BITS + i;
    if (blk * FE_BLK_BITS < total_bits) {
      uint32_t s1, s2;
      fe_block_states(x2_masks, x1_states, blk, d.c_init, s1, s2);
#pragma unroll
      for (int w = 0; w < FE_BLK_BITS / 32; w++) cw[c * FE_UL_RUNW + i * (FE_BLK_BITS / 32) + w] = fe_x1_next32(s1) ^ fe_x2_next32(s2);
    }
  }
  /* first de-interleaved bit index that needs the per-bit path, rounded down to a symbol */
  uint32_t first = 0xFFFFFFFFu;
  if (d.nack) first = min(first, d.ack_ord[0]);
  if (d.nri) first = min(first, d.ri_ord[0]);
  if (first != 0xFFFFFFFFu) first = (first / mod) * mod;
  __syncthreads();

  const uint32_t body_end = (mod == 2) ? 16 * ((2 * d.nof_re) / 16) : 4 * (d.nof_re / 4);    /* values (QPSK) / symbols */
  const uint32_t g0 = r0 * cols * mod;                    /* first g index of this tile */
  const uint32_t shift = g0 & 7u;                         /* same 16-byte phase in shared memory as in global memory */

  for (uint32_t t = threadIdx.x; t < nsym; t += FE_THREADS) {
    const uint32_t c = t / nr, lr = t - c * nr, sh = c * rows + r0 + lr;
    const float2 q = d.sym[sh];
    int v[6];
    const bool body = mod == 2 ? 2 * sh < body_end : sh < body_end;
    const bool body_im = mod == 2 ? 2 * sh + 1 < body_end : body;
    fe_demod<MOD>(q.x, q.y, body, body_im, kc, v);
    const uint32_t o0 = ((r0 + lr) * cols + c) * mod;     /* de-interleaved index of bit 0 of this symbol */
    const uint32_t boff = (((c * rows + r0) * mod) & (FE_BLK_BITS - 1)) + lr * mod;
    if (o0 + mod <= first) {
      /* bulk: descramble (psignw with +-1: wrapping negate), into the tile at its g position */
#pragma unroll
      for (uint32_t k = 0; k < mod; k++) {
        const uint32_t b = boff + k;
        int val = v[k];
        if ((cw[c * FE_UL_RUNW + (b >> 5)] >> (b & 31)) & 1) val = fe_wrap16(-val);
        tile[shift + o0 - g0 + k] = (int16_t)val;
      }
    } else {
#pragma unroll
      for (uint32_t k = 0; k < mod; k++) {
        bool hit = false, skip = false;
        int val = v[k];
        if (d.nack) { fe_lower_bound(d.ack_ord, d.nack, o0 + k, hit); if (hit) val = 0; }
        uint32_t before = 0;
        if (d.nri) { before = fe_lower_bound(d.ri_ord, d.nri, o0 + k, hit); skip = hit; }
        const uint32_t b = boff + k;
        if ((cw[c * FE_UL_RUNW + (b >> 5)] >> (b & 31)) & 1) val = fe_wrap16(-val);
        if (!skip) d.e[o0 + k - before] = (int16_t)val;
      }
    }
  }
  __syncthreads();
  /* ---- the tile leaves in g order: elements [g0, hi) with hi = end of the tile or the first per-bit index ---- */
  const uint32_t hi = min(g0 + nsym * mod, first);
  if (hi > g0) {
    const uint32_t n = hi - g0;
    const uint32_t head = min(n, (8u - shift) & 7u);      /* int16 until the global address is 16-byte aligned */
    for (uint32_t i = threadIdx.x; i < head; i += FE_THREADS) d.e[g0 + i] = tile[shift + i];
    const uint32_t nvec = (n - head) / 8;
    const uint4 *src = reinterpret_cast<const uint4 *>(tile + shift + head);
    uint4 *dst = reinterpret_cast<uint4 *>(d.e + g0 + head);
    for (uint32_t i = threadIdx.x; i < nvec; i += FE_THREADS) dst[i] = src[i];
    for (uint32_t i = head + 8 * nvec + threadIdx.x; i < n; i += FE_THREADS) d.e[g0 + i] = tile[shift + i];
  }
  __syncthreads();
  if (tile_idx == 0 && d.nri && threadIdx.x == 0) {
    /* g[0] quirk: every RI position scatters to index 0 after the rightful owner did (sch.c:559, vector.c:119-123); the
     * last one in q order stays */
    const uint32_t p = d.last_ri_pos, ps = p / mod;
    const float2 q = d.sym[ps];
    int t[6];
    const bool body = mod == 2 ? 2 * ps < body_end : ps < body_end;
    const bool body_im = mod == 2 ? 2 * ps + 1 < body_end : body;
    fe_demod<MOD>(q.x, q.y, body, body_im, kc, t);
    const int val = fe_seq_bit(x2_masks, x1_states, d.c_init, p) ? fe_wrap16(-fe_pick<MOD>(t, p - ps * mod)) : fe_pick<MOD>(t, p - ps * mod);
    d.e[0] = (int16_t)val;
  }
}

/* ---- DL tile: registers -> packed words -> shared-memory tile -> 16-byte stores ---- */
template <int MOD, bool CSI>
__device__ __forceinline__ void fe_tile_dl(const FecFeDesc &d, uint32_t s0, float cmax, const uint32_t *x2_masks,
                                           const uint32_t *x1_states, const FeConst &kc, uint32_t *cw, uint32_t *tile)
{
  constexpr uint32_t mod = MOD;
  constexpr int NP = FE_SEG / 2 / FE_THREADS;                /* symbol pairs per thread */
  const uint32_t nsym = min((uint32_t)FE_SEG, d.nof_re - s0);
  const uint32_t nbits = nsym * mod, p0 = s0 * mod;
  const uint32_t body_end = (mod == 2) ? 16 * ((2 * d.nof_re) / 16) : 4 * (d.nof_re / 4);

  /* symbols (and weights) of all pairs of this thread first: their HBM latency hides behind the sequence generation */
  float4 q[NP];
  float cs[NP][2];
#pragma unroll
  for (int r = 0; r < NP; r++) {
    const uint32_t ls = 2 * (threadIdx.x + r * FE_THREADS);
    q[r] = make_float4(0.f, 0.f, 0.f, 0.f);
    cs[r][0] = cs[r][1] = 0.f;
    if (ls + 1 < nsym) q[r] = *reinterpret_cast<const float4 *>(d.sym + s0 + ls);
    else if (ls < nsym) { const float2 t = d.sym[s0 + ls]; q[r].x = t.x; q[r].y = t.y; }
    if (CSI) {
      if (ls < nsym) cs[r][0] = d.csi[s0 + ls];
      if (ls + 1 < nsym) cs[r][1] = d.csi[s0 + ls + 1];
    }
  }
  {
    const uint32_t nblk = (nbits + FE_BLK_BITS - 1) / FE_BLK_BITS;
    if (threadIdx.x < nblk) {
      uint32_t s1, s2;
      fe_block_states(x2_masks, x1_states, p0 / FE_BLK_BITS + threadIdx.x, d.c_init, s1, s2);
#pragma unroll
      for (int w = 0; w < FE_BLK_BITS / 32; w++) cw[threadIdx.x * (FE_BLK_BITS / 32) + w] = fe_x1_next32(s1) ^ fe_x2_next32(s2);
    }
    if (threadIdx.x == 0) cw[nblk * (FE_BLK_BITS / 32)] = 0;
  }
  __syncthreads();
  const float cscale = CSI ? __fdiv_rn(32767.0f, cmax) : 0.f;

#pragma unroll
  for (int r = 0; r < NP; r++) {
    const uint32_t pr = threadIdx.x + r * FE_THREADS, ls = 2 * pr, s = s0 + ls;
    if (ls >= nsym) break;
    const bool two = ls + 1 < nsym;
    int v[12];
#pragma unroll
    for (int k = 0; k < 12; k++) v[k] = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const uint32_t sh = s + h;
      const bool body = mod == 2 ? 2 * sh < body_end : sh < body_end;
      const bool body_im = mod == 2 ? 2 * sh + 1 < body_end : body;
      fe_demod<MOD>(h ? q[r].z : q[r].x, h ? q[r].w : q[r].y, body, body_im, kc, v + 6 * h);
    }
    /* descrambling: the 2*mod sequence bits of the pair sit at bit ls*mod of the tile's sequence words */
    const uint32_t b0 = ls * mod;
    const uint32_t bits = __funnelshift_r(cw[b0 >> 5], cw[(b0 >> 5) + 1], b0 & 31);
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
      for (uint32_t k = 0; k < mod; k++)
        if ((bits >> (h * mod + k)) & 1) v[6 * h + k] = CSI ? fe_wrap16(-v[6 * h + k]) : -v[6 * h + k];
    if (CSI) {
      /* csi_correction: SSE body over whole groups (QPSK / 64QAM: symbol pairs, 16QAM: symbols), scalar tail */
      if (mod == 6 || mod == 2) {
        if (two) {
          const int w1 = fe_sat16_rn(__fmul_rn(cs[r][0], cscale)), w3 = fe_sat16_rn(__fmul_rn(cs[r][1], cscale));
          if (mod == 6) {
#pragma unroll
            for (int k = 0; k < 4; k++) v[k] = fe_mulhi16(v[k], w1);
            v[4] = fe_mulhi16(v[4], w3); v[5] = fe_mulhi16(v[5], w3);
            v[6] = fe_mulhi16(v[6], w1); v[7] = fe_mulhi16(v[7], w1);
#pragma unroll
            for (int k = 2; k < 6; k++) v[6 + k] = fe_mulhi16(v[6 + k], w3);
          } else {
            v[0] = fe_mulhi16(v[0], w3); v[1] = fe_mulhi16(v[1], w3);
            v[6] = fe_mulhi16(v[6], w1); v[7] = fe_mulhi16(v[7], w1);
          }
        } else {
          const float w = __fdiv_rn(cs[r][0], cmax);
#pragma unroll
          for (uint32_t k = 0; k < mod; k++) v[k] = fe_c16((double)__fmul_rn((float)v[k], w));
        }
      } else if (mod == 4) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int w = fe_sat16_rn(__fmul_rn(cs[r][h], cscale));
#pragma unroll
          for (int k = 0; k < 4; k++) v[6 * h + k] = fe_mulhi16(v[6 * h + k], w);
        }
      }
    }
    /* 2*mod int16 = mod words (mod 1: one word, upper half unused by a lone symbol) */
    if (mod == 1) {
      tile[pr] = __byte_perm(v[0], v[6], 0x5410);
    } else {
      uint32_t w[MOD];
#pragma unroll
      for (uint32_t k = 0; k < mod / 2; k++) {
        w[k] = __byte_perm(v[2 * k], v[2 * k + 1], 0x5410);
        w[mod / 2 + k] = __byte_perm(v[6 + 2 * k], v[6 + 2 * k + 1], 0x5410);
      }
      if (mod == 2) { *reinterpret_cast<uint2 *>(tile + 2 * pr) = make_uint2(w[0], w[1]); }
      else {
#pragma unroll
        for (uint32_t k = 0; k < mod / 2; k++) *reinterpret_cast<uint2 *>(tile + mod * pr + 2 * k) = make_uint2(w[2 * k], w[2 * k + 1]);
      }
    }
  }
  __syncthreads();
  /* tile -> HBM: p0*2 bytes is a multiple of 16 (p0 = 2048*mod*blockIdx.x) */
  int16_t *dst = d.e + p0;
  const uint32_t n16 = nbits / 8;
  for (uint32_t i = threadIdx.x; i < n16; i += FE_THREADS) reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(tile)[i];
  for (uint32_t i = n16 * 8 + threadIdx.x; i < nbits; i += FE_THREADS) dst[i] = reinterpret_cast<const int16_t *>(tile)[i];
}

__global__ void __launch_bounds__(FE_THREADS, 4)
fec_fe_dl_kernel(const FecFeDesc *descs, const float *csi_max, const uint32_t *x2_masks, const uint32_t *x1_states, FeConst kc)
{
  const FecFeDesc d = descs[blockIdx.y];
  const uint32_t s0 = blockIdx.x * FE_SEG;
  if (d.cols || s0 >= d.nof_re) return;
  __shared__ uint32_t cw[FE_SEG * 6 / 32 + 1];
  __shared__ __align__(16) uint32_t tile[FE_SEG * 6 / 2];
  if (d.csi) {
    const float cmax = csi_max[blockIdx.y];
    switch (d.mod) {
    case 6: fe_tile_dl<6, true>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
    case 4: fe_tile_dl<4, true>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
    default: fe_tile_dl<2, true>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
    }
  } else {
    switch (d.mod) {
    case 6: fe_tile_dl<6, false>(d, s0, 1.f, x2_masks, x1_states, kc, cw, tile); break;
    case 4: fe_tile_dl<4, false>(d, s0, 1.f, x2_masks, x1_states, kc, cw, tile); break;
    case 2: fe_tile_dl<2, false>(d, s0, 1.f, x2_masks, x1_states, kc, cw, tile); break;
    default: fe_tile_dl<1, false>(d, s0, 1.f, x2_masks, x1_states, kc, cw, tile); break;
    }
  }
}

/* ---- 8-bit LLR mode (q->llr_is_8bit, pdsch.c:795-813): srslte_demod_soft_demodulate_b (demod_soft.c:458-478: SSE bodies
 * over 8 symbols — 16 values for QPSK — with two saturating packs, scalar tails with C conversions), srslte_scrambling_sb_offset
 * (psignb: wrapping int8 negate) and the scalar int8 branch of csi_correction (pdsch.c:707-714). Downlink only. One byte per
 * LLR leaves for HBM, so this is not the bandwidth-critical form of the kernel: one symbol per thread and pass, the LLR bytes
 * staged in shared memory so that the stores are 16-byte lines. ---- */
__device__ __forceinline__ int fe_wrap8(int v) { return (int)(signed char)v; }
__device__ __forceinline__ int fe_abs8(int v) { return fe_wrap8(v < 0 ? -v : v); }        /* pabsb */
__device__ __forceinline__ int fe_sat8(int v16) { return max(-128, min(127, v16)); }      /* packsswb on a packssdw result */
__device__ __forceinline__ int fe_c8(double v)                                             /* cvttsd2si to 32 bits, low byte kept */
{
  if (!(v > -2147483649.0 && v < 2147483648.0)) return 0;
  return (int)(signed char)__double2int_rz(v);
}
struct FeConst8 { float qpsk_scale; double sqrt2, off16, off64a, off64b; int ioff16, ioff64a, ioff64b; };

template <int mod>
__device__ __forceinline__ void fe_demod8(float re, float im, bool body, bool body_im, const FeConst8 &c, int *v)
{
  if (mod == 6) {
    if (body) {
      const int yr = fe_sat8(fe_sat16_rn(__fmul_rn(re, -40.0f))), yi = fe_sat8(fe_sat16_rn(__fmul_rn(im, -40.0f)));
      const int ar = fe_wrap8(fe_abs8(yr) - c.ioff64a), ai = fe_wrap8(fe_abs8(yi) - c.ioff64a);
      v[0] = yr; v[1] = yi; v[2] = ar; v[3] = ai;
      v[4] = fe_wrap8(fe_abs8(ar) - c.ioff64b); v[5] = fe_wrap8(fe_abs8(ai) - c.ioff64b);
    } else {
      const int yr = fe_c8((double)__fmul_rn(40.0f, re)), yi = fe_c8((double)__fmul_rn(40.0f, im));
      v[0] = fe_c8(-(double)yr); v[1] = fe_c8(-(double)yi);
      v[2] = fe_c8((double)abs(yr) - c.off64a); v[3] = fe_c8((double)abs(yi) - c.off64a);
      v[4] = fe_c8((double)abs(v[2]) - c.off64b); v[5] = fe_c8((double)abs(v[3]) - c.off64b);
    }
  } else if (mod == 4) {
    if (body) {
      const int yr = fe_sat8(fe_sat16_rn(__fmul_rn(re, -30.0f))), yi = fe_sat8(fe_sat16_rn(__fmul_rn(im, -30.0f)));
      v[0] = yr; v[1] = yi;
      v[2] = fe_wrap8(fe_abs8(yr) - c.ioff16); v[3] = fe_wrap8(fe_abs8(yi) - c.ioff16);
    } else {
      const int yr = fe_c8((double)__fmul_rn(30.0f, re)), yi = fe_c8((double)__fmul_rn(30.0f, im));
      v[0] = fe_wrap8(-yr); v[1] = fe_wrap8(-yi);
      v[2] = fe_c8((double)abs(yr) - c.off16); v[3] = fe_c8((double)abs(yi) - c.off16);
    }
  } else if (mod == 2) {
    const float pr = __fmul_rn(re, c.qpsk_scale), pi = __fmul_rn(im, c.qpsk_scale);
    v[0] = body ? fe_sat8(fe_sat16_rz(pr)) : fe_c8((double)pr);
    v[1] = body_im ? fe_sat8(fe_sat16_rz(pi)) : fe_c8((double)pi);
  } else {
    const float s = __fmul_rn(-20.0f, __fadd_rn(re, im));
    v[0] = fe_c8((double)s / c.sqrt2);
  }
}

template <int MOD>
__device__ __forceinline__ void fe_tile_dl8(const FecFeDesc &d, uint32_t s0, float cmax, const uint32_t *x2_masks,
                                            const uint32_t *x1_states, const FeConst8 &kc, uint32_t *cw, signed char *tile)
{
  constexpr uint32_t mod = MOD;
  const uint32_t nsym = min((uint32_t)FE_SEG, d.nof_re - s0);
  const uint32_t nbits = nsym * mod, p0 = s0 * mod;
  const uint32_t body_end = (mod == 2) ? 16 * ((2 * d.nof_re) / 16) : 8 * (d.nof_re / 8);     /* values (QPSK) / symbols */
  {
    const uint32_t nblk = (nbits + FE_BLK_BITS - 1) / FE_BLK_BITS;
    if (threadIdx.x < nblk) {
      uint32_t s1, s2;
      fe_block_states(x2_masks, x1_states, p0 / FE_BLK_BITS + threadIdx.x, d.c_init, s1, s2);
#pragma unroll
      for (int w = 0; w < FE_BLK_BITS / 32; w++) cw[threadIdx.x * (FE_BLK_BITS / 32) + w] = fe_x1_next32(s1) ^ fe_x2_next32(s2);
    }
  }
  __syncthreads();
  for (uint32_t ls = threadIdx.x; ls < nsym; ls += FE_THREADS) {
    const uint32_t sh = s0 + ls;
    const float2 q = d.sym[sh];
    const bool body = mod == 2 ? 2 * sh < body_end : sh < body_end;
    const bool body_im = mod == 2 ? 2 * sh + 1 < body_end : body;
    int v[6];
    fe_demod8<MOD>(q.x, q.y, body, body_im, kc, v);
    const float w = d.csi ? __fdiv_rn(d.csi[sh], cmax) : 1.f;
#pragma unroll
    for (uint32_t k = 0; k < mod; k++) {
      const uint32_t b = ls * mod + k;
      int val = v[k];
      if ((cw[b >> 5] >> (b & 31)) & 1) val = fe_wrap8(-val);
      if (d.csi) val = fe_c8((double)__fmul_rn((float)val, w));
      tile[b] = (signed char)val;
    }
  }
  __syncthreads();
  /* tile -> HBM: p0 bytes is a multiple of 16 (p0 = 2048*mod*blockIdx.x) */
  signed char *dst = reinterpret_cast<signed char *>(d.e) + p0;
  const uint32_t n16 = nbits / 16;
  for (uint32_t i = threadIdx.x; i < n16; i += FE_THREADS) reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(tile)[i];
  for (uint32_t i = n16 * 16 + threadIdx.x; i < nbits; i += FE_THREADS) dst[i] = tile[i];
}

__global__ void __launch_bounds__(FE_THREADS)
fec_fe_dl8_kernel(const FecFeDesc *descs, const float *csi_max, const uint32_t *x2_masks, const uint32_t *x1_states, FeConst8 kc)
{
  const FecFeDesc d = descs[blockIdx.y];
  const uint32_t s0 = blockIdx.x * FE_SEG;
  if (d.cols || s0 >= d.nof_re) return;
  __shared__ uint32_t cw[FE_SEG * 6 / 32 + 1];
  __shared__ __align__(16) signed char tile[FE_SEG * 6];
  const float cmax = d.csi ? csi_max[blockIdx.y] : 1.f;
  switch (d.mod) {
  case 6: fe_tile_dl8<6>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
  case 4: fe_tile_dl8<4>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
  case 2: fe_tile_dl8<2>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
  default: fe_tile_dl8<1>(d, s0, cmax, x2_masks, x1_states, kc, cw, tile); break;
  }
}

__global__ void __launch_bounds__(FE_THREADS)
fec_fe_ul_kernel(const FecFeDesc *descs, const uint32_t *x2_masks, const uint32_t *x1_states, FeConst kc)
{
  const FecFeDesc d = descs[blockIdx.y];
  if (!d.cols) return;
  __shared__ uint32_t cw[FE_UL_MAXCOLS * FE_UL_RUNW];
  __shared__ __align__(16) int16_t tile[FE_SEG * 6 + 8];
  const uint32_t rt = min((uint32_t)FE_SEG / d.cols, (uint32_t)FE_UL_RT);
  for (uint32_t t = blockIdx.x; t * rt < d.rows; t += gridDim.x) {        /* grid-stride: any cols / rows combination is covered */
    switch (d.mod) {
    case 6: fe_tile_ul<6>(d, t, x2_masks, x1_states, kc, cw, tile); break;
    case 4: fe_tile_ul<4>(d, t, x2_masks, x1_states, kc, cw, tile); break;
    default: fe_tile_ul<2>(d, t, x2_masks, x1_states, kc, cw, tile); break;
    }
    __syncthreads();
  }
}

/* UL control information hand-back: raw (pre-descrambling) q LLRs at the HARQ-ACK and RI positions in the caller's order
 * (what srslte_uci_decode_ack_ri reads, sch.c:903-930), then the first n_cqi LLRs of g (srslte_uci_decode_cqi_pusch,
 * sch.c:951-960). Runs after fec_fe_kernel. */
__global__ void __launch_bounds__(FE_THREADS) fec_fe_uci_kernel(const FecFeDesc *descs, FeConst kc)
{
  const FecFeDesc d = descs[blockIdx.x];
  if (!d.uci_out) return;
  const uint32_t mod = d.mod;
  const uint32_t body_end = (mod == 2) ? 16 * ((2 * d.nof_re) / 16) : 4 * (d.nof_re / 4);
  for (uint32_t i = threadIdx.x; i < d.nuci; i += FE_THREADS) {
    const uint32_t p = d.uci_pos[i], s = p / mod;
    const float2 q = d.sym[s];
    int t[6] = {0, 0, 0, 0, 0, 0};
    const bool body = mod == 2 ? 2 * s < body_end : s < body_end;
    const bool body_im = mod == 2 ? 2 * s + 1 < body_end : body;
    switch (mod) {
    case 6: fe_demod<6>(q.x, q.y, body, body_im, kc, t); break;
    case 4: fe_demod<4>(q.x, q.y, body, body_im, kc, t); break;
    case 2: fe_demod<2>(q.x, q.y, body, body_im, kc, t); break;
    default: fe_demod<1>(q.x, q.y, body, body_im, kc, t); break;
    }
    d.uci_out[i] = (int16_t)fe_pick<6>(t, p - s * mod);
  }
  for (uint32_t i = threadIdx.x; i < d.n_cqi; i += FE_THREADS) d.uci_out[d.nuci + i] = d.e[i];
}

cudaError_t fec_launch_frontend(const FecFeDesc *descs, float *csi_max, const uint32_t *x2_masks, const uint32_t *x1_states,
                                uint32_t ntb, uint32_t max_re, int any_csi, int any_uci, int kinds, cudaStream_t st, int llr8)
{
  if (!ntb || !max_re) return cudaSuccess;
  if (llr8) {
    /* 8-bit LLR mode: downlink only (the caller has refused uplink transport blocks) */
    FeConst8 k8;
    k8.qpsk_scale = (float)(-20 * sqrt(2.0));
    k8.sqrt2 = sqrt(2.0);
    k8.off16 = 2 * 30 / sqrt(10.0); k8.off64a = 4 * 40 / sqrt(42.0); k8.off64b = 2 * 40 / sqrt(42.0);
    k8.ioff16 = (signed char)k8.off16; k8.ioff64a = (signed char)k8.off64a; k8.ioff64b = (signed char)k8.off64b;
    if (any_csi) fec_fe_csimax_kernel<<<ntb, FE_THREADS, 0, st>>>(descs, csi_max);
    dim3 g8((max_re + FE_SEG - 1) / FE_SEG, ntb);
    fec_fe_dl8_kernel<<<g8, FE_THREADS, 0, st>>>(descs, csi_max, x2_masks, x1_states, k8);
    return cudaGetLastError();
  }
  FeConst kc;
  kc.qpsk_scale = (float)(-100 * sqrt(2.0));
  kc.sqrt2 = sqrt(2.0);
  kc.off16 = 2 * 400 / sqrt(10.0); kc.off64a = 4 * 700 / sqrt(42.0); kc.off64b = 2 * 700 / sqrt(42.0);
  kc.ioff16 = (short)kc.off16; kc.ioff64a = (short)kc.off64a; kc.ioff64b = (short)kc.off64b;
  if (any_csi) fec_fe_csimax_kernel<<<ntb, FE_THREADS, 0, st>>>(descs, csi_max);
  dim3 grid((max_re + FE_SEG - 1) / FE_SEG, ntb);
  if (kinds & 1) fec_fe_dl_kernel<<<grid, FE_THREADS, 0, st>>>(descs, csi_max, x2_masks, x1_states, kc);
  /* UL tiles are FE_SEG / cols whole rows: up to one tile more than nof_re / FE_SEG (cols does not divide FE_SEG) */
  dim3 grid_ul(grid.x + 1, ntb);
  if (kinds & 2) fec_fe_ul_kernel<<<grid_ul, FE_THREADS, 0, st>>>(descs, x2_masks, x1_states, kc);
  if (any_uci) fec_fe_uci_kernel<<<ntb, FE_THREADS, 0, st>>>(descs, kc);
  return cudaGetLastError();
}
