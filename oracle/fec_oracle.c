/*
 * fec_oracle.c — scalar C restatement of the reference decode path (see fec_oracle.h for status).
 * TEST INFRASTRUCTURE ONLY; never linked into the product.
 *
 * Written from the algorithm description in SURVEY.md section 8 and 3GPP TS 36.212 (5.1.1 CRC, 5.1.2
 * segmentation, 5.1.3.2 turbo code/QPP, 5.1.4.1 rate matching); each function names the reference
 * file:line whose behaviour it restates and is checked bit-for-bit against it.
 */
#define _GNU_SOURCE
#include "fec_oracle.h"
#include "lte_qpp_table.h"
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ sizes / segmentation ---- */

int orc_cb_index(int K)
{   /* first table entry >= K (cbsegm.c:116-127) */
  for (int i = 0; i < LTE_NOF_CB_SIZES; i++) if (lte_qpp_table[i][0] >= K) return i;
  return -1;
}
int orc_cb_size(int idx) { return (idx >= 0 && idx < LTE_NOF_CB_SIZES) ? lte_qpp_table[idx][0] : -1; }

int orc_subblocks(int K)
{   /* turbodecoder.c:364-376 with LV_HAVE_AVX2 */
  if (K % 16 == 0 && K > 800) return 16;
  if (K % 8 == 0 && K > 400) return 8;
  return 0;
}

/* decoder class of the 8-bit LLR path (turbodecoder.c:392-406, AVX2 build): 32 / 16 sub-blocks run in int8 arithmetic,
 * 8 / 0 fall back to the 16-bit decoders on widened inputs (turbodecoder.c:408-452) */
int orc_subblocks8(int K)
{
  if (!(K % 32) && K > 2048) return 32;
  if (!(K % 16) && K > 800) return 16;
  if (!(K % 8) && K > 400) return 8;
  return 0;
}

int orc_cbsegm(uint32_t tbs, uint32_t *o)
{   /* cbsegm.c:58-108; o = {F,C,K1,K2,K1_idx,K2_idx,C1,C2,tbs} */
  memset(o, 0, 9 * sizeof(uint32_t));
  if (tbs == 0) return 0;
  uint32_t B = tbs + 24, C, Bp;
  if (B <= 6144) { C = 1; Bp = B; }
  else { C = (uint32_t)ceilf((float)B / (6144 - 24)); Bp = B + 24 * C; }
  int i1 = orc_cb_index((int)((Bp - 1) / C + 1));
  if (i1 < 0) return -1;
  uint32_t K1 = lte_qpp_table[i1][0], K2 = 0, K2i = 0, C1 = 1, C2 = 0;
  if (C > 1) {
    K2i = i1 - 1; K2 = lte_qpp_table[K2i][0];
    C2 = (C * K1 - Bp) / (K1 - K2); C1 = C - C2;
  }
  o[0] = C1 * K1 + C2 * K2 - Bp; o[1] = C; o[2] = K1; o[3] = K2; o[4] = i1; o[5] = K2i; o[6] = C1; o[7] = C2; o[8] = tbs;
  return 0;
}

/* ------------------------------------------------------------------ CRC ---- */
/* MSB-first, zero init, no reflection, no final xor (crc.c:74-101,144-158). Like
 * srslte_crc_checksum_byte, only the nbits/8 WHOLE bytes are consumed ("len is multiple of 8",
 * crc.c:143); every length on this path (K, tbs, tbs+24) is a multiple of 8. */
uint32_t orc_crc(int type, const uint8_t *bytes, int nbits)
{
  const uint32_t poly = type ? 0x1800063u : 0x1864CFBu;
  uint32_t r = 0;
  nbits &= ~7;
  for (int i = 0; i < nbits; i++) {
    uint32_t bit = (bytes[i >> 3] >> (7 - (i & 7))) & 1u;
    r = (r << 1) | bit;            /* message bit enters at the bottom of a 25-bit register */
    if (r & 0x1000000u) r ^= poly;
  }
  for (int i = 0; i < 24; i++) {   /* flush 24 zero bits = multiply by x^24 */
    r <<= 1;
    if (r & 0x1000000u) r ^= poly;
  }
  return r & 0xFFFFFFu;
}

/* ------------------------------------------------------------------ QPP interleaver ---- */
int orc_qpp(int K, uint16_t *fwd, uint16_t *rev)
{
  int idx = orc_cb_index(K);
  if (idx < 0 || lte_qpp_table[idx][0] != K) return -1;
  uint64_t f1 = lte_qpp_table[idx][1], f2 = lte_qpp_table[idx][2];
  for (uint64_t i = 0; i < (uint64_t)K; i++) {
    uint64_t j = (f1 * i + f2 * i * i) % (uint64_t)K;
    fwd[i] = (uint16_t)j; rev[j] = (uint16_t)i;
  }
  return 0;
}

int orc_sb_pos(int x, int K, int P)
{
  if (P <= 1) return x;
  int L = K / P;
  return (x % L) * P + x / L;
}

int orc_qpp_sb(int K, int P, uint16_t *fwd, uint16_t *rev)
{   /* tc_interl_lte.c:104-112: tables re-indexed into the sb domain */
  static __thread uint16_t f[ORC_MAX_K], r[ORC_MAX_K];
  if (orc_qpp(K, f, r)) return -1;
  if (P <= 1) { memcpy(fwd, f, 2 * K); memcpy(rev, r, 2 * K); return 0; }
  int L = K / P;
  for (int pos = 0; pos < K; pos++) {
    int nat = (pos % P) * L + pos / P;
    fwd[pos] = (uint16_t)orc_sb_pos(f[nat], K, P);
    rev[pos] = (uint16_t)orc_sb_pos(r[nat], K, P);
  }
  return 0;
}

/* ------------------------------------------------------------------ rate de-matching ---- */
static const uint8_t RM_COLPERM[32] = {0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                                       1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};

/* tab[i] = natural position (3*idx+stream, tail included) of the i-th transmitted bit for this rv,
 * i in [0, 3K+12). 36.212 5.1.4.1.1-2; restates rm_turbo.c:164-238. */
int orc_rm_table(int K, int rv, uint16_t *tab)
{
  const int D = K + 4, R = (D - 1) / 32 + 1, KP = 32 * R, ND = KP - D, NCB = 3 * KP;
  const int k0 = R * (2 * (int)ceilf((float)NCB / (float)(8 * R)) * rv + 2);
  int n = 0;
  for (int j = 0; n < 3 * D; j++) {
    int w = (k0 + j) % NCB, y, stream;
    if (w < KP)                { stream = 0; y = (w % R) * 32 + RM_COLPERM[w / R]; }
    else if (((w - KP) & 1) == 0) { int m = (w - KP) / 2; stream = 1; y = (m % R) * 32 + RM_COLPERM[m / R]; }
    else { int m = (w - KP) / 2; stream = 2; y = (RM_COLPERM[m / R] + 32 * (m % R) + 1) % KP; }
    if (y < ND) continue;          /* <NULL> dummy bit */
    tab[n++] = (uint16_t)(3 * (y - ND) + stream);
  }
  return 0;
}

/* natural coded position t -> index in the reference soft buffer (rm_turbo.c:250-264) */
int orc_rm_soft_index(int t, int K, int P)
{
  if (P == 0) return t;
  if (t < 3 * K) return (t % 3) * (K + 32) + orc_sb_pos(t / 3, K, P);
  return 3 * (K + 32) + (t - 3 * K);
}

int orc_rm_rx(const int16_t *in, int E, int K, int rv, int sb_layout, int16_t *soft)
{   /* soft[deinter[i % (3K+12)]] += e[i], wrapping int16 (rm_turbo.c:378-430) */
  static __thread uint16_t tab[3 * ORC_MAX_K + 12];
  if (rv < 0 || rv > 3 || orc_cb_index(K) < 0 || orc_cb_size(orc_cb_index(K)) != K) return -2;
  orc_rm_table(K, rv, tab);
  int P = sb_layout ? orc_subblocks(K) : 0, n = 3 * K + 12;
  for (int i = 0; i < E; i++) {
    int pos = orc_rm_soft_index(tab[i % n], K, P);
    soft[pos] = (int16_t)(soft[pos] + in[i]);
  }
  return 0;
}

/* ------------------------------------------------------------------ constituent MAP decoders ---- */
#define NEG_INF (-10000)
/* test instrumentation: how often a saturating op clipped / peak intermediate magnitude / peak MAP input magnitude
 * since the last orc_dbg_reset() (thread-local). Used to validate the range proof behind the packed CUDA fast path. */
static __thread long orc_sat_count = 0;
static __thread int  orc_peak_val = 0, orc_peak_in = 0;
/* largest spread max_i o[i] - min_i o[i] of the 8 state metrics seen at a normalisation point, per recursion
 * (0 beta warm-up, 1 beta, 2 alpha warm-up, 3 alpha): the quantity the packed kernel's run-time range certificate bounds */
static __thread int  orc_phase = 0, orc_spread[4] = {0, 0, 0, 0};
/* the quantities the packed CUDA kernel's run-time range certificate is computed from (fec_tdec_packed.cu, pk_certificate):
 * orc_mag[phase]: largest |state metric| right after a normalisation and at the start of a main pass (sub-block 0's known
 * alpha start state excluded); orc_llr_range: largest (max - min) over one sub-block of the two LLR maxima m0 / m1 */
static __thread int  orc_mag[4] = {0, 0, 0, 0}, orc_llr_range = 0;
void orc_dbg_reset(void) { orc_sat_count = 0; orc_peak_val = 0; orc_peak_in = 0; orc_llr_range = 0; for (int i = 0; i < 4; i++) orc_spread[i] = orc_mag[i] = 0; }
int  orc_dbg_spread(int phase) { return orc_spread[phase & 3]; }
int  orc_dbg_mag(int phase) { return orc_mag[phase & 3]; }
int  orc_dbg_llr_range(void) { return orc_llr_range; }
static void orc_track_mag(const int16_t o[8])
{
  for (int i = 0; i < 8; i++) { int a = o[i] < 0 ? -(int)o[i] : o[i]; if (a > orc_mag[orc_phase]) orc_mag[orc_phase] = a; }
}
long orc_dbg_sat_count(void) { return orc_sat_count; }
int  orc_dbg_peak_val(void) { return orc_peak_val; }
int  orc_dbg_peak_in(void) { return orc_peak_in; }
static inline int16_t sat16(int v)
{
  int a = v < 0 ? -v : v;
  if (a > orc_peak_val) orc_peak_val = a;
  if (v > 32767) { orc_sat_count++; return 32767; }
  if (v < -32768) { orc_sat_count++; return -32768; }
  return (int16_t)v;
}
static inline int16_t adds(int16_t a, int16_t b) { return sat16((int)a + (int)b); }
static inline int16_t subs(int16_t a, int16_t b) { return sat16((int)a - (int)b); }
static inline int16_t wrap16(int v) { return (int16_t)(uint16_t)v; }
static inline int16_t maxs(int16_t a, int16_t b) { return a > b ? a : b; }

/* backward (beta) step of the windowed decoder: o <- max(m,n) (turbodecoder_win.h:397-418) */
static void win_beta_step(int16_t o[8], int16_t x, int16_t y, int saturate)
{
  int16_t xy, m[8], n[8];
  if (saturate) {
    xy = adds(x, y);
    m[0] = adds(o[4], xy); m[1] = o[4];          m[2] = adds(o[5], y);  m[3] = adds(o[5], x);
    m[4] = adds(o[6], x);  m[5] = adds(o[6], y); m[6] = o[7];           m[7] = adds(o[7], xy);
    n[0] = o[0];           n[1] = adds(o[0], xy); n[2] = adds(o[1], x);  n[3] = adds(o[1], y);
    n[4] = adds(o[2], y);  n[5] = adds(o[2], x); n[6] = adds(o[3], xy); n[7] = o[3];
  } else {   /* tail trellis uses plain (wrapping) adds: turbodecoder_win.h:235-242,263-307 */
    xy = wrap16(x + y);
    m[0] = wrap16(o[4] + xy); m[1] = o[4];             m[2] = wrap16(o[5] + y);  m[3] = wrap16(o[5] + x);
    m[4] = wrap16(o[6] + x);  m[5] = wrap16(o[6] + y); m[6] = o[7];              m[7] = wrap16(o[7] + xy);
    n[0] = o[0];              n[1] = wrap16(o[0] + xy); n[2] = wrap16(o[1] + x);  n[3] = wrap16(o[1] + y);
    n[4] = wrap16(o[2] + y);  n[5] = wrap16(o[2] + x); n[6] = wrap16(o[3] + xy); n[7] = o[3];
  }
  for (int i = 0; i < 8; i++) o[i] = maxs(m[i], n[i]);
}

static void win_alpha_branches(const int16_t o[8], int16_t x, int16_t y, int16_t m[8], int16_t n[8])
{   /* turbodecoder_win.h:521-539: m = hypothesis bit 0, n = hypothesis bit 1 */
  int16_t xy = adds(x, y);
  m[0] = o[0];           m[1] = adds(o[3], y); m[2] = adds(o[4], y); m[3] = o[7];
  m[4] = o[1];           m[5] = adds(o[2], y); m[6] = adds(o[5], y); m[7] = o[6];
  n[0] = adds(o[1], xy); n[1] = adds(o[2], x); n[2] = adds(o[5], x); n[3] = adds(o[6], xy);
  n[4] = adds(o[0], xy); n[5] = adds(o[3], x); n[6] = adds(o[4], x); n[7] = adds(o[7], xy);
}

static inline void win_normalize(int k, int16_t o[8])
{   /* turbodecoder_win.h:244-261 with normalize_period 2 */
  if ((k % 2) == 0 && k != 0) {
    int hi = o[0], lo = o[0];
    for (int i = 1; i < 8; i++) { if (o[i] > hi) hi = o[i]; if (o[i] < lo) lo = o[i]; }
    if (hi - lo > orc_spread[orc_phase]) orc_spread[orc_phase] = hi - lo;
    for (int i = 1; i < 8; i++) o[i] = subs(o[i], o[0]);
    o[0] = 0;
    orc_track_mag(o);
  }
}

/* windowed max-log-MAP, P sub-blocks of L=K/P, natural indexing (lane d, step k <-> d*L+k).
 * Restates tdec_win{sse16,avx16}_dec = beta then alpha (turbodecoder_win.h:310-435,438-586). */
static void map_windowed(int K, int P, const int16_t *in, const int16_t *app, const int16_t *par,
                         const int16_t *tail_x, const int16_t *tail_p, int16_t *out)
{
  const int L = K / P, W = 40;
  static __thread int16_t X[ORC_MAX_K];
  static __thread int16_t B[16][ORC_MAX_K / 8 + 1][8];   /* beta store, (L+1) x 8 per lane; L <= 768 (K=6144,P=8 manual) */
  int16_t warm[16][8], o[8];
  for (int i = 0; i < K; i++) {
    X[i] = app ? adds(app[i], in[i]) : in[i];
    int ax = X[i] < 0 ? -X[i] : X[i], ay = par[i] < 0 ? -par[i] : par[i];
    if (ax > orc_peak_in) orc_peak_in = ax;
    if (ay > orc_peak_in) orc_peak_in = ay;
  }

  /* beta warm-up over the FIRST 40 symbols of every sub-block */
  orc_phase = 0;
  for (int d = 0; d < P; d++) {
    for (int i = 0; i < 8; i++) o[i] = NEG_INF;
    for (int k = W - 1; k >= 0; k--) { win_beta_step(o, X[d * L + k], par[d * L + k], 1); win_normalize(k, o); }
    memcpy(warm[d], o, sizeof(o));
  }
  orc_phase = 1;
  for (int d = 0; d < P; d++) {
    if (d < P - 1) memcpy(o, warm[d + 1], sizeof(o));
    else {   /* last lane: known end state through the 3 tail steps, no apriori, wrapping adds */
      o[0] = 0; for (int i = 1; i < 8; i++) o[i] = NEG_INF;
      for (int k = 2; k >= 0; k--) win_beta_step(o, tail_x[k], tail_p[k], 0);
    }
    orc_track_mag(o);
    memcpy(B[d][L], o, sizeof(o));
    for (int k = L - 1; k >= 0; k--) {
      win_beta_step(o, X[d * L + k], par[d * L + k], 1);
      memcpy(B[d][k], o, sizeof(o));     /* stored BEFORE normalisation */
      win_normalize(k, o);
    }
  }
  /* alpha warm-up over the LAST 40 symbols of every sub-block */
  orc_phase = 2;
  for (int d = 0; d < P; d++) {
    int16_t m[8], n[8];
    for (int i = 0; i < 8; i++) o[i] = NEG_INF;
    for (int k = 0; k < W; k++) {
      int idx = d * L + L - W + k;
      win_alpha_branches(o, X[idx], par[idx], m, n);
      for (int i = 0; i < 8; i++) o[i] = maxs(m[i], n[i]);
      win_normalize(k, o);
    }
    memcpy(warm[d], o, sizeof(o));
  }
  orc_phase = 3;
  for (int d = 0; d < P; d++) {
    int16_t m[8], n[8];
    if (d > 0) { memcpy(o, warm[d - 1], sizeof(o)); orc_track_mag(o); }
    else { o[0] = 0; for (int i = 1; i < 8; i++) o[i] = NEG_INF; }
    int vhi = -40000, vlo = 40000;
    for (int k = 0; k < L; k++) {
      int idx = d * L + k;
      win_alpha_branches(o, X[idx], par[idx], m, n);
      const int16_t *b = B[d][k + 1];
      int16_t m0 = adds(b[0], m[0]), m1 = adds(b[0], n[0]);
      for (int i = 1; i < 8; i++) { m0 = maxs(m0, adds(b[i], m[i])); m1 = maxs(m1, adds(b[i], n[i])); }
      if (m0 > vhi) vhi = m0;
      if (m1 > vhi) vhi = m1;
      if (m0 < vlo) vlo = m0;
      if (m1 < vlo) vlo = m1;
      if (vhi - vlo > orc_llr_range) orc_llr_range = vhi - vlo;
      int16_t r = subs(m1, m0);
      if (P == 8) r = (int16_t)(r >> 1);   /* divide_output exists for sse16 only (turbodecoder_win.h:61,565-567) */
      out[idx] = r;
      for (int i = 0; i < 8; i++) o[i] = maxs(m[i], n[i]);
      win_normalize(k, o);
    }
  }
}

/* ---- 8-bit windowed decoders tdec_win{sse8,avx8}_dec (turbodecoder_win.h:97-178 for the parameter
 * set): int8 saturating arithmetic, INF = 0 (all start states 0), normalisation after EVERY step k != 0 by the maximum of
 * the 8 states (normalize_max, normalize_period 1, :244-254), output (m1 - m0) >> 1 per byte (divide_output, :565-567),
 * tail trellis with the scalar add that saturates upwards only (:236-243). Values are int8 carried in int16. ---- */
static inline int16_t sat8(int v) { return (int16_t)(v > 127 ? 127 : (v < -128 ? -128 : v)); }
static inline int16_t adds8(int16_t a, int16_t b) { return sat8((int)a + (int)b); }
static inline int16_t subs8(int16_t a, int16_t b) { return sat8((int)a - (int)b); }
static inline int16_t sadd8_tail(int16_t x, int16_t y) { int z = x + y; return (int16_t)(z > 127 ? 127 : (int8_t)z); }

static void win8_beta_step(int16_t o[8], int16_t x, int16_t y, int tail)
{
  int16_t xy, m[8], n[8];
#define A8(a, b) (tail ? sadd8_tail(a, b) : adds8(a, b))
  xy = A8(x, y);
  m[0] = A8(o[4], xy); m[1] = o[4];         m[2] = A8(o[5], y);  m[3] = A8(o[5], x);
  m[4] = A8(o[6], x);  m[5] = A8(o[6], y);  m[6] = o[7];         m[7] = A8(o[7], xy);
  n[0] = o[0];         n[1] = A8(o[0], xy); n[2] = A8(o[1], x);  n[3] = A8(o[1], y);
  n[4] = A8(o[2], y);  n[5] = A8(o[2], x);  n[6] = A8(o[3], xy); n[7] = o[3];
#undef A8
  for (int i = 0; i < 8; i++) o[i] = maxs(m[i], n[i]);
}

static void win8_alpha_branches(const int16_t o[8], int16_t x, int16_t y, int16_t m[8], int16_t n[8])
{
  int16_t xy = adds8(x, y);
  m[0] = o[0];            m[1] = adds8(o[3], y); m[2] = adds8(o[4], y); m[3] = o[7];
  m[4] = o[1];            m[5] = adds8(o[2], y); m[6] = adds8(o[5], y); m[7] = o[6];
  n[0] = adds8(o[1], xy); n[1] = adds8(o[2], x); n[2] = adds8(o[5], x); n[3] = adds8(o[6], xy);
  n[4] = adds8(o[0], xy); n[5] = adds8(o[3], x); n[6] = adds8(o[4], x); n[7] = adds8(o[7], xy);
}

static inline void win8_normalize(int k, int16_t o[8])
{
  if (k != 0) {
    int16_t m = o[0];
    for (int i = 1; i < 8; i++) m = maxs(m, o[i]);
    for (int i = 0; i < 8; i++) o[i] = subs8(o[i], m);
  }
}

static void map_windowed8(int K, int P, const int16_t *in, const int16_t *app, const int16_t *par,
                          const int16_t *tail_x, const int16_t *tail_p, int16_t *out)
{
  const int L = K / P, W = 40;
  static __thread int16_t X[ORC_MAX_K];
  static __thread int16_t B[32][ORC_MAX_K / 16 + 1][8];
  int16_t warm[32][8], o[8], m[8], n[8];
  for (int i = 0; i < K; i++) X[i] = app ? adds8(app[i], in[i]) : in[i];
  for (int d = 0; d < P; d++) {                 /* beta warm-up over the first 40 symbols of every sub-block (:321-384) */
    memset(o, 0, sizeof(o));
    for (int k = W - 1; k >= 0; k--) { win8_beta_step(o, X[d * L + k], par[d * L + k], 0); win8_normalize(k, o); }
    memcpy(warm[d], o, sizeof(o));
  }
  for (int d = 0; d < P; d++) {
    if (d < P - 1) memcpy(o, warm[d + 1], sizeof(o));
    else { memset(o, 0, sizeof(o)); for (int k = 2; k >= 0; k--) win8_beta_step(o, tail_x[k], tail_p[k], 1); }
    memcpy(B[d][L], o, sizeof(o));
    for (int k = L - 1; k >= 0; k--) {
      win8_beta_step(o, X[d * L + k], par[d * L + k], 0);
      memcpy(B[d][k], o, sizeof(o));            /* stored before normalisation (:419-432) */
      win8_normalize(k, o);
    }
  }
  for (int d = 0; d < P; d++) {                 /* alpha warm-up over the last 40 symbols (:459-506) */
    memset(o, 0, sizeof(o));
    for (int k = 0; k < W; k++) {
      int idx = d * L + L - W + k;
      win8_alpha_branches(o, X[idx], par[idx], m, n);
      for (int i = 0; i < 8; i++) o[i] = maxs(m[i], n[i]);
      win8_normalize(k, o);
    }
    memcpy(warm[d], o, sizeof(o));
  }
  for (int d = 0; d < P; d++) {
    if (d > 0) memcpy(o, warm[d - 1], sizeof(o)); else memset(o, 0, sizeof(o));   /* known state: (0, -INF x 7) with INF = 0 */
    for (int k = 0; k < L; k++) {
      int idx = d * L + k;
      win8_alpha_branches(o, X[idx], par[idx], m, n);
      const int16_t *b = B[d][k + 1];
      int16_t m0 = adds8(b[0], m[0]), m1 = adds8(b[0], n[0]);
      for (int i = 1; i < 8; i++) { m0 = maxs(m0, adds8(b[i], m[i])); m1 = maxs(m1, adds8(b[i], n[i])); }
      out[idx] = (int16_t)(subs8(m1, m0) >> 1);
      for (int i = 0; i < 8; i++) o[i] = maxs(m[i], n[i]);
      win8_normalize(k, o);
    }
  }
}

/* non-windowed SSE decoder, wrapping int16 (turbodecoder_sse.c:105-206,211-297,300-353) */
static void map_sse(int K, const int16_t *in, const int16_t *app, const int16_t *par,
                    const int16_t *tail_x, const int16_t *tail_p, int16_t *out)
{
  static __thread int16_t g0[ORC_MAX_K + 3], g1[ORC_MAX_K + 3], A[ORC_MAX_K + 4][8];
  static const int SP[8] = {1, 2, 5, 6, 0, 3, 4, 7}, SN[8] = {0, 3, 4, 7, 1, 2, 5, 6};
  static const int BP[8] = {4, 0, 1, 5, 6, 2, 3, 7}, BN[8] = {0, 4, 5, 1, 2, 6, 7, 3};
  static const int GA[8] = {1, 1, 0, 0, 0, 0, 1, 1}, GB[8] = {1, 0, 0, 1, 1, 0, 0, 1};
  for (int k = 0; k < K; k++) {
    int16_t x = app ? wrap16(app[k] + in[k]) : in[k];
    g1[k] = (int16_t)(wrap16(x + par[k]) >> 1);
    g0[k] = (int16_t)(wrap16(x - par[k]) >> 1);
  }
  for (int k = 0; k < 3; k++) {   /* tail: C integer division, truncating (turbodecoder_sse.c:349-352) */
    g0[K + k] = (int16_t)(((int)tail_x[k] - (int)tail_p[k]) / 2);
    g1[K + k] = (int16_t)(((int)tail_x[k] + (int)tail_p[k]) / 2);
  }
  int16_t a[8], t[8];
  a[0] = 0; for (int i = 1; i < 8; i++) a[i] = NEG_INF;
  memcpy(A[0], a, sizeof(a));
  for (int k = 0; k < K; k++) {
    int16_t ap[8], an[8];
    for (int i = 0; i < 8; i++) { int16_t g = GA[i] ? g1[k] : g0[k]; ap[i] = wrap16(a[i] + g); an[i] = wrap16(a[i] - g); }
    for (int j = 0; j < 8; j++) t[j] = maxs(ap[SP[j]], an[SN[j]]);
    memcpy(a, t, sizeof(a));
    memcpy(A[k + 1], a, sizeof(a));          /* stored before normalisation */
    if ((k % 4) == 3) { int16_t z = a[0]; for (int i = 0; i < 8; i++) a[i] = wrap16(a[i] - z); }
  }
  int16_t b[8];
  b[0] = 0; for (int i = 1; i < 8; i++) b[i] = NEG_INF;
  int cnt = 0;
  for (int k = K + 2; k >= 0; k--) {
    int16_t bp[8], bn[8], sp[8], sn[8];
    for (int i = 0; i < 8; i++) { int16_t g = GB[i] ? g1[k] : g0[k]; bp[i] = wrap16(b[i] + g); bn[i] = wrap16(b[i] - g); }
    for (int j = 0; j < 8; j++) { sp[j] = bp[BP[j]]; sn[j] = bn[BN[j]]; b[j] = maxs(sp[j], sn[j]); }
    if (k < K) {
      int16_t mp = wrap16(sp[0] + A[k][0]), mn = wrap16(sn[0] + A[k][0]);
      for (int j = 1; j < 8; j++) { mp = maxs(mp, wrap16(sp[j] + A[k][j])); mn = maxs(mn, wrap16(sn[j] + A[k][j])); }
      out[k] = wrap16((int)mp - (int)mn);    /* hMax() returns 0x7FFF-max, so hMax(bn)-hMax(bp) = max(bp)-max(bn) (turbodecoder_sse.c:97-102,178) */
      if (++cnt == 4) { cnt = 0; int16_t z = b[0]; for (int i = 0; i < 8; i++) b[i] = wrap16(b[i] - z); }
    }
  }
}

void orc_map(int K, int P, const int16_t *x, const int16_t *app, const int16_t *par,
             const int16_t *tail_x, const int16_t *tail_p, int16_t *out)
{
  if (P == 0) map_sse(K, x, app, par, tail_x, tail_p, out);
  else map_windowed(K, P, x, app, par, tail_x, tail_p, out);
}

/* ------------------------------------------------------------------ half-iteration driver ---- */
static int tdec_setup(orc_tdec_t *h, int K)
{
  int idx = orc_cb_index(K);
  if (K > ORC_MAX_K || idx < 0 || orc_cb_size(idx) != K) return -1;
  h->K = K; h->P = orc_subblocks(K); h->n_iter = 0; h->llr8 = 0;
  return orc_qpp(K, h->fwd, h->rev);
}

/* srslte_vec_sub_bbb as the 8-bit iteration driver calls it (vector_simd.c:165-192, aligned buffers, AVX2): saturating in
 * blocks of 32 array elements, the K mod 32 elements behind the last full block by the scalar (wrapping) loop. The arrays
 * are in sub-block order: natural bit (d, k) sits at k*P + d. */
static void vec_sub8(const orc_tdec_t *h, int16_t *z, const int16_t *x, const int16_t *y)
{
  const int K = h->K, L = K / h->P, full = K - (K % 32);
  for (int d = 0; d < h->P; d++)
    for (int k = 0; k < L; k++) {
      int i = d * L + k, pos = k * h->P + d;
      z[i] = pos < full ? subs8(x[i], y[i]) : (int16_t)(int8_t)(x[i] - y[i]);
    }
}

static void tdec_tails(orc_tdec_t *h, const int16_t *t)
{   /* turbodecoder_iter.h:61-70 / turbodecoder_win.h:667-673 / turbodecoder_sse.c:486-492 */
  for (int i = 0; i < 3; i++) {
    h->t_syst[i] = t[2 * i]; h->t_p0[i] = t[2 * i + 1];
    h->t_app2[i] = t[6 + 2 * i]; h->t_p1[i] = t[6 + 2 * i + 1];
  }
}

int orc_tdec_new_cb_natural(orc_tdec_t *h, int K, const int16_t *in)
{
  if (tdec_setup(h, K)) return -1;
  for (int i = 0; i < K; i++) { h->syst[i] = in[3 * i]; h->p0[i] = in[3 * i + 1]; h->p1[i] = in[3 * i + 2]; }
  tdec_tails(h, in + 3 * K);
  return 0;
}

int orc_tdec_new_cb_soft(orc_tdec_t *h, int K, const int16_t *soft)
{
  if (tdec_setup(h, K)) return -1;
  if (h->P == 0) return orc_tdec_new_cb_natural(h, K, soft);
  for (int x = 0; x < K; x++) {
    int pos = orc_sb_pos(x, K, h->P);
    h->syst[x] = soft[pos]; h->p0[x] = soft[K + 32 + pos]; h->p1[x] = soft[2 * (K + 32) + pos];
  }
  tdec_tails(h, soft + 3 * (K + 32));
  return 0;
}

void orc_tdec_iteration(orc_tdec_t *h, uint8_t *out_bytes, int16_t *llr_nat)
{   /* run_tdec_iteration_16bit (turbodecoder_iter.h:73-147) + tdec_decision_byte (turbodecoder.c:353-360) */
  const int K = h->K, n = h->n_iter;
  if (h->llr8) {     /* run_tdec_iteration_8bit: same driver, int8 decoders and saturating vector subtraction */
    if ((n % 2) == 0) {
      if (n) vec_sub8(h, h->app1, h->app1, h->ext1);
      map_windowed8(K, h->P, h->syst, n ? h->app1 : NULL, h->p0, h->t_syst, h->t_p0, h->ext1);
    } else {
      if (n > 1) vec_sub8(h, h->ext1, h->ext1, h->app1);
      for (int i = 0; i < K; i++) h->app2[h->rev[i]] = h->ext1[i];
      map_windowed8(K, h->P, h->app2, NULL, h->p1, h->t_app2, h->t_p1, h->ext2);
      for (int i = 0; i < K; i++) h->app1[h->fwd[i]] = h->ext2[i];
    }
  } else if ((n % 2) == 0) {
    if (n) for (int i = 0; i < K; i++) h->app1[i] = wrap16(h->app1[i] - h->ext1[i]);
    orc_map(K, h->P, h->syst, n ? h->app1 : NULL, h->p0, h->t_syst, h->t_p0, h->ext1);
  } else {
    if (n > 1) for (int i = 0; i < K; i++) h->ext1[i] = wrap16(h->ext1[i] - h->app1[i]);
    for (int i = 0; i < K; i++) h->app2[h->rev[i]] = h->ext1[i];
    orc_map(K, h->P, h->app2, NULL, h->p1, h->t_app2, h->t_p1, h->ext2);
    for (int i = 0; i < K; i++) h->app1[h->fwd[i]] = h->ext2[i];
  }
  h->n_iter++;
  const int16_t *A = (h->n_iter % 2) ? h->ext1 : h->app1;
  if (out_bytes) {
    memset(out_bytes, 0, K / 8);
    for (int x = 0; x < K; x++) if (A[x] > 0) out_bytes[x >> 3] |= (uint8_t)(0x80u >> (x & 7));
  }
  if (llr_nat) memcpy(llr_nat, A, sizeof(int16_t) * K);
}

int orc_tdec_halfits(int K, const int16_t *in_nat, int nof, uint8_t *out_bytes, int16_t *llr_nat)
{
  orc_tdec_t *h = malloc(sizeof(*h));
  int r = orc_tdec_new_cb_natural(h, K, in_nat);
  for (int n = 0; n < nof && !r; n++)
    orc_tdec_iteration(h, out_bytes + (size_t)n * (K / 8), llr_nat ? llr_nat + (size_t)n * K : NULL);
  free(h);
  return r;
}

/* 8-bit LLR entry points (srslte_tdec_iteration_8bit, turbodecoder.c:408-452,536-543): sub-block classes 32 / 16 decode
 * in int8; classes 8 / 0 widen the input to int16 once (convert_8_to_16) and run the 16-bit decoders of that class. */
static void tdec8_class(orc_tdec_t *h)
{
  h->P = orc_subblocks8(h->K);
  h->llr8 = h->P >= 16;
}

int orc_tdec8_new_cb_natural(orc_tdec_t *h, int K, const int8_t *in)
{
  if (tdec_setup(h, K)) return -1;
  tdec8_class(h);
  int16_t t[12];
  for (int i = 0; i < K; i++) { h->syst[i] = in[3 * i]; h->p0[i] = in[3 * i + 1]; h->p1[i] = in[3 * i + 2]; }
  for (int i = 0; i < 12; i++) t[i] = in[3 * K + i];
  tdec_tails(h, t);
  return 0;
}

int orc_tdec8_new_cb_soft(orc_tdec_t *h, int K, const int8_t *soft)
{
  if (tdec_setup(h, K)) return -1;
  tdec8_class(h);
  if (h->P == 0) return orc_tdec8_new_cb_natural(h, K, soft);
  int16_t t[12];
  for (int x = 0; x < K; x++) {
    int pos = orc_sb_pos(x, K, h->P);
    h->syst[x] = soft[pos]; h->p0[x] = soft[K + 32 + pos]; h->p1[x] = soft[2 * (K + 32) + pos];
  }
  for (int i = 0; i < 12; i++) t[i] = soft[3 * (K + 32) + i];
  tdec_tails(h, t);
  return 0;
}

int orc_tdec_halfits8(int K, const int8_t *in_nat, int nof, uint8_t *out_bytes, int16_t *llr_nat)
{
  orc_tdec_t *h = malloc(sizeof(*h));
  int r = orc_tdec8_new_cb_natural(h, K, in_nat);
  for (int n = 0; n < nof && !r; n++)
    orc_tdec_iteration(h, out_bytes + (size_t)n * (K / 8), llr_nat ? llr_nat + (size_t)n * K : NULL);
  free(h);
  return r;
}

int orc_rm_rx8(const int8_t *in, int E, int K, int rv, int sb_layout, int8_t *soft)
{   /* soft[deinter[i % (3K+12)]] += e[i], wrapping int8, table of the 8-bit decoder class (rm_turbo.c:432-469,543-631) */
  static __thread uint16_t tab[3 * ORC_MAX_K + 12];
  if (rv < 0 || rv > 3 || orc_cb_index(K) < 0 || orc_cb_size(orc_cb_index(K)) != K) return -2;
  orc_rm_table(K, rv, tab);
  int P = sb_layout ? orc_subblocks8(K) : 0, n = 3 * K + 12;
  for (int i = 0; i < E; i++) {
    int pos = orc_rm_soft_index(tab[i % n], K, P);
    soft[pos] = (int8_t)(soft[pos] + in[i]);
  }
  return 0;
}

static int decode_cb_h(orc_tdec_t *h, int max_its, int crc_type, int crc_bits, uint8_t *out_bytes, int *nits)
{   /* sch.c:361-391 */
  int n = 0, ok = 0;
  do {
    orc_tdec_iteration(h, out_bytes, NULL);
    n++;
    if (!orc_crc(crc_type, out_bytes, crc_bits)) ok = 1;
  } while (n < max_its && !ok);
  *nits = n;
  return ok;
}

int orc_decode_cb(const int16_t *soft, int K, int max_its, int crc_type, uint8_t *out_bytes, int *nits)
{
  orc_tdec_t *h = malloc(sizeof(*h));
  if (orc_tdec_new_cb_soft(h, K, soft)) { free(h); return -1; }
  int ok = decode_cb_h(h, max_its, crc_type, K, out_bytes, nits);
  free(h);
  return ok;
}

/* ------------------------------------------------------------------ transport block ---- */
typedef struct { orc_softbuffer_t sb; orc_tdec_t dec; } orc_tb_t;

void *orc_tb_new(void) { orc_tb_t *t = calloc(1, sizeof(orc_tb_t)); if (t) t->sb.max_cb = ORC_MAX_CB; return t; }
void  orc_tb_free(void *p) { free(p); }

void orc_tb_reset(void *p, uint32_t tbs)
{
  orc_tb_t *t = p;
  int n = (int)((tbs + 24) / (6144 - 24) + 1);
  if (n > t->sb.max_cb) n = t->sb.max_cb;
  for (int i = 0; i < n; i++) memset(t->sb.soft[i], 0, sizeof(t->sb.soft[i]));
  memset(t->sb.cb_crc, 0, sizeof(t->sb.cb_crc));
}

void orc_tb_get_soft(void *p, int cb, int16_t *out) { memcpy(out, ((orc_tb_t *)p)->sb.soft[cb], sizeof(int16_t) * ORC_SOFT_SIZE); }

/* decode_tb + decode_tb_cb (sch.c:307-498). Returns 0 OK, -1 CRC fail, -2 invalid. data needs
 * tbs/8 + 3 bytes plus K/8 slack for the last CB's write. */
static int tb_decode_any(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const void *e_any, int llr8,
                         uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *cb_its)
{
  orc_tb_t *t = p;
  const int16_t *e_bits = e_any;
  const int8_t *e_bits_b = e_any;
  uint32_t s[9];
  if (!t || !data || !e_bits) return -2;
  if (orc_cbsegm(tbs, s)) return -2;
  const uint32_t F = s[0], C = s[1], K1 = s[2], K2 = s[3], C1 = s[6];
  if (tbs == 0 || C == 0) return 0;
  if (F) return -2;
  if (C > (uint32_t)t->sb.max_cb) return -2;
  data[tbs / 8] = 0; data[tbs / 8 + 1] = 0; data[tbs / 8 + 2] = 0;
  for (uint32_t cb = 0; cb < C; cb++) {
    uint32_t K = cb < C1 ? K1 : K2, rlen = C == 1 ? K : K - 24;
    if (cb_its) cb_its[cb] = 0;
    if (!t->sb.cb_crc[cb]) {
      uint32_t Gp = nof_e_bits / Qm, gamma = Gp % C, n_e = Qm * (Gp / C), rp = cb * n_e, E = n_e;
      if (cb > C - gamma) { E = n_e + Qm; rp = (C - gamma) * n_e + (cb - (C - gamma)) * E; }   /* sch.c:339 quirk */
      if (llr8) {      /* the soft buffer's memory holds int8 values (sch.c:344-348,362-364) */
        if (orc_rm_rx8(e_bits_b + rp, (int)E, (int)K, (int)rv, 1, (int8_t *)t->sb.soft[cb])) return -1;
        orc_tdec8_new_cb_soft(&t->dec, (int)K, (const int8_t *)t->sb.soft[cb]);
      } else {
        if (orc_rm_rx(e_bits + rp, (int)E, (int)K, (int)rv, 1, t->sb.soft[cb])) return -1;
        orc_tdec_new_cb_soft(&t->dec, (int)K, t->sb.soft[cb]);
      }
      int nits = 0;
      int ok = decode_cb_h(&t->dec, (int)max_its, C > 1 ? 1 : 0, C > 1 ? (int)K : (int)tbs + 24, &data[cb * rlen / 8], &nits);
      if (ok) t->sb.cb_crc[cb] = 1;
      if (cb_its) cb_its[cb] = (uint32_t)nits;
    } else {
      memcpy(&data[cb * rlen / 8], t->sb.data[cb], rlen / 8);
    }
  }
  t->sb.tb_crc = 1;
  for (uint32_t i = 0; i < C; i++) if (!t->sb.cb_crc[i]) { t->sb.tb_crc = 0; break; }
  if (!t->sb.tb_crc)
    for (uint32_t i = 0; i < C; i++) if (t->sb.cb_crc[i]) {
      uint32_t K = i < C1 ? K1 : K2, rlen = C == 1 ? K : K - 24;
      memcpy(t->sb.data[i], &data[i * rlen / 8], rlen / 8);
    }
  for (uint32_t i = 0; i < C && cb_crc_out; i++) cb_crc_out[i] = t->sb.cb_crc[i];
  if (!t->sb.tb_crc) return -1;
  uint32_t par_rx = orc_crc(0, data, (int)tbs);
  uint32_t par_tx = ((uint32_t)data[tbs / 8] << 16) | ((uint32_t)data[tbs / 8 + 1] << 8) | data[tbs / 8 + 2];
  return (par_rx == par_tx && par_rx) ? 0 : -1;
}

int orc_tb_decode(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const int16_t *e_bits,
                  uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *cb_its)
{
  return tb_decode_any(p, tbs, Qm, rv, nof_e_bits, e_bits, 0, max_its, data, cb_crc_out, cb_its);
}

/* the same with q->llr_is_8bit set: int8 LLRs, int8 soft buffers, srslte_tdec_iteration_8bit */
int orc_tb_decode8(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const int8_t *e_bits,
                   uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *cb_its)
{
  return tb_decode_any(p, tbs, Qm, rv, nof_e_bits, e_bits, 1, max_its, data, cb_crc_out, cb_its);
}

/* ------------------------------------------------------------------ multithreaded timing harness ---- */
typedef struct {
  int tid, nthreads, ntb, reps;
  uint32_t tbs, Qm, rv, nof_e_bits, max_its;
  const int16_t *e;
  pthread_barrier_t *bar;
  uint64_t n_tb, n_ok, its;
  struct timespec t0, t1;
} orc_mt_t;

/* optional per-TB results of the last pass, 3 words per TB: return value, sum of half-iterations / C (sch.c:420), FNV-1a of
 * the tbs/8 + 3 output bytes (NULL: off) */
static uint32_t *g_mt_results;
void orc_bench_set_results(uint32_t *buf) { g_mt_results = buf; }
static uint32_t orc_fnv1a(const uint8_t *p, size_t n) { uint32_t h = 2166136261u; for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; } return h; }

static void *orc_mt_worker(void *a_)
{
  orc_mt_t *a = a_;
  void *tb = orc_tb_new();
  uint8_t *data = malloc(a->tbs / 8 + 3 + 1024), crc[ORC_MAX_CB];
  uint32_t its[ORC_MAX_CB];
  pthread_barrier_wait(a->bar);
  clock_gettime(CLOCK_MONOTONIC, &a->t0);
  for (int r = 0; r < a->reps; r++)
    for (int i = a->tid; i < a->ntb; i += a->nthreads) {
      orc_tb_reset(tb, a->tbs);
      int ret = orc_tb_decode(tb, a->tbs, a->Qm, a->rv, a->nof_e_bits, a->e + (size_t)i * a->nof_e_bits, a->max_its, data, crc, its);
      a->n_tb++; if (ret == 0) a->n_ok++;
      uint32_t sum = 0, nz = 0;
      for (int c = 0; c < ORC_MAX_CB; c++) { a->its += its[c]; sum += its[c]; }
      if (g_mt_results) {
        uint32_t seg[9]; orc_cbsegm(a->tbs, seg); nz = seg[1] ? seg[1] : 1;
        g_mt_results[3 * i] = (uint32_t)ret; g_mt_results[3 * i + 1] = sum / nz; g_mt_results[3 * i + 2] = orc_fnv1a(data, a->tbs / 8 + 3);
      }
    }
  clock_gettime(CLOCK_MONOTONIC, &a->t1);
  free(data); orc_tb_free(tb);
  return NULL;
}

/* ------------------------------------------------------------------ transmit side ---- */
/* One constituent encoder of 36.212 5.1.3.2.1 (g0 = 1 + D^2 + D^3 feedback, g1 = 1 + D + D^3), restating the bitwise
 * srslte_tcod_encode (turbocoder.c:82-165): u[K] -> z[K], then the 3 trellis-termination steps -> xt[3], zt[3]. */
static void rsc_encode(const uint8_t *u, int K, uint8_t *z, uint8_t *xt, uint8_t *zt)
{
  int r1 = 0, r2 = 0, r3 = 0;
  for (int k = 0; k < K; k++) {
    int fb = u[k] ^ r2 ^ r3;
    z[k] = (uint8_t)(fb ^ r1 ^ r3);
    r3 = r2; r2 = r1; r1 = fb;
  }
  for (int k = 0; k < 3; k++) {            /* switch in the lower position: the input is the feedback itself */
    xt[k] = (uint8_t)(r2 ^ r3);
    zt[k] = (uint8_t)(r1 ^ r3);
    r3 = r2; r2 = r1; r1 = 0;
  }
}

/* bits[K] (one per byte) -> coded[3K+12] (one per byte) in the order srslte_tcod_encode emits:
 * [x0 z0 z'0 x1 z1 z'1 ...] then xK zK xK+1 zK+1 xK+2 zK+2 x'K z'K x'K+1 z'K+1 x'K+2 z'K+2 (turbocoder.c:82-165) */
int orc_encode_cb(int K, const uint8_t *bits, uint8_t *coded)
{
  static __thread uint8_t ui[ORC_MAX_K], z0[ORC_MAX_K], z1[ORC_MAX_K];
  static __thread uint16_t fwd[ORC_MAX_K], rev[ORC_MAX_K];
  uint8_t xt0[3], zt0[3], xt1[3], zt1[3];
  if (orc_qpp(K, fwd, rev)) return -1;
  for (int i = 0; i < K; i++) ui[i] = bits[fwd[i]];
  rsc_encode(bits, K, z0, xt0, zt0);
  rsc_encode(ui, K, z1, xt1, zt1);
  for (int i = 0; i < K; i++) { coded[3 * i] = bits[i]; coded[3 * i + 1] = z0[i]; coded[3 * i + 2] = z1[i]; }
  for (int k = 0; k < 3; k++) {
    coded[3 * K + 2 * k] = xt0[k]; coded[3 * K + 2 * k + 1] = zt0[k];
    coded[3 * K + 6 + 2 * k] = xt1[k]; coded[3 * K + 6 + 2 * k + 1] = zt1[k];
  }
  return 0;
}

/* encode_tb_off + srslte_tcod_encode_lut + srslte_rm_turbo_tx_lut (sch.c:188-296, turbocoder.c:196-330,
 * rm_turbo.c:332-376): data[tbs/8] -> e_bits, packed MSB first, (nof_e_bits+7)/8 bytes. Stateless: a retransmission
 * (rv != 0) re-derives the circular buffer the reference keeps in its tx soft buffer. 0 / -1 / -2 as the reference. */
int orc_tb_encode(uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const uint8_t *data, uint8_t *e_bits)
{
  uint32_t s[9];
  if (!data || !e_bits || rv > 3 || Qm == 0) return -2;
  if (orc_cbsegm(tbs, s)) return -2;
  const uint32_t F = s[0], C = s[1], K1 = s[2], K2 = s[3], C2 = s[7];
  if (F) return -1;                                  /* "filler bits are not supported" sch.c:202-205 */
  if (C > ORC_MAX_CB) return -1;
  memset(e_bits, 0, (nof_e_bits + 7) / 8);
  if (C == 0) return 0;
  static __thread uint8_t tb[ORC_MAX_CB * ORC_MAX_K / 8 + 8], cbb[ORC_MAX_K / 8];
  static __thread uint8_t bits[ORC_MAX_K], coded[3 * ORC_MAX_K + 12];
  static __thread uint16_t tab[3 * ORC_MAX_K + 12];
  const uint32_t par = orc_crc(0, data, (int)tbs);   /* TB CRC-24A, attached behind the data (36.212 5.3.2.1) */
  memcpy(tb, data, tbs / 8);
  tb[tbs / 8] = (uint8_t)(par >> 16); tb[tbs / 8 + 1] = (uint8_t)(par >> 8); tb[tbs / 8 + 2] = (uint8_t)par;
  const uint32_t Gp = nof_e_bits / Qm, gamma = Gp % C;
  uint32_t rp = 0, wp = 0;
  for (uint32_t i = 0; i < C; i++) {
    const uint32_t K = i < C2 ? K2 : K1, rlen = C > 1 ? K - 24 : K;
    const uint32_t n_e = (i <= C - gamma - 1) ? Qm * (Gp / C) : Qm * ((Gp + C - 1) / C);      /* sch.c:241-245 */
    memcpy(cbb, tb + rp / 8, rlen / 8);
    if (C > 1) {
      const uint32_t c = orc_crc(1, cbb, (int)rlen);                                          /* CB CRC-24B */
      cbb[rlen / 8] = (uint8_t)(c >> 16); cbb[rlen / 8 + 1] = (uint8_t)(c >> 8); cbb[rlen / 8 + 2] = (uint8_t)c;
    }
    for (uint32_t k = 0; k < K; k++) bits[k] = (cbb[k >> 3] >> (7 - (k & 7))) & 1u;
    if (orc_encode_cb((int)K, bits, coded)) return -1;
    orc_rm_table((int)K, (int)rv, tab);
    const uint32_t N = 3 * K + 12;
    for (uint32_t m = 0; m < n_e; m++) {
      const uint32_t o = wp + m;
      if (coded[tab[m % N]]) e_bits[o >> 3] |= (uint8_t)(0x80u >> (o & 7));
    }
    rp += rlen; wp += n_e;
  }
  return 0;
}

/* out3 = {TBs decoded, TBs OK, total CB half-iterations}; returns wall seconds */
double orc_bench_tb_mt(int nthreads, int ntb, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                       const int16_t *e_bits_all, uint32_t max_its, int reps, uint64_t *out3)
{
  pthread_t th[256]; orc_mt_t a[256]; pthread_barrier_t bar;
  if (nthreads > 256) nthreads = 256;
  pthread_barrier_init(&bar, NULL, nthreads);
  for (int t = 0; t < nthreads; t++) {
    a[t] = (orc_mt_t){.tid = t, .nthreads = nthreads, .ntb = ntb, .reps = reps, .tbs = tbs, .Qm = Qm, .rv = rv,
                      .nof_e_bits = nof_e_bits, .max_its = max_its, .e = e_bits_all, .bar = &bar};
    pthread_create(&th[t], NULL, orc_mt_worker, &a[t]);
  }
  double t0 = 1e300, t1 = 0;
  out3[0] = out3[1] = out3[2] = 0;
  for (int t = 0; t < nthreads; t++) {
    pthread_join(th[t], NULL);
    double s = a[t].t0.tv_sec + 1e-9 * a[t].t0.tv_nsec, e = a[t].t1.tv_sec + 1e-9 * a[t].t1.tv_nsec;
    if (s < t0) t0 = s;
    if (e > t1) t1 = e;
    out3[0] += a[t].n_tb; out3[1] += a[t].n_ok; out3[2] += a[t].its;
  }
  pthread_barrier_destroy(&bar);
  return t1 - t0;
}

/* ================================================================== front end ====
 * Soft demodulation, scrambling sequence, descrambling, CSI weighting and the UL-SCH channel de-interleaver: the
 * stages between the equaliser and decode_tb. Scalar restatement of what the reference's AVX2 build computes,
 * INCLUDING its quirks (SIMD body vs scalar tail rounding, swapped CSI weights inside a symbol pair, g[0]
 * overwritten by RI positions); checked bit for bit against oracle/_ref in tests/test_oracle_frontend.py. */

/* cvtps2dq / cvttps2dq followed by packssdw: out-of-range and NaN give the "integer indefinite" 0x80000000 -> -32768 */
static int16_t fe_cvt_sat16(float v, int truncate)
{
  if (!(v >= -2147483648.0f && v < 2147483648.0f)) return -32768;
  long r = truncate ? (long)truncf(v) : lrintf(v);          /* default rounding mode: nearest even */
  return (int16_t)(r > 32767 ? 32767 : r < -32768 ? -32768 : r);
}
/* C conversion float/double -> short as gcc emits it on x86-64: cvttss2si to 32 bits, low half kept */
static int16_t fe_cvt_c16(double v)
{
  if (!(v > -2147483649.0 && v < 2147483648.0)) return 0;
  return (int16_t)(int32_t)v;
}
static int16_t fe_abs16(int16_t v) { return (int16_t)(v < 0 ? -v : v); }   /* pabsw: |-32768| = -32768 */

/* srslte_demod_soft_demodulate_s (demod_soft.c:437-456) -> demod_{bpsk,qpsk,16qam,64qam}_lte_s (:56-60,72-74,98-138,206-219,
 * 238-301,372-390) */
int orc_demod_s(int mod_bits, const float *sym, int16_t *llr, int nsym)
{
  switch (mod_bits) {
  case 1:
    for (int i = 0; i < nsym; i++) {
      const float s = -100.0f * (sym[2 * i] + sym[2 * i + 1]);
      llr[i] = fe_cvt_c16((double)s / sqrt(2));
    }
    return 0;
  case 2: {
    /* srslte_vec_convert_fi_simd (vector_simd.c:394-429): 16 values per AVX2 step with truncation + saturation, scalar tail */
    const float scale = (float)(-100 * sqrt(2));
    const int len = 2 * nsym;
    int i = 0;
    for (; i < len - 16 + 1; i += 16)
      for (int k = 0; k < 16; k++) llr[i + k] = fe_cvt_sat16(sym[i + k] * scale, 1);
    for (; i < len; i++) llr[i] = fe_cvt_c16(sym[i] * scale);
    return 0;
  }
  case 4: {
    const int16_t off = (int16_t)(2 * 400 / sqrt(10));
    int i = 0;
    for (; i < 4 * (nsym / 4); i++) {
      const int16_t re = fe_cvt_sat16(sym[2 * i] * -400.0f, 0), im = fe_cvt_sat16(sym[2 * i + 1] * -400.0f, 0);
      llr[4 * i + 0] = re; llr[4 * i + 1] = im;
      llr[4 * i + 2] = (int16_t)(fe_abs16(re) - off); llr[4 * i + 3] = (int16_t)(fe_abs16(im) - off);
    }
    for (; i < nsym; i++) {
      const short yre = fe_cvt_c16(400 * sym[2 * i]), yim = fe_cvt_c16(400 * sym[2 * i + 1]);
      llr[4 * i + 0] = (int16_t)-yre; llr[4 * i + 1] = (int16_t)-yim;
      llr[4 * i + 2] = fe_cvt_c16(abs(yre) - 2 * 400 / sqrt(10));
      llr[4 * i + 3] = fe_cvt_c16(abs(yim) - 2 * 400 / sqrt(10));
    }
    return 0;
  }
  case 6: {
    const int16_t off1 = (int16_t)(4 * 700 / sqrt(42)), off2 = (int16_t)(2 * 700 / sqrt(42));
    int i = 0;
    for (; i < 4 * (nsym / 4); i++) {
      for (int c = 0; c < 2; c++) {
        const int16_t y = fe_cvt_sat16(sym[2 * i + c] * -700.0f, 0);
        const int16_t a1 = (int16_t)(fe_abs16(y) - off1);
        const int16_t a2 = (int16_t)(fe_abs16(a1) - off2);
        llr[6 * i + c] = y; llr[6 * i + 2 + c] = a1; llr[6 * i + 4 + c] = a2;
      }
    }
    for (; i < nsym; i++) {
      for (int c = 0; c < 2; c++) {
        const float y = fe_cvt_c16(700 * sym[2 * i + c]);
        llr[6 * i + c] = fe_cvt_c16(-y);
        llr[6 * i + 2 + c] = fe_cvt_c16(abs((int)y) - 4 * 700 / sqrt(42));
        llr[6 * i + 4 + c] = fe_cvt_c16(abs(llr[6 * i + 2 + c]) - 2 * 700 / sqrt(42));
      }
    }
    return 0;
  }
  default:
    return -1;
  }
}

/* srslte_sequence_set_LTE_pr (sequence.c:54-86), 36.211 7.2: c(n) = x1(n+1600) ^ x2(n+1600), x1(0)=1, x2 from c_init;
 * written with two 31-bit shift registers */
void orc_sequence(uint32_t c_init, uint32_t len, uint8_t *c)
{
  uint32_t x1 = 1, x2 = c_init & 0x7FFFFFFFu;
  for (uint32_t n = 0; n < 1600 + len; n++) {
    if (n >= 1600) c[n - 1600] = (uint8_t)((x1 ^ x2) & 1);
    const uint32_t f1 = ((x1 >> 3) ^ x1) & 1, f2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1;
    x1 = (x1 >> 1) | (f1 << 30);
    x2 = (x2 >> 1) | (f2 << 30);
  }
}

/* sequences.c:64-73 */
uint32_t orc_cinit_pdsch(uint32_t rnti, uint32_t q, uint32_t nslot, uint32_t cell_id) { return (rnti << 14) + (q << 13) + ((nslot / 2) << 9) + cell_id; }
uint32_t orc_cinit_pusch(uint32_t rnti, uint32_t nslot, uint32_t cell_id) { return (rnti << 14) + ((nslot / 2) << 9) + cell_id; }

/* srslte_scrambling_s_offset (scrambling.c:49-52) -> srslte_vec_neg_sss (vector_simd.c:223-250): psignw with +-1, wrapping */
void orc_descramble_s(const uint8_t *c, int16_t *llr, uint32_t len)
{
  for (uint32_t i = 0; i < len; i++) if (c[i]) llr[i] = (int16_t)-llr[i];
}

/* csi_correction, 16-bit branch of the SSE build (pdsch.c:675-776) */
void orc_csi_s(int mod_bits, int16_t *e, const float *csi, uint32_t nof_bits)
{
  const uint32_t qm = (uint32_t)mod_bits, nsym = nof_bits / qm;
  float csi_max = 1.0f;
  if (nsym) { csi_max = -INFINITY; for (uint32_t i = 0; i < nsym; i++) if (csi[i] > csi_max) csi_max = csi[i]; }  /* srslte_vec_max_fi */
  const float scale = 32767 / csi_max;
  uint32_t i = 0;        /* bit index */
#define CSI_W(s) fe_cvt_sat16(csi[s] * scale, 0)
#define CSI_MULHI(x, w) ((int16_t)(((int32_t)(x) * (int32_t)(w)) >> 16))
  if (qm == 2) {
    for (; nof_bits >= 4 && i < nof_bits - 3; i += 4) {   /* (the reference's unsigned bound underflows below 4 bits) */
      /* two symbols per step; the blend hands each symbol the OTHER one's weight */
      const uint32_t s = i / 2;
      const int16_t w1 = CSI_W(s), w2 = CSI_W(s + 1);
      e[i] = CSI_MULHI(e[i], w2); e[i + 1] = CSI_MULHI(e[i + 1], w2);
      e[i + 2] = CSI_MULHI(e[i + 2], w1); e[i + 3] = CSI_MULHI(e[i + 3], w1);
    }
  } else if (qm == 4) {
    for (; nof_bits >= 4 && i < nof_bits - 3; i += 4) {
      const int16_t w = CSI_W(i / 4);
      for (int k = 0; k < 4; k++) e[i + k] = CSI_MULHI(e[i + k], w);
    }
  } else if (qm == 6) {
    for (; nof_bits >= 12 && i < nof_bits - 11; i += 12) {  /* two symbols per step; the middle quad takes [w3 w3 w1 w1] */
      const uint32_t s = i / 6;
      const int16_t w1 = CSI_W(s), w3 = CSI_W(s + 1);
      for (int k = 0; k < 4; k++) e[i + k] = CSI_MULHI(e[i + k], w1);
      e[i + 4] = CSI_MULHI(e[i + 4], w3); e[i + 5] = CSI_MULHI(e[i + 5], w3);
      e[i + 6] = CSI_MULHI(e[i + 6], w1); e[i + 7] = CSI_MULHI(e[i + 7], w1);
      for (int k = 8; k < 12; k++) e[i + k] = CSI_MULHI(e[i + k], w3);
    }
  }
  i /= qm;
  for (; i < nsym; i++) {
    const float w = csi[i] / csi_max;
    for (uint32_t k = 0; k < qm; k++) e[qm * i + k] = fe_cvt_c16((float)e[qm * i + k] * w);
  }
#undef CSI_W
#undef CSI_MULHI
}

/* ---- 8-bit LLR front end (q->llr_is_8bit: pdsch.c:795-813, 707-714) ---------------------------------------------------- */
/* C conversion float/double -> int8_t as gcc emits it on x86-64: cvttss2si / cvttsd2si to 32 bits, low byte kept */
static int8_t fe_cvt_c8(double v)
{
  if (!(v > -2147483649.0 && v < 2147483648.0)) return 0;
  return (int8_t)(int32_t)v;
}
/* cvtps2dq (nearest even) or cvttps2dq, then packssdw and packsswb: saturation to int16, then to int8 */
static int8_t fe_cvt_sat8(float v, int truncate)
{
  const int16_t w = fe_cvt_sat16(v, truncate);
  return (int8_t)(w > 127 ? 127 : w < -128 ? -128 : w);
}
static int8_t fe_abs8(int8_t v) { return (int8_t)(v < 0 ? -v : v); }        /* pabsb: |-128| = -128 */

/* srslte_demod_soft_demodulate_b (demod_soft.c:458-478) -> demod_{bpsk,qpsk,16qam,64qam}_lte_b (:49-53,67-69,136-187,209-224,
 * 305-372,397-413): SSE bodies over 8 symbols (16 values for QPSK), scalar tails */
int orc_demod_b(int mod_bits, const float *sym, int8_t *llr, int nsym)
{
  switch (mod_bits) {
  case 1:
    for (int i = 0; i < nsym; i++) {
      const float s = (float)(int8_t)-20 * (sym[2 * i] + sym[2 * i + 1]);
      llr[i] = fe_cvt_c8((double)s / sqrt(2));
    }
    return 0;
  case 2: {
    /* srslte_vec_convert_fb_simd (vector_simd.c:433-496): 16 values per SSE step, truncation + two saturating packs */
    const float scale = (float)(-20 * sqrt(2));
    const int len = 2 * nsym;
    int i = 0;
    for (; i < len - 16 + 1; i += 16)
      for (int k = 0; k < 16; k++) llr[i + k] = fe_cvt_sat8(sym[i + k] * scale, 1);
    for (; i < len; i++) llr[i] = fe_cvt_c8(sym[i] * scale);
    return 0;
  }
  case 4: {
    const int8_t off = (int8_t)(2 * 30 / sqrt(10));
    int i = 0;
    for (; i < 8 * (nsym / 8); i++)
      for (int c = 0; c < 2; c++) {
        const int8_t y = fe_cvt_sat8(sym[2 * i + c] * -30.0f, 0);
        llr[4 * i + c] = y; llr[4 * i + 2 + c] = (int8_t)(fe_abs8(y) - off);
      }
    for (; i < nsym; i++)
      for (int c = 0; c < 2; c++) {
        const short y = fe_cvt_c8(30 * sym[2 * i + c]);
        llr[4 * i + c] = (int8_t)-y;
        llr[4 * i + 2 + c] = fe_cvt_c8(abs(y) - 2 * 30 / sqrt(10));
      }
    return 0;
  }
  case 6: {
    const int8_t off1 = (int8_t)(4 * 40 / sqrt(42)), off2 = (int8_t)(2 * 40 / sqrt(42));
    int i = 0;
    for (; i < 8 * (nsym / 8); i++)
      for (int c = 0; c < 2; c++) {
        const int8_t y = fe_cvt_sat8(sym[2 * i + c] * -40.0f, 0);
        const int8_t a1 = (int8_t)(fe_abs8(y) - off1);
        llr[6 * i + c] = y; llr[6 * i + 2 + c] = a1; llr[6 * i + 4 + c] = (int8_t)(fe_abs8(a1) - off2);
      }
    for (; i < nsym; i++)
      for (int c = 0; c < 2; c++) {
        const float y = fe_cvt_c8(40 * sym[2 * i + c]);
        llr[6 * i + c] = fe_cvt_c8(-y);
        llr[6 * i + 2 + c] = fe_cvt_c8(abs((int)y) - 4 * 40 / sqrt(42));
        llr[6 * i + 4 + c] = fe_cvt_c8(abs(llr[6 * i + 2 + c]) - 2 * 40 / sqrt(42));
      }
    return 0;
  }
  default:
    return -1;
  }
}

/* srslte_scrambling_sb_offset (scrambling.c:53-56) -> srslte_vec_neg_bbb (vector_simd.c:252-279): psignb with +-1, wrapping */
void orc_descramble_b(const uint8_t *c, int8_t *llr, uint32_t len)
{
  for (uint32_t i = 0; i < len; i++) if (c[i]) llr[i] = (int8_t)-llr[i];
}

/* csi_correction, 8-bit branch (pdsch.c:707-714): scalar throughout */
void orc_csi_b(int mod_bits, int8_t *e, const float *csi, uint32_t nof_bits)
{
  const uint32_t qm = (uint32_t)mod_bits, nsym = nof_bits / qm;
  float csi_max = 1.0f;
  if (nsym) { csi_max = -INFINITY; for (uint32_t i = 0; i < nsym; i++) if (csi[i] > csi_max) csi_max = csi[i]; }
  for (uint32_t i = 0; i < nsym; i++) {
    const float w = csi[i] / csi_max;
    for (uint32_t k = 0; k < qm; k++) e[qm * i + k] = fe_cvt_c8((float)e[qm * i + k] * w);
  }
}

/* ulsch_deinterleave (sch.c:860-881) = ulsch_interleave_gen (:550-568) + srslte_vec_lut_sis (vector.c:119-123): q is
 * column-major (cols SC-FDMA symbols x rows), g row-major without the RI positions; RI positions get lut 0, so g[0]
 * ends up holding the LAST RI position's value (scatter in ascending q order). g must be pre-filled by the caller if the
 * untouched tail matters (the reference leaves it as it was). */
void orc_ulsch_deinterleave(const int16_t *q, uint32_t Qm, uint32_t Hp, uint32_t cols, int16_t *g, const uint32_t *ri_pos, uint32_t nri)
{
  const uint32_t rows = Hp / cols, n = Hp * Qm;
  uint8_t *present = calloc(n + 1, 1);
  uint32_t *lut = calloc(n + 1, sizeof(uint32_t));
  for (uint32_t i = 0; i < nri; i++) present[ri_pos[i]] = 1;
  uint32_t idx = 0;
  for (uint32_t j = 0; j < rows; j++)
    for (uint32_t i = 0; i < cols; i++)
      for (uint32_t k = 0; k < Qm; k++) {
        const uint32_t p = (i * rows + j) * Qm + k;
        lut[p] = present[p] ? 0 : idx++;
      }
  for (uint32_t p = 0; p < n; p++) g[lut[p]] = q[p];
  free(present); free(lut);
}
