/*
 * ref_shim.c — thin ctypes-friendly driver over the UNMODIFIED reference objects (oracle/_ref only).
 *
 * TEST INFRASTRUCTURE. This file is compiled against the reference's own headers under
 * /root/reference/lib/include by oracle/Makefile (target `ref`) and linked with the reference's own
 * objects into oracle/_ref/libsrslte_ref.so. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load that library. Nothing here is product code.
 *
 * Every entry point calls the reference's public API:
 *   srslte_tdec_*            lib/include/srslte/phy/fec/turbodecoder.h:102-140
 *   srslte_rm_turbo_rx_lut_  lib/include/srslte/phy/fec/rm_turbo.h:51-94
 *   srslte_dlsch_encode/decode, srslte_sch_*   lib/include/srslte/phy/phch/sch.h:83-162
 *   srslte_crc_*, srslte_cbsegm*, srslte_softbuffer_rx_*, srslte_tcod_*
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <sched.h>
#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "srslte/phy/common/phy_common.h"
#include "srslte/phy/fec/cbsegm.h"
#include "srslte/phy/fec/crc.h"
#include "srslte/phy/fec/rm_turbo.h"
#include "srslte/phy/fec/softbuffer.h"
#include "srslte/phy/fec/tc_interl.h"
#include "srslte/phy/fec/turbocoder.h"
#include "srslte/phy/fec/turbodecoder.h"
#include "srslte/phy/phch/pdsch_cfg.h"
#include "srslte/phy/phch/sch.h"
#include "srslte/phy/utils/bit.h"
#include "srslte/phy/utils/vector.h"

extern bool decode_tb_cb(srslte_sch_t *q, srslte_softbuffer_rx_t *softbuffer, srslte_cbsegm_t *cb_segm,
                         uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, void *e_bits, uint8_t *data);

static srslte_tdec_t g_tdec;      /* AUTO decoder for the direct tdec entry points */
static srslte_tcod_t g_tcod;
static srslte_crc_t  g_crc[2];    /* 0: CRC24A, 1: CRC24B */
static int           g_init = 0;

int ref_init(void)
{
  if (g_init) return 0;
  srslte_rm_turbo_gentables();
  if (srslte_tdec_init(&g_tdec, 6144)) return -1;
  if (srslte_tcod_init(&g_tcod, 6144)) return -1;
  if (srslte_crc_init(&g_crc[0], SRSLTE_LTE_CRC24A, 24)) return -1;
  if (srslte_crc_init(&g_crc[1], SRSLTE_LTE_CRC24B, 24)) return -1;
  g_init = 1;
  return 0;
}

int ref_cb_index(int K) { return srslte_cbsegm_cbindex((uint32_t)K); }
int ref_cb_size(int idx) { return srslte_cbsegm_cbsize((uint32_t)idx); }
int ref_subblocks(int K) { return (int)srslte_tdec_autoimp_get_subblocks((uint32_t)K); }

/* out[9] = {F,C,K1,K2,K1_idx,K2_idx,C1,C2,tbs} */
int ref_cbsegm(uint32_t tbs, uint32_t *out)
{
  srslte_cbsegm_t s;
  memset(&s, 0, sizeof(s));
  int r = srslte_cbsegm(&s, tbs);
  out[0] = s.F; out[1] = s.C; out[2] = s.K1; out[3] = s.K2; out[4] = s.K1_idx; out[5] = s.K2_idx;
  out[6] = s.C1; out[7] = s.C2; out[8] = s.tbs;
  return r;
}

uint32_t ref_crc(int type, const uint8_t *bytes, int nbits)
{
  return srslte_crc_checksum_byte(&g_crc[type ? 1 : 0], (uint8_t *)bytes, nbits);
}

/* QPP tables exactly as the decoder object holds them (P = 1, 8, 16) */
int ref_interleaver(int K, int P, uint16_t *fwd, uint16_t *rev)
{
  srslte_tc_interl_t t;
  if (srslte_tc_interl_init(&t, (uint32_t)K)) return -1;
  int r = srslte_tc_interl_LTE_gen_interl(&t, (uint32_t)K, (uint32_t)(P ? P : 1));
  memcpy(fwd, t.forward, sizeof(uint16_t) * K);
  memcpy(rev, t.reverse, sizeof(uint16_t) * K);
  srslte_tc_interl_free(&t);
  return r;
}

/* rate de-matching + combine into a caller-held 18600-entry soft buffer */
int ref_rm_rx(const int16_t *in, int E, int cb_idx, int rv, int sb_layout, int16_t *soft)
{
  return srslte_rm_turbo_rx_lut_((int16_t *)in, soft, (uint32_t)E, (uint32_t)cb_idx, (uint32_t)rv, sb_layout != 0);
}

/* Turbo-encode one code block: bits_in[K] (one bit per byte) -> coded[3K+12] (one bit per byte),
 * order [s0 p0 p'0 s1 p1 p'1 ...] + 12 tail, as srslte_tcod_encode emits it. */
int ref_encode_cb(int K, const uint8_t *bits_in, uint8_t *coded)
{
  return srslte_tcod_encode(&g_tcod, (uint8_t *)bits_in, coded, (uint32_t)K);
}

/* decision source in natural order after the current half-iteration */
static void llr_natural(srslte_tdec_t *h, int K, int16_t *out)
{
  int16_t *A = (h->n_iter % 2) ? (int16_t *)h->ext1 : (int16_t *)h->app1;
  int P = 0;
  if (h->dec_type == SRSLTE_TDEC_AUTO) P = (int)srslte_tdec_autoimp_get_subblocks((uint32_t)K);
  else P = h->nof_blocks16[0] > 1 ? h->nof_blocks16[0] : 0;
  if (P == 0) { memcpy(out, A, sizeof(int16_t) * K); return; }
  int L = K / P;
  for (int x = 0; x < K; x++) out[x] = A[(x % L) * P + x / L];
}

/* Run nof half-iterations through srslte_tdec_new_cb + srslte_tdec_iteration on a NATURAL-layout
 * input (3K+12 int16, as turbodecoder_test feeds it: force_not_sb). mode 0 = AUTO, else a
 * srslte_tdec_impl_type_t for srslte_tdec_init_manual. Snapshots per half-iteration:
 * out_bytes[nof][K/8], llr_nat[nof][K] (may be NULL). */
int ref_tdec_halfits(int K, const int16_t *in_nat, int nof, int mode, uint8_t *out_bytes, int16_t *llr_nat)
{
  srslte_tdec_t local, *h = &g_tdec;
  if (mode != 0) {
    if (srslte_tdec_init_manual(&local, 6144, (srslte_tdec_impl_type_t)mode)) return -1;
    h = &local;
  }
  srslte_tdec_force_not_sb(h);
  int16_t *in = srslte_vec_malloc(sizeof(int16_t) * (3 * K + 12 + 64));
  memcpy(in, in_nat, sizeof(int16_t) * (3 * K + 12));
  int r = srslte_tdec_new_cb(h, (uint32_t)K);
  if (!r) {
    for (int n = 0; n < nof; n++) {
      srslte_tdec_iteration(h, in, out_bytes + (size_t)n * (K / 8));
      if (llr_nat) llr_natural(h, K, llr_nat + (size_t)n * K);
    }
  }
  free(in);
  h->force_not_sb = false;
  if (mode != 0) srslte_tdec_free(&local);
  return r;
}

/* n code blocks of one size through ONE decoder object (mode as above: 0 AUTO, SRSLTE_TDEC_GENERIC = 1 the full-length
 * int16 max-log-MAP of turbodecoder_gen.c:59-203, ...), nof half-iterations each via srslte_tdec_run_all; natural-layout
 * inputs in_all[n][3K+12], decisions out_bytes[n][K/8]. For BER/BLER statistics (tests/golden/bler_gen.npz). */
int ref_tdec_batch(int K, const int16_t *in_all, int n, int nof, int mode, uint8_t *out_bytes)
{
  srslte_tdec_t local, *h = &g_tdec;
  if (mode != 0) {
    if (srslte_tdec_init_manual(&local, 6144, (srslte_tdec_impl_type_t)mode)) return -1;
    h = &local;
  }
  srslte_tdec_force_not_sb(h);
  int16_t *in = srslte_vec_malloc(sizeof(int16_t) * (3 * K + 12 + 64));
  int r = 0;
  for (int i = 0; i < n && !r; i++) {
    memcpy(in, in_all + (size_t)i * (3 * K + 12), sizeof(int16_t) * (3 * K + 12));
    r = srslte_tdec_run_all(h, in, out_bytes + (size_t)i * (K / 8), (uint32_t)nof, (uint32_t)K);
  }
  free(in);
  h->force_not_sb = false;
  if (mode != 0) srslte_tdec_free(&local);
  return r;
}

/* ---- 8-bit LLR path ---- */
int ref_subblocks8(int K) { return (int)srslte_tdec_autoimp_get_subblocks_8bit((uint32_t)K); }

int ref_rm_rx8(const int8_t *in, int E, int cb_idx, int rv, int8_t *soft)
{
  return srslte_rm_turbo_rx_lut_8bit((int8_t *)in, soft, (uint32_t)E, (uint32_t)cb_idx, (uint32_t)rv);
}

/* nof half-iterations through srslte_tdec_new_cb + srslte_tdec_iteration_8bit on a NATURAL-layout int8 input
 * (force_not_sb). Snapshots per half-iteration: out_bytes[nof][K/8], llr_nat[nof][K] (int16, may be NULL). */
int ref_tdec_halfits8(int K, const int8_t *in_nat, int nof, uint8_t *out_bytes, int16_t *llr_nat)
{
  srslte_tdec_t *h = &g_tdec;
  srslte_tdec_force_not_sb(h);
  int8_t *in = srslte_vec_malloc(3 * K + 12 + 64);
  memcpy(in, in_nat, 3 * K + 12);
  int r = srslte_tdec_new_cb(h, (uint32_t)K);
  if (!r) {
    for (int n = 0; n < nof; n++) {
      srslte_tdec_iteration_8bit(h, in, out_bytes + (size_t)n * (K / 8));
      if (llr_nat) {
        int16_t *o = llr_nat + (size_t)n * K;
        if (h->current_llr_type == SRSLTE_TDEC_16) llr_natural(h, K, o);
        else {
          int8_t *A = (h->n_iter % 2) ? (int8_t *)h->ext1 : (int8_t *)h->app1;
          int P = ref_subblocks8(K), L = K / P;
          for (int x = 0; x < K; x++) o[x] = A[(x % L) * P + x / L];
        }
      }
    }
  }
  free(in);
  h->force_not_sb = false;
  return r;
}

/* Decode ONE code block from a soft buffer as sch.c:356-391 does (sb layout when P>0), CRC early stop.
 * crc_type 0 = CRC24A, 1 = CRC24B. soft is modified in place exactly as the reference does (trap 7).
 * Returns 1 if CRC OK else 0; *nits = half-iterations run. */
int ref_decode_cb(int16_t *soft, int K, int max_its, int crc_type, uint8_t *out_bytes, int *nits)
{
  int16_t *al = srslte_vec_malloc(sizeof(int16_t) * SOFTBUFFER_SIZE);   /* the decoder uses aligned SIMD loads */
  memcpy(al, soft, sizeof(int16_t) * SOFTBUFFER_SIZE);
  srslte_tdec_new_cb(&g_tdec, (uint32_t)K);
  int n = 0, ok = 0;
  do {
    srslte_tdec_iteration(&g_tdec, al, out_bytes);
    n++;
    if (!srslte_crc_checksum_byte(&g_crc[crc_type ? 1 : 0], out_bytes, K)) ok = 1;
  } while (n < max_its && !ok);
  *nits = n;
  memcpy(soft, al, sizeof(int16_t) * SOFTBUFFER_SIZE);
  free(al);
  return ok;
}

/* ------------------------------------------------------------------ transport-block level ---- */

typedef struct {
  srslte_sch_t           sch;
  srslte_softbuffer_rx_t rx;
  srslte_softbuffer_tx_t tx;
  srslte_pdsch_cfg_t     cfg;
  uint8_t               *packed;   /* packed e_bits scratch for the encoder */
  uint32_t               cb_its[SRSLTE_MAX_CODEBLOCKS];
} ref_tb_t;

void *ref_tb_new(int nof_prb)
{
  ref_tb_t *t = calloc(1, sizeof(ref_tb_t));
  if (!t) return NULL;
  if (srslte_sch_init(&t->sch)) { free(t); return NULL; }
  if (srslte_softbuffer_rx_init(&t->rx, (uint32_t)nof_prb)) { free(t); return NULL; }
  if (srslte_softbuffer_tx_init(&t->tx, (uint32_t)nof_prb)) { free(t); return NULL; }
  t->packed = srslte_vec_malloc(200000);
  return t;
}

void ref_tb_free(void *p)
{
  ref_tb_t *t = p;
  if (!t) return;
  /* NB: srslte_sch_free() tears down process-global RM tables (sch.c:157); regenerate lazily */
  srslte_sch_free(&t->sch);
  srslte_rm_turbo_gentables();
  srslte_softbuffer_rx_free(&t->rx);
  srslte_softbuffer_tx_free(&t->tx);
  free(t->packed);
  free(t);
}

int ref_tb_max_cb(void *p) { return (int)((ref_tb_t *)p)->rx.max_cb; }

/* new transmission: zero soft buffers + cb_crc for this TBS (softbuffer.c:128-151) */
void ref_tb_reset(void *p, uint32_t tbs) { srslte_softbuffer_rx_reset_tbs(&((ref_tb_t *)p)->rx, tbs); }

static int fill_cfg(ref_tb_t *t, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits)
{
  memset(&t->cfg, 0, sizeof(t->cfg));
  if (srslte_cbsegm(&t->cfg.cb_segm[0], tbs)) return -1;
  t->cfg.grant.tb_en[0] = true;
  t->cfg.grant.Qm[0] = Qm;
  t->cfg.grant.mcs[0].tbs = (int)tbs;
  t->cfg.nof_layers = 1;
  t->cfg.rv[0] = rv;
  t->cfg.nbits[0].nof_bits = nof_e_bits;
  return 0;
}

/* data: tbs/8 bytes (+3 scratch bytes for the TB CRC the encoder appends). e_bits: one bit per byte. */
int ref_tb_encode(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, uint8_t *data,
                  uint8_t *e_bits_unpacked)
{
  ref_tb_t *t = p;
  if (fill_cfg(t, tbs, Qm, rv, nof_e_bits)) return -1;
  memset(t->packed, 0, nof_e_bits / 8 + 8);
  int r = srslte_dlsch_encode(&t->sch, &t->cfg, &t->tx, data, t->packed);
  srslte_bit_unpack_vector(t->packed, e_bits_unpacked, (int)nof_e_bits);
  return r;
}

/* srslte_dlsch_decode as the apps call it. data needs tbs/8 + 3 + 768 bytes. cb_crc_out[C]. */
int ref_tb_decode(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, int16_t *e_bits,
                  uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *avg_its)
{
  ref_tb_t *t = p;
  if (fill_cfg(t, tbs, Qm, rv, nof_e_bits)) return -2;
  srslte_sch_set_max_noi(&t->sch, max_its);
  int r = srslte_dlsch_decode(&t->sch, &t->cfg, &t->rx, e_bits, data);
  for (uint32_t i = 0; i < t->cfg.cb_segm[0].C && cb_crc_out; i++) cb_crc_out[i] = t->rx.cb_crc[i];
  if (avg_its) *avg_its = srslte_sch_last_noi(&t->sch);
  return r;
}

/* srslte_dlsch_decode with q->llr_is_8bit set: int8 e_bits, the soft buffers hold int8 */
int ref_tb_decode8(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, int8_t *e_bits,
                   uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *avg_its)
{
  ref_tb_t *t = p;
  if (fill_cfg(t, tbs, Qm, rv, nof_e_bits)) return -2;
  srslte_sch_set_max_noi(&t->sch, max_its);
  t->sch.llr_is_8bit = true;
  int r = srslte_dlsch_decode(&t->sch, &t->cfg, &t->rx, (int16_t *)e_bits, data);
  t->sch.llr_is_8bit = false;
  for (uint32_t i = 0; i < t->cfg.cb_segm[0].C && cb_crc_out; i++) cb_crc_out[i] = t->rx.cb_crc[i];
  if (avg_its) *avg_its = srslte_sch_last_noi(&t->sch);
  return r;
}

/* Same contract, but the per-CB loop of sch.c:307-422 is driven here from the reference's own
 * primitives so that the per-CB half-iteration counts (which the reference only averages) are visible.
 * Used to pin per-CB iteration parity; cross-checked against ref_tb_decode in tests. */
int ref_tb_decode_detail(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, int16_t *e_bits,
                         uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *cb_its)
{
  ref_tb_t *t = p;
  if (fill_cfg(t, tbs, Qm, rv, nof_e_bits)) return -2;
  srslte_cbsegm_t *s = &t->cfg.cb_segm[0];
  if (s->F) return -2;
  if (s->C > t->rx.max_cb) return -2;
  data[tbs / 8 + 0] = 0; data[tbs / 8 + 1] = 0; data[tbs / 8 + 2] = 0;
  for (uint32_t cb = 0; cb < s->C; cb++) {
    uint32_t K = cb < s->C1 ? s->K1 : s->K2, Kidx = cb < s->C1 ? s->K1_idx : s->K2_idx;
    uint32_t rlen = s->C == 1 ? K : K - 24;
    cb_its[cb] = 0;
    if (!t->rx.cb_crc[cb]) {
      uint32_t Gp = nof_e_bits / Qm, gamma = Gp % s->C, n_e = Qm * (Gp / s->C);
      uint32_t rp = cb * n_e, E = n_e;
      if (cb > s->C - gamma) { E = n_e + Qm; rp = (s->C - gamma) * n_e + (cb - (s->C - gamma)) * E; }
      srslte_rm_turbo_rx_lut(&e_bits[rp], t->rx.buffer_f[cb], E, Kidx, rv);
      srslte_tdec_new_cb(&t->sch.decoder, K);
      bool ok = false; uint32_t n = 0;
      do {
        srslte_tdec_iteration(&t->sch.decoder, t->rx.buffer_f[cb], &data[cb * rlen / 8]);
        n++;
        uint32_t c = s->C > 1 ? srslte_crc_checksum_byte(&t->sch.crc_cb, &data[cb * rlen / 8], K)
                              : srslte_crc_checksum_byte(&t->sch.crc_tb, &data[cb * rlen / 8], tbs + 24);
        if (!c) { t->rx.cb_crc[cb] = true; ok = true; }
      } while (n < max_its && !ok);
      cb_its[cb] = n;
    } else {
      memcpy(&data[cb * rlen / 8], t->rx.data[cb], rlen / 8);
    }
  }
  t->rx.tb_crc = true;
  for (uint32_t i = 0; i < s->C && t->rx.tb_crc; i++) t->rx.tb_crc = t->rx.cb_crc[i];
  if (!t->rx.tb_crc) {
    for (uint32_t i = 0; i < s->C; i++) if (t->rx.cb_crc[i]) {
      uint32_t K = i < s->C1 ? s->K1 : s->K2, rlen = s->C == 1 ? K : K - 24;
      memcpy(t->rx.data[i], &data[i * rlen / 8], rlen / 8);
    }
  }
  for (uint32_t i = 0; i < s->C; i++) cb_crc_out[i] = t->rx.cb_crc[i];
  if (!t->rx.tb_crc) return -1;
  uint32_t par_rx = srslte_crc_checksum_byte(&t->sch.crc_tb, data, tbs);
  uint32_t par_tx = ((uint32_t)data[tbs / 8] << 16) | ((uint32_t)data[tbs / 8 + 1] << 8) | data[tbs / 8 + 2];
  return (par_rx == par_tx && par_rx) ? 0 : -1;
}

void ref_tb_get_soft(void *p, int cb, int16_t *out) { memcpy(out, ((ref_tb_t *)p)->rx.buffer_f[cb], 18600 * sizeof(int16_t)); }
void ref_tb_set_soft(void *p, int cb, const int16_t *in) { memcpy(((ref_tb_t *)p)->rx.buffer_f[cb], in, 18600 * sizeof(int16_t)); }

/* ------------------------------------------------------------------ multithreaded CPU baseline ---- */
/* The reference's own decode path (srslte_dlsch_decode, AVX2 AUTO decoders), one srslte_sch_t + private
 * soft buffers per thread, TBs statically partitioned. */
typedef struct {
  int tid, nthreads, ntb, reps;
  uint32_t tbs, Qm, rv, nof_e_bits, max_its;
  const int16_t *e_bits_all;
  ref_tb_t *tb;
  pthread_barrier_t *bar;
  uint64_t n_ok, n_tb, sum_avg_its;
  struct timespec t0, t1;
} mt_arg_t;

/* optional per-TB results of the last pass, 3 words per TB: decode_tb return value, srslte_sch_last_noi, FNV-1a of the
 * tbs/8 + 3 output bytes — so that a bench can compare what it decoded with what the reference decodes (NULL: off) */
static uint32_t *g_mt_results;
void ref_bench_set_results(uint32_t *buf) { g_mt_results = buf; }
static uint32_t fnv1a(const uint8_t *p, size_t n) { uint32_t h = 2166136261u; for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; } return h; }

static void *mt_worker(void *a_)
{
  mt_arg_t *a = a_;
  cpu_set_t set; CPU_ZERO(&set); CPU_SET(a->tid, &set);
  pthread_setaffinity_np(pthread_self(), sizeof(set), &set);   /* best effort */
  uint8_t *data = malloc(a->tbs / 8 + 3 + 1024);
  int16_t *e = srslte_vec_malloc(sizeof(int16_t) * (a->nof_e_bits + 64));
  uint8_t crc[SRSLTE_MAX_CODEBLOCKS]; uint32_t avg;
  pthread_barrier_wait(a->bar);
  clock_gettime(CLOCK_MONOTONIC, &a->t0);
  for (int r = 0; r < a->reps; r++) {
    for (int i = a->tid; i < a->ntb; i += a->nthreads) {
      memcpy(e, a->e_bits_all + (size_t)i * a->nof_e_bits, sizeof(int16_t) * a->nof_e_bits);
      srslte_softbuffer_rx_reset_tbs(&a->tb->rx, a->tbs);
      int ret = ref_tb_decode(a->tb, a->tbs, a->Qm, a->rv, a->nof_e_bits, e, a->max_its, data, crc, &avg);
      a->n_tb++; a->sum_avg_its += avg;
      if (ret == 0) a->n_ok++;
      if (g_mt_results) { g_mt_results[3 * i] = (uint32_t)ret; g_mt_results[3 * i + 1] = avg; g_mt_results[3 * i + 2] = fnv1a(data, a->tbs / 8 + 3); }
    }
  }
  clock_gettime(CLOCK_MONOTONIC, &a->t1);
  free(data); free(e);
  return NULL;
}

/* Returns wall seconds (latest finish - earliest start); out[0]=TBs decoded, out[1]=TBs CRC-OK,
 * out[2]=sum over TBs of the reference's (integer) mean half-iterations per CB. */
double ref_bench_tb_mt(int nthreads, int ntb, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                       const int16_t *e_bits_all, uint32_t max_its, int reps, int nof_prb, uint64_t *out)
{
  pthread_t *th = calloc(nthreads, sizeof(pthread_t));
  mt_arg_t *args = calloc(nthreads, sizeof(mt_arg_t));
  pthread_barrier_t bar;
  pthread_barrier_init(&bar, NULL, nthreads);
  for (int t = 0; t < nthreads; t++) {
    args[t] = (mt_arg_t){.tid = t, .nthreads = nthreads, .ntb = ntb, .reps = reps, .tbs = tbs, .Qm = Qm, .rv = rv,
                         .nof_e_bits = nof_e_bits, .max_its = max_its, .e_bits_all = e_bits_all, .bar = &bar};
    args[t].tb = ref_tb_new(nof_prb);
  }
  for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, mt_worker, &args[t]);
  for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
  double t0 = 1e300, t1 = 0;
  out[0] = out[1] = out[2] = 0;
  for (int t = 0; t < nthreads; t++) {
    double a = args[t].t0.tv_sec + 1e-9 * args[t].t0.tv_nsec, b = args[t].t1.tv_sec + 1e-9 * args[t].t1.tv_nsec;
    if (a < t0) t0 = a;
    if (b > t1) t1 = b;
    out[0] += args[t].n_tb; out[1] += args[t].n_ok; out[2] += args[t].sum_avg_its;
  }
  for (int t = 0; t < nthreads; t++) ref_tb_free(args[t].tb);
  pthread_barrier_destroy(&bar);
  free(th); free(args);
  return t1 - t0;
}
