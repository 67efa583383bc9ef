/*
 * ref_shim_fe.c — driver over the UNMODIFIED reference for the stages right in front of decode_tb
 * (SURVEY.md 8f rows f1/f2): soft demodulation, scrambling sequence + descrambling, CSI weighting, the whole
 * PDSCH codeword decode, the UL-SCH channel de-interleaver and srslte_ulsch_uci_decode.
 *
 * TEST INFRASTRUCTURE (oracle/_ref only). The reference's pdsch.c is compiled into this translation unit from
 * where it lies (#include below, nothing copied) because csi_correction() and srslte_pdsch_codeword_decode() are
 * file-static there (pdsch.c:675-776, 778-834).
 */
#define _GNU_SOURCE
#include "phch/pdsch.c"          /* -I$(REF)/lib/src/phy: the reference source itself */

#include "srslte/phy/modem/demod_soft.h"
#include "srslte/phy/phch/pusch_cfg.h"
#include "srslte/phy/phch/uci.h"
#include "srslte/phy/scrambling/scrambling.h"

extern void ulsch_deinterleave(int16_t *q_bits, uint32_t Qm, uint32_t H_prime_total, uint32_t N_pusch_symbs, int16_t *g_bits,
                               srslte_uci_bit_t *ri_bits, uint32_t nof_ri_bits, uint8_t *ri_present, uint32_t *inteleaver_lut);

static srslte_mod_t mod_of(int bits)
{
  return bits == 1 ? SRSLTE_MOD_BPSK : bits == 2 ? SRSLTE_MOD_QPSK : bits == 4 ? SRSLTE_MOD_16QAM : SRSLTE_MOD_64QAM;
}

/* srslte_demod_soft_demodulate_s (demod_soft.c:437-456); sym must be 32-byte aligned like srslte_vec_malloc memory */
int ref_demod_s(int mod_bits, const float *sym, int16_t *llr, int nsym)
{
  return srslte_demod_soft_demodulate_s(mod_of(mod_bits), (const cf_t *)sym, llr, nsym);
}

/* srslte_sequence_LTE_pr (sequence.c:128-140): c[len], one bit per byte */
int ref_sequence(uint32_t c_init, uint32_t len, uint8_t *c)
{
  srslte_sequence_t s;
  memset(&s, 0, sizeof(s));
  if (srslte_sequence_LTE_pr(&s, len, c_init)) return -1;
  memcpy(c, s.c, len);
  srslte_sequence_free(&s);
  return 0;
}

uint32_t ref_cinit_pdsch(uint32_t rnti, uint32_t q, uint32_t nslot, uint32_t cell_id)
{
  /* the seed srslte_sequence_pdsch hands to srslte_sequence_LTE_pr (sequences.c:64-66), read back by brute force would
   * be silly: the expression is evaluated by the reference itself through a 1-bit sequence compare in the tests; here
   * only the arguments are forwarded */
  srslte_sequence_t s, t;
  memset(&s, 0, sizeof(s)); memset(&t, 0, sizeof(t));
  srslte_sequence_pdsch(&s, (uint16_t)rnti, (int)q, nslot, cell_id, 64);
  const uint32_t guess = (rnti << 14) + (q << 13) + ((nslot / 2) << 9) + cell_id;
  srslte_sequence_LTE_pr(&t, 64, guess);
  const int same = memcmp(s.c, t.c, 64) == 0;
  srslte_sequence_free(&s); srslte_sequence_free(&t);
  return same ? guess : 0xFFFFFFFFu;
}

uint32_t ref_cinit_pusch(uint32_t rnti, uint32_t nslot, uint32_t cell_id)
{
  srslte_sequence_t s, t;
  memset(&s, 0, sizeof(s)); memset(&t, 0, sizeof(t));
  srslte_sequence_pusch(&s, (uint16_t)rnti, nslot, cell_id, 64);
  const uint32_t guess = (rnti << 14) + ((nslot / 2) << 9) + cell_id;
  srslte_sequence_LTE_pr(&t, 64, guess);
  const int same = memcmp(s.c, t.c, 64) == 0;
  srslte_sequence_free(&s); srslte_sequence_free(&t);
  return same ? guess : 0xFFFFFFFFu;
}

/* srslte_scrambling_s_offset (scrambling.c:49-52) with a freshly generated sequence */
int ref_descramble_s(uint32_t c_init, int16_t *llr, uint32_t len)
{
  srslte_sequence_t s;
  memset(&s, 0, sizeof(s));
  if (srslte_sequence_LTE_pr(&s, len, c_init)) return -1;
  srslte_scrambling_s_offset(&s, llr, 0, (int)len);
  srslte_sequence_free(&s);
  return 0;
}

/* csi_correction (pdsch.c:675-776), 16-bit LLRs */
int ref_csi_s(int mod_bits, int16_t *llr, float *csi, uint32_t nof_bits)
{
  static srslte_pdsch_t q;
  static srslte_pdsch_cfg_t cfg;
  memset(&q, 0, sizeof(q)); memset(&cfg, 0, sizeof(cfg));
  q.csi[0] = csi; q.llr_is_8bit = false; q.csi_enabled = true;
  cfg.grant.mcs[0].mod = mod_of(mod_bits);
  cfg.nbits[0].nof_bits = nof_bits;
  csi_correction(&q, &cfg, 0, 0, llr);
  return 0;
}

/* ---- the 8-bit LLR variants of the three stages (q->llr_is_8bit, pdsch.c:795-813) ---- */
int ref_demod_b(int mod_bits, const float *sym, int8_t *llr, int nsym)
{
  return srslte_demod_soft_demodulate_b(mod_of(mod_bits), (const cf_t *)sym, llr, nsym);
}

int ref_descramble_b(uint32_t c_init, int8_t *llr, uint32_t len)
{
  srslte_sequence_t s;
  memset(&s, 0, sizeof(s));
  if (srslte_sequence_LTE_pr(&s, len, c_init)) return -1;
  srslte_scrambling_sb_offset(&s, llr, 0, (int)len);
  srslte_sequence_free(&s);
  return 0;
}

int ref_csi_b(int mod_bits, int8_t *llr, float *csi, uint32_t nof_bits)
{
  static srslte_pdsch_t q;
  static srslte_pdsch_cfg_t cfg;
  memset(&q, 0, sizeof(q)); memset(&cfg, 0, sizeof(cfg));
  q.csi[0] = csi; q.llr_is_8bit = true; q.csi_enabled = true;
  cfg.grant.mcs[0].mod = mod_of(mod_bits);
  cfg.nbits[0].nof_bits = nof_bits;
  csi_correction(&q, &cfg, 0, 0, llr);
  return 0;
}

/* ---- the whole PDSCH codeword decode: srslte_pdsch_codeword_decode (pdsch.c:778-834) ---- */
typedef struct {
  srslte_pdsch_t          q;
  srslte_pdsch_cfg_t      cfg;
  srslte_softbuffer_rx_t  rx;
} ref_cw_t;

void *ref_cw_new(int nof_prb, uint32_t cell_id)
{
  ref_cw_t *t = calloc(1, sizeof(ref_cw_t));
  if (!t) return NULL;
  t->q.cell.id = cell_id; t->q.cell.nof_prb = (uint32_t)nof_prb; t->q.cell.nof_ports = 1;
  t->q.is_ue = true;
  t->q.users = calloc(1, sizeof(srslte_pdsch_user_t *));        /* no pre-generated sequences: tmp_seq path (pdsch.c:629) */
  t->q.d[0] = srslte_vec_malloc(sizeof(cf_t) * 20000);
  t->q.e[0] = srslte_vec_malloc(sizeof(int16_t) * 6 * 20000);
  t->q.csi[0] = srslte_vec_malloc(sizeof(float) * 20000);
  if (srslte_sequence_init(&t->q.tmp_seq, 6 * 20000)) return NULL;
  if (srslte_sch_init(&t->q.dl_sch)) return NULL;
  if (srslte_softbuffer_rx_init(&t->rx, (uint32_t)nof_prb)) return NULL;
  return t;
}

void ref_cw_free(void *p)
{
  ref_cw_t *t = p;
  if (!t) return;
  srslte_sch_free(&t->q.dl_sch);
  srslte_rm_turbo_gentables();
  srslte_softbuffer_rx_free(&t->rx);
  srslte_sequence_free(&t->q.tmp_seq);
  free(t->q.d[0]); free(t->q.e[0]); free(t->q.csi[0]); free(t->q.users);
  free(t);
}

void ref_cw_reset(void *p, uint32_t tbs) { srslte_softbuffer_rx_reset_tbs(&((ref_cw_t *)p)->rx, tbs); }

/* symbols: nof_re complex floats; csi NULL = disabled. Returns what pdsch.c:816-828 makes of the decode: 0 and *ack. e_out
 * (optional) receives the LLRs handed to srslte_dlsch_decode2. */
int ref_cw_decode(void *p, uint32_t tbs, int mod_bits, uint32_t rv, uint32_t nof_re, uint32_t rnti, uint32_t sf_idx,
                  const float *symbols, const float *csi, uint32_t max_its, uint8_t *data, int *ack, int16_t *e_out,
                  uint8_t *cb_crc_out)
{
  ref_cw_t *t = p;
  memset(&t->cfg, 0, sizeof(t->cfg));
  if (srslte_cbsegm(&t->cfg.cb_segm[0], tbs)) return -2;
  t->cfg.grant.tb_en[0] = true;
  t->cfg.grant.Qm[0] = (uint32_t)mod_bits;
  t->cfg.grant.mcs[0].tbs = (int)tbs;
  t->cfg.grant.mcs[0].mod = mod_of(mod_bits);
  t->cfg.nof_layers = 1;
  t->cfg.rv[0] = rv;
  t->cfg.sf_idx = sf_idx;
  t->cfg.nbits[0].nof_bits = nof_re * (uint32_t)mod_bits;
  t->cfg.nbits[0].nof_re = nof_re;
  memcpy(t->q.d[0], symbols, sizeof(cf_t) * nof_re);
  t->q.csi_enabled = csi != NULL;
  if (csi) memcpy(t->q.csi[0], csi, sizeof(float) * nof_re);
  t->q.llr_is_8bit = false;
  srslte_sch_set_max_noi(&t->q.dl_sch, max_its);
  bool a = false;
  int r = srslte_pdsch_codeword_decode(&t->q, &t->cfg, &t->q.dl_sch, &t->rx, (uint16_t)rnti, data, 0, 0, &a);
  if (ack) *ack = a;
  if (e_out) memcpy(e_out, t->q.e[0], sizeof(int16_t) * nof_re * (uint32_t)mod_bits);
  for (uint32_t i = 0; i < t->cfg.cb_segm[0].C && cb_crc_out; i++) cb_crc_out[i] = t->rx.cb_crc[i];
  return r;
}

/* ---- UL-SCH: the calls of srslte_pusch_decode from the soft demodulator on (pusch.c:617-654) ---- */
typedef struct {
  srslte_sch_t            sch;
  srslte_pusch_cfg_t      cfg;
  srslte_softbuffer_rx_t  rx;
  int16_t                *q_bits, *g_bits;
} ref_ul_t;

void *ref_ul_new(int nof_prb)
{
  ref_ul_t *t = calloc(1, sizeof(ref_ul_t));
  if (!t) return NULL;
  if (srslte_sch_init(&t->sch)) return NULL;
  if (srslte_softbuffer_rx_init(&t->rx, (uint32_t)nof_prb)) return NULL;
  t->q_bits = srslte_vec_malloc(sizeof(int16_t) * 6 * 20000);
  t->g_bits = srslte_vec_malloc(sizeof(int16_t) * 6 * 20000);
  return t;
}

void ref_ul_free(void *p)
{
  ref_ul_t *t = p;
  if (!t) return;
  srslte_sch_free(&t->sch);
  srslte_rm_turbo_gentables();
  srslte_softbuffer_rx_free(&t->rx);
  free(t->q_bits); free(t->g_bits);
  free(t);
}

void ref_ul_reset(void *p, uint32_t tbs) { srslte_softbuffer_rx_reset_tbs(&((ref_ul_t *)p)->rx, tbs); }

/* uci[0..2] = O_ack, O_ri (0/1), O_cqi; I_offset = {ack, ri, cqi}. out_uci[0..3] = ack, ack_2, ri, cqi_ack; cqi bits to
 * out_cqi. q_out / g_out (optional): LLRs after descrambling / after the channel de-interleaver. */
int ref_ul_decode(void *p, uint32_t tbs, int mod_bits, uint32_t rv, uint32_t L_prb, uint32_t nof_symb, uint32_t c_init,
                  const float *symbols, const uint32_t *uci_len, const uint32_t *I_offset, uint32_t max_its,
                  uint8_t *data, uint8_t *out_uci, uint8_t *out_cqi, int16_t *q_out, int16_t *g_out, uint8_t *cb_crc_out,
                  uint32_t *ack_pos_out, uint32_t *ri_pos_out, uint32_t *pos_counts)
{
  ref_ul_t *t = p;
  memset(&t->cfg, 0, sizeof(t->cfg));
  if (tbs && srslte_cbsegm(&t->cfg.cb_segm, tbs)) return -2;
  const uint32_t nof_re = L_prb * 12 * nof_symb;
  t->cfg.grant.Qm = (uint32_t)mod_bits;
  t->cfg.grant.mcs.tbs = (int)tbs;
  t->cfg.grant.mcs.mod = mod_of(mod_bits);
  t->cfg.grant.L_prb = L_prb;
  t->cfg.grant.M_sc = L_prb * 12;
  t->cfg.grant.M_sc_init = L_prb * 12;
  t->cfg.rv = rv;
  t->cfg.cp = SRSLTE_CP_NORM;
  t->cfg.nbits.nof_re = nof_re;
  t->cfg.nbits.nof_symb = nof_symb;
  t->cfg.nbits.nof_bits = nof_re * (uint32_t)mod_bits;
  t->cfg.uci_cfg.I_offset_ack = I_offset[0]; t->cfg.uci_cfg.I_offset_ri = I_offset[1]; t->cfg.uci_cfg.I_offset_cqi = I_offset[2];
  srslte_sch_set_max_noi(&t->sch, max_its);

  srslte_uci_data_t uci;
  memset(&uci, 0, sizeof(uci));
  uci.uci_ack_len = uci_len[0]; uci.uci_ri_len = uci_len[1]; uci.uci_cqi_len = uci_len[2];

  srslte_demod_soft_demodulate_s(t->cfg.grant.mcs.mod, (const cf_t *)symbols, t->q_bits, (int)nof_re);   /* pusch.c:620 */
  srslte_sequence_t seq;
  memset(&seq, 0, sizeof(seq));
  if (srslte_sequence_LTE_pr(&seq, t->cfg.nbits.nof_bits, c_init)) return -1;
  if (pos_counts) {
    /* which q positions does the reference treat as HARQ-ACK / RI?  Run its own routine on scratch copies with one kind
     * enabled at a time and read sch.ack_ri_bits back (entries still holding the sentinel were not written) */
    pos_counts[0] = pos_counts[1] = 0;
    int16_t *scratch = srslte_vec_malloc(sizeof(int16_t) * t->cfg.nbits.nof_bits);
    for (int kind = 0; kind < 2; kind++) {
      srslte_uci_data_t u2 = uci;
      if (kind == 0) u2.uci_ri_len = 0; else u2.uci_ack_len = 0;
      if ((kind == 0 && !u2.uci_ack_len) || (kind == 1 && !u2.uci_ri_len)) continue;
      for (int i = 0; i < 12 * 288; i++) t->sch.ack_ri_bits[i].position = 0xFFFFFFFFu;
      memcpy(scratch, t->q_bits, sizeof(int16_t) * t->cfg.nbits.nof_bits);
      if (srslte_ulsch_uci_decode_ri_ack(&t->sch, &t->cfg, &t->rx, scratch, seq.c, &u2)) return -1;
      uint32_t n = 0;
      while (n < 12 * 288 && t->sch.ack_ri_bits[n].position != 0xFFFFFFFFu) {
        (kind == 0 ? ack_pos_out : ri_pos_out)[n] = t->sch.ack_ri_bits[n].position;
        n++;
      }
      pos_counts[kind] = n;
    }
    free(scratch);
  }
  if (srslte_ulsch_uci_decode_ri_ack(&t->sch, &t->cfg, &t->rx, t->q_bits, seq.c, &uci)) return -1;       /* pusch.c:636 */
  srslte_scrambling_s_offset(&seq, t->q_bits, 0, (int)t->cfg.nbits.nof_bits);                            /* pusch.c:653 */
  if (q_out) memcpy(q_out, t->q_bits, sizeof(int16_t) * t->cfg.nbits.nof_bits);
  int r = srslte_ulsch_uci_decode(&t->sch, &t->cfg, &t->rx, t->q_bits, t->g_bits, data, &uci);           /* pusch.c:657 */
  if (g_out) memcpy(g_out, t->g_bits, sizeof(int16_t) * t->cfg.nbits.nof_bits);
  if (out_uci) { out_uci[0] = uci.uci_ack; out_uci[1] = uci.uci_ack_2; out_uci[2] = uci.uci_ri; out_uci[3] = uci.cqi_ack; }
  if (out_cqi) memcpy(out_cqi, uci.uci_cqi, sizeof(uci.uci_cqi));
  for (uint32_t i = 0; tbs && i < t->cfg.cb_segm.C && cb_crc_out; i++) cb_crc_out[i] = t->rx.cb_crc[i];
  srslte_sequence_free(&seq);
  return r;
}

/* ulsch_deinterleave alone (sch.c:860-881); ri_pos = q positions occupied by RI bits */
int ref_ulsch_deinterleave(const int16_t *q_bits, uint32_t Qm, uint32_t H_prime_total, uint32_t nof_symb, int16_t *g_bits,
                           const uint32_t *ri_pos, uint32_t nof_ri_bits)
{
  const uint32_t n = H_prime_total * Qm;
  srslte_uci_bit_t *ri = calloc(nof_ri_bits + 1, sizeof(srslte_uci_bit_t));
  uint8_t *present = calloc(n + 16, 1);
  uint32_t *lut = calloc(n + 16, sizeof(uint32_t));
  int16_t *q = srslte_vec_malloc(sizeof(int16_t) * (n + 16));
  memcpy(q, q_bits, sizeof(int16_t) * n);
  for (uint32_t i = 0; i < nof_ri_bits; i++) ri[i].position = ri_pos[i];
  ulsch_deinterleave(q, Qm, H_prime_total, nof_symb, g_bits, ri, nof_ri_bits, present, lut);
  free(ri); free(present); free(lut); free(q);
  return 0;
}

/* srslte_mod_modulate (mod.c) with the LTE tables (modem_table.c / lte_tables.c): bits one per byte -> cf_t */
int ref_modulate(int mod_bits, const uint8_t *bits, float *symbols, int nbits)
{
  srslte_modem_table_t tab;
  if (srslte_modem_table_lte(&tab, mod_of(mod_bits))) return -1;
  int n = srslte_mod_modulate(&tab, (uint8_t *)bits, (cf_t *)symbols, (uint32_t)nbits);
  srslte_modem_table_free(&tab);
  return n;
}
