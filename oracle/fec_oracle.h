/*
 * fec_oracle.h — CPU restatement of the reference's transport-block decode path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is product code: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg may load it, and only as the checker. The product library
 * (libsrslte_fec_cuda.so) never links or calls it and has no CPU fallback.
 *
 * Parity status: PINNED. Every function here is checked (tests/test_oracle_vs_ref.py, run wherever
 * /root/reference is mounted) against the unmodified reference objects in oracle/_ref, and on any box
 * against golden vectors under tests/golden/ that were produced by those objects (oracle/gen_golden.py).
 *
 * All arrays are in NATURAL bit order (index x = bit position in the code block); the reference's
 * sub-block-interleaved ("sb") layout is only an addressing convention (element k of sub-block d at
 * k*P+d) and is converted at the edges by orc_sb_pos().
 */
#ifndef FEC_ORACLE_H
#define FEC_ORACLE_H
#include <stdint.h>

#define ORC_MAX_K       6144
#define ORC_SOFT_SIZE   18600      /* reference softbuffer.h:55 */
#define ORC_MAX_CB      13

int      orc_cb_index(int K);                      /* cbsegm.c:116-127 */
int      orc_cb_size(int idx);                     /* cbsegm.c:134-141 */
int      orc_subblocks(int K);                     /* turbodecoder.c:364-376 (AVX2 build) */
int      orc_cbsegm(uint32_t tbs, uint32_t *out9); /* cbsegm.c:58-108 */
uint32_t orc_crc(int type, const uint8_t *bytes, int nbits);        /* crc.c:144-158, crc.h:68-80 */
int      orc_qpp(int K, uint16_t *fwd, uint16_t *rev);              /* tc_interl_lte.c:78-103 */
int      orc_qpp_sb(int K, int P, uint16_t *fwd, uint16_t *rev);    /* tc_interl_lte.c:104-112 */
int      orc_sb_pos(int x, int K, int P);                           /* natural x -> sb position */
int      orc_rm_table(int K, int rv, uint16_t *tab);                /* rm_turbo.c:164-238 (natural) */
int      orc_rm_soft_index(int t, int K, int P);                    /* rm_turbo.c:250-264 */
int      orc_rm_rx(const int16_t *in, int E, int K, int rv, int sb_layout, int16_t *soft); /* rm_turbo.c:378-430 */

/* instrumentation (thread-local): saturation clips, peak |intermediate|, peak |MAP input| since orc_dbg_reset() */
void orc_dbg_reset(void);
long orc_dbg_sat_count(void);
int  orc_dbg_peak_val(void);
int  orc_dbg_peak_in(void);
int  orc_dbg_mag(int phase);      /* largest |state metric| after a normalisation / at the start of a main pass (certificate input) */
int  orc_dbg_llr_range(void);     /* largest max-min of the LLR maxima m0/m1 over one sub-block (certificate input) */
int  orc_dbg_spread(int phase);   /* largest state-metric spread at a normalisation point: 0 beta warm-up, 1 beta, 2 alpha warm-up, 3 alpha */

/* one constituent MAP decoder run, natural order. tail_x/tail_p: the 3 tail LLRs of this decoder. */
void orc_map(int K, int P, const int16_t *x, const int16_t *app, const int16_t *par,
             const int16_t *tail_x, const int16_t *tail_p, int16_t *out);

typedef struct {
  int      K, P, n_iter, llr8;   /* llr8: int8 decoders of the 8-bit LLR path (values carried in int16) */
  int16_t  syst[ORC_MAX_K + 4], p0[ORC_MAX_K + 4], p1[ORC_MAX_K + 4];
  int16_t  t_syst[3], t_p0[3], t_app2[3], t_p1[3];
  int16_t  app1[ORC_MAX_K], app2[ORC_MAX_K], ext1[ORC_MAX_K], ext2[ORC_MAX_K];
  uint16_t fwd[ORC_MAX_K], rev[ORC_MAX_K];
} orc_tdec_t;

int  orc_tdec_new_cb_natural(orc_tdec_t *h, int K, const int16_t *in_nat);        /* [s p p']xK + 12 tail */
int  orc_tdec_new_cb_soft(orc_tdec_t *h, int K, const int16_t *soft);             /* reference soft-buffer layout */
void orc_tdec_iteration(orc_tdec_t *h, uint8_t *out_bytes, int16_t *llr_nat);     /* turbodecoder_iter.h:73-147 */
int  orc_tdec_halfits(int K, const int16_t *in_nat, int nof, uint8_t *out_bytes, int16_t *llr_nat);
int  orc_decode_cb(const int16_t *soft, int K, int max_its, int crc_type, uint8_t *out_bytes, int *nits);
/* 8-bit LLR path */
int  orc_subblocks8(int K);                                                         /* turbodecoder.c:392-406 */
int  orc_tdec8_new_cb_natural(orc_tdec_t *h, int K, const int8_t *in_nat);
int  orc_tdec8_new_cb_soft(orc_tdec_t *h, int K, const int8_t *soft);
int  orc_tdec_halfits8(int K, const int8_t *in_nat, int nof, uint8_t *out_bytes, int16_t *llr_nat);
int  orc_rm_rx8(const int8_t *in, int E, int K, int rv, int sb_layout, int8_t *soft);   /* rm_turbo.c:432-469 */

typedef struct {
  int      max_cb;
  int16_t  soft[ORC_MAX_CB][ORC_SOFT_SIZE];
  uint8_t  data[ORC_MAX_CB][ORC_MAX_K / 8];
  uint8_t  cb_crc[ORC_MAX_CB];
  uint8_t  tb_crc;
} orc_softbuffer_t;

void *orc_tb_new(void);
void  orc_tb_free(void *p);
void  orc_tb_reset(void *p, uint32_t tbs);                                        /* softbuffer.c:128-151 */
int   orc_tb_decode(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const int16_t *e_bits,
                    uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *cb_its); /* sch.c:307-498 */
int   orc_tb_decode8(void *p, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const int8_t *e_bits,
                     uint32_t max_its, uint8_t *data, uint8_t *cb_crc_out, uint32_t *cb_its);
void  orc_tb_get_soft(void *p, int cb, int16_t *out);

/* front end in front of decode_tb: see the section header in fec_oracle.c */
int      orc_demod_s(int mod_bits, const float *sym, int16_t *llr, int nsym);                 /* demod_soft.c:437-456 */
void     orc_sequence(uint32_t c_init, uint32_t len, uint8_t *c);                              /* sequence.c:54-86 */
uint32_t orc_cinit_pdsch(uint32_t rnti, uint32_t q, uint32_t nslot, uint32_t cell_id);         /* sequences.c:64-66 */
uint32_t orc_cinit_pusch(uint32_t rnti, uint32_t nslot, uint32_t cell_id);                     /* sequences.c:71-73 */
void     orc_descramble_s(const uint8_t *c, int16_t *llr, uint32_t len);                       /* scrambling.c:49-52 */
void     orc_csi_s(int mod_bits, int16_t *e, const float *csi, uint32_t nof_bits);             /* pdsch.c:675-776 */
void     orc_ulsch_deinterleave(const int16_t *q, uint32_t Qm, uint32_t Hp, uint32_t cols, int16_t *g,
                                const uint32_t *ri_pos, uint32_t nri);                         /* sch.c:860-881 */

/* transmit side: turbo encoder + rate matching of srslte_dlsch_encode / srslte_ulsch_encode's data part */
int      orc_encode_cb(int K, const uint8_t *bits, uint8_t *coded);                              /* turbocoder.c:82-165 */
int      orc_tb_encode(uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits, const uint8_t *data,
                       uint8_t *e_bits);                                                        /* sch.c:188-296 */

/* multithreaded timing harness for bench.py's cpu_baseline leg ("port") */
double orc_bench_tb_mt(int nthreads, int ntb, uint32_t tbs, uint32_t Qm, uint32_t rv, uint32_t nof_e_bits,
                       const int16_t *e_bits_all, uint32_t max_its, int reps, uint64_t *out3);
/* 8-bit LLR front end (pdsch.c:795-813 with q->llr_is_8bit) */
int      orc_demod_b(int mod_bits, const float *sym, int8_t *llr, int nsym);
void     orc_descramble_b(const uint8_t *c, int8_t *llr, uint32_t len);
void     orc_csi_b(int mod_bits, int8_t *e, const float *csi, uint32_t nof_bits);
#endif
