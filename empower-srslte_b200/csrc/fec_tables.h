/*
 * fec_tables.h — host-side integer bookkeeping of the decode path (plain C, no device code):
 * code-block sizes and segmentation, decoder class per K, device data-layout geometry, and the
 * generators of the per-K lookup tables the kernels read from HBM.
 */
#ifndef FEC_TABLES_H
#define FEC_TABLES_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEC_NOF_K        188
#define FEC_MAX_K        6144
#define FEC_R24A_LEN     (13 * 6144 + 64)   /* CRC power table long enough for the largest transport block */

/* Device layout of one code block ("geometry"). The reference's sub-block-interleaved soft buffer
 * (rm_turbo.c:250-264: element k of sub-block d at k*P+d) becomes, on the device, NL contiguous lanes of
 * Lp int16 per stream so that one thread walks one sub-block with unit stride:
 *   stream s (0 syst, 1 parity, 2 parity'), lane d, step k  ->  s*SS + d*Lp + k ;  12 tail LLRs at 3*SS.
 * The extrinsic array E lives in shared memory with lane stride LpE (LpE/2 odd: conflict-free gathers). */
typedef struct {
  int K;      /* code block size */
  int idx;    /* index in the 188-entry table */
  int P;      /* decoder class as the reference's AUTO mode picks it: 0 (K<=400), 8, 16 */
  int NL;     /* lanes = P, or 1 for the non-windowed class */
  int L;      /* K / NL */
  int Lp;     /* lane stride in HBM, multiple of 8 */
  int LpE;    /* lane stride of E in shared memory */
  int SS;     /* stream stride NL*Lp */
  int f1, f2; /* QPP parameters */
} fec_geom_t;

int  fec_k_index(uint32_t K);            /* first table index with K_i >= K, -1 if none (cbsegm.c:116-127) */
int  fec_k_size(uint32_t idx);           /* K_i or -1 */
int  fec_k_subblocks(uint32_t K);        /* 16 / 8 / 0 (turbodecoder.c:364-376, AVX2 build) */
int  fec_segm(uint32_t tbs, uint32_t out9[9]);   /* cbsegm.c:58-108 */
int  fec_geom(uint32_t K, fec_geom_t *g);        /* -1 if K is not an LTE size */

/* step table, 2 parities x NL x Lp entries of (natural_position << 16 | E_address); pads = 0 */
void fec_gen_step_table(const fec_geom_t *g, uint32_t *tab);
/* rate-matching rank table in device soft-buffer layout (3*SS+16 entries, pad = 0xFFFF) and the 4 rv
 * start ranks: transmitted index of soft position o for rv is (rank[o] - off[rv]) mod (3K+12). */
void fec_gen_rm_rank(const fec_geom_t *g, uint16_t *rank, uint32_t off[4]);
/* natural coded position t (0..3K+11) -> device soft-buffer offset, and -> reference soft-buffer offset */
int  fec_soft_offset_dev(const fec_geom_t *g, int t);
int  fec_soft_offset_ref(const fec_geom_t *g, int t, int sb_layout);
/* R[m] = x^(m+24) mod g(x): CRC of a message is the XOR of R[nbits-1-i] over its set bits i */
void fec_gen_crc_powers(uint32_t poly, uint32_t *R, int n);

#ifdef __cplusplus
}
#endif
#endif
